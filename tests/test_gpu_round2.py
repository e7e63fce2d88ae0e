"""GPU parity tests, second set (run on the B200 box): production kernels at production size against the unmodified
reference build, non-finite input, the reference module test's 216 block shapes, concurrent streams, pageable host memory
through the staged copies, the sharded C ABI, and the reference's own driver relinked against this library."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

import oracle_libs as ol
from cvxcompress_b200 import synth

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
BIN = os.path.join(HERE, "cabi", "bin")


@pytest.fixture()
def gpu():
    import torch
    from cvxcompress_b200 import _lib, api
    assert torch.cuda.is_available()
    _lib.lib()
    api.set_transform("exact")
    yield api
    api.set_transform("exact")


def _streams_equal(a, b):
    return len(a) == len(b) and bytes(a[:len(a) - 7]) == bytes(b[:len(b) - 7])


def _snr(v, w):
    import torch
    e = float(torch.sqrt(torch.sum((v.double() - w.double()) ** 2) / torch.sum(v.double() ** 2)).item())
    return -20.0 * np.log10(e)


# ---- the kernels the bench times, at the size the bench times them ------------------------------------------------
@pytest.mark.skipif(not ol.have_ref(), reason="oracle/_ref (reference build) did not travel")
@pytest.mark.timeout(600)
def test_lifting_production_kernels_against_reference_at_config3_size(gpu):
    """k_encode32ws / k_decode32ws (LIFTING, what bench.py times) on config-3 G2 data, against the UNMODIFIED reference on
    the first 320 planes of the same volume (1056x1248x320 = 421.7 MCells, compressed as a volume of its own by both):
    ratio within 0.1 %, SNR within 0.1 dB, the reference decodes the GPU stream, the GPU decodes the reference's
    all-threads (permuted block order) stream -- north_star's four correctness parts at production block counts
    (12 870 blocks, 87 hand-overs per CTA).  Then the full 960-plane volume: round trip at the same SNR."""
    import torch
    api = gpu
    nx, ny, nzs, nz = 1056, 1248, 320, 960
    threads = len(os.sched_getaffinity(0))
    d = synth.g2_torch(nx, ny, nzs, "cuda", z0=0, nz_total=nz)
    vol = d.cpu().numpy()
    api.set_transform("lifting")
    s_gpu, ln_gpu, _ = api.compress_device(d, 32, 32, 32, 1e-2)
    rec_gpu = api.decompress_device(s_gpu, ln_gpu, nx, ny, nzs)
    s_ref, _ = ol.ref_compress(vol, 32, 32, 32, 1e-2, threads=threads)
    rec_ref = torch.from_numpy(ol.ref_decompress(s_ref, nx, ny, nzs, threads=threads)).cuda()
    # ratio and SNR
    assert abs(ln_gpu - len(s_ref)) <= 1e-3 * len(s_ref), (ln_gpu, len(s_ref))
    snr_gpu, snr_ref = _snr(d, rec_gpu), _snr(d, rec_ref)
    assert abs(snr_gpu - snr_ref) < 0.1, (snr_gpu, snr_ref)
    vmax = float(d.abs().max().item())
    # the reference decodes the GPU stream: same reconstruction as the GPU's own decode up to the LIFTING tolerance
    h_gpu = s_gpu[:ln_gpu].cpu().numpy()
    ref_of_gpu = torch.from_numpy(ol.ref_decompress(h_gpu, nx, ny, nzs, threads=threads)).cuda()
    assert float((ref_of_gpu - rec_gpu).abs().max().item()) <= 2e-5 * vmax
    # the GPU decodes the reference's multi-thread stream
    sd = torch.zeros(len(s_ref) + 64, dtype=torch.uint8, device="cuda")
    sd[:len(s_ref)] = torch.from_numpy(s_ref).cuda()
    gpu_of_ref = api.decompress_device(sd, len(s_ref), nx, ny, nzs)
    assert float((gpu_of_ref - rec_ref).abs().max().item()) <= 2e-5 * vmax
    # EXACT mode on the same slab: byte-identical payload per block (the reference's order is permuted, its bytes are not)
    api.set_transform("exact")
    s_ex, ln_ex, _ = api.compress_device(d, 32, 32, 32, 1e-2)
    assert ln_ex == len(s_ref)
    pa, rawa, _ = ol.block_payloads(s_ex[:ln_ex].cpu().numpy())
    pb, rawb, _ = ol.block_payloads(s_ref)
    assert len(pa) == len(pb) and np.array_equal(rawa, rawb) and all(x == y for x, y in zip(pa, pb))
    del rec_gpu, rec_ref, ref_of_gpu, gpu_of_ref, sd
    # the whole volume
    api.set_transform("lifting")
    dfull = synth.g2_torch(nx, ny, nz, "cuda", z0=0, nz_total=nz)
    s, ln, _ = api.compress_device(dfull, 32, 32, 32, 1e-2)
    rec = api.decompress_device(s, ln, nx, ny, nz)
    assert abs(_snr(dfull, rec) - snr_ref) < 0.2           # different RMS (whole volume vs slab), same quantiser


# ---- out-of-range and non-finite input (Run_Length_Encode_Slow.cpp:203-205: cvttss2si gives 0x80000000) --------------
@pytest.mark.parametrize("mode", ["exact", "lifting"])
def test_out_of_range_values(gpu, mode):
    """Finite values whose scaled coefficients leave the int32 range: the quantiser's conversion saturates to 0x80000000
    (x86 cvttss2si) and the coder must fall through to VLESC4, carrying the scaled float itself."""
    import torch
    api = gpu
    nx, ny, nz = 96, 64, 64
    vol = synth.g2(nx, ny, nz).copy()
    vol[10, 10, 10] = 3e30
    vol[50, 60, 33] = -2.5e19
    vol[20:22, 8:16, 32:64] *= 1e15
    want, _ = ol.oracle_compress(vol, 32, 32, 32, 1e-2)
    toks = ol.parse_stream(want)
    d = torch.from_numpy(vol).cuda()
    if mode == "exact":
        s, ln, _ = api.compress_device(d, 32, 32, 32, 1e-2)
        assert _streams_equal(s[:ln].cpu().numpy(), want)
        rec = api.decompress_device(s, ln, nx, ny, nz).cpu().numpy()
        ref = ol.oracle_decompress(want, nx, ny, nz)
        assert np.array_equal(rec.view(np.uint32), ref.view(np.uint32))
    else:
        api.set_transform("lifting")
        s, ln, _ = api.compress_device(d, 32, 32, 32, 1e-2)
        assert abs(ln - len(want)) <= 0.01 * len(want)
        rec = api.decompress_device(s, ln, nx, ny, nz).cpu().numpy()
        ref = ol.oracle_decompress(s[:ln].cpu().numpy(), nx, ny, nz)       # the oracle decodes the GPU's stream
        assert np.all(np.isfinite(rec)) and np.max(np.abs(rec - ref)) <= 2e-5 * np.max(np.abs(ref))
    assert toks["nnn"] == 12


@pytest.mark.parametrize("mode", ["exact", "lifting"])
def test_nan_and_inf_input(gpu, mode):
    """NaN / Inf cells poison their own block (Inf - Inf inside the filter bank makes NaN) and change the global RMS to NaN,
    hence glob_mulfac to 1 (CvxCompress.cpp:291-295); the reference's driver asserts its input has none
    (Test_With_Generated_Input.cpp:66-68).  NaN bit patterns differ between x86 and the GPU, so byte identity is not
    defined for poisoned blocks; what must hold: no hang, same header, every clean block byte-identical, the poisoned
    cells non-finite at the same places in the GPU's and the oracle's decode of the GPU stream."""
    import torch
    api = gpu
    nx, ny, nz = 96, 64, 64
    vol = synth.g2(nx, ny, nz).copy()
    vol[3, 5, 7] = np.nan                # block (0, 0, 0)
    vol[40, 33, 90] = np.inf             # block (2, 1, 1)
    vol[41, 2, 1] = -np.inf              # block (0, 0, 1)
    poisoned = {0, 6 + 3 + 2, 6}
    want, _ = ol.oracle_compress(vol, 32, 32, 32, 1e-2)
    d = torch.from_numpy(vol).cuda()
    api.set_transform(mode)
    s, ln, _ = api.compress_device(d, 32, 32, 32, 1e-2)
    got = s[:ln].cpu().numpy()
    assert bytes(got[:32]) == bytes(want[:32])                      # header: glob_mulfac == 1.0
    if mode == "exact":
        pa, rawa, _ = ol.block_payloads(got)
        pb, rawb, _ = ol.block_payloads(want)
        for i in range(len(pa)):
            if i not in poisoned:
                assert pa[i] == pb[i] and rawa[i] == rawb[i], i
    rec = api.decompress_device(s, ln, nx, ny, nz).cpu().numpy()
    ref = ol.oracle_decompress(got, nx, ny, nz)                     # the oracle decodes the GPU's stream
    assert np.array_equal(np.isfinite(rec), np.isfinite(ref))
    ok = np.isfinite(ref)
    assert np.max(np.abs(rec[ok] - ref[ok])) <= (0 if mode == "exact" else 2e-5 * np.max(np.abs(ref[ok])))
    clean = np.s_[32:64, 32:64, 0:32]                                # a block without a poisoned cell stays finite
    assert np.all(np.isfinite(rec[clean]))


# ---- all 216 block shapes of the reference's module test (CvxCompress.cpp:710-774) -------------------------------------
@pytest.mark.timeout(900)
def test_all_216_block_shapes_stream_identical(gpu):
    """Every (bx, by, bz) in {8..256}^3 on one small volume, EXACT mode: the stream equals the oracle's byte for byte and
    decodes to the oracle's reconstruction bit for bit.  Crosses every dispatch boundary (32^3 kernels, shared-memory
    kernels for other shapes up to 32768 cells, large-block path)."""
    import torch
    api = gpu
    nx, ny, nz = 72, 40, 36
    vol = synth.g2(nx, ny, nz)
    d = torch.from_numpy(vol).cuda()
    sizes = (8, 16, 32, 64, 128, 256)
    bad = []
    for bz in sizes:
        for by in sizes:
            for bx in sizes:
                want, _ = ol.oracle_compress(vol, bx, by, bz, 1e-2)
                s, ln, _ = api.compress_device(d, bx, by, bz, 1e-2)
                got = s[:ln].cpu().numpy()
                if not _streams_equal(got, want):
                    bad.append(("stream", bx, by, bz))
                    continue
                rec = api.decompress_device(s, ln, nx, ny, nz).cpu().numpy()
                ref = ol.oracle_decompress(want, nx, ny, nz)
                if not np.array_equal(rec.view(np.uint32), ref.view(np.uint32)):
                    bad.append(("decode", bx, by, bz))
    assert not bad, bad


# ---- two streams of one host thread at once (the default workspace is per stream) ----------------------------
def test_two_cuda_streams_concurrently(gpu):
    import torch
    api = gpu
    vols = [synth.g2(160, 96, 128), synth.g3(160 * 96 * 128, seed=7, lo=-1.0, hi=1.0).reshape(128, 96, 160)]
    want = [ol.oracle_compress(v, 32, 32, 32, 1e-2)[0] for v in vols]
    dv = [torch.from_numpy(v).cuda() for v in vols]
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    cap = [api.max_compressed_bytes(160, 96, 128, 32, 32, 32) + 8 for _ in vols]
    outs = [torch.empty(c, dtype=torch.uint8, device="cuda") for c in cap]
    lens = [torch.zeros(1, dtype=torch.int64, device="cuda") for _ in vols]
    stat = [torch.zeros(1, dtype=torch.int32, device="cuda") for _ in vols]
    torch.cuda.synchronize()
    for rep in range(6):
        for i in (0, 1):
            with torch.cuda.stream(streams[i]):
                for _ in range(3):       # several calls back to back per stream, interleaved with the other stream's
                    api.compress_device_async(dv[i], 32, 32, 32, 1e-2, False, outs[i], lens[i], stat[i])
        torch.cuda.synchronize()
        for i in (0, 1):
            ln = int(lens[i].item())
            assert int(stat[i].item()) == 0
            assert _streams_equal(outs[i][:ln].cpu().numpy(), want[i]), (rep, i)


# ---- pageable host memory through the staged copies (cvx_stage.cpp) -------------------------------------------------
@pytest.mark.skipif(not ol.have_ref(), reason="oracle/_ref (reference build) did not travel")
def test_pageable_host_buffers_through_class_api(gpu):
    """What a caller of the reference does (Test_With_Generated_Input.cpp:42-53): plain malloc'd buffers.  48 MB volume,
    so the multi-threaded bounce-buffer path runs (>= 4 MiB), both directions; result byte-identical to the reference
    at num_threads = 1 (natural block order) and to the device-resident path."""
    from cvxcompress_b200 import _lib
    L = _lib.lib()
    nx, ny, nz = 352, 208, 164                      # nz not a multiple of the block size; 12.0 M cells
    vol = synth.g2(nx, ny, nz)
    want, _ = ol.ref_compress(vol, 32, 32, 32, 1e-2, threads=1)
    comp = np.zeros(vol.size, np.uint8)             # N bytes, as the reference's driver sizes it
    ln = C.c_long(0)
    L.cvx_compress_th(1e-2, vol.ctypes.data, nx, ny, nz, 32, 32, 32, False, comp.ctypes.data, 4, C.byref(ln))
    assert _streams_equal(comp[:ln.value], want)
    assert not comp[ln.value - 7:ln.value + 64].any()        # the 7 slack bytes and everything after stay untouched
    out = np.full(vol.size, 7.0, np.float32)
    L.cvx_decompress_inplace_th(out.ctypes.data, nx, ny, nz, comp.ctypes.data, 4, ln.value)
    ref = ol.ref_decompress(want, nx, ny, nz)
    assert np.array_equal(out.view(np.uint32), ref.reshape(-1).view(np.uint32))
    # compressed_length is never read by the reference's Decompress: 0 must work too (CvxCompress.cpp:463-571)
    out2 = np.zeros(vol.size, np.float32)
    L.cvx_decompress_inplace_th(out2.ctypes.data, nx, ny, nz, comp.ctypes.data, 4, 0)
    assert np.array_equal(out2.view(np.uint32), out.view(np.uint32))


# ---- the sharded C ABI, driven from C++ without torch or MPI --------------------------------------------------------
def _driver(*args):
    exe = os.path.join(BIN, "shard_driver")
    if not os.path.exists(exe):
        pytest.skip("tests/cabi/bin/shard_driver not built (python -c 'import __graft_entry__ as g; g.build()')")
    r = subprocess.run([exe] + [str(a) for a in args], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    return r.returncode, r.stdout


@pytest.mark.parametrize("case", [(96, 80, 200, 32, 32, 32, 0.01, 0, "exact"), (96, 80, 200, 32, 32, 32, 0.01, 0, "lifting"),
                                  (100, 60, 150, 16, 16, 64, 0.01, 1, "exact"), (70, 50, 90, 8, 16, 8, 1e-7, 0, "exact")],
                         ids=lambda c: "%dx%dx%d-b%dx%dx%d-s%g-l%d-%s" % c)
def test_sharded_c_abi_world1(gpu, case):
    rc, out = _driver(1, *case)
    assert rc == 0 and "OK" in out, out


@pytest.mark.parametrize("world", [2, 4, 8])
def test_sharded_c_abi_multi_gpu(gpu, world):
    """One process per GPU (fork, NCCL unique id through a file): every rank's header / table slice / payload shard equals
    the corresponding bytes of the single-GPU stream, for every world size."""
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip("%d GPUs needed" % world)
    for case in [(96, 80, 320, 32, 32, 32, 0.01, 0, "exact"), (96, 80, 320, 32, 32, 32, 0.01, 0, "lifting"),
                 (100, 60, 600, 16, 16, 64, 0.01, 1, "exact")]:
        rc, out = _driver(world, *case)
        assert rc == 0 and "ALL OK" in out, out


def test_sharded_python_binding_world1(gpu):
    """sharded.compress_slab / decompress_slab over the C entry points (world 1: no NCCL): same stream as the whole-volume call."""
    import torch
    from cvxcompress_b200 import sharded
    api = gpu
    vol = synth.g2(96, 64, 100)
    d = torch.from_numpy(vol).cuda()
    ss = sharded.compress_slab(d, 96, 64, 100, 32, 32, 32, 1e-2)
    full = sharded.assemble(ss).cpu().numpy()
    s, ln, _ = api.compress_device(d, 32, 32, 32, 1e-2)
    assert ss.total_length() == ln and _streams_equal(full, s[:ln].cpu().numpy())
    rec = sharded.decompress_slab(ss)
    assert torch.equal(rec, api.decompress_device(s, ln, 96, 64, 100))
    # a payload buffer that is too small is reported, not silently truncated
    tiny = torch.empty(1024, dtype=torch.uint8, device="cuda")
    bad = sharded.compress_slab(d, 96, 64, 100, 32, 32, 32, 1e-2, payload=tiny)
    with pytest.raises(RuntimeError):
        sharded.assemble(bad)
    # noise at a tiny scale compresses worse than 4:1 (raw blocks): the default payload buffer must hold it
    noise = torch.from_numpy(synth.g3(96 * 64 * 100, seed=3, lo=-1.0, hi=1.0).reshape(100, 64, 96)).cuda()
    ssn = sharded.compress_slab(noise, 96, 64, 100, 32, 32, 32, 1e-9)
    want, _ = ol.oracle_compress(noise.cpu().numpy(), 32, 32, 32, 1e-9)
    assert _streams_equal(sharded.assemble(ssn).cpu().numpy(), want)


# ---- the reference's own driver, relinked ------------------------------------------------
@pytest.mark.timeout(900)
def test_reference_driver_relinked_against_this_library(gpu):
    """tests/cabi/bin/ref_Test_With_Generated_Input = /root/reference/Test_With_Generated_Input.cpp, unmodified, compiled
    against include/CvxCompress.hxx and linked with libcvxcompress.so.  Its own asserts (error < 2e-4, SNR > 75 dB,
    Test_With_Generated_Input.cpp:121-122) must hold and it must print the README's compressed lengths."""
    exe = os.path.join(BIN, "ref_Test_With_Generated_Input")
    if not os.path.exists(exe):
        pytest.skip("tests/cabi/bin/ref_Test_With_Generated_Input not built (needs /root/reference at build time)")
    env = dict(os.environ)
    env.pop("CVX_GPU_TRANSFORM", None)               # the drop-in default: EXACT
    r = subprocess.run([exe], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=800, env=env)
    assert r.returncode == 0, r.stdout[-2000:]
    lens = [int(x) for x in re.findall(r"compressed length in bytes = (\d+)", r.stdout)]
    assert lens == [178789, 1430039, 4826289], r.stdout[-2000:]
    snrs = [float(x) for x in re.findall(r"SNR = ([0-9.]+) dB", r.stdout)]
    assert len(snrs) == 3 and all(abs(s - 75.6) < 0.05 for s in snrs), snrs


# ---- large blocks (64^3 .. 256^3 and anisotropic shapes with a 128 / 256 side): cvx_bigblock.inl ----------------------
@pytest.mark.timeout(600)
@pytest.mark.parametrize("shape", [(64, 64, 64), (128, 128, 128), (256, 256, 256), (256, 8, 16), (8, 128, 64), (128, 128, 1), (16, 64, 256)],
                         ids=lambda s: "%dx%dx%d" % s)
def test_large_blocks_full_pipeline(gpu, shape):
    """Cubes the reference publishes throughput for (README.md:38-40) and shapes whose lines run in shared memory
    (128, 256): EXACT byte-identical to the oracle; LIFTING decodes with the oracle to the same volume within the LIFTING
    tolerance, at the same ratio and SNR."""
    import torch
    api = gpu
    bx, by, bz = shape
    nx, ny, nz = 200, 136, (1 if bz == 1 else 144)
    if bx * by * bz >= 128 ** 3:
        nx, ny, nz = 300, 260, 270                      # 2 x 2 x 2 (128^3) or 2 x 2 x 2 partial (256^3) blocks
    vol = synth.g2(nx, ny, nz)
    d = torch.from_numpy(vol).cuda()
    want, _ = ol.oracle_compress(vol, bx, by, bz, 1e-2)
    ref = ol.oracle_decompress(want, nx, ny, nz)
    s, ln, _ = api.compress_device(d, bx, by, bz, 1e-2)
    assert _streams_equal(s[:ln].cpu().numpy(), want)
    rec = api.decompress_device(s, ln, nx, ny, nz).cpu().numpy()
    assert np.array_equal(rec.view(np.uint32), ref.view(np.uint32))
    api.set_transform("lifting")
    s2, ln2, _ = api.compress_device(d, bx, by, bz, 1e-2)
    assert abs(ln2 - ln) <= 0.005 * ln
    rec2 = api.decompress_device(s2, ln2, nx, ny, nz).cpu().numpy()
    ref2 = ol.oracle_decompress(s2[:ln2].cpu().numpy(), nx, ny, nz)
    assert np.max(np.abs(rec2 - ref2)) <= 2e-5 * np.max(np.abs(vol))
    e1, snr1 = synth.quality(vol, ref)
    e2, snr2 = synth.quality(vol, rec2)
    assert abs(snr1 - snr2) < 0.1
    # local RMS through the same path
    s3, ln3, _ = api.compress_device(d, bx, by, bz, 1e-2, True)
    ref3 = ol.oracle_decompress(s3[:ln3].cpu().numpy(), nx, ny, nz)
    rec3 = api.decompress_device(s3, ln3, nx, ny, nz).cpu().numpy()
    assert np.max(np.abs(rec3 - ref3)) <= 2e-5 * np.max(np.abs(vol))
    # the many-CTA coder of the large-block path (k_stage_measure / plan / emit): EXACT streams byte-identical with local RMS,
    # and on noise at a tiny scale (wide tokens, raw blocks), for both RMS kinds
    api.set_transform("exact")
    rng = np.random.default_rng(11)
    noisy = (vol + rng.standard_normal(vol.shape).astype(np.float32) * np.float32(0.3)).astype(np.float32)
    dn = torch.from_numpy(noisy).cuda()
    combos = ((vol, d, 1e-2, True), (noisy, dn, 1e-3, False), (noisy, dn, 1e-7, False), (noisy, dn, 1e-7, True))
    if bx * by * bz >= 128 ** 3:
        combos = combos[3:4]                            # the 21 MCell volume: one oracle run instead of four
    for v, dv, scale, local in combos:
        w, _ = ol.oracle_compress(v, bx, by, bz, scale, local)
        s4, ln4, _ = api.compress_device(dv, bx, by, bz, scale, local)
        assert _streams_equal(s4[:ln4].cpu().numpy(), w), (shape, scale, local)
        rec4 = api.decompress_device(s4, ln4, nx, ny, nz).cpu().numpy()
        assert np.array_equal(rec4.view(np.uint32), ol.oracle_decompress(w, nx, ny, nz).view(np.uint32)), (shape, scale, local)


@pytest.mark.parametrize("shape", [(64, 64, 64), (128, 128, 128), (256, 256, 256), (256, 16, 8)], ids=lambda s: "%dx%dx%d" % s)
@pytest.mark.parametrize("mode", ["exact", "lifting"])
def test_large_block_wavelet_only(gpu, shape, mode):
    """cvxgpu_wavelet_blocks_device on large dense blocks: EXACT bit-identical to the oracle's transform, LIFTING within
    1e-5 rms-relative (the reference's own Fast-vs-Slow criterion, CvxCompress.cpp:597); forward + inverse returns the input."""
    import torch
    api = gpu
    bx, by, bz = shape
    n = bx * by * bz
    nblk = 3 if n <= 128 ** 3 else 2
    src = synth.g3(n * nblk, seed=11, lo=-1.0, hi=1.0)
    want = src.copy().reshape(nblk, n)
    for k in range(nblk):
        ol.oracle().cvxo_wavelet_forward(want[k], bx, by, bz, 0)
    api.set_transform(mode)
    d = torch.from_numpy(src).cuda().view(nblk, bz, by, bx)
    api.wavelet_blocks_device(d, bx, by, bz, inverse=False)
    got = d.cpu().numpy().reshape(nblk, n)
    if mode == "exact":
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    else:
        rel = np.sqrt(np.mean((got.astype(np.float64) - want) ** 2) / np.mean(want.astype(np.float64) ** 2))
        assert rel < 1e-5, rel
    api.wavelet_blocks_device(d, bx, by, bz, inverse=True)
    back = d.cpu().numpy().reshape(-1)
    assert np.max(np.abs(back - src)) < 1e-5


@pytest.mark.parametrize("nx", [96, 101, 102, 70])
@pytest.mark.parametrize("shift", [0, 1])
def test_ws_decoder_row_stores_any_width_and_alignment(gpu, nx, shift):
    """k_decode32ws stores rows straight from registers (Copy_From_Block clipping, Block_Copy.cpp:136-212): odd nx, an odd
    plane size, a clipped last column and a volume that is only 4-byte aligned must all equal the fused decoder bit for bit
    and leave the cells around the volume untouched."""
    import torch
    from cvxcompress_b200 import _lib
    api = gpu
    L = _lib.lib()
    ny, nz = 45, 70
    api.set_transform("lifting")
    vol = torch.from_numpy(synth.g2(nx, ny, nz)).cuda()
    s, ln, _ = api.compress_device(vol, 32, 32, 32, 1e-2)
    n = nx * ny * nz
    outs = []
    for ws in (1, 0):
        L.cvx_debug_set_ws(ws)
        buf = torch.full((n + 8,), 123.0, dtype=torch.float32, device="cuda")
        out = buf[shift:shift + n].view(nz, ny, nx)
        api.decompress_device(s, ln, nx, ny, nz, out=out)
        torch.cuda.synchronize()
        assert float(buf[:shift].sum().item()) == 123.0 * shift and torch.all(buf[shift + n:] == 123.0)
        outs.append(out.clone())
    L.cvx_debug_set_ws(1)
    assert torch.equal(outs[0], outs[1])
    assert _snr(vol, outs[0]) > 40


# ---- the sparse-coder / windowed-parser kernels of the 8192 .. 32768-cell shapes (k_encode_sparse, k_decode_win) ----------
@pytest.mark.timeout(600)
@pytest.mark.parametrize("shape", [(16, 16, 64), (64, 16, 16), (32, 16, 16), (64, 32, 16)],
                         ids=lambda s: "%dx%dx%d" % s)
@pytest.mark.parametrize("local", [False, True], ids=["global", "local"])
def test_mid_size_shapes_sparse_dense_raw_and_foreign_order(gpu, shape, local):
    """BASELINE config 4's block shape and its siblings, many blocks per CTA (ragged volume: partial blocks on every side):
    EXACT streams byte-identical to the oracle and reconstructions bit-identical -- on smooth data (sparse list coder), on
    data + noise at a small scale (list overflow -> dense coder, long payloads -> several parser windows, raw blocks), and,
    where the reference build travelled, on the reference's multi-threaded (permuted block order) stream; LIFTING within the
    quality tolerance of EXACT."""
    import torch
    api = gpu
    bx, by, bz = shape
    nx, ny, nz = 168, 150, 132
    rng = np.random.default_rng(7)
    smooth = synth.g2(nx, ny, nz)
    noisy = (smooth + rng.standard_normal(smooth.shape).astype(np.float32) * np.float32(0.3)).astype(np.float32)
    for vol, scale in ((smooth, 1e-2), (noisy, 1e-3), (noisy, 1e-7)):
        d = torch.from_numpy(vol).cuda()
        want, _ = ol.oracle_compress(vol, bx, by, bz, scale, local)
        s, ln, _ = api.compress_device(d, bx, by, bz, scale, local)
        got = s[:ln].cpu().numpy()
        assert _streams_equal(got, want), (shape, local, scale)
        rec = api.decompress_device(s, ln, nx, ny, nz).cpu().numpy()
        ref = ol.oracle_decompress(want, nx, ny, nz)
        assert np.array_equal(rec.view(np.uint32), ref.view(np.uint32)), (shape, local, scale)
        if ol.have_ref():
            s_ref, _ = ol.ref_compress(vol, bx, by, bz, scale, local, threads=4)
            rec2 = api.decompress_device(torch.from_numpy(np.ascontiguousarray(s_ref)).cuda(), len(s_ref), nx, ny, nz).cpu().numpy()
            assert np.array_equal(rec2.view(np.uint32), ref.view(np.uint32)), (shape, local, scale, "foreign order")
    # LIFTING: same quality, decodable by the oracle
    d = torch.from_numpy(smooth).cuda()
    s0, ln0, _ = api.compress_device(d, bx, by, bz, 1e-2, local)
    _, snr0 = synth.quality(smooth, api.decompress_device(s0, ln0, nx, ny, nz).cpu().numpy())
    api.set_transform("lifting")
    s1, ln1, _ = api.compress_device(d, bx, by, bz, 1e-2, local)
    rec1 = api.decompress_device(s1, ln1, nx, ny, nz).cpu().numpy()
    ref1 = ol.oracle_decompress(s1[:ln1].cpu().numpy(), nx, ny, nz)
    api.set_transform("exact")
    _, snr1 = synth.quality(smooth, rec1)
    assert abs(snr1 - snr0) < 0.1 and abs(ln1 - ln0) <= 0.005 * ln0
    assert float(np.max(np.abs(rec1 - ref1))) <= 2e-5 * float(np.max(np.abs(smooth)))

"""Pins the CPU restatement (oracle/cvx_oracle.c) against
  (a) the reference's own golden numbers (README.md:72-99, Test_With_Generated_Input.cpp:121-122), and
  (b) the unmodified reference built into oracle/_ref (when present): stage by stage and end to end.
No GPU, no product code."""
import ctypes as C

import numpy as np
import pytest

import oracle_libs as ol
from cvxcompress_b200 import synth

needs_ref = pytest.mark.skipif(not ol.have_ref(), reason="oracle/_ref not built (needs /root/reference)")

SHAPES = [(8, 8, 8), (16, 16, 16), (32, 32, 32), (16, 16, 64), (64, 8, 16), (8, 32, 1), (32, 32, 1),
          (64, 64, 8), (128, 8, 8), (8, 8, 256), (256, 8, 1)]


def test_golden_c1_length_and_snr():
    """Config 1: 352x416x320, 32^3, scale 1e-2, global RMS -> 178789 bytes, error 1.663531e-04, 75.6 dB."""
    nx, ny, nz = 352, 416, 320
    vol = synth.g1(nx, ny, nz)
    stream, ratio = ol.oracle_compress(vol, 32, 32, 32, 1e-2)
    assert len(stream) == 178789
    assert abs(ratio - 1048.35) < 0.01
    s = ol.parse_stream(stream)
    assert abs(s["mulfac"] - 141.421356) < 1e-3
    out = ol.oracle_decompress(stream, nx, ny, nz)
    err, snr = synth.quality(vol, out)
    assert abs(err - 1.663531e-04) < 5e-10 * 1e3
    assert err < 2e-4 and snr > 75.0            # the reference's own asserts
    assert abs(snr - 75.6) < 0.05
    pays, raw, _ = ol.block_payloads(stream)
    assert all(len(p) == 117 for p in pays) and not raw.any()


@needs_ref
@pytest.mark.parametrize("shape", SHAPES)
def test_wavelet_matches_reference_fast_bitexact(shape):
    bx, by, bz = shape
    n = bx * by * bz
    blk = ol.aligned_f32(n)
    blk[:] = synth.g3(n, seed=7 + n, lo=-1.0, hi=1.0)
    tmp = ol.aligned_f32(8 * 256 + 64)
    a = blk.copy()
    ol.oracle().cvxo_wavelet_forward(a, bx, by, bz, 0)
    r = ol.aligned_f32(n); r[:] = blk
    ol.ref().fwd_fast(r.ctypes.data, tmp.ctypes.data, bx, by, bz)
    assert np.array_equal(a.view(np.uint32), r.view(np.uint32))
    ol.oracle().cvxo_wavelet_inverse(a, bx, by, bz, 0)
    ol.ref().inv_fast(r.ctypes.data, tmp.ctypes.data, bx, by, bz)
    assert np.array_equal(a.view(np.uint32), r.view(np.uint32))
    assert np.abs(a - blk).max() < 2e-6


@needs_ref
@pytest.mark.parametrize("shape", [(8, 8, 8), (32, 32, 32), (16, 16, 64)])
def test_wavelet_slow_association_matches_reference_slow(shape):
    bx, by, bz = shape
    n = bx * by * bz
    blk = synth.g3(n, seed=11, lo=-1.0, hi=1.0)
    a = blk.copy()
    ol.oracle().cvxo_wavelet_forward(a, bx, by, bz, 1)
    r = blk.copy()
    work = np.zeros(1024, dtype=np.float32)
    ol.ref().fwd_slow(r, work, bx, by, bz, 0, 0, 0, bx, by, bz)
    assert np.array_equal(a.view(np.uint32), r.view(np.uint32))
    ol.oracle().cvxo_wavelet_inverse(a, bx, by, bz, 1)
    ol.ref().inv_slow(r, work, bx, by, bz, 0, 0, 0, bx, by, bz)
    assert np.array_equal(a.view(np.uint32), r.view(np.uint32))


def _rle_cases():
    rng = np.random.default_rng(5)
    n = 4096
    cases = {}
    cases["zeros"] = np.zeros(n, np.float32)
    v = np.zeros(n, np.float32); v[[0, 9, 300, 301, 4095]] = [3, -7, 200, -40000, 1e7]; cases["sparse"] = v
    cases["bytes"] = rng.integers(-124, 125, n).astype(np.float32)
    cases["shorts"] = rng.integers(-32768, 32768, n).astype(np.float32)
    cases["int3"] = rng.integers(-8388608, 8388608, n).astype(np.float32)
    cases["floats"] = (rng.standard_normal(n) * 1e9).astype(np.float32)
    mix = rng.standard_normal(n).astype(np.float32) * np.exp(rng.uniform(-3, 12, n)).astype(np.float32)
    mix[rng.random(n) < 0.5] = 0; cases["mixed"] = mix
    esc = rng.choice(np.array([125, 126, 127, -125, -126, -127, -128, 124, -124, 255, 256, 257], np.float32), n)
    esc[rng.random(n) < 0.3] = 0; cases["escapes"] = esc
    runs = np.zeros(n, np.float32)
    pos = 0
    for r in [1, 2, 3, 254, 255, 256, 257, 1000]:
        pos += r; runs[pos] = 5; pos += 1
    cases["runlens"] = runs
    nanv = mix.copy(); nanv[5] = np.nan; nanv[77] = np.inf; nanv[78] = -np.inf; nanv[100] = 3e9; nanv[101] = -2147483648.0
    cases["nonfinite"] = nanv
    g = rng.integers(1, 125, n).astype(np.float32); g[::8] = 300; g[1::8] = 301; g[2::8] = -400; g[3::8] = 4000
    cases["mode_s8_edge"] = g        # 4 bytes + 4 shorts: nb + 3*(8-nb) = 16 -> elementwise
    g2 = g.copy(); g2[4::8] = 999; cases["mode_s8"] = g2     # 3 bytes + 5 shorts: 18 > 17 -> VLESC2_8x
    g3 = g2.copy(); g3[5::8] = 70000; cases["mode_t8"] = g3
    return cases


@needs_ref
@pytest.mark.parametrize("name", list(_rle_cases().keys()))
@pytest.mark.parametrize("mulfac", [1.0, 0.37, 141.421356])
def test_rle_matches_reference(name, mulfac):
    vals = _rle_cases()[name]
    n = len(vals)
    av = ol.aligned_f32(n); av[:] = vals
    dst_ref = ol.aligned_u8(5 * n + 64)
    pos = C.c_int(0)
    ol.ref().rle_encode(np.float32(mulfac), av.ctypes.data, n, dst_ref.ctypes.data, C.byref(pos))
    dst = np.zeros(5 * n + 64, np.uint8)
    m = ol.oracle().cvxo_rle_encode(np.float32(mulfac), av, n, dst)
    assert m == pos.value
    assert bytes(dst[:m]) == bytes(dst_ref[:m])
    out = np.zeros(n, np.float32)
    used = ol.oracle().cvxo_rle_decode(np.float32(mulfac), out, n, dst)
    assert used == m
    out_ref = np.zeros(n, np.float32)
    ol.ref().rle_decode(np.float32(mulfac), out_ref, n, dst_ref.ctypes.data)
    assert np.array_equal(out.view(np.uint32), out_ref.view(np.uint32))


CONFIGS = [  # nx, ny, nz, bx, by, bz, scale, local, generator
    (96, 64, 64, 32, 32, 32, 1e-2, False, "g2"),
    (151, 101, 51, 32, 32, 32, 1e-2, False, "g2"),        # partial blocks (Read_Raw_Volume size)
    (151, 101, 51, 16, 16, 64, 1e-2, True, "g2"),         # anisotropic + local RMS
    (37, 41, 43, 8, 8, 8, 1e-3, False, "g2"),             # nx not a multiple of 4
    (160, 96, 5, 32, 32, 1, 1e-2, False, "g2"),           # bz == 1 (2-D mode)
    (64, 64, 64, 32, 32, 32, 1e-2, False, "zero"),        # all-zero volume
    (64, 64, 64, 32, 32, 32, 1e-8, False, "noise"),       # raw fallback
    (64, 64, 64, 32, 32, 32, 1e-5, True, "noise"),        # low ratio, local
    (70, 64, 64, 64, 64, 64, 1e-2, False, "g2"),          # 64^3 blocks
]


def _make(gen, nx, ny, nz):
    if gen == "g2":
        return synth.g2(nx, ny, nz)
    if gen == "zero":
        return np.zeros((nz, ny, nx), np.float32)
    return synth.g3(nx * ny * nz, seed=3, lo=-1.0, hi=1.0).reshape(nz, ny, nx)


@needs_ref
@pytest.mark.parametrize("cfg", CONFIGS)
def test_stream_matches_reference_bytewise(cfg):
    nx, ny, nz, bx, by, bz, scale, local, gen = cfg
    vol = _make(gen, nx, ny, nz)
    s_ref, r_ref = ol.ref_compress(vol, bx, by, bz, scale, local, threads=1)
    s_or, r_or = ol.oracle_compress(vol, bx, by, bz, scale, local)
    assert len(s_ref) == len(s_or)
    n = len(s_ref) - 7     # the 7 slack bytes are never written by either side
    assert bytes(s_ref[:n]) == bytes(s_or[:n])
    assert r_ref == pytest.approx(r_or, rel=1e-6)
    a = ol.ref_decompress(s_ref, nx, ny, nz, threads=2)
    b = ol.oracle_decompress(s_ref, nx, ny, nz)
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    # the reference with several threads permutes block order only
    s_mt, _ = ol.ref_compress(vol, bx, by, bz, scale, local, threads=4)
    p1, raw1, _ = ol.block_payloads(s_or)
    p2, raw2, _ = ol.block_payloads(s_mt)
    assert p1 == p2 and np.array_equal(raw1, raw2)
    c = ol.oracle_decompress(s_mt, nx, ny, nz)
    assert np.array_equal(a.view(np.uint32), c.view(np.uint32))


@needs_ref
def test_copy_and_rms_match_reference():
    nx, ny, nz = 35, 37, 39
    vol = synth.g2(nx, ny, nz)
    for (bx, by, bz, x0, y0, z0) in [(32, 32, 32, 0, 0, 0), (32, 32, 32, 32, 32, 32), (8, 16, 8, 32, 32, 32), (16, 8, 32, 24, 30, 8)]:
        n = bx * by * bz
        a = np.full(n, 9.0, np.float32)
        ol.oracle().cvxo_copy_to_block(vol.reshape(-1), x0, y0, z0, nx, ny, nz, a, bx, by, bz)
        r = ol.aligned_f32(n); r[:] = 9.0
        ol.ref().copy_to_block(vol.reshape(-1), x0, y0, z0, nx, ny, nz, r.ctypes.data, bx, by, bz)
        assert np.array_equal(a, r)
        va = np.full(nx * ny * nz, -1.0, np.float32); vr = va.copy()
        ol.oracle().cvxo_copy_from_block(a, bx, by, bz, va, x0, y0, z0, nx, ny, nz)
        ol.ref().copy_from_block(r.ctypes.data, bx, by, bz, vr, x0, y0, z0, nx, ny, nz)
        assert np.array_equal(va, vr)
    # global RMS vs naive double loop, rel 1e-5 (CvxCompress.cpp:1101-1131 uses 37x41x43)
    v = synth.g2(37, 41, 43)
    rms = ol.oracle().cvxo_global_rms(v.reshape(-1), 37, 41, 43)
    naive = np.sqrt(np.mean(v.astype(np.float64) ** 2))
    assert abs(rms - naive) / naive < 1e-6


def test_block_size_validity():
    o = ol.oracle()
    assert o.cvxo_is_valid_block_size(32, 32, 32) and o.cvxo_is_valid_block_size(8, 256, 1)
    for bad in [(4, 8, 8), (8, 8, 512), (24, 8, 8), (8, 8, 2), (8, 8, 0), (8, 1, 8)]:
        assert not o.cvxo_is_valid_block_size(*bad)

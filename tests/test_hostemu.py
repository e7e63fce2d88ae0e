"""CPU checks of the device code's per-thread phases (cvxcompress_b200/csrc/*.cuh compiled for the
host by tests/hostemu) against the oracle: line transforms (EXACT bit-exact, LIFT within tolerance),
the segmented RLE encoder for several thread counts, and the speculative chunk parser."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import oracle_libs as ol
from cvxcompress_b200 import synth
from test_oracle import _rle_cases

HERE = os.path.dirname(os.path.abspath(__file__))
EMU_DIR = os.path.join(HERE, "hostemu")
EMU_SO = os.path.join(EMU_DIR, "libcvx_hostemu.so")
_f32p = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")


def _build():
    srcs = [os.path.join(EMU_DIR, "hostemu.cu")] + [os.path.join(ol.ROOT, "cvxcompress_b200", "csrc", f)
                                                    for f in ("cvx_wavelet.cuh", "cvx_rle.cuh")]
    if os.path.exists(EMU_SO) and all(os.path.getmtime(EMU_SO) >= os.path.getmtime(s) for s in srcs):
        return
    subprocess.check_call(["nvcc", "-O2", "-std=c++17", "-arch=sm_100a", "-diag-suppress", "20013,20015",
                           "-Xcompiler", "-fPIC,-ffp-contract=off", "--shared", "-o", EMU_SO, srcs[0]])


@pytest.fixture(scope="module")
def emu():
    _build()
    L = C.CDLL(EMU_SO)
    L.emu_line.argtypes = [_f32p, C.c_int, C.c_int, C.c_int]
    L.emu_wavelet.argtypes = [_f32p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
    L.emu_rle_encode.argtypes = [C.c_float, _f32p, C.c_int, C.c_int, _u8p]
    L.emu_rle_decode.argtypes = [C.c_float, _f32p, C.c_int, C.c_void_p, C.c_long, C.c_long, C.c_int]
    L.emu_rle_decode8.argtypes = [C.c_float, _f32p, C.c_int, C.c_void_p, C.c_long, C.c_long, C.c_int, C.c_long]
    L.emu_quantise.argtypes = [C.c_float]
    L.emu_encode32_lifting_zyx.argtypes = [_f32p]
    return L


@pytest.mark.parametrize("shape", [(8, 8, 8), (16, 16, 16), (32, 32, 32), (16, 16, 64), (64, 32, 8), (32, 32, 1),
                                   (128, 8, 8), (8, 8, 256)])
def test_exact_lines_bitexact_vs_oracle(emu, shape):
    bx, by, bz = shape
    n = bx * by * bz
    blk = synth.g3(n, seed=21 + n, lo=-1.0, hi=1.0)
    a = blk.copy(); b = blk.copy()
    assert emu.emu_wavelet(a, bx, by, bz, 1, 0) == 0
    ol.oracle().cvxo_wavelet_forward(b, bx, by, bz, 0)
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    assert emu.emu_wavelet(a, bx, by, bz, 1, 1) == 0
    ol.oracle().cvxo_wavelet_inverse(b, bx, by, bz, 0)
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))


@pytest.mark.parametrize("shape", [(8, 8, 8), (16, 16, 16), (32, 32, 32), (16, 16, 64), (32, 32, 1)])
def test_lift_lines_within_tolerance(emu, shape):
    """north_star: coefficients match Wavelet_Transform_Slow within 1e-6 relative (to the block magnitude;
    the reference's own Fast-vs-Slow criterion is rms-relative 1e-5, CvxCompress.cpp:586-598)."""
    bx, by, bz = shape
    n = bx * by * bz
    blk = synth.g3(n, seed=5 + n, lo=-1.0, hi=1.0)
    a = blk.copy(); b = blk.copy()
    assert emu.emu_wavelet(a, bx, by, bz, 0, 0) == 0
    ol.oracle().cvxo_wavelet_forward(b, bx, by, bz, 1)       # Slow association
    rms_rel = np.sqrt(np.mean((a - b).astype(np.float64) ** 2)) / np.sqrt(np.mean(b.astype(np.float64) ** 2))
    max_rel = np.abs(a - b).max() / np.abs(b).max()
    assert rms_rel < 1e-6, rms_rel
    assert max_rel < 2e-6, max_rel
    # inverse of the oracle's coefficients reproduces the oracle's inverse
    c = b.copy(); d = b.copy()
    assert emu.emu_wavelet(c, bx, by, bz, 0, 1) == 0
    ol.oracle().cvxo_wavelet_inverse(d, bx, by, bz, 1)
    assert np.abs(c - d).max() < 2e-6
    assert np.abs(c - blk).max() < 3e-6


@pytest.mark.parametrize("name", list(_rle_cases().keys()))
@pytest.mark.parametrize("nthreads", [1, 8, 64, 512])
def test_segmented_encoder_bytes_equal_oracle(emu, name, nthreads):
    vals = _rle_cases()[name]
    n = len(vals)
    for mulfac in (1.0, 0.37):
        ref = np.zeros(5 * n + 64, np.uint8)
        m = ol.oracle().cvxo_rle_encode(np.float32(mulfac), vals, n, ref)
        out = np.zeros(5 * n + 64, np.uint8)
        k = emu.emu_rle_encode(np.float32(mulfac), vals, n, nthreads, out)
        assert k == m
        assert bytes(out[:k]) == bytes(ref[:m])


@pytest.mark.parametrize("name", list(_rle_cases().keys()))
@pytest.mark.parametrize("off", [0, 5, 16, 31])
def test_speculative_decoder_equals_oracle(emu, name, off):
    vals = _rle_cases()[name]
    n = len(vals)
    mulfac = np.float32(0.37)
    enc = np.zeros(5 * n + 64, np.uint8)
    m = ol.oracle().cvxo_rle_encode(mulfac, vals, n, enc)
    want = np.zeros(n, np.float32)
    ol.oracle().cvxo_rle_decode(mulfac, want, n, enc)
    rng = np.random.default_rng(off)
    buf = ol.aligned_u8(off + m + 4096)
    buf[:] = rng.integers(0, 256, len(buf), dtype=np.uint8)       # adversarial garbage around the payload
    buf[off:off + m] = enc[:m]
    for nthreads in (4, 64, 512):
        got = np.full(n, 7.0, np.float32)
        rc = emu.emu_rle_decode(mulfac, got, n, buf.ctypes.data, off, len(buf), nthreads)
        assert rc >= 0
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("name", list(_rle_cases().keys()))
@pytest.mark.parametrize("off", [0, 5, 16, 27])
def test_chunk8_decoder_equals_oracle(emu, name, off):
    """the 8-byte chunk formulation used by k_decode32, with exact, wrong (too short) and unknown length hints"""
    vals = _rle_cases()[name]
    n = len(vals)
    mulfac = np.float32(0.37)
    enc = np.zeros(5 * n + 64, np.uint8)
    m = ol.oracle().cvxo_rle_encode(mulfac, vals, n, enc)
    want = np.zeros(n, np.float32)
    ol.oracle().cvxo_rle_decode(mulfac, want, n, enc)
    rng = np.random.default_rng(off + 100)
    buf = ol.aligned_u8(off + m + 16384)
    buf[:] = rng.integers(0, 256, len(buf), dtype=np.uint8)
    buf[off:off + m] = enc[:m]
    for nch in (2, 16, 1024):
        for hint in (m, max(1, m // 3), 0):
            got = np.full(n, 7.0, np.float32)
            rc = emu.emu_rle_decode8(mulfac, got, n, buf.ctypes.data, off, len(buf), nch, hint)
            assert rc >= 0, (nch, hint, rc)
            assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), (nch, hint)


def test_quantise_matches_oracle(emu):
    for f in [0.0, 0.99, -0.99, 1.0, -1.5, 124.9, 125.0, 2147483520.0, 2147483648.0, -2147483648.0, -2147483904.0,
              float("nan"), float("inf"), float("-inf"), 3e9, -3e9, 8388607.5]:
        assert emu.emu_quantise(np.float32(f)) == ol.oracle().cvxo_quantise(np.float32(f)), f


def test_encoders_zyx_order_with_folded_scalings_within_lifting_tolerance(emu):
    """What k_encode32ws / k_encode32<LIFTING> compute -- passes z, y, x, bare lifting steps on the z and y lines, one multiply
    on the x rows carrying all three axes' scalings -- against the oracle's forward transform (reference order x, y, z, direct
    form): rms-relative <= 1e-6 and max-norm <= 2e-6 * max|c|, the LIFTING contract of DESIGN.md section 5."""
    for seed in (3, 4):
        blk = synth.g3(32768, seed=seed, lo=-1.0, hi=1.0)
        a = blk.copy(); b = blk.copy()
        assert emu.emu_encode32_lifting_zyx(a) == 0
        ol.oracle().cvxo_wavelet_forward(b, 32, 32, 32, 1)
        d = a.astype(np.float64) - b.astype(np.float64)
        assert np.sqrt(np.mean(d * d)) <= 1e-6 * np.sqrt(np.mean(b.astype(np.float64) ** 2))
        assert np.max(np.abs(d)) <= 2e-6 * np.max(np.abs(b))

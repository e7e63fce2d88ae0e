"""cvxgpu_validate_stream: the bounds-checked walk accepts every stream the reference and the oracle
write -- any block order, local RMS, raw blocks, 2-D blocks, the committed golden streams -- and names the first offending
block of a damaged one.  Host logic only: runs without a GPU.  The streams come from the checkers (oracle / oracle/_ref);
the thing tested is the product library's validator."""
import os

import numpy as np
import pytest

import oracle_libs as ol
from cvxcompress_b200 import _lib, api, synth

HERE = os.path.dirname(os.path.abspath(__file__))


def _vol(nx, ny, nz, seed=0):
    return synth.g2(nx, ny, nz) + np.random.default_rng(seed).standard_normal((nz, ny, nx)).astype(np.float32) * 1e-3


def _facts_agree(info, stream):
    s = ol.parse_stream(stream)
    assert (info["nx"], info["ny"], info["nz"], info["bx"], info["by"], info["bz"]) == (s["nx"], s["ny"], s["nz"], s["bx"], s["by"], s["bz"])
    assert info["nblocks"] == s["nnn"] and info["use_local_rms"] == (s["flags"] & 1)
    assert info["payload_bytes"] == len(s["payload"])
    assert info["raw_blocks"] == int((s["offs"] >> np.uint64(63)).sum())
    assert info["used_payload_bytes"] == info["payload_bytes"]          # the reference packs blocks back to back
    payloads, _, _ = ol.block_payloads(stream)
    assert info["min_block_bytes"] == min(map(len, payloads)) and info["max_block_bytes"] == max(map(len, payloads))


@pytest.mark.parametrize("dims,blk,scale,local", [
    ((96, 80, 72), (32, 32, 32), 1e-2, False),
    ((70, 45, 33), (16, 8, 32), 1e-3, True),          # ragged edges, local RMS
    ((64, 64, 5), (32, 32, 1), 1e-2, False),          # 2-D blocks
    ((40, 40, 40), (8, 8, 8), 1e-4, True),
    ((130, 20, 20), (128, 16, 16), 1e-2, False),
])
def test_oracle_streams_validate(dims, blk, scale, local):
    nx, ny, nz = dims
    stream, _ = ol.oracle_compress(_vol(nx, ny, nz), *blk, scale, local)
    info = api.validate_stream(stream)
    _facts_agree(info, stream)
    assert info["natural_order"] == 1 and info["tokens"] > 0
    lvl0 = api.validate_stream(stream, level=0)
    assert lvl0["used_payload_bytes"] == -1 and lvl0["tokens"] == -1 and lvl0["raw_blocks"] == info["raw_blocks"]


@pytest.mark.skipif(not ol.have_ref(), reason="oracle/_ref not built")
@pytest.mark.parametrize("threads", [1, 4])
def test_reference_streams_validate_in_any_block_order(threads):
    vol = _vol(96, 96, 64, seed=3)
    stream, _ = ol.ref_compress(vol, 32, 32, 32, 1e-2, False, threads=threads)
    info = api.validate_stream(stream)
    _facts_agree(info, stream)
    if threads == 1:
        assert info["natural_order"] == 1


def test_raw_blocks_validate():
    # white noise at a tiny scale: every coefficient needs an escape, the encoder falls back to raw fp32 (CvxCompress.cpp:350-360)
    vol = np.random.default_rng(1).standard_normal((16, 16, 16)).astype(np.float32) * 1e6
    stream, _ = ol.oracle_compress(vol, 8, 8, 8, 1e-9, False)
    info = api.validate_stream(stream)
    assert info["raw_blocks"] == info["nblocks"] == 8
    assert info["min_block_bytes"] == info["max_block_bytes"] == 4 * 512
    # a raw block that would pass the payload end is caught at level 0
    bad = stream[:len(stream) - 16].copy()
    with pytest.raises(_lib.CvxError, match="raw block 7"):
        api.validate_stream(bad, level=0)


def test_golden_streams_validate():
    z = np.load(os.path.join(HERE, "golden", "small_streams.npz"))
    n = 0
    for k in z.files:
        a = z[k]
        if a.dtype == np.uint8 and a.ndim == 1 and len(a) >= 39:
            # make_golden.py stores the written bytes only; compressed_length counts 7 more (CvxCompress.cpp:421-422)
            full = np.concatenate([a, np.zeros(7, np.uint8)])
            _facts_agree(api.validate_stream(full), full)
            with pytest.raises(_lib.CvxError):          # without the slack the last block no longer fits
                api.validate_stream(a)
            n += 1
    assert n == 5


def _good():
    stream, _ = ol.oracle_compress(_vol(64, 64, 64, seed=5), 32, 32, 32, 1e-2, True)
    return stream.copy(), ol.parse_stream(stream)


def test_damaged_streams_are_refused():
    good, s = _good()
    api.validate_stream(good)
    E = _lib.CvxError

    with pytest.raises(E, match="shorter than its 32-byte header"):
        api.validate_stream(good[:20])
    with pytest.raises(E, match="cannot hold"):
        api.validate_stream(good[:32 + 12 * s["nnn"]])

    b = good.copy(); b[12:16] = np.frombuffer(np.int32(24).tobytes(), np.uint8)        # bx = 24
    with pytest.raises(E, match="bad header"):
        api.validate_stream(b)
    b = good.copy(); b[28] = 3
    with pytest.raises(E, match="flags"):
        api.validate_stream(b)
    b = good.copy(); b[24:28] = 0                                                     # glob_mulfac = 0
    with pytest.raises(E, match="multiplier"):
        api.validate_stream(b)
    b = good.copy(); b[32 + 8 * s["nnn"] + 4 * 3:][:4] = np.frombuffer(np.float32(np.nan).tobytes(), np.uint8)
    with pytest.raises(E, match="block 3: multiplier"):
        api.validate_stream(b)

    b = good.copy(); b[32 + 8 * 5:][:8] = np.frombuffer(np.uint64(len(s["payload"]) + 1).tobytes(), np.uint8)
    with pytest.raises(E, match="block 5: offset"):
        api.validate_stream(b, level=0)
    b = good.copy(); b[32 + 8 * 5:][:8] = b[32 + 8 * 2:][:8]
    with pytest.raises(E, match="share payload offset"):
        api.validate_stream(b)
    api.validate_stream(b, level=0)                      # level 0 does not compare blocks with each other

    # truncation inside the last block: level 0 cannot see it, level 1 does
    cut = good[:len(good) - 40]
    api.validate_stream(cut, level=0)
    with pytest.raises(E, match="block 7"):
        api.validate_stream(cut)


def test_run_overshoot_and_cut_tokens_are_refused():
    good, s = _good()
    pb = s["payload_base"]
    offs = (s["offs"] & np.uint64(0x7FFFFFFFFFFFFFFF)).astype(np.int64)
    pay = good[pb:]
    # find a RLESC1 token (byte 127) that starts a token in block 0 by walking it with the published token table
    size = {127: 2, 125: 4, 0x83: 3, 0x81: 4, 0x80: 5, 0x82: 17, 126: 25}
    pos, hit = int(offs[0]), None
    while pos < int(offs[1]):
        c = int(pay[pos])
        if c == 127 and hit is None:
            hit = pos
        pos += size.get(c, 1)
    assert pos == int(offs[1]) and hit is not None
    b = good.copy(); b[pb + hit + 1] = 255 if b[pb + hit + 1] != 255 else 254          # lengthen / shorten one zero run
    with pytest.raises(_lib.CvxError, match="block 0"):
        api.validate_stream(b)
    # turn the block's last byte into the head of a 25-byte token: it is cut off by the next block's first byte
    b = good.copy(); b[pb + int(offs[1]) - 1] = 126
    with pytest.raises(_lib.CvxError, match="block 0"):
        api.validate_stream(b)


def test_safe_decode_switch_is_host_logic():
    L = _lib.lib()
    prev = api.set_safe_decode(1)
    try:
        assert L.cvxgpu_get_safe_decode() == 1
        assert api.set_safe_decode(None) == 1 and L.cvxgpu_get_safe_decode() == -1
        assert L.cvxgpu_set_safe_decode(2) == _lib.ERR_ARG and b"safe-decode" in L.cvxgpu_last_error()
        assert L.cvxgpu_get_safe_decode() == -1
    finally:
        api.set_safe_decode(prev)


@pytest.mark.gpu
def test_safe_decode_refuses_a_damaged_host_stream_before_the_gpu():
    import ctypes as C
    L = _lib.lib()
    nx, ny, nz = 64, 64, 64
    good, s = _good()
    want = ol.oracle_decompress(good, nx, ny, nz)
    bad = good.copy()
    bad[s["payload_base"] + int(s["offs"][1] & np.uint64(0x7FFFFFFFFFFFFFFF)) - 1] = 126      # cut 25-byte token ends block 0
    out = np.full(nx * ny * nz, 7.0, dtype=np.float32)
    prev = api.set_safe_decode(1)
    try:
        rc = L.cvxgpu_decompress(out.ctypes.data, nx, ny, nz, bad.ctypes.data, len(bad))
        assert rc == _lib.ERR_STREAM and b"block 0" in L.cvxgpu_last_error()
        assert (out == 7.0).all()                                   # nothing was decoded
        rc = L.cvxgpu_decompress(out.ctypes.data, nx, ny, nz, good.ctypes.data, len(good))
        assert rc == 0 and np.array_equal(out.view(np.uint32), want.reshape(-1).view(np.uint32))
        api.set_safe_decode(None)                                   # reference behaviour: decoded without a walk, and it terminates
        assert L.cvxgpu_decompress(out.ctypes.data, nx, ny, nz, bad.ctypes.data, len(bad)) == 0
    finally:
        api.set_safe_decode(prev)


def test_threaded_walk_equals_serial_walk(monkeypatch):
    vol = _vol(160, 160, 96, seed=9)
    stream, _ = ol.oracle_compress(vol, 8, 8, 8, 1e-3, False)          # 4800 blocks: enough for every host thread
    monkeypatch.setenv("CVX_GPU_VALIDATE_THREADS", "1")
    serial = api.validate_stream(stream)
    monkeypatch.setenv("CVX_GPU_VALIDATE_THREADS", "7")
    assert api.validate_stream(stream) == serial
    monkeypatch.delenv("CVX_GPU_VALIDATE_THREADS")
    assert api.validate_stream(stream) == serial
    _facts_agree(serial, stream)
    # two damaged blocks: the lowest one is reported whatever thread meets it first
    s = ol.parse_stream(stream)
    offs = (s["offs"] & np.uint64(0x7FFFFFFFFFFFFFFF)).astype(np.int64)
    bad = stream.copy()
    for blk in (4100, 1234):
        bad[s["payload_base"] + int(offs[blk + 1]) - 1] = 126
    for threads in ("1", "7", "64"):
        monkeypatch.setenv("CVX_GPU_VALIDATE_THREADS", threads)
        with pytest.raises(_lib.CvxError, match="block 1234:"):
            api.validate_stream(bad)


def test_overflowing_header_is_refused_not_fatal():
    """A crafted header whose block count overflows 64 bits (16777216 x 16777216 x 8388608 cells in 8^3 blocks: nblk = 2^62,
    8 * nblk wraps to 0) must come back as CVXGPU_ERR_STREAM -- no allocation, no exception across the C ABI."""
    buf = np.zeros(96, np.uint8)
    hdr = np.array([16777216, 16777216, 8388608, 8, 8, 8], np.int32)
    buf[:24] = hdr.view(np.uint8)
    buf[24:28] = np.frombuffer(np.float32(1.0).tobytes(), np.uint8)
    with pytest.raises(_lib.CvxError, match="cannot hold"):
        api.validate_stream(buf)
    # dimensions next to INT_MAX: nx + bx - 1 overflows a 32-bit int
    hdr = np.array([2147483647, 2147483647, 2147483647, 256, 256, 256], np.int32)
    buf[:24] = hdr.view(np.uint8)
    with pytest.raises(_lib.CvxError, match="cannot hold"):
        api.validate_stream(buf)

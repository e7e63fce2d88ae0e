"""ctypes bindings to the two CHECKERS (never to the product):

* ``oracle/libcvx_oracle.so``  -- the plain-C restatement (oracle/cvx_oracle.c)
* ``oracle/_ref/libcvxcompress_ref.so`` -- the unmodified reference built by oracle/Makefile
  (exports cvx_* from CvxCompress.hxx:139-192 plus the internal C++ stage functions).

Test infrastructure only; imported by tests/, __graft_entry__.smoke() and bench.py's CPU legs.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_SO = os.path.join(ORACLE_DIR, "libcvx_oracle.so")
REF_SO = os.path.join(ORACLE_DIR, "_ref", "libcvxcompress_ref.so")

_f32p = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")


def build_oracle(ref=True):
    """(Re)build the checkers.  The reference build needs /root/reference (absent on the GPU box)."""
    subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "oracle"])
    if ref and os.path.isdir("/root/reference") and not os.path.exists(REF_SO):
        subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "ref"])


_oracle = None
_ref = None


def oracle():
    global _oracle
    if _oracle is None:
        if not os.path.exists(ORACLE_SO):
            build_oracle(ref=False)
        L = C.CDLL(ORACLE_SO)
        L.cvxo_wavelet_forward.argtypes = [_f32p, C.c_int, C.c_int, C.c_int, C.c_int]
        L.cvxo_wavelet_inverse.argtypes = [_f32p, C.c_int, C.c_int, C.c_int, C.c_int]
        L.cvxo_copy_to_block.argtypes = [_f32p] + [C.c_int] * 6 + [_f32p] + [C.c_int] * 3
        L.cvxo_copy_from_block.argtypes = [_f32p] + [C.c_int] * 3 + [_f32p] + [C.c_int] * 6
        L.cvxo_global_rms.argtypes = [_f32p, C.c_int, C.c_int, C.c_int]
        L.cvxo_global_rms.restype = C.c_float
        L.cvxo_local_rms.argtypes = [_f32p, C.c_int, C.c_int, C.c_int]
        L.cvxo_local_rms.restype = C.c_float
        L.cvxo_rle_encode.argtypes = [C.c_float, _f32p, C.c_int, _u8p]
        L.cvxo_rle_encode.restype = C.c_int
        L.cvxo_rle_decode.argtypes = [C.c_float, _f32p, C.c_int, _u8p]
        L.cvxo_rle_decode.restype = C.c_int
        L.cvxo_quantise.argtypes = [C.c_float]
        L.cvxo_quantise.restype = C.c_int
        L.cvxo_compress.argtypes = [C.c_float, _f32p] + [C.c_int] * 7 + [_u8p, C.POINTER(C.c_long)]
        L.cvxo_compress.restype = C.c_float
        L.cvxo_decompress.argtypes = [_f32p, C.c_int, C.c_int, C.c_int, _u8p]
        L.cvxo_decompress.restype = C.c_int
        L.cvxo_is_valid_block_size.argtypes = [C.c_int] * 3
        L.cvxo_max_compressed_bytes.argtypes = [C.c_int] * 7
        L.cvxo_max_compressed_bytes.restype = C.c_long
        L.cvxo_mirr.argtypes = [C.c_int, C.c_int]
        L.cvxo_mirr_sl.argtypes = [C.c_int, C.c_int]
        L.cvxo_mirr_sh.argtypes = [C.c_int, C.c_int, C.c_int]
        _oracle = L
    return _oracle


def have_ref():
    return os.path.exists(REF_SO)


def ref():
    """The unmodified reference library (None-safe: raises if it was never built)."""
    global _ref
    if _ref is None:
        if not os.path.exists(REF_SO):
            raise FileNotFoundError(REF_SO + " (run `make -C oracle ref` where /root/reference exists)")
        L = C.CDLL(REF_SO)
        L.cvx_compress_th.argtypes = [C.c_float, _f32p] + [C.c_int] * 6 + [C.c_bool, _u8p, C.c_int, C.POINTER(C.c_long)]
        L.cvx_compress_th.restype = C.c_float
        L.cvx_decompress_inplace_th.argtypes = [_f32p, C.c_int, C.c_int, C.c_int, _u8p, C.c_int, C.c_long]
        L.cvx_decompress_inplace_th.restype = None
        # internal stage functions (Itanium-mangled C++), see SURVEY.md 8(b)
        L.fwd_fast = L._Z30Wavelet_Transform_Fast_ForwardPDv8_fS0_iii
        L.inv_fast = L._Z30Wavelet_Transform_Fast_InversePDv8_fS0_iii
        for f in (L.fwd_fast, L.inv_fast):
            f.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
            f.restype = None
        L.fwd_slow = L._Z30Wavelet_Transform_Slow_ForwardPfS_iiiiiiiii
        L.inv_slow = L._Z30Wavelet_Transform_Slow_InversePfS_iiiiiiiii
        for f in (L.fwd_slow, L.inv_slow):
            f.argtypes = [_f32p, _f32p] + [C.c_int] * 9
            f.restype = None
        L.rle_encode = L._Z22Run_Length_Encode_SlowfPfiPmRi
        L.rle_encode.argtypes = [C.c_float, C.c_void_p, C.c_int, C.c_void_p, C.POINTER(C.c_int)]
        L.rle_encode.restype = None
        L.rle_decode = L._Z22Run_Length_Decode_SlowfPfiPm
        L.rle_decode.argtypes = [C.c_float, _f32p, C.c_int, C.c_void_p]
        L.rle_decode.restype = C.c_int
        L.copy_to_block = L._Z13Copy_To_BlockPfiiiiiiPDv4_fiii
        L.copy_to_block.argtypes = [_f32p] + [C.c_int] * 6 + [C.c_void_p] + [C.c_int] * 3
        L.copy_to_block.restype = None
        L.copy_from_block = L._Z15Copy_From_BlockPDv4_fiiiPfiiiiii
        L.copy_from_block.argtypes = [C.c_void_p] + [C.c_int] * 3 + [_f32p] + [C.c_int] * 6
        L.copy_from_block.restype = None
        _ref = L
    return _ref


def aligned_f32(n, align=64):
    """The reference's AVX kernels use aligned loads on block/tmp buffers."""
    raw = np.zeros(n * 4 + align, dtype=np.uint8)
    off = (-raw.ctypes.data) % align
    return raw[off:off + n * 4].view(np.float32)


def aligned_u8(n, align=64):
    raw = np.zeros(n + align, dtype=np.uint8)
    off = (-raw.ctypes.data) % align
    return raw[off:off + n]


# ---- convenience wrappers ------------------------------------------------------------------
def ref_compress(vol, bx, by, bz, scale, local=False, threads=1):
    nz, ny, nx = vol.shape
    cap = oracle().cvxo_max_compressed_bytes(nx, ny, nz, bx, by, bz, int(local)) + 64
    out = aligned_u8(cap)
    ln = C.c_long(0)
    ratio = ref().cvx_compress_th(scale, np.ascontiguousarray(vol).reshape(-1), nx, ny, nz, bx, by, bz,
                                  bool(local), out, threads, C.byref(ln))
    return out[:ln.value].copy(), float(ratio)


def ref_decompress(stream, nx, ny, nz, threads=1):
    vol = np.zeros(nx * ny * nz, dtype=np.float32)
    buf = aligned_u8(len(stream) + 64)
    buf[:len(stream)] = stream
    ref().cvx_decompress_inplace_th(vol, nx, ny, nz, buf, threads, len(stream))
    return vol.reshape(nz, ny, nx)


def oracle_compress(vol, bx, by, bz, scale, local=False):
    nz, ny, nx = vol.shape
    cap = oracle().cvxo_max_compressed_bytes(nx, ny, nz, bx, by, bz, int(local)) + 64
    out = np.zeros(cap, dtype=np.uint8)
    ln = C.c_long(0)
    ratio = oracle().cvxo_compress(scale, np.ascontiguousarray(vol).reshape(-1), nx, ny, nz, bx, by, bz,
                                   int(local), out, C.byref(ln))
    return out[:ln.value].copy(), float(ratio)


def oracle_decompress(stream, nx, ny, nz):
    vol = np.zeros(nx * ny * nz, dtype=np.float32)
    buf = np.zeros(len(stream) + 64, dtype=np.uint8)
    buf[:len(stream)] = stream
    rc = oracle().cvxo_decompress(vol, nx, ny, nz, buf)
    assert rc == 0
    return vol.reshape(nz, ny, nx)


def parse_stream(stream):
    """Split a stream into header fields, offsets, mulfacs and the payload view."""
    hdr = np.frombuffer(stream[:32].tobytes(), dtype=np.int32)
    nx, ny, nz, bx, by, bz = (int(v) for v in hdr[:6])
    mulfac = float(np.frombuffer(stream[24:28].tobytes(), dtype=np.float32)[0])
    flags = int(hdr[7])
    nnn = -(-nx // bx) * -(-ny // by) * -(-nz // bz)
    offs = np.frombuffer(stream[32:32 + 8 * nnn].tobytes(), dtype=np.uint64)
    p = 32 + 8 * nnn
    bmf = None
    if flags & 1:
        bmf = np.frombuffer(stream[p:p + 4 * nnn].tobytes(), dtype=np.float32)
        p += 4 * nnn
    return dict(nx=nx, ny=ny, nz=nz, bx=bx, by=by, bz=bz, mulfac=mulfac, flags=flags, nnn=nnn,
                offs=offs, blkmulfac=bmf, payload=stream[p:len(stream) - 7], payload_base=p)


def block_payloads(stream):
    """Per-block payload bytes irrespective of the order blocks were laid out in."""
    s = parse_stream(stream)
    offs = s["offs"] & np.uint64(0x7FFFFFFFFFFFFFFF)
    raw = (s["offs"] >> np.uint64(63)).astype(bool)
    order = np.argsort(offs, kind="stable")
    ends = np.empty(s["nnn"], dtype=np.int64)
    so = offs[order].astype(np.int64)
    ends[order] = np.append(so[1:], len(s["payload"]))
    return [bytes(s["payload"][int(offs[i]):int(ends[i])]) for i in range(s["nnn"])], raw, s

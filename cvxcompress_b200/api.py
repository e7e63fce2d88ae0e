"""Host-side mirror of the reference interface (CvxCompress.hxx:19-135) over the C ABI.

``CvxCompress`` keeps the reference's method names and argument meaning for host (numpy) buffers:
it calls the library's cvx_compress_th / cvx_decompress_inplace_th, i.e. exactly what a program
linked against the reference would call.  ``compress_device`` / ``decompress_device`` are the
device-resident fast path (torch CUDA tensors -> cvxgpu_*_device), the one bench.py times.
"""
import ctypes as C

import numpy as np

from . import _lib


def set_transform(mode):
    """'exact' (bit-identical to the reference, default) or 'lifting' (fast, ~4e-7 rms-relative)."""
    m = {"exact": _lib.TRANSFORM_EXACT, "lifting": _lib.TRANSFORM_LIFTING}[mode] if isinstance(mode, str) else int(mode)
    _lib.check(_lib.lib().cvxgpu_set_transform(m))


def get_transform():
    return "exact" if _lib.lib().cvxgpu_get_transform() == _lib.TRANSFORM_EXACT else "lifting"


def max_compressed_bytes(nx, ny, nz, bx, by, bz, use_local_rms=False):
    n = _lib.lib().cvxgpu_max_compressed_bytes(nx, ny, nz, bx, by, bz, int(bool(use_local_rms)))
    if n < 0:
        raise ValueError("invalid dimensions / block size")
    return n


def num_blocks(nx, ny, nz, bx, by, bz):
    return _lib.lib().cvxgpu_num_blocks(nx, ny, nz, bx, by, bz)


def parse_header(stream):
    """(nx, ny, nz, bx, by, bz, glob_mulfac, flags) of a host stream (numpy uint8 / bytes)."""
    buf = np.frombuffer(bytes(stream[:32]), dtype=np.uint8) if not isinstance(stream, np.ndarray) else np.ascontiguousarray(stream[:32])
    v = [C.c_int() for _ in range(6)]
    gm, fl = C.c_float(), C.c_int()
    _lib.check(_lib.lib().cvxgpu_parse_header(buf.ctypes.data, len(buf), *[C.byref(x) for x in v], C.byref(gm), C.byref(fl)))
    return tuple(x.value for x in v) + (gm.value, fl.value)


def validate_stream(stream, length=None, level=1):
    """Bounds-checked walk of a HOST stream (numpy uint8 / bytes) before it is handed to a decoder: header, tables and,
    at level 1, every token of every block (cvxgpu_validate_stream).  Returns the stream's facts as a dict; raises
    CvxError (code CVXGPU_ERR_STREAM, message naming the first offending block) on a malformed stream.  No GPU needed."""
    buf = np.frombuffer(bytes(stream), dtype=np.uint8) if not isinstance(stream, np.ndarray) else np.ascontiguousarray(stream).view(np.uint8).reshape(-1)
    n = len(buf) if length is None else int(length)
    if n > len(buf):
        raise ValueError("length %d exceeds the %d-byte buffer" % (n, len(buf)))
    info = _lib.StreamInfo()
    _lib.check(_lib.lib().cvxgpu_validate_stream(buf.ctypes.data, n, int(level), C.byref(info)))
    return info.as_dict()


def set_safe_decode(level):
    """-1 / None: off (the reference's behaviour); 0 or 1: the host-pointer Decompress path validates a host stream at that
    level before staging it (cvxgpu_set_safe_decode).  Returns the previous level."""
    prev = _lib.lib().cvxgpu_get_safe_decode()
    _lib.check(_lib.lib().cvxgpu_set_safe_decode(-1 if level is None else int(level)))
    return prev


class CvxCompress:
    """Python mirror of the reference class for HOST buffers (numpy)."""

    Min_BX = Min_BY = Min_BZ = 8
    Max_BX = Max_BY = Max_BZ = 256

    def Is_Valid_Block_Size(self, bx, by, bz):
        return bool(_lib.lib().cvxgpu_is_valid_block_size(bx, by, bz))

    def Compress(self, scale, vol, nx, ny, nz, bx, by, bz, use_local_RMS=False, compressed=None, num_threads=1):
        """Returns (ratio, compressed_length).  `vol`: float32 numpy array of nz*ny*nx cells (x fastest);
        `compressed`: writable uint8/uint32 numpy buffer (allocated when None and returned as third item)."""
        vol = np.ascontiguousarray(vol, dtype=np.float32)
        assert vol.size == nx * ny * nz
        if not self.Is_Valid_Block_Size(bx, by, bz):
            raise ValueError("invalid block size %dx%dx%d" % (bx, by, bz))     # the reference asserts (CvxCompress.cpp:246-248)
        own = compressed is None
        if own:
            compressed = np.empty(max_compressed_bytes(nx, ny, nz, bx, by, bz, use_local_RMS) + 8, dtype=np.uint8)
        ln = C.c_long(0)
        ratio = _lib.lib().cvx_compress_th(float(scale), vol.ctypes.data, nx, ny, nz, bx, by, bz, bool(use_local_RMS),
                                           compressed.ctypes.data, int(num_threads), C.byref(ln))
        if own:
            return float(ratio), ln.value, compressed
        return float(ratio), ln.value

    def Decompress(self, vol, nx, ny, nz, compressed, compressed_length, num_threads=1):
        """In-place variant (CvxCompress.hxx:105-113): fills `vol` (float32 numpy, nz*ny*nx)."""
        assert vol.dtype == np.float32 and vol.flags["C_CONTIGUOUS"] and vol.size == nx * ny * nz
        hx, hy, hz = parse_header(compressed)[:3]
        if (hx, hy, hz) != (nx, ny, nz):
            raise ValueError("stream holds %dx%dx%d, expected %dx%dx%d" % (hx, hy, hz, nx, ny, nz))   # reference: printf + assert
        compressed = np.ascontiguousarray(compressed)
        _lib.lib().cvx_decompress_inplace_th(vol.ctypes.data, nx, ny, nz, compressed.ctypes.data, int(num_threads),
                                             int(compressed_length))
        return vol

    def Decompress_outofplace(self, compressed, compressed_length):
        """CvxCompress.hxx:97-103: returns (vol[nz,ny,nx], nx, ny, nz)."""
        nx, ny, nz = parse_header(compressed)[:3]
        vol = np.empty((nz, ny, nx), dtype=np.float32)
        self.Decompress(vol, nx, ny, nz, compressed, compressed_length)
        return vol, nx, ny, nz


# ---- device-resident path (torch CUDA tensors) --------------------------------------------------
def _stream_ptr():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def compress_device(vol, bx, by, bz, scale, use_local_rms=False, out=None):
    """vol: CUDA float32 tensor [nz, ny, nx].  Returns (stream uint8 CUDA tensor, length, ratio)."""
    import torch
    assert vol.is_cuda and vol.dtype == torch.float32 and vol.is_contiguous() and vol.dim() == 3
    nz, ny, nx = vol.shape
    cap = max_compressed_bytes(nx, ny, nz, bx, by, bz, use_local_rms)
    if out is None:
        out = torch.empty(cap + 8, dtype=torch.uint8, device=vol.device)
    else:
        cap = min(cap, out.numel())
    ln, ratio = C.c_long(0), C.c_float(0)
    with torch.cuda.device(vol.device):
        _lib.check(_lib.lib().cvxgpu_compress_device(None, vol.data_ptr(), nx, ny, nz, bx, by, bz, float(scale),
                                                     int(bool(use_local_rms)), out.data_ptr(), cap, C.byref(ln), C.byref(ratio),
                                                     _stream_ptr()))
    return out, ln.value, ratio.value


def compress_device_async(vol, bx, by, bz, scale, use_local_rms, out, d_length, d_status):
    """No host synchronisation: length -> d_length (int64 CUDA tensor[1]), status -> d_status (int32[1])."""
    import torch
    nz, ny, nx = vol.shape
    with torch.cuda.device(vol.device):
        _lib.check(_lib.lib().cvxgpu_compress_device_async(None, vol.data_ptr(), nx, ny, nz, bx, by, bz, float(scale),
                                                           int(bool(use_local_rms)), out.data_ptr(), out.numel(),
                                                           d_length.data_ptr(), d_status.data_ptr(), _stream_ptr()))


def decompress_device(stream, length, nx, ny, nz, out=None):
    """stream: CUDA uint8 tensor.  Returns the CUDA float32 volume [nz, ny, nx]."""
    import torch
    assert stream.is_cuda and stream.dtype == torch.uint8
    if out is None:
        out = torch.empty((nz, ny, nx), dtype=torch.float32, device=stream.device)
    with torch.cuda.device(stream.device):
        _lib.check(_lib.lib().cvxgpu_decompress_device(None, out.data_ptr(), nx, ny, nz, stream.data_ptr(), int(length), _stream_ptr()))
    return out


def wavelet_blocks_device(blocks, bx, by, bz, inverse=False):
    """In-place forward / inverse transform of a CUDA float32 tensor holding dense blocks [nblocks, bz, by, bx]."""
    import torch
    assert blocks.is_cuda and blocks.dtype == torch.float32 and blocks.is_contiguous()
    nblocks = blocks.numel() // (bx * by * bz)
    with torch.cuda.device(blocks.device):
        _lib.check(_lib.lib().cvxgpu_wavelet_blocks_device(None, blocks.data_ptr(), nblocks, bx, by, bz, int(bool(inverse)), _stream_ptr()))
    return blocks

"""torch custom ops over the device-pointer C ABI: the callers of this path are GPU wave-propagation
codes that checkpoint wavefields already resident in HBM, which the reference's host-pointer class (CvxCompress.hxx:32-45)
cannot serve.

    stream, length = torch.ops.cvx.compress(vol, bx, by, bz, scale, use_local_rms)
    length         = torch.ops.cvx.compress_into(vol, bx, by, bz, scale, use_local_rms, out)     # caller-sized buffer
    vol            = torch.ops.cvx.decompress(stream, nx, ny, nz)
    vol            = torch.ops.cvx.decompress_known(stream, nx, ny, nz, bx, by, bz, use_local_rms)   # no host sync

`compress` allocates the worst case (every block raw: more than the volume itself); a checkpoint loop that knows its data
passes its own buffer to `compress_into` -- a quarter of the volume is ample for seismic wavefields -- and an overflow
comes back as a negative length, nothing is written past the buffer.  `decompress_known` takes the header fields from the
caller instead of reading the stream's first 32 bytes back, so the decode side is free of host synchronisation too and
both directions can be captured in one CUDA graph.

Both are stream-ordered on torch's current CUDA stream.  `compress` does not synchronise with the host: `length` is a
1-element int64 CUDA tensor (compressed_length, CvxCompress.cpp:421-422; negative = cvxgpu error code, e.g. the stream did
not fit), `stream` a uint8 CUDA tensor of the worst-case capacity whose first `length` bytes are the reference-format
stream.  `decompress` reads nx,ny,nz-matching streams only (32-byte header read, then asynchronous); reads are bounded by
the tensor's size.  CUDA tensors only: there is no CPU implementation, a CPU tensor raises.

torch is plumbing here (tensors, streams, dispatch); the work is libcvxcompress.so's kernels."""
import torch

from . import _lib, api

_lib_def = torch.library.Library("cvx", "DEF")
_lib_def.define("compress(Tensor vol, int bx, int by, int bz, float scale, bool use_local_rms) -> (Tensor, Tensor)")
_lib_def.define("decompress(Tensor stream, int nx, int ny, int nz) -> Tensor")
_lib_def.define("compress_into(Tensor vol, int bx, int by, int bz, float scale, bool use_local_rms, Tensor(a!) out) -> Tensor")
_lib_def.define("decompress_known(Tensor stream, int nx, int ny, int nz, int bx, int by, int bz, bool use_local_rms) -> Tensor")


def _check_vol(vol):
    if vol.dim() != 3 or vol.dtype != torch.float32:
        raise ValueError("cvx::compress: vol must be a float32 tensor [nz, ny, nx]")


def _compress_cuda(vol, bx, by, bz, scale, use_local_rms):
    _check_vol(vol)
    vol = vol.contiguous()
    nz, ny, nx = vol.shape
    cap = api.max_compressed_bytes(nx, ny, nz, bx, by, bz, use_local_rms)
    if cap < 0:
        raise ValueError("cvx::compress: invalid block size %dx%dx%d" % (bx, by, bz))
    out = torch.empty(cap + 8, dtype=torch.uint8, device=vol.device)
    d_length = torch.zeros(1, dtype=torch.int64, device=vol.device)
    d_status = torch.zeros(1, dtype=torch.int32, device=vol.device)
    api.compress_device_async(vol, bx, by, bz, scale, use_local_rms, out, d_length, d_status)
    # a failed encode (capacity) reports through length: negative error code instead of a byte count, still without a host sync
    d_length = torch.where(d_status != 0, d_status.to(torch.int64), d_length)
    return out, d_length


def _compress_into_cuda(vol, bx, by, bz, scale, use_local_rms, out):
    _check_vol(vol)
    if out.dtype != torch.uint8 or out.dim() != 1 or not out.is_contiguous() or out.data_ptr() % 8:
        raise ValueError("cvx::compress_into: out must be a contiguous, 8-byte aligned 1-D uint8 tensor")
    vol = vol.contiguous()
    nz, ny, nx = vol.shape
    if api.max_compressed_bytes(nx, ny, nz, bx, by, bz, use_local_rms) < 0:
        raise ValueError("cvx::compress_into: invalid block size %dx%dx%d" % (bx, by, bz))
    d_length = torch.zeros(1, dtype=torch.int64, device=vol.device)
    d_status = torch.zeros(1, dtype=torch.int32, device=vol.device)
    api.compress_device_async(vol, bx, by, bz, scale, use_local_rms, out, d_length, d_status)
    return torch.where(d_status != 0, d_status.to(torch.int64), d_length)


def _decompress_known_cuda(stream, nx, ny, nz, bx, by, bz, use_local_rms):
    import ctypes as C
    if stream.dtype != torch.uint8 or stream.dim() != 1:
        raise ValueError("cvx::decompress_known: stream must be a 1-D uint8 tensor")
    stream = stream.contiguous()
    out = torch.empty((nz, ny, nx), dtype=torch.float32, device=stream.device)
    with torch.cuda.device(stream.device):
        _lib.check(_lib.lib().cvxgpu_decompress_slab_device(
            None, out.data_ptr(), nx, ny, nz, bx, by, bz, int(bool(use_local_rms)), 0, 1, stream.data_ptr(), stream.numel(),
            C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    return out


def _decompress_cuda(stream, nx, ny, nz):
    if stream.dtype != torch.uint8 or stream.dim() != 1:
        raise ValueError("cvx::decompress: stream must be a 1-D uint8 tensor")
    stream = stream.contiguous()
    return api.decompress_device(stream, stream.numel(), nx, ny, nz)


def _no_cpu(*args, **kwargs):
    raise RuntimeError("cvx ops run on CUDA tensors only: this library has no CPU implementation")


_lib_impl = torch.library.Library("cvx", "IMPL")
_lib_impl.impl("compress", _compress_cuda, "CUDA")
_lib_impl.impl("decompress", _decompress_cuda, "CUDA")
_lib_impl.impl("compress_into", _compress_into_cuda, "CUDA")
_lib_impl.impl("decompress_known", _decompress_known_cuda, "CUDA")
_lib_impl.impl("compress_into", _no_cpu, "CPU")
_lib_impl.impl("decompress_known", _no_cpu, "CPU")
_lib_impl.impl("compress", _no_cpu, "CPU")
_lib_impl.impl("decompress", _no_cpu, "CPU")


@torch.library.register_fake("cvx::compress")
def _compress_fake(vol, bx, by, bz, scale, use_local_rms):
    _check_vol(vol)
    nz, ny, nx = vol.shape
    cap = api.max_compressed_bytes(nx, ny, nz, bx, by, bz, use_local_rms)
    return vol.new_empty(cap + 8, dtype=torch.uint8), vol.new_empty(1, dtype=torch.int64)


@torch.library.register_fake("cvx::decompress")
def _decompress_fake(stream, nx, ny, nz):
    return stream.new_empty((nz, ny, nx), dtype=torch.float32)


@torch.library.register_fake("cvx::compress_into")
def _compress_into_fake(vol, bx, by, bz, scale, use_local_rms, out):
    _check_vol(vol)
    return vol.new_empty(1, dtype=torch.int64)


@torch.library.register_fake("cvx::decompress_known")
def _decompress_known_fake(stream, nx, ny, nz, bx, by, bz, use_local_rms):
    return stream.new_empty((nz, ny, nx), dtype=torch.float32)

"""ctypes binding of the product library ``cvxcompress_b200/lib/libcvxcompress.so``
(C ABI: include/cvx_gpu.h; reference-compatible C functions: include/CvxCompress.hxx).

There is no fallback: if the library is missing this module raises, and every compute entry
point fails loudly when no CUDA device is present.
"""
import ctypes as C
import os
import subprocess

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
# CVXCOMPRESS_LIB: developer override (e.g. the `make TIMING=1 OUT=../lib_timing` build)
LIB_PATH = os.environ.get("CVXCOMPRESS_LIB") or os.path.join(PKG_DIR, "lib", "libcvxcompress.so")
CSRC_DIR = os.path.join(PKG_DIR, "csrc")

OK, ERR_ARG, ERR_CUDA, ERR_CAPACITY, ERR_STREAM = 0, -1, -2, -3, -4
TRANSFORM_EXACT, TRANSFORM_LIFTING = 0, 1

# every symbol include/cvx_gpu.h and include/CvxCompress.hxx (extern "C" part) declare
C_SYMBOLS = [
    "cvxgpu_last_error", "cvxgpu_device_count", "cvxgpu_set_transform", "cvxgpu_get_transform",
    "cvxgpu_workspace_create", "cvxgpu_workspace_destroy", "cvxgpu_is_valid_block_size", "cvxgpu_num_blocks",
    "cvxgpu_max_compressed_bytes", "cvxgpu_compress_device", "cvxgpu_compress_device_async",
    "cvxgpu_decompress_device", "cvxgpu_parse_header", "cvxgpu_compress", "cvxgpu_decompress",
    "cvxgpu_sumsq_device", "cvxgpu_encode_blocks_device", "cvxgpu_write_header_device",
    "cvxgpu_offsets_add_device", "cvxgpu_decode_blocks_device", "cvxgpu_wavelet_blocks_device",
    "cvx_compress", "cvx_decompress_outofplace", "cvx_decompress_inplace", "cvx_compress_th",
    "cvx_decompress_inplace_th", "cvxgpu_validate_stream", "cvxgpu_set_safe_decode", "cvxgpu_get_safe_decode",
    "cvxgpu_workspace_release_thread_default", "cvxgpu_sumsq_rows_device", "cvxgpu_sum_rows_device",
    "cvxgpu_shard_range", "cvxgpu_shard_unique_id", "cvxgpu_shard_create", "cvxgpu_shard_destroy", "cvxgpu_shard_uses_peer_memory",
    "cvxgpu_compress_sharded", "cvxgpu_decompress_sharded", "cvxgpu_decompress_slab_device",
    "cvxgpu_sumsq_rows_static_device",
]


class CvxError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("cvxgpu error %d: %s" % (code, msg))
        self.code = code


class StreamInfo(C.Structure):
    """cvxgpu_stream_info (include/cvx_gpu.h)."""
    _fields_ = [(n, C.c_int) for n in ("nx", "ny", "nz", "bx", "by", "bz", "use_local_rms")] + [("glob_mulfac", C.c_float)] + \
               [(n, C.c_long) for n in ("nblocks", "raw_blocks", "table_bytes", "payload_bytes", "used_payload_bytes",
                                        "min_block_bytes", "max_block_bytes", "tokens")] + [("natural_order", C.c_int)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


def build(force=False):
    """Compile the library in-tree with nvcc for sm_100a (cvxcompress_b200/csrc/Makefile)."""
    if force:
        subprocess.check_call(["make", "-s", "-C", CSRC_DIR, "clean"])
    subprocess.check_call(["make", "-s", "-C", CSRC_DIR])
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FileNotFoundError(
            LIB_PATH + " is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C cvxcompress_b200/csrc`). There is no CPU fallback.")
    try:
        import torch  # noqa: F401  (loads the CUDA runtime the library links against)
    except Exception:
        pass
    # RTLD_LOCAL: the reference build under oracle/_ref defines the same CvxCompress:: symbols; neither may interpose the other
    L = C.CDLL(LIB_PATH, mode=C.RTLD_LOCAL)
    i, l, f, p = C.c_int, C.c_long, C.c_float, C.c_void_p
    L.cvxgpu_last_error.restype = C.c_char_p
    L.cvxgpu_last_error.argtypes = []
    L.cvxgpu_device_count.argtypes = []
    L.cvxgpu_set_transform.argtypes = [i]
    L.cvxgpu_get_transform.argtypes = []
    L.cvxgpu_workspace_create.argtypes = [C.POINTER(p)]
    L.cvxgpu_workspace_destroy.argtypes = [p]
    L.cvxgpu_is_valid_block_size.argtypes = [i, i, i]
    L.cvxgpu_num_blocks.argtypes = [i] * 6
    L.cvxgpu_num_blocks.restype = l
    L.cvxgpu_max_compressed_bytes.argtypes = [i] * 7
    L.cvxgpu_max_compressed_bytes.restype = l
    L.cvxgpu_compress_device.argtypes = [p, p, i, i, i, i, i, i, f, i, p, l, C.POINTER(l), C.POINTER(f), p]
    L.cvxgpu_compress_device_async.argtypes = [p, p, i, i, i, i, i, i, f, i, p, l, p, p, p]
    L.cvxgpu_decompress_device.argtypes = [p, p, i, i, i, p, l, p]
    L.cvxgpu_parse_header.argtypes = [p, l] + [C.POINTER(i)] * 6 + [C.POINTER(f), C.POINTER(i)]
    L.cvxgpu_validate_stream.argtypes = [p, l, i, C.POINTER(StreamInfo)]
    L.cvxgpu_compress.argtypes = [p, i, i, i, i, i, i, f, i, p, C.POINTER(l), C.POINTER(f)]
    L.cvxgpu_decompress.argtypes = [p, i, i, i, p, l]
    L.cvxgpu_sumsq_device.argtypes = [p, p, l, p, p]
    L.cvxgpu_workspace_release_thread_default.argtypes = []
    L.cvxgpu_sumsq_rows_device.argtypes = [p, p, l, i, i, p, p]
    L.cvxgpu_sumsq_rows_static_device.argtypes = [p, l, i, i, p, p]
    L.cvxgpu_sum_rows_device.argtypes = [p, i, p, p]
    L.cvxgpu_shard_range.argtypes = [i, i, i, i] + [C.POINTER(i)] * 4
    L.cvxgpu_shard_unique_id.argtypes = [p]
    L.cvxgpu_shard_create.argtypes = [C.POINTER(p), p, p, i, i]
    L.cvxgpu_shard_destroy.argtypes = [p]
    L.cvxgpu_shard_uses_peer_memory.argtypes = [p]
    L.cvxgpu_compress_sharded.argtypes = [p, p, i, i, i, i, i, i, f, i, p, p, p, p, l, p, p, p]
    L.cvxgpu_decompress_sharded.argtypes = [p, p, i, i, i, i, i, i, i, p, p, p, p, p, p]
    L.cvxgpu_decompress_slab_device.argtypes = [p, p, i, i, i, i, i, i, i, i, i, p, l, p]
    L.cvxgpu_encode_blocks_device.argtypes = [p, p, i, i, i, i, i, i, f, i, p, l, p, p, p, l, p, p, p]
    L.cvxgpu_write_header_device.argtypes = [p, i, i, i, i, i, i, f, i, p, l, p]
    L.cvxgpu_offsets_add_device.argtypes = [p, l, l, p]
    L.cvxgpu_decode_blocks_device.argtypes = [p, p, i, i, i, i, i, i, f, i, p, p, p, l, p]
    L.cvxgpu_wavelet_blocks_device.argtypes = [p, p, l, i, i, i, i, p]
    # reference-compatible C functions (CvxCompress.hxx:139-192)
    L.cvx_compress.argtypes = [f, p, i, i, i, i, i, i, p, C.POINTER(l)]
    L.cvx_compress.restype = f
    L.cvx_compress_th.argtypes = [f, p, i, i, i, i, i, i, C.c_bool, p, i, C.POINTER(l)]
    L.cvx_compress_th.restype = f
    L.cvx_decompress_inplace.argtypes = [p, i, i, i, p, l]
    L.cvx_decompress_inplace.restype = None
    L.cvx_decompress_inplace_th.argtypes = [p, i, i, i, p, i, l]
    L.cvx_decompress_inplace_th.restype = None
    L.cvx_decompress_outofplace.argtypes = [C.POINTER(i), C.POINTER(i), C.POINTER(i), p, l]
    L.cvx_decompress_outofplace.restype = C.POINTER(f)
    _lib = L
    return L


def check(rc):
    if rc != OK:
        raise CvxError(rc, lib().cvxgpu_last_error().decode("utf-8", "replace"))

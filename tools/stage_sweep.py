#!/usr/bin/env python
"""Developer tool: host-pointer round trip (cvx_compress_th + cvx_decompress_inplace_th) from PAGEABLE memory on config 3,
for the CVX_GPU_STAGE_THREADS / CVX_GPU_STAGE_CHUNK_MB given in the environment.  Prints ms per direction."""
import ctypes as C, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from cvxcompress_b200 import _lib, api, synth
L = _lib.lib(); api.set_transform("lifting")
nx, ny, nz = 1056, 1248, 960
vol = synth.g2_torch(nx, ny, nz, "cuda").cpu().numpy()
out = np.zeros_like(vol); comp = np.zeros(vol.size, np.uint8); ln = C.c_long(0)
def step():
    t0 = time.perf_counter()
    L.cvx_compress_th(1e-2, vol.ctypes.data, nx, ny, nz, 32, 32, 32, False, comp.ctypes.data, 1, C.byref(ln))
    t1 = time.perf_counter()
    L.cvx_decompress_inplace_th(out.ctypes.data, nx, ny, nz, comp.ctypes.data, 1, ln.value)
    return t1 - t0, time.perf_counter() - t1
step()
r = [step() for _ in range(4)]
print("threads", os.environ.get("CVX_GPU_STAGE_THREADS", "6"), "chunk_mb", os.environ.get("CVX_GPU_STAGE_CHUNK_MB", "8"),
      "compress %.1f ms  decompress %.1f ms" % (1e3 * min(a for a, _ in r), 1e3 * min(b for _, b in r)), "cores", len(os.sched_getaffinity(0)))

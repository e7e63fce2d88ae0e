"""Times cvxgpu_validate_stream (host walk) on the full-size BASELINE config 3 stream produced on the GPU.
Run on the GPU box: python tools/validate_timing.py"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cvxcompress_b200 import api, synth  # noqa: E402

api.set_transform("lifting")
nx, ny, nz = 1056, 1248, 960
vol = synth.g2_torch(nx, ny, nz, "cuda")
s, ln, ratio = api.compress_device(vol, 32, 32, 32, 1e-2)
h = s[:ln].cpu().numpy()
for lvl in (0, 1):
    t = time.perf_counter()
    info = api.validate_stream(h, level=lvl)
    dt = time.perf_counter() - t
    print("level %d: %d stream bytes in %.2f ms (%.2f GB/s) %s" % (lvl, ln, dt * 1e3, ln / dt * 1e-9, info))

#!/usr/bin/env python
"""Small Compress/Decompress cases for compute-sanitizer (memcheck / racecheck): 32^3 kernels incl. partial blocks,
raw blocks, local RMS, both arithmetic modes, and one generic-shape case."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from cvxcompress_b200 import api, synth
for mode in ("exact", "lifting"):
    api.set_transform(mode)
    for (nx, ny, nz, b, scale, local) in ((70, 40, 40, (32, 32, 32), 1e-2, False), (64, 64, 32, (32, 32, 32), 1e-9, False),
                                          (64, 36, 64, (32, 32, 32), 1e-3, True), (40, 24, 24, (8, 8, 8), 1e-2, False)):
        vol = torch.from_numpy(synth.g2(nx, ny, nz)).cuda()
        s, ln, ratio = api.compress_device(vol, b[0], b[1], b[2], scale, local)
        out = api.decompress_device(s, ln, nx, ny, nz)
        torch.cuda.synchronize()
        err = float((out - vol).norm() / vol.norm())
        print(mode, nx, ny, nz, b, scale, local, "len", ln, "err %.2e" % err)
print("done")

#!/usr/bin/env python
"""Developer check (run on a B200): md5 of the LIFTING stream and reconstruction for a set of block shapes with the library
CVXCOMPRESS_LIB selects -- two builds whose arithmetic is meant to be bit-identical must print the same lines."""
import hashlib, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cvxcompress_b200 import api, synth
api.set_transform("lifting")
nx, ny, nz = 200, 150, 132
d = torch.from_numpy(synth.g2(nx, ny, nz)).cuda()
out = {}
for shape in [(8, 8, 8), (16, 16, 16), (16, 16, 64), (64, 16, 16), (32, 32, 16), (32, 16, 16), (8, 32, 16), (64, 64, 8), (16, 8, 1), (32, 32, 1)]:
    for local in (False, True):
        s, ln, _ = api.compress_device(d, *shape, 1e-2, local)
        rec = api.decompress_device(s, ln, nx, ny, nz)
        out["%dx%dx%d/%d" % (shape + (int(local),))] = [ln, hashlib.md5(s[:ln - 7].cpu().numpy().tobytes()).hexdigest()[:12],
                                                         hashlib.md5(rec.cpu().numpy().tobytes()).hexdigest()[:12]]
print(json.dumps(out))

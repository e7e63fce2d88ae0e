#!/usr/bin/env python
"""Developer tool: one Compress + Decompress of the config-3 volume with b^3 blocks (for an ncu launch list)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cvxcompress_b200 import api, synth
b = int(sys.argv[1]) if len(sys.argv) > 1 else 64
api.set_transform(sys.argv[2] if len(sys.argv) > 2 else "lifting")
vol = synth.g2_torch(1056, 1248, 960, torch.device("cuda", 0))
s, ln, ratio = api.compress_device(vol, b, b, b, 1e-2)
out = api.decompress_device(s, ln, 1056, 1248, 960)
torch.cuda.synchronize()
print(ln, ratio)

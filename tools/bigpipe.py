#!/usr/bin/env python
"""Developer timing (run on a B200): whole Compress / Decompress calls with large blocks on the config-3 volume."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cvxcompress_b200 import api, synth
mode = sys.argv[1] if len(sys.argv) > 1 else "lifting"
api.set_transform(mode)
NX, NY, NZ = 1056, 1248, 960
vol = synth.g2_torch(NX, NY, NZ, torch.device("cuda", 0))
res = {}
for b in (32, 64, 128, 256):
    s, ln, ratio = api.compress_device(vol, b, b, b, 1e-2)
    out = api.decompress_device(s, ln, NX, NY, NZ)
    torch.cuda.synchronize()
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    e[0].record()
    s, ln, ratio = api.compress_device(vol, b, b, b, 1e-2, out=s)
    e[1].record()
    out = api.decompress_device(s, ln, NX, NY, NZ, out=out)
    e[2].record(); torch.cuda.synchronize()
    err = float(((out - vol).double().pow(2).mean().sqrt() / vol.double().pow(2).mean().sqrt()))
    res["%d^3" % b] = {"compress_ms": e[0].elapsed_time(e[1]), "decompress_ms": e[1].elapsed_time(e[2]), "ratio": ratio, "rel_err": err}
print(json.dumps(res, indent=1))

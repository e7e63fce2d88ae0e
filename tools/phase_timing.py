#!/usr/bin/env python
"""Per-phase cycle shares of k_encode32 / k_decode32 (library built with `make TIMING=1`)."""
import ctypes as C, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cvxcompress_b200 import api, synth, _lib
mode = sys.argv[1] if len(sys.argv) > 1 else "lifting"
nz = int(sys.argv[2]) if len(sys.argv) > 2 else 256
api.set_transform(mode)
L = _lib.lib()
vol = synth.g2_torch(1056, 1248, nz, "cuda", z0=0, nz_total=960)
s, ln, _ = api.compress_device(vol, 32, 32, 32, 1e-2)
out = (C.c_ulonglong * 16)()
for kind, name, fn in ((0, "encode", lambda: api.compress_device(vol, 32, 32, 32, 1e-2, out=s)),
                       (1, "decode", lambda: api.decompress_device(s, ln, 1056, 1248, nz))):
    L.cvx_debug_phases(out, kind, 1)
    fn(); torch.cuda.synchronize()
    n = L.cvx_debug_phases(out, kind, 1)
    v = list(out)[:12]; tot = sum(v) or 1
    nblk = 33 * 39 * ((nz + 31) // 32)
    extra = list(out)[10:12]
    tot -= sum(extra)
    print(name, mode, "counters/block %.2f %.2f" % (extra[0] / nblk, extra[1] / nblk), "cycles/block %.0f" % (tot / nblk), " ".join("%d:%.1f%%" % (i, 100.0 * x / tot) for i, x in enumerate(v)))

"""Full-size check of the warp-specialised 32^3 kernels against the fused ones (LIFTING): repeated runs, device-side\ncomparison of streams and reconstructions; any difference is a race or a hand-over bug.  Run on a B200."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, numpy as np
from cvxcompress_b200 import api, synth, _lib
L = _lib.lib()
api.set_transform("lifting")
for nz in (192, 960):
    vol = synth.g2_torch(1056, 1248, nz, "cuda", z0=0, nz_total=960)
    L.cvx_debug_set_ws(0)
    s0, ln0, _ = api.compress_device(vol, 32, 32, 32, 1e-2)
    s0 = s0[:ln0 - 7].clone()
    b = api.decompress_device(s0, ln0, 1056, 1248, nz).clone()
    L.cvx_debug_set_ws(1)
    bad_enc = bad_dec = 0
    for rep in range(6):
        s, ln, _ = api.compress_device(vol, 32, 32, 32, 1e-2)
        bad_enc += int(ln != ln0 or not torch.equal(s[:ln - 7], s0))
        a = api.decompress_device(s0, ln0, 1056, 1248, nz)
        bad_dec += int((a != b).sum())
    print("nz", nz, "ws encoder streams differing from fused:", bad_enc, "of 6; ws decoder cells differing:", bad_dec,
          "err", float(((b - vol).double().pow(2).mean().sqrt() / vol.double().pow(2).mean().sqrt())))

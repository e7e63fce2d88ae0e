#!/usr/bin/env python
"""Kernel A/B harness (developer tool, run on a B200): times the encode and decode kernels of ONE library build
(CVXCOMPRESS_LIB=... selects it) on the config-3 G2 volume and checks the warp-specialised LIFTING kernels bit for bit
against the fused ones of the same build.  Prints one JSON line.

  CVXCOMPRESS_LIB=cvxcompress_b200/lib_exp/v2/libcvxcompress.so python tools/kbench.py [nz] [mode] [reps]
"""
import ctypes as C, json, math, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cvxcompress_b200 import api, synth, _lib

nz = int(sys.argv[1]) if len(sys.argv) > 1 else 960
mode = sys.argv[2] if len(sys.argv) > 2 else "lifting"
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 10
NX, NY, B, SCALE = 1056, 1248, 32, 1e-2
L = _lib.lib()
api.set_transform(mode)
dev = torch.device("cuda", 0)
vol = synth.g2_torch(NX, NY, nz, dev, z0=0, nz_total=960)
nblk = 33 * 39 * ((nz + 31) // 32)
st = lambda: C.c_void_p(torch.cuda.current_stream().cuda_stream)
d_sumsq = torch.zeros(1, dtype=torch.float64, device=dev)
offs = torch.zeros(nblk, dtype=torch.int64, device=dev)
pay = torch.empty(vol.numel() + 4096, dtype=torch.uint8, device=dev)
nbytes = torch.zeros(1, dtype=torch.int64, device=dev)
stat = torch.zeros(1, dtype=torch.int32, device=dev)
out = torch.empty_like(vol)
_lib.check(L.cvxgpu_sumsq_device(None, vol.data_ptr(), vol.numel(), d_sumsq.data_ptr(), st()))


def enc():
    _lib.check(L.cvxgpu_encode_blocks_device(None, vol.data_ptr(), NX, NY, nz, B, B, B, SCALE, 0, d_sumsq.data_ptr(), vol.numel(),
                                             offs.data_ptr(), None, pay.data_ptr(), pay.numel(), nbytes.data_ptr(), stat.data_ptr(), st()))


enc(); torch.cuda.synchronize()
pb = int(nbytes.item())
gm = 1.0 / (math.sqrt(float(d_sumsq.item()) / vol.numel()) * SCALE)


def dec():
    _lib.check(L.cvxgpu_decode_blocks_device(None, out.data_ptr(), NX, NY, nz, B, B, B, gm, 0, offs.data_ptr(), None, pay.data_ptr(), pb, st()))


def timeit(fn):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


res = {"lib": os.environ.get("CVXCOMPRESS_LIB", "default"), "nz": nz, "mode": mode, "payload": pb}
res["enc_ms"] = timeit(enc)
dec(); torch.cuda.synchronize()
res["dec_ms"] = timeit(dec)
abytes = 4 * vol.numel() + pb + 8 * nblk
res["enc_gbs"] = abytes / res["enc_ms"] / 1e6
res["dec_gbs"] = abytes / res["dec_ms"] / 1e6
res["err"] = float(((out - vol).double().pow(2).mean().sqrt() / vol.double().pow(2).mean().sqrt()))
if mode == "lifting" and hasattr(L, "cvx_debug_set_ws"):
    # bitwise against the fused kernels of the same build, several repetitions (races show up run to run)
    pay_ws, offs_ws, out_ws = pay[:pb].clone(), offs.clone(), out.clone()
    L.cvx_debug_set_ws(0)
    enc(); dec(); torch.cuda.synchronize()
    pay_f, offs_f, out_f, pb_f = pay[:pb].clone(), offs.clone(), out.clone(), int(nbytes.item())
    L.cvx_debug_set_ws(1)
    bad_e = bad_d = 0
    for _ in range(4):
        enc(); torch.cuda.synchronize()
        ok = int(nbytes.item()) == pb_f and torch.equal(pay[:pb_f], pay_f) and torch.equal(offs, offs_f)
        bad_e += 0 if ok else 1
        dec(); torch.cuda.synchronize()
        bad_d += int((out != out_f).sum().item() != 0)
    res["ws_vs_fused_bad_enc"] = bad_e
    res["ws_vs_fused_bad_dec"] = bad_d
print(json.dumps(res))

"""Generates tests/golden/golden.json and tests/golden/small_streams.npz with the UNMODIFIED reference
(oracle/_ref/libcvxcompress_ref.so, built from /root/reference by oracle/Makefile).  Run in the build
container (the reference sources do not exist on the GPU box):

    make -C oracle ref && python tests/golden/make_golden.py
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle_libs as ol                    # noqa: E402
from cvxcompress_b200 import synth          # noqa: E402


def sha(b):
    return hashlib.sha256(bytes(b)).hexdigest()


def main():
    out = {}
    small = {}
    # config 1: the reference's own golden (README.md:72-81)
    vol = synth.g1(352, 416, 320)
    s, ratio = ol.ref_compress(vol, 32, 32, 32, 1e-2, False, threads=1)
    rec = ol.ref_decompress(s, 352, 416, 320, threads=4)
    err, snr = synth.quality(vol, rec)
    out["c1_g1_352x416x320_b32_s1e-2"] = dict(length=len(s), ratio=ratio, error=err, snr=snr,
                                             sha256_written=sha(s[:len(s) - 7]), sha256_decoded=sha(rec.tobytes()))
    # small cases whose full streams are committed
    cases = [("g2", 151, 101, 51, 32, 32, 32, 1e-2, False), ("g2", 151, 101, 51, 16, 16, 64, 1e-2, True),
             ("g2", 37, 41, 43, 8, 8, 8, 1e-2, False), ("g2", 160, 96, 5, 32, 32, 1, 1e-2, False),
             ("g2", 70, 50, 40, 16, 16, 16, 1e-4, False)]
    for kind, nx, ny, nz, bx, by, bz, scale, local in cases:
        vol = synth.g2(nx, ny, nz)
        s, ratio = ol.ref_compress(vol, bx, by, bz, scale, local, threads=1)
        rec = ol.ref_decompress(s, nx, ny, nz, threads=2)
        err, snr = synth.quality(vol, rec)
        key = "%s_%dx%dx%d_b%dx%dx%d_s%g_%s" % (kind, nx, ny, nz, bx, by, bz, scale, "local" if local else "global")
        out[key] = dict(length=len(s), ratio=ratio, error=err, snr=snr, sha256_written=sha(s[:len(s) - 7]),
                        sha256_decoded=sha(rec.tobytes()), nx=nx, ny=ny, nz=nz, bx=bx, by=by, bz=bz, scale=scale, local=local)
        small[key] = s[:len(s) - 7].copy()
    with open(os.path.join(HERE, "golden.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    np.savez_compressed(os.path.join(HERE, "small_streams.npz"), **small)
    for k, v in out.items():
        print(k, v["length"], "%.2f" % v["ratio"], "%.6e" % v["error"], "%.1f" % v["snr"])


if __name__ == "__main__":
    main()

"""cvxfile (cvxcompress_b200/csrc/cvxfile.cpp): the file driver over the class API, counterpart of the reference's
Test_Compression.cpp:53-230.  `info` is host logic and runs without a GPU; compress / decompress are
GPU tests: the stream the tool writes must equal the oracle's byte for byte (default EXACT transform) and the volume it writes
back must equal the oracle's decode of that stream bit for bit."""
import os
import re
import subprocess

import numpy as np
import pytest

import oracle_libs as ol
from cvxcompress_b200 import _lib, synth

TOOL = os.path.join(os.path.dirname(_lib.LIB_PATH), "cvxfile")


def _run(*args, env=None):
    _lib.build()
    e = dict(os.environ)
    e.pop("CVX_GPU_TRANSFORM", None)
    e.update(env or {})
    return subprocess.run([TOOL] + [str(a) for a in args], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, env=e, timeout=300)


def test_usage_and_missing_file():
    assert _run().returncode == 2
    assert _run("frobnicate", "x").returncode == 2
    r = _run("info", "/nonexistent/stream.cvx")
    assert r.returncode == 1 and "cannot open" in r.stderr


def test_info_needs_no_gpu(tmp_path):
    vol = synth.g2(70, 50, 40)
    stream, ratio = ol.oracle_compress(vol, 16, 16, 16, 1e-3, True)
    f = tmp_path / "a.cvx"
    stream.tofile(f)
    r = _run("info", f)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "70 x 50 x 40 cells, 16 x 16 x 16 blocks, local RMS" in r.stdout
    assert "60 blocks (0 raw)" in r.stdout and "natural block order" in r.stdout and "well formed (level 1)" in r.stdout
    assert abs(float(re.search(r"ratio ([0-9.]+):1", r.stdout).group(1)) - ratio) < 0.01
    assert "well formed (level 0)" in _run("info", f, "--shallow").stdout
    # a stream cut inside its last block: refused, the first offending block is named
    stream[:len(stream) - 50].tofile(f)
    r = _run("info", f)
    assert r.returncode == 1 and "INVALID: block 59" in r.stdout
    # decompress refuses it before touching the GPU
    r = _run("decompress", f, tmp_path / "a.raw")
    assert r.returncode == 1 and "not a well-formed stream" in r.stderr and not (tmp_path / "a.raw").exists()


def test_compress_rejects_bad_arguments_before_the_gpu(tmp_path):
    f = tmp_path / "v.raw"
    np.zeros(8 * 8 * 8, np.float32).tofile(f)
    assert _run("compress", f, 8, 8, 8, 1e-2, 24, 8, 8, tmp_path / "v.cvx").returncode == 2     # invalid block size
    assert _run("compress", f, 8, 8, 8, 1e-2, 8, 8).returncode == 2                              # too few arguments


@pytest.mark.gpu
@pytest.mark.parametrize("dims,blk,scale,local", [
    ((151, 101, 51), (32, 32, 32), 1e-2, False),      # the reference driver's volume size (Read_Raw_Volume.cpp:13-15)
    ((70, 50, 40), (16, 16, 16), 1e-4, True),
])
def test_compress_then_decompress_files(tmp_path, dims, blk, scale, local):
    nx, ny, nz = dims
    vol = synth.g2(nx, ny, nz)
    raw, cvx, back = tmp_path / "v.raw", tmp_path / "v.cvx", tmp_path / "back.raw"
    vol.tofile(raw)
    args = ["compress", raw, nx, ny, nz, scale, *blk, cvx, 2] + (["--local"] if local else [])
    r = _run(*args)
    assert r.returncode == 0, r.stdout + r.stderr
    assert ("Using %s RMS." % ("local" if local else "global")) in r.stdout and r.stdout.count("Compressing.") == 2
    want, want_ratio = ol.oracle_compress(vol, *blk, scale, local)
    got = np.fromfile(cvx, dtype=np.uint8)
    assert len(got) == len(want) and bytes(got[:-7]) == bytes(want[:-7]) and not got[-7:].any()
    assert abs(float(re.search(r"ratio is ([0-9.]+):1", r.stdout).group(1)) - want_ratio) < 0.01
    ref = ol.oracle_decompress(want, nx, ny, nz)
    err = float(re.search(r"Error is ([0-9.e+-]+)", r.stdout).group(1))
    want_err = float(np.sqrt(np.mean((vol.astype(np.float64) - ref) ** 2)) / np.sqrt(np.mean(vol.astype(np.float64) ** 2)))
    assert abs(err - want_err) <= 2e-3 * want_err

    assert _run("info", cvx).returncode == 0
    r = _run("decompress", cvx, back)
    assert r.returncode == 0, r.stdout + r.stderr
    rec = np.fromfile(back, dtype=np.float32).reshape(nz, ny, nx)
    assert np.array_equal(rec.view(np.uint32), ref.view(np.uint32))


@pytest.mark.gpu
def test_wrong_file_size_is_refused(tmp_path):
    f = tmp_path / "v.raw"
    np.zeros(1000, np.float32).tofile(f)
    r = _run("compress", f, 16, 16, 16, 1e-2, 8, 8, 8, tmp_path / "v.cvx")
    assert r.returncode == 1 and "need" in r.stderr

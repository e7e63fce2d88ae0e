"""The C-ABI library loads and exports every symbol the headers declare (no compute calls: runs without a GPU)."""
import ctypes as C
import os
import re
import subprocess

import pytest

from cvxcompress_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_c_symbols():
    names = set()
    for hdr in ("cvx_gpu.h", "CvxCompress.hxx"):
        txt = open(os.path.join(ROOT, "include", hdr)).read()
        txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
        txt = re.sub(r"//[^\n]*", "", txt)
        names |= set(re.findall(r"\b(cvxgpu_[a-z_0-9]+|cvx_[a-z_]+)\s*\(", txt))
    return sorted(names)


def test_library_exports_every_declared_symbol():
    _lib.build()
    L = C.CDLL(_lib.LIB_PATH)
    declared = _declared_c_symbols()
    assert len(declared) >= 25
    missing = [n for n in declared if not hasattr(L, n)]
    assert not missing, missing
    assert sorted(set(_lib.C_SYMBOLS)) == declared      # the ctypes binding covers exactly the declared surface


def test_class_symbols_exported_with_reference_mangling():
    """the Itanium-mangled C++ symbols a program compiled against the reference's CvxCompress.hxx links to"""
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], stdout=subprocess.PIPE, text=True, check=True).stdout
    for sym in ("_ZN11CvxCompressC1Ev", "_ZN11CvxCompressD1Ev",
                "_ZN11CvxCompress8CompressEfPfiiiiiibPjiRl",       # Compress(float,float*,int x6,bool,unsigned*,int,long&)
                "_ZN11CvxCompress8CompressEfPfiiiiiibPjRl", "_ZN11CvxCompress8CompressEfPfiiiiiiPjRl",
                "_ZN11CvxCompress8CompressEfPfiiiiiiPjiRl",
                "_ZN11CvxCompress10DecompressERiS0_S0_Pjl", "_ZN11CvxCompress10DecompressEPfiiiPjil",
                "_ZN11CvxCompress10DecompressEPfiiiPjl", "_ZN11CvxCompress19Is_Valid_Block_SizeEiii",
                "_ZN11CvxCompress16Run_Module_TestsEbb", "_ZTV11CvxCompress"):
        assert sym in out, sym


def test_host_only_entry_points():
    L = _lib.lib()
    assert L.cvxgpu_is_valid_block_size(32, 32, 32) == 1
    assert L.cvxgpu_is_valid_block_size(32, 32, 1) == 1           # 2-D mode (CvxCompress.cpp:246-248)
    assert L.cvxgpu_is_valid_block_size(24, 32, 32) == 0
    assert L.cvxgpu_is_valid_block_size(4, 8, 8) == 0
    assert L.cvxgpu_is_valid_block_size(512, 8, 8) == 0
    assert L.cvxgpu_num_blocks(1056, 1248, 960, 32, 32, 32) == 38610
    assert L.cvxgpu_max_compressed_bytes(64, 64, 64, 32, 32, 32, 0) == 32 + 8 * 8 + 4 * 32768 * 8 + 7
    assert L.cvxgpu_max_compressed_bytes(64, 64, 64, 32, 32, 32, 1) == 32 + 12 * 8 + 4 * 32768 * 8 + 7
    assert L.cvxgpu_set_transform(7) == _lib.ERR_ARG
    assert b"transform" in L.cvxgpu_last_error()


def test_no_cpu_fallback_without_a_device():
    """without a CUDA device every compute entry fails loudly with CVXGPU_ERR_CUDA"""
    L = _lib.lib()
    if L.cvxgpu_device_count() > 0:
        pytest.skip("a GPU is present")
    import numpy as np
    vol = np.zeros(8 * 8 * 8, np.float32)
    out = np.zeros(8192, np.uint8)
    ln, ratio = C.c_long(0), C.c_float(0)
    rc = L.cvxgpu_compress(vol.ctypes.data, 8, 8, 8, 8, 8, 8, 1e-2, 0, out.ctypes.data, C.byref(ln), C.byref(ratio))
    assert rc == _lib.ERR_CUDA
    assert L.cvxgpu_last_error()


def test_product_and_reference_libraries_do_not_interpose():
    """both libraries define CvxCompress::*; each is linked -Bsymbolic and loaded RTLD_LOCAL so that the reference
    build (the checker) never runs product code and vice versa, whatever the load order"""
    import oracle_libs as ol
    for path in (_lib.LIB_PATH,) + ((ol.REF_SO,) if ol.have_ref() else ()):
        dyn = subprocess.run(["readelf", "-d", path], stdout=subprocess.PIPE, text=True, check=True).stdout
        assert "SYMBOLIC" in dyn, path
    if not ol.have_ref():
        pytest.skip("oracle/_ref not built")
    _lib.lib()                                     # product library first: the order bench.py uses
    import numpy as np
    from cvxcompress_b200 import synth
    vol = synth.g2(24, 16, 16)
    s, _ = ol.ref_compress(vol, 8, 8, 8, 1e-2)     # would abort (no GPU) or run GPU code if interposed
    want, _ = ol.oracle_compress(vol, 8, 8, 8, 1e-2, False)
    assert bytes(s[:len(s) - 7]) == bytes(want[:len(want) - 7])

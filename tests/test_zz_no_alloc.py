"""cvxgpu_sumsq_rows_static_device: the per-block-row sum of squares with its scratch in a static device array of the
library (no cudaMalloc, no workspace), through the C ABI, against cvxgpu_sumsq_rows_device.  Same kernels and summation
order, so the tolerance is the one the EXACT-mode tests use: bit identity.  Collected last on purpose;
passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

MAX_ROWS = 4096          # CVXGPU_STATIC_MAX_ROWS (include/cvx_gpu.h)


def test_static_entry_refuses_more_rows_than_its_scratch_holds():
    """Size check comes before any CUDA call: runs without a GPU.  The pointers are never dereferenced."""
    from cvxcompress_b200 import _lib
    L = _lib.lib()
    fake = C.c_void_p(0x1000)
    rc = L.cvxgpu_sumsq_rows_static_device(fake, 64, 1, MAX_ROWS + 1, fake, None)
    assert rc == -3                                              # CVXGPU_ERR_CAPACITY
    msg = L.cvxgpu_last_error().decode()
    assert "4097 block rows" in msg and "4096" in msg and "cvxgpu_sumsq_rows_device" in msg
    assert L.cvxgpu_sumsq_rows_static_device(None, 64, 1, 8, fake, None) == -1          # CVXGPU_ERR_ARG
    assert L.cvxgpu_sumsq_rows_static_device(fake, 64, 0, 8, fake, None) == -1


@pytest.mark.gpu
def test_static_entry_matches_workspace_entry_on_two_calls_in_a_row():
    import torch
    from cvxcompress_b200 import _lib
    assert torch.cuda.is_available()
    L = _lib.lib()
    dev = torch.device("cuda", 0)
    st = lambda: C.c_void_p(torch.cuda.current_stream().cuda_stream)
    g = torch.Generator(device="cpu").manual_seed(20261105)
    # (plane cells, bz, nz, element offset of the base pointer): ragged last row, a base that is not 16-byte aligned,
    # a row shorter than one group per thread, many rows -- and the first case again at the end (stale partials of a
    # LARGER call sit in the array by then)
    cases = [(352 * 416, 32, 70, 0), (1000 * 77, 8, 333, 1), (37, 4, 9, 3), (96 * 64, 1, 2500, 0), (352 * 416, 32, 70, 0)]
    vols, wants, gots = [], [], []
    for plane, bz, nz, off in cases:
        v = (torch.randn(plane * nz + off, generator=g) * 1e3).to(dev)
        rows = (nz + bz - 1) // bz
        vols.append(v)
        wants.append(torch.zeros(rows, dtype=torch.float64, device=dev))
        gots.append(torch.full((rows,), -1.0, dtype=torch.float64, device=dev))
    # the existing entry point first (grows its workspace with cudaMalloc as it likes)
    for (plane, bz, nz, off), v, w in zip(cases, vols, wants):
        _lib.check(L.cvxgpu_sumsq_rows_device(None, v.data_ptr() + 4 * off, plane, bz, nz, w.data_ptr(), st()))
    torch.cuda.synchronize()
    # the static entry point, call after call on one stream, every buffer allocated beforehand: the device's free memory
    # must not move from the second call on (the first may load the module lazily)
    free_after_first = None
    for k, ((plane, bz, nz, off), v, o) in enumerate(zip(cases, vols, gots)):
        _lib.check(L.cvxgpu_sumsq_rows_static_device(v.data_ptr() + 4 * off, plane, bz, nz, o.data_ptr(), st()))
        if k == 0:
            torch.cuda.synchronize()
            free_after_first = torch.cuda.mem_get_info(dev)[0]
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(dev)[0] == free_after_first
    for k, (w, o) in enumerate(zip(wants, gots)):
        assert torch.equal(w.view(torch.int64), o.view(torch.int64)), k
    # and both agree with numpy's fp64 sum to rounding (the order of summation differs)
    plane, bz, nz, off = cases[0]
    h = vols[0][off:].cpu().numpy().astype(np.float64).reshape(nz, plane)
    ref = np.array([np.sum(h[r * bz:(r + 1) * bz] ** 2) for r in range((nz + bz - 1) // bz)])
    assert np.allclose(gots[0].cpu().numpy(), ref, rtol=1e-12)
    # the limit itself is usable: CVXGPU_STATIC_MAX_ROWS rows of one plane each
    plane = 515
    v = torch.randn(plane * MAX_ROWS, generator=g).to(dev)
    w = torch.zeros(MAX_ROWS, dtype=torch.float64, device=dev)
    o = torch.zeros(MAX_ROWS, dtype=torch.float64, device=dev)
    _lib.check(L.cvxgpu_sumsq_rows_device(None, v.data_ptr(), plane, 1, MAX_ROWS, w.data_ptr(), st()))
    _lib.check(L.cvxgpu_sumsq_rows_static_device(v.data_ptr(), plane, 1, MAX_ROWS, o.data_ptr(), st()))
    torch.cuda.synchronize()
    assert torch.equal(w.view(torch.int64), o.view(torch.int64))

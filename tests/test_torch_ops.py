"""torch.ops.cvx.compress / decompress (cvxcompress_b200/torch_ops.py): registration, shape inference
and the absence of a CPU path run anywhere; the GPU tests compare the ops with the oracle byte for byte."""
import numpy as np
import pytest
import torch

import oracle_libs as ol
from cvxcompress_b200 import api, synth
from cvxcompress_b200 import torch_ops  # noqa: F401  (registers the ops)


def test_ops_are_registered_with_fake_kernels():
    from torch._subclasses.fake_tensor import FakeTensorMode
    assert hasattr(torch.ops.cvx, "compress") and hasattr(torch.ops.cvx, "decompress")
    with FakeTensorMode():
        vol = torch.empty((40, 50, 70), dtype=torch.float32)
        stream, length = torch.ops.cvx.compress(vol, 16, 16, 16, 1e-3, True)
        assert stream.dtype == torch.uint8 and stream.numel() == api.max_compressed_bytes(70, 50, 40, 16, 16, 16, True) + 8
        assert length.dtype == torch.int64 and length.shape == (1,)
        back = torch.ops.cvx.decompress(stream, 70, 50, 40)
        assert back.shape == (40, 50, 70) and back.dtype == torch.float32


def test_cpu_tensors_are_refused():
    with pytest.raises(RuntimeError, match="no CPU implementation"):
        torch.ops.cvx.compress(torch.zeros((8, 8, 8)), 8, 8, 8, 1e-2, False)
    with pytest.raises(RuntimeError, match="no CPU implementation"):
        torch.ops.cvx.decompress(torch.zeros(64, dtype=torch.uint8), 8, 8, 8)


@pytest.mark.gpu
@pytest.mark.parametrize("dims,blk,scale,local", [((96, 80, 72), (32, 32, 32), 1e-2, False), ((70, 50, 40), (16, 16, 16), 1e-3, True)])
def test_ops_equal_the_oracle(dims, blk, scale, local):
    nx, ny, nz = dims
    api.set_transform("exact")
    vol = synth.g2(nx, ny, nz)
    d = torch.from_numpy(vol).cuda()
    stream, length = torch.ops.cvx.compress(d, *blk, scale, local)
    ln = int(length.item())
    want, _ = ol.oracle_compress(vol, *blk, scale, local)
    got = stream[:ln].cpu().numpy()
    assert ln == len(want) and bytes(got[:-7]) == bytes(want[:-7])
    rec = torch.ops.cvx.decompress(stream, nx, ny, nz)                 # the whole capacity tensor: reads are bounded by it
    ref = ol.oracle_decompress(want, nx, ny, nz)
    assert np.array_equal(rec.cpu().numpy().view(np.uint32), ref.view(np.uint32))
    # stream-ordered on a side stream, no host synchronisation between the two ops
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        s2, l2 = torch.ops.cvx.compress(d, *blk, scale, local)
        r2 = torch.ops.cvx.decompress(s2, nx, ny, nz)
    side.synchronize()
    assert int(l2.item()) == ln and torch.equal(s2[:ln - 7], stream[:ln - 7]) and torch.equal(r2, rec)


@pytest.mark.gpu
def test_op_rejects_bad_block_size_and_wrong_dims():
    d = torch.zeros((16, 16, 16), device="cuda")
    with pytest.raises(ValueError, match="block size"):
        torch.ops.cvx.compress(d, 24, 8, 8, 1e-2, False)
    stream, _ = torch.ops.cvx.compress(d, 8, 8, 8, 1e-2, False)
    with pytest.raises(Exception, match="expects|holds"):
        torch.ops.cvx.decompress(stream, 16, 16, 32)


@pytest.mark.gpu
def test_compress_op_replays_in_a_cuda_graph():
    """Compress has no host synchronisation, so a small-volume checkpoint loop can be captured once and replayed
    (six launches per call otherwise dominate small volumes)."""
    api.set_transform("exact")
    nx, ny, nz = 96, 64, 64
    a = torch.from_numpy(synth.g2(nx, ny, nz)).cuda()
    b = torch.from_numpy(synth.g2(nx, ny, nz)[::-1].copy()).cuda()
    static_in = a.clone()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):                       # warm-up outside capture: workspace allocations happen here
        for _ in range(2):
            torch.ops.cvx.compress(static_in, 32, 32, 32, 1e-2, False)
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        stream, length = torch.ops.cvx.compress(static_in, 32, 32, 32, 1e-2, False)
    for vol in (a, b, a):
        static_in.copy_(vol)
        g.replay()
        torch.cuda.synchronize()
        want, _ = ol.oracle_compress(vol.cpu().numpy(), 32, 32, 32, 1e-2, False)
        ln = int(length.item())
        assert ln == len(want) and bytes(stream[:ln - 7].cpu().numpy()) == bytes(want[:-7])


@pytest.mark.gpu
def test_round_trip_in_one_cuda_graph_with_caller_sized_buffer():
    """compress_into (a buffer of a quarter of the volume, not the worst case) + decompress_known (header by value): neither
    synchronises with the host, so a whole checkpoint round trip replays from one CUDA graph."""
    api.set_transform("exact")
    nx, ny, nz = 96, 64, 64
    a = torch.from_numpy(synth.g2(nx, ny, nz)).cuda()
    b = torch.from_numpy(synth.g2(nx, ny, nz)[::-1].copy()).cuda()
    static_in = a.clone()
    buf = torch.empty(a.numel() + 4096, dtype=torch.uint8, device="cuda")      # N bytes: a quarter of the volume
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        torch.ops.cvx.compress_into(static_in, 32, 32, 32, 1e-2, False, buf)
        torch.ops.cvx.decompress_known(buf, nx, ny, nz, 32, 32, 32, False)
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        length = torch.ops.cvx.compress_into(static_in, 32, 32, 32, 1e-2, False, buf)
        back = torch.ops.cvx.decompress_known(buf, nx, ny, nz, 32, 32, 32, False)
    for vol in (a, b):
        static_in.copy_(vol)
        g.replay()
        torch.cuda.synchronize()
        want, _ = ol.oracle_compress(vol.cpu().numpy(), 32, 32, 32, 1e-2, False)
        ln = int(length.item())
        assert ln == len(want) and bytes(buf[:ln - 7].cpu().numpy()) == bytes(want[:-7])
        ref = ol.oracle_decompress(want, nx, ny, nz)
        assert np.array_equal(back.cpu().numpy().view(np.uint32), ref.view(np.uint32))
    # a buffer that is too small: negative length (CVXGPU_ERR_CAPACITY), nothing written past it
    small = torch.zeros(4096 + 64, dtype=torch.uint8, device="cuda")
    guard = small[4096:].clone()
    ln = torch.ops.cvx.compress_into(a, 32, 32, 32, 1e-2, False, small[:4096])
    assert int(ln.item()) == -3 and torch.equal(small[4096:], guard)

"""z-slab sharding (cvxcompress_b200/sharded.py) on CPU: world_size 2 and 3 over gloo.

The product compute backend needs a GPU; here the host logic -- slab ranges, the two all-gathers, the rank-order
sum of squares, offset bases, stream assembly, per-rank decode of a slab -- runs with a CHECKER backend made of the
oracle's per-stage functions, and the assembled stream must equal the oracle's single-process stream byte for byte.
"""
import ctypes as C
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle_libs as ol
from cvxcompress_b200 import sharded, synth


class OracleBackend:
    """sharded.py's backend interface on the CPU, built from oracle/ stage functions (test infrastructure)."""

    def sumsq(self, slab):
        v = slab.numpy().astype(np.float64)
        return torch.tensor([float(np.sum(v * v))], dtype=torch.float64)

    def sumsq_rows(self, slab, g):
        v = slab.numpy().astype(np.float64)
        rows = [float(np.sum(v[i * g.bz:(i + 1) * g.bz] ** 2)) for i in range(g.b1 - g.b0)]
        return torch.tensor(rows, dtype=torch.float64)

    @staticmethod
    def _glob_mulfac(g, scale, local, sumsq_total):
        rms = np.float32(1.0) if local else np.float32(np.sqrt(float(sumsq_total.item()) / g.ncells))
        m = np.float32(1.0) / (rms * np.float32(scale)) if rms != 0 else np.float32(1.0)       # CvxCompress.cpp:291-295
        return m if np.isfinite(m) else np.float32(1.0)

    def header(self, g, scale, local, sumsq_total):
        h = np.zeros(8, np.int32)
        h[:6] = (g.nx, g.ny, g.nz, g.bx, g.by, g.bz)
        h[6] = np.float32(self._glob_mulfac(g, scale, local, sumsq_total)).view(np.int32)
        h[7] = 1 if local else 0
        return torch.from_numpy(h.view(np.uint8).copy())

    def encode_blocks(self, slab, g, scale, local, sumsq_total, offs, bmf, payload, nbytes, status):
        O = ol.oracle()
        vol = np.ascontiguousarray(slab.numpy()).reshape(-1)
        B = g.block_cells
        gm = self._glob_mulfac(g, scale, local, sumsq_total)
        out = payload.numpy()
        pos, k = 0, 0
        blk = np.zeros(B, np.float32)
        enc = np.zeros(5 * B + 64, np.uint8)
        nbz = -(-g.nz_slab // g.bz)
        for iz in range(nbz):
            for iy in range(g.nby):
                for ix in range(g.nbx):
                    O.cvxo_copy_to_block(vol, ix * g.bx, iy * g.by, iz * g.bz, g.nx, g.ny, g.nz_slab, blk, g.bx, g.by, g.bz)
                    O.cvxo_wavelet_forward(blk, g.bx, g.by, g.bz, 0)
                    mf = gm
                    if local:
                        rms = np.float32(O.cvxo_local_rms(blk, g.bx, g.by, g.bz))
                        mf = np.float32(1.0) / (rms * np.float32(scale)) if rms != 0 else np.float32(1.0)
                        bmf[k] = float(mf)
                    n = O.cvxo_rle_encode(np.float32(mf), blk, B, enc)
                    if n > 4 * B:                                                  # raw fallback, CvxCompress.cpp:350-360
                        out[pos:pos + 4 * B] = blk.view(np.uint8)
                        offs[k] = pos | (-1 << 63)
                        pos += 4 * B
                    else:
                        out[pos:pos + n] = enc[:n]
                        offs[k] = pos
                        pos += n
                    k += 1
        nbytes[0] = pos

    def decode_blocks(self, slab, g, glob_mulfac, local, offs, bmf, payload, payload_bytes):
        O = ol.oracle()
        vol = slab.numpy().reshape(-1)
        B = g.block_cells
        pay = np.concatenate([payload.numpy()[:payload_bytes], np.zeros(64, np.uint8)])
        blk = np.zeros(B, np.float32)
        k = 0
        nbz = -(-g.nz_slab // g.bz)
        for iz in range(nbz):
            for iy in range(g.nby):
                for ix in range(g.nbx):
                    o = int(offs[k])
                    raw = o < 0
                    o &= (1 << 63) - 1
                    if raw:
                        blk[:] = pay[o:o + 4 * B].view(np.float32)
                    else:
                        mf = np.float32(bmf[k]) if local else np.float32(glob_mulfac)
                        O.cvxo_rle_decode(mf, blk, B, pay[o:])
                    O.cvxo_wavelet_inverse(blk, g.bx, g.by, g.bz, 0)
                    O.cvxo_copy_from_block(blk, g.bx, g.by, g.bz, vol, ix * g.bx, iy * g.by, iz * g.bz, g.nx, g.ny, g.nz_slab)
                    k += 1


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, case, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        nx, ny, nz, bx, by, bz, scale, local = case
        vol = synth.g2(nx, ny, nz)
        g = sharded.Geometry(nx, ny, nz, bx, by, bz, rank, world)
        slab = torch.from_numpy(np.ascontiguousarray(vol[g.z0:g.z1]))
        be = OracleBackend()
        cap = 4 * g.nblk_slab * g.block_cells + 5 * g.block_cells + 4096          # every block raw + coder overshoot
        ss = sharded.compress_slab(slab, nx, ny, nz, bx, by, bz, scale, local, payload=torch.zeros(cap, dtype=torch.uint8), backend=be)
        ss.check()
        full = sharded.assemble(ss).numpy()
        want, _ = ol.oracle_compress(vol, bx, by, bz, scale, local)
        ok_stream = len(full) == len(want) and bytes(full[:len(full) - 7]) == bytes(want[:len(want) - 7])
        # every rank decodes its own slab from its shard, and from the assembled full stream
        rec = sharded.decompress_slab(ss, backend=be).numpy()
        ref = ol.oracle_decompress(want, nx, ny, nz)[g.z0:g.z1]
        ok_dec = np.array_equal(rec.view(np.uint32), ref.view(np.uint32))
        rec2, _ = sharded.decompress_slab_from_stream(torch.from_numpy(full), len(full), backend=be)
        ok_dec2 = np.array_equal(rec2.numpy().view(np.uint32), ref.view(np.uint32))
        q.put((rank, ok_stream, ok_dec, ok_dec2, ss.total_length(), len(want)))
    finally:
        dist.destroy_process_group()


CASES = [
    (40, 24, 80, 8, 8, 8, 1e-2, False),        # 10 block rows: 5 + 5, or 3 + 3 + 4
    (40, 24, 44, 8, 8, 16, 1e-3, True),        # local RMS, partial last block row, 3 block rows: uneven split
    (37, 19, 24, 16, 8, 8, 1e-7, False),       # raw-fallback blocks (tiny scale), nx not a multiple of 4
]


@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("case", CASES, ids=lambda c: "%dx%dx%d-b%dx%dx%d-s%g-l%d" % c)
def test_sharded_stream_equals_single_process_oracle(world, case):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, case, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, ok_stream, ok_dec, ok_dec2, total, want_len in sorted(res):
        assert total == want_len, (rank, total, want_len)
        assert ok_stream, "rank %d: assembled stream differs from the single-process oracle stream" % rank
        assert ok_dec, "rank %d: slab decoded from its shard differs" % rank
        assert ok_dec2, "rank %d: slab decoded from the assembled stream differs" % rank


def test_slab_ranges_cover_all_block_rows():
    for nbz in (1, 3, 30, 40):
        for world in (1, 2, 4, 8):
            cover = []
            for r in range(world):
                b0, b1 = sharded.slab_range(nbz, r, world)
                cover += list(range(b0, b1))
            assert cover == list(range(nbz))
    # BASELINE config 3: nbz = 30 -> 8 ranks own 3 or 4 block rows
    sizes = [sharded.slab_range(30, r, 8)[1] - sharded.slab_range(30, r, 8)[0] for r in range(8)]
    assert sorted(set(sizes)) == [3, 4] and sum(sizes) == 30

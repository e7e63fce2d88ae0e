"""z-slab sharding of Compress / Decompress over the ranks of a torch.distributed group.

Blocks are independent (zero padding + in-block mirror extension, Block_Copy.cpp:92-115,
Wavelet_Transform_Slow.cpp:59-67), so any partition of whole blocks is valid.  Rank r of G owns the
block rows [r*nbz//G, (r+1)*nbz//G): a contiguous iBlk range and a contiguous z-slab of the volume.
The reference parallelises ONE volume over its workers (CvxCompress.cpp:318, :528); the only couplings are two scalars:

  * the global RMS (CvxCompress.cpp:73-117, :249)       -> all-gather of the per-BLOCK-ROW fp64 sums of squares,
                                                           added in row order on every rank: glob_mulfac, hence
                                                           the whole stream, is identical for every G
  * the payload offsets (CvxCompress.cpp:365-390, :302) -> all-gather of one int64 payload size per rank,
                                                           exclusive prefix = the rank's base offset

Decompress needs no collective.  No data-path collective exists; `assemble()` (optional) gathers the
shards into one contiguous reference-format stream.

Two drivers of the same protocol:
  * the product path (CUDA slabs, no `backend=`): `NcclShard`, a ctypes binding of the library's own C entry points
    cvxgpu_compress_sharded / cvxgpu_decompress_sharded (csrc/cvx_shard.cpp: kernels + two ncclAllGather on the
    caller's stream, no host synchronisation, CUDA-graph capturable).  torch.distributed only ships the 128-byte
    NCCL unique id once;
  * the host-logic path (`backend=` given): the same steps over torch.distributed collectives with the compute behind a
    small backend interface -- tests/ plug in a CPU checker backend to exercise geometry, ordering and assembly under
    gloo without a GPU.  Nothing here imports oracle/.
"""
import ctypes as C

import torch
import torch.distributed as dist


def slab_range(nbz, rank, world):
    """Block rows [b0, b1) owned by `rank` (floor(r*nbz/G) .. floor((r+1)*nbz/G))."""
    return (rank * nbz) // world, ((rank + 1) * nbz) // world


class Geometry:
    """Global volume / block grid and this rank's share of it."""

    def __init__(self, nx, ny, nz, bx, by, bz, rank, world):
        self.nx, self.ny, self.nz, self.bx, self.by, self.bz = nx, ny, nz, bx, by, bz
        self.rank, self.world = rank, world
        self.nbx, self.nby, self.nbz = -(-nx // bx), -(-ny // by), -(-nz // bz)
        self.nblk = self.nbx * self.nby * self.nbz
        self.b0, self.b1 = slab_range(self.nbz, rank, world)
        self.z0 = min(self.b0 * bz, nz)
        self.z1 = min(self.b1 * bz, nz)
        self.nz_slab = self.z1 - self.z0
        self.blk_first = self.b0 * self.nbx * self.nby
        self.nblk_slab = (self.b1 - self.b0) * self.nbx * self.nby
        self.ncells = nx * ny * nz
        self.block_cells = bx * by * bz

    def table_bytes(self, local):
        return 8 * self.nblk + (4 * self.nblk if local else 0)


class CudaBackend:
    """The product compute: sm_100a kernels through the C ABI.  Fails loudly without the library / a GPU."""

    def __init__(self):
        from . import _lib
        self._lib = _lib
        self.L = _lib.lib()

    @staticmethod
    def _st():
        return C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def sumsq(self, slab):
        out = torch.zeros(1, dtype=torch.float64, device=slab.device)
        if slab.numel():
            self._lib.check(self.L.cvxgpu_sumsq_device(None, slab.data_ptr(), slab.numel(), out.data_ptr(), self._st()))
        return out

    def sumsq_rows(self, slab, g):
        """fp64 sum of squares of every block row of the slab (cvxgpu_sumsq_rows_device): tensor[b1 - b0]."""
        n = g.b1 - g.b0
        out = torch.zeros(max(n, 1), dtype=torch.float64, device=slab.device)
        if slab.numel():
            self._lib.check(self.L.cvxgpu_sumsq_rows_device(None, slab.data_ptr(), g.nx * g.ny, g.bz, g.nz_slab, out.data_ptr(), self._st()))
        return out[:n]

    def header(self, g, scale, local, sumsq_total):
        h = torch.zeros(32, dtype=torch.uint8, device=sumsq_total.device)
        self._lib.check(self.L.cvxgpu_write_header_device(h.data_ptr(), g.nx, g.ny, g.nz, g.bx, g.by, g.bz, float(scale),
                                                          int(local), sumsq_total.data_ptr(), g.ncells, self._st()))
        return h

    def encode_blocks(self, slab, g, scale, local, sumsq_total, offs, bmf, payload, nbytes, status):
        self._lib.check(self.L.cvxgpu_encode_blocks_device(
            None, slab.data_ptr(), g.nx, g.ny, g.nz_slab, g.bx, g.by, g.bz, float(scale), int(local),
            sumsq_total.data_ptr(), g.ncells, offs.data_ptr(), bmf.data_ptr() if bmf is not None else None,
            payload.data_ptr(), payload.numel(), nbytes.data_ptr(), status.data_ptr(), self._st()))

    def decode_blocks(self, slab, g, glob_mulfac, local, offs, bmf, payload, payload_bytes):
        self._lib.check(self.L.cvxgpu_decode_blocks_device(
            None, slab.data_ptr(), g.nx, g.ny, g.nz_slab, g.bx, g.by, g.bz, float(glob_mulfac), int(local),
            offs.data_ptr(), bmf.data_ptr() if bmf is not None else None, payload.data_ptr(), int(payload_bytes), self._st()))


class ShardedStream:
    """One rank's share of a compressed stream: the 32-byte header (same on every rank), this rank's
    slice of the offset table (offsets already global: local offset + base) and of the mulfac table,
    and this rank's payload bytes, which start at byte `base` of the global payload."""

    def __init__(self, geom, local, header, offs, bmf, payload, nbytes, bases, status):
        self.geom, self.local = geom, local
        self.header, self.offs, self.bmf, self.payload = header, offs, bmf, payload
        self.nbytes = nbytes          # int64[1] device: this rank's payload size
        self.bases = bases            # int64[world + 1] device: exclusive prefix of all ranks' payload sizes
        self.status = status          # int32[1] device: 0 or CVXGPU_ERR_CAPACITY

    def total_length(self):
        """compressed_length = 32 + 8*nnn [+ 4*nnn] + payload + 7 (CvxCompress.cpp:421-422).  Synchronises."""
        return 32 + self.geom.table_bytes(self.local) + int(self.bases[-1].item()) + 7

    def host_meta(self):
        """(glob_mulfac, payload bytes of this rank) as host values; read once (synchronises) and cached so that
        repeated decodes of the same stream stay asynchronous."""
        if getattr(self, "_meta", None) is None:
            b = self.bases.cpu()
            self._meta = (float(self.header[24:28].view(torch.float32).item()), int(b[self.geom.rank + 1] - b[self.geom.rank]))
        return self._meta

    def check(self):
        if int(self.status.item()) != 0:
            raise RuntimeError("rank %d: payload buffer too small for its slab" % self.geom.rank)


def _world(group):
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


def _all_gather_scalar(x, world, group):
    """x: tensor[1] on this rank -> tensor[world] (rank order)."""
    if world == 1:
        return x.clone()
    out = torch.empty(world, dtype=x.dtype, device=x.device)
    dist.all_gather_into_tensor(out, x, group=group)
    return out


class NcclShard:
    """The product multi-GPU driver: this rank's cvxgpu_shard (NCCL communicator created inside the library from a unique
    id that rank 0 draws and torch.distributed ships once), and one C call per direction."""

    _cache = {}

    def __init__(self, group=None):
        from . import _lib
        self._lib = _lib
        self.L = _lib.lib()
        self.rank, self.world = _world(group)
        self.device = torch.device("cuda", torch.cuda.current_device())
        ident = torch.zeros(128, dtype=torch.uint8)
        if self.world > 1:
            if self.rank == 0:
                buf = (C.c_ubyte * 128)()
                _lib.check(self.L.cvxgpu_shard_unique_id(buf))
                ident = torch.frombuffer(bytearray(buf), dtype=torch.uint8).clone()
            on_gpu = dist.get_backend(group) == "nccl"
            t = ident.to(self.device) if on_gpu else ident
            dist.broadcast(t, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
            ident = t.cpu()
        self.handle = C.c_void_p()
        idbuf = (C.c_ubyte * 128).from_buffer_copy(bytes(ident.numpy().tobytes()))
        with torch.cuda.device(self.device):
            _lib.check(self.L.cvxgpu_shard_create(C.byref(self.handle), idbuf if self.world > 1 else None, None, self.rank, self.world))

    @classmethod
    def get(cls, group=None):
        key = (id(group), torch.cuda.current_device())
        if key not in cls._cache:
            cls._cache[key] = cls(group)
        return cls._cache[key]

    def close(self):
        if self.handle:
            self.L.cvxgpu_shard_destroy(self.handle)
            self.handle = C.c_void_p()

    @staticmethod
    def _st():
        return C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def compress(self, slab, g, scale, local, payload=None, out=None):
        dev = slab.device
        n = g.nblk_slab
        if out is None:
            header = torch.zeros(32, dtype=torch.uint8, device=dev)
            offs = torch.zeros(max(n, 1), dtype=torch.int64, device=dev)
            bmf = torch.zeros(max(n, 1), dtype=torch.float32, device=dev) if local else None
            bases = torch.zeros(self.world + 1, dtype=torch.int64, device=dev)
            status = torch.zeros(1, dtype=torch.int32, device=dev)
        else:                                   # reuse a previous ShardedStream's buffers (steady-state loops, graph capture)
            header, offs, bmf, bases, status = out.header, out._offs_buf, out._bmf_buf, out.bases, out.status
            payload = out.payload if payload is None else payload
        if payload is None:
            payload = torch.empty(default_payload_capacity(g, local), dtype=torch.uint8, device=dev)
        self._lib.check(self.L.cvxgpu_compress_sharded(
            self.handle, slab.data_ptr() if n else None, g.nx, g.ny, g.nz, g.bx, g.by, g.bz, float(scale), int(local),
            header.data_ptr(), offs.data_ptr(), bmf.data_ptr() if local else None, payload.data_ptr(), payload.numel(),
            bases.data_ptr(), status.data_ptr(), self._st()))
        ss = ShardedStream(g, local, header, offs[:n], bmf[:n] if local else None, payload, None, bases, status)
        ss._offs_buf, ss._bmf_buf, ss.driver = offs, bmf, self
        return ss

    def decompress(self, ss, out):
        g = ss.geom
        if g.nblk_slab == 0:
            return out
        self._lib.check(self.L.cvxgpu_decompress_sharded(
            self.handle, out.data_ptr(), g.nx, g.ny, g.nz, g.bx, g.by, g.bz, int(ss.local), ss.header.data_ptr(),
            ss.offs.data_ptr(), ss.bmf.data_ptr() if ss.local else None, ss.payload.data_ptr(), ss.bases.data_ptr(), self._st()))
        return out


def default_payload_capacity(g, local):
    """Bytes reserved for a rank's payload when the caller gives no buffer: the worst case (every block of the slab raw,
    cvxgpu_max_compressed_bytes minus header and tables) so that no data set can overflow it."""
    return 4 * g.block_cells * max(g.nblk_slab, 1) + 64


def compress_slab(slab, nx, ny, nz, bx, by, bz, scale, use_local_rms=False, payload=None, backend=None, group=None, out=None):
    """Compress this rank's z-slab (tensor [nz_slab, ny, nx]) of the global nx*ny*nz volume.
    Everything is enqueued on the current stream; no host synchronisation.  Returns a ShardedStream.
    `payload`: this rank's payload buffer (uint8); when omitted the worst case is allocated.  Call .check() (or
    assemble(), which does) before trusting a stream written into a smaller buffer."""
    rank, world = _world(group)
    g = Geometry(nx, ny, nz, bx, by, bz, rank, world)
    assert tuple(slab.shape) == (g.nz_slab, ny, nx), "rank %d expects a slab of %d planes" % (rank, g.nz_slab)
    local = bool(use_local_rms)
    if backend is None:
        return NcclShard.get(group).compress(slab, g, scale, local, payload=payload, out=out)
    be = backend
    dev = slab.device
    # (1) global RMS: per-block-row sums of squares, all-gather (padded to the largest rank), added in ROW order
    if local:
        sumsq_total = torch.zeros(1, dtype=torch.float64, device=dev)
    else:
        counts = [slab_range(g.nbz, r, world)[1] - slab_range(g.nbz, r, world)[0] for r in range(world)]
        seg = max(max(counts), 1)
        mine = torch.zeros(seg, dtype=torch.float64, device=dev)
        rows = be.sumsq_rows(slab, g)
        mine[:rows.numel()] = rows
        if world > 1:
            allrows = torch.empty(world * seg, dtype=torch.float64, device=dev)
            dist.all_gather_into_tensor(allrows, mine, group=group)
        else:
            allrows = mine
        sumsq_total = torch.zeros(1, dtype=torch.float64, device=dev)
        for r in range(world):
            for i in range(counts[r]):
                sumsq_total += allrows[r * seg + i]
    header = be.header(g, scale, local, sumsq_total)
    # (2) this rank's blocks -> local table slice + local payload
    n = g.nblk_slab
    offs = torch.zeros(max(n, 1), dtype=torch.int64, device=dev)
    bmf = torch.zeros(max(n, 1), dtype=torch.float32, device=dev) if local else None
    if payload is None:
        payload = torch.empty(default_payload_capacity(g, local), dtype=torch.uint8, device=dev)
    nbytes = torch.zeros(1, dtype=torch.int64, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    if n:
        be.encode_blocks(slab, g, scale, local, sumsq_total, offs, bmf, payload, nbytes, status)
    # (3) payload sizes -> exclusive prefix -> base of this rank; offsets become global.  Adding a base < 2^63 as
    #     int64 leaves the raw flag (bit 63, CvxCompress.cpp:382-384) untouched.
    sizes = _all_gather_scalar(nbytes, world, group)
    bases = torch.zeros(world + 1, dtype=torch.int64, device=dev)
    bases[1:] = torch.cumsum(sizes, 0)
    if n:
        offs[:n] += bases[rank]
    ss = ShardedStream(g, local, header, offs[:n], bmf[:n] if local else None, payload, nbytes, bases, status)
    ss.driver = None
    return ss


def decompress_slab(ss, out=None, backend=None):
    """Decode this rank's blocks of a ShardedStream into its z-slab [nz_slab, ny, nx].  No collective; on the product
    path no host synchronisation either (glob_mulfac, the payload extent and the offset bias are read on the device)."""
    g = ss.geom
    dev = ss.payload.device
    if out is None:
        out = torch.empty((g.nz_slab, g.ny, g.nx), dtype=torch.float32, device=dev)
    if g.nblk_slab == 0:
        return out
    if backend is None and getattr(ss, "driver", None) is not None:
        return ss.driver.decompress(ss, out)
    ss.check()
    be = backend or CudaBackend()
    glob_mulfac, nbytes = ss.host_meta()                     # the piecewise C ABI takes both by value
    offs_local = ss.offs - ss.bases[g.rank]
    be.decode_blocks(out, g, glob_mulfac, ss.local, offs_local, ss.bmf, ss.payload, nbytes)
    return out


def decompress_slab_from_stream(stream, length, rank=None, world=None, out=None, backend=None, group=None, header=None):
    """Decode this rank's iBlk range from a FULL reference-format stream (any block order, e.g. one written by
    the reference with several threads, CvxCompress.cpp:365-390) held by every rank.  No collective.
    `header`: (nx, ny, nz, bx, by, bz, use_local_rms) when the caller knows it -- the call then does not read the
    stream's first 32 bytes back (no host synchronisation; glob_mulfac is read on the device)."""
    r, w = _world(group)
    rank = r if rank is None else rank
    world = w if world is None else world
    if header is None:
        hdr = stream[:32].cpu()
        nx, ny, nz, bx, by, bz = (int(v) for v in hdr[:24].view(torch.int32))
        local = bool(int(hdr[28:32].view(torch.int32)[0]) & 1)
    else:
        nx, ny, nz, bx, by, bz, local = header
        local = bool(local)
    g = Geometry(nx, ny, nz, bx, by, bz, rank, world)
    if out is None:
        out = torch.empty((g.nz_slab, ny, nx), dtype=torch.float32, device=stream.device)
    if g.nblk_slab == 0:
        return out, g
    if backend is None:
        from . import _lib
        with torch.cuda.device(stream.device):
            _lib.check(_lib.lib().cvxgpu_decompress_slab_device(
                None, out.data_ptr(), nx, ny, nz, bx, by, bz, int(local), rank, world, stream.data_ptr(), int(length),
                C.c_void_p(torch.cuda.current_stream().cuda_stream)))
        return out, g
    be = backend
    glob_mulfac = float(stream[24:28].cpu().view(torch.float32)[0])
    tb = g.table_bytes(local)
    offs = stream[32:32 + 8 * g.nblk].view(torch.int64)[g.blk_first:g.blk_first + g.nblk_slab]
    bmf = None
    if local:
        bmf = stream[32 + 8 * g.nblk:32 + 12 * g.nblk].view(torch.float32)[g.blk_first:g.blk_first + g.nblk_slab]
    payload = stream[32 + tb:]
    be.decode_blocks(out, g, glob_mulfac, local, offs, bmf, payload, int(length) - 32 - tb)
    return out, g


def assemble(ss, group=None):
    """Gather a ShardedStream into ONE contiguous reference-format stream (uint8 tensor of compressed_length
    bytes; the last 7 are slack) on every rank.  Optional: costs one padded all-gather of the payloads."""
    g = ss.geom
    rank, world = _world(group)
    dev = ss.payload.device
    ss.check()                      # a payload buffer that was too small leaves a truncated shard: refuse to assemble it
    bases = ss.bases.cpu()
    total = int(bases[-1])
    tb = g.table_bytes(ss.local)
    out = torch.zeros(32 + tb + total + 7, dtype=torch.uint8, device=dev)
    out[:32] = ss.header
    counts = [Geometry(g.nx, g.ny, g.nz, g.bx, g.by, g.bz, r, world).nblk_slab for r in range(world)]
    sizes = [int(bases[r + 1] - bases[r]) for r in range(world)]
    maxn, maxb = max(max(counts), 1), max(max(sizes), 1)

    def gather(x, n, dtype):
        pad = torch.zeros(n, dtype=dtype, device=dev)
        pad[:x.numel()] = x
        if world == 1:
            return [pad]
        buf = torch.empty(world * n, dtype=dtype, device=dev)
        dist.all_gather_into_tensor(buf, pad, group=group)
        return list(buf.view(world, n))

    offs_all = gather(ss.offs, maxn, torch.int64)
    bmf_all = gather(ss.bmf, maxn, torch.float32) if ss.local else None
    pay_all = gather(ss.payload[:sizes[rank]], maxb, torch.uint8)
    tab = out[32:32 + 8 * g.nblk].view(torch.int64)
    mf = out[32 + 8 * g.nblk:32 + 12 * g.nblk].view(torch.float32) if ss.local else None
    first = 0
    for r in range(world):
        tab[first:first + counts[r]] = offs_all[r][:counts[r]]
        if ss.local:
            mf[first:first + counts[r]] = bmf_all[r][:counts[r]]
        out[32 + tb + int(bases[r]):32 + tb + int(bases[r + 1])] = pay_all[r][:sizes[r]]
        first += counts[r]
    return out

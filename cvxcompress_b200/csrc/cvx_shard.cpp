// cvx_shard.cpp -- z-slab sharding of Compress / Decompress over several GPUs behind the C ABI (include/cvx_gpu.h,
// SURVEY.md 8(e)).  One process (or host thread) per GPU; rank r of G owns the block rows [r*nbz/G, (r+1)*nbz/G) of the
// volume: a contiguous iBlk range and a contiguous z-slab of memory.  The reference parallelises ONE volume over its
// workers (CvxCompress.cpp:318 `omp parallel for` over the blocks, :528 for decode); the only couplings are two scalars:
//   * global RMS (CvxCompress.cpp:73-117)           -> all-gather of the per-BLOCK-ROW sums of squares, added in row order
//                                                      on every rank: glob_mulfac and the stream are identical for any G
//   * payload offsets (CvxCompress.cpp:365-390,:302) -> all-gather of one int64 payload size per rank, exclusive prefix
// The two scalars travel through PEER MEMORY when the ranks can map each other's HBM (one process per GPU on one NVSwitch
// node: CUDA IPC): the producing side pushes its values into every rank's mailbox and the consuming kernel spins on its
// own mailbox's flags (cvx_kernels.cu, ShardMailbox) -- no collective launch on the path; otherwise by two ncclAllGather.
// Everything is enqueued on the caller's stream (kernels and, in the fallback, the two ncclAllGather calls): no host synchronisation, so
// the sequence can be captured in a CUDA graph.  NCCL is bound at run time (dlsym on the process, then libnccl.so.2),
// so the library has no link-time dependency on it and shares the communicator of a host application that already
// has one.  Decompress needs no collective.
#include "../../include/cvx_gpu.h"
#include "cvx_internal.h"

#include <dlfcn.h>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <vector>

using namespace cvx;

int cvx_internal_fail(int code, const char* fmt, ...);      // cvx_api.cpp: sets cvxgpu_last_error() of this thread
int cvx_internal_exact_mode();
#define FAIL cvx_internal_fail
#define CU(call)                                                                                       \
    do {                                                                                               \
        cudaError_t e__ = (call);                                                                      \
        if (e__ != cudaSuccess) return FAIL(CVXGPU_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e__)); \
    } while (0)

// ---- NCCL, bound at run time ---------------------------------------------------------------------------------
namespace {
typedef struct ncclComm* ncclComm_t;
struct ncclUniqueId { char internal[128]; };
constexpr int kNcclUint8 = 1, kNcclInt32 = 2, kNcclInt64 = 4, kNcclFloat64 = 8;         // ncclDataType_t values (nccl.h)
struct Nccl {
    int (*GetUniqueId)(ncclUniqueId*) = nullptr;
    int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    bool ok = false;
    char why[256] = "";
};
Nccl& nccl()
{
    static Nccl N;
    static std::once_flag once;
    std::call_once(once, [] {
        void* h = nullptr;
        auto sym = [&](const char* name) -> void* {
            void* p = dlsym(RTLD_DEFAULT, name);            // a host application (e.g. torch) that already loaded NCCL
            if (!p) {
                if (!h) {
                    const char* env = getenv("CVX_GPU_NCCL_LIB");
                    h = dlopen(env && *env ? env : "libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
                    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
                }
                if (h) p = dlsym(h, name);
            }
            return p;
        };
        N.GetUniqueId = reinterpret_cast<decltype(N.GetUniqueId)>(sym("ncclGetUniqueId"));
        N.CommInitRank = reinterpret_cast<decltype(N.CommInitRank)>(sym("ncclCommInitRank"));
        N.CommDestroy = reinterpret_cast<decltype(N.CommDestroy)>(sym("ncclCommDestroy"));
        N.AllGather = reinterpret_cast<decltype(N.AllGather)>(sym("ncclAllGather"));
        N.GetErrorString = reinterpret_cast<decltype(N.GetErrorString)>(sym("ncclGetErrorString"));
        N.ok = N.GetUniqueId && N.CommInitRank && N.CommDestroy && N.AllGather;
        if (!N.ok) snprintf(N.why, sizeof(N.why), "NCCL not found (dlsym on the process, libnccl.so.2; set CVX_GPU_NCCL_LIB): %s", dlerror() ? dlerror() : "symbol missing");
    });
    return N;
}
int nccl_fail(const char* what, int rc)
{
    Nccl& N = nccl();
    return FAIL(CVXGPU_ERR_CUDA, "%s: NCCL error %d (%s)", what, rc, N.GetErrorString ? N.GetErrorString(rc) : "?");
}
}  // namespace

struct cvxgpu_shard {
    int rank = 0, world = 1;
    ncclComm_t comm = nullptr;
    bool own_comm = false;
    cvxgpu_workspace* ws = nullptr;
    // device scratch of the protocol
    double* d_rows_send = nullptr;      // seg doubles
    double* d_rows_recv = nullptr;      // world * seg doubles
    int* d_cnt = nullptr;               // world ints: block rows of every rank
    int seg = 0;
    long* d_sizes = nullptr;            // world int64: payload bytes of every rank
    long* d_nbytes = nullptr;           // [0] this rank's payload bytes (encode output) [1],[2] extent scratch of decode
    // geometry the scratch was sized for
    int g_nz = -1, g_bz = -1;
    // peer-memory exchange (cvx_kernels.cu: ShardMailbox): this rank's mailbox, every rank's mailbox as mapped here, epochs
    bool p2p = false;
    int mcap = 0;
    unsigned long long* d_mailbox = nullptr;
    unsigned long long** d_peers = nullptr;
    unsigned long long* d_epoch = nullptr;      // [0] row sums, [1] payload sizes
    std::vector<void*> opened;
};

extern "C" int cvxgpu_shard_range(int nz, int bz, int rank, int world, int* z0, int* nz_slab, int* brow0, int* nbrows)
{
    if (nz <= 0 || bz <= 0 || world <= 0 || rank < 0 || rank >= world) return FAIL(CVXGPU_ERR_ARG, "cvxgpu_shard_range: bad argument");
    const long nbz = ((long)nz + bz - 1) / bz;
    const long b0 = (long)rank * nbz / world, b1 = ((long)rank + 1) * nbz / world;
    const long za = b0 * bz < nz ? b0 * bz : nz, zb = b1 * bz < nz ? b1 * bz : nz;
    if (z0) *z0 = (int)za;
    if (nz_slab) *nz_slab = (int)(zb - za);
    if (brow0) *brow0 = (int)b0;
    if (nbrows) *nbrows = (int)(b1 - b0);
    return CVXGPU_OK;
}

extern "C" int cvxgpu_shard_unique_id(void* id128)
{
    if (!id128) return FAIL(CVXGPU_ERR_ARG, "cvxgpu_shard_unique_id: null pointer");
    Nccl& N = nccl();
    if (!N.ok) return FAIL(CVXGPU_ERR_CUDA, "%s", N.why);
    ncclUniqueId id;
    const int rc = N.GetUniqueId(&id);
    if (rc) return nccl_fail("ncclGetUniqueId", rc);
    memcpy(id128, &id, 128);
    return CVXGPU_OK;
}

// 1: the per-rank scalars of a sharded Compress travel through peer-memory mailboxes; 0: by ncclAllGather (or world == 1)
extern "C" int cvxgpu_shard_uses_peer_memory(const cvxgpu_shard* sh) { return sh && sh->world > 1 && sh->p2p ? 1 : 0; }

extern "C" int cvxgpu_shard_destroy(cvxgpu_shard* sh)
{
    if (!sh) return CVXGPU_OK;
    if (sh->d_rows_send) cudaFree(sh->d_rows_send);
    if (sh->d_rows_recv) cudaFree(sh->d_rows_recv);
    if (sh->d_cnt) cudaFree(sh->d_cnt);
    if (sh->d_sizes) cudaFree(sh->d_sizes);
    if (sh->d_nbytes) cudaFree(sh->d_nbytes);
    for (void* q : sh->opened) cudaIpcCloseMemHandle(q);
    if (sh->d_mailbox) cudaFree(sh->d_mailbox);
    if (sh->d_peers) cudaFree(sh->d_peers);
    if (sh->d_epoch) cudaFree(sh->d_epoch);
    if (sh->ws) cvxgpu_workspace_destroy(sh->ws);
    if (sh->own_comm && sh->comm) nccl().CommDestroy(sh->comm);
    delete sh;
    return CVXGPU_OK;
}

// Map every rank's mailbox into this process (CUDA IPC over NVLink / NVSwitch peer access).  Collective over the shard's
// communicator: all ranks use the mailboxes or none does (a rank that cannot map a peer -- several ranks in ONE process,
// no peer access, CVX_GPU_SHARD_EXCHANGE=nccl -- votes no).  Never fails the shard: NCCL remains the fallback.
static void shard_setup_mailboxes(cvxgpu_shard* sh)
{
    Nccl& N = nccl();
    const int W = sh->world, cap = 64;
    const char* env = getenv("CVX_GPU_SHARD_EXCHANGE");
    int ok = !(env && !strcmp(env, "nccl"));
    const size_t bytes = shard_mailbox_bytes(W, cap);
    unsigned char* d_h = nullptr;                // [0, 64): my handle, [64, 64 + 64 W): everybody's, then W ints of votes
    std::vector<unsigned char> h((size_t)64 * W);
    std::vector<int> votes((size_t)W, 0);
    cudaIpcMemHandle_t mine;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handle size");
    memset(&mine, 0, sizeof(mine));
    if (cudaMalloc(&sh->d_mailbox, bytes) != cudaSuccess || cudaMemset(sh->d_mailbox, 0, bytes) != cudaSuccess ||
        cudaMalloc(&sh->d_epoch, 16) != cudaSuccess || cudaMemset(sh->d_epoch, 0, 16) != cudaSuccess ||
        cudaMalloc(&sh->d_peers, sizeof(void*) * W) != cudaSuccess || cudaIpcGetMemHandle(&mine, sh->d_mailbox) != cudaSuccess) { cudaGetLastError(); ok = 0; }
    if (cudaMalloc(&d_h, 64 + 64 * (size_t)W + 8 * (size_t)W) != cudaSuccess) { cudaGetLastError(); return; }   // nothing collective has started yet
    cudaMemcpy(d_h, &mine, 64, cudaMemcpyHostToDevice);
    bool coll = N.AllGather(d_h, d_h + 64, 64, kNcclUint8, sh->comm, 0) == 0 && cudaStreamSynchronize(0) == cudaSuccess;
    if (coll) cudaMemcpy(h.data(), d_h + 64, (size_t)64 * W, cudaMemcpyDeviceToHost);
    std::vector<unsigned long long*> peers((size_t)W, nullptr);
    if (coll && ok) {
        for (int r = 0; r < W && ok; ++r) {
            if (r == sh->rank) { peers[(size_t)r] = sh->d_mailbox; continue; }
            cudaIpcMemHandle_t hr;
            memcpy(&hr, h.data() + (size_t)64 * r, 64);
            void* q = nullptr;
            if (cudaIpcOpenMemHandle(&q, hr, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); ok = 0; break; }
            sh->opened.push_back(q);
            peers[(size_t)r] = static_cast<unsigned long long*>(q);
        }
    }
    // the vote: one int per rank
    int* d_v = reinterpret_cast<int*>(d_h + 64 + 64 * (size_t)W);
    const int myvote = (coll && ok) ? 1 : 0;
    cudaMemcpy(d_v + W, &myvote, 4, cudaMemcpyHostToDevice);
    bool all = coll && N.AllGather(d_v + W, d_v, 1, kNcclInt32, sh->comm, 0) == 0 && cudaStreamSynchronize(0) == cudaSuccess;
    if (all) {
        cudaMemcpy(votes.data(), d_v, 4 * (size_t)W, cudaMemcpyDeviceToHost);
        for (int r = 0; r < W; ++r) all = all && votes[(size_t)r] == 1;
    }
    if (all) all = cudaMemcpy(sh->d_peers, peers.data(), sizeof(void*) * W, cudaMemcpyHostToDevice) == cudaSuccess;
    cudaFree(d_h);
    sh->p2p = all;
    sh->mcap = cap;
}

extern "C" int cvxgpu_shard_create(cvxgpu_shard** out, const void* id128, void* nccl_comm, int rank, int world)
{
    if (!out) return FAIL(CVXGPU_ERR_ARG, "cvxgpu_shard_create: null pointer");
    *out = nullptr;
    if (world < 1 || rank < 0 || rank >= world) return FAIL(CVXGPU_ERR_ARG, "cvxgpu_shard_create: rank %d of %d", rank, world);
    if (world > 1 && !id128 && !nccl_comm) return FAIL(CVXGPU_ERR_ARG, "cvxgpu_shard_create: %d ranks need a NCCL unique id or a communicator", world);
    cvxgpu_shard* sh = new cvxgpu_shard();
    sh->rank = rank; sh->world = world;
    int rc = cvxgpu_workspace_create(&sh->ws);
    if (rc) { cvxgpu_shard_destroy(sh); return rc; }
    if (world > 1) {
        Nccl& N = nccl();
        if (!N.ok) { cvxgpu_shard_destroy(sh); return FAIL(CVXGPU_ERR_CUDA, "%s", N.why); }
        if (nccl_comm) {
            sh->comm = reinterpret_cast<ncclComm_t>(nccl_comm);
        } else {
            ncclUniqueId id;
            memcpy(&id, id128, 128);
            const int nrc = N.CommInitRank(&sh->comm, world, id, rank);
            if (nrc) { sh->comm = nullptr; cvxgpu_shard_destroy(sh); return nccl_fail("ncclCommInitRank", nrc); }
            sh->own_comm = true;
        }
    }
    cudaError_t e;
    if ((e = cudaMalloc(&sh->d_cnt, sizeof(int) * world)) != cudaSuccess || (e = cudaMalloc(&sh->d_sizes, sizeof(long) * world)) != cudaSuccess ||
        (e = cudaMalloc(&sh->d_nbytes, sizeof(long) * 4)) != cudaSuccess || (e = cudaMemset(sh->d_nbytes, 0, sizeof(long) * 4)) != cudaSuccess) {
        cvxgpu_shard_destroy(sh);
        return FAIL(CVXGPU_ERR_CUDA, "cvxgpu_shard_create: %s", cudaGetErrorString(e));
    }
    if (world > 1) shard_setup_mailboxes(sh);           // optional: on any failure the two scalars travel by ncclAllGather
    *out = sh;
    return CVXGPU_OK;
}

// scratch of the row-sum exchange for a volume of nz planes in block rows of bz (called outside stream capture the first time)
static int shard_prepare(cvxgpu_shard* sh, int nz, int bz)
{
    if (sh->g_nz == nz && sh->g_bz == bz) return CVXGPU_OK;
    std::vector<int> cnt((size_t)sh->world);
    int seg = 1;
    for (int r = 0; r < sh->world; ++r) {
        int nb = 0;
        cvxgpu_shard_range(nz, bz, r, sh->world, nullptr, nullptr, nullptr, &nb);
        cnt[(size_t)r] = nb;
        if (nb > seg) seg = nb;
    }
    if (seg > sh->seg) {
        if (sh->d_rows_send) cudaFree(sh->d_rows_send);
        if (sh->d_rows_recv) cudaFree(sh->d_rows_recv);
        sh->d_rows_send = sh->d_rows_recv = nullptr; sh->seg = 0;
        CU(cudaMalloc(&sh->d_rows_send, sizeof(double) * (size_t)seg));
        CU(cudaMalloc(&sh->d_rows_recv, sizeof(double) * (size_t)seg * sh->world));
        CU(cudaMemset(sh->d_rows_send, 0, sizeof(double) * (size_t)seg));
        sh->seg = seg;
    }
    CU(cudaMemcpy(sh->d_cnt, cnt.data(), sizeof(int) * sh->world, cudaMemcpyHostToDevice));
    CU(ws_reserve_rows(sh->ws, seg));
    sh->g_nz = nz; sh->g_bz = bz;
    return CVXGPU_OK;
}

extern "C" int cvxgpu_compress_sharded(cvxgpu_shard* sh, const float* d_slab, int nx, int ny, int nz, int bx, int by, int bz,
                                       float scale, int use_local_rms, void* d_header, unsigned long long* d_offs,
                                       float* d_blkmulfac, void* d_payload, long payload_capacity, long* d_bases,
                                       int* d_status, void* cuda_stream)
{
    if (!sh) return FAIL(CVXGPU_ERR_ARG, "cvxgpu_compress_sharded: null shard");
    if (nx <= 0 || ny <= 0 || nz <= 0 || !valid_block_size(bx, by, bz)) return FAIL(CVXGPU_ERR_ARG, "cvxgpu_compress_sharded: bad dimensions / block size");
    const int local = use_local_rms ? 1 : 0;
    int z0, nzs, b0, nb;
    cvxgpu_shard_range(nz, bz, sh->rank, sh->world, &z0, &nzs, &b0, &nb);
    if (!d_header || !d_bases || !d_status || (nzs > 0 && (!d_slab || !d_offs || !d_payload || (local && !d_blkmulfac))))
        return FAIL(CVXGPU_ERR_ARG, "cvxgpu_compress_sharded: null pointer");
    int rc = shard_prepare(sh, nz, bz);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)cuda_stream;
    cvxgpu_workspace* ws = sh->ws;
    Nccl& N = nccl();
    const Shape gs = make_shape(nx, ny, nz, bx, by, bz);
    const long ncells_total = (long)nx * ny * nz;
    // (1) global RMS: this rank's block-row sums -> all ranks' -> the same fixed-order total everywhere -> header, glob_mulfac
    const bool mailbox = sh->world > 1 && sh->p2p && sh->seg <= sh->mcap;      // else the two scalars travel by ncclAllGather
    const double* d_rows = nullptr;
    int nseg = 0;
    if (!local) {
        if (nzs > 0) CU(launch_sumsq_rows(ws, d_slab, (long)nx * ny, bz, nzs, sh->world > 1 ? sh->d_rows_send : ws->d_rowsum, st));
        d_rows = ws->d_rowsum;
        if (mailbox) {
            CU(launch_shard_push(sh->d_rows_send, nb, 0, sh->d_peers, sh->world, sh->mcap, sh->rank, sh->d_epoch, st));
        } else if (sh->world > 1) {
            const int nrc = N.AllGather(sh->d_rows_send, sh->d_rows_recv, (size_t)sh->seg, kNcclFloat64, sh->comm, st);
            if (nrc) return nccl_fail("ncclAllGather(row sums)", nrc);
            d_rows = sh->d_rows_recv; nseg = sh->world;
        }
    }
    if (mailbox && !local) {
        CU(launch_rms_header_mailbox(sh->d_mailbox, sh->world, sh->mcap, sh->d_epoch, sh->d_cnt, ws->d_sumsq, d_header, gs, scale, local,
                                     ncells_total, ws->d_mulfac, d_status, st));
    } else {
        CU(launch_rms_header(d_rows, nseg, sh->seg, sh->d_cnt, nb, ws->d_sumsq, d_header, gs, scale, local, ncells_total, ws->d_mulfac, st));
    }
    // (2) this rank's blocks -> its table slice (local offsets) and payload; launch_encode clears *d_status and the byte count
    if (nzs > 0) {
        const Shape ss = make_shape(nx, ny, nzs, bx, by, bz);
        CU(launch_encode(ws, d_slab, ss, scale, local, ws->d_mulfac, d_offs, d_blkmulfac, d_payload, payload_capacity,
                         sh->d_nbytes, d_status, cvx_internal_exact_mode(), st));
    } else {
        CU(cudaMemsetAsync(d_status, 0, sizeof(int), st));
        CU(cudaMemsetAsync(sh->d_nbytes, 0, sizeof(long), st));
    }
    // (3) payload sizes -> exclusive prefix -> global offsets
    const long nblk_slab = (long)nb * gs.nbx * gs.nby;
    if (mailbox) {
        CU(launch_shard_push(sh->d_nbytes, 1, 1, sh->d_peers, sh->world, sh->mcap, sh->rank, sh->d_epoch + 1, st));
        CU(launch_shard_fixup_mailbox(sh->d_mailbox, sh->world, sh->mcap, sh->d_epoch + 1, sh->rank, d_bases, d_offs, nblk_slab, d_status, st));
    } else if (sh->world > 1) {
        const int nrc = N.AllGather(sh->d_nbytes, sh->d_sizes, 1, kNcclInt64, sh->comm, st);
        if (nrc) return nccl_fail("ncclAllGather(payload sizes)", nrc);
        CU(launch_shard_fixup(sh->d_sizes, sh->world, sh->rank, d_bases, d_offs, nblk_slab, st));
    } else {
        CU(launch_shard_fixup(sh->d_nbytes, 1, 0, d_bases, d_offs, 0, st));
    }
    return CVXGPU_OK;
}

extern "C" int cvxgpu_decompress_sharded(cvxgpu_shard* sh, float* d_slab, int nx, int ny, int nz, int bx, int by, int bz,
                                         int use_local_rms, const void* d_header, const unsigned long long* d_offs,
                                         const float* d_blkmulfac, const void* d_payload, const long* d_bases, void* cuda_stream)
{
    if (!sh) return FAIL(CVXGPU_ERR_ARG, "cvxgpu_decompress_sharded: null shard");
    if (nx <= 0 || ny <= 0 || nz <= 0 || !valid_block_size(bx, by, bz)) return FAIL(CVXGPU_ERR_ARG, "cvxgpu_decompress_sharded: bad dimensions / block size");
    const int local = use_local_rms ? 1 : 0;
    int z0, nzs, b0, nb;
    cvxgpu_shard_range(nz, bz, sh->rank, sh->world, &z0, &nzs, &b0, &nb);
    if (nzs <= 0) return CVXGPU_OK;
    if (!d_slab || !d_header || !d_offs || !d_payload || !d_bases || (local && !d_blkmulfac))
        return FAIL(CVXGPU_ERR_ARG, "cvxgpu_decompress_sharded: null pointer");
    if (reinterpret_cast<uintptr_t>(d_payload) & 15) return FAIL(CVXGPU_ERR_ARG, "cvxgpu_decompress_sharded: the payload shard must be 16-byte aligned");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    CU(launch_shard_extent(d_bases, sh->rank, sh->d_nbytes + 1, st));
    const Shape ss = make_shape(nx, ny, nzs, bx, by, bz);
    const float* d_gm = reinterpret_cast<const float*>(static_cast<const unsigned char*>(d_header) + 24);
    CU(launch_decode(sh->ws, d_slab, ss, 1.0f, local, d_offs, d_blkmulfac, d_payload, 0, cvx_internal_exact_mode(), st, d_gm,
                     sh->d_nbytes + 1, sh->d_nbytes + 2));
    return CVXGPU_OK;
}

// every rank holds the FULL stream (e.g. one written by the reference with all host cores, arbitrary block order,
// CvxCompress.cpp:365-390) and decodes its own iBlk range into its z-slab.  No collective, no host synchronisation.
extern "C" int cvxgpu_decompress_slab_device(cvxgpu_workspace* wsin, float* d_slab, int nx, int ny, int nz, int bx, int by, int bz,
                                             int use_local_rms, int rank, int world, const void* d_stream, long length,
                                             void* cuda_stream)
{
    if (nx <= 0 || ny <= 0 || nz <= 0 || !valid_block_size(bx, by, bz)) return FAIL(CVXGPU_ERR_ARG, "cvxgpu_decompress_slab_device: bad dimensions / block size");
    int z0, nzs, b0, nb;
    int rc = cvxgpu_shard_range(nz, bz, rank, world, &z0, &nzs, &b0, &nb);
    if (rc) return rc;
    if (nzs <= 0) return CVXGPU_OK;
    if (!d_slab || !d_stream) return FAIL(CVXGPU_ERR_ARG, "cvxgpu_decompress_slab_device: null pointer");
    if (reinterpret_cast<uintptr_t>(d_stream) & 7) return FAIL(CVXGPU_ERR_ARG, "compressed stream must be 8-byte aligned");
    const int local = use_local_rms ? 1 : 0;
    const Shape gs = make_shape(nx, ny, nz, bx, by, bz);
    const long tables = 8 * gs.nblk + (local ? 4 * gs.nblk : 0);
    const long payload_bytes = length > 0 ? length - 32 - tables : (1L << 62);
    if (payload_bytes < 0) return FAIL(CVXGPU_ERR_STREAM, "stream length %ld is shorter than header + tables", length);
    cvxgpu_workspace* ws = wsin;
    if (!ws) { rc = cvx_internal_default_ws(&ws, cuda_stream); if (rc) return rc; }
    const unsigned char* base = static_cast<const unsigned char*>(d_stream);
    const long blk_first = (long)b0 * gs.nbx * gs.nby;
    const unsigned long long* d_offs = reinterpret_cast<const unsigned long long*>(base + 32) + blk_first;
    const float* d_bmf = local ? reinterpret_cast<const float*>(base + 32 + 8 * gs.nblk) + blk_first : nullptr;
    const Shape ss = make_shape(nx, ny, nzs, bx, by, bz);
    CU(launch_decode(ws, d_slab, ss, 1.0f, local, d_offs, d_bmf, base + 32 + tables, payload_bytes, cvx_internal_exact_mode(),
                     (cudaStream_t)cuda_stream, reinterpret_cast<const float*>(base + 24)));
    return CVXGPU_OK;
}

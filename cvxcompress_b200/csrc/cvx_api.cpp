// cvx_api.cpp -- the C ABI of include/cvx_gpu.h: argument checks, workspace, stream layout
// (header / offset table / mulfac table / payload, CvxCompress.cpp:284-315 and :473-517) and the
// host-pointer staging used by the CvxCompress class.  All arithmetic is in cvx_kernels.cu;
// there is no CPU implementation of any stage here.
#include "../../include/cvx_gpu.h"
#include "cvx_internal.h"

#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <mutex>
#include <thread>
#include <vector>

using namespace cvx;

// ---- errors -------------------------------------------------------------------------------
static thread_local char g_err[512] = "";

static int fail(int code, const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}
#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e__ = (call);                                                                  \
        if (e__ != cudaSuccess) return fail(CVXGPU_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e__)); \
    } while (0)

extern "C" const char* cvxgpu_last_error(void) { return g_err; }

extern "C" int cvxgpu_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

// ---- transform mode -------------------------------------------------------------------------
static std::atomic<int> g_transform{-1};

extern "C" int cvxgpu_get_transform(void)
{
    int m = g_transform.load();
    if (m < 0) {
        const char* e = getenv("CVX_GPU_TRANSFORM");
        m = (e && (!strcmp(e, "lifting") || !strcmp(e, "lift") || !strcmp(e, "1"))) ? CVXGPU_TRANSFORM_LIFTING
                                                                                  : CVXGPU_TRANSFORM_EXACT;
        g_transform.store(m);
    }
    return m;
}
extern "C" int cvxgpu_set_transform(int mode)
{
    if (mode != CVXGPU_TRANSFORM_EXACT && mode != CVXGPU_TRANSFORM_LIFTING) return fail(CVXGPU_ERR_ARG, "unknown transform mode %d", mode);
    g_transform.store(mode);
    return CVXGPU_OK;
}
// ---- safe decode: -1 off (reference behaviour: the stream is trusted), 0 / 1 = cvxgpu_validate_stream level run by
// cvxgpu_decompress on HOST streams before they are staged.  Default from CVX_GPU_SAFE_DECODE ("0", "1"; unset = off).
static std::atomic<int> g_safe_decode{-2};
extern "C" int cvxgpu_get_safe_decode(void)
{
    int m = g_safe_decode.load();
    if (m < -1) {
        const char* e = getenv("CVX_GPU_SAFE_DECODE");
        m = !e || !*e ? -1 : (!strcmp(e, "0") ? 0 : (!strcmp(e, "off") ? -1 : 1));
        g_safe_decode.store(m);
    }
    return m;
}
extern "C" int cvxgpu_set_safe_decode(int level)
{
    if (level < -1 || level > 1) return fail(CVXGPU_ERR_ARG, "safe-decode level %d: use -1 (off), 0 or 1", level);
    g_safe_decode.store(level);
    return CVXGPU_OK;
}
static int exact_mode() { return cvxgpu_get_transform() == CVXGPU_TRANSFORM_EXACT ? 1 : 0; }

// ---- workspace ------------------------------------------------------------------------------
extern "C" int cvxgpu_workspace_create(cvxgpu_workspace** out)
{
    if (!out) return fail(CVXGPU_ERR_ARG, "null workspace pointer");
    *out = nullptr;
    int dev = 0;
    CU(cudaGetDevice(&dev));
    CvxRelaxCapture relax;               // a default workspace may be born under a stream capture (first call on the capture stream)
    cvxgpu_workspace* ws = new cvxgpu_workspace();
    ws->device = dev;
    cudaDeviceProp prop;
    cudaError_t e = cudaGetDeviceProperties(&prop, dev);
    if (e != cudaSuccess) { delete ws; return fail(CVXGPU_ERR_CUDA, "cudaGetDeviceProperties: %s", cudaGetErrorString(e)); }
    ws->sm_count = prop.multiProcessorCount;
    ws->n_partial = ws->sm_count * 8;
    if ((e = cudaMalloc(&ws->d_scalars, 256)) != cudaSuccess ||
        (e = cudaMalloc(&ws->d_partial, sizeof(double) * ws->n_partial)) != cudaSuccess ||
        (e = cudaMallocHost(&ws->h_pinned, 256)) != cudaSuccess) {
        cvxgpu_workspace_destroy(ws);
        return fail(CVXGPU_ERR_CUDA, "workspace allocation: %s", cudaGetErrorString(e));
    }
    ws->d_sumsq = reinterpret_cast<double*>(ws->d_scalars);
    ws->d_mulfac = reinterpret_cast<float*>(ws->d_scalars + 8);
    ws->d_ticket = reinterpret_cast<unsigned int*>(ws->d_scalars + 16);
    ws->d_payload_bytes = reinterpret_cast<long*>(ws->d_scalars + 24);
    ws->d_length = reinterpret_cast<long*>(ws->d_scalars + 32);
    ws->d_err = reinterpret_cast<int*>(ws->d_scalars + 40);
    *out = ws;
    return CVXGPU_OK;
}

extern "C" int cvxgpu_workspace_destroy(cvxgpu_workspace* ws)
{
    if (!ws) return CVXGPU_OK;
    if (ws->d_scalars) cudaFree(ws->d_scalars);
    if (ws->d_partial) cudaFree(ws->d_partial);
    if (ws->d_status) cudaFree(ws->d_status);
    if (ws->d_scratch) cudaFree(ws->d_scratch);
    if (ws->h_pinned) cudaFreeHost(ws->h_pinned);
    if (ws->d_stage_vol) cudaFree(ws->d_stage_vol);
    if (ws->d_stage_stream) cudaFree(ws->d_stage_stream);
    if (ws->h_stage) cudaFreeHost(ws->h_stage);
    if (ws->priv_stream) cudaStreamDestroy(ws->priv_stream);
    if (ws->d_rowpartial) cudaFree(ws->d_rowpartial);
    if (ws->d_rowsum) cudaFree(ws->d_rowsum);
    delete ws;
    return CVXGPU_OK;
}

// Per-host-thread default workspaces, one per (device, CUDA stream): the look-back status words, the ticket and the
// scalar slab of a workspace belong to ONE call in flight, and two calls of one host thread on different streams may
// overlap on the GPU (double-buffered checkpointing, a captured graph replayed next to eager work), so they must not
// share them.  Calls on the same stream are ordered by the stream.  cvxgpu_workspace_release_thread_default() frees them.
struct TlsWorkspaces {
    struct Ent { int dev; cudaStream_t st; cvxgpu_workspace* ws; };
    std::vector<Ent> v;
    ~TlsWorkspaces()
    {
        // the CUDA context may already be gone at thread/process exit: leak device memory, free nothing
    }
};
static thread_local TlsWorkspaces g_tls;

static int get_ws(cvxgpu_workspace* in, cvxgpu_workspace** out, void* cuda_stream)
{
    if (in) { *out = in; return CVXGPU_OK; }
    int dev = 0;
    CU(cudaGetDevice(&dev));
    const cudaStream_t st = (cudaStream_t)cuda_stream;
    for (auto& e : g_tls.v)
        if (e.dev == dev && e.st == st) { *out = e.ws; return CVXGPU_OK; }
    cvxgpu_workspace* ws = nullptr;
    int rc = cvxgpu_workspace_create(&ws);
    if (rc != CVXGPU_OK) return rc;
    g_tls.v.push_back({dev, st, ws});
    *out = ws;
    return CVXGPU_OK;
}

extern "C" int cvxgpu_workspace_release_thread_default(void)
{
    int rc = CVXGPU_OK;
    for (auto& e : g_tls.v) {
        int dev = 0;
        cudaGetDevice(&dev);
        if (dev != e.dev) cudaSetDevice(e.dev);
        if (cudaDeviceSynchronize() != cudaSuccess) rc = CVXGPU_ERR_CUDA;
        cvxgpu_workspace_destroy(e.ws);
        if (dev != e.dev) cudaSetDevice(dev);
    }
    g_tls.v.clear();
    return rc;
}

// for the other translation units of the library (cvx_shard.cpp)
int cvx_internal_fail(int code, const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}
int cvx_internal_exact_mode() { return exact_mode(); }
int cvx_internal_default_ws(cvxgpu_workspace** out, void* cuda_stream) { return get_ws(nullptr, out, cuda_stream); }

// ---- shape helpers --------------------------------------------------------------------------
extern "C" int cvxgpu_is_valid_block_size(int bx, int by, int bz) { return valid_block_size(bx, by, bz) ? 1 : 0; }

extern "C" long cvxgpu_num_blocks(int nx, int ny, int nz, int bx, int by, int bz)
{
    if (nx <= 0 || ny <= 0 || nz <= 0 || !valid_block_size(bx, by, bz)) return -1;
    return make_shape(nx, ny, nz, bx, by, bz).nblk;
}

extern "C" long cvxgpu_max_compressed_bytes(int nx, int ny, int nz, int bx, int by, int bz, int use_local_rms)
{
    long nnn = cvxgpu_num_blocks(nx, ny, nz, bx, by, bz);
    if (nnn < 0) return -1;
    return 32 + 8 * nnn + (use_local_rms ? 4 * nnn : 0) + 4L * bx * by * bz * nnn + 7;
}

static int check_dims(int nx, int ny, int nz, int bx, int by, int bz)
{
    if (nx <= 0 || ny <= 0 || nz <= 0) return fail(CVXGPU_ERR_ARG, "volume dimensions must be positive (got %d x %d x %d)", nx, ny, nz);
    if (!valid_block_size(bx, by, bz))
        return fail(CVXGPU_ERR_ARG, "invalid block size %d x %d x %d (powers of two in 8..256; bz may be 1)", bx, by, bz);
    return CVXGPU_OK;
}

// ---- pieces ---------------------------------------------------------------------------------
extern "C" int cvxgpu_sumsq_device(cvxgpu_workspace* wsin, const float* d_vol, long ncells, double* d_sumsq, void* cuda_stream)
{
    cvxgpu_workspace* ws;
    int rc = get_ws(wsin, &ws, cuda_stream);
    if (rc) return rc;
    if (!d_vol || !d_sumsq || ncells <= 0) return fail(CVXGPU_ERR_ARG, "cvxgpu_sumsq_device: bad argument");
    CU(launch_sumsq(ws, d_vol, ncells, d_sumsq, (cudaStream_t)cuda_stream));
    return CVXGPU_OK;
}

extern "C" int cvxgpu_sumsq_rows_device(cvxgpu_workspace* wsin, const float* d_vol, long plane_cells, int bz, int nz,
                                        double* d_rowsums, void* cuda_stream)
{
    cvxgpu_workspace* ws;
    int rc = get_ws(wsin, &ws, cuda_stream);
    if (rc) return rc;
    if (!d_vol || !d_rowsums || plane_cells <= 0 || bz <= 0 || nz <= 0) return fail(CVXGPU_ERR_ARG, "cvxgpu_sumsq_rows_device: bad argument");
    CU(launch_sumsq_rows(ws, d_vol, plane_cells, bz, nz, d_rowsums, (cudaStream_t)cuda_stream));
    return CVXGPU_OK;
}

extern "C" int cvxgpu_sum_rows_device(const double* d_rowsums, int nrows, double* d_total, void* cuda_stream)
{
    if (!d_rowsums || !d_total || nrows < 0) return fail(CVXGPU_ERR_ARG, "cvxgpu_sum_rows_device: bad argument");
    CU(launch_sum_rows(d_rowsums, 0, 0, nullptr, nrows, d_total, (cudaStream_t)cuda_stream));
    return CVXGPU_OK;
}

extern "C" int cvxgpu_write_header_device(void* d_stream, int nx, int ny, int nz, int bx, int by, int bz, float scale,
                                          int use_local_rms, const double* d_sumsq_total, long ncells_total, void* cuda_stream)
{
    int rc = check_dims(nx, ny, nz, bx, by, bz);
    if (rc) return rc;
    if (!d_stream || (!use_local_rms && !d_sumsq_total)) return fail(CVXGPU_ERR_ARG, "cvxgpu_write_header_device: null pointer");
    Shape s = make_shape(nx, ny, nz, bx, by, bz);
    CU(launch_write_header(d_stream, s, scale, use_local_rms ? 1 : 0, d_sumsq_total, ncells_total, nullptr, (cudaStream_t)cuda_stream));
    return CVXGPU_OK;
}

extern "C" int cvxgpu_offsets_add_device(unsigned long long* d_offs, long n, long base, void* cuda_stream)
{
    if (!d_offs || n < 0 || base < 0) return fail(CVXGPU_ERR_ARG, "cvxgpu_offsets_add_device: bad argument");
    CU(launch_offsets_add(d_offs, n, base, (cudaStream_t)cuda_stream));
    return CVXGPU_OK;
}

extern "C" int cvxgpu_encode_blocks_device(cvxgpu_workspace* wsin, const float* d_vol, int nx, int ny, int nz_slab,
                                           int bx, int by, int bz, float scale, int use_local_rms,
                                           const double* d_sumsq_total, long ncells_total,
                                           unsigned long long* d_offs, float* d_blkmulfac,
                                           void* d_payload, long payload_capacity,
                                           long* d_payload_bytes, int* d_status, void* cuda_stream)
{
    cvxgpu_workspace* ws;
    int rc = get_ws(wsin, &ws, cuda_stream);
    if (rc) return rc;
    rc = check_dims(nx, ny, nz_slab, bx, by, bz);
    if (rc) return rc;
    const int local = use_local_rms ? 1 : 0;
    if (!d_vol || !d_offs || !d_payload || !d_payload_bytes || !d_status || (local && !d_blkmulfac) || (!local && !d_sumsq_total))
        return fail(CVXGPU_ERR_ARG, "cvxgpu_encode_blocks_device: null pointer");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    Shape s = make_shape(nx, ny, nz_slab, bx, by, bz);
    CU(launch_write_header(nullptr, s, scale, local, d_sumsq_total, ncells_total, ws->d_mulfac, st));      // launch_encode clears *d_status
    CU(launch_encode(ws, d_vol, s, scale, local, ws->d_mulfac, d_offs, d_blkmulfac, d_payload, payload_capacity,
                     d_payload_bytes, d_status, exact_mode(), st));
    return CVXGPU_OK;
}

extern "C" int cvxgpu_decode_blocks_device(cvxgpu_workspace* wsin, float* d_vol_slab, int nx, int ny, int nz_slab,
                                           int bx, int by, int bz, float glob_mulfac, int use_local_rms,
                                           const unsigned long long* d_offs, const float* d_blkmulfac,
                                           const void* d_payload, long payload_bytes, void* cuda_stream)
{
    cvxgpu_workspace* ws;
    int rc = get_ws(wsin, &ws, cuda_stream);
    if (rc) return rc;
    rc = check_dims(nx, ny, nz_slab, bx, by, bz);
    if (rc) return rc;
    const int local = use_local_rms ? 1 : 0;
    if (!d_vol_slab || !d_offs || !d_payload || (local && !d_blkmulfac) || payload_bytes < 0)
        return fail(CVXGPU_ERR_ARG, "cvxgpu_decode_blocks_device: bad argument");
    Shape s = make_shape(nx, ny, nz_slab, bx, by, bz);
    CU(launch_decode(ws, d_vol_slab, s, glob_mulfac, local, d_offs, d_blkmulfac, d_payload, payload_bytes, exact_mode(),
                     (cudaStream_t)cuda_stream));
    return CVXGPU_OK;
}

extern "C" int cvxgpu_wavelet_blocks_device(cvxgpu_workspace* wsin, float* d_blocks, long nblocks, int bx, int by, int bz,
                                            int inverse, void* cuda_stream)
{
    cvxgpu_workspace* ws;
    int rc = get_ws(wsin, &ws, cuda_stream);
    if (rc) return rc;
    if (!valid_block_size(bx, by, bz)) return fail(CVXGPU_ERR_ARG, "invalid block size %d x %d x %d", bx, by, bz);
    if (!d_blocks || nblocks < 0) return fail(CVXGPU_ERR_ARG, "cvxgpu_wavelet_blocks_device: bad argument");
    CU(launch_wavelet(ws, d_blocks, nblocks, bx, by, bz, inverse ? 1 : 0, exact_mode(), (cudaStream_t)cuda_stream));
    return CVXGPU_OK;
}

// ---- whole-volume calls ---------------------------------------------------------------------
static int compress_enqueue(cvxgpu_workspace* ws, const float* d_vol, const Shape& s, float scale, int local,
                            void* d_stream, long capacity, long* d_length, int* d_status, cudaStream_t st)
{
    const long tables = 8 * s.nblk + (local ? 4 * s.nblk : 0);
    if (capacity < 32 + tables + 7) return fail(CVXGPU_ERR_CAPACITY, "capacity %ld is smaller than header + tables (%ld bytes)", capacity, 32 + tables + 7);
    if (reinterpret_cast<uintptr_t>(d_stream) & 7) return fail(CVXGPU_ERR_ARG, "compressed stream must be 8-byte aligned");
    unsigned char* base = reinterpret_cast<unsigned char*>(d_stream);
    unsigned long long* d_offs = reinterpret_cast<unsigned long long*>(base + 32);          // CvxCompress.cpp:302
    float* d_bmf = local ? reinterpret_cast<float*>(base + 32 + 8 * s.nblk) : nullptr;     // :308
    unsigned char* d_payload = base + 32 + tables;                                          // :309 / :314
    const long ncells = (long)s.nx * s.ny * s.nz;
    // per-block-row sums added in row order: the same arithmetic the sharded path performs for any number of GPUs
    if (!local) CU(launch_sumsq_rows(ws, d_vol, (long)s.nx * s.ny, s.bz, s.nz, nullptr, st));
    CU(launch_rms_header(local ? nullptr : ws->d_rowsum, 0, 0, nullptr, s.nbz, ws->d_sumsq, d_stream, s, scale, local, ncells, ws->d_mulfac, st));
    CU(launch_encode(ws, d_vol, s, scale, local, ws->d_mulfac, d_offs, d_bmf, d_payload, capacity - 32 - tables - 7,
                     ws->d_payload_bytes, d_status, exact_mode(), st));
    CU(launch_finish_length(ws->d_payload_bytes, tables, d_length, st));
    return CVXGPU_OK;
}

extern "C" int cvxgpu_compress_device_async(cvxgpu_workspace* wsin, const float* d_vol, int nx, int ny, int nz,
                                            int bx, int by, int bz, float scale, int use_local_rms,
                                            void* d_stream, long capacity, long* d_length, int* d_status, void* cuda_stream)
{
    cvxgpu_workspace* ws;
    int rc = get_ws(wsin, &ws, cuda_stream);
    if (rc) return rc;
    rc = check_dims(nx, ny, nz, bx, by, bz);
    if (rc) return rc;
    if (!d_vol || !d_stream || !d_length || !d_status) return fail(CVXGPU_ERR_ARG, "cvxgpu_compress_device_async: null pointer");
    Shape s = make_shape(nx, ny, nz, bx, by, bz);
    return compress_enqueue(ws, d_vol, s, scale, use_local_rms ? 1 : 0, d_stream, capacity, d_length, d_status, (cudaStream_t)cuda_stream);
}

extern "C" int cvxgpu_compress_device(cvxgpu_workspace* wsin, const float* d_vol, int nx, int ny, int nz,
                                      int bx, int by, int bz, float scale, int use_local_rms,
                                      void* d_stream, long capacity, long* length, float* ratio, void* cuda_stream)
{
    cvxgpu_workspace* ws;
    int rc = get_ws(wsin, &ws, cuda_stream);
    if (rc) return rc;
    rc = check_dims(nx, ny, nz, bx, by, bz);
    if (rc) return rc;
    if (!d_vol || !d_stream || !length) return fail(CVXGPU_ERR_ARG, "cvxgpu_compress_device: null pointer");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    Shape s = make_shape(nx, ny, nz, bx, by, bz);
    rc = compress_enqueue(ws, d_vol, s, scale, use_local_rms ? 1 : 0, d_stream, capacity, ws->d_length, ws->d_err, st);
    if (rc) return rc;
    CU(cudaMemcpyAsync(ws->h_pinned, ws->d_length, 16, cudaMemcpyDeviceToHost, st));       // length (8) + err (4 of the next 8)
    CU(cudaStreamSynchronize(st));
    long len;
    int err;
    memcpy(&len, ws->h_pinned, 8);
    memcpy(&err, ws->h_pinned + 8, 4);
    *length = len;
    if (ratio) *ratio = (float)(((double)nx * (double)ny * (double)nz * 4.0) / (double)len);      // CvxCompress.cpp:425
    if (err != 0) return fail(CVXGPU_ERR_CAPACITY, "compressed stream needs %ld bytes but capacity is %ld", len, capacity);
    return CVXGPU_OK;
}

extern "C" int cvxgpu_parse_header(const void* h_stream, long length, int* nx, int* ny, int* nz,
                                   int* bx, int* by, int* bz, float* glob_mulfac, int* flags)
{
    if (!h_stream || length < 32) return fail(CVXGPU_ERR_STREAM, "stream shorter than its 32-byte header");
    int32_t h[8];
    memcpy(h, h_stream, 32);
    if (h[0] <= 0 || h[1] <= 0 || h[2] <= 0 || !valid_block_size(h[3], h[4], h[5]))
        return fail(CVXGPU_ERR_STREAM, "bad header: volume %d x %d x %d, block %d x %d x %d", h[0], h[1], h[2], h[3], h[4], h[5]);
    if (nx) *nx = h[0];
    if (ny) *ny = h[1];
    if (nz) *nz = h[2];
    if (bx) *bx = h[3];
    if (by) *by = h[4];
    if (bz) *bz = h[5];
    if (glob_mulfac) memcpy(glob_mulfac, &h[6], 4);
    if (flags) *flags = h[7];
    return CVXGPU_OK;
}

// ---- stream validator (host memory, no GPU): SURVEY.md 8(f) row 3 -----------------------------
// Token sizes and value counts of the reference's code table (Run_Length_Escape_Codes.hxx:8-14; decoder
// Run_Length_Encode_Slow.cpp:416-512).  Returns the token's byte count and sets *nvals; needs p[0..3] readable
// only as far as `avail` says.
static inline int token_extent(const unsigned char* p, long avail, long* nvals)
{
    const int c = (signed char)p[0];
    switch (c) {
    case 127:  if (avail < 2) return 0; *nvals = p[1]; return 2;                                             // RLESC1
    case 125:  if (avail < 4) return 0; *nvals = (long)p[1] | ((long)p[2] << 8) | ((long)p[3] << 16); return 4;  // RLESC3
    case -125: *nvals = 1; return 3;   // VLESC2
    case -127: *nvals = 1; return 4;   // VLESC3
    case -128: *nvals = 1; return 5;   // VLESC4
    case -126: *nvals = 8; return 17;  // VLESC2_8x
    case 126:  *nvals = 8; return 25;  // VLESC3_8x
    default:   *nvals = 1; return 1;   // literal -124..124
    }
}

static int validate_stream_body(const void* h_stream, long length, int level, cvxgpu_stream_info* info);

extern "C" int cvxgpu_validate_stream(const void* h_stream, long length, int level, cvxgpu_stream_info* info)
{
    // nothing may unwind through the C ABI: an allocation failure on a hostile header is a malformed stream
    try {
        return validate_stream_body(h_stream, length, level, info);
    } catch (const std::exception& e) {
        return fail(CVXGPU_ERR_STREAM, "stream validation aborted: %s", e.what());
    } catch (...) {
        return fail(CVXGPU_ERR_STREAM, "stream validation aborted");
    }
}

static int validate_stream_body(const void* h_stream, long length, int level, cvxgpu_stream_info* info)
{
    cvxgpu_stream_info local_info;
    cvxgpu_stream_info& I = info ? *info : local_info;
    memset(&I, 0, sizeof(I));
    I.used_payload_bytes = I.min_block_bytes = I.max_block_bytes = I.tokens = -1;
    int flags;
    int rc = cvxgpu_parse_header(h_stream, length, &I.nx, &I.ny, &I.nz, &I.bx, &I.by, &I.bz, &I.glob_mulfac, &flags);
    if (rc) return rc;
    I.use_local_rms = flags & 1;
    if (flags & ~1) return fail(CVXGPU_ERR_STREAM, "header flags 0x%x: only bit 0 (local RMS) is defined", flags);
    if (!std::isfinite(I.glob_mulfac) || I.glob_mulfac == 0.0f)
        return fail(CVXGPU_ERR_STREAM, "global multiplier %g is not a finite non-zero number", (double)I.glob_mulfac);
    // the header is untrusted: the block count is formed in 128 bits and must fit the stream before anything is sized by it
    {
        const long nbx = ((long)I.nx + I.bx - 1) / I.bx, nby = ((long)I.ny + I.by - 1) / I.by, nbz = ((long)I.nz + I.bz - 1) / I.bz;
        const __int128 nnn = (__int128)nbx * nby * nbz;
        const long room = length - 39;                          // header + 7 slack bytes
        if (room < 0 || nnn > (__int128)(room / (I.use_local_rms ? 12 : 8)))
            return fail(CVXGPU_ERR_STREAM, "stream of %ld bytes cannot hold the tables of the %ld x %ld x %ld blocks its header claims", length, nbx, nby, nbz);
    }
    const Shape s = make_shape(I.nx, I.ny, I.nz, I.bx, I.by, I.bz);
    const long B = (long)I.bx * I.by * I.bz;
    I.nblocks = s.nblk;
    I.table_bytes = 8 * s.nblk + (I.use_local_rms ? 4 * s.nblk : 0);
    I.payload_bytes = length - 32 - I.table_bytes - 7;      // CvxCompress.cpp:421-422: 7 bytes of slack end the stream
    if (I.payload_bytes < 0)
        return fail(CVXGPU_ERR_STREAM, "stream of %ld bytes cannot hold its header, %ld table bytes and the 7-byte slack",
                    length, I.table_bytes);
    const unsigned char* base = static_cast<const unsigned char*>(h_stream);
    const unsigned char* payload = base + 32 + I.table_bytes;
    const unsigned long long RAW = 1ULL << 63;
    std::vector<unsigned long long> starts((size_t)s.nblk);
    I.natural_order = 1;
    unsigned long long prev = 0;
    for (long b = 0; b < s.nblk; ++b) {
        unsigned long long e;
        memcpy(&e, base + 32 + 8 * b, 8);
        const unsigned long long off = e & ~RAW;
        if (off >= (unsigned long long)I.payload_bytes && !(off == 0 && I.payload_bytes == 0))
            return fail(CVXGPU_ERR_STREAM, "block %ld: offset %llu lies outside the %ld payload bytes", b, off, I.payload_bytes);
        if (e & RAW) {
            ++I.raw_blocks;
            if (off + 4ULL * B > (unsigned long long)I.payload_bytes)
                return fail(CVXGPU_ERR_STREAM, "raw block %ld: %ld bytes at offset %llu pass the payload end %ld", b, 4 * B, off, I.payload_bytes);
        }
        if (b && off <= prev) I.natural_order = 0;
        prev = off;
        starts[(size_t)b] = off;
        if (I.use_local_rms) {
            float m;
            memcpy(&m, base + 32 + 8 * s.nblk + 4 * b, 4);
            if (!std::isfinite(m) || m == 0.0f)
                return fail(CVXGPU_ERR_STREAM, "block %ld: multiplier %g is not a finite non-zero number", b, (double)m);
        }
    }
    if (level <= 0) return CVXGPU_OK;

    // every block must end at or before the next block's first byte (payload order is free: CvxCompress.cpp:382-384)
    std::vector<unsigned long long> sorted(starts);
    std::sort(sorted.begin(), sorted.end());
    for (size_t i = 1; i < sorted.size(); ++i)
        if (sorted[i] == sorted[i - 1])
            return fail(CVXGPU_ERR_STREAM, "two blocks share payload offset %llu", sorted[i]);
    // blocks are independent: walk contiguous iBlk ranges on the host's cores (a serial walk of BASELINE config 3's
    // 122 MB stream takes 0.2 s, longer than decoding it); the lowest offending block is the one reported
    struct Part { long used = 0, tokens = 0, minb = 1L << 62, maxb = 0, bad = -1; char msg[224]; };
    auto walk = [&](long b0, long b1, Part& P) {
        auto bad = [&](long b, const char* fmt, long a1, long a2, long a3, long a4) {
            P.bad = b;
            snprintf(P.msg, sizeof(P.msg), fmt, b, a1, a2, a3, a4);
        };
        for (long b = b0; b < b1; ++b) {
            unsigned long long e;
            memcpy(&e, base + 32 + 8 * b, 8);
            const long off = (long)(e & ~RAW);
            const auto nxt = std::upper_bound(sorted.begin(), sorted.end(), (unsigned long long)off);
            const long limit = nxt == sorted.end() ? I.payload_bytes : (long)*nxt;
            long bytes;
            if (e & RAW) {
                bytes = 4 * B;
            } else {
                long pos = off, vals = 0;
                while (vals < B) {
                    if (pos >= limit)
                        return bad(b, "block %ld: %ld of %ld values decoded when its payload [%ld, %ld) ends", vals, B, off, limit);
                    long nv = 0;
                    const int tb = token_extent(payload + pos, limit - pos, &nv);
                    if (tb == 0 || pos + tb > limit)
                        return bad(b, "block %ld: token 0x%02lx at payload byte %ld is cut off at %ld", (long)payload[pos], pos, limit, 0);
                    pos += tb;
                    vals += nv;
                    ++P.tokens;
                }
                if (vals != B) return bad(b, "block %ld: tokens produce %ld values, the block holds %ld", vals, B, 0, 0);
                bytes = pos - off;
            }
            if (off + bytes > limit)
                return bad(b, "block %ld: %ld bytes at offset %ld run into the next block at %ld", bytes, off, limit, 0);
            P.used += bytes;
            if (bytes < P.minb) P.minb = bytes;
            if (bytes > P.maxb) P.maxb = bytes;
        }
    };
    long nthreads = (long)std::thread::hardware_concurrency();
    if (const char* e = getenv("CVX_GPU_VALIDATE_THREADS")) nthreads = atol(e);
    nthreads = std::max(1L, std::min({nthreads, 64L, (s.nblk + 63) / 64}));
    std::vector<Part> parts((size_t)nthreads);
    std::vector<std::thread> pool;
    for (long t = 1; t < nthreads; ++t)
        pool.emplace_back(walk, s.nblk * t / nthreads, s.nblk * (t + 1) / nthreads, std::ref(parts[(size_t)t]));
    walk(0, s.nblk / nthreads, parts[0]);
    for (auto& th : pool) th.join();
    I.used_payload_bytes = I.tokens = I.max_block_bytes = 0;
    I.min_block_bytes = 1L << 62;
    for (const Part& P : parts) {
        I.used_payload_bytes += P.used;
        I.tokens += P.tokens;
        I.min_block_bytes = std::min(I.min_block_bytes, P.minb);
        I.max_block_bytes = std::max(I.max_block_bytes, P.maxb);
    }
    for (const Part& P : parts)             // parts are in iBlk order: the first bad part holds the lowest bad block
        if (P.bad >= 0) return fail(CVXGPU_ERR_STREAM, "%s", P.msg);
    return CVXGPU_OK;
}

// The reference never reads compressed_length (CvxCompress.cpp:463-571), so a caller may pass 0 or garbage <= 0.  The
// stream's extent follows from its own tables: the block with the largest payload offset ends the payload, a raw block
// being 4*B bytes and a coded one as long as its tokens say (walked here, one block).  Host memory, header trusted as
// the reference trusts it.
static long derive_stream_length(const void* h_stream)
{
    int nx, ny, nz, bx, by, bz, flags;
    if (cvxgpu_parse_header(h_stream, 32, &nx, &ny, &nz, &bx, &by, &bz, nullptr, &flags)) return -1;
    const Shape s = make_shape(nx, ny, nz, bx, by, bz);
    const long B = (long)bx * by * bz;
    const long tables = 8 * s.nblk + ((flags & 1) ? 4 * s.nblk : 0);
    const unsigned char* base = static_cast<const unsigned char*>(h_stream);
    const unsigned long long RAW = 1ULL << 63;
    unsigned long long best = 0;
    bool best_raw = false;
    for (long b = 0; b < s.nblk; ++b) {
        unsigned long long e;
        memcpy(&e, base + 32 + 8 * b, 8);
        if ((e & ~RAW) >= best) { best = e & ~RAW; best_raw = (e & RAW) != 0; }
    }
    long bytes = 4 * B;
    if (!best_raw) {
        const unsigned char* p = base + 32 + tables + best;
        long pos = 0, vals = 0;
        while (vals < B) {
            long nv = 0;
            const int tb = token_extent(p + pos, 1L << 40, &nv);
            pos += tb;
            vals += nv;
            if (pos > 5 * B + 64) return -1;
        }
        bytes = pos;
    }
    return 32 + tables + (long)best + bytes + 7;
}

static int decompress_with_header(cvxgpu_workspace* ws, float* d_vol, int nx, int ny, int nz, const int32_t* h,
                                  const void* d_stream, long length, cudaStream_t st)
{
    int hx, hy, hz, bx, by, bz, flags;
    float gm;
    int rc = cvxgpu_parse_header(h, 32, &hx, &hy, &hz, &bx, &by, &bz, &gm, &flags);
    if (rc) return rc;
    if (hx != nx || hy != ny || hz != nz)        // CvxCompress.cpp:473-488 prints and asserts
        return fail(CVXGPU_ERR_STREAM, "stream holds a %d x %d x %d volume, caller expects %d x %d x %d", hx, hy, hz, nx, ny, nz);
    const int local = flags & 1;
    Shape s = make_shape(nx, ny, nz, bx, by, bz);
    const long tables = 8 * s.nblk + (local ? 4 * s.nblk : 0);
    // `length` is informative in the reference (never read); here it bounds the reads when given
    long payload_bytes = length > 0 ? length - 32 - tables : (1L << 62);
    if (payload_bytes < 0) return fail(CVXGPU_ERR_STREAM, "stream length %ld is shorter than header + tables", length);
    const unsigned char* base = reinterpret_cast<const unsigned char*>(d_stream);
    const unsigned long long* d_offs = reinterpret_cast<const unsigned long long*>(base + 32);
    const float* d_bmf = local ? reinterpret_cast<const float*>(base + 32 + 8 * s.nblk) : nullptr;
    CU(launch_decode(ws, d_vol, s, gm, local, d_offs, d_bmf, base + 32 + tables, payload_bytes, exact_mode(), st));
    return CVXGPU_OK;
}

extern "C" int cvxgpu_decompress_device(cvxgpu_workspace* wsin, float* d_vol, int nx, int ny, int nz,
                                        const void* d_stream, long length, void* cuda_stream)
{
    cvxgpu_workspace* ws;
    int rc = get_ws(wsin, &ws, cuda_stream);
    if (rc) return rc;
    if (!d_vol || !d_stream) return fail(CVXGPU_ERR_ARG, "cvxgpu_decompress_device: null pointer");
    if (reinterpret_cast<uintptr_t>(d_stream) & 7) return fail(CVXGPU_ERR_ARG, "compressed stream must be 8-byte aligned");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    CU(cudaMemcpyAsync(ws->h_pinned + 64, d_stream, 32, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    int32_t h[8];
    memcpy(h, ws->h_pinned + 64, 32);
    return decompress_with_header(ws, d_vol, nx, ny, nz, h, d_stream, length, st);
}

// ---- host-pointer entry points ----------------------------------------------------------------
// 0: host (pageable or pinned), 1: device-accessible (device or managed)
static int pointer_on_device(const void* p)
{
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return 0; }
    return (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged) ? 1 : 0;
}

static int reserve(unsigned char** p, size_t* have, size_t want)
{
    if (want <= *have) return CVXGPU_OK;
    if (*p) cudaFree(*p);
    *p = nullptr; *have = 0;
    CU(cudaMalloc(p, want));
    *have = want;
    return CVXGPU_OK;
}

// the host-pointer entry points run on a private non-blocking stream of the calling thread's workspace: the legacy
// stream 0 would serialise against every other stream of the host application
static int priv_stream(cvxgpu_workspace* ws, cudaStream_t* st)
{
    if (!ws->priv_stream) CU(cudaStreamCreateWithFlags(&ws->priv_stream, cudaStreamNonBlocking));
    *st = ws->priv_stream;
    return CVXGPU_OK;
}

extern "C" int cvxgpu_compress(const float* vol, int nx, int ny, int nz, int bx, int by, int bz, float scale,
                               int use_local_rms, void* stream_out, long* length, float* ratio)
{
    cvxgpu_workspace* ws;
    int rc = get_ws(nullptr, &ws, nullptr);
    if (rc) return rc;
    rc = check_dims(nx, ny, nz, bx, by, bz);
    if (rc) return rc;
    if (!vol || !stream_out || !length) return fail(CVXGPU_ERR_ARG, "cvxgpu_compress: null pointer");
    const int local = use_local_rms ? 1 : 0;
    const size_t vol_bytes = (size_t)nx * ny * nz * 4;
    const long max_bytes = cvxgpu_max_compressed_bytes(nx, ny, nz, bx, by, bz, local);
    const int vol_dev = pointer_on_device(vol), out_dev = pointer_on_device(stream_out);
    cudaStream_t st;
    rc = priv_stream(ws, &st);
    if (rc) return rc;
    const float* d_vol = vol;
    if (!vol_dev) {
        rc = reserve(&ws->d_stage_vol, &ws->stage_vol_bytes, vol_bytes);
        if (rc) return rc;
        rc = cvx_stage_upload(ws, ws->d_stage_vol, vol, vol_bytes, st);      // pinned: one DMA; pageable: threaded bounce buffers
        if (rc) return rc;
        d_vol = reinterpret_cast<const float*>(ws->d_stage_vol);
    }
    if (out_dev) {
        // the class API has no capacity argument (CvxCompress.hxx:32-45): the caller's buffer is trusted
        return cvxgpu_compress_device(ws, d_vol, nx, ny, nz, bx, by, bz, scale, local, stream_out, max_bytes, length, ratio, st);
    }
    // first try a stream of at most a quarter of the volume (+ tables); if a block mix needs more, retry with the worst case
    const Shape s = make_shape(nx, ny, nz, bx, by, bz);
    long cap = 32 + 8 * s.nblk + (local ? 4 * s.nblk : 0) + (long)(vol_bytes / 4) + 4096;
    if (cap > max_bytes) cap = max_bytes;
    for (int attempt = 0; attempt < 2; ++attempt) {
        rc = reserve(&ws->d_stage_stream, &ws->stage_stream_bytes, (size_t)cap + 64);
        if (rc) return rc;
        rc = cvxgpu_compress_device(ws, d_vol, nx, ny, nz, bx, by, bz, scale, local, ws->d_stage_stream, cap, length, ratio, st);
        if (rc != CVXGPU_ERR_CAPACITY || cap == max_bytes) break;
        cap = max_bytes;
    }
    if (rc) return rc;
    // only [0, length - 7) is written: the final 7 slack bytes are never touched (CvxCompress.cpp:421)
    return cvx_stage_download(ws, stream_out, ws->d_stage_stream, (size_t)(*length - 7), st);
}

extern "C" int cvxgpu_decompress(float* vol, int nx, int ny, int nz, const void* stream_in, long length)
{
    cvxgpu_workspace* ws;
    int rc = get_ws(nullptr, &ws, nullptr);
    if (rc) return rc;
    if (!vol || !stream_in) return fail(CVXGPU_ERR_ARG, "cvxgpu_decompress: null pointer");
    const int vol_dev = pointer_on_device(vol), in_dev = pointer_on_device(stream_in);
    cudaStream_t st;
    rc = priv_stream(ws, &st);
    if (rc) return rc;
    if (in_dev && vol_dev) {
        rc = cvxgpu_decompress_device(ws, vol, nx, ny, nz, stream_in, length, st);
        if (rc) return rc;
        CU(cudaStreamSynchronize(st));
        return CVXGPU_OK;
    }
    int32_t h[8];
    if (in_dev) {
        if (reinterpret_cast<uintptr_t>(stream_in) & 7) return fail(CVXGPU_ERR_ARG, "compressed stream must be 8-byte aligned");
        CU(cudaMemcpy(h, stream_in, 32, cudaMemcpyDeviceToHost));
    } else {
        memcpy(h, stream_in, 32);
    }
    int hx, hy, hz, bx, by, bz, flags;
    rc = cvxgpu_parse_header(h, 32, &hx, &hy, &hz, &bx, &by, &bz, nullptr, &flags);
    if (rc) return rc;
    const void* d_stream = stream_in;
    if (!in_dev) {
        // the reference never reads compressed_length (CvxCompress.cpp:463-571): callers may pass 0 or garbage <= 0, and the
        // extent of the stream then follows from its own tables
        long len = length;
        if (len <= 0) {
            len = derive_stream_length(stream_in);
            if (len <= 0) return fail(CVXGPU_ERR_STREAM, "cvxgpu_decompress: cannot derive the stream length from its tables");
            length = len;
        }
        // opt-in Decompress_Safe (commented out in the reference, Test_With_Generated_Input.cpp:89): refuse a malformed
        // host stream before a byte of it reaches the GPU
        const int safe = cvxgpu_get_safe_decode();
        if (safe >= 0) {
            rc = cvxgpu_validate_stream(stream_in, len, safe, nullptr);
            if (rc) return rc;
        }
        rc = reserve(&ws->d_stage_stream, &ws->stage_stream_bytes, (size_t)len + 64);
        if (rc) return rc;
        rc = cvx_stage_upload(ws, ws->d_stage_stream, stream_in, (size_t)(len - 7 > 32 ? len - 7 : len), st);
        if (rc) return rc;
        d_stream = ws->d_stage_stream;
    }
    const size_t vol_bytes = (size_t)nx * ny * nz * 4;
    float* d_vol = vol;
    if (!vol_dev) {
        rc = reserve(&ws->d_stage_vol, &ws->stage_vol_bytes, vol_bytes);
        if (rc) return rc;
        d_vol = reinterpret_cast<float*>(ws->d_stage_vol);
    }
    rc = decompress_with_header(ws, d_vol, nx, ny, nz, h, d_stream, length, st);
    if (rc) return rc;
    if (!vol_dev) return cvx_stage_download(ws, vol, d_vol, vol_bytes, st);
    CU(cudaStreamSynchronize(st));
    return CVXGPU_OK;
}

// 32 header bytes of a stream that may live in host or device memory (used by the class's
// out-of-place Decompress, which must size its result before decoding)
int cvx_internal_fetch_header(const void* stream, unsigned char* hdr)
{
    if (!stream) return fail(CVXGPU_ERR_ARG, "null stream");
    if (pointer_on_device(stream)) {
        CU(cudaMemcpy(hdr, stream, 32, cudaMemcpyDeviceToHost));
    } else {
        memcpy(hdr, stream, 32);
    }
    return CVXGPU_OK;
}

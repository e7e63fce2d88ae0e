// cvx_kernels.cu -- sm_100a kernels of the CvxCompress hot path and their launchers.
//
//   k_sumsq / k_rms_finish   Compute_Global_RMS              CvxCompress.cpp:73-117
//   k_encode_fused           Copy_To_Block + Wavelet_Transform_Fast_Forward + Compute_Local_RMS +
//                            Run_Length_Encode_Slow + raw fallback + offset table
//                            (CvxCompress.cpp:318-420), one CTA per block, block in shared memory
//   k_decode_fused           Run_Length_Decode_Slow + Wavelet_Transform_Fast_Inverse + Copy_From_Block
//                            (CvxCompress.cpp:528-568)
//   k_wavelet_fused          forward or inverse transform of dense blocks (module-test sweep)
//   k_stage_*                the same pipeline staged through global scratch for block shapes
//                            that do not fit shared memory (any dimension > 64 or > 32768 cells)
//
// Data layout in HBM: the volume is nx-fastest fp32; the stream is the reference's byte
// format.  A block lives in shared memory as a dense, XOR-swizzled array (cvx_device.cuh).
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>
#include <string.h>
#include <stdlib.h>
#include "cvx_wavelet.cuh"
#include "cvx_rle.cuh"
#include "cvx_device.cuh"
#include "cvx_codec.cuh"
#include "cvx_fused32.cuh"
#include "cvx_internal.h"
#include "../../include/cvx_gpu.h"
#include <cudaTypedefs.h>

namespace cvx {

// ------------------------------------------------------------------------------------------
// Global RMS: per-CTA fp64 partial sums of v*v, then a fixed-order final sum (deterministic).
// ------------------------------------------------------------------------------------------
constexpr int SUMSQ_T = 256;

__global__ void __launch_bounds__(SUMSQ_T) k_sumsq(const float* __restrict__ vol, long n, double* __restrict__ partial)
{
    const long nthreads = (long)gridDim.x * SUMSQ_T;
    const long gtid = (long)blockIdx.x * SUMSQ_T + threadIdx.x;
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
    // peel to 16-byte alignment, then float4
    const uintptr_t addr = reinterpret_cast<uintptr_t>(vol);
    long head = ((16 - (addr & 15)) & 15) >> 2;
    if (head > n) head = n;
    if (gtid < head) { double d = vol[gtid]; a0 += d * d; }
    const float4* v4 = reinterpret_cast<const float4*>(vol + head);
    const long n4 = (n - head) >> 2;
    for (long i = gtid; i < n4; i += nthreads) {
        float4 q = __ldg(&v4[i]);
        double d0 = q.x, d1 = q.y, d2 = q.z, d3 = q.w;
        a0 += d0 * d0; a1 += d1 * d1; a2 += d2 * d2; a3 += d3 * d3;
    }
    const long tail0 = head + (n4 << 2);
    if (gtid < n - tail0) { double d = vol[tail0 + gtid]; a1 += d * d; }
    double s = (a0 + a1) + (a2 + a3);
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
    __shared__ double sw[SUMSQ_T / 32];
    if (lane_id() == 0) sw[warp_id()] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < SUMSQ_T / 32; ++w) t += sw[w];
        partial[blockIdx.x] = t;
    }
}

__global__ void k_sumsq_finish(const double* __restrict__ partial, int n, double* __restrict__ out)
{
    // one warp, fixed order: lane-strided partial sums, then a shuffle tree
    double s = 0.0;
    for (int i = threadIdx.x; i < n; i += 32) s += partial[i];
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
    if (threadIdx.x == 0) *out = s;
}

// ---- the same reduction, per block row (z range of one row of blocks), independent of how the volume is cut into
// slabs: the global RMS of a volume sharded over G GPUs must not depend on G, so every block row's sum
// is formed by a fixed number of CTAs in a fixed order relative to the ROW start (no alignment peel: rows that do not
// start on a 16-byte boundary take scalar loads into the same four accumulators), and the rows are added in row order.
constexpr int SUMSQ_ROW_CTAS = 128;

__global__ void __launch_bounds__(SUMSQ_T) k_sumsq_rows(const float* __restrict__ vol, long plane_cells, int bz, int nz,
                                                        double* __restrict__ partial)
{
    const int row = blockIdx.y;
    const int z0 = row * bz;
    const int nzr = min(bz, nz - z0);
    const long n = plane_cells * nzr;
    const float* base = vol + plane_cells * (long)z0;
    constexpr long NT = (long)SUMSQ_ROW_CTAS * SUMSQ_T;
    const long gtid = (long)blockIdx.x * SUMSQ_T + threadIdx.x;
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
    const long n4 = n >> 2;
    // element group g (4 cells) belongs to thread g mod NT and is added in increasing g: the order is a property of the
    // row alone.  Four groups are in flight per thread (a slab of a few rows on one GPU still has to fill the HBM pipes).
    if ((reinterpret_cast<uintptr_t>(base) & 15) == 0) {
        const float4* v4 = reinterpret_cast<const float4*>(base);
        long i = gtid;
        for (; i + 3 * NT < n4; i += 4 * NT) {
            const float4 q0 = __ldg(&v4[i]), q1 = __ldg(&v4[i + NT]), q2 = __ldg(&v4[i + 2 * NT]), q3 = __ldg(&v4[i + 3 * NT]);
#define CVX_ACC(q) { const double d0 = q.x, d1 = q.y, d2 = q.z, d3 = q.w; a0 += d0 * d0; a1 += d1 * d1; a2 += d2 * d2; a3 += d3 * d3; }
            CVX_ACC(q0) CVX_ACC(q1) CVX_ACC(q2) CVX_ACC(q3)
        }
        for (; i < n4; i += NT) { const float4 q = __ldg(&v4[i]); CVX_ACC(q) }
#undef CVX_ACC
    } else {
        for (long i = gtid; i < n4; i += NT) {
            const double d0 = __ldg(base + 4 * i), d1 = __ldg(base + 4 * i + 1), d2 = __ldg(base + 4 * i + 2), d3 = __ldg(base + 4 * i + 3);
            a0 += d0 * d0; a1 += d1 * d1; a2 += d2 * d2; a3 += d3 * d3;
        }
    }
    if (gtid < n - (n4 << 2)) { const double d = base[(n4 << 2) + gtid]; a0 += d * d; }
    double s = (a0 + a1) + (a2 + a3);
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
    __shared__ double sw[SUMSQ_T / 32];
    if (lane_id() == 0) sw[warp_id()] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < SUMSQ_T / 32; ++w) t += sw[w];
        partial[(long)row * SUMSQ_ROW_CTAS + blockIdx.x] = t;
    }
}

// one warp per block row: its SUMSQ_ROW_CTAS partials in a fixed order
__global__ void k_sumsq_rows_finish(const double* __restrict__ partial, double* __restrict__ rowsum)
{
    const int row = blockIdx.x;
    double s = 0.0;
    for (int i = threadIdx.x; i < SUMSQ_ROW_CTAS; i += 32) s += partial[(long)row * SUMSQ_ROW_CTAS + i];
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
    if (threadIdx.x == 0) rowsum[row] = s;
}

// total = rows added one after the other in row order.  `rows` may be the concatenation of per-rank segments of
// `seg` doubles each (the receive buffer of an all-gather): rank r contributes its first cnt[r] entries.
__global__ void k_sum_rows(const double* __restrict__ rows, int nseg, int seg, const int* __restrict__ cnt, int n_single,
                           double* __restrict__ total)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    double s = 0.0;
    if (nseg <= 0) {
        for (int i = 0; i < n_single; ++i) s += rows[i];
    } else {
        for (int r = 0; r < nseg; ++r)
            for (int i = 0; i < cnt[r]; ++i) s += rows[(long)r * seg + i];
    }
    *total = s;
}

// glob_mulfac = 1 / (rms * scale), 1 when rms == 0 or the result is not finite (CvxCompress.cpp:291-295);
// rms = (float)sqrt(sumsq / N) (:115-116).  With local RMS the global rms is taken as 1 (:249).
__device__ __forceinline__ float glob_mulfac_of(const double* sumsq, long ncells, float scale, int use_local)
{
    float rms = 1.0f;
    if (!use_local) rms = (float)sqrt(*sumsq / (double)ncells);
    float m = rms != 0.0f ? __fdiv_rn(1.0f, __fmul_rn(rms, scale)) : 1.0f;
    if (!isfinite(m)) m = 1.0f;
    return m;
}

__global__ void k_write_header(int32_t* hdr, int nx, int ny, int nz, int bx, int by, int bz, float scale, int use_local,
                               const double* sumsq, long ncells, float* mulfac_out)
{
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        float m = glob_mulfac_of(sumsq, ncells, scale, use_local);
        if (hdr) {
            hdr[0] = nx; hdr[1] = ny; hdr[2] = nz; hdr[3] = bx; hdr[4] = by; hdr[5] = bz;
            hdr[6] = __float_as_int(m);
            hdr[7] = use_local ? 1 : 0;
        }
        if (mulfac_out) *mulfac_out = m;
    }
}

// k_sum_rows and k_write_header in one launch (one graph node, one kernel-to-kernel latency less per Compress): the row
// sums are added in row order, then the header and glob_mulfac follow from the total.  rows == nullptr: the total is
// already in *sumsq (local RMS: not read at all).
__global__ void k_rms_header(const double* __restrict__ rows, int nseg, int seg, const int* __restrict__ cnt, int n_single,
                             double* __restrict__ sumsq, int32_t* hdr, int nx, int ny, int nz, int bx, int by, int bz, float scale,
                             int use_local, long ncells, float* mulfac_out)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    if (rows && !use_local) {
        double s = 0.0;
        if (nseg <= 0) {
            for (int i = 0; i < n_single; ++i) s += rows[i];
        } else {
            for (int r = 0; r < nseg; ++r)
                for (int i = 0; i < cnt[r]; ++i) s += rows[(long)r * seg + i];
        }
        *sumsq = s;
    }
    const float m = glob_mulfac_of(sumsq, ncells, scale, use_local);
    if (hdr) {
        hdr[0] = nx; hdr[1] = ny; hdr[2] = nz; hdr[3] = bx; hdr[4] = by; hdr[5] = bz;
        hdr[6] = __float_as_int(m);
        hdr[7] = use_local ? 1 : 0;
    }
    if (mulfac_out) *mulfac_out = m;
}

// everything an encode launch needs cleared, in one launch: the look-back words, the ticket, the status word and the byte count
__global__ void k_encode_prep(unsigned long long* __restrict__ status, long nblk, unsigned int* ticket, int* err, long* payload_bytes)
{
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nblk) status[i] = 0ull;
    if (i == 0) { *ticket = 0u; if (err) *err = 0; if (payload_bytes) *payload_bytes = 0; }
}

__global__ void k_offsets_add(unsigned long long* offs, long n, unsigned long long base)
{
    long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) offs[i] += base;          // bit 63 (raw flag) is untouched: base < 2^63
}

// z-slab sharding: sizes[r] = payload bytes of rank r (all-gathered).  bases = exclusive prefix
// (bases[world] = total payload), and this rank's table slice becomes global by adding bases[rank] (bit 63, the raw
// flag, is untouched: the sum stays below 2^63).
__global__ void k_shard_fixup(const long* __restrict__ sizes, int world, int rank, long* __restrict__ bases,
                              unsigned long long* __restrict__ offs, long n)
{
    long base = 0;
    for (int r = 0; r < rank; ++r) base += sizes[r];
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) offs[i] += (unsigned long long)base;
    if (i == 0) {
        long acc = 0;
        for (int r = 0; r < world; ++r) { bases[r] = acc; acc += sizes[r]; }
        bases[world] = acc;
    }
}
// out[0] = bases[rank + 1] - bases[rank] (this rank's payload bytes), out[1] = bases[rank]
__global__ void k_shard_extent(const long* __restrict__ bases, int rank, long* __restrict__ out)
{
    if (threadIdx.x == 0 && blockIdx.x == 0) { out[0] = bases[rank + 1] - bases[rank]; out[1] = bases[rank]; }
}

// ---- peer-memory exchange of the two scalars-per-rank of a sharded Compress (instead of two ncclAllGather launches) ----
// Every rank owns a MAILBOX in its HBM that all peers of the node can write through NVLink (CUDA IPC mapping, set up once
// per shard).  A push writes this rank's few values into its slot of EVERY rank's mailbox and then, after a system-scope
// fence, the call's epoch into its flag there; a consumer spins on the flags of its own mailbox (its own HBM) until all
// ranks' epochs have arrived.  Slots and flags are double-buffered by epoch parity: a rank can be at most one call ahead
// of the slowest reader of its previous values (it needs everybody's flags of call k before it can push call k + 1).
// The epoch lives in device memory and is advanced by the push kernel, so a captured CUDA graph replays correctly.
struct ShardMailbox {                       // layout parameters travel by value; the memory is one allocation per rank
    unsigned long long* base;               // [2 parities][2 kinds] x { values[world][cap], flags[world] }
    int world, cap;
    __host__ __device__ size_t kind_words() const { return (size_t)world * cap + world; }
    __device__ unsigned long long* values(int parity, int kind, int r) const { return base + (size_t)(parity * 2 + kind) * kind_words() + (size_t)r * cap; }
    __device__ unsigned long long* flag(int parity, int kind, int r) const { return base + (size_t)(parity * 2 + kind) * kind_words() + (size_t)world * cap + r; }
};
__device__ __forceinline__ void st_sys_u64(unsigned long long* p, unsigned long long v) { asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory"); }
__device__ __forceinline__ unsigned long long ld_sys_u64(const unsigned long long* p)
{
    unsigned long long v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
// src: n 64-bit values of this rank; peers[r] = rank r's mailbox base as mapped into this process
__global__ void __launch_bounds__(128) k_shard_push(const unsigned long long* __restrict__ src, int n, int kind, unsigned long long* const* __restrict__ peers,
                                                    int world, int cap, int rank, unsigned long long* epoch)
{
    __shared__ unsigned long long s_e;
    if (threadIdx.x == 0) s_e = *epoch + 1ull;
    __syncthreads();
    const unsigned long long e = s_e;
    const int parity = (int)(e & 1ull);
    for (int i = threadIdx.x; i < world * n; i += blockDim.x) {
        const int r = i / n, k = i - r * n;
        ShardMailbox mb; mb.base = peers[r]; mb.world = world; mb.cap = cap;
        st_sys_u64(mb.values(parity, kind, rank) + k, src[k]);
    }
    __threadfence_system();
    __syncthreads();
    for (int r = threadIdx.x; r < world; r += blockDim.x) {
        ShardMailbox mb; mb.base = peers[r]; mb.world = world; mb.cap = cap;
        st_sys_u64(mb.flag(parity, kind, rank), e);
    }
    if (threadIdx.x == 0) *epoch = e;
}
// all ranks' values of the current epoch have arrived in MY mailbox; false after ~4 s (a peer died: report, do not hang)
__device__ __forceinline__ bool shard_wait_all(const ShardMailbox& mb, int kind, unsigned long long e)
{
    __shared__ int s_bad;
    if (threadIdx.x == 0) s_bad = 0;
    __syncthreads();
    const int parity = (int)(e & 1ull);
    for (int r = threadIdx.x; r < mb.world; r += blockDim.x) {
        const long long t0 = clock64();
        while (ld_sys_u64(mb.flag(parity, kind, r)) != e)
            if (clock64() - t0 > 8000000000ll) { s_bad = 1; break; }
    }
    __threadfence_system();
    __syncthreads();
    return s_bad == 0;
}
// the mailbox variant of k_rms_header: waits for every rank's block-row sums, adds them in row order, writes header + glob_mulfac
__global__ void __launch_bounds__(64) k_rms_header_mailbox(ShardMailbox mb, const unsigned long long* epoch, const int* __restrict__ cnt,
                                                           double* __restrict__ sumsq, int32_t* hdr, int nx, int ny, int nz, int bx, int by, int bz,
                                                           float scale, int use_local, long ncells, float* mulfac_out, int* d_status)
{
    const unsigned long long e = *epoch;
    const bool ok = shard_wait_all(mb, 0, e);
    if (threadIdx.x != 0) return;
    if (!ok) *d_status = -2;
    double s = 0.0;
    for (int r = 0; r < mb.world; ++r)
        for (int i = 0; i < cnt[r]; ++i) s += __longlong_as_double((long long)ld_sys_u64(mb.values((int)(e & 1ull), 0, r) + i));
    *sumsq = s;
    const float m = glob_mulfac_of(sumsq, ncells, scale, use_local);
    hdr[0] = nx; hdr[1] = ny; hdr[2] = nz; hdr[3] = bx; hdr[4] = by; hdr[5] = bz;
    hdr[6] = __float_as_int(m);
    hdr[7] = use_local ? 1 : 0;
    *mulfac_out = m;
}
// the mailbox variant of k_shard_fixup: waits for every rank's payload size
__global__ void __launch_bounds__(256) k_shard_fixup_mailbox(ShardMailbox mb, const unsigned long long* epoch, int rank, long* __restrict__ bases,
                                                             unsigned long long* __restrict__ offs, long n, int* d_status)
{
    const unsigned long long e = *epoch;
    const bool ok = shard_wait_all(mb, 1, e);
    const int parity = (int)(e & 1ull);
    long base = 0;
    for (int r = 0; r < rank; ++r) base += (long)ld_sys_u64(mb.values(parity, 1, r));
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) offs[i] += (unsigned long long)base;
    if (i == 0) {
        if (!ok) *d_status = -2;
        long acc = 0;
        for (int r = 0; r < mb.world; ++r) { bases[r] = acc; acc += (long)ld_sys_u64(mb.values(parity, 1, r)); }
        bases[mb.world] = acc;
    }
}

// length = 32 + 8*nnn [+ 4*nnn] + payload + 7  (CvxCompress.cpp:421-422)
__global__ void k_finish_length(const long* payload_bytes, long table_bytes, long* length)
{
    if (threadIdx.x == 0 && blockIdx.x == 0) *length = 32 + table_bytes + *payload_bytes + 7;
}

// ------------------------------------------------------------------------------------------
// Transform passes on a block in shared memory.  One thread owns one line, in registers.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ int physr(int i, int sh) { return i ^ (((i >> sh) & 7) << 2); }

// x lines: the thread's row is contiguous -> LDS.128 / STS.128, conflict-free by the swizzle
template <int N, bool EXACT, bool FWD, int T>
__device__ __forceinline__ void pass_x(float* sm, int nrows)
{
    constexpr int SH = Swz<N>::SH;
    for (int r = threadIdx.x; r < nrows; r += T) {
        const int base = r * N;
        float in[N], v[N];
#pragma unroll
        for (int c = 0; c < N / 4; ++c) {
            float4 q = *reinterpret_cast<const float4*>(&sm[physr(base + 4 * c, SH)]);
            in[4 * c] = q.x; in[4 * c + 1] = q.y; in[4 * c + 2] = q.z; in[4 * c + 3] = q.w;
        }
        if (FWD) {
#pragma unroll
            for (int i = 0; i < N; ++i) v[i] = in[i];
            fwd_line<N, EXACT>(v);
#pragma unroll
            for (int c = 0; c < N / 4; ++c) {
                float4 q;
                q.x = v[sub_to_reg<N, EXACT>(4 * c)]; q.y = v[sub_to_reg<N, EXACT>(4 * c + 1)];
                q.z = v[sub_to_reg<N, EXACT>(4 * c + 2)]; q.w = v[sub_to_reg<N, EXACT>(4 * c + 3)];
                *reinterpret_cast<float4*>(&sm[physr(base + 4 * c, SH)]) = q;
            }
        } else {
#pragma unroll
            for (int i = 0; i < N; ++i) v[sub_to_reg<N, EXACT>(i)] = in[i];
            inv_line<N, EXACT>(v);
#pragma unroll
            for (int c = 0; c < N / 4; ++c) {
                float4 q = make_float4(v[4 * c], v[4 * c + 1], v[4 * c + 2], v[4 * c + 3]);
                *reinterpret_cast<float4*>(&sm[physr(base + 4 * c, SH)]) = q;
            }
        }
    }
}

// strided lines (y: stride BX, z: stride BX*BY): consecutive lanes own consecutive x
template <int N, bool EXACT, bool FWD, int T>
__device__ __forceinline__ void pass_strided(float* sm, int nlines, int lbx, int group_stride, int stride, int sh)
{
    // line l: x = l & (BX-1), g = l >> lbx; first element at g * group_stride + x
    for (int l = threadIdx.x; l < nlines; l += T) {
        const int base = (l >> lbx) * group_stride + (l & ((1 << lbx) - 1));
        float v[N];
        if (FWD) {
#pragma unroll
            for (int j = 0; j < N; ++j) v[j] = sm[physr(base + j * stride, sh)];
            fwd_line<N, EXACT>(v);
#pragma unroll
            for (int r = 0; r < N; ++r) sm[physr(base + reg_to_sub<N, EXACT>(r) * stride, sh)] = v[r];
        } else {
#pragma unroll
            for (int r = 0; r < N; ++r) v[r] = sm[physr(base + reg_to_sub<N, EXACT>(r) * stride, sh)];
            inv_line<N, EXACT>(v);
#pragma unroll
            for (int j = 0; j < N; ++j) sm[physr(base + j * stride, sh)] = v[j];
        }
    }
}

template <int MAXN, bool EXACT, bool FWD, int T>
__device__ __forceinline__ void pass_x_n(float* sm, int n, int nrows)
{
    switch (n) {
    case 8: pass_x<8, EXACT, FWD, T>(sm, nrows); break;
    case 16: pass_x<16, EXACT, FWD, T>(sm, nrows); break;
    case 32: pass_x<32, EXACT, FWD, T>(sm, nrows); break;
    case 64: if constexpr (MAXN >= 64) pass_x<64, EXACT, FWD, T>(sm, nrows); break;
    default: break;
    }
}
template <int MAXN, bool EXACT, bool FWD, int T>
__device__ __forceinline__ void pass_strided_n(float* sm, int n, int nlines, int lbx, int group_stride, int stride, int sh)
{
    switch (n) {
    case 8: pass_strided<8, EXACT, FWD, T>(sm, nlines, lbx, group_stride, stride, sh); break;
    case 16: pass_strided<16, EXACT, FWD, T>(sm, nlines, lbx, group_stride, stride, sh); break;
    case 32: pass_strided<32, EXACT, FWD, T>(sm, nlines, lbx, group_stride, stride, sh); break;
    case 64: if constexpr (MAXN >= 64) pass_strided<64, EXACT, FWD, T>(sm, nlines, lbx, group_stride, stride, sh); break;
    default: break;
    }
}

struct BlkShape {
    int bx, by, bz;       // block dims (bz may be 1)
    int lbx, lby;         // log2
    int B;                // cells
    int sh;               // swizzle shift
};

// forward: x, y (per z-slice) then z (Wavelet_Transform_Fast.cpp:24-145); inverse: same axis order
// with the synthesis operator (:148-279)
template <int MAXN, bool EXACT, bool FWD, int T>
__device__ __forceinline__ void transform_block(float* sm, const BlkShape& s)
{
    pass_x_n<MAXN, EXACT, FWD, T>(sm, s.bx, s.by * s.bz);
    __syncthreads();
    pass_strided_n<MAXN, EXACT, FWD, T>(sm, s.by, s.bx * s.bz, s.lbx, s.bx * s.by, s.bx, s.sh);
    __syncthreads();
    if (s.bz > 1) {
        pass_strided_n<MAXN, EXACT, FWD, T>(sm, s.bz, s.bx * s.by, s.lbx + s.lby, 0, s.bx * s.by, s.sh);
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------
// Copy_To_Block / Copy_From_Block (Block_Copy.cpp:21-116, :136-212) between the volume and the
// swizzled shared-memory block.  vec_ok: nx % 4 == 0 and the volume base is 16-byte aligned, so
// every block row starts on a 16-byte boundary and 128-bit accesses are legal.
// ------------------------------------------------------------------------------------------
struct VolView {
    int nx, ny, nz;
    int vec_ok;
};

template <int T>
__device__ __forceinline__ void gather_block(float* sm, const float* __restrict__ vol, const VolView& g,
                                             int x0, int y0, int z0, const BlkShape& s)
{
    if (g.vec_ok) {
        const int lcpr = s.lbx - 2;
        const int nchunks = s.B >> 2;
#pragma unroll 4
        for (int idx = threadIdx.x; idx < nchunks; idx += T) {
            const int c = idx & ((1 << lcpr) - 1);
            const int r = idx >> lcpr;
            const int gy = y0 + (r & (s.by - 1)), gz = z0 + (r >> s.lby), gx = x0 + 4 * c;
            float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
            if (gy < g.ny && gz < g.nz && gx < g.nx)
                q = __ldg(reinterpret_cast<const float4*>(vol + ((long)gz * g.ny + gy) * g.nx + gx));
            *reinterpret_cast<float4*>(&sm[physr(4 * idx, s.sh)]) = q;
        }
    } else {
#pragma unroll 4
        for (int i = threadIdx.x; i < s.B; i += T) {
            const int x = i & (s.bx - 1);
            const int r = i >> s.lbx;
            const int gy = y0 + (r & (s.by - 1)), gz = z0 + (r >> s.lby), gx = x0 + x;
            float q = 0.f;
            if (gy < g.ny && gz < g.nz && gx < g.nx) q = __ldg(vol + ((long)gz * g.ny + gy) * g.nx + gx);
            sm[physr(i, s.sh)] = q;
        }
    }
}

template <int T>
__device__ __forceinline__ void scatter_block(const float* sm, float* __restrict__ vol, const VolView& g,
                                              int x0, int y0, int z0, const BlkShape& s)
{
    if (g.vec_ok) {
        const int lcpr = s.lbx - 2;
        const int nchunks = s.B >> 2;
#pragma unroll 4
        for (int idx = threadIdx.x; idx < nchunks; idx += T) {
            const int c = idx & ((1 << lcpr) - 1);
            const int r = idx >> lcpr;
            const int gy = y0 + (r & (s.by - 1)), gz = z0 + (r >> s.lby), gx = x0 + 4 * c;
            if (gy < g.ny && gz < g.nz && gx < g.nx) {
                float4 q = *reinterpret_cast<const float4*>(&sm[physr(4 * idx, s.sh)]);
                *reinterpret_cast<float4*>(vol + ((long)gz * g.ny + gy) * g.nx + gx) = q;
            }
        }
    } else {
#pragma unroll 4
        for (int i = threadIdx.x; i < s.B; i += T) {
            const int x = i & (s.bx - 1);
            const int r = i >> s.lbx;
            const int gy = y0 + (r & (s.by - 1)), gz = z0 + (r >> s.lby), gx = x0 + x;
            if (gy < g.ny && gz < g.nz && gx < g.nx) vol[((long)gz * g.ny + gy) * g.nx + gx] = sm[physr(i, s.sh)];
        }
    }
}

// sum of squares of the block's coefficients over the CTA (Compute_Local_RMS, CvxCompress.cpp:119-142;
// the summation order differs from the reference's 8 serial lanes, so blkmulfac may differ in the
// last bits -- the value travels in the stream, so decoding is unaffected)
// EXACT: the reference's own order -- 8 lanes, lane j adds c[8i+j]^2 serially (mul then add, no FMA:
// the shipped build is -mavx only), then ((a0+a1)+(a2+a3)) + ((a4+a5)+(a6+a7)) from the two hadd
// steps and the 128-bit halves add (CvxCompress.cpp:124-139).  Bit-identical blkmulfac, at the cost
// of a serial chain of B/8 dependent FADDs on 8 threads.  `get` maps a raster index to the value.
template <int T, class Get>
__device__ __forceinline__ float cta_sumsq_exact(int B, float* s_red, Get get)
{
    if (threadIdx.x < 8) {
        float acc = 0.f;
        const int n = B >> 3;
#pragma unroll 8
        for (int i = 0; i < n; ++i) {
            float v = get(8 * i + (int)threadIdx.x);
            acc = __fadd_rn(acc, __fmul_rn(v, v));
        }
        s_red[threadIdx.x] = acc;
    }
    __syncthreads();
    const float lo = __fadd_rn(__fadd_rn(s_red[0], s_red[1]), __fadd_rn(s_red[2], s_red[3]));
    const float hi = __fadd_rn(__fadd_rn(s_red[4], s_red[5]), __fadd_rn(s_red[6], s_red[7]));
    __syncthreads();
    return __fadd_rn(lo, hi);
}

template <int T, class Bar = CtaBar>
__device__ __forceinline__ float cta_sumsq(const float* sm, int B, float* s_red)
{
    float a = 0.f;
    for (int i = threadIdx.x; i < (B >> 2); i += T) {
        float4 q = reinterpret_cast<const float4*>(sm)[i];     // any order: the swizzle is a permutation
        a = __fmaf_rn(q.x, q.x, a); a = __fmaf_rn(q.y, q.y, a); a = __fmaf_rn(q.z, q.z, a); a = __fmaf_rn(q.w, q.w, a);
    }
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) a += __shfl_xor_sync(0xffffffffu, a, d);
    if (lane_id() == 0) s_red[warp_id()] = a;
    Bar::sync();
    float t = 0.f;
    for (int w = 0; w < (T + 31) / 32; ++w) t += s_red[w];
    Bar::sync();
    return t;
}

// ------------------------------------------------------------------------------------------
// Fused encode: persistent CTAs take blocks in natural order from an atomic ticket.
// ------------------------------------------------------------------------------------------
struct EncParams {
    const float* vol;
    VolView g;
    BlkShape s;
    int nbx, nby;
    long nblk;
    float scale;
    int use_local;
    const float* glob_mulfac;        // device scalar
    unsigned long long* offs;        // nblk
    float* blkmulfac;                // nblk or null
    uint8_t* payload;
    long cap;
    unsigned long long* status;      // nblk look-back words, zeroed
    unsigned int* ticket;            // zeroed
    long* payload_bytes;             // out
    int* err;                        // out: set to CVXGPU_ERR_CAPACITY on overflow
    int nseg, gps;                   // RLE segmentation of a block
};

constexpr int SEG_CAP = 1024;

// raw fallback payload: B fp32 coefficients in raster order (CvxCompress.cpp:355-360)
template <int T, class Phys>
__device__ __forceinline__ void store_raw(const float* coef, long B, uint8_t* dst, Phys phys)
{
    if ((reinterpret_cast<uintptr_t>(dst) & 3) == 0) {
        float* d = reinterpret_cast<float*>(dst);
        for (long i = threadIdx.x; i < B; i += T) d[i] = coef[phys(i)];
    } else {
        for (long i = threadIdx.x; i < B; i += T) {
            uint32_t u = __float_as_uint(coef[phys(i)]);
            dst[4 * i] = (uint8_t)u; dst[4 * i + 1] = (uint8_t)(u >> 8);
            dst[4 * i + 2] = (uint8_t)(u >> 16); dst[4 * i + 3] = (uint8_t)(u >> 24);
        }
    }
}

// shared tail of the fused and the staged encoders: plan, look-back, table entry, emit
template <int T, class Get8, class Phys>
__device__ __forceinline__ void encode_tail(const EncParams& p, long t, const float* coef, float mulfac,
                                            Get8 get8, Phys phys, int4* s_seg, int* scratch, unsigned long long* s_off)
{
    const long B = p.s.B;
    int len = cta_rle_plan<T>(p.nseg, p.gps, mulfac, get8, s_seg, scratch);
    const bool raw = (long)len > 4 * B;                   // CvxCompress.cpp:350
    const unsigned long long nbytes = raw ? (unsigned long long)(4 * B) : (unsigned long long)len;
    if (warp_id() == 0) {
        unsigned long long excl = lookback_exclusive(p.status, t, nbytes);
        if (lane_id() == 0) {
            *s_off = excl;
            p.offs[t] = excl | (raw ? 0x8000000000000000ull : 0ull);      // :382-384
            if (t == p.nblk - 1) *p.payload_bytes = (long)(excl + nbytes);
        }
    }
    __syncthreads();
    const unsigned long long off = *s_off;
    if ((long)(off + nbytes) > p.cap) {
        if (threadIdx.x == 0) *p.err = -3;
    } else {
        uint8_t* dst = p.payload + off;
        if (raw) store_raw<T>(coef, B, dst, phys);
        else cta_rle_emit<T>(p.nseg, p.gps, mulfac, get8, s_seg, [&](int pos, uint8_t b) { dst[pos] = b; });
    }
    __syncthreads();
}

template <int MAXN, bool EXACT, int T, int MINB>
__global__ void __launch_bounds__(T, MINB) k_encode_fused(const EncParams p)
{
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    float* sm = reinterpret_cast<float*>(smem_raw);
    int4* s_seg = reinterpret_cast<int4*>(smem_raw + (size_t)p.s.B * 4);
    __shared__ int scratch[72];
    __shared__ float s_red[32];
    __shared__ unsigned int s_ticket;
    __shared__ unsigned long long s_off;
    const BlkShape s = p.s;
    const float gmul = p.use_local ? 1.0f : *p.glob_mulfac;
    for (;;) {
        if (threadIdx.x == 0) s_ticket = atomicAdd(p.ticket, 1u);
        __syncthreads();
        const long t = s_ticket;
        if (t >= p.nblk) break;
        const int iix = (int)(t % p.nbx);
        const long q = t / p.nbx;
        const int iiy = (int)(q % p.nby), iiz = (int)(q / p.nby);
        gather_block<T>(sm, p.vol, p.g, iix * s.bx, iiy * s.by, iiz * s.bz, s);
        __syncthreads();
        transform_block<MAXN, EXACT, true, T>(sm, s);
        float mulfac = gmul;
        if (p.use_local) {
            const int shl = s.sh;
            float ss = EXACT ? cta_sumsq_exact<T>(s.B, s_red, [&](int i) { return sm[physr(i, shl)]; })
                             : cta_sumsq<T>(sm, s.B, s_red);
            float rms = sqrtf(__fdiv_rn(ss, (float)s.B));
            mulfac = rms != 0.0f ? __fdiv_rn(1.0f, __fmul_rn(rms, p.scale)) : 1.0f;     // CvxCompress.cpp:346
            if (threadIdx.x == 0) p.blkmulfac[t] = mulfac;
        }
        const int sh = s.sh;
        auto get8 = [&](int g, float* c) {
            const int i = g * 8;
            float4 a = *reinterpret_cast<const float4*>(&sm[physr(i, sh)]);
            float4 b = *reinterpret_cast<const float4*>(&sm[physr(i + 4, sh)]);
            c[0] = a.x; c[1] = a.y; c[2] = a.z; c[3] = a.w; c[4] = b.x; c[5] = b.y; c[6] = b.z; c[7] = b.w;
        };
        encode_tail<T>(p, t, sm, mulfac, get8, [sh](long i) { return (long)physr((int)i, sh); }, s_seg, scratch, &s_off);
    }
}

#include "cvx_encode32.inl"
#include "cvx_encode32ws.inl"
#include "cvx_encode_sparse.inl"

// ------------------------------------------------------------------------------------------
// Fused decode
// ------------------------------------------------------------------------------------------
struct DecParams {
    float* vol;
    VolView g;
    BlkShape s;
    int nbx, nby;
    long nblk;
    float glob_mulfac;
    int use_local;
    const unsigned long long* offs;
    const float* blkmulfac;
    const uint8_t* payload;
    long payload_bytes;
    // when non-null these device scalars replace the two by-value fields above (a decode enqueued behind the encode that
    // produces them: no host round trip, CUDA-graph capturable)
    const float* gm_ptr;
    const long* pbytes_ptr;
    const long* bias_ptr;            // table offsets are biased by *bias_ptr (a rank's slice of a sharded stream: global offsets)
};
__device__ __forceinline__ DecParams dec_resolve(const DecParams& in)
{
    DecParams p = in;
    if (in.gm_ptr) p.glob_mulfac = *in.gm_ptr;
    if (in.pbytes_ptr) p.payload_bytes = *in.pbytes_ptr;
    if (in.bias_ptr) { const long b = *in.bias_ptr; p.payload -= b; p.payload_bytes += b; }
    return p;
}

template <int T, class PutV, class Phys>
__device__ __forceinline__ void decode_head(const DecParams& p, long t, float* coef, PutV put, Phys phys,
                                            uint32_t* s_win, int* s_ext, int* scratch)
{
    const long B = p.s.B;
    unsigned long long o = p.offs[t];
    const bool raw = (o >> 63) != 0;                               // CvxCompress.cpp:545-547
    o &= 0x7FFFFFFFFFFFFFFFull;
    const uint8_t* lo = p.payload;
    const uint8_t* hi = p.payload + p.payload_bytes;
    const uint8_t* src = p.payload + o;
    if (raw) {                                                     // :555
        if ((reinterpret_cast<uintptr_t>(src) & 3) == 0 && src + 4 * B <= hi) {
            const float* f = reinterpret_cast<const float*>(src);
            for (long i = threadIdx.x; i < B; i += T) coef[phys(i)] = f[i];
        } else {
            for (long i = threadIdx.x; i < B; i += T) {
                uint32_t u = 0;
                for (int k = 0; k < 4; ++k) {
                    const uint8_t* b = src + 4 * i + k;
                    if (b < hi) u |= (uint32_t)(*b) << (8 * k);
                }
                coef[phys(i)] = __uint_as_float(u);
            }
        }
    } else {
        float mulfac = p.use_local ? p.blkmulfac[t] : p.glob_mulfac;
        float scalefac = __fdiv_rn(1.0f, mulfac);                  // Run_Length_Encode_Slow.cpp:392
        cta_rle_decode<T>(lo, hi, src, (int)B, scalefac, put, s_win, s_ext, scratch);
    }
    __syncthreads();
}

template <int MAXN, bool EXACT, int T, int MINB>
__global__ void __launch_bounds__(T, MINB) k_decode_fused(const DecParams pin)
{
    const DecParams p = dec_resolve(pin);
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    float* sm = reinterpret_cast<float*>(smem_raw);
    uint32_t* s_win = reinterpret_cast<uint32_t*>(smem_raw + (size_t)p.s.B * 4);
    int* s_ext = reinterpret_cast<int*>(s_win + (T + 2) * 4);
    __shared__ int scratch[72];
    const BlkShape s = p.s;
    const int sh = s.sh;
    for (long t = blockIdx.x; t < p.nblk; t += gridDim.x) {
        for (int i = threadIdx.x; i < (s.B >> 2); i += T) reinterpret_cast<float4*>(sm)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        __syncthreads();
        decode_head<T>(p, t, sm, [&](int i, float v) { sm[physr(i, sh)] = v; },
                       [sh](long i) { return (long)physr((int)i, sh); }, s_win, s_ext, scratch);
        transform_block<MAXN, EXACT, false, T>(sm, s);
        const int iix = (int)(t % p.nbx);
        const long q = t / p.nbx;
        const int iiy = (int)(q % p.nby), iiz = (int)(q / p.nby);
        scatter_block<T>(sm, p.vol, p.g, iix * s.bx, iiy * s.by, iiz * s.bz, s);
        __syncthreads();
    }
}

#include "cvx_decode32.inl"
#include "cvx_decode32ws.inl"
#include "cvx_decode_win.inl"

// forward / inverse transform of dense blocks: block k is the z-range [k*bz, (k+1)*bz) of a
// bx * by * (bz*nblocks) volume
template <int MAXN, bool EXACT, bool FWD, int T, int MINB>
__global__ void __launch_bounds__(T, MINB) k_wavelet_fused(float* blocks, long nblocks, const BlkShape s)
{
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    float* sm = reinterpret_cast<float*>(smem_raw);
    VolView g; g.nx = s.bx; g.ny = s.by; g.nz = s.bz; g.vec_ok = 1;
    for (long k = blockIdx.x; k < nblocks; k += gridDim.x) {
        float* blk = blocks + k * (long)s.B;
        gather_block<T>(sm, blk, g, 0, 0, 0, s);
        __syncthreads();
        transform_block<MAXN, EXACT, FWD, T>(sm, s);
        scatter_block<T>(sm, blk, g, 0, 0, 0, s);
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------
// Large-block path (any dimension > 64, or more than 32768 cells): blocks of a batch are dense arrays in global
// scratch; the three line passes are cvx_bigblock.inl, the coder / parser below work on the scratch blocks.
// ------------------------------------------------------------------------------------------
#include "cvx_bigblock.inl"

constexpr int STAGE_T = 512;

__global__ void __launch_bounds__(STAGE_T, 1) k_stage_encode(const EncParams p, const float* scratch, long blk_first, int nb)
{
    __shared__ int4 s_seg[SEG_CAP];
    __shared__ int scratch_i[72];
    __shared__ float s_red[32];
    __shared__ unsigned long long s_off;
    const float gmul = p.use_local ? 1.0f : *p.glob_mulfac;
    // blocks of the batch are taken in order: CTA b encodes block blk_first + b (+ k * gridDim.x)
    for (int k = blockIdx.x; k < nb; k += gridDim.x) {
        const long t = blk_first + k;
        const float* coef = scratch + (long)k * p.s.B;
        float mulfac = gmul;
        if (p.use_local) {
            // large-block path: always the reference's summation order (the multiplier travels in the stream)
            float ss = cta_sumsq_exact<STAGE_T>(p.s.B, s_red, [&](int i) { return __ldg(coef + i); });
            float rms = sqrtf(__fdiv_rn(ss, (float)p.s.B));
            mulfac = rms != 0.0f ? __fdiv_rn(1.0f, __fmul_rn(rms, p.scale)) : 1.0f;
            if (threadIdx.x == 0) p.blkmulfac[t] = mulfac;
        }
        auto get8 = [&](int g, float* c) {
            const float4* q = reinterpret_cast<const float4*>(coef + (long)g * 8);
            float4 a = q[0], b = q[1];
            c[0] = a.x; c[1] = a.y; c[2] = a.z; c[3] = a.w; c[4] = b.x; c[5] = b.y; c[6] = b.z; c[7] = b.w;
        };
        encode_tail<STAGE_T>(p, t, coef, mulfac, get8, [](long i) { return i; }, s_seg, scratch_i, &s_off);
    }
}

// ---- large blocks, coder on many CTAs -------------------------------------------------------------------------------
// k_stage_encode gives a block to ONE CTA: at 256^3 (64 MiB of coefficients) every thread walks 32768 cells twice, and a
// volume has fewer such blocks than the GPU has SMs -- 315 ms per Compress of the config-3 volume, no faster than the
// reference's host threads.  The segment coder is already a parallel decomposition (cvx_codec.cuh: measure every segment,
// scan the zero-run carries and the byte counts, emit every segment at its offset); here its three steps are three
// launches over the whole BATCH with segments of STG_V cells:
//   k_stage_mulfac   (local RMS only) one CTA per block: the reference's summation order -> block multiplier
//   k_stage_measure  one thread per segment, grid-wide: { has, lead, trail, inner } into global memory
//   k_stage_plan     one CTA per block: carry + byte-count scans over the block's segments (cta_rle_plan on the global array),
//                    raw decision (CvxCompress.cpp:350), payload offset by the decoupled look-back, table entry
//   k_stage_emit     one thread per segment, grid-wide: tokens (or raw coefficients) at the segment's final address
// Byte-identical to k_stage_encode: the segmentation does not show in the stream.
constexpr int STG_V = 256;                       // cells per segment (32 groups of eight)
struct StageCoder {
    const float* scratch;            // nb dense blocks of coefficients
    int4* seg;                       // nb * (B / STG_V) segment records
    float* mulfac;                   // nb block multipliers
    unsigned long long* blkoff;      // nb payload offsets (bit 63: raw, all ones: does not fit)
    long blk_first;
    int nb;
};
__device__ __forceinline__ void stage_get8(const float* coef, long g, float* c)
{
    const float4* q = reinterpret_cast<const float4*>(coef + g * 8);
    const float4 a = q[0], b = q[1];
    c[0] = a.x; c[1] = a.y; c[2] = a.z; c[3] = a.w; c[4] = b.x; c[5] = b.y; c[6] = b.z; c[7] = b.w;
}
__global__ void __launch_bounds__(STAGE_T, 1) k_stage_mulfac(const EncParams p, const StageCoder sc)
{
    __shared__ float s_red[32];
    const float gmul = p.use_local ? 1.0f : *p.glob_mulfac;
    for (int k = blockIdx.x; k < sc.nb; k += gridDim.x) {
        float mulfac = gmul;
        if (p.use_local) {
            // large-block path: always the reference's summation order (the multiplier travels in the stream)
            const float* coef = sc.scratch + (long)k * p.s.B;
            const float ss = cta_sumsq_exact<STAGE_T>(p.s.B, s_red, [&](int i) { return __ldg(coef + i); });
            const float rms = sqrtf(__fdiv_rn(ss, (float)p.s.B));
            mulfac = rms != 0.0f ? __fdiv_rn(1.0f, __fmul_rn(rms, p.scale)) : 1.0f;
            if (threadIdx.x == 0) p.blkmulfac[sc.blk_first + k] = mulfac;
        }
        if (threadIdx.x == 0) sc.mulfac[k] = mulfac;
    }
}
__global__ void __launch_bounds__(256) k_stage_measure(const EncParams p, const StageCoder sc)
{
    const long S = p.s.B / STG_V, total = (long)sc.nb * S;
    for (long sid = (long)blockIdx.x * 256 + threadIdx.x; sid < total; sid += (long)gridDim.x * 256) {
        const long k = sid / S, seg = sid - k * S;
        const float* coef = sc.scratch + k * p.s.B + seg * STG_V;
        const SegSum m = seg_measure(STG_V / 8, sc.mulfac[k], [&](int g, float* c) { stage_get8(coef, g, c); });
        sc.seg[sid] = make_int4(m.has, m.lead, m.trail, m.inner);
    }
}
__global__ void __launch_bounds__(STAGE_T, 1) k_stage_plan(const EncParams p, const StageCoder sc)
{
    __shared__ int scratch_i[72];
    const int S = (int)(p.s.B / STG_V);
    // blocks of the batch are taken in order: CTA b plans block blk_first + b (+ k * gridDim.x)
    for (int k = blockIdx.x; k < sc.nb; k += gridDim.x) {
        const long t = sc.blk_first + k;
        const int len = cta_rle_plan<STAGE_T>(S, STG_V / 8, 0.f, [](int, float*) {}, sc.seg + (long)k * S, scratch_i, true);
        const bool raw = (long)len > 4L * p.s.B;              // CvxCompress.cpp:350
        const unsigned long long nbytes = raw ? (unsigned long long)(4L * p.s.B) : (unsigned long long)len;
        if (warp_id() == 0) {
            const unsigned long long excl = lookback_exclusive(p.status, t, nbytes);
            if (lane_id() == 0) {
                p.offs[t] = excl | (raw ? 0x8000000000000000ull : 0ull);      // :382-384
                if (t == p.nblk - 1) *p.payload_bytes = (long)(excl + nbytes);
                const bool fits = (long)(excl + nbytes) <= p.cap;
                if (!fits) *p.err = -3;
                sc.blkoff[k] = fits ? (excl | (raw ? 0x8000000000000000ull : 0ull)) : ~0ull;
            }
        }
        __syncthreads();
    }
}
__global__ void __launch_bounds__(256) k_stage_emit(const EncParams p, const StageCoder sc)
{
    const long S = p.s.B / STG_V, total = (long)sc.nb * S;
    for (long sid = (long)blockIdx.x * 256 + threadIdx.x; sid < total; sid += (long)gridDim.x * 256) {
        const long k = sid / S, seg = sid - k * S;
        const unsigned long long bo = sc.blkoff[k];
        if (bo == ~0ull) continue;                            // does not fit: nothing is written, the call reports CAPACITY
        const float* coef = sc.scratch + k * p.s.B + seg * STG_V;
        uint8_t* dst = p.payload + (bo & 0x7FFFFFFFFFFFFFFFull);
        if (bo >> 63) {                                       // raw block: B fp32 coefficients in raster order (:355-360)
            uint8_t* d = dst + 4 * seg * STG_V;
            if ((reinterpret_cast<uintptr_t>(d) & 3) == 0) {
                for (int i = 0; i < STG_V; ++i) reinterpret_cast<float*>(d)[i] = coef[i];
            } else {
                for (int i = 0; i < STG_V; ++i) {
                    const uint32_t u = __float_as_uint(coef[i]);
                    d[4 * i] = (uint8_t)u; d[4 * i + 1] = (uint8_t)(u >> 8); d[4 * i + 2] = (uint8_t)(u >> 16); d[4 * i + 3] = (uint8_t)(u >> 24);
                }
            }
            continue;
        }
        const int4 q = sc.seg[sid];                           // { zeros carried in, byte offset inside the block payload, has, - }
        const bool last = seg == S - 1;
        if (!q.z && !last) continue;
        seg_emit(STG_V / 8, sc.mulfac[k], q.x, q.y, last, [&](int g, float* c) { stage_get8(coef, g, c); },
                 [&](int pos, uint8_t b) { dst[pos] = b; });
    }
}

__global__ void __launch_bounds__(STAGE_T, 1) k_stage_decode(const DecParams pin, float* scratch, long blk_first, int nb)
{
    const DecParams p = dec_resolve(pin);
    __shared__ uint32_t s_win[(STAGE_T + 2) * 4];
    __shared__ int s_ext[STAGE_T];
    __shared__ int scratch_i[72];
    for (int k = blockIdx.x; k < nb; k += gridDim.x) {
        const long t = blk_first + k;
        float* coef = scratch + (long)k * p.s.B;
        for (long i = threadIdx.x; i < (p.s.B >> 2); i += STAGE_T) reinterpret_cast<float4*>(coef)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        __syncthreads();
        decode_head<STAGE_T>(p, t, coef, [&](int i, float v) { coef[i] = v; }, [](long i) { return i; }, s_win, s_ext, scratch_i);
    }
}

// ------------------------------------------------------------------------------------------
// Host side: shapes, workspace growth, launchers
// ------------------------------------------------------------------------------------------
static int ilog2(int v) { int l = 0; while ((1 << l) < v) ++l; return l; }

bool valid_block_size(int bx, int by, int bz)          // CvxCompress.cpp:54-71 and the asserts :246-248
{
    auto p2 = [](int v) { return v > 0 && (v & (v - 1)) == 0; };
    if (!p2(bx) || bx < 8 || bx > 256) return false;
    if (!p2(by) || by < 8 || by > 256) return false;
    if (bz == 1) return true;
    return p2(bz) && bz >= 8 && bz <= 256;
}

bool fused_ok(int bx, int by, int bz)
{
    return bx <= 64 && by <= 64 && bz <= 64 && (long)bx * by * bz <= 32768;
}

Shape make_shape(int nx, int ny, int nz, int bx, int by, int bz)
{
    Shape s;
    s.nx = nx; s.ny = ny; s.nz = nz; s.bx = bx; s.by = by; s.bz = bz;
    s.nbx = (int)(((long)nx + bx - 1) / bx); s.nby = (int)(((long)ny + by - 1) / by); s.nbz = (int)(((long)nz + bz - 1) / bz);
    s.nblk = (long)s.nbx * s.nby * s.nbz;
    return s;
}

static BlkShape blk_shape(const Shape& s)
{
    BlkShape b;
    b.bx = s.bx; b.by = s.by; b.bz = s.bz;
    b.lbx = ilog2(s.bx); b.lby = ilog2(s.by);
    b.B = s.bx * s.by * s.bz;
    b.sh = b.lbx < 5 ? 5 : b.lbx;
    return b;
}

static VolView vol_view(const Shape& s, const void* vol)
{
    VolView g;
    g.nx = s.nx; g.ny = s.ny; g.nz = s.nz;
    g.vec_ok = ((s.nx & 3) == 0 && (reinterpret_cast<uintptr_t>(vol) & 15) == 0) ? 1 : 0;
    return g;
}

// 3-D tiled tensor map over an nx-fastest fp32 volume, box bx x by x bz, SWIZZLE_128B (the block layout of
// cvx_device.cuh), out-of-bounds cells read as zero (== Copy_To_Block's padding, Block_Copy.cpp:92-115).
// Returns false when TMA cannot address the volume (base not 16-byte aligned or nx % 4 != 0): the kernels then
// take their plain-load path.
bool make_volume_tmap(CUtensorMap* map, const void* d_vol, int nx, int ny, int nz, int bx, int by, int bz)
{
    memset(map, 0, sizeof(*map));
    if ((reinterpret_cast<uintptr_t>(d_vol) & 15) != 0 || (nx & 3) != 0 || bx * 4 > 128) return false;
    static PFN_cuTensorMapEncodeTiled_v12000 encode = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
        else
            cudaGetLastError();
    }
    if (!encode) return false;
    const cuuint64_t dims[3] = {(cuuint64_t)nx, (cuuint64_t)ny, (cuuint64_t)nz};
    const cuuint64_t strides[2] = {(cuuint64_t)nx * 4, (cuuint64_t)nx * (cuuint64_t)ny * 4};
    const cuuint32_t box[3] = {(cuuint32_t)bx, (cuuint32_t)by, (cuuint32_t)bz};
    const cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<void*>(d_vol), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS;
}

cudaError_t ws_reserve_status(cvxgpu_workspace* ws, long nblk)
{
    if (nblk <= ws->status_cap) return cudaSuccess;
    CvxRelaxCapture relax;
    if (ws->d_status) cudaFree(ws->d_status);
    ws->d_status = nullptr; ws->status_cap = 0;
    long cap = nblk + nblk / 4 + 1024;
    cudaError_t e = cudaMalloc(&ws->d_status, sizeof(unsigned long long) * (size_t)cap);
    if (e == cudaSuccess) ws->status_cap = cap;
    return e;
}

cudaError_t ws_reserve_scratch(cvxgpu_workspace* ws, size_t bytes)
{
    if (bytes <= ws->scratch_bytes) return cudaSuccess;
    CvxRelaxCapture relax;
    if (ws->d_scratch) cudaFree(ws->d_scratch);
    ws->d_scratch = nullptr; ws->scratch_bytes = 0;
    cudaError_t e = cudaMalloc(&ws->d_scratch, bytes);
    if (e == cudaSuccess) ws->scratch_bytes = bytes;
    return e;
}

cudaError_t launch_sumsq(cvxgpu_workspace* ws, const float* d_vol, long ncells, double* d_sumsq, cudaStream_t st)
{
    int grid = ws->n_partial;
    long need = (ncells / 4 + SUMSQ_T - 1) / SUMSQ_T;
    if (need < 1) need = 1;
    if (grid > need) grid = (int)need;
    k_sumsq<<<grid, SUMSQ_T, 0, st>>>(d_vol, ncells, ws->d_partial);
    k_sumsq_finish<<<1, 32, 0, st>>>(ws->d_partial, grid, d_sumsq);
    return cudaGetLastError();
}

cudaError_t ws_reserve_rows(cvxgpu_workspace* ws, int rows)
{
    if (rows <= ws->rows_cap) return cudaSuccess;
    CvxRelaxCapture relax;
    if (ws->d_rowpartial) cudaFree(ws->d_rowpartial);
    if (ws->d_rowsum) cudaFree(ws->d_rowsum);
    ws->d_rowpartial = nullptr; ws->d_rowsum = nullptr; ws->rows_cap = 0;
    const int cap = rows + 64;
    cudaError_t e = cudaMalloc(&ws->d_rowpartial, sizeof(double) * (size_t)cap * SUMSQ_ROW_CTAS);
    if (e != cudaSuccess) return e;
    e = cudaMalloc(&ws->d_rowsum, sizeof(double) * (size_t)cap);
    if (e != cudaSuccess) return e;
    ws->rows_cap = cap;
    return cudaSuccess;
}

// per-block-row sums of squares of a volume or z-slab that starts on a block-row boundary: d_rowsum[(nz + bz - 1) / bz]
cudaError_t launch_sumsq_rows(cvxgpu_workspace* ws, const float* d_vol, long plane_cells, int bz, int nz, double* d_rowsum,
                              cudaStream_t st)
{
    const int rows = (nz + bz - 1) / bz;
    cudaError_t e = ws_reserve_rows(ws, rows);
    if (e != cudaSuccess) return e;
    if (!d_rowsum) d_rowsum = ws->d_rowsum;
    for (int r0 = 0; r0 < rows; r0 += 32768) {         // gridDim.y limit
        const int nr = rows - r0 < 32768 ? rows - r0 : 32768;
        k_sumsq_rows<<<dim3(SUMSQ_ROW_CTAS, nr), SUMSQ_T, 0, st>>>(d_vol + plane_cells * (long)bz * r0, plane_cells, bz, nz - r0 * bz,
                                                                    ws->d_rowpartial + (long)r0 * SUMSQ_ROW_CTAS);
    }
    k_sumsq_rows_finish<<<rows, 32, 0, st>>>(ws->d_rowpartial, d_rowsum);
    return cudaGetLastError();
}

// d_total = rows in order; with nseg > 0 `d_rows` is nseg segments of `seg` doubles of which the first d_cnt[r] count
cudaError_t launch_sum_rows(const double* d_rows, int nseg, int seg, const int* d_cnt, int n_single, double* d_total, cudaStream_t st)
{
    k_sum_rows<<<1, 32, 0, st>>>(d_rows, nseg, seg, d_cnt, n_single, d_total);
    return cudaGetLastError();
}

cudaError_t launch_rms_header(const double* d_rows, int nseg, int seg, const int* d_cnt, int n_single, double* d_sumsq,
                              void* d_stream, const Shape& s, float scale, int use_local, long ncells_total, float* d_mulfac_out,
                              cudaStream_t st)
{
    k_rms_header<<<1, 32, 0, st>>>(d_rows, nseg, seg, d_cnt, n_single, d_sumsq, reinterpret_cast<int32_t*>(d_stream), s.nx, s.ny, s.nz,
                                   s.bx, s.by, s.bz, scale, use_local, ncells_total, d_mulfac_out);
    return cudaGetLastError();
}

cudaError_t launch_write_header(void* d_stream, const Shape& s, float scale, int use_local, const double* d_sumsq,
                                long ncells_total, float* d_mulfac_out, cudaStream_t st)
{
    k_write_header<<<1, 32, 0, st>>>(reinterpret_cast<int32_t*>(d_stream), s.nx, s.ny, s.nz, s.bx, s.by, s.bz, scale,
                                     use_local, d_sumsq, ncells_total, d_mulfac_out);
    return cudaGetLastError();
}

cudaError_t launch_finish_length(const long* d_payload_bytes, long table_bytes, long* d_length, cudaStream_t st)
{
    k_finish_length<<<1, 32, 0, st>>>(d_payload_bytes, table_bytes, d_length);
    return cudaGetLastError();
}

cudaError_t launch_shard_fixup(const long* d_sizes, int world, int rank, long* d_bases, unsigned long long* d_offs, long n, cudaStream_t st)
{
    const long nn = n > 0 ? n : 1;
    k_shard_fixup<<<(unsigned)((nn + 255) / 256), 256, 0, st>>>(d_sizes, world, rank, d_bases, d_offs, n);
    return cudaGetLastError();
}
cudaError_t launch_shard_push(const void* d_src, int n, int kind, unsigned long long* const* d_peers, int world, int cap, int rank,
                              unsigned long long* d_epoch, cudaStream_t st)
{
    k_shard_push<<<1, 128, 0, st>>>(static_cast<const unsigned long long*>(d_src), n, kind, d_peers, world, cap, rank, d_epoch);
    return cudaGetLastError();
}
cudaError_t launch_rms_header_mailbox(unsigned long long* d_mailbox, int world, int cap, const unsigned long long* d_epoch, const int* d_cnt,
                                      double* d_sumsq, void* d_header, const Shape& s, float scale, int use_local, long ncells_total,
                                      float* d_mulfac_out, int* d_status, cudaStream_t st)
{
    ShardMailbox mb; mb.base = d_mailbox; mb.world = world; mb.cap = cap;
    k_rms_header_mailbox<<<1, 64, 0, st>>>(mb, d_epoch, d_cnt, d_sumsq, reinterpret_cast<int32_t*>(d_header), s.nx, s.ny, s.nz, s.bx, s.by, s.bz,
                                           scale, use_local, ncells_total, d_mulfac_out, d_status);
    return cudaGetLastError();
}
cudaError_t launch_shard_fixup_mailbox(unsigned long long* d_mailbox, int world, int cap, const unsigned long long* d_epoch, int rank,
                                       long* d_bases, unsigned long long* d_offs, long n, int* d_status, cudaStream_t st)
{
    ShardMailbox mb; mb.base = d_mailbox; mb.world = world; mb.cap = cap;
    const long nn = n > 0 ? n : 1;
    k_shard_fixup_mailbox<<<(unsigned)((nn + 255) / 256), 256, 0, st>>>(mb, d_epoch, rank, d_bases, d_offs, n, d_status);
    return cudaGetLastError();
}
size_t shard_mailbox_bytes(int world, int cap) { return sizeof(unsigned long long) * 4 * ((size_t)world * cap + world); }

cudaError_t launch_shard_extent(const long* d_bases, int rank, long* d_out2, cudaStream_t st)
{
    k_shard_extent<<<1, 32, 0, st>>>(d_bases, rank, d_out2);
    return cudaGetLastError();
}

cudaError_t launch_offsets_add(unsigned long long* d_offs, long n, long base, cudaStream_t st)
{
    if (n <= 0) return cudaSuccess;
    k_offsets_add<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(d_offs, n, (unsigned long long)base);
    return cudaGetLastError();
}

// launch a fused kernel instance with `smem` dynamic bytes on a grid of min(nwork, SMs * occupancy)
template <class K, class... Args>
static cudaError_t launch_persistent(K kernel, int T, size_t smem, long nwork, int sm_count, cudaStream_t st, Args... args)
{
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int occ = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, T, smem);
    if (e != cudaSuccess) return e;
    if (occ < 1) return cudaErrorLaunchOutOfResources;
    long grid = (long)sm_count * occ;
    if (grid > nwork) grid = nwork;
    if (grid < 1) return cudaSuccess;
    kernel<<<(unsigned)grid, T, smem, st>>>(args...);
    return cudaGetLastError();
}

// kernel configuration by block shape: a 64-long line needs 255 registers in EXACT arithmetic (the level copy; T = 256,
// one CTA per SM) but 64 + addressing in LIFTING (in place: 128 registers, two CTAs per SM); blocks of
// >= 16384 cells use 512 threads (2 lines per thread and pass at 32^3); small blocks 128 threads
#define CVX_DISPATCH_FUSED(CALL)                                         \
    do {                                                                 \
        const int maxn = s.bx > s.by ? (s.bx > s.bz ? s.bx : s.bz) : (s.by > s.bz ? s.by : s.bz); \
        const long cells = (long)s.bx * s.by * s.bz;                     \
        if (maxn == 64) { if (exact) CALL(64, true, 256, 1); else CALL(64, false, 256, 2); }            \
        else if (cells >= 16384) { if (exact) CALL(32, true, 512, 1); else CALL(32, false, 512, 1); }   \
        else { if (exact) CALL(32, true, 128, 4); else CALL(32, false, 128, 4); }                        \
    } while (0)

// developer switch between the warp-specialised 32^3 kernels (default) and the fused ones they were derived from:
// environment CVX_GPU_NO_WS at load, or cvx_debug_set_ws() (tests compare the two bit for bit)
static int g_ws = -1;
static bool ws_enabled()
{
    if (g_ws < 0) g_ws = getenv("CVX_GPU_NO_WS") ? 0 : 1;
    return g_ws != 0;
}
extern "C" void cvx_debug_set_ws(int on) { g_ws = on ? 1 : 0; }

static void rle_segments(const BlkShape& b, int& nseg, int& gps)
{
    int V = b.B / SEG_CAP;
    if (V < b.bx) V = b.bx;
    nseg = b.B / V;
    gps = V / 8;
}

// blocks per batch of the large-block path: two blocks per SM (the parser still takes one CTA per block), between 256 MiB and
// 2 GiB of scratch -- at 256^3 (64 MiB per block) a 256 MiB batch is four blocks, i.e. four busy SMs per launch
static long stage_batch_blocks(size_t blk_bytes, int sm_count)
{
    size_t bytes = blk_bytes * (size_t)(2 * sm_count);
    if (bytes < (256ull << 20)) bytes = 256ull << 20;
    if (bytes > (2048ull << 20)) bytes = 2048ull << 20;
    return (long)(bytes / blk_bytes);
}

cudaError_t launch_encode(cvxgpu_workspace* ws, const float* d_vol, const Shape& s, float scale, int use_local,
                          const float* d_glob_mulfac, unsigned long long* d_offs, float* d_blkmulfac,
                          void* d_payload, long cap, long* d_payload_bytes, int* d_err, int exact, cudaStream_t st)
{
    cudaError_t e = ws_reserve_status(ws, s.nblk);
    if (e != cudaSuccess) return e;
    // look-back words, ticket, the caller's status word and byte count: cleared by one launch
    k_encode_prep<<<(unsigned)((s.nblk + 255) / 256), 256, 0, st>>>(ws->d_status, s.nblk, ws->d_ticket, d_err, d_payload_bytes);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    EncParams p;
    p.vol = d_vol; p.g = vol_view(s, d_vol); p.s = blk_shape(s); p.nbx = s.nbx; p.nby = s.nby; p.nblk = s.nblk;
    p.scale = scale; p.use_local = use_local; p.glob_mulfac = d_glob_mulfac; p.offs = d_offs; p.blkmulfac = d_blkmulfac;
    p.payload = reinterpret_cast<uint8_t*>(d_payload); p.cap = cap; p.status = ws->d_status; p.ticket = ws->d_ticket;
    p.payload_bytes = d_payload_bytes; p.err = d_err;
    rle_segments(p.s, p.nseg, p.gps);
    if (s.bx == 32 && s.by == 32 && s.bz == 32) {
        // the headline shape: TMA-fed persistent kernel with the sparse coder (cvx_fused32.cuh)
        CUtensorMap tmap;
        const int use_tma = make_volume_tmap(&tmap, d_vol, s.nx, s.ny, s.nz, 32, 32, 8) ? 1 : 0;
        long grid = ws->sm_count < s.nblk ? ws->sm_count : s.nblk;
        if (!exact && use_tma && ws_enabled()) {
            // LIFTING: warp-specialised kernel, the coder of block t overlaps the transform of block t+1
            e = cudaFuncSetAttribute(k_encode32ws, cudaFuncAttributeMaxDynamicSharedMemorySize, E32_SMEM);
            if (e != cudaSuccess) return e;
            k_encode32ws<<<(unsigned)grid, E32_T, E32_SMEM, st>>>(tmap, p);
            return cudaGetLastError();
        }
        auto kern = exact ? k_encode32<true> : k_encode32<false>;
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, F32_SMEM);
        if (e != cudaSuccess) return e;
        kern<<<(unsigned)grid, F32_T, F32_SMEM, st>>>(tmap, p, use_tma);
        return cudaGetLastError();
    }
    if (es_ok(s.bx, s.by, s.bz) && !getenv("CVX_GPU_NO_SPARSE")) {
        // 8192 .. 32768 cells (config 4's 16 x 16 x 64): the generic line passes + the sparse list coder (cvx_encode_sparse.inl)
        const size_t smem = es_smem_bytes(p.s.B, p.nseg);
#define CVX_CALL_ENC(MAXN, EX, T, MINB) \
    e = launch_persistent(k_encode_sparse<MAXN, EX, T, MINB>, T, smem, s.nblk, ws->sm_count, st, p)
        CVX_DISPATCH_FUSED(CVX_CALL_ENC);
#undef CVX_CALL_ENC
        return e;
    }
    if (fused_ok(s.bx, s.by, s.bz)) {
        const size_t smem = (size_t)p.s.B * 4 + sizeof(int4) * (size_t)p.nseg;
#define CVX_CALL_ENC(MAXN, EX, T, MINB) \
    e = launch_persistent(k_encode_fused<MAXN, EX, T, MINB>, T, smem, s.nblk, ws->sm_count, st, p)
        CVX_DISPATCH_FUSED(CVX_CALL_ENC);
#undef CVX_CALL_ENC
        return e;
    }
    // staged: batches of blocks through global scratch
    const size_t blk_bytes = (size_t)p.s.B * 4;
    long per_batch = stage_batch_blocks(blk_bytes, ws->sm_count);
    if (per_batch < 1) per_batch = 1;
    if (per_batch > s.nblk) per_batch = s.nblk;
    // scratch: the batch's dense blocks, then the coder's segment records and per-block multipliers / offsets
    const size_t blocks_bytes = blk_bytes * (size_t)per_batch;
    const size_t seg_bytes = (blocks_bytes / 4 / STG_V) * sizeof(int4);
    const size_t per_blk_bytes = ((size_t)per_batch * 16 + 255) & ~(size_t)255;
    e = ws_reserve_scratch(ws, blocks_bytes + seg_bytes + per_blk_bytes);
    if (e != cudaSuccess) return e;
    const bool one_cta_coder = getenv("CVX_GPU_STAGE_ONE_CTA") != nullptr;          // developer switch: the coder these kernels replace
    for (long b0 = 0; b0 < s.nblk; b0 += per_batch) {
        const int nb = (int)((s.nblk - b0) < per_batch ? (s.nblk - b0) : per_batch);
        BigParams bp;
        bp.scratch = ws->d_scratch; bp.blk_first = b0; bp.nb = nb; bp.s = p.s; bp.nbx = s.nbx; bp.nby = s.nby; bp.g = p.g;
        bp.vol_in = d_vol; bp.vol_out = nullptr;
        e = launch_big_transform(bp, true, exact != 0, ws->sm_count, st);         // gather fused into the x pass
        if (e != cudaSuccess) return e;
        const int ge = nb < ws->sm_count ? nb : ws->sm_count;
        if (one_cta_coder) {
            k_stage_encode<<<ge, STAGE_T, 0, st>>>(p, ws->d_scratch, b0, nb);
        } else {
            StageCoder sc;
            unsigned char* base = reinterpret_cast<unsigned char*>(ws->d_scratch);
            sc.scratch = ws->d_scratch;
            sc.seg = reinterpret_cast<int4*>(base + blocks_bytes);
            sc.blkoff = reinterpret_cast<unsigned long long*>(base + blocks_bytes + seg_bytes);
            sc.mulfac = reinterpret_cast<float*>(base + blocks_bytes + seg_bytes + (size_t)per_batch * 8);
            sc.blk_first = b0; sc.nb = nb;
            const long nsegs = (long)nb * (p.s.B / STG_V);
            long gs = (nsegs + 255) / 256;
            if (gs > (long)ws->sm_count * 16) gs = (long)ws->sm_count * 16;
            k_stage_mulfac<<<ge, STAGE_T, 0, st>>>(p, sc);
            k_stage_measure<<<(unsigned)gs, 256, 0, st>>>(p, sc);
            k_stage_plan<<<ge, STAGE_T, 0, st>>>(p, sc);
            k_stage_emit<<<(unsigned)gs, 256, 0, st>>>(p, sc);
        }
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

cudaError_t launch_decode(cvxgpu_workspace* ws, float* d_vol, const Shape& s, float glob_mulfac, int use_local,
                          const unsigned long long* d_offs, const float* d_blkmulfac, const void* d_payload,
                          long payload_bytes, int exact, cudaStream_t st, const float* d_glob_mulfac, const long* d_payload_bytes,
                          const long* d_offs_bias)
{
    DecParams p;
    p.gm_ptr = d_glob_mulfac; p.pbytes_ptr = d_payload_bytes; p.bias_ptr = d_offs_bias;
    p.vol = d_vol; p.g = vol_view(s, d_vol); p.s = blk_shape(s); p.nbx = s.nbx; p.nby = s.nby; p.nblk = s.nblk;
    p.glob_mulfac = glob_mulfac; p.use_local = use_local; p.offs = d_offs; p.blkmulfac = d_blkmulfac;
    p.payload = reinterpret_cast<const uint8_t*>(d_payload); p.payload_bytes = payload_bytes;
    cudaError_t e = cudaSuccess;
    if (s.bx == 32 && s.by == 32 && s.bz == 32) {
        // the headline shape: prefetched payload windows, z lines in registers, direct row stores (cvx_decode32.inl)
        // 512 threads, two lines per thread.  (A 1024-thread / 64-register / one-line-per-thread instantiation of the same
        // template was measured 12 % slower -- 3.47 vs 3.10 ms: the phases are pipe-throughput bound, not latency bound.)
        long grid = ws->sm_count < s.nblk ? ws->sm_count : s.nblk;
        if (!exact && ws_enabled()) {
            // LIFTING: warp-specialised kernel, the parse of block t+1 overlaps the transform of block t (cvx_decode32ws.inl)
            e = cudaFuncSetAttribute(k_decode32ws, cudaFuncAttributeMaxDynamicSharedMemorySize, W32_SMEM);
            if (e != cudaSuccess) return e;
            k_decode32ws<<<(unsigned)grid, W32_T, W32_SMEM, st>>>(p);
            return cudaGetLastError();
        }
        auto kern = exact ? k_decode32<true, 512> : k_decode32<false, 512>;
        const int threads = 512;
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, D32_SMEM);
        if (e != cudaSuccess) return e;
        kern<<<(unsigned)grid, threads, D32_SMEM, st>>>(p);
        return cudaGetLastError();
    }
    if (es_ok(s.bx, s.by, s.bz) && !getenv("CVX_GPU_NO_SPARSE")) {
        // 8192 .. 32768 cells (config 4's 16 x 16 x 64): the generic line passes + k_decode32's windowed parser (cvx_decode_win.inl)
#define CVX_CALL_DEC(MAXN, EX, T, MINB)                                                   \
    do {                                                                                  \
        const size_t smem = (size_t)p.s.B * 4 + 2 * (size_t)(2 * T * 8 + 32) + (size_t)(2 * T) * 4; \
        e = launch_persistent(k_decode_win<MAXN, EX, T, MINB>, T, smem, s.nblk, ws->sm_count, st, p); \
    } while (0)
        CVX_DISPATCH_FUSED(CVX_CALL_DEC);
#undef CVX_CALL_DEC
        return e;
    }
    if (fused_ok(s.bx, s.by, s.bz)) {
#define CVX_CALL_DEC(MAXN, EX, T, MINB)                                                   \
    do {                                                                                  \
        const size_t smem = (size_t)p.s.B * 4 + (size_t)(T + 2) * 16 + (size_t)T * 4;     \
        e = launch_persistent(k_decode_fused<MAXN, EX, T, MINB>, T, smem, s.nblk, ws->sm_count, st, p); \
    } while (0)
        CVX_DISPATCH_FUSED(CVX_CALL_DEC);
#undef CVX_CALL_DEC
        return e;
    }
    const size_t blk_bytes = (size_t)p.s.B * 4;
    long per_batch = stage_batch_blocks(blk_bytes, ws->sm_count);
    if (per_batch < 1) per_batch = 1;
    if (per_batch > s.nblk) per_batch = s.nblk;
    e = ws_reserve_scratch(ws, blk_bytes * (size_t)per_batch);
    if (e != cudaSuccess) return e;
    for (long b0 = 0; b0 < s.nblk; b0 += per_batch) {
        const int nb = (int)((s.nblk - b0) < per_batch ? (s.nblk - b0) : per_batch);
        const int gd = nb < ws->sm_count * 2 ? nb : ws->sm_count * 2;
        k_stage_decode<<<gd, STAGE_T, 0, st>>>(p, ws->d_scratch, b0, nb);
        BigParams bp;
        bp.scratch = ws->d_scratch; bp.blk_first = b0; bp.nb = nb; bp.s = p.s; bp.nbx = s.nbx; bp.nby = s.nby; bp.g = p.g;
        bp.vol_in = nullptr; bp.vol_out = d_vol;
        e = launch_big_transform(bp, false, exact != 0, ws->sm_count, st);        // scatter fused into the last pass
        if (e != cudaSuccess) return e;
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

cudaError_t launch_wavelet(cvxgpu_workspace* ws, float* d_blocks, long nblocks, int bx, int by, int bz, int inverse,
                           int exact, cudaStream_t st)
{
    Shape s = make_shape(bx, by, bz, bx, by, bz);
    BlkShape b = blk_shape(s);
    cudaError_t e = cudaSuccess;
    if (nblocks <= 0) return e;
    if (fused_ok(bx, by, bz)) {
        const size_t smem = (size_t)b.B * 4;
#define CVX_CALL_WAV(MAXN, EX, T, MINB)                                                                             \
    do {                                                                                                            \
        if (inverse) e = launch_persistent(k_wavelet_fused<MAXN, EX, false, T, MINB>, T, smem, nblocks, ws->sm_count, st, d_blocks, nblocks, b); \
        else e = launch_persistent(k_wavelet_fused<MAXN, EX, true, T, MINB>, T, smem, nblocks, ws->sm_count, st, d_blocks, nblocks, b);          \
    } while (0)
        CVX_DISPATCH_FUSED(CVX_CALL_WAV);
#undef CVX_CALL_WAV
        return e;
    }
    // large blocks: the three line passes in place on the caller's dense blocks, in batches that keep the int block count small
    const long per_batch = 1 << 20;
    for (long b0 = 0; b0 < nblocks; b0 += per_batch) {
        const int nb = (int)((nblocks - b0) < per_batch ? (nblocks - b0) : per_batch);
        BigParams bp;
        bp.scratch = d_blocks + b0 * b.B; bp.blk_first = 0; bp.nb = nb; bp.s = b; bp.nbx = 1; bp.nby = 1;
        bp.g.nx = bx; bp.g.ny = by; bp.g.nz = bz; bp.g.vec_ok = 1;
        bp.vol_in = nullptr; bp.vol_out = nullptr;
        e = launch_big_transform(bp, !inverse, exact != 0, ws->sm_count, st);
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}

}  // namespace cvx

// ---- cvxgpu_sumsq_rows_static_device: the per-block-row reduction for callers that may not allocate ---------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, serving loops, harnesses that forbid it).  k_sumsq_rows needs SUMSQ_ROW_CTAS fp64
// partials per block row as scratch; cvxgpu_sumsq_rows_device keeps them in a workspace that grows with cudaMalloc.  This
// variant keeps them in a file-scope device array of fixed size instead: CVXGPU_STATIC_MAX_ROWS (4096) block rows x 128
// partials x 8 bytes = 4 MiB, which every GPU that loads the library holds until the process exits.
// In host code `g_static_rowpartial` is only a stand-in for the symbol, NOT a device address (a kernel given it would fault
// or silently use host memory); the device address is looked up once per device, with that device current, by
// cudaGetSymbolAddress below.  All calls on a device share the one array: two calls that overlap on the GPU (different
// streams, no ordering between them) race on it.  Nothing in it has to start at zero -- every (row, CTA) slot a call's
// finish kernel reads was written by that call's k_sumsq_rows, a zero where a CTA had no elements -- so no clearing pass
// is needed, and what the previous call left behind is never read.
int cvx_internal_fail(int code, const char* fmt, ...);          // cvx_api.cpp: records the message cvxgpu_last_error() returns

static __device__ __align__(256) double g_static_rowpartial[(size_t)CVXGPU_STATIC_MAX_ROWS * cvx::SUMSQ_ROW_CTAS];

namespace {
constexpr int kStaticMaxDevices = 64;
double* g_static_rowpartial_addr[kStaticMaxDevices];          // device address of the array, per device ordinal (0 = not looked up)

// the array's address on the CURRENT device; no allocation (the array is part of the loaded module)
int static_rowpartial_address(double** out)
{
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return cvx_internal_fail(CVXGPU_ERR_CUDA, "cvxgpu_sumsq_rows_static_device: cudaGetDevice: %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= kStaticMaxDevices)
        return cvx_internal_fail(CVXGPU_ERR_ARG, "cvxgpu_sumsq_rows_static_device: device ordinal %d outside 0..%d", dev, kStaticMaxDevices - 1);
    double* p = __atomic_load_n(&g_static_rowpartial_addr[dev], __ATOMIC_ACQUIRE);
    if (!p) {
        void* sym = nullptr;
        e = cudaGetSymbolAddress(&sym, g_static_rowpartial);          // device `dev` is current: its copy of the array
        if (e != cudaSuccess || !sym) {
            cudaGetLastError();
            return cvx_internal_fail(CVXGPU_ERR_CUDA, "cvxgpu_sumsq_rows_static_device: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
        }
        p = static_cast<double*>(sym);
        __atomic_store_n(&g_static_rowpartial_addr[dev], p, __ATOMIC_RELEASE);   // racing threads store the same value
    }
    *out = p;
    return CVXGPU_OK;
}
}  // namespace

extern "C" int cvxgpu_sumsq_rows_static_device(const float* d_vol, long plane_cells, int bz, int nz, double* d_rowsums, void* cuda_stream)
{
    if (!d_vol || !d_rowsums || plane_cells <= 0 || bz <= 0 || nz <= 0)
        return cvx_internal_fail(CVXGPU_ERR_ARG, "cvxgpu_sumsq_rows_static_device: bad argument");
    const long rows = ((long)nz + bz - 1) / bz;
    if (rows > CVXGPU_STATIC_MAX_ROWS)
        return cvx_internal_fail(CVXGPU_ERR_CAPACITY, "cvxgpu_sumsq_rows_static_device: %ld block rows (nz = %d, bz = %d) exceed the static scratch of "
                                 "%d rows; use cvxgpu_sumsq_rows_device", rows, nz, bz, CVXGPU_STATIC_MAX_ROWS);
    double* d_partial = nullptr;
    const int rc = static_rowpartial_address(&d_partial);
    if (rc != CVXGPU_OK) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    cvx::k_sumsq_rows<<<dim3(cvx::SUMSQ_ROW_CTAS, (unsigned)rows), cvx::SUMSQ_T, 0, st>>>(d_vol, plane_cells, bz, nz, d_partial);
    cvx::k_sumsq_rows_finish<<<(unsigned)rows, 32, 0, st>>>(d_partial, d_rowsums);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cvx_internal_fail(CVXGPU_ERR_CUDA, "cvxgpu_sumsq_rows_static_device: launch: %s", cudaGetErrorString(e));
    return CVXGPU_OK;
}

// developer aid (make TIMING=1): per-phase cycle totals stamped by thread 0 of every CTA; kind 0 = encode, 1 = decode
extern "C" int cvx_debug_phases(unsigned long long* out, int kind, int reset)
{
#ifdef CVX_PHASE_TIMING
    unsigned long long h[2][16];
    if (cudaMemcpyFromSymbol(h, cvx::g_phase, sizeof(h)) != cudaSuccess) return -1;
    for (int i = 0; i < 16; ++i) out[i] = h[kind & 1][i];
    if (reset) { memset(h, 0, sizeof(h)); cudaMemcpyToSymbol(cvx::g_phase, h, sizeof(h)); }
    return 16;
#else
    (void)out; (void)kind; (void)reset;
    return 0;
#endif
}


// cvx_device.cuh -- device-side building blocks shared by the fused (block-in-shared-memory)
// kernels and the staged (block-in-global-scratch) kernels: shared-memory layout, CTA scans,
// the decoupled look-back that places each block's payload, the CTA-wide RLE encoder and the
// CTA-wide speculative RLE decoder.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "cvx_wavelet.cuh"
#include "cvx_rle.cuh"

namespace cvx {

// ------------------------------------------------------------------------------------------
// Shared-memory layout of a dense block (x fastest): the linear float index i is stored at
//     phys(i) = i ^ (((i >> SH) & 7) << 2),   SH = max(5, log2(BX))
// i.e. the 16-byte chunk index inside each 128-byte line is XORed with the row number.  This
// is the TMA SWIZZLE_128B pattern for BX = 32, so a 3-D TMA box lands directly in this layout.
// Why: a thread that owns a whole x-row reads it with LDS.128; 8 consecutive rows then hit 8
// different chunks (conflict-free), and a warp that walks 32 consecutive x of one row (y/z
// passes) still touches 32 different banks.
// ------------------------------------------------------------------------------------------
template <int BX>
struct Swz {
    static constexpr int SH = BX >= 64 ? (BX >= 256 ? 8 : BX >= 128 ? 7 : 6) : 5;
    __device__ __forceinline__ static int phys(int i) { return i ^ (((i >> SH) & 7) << 2); }
};
struct NoSwz {
    __device__ __forceinline__ static long phys(long i) { return i; }
};

__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }
__device__ __forceinline__ int warp_id() { return threadIdx.x >> 5; }

// barrier policies: the whole CTA, or a named barrier over a subset of its warps (warp-specialised kernel)
struct CtaBar {
    __device__ __forceinline__ static void sync() { __syncthreads(); }
    __device__ __forceinline__ static int sync_or(int pred) { return __syncthreads_or(pred); }
};
template <int ID, int N>
struct NamedBar {
    __device__ __forceinline__ static void sync() { asm volatile("bar.sync %0, %1;" ::"n"(ID), "n"(N) : "memory"); }
    __device__ __forceinline__ static int sync_or(int pred)
    {
        int r;
        asm volatile("{\n.reg .pred p, q;\nsetp.ne.s32 q, %1, 0;\nbar.red.or.pred p, %2, %3, q;\nselp.s32 %0, 1, 0, p;\n}"
                     : "=r"(r) : "r"(pred), "n"(ID), "n"(N) : "memory");
        return r;
    }
};

// ---- CTA scans ---------------------------------------------------------------------------
// T threads of a thread group that starts at a multiple of T (the whole CTA, or one side of a warp-specialised
// kernel); Bar is the group's barrier.
// exclusive sum over the CTA; `total` receives the CTA-wide sum.  scratch: >= 33 ints.
template <int T, class Bar = CtaBar>
__device__ __forceinline__ int cta_excl_sum(int v, int* scratch, int& total)
{
    constexpr int NW = (T + 31) / 32;
    const int gw = warp_id() % NW;
    int x = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int y = __shfl_up_sync(0xffffffffu, x, d);
        if (lane_id() >= d) x += y;
    }
    if (lane_id() == 31) scratch[gw] = x;
    Bar::sync();
    if (gw == 0) {
        int w = lane_id() < NW ? scratch[lane_id()] : 0;
        int s = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int y = __shfl_up_sync(0xffffffffu, s, d);
            if (lane_id() >= d) s += y;
        }
        if (lane_id() < NW) scratch[lane_id()] = s - w;
        if (lane_id() == 31) scratch[32] = s;
    }
    Bar::sync();
    int r = scratch[gw] + x - v;
    total = scratch[32];
    Bar::sync();
    return r;
}

// saturating variant for value counts that may come from garbage bytes (decoder)
__device__ __forceinline__ int sat_add(int a, int b)
{
    long s = (long)a + b;
    return s > (1 << 30) ? (1 << 30) : (int)s;
}
template <int T>
__device__ __forceinline__ int cta_excl_sum_sat(int v, int* scratch, int& total)
{
    constexpr int NW = (T + 31) / 32;
    int x = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int y = __shfl_up_sync(0xffffffffu, x, d);
        if (lane_id() >= d) x = sat_add(x, y);
    }
    if (lane_id() == 31) scratch[warp_id()] = x;
    __syncthreads();
    if (warp_id() == 0) {
        int w = lane_id() < NW ? scratch[lane_id()] : 0;
        int s = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int y = __shfl_up_sync(0xffffffffu, s, d);
            if (lane_id() >= d) s = sat_add(s, y);
        }
        int ex = __shfl_up_sync(0xffffffffu, s, 1);
        if (lane_id() == 0) ex = 0;
        if (lane_id() < NW) scratch[lane_id()] = ex;
        if (lane_id() == 31) scratch[32] = s;
    }
    __syncthreads();
    int ex_in_warp = __shfl_up_sync(0xffffffffu, x, 1);
    if (lane_id() == 0) ex_in_warp = 0;
    int r = sat_add(scratch[warp_id()], ex_in_warp);
    total = scratch[32];
    __syncthreads();
    return r;
}

// zero-run carry monoid: (has, run).  has = some emitter inside; run = zeros after the last
// emitter (or the whole length when !has).  combine(a, b) = b.has ? b : (a.has, a.run + b.run)
struct Carry {
    int has, run;
};
__device__ __forceinline__ Carry carry_combine(Carry a, Carry b)
{
    Carry r;
    r.has = a.has | b.has;
    r.run = b.has ? b.run : a.run + b.run;
    return r;
}
// exclusive scan of Carry over the CTA; total = combine of everything.  scratch: >= 66 ints.
template <int T>
__device__ __forceinline__ Carry cta_excl_carry(Carry v, int* scratch, Carry& total)
{
    constexpr int NW = (T + 31) / 32;
    Carry x = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        Carry y;
        y.has = __shfl_up_sync(0xffffffffu, x.has, d);
        y.run = __shfl_up_sync(0xffffffffu, x.run, d);
        if (lane_id() >= d) x = carry_combine(y, x);
    }
    if (lane_id() == 31) { scratch[2 * warp_id()] = x.has; scratch[2 * warp_id() + 1] = x.run; }
    // exclusive within the warp
    Carry e;
    e.has = __shfl_up_sync(0xffffffffu, x.has, 1);
    e.run = __shfl_up_sync(0xffffffffu, x.run, 1);
    if (lane_id() == 0) { e.has = 0; e.run = 0; }
    __syncthreads();
    if (threadIdx.x == 0) {
        Carry acc; acc.has = 0; acc.run = 0;
        for (int w = 0; w < NW; ++w) {
            Carry c; c.has = scratch[2 * w]; c.run = scratch[2 * w + 1];
            scratch[2 * w] = acc.has; scratch[2 * w + 1] = acc.run;
            acc = carry_combine(acc, c);
        }
        scratch[64] = acc.has; scratch[65] = acc.run;
    }
    __syncthreads();
    Carry wp; wp.has = scratch[2 * warp_id()]; wp.run = scratch[2 * warp_id() + 1];
    total.has = scratch[64]; total.run = scratch[65];
    Carry r = carry_combine(wp, e);
    __syncthreads();
    return r;
}

// ---- decoupled look-back over blocks in ticket order -----------------------------------------
// status word: bits 63..62 = flag (0 invalid, 1 aggregate, 2 inclusive prefix), low 62 = bytes.
#define CVX_ST_AGG (1ull << 62)
#define CVX_ST_PFX (2ull << 62)
#define CVX_ST_MASK ((1ull << 62) - 1)

// The status word carries its own payload (a byte count) and nothing else is read on the strength of it, so
// relaxed gpu-scope accesses are enough.  (ld.acquire costs an L1 invalidate -- CCTL.IVALL -- per load and
// st.release a MEMBAR.ALL.GPU; measured as 12 % of k_encode32 when the whole CTA looked back with acquire loads.)
__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long* p)
{
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_u64(unsigned long long* p, unsigned long long v)
{
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// called by warp 0 (all 32 lanes).  Publishes `len` for ticket t and returns the exclusive
// prefix (sum of the lengths of tickets < t).  Tickets are handed out by an atomic counter, so
// every smaller ticket is held by a CTA that is already running: no deadlock.
__device__ __forceinline__ unsigned long long lookback_exclusive(unsigned long long* status, long t, unsigned long long len)
{
    if (t == 0) {
        if (lane_id() == 0) st_release_u64(&status[0], CVX_ST_PFX | len);
        return 0ull;
    }
    if (lane_id() == 0) st_release_u64(&status[t], CVX_ST_AGG | len);
    unsigned long long excl = 0;
    long j = t - 1;
    for (;;) {
        long idx = j - lane_id();
        unsigned long long s = CVX_ST_PFX;     // lanes before ticket 0 read as "prefix 0"
        if (idx >= 0) {
            do { s = ld_acquire_u64(&status[idx]); } while ((s >> 62) == 0);
        }
        unsigned pfx = __ballot_sync(0xffffffffu, (s >> 62) == 2);
        int first = pfx ? __ffs(pfx) - 1 : 32;          // nearest predecessor holding a prefix
        unsigned long long v = (lane_id() <= first) ? (s & CVX_ST_MASK) : 0ull;
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
        excl += v;
        if (pfx) break;
        j -= 32;
    }
    if (lane_id() == 0) st_release_u64(&status[t], CVX_ST_PFX | (excl + len));
    return excl;
}

}  // namespace cvx

// cvx_fused32.cuh -- the 32x32x32-block kernels (BASELINE config 1/3/5: the headline shape).
//
// One persistent CTA of 512 threads per SM; a block (128 KiB fp32) lives in shared memory in the
// TMA SWIZZLE_128B layout (cvx_device.cuh) and moves as follows.
//
//   k_encode32:  TMA 3-D box load (OOB zero fill == Copy_To_Block's padding, Block_Copy.cpp:92-115)
//                -> x lines, y lines in shared memory, z lines in REGISTERS (one thread = one line,
//                all levels, Wavelet_Transform_Fast.cpp:24-145)
//                -> the block buffer is free again: the TMA load of the CTA's NEXT block is issued
//                   and overlaps everything below
//                -> sparse run-length coder: only the non-zero quantised coefficients (typically a
//                   few % of the block) are compacted, in raster order, into a list; token sizes,
//                   zero-run lengths and byte offsets are computed per list entry
//                   (Run_Length_Encode_Slow.cpp:189-294 semantics, byte-identical)
//                -> bytes staged in shared memory, block offset by decoupled look-back
//                   (CvxCompress.cpp:365-390), coalesced copy-out.
//                Blocks whose non-zero count exceeds the list (dense data, ratio < ~3) fall back to
//                the dense CTA-wide coder of cvx_codec.cuh on the same shared-memory block.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include "cvx_wavelet.cuh"
#include "cvx_rle.cuh"
#include "cvx_device.cuh"
#include "cvx_codec.cuh"

namespace cvx {

constexpr int F32_T = 512;                    // threads per CTA
constexpr int F32_B = 32768;                  // cells per block
constexpr int F32_ROWS = 1024;                // rows of 32 raster-consecutive cells
constexpr int F32_LCAP = 4096;                // sparse list capacity (entries of 8 bytes): 12.5 % non-zeros
constexpr int F32_STAGE = 24576;              // payload staging bytes, two buffers (deferred copy-out)
constexpr int F32_OFF_NZM = F32_B * 4;                        // uint32[1024]  non-zero mask per row
constexpr int F32_OFF_BASE = F32_OFF_NZM + F32_ROWS * 4;      // int32[1024]   list position of the row's first entry
constexpr int F32_OFF_LIST = F32_OFF_BASE + F32_ROWS * 4;     // uint2[LCAP]   (raster index, coefficient bits)
constexpr int F32_OFF_STAGE = F32_OFF_LIST + F32_LCAP * 8;    // bytes
constexpr int F32_SMEM = F32_OFF_STAGE + 2 * F32_STAGE;       // 221184

// ---- mbarrier / TMA (sm_90+ PTX) ---------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void fence_proxy_async()
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
// 3-D tiled box load: coordinates in elements, innermost first
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2)
{
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}

// ---- y and z lines of the 32^3 kernels.  The LIFTING encoders (passes z, y, x) run the bare lifting steps here: the
//      lines' scalings ride on the x rows' multiply (cvx_wavelet.cuh, "the scalings of all three axes in ONE multiply").
//      Everything else -- EXACT, and the LIFTING decoders -- is the plain line transform. ----
template <bool EXACT, bool FWD, class V>
__device__ __forceinline__ void f32_line_yz(V* v)
{
    if constexpr (EXACT) { if (FWD) fwd_line<32, EXACT>(v); else inv_line<32, EXACT>(v); }
    else { if (FWD) fwd_lift_levels<32, 1>(v); else inv_line<32, EXACT>(v); }
}
// factors of the thread's two x rows (row tid: y = lane, z = warp; row tid + 512: z = warp + 16) of a 512-thread group
__device__ __forceinline__ void f32_row_factors(float& fa, float& fb)
{
    const int y = threadIdx.x & 31, z = (threadIdx.x >> 5) & 15;
    const float sy = lift_scale_of_sub<32, true>(y);
    fa = __fmul_rn(sy, lift_scale_of_sub<32, true>(z));
    fb = __fmul_rn(sy, lift_scale_of_sub<32, true>(z + 16));
}

// ---- line passes on the swizzled 32^3 block: addresses reduced to a base register + immediates ----
// element (x, y, z) lives at float index  z*1024 + y*32 + (x ^ ((y & 7) << 2))

// x lines: the thread's row is 8 chunks of 16 bytes; logical chunk c sits at physical chunk c ^ (row & 7)
// `skew`: the two halves of a 512-thread group enter the pass one after the other (named barrier 7: the second half
// starts its loads when the first half's are issued), so that one half's shared-memory phase overlaps the other's math
template <bool EXACT, bool FWD>
__device__ __forceinline__ void f32_pass_x(float* sm, const bool skew = false)
{
    // both rows of the thread (r = tid and tid + 512) in one straight-line block: the two independent dependency
    // chains interleave and fill the FMA pipe where one row's deep (short) levels would leave bubbles
    float4* row0 = reinterpret_cast<float4*>(sm + threadIdx.x * 32);
    float4* row1 = row0 + (F32_T * 32) / 4;
    const int k = threadIdx.x & 7;                 // (tid + 512) & 7 == tid & 7: same swizzle for both rows
    float v[2][32];
    if (skew && (threadIdx.x & 256)) asm volatile("bar.sync 7, 512;" ::: "memory");
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        float4* row = h ? row1 : row0;
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            float4 q = row[c ^ k];
            if (FWD) {
                v[h][4 * c] = q.x; v[h][4 * c + 1] = q.y; v[h][4 * c + 2] = q.z; v[h][4 * c + 3] = q.w;
            } else {
                v[h][sub_to_reg<32, EXACT>(4 * c)] = q.x; v[h][sub_to_reg<32, EXACT>(4 * c + 1)] = q.y;
                v[h][sub_to_reg<32, EXACT>(4 * c + 2)] = q.z; v[h][sub_to_reg<32, EXACT>(4 * c + 3)] = q.w;
            }
        }
    }
    if (skew && !(threadIdx.x & 256)) asm volatile("bar.arrive 7, 512;" ::: "memory");
    if constexpr (EXACT) {
        if (FWD) { fwd_line<32, EXACT>(v[0]); fwd_line<32, EXACT>(v[1]); }
        else { inv_line<32, EXACT>(v[0]); inv_line<32, EXACT>(v[1]); }
    } else {
        // LIFTING: the two rows run side by side in packed registers between their first and last scalar operation
        float w[2][32];
        if (FWD) fwd_lift_pair<32>(v[0], v[1], w[0], w[1]); else inv_lift_pair<32>(v[0], v[1], w[0], w[1]);
#pragma unroll
        for (int i = 0; i < 32; ++i) { v[0][i] = w[0][i]; v[1][i] = w[1][i]; }
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        float4* row = h ? row1 : row0;
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            if (FWD)
                row[c ^ k] = make_float4(v[h][sub_to_reg<32, EXACT>(4 * c)], v[h][sub_to_reg<32, EXACT>(4 * c + 1)],
                                         v[h][sub_to_reg<32, EXACT>(4 * c + 2)], v[h][sub_to_reg<32, EXACT>(4 * c + 3)]);
            else
                row[c ^ k] = make_float4(v[h][4 * c], v[h][4 * c + 1], v[h][4 * c + 2], v[h][4 * c + 3]);
        }
    }
}

// y lines: thread (x = lane, z); element j at z*1024 + j*32 + (x ^ ((j & 7) << 2)): eight base
// pointers (one per j & 7) + immediate offsets
template <bool EXACT, bool FWD>
__device__ __forceinline__ void f32_pass_y(float* sm)
{
#pragma unroll 1
    for (int l = threadIdx.x; l < F32_ROWS; l += F32_T) {
        const int x = l & 31, z = l >> 5;
        float* a[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] = sm + z * 1024 + (x ^ (k << 2));
        float v[32];
        if (FWD) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = a[j & 7][j * 32];
            fwd_line<32, EXACT>(v);
#pragma unroll
            for (int r = 0; r < 32; ++r) { const int s = reg_to_sub<32, EXACT>(r); a[s & 7][s * 32] = v[r]; }
        } else {
#pragma unroll
            for (int r = 0; r < 32; ++r) { const int s = reg_to_sub<32, EXACT>(r); v[r] = a[s & 7][s * 32]; }
            inv_line<32, EXACT>(v);
#pragma unroll
            for (int j = 0; j < 32; ++j) a[j & 7][j * 32] = v[j];
        }
    }
}

// y lines, two per thread side by side in packed registers: lines (x = lane, z = warp) and (x, z + 16) are
// 16384 floats apart with the same swizzle, so both come from the same eight base pointers (FADD2/FFMA2 math)
// `skew` as in f32_pass_x: the second half of a 512-thread group starts its loads when the first half's are issued
template <bool EXACT, bool FWD>
__device__ __forceinline__ void f32_pass_y2(float* sm, const bool skew = false)
{
    const int x = threadIdx.x & 31, z = threadIdx.x >> 5;
    float* a[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) a[k] = sm + z * 1024 + (x ^ (k << 2));
    typename PairOf<EXACT>::type v[32];
    if (skew && (threadIdx.x & 256)) asm volatile("bar.sync 7, 512;" ::: "memory");
    if (FWD) {
#pragma unroll
        for (int j = 0; j < 32; ++j) p_set(v[j], a[j & 7][j * 32], a[j & 7][j * 32 + 16384]);
        if (skew && !(threadIdx.x & 256)) asm volatile("bar.arrive 7, 512;" ::: "memory");
        f32_line_yz<EXACT, true>(v);
#pragma unroll
        for (int r = 0; r < 32; ++r) {
            const int s = reg_to_sub<32, EXACT>(r);
            a[s & 7][s * 32] = p_lo(v[r]);
            a[s & 7][s * 32 + 16384] = p_hi(v[r]);
        }
    } else {
#pragma unroll
        for (int r = 0; r < 32; ++r) {
            const int s = reg_to_sub<32, EXACT>(r);
            p_set(v[r], a[s & 7][s * 32], a[s & 7][s * 32 + 16384]);
        }
        if (skew && !(threadIdx.x & 256)) asm volatile("bar.arrive 7, 512;" ::: "memory");
        f32_line_yz<EXACT, false>(v);
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            a[j & 7][j * 32] = p_lo(v[j]);
            a[j & 7][j * 32 + 16384] = p_hi(v[j]);
        }
    }
}

// ---- thread-count generic passes: T = 512 (two lines per thread) or 1024 (one line per thread, 64-register budget,
//      twice the warps to hide latencies; LIFTING only -- the EXACT level copy does not fit 64 registers) ----
template <bool EXACT, bool FWD, int T>
__device__ __forceinline__ void f32_pass_xT(float* sm, const bool skew = false)
{
    if constexpr (T == 512) {
        f32_pass_x<EXACT, FWD>(sm, skew);
    } else {
        float4* row = reinterpret_cast<float4*>(sm + threadIdx.x * 32);
        const int k = threadIdx.x & 7;
        float v[32];
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            float4 q = row[c ^ k];
            if (FWD) {
                v[4 * c] = q.x; v[4 * c + 1] = q.y; v[4 * c + 2] = q.z; v[4 * c + 3] = q.w;
            } else {
                v[sub_to_reg<32, EXACT>(4 * c)] = q.x; v[sub_to_reg<32, EXACT>(4 * c + 1)] = q.y;
                v[sub_to_reg<32, EXACT>(4 * c + 2)] = q.z; v[sub_to_reg<32, EXACT>(4 * c + 3)] = q.w;
            }
        }
        if (FWD) fwd_line<32, EXACT>(v); else inv_line<32, EXACT>(v);
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            if (FWD)
                row[c ^ k] = make_float4(v[sub_to_reg<32, EXACT>(4 * c)], v[sub_to_reg<32, EXACT>(4 * c + 1)],
                                         v[sub_to_reg<32, EXACT>(4 * c + 2)], v[sub_to_reg<32, EXACT>(4 * c + 3)]);
            else
                row[c ^ k] = make_float4(v[4 * c], v[4 * c + 1], v[4 * c + 2], v[4 * c + 3]);
        }
    }
}
template <bool EXACT, bool FWD, int T>
__device__ __forceinline__ void f32_pass_yT(float* sm, const bool skew = false)
{
    if constexpr (T == 512) {
        f32_pass_y2<EXACT, FWD>(sm, skew);
    } else {
        const int x = threadIdx.x & 31, z = threadIdx.x >> 5;
        float* a[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] = sm + z * 1024 + (x ^ (k << 2));
        float v[32];
        if (FWD) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = a[j & 7][j * 32];
            fwd_line<32, EXACT>(v);
#pragma unroll
            for (int r = 0; r < 32; ++r) { const int s = reg_to_sub<32, EXACT>(r); a[s & 7][s * 32] = v[r]; }
        } else {
#pragma unroll
            for (int r = 0; r < 32; ++r) { const int s = reg_to_sub<32, EXACT>(r); v[r] = a[s & 7][s * 32]; }
            inv_line<32, EXACT>(v);
#pragma unroll
            for (int j = 0; j < 32; ++j) a[j & 7][j * 32] = v[j];
        }
    }
}

// z line of thread-line l (x = l & 31, y = l >> 5): element j at j*1024 + y*32 + (x ^ ((y & 7) << 2))
__device__ __forceinline__ float* f32_zline(float* sm, int l)
{
    const int x = l & 31, y = l >> 5;
    return sm + y * 32 + (x ^ ((y & 7) << 2));
}


// ---- LIFTING encoders: passes in the order z, y, x, the x rows staying in registers ------------------------------------
// The three 1-D transforms commute up to rounding, and LIFTING promises a tolerance, not the reference's bits (EXACT keeps
// the reference's x, y, z order).  With x LAST a thread ends up holding two complete raster rows of coefficients: it counts
// and emits their non-zeros straight from registers (no write-back of the 96 % zeros, no ballots, no gather of the values
// out of the block buffer), and the block buffer is dead as soon as every thread has LOADED its rows, so the TMA load of
// the next block is issued before the x math and hides under it, the scan and the list emission.

// z lines in place, two per thread in packed registers: lines (x = lane, y = warp) and (x, y = warp + 16), 512 floats apart
// with the same swizzle.  `skew` as in f32_pass_x.
template <bool EXACT>
__device__ __forceinline__ void f32_pass_z2_fwd(float* sm, const bool skew = false)
{
    float* zl = f32_zline(sm, threadIdx.x);
    typename PairOf<EXACT>::type v[32];
    if (skew && (threadIdx.x & 256)) asm volatile("bar.sync 7, 512;" ::: "memory");
#pragma unroll
    for (int j = 0; j < 32; ++j) p_set(v[j], zl[j * 1024], zl[j * 1024 + 512]);
    if (skew && !(threadIdx.x & 256)) asm volatile("bar.arrive 7, 512;" ::: "memory");
    f32_line_yz<EXACT, true>(v);
#pragma unroll
    for (int r = 0; r < 32; ++r) {
        const int s = reg_to_sub<32, EXACT>(r);
        zl[s * 1024] = p_lo(v[r]);
        zl[s * 1024 + 512] = p_hi(v[r]);
    }
}

// the thread's two x rows (rows tid and tid + 512), natural order, into registers
__device__ __forceinline__ void f32_rows_load(const float* sm, float (&v)[2][32])
{
    const float4* row0 = reinterpret_cast<const float4*>(sm + threadIdx.x * 32);
    const int k = threadIdx.x & 7;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const float4* row = row0 + h * ((F32_T * 32) / 4);
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            const float4 q = row[c ^ k];
            v[h][4 * c] = q.x; v[h][4 * c + 1] = q.y; v[h][4 * c + 2] = q.z; v[h][4 * c + 3] = q.w;
        }
    }
}
// ... and back: w[h][sub_to_reg(x)] = coefficient x of row tid + 512 h (what fwd_lift_pair leaves)
__device__ __forceinline__ void f32_rows_store_coef(float* sm, const float (&w)[2][32])
{
    float4* row0 = reinterpret_cast<float4*>(sm + threadIdx.x * 32);
    const int k = threadIdx.x & 7;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        float4* row = row0 + h * ((F32_T * 32) / 4);
#pragma unroll
        for (int c = 0; c < 8; ++c)
            row[c ^ k] = make_float4(w[h][sub_to_reg<32, false>(4 * c)], w[h][sub_to_reg<32, false>(4 * c + 1)],
                                     w[h][sub_to_reg<32, false>(4 * c + 2)], w[h][sub_to_reg<32, false>(4 * c + 3)]);
    }
}
// sum of squares of the block's coefficients from the rows in registers (local RMS, LIFTING): every thread its 64 values
// in register order, warps by shuffle, warp sums in warp order -- the same in every kernel that uses it
template <int T, class Bar>
__device__ __forceinline__ float cta_sumsq_rows(const float (&w)[2][32], float* s_red)
{
    float a = 0.f;
#pragma unroll
    for (int h = 0; h < 2; ++h)
#pragma unroll
        for (int r = 0; r < 32; ++r) a = __fmaf_rn(w[h][r], w[h][r], a);
#pragma unroll
    for (int d = 16; d >= 1; d >>= 1) a += __shfl_xor_sync(0xffffffffu, a, d);
    if ((threadIdx.x & 31) == 0) s_red[(threadIdx.x & (T - 1)) >> 5] = a;
    Bar::sync();
    float t = 0.f;
    for (int i = 0; i < T / 32; ++i) t += s_red[i];
    Bar::sync();
    return t;
}
// appends (idx, c) at shared address lp when c quantises to non-zero, i.e. !(|c| <= T): predicated, no branch
__device__ __forceinline__ void f32_emit_nz(uint32_t& lp, float c, float T, unsigned idx)
{
    asm volatile("{\n\t.reg .pred p;\n\t.reg .f32 a;\n\tabs.f32 a, %1;\n\tsetp.gtu.f32 p, a, %2;\n\t"
                 "@p st.shared.v2.b32 [%0], {%3, %4};\n\t@p add.u32 %0, %0, 8;\n\t}"
                 : "+r"(lp) : "f"(c), "f"(T), "r"(idx), "r"(__float_as_uint(c)) : "memory");
}

// ---- sparse coder pieces ---------------------------------------------------------------------------
// largest T >= 0 with |fl(m * T)| < 1, so that  quantise(fl(m * v)) != 0  <=>  !(|v| <= T)
// (truncation gives 0 exactly when |f| < 1; NaN and out-of-range give 0x80000000, Run_Length_Encode_Slow.cpp:204).
// fl() is monotone, so one threshold decides.  Requires 1e-30 < |m| < 1e30.
__device__ __forceinline__ float nz_threshold(float m)
{
    const float am = fabsf(m);
    float t = __fdiv_rn(1.0f, am);
    for (int i = 0; i < 8 && !(__fmul_rn(am, t) < 1.0f); ++i) t = __uint_as_float(__float_as_uint(t) - 1u);
    for (int i = 0; i < 8; ++i) {
        const float u = __uint_as_float(__float_as_uint(t) + 1u);
        if (__fmul_rn(am, u) < 1.0f) t = u; else break;
    }
    return t;
}

// One entry of the sorted non-zero list: list[e].y = coefficient bits, list[e].x =
//   bits  0..14  raster index
//   bits 15..24  byte offset of the entry inside its chunk of 32 entries (parked by the sizing pass)
//   bits 30..31  group flag: 0 plain, 1 member of an 8-value token (written by its leader), 2 leader of a VLESC2_8x
//                group, 3 leader of a VLESC3_8x group (Run_Length_Encode_Slow.cpp:231-256), set by f32_mark_groups
__device__ __forceinline__ int run_bytes_small(int r) { return r <= 0 ? 0 : r == 1 ? 1 : r < 256 ? 2 : 4; }   // runs < 2^24
constexpr unsigned F32_IDX_MASK = 0x7FFFu;
constexpr unsigned F32_OFF_MASK = 0x3FFu;
struct EntryTok {
    int run;        // zeros between the previous non-zero and this one
    int q;          // quantised value
    float f;        // scaled value (VLESC4 payload)
    int mode;       // GM_ELEM, GM_S8, GM_T8
    int bytes;      // bytes this entry is responsible for
};
__device__ __forceinline__ EntryTok entry_token(uint2 ent, int prev_idx, float mulfac)
{
    EntryTok k;
    const int idx = (int)(ent.x & F32_IDX_MASK);
    const unsigned flag = ent.x >> 30;
    k.run = idx - prev_idx - 1;
    k.f = __fmul_rn(mulfac, __uint_as_float(ent.y));
    k.q = quantise(k.f);
    k.mode = flag == 2u ? GM_S8 : flag == 3u ? GM_T8 : GM_ELEM;
    if (flag == 0u) k.bytes = run_bytes_small(k.run) + elem_bytes(k.q);
    else if (flag == 1u) k.bytes = 0;
    else k.bytes = run_bytes_small(k.run) + (flag == 2u ? 17 : 25);
    return k;
}

// Aligned groups of 8 consecutive non-zeros that the reference codes as one 8-value token: the group's first entry
// decides (group_mode) and flags all eight.  Entry e is examined by thread e mod T.
template <int T>
__device__ __forceinline__ void f32_mark_groups(uint2* list, int n, float mulfac)
{
    for (int e = threadIdx.x & (T - 1); e + 7 < n; e += T) {
        const int idx = (int)(list[e].x & F32_IDX_MASK);
        if ((idx & 7) != 0 || (int)(list[e + 7].x & F32_IDX_MASK) != idx + 7) continue;
        int q8[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) q8[i] = quantise(__fmul_rn(mulfac, __uint_as_float(list[e + i].y)));
        const int mode = group_mode(q8);
        if (mode == GM_S8 || mode == GM_T8) {
            list[e].x |= (mode == GM_S8 ? 2u : 3u) << 30;
#pragma unroll
            for (int i = 1; i < 8; ++i) list[e + i].x |= 1u << 30;
        }
    }
}

template <class Put>
__device__ __forceinline__ int put_elem(int v, float f, int pos, Put put)
{
    if (is_b(v)) { put(pos, (uint8_t)(v & 0xFF)); return pos + 1; }
    if (is_s(v)) { put(pos, (uint8_t)VLESC2); put(pos + 1, (uint8_t)(v & 0xFF)); put(pos + 2, (uint8_t)((v >> 8) & 0xFF)); return pos + 3; }
    if (is_3(v)) {
        put(pos, (uint8_t)VLESC3); put(pos + 1, (uint8_t)(v & 0xFF)); put(pos + 2, (uint8_t)((v >> 8) & 0xFF));
        put(pos + 3, (uint8_t)((v >> 16) & 0xFF));
        return pos + 4;
    }
    const uint32_t u = __float_as_uint(f);      // VLESC4 carries the SCALED float (:267-270)
    put(pos, (uint8_t)VLESC4); put(pos + 1, (uint8_t)(u & 0xFF)); put(pos + 2, (uint8_t)((u >> 8) & 0xFF));
    put(pos + 3, (uint8_t)((u >> 16) & 0xFF)); put(pos + 4, (uint8_t)(u >> 24));
    return pos + 5;
}

// Branch-free emission of one element-mode entry (zero-run flush + value token, Run_Length_Encode_Slow.cpp:21-71 and
// :161-174) for runs below 2^24 (a 32^3 block has 2^15 cells): both tokens are composed with selects and written with
// predicated byte stores, so a warp whose lanes hold different token kinds does not serialise.
__device__ __forceinline__ void put_entry_branchfree(unsigned char* out, int pos, int run, int q, float f)
{
    // run token: none | 00 | 7F r | 7D r0 r1 r2
    const int rl = run_bytes_small(run);
    const unsigned rv = run == 1 ? 0u : run < 256 ? (0x7Fu | ((unsigned)run << 8)) : (0x7Du | ((unsigned)run << 8));
#pragma unroll
    for (int i = 0; i < 4; ++i)
        if (i < rl) out[pos + i] = (unsigned char)(rv >> (8 * i));
    pos += rl;
    // value token: q | 83 q0 q1 | 81 q0 q1 q2 | 80 f0 f1 f2 f3
    const bool b = is_b(q), s2 = is_s(q), s3 = is_3(q);
    const int el = b ? 1 : s2 ? 3 : s3 ? 4 : 5;
    const unsigned code = b ? (unsigned)(q & 0xFF) : s2 ? (unsigned)(VLESC2 & 0xFF) : s3 ? (unsigned)(VLESC3 & 0xFF) : (unsigned)(VLESC4 & 0xFF);
    const unsigned pay = (s2 || s3) ? (unsigned)q : __float_as_uint(f);
    out[pos] = (unsigned char)code;
#pragma unroll
    for (int i = 0; i < 4; ++i)
        if (i + 1 < el) out[pos + 1 + i] = (unsigned char)(pay >> (8 * i));
}

// staged payload (shared memory, 16-byte aligned) -> global, any destination alignment: 32-bit coalesced stores
template <int GT = F32_T>
__device__ __forceinline__ void copy_out(unsigned char* dst, const unsigned char* stage, int len)
{
    const int tid = threadIdx.x & (GT - 1);
    int head = (int)((4u - (unsigned)(reinterpret_cast<uintptr_t>(dst) & 3u)) & 3u);
    if (head > len) head = len;
    if (tid < head) dst[tid] = stage[tid];
    const int nw = (len - head) >> 2;
    uint32_t* d32 = reinterpret_cast<uint32_t*>(dst + head);
    const uint32_t* s32 = reinterpret_cast<const uint32_t*>(stage);
    for (int k = tid; k < nw; k += GT) {
        const uint32_t lo = s32[k];
        d32[k] = head ? __funnelshift_r(lo, s32[k + 1], 8 * head) : lo;
    }
    const int done = head + 4 * nw;
    if (tid < len - done) dst[done + tid] = stage[done + tid];
}

}  // namespace cvx

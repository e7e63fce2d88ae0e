// cvx_wavelet.cuh -- 1-D multi-level Antonini/CDF 9/7 line transforms.
//
// Two arithmetic variants of the same operator:
//   EXACT : direct-form 9/7-tap filter bank in exactly the association the reference ships
//           (generated _Ds79_AVX_n / _Us79_AVX_n, emitted by Wavelet_Transform_Slow.cpp:416-448,
//           :465-490, :613-638, :664-696): acc = a4*p4; acc = acc + a3*p3; ... ; acc = acc + a0*t0,
//           separate fp32 mul and add (the reference makefile:6-8 builds with -mavx only => no FMA).
//           Bit-identical to Wavelet_Transform_Fast_{Forward,Inverse}.
//   LIFT  : the 4-step lifting factorisation (+ two scalings) of the same filter bank with FMA,
//           in place on an interleaved line.  Equal to the filter bank to 2e-14 in fp64
//           (checked for n = 2..256); ~4e-7 rms-relative in fp32.
//
// Lines of <= 64 samples live in registers (all indices are compile-time after unrolling).
// Mirror rules: MIRR / MIRR_SL / MIRR_SH, Wavelet_Transform_Slow.cpp:59-67, :178-199.
#pragma once
#include <cuda_runtime.h>
#include <string.h>

#ifndef CVX_HD
#define CVX_HD __host__ __device__ __forceinline__
#endif

#ifdef __CUDA_ARCH__
#define CVX_MUL(a, b) __fmul_rn((a), (b))
#define CVX_ADD(a, b) __fadd_rn((a), (b))
#define CVX_FMA(a, b, c) __fmaf_rn((a), (b), (c))
#else
#include <cmath>
// host build uses -ffp-contract=off, so these stay separate roundings
#define CVX_MUL(a, b) ((a) * (b))
#define CVX_ADD(a, b) ((a) + (b))
#define CVX_FMA(a, b, c) fmaf((a), (b), (c))
#endif

namespace cvx {

// analysis / synthesis taps, Wavelet_Transform_Slow.cpp:21-30 and :136-145
#define CVX_AL0 8.526986790094000e-001f
#define CVX_AL1 3.774028556126500e-001f
#define CVX_AL2 -1.106244044184200e-001f
#define CVX_AL3 -2.384946501938001e-002f
#define CVX_AL4 3.782845550699501e-002f
#define CVX_AH0 7.884856164056601e-001f
#define CVX_AH1 -4.180922732222101e-001f
#define CVX_AH2 -4.068941760955800e-002f
#define CVX_AH3 6.453888262893799e-002f
#define CVX_SL0 7.884856164056601e-001f
#define CVX_SL1 4.180922732222101e-001f
#define CVX_SL2 -4.068941760955800e-002f
#define CVX_SL3 -6.453888262893799e-002f
#define CVX_SH0 8.526986790094000e-001f
#define CVX_SH1 -3.774028556126500e-001f
#define CVX_SH2 -1.106244044184200e-001f
#define CVX_SH3 2.384946501938001e-002f
#define CVX_SH4 3.782845550699501e-002f

// lifting constants (Daubechies-Sweldens factorisation of the 9/7 pair) and the two
// scalings that reproduce the reference's sqrt(2)-normalised taps.
#define CVX_LA -1.586134342059924f
#define CVX_LB -0.052980118572961f
#define CVX_LG 0.882911075530934f
#define CVX_LD 0.443506852043971f
#define CVX_ZL 1.1496043988602411f
#define CVX_ZH 0.8698644516247813f
#define CVX_IZL 0.8698644516247813f /* 1/ZL == ZH for this normalisation */
#define CVX_IZH 1.1496043988602411f /* 1/ZH == ZL */

CVX_HD constexpr int mirr(int v, int n)
{
    v = v < 0 ? -v : v;
    v = v >= n ? 2 * n - 2 - v : v;
    v = v < 0 ? -v : v;
    v = v >= n ? 2 * n - 2 - v : v;
    return v;
}
CVX_HD constexpr int mirr_sl(int v, int nl)
{
    v = v < 0 ? -v : v; v = v >= nl ? 2 * nl - 1 - v : v;
    v = v < 0 ? -v : v; v = v >= nl ? 2 * nl - 1 - v : v;
    v = v < 0 ? -v : v; v = v >= nl ? 2 * nl - 1 - v : v;
    return v;
}
CVX_HD constexpr int mirr_sh(int v, int nl, int nh)
{
    v -= nl;
    v = v < 0 ? -v - 1 : v; v = v >= nh ? 2 * nh - 2 - v : v;
    v = v < 0 ? -v - 1 : v; v = v >= nh ? 2 * nh - 2 - v : v;
    v = v < 0 ? -v - 1 : v; v = v >= nh ? 2 * nh - 2 - v : v;
    return nl + v;
}

// ------------------------------------------------------------------------------------------
// Value types of the register-resident line transforms: `float` = one line, `F2` = two independent
// lines side by side in one 64-bit register pair, computed with the Blackwell packed-fp32 instructions
// (add/mul/fma.rn.f32x2 -> FADD2 / FMUL2 / FFMA2: one issue slot for both lines, same IEEE rounding as
// the scalar instructions, so EXACT stays bit-exact).  The host build keeps the pair as two floats.
// ------------------------------------------------------------------------------------------
struct F2 {
    unsigned long long u;
};
CVX_HD F2 f2_make(float x, float y)
{
    F2 r;
#ifdef __CUDA_ARCH__
    asm("mov.b64 %0, {%1, %2};" : "=l"(r.u) : "f"(x), "f"(y));
#else
    unsigned int a, b;
    memcpy(&a, &x, 4); memcpy(&b, &y, 4);
    r.u = (unsigned long long)a | ((unsigned long long)b << 32);
#endif
    return r;
}
CVX_HD float f2_lo(F2 v)
{
#ifdef __CUDA_ARCH__
    float x, y;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(v.u));
    return x;
#else
    unsigned int a = (unsigned int)v.u; float x; memcpy(&x, &a, 4); return x;
#endif
}
CVX_HD float f2_hi(F2 v)
{
#ifdef __CUDA_ARCH__
    float x, y;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(v.u));
    return y;
#else
    unsigned int a = (unsigned int)(v.u >> 32); float x; memcpy(&x, &a, 4); return x;
#endif
}
CVX_HD float v_add(float a, float b) { return CVX_ADD(a, b); }
CVX_HD float v_mulc(float c, float a) { return CVX_MUL(c, a); }
CVX_HD float v_fmac(float c, float a, float b) { return CVX_FMA(c, a, b); }      // c * a + b
CVX_HD F2 v_add(F2 a, F2 b)
{
#ifdef __CUDA_ARCH__
    F2 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d.u) : "l"(a.u), "l"(b.u)); return d;
#else
    return f2_make(CVX_ADD(f2_lo(a), f2_lo(b)), CVX_ADD(f2_hi(a), f2_hi(b)));
#endif
}
CVX_HD F2 v_mulc(float c, F2 a)
{
#ifdef __CUDA_ARCH__
    F2 d; const F2 cc = f2_make(c, c);
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d.u) : "l"(cc.u), "l"(a.u)); return d;
#else
    return f2_make(CVX_MUL(c, f2_lo(a)), CVX_MUL(c, f2_hi(a)));
#endif
}
CVX_HD F2 v_fmac(float c, F2 a, F2 b)
{
#ifdef __CUDA_ARCH__
    F2 d; const F2 cc = f2_make(c, c);
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d.u) : "l"(cc.u), "l"(a.u), "l"(b.u)); return d;
#else
    return f2_make(CVX_FMA(c, f2_lo(a), f2_lo(b)), CVX_FMA(c, f2_hi(a), f2_hi(b)));
#endif
}

// S2: two lines side by side computed with SCALAR instructions.  EXACT mode uses it instead of F2 because ptxas
// (12.9) contracts mul.rn.f32x2 + add.rn.f32x2 into FFMA2 even under --fmad=false, which would break bit-exactness
// against the reference's separate multiply and add (Wavelet_Transform_Slow.cpp:418-448).
struct S2 {
    float x, y;
};
CVX_HD S2 v_add(S2 a, S2 b) { S2 r; r.x = CVX_ADD(a.x, b.x); r.y = CVX_ADD(a.y, b.y); return r; }
CVX_HD S2 v_mulc(float c, S2 a) { S2 r; r.x = CVX_MUL(c, a.x); r.y = CVX_MUL(c, a.y); return r; }
CVX_HD S2 v_fmac(float c, S2 a, S2 b) { S2 r; r.x = CVX_FMA(c, a.x, b.x); r.y = CVX_FMA(c, a.y, b.y); return r; }
template <bool EXACT> struct PairOf { typedef F2 type; };
template <> struct PairOf<true> { typedef S2 type; };
CVX_HD void p_set(F2& d, float x, float y) { d = f2_make(x, y); }
CVX_HD void p_set(S2& d, float x, float y) { d.x = x; d.y = y; }
CVX_HD float p_lo(F2 v) { return f2_lo(v); }
CVX_HD float p_hi(F2 v) { return f2_hi(v); }
CVX_HD float p_lo(S2 v) { return v.x; }
CVX_HD float p_hi(S2 v) { return v.y; }

// a thread's bundle of lines: two side by side (512-thread kernels: F2 packed or S2 scalar) or one (1024-thread kernels)
template <bool EXACT, int T> struct LineOf { typedef typename PairOf<EXACT>::type type; };
template <bool EXACT> struct LineOf<EXACT, 1024> { typedef float type; };
CVX_HD void p_set(float& d, float x, float) { d = x; }
CVX_HD float p_lo(float v) { return v; }
CVX_HD float p_hi(float v) { return v; }

// ------------------------------------------------------------------------------------------
// EXACT, register-resident.  One level on v[0..N): low half then high half (out-of-place
// semantics through the copy t[]).
// ------------------------------------------------------------------------------------------
template <int N, class V>
CVX_HD void fwd_exact_level(V* v)
{
    V t[N];
#pragma unroll
    for (int i = 0; i < N; ++i) t[i] = v[i];
    constexpr int nh = N / 2;
#pragma unroll
    for (int ix = 0; ix < nh; ++ix) {
        {
            const int i0 = 2 * ix;
            V acc = v_mulc(CVX_AL4, v_add(t[mirr(i0 - 4, N)], t[mirr(i0 + 4, N)]));
            acc = v_add(acc, v_mulc(CVX_AL3, v_add(t[mirr(i0 - 3, N)], t[mirr(i0 + 3, N)])));
            acc = v_add(acc, v_mulc(CVX_AL2, v_add(t[mirr(i0 - 2, N)], t[mirr(i0 + 2, N)])));
            acc = v_add(acc, v_mulc(CVX_AL1, v_add(t[mirr(i0 - 1, N)], t[mirr(i0 + 1, N)])));
            acc = v_add(acc, v_mulc(CVX_AL0, t[i0]));
            v[ix] = acc;
        }
        {
            const int i0 = 2 * ix + 1;
            V acc = v_mulc(CVX_AH3, v_add(t[mirr(i0 - 3, N)], t[mirr(i0 + 3, N)]));
            acc = v_add(acc, v_mulc(CVX_AH2, v_add(t[mirr(i0 - 2, N)], t[mirr(i0 + 2, N)])));
            acc = v_add(acc, v_mulc(CVX_AH1, v_add(t[mirr(i0 - 1, N)], t[mirr(i0 + 1, N)])));
            acc = v_add(acc, v_mulc(CVX_AH0, t[i0]));
            v[nh + ix] = acc;
        }
    }
}

template <int N, class V>
CVX_HD void inv_exact_level(V* v)
{
    V t[N];
#pragma unroll
    for (int i = 0; i < N; ++i) t[i] = v[i];
    constexpr int nh = N / 2, nl = N - nh;
#pragma unroll
    for (int k = 0; k < nl; ++k) {
        {   // even sample 2k
            V acc = v_mulc(CVX_SH3, v_add(t[mirr_sh(nl + k - 2, nl, nh)], t[mirr_sh(nl + k + 1, nl, nh)]));
            acc = v_add(acc, v_mulc(CVX_SL2, v_add(t[mirr_sl(k - 1, nl)], t[mirr_sl(k + 1, nl)])));
            acc = v_add(acc, v_mulc(CVX_SH1, v_add(t[mirr_sh(nl + k - 1, nl, nh)], t[mirr_sh(nl + k, nl, nh)])));
            acc = v_add(acc, v_mulc(CVX_SL0, t[k]));
            v[2 * k] = acc;
        }
        {   // odd sample 2k+1
            V acc = v_mulc(CVX_SH4, v_add(t[mirr_sh(nl + k - 2, nl, nh)], t[mirr_sh(nl + k + 2, nl, nh)]));
            acc = v_add(acc, v_mulc(CVX_SL3, v_add(t[mirr_sl(k - 1, nl)], t[mirr_sl(k + 2, nl)])));
            acc = v_add(acc, v_mulc(CVX_SH2, v_add(t[mirr_sh(nl + k - 1, nl, nh)], t[mirr_sh(nl + k + 1, nl, nh)])));
            acc = v_add(acc, v_mulc(CVX_SL1, v_add(t[mirr_sl(k, nl)], t[mirr_sl(k + 1, nl)])));
            acc = v_add(acc, v_mulc(CVX_SH0, t[nl + k]));
            v[2 * k + 1] = acc;
        }
    }
}

// levels N, N/2, ..., 2 (Wavelet_Transform_Fast.cpp:48-55) / 2, ..., N (:183-190)
template <int N, class V>
CVX_HD void fwd_exact(V* v)
{
    fwd_exact_level<N>(v);
    if constexpr (N > 2) fwd_exact<N / 2>(v);
}
template <int N, class V>
CVX_HD void inv_exact(V* v)
{
    if constexpr (N > 2) inv_exact<N / 2>(v);
    inv_exact_level<N>(v);
}

// ------------------------------------------------------------------------------------------
// LIFT, register-resident, in place on the interleaved line.  Level with stride S works on
// the samples at positions S*j, j < N/S; evens are s_j, odds are d_j.  Boundary:
// s_{m} -> s_{m-1}, d_{-1} -> d_0 (whole-sample symmetric extension == MIRR).
// ------------------------------------------------------------------------------------------
template <int N, int S, class V>
CVX_HD void fwd_lift_level(V* v)
{
    constexpr int m = N / S / 2;
#define CVX_S(j) v[S * 2 * (j)]
#define CVX_D(j) v[S * (2 * (j) + 1)]
#pragma unroll
    for (int j = 0; j < m; ++j) CVX_D(j) = v_fmac(CVX_LA, v_add(CVX_S(j), CVX_S(j + 1 < m ? j + 1 : m - 1)), CVX_D(j));
#pragma unroll
    for (int j = 0; j < m; ++j) CVX_S(j) = v_fmac(CVX_LB, v_add(CVX_D(j > 0 ? j - 1 : 0), CVX_D(j)), CVX_S(j));
#pragma unroll
    for (int j = 0; j < m; ++j) CVX_D(j) = v_fmac(CVX_LG, v_add(CVX_S(j), CVX_S(j + 1 < m ? j + 1 : m - 1)), CVX_D(j));
#pragma unroll
    for (int j = 0; j < m; ++j) CVX_S(j) = v_fmac(CVX_LD, v_add(CVX_D(j > 0 ? j - 1 : 0), CVX_D(j)), CVX_S(j));
}
template <int N, int S, class V>
CVX_HD void inv_lift_level(V* v)
{
    constexpr int m = N / S / 2;
#pragma unroll
    for (int j = 0; j < m; ++j) CVX_S(j) = v_fmac(-CVX_LD, v_add(CVX_D(j > 0 ? j - 1 : 0), CVX_D(j)), CVX_S(j));
#pragma unroll
    for (int j = 0; j < m; ++j) CVX_D(j) = v_fmac(-CVX_LG, v_add(CVX_S(j), CVX_S(j + 1 < m ? j + 1 : m - 1)), CVX_D(j));
#pragma unroll
    for (int j = 0; j < m; ++j) CVX_S(j) = v_fmac(-CVX_LB, v_add(CVX_D(j > 0 ? j - 1 : 0), CVX_D(j)), CVX_S(j));
#pragma unroll
    for (int j = 0; j < m; ++j) CVX_D(j) = v_fmac(-CVX_LA, v_add(CVX_S(j), CVX_S(j + 1 < m ? j + 1 : m - 1)), CVX_D(j));
#undef CVX_S
#undef CVX_D
}
// The two scalings of every level (low * ZL, high * ZH) are linear, so they are folded into ONE multiply per
// coefficient after the last level (forward) / before the first (inverse): the sample at interleaved position
// p = 2^(l-1) * odd is the high band of level l and carries ZL^(l-1) * ZH; position 0 carries ZL^levels.
// 20 % fewer FP32 operations per line than scaling level by level; rounding differs in the last bits only.
constexpr double lift_cpow(double b, int e) { double r = 1.0; for (int i = 0; i < e; ++i) r *= b; return r; }
constexpr int lift_ctz(int p) { int n = 0; while (!(p & 1)) { p >>= 1; ++n; } return n; }
constexpr int lift_log2(int n) { int l = 0; while (n > 1) { n >>= 1; ++l; } return l; }
constexpr float lift_scale_fwd(int p, int N)
{
    return (float)(p == 0 ? lift_cpow((double)CVX_ZL, lift_log2(N)) : lift_cpow((double)CVX_ZL, lift_ctz(p)) * (double)CVX_ZH);
}
constexpr float lift_scale_inv(int p, int N)
{
    return (float)(p == 0 ? lift_cpow((double)CVX_IZL, lift_log2(N)) : lift_cpow((double)CVX_IZL, lift_ctz(p)) * (double)CVX_IZH);
}
template <int N, int P, bool FWD, class V>
CVX_HD void lift_scale_all(V* v)
{
    constexpr float c = FWD ? lift_scale_fwd(P, N) : lift_scale_inv(P, N);      // compile-time: an immediate operand
    v[P] = v_mulc(c, v[P]);
    if constexpr (P + 1 < N) lift_scale_all<N, P + 1, FWD>(v);
}
template <int N, int S, class V>
CVX_HD void fwd_lift_levels(V* v)
{
    fwd_lift_level<N, S>(v);
    if constexpr (N / S > 2) fwd_lift_levels<N, S * 2>(v);
}
template <int N, int S, class V>
CVX_HD void inv_lift_levels(V* v)
{
    if constexpr (N / S > 2) inv_lift_levels<N, S * 2>(v);
    inv_lift_level<N, S>(v);
}
template <int N, class V>
CVX_HD void fwd_lift(V* v)
{
    fwd_lift_levels<N, 1>(v);
    lift_scale_all<N, 0, true>(v);
}
template <int N, class V>
CVX_HD void inv_lift(V* v)
{
    lift_scale_all<N, 0, false>(v);
    inv_lift_levels<N, 1>(v);
}

// Two lines that arrive and leave as SCALARS (the x rows of a block come out of LDS.128 quads, whose registers cannot be
// addressed as packed pairs): the pair (line a, line b) is formed by the first scalar operation that touches a sample
// and split again by the last one, so everything in between runs packed (FADD2/FFMA2) without a single register move.
// Same operations in the same order as fwd_lift / inv_lift on each line: bit-identical results.
//   forward: level-1 steps 1 and 2 write the packed d and s; the final scaling (scalar) unpacks.   560 -> 360 issue slots
//   inverse: the scaling packs; level-1 steps 3 and 4 (scalar) unpack.
template <int N, int P>
CVX_HD void lift_unpack_scale_fwd(const F2* v, float* oa, float* ob)
{
    constexpr float c = lift_scale_fwd(P, N);
    oa[P] = CVX_MUL(c, f2_lo(v[P]));
    ob[P] = CVX_MUL(c, f2_hi(v[P]));
    if constexpr (P + 1 < N) lift_unpack_scale_fwd<N, P + 1>(v, oa, ob);
}
template <int N, int P>
CVX_HD void lift_pack_scale_inv(const float* a, const float* b, F2* v)
{
    constexpr float c = lift_scale_inv(P, N);
    v[P] = f2_make(CVX_MUL(c, a[P]), CVX_MUL(c, b[P]));
    if constexpr (P + 1 < N) lift_pack_scale_inv<N, P + 1>(a, b, v);
}
// a, b: natural order in; oa, ob: interleaved (lifting) positions out
template <int N>
CVX_HD void fwd_lift_pair(const float* a, const float* b, float* oa, float* ob)
{
    constexpr int m = N / 2;
    F2 v[N];
#pragma unroll
    for (int j = 0; j < m; ++j) {
        const int jn = j + 1 < m ? j + 1 : m - 1;
        v[2 * j + 1] = f2_make(CVX_FMA(CVX_LA, CVX_ADD(a[2 * j], a[2 * jn]), a[2 * j + 1]),
                               CVX_FMA(CVX_LA, CVX_ADD(b[2 * j], b[2 * jn]), b[2 * j + 1]));
    }
#pragma unroll
    for (int j = 0; j < m; ++j) {
        const F2 t = v_add(v[2 * (j > 0 ? j - 1 : 0) + 1], v[2 * j + 1]);
        v[2 * j] = f2_make(CVX_FMA(CVX_LB, f2_lo(t), a[2 * j]), CVX_FMA(CVX_LB, f2_hi(t), b[2 * j]));
    }
#pragma unroll
    for (int j = 0; j < m; ++j) v[2 * j + 1] = v_fmac(CVX_LG, v_add(v[2 * j], v[2 * (j + 1 < m ? j + 1 : m - 1)]), v[2 * j + 1]);
#pragma unroll
    for (int j = 0; j < m; ++j) v[2 * j] = v_fmac(CVX_LD, v_add(v[2 * (j > 0 ? j - 1 : 0) + 1], v[2 * j + 1]), v[2 * j]);
    if constexpr (N > 2) fwd_lift_levels<N, 2>(v);
    lift_unpack_scale_fwd<N, 0>(v, oa, ob);
}
// a, b: interleaved (lifting) positions in; oa, ob: natural order out
template <int N>
CVX_HD void inv_lift_pair(const float* a, const float* b, float* oa, float* ob)
{
    constexpr int m = N / 2;
    F2 v[N];
    lift_pack_scale_inv<N, 0>(a, b, v);
    if constexpr (N > 2) inv_lift_levels<N, 2>(v);
#pragma unroll
    for (int j = 0; j < m; ++j) v[2 * j] = v_fmac(-CVX_LD, v_add(v[2 * (j > 0 ? j - 1 : 0) + 1], v[2 * j + 1]), v[2 * j]);
#pragma unroll
    for (int j = 0; j < m; ++j) v[2 * j + 1] = v_fmac(-CVX_LG, v_add(v[2 * j], v[2 * (j + 1 < m ? j + 1 : m - 1)]), v[2 * j + 1]);
#pragma unroll
    for (int j = 0; j < m; ++j) {
        const F2 t = v_add(v[2 * (j > 0 ? j - 1 : 0) + 1], v[2 * j + 1]);
        oa[2 * j] = CVX_FMA(-CVX_LB, f2_lo(t), f2_lo(v[2 * j]));
        ob[2 * j] = CVX_FMA(-CVX_LB, f2_hi(t), f2_hi(v[2 * j]));
    }
#pragma unroll
    for (int j = 0; j < m; ++j) {
        const int jn = j + 1 < m ? j + 1 : m - 1;
        oa[2 * j + 1] = CVX_FMA(-CVX_LA, CVX_ADD(oa[2 * j], oa[2 * jn]), f2_lo(v[2 * j + 1]));
        ob[2 * j + 1] = CVX_FMA(-CVX_LA, CVX_ADD(ob[2 * j], ob[2 * jn]), f2_hi(v[2 * j + 1]));
    }
}

// ---- the scalings of all three axes in ONE multiply (32^3 LIFTING encoders) -----------------------------------------
// The per-line scaling is a diagonal operator, so in a separable 3-D transform the scalings of the y and z lines commute
// with everything and can ride on the x rows' multiply: a thread that owns the x row (y, z) multiplies by
// scale_x(position) * scale_y(y) * scale_z(z) -- 64 FMUL2 fewer per thread, the z and y passes run the bare lifting steps
// (fwd_lift_levels).  Rounding differs from three separate multiplies in the last bit only.  (Measured on config 3: encode
// 2.271 -> 2.249 ms.  The same fold in the decoders, where the multiply opens the x pass and its twelve row multipliers are
// live together with both rows, was slower -- 2.170 -> 2.208 ms -- and is not used.)
// scale that lift_scale_all applies to sub-band index i of a line of N
template <int N, bool FWD, int L>
CVX_HD float lift_scale_of_sub_rec(int i, float r)
{
    constexpr float c = FWD ? lift_scale_fwd(1 << L, N) : lift_scale_inv(1 << L, N);   // compile-time: constexpr functions are host-only
    r = (i >= (N >> (L + 1)) && i < (N >> L)) ? c : r;      // indices [N / 2^(L+1), N / 2^L) sit at positions with L trailing zeros
    if constexpr ((1 << (L + 1)) < N) return lift_scale_of_sub_rec<N, FWD, L + 1>(i, r);
    else return r;
}
template <int N, bool FWD>
CVX_HD float lift_scale_of_sub(int i)
{
    constexpr float c0 = FWD ? lift_scale_fwd(0, N) : lift_scale_inv(0, N);
    return lift_scale_of_sub_rec<N, FWD, 0>(i, c0);
}
template <int N, int P>
CVX_HD void lift_unpack_scale_fwd_rs(const F2* v, float* oa, float* ob, float fa, float fb)
{
    constexpr float c = lift_scale_fwd(P, N);
    oa[P] = CVX_MUL(CVX_MUL(c, fa), f2_lo(v[P]));
    ob[P] = CVX_MUL(CVX_MUL(c, fb), f2_hi(v[P]));
    if constexpr (P + 1 < N) lift_unpack_scale_fwd_rs<N, P + 1>(v, oa, ob, fa, fb);
}
// fwd_lift_pair with the rows' extra factors fa (line a) and fb (line b) folded into the scaling
template <int N>
CVX_HD void fwd_lift_pair_rs(const float* a, const float* b, float* oa, float* ob, float fa, float fb)
{
    constexpr int m = N / 2;
    F2 v[N];
#pragma unroll
    for (int j = 0; j < m; ++j) {
        const int jn = j + 1 < m ? j + 1 : m - 1;
        v[2 * j + 1] = f2_make(CVX_FMA(CVX_LA, CVX_ADD(a[2 * j], a[2 * jn]), a[2 * j + 1]),
                               CVX_FMA(CVX_LA, CVX_ADD(b[2 * j], b[2 * jn]), b[2 * j + 1]));
    }
#pragma unroll
    for (int j = 0; j < m; ++j) {
        const F2 t = v_add(v[2 * (j > 0 ? j - 1 : 0) + 1], v[2 * j + 1]);
        v[2 * j] = f2_make(CVX_FMA(CVX_LB, f2_lo(t), a[2 * j]), CVX_FMA(CVX_LB, f2_hi(t), b[2 * j]));
    }
#pragma unroll
    for (int j = 0; j < m; ++j) v[2 * j + 1] = v_fmac(CVX_LG, v_add(v[2 * j], v[2 * (j + 1 < m ? j + 1 : m - 1)]), v[2 * j + 1]);
#pragma unroll
    for (int j = 0; j < m; ++j) v[2 * j] = v_fmac(CVX_LD, v_add(v[2 * (j > 0 ? j - 1 : 0) + 1], v[2 * j + 1]), v[2 * j]);
    if constexpr (N > 2) fwd_lift_levels<N, 2>(v);
    lift_unpack_scale_fwd_rs<N, 0>(v, oa, ob, fa, fb);
}
// interleaved position p  <->  sub-band index (low first, recursively; DC at 0)
CVX_HD constexpr int lift_pos_to_idx(int p, int N)
{
    if (p == 0) return 0;
    int s = p & (-p);
    return N / (2 * s) + (p / s - 1) / 2;
}

// inverse of lift_pos_to_idx: sub-band index i -> interleaved position
CVX_HD constexpr int lift_idx_to_pos(int i, int N)
{
    if (i == 0) return 0;
    int h = 1;
    while (2 * h <= i) h *= 2;
    return (N / (2 * h)) * (2 * (i - h) + 1);
}

// ------------------------------------------------------------------------------------------
// Uniform front end.  v[] holds the line in NATURAL order before fwd / after inv, and in
// SUB-BAND order after fwd / before inv for EXACT; for LIFT the sub-band side is addressed
// through sub_pos<>(i): the register that holds sub-band coefficient i.
// ------------------------------------------------------------------------------------------
template <int N, bool EXACT, class V>
CVX_HD void fwd_line(V* v)
{
    if constexpr (EXACT) fwd_exact<N>(v); else fwd_lift<N>(v);
}
template <int N, bool EXACT, class V>
CVX_HD void inv_line(V* v)
{
    if constexpr (EXACT) inv_exact<N>(v); else inv_lift<N>(v);
}
// S2: one line after the other (keeps the live register set at one line + its level copy)
template <int N, bool EXACT>
CVX_HD void fwd_line(S2* v)
{
    float a[N];
#pragma unroll
    for (int i = 0; i < N; ++i) a[i] = v[i].x;
    fwd_line<N, EXACT>(a);
#pragma unroll
    for (int i = 0; i < N; ++i) { v[i].x = a[i]; a[i] = v[i].y; }
    fwd_line<N, EXACT>(a);
#pragma unroll
    for (int i = 0; i < N; ++i) v[i].y = a[i];
}
template <int N, bool EXACT>
CVX_HD void inv_line(S2* v)
{
    float a[N];
#pragma unroll
    for (int i = 0; i < N; ++i) a[i] = v[i].x;
    inv_line<N, EXACT>(a);
#pragma unroll
    for (int i = 0; i < N; ++i) { v[i].x = a[i]; a[i] = v[i].y; }
    inv_line<N, EXACT>(a);
#pragma unroll
    for (int i = 0; i < N; ++i) v[i].y = a[i];
}
// register index r -> sub-band index held there
template <int N, bool EXACT>
CVX_HD constexpr int reg_to_sub(int r)
{
    return EXACT ? r : lift_pos_to_idx(r, N);
}

// sub-band index i -> register that holds it
template <int N, bool EXACT>
CVX_HD constexpr int sub_to_reg(int i)
{
    return EXACT ? i : lift_idx_to_pos(i, N);
}

// ------------------------------------------------------------------------------------------
// Generic-length EXACT transforms on a strided line in memory with a caller-provided
// temporary of n floats (any power of two up to 256).  Used by the staged large-block path.
// ------------------------------------------------------------------------------------------
__host__ __device__ inline void fwd_exact_level_mem(float* p, long stride, float* t, int n)
{
    for (int i = 0; i < n; ++i) t[i] = p[i * stride];
    int nh = n / 2;
    for (int ix = 0; ix < nh; ++ix) {
        int i0 = 2 * ix;
        float acc = CVX_MUL(CVX_AL4, CVX_ADD(t[mirr(i0 - 4, n)], t[mirr(i0 + 4, n)]));
        acc = CVX_ADD(acc, CVX_MUL(CVX_AL3, CVX_ADD(t[mirr(i0 - 3, n)], t[mirr(i0 + 3, n)])));
        acc = CVX_ADD(acc, CVX_MUL(CVX_AL2, CVX_ADD(t[mirr(i0 - 2, n)], t[mirr(i0 + 2, n)])));
        acc = CVX_ADD(acc, CVX_MUL(CVX_AL1, CVX_ADD(t[mirr(i0 - 1, n)], t[mirr(i0 + 1, n)])));
        acc = CVX_ADD(acc, CVX_MUL(CVX_AL0, t[i0]));
        p[ix * stride] = acc;
        i0 = 2 * ix + 1;
        acc = CVX_MUL(CVX_AH3, CVX_ADD(t[mirr(i0 - 3, n)], t[mirr(i0 + 3, n)]));
        acc = CVX_ADD(acc, CVX_MUL(CVX_AH2, CVX_ADD(t[mirr(i0 - 2, n)], t[mirr(i0 + 2, n)])));
        acc = CVX_ADD(acc, CVX_MUL(CVX_AH1, CVX_ADD(t[mirr(i0 - 1, n)], t[mirr(i0 + 1, n)])));
        acc = CVX_ADD(acc, CVX_MUL(CVX_AH0, t[i0]));
        p[(nh + ix) * stride] = acc;
    }
}
__host__ __device__ inline void inv_exact_level_mem(float* p, long stride, float* t, int n)
{
    for (int i = 0; i < n; ++i) t[i] = p[i * stride];
    int nh = n / 2, nl = n - nh;
    for (int k = 0; k < nl; ++k) {
        float acc = CVX_MUL(CVX_SH3, CVX_ADD(t[mirr_sh(nl + k - 2, nl, nh)], t[mirr_sh(nl + k + 1, nl, nh)]));
        acc = CVX_ADD(acc, CVX_MUL(CVX_SL2, CVX_ADD(t[mirr_sl(k - 1, nl)], t[mirr_sl(k + 1, nl)])));
        acc = CVX_ADD(acc, CVX_MUL(CVX_SH1, CVX_ADD(t[mirr_sh(nl + k - 1, nl, nh)], t[mirr_sh(nl + k, nl, nh)])));
        acc = CVX_ADD(acc, CVX_MUL(CVX_SL0, t[k]));
        p[(2 * k) * stride] = acc;
        acc = CVX_MUL(CVX_SH4, CVX_ADD(t[mirr_sh(nl + k - 2, nl, nh)], t[mirr_sh(nl + k + 2, nl, nh)]));
        acc = CVX_ADD(acc, CVX_MUL(CVX_SL3, CVX_ADD(t[mirr_sl(k - 1, nl)], t[mirr_sl(k + 2, nl)])));
        acc = CVX_ADD(acc, CVX_MUL(CVX_SH2, CVX_ADD(t[mirr_sh(nl + k - 1, nl, nh)], t[mirr_sh(nl + k + 1, nl, nh)])));
        acc = CVX_ADD(acc, CVX_MUL(CVX_SL1, CVX_ADD(t[mirr_sl(k, nl)], t[mirr_sl(k + 1, nl)])));
        acc = CVX_ADD(acc, CVX_MUL(CVX_SH0, t[nl + k]));
        p[(2 * k + 1) * stride] = acc;
    }
}
__host__ __device__ inline void fwd_exact_mem(float* p, long stride, float* t, int dim)
{
    for (int n = dim; n >= 2; n >>= 1) fwd_exact_level_mem(p, stride, t, n);
}
__host__ __device__ inline void inv_exact_mem(float* p, long stride, float* t, int dim)
{
    for (int n = 2; n <= dim; n <<= 1) inv_exact_level_mem(p, stride, t, n);
}

}  // namespace cvx

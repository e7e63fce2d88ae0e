// cvx_rle.cuh -- quantiser + byte-oriented run-length / escape coder, written as per-thread
// phases so that the same code runs inside the CUDA kernels and in the host emulation used
// by the CPU unit tests (tests/hostemu).
//
// Format and encoder decision procedure: Run_Length_Encode_Slow.cpp:189-294 (TMJ_AVX_RLE
// path), flush rule :21-71, escape codes Run_Length_Escape_Codes.hxx:8-14, decoder :388-527.
//
// Parallel formulation (the reference is byte-serial):
//   encode : a thread owns a run of consecutive groups-of-8 in raster order and makes two
//            walks over them.  Walk 1 (seg_measure) yields {has, lead, trail, inner}:
//            whether it emits any non-run token, the zeros before its first emitter, the
//            zeros after its last one, and the bytes from first to last emitter.  A segmented
//            scan over threads gives the zero run carried INTO each thread and, from that,
//            each thread's byte count; an exclusive scan gives its byte offset.  Walk 2
//            (seg_emit) writes the bytes.  Zero runs therefore span threads exactly as the
//            reference's `rle` counter spans groups.
//   decode : every 16-byte chunk of the payload is walked speculatively from the chunk
//            before it (tokens self-synchronise within a few bytes), the chain of
//            exit-state == next entry-state is verified and repaired until consistent, then
//            a scan of values-per-chunk places every token (chunk_walk / chunk_count /
//            chunk_expand).
#pragma once
#include <stdint.h>
#include <limits.h>
#include <string.h>
#include "cvx_wavelet.cuh"

namespace cvx {

enum : int {
    RLESC1 = 127, VLESC3_8x = 126, RLESC3 = 125,
    VLESC2 = -125, VLESC2_8x = -126, VLESC3 = -127, VLESC4 = -128
};

// _mm256_cvttps_epi32 semantics (Run_Length_Encode_Slow.cpp:204): truncate toward zero;
// NaN and |f| >= 2^31 give 0x80000000 (CUDA's cvt.rzi saturates and maps NaN to 0, so guard).
CVX_HD int quantise(float f)
{
#ifdef __CUDA_ARCH__
    int q = __float2int_rz(f);
    return (fabsf(f) < 2147483648.0f) ? q : INT_MIN;
#else
    return (fabsf(f) < 2147483648.0f) ? (int)f : INT_MIN;
#endif
}

// classes are tested on (float)q in the reference (:205-245); for these bounds the integer
// compares below are equivalent because every bound is exactly representable.
CVX_HD bool is_b(int q) { return q > VLESC2 && q < RLESC3; }       // 1-byte literal (zero included)
CVX_HD bool is_s(int q) { return q >= -32768 && q <= 32767; }
CVX_HD bool is_3(int q) { return q >= -8388608 && q <= 8388607; }
CVX_HD int elem_bytes(int q) { return is_b(q) ? 1 : is_s(q) ? 3 : is_3(q) ? 4 : 5; }

// bytes of the token(s) that flush a zero run of length r (EncodeRLE_Slow :21-71).
// r >= 2^24 (only an all-zero 256^3 block) would wrap in the reference's 24-bit field and be
// undecodable; we split it into 0xFFFFFF-long RLESC3 tokens instead (documented deviation).
CVX_HD int run_bytes(int r)
{
    int n = 0;
    while (r > 0xFFFFFF) { n += 4; r -= 0xFFFFFF; }
    return n + (r <= 0 ? 0 : r == 1 ? 1 : r < 256 ? 2 : 4);
}

enum : int { GM_ZERO8 = 0, GM_ELEM = 1, GM_S8 = 2, GM_T8 = 3 };

// group-of-8 mode selection, Run_Length_Encode_Slow.cpp:206-257
CVX_HD int group_mode(const int* q)
{
    int Z = 0, nb = 0, ns = 0, n3 = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        Z += (q[j] == 0); nb += is_b(q[j]); ns += is_s(q[j]); n3 += is_3(q[j]);
    }
    if (Z == 8) return GM_ZERO8;
    if (Z != 0) return GM_ELEM;
    if (nb == 8) return GM_ELEM;                                   // 8 literals == elementwise
    if (ns == 8 && (nb + (8 - nb) * 3) > 17) return GM_S8;
    if (n3 == 8 && (nb + (ns - nb) * 3 + (8 - ns) * 4) > 25) return GM_T8;
    return GM_ELEM;
}

struct SegSum {
    int has;     // emits at least one non-run token
    int lead;    // zeros before the first emitter (== segment length if !has)
    int trail;   // zeros after the last emitter   (== segment length if !has)
    int inner;   // bytes from the first emitter through the last one, interior runs included
};

// Walk 1.  `get8(g, f)` loads the 8 coefficients of group g (0-based within the segment).
template <class Get8>
CVX_HD SegSum seg_measure(int ngroups, float mulfac, Get8 get8)
{
    SegSum s; s.has = 0; s.lead = 0; s.trail = 0; s.inner = 0;
    int run = 0;
    for (int g = 0; g < ngroups; ++g) {
        float c[8]; int q[8];
        get8(g, c);
#pragma unroll
        for (int j = 0; j < 8; ++j) q[j] = quantise(CVX_MUL(mulfac, c[j]));
        int mode = group_mode(q);
        if (mode == GM_ZERO8) { run += 8; continue; }
        if (mode == GM_S8 || mode == GM_T8) {
            if (!s.has) { s.has = 1; s.lead = run; } else s.inner += run_bytes(run);
            run = 0;
            s.inner += (mode == GM_S8) ? 17 : 25;
            continue;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            if (q[j] == 0) { ++run; continue; }
            if (!s.has) { s.has = 1; s.lead = run; } else s.inner += run_bytes(run);
            run = 0;
            s.inner += elem_bytes(q[j]);
        }
    }
    s.trail = run;
    if (!s.has) s.lead = run;
    return s;
}

template <class Put>
CVX_HD int put_run(int r, int pos, Put put)
{
    while (r > 0xFFFFFF) {
        put(pos, (uint8_t)RLESC3); put(pos + 1, 0xFF); put(pos + 2, 0xFF); put(pos + 3, 0xFF);
        pos += 4; r -= 0xFFFFFF;
    }
    if (r <= 0) return pos;
    if (r == 1) { put(pos, 0); return pos + 1; }
    if (r < 256) { put(pos, (uint8_t)RLESC1); put(pos + 1, (uint8_t)r); return pos + 2; }
    put(pos, (uint8_t)RLESC3); put(pos + 1, (uint8_t)(r & 0xFF)); put(pos + 2, (uint8_t)((r >> 8) & 0xFF));
    put(pos + 3, (uint8_t)((r >> 16) & 0xFF));
    return pos + 4;
}

// Walk 2.  carry_in = zeros immediately preceding this segment; pos = byte offset of this
// segment's first byte.  Returns the end position.  If `flush_tail`, the trailing run is
// emitted too (last segment of the block, Run_Length_Encode_Slow.cpp:294).
template <class Get8, class Put>
CVX_HD int seg_emit(int ngroups, float mulfac, int carry_in, int pos, bool flush_tail, Get8 get8, Put put)
{
    int run = carry_in;
    for (int g = 0; g < ngroups; ++g) {
        float c[8]; float f[8]; int q[8];
        get8(g, c);
#pragma unroll
        for (int j = 0; j < 8; ++j) { f[j] = CVX_MUL(mulfac, c[j]); q[j] = quantise(f[j]); }
        int mode = group_mode(q);
        if (mode == GM_ZERO8) { run += 8; continue; }
        if (mode == GM_S8) {
            pos = put_run(run, pos, put); run = 0;
            put(pos++, (uint8_t)VLESC2_8x);
#pragma unroll
            for (int j = 0; j < 8; ++j) { put(pos++, (uint8_t)(q[j] & 0xFF)); put(pos++, (uint8_t)((q[j] >> 8) & 0xFF)); }
            continue;
        }
        if (mode == GM_T8) {
            pos = put_run(run, pos, put); run = 0;
            put(pos++, (uint8_t)VLESC3_8x);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                put(pos++, (uint8_t)(q[j] & 0xFF)); put(pos++, (uint8_t)((q[j] >> 8) & 0xFF));
                put(pos++, (uint8_t)((q[j] >> 16) & 0xFF));
            }
            continue;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            if (q[j] == 0) { ++run; continue; }
            pos = put_run(run, pos, put); run = 0;
            int v = q[j];
            if (is_b(v)) {
                put(pos++, (uint8_t)(v & 0xFF));
            } else if (is_s(v)) {
                put(pos++, (uint8_t)VLESC2); put(pos++, (uint8_t)(v & 0xFF)); put(pos++, (uint8_t)((v >> 8) & 0xFF));
            } else if (is_3(v)) {
                put(pos++, (uint8_t)VLESC3); put(pos++, (uint8_t)(v & 0xFF)); put(pos++, (uint8_t)((v >> 8) & 0xFF));
                put(pos++, (uint8_t)((v >> 16) & 0xFF));
            } else {   // VLESC4 carries the SCALED float, not q (:267-270)
#ifdef __CUDA_ARCH__
                uint32_t u = __float_as_uint(f[j]);
#else
                uint32_t u; memcpy(&u, &f[j], 4);
#endif
                put(pos++, (uint8_t)VLESC4); put(pos++, (uint8_t)(u & 0xFF)); put(pos++, (uint8_t)((u >> 8) & 0xFF));
                put(pos++, (uint8_t)((u >> 16) & 0xFF)); put(pos++, (uint8_t)(u >> 24));
            }
        }
    }
    if (flush_tail) pos = put_run(run, pos, put);
    return pos;
}

// ------------------------------------------------------------------------------------------
// Decoder pieces
// ------------------------------------------------------------------------------------------
// total length in bytes of the token whose first byte is b (Run_Length_Encode_Slow.cpp:415-513)
CVX_HD int token_len(uint32_t b)
{
    // 0x7D RLESC3 4, 0x7E VLESC3_8x 25, 0x7F RLESC1 2, 0x80 VLESC4 5, 0x81 VLESC3 4, 0x82 VLESC2_8x 17, 0x83 VLESC2 3, else 1:
    // an 8-entry byte table indexed by min(b - 0x7D, 7) (PRMT on the device)
    uint32_t i = b - 0x7Du;
    i = i < 7u ? i : 7u;
#ifdef __CUDA_ARCH__
    return (int)(__byte_perm(0x05021904u, 0x01031104u, i) & 0xFFu);
#else
    const uint64_t tbl = 0x0103110405021904ull;
    return (int)((tbl >> (8u * i)) & 0xFFu);
#endif
}

// Walk one 16-byte chunk.  d = bytes until the next token start (0 == a token starts at
// byte 0 of the chunk).  On return d is the same quantity for the next chunk; bit p of the
// result marks a token start at byte p.
CVX_HD uint32_t chunk_walk(const uint32_t w[4], int& d)
{
    uint32_t mask = 0;
#pragma unroll
    for (int p = 0; p < 16; ++p) {
        uint32_t b = (w[p >> 2] >> (8 * (p & 3))) & 0xFFu;
        int L = token_len(b);
        bool st = (d == 0);
        mask |= (st ? 1u : 0u) << p;
        d = st ? L : d;
        d -= 1;
    }
    return mask;
}

// values produced by the tokens that start in this chunk.  getb(k) returns payload byte k
// (absolute index in the staged window); base = index of this chunk's first byte.
template <class GetB>
CVX_HD int chunk_count(uint32_t mask, const uint32_t w[4], int base, GetB getb)
{
    int n = 0;
    while (mask) {
#ifdef __CUDA_ARCH__
        int p = __ffs(mask) - 1;
#else
        int p = __builtin_ctz(mask);
#endif
        mask &= mask - 1;
        int c = (int)(int8_t)((w[p >> 2] >> (8 * (p & 3))) & 0xFFu);
        if (c == RLESC1) n += (int)getb(base + p + 1);
        else if (c == RLESC3) n += (int)getb(base + p + 1) | ((int)getb(base + p + 2) << 8) | ((int)getb(base + p + 3) << 16);
        else if (c == VLESC2_8x || c == VLESC3_8x) n += 8;
        else n += 1;
    }
    return n;
}

// write the values of the tokens that start in this chunk; vi = raster index of the first
// value, num = block size.  Zeros of runs are NOT written (the block is pre-zeroed).
// put(i, val) stores coefficient i.  scalefac = 1.0f / mulfac (:392).
template <class GetB, class PutV>
CVX_HD void chunk_expand(uint32_t mask, const uint32_t w[4], int base, int vi, int num, float scalefac, GetB getb, PutV put)
{
    while (mask && vi < num) {
#ifdef __CUDA_ARCH__
        int p = __ffs(mask) - 1;
#else
        int p = __builtin_ctz(mask);
#endif
        mask &= mask - 1;
        int c = (int)(int8_t)((w[p >> 2] >> (8 * (p & 3))) & 0xFFu);
        int k = base + p + 1;
        if (c > VLESC2 && c < RLESC3) {
            if (c != 0) put(vi, CVX_MUL((float)c, scalefac));
            vi += 1;
        } else if (c == RLESC1) {
            vi += (int)getb(k);
        } else if (c == RLESC3) {
            vi += (int)getb(k) | ((int)getb(k + 1) << 8) | ((int)getb(k + 2) << 16);
        } else if (c == VLESC2) {
            int q = (int)(int16_t)(getb(k) | (getb(k + 1) << 8));
            put(vi, CVX_MUL((float)q, scalefac)); vi += 1;
        } else if (c == VLESC3) {
            int q = (int)((uint32_t)getb(k) << 8 | (uint32_t)getb(k + 1) << 16 | (uint32_t)getb(k + 2) << 24) >> 8;
            put(vi, CVX_MUL((float)q, scalefac)); vi += 1;
        } else if (c == VLESC4) {
            uint32_t u = (uint32_t)getb(k) | (uint32_t)getb(k + 1) << 8 | (uint32_t)getb(k + 2) << 16 | (uint32_t)getb(k + 3) << 24;
#ifdef __CUDA_ARCH__
            float f = __uint_as_float(u);
#else
            float f; memcpy(&f, &u, 4);
#endif
            put(vi, CVX_MUL(f, scalefac)); vi += 1;
        } else if (c == VLESC2_8x) {
            for (int j = 0; j < 8 && vi + j < num; ++j) {
                int q = (int)(int16_t)(getb(k + 2 * j) | (getb(k + 2 * j + 1) << 8));
                put(vi + j, CVX_MUL(scalefac, (float)q));
            }
            vi += 8;
        } else {   // VLESC3_8x
            for (int j = 0; j < 8 && vi + j < num; ++j) {
                int q = (int)((uint32_t)getb(k + 3 * j) << 8 | (uint32_t)getb(k + 3 * j + 1) << 16 | (uint32_t)getb(k + 3 * j + 2) << 24) >> 8;
                put(vi + j, CVX_MUL(scalefac, (float)q));
            }
            vi += 8;
        }
    }
}

// ------------------------------------------------------------------------------------------
// 8-byte chunk decoder pieces (k_decode32).  A = the chunk's 8 payload bytes (little endian),
// N = the 8 bytes that follow it (operands of tokens that start near the chunk's end).  Every byte
// position is a compile-time constant after unrolling, so operand extraction is a static shift.
// ------------------------------------------------------------------------------------------
// low 32 bits of the 128-bit value (N:A) >> 8*(p+1): the up-to-4 operand bytes of a token starting at byte p
CVX_HD uint32_t chunk8_operand(uint64_t A, uint64_t N, int p)
{
    const int s = 8 * (p + 1);
    if (s >= 64) return (uint32_t)N;
    uint64_t x = A >> s;
    if (s > 32) x |= N << (64 - s);
    return (uint32_t)x;
}

// Walk one chunk: d = bytes until the next token start on entry and on return; bit p of the result marks a
// token start at byte p.  COUNT: n += number of values produced by the tokens that start in this chunk.
template <bool COUNT>
CVX_HD uint32_t chunk8_walk(uint64_t A, uint64_t N, int& d, int& n)
{
    uint32_t mask = 0;
#pragma unroll
    for (int p = 0; p < 8; ++p) {
        const uint32_t b = (uint32_t)(A >> (8 * p)) & 0xFFu;
        const int L = token_len(b);
        const bool st = (d == 0);
        mask |= (st ? 1u : 0u) << p;
        if (COUNT) {
            const uint32_t op = chunk8_operand(A, N, p);
            int c = 1;
            c = (b == 0x7Fu) ? (int)(op & 0xFFu) : c;                 // RLESC1
            c = (b == 0x7Du) ? (int)(op & 0xFFFFFFu) : c;             // RLESC3
            c = (b == 0x7Eu || b == 0x82u) ? 8 : c;                   // VLESC3_8x, VLESC2_8x
            n += st ? c : 0;
        }
        d = st ? L : d;
        d -= 1;
    }
    return mask;
}

// the tokens that are neither a literal nor RLESC1: kept out of line so that the unrolled fast path stays small
// (the expansion is instruction-cache bound otherwise).  k = window index of the token's first operand byte.
#ifdef __CUDACC__
#define CVX_HD_NOINLINE __host__ __device__ __noinline__
#else
#define CVX_HD_NOINLINE
#endif
template <class GetB, class PutV>
CVX_HD_NOINLINE int chunk8_rare(int c, uint32_t op, int k, int vi, int num, float scalefac, GetB getb, PutV put)
{
    if (c == RLESC3) return (int)(op & 0xFFFFFFu);
    if (c == VLESC2_8x || c == VLESC3_8x) {
        for (int j = 0; j < 8 && vi + j < num; ++j) {
            int v;
            if (c == VLESC2_8x) v = (int)(int16_t)(getb(k + 2 * j) | (getb(k + 2 * j + 1) << 8));
            else v = (int)((uint32_t)getb(k + 3 * j) << 8 | (uint32_t)getb(k + 3 * j + 1) << 16 | (uint32_t)getb(k + 3 * j + 2) << 24) >> 8;
            put(vi + j, CVX_MUL(scalefac, (float)v));
        }
        return 8;
    }
    float fv;
    if (c == VLESC4) {
#ifdef __CUDA_ARCH__
        fv = __uint_as_float(op);
#else
        memcpy(&fv, &op, 4);
#endif
    } else {
        fv = (float)((c == VLESC2) ? (int)(int16_t)(op & 0xFFFFu) : ((int)(op << 8) >> 8));
    }
    if (vi < num) put(vi, CVX_MUL(fv, scalefac));
    return 1;
}

// Write the values of the tokens that start in this chunk (Run_Length_Encode_Slow.cpp:415-513).  vi = raster index
// of the first value; zeros are not written (pre-zeroed destination).  getb(k) = byte k of the staged window,
// base = window index of this chunk's first byte.  getb / put must be cheap to copy (they are passed by value).
template <class GetB, class PutV>
CVX_HD int chunk8_expand(uint32_t mask, uint64_t A, uint64_t N, int base, int vi, int num, float scalefac, GetB getb, PutV put)
{
#pragma unroll
    for (int p = 0; p < 8; ++p) {
        const bool st = ((mask >> p) & 1u) != 0;
        const uint32_t b = (uint32_t)(A >> (8 * p)) & 0xFFu;
        const int c = (int)(int8_t)b;
        const bool lit = (uint32_t)(c + 124) <= 248u;                 // -124 .. 124: one value, the byte itself
        const uint32_t op = chunk8_operand(A, N, p);
        // literal, VLESC2 and VLESC3 (the head of a block is dense in the latter two) without divergence
        const bool v2 = (c == VLESC2), v3 = (c == VLESC3);
        int q = c;
        q = v2 ? (int)(int16_t)(op & 0xFFFFu) : q;
        q = v3 ? ((int)(op << 8) >> 8) : q;
        const bool single = lit || v2 || v3;
        if (st && single && q != 0 && vi < num) put(vi, CVX_MUL((float)q, scalefac));
        int adv = (c == RLESC1) ? (int)(op & 0xFFu) : 1;
        if (st && !single && c != RLESC1) adv = chunk8_rare(c, op, base + p + 1, vi, num, scalefac, getb, put);
        vi += st ? adv : 0;
    }
    return vi;                            // raster index after the chunk's last token
}

}  // namespace cvx

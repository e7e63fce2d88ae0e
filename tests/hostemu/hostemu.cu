// Host emulation of the per-thread device phases in cvxcompress_b200/csrc/*.cuh.
// Compiled by nvcc for the HOST only (the same __host__ __device__ code the kernels inline),
// so the CPU test-suite can check the arithmetic and the segment/scan formulation against
// the oracle without a GPU.  Test infrastructure; not part of the product library.
#include <cstring>
#include <vector>
#include <algorithm>
#include "../../cvxcompress_b200/csrc/cvx_wavelet.cuh"
#include "../../cvxcompress_b200/csrc/cvx_rle.cuh"

using namespace cvx;

template <int N, bool EXACT>
static void fwd_line_sub(float* line)
{
    float v[N];
    for (int i = 0; i < N; ++i) v[i] = line[i];
    fwd_line<N, EXACT>(v);
    for (int r = 0; r < N; ++r) line[reg_to_sub<N, EXACT>(r)] = v[r];
}
template <int N, bool EXACT>
static void inv_line_sub(float* line)
{
    float v[N];
    for (int r = 0; r < N; ++r) v[r] = line[reg_to_sub<N, EXACT>(r)];
    inv_line<N, EXACT>(v);
    for (int i = 0; i < N; ++i) line[i] = v[i];
}

extern "C" int emu_line(float* line, int n, int exact, int inverse)
{
#define CASE(N)                                                                                  \
    case N:                                                                                      \
        if (exact) { if (inverse) inv_line_sub<N, true>(line); else fwd_line_sub<N, true>(line); } \
        else       { if (inverse) inv_line_sub<N, false>(line); else fwd_line_sub<N, false>(line); } \
        return 0;
    switch (n) { CASE(2) CASE(4) CASE(8) CASE(16) CASE(32) CASE(64) default: break; }
#undef CASE
    if (!exact) return -1;
    std::vector<float> t(n);
    if (inverse) inv_exact_mem(line, 1, t.data(), n); else fwd_exact_mem(line, 1, t.data(), n);
    return 0;
}

// dense block transform built from emu_line with the reference's pass order
extern "C" int emu_wavelet(float* b, int bx, int by, int bz, int exact, int inverse)
{
    std::vector<float> line(256);
    auto run = [&](long base, long stride, int n) {
        for (int i = 0; i < n; ++i) line[i] = b[base + i * stride];
        int rc = emu_line(line.data(), n, exact, inverse);
        for (int i = 0; i < n; ++i) b[base + i * stride] = line[i];
        return rc;
    };
    int rc = 0;
    for (int z = 0; z < bz; ++z) {
        for (int y = 0; y < by; ++y) rc |= run(((long)z * by + y) * bx, 1, bx);
        for (int x = 0; x < bx; ++x) rc |= run((long)z * by * bx + x, bx, by);
    }
    if (bz > 1)
        for (int y = 0; y < by; ++y)
            for (int x = 0; x < bx; ++x) rc |= run((long)y * bx + x, (long)bx * by, bz);
    return rc;
}

// encode with `nthreads` emulated threads, each owning num/8/nthreads consecutive groups
extern "C" int emu_rle_encode(float mulfac, const float* vals, int num, int nthreads, uint8_t* dst)
{
    int ng_total = num / 8;
    if (ng_total % nthreads) return -1;
    int ng = ng_total / nthreads, V = ng * 8;
    std::vector<SegSum> s(nthreads);
    for (int t = 0; t < nthreads; ++t) {
        const float* base = vals + (long)t * V;
        s[t] = seg_measure(ng, mulfac, [&](int g, float* c) { for (int j = 0; j < 8; ++j) c[j] = base[g * 8 + j]; });
    }
    // segmented scan: zeros carried into each thread
    std::vector<int> carry(nthreads + 1), off(nthreads + 1);
    carry[0] = 0;
    for (int t = 0; t < nthreads; ++t) carry[t + 1] = s[t].has ? s[t].trail : carry[t] + V;
    off[0] = 0;
    for (int t = 0; t < nthreads; ++t) {
        int nb = s[t].has ? run_bytes(carry[t] + s[t].lead) + s[t].inner : 0;
        if (t == nthreads - 1) nb += run_bytes(carry[t + 1]);
        off[t + 1] = off[t] + nb;
    }
    for (int t = 0; t < nthreads; ++t) {
        const float* base = vals + (long)t * V;
        if (!s[t].has && t != nthreads - 1) continue;
        int end = seg_emit(ng, mulfac, carry[t], off[t], t == nthreads - 1,
                           [&](int g, float* c) { for (int j = 0; j < 8; ++j) c[j] = base[g * 8 + j]; },
                           [&](int pos, uint8_t b) { dst[pos] = b; });
        if (end != off[t + 1]) return -2;
    }
    return off[nthreads];
}

// decode: `buf` is a 16-byte aligned buffer of buflen bytes, the block payload starts at
// byte `off`.  Emulates the windowed speculative parse with `nthreads` chunks per window.
// Returns the number of repair rounds (>= 0) or a negative error.
extern "C" int emu_rle_decode(float mulfac, float* vals, int num, const uint8_t* buf, long off, long buflen, int nthreads)
{
    float scalefac = 1.0f / mulfac;
    for (int i = 0; i < num; ++i) vals[i] = 0.0f;
    long c0 = off >> 4;                       // first chunk
    int entry = (int)(off & 15);              // bytes until the first token start
    int produced = 0, repairs = 0;
    auto getb = [&](long k) -> uint32_t { return (k >= 0 && k < buflen) ? buf[k] : 0u; };
    auto load = [&](long chunk, uint32_t w[4]) {
        for (int i = 0; i < 4; ++i) {
            uint32_t x = 0;
            for (int j = 0; j < 4; ++j) x |= getb(chunk * 16 + i * 4 + j) << (8 * j);
            w[i] = x;
        }
    };
    while (produced < num) {
        std::vector<int> ent(nthreads), ext(nthreads), cnt(nthreads);
        std::vector<uint32_t> mask(nthreads);
        std::vector<uint32_t> W(nthreads * 4);
        for (int t = 0; t < nthreads; ++t) {
            load(c0 + t, &W[t * 4]);
            int d;
            if (t == 0) d = entry;
            else { uint32_t pw[4]; load(c0 + t - 1, pw); d = 0; chunk_walk(pw, d); }   // speculative entry
            ent[t] = d;
            mask[t] = chunk_walk(&W[t * 4], d);
            ext[t] = d;
        }
        for (;;) {   // verify the chain, repair mismatches
            bool bad = false;
            std::vector<int> fix(nthreads, -1);
            for (int t = 1; t < nthreads; ++t) if (ent[t] != ext[t - 1]) { fix[t] = ext[t - 1]; bad = true; }
            if (!bad) break;
            ++repairs;
            for (int t = 1; t < nthreads; ++t) if (fix[t] >= 0) { int d = fix[t]; ent[t] = d; mask[t] = chunk_walk(&W[t * 4], d); ext[t] = d; }
        }
        int vi = produced;
        for (int t = 0; t < nthreads; ++t) {
            long base = (c0 + t) * 16;
            int n = chunk_count(mask[t], &W[t * 4], 0, [&](int k) { return getb(base + k); });
            chunk_expand(mask[t], &W[t * 4], 0, vi, num, scalefac, [&](int k) { return getb(base + k); },
                         [&](int i, float v) { vals[i] = v; });
            vi = std::min(vi + std::min(n, num), num);
        }
        produced = vi;
        entry = ext[nthreads - 1];
        c0 += nthreads;
        if (c0 * 16 > buflen + 64 && produced < num) return -3;   // truncated stream
    }
    return repairs;
}

// decode with the 8-byte chunk formulation of k_decode32: windows of `nch_max` chunks, speculation from the two
// previous chunks, chain verification/repair, clamped counts, exclusive scan, expansion.  `hint` = payload bytes
// of this block when known (limits the first window), 0 = unknown.  Returns repair rounds or a negative error.
extern "C" int emu_rle_decode8(float mulfac, float* vals, int num, const uint8_t* buf, long off, long buflen, int nch_max, long hint)
{
    float scalefac = 1.0f / mulfac;
    for (int i = 0; i < num; ++i) vals[i] = 0.0f;
    long cb = off & ~15L;                     // window base (16-byte aligned, as the cp.async staging needs)
    int entry = (int)(off - cb);
    int produced = 0, repairs = 0;
    auto getb = [&](long k) -> uint32_t { return (k >= 0 && k < buflen) ? buf[k] : 0u; };
    auto load64 = [&](long byte0) { uint64_t x = 0; for (int j = 0; j < 8; ++j) x |= (uint64_t)getb(byte0 + j) << (8 * j); return x; };
    while (produced < num) {
        int nch = nch_max;
        if (hint > 0) { long want = (entry + hint + 7) >> 3; want += want & 1; if (want < nch) nch = (int)want; }
        std::vector<int> ent(nch), ext(nch), cnt(nch);
        std::vector<uint32_t> mask(nch);
        std::vector<uint64_t> A(nch), N(nch);
        for (int c = 0; c < nch; ++c) {
            A[c] = load64(cb + 8L * c); N[c] = load64(cb + 8L * c + 8);
            int d, dummy = 0;
            if (c == 0) d = entry;
            else if (c == 1) { d = entry; chunk8_walk<false>(load64(cb), 0, d, dummy); }
            else { d = 0; chunk8_walk<false>(load64(cb + 8L * (c - 2)), 0, d, dummy); chunk8_walk<false>(load64(cb + 8L * (c - 1)), 0, d, dummy); }
            ent[c] = d; cnt[c] = 0;
            mask[c] = chunk8_walk<true>(A[c], N[c], d, cnt[c]);
            ext[c] = d;
        }
        for (;;) {
            bool bad = false;
            std::vector<int> fix(nch, -1);
            for (int c = 1; c < nch; ++c) if (ent[c] != ext[c - 1]) { fix[c] = ext[c - 1]; bad = true; }
            if (!bad) break;
            ++repairs;
            for (int c = 1; c < nch; ++c) if (fix[c] >= 0) { int d = fix[c]; ent[c] = d; cnt[c] = 0; mask[c] = chunk8_walk<true>(A[c], N[c], d, cnt[c]); ext[c] = d; }
        }
        long vi = produced;
        for (int c = 0; c < nch; ++c) {
            int n = cnt[c] > (1 << 20) ? (1 << 20) : cnt[c];
            if (vi < num)
                chunk8_expand(mask[c], A[c], N[c], 8 * c, (int)vi, num, scalefac, [&](int k) { return getb(cb + k); },
                              [&](int i, float v) { vals[i] = v; });
            vi += n;
        }
        produced = vi > (1 << 30) ? (1 << 30) : (int)vi;
        entry = ext[nch - 1];
        cb += 8L * nch;
        hint = 0;
        if (cb > buflen + 64 && produced < num) return -3;
    }
    return repairs;
}

extern "C" int emu_token_len(int b) { return token_len((uint32_t)(b & 0xFF)); }
extern "C" int emu_quantise(float f) { return quantise(f); }

// The 32^3 LIFTING encoders' arithmetic (cvx_fused32.cuh, "LIFTING encoders"): passes in the order z, y, x, bare lifting steps on
// the z and y lines, and the x rows' multiply carrying the scalings of all three axes (fwd_lift_pair_rs with the row factors of
// f32_row_factors).  In: a dense 32^3 block, natural order.  Out: coefficients in sub-band order (what the oracle's forward
// transform leaves), so the caller can compare the two within the LIFTING tolerance.
extern "C" int emu_encode32_lifting_zyx(float* b)
{
    constexpr int N = 32;
    auto at = [&](int x, int y, int z) -> float& { return b[((long)z * N + y) * N + x]; };
    float v[N];
    // z lines, then y lines: bare lifting steps, results stored by sub-band index (reg_to_sub) as the kernels do
    for (int y = 0; y < N; ++y)
        for (int x = 0; x < N; ++x) {
            for (int z = 0; z < N; ++z) v[z] = at(x, y, z);
            fwd_lift_levels<N, 1>(v);
            for (int r = 0; r < N; ++r) at(x, y, reg_to_sub<N, false>(r)) = v[r];
        }
    for (int z = 0; z < N; ++z)
        for (int x = 0; x < N; ++x) {
            for (int y = 0; y < N; ++y) v[y] = at(x, y, z);
            fwd_lift_levels<N, 1>(v);
            for (int r = 0; r < N; ++r) at(x, reg_to_sub<N, false>(r), z) = v[r];
        }
    // x rows in pairs (row (y, z) and row (y, z + 16), as one thread holds them), scalings of y and z folded in
    for (int z = 0; z < 16; ++z)
        for (int y = 0; y < N; ++y) {
            float a[N], c[N], oa[N], oc[N];
            for (int x = 0; x < N; ++x) { a[x] = at(x, y, z); c[x] = at(x, y, z + 16); }
            const float sy = lift_scale_of_sub<N, true>(y);
            const float fa = sy * lift_scale_of_sub<N, true>(z), fc = sy * lift_scale_of_sub<N, true>(z + 16);
            fwd_lift_pair_rs<N>(a, c, oa, oc, fa, fc);
            for (int x = 0; x < N; ++x) { at(x, y, z) = oa[sub_to_reg<N, false>(x)]; at(x, y, z + 16) = oc[sub_to_reg<N, false>(x)]; }
        }
    return 0;
}

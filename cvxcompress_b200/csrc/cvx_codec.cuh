// cvx_codec.cuh -- CTA-wide quantise + run-length encode / decode built on the per-thread
// phases of cvx_rle.cuh (whose byte-exactness against Run_Length_Encode_Slow.cpp:189-294 and
// :388-527 is checked on the CPU by tests/test_hostemu.py).
#pragma once
#include "cvx_device.cuh"

namespace cvx {

// ------------------------------------------------------------------------------------------
// Encode, phase 1: measure every segment, scan, and leave per segment
//     s_seg[seg] = { zeros carried in, byte offset inside the block payload, has, - }.
// A segment is `gps` consecutive groups of 8 coefficients in raster order; get8(g, c) loads
// group g (block-wide group index).  Returns the block's encoded length in bytes.
// scratch: >= 66 ints.
// ------------------------------------------------------------------------------------------
// `measured`: s_seg already holds every segment's { has, lead, trail, inner } (written by another kernel: the large-block
// path measures the segments of a whole batch grid-wide); s_seg may then be global memory.
template <int T, class Get8>
__device__ __forceinline__ int cta_rle_plan(int nseg, int gps, float mulfac, Get8 get8, int4* s_seg, int* scratch, const bool measured = false)
{
    const int tid = threadIdx.x;
    if (!measured) {
        for (int seg = tid; seg < nseg; seg += T) {
            const int g0 = seg * gps;
            SegSum s = seg_measure(gps, mulfac, [&](int g, float* c) { get8(g0 + g, c); });
            s_seg[seg] = make_int4(s.has, s.lead, s.trail, s.inner);
        }
    }
    __syncthreads();
    const int V = gps * 8;
    const int K = (nseg + T - 1) / T;
    const int s0 = min(nseg, tid * K), s1 = min(nseg, s0 + K);
    Carry agg; agg.has = 0; agg.run = 0;
    for (int s = s0; s < s1; ++s) {
        int4 q = s_seg[s];
        Carry c; c.has = q.x; c.run = q.x ? q.z : V;
        agg = carry_combine(agg, c);
    }
    Carry total;
    Carry pre = cta_excl_carry<T>(agg, scratch, total);
    int bytes = 0;
    Carry c = pre;
    for (int s = s0; s < s1; ++s) {
        int4 q = s_seg[s];
        int cin = c.run;
        int b = q.x ? run_bytes(cin + q.y) + q.w : 0;
        if (s == nseg - 1) b += run_bytes(total.run);         // final flush, Run_Length_Encode_Slow.cpp:294
        Carry me; me.has = q.x; me.run = q.x ? q.z : V;
        c = carry_combine(c, me);
        s_seg[s] = make_int4(cin, b, q.x, 0);
        bytes += b;
    }
    int len;
    int off = cta_excl_sum<T>(bytes, scratch, len);
    for (int s = s0; s < s1; ++s) {
        int4 q = s_seg[s];
        int b = q.y;
        q.y = off;
        s_seg[s] = q;
        off += b;
    }
    __syncthreads();
    return len;
}

// Encode, phase 2: write the bytes.  put(pos, byte) stores payload byte `pos` of this block.
template <int T, class Get8, class Put>
__device__ __forceinline__ void cta_rle_emit(int nseg, int gps, float mulfac, Get8 get8, const int4* s_seg, Put put)
{
    for (int seg = threadIdx.x; seg < nseg; seg += T) {
        int4 q = s_seg[seg];
        const bool last = (seg == nseg - 1);
        if (!q.z && !last) continue;
        const int g0 = seg * gps;
        seg_emit(gps, mulfac, q.x, q.y, last, [&](int g, float* c) { get8(g0 + g, c); }, put);
    }
}

// ------------------------------------------------------------------------------------------
// Decode one block's payload with the whole CTA.  The payload has no length field, so it is
// consumed in windows of T 16-byte chunks until `num` values exist.  [lo, hi) is the byte range
// of the stream that may be read; bytes outside read as 0.  The destination must be pre-zeroed
// (runs write nothing).  put(i, v) stores coefficient i.
// s_win: (T + 2) * 4 words; s_ext: T ints; scratch: >= 33 ints.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 load_chunk_guarded(const uint8_t* a, const uint8_t* lo, const uint8_t* hi)
{
    if (a >= lo && a + 16 <= hi) return *reinterpret_cast<const uint4*>(a);
    uint32_t w[4] = {0u, 0u, 0u, 0u};
    if (a + 16 > lo && a < hi) {
        for (int k = 0; k < 16; ++k) {
            const uint8_t* p = a + k;
            if (p >= lo && p < hi) w[k >> 2] |= (uint32_t)(*p) << (8 * (k & 3));
        }
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
}

template <int T, class PutV>
__device__ __forceinline__ void cta_rle_decode(const uint8_t* lo, const uint8_t* hi, const uint8_t* src, int num,
                                               float scalefac, PutV put, uint32_t* s_win, int* s_ext, int* scratch)
{
    const int tid = threadIdx.x;
    const uint8_t* cbase = reinterpret_cast<const uint8_t*>(reinterpret_cast<uintptr_t>(src) & ~(uintptr_t)15);
    int entry = (int)(src - cbase);       // bytes until the first token start
    int produced = 0;
    while (produced < num) {
        uint4 c = load_chunk_guarded(cbase + 16 * (long)tid, lo, hi);
        reinterpret_cast<uint4*>(s_win)[tid] = c;
        if (tid < 2) reinterpret_cast<uint4*>(s_win)[T + tid] = load_chunk_guarded(cbase + 16 * (long)(T + tid), lo, hi);
        __syncthreads();
        uint32_t w[4] = {c.x, c.y, c.z, c.w};
        int d;
        if (tid == 0) {
            d = entry;
        } else {                              // speculate: a token starts at byte 0 of the previous chunk
            uint4 pc = reinterpret_cast<const uint4*>(s_win)[tid - 1];
            uint32_t pw[4] = {pc.x, pc.y, pc.z, pc.w};
            d = 0;
            chunk_walk(pw, d);
        }
        int ent = d;
        uint32_t mask = chunk_walk(w, d);
        int ext = d;
        for (;;) {                            // verify the chain exit(t-1) == entry(t); repair until consistent
            s_ext[tid] = ext;
            __syncthreads();
            int want = tid > 0 ? s_ext[tid - 1] : ent;
            bool bad = (want != ent);
            int any = __syncthreads_or(bad ? 1 : 0);
            if (!any) break;
            if (bad) { d = want; ent = want; mask = chunk_walk(w, d); ext = d; }
        }
        auto getb = [&](int k) -> uint32_t { return (s_win[k >> 2] >> (8 * (k & 3))) & 0xFFu; };
        const int base = tid * 16;
        int n = chunk_count(mask, w, base, getb);
        n = n > (1 << 26) ? (1 << 26) : n;
        int total;
        int vi = cta_excl_sum_sat<T>(n, scratch, total);
        vi = sat_add(vi, produced);
        if (vi < num) chunk_expand(mask, w, base, vi, num, scalefac, getb, put);
        entry = s_ext[T - 1];
        produced = sat_add(produced, total);
        cbase += 16 * (long)T;
        __syncthreads();
        if (cbase >= hi + 64) break;          // truncated stream: stop instead of spinning
    }
}

}  // namespace cvx

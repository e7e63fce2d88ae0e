// cvx_decode32.inl -- k_decode32, included by cvx_kernels.cu inside namespace cvx after DecParams.
//
// One persistent CTA of 512 threads per SM; per 32^3 block:
//   payload window (8 KiB, cp.async-prefetched while the PREVIOUS block is transformed)
//   -> speculative token parse over 8-byte chunks, two chunks per thread (Run_Length_Decode_Slow,
//      Run_Length_Encode_Slow.cpp:388-527): walk, verify/repair the chain, count, scan, expand into the
//      pre-zeroed shared-memory block (TMA SWIZZLE_128B layout, cvx_device.cuh)
//   -> inverse x lines, inverse y lines in shared memory (Wavelet_Transform_Fast.cpp:148-279 order)
//   -> inverse z lines in REGISTERS; the block buffer is zeroed for the next block at once and the
//      results go straight to the volume as 128-byte row stores (Copy_From_Block clipping,
//      Block_Copy.cpp:136-212) -- no write-back to shared memory, no alignment requirement on the volume.
// A block's payload length is not stored in the stream; when the next table entry is larger (natural block
// order) the difference bounds the first window, otherwise the full window is parsed.  Either way parsing
// continues window by window until 32768 values exist, so any block order decodes.

constexpr int D32_T = 512;
constexpr int D32_CH = 1024;                          // 8-byte chunks per window, two per thread
constexpr int D32_MINWIN = 4080;                      // bytes of a first window parsed regardless of the table's length hint (+ <= 15 of misalignment: 512 chunks)
constexpr int D32_WIN = D32_CH * 8 + 32;              // + look-ahead for tokens straddling the window end
constexpr int D32_OFF_WIN = F32_B * 4;                // two windows
constexpr int D32_OFF_EXT = D32_OFF_WIN + 2 * D32_WIN;
constexpr int D32_SMEM = D32_OFF_EXT + D32_CH * 4;    // 151 648 bytes

__device__ __forceinline__ void cp_async16_zfill(void* smem_dst, const void* gsrc, int src_bytes)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// stage the window [cbase, cbase + D32_WIN) (cbase 16-byte aligned); bytes at or beyond `hi` read as zero
__device__ __forceinline__ void d32_issue_window(unsigned char* win, const uint8_t* cbase, const uint8_t* hi, int nthreads, int win_bytes = D32_WIN)
{
    for (int g = threadIdx.x; g < win_bytes / 16; g += nthreads) {
        const uint8_t* a = cbase + 16 * g;
        long left = hi - a;
        int n = left >= 16 ? 16 : (left > 0 ? (int)left : 0);
        cp_async16_zfill(win + 16 * g, n > 0 ? a : cbase, n);
    }
    cp_async_commit();
}

// exclusive sum of a pair of 32-bit counts packed in 64 bits (each count <= 2^20, 512 threads: no carry between halves)
template <int T, class Bar>
__device__ __forceinline__ unsigned long long d32_excl_sum64(unsigned long long v, unsigned long long* scratch64, unsigned long long& total)
{
    static_assert((T & (T - 1)) == 0, "group size is a power of two");
    const int gw = warp_id() & (T / 32 - 1);          // warp index inside the group (groups start at multiples of T/32 warps)
    unsigned long long x = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        unsigned long long y = __shfl_up_sync(0xffffffffu, x, d);
        if (lane_id() >= d) x += y;
    }
    if (lane_id() == 31) scratch64[gw] = x;
    Bar::sync();
    unsigned long long w = lane_id() < T / 32 ? scratch64[lane_id()] : 0ull;
    unsigned long long s = w;
#pragma unroll
    for (int d = 1; d < T / 32; d <<= 1) {
        unsigned long long y = __shfl_up_sync(0xffffffffu, s, d);
        if (lane_id() >= d) s += y;
    }
    total = __shfl_sync(0xffffffffu, s, T / 32 - 1);
    const unsigned long long wbase = __shfl_sync(0xffffffffu, s - w, gw);
    Bar::sync();
    return wbase + x - v;
}

struct D32Src {
    const uint8_t* src;     // first payload byte of the block
    long hint;              // payload bytes of the block when the table tells, else 0
    int raw;
    float scalefac;
};

// table entry of block t -> where its payload is (CvxCompress.cpp:545-548) and, when the next entry is larger, how long.
// Two steps so that a caller can issue the table loads a whole block ahead of their use (k_decode32ws parser).
struct D32Raw {
    unsigned long long o, o1;
};
__device__ __forceinline__ D32Raw d32_locate_issue(const DecParams& p, long t)
{
    D32Raw r;
    r.o = 0; r.o1 = 0;
    // volatile: issued where written, so the latency hides under whatever follows instead of in front of the first use
    if (t < p.nblk) asm volatile("ld.global.nc.u64 %0, [%1];" : "=l"(r.o) : "l"(p.offs + t));
    if (t + 1 < p.nblk) asm volatile("ld.global.nc.u64 %0, [%1];" : "=l"(r.o1) : "l"(p.offs + t + 1));
    return r;
}
// maxch: 8-byte chunks of a window (two per thread of the parsing group; D32_CH for the 512-thread kernels)
__device__ __forceinline__ D32Src d32_locate_finish(const DecParams& p, long t, const D32Raw& w, int maxch = D32_CH)
{
    D32Src r;
    unsigned long long o = w.o;
    r.raw = (int)(o >> 63);
    o &= 0x7FFFFFFFFFFFFFFFull;
    r.src = p.payload + o;
    long nxt = (t + 1 < p.nblk) ? (long)(w.o1 & 0x7FFFFFFFFFFFFFFFull) : p.payload_bytes;
    // The table tells a block's payload length only when the blocks lie in natural order; a stream written by the
    // reference with several threads has them permuted (CvxCompress.cpp:365-390) and the difference to the next entry is
    // then meaningless (negative, tiny or huge).  So a first window of at least D32_MINWIN bytes is parsed whatever the
    // table says: that is the same one (k_decode32) or two (k_decode32ws parser) chunks per thread a typical ~3 KiB
    // payload takes anyway, and it serves a block of any order; a payload that turns out longer continues window by window.
    const long h = nxt - (long)o;
    const long minwin = D32_MINWIN < maxch * 8 - 16 ? D32_MINWIN : maxch * 8 - 16;
    r.hint = (h > minwin && h <= maxch * 8 - 16) ? h : minwin;
    const float mulfac = p.use_local ? p.blkmulfac[t] : p.glob_mulfac;
    r.scalefac = __fdiv_rn(1.0f, mulfac);                        // Run_Length_Encode_Slow.cpp:392
    return r;
}
__device__ __forceinline__ D32Src d32_locate(const DecParams& p, long t, int maxch = D32_CH)
{
    return d32_locate_finish(p, t, d32_locate_issue(p, t), maxch);
}

// per-thread result of walking its two chunks of a window
struct D32Walk {
    uint64_t A[2], N[2];
    uint32_t mask[2];
    int ent[2], ext[2], cnt[2];
};

__device__ __forceinline__ int d32_nchunks(int entry, long hint, int maxch = D32_CH)
{
    int nch = maxch;
    if (hint > 0) {
        long want = (entry + hint + 7) >> 3;
        want += want & 1;                 // even: the next window base stays 16-byte aligned
        if (want < nch) nch = (int)want;
    }
    return nch;
}

// Speculative walk of this thread's chunks c = tid, tid + 512 of the staged window: entry state guessed from the two
// chunks before it, then the chunk itself with value counting.  Straight-line code (selects, clamped loads, no
// branches) so that the compiler can interleave it with the packed z-line arithmetic of the PREVIOUS block, whose
// FFMA2/FADD2 stream leaves half of the issue slots free.
template <int T>
__device__ __forceinline__ void d32_walk_one(const uint64_t* w64, int entry, int nch, int j, D32Walk& w)
{
    const int c = threadIdx.x + T * j;
    const bool act = c < nch;
    const uint64_t P2 = w64[c >= 2 ? c - 2 : 0], P1 = w64[c >= 1 ? c - 1 : 0];
    w.A[j] = w64[c]; w.N[j] = w64[c + 1];
    int d = 0, dummy = 0;
    chunk8_walk<false>(P2, 0ull, d, dummy);
    d = (c == 1) ? entry : d;             // chunk 1: exact entry through chunk 0
    chunk8_walk<false>(P1, 0ull, d, dummy);
    d = (c == 0) ? entry : d;
    w.ent[j] = d;
    int cnt = 0;
    const uint32_t m = chunk8_walk<true>(w.A[j], w.N[j], d, cnt);
    w.mask[j] = act ? m : 0u;
    w.cnt[j] = act ? cnt : 0;
    w.ext[j] = act ? d : 0;
}
template <int T>
__device__ __forceinline__ void d32_walk(const unsigned char* win, int entry, int nch, D32Walk& w)
{
    const uint64_t* w64 = reinterpret_cast<const uint64_t*>(win);
    // chunks 512..1023 only exist for payloads above 4 KiB (or an unknown length): CTA-uniform branch, placed first so
    // that the always-executed half below stays in one basic block with the arithmetic that follows the call
    w.A[1] = 0; w.N[1] = 0; w.mask[1] = 0; w.cnt[1] = 0; w.ext[1] = 0; w.ent[1] = 0;
    if constexpr (T < D32_CH) { if (nch > T) d32_walk_one<T>(w64, entry, nch, 1, w); }
    d32_walk_one<T>(w64, entry, nch, 0, w);
}

// parse the block whose first window is staged in `win` and already walked (`w`); writes coefficients through put(i, v)
// num: cells of the block; maxch: chunks of a window (<= 2 T)
template <int T, class Bar, class PutV>
__device__ __forceinline__ void d32_parse(const uint8_t* hi, const D32Src& b, unsigned char* win, int* s_ext,
                                          unsigned long long* scratch64, D32Walk& w, PutV put, const int num = F32_B, const int maxch = D32_CH)
{
    const int tid = threadIdx.x;
    const uint8_t* cbase = reinterpret_cast<const uint8_t*>(reinterpret_cast<uintptr_t>(b.src) & ~(uintptr_t)15);
    int entry = (int)(b.src - cbase);
    long hint = b.hint;
    int produced = 0;
    auto getb = [win](int k) -> uint32_t { return (uint32_t)win[k]; };
    for (bool first = true;; first = false) {
        const int nch = d32_nchunks(entry, hint, maxch);
        if (!first) {                     // rare: payload longer than one window, or a wrong hint
            Bar::sync();
            d32_issue_window(win, cbase, hi, T, maxch * 8 + 32);
            cp_async_wait_all();
            Bar::sync();
            d32_walk<T>(win, entry, nch, w);
        }
        for (;;) {                        // verify exit(c-1) == entry(c); repair until the chain is consistent
#pragma unroll
            for (int j = 0; j < 2; ++j) { const int c = tid + T * j; if (c < nch) s_ext[c] = w.ext[j]; }
            Bar::sync();
            bool bad[2];
            int want[2];
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int c = tid + T * j;
                want[j] = (c > 0 && c < nch) ? s_ext[c - 1] : w.ent[j];
                bad[j] = want[j] != w.ent[j];
            }
            const int last_ext = s_ext[nch - 1];
            const int any = Bar::sync_or((bad[0] || bad[1]) ? 1 : 0);
#ifdef CVX_PHASE_TIMING
            if (tid == 0) atomicAdd(&g_phase[1][10], 1ull);
            if (bad[0] || bad[1]) atomicAdd(&g_phase[1][11], 1ull);
#endif
            if (!any) { entry = last_ext; break; }
#pragma unroll
            for (int j = 0; j < 2; ++j)
                if (bad[j]) { int d = want[j]; w.ent[j] = d; w.cnt[j] = 0; w.mask[j] = chunk8_walk<true>(w.A[j], w.N[j], d, w.cnt[j]); w.ext[j] = d; }
        }
        const unsigned c0 = (unsigned)min(w.cnt[0], 1 << 20), c1 = (unsigned)min(w.cnt[1], 1 << 20);
        unsigned long long total;
        const unsigned long long ex = d32_excl_sum64<T, Bar>((unsigned long long)c0 | ((unsigned long long)c1 << 32), scratch64, total);
        const long tot0 = (long)(total & 0xFFFFFFFFull), tot1 = (long)(total >> 32);
        long vi0 = (long)produced + (long)(ex & 0xFFFFFFFFull);
        long vi1 = (long)produced + tot0 + (long)(ex >> 32);
        if (w.mask[0] && vi0 < num) chunk8_expand(w.mask[0], w.A[0], w.N[0], 8 * tid, (int)vi0, num, b.scalefac, getb, put);
        if (w.mask[1] && vi1 < num) chunk8_expand(w.mask[1], w.A[1], w.N[1], 8 * (tid + T), (int)vi1, num, b.scalefac, getb, put);
        long np = (long)produced + tot0 + tot1;
        produced = np > (1 << 30) ? (1 << 30) : (int)np;
        if (produced >= num) break;
        cbase += 8 * nch;
        hint = 0;
        if (cbase >= hi + 64) break;      // truncated stream: stop instead of spinning
    }
}

template <bool EXACT, int T>
__global__ void __launch_bounds__(T, 1) k_decode32(const DecParams pin)
{
    const DecParams p = dec_resolve(pin);
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    float* sm = reinterpret_cast<float*>(smem_raw);
    unsigned char* s_win = smem_raw + D32_OFF_WIN;
    int* s_ext = reinterpret_cast<int*>(smem_raw + D32_OFF_EXT);
    __shared__ unsigned long long scratch64[T / 32];
    constexpr int LPT = 1024 / T;                                   // z lines (and x rows, y lines) per thread

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint8_t* hi = p.payload + p.payload_bytes;
    const long plane = (long)p.g.nx * p.g.ny;

    long t = blockIdx.x;
    if (t >= p.nblk) return;
    D32Src cur = d32_locate(p, t);
    int buf = 0;
    d32_issue_window(s_win, reinterpret_cast<const uint8_t*>(reinterpret_cast<uintptr_t>(cur.src) & ~(uintptr_t)15), hi, T);
    {
        float4* z4 = reinterpret_cast<float4*>(sm);
#pragma unroll
        for (int i = 0; i < F32_B / 4 / T; ++i) z4[tid + T * i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    D32Walk wk;
    cp_async_wait_all();
    __syncthreads();                                                // first window staged, block buffer zeroed
    {
        const int entry = (int)(reinterpret_cast<uintptr_t>(cur.src) & 15);
        d32_walk<T>(s_win, entry, d32_nchunks(entry, cur.hint), wk);
    }
    CVX_PH_DECL;
    for (;;) {
        const long tn = t + gridDim.x;
        D32Src nxt = cur;
        if (tn < p.nblk) nxt = d32_locate(p, tn);                   // table reads overlap the parse below
        CVX_PH(0);
        if (cur.raw) {                                              // CvxCompress.cpp:555: B fp32 coefficients
            __syncthreads();                                        // every thread has zeroed its cells of the previous block
            for (int i = tid; i < F32_B; i += T) {
                uint32_t u = 0;
                const uint8_t* s = cur.src + 4L * i;
                if ((reinterpret_cast<uintptr_t>(cur.src) & 3) == 0 && s + 4 <= hi) u = *reinterpret_cast<const uint32_t*>(s);
                else for (int k = 0; k < 4; ++k) if (s + k < hi) u |= (uint32_t)s[k] << (8 * k);
                sm[physr(i, 5)] = __uint_as_float(u);
            }
        } else {
            d32_parse<T, CtaBar>(hi, cur, s_win + buf * D32_WIN, s_ext, scratch64, wk, [sm](int i, float v) { sm[physr(i, 5)] = v; });
        }
        __syncthreads();
        CVX_PH(1);
        if (tn < p.nblk)                                            // next block's payload streams in under the transform
            d32_issue_window(s_win + (buf ^ 1) * D32_WIN,
                             reinterpret_cast<const uint8_t*>(reinterpret_cast<uintptr_t>(nxt.src) & ~(uintptr_t)15), hi, T);
        // warp w owns the x rows AND the y lines of the z slices w and w + 16: no CTA barrier between the two passes,
        // so warps drift apart and one warp's shared-memory traffic overlaps another's arithmetic
        f32_pass_xT<EXACT, false, T>(sm);
        __syncwarp();
        CVX_PH(2);
        f32_pass_yT<EXACT, false, T>(sm);
        cp_async_wait_all();                                        // this thread's part of the next window has landed ...
        __syncthreads();                                            // ... and everybody's; all y lines are written
        CVX_PH(3);
        typename LineOf<EXACT, T>::type v[32];                      // z line(s): (x = lane, y = warp) [and (x, y + 16), packed]
        {
            // each thread reads its two z lines and leaves zeros behind: the buffer is clean for the next block's
            // expansion without another barrier or a separate fill pass
            float* zl = f32_zline(sm, tid);
#pragma unroll
            for (int r = 0; r < 32; ++r) {
                float* e = zl + reg_to_sub<32, EXACT>(r) * 1024;
                p_set(v[r], e[0], e[512]);
            }
#pragma unroll
            for (int r = 0; r < 32; ++r) { zl[r * 1024] = 0.f; if (LPT == 2) zl[r * 1024 + 512] = 0.f; }
        }
        CVX_PH(4);
        const int iix = (int)(t % p.nbx);
        const long q = t / p.nbx;
        const int iiy = (int)(q % p.nby), iiz = (int)(q / p.nby);
        const int gx = iix * 32 + lane, gz0 = iiz * 32;
        const int nzl = min(32, p.g.nz - gz0);
        {
            // walk the NEXT block's chunks here: integer work that fills the issue slots the packed z-line arithmetic leaves
            // free (no branch in between, so both streams are scheduled together)
            const int entry = (int)(reinterpret_cast<uintptr_t>(nxt.src) & 15);
            d32_walk<T>(s_win + (buf ^ 1) * D32_WIN, entry, d32_nchunks(entry, nxt.hint), wk);
        }
        inv_line<32, EXACT>(v);
#pragma unroll
        for (int k = 0; k < LPT; ++k) {
            const int gy = iiy * 32 + warp + 16 * k;
            if (gx < p.g.nx && gy < p.g.ny) {
                float* dp = p.vol + ((long)gz0 * p.g.ny + gy) * p.g.nx + gx;
                if (nzl == 32) {                                    // interior block: 32 unconditional row stores
#pragma unroll
#ifdef CVX_EXP_NOSTORE
                    for (int j = 0; j < 32; ++j) { const float o = k ? p_hi(v[j]) : p_lo(v[j]); if (o != o) *dp = o; dp += plane; }
#else
                    for (int j = 0; j < 32; ++j) { *dp = k ? p_hi(v[j]) : p_lo(v[j]); dp += plane; }
#endif
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) { if (j < nzl) *dp = k ? p_hi(v[j]) : p_lo(v[j]); dp += plane; }
                }
            }
        }
        CVX_PH(5);
        if (tn >= p.nblk) break;
        t = tn;
        cur = nxt;
        buf ^= 1;
    }
    CVX_PH_FLUSH(1);
}

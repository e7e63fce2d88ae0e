// cvx_decode32ws.inl -- k_decode32ws: warp-specialised 32^3 decoder (LIFTING arithmetic), included after cvx_decode32.inl.
//
// The fused decoder's phases each saturate ONE pipe -- the token parse the (half-rate) integer pipe, the line
// transforms the FMA pipe, the row stores the load/store unit -- and in k_decode32 they run one after the other.
// Here a CTA has 24 warps: warps 0..15 (the transform group) run expand -> x -> y -> z -> store of block t while
// warps 16..23 (the parser group) turn the next blocks' payloads (up to two ahead: two list buffers) into compact (raster index, value) lists, so the
// integer work of one block overlaps the floating-point work of its predecessor.
//   parser:    payload window (cp.async) -> contiguous chunk ranges per thread (speculation only at the range start,
//              exact state inside it) -> chain check over 128 threads -> value / entry counts -> scan -> list
//   transform: scatter the list into its zeroed block buffer, then exactly the k_decode32 pipeline
// Hand-over by two named barriers (FULL: list ready, EMPTY: list consumed).  A block the parser cannot serve (raw,
// unknown payload length = foreign block order, payload or list larger than the staging) is flagged and decoded
// in line by the transform group with the k_decode32 routines on a named barrier.

constexpr int W32_TT = 512;                          // transform threads (warps 0..15)
constexpr int W32_PT = 256;                          // parser threads (warps 16..19)
constexpr int W32_T = W32_TT + W32_PT;
constexpr int W32_LCAP = 4096;                       // entries (8 bytes each) of each of the TWO lists: the parser may run two blocks ahead
constexpr int W32_OFF_WIN = F32_B * 4;               // two parser windows, then the in-line window
constexpr int W32_OFF_EXT = W32_OFF_WIN + 3 * D32_WIN;
constexpr int W32_OFF_MASK = W32_OFF_EXT + D32_CH * 4;          // uint8[1024] token-start masks of the parser
constexpr int W32_OFF_LIST = W32_OFF_MASK + D32_CH;
constexpr int W32_SMEM = W32_OFF_LIST + 2 * W32_LCAP * 8;       // 226 432 bytes

typedef NamedBar<1, W32_TT> TBar;                    // transform group
typedef NamedBar<2, W32_PT> PBar;                    // parser group
__device__ __forceinline__ void w32_arrive(int id) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "n"(W32_T) : "memory"); }
__device__ __forceinline__ void w32_wait(int id) { asm volatile("bar.sync %0, %1;" ::"r"(id), "n"(W32_T) : "memory"); }
constexpr int W32_BAR_FULL = 3, W32_BAR_EMPTY = 5;   // + list buffer index (0 / 1)

#ifdef CVX_PHASE_TIMING
#define W32_PH_DECL long long ph_last = clock64(); unsigned long long ph_acc[12] = {0}
#define W32_PH(i) do { if (ptid == 0) { long long c_ = clock64(); ph_acc[i] += (unsigned long long)(c_ - ph_last); ph_last = c_; } } while (0)
#define W32_PH_FLUSH do { if (ptid == 0) for (int i_ = 6; i_ < 10; ++i_) atomicAdd(&g_phase[1][i_], ph_acc[i_]); } while (0)
#else
#define W32_PH_DECL
#define W32_PH(i)
#define W32_PH_FLUSH
#endif

// speculative entry state of chunk c (two chunks of look-back, exact through chunk 0)
__device__ __forceinline__ int w32_spec_entry(const uint64_t* w64, int c, int entry)
{
    int d = 0, dummy = 0;
    chunk8_walk<false>(w64[c >= 2 ? c - 2 : 0], 0ull, d, dummy);
    d = (c == 1) ? entry : d;
    chunk8_walk<false>(w64[c >= 1 ? c - 1 : 0], 0ull, d, dummy);
    return (c == 0) ? entry : d;
}

// walk of one chunk with BOTH counts in one pass: n += values, ne += list entries (values the reference decoder stores,
// zeros of runs excluded) of the tokens that start in the chunk -- the byte, operand and token-start work is shared
// (a separate entry-count pass over the masks cost 12 % more parser instructions: decode 2.277 -> 2.249 ms)
__device__ __forceinline__ uint32_t w32_walk_count(uint64_t A, uint64_t N, int& d, int& n, int& ne)
{
    uint32_t mask = 0;
#pragma unroll
    for (int p = 0; p < 8; ++p) {
        const uint32_t b = (uint32_t)(A >> (8 * p)) & 0xFFu;
        const int L = token_len(b);
        const bool st = (d == 0);
        mask |= (st ? 1u : 0u) << p;
        const uint32_t op = chunk8_operand(A, N, p);
        const bool is8 = (b == 0x7Eu || b == 0x82u);                  // VLESC3_8x, VLESC2_8x
        int c = 1;
        c = (b == 0x7Fu) ? (int)(op & 0xFFu) : c;                     // RLESC1
        c = (b == 0x7Du) ? (int)(op & 0xFFFFFFu) : c;                 // RLESC3
        c = is8 ? 8 : c;
        // entries: a literal / VLESC2 / VLESC3 whose value is not zero, a VLESC4, eight for an 8-value token, none for runs
        int e = (b != 0u && (b - 0x7Du) > 6u) ? 1 : 0;                // literal other than 0 (0x7D..0x83 are the escape codes)
        e = (b == 0x83u) ? ((op & 0xFFFFu) != 0u ? 1 : 0) : e;        // VLESC2
        e = (b == 0x81u) ? ((op & 0xFFFFFFu) != 0u ? 1 : 0) : e;      // VLESC3
        e = (b == 0x80u) ? 1 : e;                                     // VLESC4
        e = is8 ? 8 : e;
        n += st ? c : 0;
        ne += st ? e : 0;
        d = st ? L : d;
        d -= 1;
    }
    return mask;
}

// VLESC4 and the 8-value tokens, out of line: writes 1 or 8 entries at `out`
__device__ __noinline__ void w32_rare_to_list(int c, uint32_t op, int k, int vi, float scalefac, const unsigned char* win, uint2* out)
{
    if (c == VLESC4) { *out = make_uint2((unsigned)vi, __float_as_uint(__fmul_rn(__uint_as_float(op), scalefac))); return; }
    for (int j = 0; j < 8; ++j) {
        int v;
        if (c == VLESC2_8x) v = (int)(int16_t)((uint32_t)win[k + 2 * j] | ((uint32_t)win[k + 2 * j + 1] << 8));
        else v = (int)((uint32_t)win[k + 3 * j] << 8 | (uint32_t)win[k + 3 * j + 1] << 16 | (uint32_t)win[k + 3 * j + 2] << 24) >> 8;
        out[j] = make_uint2((unsigned)(vi + j), __float_as_uint(__fmul_rn(scalefac, (float)v)));
    }
}

// expansion of one chunk into list entries (same token semantics as chunk8_expand); returns the advanced raster index
__device__ __forceinline__ int w32_expand_list(uint32_t mask, uint64_t A, uint64_t N, int base, int vi, float scalefac,
                                               const unsigned char* win, uint2*& out)
{
#pragma unroll
    for (int p = 0; p < 8; ++p) {
        const bool st = ((mask >> p) & 1u) != 0;
        const int c = (int)(int8_t)((uint32_t)(A >> (8 * p)) & 0xFFu);
        const uint32_t op = chunk8_operand(A, N, p);
        const bool lit = (uint32_t)(c + 124) <= 248u;
        const bool v2 = (c == VLESC2), v3 = (c == VLESC3);
        int q = c;
        q = v2 ? (int)(int16_t)(op & 0xFFFFu) : q;
        q = v3 ? ((int)(op << 8) >> 8) : q;
        const bool single = lit || v2 || v3;
        if (st && single && q != 0) *out++ = make_uint2((unsigned)vi, __float_as_uint(__fmul_rn((float)q, scalefac)));
        int adv = (c == RLESC1) ? (int)(op & 0xFFu) : 1;
        adv = (c == RLESC3) ? (int)(op & 0xFFFFFFu) : adv;
        const bool is8 = (c == VLESC2_8x || c == VLESC3_8x);
        if (st && (c == VLESC4 || is8)) {
            w32_rare_to_list(c, op, base + p + 1, vi, scalefac, win, out);
            out += is8 ? 8 : 1;
        }
        adv = is8 ? 8 : adv;
        vi += st ? adv : 0;
    }
    return vi;
}

// The parser group's work on one block.  Returns (to every parser thread) the number of list entries, or -1 when
// the block must be decoded in line.
__device__ __forceinline__ int w32_parse_to_list(const D32Src& b, const unsigned char* win, unsigned char* s_mask, int* s_ext,
                                                 unsigned long long* scratch64, uint2* list)
{
    const int ptid = threadIdx.x - W32_TT;
    const uint64_t* w64 = reinterpret_cast<const uint64_t*>(win);
    const int entry = (int)(reinterpret_cast<uintptr_t>(b.src) & 15);
    const int nch = d32_nchunks(entry, b.hint);
    const int K = (nch + W32_PT - 1) / W32_PT;
    const int c0 = min(nch, ptid * K), c1 = min(nch, c0 + K);
    const bool active = c0 < c1;
    int ent = active ? w32_spec_entry(w64, c0, entry) : 0;
    int ext = 0, cntv = 0, cnte = 0;
    auto walk_range = [&](int d) {
        cntv = 0; cnte = 0;
        for (int c = c0; c < c1; ++c) {
            int n = 0;
            const uint32_t m = w32_walk_count(w64[c], w64[c + 1], d, n, cnte);
            s_mask[c] = (unsigned char)m;
            cntv = min(cntv + min(n, 1 << 20), 1 << 20);
        }
        ext = d;
    };
    walk_range(ent);
    for (;;) {                             // chain check over the ranges; inactive threads sit at the tail
        s_ext[ptid] = ext;
        PBar::sync();
        const int want = (ptid > 0 && active) ? s_ext[ptid - 1] : ent;
        const bool bad = want != ent;
        if (!PBar::sync_or(bad ? 1 : 0)) break;
        if (bad) { ent = want; walk_range(ent); }
    }
    unsigned long long total;
    const unsigned long long ex = d32_excl_sum64<W32_PT, PBar>((unsigned long long)(unsigned)cntv | ((unsigned long long)(unsigned)cnte << 32), scratch64, total);
    const long totv = (long)(total & 0xFFFFFFFFull), tote = (long)(total >> 32);
    if (totv < F32_B || tote > W32_LCAP) return -1;          // wrong length hint (foreign block order) or a dense block
    int vi = (int)(ex & 0xFFFFFFFFull);
    uint2* out = list + (int)(ex >> 32);
    for (int c = c0; c < c1; ++c) vi = w32_expand_list((uint32_t)s_mask[c], w64[c], w64[c + 1], 8 * c, vi, b.scalefac, win, out);
    return (int)tote;
}


// The block buffer must be all zero before a block's list is scattered into it.  Instead of 64 zero stores per transform
// thread (1024 shared-memory store wavefronts per block on the critical group) one thread asks the bulk-copy engine to
// overwrite the buffer from a 128 KiB page of zeros that stays in L2: asynchronous, no LSU slots, hidden under the z math
// and the row stores.
__device__ __align__(128) unsigned char g_zero_page[F32_B * 4];
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

__global__ void __launch_bounds__(W32_T, 1) k_decode32ws(const DecParams pin)
{
    const DecParams p = dec_resolve(pin);
    constexpr bool EXACT = false;
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    float* sm = reinterpret_cast<float*>(smem_raw);
    unsigned char* s_win = smem_raw + W32_OFF_WIN;
    int* s_ext = reinterpret_cast<int*>(smem_raw + W32_OFF_EXT);
    unsigned char* s_mask = smem_raw + W32_OFF_MASK;
    uint2* s_lists = reinterpret_cast<uint2*>(smem_raw + W32_OFF_LIST);
    __shared__ unsigned long long scratch64[W32_TT / 32];
    __shared__ unsigned long long pscratch64[W32_PT / 32];
    __shared__ int s_ctl[2];                                       // per list buffer: entries in the list, or -1: decode in line
    __shared__ int p_ext[W32_PT];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint8_t* hi = p.payload + p.payload_bytes;
    if ((long)blockIdx.x >= p.nblk) return;
    __shared__ uint64_t s_zbar;                                     // completion of the block buffer's bulk zero fill
    if (tid == 0) mbar_init(&s_zbar, 1);
    __syncthreads();

    if (tid >= W32_TT) {
        // ------------------------------ parser group ------------------------------
        asm volatile("setmaxnreg.dec.sync.aligned.u32 48;");
        const int ptid = tid - W32_TT;
        int it = 0;
        W32_PH_DECL;
        // payload windows are double-buffered: the window of block t+1 is requested before block t is parsed
        auto servable_of = [&](const D32Src& b) {
            return !b.raw && b.hint > 0 && (int)(reinterpret_cast<uintptr_t>(b.src) & 15) + b.hint <= D32_CH * 8;
        };
        auto request = [&](const D32Src& b, unsigned char* w) {
            if (servable_of(b)) {
                const uint8_t* cbase = reinterpret_cast<const uint8_t*>(reinterpret_cast<uintptr_t>(b.src) & ~(uintptr_t)15);
                for (int g = ptid; g < D32_WIN / 16; g += W32_PT) {
                    const uint8_t* a = cbase + 16 * g;
                    long left = hi - a;
                    int nb = left >= 16 ? 16 : (left > 0 ? (int)left : 0);
                    cp_async16_zfill(w + 16 * g, nb > 0 ? a : cbase, nb);
                }
            }
            cp_async_commit();                                     // one group per block, empty or not
        };
        D32Src b = d32_locate(p, blockIdx.x);
        request(b, s_win);
        D32Raw nraw = d32_locate_issue(p, (long)blockIdx.x + gridDim.x);      // table entries are read one block ahead of their use
        for (long t = blockIdx.x; t < p.nblk; t += gridDim.x, ++it) {
            unsigned char* win = s_win + (it & 1) * D32_WIN;
            D32Src bn = b;
            if (t + gridDim.x < p.nblk) {
                bn = d32_locate_finish(p, t + gridDim.x, nraw);
                request(bn, s_win + ((it + 1) & 1) * D32_WIN);     // that window's last reader finished before the previous PBar
            } else {
                cp_async_commit();
            }
            nraw = d32_locate_issue(p, t + 2L * gridDim.x);
            const bool servable = servable_of(b);
            int n = -1;
            asm volatile("cp.async.wait_group 1;" ::: "memory");   // this block's window has landed (this thread's part ...
            PBar::sync();                                          // ... and everybody else's)
            W32_PH(6);
            const int lb = it & 1;
            if (it >= 2) w32_wait(W32_BAR_EMPTY + lb);             // the list written two blocks ago has been scattered
            W32_PH(7);
            if (servable) n = w32_parse_to_list(b, win, s_mask, p_ext, pscratch64, s_lists + lb * W32_LCAP);
            if (ptid == 0) s_ctl[lb] = n;
            __threadfence_block();
            PBar::sync();                                          // every parser thread is done with the window: the next load may overwrite it
            w32_arrive(W32_BAR_FULL + lb);
            b = bn;
            W32_PH(8);
        }
        W32_PH_FLUSH;
        for (int k = it >= 2 ? it - 2 : 0; k < it; ++k) w32_wait(W32_BAR_EMPTY + (k & 1));     // match the transform group's last arrivals
        return;
    }
    // ------------------------------ transform group ------------------------------
    asm volatile("setmaxnreg.inc.sync.aligned.u32 96;");
    const long plane = (long)p.g.nx * p.g.ny;
    {
        float4* z4 = reinterpret_cast<float4*>(sm);
#pragma unroll
        for (int i = 0; i < F32_B / 4 / W32_TT; ++i) z4[tid + W32_TT * i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    unsigned char* win2 = s_win + 2 * D32_WIN;                      // window of the in-line path
    CVX_PH_DECL;
    int tb = 0;                                                     // list buffer of the block at hand
    uint32_t zparity = 0;
    for (long t = blockIdx.x; t < p.nblk; t += gridDim.x, tb ^= 1) {
        CVX_PH(5);
        w32_wait(W32_BAR_FULL + tb);                                // also orders every thread's zeroing before the scatter
        if (t != (long)blockIdx.x) { mbar_wait(&s_zbar, zparity); zparity ^= 1u; }      // the bulk zero fill issued in the previous iteration has landed
        CVX_PH(0);
        const int n = s_ctl[tb];
        const uint2* s_list = s_lists + tb * W32_LCAP;
        if (n >= 0) {
            for (int i = tid; i < n; i += W32_TT) {
                const uint2 e = s_list[i];
                if (e.x < (unsigned)F32_B) sm[physr((int)e.x, 5)] = __uint_as_float(e.y);
            }
        } else {
            const D32Src cur = d32_locate(p, t);
            if (cur.raw) {                                          // CvxCompress.cpp:555: B fp32 coefficients
                for (int i = tid; i < F32_B; i += W32_TT) {
                    uint32_t u = 0;
                    const uint8_t* s = cur.src + 4L * i;
                    if ((reinterpret_cast<uintptr_t>(cur.src) & 3) == 0 && s + 4 <= hi) u = *reinterpret_cast<const uint32_t*>(s);
                    else for (int k = 0; k < 4; ++k) if (s + k < hi) u |= (uint32_t)s[k] << (8 * k);
                    sm[physr(i, 5)] = __uint_as_float(u);
                }
            } else {
                const int entry = (int)(reinterpret_cast<uintptr_t>(cur.src) & 15);
                for (int g = tid; g < D32_WIN / 16; g += W32_TT) {
                    const uint8_t* a = cur.src - entry + 16 * g;
                    long left = hi - a;
                    int nb = left >= 16 ? 16 : (left > 0 ? (int)left : 0);
                    cp_async16_zfill(win2 + 16 * g, nb > 0 ? a : cur.src - entry, nb);
                }
                cp_async_commit();
                cp_async_wait_all();
                TBar::sync();
                D32Walk wk;
                d32_walk<W32_TT>(win2, entry, d32_nchunks(entry, cur.hint), wk);
                d32_parse<W32_TT, TBar>(hi, cur, win2, s_ext, scratch64, wk, [sm](int i, float v) { sm[physr(i, 5)] = v; });
            }
        }
        TBar::sync();                                               // block complete, s_ctl and the list consumed
        w32_arrive(W32_BAR_EMPTY + tb);
        CVX_PH(1);
        f32_pass_xT<EXACT, false, W32_TT>(sm, true);
        __syncwarp();
        f32_pass_yT<EXACT, false, W32_TT>(sm);                      // (skewing this pass as well was measured: no change)
        TBar::sync();
        CVX_PH(2);
        typename LineOf<EXACT, W32_TT>::type v[32];
        {
            float* zl = f32_zline(sm, tid);
            if (warp >= 8) asm volatile("bar.sync 7, 512;" ::: "memory");       // skew: see f32_pass_x
#pragma unroll
            for (int r = 0; r < 32; ++r) {
                float* e = zl + reg_to_sub<32, EXACT>(r) * 1024;
                p_set(v[r], e[0], e[512]);
            }
            if (warp < 8) asm volatile("bar.arrive 7, 512;" ::: "memory");
            // every thread's z lines are in registers -> the buffer may be overwritten: warps 0..14 arrive (no wait), warp 15
            // (last to start its loads, see the skew) waits and one of its lanes issues the zero fill when there is a next block
            if (t + gridDim.x < p.nblk) {
                if (warp != 15) asm volatile("bar.arrive 8, 512;" ::: "memory");
                else {
                    asm volatile("bar.sync 8, 512;" ::: "memory");
                    if (lane == 0) {
                        fence_proxy_async();
                        mbar_expect_tx(&s_zbar, F32_B * 4);
#pragma unroll
                        for (int b = 0; b < 4; ++b) bulk_g2s(sm + b * 8192, g_zero_page + b * 32768, 32768, &s_zbar);
                    }
                }
            }
        }
        const int iix = (int)(t % p.nbx);
        const long q = t / p.nbx;
        const int iiy = (int)(q % p.nby), iiz = (int)(q / p.nby);
        const int gx = iix * 32 + lane, gz0 = iiz * 32;
        const int nzl = min(32, p.g.nz - gz0);
        inv_line<32, EXACT>(v);
        CVX_PH(3);
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const int gy = iiy * 32 + warp + 16 * k;
            if (gx < p.g.nx && gy < p.g.ny) {
                float* dp = p.vol + ((long)gz0 * p.g.ny + gy) * p.g.nx + gx;
                if (nzl == 32 && plane < (1L << 30)) {
                    // plane j of the block at byte offset pb * j: ptxas turns this into shift / add chains, about three
                    // instructions per row store.  (Forcing one IMAD.WIDE with an immediate j per store through PTX was
                    // measured SLOWER: decode 2.28 -> 2.55 ms.)
                    const unsigned pb = (unsigned)plane * 4u;
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        *reinterpret_cast<float*>(reinterpret_cast<char*>(dp) + (unsigned long long)pb * (unsigned)j) = k ? p_hi(v[j]) : p_lo(v[j]);
                } else if (nzl == 32) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) { *dp = k ? p_hi(v[j]) : p_lo(v[j]); dp += plane; }
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) { if (j < nzl) *dp = k ? p_hi(v[j]) : p_lo(v[j]); dp += plane; }
                }
            }
        }
    }
    CVX_PH(5);
    CVX_PH_FLUSH(1);
}

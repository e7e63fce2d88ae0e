// cvx_decode_win.inl -- k_decode_win: the fused decoder of the block shapes other than 32^3 that hold 8192 .. 32768 cells
// (BASELINE config 4: 16 x 16 x 64), included by cvx_kernels.cu after cvx_decode32.inl.
//
// k_decode_fused parses a block with cta_rle_decode (cvx_codec.cuh): the payload is fetched with plain loads when the block
// is already waiting for it, 16-byte chunks are walked and expanded behind five CTA barriers, and nothing overlaps -- on
// config 4 two thirds of the kernel's stall samples sat there.  This kernel keeps k_decode_fused's line passes (any shape,
// EXACT or LIFTING) and takes the payload side from k_decode32: persistent CTAs, the NEXT block's payload window streams in
// by cp.async under the transform, 8-byte chunks (two per thread), the speculative walk of the next window hoisted behind
// the scatter, parse continuing window by window for long payloads or foreign block order (d32_parse with the block's cell
// count).  The scatter to the volume leaves zeros behind, so the buffer is clean for the next block's expansion.

template <int MAXN, bool EXACT, int T, int MINB>
__global__ void __launch_bounds__(T, MINB) k_decode_win(const DecParams pin)
{
    const DecParams p = dec_resolve(pin);
    constexpr int CH = 2 * T;                                       // 8-byte chunks per window, two per thread
    constexpr int WIN = CH * 8 + 32;                                // + look-ahead for tokens straddling the window end
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const BlkShape s = p.s;
    const int B = s.B, sh = s.sh;
    float* sm = reinterpret_cast<float*>(smem_raw);
    unsigned char* s_win = smem_raw + (size_t)B * 4;                // two windows
    int* s_ext = reinterpret_cast<int*>(s_win + 2 * WIN);
    __shared__ unsigned long long scratch64[T / 32];

    const int tid = threadIdx.x;
    const uint8_t* hi = p.payload + p.payload_bytes;
    long t = blockIdx.x;
    if (t >= p.nblk) return;
    auto win_base = [](const D32Src& b) { return reinterpret_cast<const uint8_t*>(reinterpret_cast<uintptr_t>(b.src) & ~(uintptr_t)15); };
    D32Src cur = d32_locate(p, t, CH);
    int buf = 0;
    d32_issue_window(s_win, win_base(cur), hi, T, WIN);
    for (int i = tid; i < (B >> 2); i += T) reinterpret_cast<float4*>(sm)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    D32Walk wk;
    cp_async_wait_all();
    __syncthreads();                                                // first window staged, block buffer zeroed
    {
        const int entry = (int)(reinterpret_cast<uintptr_t>(cur.src) & 15);
        d32_walk<T>(s_win, entry, d32_nchunks(entry, cur.hint, CH), wk);
    }
    for (;;) {
        const long tn = t + gridDim.x;
        D32Src nxt = cur;
        if (tn < p.nblk) nxt = d32_locate(p, tn, CH);               // table reads overlap the parse below
        if (cur.raw) {                                              // CvxCompress.cpp:555: B fp32 coefficients
            __syncthreads();                                        // every thread has zeroed its cells of the previous block
            for (int i = tid; i < B; i += T) {
                uint32_t u = 0;
                const uint8_t* sp = cur.src + 4L * i;
                if ((reinterpret_cast<uintptr_t>(cur.src) & 3) == 0 && sp + 4 <= hi) u = *reinterpret_cast<const uint32_t*>(sp);
                else for (int k = 0; k < 4; ++k) if (sp + k < hi) u |= (uint32_t)sp[k] << (8 * k);
                sm[physr(i, sh)] = __uint_as_float(u);
            }
        } else {
            d32_parse<T, CtaBar>(hi, cur, s_win + buf * WIN, s_ext, scratch64, wk, [sm, sh](int i, float v) { sm[physr(i, sh)] = v; }, B, CH);
        }
        __syncthreads();
        if (tn < p.nblk) d32_issue_window(s_win + (buf ^ 1) * WIN, win_base(nxt), hi, T, WIN);     // streams in under the transform
        transform_block<MAXN, EXACT, false, T>(sm, s);              // ends with a CTA barrier
        {
            // Copy_From_Block (Block_Copy.cpp:136-212) with zeros left behind: each thread clears exactly the cells it read
            const int iix = (int)(t % p.nbx);
            const long q = t / p.nbx;
            const int x0 = iix * s.bx, y0 = (int)(q % p.nby) * s.by, z0 = (int)(q / p.nby) * s.bz;
            if (p.g.vec_ok) {
                const int lcpr = s.lbx - 2;
#pragma unroll 4
                for (int idx = tid; idx < (B >> 2); idx += T) {
                    const int c = idx & ((1 << lcpr) - 1), r = idx >> lcpr;
                    const int gy = y0 + (r & (s.by - 1)), gz = z0 + (r >> s.lby), gx = x0 + 4 * c;
                    float4* cell = reinterpret_cast<float4*>(&sm[physr(4 * idx, sh)]);
                    if (gy < p.g.ny && gz < p.g.nz && gx < p.g.nx)
                        *reinterpret_cast<float4*>(p.vol + ((long)gz * p.g.ny + gy) * p.g.nx + gx) = *cell;
                    *cell = make_float4(0.f, 0.f, 0.f, 0.f);
                }
            } else {
#pragma unroll 4
                for (int i = tid; i < B; i += T) {
                    const int x = i & (s.bx - 1), r = i >> s.lbx;
                    const int gy = y0 + (r & (s.by - 1)), gz = z0 + (r >> s.lby), gx = x0 + x;
                    float* cell = &sm[physr(i, sh)];
                    if (gy < p.g.ny && gz < p.g.nz && gx < p.g.nx) p.vol[((long)gz * p.g.ny + gy) * p.g.nx + gx] = *cell;
                    *cell = 0.f;
                }
            }
        }
        if (tn >= p.nblk) break;
        cp_async_wait_all();                                        // this thread's part of the next window has landed ...
        __syncthreads();                                            // ... and everybody's; the block buffer is all zero
        {
            const int entry = (int)(reinterpret_cast<uintptr_t>(nxt.src) & 15);
            d32_walk<T>(s_win + (buf ^ 1) * WIN, entry, d32_nchunks(entry, nxt.hint, CH), wk);
        }
        t = tn;
        cur = nxt;
        buf ^= 1;
    }
}

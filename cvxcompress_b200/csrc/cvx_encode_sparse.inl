// cvx_encode_sparse.inl -- k_encode_sparse: the fused encoder of the block shapes other than 32^3 that hold 8192 .. 32768
// cells (BASELINE config 4: 16 x 16 x 64), included by cvx_kernels.cu after cvx_encode32.inl.
//
// k_encode_fused codes every block with the dense segment coder (cvx_codec.cuh): two passes over ALL coefficients in
// groups of eight, divergent, plus a look-back that the whole CTA waits for -- on config 4 that was 63 % of the kernel's
// instructions and 74 % of its stall samples for data that quantises to ~95 % zeros.  This kernel keeps k_encode_fused's
// gather and line passes (any shape, EXACT or LIFTING) and replaces the tail with k_encode32's sparse coder, generalised
// from 1024 rows of 32 cells to B / 32 rows:
//   one ballot per row of 32 raster-consecutive cells -> row masks -> scan -> (raster index, coefficient) list of the
//   non-zeros -> group-of-8 marking, token sizes, early aggregate publish -> branch-free emission into a staging buffer
//   -> look-back and copy-out deferred by one block (by then the neighbouring tickets have published).
// Byte-identical to the dense coder (and so to Run_Length_Encode_Slow.cpp:189-294); a block with more non-zeros than the
// list holds (18.75 % of its cells), or whose multiplier is outside the sparse path's range, is coded by the dense coder, its
// payload staged and its copy-out deferred in the same way.

constexpr int ES_MAXROWS = 1024;          // rows of 32 cells: blocks of up to 32768 cells

__host__ __device__ inline int es_lcap(int B) { const int c = 3 * (B / 16); return c < 4096 ? c : 4096; }      // list entries (8 bytes each): 18.75 % of the cells, at most 128 chunks of 32
__host__ __device__ inline int es_stage(int B) { return 3 * (B / 8); }          // bytes of each of the two staging buffers
__host__ __device__ inline int es_list_bytes(int B, int nseg) { const int a = es_lcap(B) * 8, b = nseg * 16; return a > b ? a : b; }
static size_t es_smem_bytes(int B, int nseg) { return (size_t)B * 4 + (size_t)(B / 32) * 8 + (size_t)es_list_bytes(B, nseg) + 2 * (size_t)es_stage(B); }
static bool es_ok(int bx, int by, int bz)
{
    const long B = (long)bx * by * bz;
    return bx <= 64 && by <= 64 && bz <= 64 && B >= 8192 && B <= 32768 && !(bx == 32 && by == 32 && bz == 32);
}

template <int MAXN, bool EXACT, int T, int MINB>
__global__ void __launch_bounds__(T, MINB) k_encode_sparse(const EncParams p)
{
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const BlkShape s = p.s;
    const int B = s.B, sh = s.sh, NR = B >> 5, LCAP = es_lcap(B), STAGE = es_stage(B);
    float* sm = reinterpret_cast<float*>(smem_raw);
    uint32_t* s_nzm = reinterpret_cast<uint32_t*>(smem_raw + (size_t)B * 4);
    int* s_base = reinterpret_cast<int*>(s_nzm + NR);
    uint2* s_list = reinterpret_cast<uint2*>(s_base + NR);
    unsigned char* s_stage = reinterpret_cast<unsigned char*>(s_list) + es_list_bytes(B, p.nseg);
    __shared__ int scratch[72];
    __shared__ float s_red[32];
    __shared__ unsigned int s_ticket[2];
    __shared__ unsigned long long s_lb[T / 32 + 2];
    __shared__ int s_chunk[ES_MAXROWS * 32 / 8 / 32];
    __shared__ unsigned short s_dq[ES_MAXROWS];
    __shared__ int s_dqn;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned lt = (1u << lane) - 1u;
    const float gmul = p.use_local ? 1.0f : *p.glob_mulfac;


    // the block whose payload sits in a staging buffer awaiting its offset (deferred copy-out)
    long pend_t = -1;
    int pend_len = 0, buf = 0;
    auto complete_pending = [&]() {
        if (pend_t < 0) return;
        const unsigned long long excl = wide_lookback<T, CtaBar>(p, pend_t, (unsigned long long)pend_len, false, s_lb);
        if ((long)(excl + (unsigned long long)pend_len) <= p.cap) copy_out<T>(p.payload + excl, s_stage + (buf ^ 1) * STAGE, pend_len);
        else if (tid == 0) *p.err = -3;
        pend_t = -1;
    };

    for (int it = 0;; ++it) {
        // the ticket is drawn when the block starts, not a block ahead as in k_encode32: block costs differ by 3 x here (dense
        // blocks), and a ticket parked behind a slow block holds up the look-back of every CTA that is ahead of it
        if (tid == 0) s_ticket[it & 1] = atomicAdd(p.ticket, 1u);
        __syncthreads();
        const long t = s_ticket[it & 1];
        if (t >= p.nblk) break;
        {
            const int iix = (int)(t % p.nbx);
            const long q = t / p.nbx;
            gather_block<T>(sm, p.vol, p.g, iix * s.bx, (int)(q % p.nby) * s.by, (int)(q / p.nby) * s.bz, s);
        }
        __syncthreads();
        transform_block<MAXN, EXACT, true, T>(sm, s);
        float mulfac = gmul;
        if (p.use_local) {                                           // Compute_Local_RMS, CvxCompress.cpp:119-142, :343-348
            const float ss = EXACT ? cta_sumsq_exact<T>(B, s_red, [&](int i) { return sm[physr(i, sh)]; })
                                   : cta_sumsq<T>(sm, B, s_red);
            const float rms = sqrtf(__fdiv_rn(ss, (float)B));
            mulfac = rms != 0.0f ? __fdiv_rn(1.0f, __fmul_rn(rms, p.scale)) : 1.0f;
            if (tid == 0) p.blkmulfac[t] = mulfac;
        }
        // ---- which coefficients quantise to non-zero: one compare per cell, one ballot per row of 32 raster-consecutive cells ----
        const bool sparse_ok = fabsf(mulfac) > 1e-30f && fabsf(mulfac) < 1e30f;
        const float Tz = sparse_ok ? nz_threshold(mulfac) : 0.0f;
        for (int row = warp; row < NR; row += T / 32) {
            const float c = sm[physr(row * 32 + lane, sh)];
            const unsigned m = __ballot_sync(0xffffffffu, !(fabsf(c) <= Tz));
            if (lane == 0) s_nzm[row] = m;
        }
        __syncthreads();
        // rows RPT * tid .. + RPT - 1 -> list positions (exclusive scan of the row populations)
        const int RPT = NR / T;                                      // 1, 2 or 4 (es_ok shapes with the dispatch's T)
        int n = LCAP + 1;
        if (sparse_ok) {
            int cnt = 0;
            for (int k = 0; k < RPT; ++k) cnt += __popc(s_nzm[RPT * tid + k]);
            int base = cta_excl_sum<T>(cnt, scratch, n);
            for (int k = 0; k < RPT; ++k) { s_base[RPT * tid + k] = base; base += __popc(s_nzm[RPT * tid + k]); }
        }
        if (n > LCAP) {
            // ---- dense block: the segment coder of cvx_codec.cuh on the shared-memory block (also raw fallback,
            //      CvxCompress.cpp:350-360), with the same early publish / staged payload / deferred copy-out as below:
            //      a CTA that waited here for its predecessors' byte counts stalled every CTA behind it ----
            __syncthreads();
            auto get8 = [&](int g, float* c) {
                const int i = g * 8;
                float4 a = *reinterpret_cast<const float4*>(&sm[physr(i, sh)]);
                float4 b = *reinterpret_cast<const float4*>(&sm[physr(i + 4, sh)]);
                c[0] = a.x; c[1] = a.y; c[2] = a.z; c[3] = a.w; c[4] = b.x; c[5] = b.y; c[6] = b.z; c[7] = b.w;
            };
            int4* s_seg = reinterpret_cast<int4*>(s_list);
            const int dlen = cta_rle_plan<T>(p.nseg, p.gps, mulfac, get8, s_seg, scratch);
            const bool raw = (long)dlen > 4L * B;                     // CvxCompress.cpp:350
            const int nbytes = raw ? 4 * B : dlen;
            if (tid == 0) publish_aggregate(p, t, (unsigned long long)nbytes);
            const bool ddirect = raw || nbytes > STAGE - 8;
            if (ddirect) {
                const unsigned long long excl = wide_lookback<T, CtaBar>(p, t, (unsigned long long)nbytes, raw, s_lb);
                if ((long)(excl + (unsigned long long)nbytes) > p.cap) {
                    if (tid == 0) *p.err = -3;
                } else {
                    uint8_t* dst = p.payload + excl;
                    if (raw) store_raw<T>(sm, B, dst, [sh](long i) { return (long)physr((int)i, sh); });
                    else cta_rle_emit<T>(p.nseg, p.gps, mulfac, get8, s_seg, [&](int pos, uint8_t b) { dst[pos] = b; });
                }
            } else {
                unsigned char* out = s_stage + buf * STAGE;
                cta_rle_emit<T>(p.nseg, p.gps, mulfac, get8, s_seg, [&](int pos, uint8_t b) { out[pos] = b; });
            }
            complete_pending();
            if (!ddirect) { pend_t = t; pend_len = nbytes; buf ^= 1; }
            __syncthreads();
            continue;
        }
        // ---- sparse block: (index, value) list in raster order.  A row with up to 4 non-zeros is unpacked by its thread;
        //      denser rows (the low-frequency head of the block) are queued and compacted by whole warps, one lane per cell ----
        if (tid == 0) s_dqn = 0;
        __syncthreads();                 // s_base complete
        for (int row = tid; row < NR; row += T) {
            unsigned m = s_nzm[row];
            if (__popc(m) > 4) {
                s_dq[atomicAdd(&s_dqn, 1)] = (unsigned short)row;
            } else {
                int at = s_base[row];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    if (m) {
                        const int x = __ffs(m) - 1;
                        m &= m - 1;
                        s_list[at++] = make_uint2((unsigned)(row * 32 + x), __float_as_uint(sm[physr(row * 32 + x, sh)]));
                    }
                }
            }
        }
        __syncthreads();
        for (int i = warp; i < s_dqn; i += T / 32) {
            const int rr = s_dq[i];
            const unsigned mm = s_nzm[rr];
            if ((mm >> lane) & 1u)
                s_list[s_base[rr] + __popc(mm & lt)] = make_uint2((unsigned)(rr * 32 + lane), __float_as_uint(sm[physr(rr * 32 + lane, sh)]));
        }
        __syncthreads();                 // list complete
        f32_mark_groups<T>(s_list, n, mulfac);
        __syncthreads();
        // ---- tokens per list entry: chunks of 32 consecutive entries dealt round-robin to the warps; byte offsets by a
        //      shuffle scan inside the chunk + one scan of the (<= 128) chunk totals ----
        const int nchunks = (n + 31) >> 5;
        for (int c = warp; c < nchunks; c += T / 32) {
            const int e = 32 * c + lane;
            int bytes = 0;
            uint2 ent = make_uint2(0u, 0u);
            if (e < n) {
                ent = s_list[e];
                const int prev = e > 0 ? (int)(s_list[e - 1].x & F32_IDX_MASK) : -1;
                bytes = entry_token(ent, prev, mulfac).bytes;
            }
            int incl = bytes;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int y = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d) incl += y;
            }
            if (e < n) s_list[e].x = ent.x | ((unsigned)(incl - bytes) << 15);      // offset inside the chunk, next to the index
            if (lane == 31) s_chunk[c] = incl;
        }
        __syncthreads();
        int len;
        int cpre[4];                     // lane l: exclusive byte offsets of chunks 4l .. 4l+3 (every warp computes all of them)
        {
            int cs[4], tot = 0;
#pragma unroll
            for (int i = 0; i < 4; ++i) { cs[i] = (4 * lane + i < nchunks) ? s_chunk[4 * lane + i] : 0; tot += cs[i]; }
            int sc = tot;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int y = __shfl_up_sync(0xffffffffu, sc, d);
                if (lane >= d) sc += y;
            }
            len = __shfl_sync(0xffffffffu, sc, 31);
            cpre[0] = sc - tot;
#pragma unroll
            for (int i = 1; i < 4; ++i) cpre[i] = cpre[i - 1] + cs[i - 1];
        }
        const int tail_run = B - 1 - (n > 0 ? (int)(s_list[n - 1].x & F32_IDX_MASK) : -1);          // final flush (:294)
        const int tail_pos = len;
        len += run_bytes(tail_run);
        if (tid == 0) publish_aggregate(p, t, (unsigned long long)len);
        const bool direct = len > STAGE - 8;
        unsigned long long excl_now = 0;
        if (direct) excl_now = wide_lookback<T, CtaBar>(p, t, (unsigned long long)len, false, s_lb);
        const bool fits = !direct || (long)(excl_now + (unsigned long long)len) <= p.cap;
        unsigned char* out = direct ? p.payload + excl_now : s_stage + buf * STAGE;
        if (fits) {
            auto put = [&](int ps, uint8_t b) { out[ps] = b; };
            for (int c = warp; c < nchunks; c += T / 32) {
                const int src = c >> 2;
                const int b0 = __shfl_sync(0xffffffffu, cpre[0], src), b1 = __shfl_sync(0xffffffffu, cpre[1], src);
                const int b2 = __shfl_sync(0xffffffffu, cpre[2], src), b3 = __shfl_sync(0xffffffffu, cpre[3], src);
                const int cbase = (c & 2) ? ((c & 1) ? b3 : b2) : ((c & 1) ? b1 : b0);
                const int e = 32 * c + lane;
                if (e >= n) continue;
                const uint2 ent = s_list[e];
                const int prev = e > 0 ? (int)(s_list[e - 1].x & F32_IDX_MASK) : -1;
                const EntryTok k = entry_token(ent, prev, mulfac);
                if (k.bytes == 0) continue;
                int pos = cbase + (int)((ent.x >> 15) & F32_OFF_MASK);
                if (k.mode == GM_ELEM) {
                    put_entry_branchfree(out, pos, k.run, k.q, k.f);
                } else {
                    pos = put_run(k.run, pos, put);
                    const int w = k.mode == GM_S8 ? 2 : 3;
                    put(pos++, (uint8_t)(k.mode == GM_S8 ? VLESC2_8x : VLESC3_8x));
                    for (int i = 0; i < 8; ++i) {
                        const int q = quantise(__fmul_rn(mulfac, __uint_as_float(s_list[e + i].y)));
                        put(pos++, (uint8_t)(q & 0xFF)); put(pos++, (uint8_t)((q >> 8) & 0xFF));
                        if (w == 3) put(pos++, (uint8_t)((q >> 16) & 0xFF));
                    }
                }
            }
            if (tid == 0) put_run(tail_run, tail_pos, put);
        } else if (tid == 0) {
            *p.err = -3;
        }
        // the PREVIOUS staged block gets its offset now (its neighbours published a block time ago) and is copied out
        complete_pending();
        if (!direct) { pend_t = t; pend_len = len; buf ^= 1; }
        __syncthreads();                 // staging buffer complete before the next block's copy-out reads it; block buffer free
    }
    __syncthreads();
    complete_pending();
}

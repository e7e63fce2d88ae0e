// cvx_encode32.inl -- k_encode32, included by cvx_kernels.cu inside namespace cvx after EncParams,
// gather_block, cta_sumsq*, encode_tail (design notes: cvx_fused32.cuh).

// ---- placing block payloads in natural (ticket) order without stalling the pipeline -------------------------
// A block's byte count is published as an AGGREGATE the moment it is known; its exclusive prefix (= payload
// offset) is only needed for the copy-out, which is deferred by one block: by then every ticket in flight next
// to it has published, so the look-back below finds its words ready instead of spinning on the slowest CTA of the
// wave (measured: 30 % of the kernel waited in the undeferred version).  The look-back reads F32_T status words
// per round with the whole CTA (one L2 round trip for a window deeper than the grid).
__device__ __forceinline__ void publish_aggregate(const EncParams& p, long t, unsigned long long len)
{
    st_release_u64(&p.status[t], (t == 0 ? CVX_ST_PFX : CVX_ST_AGG) | len);
}

// all GT threads of the group; returns the exclusive prefix of ticket t to every thread.  scratch64: >= GT/32 words.
template <int GT = F32_T, class Bar = CtaBar>
__device__ __forceinline__ unsigned long long wide_lookback(const EncParams& p, long t, unsigned long long len, bool raw,
                                                            unsigned long long* scratch64)
{
    static_assert((GT & (GT - 1)) == 0, "group size is a power of two");
    const int gtid = threadIdx.x & (GT - 1), gwarp = gtid >> 5;
    unsigned long long excl = 0;
    if (t > 0) {
        long j = t - 1;
        for (;;) {
            const long idx = j - gtid;
            unsigned long long s = CVX_ST_PFX;                      // before ticket 0: prefix 0
            if (idx >= 0) {
                while (((s = ld_acquire_u64(&p.status[idx])) >> 62) == 0) __nanosleep(64);      // not published yet: poll gently, the SM's other warps have work
            }
            const unsigned pfx = __ballot_sync(0xffffffffu, (s >> 62) == 2);
            const int first = pfx ? __ffs(pfx) - 1 : 32;            // nearest predecessor in this warp holding a prefix
            unsigned long long v = (lane_id() <= first) ? (s & CVX_ST_MASK) : 0ull;
#pragma unroll
            for (int d = 16; d >= 1; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
            if (lane_id() == 0) scratch64[gwarp] = v | (pfx ? (1ull << 63) : 0ull);
            Bar::sync();
            bool done = false;
            for (int w = 0; w < GT / 32; ++w) {                  // warps in look-back order; stop at the first prefix
                const unsigned long long x = scratch64[w];
                excl += x & ~(1ull << 63);
                if (x >> 63) { done = true; break; }
            }
            Bar::sync();
            if (done) break;
            j -= GT;
        }
        if (gtid == 0) st_release_u64(&p.status[t], CVX_ST_PFX | (excl + len));
    }
    if (gtid == 0) {
        p.offs[t] = excl | (raw ? 0x8000000000000000ull : 0ull);      // CvxCompress.cpp:382-384
        if (t == p.nblk - 1) *p.payload_bytes = (long)(excl + len);
    }
    return excl;
}

// optional per-phase cycle accounting (make TIMING=1): thread 0 stamps clock64() at phase boundaries
#ifdef CVX_PHASE_TIMING
__device__ unsigned long long g_phase[2][16];
#define CVX_PH_DECL long long ph_last = clock64(); unsigned long long ph_acc[12] = {0}
#define CVX_PH(i) do { if (threadIdx.x == 0) { long long c_ = clock64(); ph_acc[i] += (unsigned long long)(c_ - ph_last); ph_last = c_; } } while (0)
#define CVX_PH_FLUSH(k) do { if (threadIdx.x == 0) for (int i_ = 0; i_ < 12; ++i_) atomicAdd(&g_phase[k][i_], ph_acc[i_]); } while (0)
#else
#define CVX_PH_DECL
#define CVX_PH(i)
#define CVX_PH_FLUSH(k)
#endif

template <bool EXACT>
__global__ void __launch_bounds__(F32_T, 1) k_encode32(const __grid_constant__ CUtensorMap tmap, const EncParams p, const int use_tma)
{
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    float* sm = reinterpret_cast<float*>(smem_raw);
    uint32_t* s_nzm = reinterpret_cast<uint32_t*>(smem_raw + F32_OFF_NZM);
    int* s_base = reinterpret_cast<int*>(smem_raw + F32_OFF_BASE);
    uint2* s_list = reinterpret_cast<uint2*>(smem_raw + F32_OFF_LIST);
    unsigned char* s_stage = smem_raw + F32_OFF_STAGE;
    __shared__ uint64_t s_bar;
    __shared__ int scratch[72];
    __shared__ float s_red[32];
    __shared__ unsigned int s_ticket[2];
    __shared__ unsigned long long s_off;
    __shared__ unsigned long long s_lb[F32_T / 32 + 2];
    __shared__ int s_chunk[F32_LCAP / 32];
    __shared__ unsigned short s_dq[F32_ROWS];
    __shared__ int s_dqn;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned lt = (1u << lane) - 1u;
    const BlkShape s = p.s;
    const float gmul = p.use_local ? 1.0f : *p.glob_mulfac;

    auto issue_load = [&](unsigned int t) {      // one thread: TMA of block t into the block buffer, 4 boxes of 32x32x8
        const int iix = (int)(t % p.nbx);
        const unsigned int q = t / p.nbx;
        const int iiy = (int)(q % p.nby), iiz = (int)(q / p.nby);
        mbar_expect_tx(&s_bar, F32_B * 4);
#pragma unroll
        for (int b = 0; b < 4; ++b) tma_load_3d(sm + b * 8192, &tmap, &s_bar, iix * 32, iiy * 32, iiz * 32 + b * 8);
    };

    if (tid == 0) {
        mbar_init(&s_bar, 1);
        const unsigned int t0 = atomicAdd(p.ticket, 1u);
        s_ticket[0] = t0;
        if (use_tma && t0 < p.nblk) { fence_proxy_async(); issue_load(t0); }
    }
    __syncthreads();

    // the block whose payload sits in a staging buffer awaiting its offset (deferred copy-out)
    long pend_t = -1;
    int pend_len = 0, buf = 0;
    auto complete_pending = [&]() {
        if (pend_t < 0) return;
        const unsigned long long excl = wide_lookback(p, pend_t, (unsigned long long)pend_len, false, s_lb);
        if ((long)(excl + (unsigned long long)pend_len) <= p.cap) copy_out(p.payload + excl, s_stage + (buf ^ 1) * F32_STAGE, pend_len);
        else if (tid == 0) *p.err = -3;
        pend_t = -1;
    };

    uint32_t parity = 0;
    float row_fa = 1.0f, row_fb = 1.0f;      // LIFTING: y / z scalings of the thread's two x rows
    if constexpr (!EXACT) f32_row_factors(row_fa, row_fb);
    CVX_PH_DECL;
    for (int it = 0;; ++it) {
        const long t = s_ticket[it & 1];
        if (t >= p.nblk) break;
        // the NEXT block's ticket is drawn now: the atomic's round trip hides under the transform instead of sitting
        // in front of the prefetch
        unsigned int nt = 0;
        if (tid == 0) nt = atomicAdd(p.ticket, 1u);
        if (use_tma) {
            mbar_wait(&s_bar, parity);
            parity ^= 1u;
            CVX_PH(0);
        } else {
            const int iix = (int)(t % p.nbx);
            const long q = t / p.nbx;
            gather_block<F32_T>(sm, p.vol, p.g, iix * 32, (int)(q % p.nby) * 32, (int)(q / p.nby) * 32, s);
            __syncthreads();
        }
        float mulfac;
        bool sparse_ok;
        if constexpr (EXACT) {
            f32_pass_x<EXACT, true>(sm);
            __syncwarp();                    // warp w owns the x rows and the y lines of z slices w and w + 16: no CTA barrier needed
            CVX_PH(1);
            f32_pass_y2<EXACT, true>(sm);
            __syncthreads();
            CVX_PH(2);
            // z lines stay in registers: thread-line l = tid + 512 k is (x = lane, y = warp + 16 k)
            typename PairOf<EXACT>::type vz[32];
            {
                const float* zl = f32_zline(sm, tid);
    #pragma unroll
                for (int j = 0; j < 32; ++j) p_set(vz[j], zl[j * 1024], zl[j * 1024 + 512]);
                fwd_line<32, EXACT>(vz);
            }
    #define V_(k, r) ((k) ? p_hi(vz[r]) : p_lo(vz[r]))
            // The coefficients go back to the block buffer (one conflict-free store per cell); the coder below works on
            // row masks and gathers the few non-zero values from there, which costs far fewer integer instructions
            // (half-rate pipe) than compacting them out of the registers row by row.
            mulfac = gmul;
            if (p.use_local) {                                       // Compute_Local_RMS, CvxCompress.cpp:119-142, :343-348
    #pragma unroll
                for (int k = 0; k < 2; ++k) {
                    float* zl = f32_zline(sm, tid + F32_T * k);
    #pragma unroll
                    for (int r = 0; r < 32; ++r) zl[reg_to_sub<32, EXACT>(r) * 1024] = V_(k, r);
                }
                __syncthreads();
                const float ss = EXACT ? cta_sumsq_exact<F32_T>(F32_B, s_red, [&](int i) { return sm[physr(i, 5)]; })
                                       : cta_sumsq<F32_T>(sm, F32_B, s_red);
                const float rms = sqrtf(__fdiv_rn(ss, (float)F32_B));
                mulfac = rms != 0.0f ? __fdiv_rn(1.0f, __fmul_rn(rms, p.scale)) : 1.0f;
                if (tid == 0) p.blkmulfac[t] = mulfac;
            }
            // ---- which coefficients quantise to non-zero: one compare per cell, one ballot per row ----
            sparse_ok = fabsf(mulfac) > 1e-30f && fabsf(mulfac) < 1e30f;
            const float T = sparse_ok ? nz_threshold(mulfac) : 0.0f;
    #pragma unroll
            for (int k = 0; k < 2; ++k) {
                float* zl = f32_zline(sm, tid + F32_T * k);
    #pragma unroll
                for (int r = 0; r < 32; ++r) {
                    const float c = V_(k, r);
                    if (!p.use_local) zl[reg_to_sub<32, EXACT>(r) * 1024] = c;
                    const unsigned m = __ballot_sync(0xffffffffu, !(fabsf(c) <= T));
                    if (lane == 0) s_nzm[reg_to_sub<32, EXACT>(r) * 32 + warp + 16 * k] = m;
                }
            }
        } else {
            // LIFTING: passes in the order z, y, x (cvx_fused32.cuh, "LIFTING encoders") -- the same arithmetic on the same
            // threads as k_encode32ws, so that the two kernels stay bit-identical; here the rows go back to the block buffer
            // and the rest of the block is coded as in EXACT mode
            f32_pass_z2_fwd<EXACT>(sm);
            __syncthreads();
            CVX_PH(1);
            f32_pass_y2<EXACT, true>(sm);
            __syncwarp();                // warp w owns the y lines and the x rows of z slices w and w + 16
            CVX_PH(2);
            float w[2][32];
            {
                float v[2][32];
                f32_rows_load(sm, v);
                fwd_lift_pair_rs<32>(v[0], v[1], w[0], w[1], row_fa, row_fb);
            }
            f32_rows_store_coef(sm, w);
            mulfac = gmul;
            if (p.use_local) {
                const float ss = cta_sumsq_rows<F32_T, CtaBar>(w, s_red);
                const float rms = sqrtf(__fdiv_rn(ss, (float)F32_B));
                mulfac = rms != 0.0f ? __fdiv_rn(1.0f, __fmul_rn(rms, p.scale)) : 1.0f;
                if (tid == 0) p.blkmulfac[t] = mulfac;
            }
            sparse_ok = fabsf(mulfac) > 1e-30f && fabsf(mulfac) < 1e30f;
            const float T = sparse_ok ? nz_threshold(mulfac) : 0.0f;
            unsigned m0 = 0u, m1 = 0u;
#pragma unroll
            for (int x = 0; x < 32; ++x) {
                m0 |= (!(fabsf(w[0][sub_to_reg<32, EXACT>(x)]) <= T) ? 1u : 0u) << x;
                m1 |= (!(fabsf(w[1][sub_to_reg<32, EXACT>(x)]) <= T) ? 1u : 0u) << x;
            }
            s_nzm[tid] = m0;
            s_nzm[tid + F32_T] = m1;
        }
        __syncthreads();
        CVX_PH(3);
        // rows 2*tid, 2*tid+1 -> list positions (exclusive scan of the row populations)
        int n = F32_LCAP + 1;
        if (sparse_ok) {
            const int c0 = __popc(s_nzm[2 * tid]), c1 = __popc(s_nzm[2 * tid + 1]);
            const int base = cta_excl_sum<F32_T>(c0 + c1, scratch, n);
            s_base[2 * tid] = base;
            s_base[2 * tid + 1] = base + c0;
        }
        if (n > F32_LCAP) {
            // ---- dense block: the CTA-wide coder on the shared-memory block (also raw fallback, CvxCompress.cpp:350-360) ----
            __syncthreads();
            auto get8 = [&](int g, float* c) {
                const int i = g * 8;
                float4 a = *reinterpret_cast<const float4*>(&sm[physr(i, 5)]);
                float4 b = *reinterpret_cast<const float4*>(&sm[physr(i + 4, 5)]);
                c[0] = a.x; c[1] = a.y; c[2] = a.z; c[3] = a.w; c[4] = b.x; c[5] = b.y; c[6] = b.z; c[7] = b.w;
            };
            encode_tail<F32_T>(p, t, sm, mulfac, get8, [](long i) { return (long)physr((int)i, 5); },
                               reinterpret_cast<int4*>(s_list), scratch, &s_off);
            if (tid == 0) {
                s_ticket[(it + 1) & 1] = nt;
                if (use_tma && nt < p.nblk) { fence_proxy_async(); issue_load(nt); }
            }
            __syncthreads();
            continue;
        }
        // ---- sparse block: (index, value) list in raster order.  Thread t owns rows t and t + 512: a row with up to 4
        //      non-zeros is unpacked by its thread; denser rows (they cluster in the low-z head of the block) are queued
        //      and compacted by whole warps, one lane per cell, dealt round-robin so that no warp carries the head alone ----
        if (tid == 0) s_dqn = 0;
        __syncthreads();                 // s_base complete
        CVX_PH(4);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int row = tid + F32_T * h;
            unsigned m = s_nzm[row];
            if (__popc(m) > 4) {
                s_dq[atomicAdd(&s_dqn, 1)] = (unsigned short)row;
            } else {
                const float* rp = sm + row * 32;
                const int sw = (row & 7) << 2;
                int at = s_base[row];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    if (m) {
                        const int x = __ffs(m) - 1;
                        m &= m - 1;
                        s_list[at++] = make_uint2((unsigned)(row * 32 + x), __float_as_uint(rp[x ^ sw]));
                    }
                }
            }
        }
        __syncthreads();
        for (int i = warp; i < s_dqn; i += F32_T / 32) {
            const int rr = s_dq[i];
            const unsigned mm = s_nzm[rr];
            if ((mm >> lane) & 1u)
                s_list[s_base[rr] + __popc(mm & lt)] = make_uint2((unsigned)(rr * 32 + lane), __float_as_uint(sm[rr * 32 + (lane ^ ((rr & 7) << 2))]));
        }
        __syncthreads();                 // list complete; nobody reads the block buffer any more
        if (tid == 0) {                  // the next block streams in while this one is coded
            s_ticket[(it + 1) & 1] = nt;
            if (use_tma && nt < p.nblk) { fence_proxy_async(); issue_load(nt); }
        }
        f32_mark_groups<F32_T>(s_list, n, mulfac);
        __syncthreads();
        CVX_PH(5);
        // ---- tokens per list entry: chunks of 32 consecutive entries are dealt round-robin to the warps (the dense
        //      head of a block -- group-of-8 candidates, wide values -- is spread over all of them); byte offsets by a
        //      shuffle scan inside the chunk + one scan of the (<= 128) chunk totals ----
        const int nchunks = (n + 31) >> 5;
        for (int c = warp; c < nchunks; c += F32_T / 32) {
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
            // park the entry's offset inside its chunk next to the index (read back by the emit pass)
            if (e < n) s_list[e].x = ent.x | ((unsigned)(incl - bytes) << 15);
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
        CVX_PH(6);
        const int tail_run = F32_B - 1 - (n > 0 ? (int)(s_list[n - 1].x & F32_IDX_MASK) : -1);      // final flush (:294)
        const int tail_pos = len;
        len += run_bytes(tail_run);
        if (tid == 0) publish_aggregate(p, t, (unsigned long long)len);
        const bool direct = len > F32_STAGE - 8;
        unsigned long long excl_now = 0;
        if (direct) excl_now = wide_lookback(p, t, (unsigned long long)len, false, s_lb);     // rare: ratio < ~5
        const bool fits = !direct || (long)(excl_now + (unsigned long long)len) <= p.cap;
        unsigned char* out = direct ? p.payload + excl_now : s_stage + buf * F32_STAGE;
        if (fits) {
            auto put = [&](int ps, uint8_t b) { out[ps] = b; };
            for (int c = warp; c < nchunks; c += F32_T / 32) {
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
        CVX_PH(7);
        complete_pending();
        CVX_PH(8);
        if (!direct) { pend_t = t; pend_len = len; buf ^= 1; }
        __syncthreads();
        CVX_PH(9);
    }
    __syncthreads();
    complete_pending();
    CVX_PH_FLUSH(0);
}
#undef V_

// cvx_encode32ws_transform.inl -- body of k_encode32ws for the transform group (threads 0..511); see cvx_encode32ws.inl.
// Same pipeline as k_encode32 up to the non-zero list; barriers are the group's named barrier.
{
    asm volatile("setmaxnreg.inc.sync.aligned.u32 96;");
    const int warp = tid >> 5;
    uint32_t parity = 0;
    int msgs = 0;                        // messages posted so far; message i uses list buffer i & 1
    CVX_PH_DECL;
    for (int it = 0;; ++it) {
        const long t = s_ticket[it & 1];
        if (t >= p.nblk) break;
        unsigned int nt = 0;
        if (tid == 0) nt = atomicAdd(p.ticket, 1u);            // the next block's ticket: round trip hidden under the transform
        mbar_wait(&s_bar[warp >> 3], parity);                  // the boxes holding z slices warp and warp + 16
        mbar_wait(&s_bar[2 + (warp >> 3)], parity);
        parity ^= 1u;
        CVX_PH(0);
        f32_pass_x<EXACT, true>(sm);
        __syncwarp();                    // warp w owns the x rows and the y lines of z slices w and w + 16
        CVX_PH(1);
        f32_pass_y2<EXACT, true>(sm);
        ETBar::sync();
        CVX_PH(2);
        typename PairOf<EXACT>::type vz[32];                   // z lines in registers: (x = lane, y = warp + 16 k)
        {
            if (warp >= 8) asm volatile("bar.sync 7, 512;" ::: "memory");      // second half starts its loads when the first half's are issued
            const float* zl = f32_zline(sm, tid);
#pragma unroll
            for (int j = 0; j < 32; ++j) p_set(vz[j], zl[j * 1024], zl[j * 1024 + 512]);
            if (warp < 8) asm volatile("bar.arrive 7, 512;" ::: "memory");
            fwd_line<32, EXACT>(vz);
        }
        float mulfac = gmul;
        if (p.use_local) {               // Compute_Local_RMS, CvxCompress.cpp:119-142, :343-348 (LIFTING: parallel fp32 reduction)
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                float* zl = f32_zline(sm, tid + E32_TT * k);
#pragma unroll
                for (int r = 0; r < 32; ++r) zl[reg_to_sub<32, EXACT>(r) * 1024] = k ? p_hi(vz[r]) : p_lo(vz[r]);
            }
            ETBar::sync();
            const float ss = cta_sumsq<E32_TT, ETBar>(sm, F32_B, s_red);
            const float rms = sqrtf(__fdiv_rn(ss, (float)F32_B));
            mulfac = rms != 0.0f ? __fdiv_rn(1.0f, __fmul_rn(rms, p.scale)) : 1.0f;
            if (tid == 0) p.blkmulfac[t] = mulfac;
        }
        const bool sparse_ok = fabsf(mulfac) > 1e-30f && fabsf(mulfac) < 1e30f;
        const float T = sparse_ok ? nz_threshold(mulfac) : 0.0f;
        if (!p.use_local) {              // coefficients back to the block buffer, one ballot per row
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                float* zl = f32_zline(sm, tid + E32_TT * k);
#pragma unroll
                for (int r = 0; r < 32; ++r) {
                    const float c = k ? p_hi(vz[r]) : p_lo(vz[r]);
                    zl[reg_to_sub<32, EXACT>(r) * 1024] = c;
                    const unsigned m = __ballot_sync(0xffffffffu, !(fabsf(c) <= T));
                    if (lane == 0) s_nzm[reg_to_sub<32, EXACT>(r) * 32 + warp + 16 * k] = m;
                }
            }
        } else {                         // local RMS: the coefficients are already there
#pragma unroll
            for (int k = 0; k < 2; ++k) {
#pragma unroll
                for (int r = 0; r < 32; ++r) {
                    const float c = k ? p_hi(vz[r]) : p_lo(vz[r]);
                    const unsigned m = __ballot_sync(0xffffffffu, !(fabsf(c) <= T));
                    if (lane == 0) s_nzm[reg_to_sub<32, EXACT>(r) * 32 + warp + 16 * k] = m;
                }
            }
        }
        ETBar::sync();
        CVX_PH(3);
        int n = F32_LCAP + 1;
        if (sparse_ok) {                 // rows 2*tid, 2*tid+1 -> list positions
            const int c0 = __popc(s_nzm[2 * tid]), c1 = __popc(s_nzm[2 * tid + 1]);
            const int base = cta_excl_sum<E32_TT, ETBar>(c0 + c1, scratch, n);
            s_base[2 * tid] = base;
            s_base[2 * tid + 1] = base + c0;
        }
        if (tid == 0) s_dqn = 0;
        const int lb = msgs & 1;
        uint2* s_list = s_lists + lb * F32_LCAP;
        if (msgs >= 2) e32_wait(E32_BAR_EMPTY + lb);           // the coder has read this buffer's previous list
        else ETBar::sync();                                    // s_base, s_dqn visible
        CVX_PH(4);
        if (n > F32_LCAP) {
            if (tid == 0) { s_msg[lb].t = t; s_msg[lb].n = -1; s_msg[lb].mulfac = mulfac; }
            e32_arrive(E32_BAR_FULL + lb);
            dense_block(t, mulfac, s_list);                    // together with the coder group; ends with a CTA barrier
            if (tid == 0) {
                s_ticket[(it + 1) & 1] = nt;
                if (nt < p.nblk) { fence_proxy_async(); issue_load(nt); }
            }
            ++msgs;
            ETBar::sync();
            continue;
        }
        // ---- (index, value) list in raster order: rows with up to 4 non-zeros by their thread, denser rows by whole warps ----
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int row = tid + E32_TT * h;
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
        ETBar::sync();
        for (int i = warp; i < s_dqn; i += E32_TT / 32) {
            const int rr = s_dq[i];
            const unsigned mm = s_nzm[rr];
            if ((mm >> lane) & 1u)
                s_list[s_base[rr] + __popc(mm & lt)] = make_uint2((unsigned)(rr * 32 + lane), __float_as_uint(sm[rr * 32 + (lane ^ ((rr & 7) << 2))]));
        }
        ETBar::sync();                   // list complete; nobody reads the block buffer any more
        if (tid == 0) {
            s_msg[lb].t = t; s_msg[lb].n = n; s_msg[lb].mulfac = mulfac;
            s_ticket[(it + 1) & 1] = nt;
            if (nt < p.nblk) { fence_proxy_async(); issue_load(nt); }
        }
        e32_arrive(E32_BAR_FULL + lb);
        ++msgs;
        ETBar::sync();                   // s_ticket visible
        CVX_PH(5);
    }
    // no more blocks: tell the coder, then absorb its last EMPTY arrival so that no barrier is left half-way
    {
        const int lb = msgs & 1;
        if (msgs >= 2) e32_wait(E32_BAR_EMPTY + lb);
        if (tid == 0) { s_msg[lb].t = -1; s_msg[lb].n = 0; s_msg[lb].mulfac = 0.f; }
        e32_arrive(E32_BAR_FULL + lb);
        if (msgs >= 1) e32_wait(E32_BAR_EMPTY + ((msgs - 1) & 1));
    }
    CVX_PH_FLUSH(0);
}

// cvx_encode32ws_transform.inl -- body of k_encode32ws for the transform group (threads 0..511); see cvx_encode32ws.inl.
// Passes in the order z, y, x (cvx_fused32.cuh, "LIFTING encoders"): the x rows stay in registers, the block buffer is
// free once every thread has loaded its two rows, and the non-zero list is emitted straight from the registers.
{
    asm volatile("setmaxnreg.inc.sync.aligned.u32 96;");
    uint32_t parity = 0;
    float row_fa, row_fb;                // y / z scalings of the thread's two x rows
    f32_row_factors(row_fa, row_fb);
    int msgs = 0;                        // messages posted so far; message i uses list buffer i & 1
    CVX_PH_DECL;
    for (int it = 0;; ++it) {
        const long t = s_ticket[it & 1];
        if (t >= p.nblk) break;
        unsigned int nt = 0;
        if (tid == 0) nt = atomicAdd(p.ticket, 1u);            // the next block's ticket: round trip hidden under the transform
        mbar_wait(&s_bar[0], parity);                          // the whole block (a z line crosses all four boxes)
        parity ^= 1u;
        CVX_PH(0);
        f32_pass_z2_fwd<EXACT>(sm, true);
        ETBar::sync();
        CVX_PH(1);
        f32_pass_y2<EXACT, true>(sm, true);      // skewed halves: one half's loads / stores overlap the other's math
        __syncwarp();                    // warp w owns the y lines and the x rows of z slices w and w + 16
        CVX_PH(2);
        float w[2][32];                  // coefficients of rows tid and tid + 512: w[h][sub_to_reg(x)]
        {
            float v[2][32];
            f32_rows_load(sm, v);
            ETBar::sync();               // every row is in registers: nobody reads the block buffer any more
            if (tid == 0) {              // the next block streams in under the x math, the scan and the list emission
                s_ticket[(it + 1) & 1] = nt;
                if (nt < p.nblk) { fence_proxy_async(); issue_load(nt); }
            }
            fwd_lift_pair_rs<32>(v[0], v[1], w[0], w[1], row_fa, row_fb);
        }
        float mulfac = gmul;
        if (p.use_local) {               // Compute_Local_RMS, CvxCompress.cpp:119-142, :343-348 (LIFTING: parallel fp32 reduction)
            const float ss = cta_sumsq_rows<E32_TT, ETBar>(w, s_red);
            const float rms = sqrtf(__fdiv_rn(ss, (float)F32_B));
            mulfac = rms != 0.0f ? __fdiv_rn(1.0f, __fmul_rn(rms, p.scale)) : 1.0f;
            if (tid == 0) p.blkmulfac[t] = mulfac;
        }
        const bool sparse_ok = fabsf(mulfac) > 1e-30f && fabsf(mulfac) < 1e30f;
        const float T = sparse_ok ? nz_threshold(mulfac) : 0.0f;
        CVX_PH(3);
        // ---- list positions: rows are raster-consecutive in tid (row tid, then row tid + 512): one scan of both counts ----
        int n = F32_LCAP + 1, b0 = 0, b1 = 0;
        if (sparse_ok) {
            int c0 = 0, c1 = 0;
#pragma unroll
            for (int r = 0; r < 32; ++r) { c0 += !(fabsf(w[0][r]) <= T) ? 1 : 0; c1 += !(fabsf(w[1][r]) <= T) ? 1 : 0; }
            int tot;
            const int ex = cta_excl_sum<E32_TT, ETBar>(c0 | (c1 << 16), scratch, tot);     // <= 16384 per half: no carry
            const int tot0 = tot & 0xFFFF;
            n = tot0 + (tot >> 16);
            b0 = ex & 0xFFFF;
            b1 = tot0 + (ex >> 16);
        }
        const int lb = msgs & 1;
        uint2* s_list = s_lists + lb * F32_LCAP;
        if (msgs >= 2) e32_wait(E32_BAR_EMPTY + lb);           // the coder has read this buffer's previous list
        CVX_PH(4);
        if (n > F32_LCAP) {
            // dense block (or a multiplier outside the sparse path's range): the generic coder wants the coefficients in the
            // block buffer, which may already be receiving the next block -- let that load land, overwrite it with the rows
            // from the registers, code together with the coder group, then fetch the next block again
            ETBar::sync();                                     // s_ticket visible on the !sparse_ok way here
            const unsigned int next_t = s_ticket[(it + 1) & 1];
            if (next_t < p.nblk) { mbar_wait(&s_bar[0], parity); parity ^= 1u; }
            f32_rows_store_coef(sm, w);
            if (tid == 0) { s_msg[lb].t = t; s_msg[lb].n = -1; s_msg[lb].mulfac = mulfac; }
            e32_arrive(E32_BAR_FULL + lb);                     // orders every thread's rows before the coder group's reads
            ETBar::sync();                                     // ... and before this group's
            dense_block(t, mulfac, s_list);                    // together with the coder group; ends with a CTA barrier
            if (tid == 0 && next_t < p.nblk) { fence_proxy_async(); issue_load(next_t); }
            ++msgs;
            continue;
        }
        // ---- (index, value) list in raster order, straight from the registers: 64 predicated stores per thread ----
        {
            uint32_t lp0 = smem_u32(s_list + b0), lp1 = smem_u32(s_list + b1);
            const unsigned i0 = (unsigned)tid * 32u, i1 = i0 + (unsigned)E32_TT * 32u;
#pragma unroll
            for (int x = 0; x < 32; ++x) {
                f32_emit_nz(lp0, w[0][sub_to_reg<32, EXACT>(x)], T, i0 | (unsigned)x);
                f32_emit_nz(lp1, w[1][sub_to_reg<32, EXACT>(x)], T, i1 | (unsigned)x);
            }
        }
        if (tid == 0) { s_msg[lb].t = t; s_msg[lb].n = n; s_msg[lb].mulfac = mulfac; }
        e32_arrive(E32_BAR_FULL + lb);   // every thread's entries precede its own arrival: no group barrier needed
        ++msgs;
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

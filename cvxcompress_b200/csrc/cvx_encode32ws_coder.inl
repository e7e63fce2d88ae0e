// cvx_encode32ws_coder.inl -- body of k_encode32ws for the coder group (threads 512..767); see cvx_encode32ws.inl.
// Token sizing, look-back, emission and deferred copy-out of k_encode32, on the group's named barrier.
{
    asm volatile("setmaxnreg.dec.sync.aligned.u32 48;");
    const int ctid = tid - E32_TT, cwarp = ctid >> 5;
    constexpr int CW = E32_CT / 32;
    // the block whose payload sits in a staging buffer awaiting its offset (deferred copy-out)
    long pend_t = -1;
    int pend_len = 0, buf = 0;
    auto complete_pending = [&]() {
        if (pend_t < 0) return;
        const unsigned long long excl = wide_lookback<E32_CT, ECBar>(p, pend_t, (unsigned long long)pend_len, false, s_lb);
        if ((long)(excl + (unsigned long long)pend_len) <= p.cap) copy_out<E32_CT>(p.payload + excl, s_stage + (buf ^ 1) * E32_STAGE, pend_len);
        else if (ctid == 0) *p.err = -3;
        pend_t = -1;
    };
    E32_PH_DECL;
    for (int i = 0;; ++i) {
        const int lb = i & 1;
        e32_wait(E32_BAR_FULL + lb);
        E32_PH(6);
        const long t = s_msg[lb].t;
        const int n = s_msg[lb].n;
        const float mulfac = s_msg[lb].mulfac;
        if (t < 0) break;
        uint2* s_list = s_lists + lb * F32_LCAP;
        if (n < 0) {                     // dense block: all 768 threads
            dense_block(t, mulfac, s_list);
            e32_arrive(E32_BAR_EMPTY + lb);
            continue;
        }
        f32_mark_groups<E32_CT>(s_list, n, mulfac);
        ECBar::sync();
        E32_PH(7);
        // ---- tokens per list entry: chunks of 32 consecutive entries dealt round-robin to the warps; byte offsets by a
        //      shuffle scan inside the chunk + one scan of the (<= 128) chunk totals ----
        const int nchunks = (n + 31) >> 5;
        for (int c = cwarp; c < nchunks; c += CW) {
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
        ECBar::sync();
        int len;
        int cpre[4];                     // lane l: exclusive byte offsets of chunks 4l .. 4l+3 (every warp computes all of them)
        {
            int cs[4], tot = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) { cs[k] = (4 * lane + k < nchunks) ? s_chunk[4 * lane + k] : 0; tot += cs[k]; }
            int sc = tot;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int y = __shfl_up_sync(0xffffffffu, sc, d);
                if (lane >= d) sc += y;
            }
            len = __shfl_sync(0xffffffffu, sc, 31);
            cpre[0] = sc - tot;
#pragma unroll
            for (int k = 1; k < 4; ++k) cpre[k] = cpre[k - 1] + cs[k - 1];
        }
        const int tail_run = F32_B - 1 - (n > 0 ? (int)(s_list[n - 1].x & F32_IDX_MASK) : -1);      // final flush (:294)
        const int tail_pos = len;
        len += run_bytes(tail_run);
        if (ctid == 0) publish_aggregate(p, t, (unsigned long long)len);
        E32_PH(8);
        const bool direct = len > E32_STAGE - 8;
        unsigned long long excl_now = 0;
        if (direct) excl_now = wide_lookback<E32_CT, ECBar>(p, t, (unsigned long long)len, false, s_lb);     // rare: ratio < ~10
        const bool fits = !direct || (long)(excl_now + (unsigned long long)len) <= p.cap;
        unsigned char* out = direct ? p.payload + excl_now : s_stage + buf * E32_STAGE;
        if (fits) {
            auto put = [&](int ps, uint8_t b) { out[ps] = b; };
            for (int c = cwarp; c < nchunks; c += CW) {
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
                    for (int j = 0; j < 8; ++j) {
                        const int q = quantise(__fmul_rn(mulfac, __uint_as_float(s_list[e + j].y)));
                        put(pos++, (uint8_t)(q & 0xFF)); put(pos++, (uint8_t)((q >> 8) & 0xFF));
                        if (w == 3) put(pos++, (uint8_t)((q >> 16) & 0xFF));
                    }
                }
            }
            if (ctid == 0) put_run(tail_run, tail_pos, put);
        } else if (ctid == 0) {
            *p.err = -3;
        }
        e32_arrive(E32_BAR_EMPTY + lb);  // this thread has finished reading the list
        E32_PH(9);
        // the PREVIOUS staged block gets its offset now (its neighbours published a block time ago) and is copied out
        complete_pending();
        E32_PH(10);
        if (!direct) { pend_t = t; pend_len = len; buf ^= 1; }
        ECBar::sync();                   // staging buffer complete before the next block's copy-out reads it
    }
    complete_pending();
    E32_PH_FLUSH;
}

            const int wasted = sh.wasted;
            const bool lean = sh.lean != 0;
            const int sbps = a.bps - wasted;
            const int lpc_order = sh.lpc_order;
            const int lshift = sh.shift;
            if (WIDE && !lean) {
                fat_search<HISTN>(sh, pcm, tabs, tid, nv, nsamp, wasted, po_max, kmax);
            } else {
                // ---- K5: LPC residuals (kept in registers) and their |r| sums ----
                const bool lpc_try0 = sh.lpc_ok != 0;
                if (lpc_try0) {
                    int32_t s[HISTN + FRX_SPT], q[HISTN];
                    read_pcm<HISTN>(pcm, tid, wasted, s);
#pragma unroll
                    for (int j = 0; j < HISTN; j++) q[j] = sh.qlp[j];
                    uint32_t ab;
                    if (!WIDE || sh.lpc_narrow) ab = lpc_residuals_park<0, HISTN>(s, q, lshift, res, tid, plain, nv, lpc_order - g0);
                    else ab = lpc_residuals_park<1, HISTN>(s, q, lshift, res, tid, plain, nv, lpc_order - g0);
                    tabs[5 * FRX_ENC_THREADS + tid] = ab;
                }
                compute_sync();  // SC
                DBG_PH(8);
                if (lpc_try0) {
                    if (full) { if (warp <= po_max) analyse_level(sh, tabs + 5 * FRX_ENC_THREADS, 1, warp, lpc_order, kmax, 0); }
                    else if (warp == 1) analyse_partitions(sh, tabs + 5 * FRX_ENC_THREADS, 1, nsamp, lpc_order, po_max, kmax, 0);
                }
                compute_sync();  // SD
                DBG_PH(12);
            }
            const int fixed_order = sh.fixed_order;
            const bool is_const = sh.is_const != 0;
            const bool lpc_try = sh.lpc_ok != 0;

            // ---- choose ----
            const uint64_t verb_bits = 8ull + (uint64_t)wasted + (uint64_t)nsamp * (uint64_t)sbps;
            int type = 1, order = 0, cand = 0, po = 0;
            bool rice2 = false;
            uint64_t best_bits = verb_bits;
            if (is_const) { type = 0; best_bits = 8ull + wasted + sbps; }
            else {
                if (fixed_order >= 0) {
                    int lp; uint64_t lb; bool l2;
                    pick_level(sh, 0, po_max, lp, lb, l2);
                    const uint64_t b = 8ull + wasted + (uint64_t)fixed_order * sbps + lb;
                    if (b < best_bits) { best_bits = b; type = 2; order = fixed_order; cand = 0; po = lp; rice2 = l2; }
                }
                if (lpc_try) {
                    int lp; uint64_t lb; bool l2;
                    pick_level(sh, 1, po_max, lp, lb, l2);
                    const uint64_t b = 8ull + wasted + 9ull + (uint64_t)lpc_order * (uint64_t)(sh.precision + sbps) + lb;
                    if (b < best_bits) { best_bits = b; type = 3; order = lpc_order; cand = 1; po = lp; rice2 = l2; }
                }
            }
            // ---- K7a: winner's residuals (zig-zag folded), exact bit count per thread, block scan ----
            const uint32_t sub_start = bitpos;
            const uint32_t smask = sbps >= 32 ? 0xFFFFFFFFu : ((1u << sbps) - 1u);
            int psize = nsamp;
            uint32_t kk = 0;
            uint32_t mybits = 0;
            const int skip = (tid == 0) ? order : 0;
            if (type >= 2) {
                psize = nsamp >> po;
                kk = sh.kparam[cand][(1 << po) - 1 + min(g0 / psize, (1 << po) - 1)];
                int32_t r[FRX_SPT];
                if (type == 2 || (WIDE && !lean)) {
                    // fixed winner (4 ops per sample) or a fat-path LPC winner: recompute from the PCM, park in the residual buffer
                    int32_t s[HISTN + FRX_SPT];
                    read_pcm<HISTN>(pcm, tid, wasted, s);
                    if (type == 2) {
                        if (WIDE && !lean) (void)fixed_residuals<true, HISTN>(order, s, r);
                        else (void)fixed_residuals<false, HISTN>(order, s, r);
                    } else if (WIDE) {
                        int32_t q[HISTN];
#pragma unroll
                        for (int j = 0; j < HISTN; j++) q[j] = sh.qlp[j];
                        (void)lpc_residuals_dispatch<1, HISTN>(order, s, q, lshift, r);
                    }
#pragma unroll
                    for (int i = 0; i < FRX_SPT; i++) res[RES_IDX(i, tid)] = r[i];
                } else {
                    // lean LPC winner: K5 left its residuals in the buffer
#pragma unroll
                    for (int i = 0; i < FRX_SPT; i++) r[i] = res[RES_IDX(i, tid)];
                }
                if (plain) {
                    uint32_t qs = 0;
#pragma unroll
                    for (int i = 0; i < FRX_SPT; i++) qs += zigzag(r[i]) >> kk;
                    mybits = qs + (uint32_t)FRX_SPT * (1u + kk);
                } else {
#pragma unroll
                    for (int i = 0; i < FRX_SPT; i++)
                        if (i >= skip && i < nv) mybits += (zigzag(r[i]) >> kk) + 1u + kk;
                }
                if (nv > 0 && g0 % psize == 0) mybits += rice2 ? 5u : 4u;
                if (tid == 0) {
                    mybits += 8u + (uint32_t)wasted + (uint32_t)order * (uint32_t)sbps + 6u;
                    if (type == 3) mybits += 9u + (uint32_t)order * (uint32_t)sh.precision;
                }
            } else if (type == 1) {
                mybits = (uint32_t)nv * (uint32_t)sbps;
                if (tid == 0) mybits += 8u + (uint32_t)wasted;
            } else {
                if (tid == 0) mybits = 8u + (uint32_t)wasted + (uint32_t)sbps;
            }
            uint32_t incl = mybits;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t n = __shfl_up_sync(FULLMASK, incl, o);
                if (lane >= o) incl += n;
            }
            if (lane == 31) sh.scan[warp] = incl;
            compute_sync();  // SE
            DBG_PH(13);
            uint32_t wbase = 0, total = 0;
#pragma unroll
            for (int w2 = 0; w2 < NWARPS; w2++) {
                const uint32_t v = sh.scan[w2];
                if (w2 < warp) wbase += v;
                total += v;
            }
            if (type >= 2 && (uint64_t)total >= verb_bits) {
                // the Rice-coded subframe would not beat VERBATIM: write it verbatim (keeps the frame inside its buffer)
                type = 1;
                mybits = (uint32_t)nv * (uint32_t)sbps;
                if (tid == 0) mybits += 8u + (uint32_t)wasted;
                incl = mybits;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t n = __shfl_up_sync(FULLMASK, incl, o);
                    if (lane >= o) incl += n;
                }
                if (lane == 31) sh.scan2[warp] = incl;
                compute_sync();
                wbase = 0; total = 0;
#pragma unroll
                for (int w2 = 0; w2 < NWARPS; w2++) {
                    const uint32_t v = sh.scan2[w2];
                    if (w2 < warp) wbase += v;
                    total += v;
                }
            }
            const uint32_t start = sub_start + wbase + incl - mybits;
            bitpos = sub_start + total;
            if (ch == n_channels - 1 && tid == 0) {
                // the frame's size is final: hand it to the control warp so the look-back overlaps packing and the CRC
                const uint32_t nb = ((bitpos + 7u) >> 3) + 2u;
                if (f != 0) frx_st_relaxed_u64(&a.desc[f], (1ull << 62) | (uint64_t)nb);
                DBG_TRACE(f, 1);   // flag A: size known, prefix not yet
                sh.frame_bytes[slot] = nb;
                mbar_arrive(&sh.sized[slot]);
            }
            // ---- K7b: pack ----
            if (type >= 2) {
                // every header field sits at a position known in closed form, so the lanes of warp 0 write them side by side
                const uint32_t precision = (uint32_t)sh.precision;
                const uint32_t p_warm = sub_start + 8u + (uint32_t)wasted;
                const uint32_t p_lpc = p_warm + (uint32_t)order * (uint32_t)sbps;
                const uint32_t p_rice = p_lpc + (type == 3 ? 9u + (uint32_t)order * precision : 0u);
                if (tid == 0) {
                    // subframe header: 0 | type(6) | wasted flag, then the wasted-bits count in unary
                    const uint32_t tcode = type == 2 ? (8u | (uint32_t)order) : (32u | (uint32_t)(order - 1));
                    put_field(bits, sub_start, (tcode << 1) | (wasted ? 1u : 0u), 8u);
                    if (wasted) put_field(bits, sub_start + 8u + (uint32_t)wasted - 1u, 1u, 1u);
                    if (type == 3) put_field(bits, p_lpc, ((precision - 1u) << 5) | ((uint32_t)sh.shift & 31u), 9u);
                    put_field(bits, p_rice, ((rice2 ? 1u : 0u) << 4) | (uint32_t)po, 6u);
                }
                if (tid < order) {
                    put_field(bits, p_warm + (uint32_t)tid * (uint32_t)sbps, (uint32_t)(pcm[PCM_IDX(tid)] >> wasted) & smask, (uint32_t)sbps);   // warm-up
                    if (type == 3) put_field(bits, p_lpc + 9u + (uint32_t)tid * precision, (uint32_t)sh.qlp[tid] & ((1u << precision) - 1u), precision);
                }
                if (mybits) {
                    uint32_t p = tid == 0 ? p_rice + 6u : start;
                    if (g0 % psize == 0) {
                        const uint32_t pb = rice2 ? 5u : 4u;
                        if (kk) put_field(bits, p, kk, pb);
                        p += pb;
                    }
                    pack_rice(bits, p, res, tid, kk, skip, nv, plain);
                }
            } else if (mybits) {
                BitW bw;
                bw.init(bits, start);
                if (tid == 0) {
                    bw.put((type == 0 ? 0u : 2u) | (wasted ? 1u : 0u), 8);
                    if (wasted) { if (wasted > 1) bw.skip((uint32_t)wasted - 1u); bw.put(1u, 1); }
                }
                if (type == 0) {
                    if (tid == 0) bw.put((uint32_t)(pcm[PCM_IDX(0)] >> wasted) & smask, sbps);
                } else {
                    for (int i = 0; i < nv; i++) bw.put((uint32_t)(pcm[PCM_IDX(g0 + i)] >> wasted) & smask, sbps);
                }
                bw.finish();
            }
            // the PCM buffer is rewritten by the next channel's K1: every reader of it (verbatim / warm-up packing) must be done
            if (ch + 1 < n_channels) compute_sync();
        }  // channels
        const uint32_t crc_xp = crc_shift_factor(a, (bitpos + 7u) >> 3, tid);
        compute_sync();  // Sa: every subframe is packed
        DBG_PH(14);

        // ---- frame footer: CRC-16 of this frame; copy-out of the PREVIOUS frame (its byte offset has had a frame time to arrive) ----
        const uint32_t body_bytes = (bitpos + 7u) >> 3;
        const long long cc0 = DBG_CLK();
        crc_frame(a, sh, bits, sh.crc_part[buf], body_bytes, tid, crc_xp);

        if (pend) {
            const int pslot = (int)((fseq - 1u) & (FD_SLOTS - 1));
            uint32_t *pbits = bits0 + (size_t)(buf ^ 1) * a.frame_words;
            const long long cp0 = DBG_CLK();
            mbar_wait(&sh.placed[pslot], ((fseq - 1u) / FD_SLOTS) & 1u);
            if (tid == 0) { DBG_TRACE(f, 7); }
            copy_out_frame(a, sh, pbits, sh.crc_part[buf ^ 1], sh.out_offset[pslot], pend_body, tid);
            compute_sync();  // all reads of the previous frame's buffer are done
            for (uint32_t j = tid; j <= (pend_body >> 2); j += FRX_ENC_THREADS) pbits[j] = 0u;
        }
        pend = true;
        pend_body = body_bytes;
        fseq++;
        compute_sync();  // crc_part / cleared buffer visible; the next frame's header words go into a clean buffer
        DBG_PH(5);
    }
    if (pend) {   // flush the last frame
        const int pslot = (int)((fseq - 1u) & (FD_SLOTS - 1));
        const int pbuf = (int)((fseq - 1u) & 1u);
        mbar_wait(&sh.placed[pslot], ((fseq - 1u) / FD_SLOTS) & 1u);
        copy_out_frame(a, sh, bits0 + (size_t)pbuf * a.frame_words, sh.crc_part[pbuf], sh.out_offset[pslot], pend_body, tid);
    }
}


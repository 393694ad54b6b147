template <int DT>
__global__ void __launch_bounds__(32 * DEC_WARPS, 20) frx_decode_kernel(const DecArgs a) {
    typedef typename FrxElem<DT>::T T;
    __shared__ DecWarpSmem wsm[DEC_WARPS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    DecWarpSmem &ws = wsm[warp];
    const uint32_t f = (blockIdx.x * DEC_WARPS + warp) * 32u + (uint32_t)lane;
    bool bad = false, live = f < a.total_frames;
    int status = FRX_E_BITSTREAM;
    int ti = 0;
    frx_dtile t;
    memset(&t, 0, sizeof(t));
    uint32_t fj = 0;
    uint64_t pos = 0, end = 0;
    long long s0 = 0;
    int nsamp = 0, channels = 0, bps = 0;
    uint32_t stereo = 0;
    FrxDequant dq;
    dq.range_f = 0.f; dq.dmin_f = 0.f;
    BitR br;
    if (live) {
        ti = dec_find_tile(a.dprefix, a.n_tiles, f);
        t = a.tiles[ti];
        fj = f - a.dprefix[ti];
        pos = a.frame_pos[f];
        if (pos == FRX_NOPOS || pos < t.byte_offset || pos >= t.byte_offset + t.byte_size) bad = true;
        end = t.byte_offset + t.byte_size;
        if (!bad && fj + 1 < t.n_frames) {
            const uint64_t nx = a.frame_pos[f + 1];
            if (nx != FRX_NOPOS && nx > pos && nx <= end) end = nx;
        }
        dq.range_f = (float)(t.dmax - t.dmin);
        dq.dmin_f = (float)t.dmin;
    }
    br.init(a.bytes, a.n_bytes, pos, &ws.ring[0][lane], live && !bad);   // warp-wide: fills the input rings
    if (live && !bad) {
        const long long tile_n = (long long)t.h * t.w;
        s0 = (long long)fj * a.blocksize;
        const int expect = (int)min((long long)a.blocksize, tile_n - s0);
        // ---- K10: frame header ----
        const uint32_t sync = br.get(16);
        if (sync != 0xFFF8u) bad = true;
        const uint32_t bs_code = br.get(4), sr_code = br.get(4), ch_code = br.get(4), sz_code = br.get(3);
        br.get(1);
        {
            const uint32_t x = br.get(8);
            const int extra = !(x & 0x80u) ? 0 : (x & 0xE0u) == 0xC0u ? 1 : (x & 0xF0u) == 0xE0u ? 2 : (x & 0xF8u) == 0xF0u ? 3 : (x & 0xFCu) == 0xF8u ? 4 : 5;
            for (int e = 0; e < extra; e++) br.get(8);
        }
        if (bs_code == 1) nsamp = 192;
        else if (bs_code >= 2 && bs_code <= 5) nsamp = 576 << (bs_code - 2);
        else if (bs_code == 6) nsamp = (int)br.get(8) + 1;
        else if (bs_code == 7) nsamp = (int)br.get(16) + 1;
        else if (bs_code >= 8) nsamp = 256 << (bs_code - 8);
        else { nsamp = 0; bad = true; }
        if (sr_code == 12u) br.get(8); else if (sr_code == 13u || sr_code == 14u) br.get(16);
        br.get(8);  // CRC-8 (validated by discovery for foreign files; the CRC-16 pass covers the header too)
        bps = sz_code ? frx_size_code_bps(sz_code) : t.bps;
        channels = ch_code < 8u ? (int)ch_code + 1 : 2;
        if (nsamp != expect || bps != t.bps || channels != t.channels || ch_code > 10u) bad = true;
        if (ch_code >= 8u && !bad) {
            stereo = ch_code - 7u;                       // 1 left/side, 2 side/right, 3 mid/side
            if (!a.stereo_scratch || bps >= 32) { bad = true; status = FRX_E_UNSUPPORTED; }
        }
    }
    int my_n = (live && !bad) ? nsamp : 0;
    const int my_ch = (live && !bad) ? channels : 0;
    const int nmax = __reduce_max_sync(0xFFFFFFFFu, my_n), chmax = __reduce_max_sync(0xFFFFFFFFu, my_ch);
    const bool any_stereo = __any_sync(0xFFFFFFFFu, stereo != 0u && my_n > 0);
    const int w = my_n ? t.w : (1 << 30);
    const int row0 = my_n ? (int)(s0 / t.w) : 0, col0 = my_n ? (int)(s0 - (long long)row0 * t.w) : 0;

    for (int c = 0; c < chmax; c++) {
        SubF sf;
        sf.type = 3; sf.order = 0; sf.shift = 0; sf.wasted = 0; sf.sbps = bps; sf.cval = 0;
        sf.rs.k = 0; sf.rs.left = 0; sf.rs.pbits = 4; sf.rs.esc = 15u; sf.rs.rawbits = -1; sf.rs.psize = 1;
        int32_t coef[DEC_MAXORD], h[DEC_MAXORD];
#pragma unroll
        for (int j = 0; j < DEC_MAXORD; j++) { coef[j] = 0; h[j] = 0; }
        bool wide = false;
        const unsigned long long base_addr = (unsigned long long)((T *)a.out + t.out_offset + (long long)c * t.out_band_stride);
        if (c < my_ch && !bad) {
            // ---- K11 front: subframe header, warm-up samples, predictor ----
            const uint32_t hdr = br.get(8);
            const uint32_t type = (hdr >> 1) & 0x3Fu;
            if (hdr & 0x80u) bad = true;
            if (hdr & 1u) sf.wasted = (int)br.unary() + 1;
            const int side = ((stereo == 1u && c == 1) || (stereo == 2u && c == 0) || (stereo == 3u && c == 1)) ? 1 : 0;
            sf.sbps = bps + side - sf.wasted;
            if (sf.sbps <= 0) bad = true;
            if (bad) {}
            else if (type == 0) { sf.type = 0; sf.cval = br.sget(sf.sbps); }
            else if (type == 1) sf.type = 1;
            else {
                bool lpc = false;
                int order = 0, prec = 3;
                if (type >= 8 && type <= 12) order = (int)type - 8;
                else if (type >= 32) { order = (int)(type & 31u) + 1; lpc = true; }
                else bad = true;
                if (order > nsamp) bad = true;
                if (!bad && order > DEC_MAXORD && stereo) { bad = true; status = FRX_E_UNSUPPORTED; }
                if (!bad && order > DEC_MAXORD) {
                    // rare long predictors: this lane restores its whole subframe on its own, then idles through the lockstep part
                    int32_t coefl[32], warml[32];
                    for (int i = 0; i < order; i++) warml[i] = br.sget(sf.sbps);
                    prec = (int)br.get(4) + 1;
                    const int shift = br.sget(5);
                    if (prec == 16 || shift < 0) bad = true;
                    for (int j = 0; j < order; j++) coefl[j] = br.sget(prec);
                    const uint32_t method = br.get(2), po = br.get(4);
                    if (method > 1u || (po > 0 && ((uint32_t)nsamp >> po) << po != (uint32_t)nsamp) || ((uint32_t)nsamp >> po) < (uint32_t)order) bad = true;
                    if (!bad) {
                        RiceState rs;
                        rs.pbits = method ? 5 : 4; rs.esc = method ? 31u : 15u; rs.psize = (uint32_t)nsamp >> po; rs.rawbits = -1;
                        rs.k = br.get(rs.pbits);
                        rs.left = rs.psize - (uint32_t)order;
                        if (rs.k == rs.esc) rs.rawbits = (int)br.get(5);
                        BitR brl = br;     // only the copy's address escapes into the out-of-line routine: br itself stays in registers
                        restore_subframe_generic<DT>(brl, rs, (T *)base_addr, t.out_row_stride, t.w, s0, coefl, order, shift, warml, nsamp, sf.wasted,
                                                     bps, dq, a.pcm_mode);
                        br = brl;
                    }
                } else if (!bad) {
#pragma unroll
                    for (int i = 0; i < DEC_MAXORD; i++) if (i < order) h[i] = br.sget(sf.sbps);
                    if (lpc) {
                        prec = (int)br.get(4) + 1;
                        sf.shift = br.sget(5);
                        if (prec == 16 || sf.shift < 0) bad = true;
#pragma unroll
                        for (int j = 0; j < DEC_MAXORD; j++) if (j < order) coef[j] = br.sget(prec);
                    } else {
                        if (order == 1) { coef[0] = 1; }
                        else if (order == 2) { coef[0] = 2; coef[1] = -1; }
                        else if (order == 3) { coef[0] = 3; coef[1] = -3; coef[2] = 1; }
                        else if (order == 4) { coef[0] = 4; coef[1] = -6; coef[2] = 4; coef[3] = -1; }
                    }
                    const uint32_t method = br.get(2), po = br.get(4);
                    if (method > 1u || (po > 0 && ((uint32_t)nsamp >> po) << po != (uint32_t)nsamp) || ((uint32_t)nsamp >> po) < (uint32_t)order) bad = true;
                    if (!bad) {
                        sf.type = 2; sf.order = order;
                        sf.rs.pbits = method ? 5 : 4; sf.rs.esc = method ? 31u : 15u; sf.rs.psize = (uint32_t)nsamp >> po;
                        // the first partition is short by `order`: prime it here
                        sf.rs.k = br.get(sf.rs.pbits);
                        sf.rs.left = sf.rs.psize - (uint32_t)order;
                        sf.rs.rawbits = -1;
                        if (sf.rs.k == sf.rs.esc) sf.rs.rawbits = (int)br.get(5);
                        wide = order > 0 && (sf.sbps + prec + ilog2_u32((uint32_t)order) > 32);
                    }
                }
            }
        }
        const int lane_n = (sf.type != 3 && !bad) ? my_n : 0;
        const int omax = __reduce_max_sync(0xFFFFFFFFu, (uint32_t)(sf.type == 2 && lane_n ? sf.order : 0));
        const bool anywide = __any_sync(0xFFFFFFFFu, wide && lane_n);
        // ---- K11 back + K12: lockstep restore, dequantise, coalesced stores ----
        if (any_stereo) {
            // warps holding decorrelated stereo frames take one general instantiation (16 taps, 64-bit accumulate)
            const uint32_t st_mode = stereo ? (c == 0 ? 1u : 1u + stereo) : 0u;
            const unsigned long long scratch_addr = (unsigned long long)(a.stereo_scratch + (size_t)(live ? f : 0u) * (size_t)a.blocksize);
            run_chunks<DT, 16, true, true>(br, sf, coef, h, ws, lane, lane_n, nmax, base_addr, t.out_row_stride, w, row0, col0, bps, dq, a.pcm_mode,
                                           st_mode, scratch_addr, (long long)t.out_band_stride * (long long)sizeof(T));
        } else if (omax <= 4) {
            if (anywide) run_chunks<DT, 4, true, false>(br, sf, coef, h, ws, lane, lane_n, nmax, base_addr, t.out_row_stride, w, row0, col0, bps, dq, a.pcm_mode);
            else run_chunks<DT, 4, false, false>(br, sf, coef, h, ws, lane, lane_n, nmax, base_addr, t.out_row_stride, w, row0, col0, bps, dq, a.pcm_mode);
        } else if (omax <= 8) {
            if (anywide) run_chunks<DT, 8, true, false>(br, sf, coef, h, ws, lane, lane_n, nmax, base_addr, t.out_row_stride, w, row0, col0, bps, dq, a.pcm_mode);
            else run_chunks<DT, 8, false, false>(br, sf, coef, h, ws, lane, lane_n, nmax, base_addr, t.out_row_stride, w, row0, col0, bps, dq, a.pcm_mode);
        } else {
            if (anywide) run_chunks<DT, 16, true, false>(br, sf, coef, h, ws, lane, lane_n, nmax, base_addr, t.out_row_stride, w, row0, col0, bps, dq, a.pcm_mode);
            else run_chunks<DT, 16, false, false>(br, sf, coef, h, ws, lane, lane_n, nmax, base_addr, t.out_row_stride, w, row0, col0, bps, dq, a.pcm_mode);
        }
        if (bad) my_n = 0;
    }
    if (!live) return;
    // ---- footer: byte align; the frame must end where the next one starts (CRC-16: frx_crc_check_kernel) ----
    const uint64_t fsize = (uint64_t)((br.consumed() + 7u) >> 3) + 2u;
    if (!bad && pos + fsize > end) bad = true;
    if (!bad && fj + 1 < t.n_frames && a.frame_pos[f + 1] != FRX_NOPOS && pos + fsize != a.frame_pos[f + 1]) bad = true;
    if (a.frame_len) a.frame_len[f] = bad ? 0u : (uint32_t)fsize;
    if (bad) atomicExch(&a.tile_status[ti], status);
}

// ---- CRC-16 of every decoded frame: one warp per frame, coalesced reads.  Lane l takes a contiguous run of Lq 16-byte
// quads (the run of lane 0 may be short: virtual leading zeros do not change a CRC); the 32 chunk CRCs are moved to
// their place with x^(128 Lq j) mod P -- the encoder's table -- and xor-ed.  A frame including its stored CRC gives 0.
__global__ void __launch_bounds__(128) frx_crc_check_kernel(const DecArgs a, const uint16_t *xpow) {
    __shared__ uint16_t crc_tab[4][256];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) (&crc_tab[0][0])[i] = a.crc_tab[i];
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t f = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; f < a.total_frames; f += nwarps) {
        const uint32_t fsize = a.frame_len[f];
        if (fsize == 0) continue;                       // the decoder already flagged it
        const uint8_t *p = a.bytes + a.frame_pos[f];
        uint32_t crc = 0;                               // running CRC over the passes (frames above 32 * 48 quads)
        uint32_t done = 0;
        while (done < fsize) {
            const uint32_t pass = min(fsize - done, 32u * 16u * (uint32_t)FRX_CRC_MAXLQ);
            const uint32_t Lq = (pass + 511u) / 512u, C = Lq * 16u;
            // lane l covers pass bytes [pass - (32 - l) C, pass - (31 - l) C), clipped at 0
            const long long b1 = (long long)pass - (long long)(31 - lane) * C;
            long long b0 = b1 - C;
            if (b0 < 0) b0 = 0;
            uint32_t c = 0;
            if (b1 > 0) {
                const uint8_t *q = p + done + b0;
                long long n = b1 - b0;
                while (n > 0 && ((uintptr_t)q & 3u)) { c = ((c << 8) ^ crc_tab[0][((c >> 8) ^ *q) & 0xFFu]) & 0xFFFFu; q++; n--; }
                for (; n >= 4; n -= 4, q += 4) {
                    const uint32_t x = frx_bswap32(__ldg(reinterpret_cast<const uint32_t *>(q))) ^ (c << 16);
                    c = (uint32_t)crc_tab[3][x >> 24] ^ crc_tab[2][(x >> 16) & 0xFFu] ^ crc_tab[1][(x >> 8) & 0xFFu] ^ crc_tab[0][x & 0xFFu];
                }
                for (; n > 0; n--, q++) c = ((c << 8) ^ crc_tab[0][((c >> 8) ^ *q) & 0xFFu]) & 0xFFFFu;
            }
            if (c && lane != 31) c = frx_gf16_mulmod(c, (uint32_t)__ldg(&xpow[(Lq - 1) * 256 + (31 - lane)]));
            c = __reduce_xor_sync(0xFFFFFFFFu, c);
            // crc <- crc * x^(8 pass) + c ; x^(8 pass) = x^(128 Lq 32) when the pass is full, and a full pass is the only one followed by another
            if (done) {
                // previous passes were full: shift the running CRC by this pass's length, 8 bits at a time is too slow -> square-and-multiply
                uint32_t e = pass * 8u, base = 2u, acc = 1u;
                while (e) { if (e & 1u) acc = frx_gf16_mulmod(acc, base); base = frx_gf16_mulmod(base, base); e >>= 1; }
                crc = frx_gf16_mulmod(crc, acc) ^ c;
            } else crc = c;
            done += pass;
        }
        if (crc != 0 && lane == 0) {
            const int ti = dec_find_tile(a.dprefix, a.n_tiles, f);
            atomicExch(&a.tile_status[ti], FRX_E_BITSTREAM);
        }
    }
}



// ==========================================================================================
// The frame encoder.  WIDE: PCM wider than 16 bits (RICE2 parameters, 64-bit LPC accumulate when needed).  H: history
// samples of the register windows = the largest LPC order of the compression level (4 when the level has no LPC stage).
// ==========================================================================================
template <bool WIDE, int H>
__global__ void __launch_bounds__(ENC_CTA, 1) frx_encode_kernel(const EncArgs a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    uint16_t (*crc_tab)[256] = reinterpret_cast<uint16_t (*)[256]>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned char *wbase = smem_raw + 2048 + (size_t)warp * WARP_BYTES;
    WarpSh &ws = *reinterpret_cast<WarpSh *>(wbase);
    uint32_t *buf = reinterpret_cast<uint32_t *>(wbase + WARP_SH_BYTES);
    uint32_t *pcm = buf + SLACK_W;
    for (int i = tid; i < 1024; i += ENC_CTA) (&crc_tab[0][0])[i] = a.crc_tab[i];
    if (lane == 0) {
        mbar_init(&ws.mbar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    uint32_t phase = 0;
    int ti = 0;
    const bool stereo = a.stereo != 0;

    for (;;) {
        uint32_t f = 0;
        if (lane == 0) f = atomicAdd(a.ticket, 1u);
        f = __shfl_sync(FULLMASK, f, 0);
        if (f >= a.total_frames) break;
        build_frame_desc(a, ws.fd, f, ti, lane);
        const int nsamp = ws.fd.nsamp;
        const int n_ch = ws.fd.t.n_channels;
        BufState bs;
        bs.kind = BUF_JUNK; bs.vch = -1; bs.type = 0; bs.order = 0;

        // ---- sweep 1: search every subframe, exact sizes ----
        const int nsub = n_ch;
        int sub0 = 0, sub1 = 1;     // decision records of the first two subframes (the stereo choice permutes them)
        if (n_ch == 2 && stereo) {
            // left, right, mid, side; the cheapest of L/R, L/S, S/R, M/S (libFLAC's exhaustive stereo decision)
            analyse_channel<WIDE, H>(a, ws, buf, 0, 0, ws.fd.hdr_bits, phase, bs, lane);
            analyse_channel<WIDE, H>(a, ws, buf, 1, 1, ws.fd.hdr_bits, phase, bs, lane);
            analyse_channel<WIDE, H>(a, ws, buf, 8, 2, ws.fd.hdr_bits, phase, bs, lane);
            analyse_channel<WIDE, H>(a, ws, buf, 9, 3, ws.fd.hdr_bits, phase, bs, lane);
            const uint32_t bL = ws.dec[0].bits, bR = ws.dec[1].bits, bM = ws.dec[2].bits, bS = ws.dec[3].bits;
            uint32_t best = bL + bR; int ca = 1; sub0 = 0; sub1 = 1;
            if (bL + bS < best) { best = bL + bS; ca = 8; sub0 = 0; sub1 = 3; }
            if (bR + bS < best) { best = bR + bS; ca = 9; sub0 = 3; sub1 = 1; }
            if (bM + bS < best) { best = bM + bS; ca = 10; sub0 = 2; sub1 = 3; }
            if (lane == 0) make_frame_header(a, ws.fd, ca);
            __syncwarp();
        } else {
            for (int c = 0; c < n_ch; c++)
                analyse_channel<WIDE, H>(a, ws, buf, c, c, c == 0 ? ws.fd.hdr_bits : 32u, phase, bs, lane);
        }
        uint32_t frame_bits = ws.fd.hdr_bits;
        for (int c = 0; c < nsub; c++) frame_bits += ws.dec[c == 0 ? sub0 : (c == 1 ? sub1 : c)].bits;
        const uint32_t body_bytes = (frame_bits + 7u) >> 3;
        const uint32_t nbytes = body_bytes + 2u;
        if (lane == 0) {
            // the frame's size is final: publish it, the look-back of the frames behind this one can proceed
            if (f != 0) frx_st_relaxed_u64(&a.desc[f], (1ull << 62) | (uint64_t)nbytes);
            if (a.frame_size) a.frame_size[f] = nbytes;
        }

        // ---- sweep 2: pack ----
        Emitter em;
        em.P = 0; em.carry = 0;
        {
            uint32_t val[1], len[1], zer[1] = { 0 };
            const uint32_t hbits = ws.fd.hdr_bits;
            val[0] = lane < 4 ? ws.fd.hdr[lane] : 0u;
            len[0] = (uint32_t)max(0, min(32, (int)hbits - 32 * lane));
            if (len[0] < 32u && len[0] > 0u) val[0] >>= (32u - len[0]);
            emit_fields<1>(buf, em, val, len, zer, lane);
        }
        bool have_off = false, overflow = false;
        uint64_t off = 0;
        uint32_t done = 0;          // bytes of the frame already copied out (multi-channel frames go out subframe by subframe)
        uint32_t crc_run = 0;
        auto get_offset = [&]() {
            LookBack lb;
            lb.pos = (long long)f - 1; lb.exclusive = 0; lb.blocked = false;
            while (!lookback_step(a, lb, lane)) { if (lb.blocked) __nanosleep(100); }
            off = lb.exclusive;
            if (lane == 0) {
                frx_st_relaxed_u64(&a.desc[f], (2ull << 62) | (off + (uint64_t)nbytes));
                if (a.frame_offset) a.frame_offset[f] = off;
                if (f == a.total_frames - 1 && a.total_bytes) *a.total_bytes = off + nbytes;
            }
            have_off = true;
            if (off + nbytes > a.capacity) {
                overflow = true;
                if (lane == 0) atomicExch(a.status, FRX_E_OVERFLOW);
            }
        };
        for (int c = 0; c < nsub; c++) {
            const SubDec &d = ws.dec[c == 0 ? sub0 : (c == 1 ? sub1 : c)];
            materialise<WIDE, H>(a, ws, buf, d, phase, bs, lane);
            em = pack_subframe_header(ws, buf, em, d, pcm, lane);
            if (d.type >= 2) em = d.runs ? pack_runs(buf, em, d, pcm, lane) : pack_rice(buf, em, d, pcm, nsamp, lane);
            else if (d.type == 1) em = pack_verbatim(buf, em, d, pcm, nsamp, lane);
            bs.kind = BUF_JUNK;                     // the packed bits now lie over the samples
            if (c + 1 < nsub) {
                // the next subframe needs the whole buffer: send the complete words of the stream on their way
                if (!have_off) get_offset();
                const uint32_t nw = em.P >> 5;
                __syncwarp();
                const uint32_t cs = crc_words(a, crc_tab, buf, nw, lane);
                crc_run = (done ? frx_gf16_mulmod(crc_run, gf16_xpow32(nw)) : 0u) ^ cs;
                if (!overflow) copy_words_out(a.bytes + off + done, buf, nw, lane);
                done += nw * 4u;
                em.P &= 31u;
                __syncwarp();
            }
        }
        // ---- frame footer: zero padding to a byte boundary, CRC-16 ----
        {
            const uint32_t p0 = em.P;
            em.P = (em.P + 7u) & ~7u;
            if ((p0 & 31u) && !(em.P & 31u)) {       // the padding completes the last word: it is still in the carry
                if (lane == 0) buf[p0 >> 5] = em.carry;
                em.carry = 0;
            }
        }
        const uint32_t nw = em.P >> 5, tb = (em.P & 31u) >> 3;
        __syncwarp();
        uint32_t crc = crc_words(a, crc_tab, buf, nw, lane);
        if (done) crc ^= frx_gf16_mulmod(crc_run, gf16_xpow32(nw));
        for (uint32_t b = 0; b < tb; b++) {
            const uint32_t byte = (em.carry >> (24 - 8 * b)) & 0xFFu;
            crc = ((crc << 8) ^ crc_tab[0][((crc >> 8) ^ byte) & 0xFFu]) & 0xFFFFu;
        }
        if (!have_off) get_offset();
        if (!overflow) {
            uint8_t *dst = a.bytes + off + done;
            copy_words_out(dst, buf, nw, lane);
            if (lane < (int)tb) dst[nw * 4u + lane] = (uint8_t)((em.carry >> (24 - 8 * lane)) & 0xFFu);
            if (lane == 4) dst[nw * 4u + tb] = (uint8_t)(crc >> 8);
            if (lane == 5) dst[nw * 4u + tb + 1u] = (uint8_t)crc;
        }
        __syncwarp();
    }
}

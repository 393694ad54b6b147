
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

    // Multi-channel frames are STREAMED: a channel is packed right after its analysis (the buffer holds one channel at a time, so
    // packing later would mean staging and predicting it a second time) and its words go to this warp's slot in global memory
    // (the slots of all resident warps together stay in L2).  The frame's size is only known -- and published -- when its last
    // channel has been analysed, so its byte offset is looked up one frame later, just before the NEXT frame needs the slot:
    // a whole channel of slack in which the frames in front publish their sizes.  Single-channel frames (and stereo frames,
    // whose channel assignment needs all four analyses) keep the two sweeps and copy out at once.
    uint8_t *slot = a.slots + (size_t)(blockIdx.x * ENC_WARPS + warp) * a.slot_stride;
    bool pend = false;
    uint32_t pend_f = 0, pend_nbytes = 0;
    auto look_back = [&](uint32_t ff, uint32_t nb, bool &ovf) -> uint64_t {
        LookBack lb;
        lb.pos = (long long)ff - 1; lb.exclusive = 0; lb.blocked = false;
        while (!lookback_step(a, lb, lane)) { if (lb.blocked) __nanosleep(100); }
        const uint64_t off = lb.exclusive;
        if (lane == 0) {
            frx_st_relaxed_u64(&a.desc[ff], (2ull << 62) | (off + (uint64_t)nb));
            if (a.frame_offset) a.frame_offset[ff] = off;
            if (ff == a.total_frames - 1 && a.total_bytes) *a.total_bytes = off + nb;
        }
        ovf = off + nb > a.capacity;
        if (ovf && lane == 0) atomicExch(a.status, FRX_E_OVERFLOW);
        return off;
    };
    auto finish_pending = [&]() {
        bool ovf;
        const uint64_t off = look_back(pend_f, pend_nbytes, ovf);
        if (!ovf) copy_slot_out(a.bytes + off, slot, pend_nbytes, lane);
        pend = false;
    };

    for (;;) {
        uint32_t f = 0;
        if (lane == 0) f = atomicAdd(a.ticket, 1u);
        f = __shfl_sync(FULLMASK, f, 0);
        if (f >= a.total_frames) break;
        build_frame_desc(a, ws.fd, f, ti, lane);
        const int nsamp = ws.fd.nsamp;
        const int n_ch = ws.fd.t.n_channels;
        const bool two_sweeps = n_ch == 1 || (n_ch == 2 && stereo);
        BufState bs;
        bs.kind = BUF_JUNK; bs.vch = -1; bs.type = 0; bs.order = 0;
        Emitter em;
        em.P = 0; em.carry = 0;
        uint32_t done = 0;          // bytes of the frame already written out
        uint32_t crc_run = 0;
        uint32_t nbytes = 0;
        bool overflow = false;
        uint8_t *outp = slot;
        auto emit_frame_header = [&]() {
            uint32_t val[1], len[1], zer[1] = { 0 };
            const uint32_t hbits = ws.fd.hdr_bits;
            val[0] = lane < 4 ? ws.fd.hdr[lane] : 0u;
            len[0] = (uint32_t)max(0, min(32, (int)hbits - 32 * lane));
            if (len[0] < 32u && len[0] > 0u) val[0] >>= (32u - len[0]);
            emit_fields<1>(buf, em, val, len, zer, lane);
        };
        auto publish_size = [&](uint32_t frame_bits) {
            nbytes = ((frame_bits + 7u) >> 3) + 2u;
            if (lane == 0) {
                // the frame's size is final: the look-back of the frames behind this one can proceed
                if (f != 0) frx_st_relaxed_u64(&a.desc[f], (1ull << 62) | (uint64_t)nbytes);
                if (a.frame_size) a.frame_size[f] = nbytes;
            }
        };
        auto pack_subframe = [&](const SubDec &d) {
            em = pack_subframe_header(ws, buf, em, d, pcm, lane);
            if (d.type >= 2) em = d.runs ? pack_runs(buf, em, d, pcm, lane) : pack_rice(buf, em, d, pcm, nsamp, lane);
            else if (d.type == 1) em = pack_verbatim(buf, em, d, pcm, nsamp, lane);
            bs.kind = BUF_JUNK;                     // the packed bits now lie over the samples
        };
        auto flush_words = [&]() {                   // the next channel needs the whole buffer: the complete words leave
            const uint32_t nw = em.P >> 5;
            __syncwarp();
            const uint32_t cs = crc_words(a, crc_tab, buf, nw, lane);
            crc_run = (done ? frx_gf16_mulmod(crc_run, gf16_xpow32(nw)) : 0u) ^ cs;
            if (!overflow) copy_words_out(outp + done, buf, nw, lane);
            done += nw * 4u;
            em.P &= 31u;
            __syncwarp();
        };

        if (two_sweeps) {
            // ---- sweep 1: search every subframe, exact sizes ----
            int sub0 = 0, sub1 = 1;     // decision records of the two subframes (the stereo choice permutes them)
            if (n_ch == 2) {
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
                publish_size(ws.fd.hdr_bits + ws.dec[sub0].bits + ws.dec[sub1].bits);
            } else {
                analyse_channel<WIDE, H>(a, ws, buf, 0, 0, ws.fd.hdr_bits, phase, bs, lane);
                publish_size(ws.fd.hdr_bits + ws.dec[0].bits);
            }
            if (pend) finish_pending();
            // ---- sweep 2: pack; the look-back runs after it (the sizes in front have had a pack time to arrive) ----
            emit_frame_header();
            for (int c = 0; c < n_ch; c++) {
                const SubDec &d = ws.dec[c == 0 ? sub0 : sub1];
                materialise<WIDE, H>(a, ws, buf, d, phase, bs, lane);
                pack_subframe(d);
                if (c + 1 < n_ch) {
                    if (outp == slot) { const uint64_t off = look_back(f, nbytes, overflow); outp = a.bytes + off; }
                    flush_words();
                }
            }
        } else {
            // ---- streamed: analyse, pack and send each channel on its way ----
            for (int c = 0; c < n_ch; c++) {
                analyse_channel<WIDE, H>(a, ws, buf, c, 0, c == 0 ? ws.fd.hdr_bits : (em.P & 31u), phase, bs, lane);
                const SubDec &d = ws.dec[0];
                if (c == 0) emit_frame_header();
                if (c + 1 == n_ch) publish_size(done * 8u + em.P + d.bits);
                materialise<WIDE, H>(a, ws, buf, d, phase, bs, lane);      // (a no-op unless the subframe fell back to VERBATIM)
                pack_subframe(d);
                if (c + 1 < n_ch) {
                    if (pend) finish_pending();                             // the slot is about to be written
                    flush_words();
                }
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
        if (two_sweeps && outp == slot) { const uint64_t off = look_back(f, nbytes, overflow); outp = a.bytes + off; }
        if (!overflow) {
            uint8_t *dst = outp + done;
            copy_words_out(dst, buf, nw, lane);
            if (lane < (int)tb) dst[nw * 4u + lane] = (uint8_t)((em.carry >> (24 - 8 * lane)) & 0xFFu);
            if (lane == 4) dst[nw * 4u + tb] = (uint8_t)(crc >> 8);
            if (lane == 5) dst[nw * 4u + tb + 1u] = (uint8_t)crc;
        }
        if (!two_sweeps) { pend = true; pend_f = f; pend_nbytes = nbytes; }
        __syncwarp();
    }
    if (pend) finish_pending();
}


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
    // Phase alignment for the narrow kernels: with the frame warps of an SM each in a phase of its own, the 8/16-bit
    // workloads ran 1.6-2.3x slower (instruction fetch: 18 % of the stall samples were no_instruction, all phases stretched,
    // and the frames drifted so far apart that nearly every look-back had to wait); in step they share the instruction cache
    // and publish their sizes together.  The wide kernels (float32 / 32-bit) lose a few per cent to the waiting at the
    // barriers instead (their load phase is exposed), so they run free (profiles/r2_encode_experiments.md).
    constexpr bool ALIGNED = !WIDE;

    // Where the bytes go.  A frame's place in the output is the sum of the sizes of all frames in front of it (decoupled
    // look-back over the descriptors), and a frame whose predecessors have not all published their sizes must not hold up the
    // warp: it is PARKED in one of the warp's two slots in global memory (L2-resident) and the warp goes on to its next frame;
    // the parked frame is sent to its final place as soon as a (non-blocking) look-back completes, at the latest when the next
    // frame is finished -- a whole frame of slack, so that the blocking wait is practically never taken.
    // Multi-channel frames are STREAMED: a channel is packed right after its analysis (the buffer holds one channel at a time,
    // packing later would mean staging and predicting it a second time) and the complete words go to the current slot; the
    // frame's size is published when its last channel has been analysed.  Single-channel frames and stereo frames (whose
    // channel assignment needs all four analyses) run two sweeps: analyse, publish the size, pack.
    uint8_t *slots2 = a.slots + (size_t)(blockIdx.x * ENC_WARPS + warp) * 2u * a.slot_stride;
    int cur = 0;                          // the slot the next streamed / parked frame uses
    bool pend = false;                    // a parked frame waits in pend_slot
    uint32_t pend_f = 0, pend_nbytes = 0;
    uint8_t *pend_slot = slots2;
    LookBack plb;                         // its look-back, resumed where it stopped
    plb.pos = -1; plb.exclusive = 0; plb.blocked = false;
    auto lb_run = [&](LookBack &lb, bool block) -> bool {      // true: lb.exclusive is the byte offset
        for (;;) {
            if (lookback_step(a, lb, lane)) return true;
            if (lb.blocked) {
                if (!block) return false;
                DBG_ADD(1, 1);
                __nanosleep(100);
            }
        }
    };
    auto lb_commit = [&](uint32_t ff, uint32_t nb, uint64_t off) -> bool {     // publishes the inclusive prefix; true on overflow
        if (lane == 0) {
            frx_st_relaxed_u64(&a.desc[ff], (2ull << 62) | (off + (uint64_t)nb));
            if (a.frame_offset) a.frame_offset[ff] = off;
            if (ff == a.total_frames - 1 && a.total_bytes) *a.total_bytes = off + nb;
        }
        const bool ovf = off + nb > a.capacity;
        if (ovf && lane == 0) atomicExch(a.status, FRX_E_OVERFLOW);
        return ovf;
    };
    auto finish_pending = [&](bool block) {
        const long long tl0 = DBG_CLK();
        const bool ok = lb_run(plb, block);
        if (block) DBG_ADD(0, DBG_CLK() - tl0);
        if (!ok) return;
        const bool ovf = lb_commit(pend_f, pend_nbytes, plb.exclusive);
        const long long ts0 = DBG_CLK();
        if (!ovf) copy_slot_out(a.bytes + plb.exclusive, pend_slot, pend_nbytes, lane);
        DBG_ADD(2, DBG_CLK() - ts0);
        pend = false;
    };

    for (;;) {
        uint32_t f = 0;
        if (lane == 0) f = atomicAdd(a.ticket, 1u);
        f = __shfl_sync(FULLMASK, f, 0);
        if (ALIGNED) {
            if (!group_or(f < a.total_frames)) break;          // the group's frames start together
            if (f >= a.total_frames) {
                for (int k = 0; k < a.bars_per_frame; k++) group_bar();
                continue;
            }
        } else if (f >= a.total_frames) break;
        const long long tf0 = DBG_CLK();
        build_frame_desc(a, ws.fd, f, ti, lane);
        const int nsamp = ws.fd.nsamp;
        const int n_ch = ws.fd.t.n_channels;
        const bool two_sweeps = n_ch == 1 || (n_ch == 2 && stereo);
        BufState bs;
        bs.kind = BUF_JUNK; bs.vch = -1; bs.type = 0; bs.order = 0;
        int nbar = 0;
        Emitter em;
        em.P = 0; em.carry = 0;
        uint8_t *const slot = slots2 + (size_t)cur * a.slot_stride;
        uint32_t done = 0;          // bytes of the frame already in the slot
        uint32_t crc_run = 0;
        uint32_t nbytes = 0;
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
            ALIGN_BAR(nbar);
        };
        auto flush_words = [&]() {                   // the next channel needs the whole buffer: the complete words go to the slot
            const uint32_t nw = em.P >> 5;
            __syncwarp();
            const uint32_t cs = crc_words(a, crc_tab, buf, nw, lane);
            crc_run = (done ? frx_gf16_mulmod(crc_run, gf16_xpow32(a, nw)) : 0u) ^ cs;
            copy_words_out(slot + done, buf, nw, lane);
            done += nw * 4u;
            em.P &= 31u;
            __syncwarp();
        };

        if (two_sweeps) {
            // ---- sweep 1: search every subframe, exact sizes ----
            int sub0 = 0, sub1 = 1;     // decision records of the two subframes (the stereo choice permutes them)
            if (n_ch == 2) {
                // left, right, mid, side; the cheapest of L/R, L/S, S/R, M/S (libFLAC's exhaustive stereo decision)
                nbar += analyse_channel<WIDE, H>(a, ws, buf, 0, 0, ws.fd.hdr_bits, phase, bs, lane);
                nbar += analyse_channel<WIDE, H>(a, ws, buf, 1, 1, ws.fd.hdr_bits, phase, bs, lane);
                if (pend) finish_pending(false);
                nbar += analyse_channel<WIDE, H>(a, ws, buf, 8, 2, ws.fd.hdr_bits, phase, bs, lane);
                nbar += analyse_channel<WIDE, H>(a, ws, buf, 9, 3, ws.fd.hdr_bits, phase, bs, lane);
                const uint32_t bL = ws.dec[0].bits, bR = ws.dec[1].bits, bM = ws.dec[2].bits, bS = ws.dec[3].bits;
                uint32_t best = bL + bR; int ca = 1; sub0 = 0; sub1 = 1;
                if (bL + bS < best) { best = bL + bS; ca = 8; sub0 = 0; sub1 = 3; }
                if (bR + bS < best) { best = bR + bS; ca = 9; sub0 = 3; sub1 = 1; }
                if (bM + bS < best) { best = bM + bS; ca = 10; sub0 = 2; sub1 = 3; }
                if (lane == 0) make_frame_header(a, ws.fd, ca);
                __syncwarp();
                publish_size(ws.fd.hdr_bits + ws.dec[sub0].bits + ws.dec[sub1].bits);
            } else {
                nbar += analyse_channel<WIDE, H>(a, ws, buf, 0, 0, ws.fd.hdr_bits, phase, bs, lane);
                publish_size(ws.fd.hdr_bits + ws.dec[0].bits);
            }
            if (pend) finish_pending(false);
            // ---- sweep 2: pack ----
            emit_frame_header();
            for (int c = 0; c < n_ch; c++) {
                const SubDec &d = ws.dec[c == 0 ? sub0 : sub1];
                materialise<WIDE, H>(a, ws, buf, d, phase, bs, lane);
                pack_subframe(d);
                if (c + 1 < n_ch) flush_words();
            }
        } else {
            // ---- streamed: analyse, pack and send each channel on its way ----
            for (int c = 0; c < n_ch; c++) {
                nbar += analyse_channel<WIDE, H>(a, ws, buf, c, 0, c == 0 ? ws.fd.hdr_bits : (em.P & 31u), phase, bs, lane);
                const SubDec &d = ws.dec[0];
                if (c == 0) emit_frame_header();
                if (c + 1 == n_ch) publish_size(done * 8u + em.P + d.bits);
                if (pend) finish_pending(false);
                materialise<WIDE, H>(a, ws, buf, d, phase, bs, lane);      // (a no-op unless the subframe fell back to VERBATIM)
                pack_subframe(d);
                if (c + 1 < n_ch) flush_words();
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
        if (done) crc ^= frx_gf16_mulmod(crc_run, gf16_xpow32(a, nw));
        for (uint32_t b = 0; b < tb; b++) {
            const uint32_t byte = (em.carry >> (24 - 8 * b)) & 0xFFu;
            crc = ((crc << 8) ^ crc_tab[0][((crc >> 8) ^ byte) & 0xFFu]) & 0xFFFFu;
        }
        // ---- out: to its final place when the offset is there, else parked ----
        if (pend) finish_pending(true);              // (the frame before this one: it has had this whole frame's time)
        LookBack lb;
        lb.pos = (long long)f - 1; lb.exclusive = 0; lb.blocked = false;
        const bool have = lb_run(lb, false);
        bool skip = false;
        uint8_t *dst = slot + done;
        if (have) {
            skip = lb_commit(f, nbytes, lb.exclusive);
            if (!skip && done) copy_slot_out(a.bytes + lb.exclusive, slot, done, lane);
            dst = a.bytes + lb.exclusive + done;
        }
        if (!skip) {
            copy_words_out(dst, buf, nw, lane);
            if (lane < (int)tb) dst[nw * 4u + lane] = (uint8_t)((em.carry >> (24 - 8 * lane)) & 0xFFu);
            if (lane == 4) dst[nw * 4u + tb] = (uint8_t)(crc >> 8);
            if (lane == 5) dst[nw * 4u + tb + 1u] = (uint8_t)crc;
        }
        if (!have) {
            pend = true; pend_f = f; pend_nbytes = nbytes; pend_slot = slot; plb = lb;
            cur ^= 1;
            DBG_ADD(12, 1);
        }
        __syncwarp();
        if (ALIGNED) { while (nbar < a.bars_per_frame) ALIGN_BAR(nbar); }      // the same number of barriers for every frame
        DBG_ADD(3, DBG_CLK() - tf0); DBG_ADD(4, 1);
    }
    if (pend) finish_pending(true);
}

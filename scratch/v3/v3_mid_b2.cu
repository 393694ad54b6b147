// ------------------------------------------------------------------------------------------
// control warp pieces
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ int frame_header_bytes(uint8_t *hb, int nsamp, int sr_code, int sample_rate, int channels, int bps, uint32_t fj) {
    int n = 0, bs_code, bs_extra = 0;
    switch (nsamp) {
        case 192: bs_code = 1; break; case 576: bs_code = 2; break; case 1152: bs_code = 3; break;
        case 2304: bs_code = 4; break; case 4608: bs_code = 5; break; case 256: bs_code = 8; break;
        case 512: bs_code = 9; break; case 1024: bs_code = 10; break; case 2048: bs_code = 11; break;
        case 4096: bs_code = 12; break; case 8192: bs_code = 13; break; case 16384: bs_code = 14; break;
        case 32768: bs_code = 15; break;
        default: if (nsamp <= 256) { bs_code = 6; bs_extra = 1; } else { bs_code = 7; bs_extra = 2; }
    }
    const int sz_code = bps == 16 ? 4 : (bps == 32 ? 7 : (bps == 8 ? 1 : (bps == 24 ? 6 : 0)));
    hb[n++] = 0xFF; hb[n++] = 0xF8;
    hb[n++] = (uint8_t)((bs_code << 4) | sr_code);
    hb[n++] = (uint8_t)(((channels - 1) << 4) | (sz_code << 1));
    const uint32_t v = fj;
    if (v < 0x80u) hb[n++] = (uint8_t)v;
    else if (v < 0x800u) { hb[n++] = (uint8_t)(0xC0 | (v >> 6)); hb[n++] = (uint8_t)(0x80 | (v & 0x3F)); }
    else if (v < 0x10000u) { hb[n++] = (uint8_t)(0xE0 | (v >> 12)); hb[n++] = (uint8_t)(0x80 | ((v >> 6) & 0x3F)); hb[n++] = (uint8_t)(0x80 | (v & 0x3F)); }
    else if (v < 0x200000u) { hb[n++] = (uint8_t)(0xF0 | (v >> 18)); hb[n++] = (uint8_t)(0x80 | ((v >> 12) & 0x3F)); hb[n++] = (uint8_t)(0x80 | ((v >> 6) & 0x3F)); hb[n++] = (uint8_t)(0x80 | (v & 0x3F)); }
    else if (v < 0x4000000u) { hb[n++] = (uint8_t)(0xF8 | (v >> 24)); hb[n++] = (uint8_t)(0x80 | ((v >> 18) & 0x3F)); hb[n++] = (uint8_t)(0x80 | ((v >> 12) & 0x3F)); hb[n++] = (uint8_t)(0x80 | ((v >> 6) & 0x3F)); hb[n++] = (uint8_t)(0x80 | (v & 0x3F)); }
    else { hb[n++] = (uint8_t)(0xFC | (v >> 30)); hb[n++] = (uint8_t)(0x80 | ((v >> 24) & 0x3F)); hb[n++] = (uint8_t)(0x80 | ((v >> 18) & 0x3F)); hb[n++] = (uint8_t)(0x80 | ((v >> 12) & 0x3F)); hb[n++] = (uint8_t)(0x80 | ((v >> 6) & 0x3F)); hb[n++] = (uint8_t)(0x80 | (v & 0x3F)); }
    if (bs_extra == 1) hb[n++] = (uint8_t)(nsamp - 1);
    else if (bs_extra == 2) { hb[n++] = (uint8_t)((nsamp - 1) >> 8); hb[n++] = (uint8_t)(nsamp - 1); }
    if (sr_code == 13) { hb[n++] = (uint8_t)(sample_rate >> 8); hb[n++] = (uint8_t)sample_rate; }
    uint32_t c = 0;
    for (int i = 0; i < n; i++) c = frx_crc8_byte(c, hb[i]);
    hb[n++] = (uint8_t)c;
    for (int i = n; i < 16; i++) hb[i] = 0;
    return n;
}

// control warp: everything the compute warps need to know about frame f.  The lanes probe 32 tile boundaries at once
// and fetch the tile and its quantiser words side by side, so the whole descriptor costs two memory round trips.
__device__ __noinline__ void build_frame_desc(const EncArgs &a, FrameDesc &d, uint32_t f, int &ti, int lane) {
    {
        const int cand = ti + 1 + lane;
        const uint32_t pv = cand <= a.n_tiles ? __ldg(&a.frame_prefix[cand]) : 0xFFFFFFFFu;
        const uint32_t m = __ballot_sync(FULLMASK, f < pv);
        if (m) ti += __ffs((int)m) - 1;
        else {   // long jump: bisection (frame_prefix[n_tiles] = total frames > f)
            int lo = ti + 32, hi = a.n_tiles;
            while (hi - lo > 1) {
                const int mid = (lo + hi) >> 1;
                if (__ldg(&a.frame_prefix[mid]) <= f) lo = mid; else hi = mid;
            }
            ti = lo;
        }
    }
    constexpr int TW = (int)(sizeof(frx_tile) / 4), QW = (int)(sizeof(FrxQuant) / 4);
    static_assert(TW + QW < 31, "descriptor words must fit one warp");
    uint32_t wv = 0;
    if (lane < TW) wv = __ldg(reinterpret_cast<const uint32_t *>(a.tiles + ti) + lane);
    else if (lane < TW + QW) wv = __ldg(reinterpret_cast<const uint32_t *>(a.quant + ti) + (lane - TW));
    else if (lane == 31) wv = __ldg(&a.frame_prefix[ti]);
    if (lane < TW) reinterpret_cast<uint32_t *>(&d.t)[lane] = wv;
    else if (lane < TW + QW) reinterpret_cast<uint32_t *>(&d.q)[lane - TW] = wv;
    const uint32_t fj = f - __shfl_sync(FULLMASK, wv, 31);
    __syncwarp();
    if (lane == 0) {
        const frx_tile t = d.t;
        const long long tile_n = (long long)t.h * t.w;
        const uint32_t s0 = fj * (uint32_t)a.blocksize;
        const int nsamp = (int)min((long long)a.blocksize, tile_n - (long long)s0);
        d.f = f; d.fj = fj; d.s0 = s0; d.nsamp = nsamp; d.tile_idx = ti;
        {
            const int tz = __ffs(nsamp) - 1;               // trailing zeros of the block size
            int po_max = min(a.max_po, max(0, tz - 4));    // partitions stay multiples of 16 samples
            while (po_max > 0 && (nsamp >> po_max) <= FRX_MAX_LPC) po_max--;
            d.po_max = (int16_t)po_max;
            d.max_lpc = (int16_t)min(a.max_lpc_order, nsamp - 1);
        }
        const int epv = 16 / frx_dtype_size(a.dtype);            // elements per 16 bytes
        d.tma = (a.tma_ok && (t.w % epv) == 0 && (t.col_off % epv) == 0) ? 1 : 0;
        uint8_t hb[16];
        const int n = frame_header_bytes(hb, nsamp, a.sr_code, a.sample_rate, t.n_channels, a.bps, fj);
#pragma unroll
        for (int k = 0; k < 4; k++)
            d.hdr[k] = ((uint32_t)hb[4 * k] << 24) | ((uint32_t)hb[4 * k + 1] << 16) | ((uint32_t)hb[4 * k + 2] << 8) | hb[4 * k + 3];
        d.hdr_bits = (uint32_t)n * 8u;
    }
    __syncwarp();
}

// K8: decoupled look-back over the frame sizes (whole control warp).  A descriptor word carries flag and value
// together (A: the frame's size, P: the inclusive byte prefix), so relaxed device-scope accesses are enough.  The compute
// warps publish A themselves as soon as the size is final; here 512 descriptors are examined per round trip (16
// consecutive ones per lane, nearest first).  When the walk meets a frame that has not published yet, only that single
// word is polled -- re-reading the window would hammer the L2 slice the publishers store to -- and the walk is resumable,
// so the control warp keeps feeding the raw-sample pipeline while it waits.
#define LB_PER_LANE 16
struct LookBack {
    long long pos;        // nearest descriptor not yet accounted for
    uint64_t exclusive;   // bytes summed so far
    bool blocked;         // desc[pos] had not been published at the last visit
};
// one step; true when the exclusive prefix is complete
__device__ __noinline__ bool lookback_step(const EncArgs &a, LookBack &lb, int lane) {
    const uint64_t FLAG_P = 2ull << 62, VMASK = (1ull << 62) - 1;
    if (lb.pos < 0) return true;
    if (lb.blocked) {
        uint64_t w = 0;
        if (lane == 0) w = frx_ld_relaxed_u64(&a.desc[lb.pos]);
        w = __shfl_sync(FULLMASK, w, 0);
        if ((w >> 62) == 0ull) return false;
        lb.blocked = false;
    }
    if (lane == 0) DBG_ADD(3, 1);
    uint64_t v[LB_PER_LANE];
    const long long base = lb.pos - (long long)lane * LB_PER_LANE;
#pragma unroll
    for (int k = 0; k < LB_PER_LANE; k++) {
        const long long idx = base - k;
        v[k] = idx >= 0 ? frx_ld_relaxed_u64(&a.desc[idx]) : FLAG_P;
    }
    // per lane: walk its 16 entries nearest-first up to the first inclusive prefix (P) or the first one not ready
    uint64_t sum = 0;
    int state = 0;                     // 0: all 16 ready, no P   1: stopped at a P   2: stopped at a not-ready entry
    int taken = 0;
#pragma unroll
    for (int k = 0; k < LB_PER_LANE; k++) {
        if (state == 0) {
            const uint32_t flag = (uint32_t)(v[k] >> 62);
            if (flag == 0u) state = 2;
            else { sum += v[k] & VMASK; taken++; if (flag == 2u) state = 1; }
        }
    }
    const uint32_t mP = __ballot_sync(FULLMASK, state == 1), mN = __ballot_sync(FULLMASK, state == 2);
    const int firstP = mP ? (__ffs((int)mP) - 1) : 32, firstN = mN ? (__ffs((int)mN) - 1) : 32;
    // lanes before the first stopper contribute everything; the stopper contributes what it took
    const int stop = min(firstP, firstN);
    lb.exclusive += warp_sum_u64((lane <= stop) ? sum : 0ull);
    if (firstP < firstN) return true;  // reached an inclusive prefix
    if (firstN < 32) {
        const int tk = __shfl_sync(FULLMASK, taken, firstN);
        lb.pos -= (long long)firstN * LB_PER_LANE + tk;      // the frame that has not published its size yet
        lb.blocked = true;
        if (lane == 0) DBG_ADD(4, 1);
    } else lb.pos -= 32 * LB_PER_LANE;
    return false;
}

// The control warp's event loop.  A CTA works on one frame and has at most the next one staged; the next ticket is
// taken only when the current frame's size has been published, because a ticket that is reserved but not yet sized
// holds up the look-back of every later frame (and convoys of such tickets serialise the whole grid).
__device__ __noinline__ void control_warp(const EncArgs &a, EncSmall &sh, unsigned char *raw, int lane) {
    const int esz = frx_dtype_size(a.dtype);
    uint32_t unit_iss = 0, fseq_iss = 0, fseq_lb = 0;
    int ch_iss = 0, nch_cur = 0, ti = 0;
    bool have_frame = false, want_frame = true, stop_issued = false, lb_active = false;
    LookBack lb;
    lb.pos = -1; lb.exclusive = 0; lb.blocked = false;
    for (;;) {
        bool progress = false;
        // ---- 1. a frame's size became known: get the next frame going, then start this frame's look-back ----
        if (!lb_active && fseq_lb < fseq_iss) {
            const int slot = (int)(fseq_lb & (FD_SLOTS - 1));
            if (mbar_test(&sh.sized[slot], (fseq_lb / FD_SLOTS) & 1u)) {
                if (!stop_issued) want_frame = true;
                lb.pos = (long long)sh.fd[slot].f - 1;
                lb.exclusive = 0; lb.blocked = false;
                lb_active = true;
                if (lane == 0) DBG_TRACE(sh.fd[slot].f, 2);
                progress = true;
            }
        }
        // ---- 2. next unit (frame, channel) into the raw stage ring ----
        if (have_frame || want_frame) {
            const int stage = 0;
            if (mbar_test(&sh.empty[stage], (unit_iss & 1u) ^ 1u)) {
                progress = true;
                if (!have_frame) {
                    const int slot = (int)(fseq_iss & (FD_SLOTS - 1));
                    uint32_t ticket = 0;
                    if (lane == 0) ticket = atomicAdd(a.ticket, 1u);
                    ticket = __shfl_sync(FULLMASK, ticket, 0);
                    want_frame = false;
                    if (ticket >= a.total_frames) {
                        if (lane == 0) { sh.fd[slot].f = STOP_FRAME; mbar_arrive(&sh.full[stage]); }
                        __syncwarp();
                        stop_issued = true;
                    } else {
                        build_frame_desc(a, sh.fd[slot], ticket, ti, lane);
                        if (lane == 0) { DBG_TRACE(ticket, 0); DBG_TRACEV(ticket, 6, blockIdx.x); }
                        nch_cur = sh.fd[slot].t.n_channels;
                        have_frame = true;
                        ch_iss = 0;
                    }
                }
                if (have_frame) {
                    const FrameDesc &d = sh.fd[fseq_iss & (FD_SLOTS - 1)];
                    unsigned char *dst = raw;
                    if (d.tma) {
                        const long long w = d.t.w;
                        const long long s0 = d.s0, s1 = s0 + d.nsamp;
                        const long long row0 = s0 / w, row1 = (s1 - 1) / w;
                        const unsigned char *band = (const unsigned char *)a.raster +
                            ((long long)(d.t.first_band + ch_iss) * a.band_stride + (long long)d.t.row_off * a.row_stride + d.t.col_off) * esz;
                        if (lane == 0) mbar_arrive_expect_tx(&sh.full[stage], (uint32_t)d.nsamp * (uint32_t)esz);
                        __syncwarp();
                        for (long long r = row0 + lane; r <= row1; r += 32) {
                            const long long b0 = max(s0, r * w), b1 = min(s1, (r + 1) * w);
                            bulk_g2s(dst + (b0 - s0) * esz, band + (r * a.row_stride + (b0 - r * w)) * esz, (uint32_t)((b1 - b0) * esz), &sh.full[stage]);
                        }
                    } else if (lane == 0) {
                        mbar_arrive(&sh.full[stage]);
                    }
                    __syncwarp();
                    unit_iss++;
                    if (++ch_iss == nch_cur) { have_frame = false; fseq_iss++; }
                }
            }
        }
        // ---- 3. the look-back in flight ----
        if (lb_active) {
            if (lookback_step(a, lb, lane)) {
                const int slot = (int)(fseq_lb & (FD_SLOTS - 1));
                if (lane == 0) {
                    const uint32_t f = sh.fd[slot].f;
                    const uint32_t nbytes = sh.frame_bytes[slot];
                    const uint64_t off = lb.exclusive;
                    frx_st_relaxed_u64(&a.desc[f], (2ull << 62) | (off + (uint64_t)nbytes));
                    sh.out_offset[slot] = off;
                    if (a.frame_size) a.frame_size[f] = nbytes;
                    if (a.frame_offset) a.frame_offset[f] = off;
                    if (f == a.total_frames - 1 && a.total_bytes) *a.total_bytes = off + nbytes;
                    DBG_TRACE(f, 3);
                    DBG_ADD(6, 1);
                    mbar_arrive(&sh.placed[slot]);
                }
                __syncwarp();
                lb_active = false;
                fseq_lb++;
                progress = true;
            } else if (!lb.blocked) progress = true;
        }
        if (stop_issued && !lb_active && fseq_lb == fseq_iss) break;
        if (!progress) __nanosleep(200);
    }
}

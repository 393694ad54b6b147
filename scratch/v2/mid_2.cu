
// ------------------------------------------------------------------------------------------
// frame header, frame descriptor
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ int frame_header_bytes(uint8_t *hb, int nsamp, int sr_code, int sample_rate, int chan_code, int bps, uint32_t fj) {
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
    hb[n++] = (uint8_t)((chan_code << 4) | (sz_code << 1));
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
// header words of the frame for a channel-assignment code (0..7 independent, 8 left/side, 9 side/right, 10 mid/side)
__device__ __noinline__ void make_frame_header(const EncArgs &a, FrameDesc &d, int chan_code) {
    uint8_t hb[16];
    const int n = frame_header_bytes(hb, d.nsamp, a.sr_code, a.sample_rate, chan_code, a.bps, d.fj);
#pragma unroll
    for (int k = 0; k < 4; k++)
        d.hdr[k] = ((uint32_t)hb[4 * k] << 24) | ((uint32_t)hb[4 * k + 1] << 16) | ((uint32_t)hb[4 * k + 2] << 8) | hb[4 * k + 3];
    d.hdr_bits = (uint32_t)n * 8u;
}

// Everything the warp needs to know about frame f.  The lanes probe 32 tile boundaries at once and fetch the tile and its
// quantiser words side by side, so the whole descriptor costs two memory round trips.
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
        const long long s0 = (long long)fj * (long long)a.blocksize;
        const int nsamp = (int)min((long long)a.blocksize, tile_n - s0);
        d.f = f; d.fj = fj; d.s0 = s0; d.nsamp = nsamp; d.tile_idx = ti;
        {
            const int tz = __ffs(nsamp) - 1;               // trailing zeros of the block size
            int po_max = min(a.max_po, max(0, tz - 4));    // partitions stay multiples of 16 samples
            while (po_max > 0 && (nsamp >> po_max) <= FRX_MAX_LPC) po_max--;
            if (nsamp != FRX_MAX_BLOCK && po_max > 5) po_max = 5;   // short frames: the partition bins hold 32 sums
            d.po_max = (int16_t)po_max;
            d.max_lpc = (int16_t)min(a.max_lpc_order, nsamp - 1);
        }
        const int epv = 16 / frx_dtype_size(a.dtype);            // elements per 16 bytes
        d.tma = (a.tma_ok && (t.w % epv) == 0 && (t.col_off % epv) == 0 && (a.blocksize % epv) == 0) ? 1 : 0;
        make_frame_header(a, d, t.n_channels - 1);
    }
    __syncwarp();
}

// K8: decoupled look-back over the frame sizes (whole warp).  A descriptor word carries flag and value together (A: the
// frame's size, P: the inclusive byte prefix), so relaxed device-scope accesses are enough.  512 descriptors are examined
// per round trip, as 16 rows of 32 consecutive words (one coalesced 256-byte load per row, all 16 in flight together); the
// rows are then walked nearest-first with two ballots each.  When the walk meets a frame that has not published yet, only
// that single word is polled.
#define LB_ROWS 16
struct LookBack {
    long long pos;        // nearest descriptor not yet accounted for
    uint64_t exclusive;   // bytes summed so far
    bool blocked;         // desc[pos] had not been published at the last visit
};
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
    uint64_t v[LB_ROWS];
    const long long base = lb.pos - lane;
#pragma unroll
    for (int k = 0; k < LB_ROWS; k++) {
        const long long idx = base - 32 * k;
        v[k] = idx >= 0 ? frx_ld_relaxed_u64(&a.desc[idx]) : FLAG_P;      // in front of frame 0: an inclusive prefix of 0
    }
    uint64_t sum = 0;
    int result = 0;                    // 0: all 512 were sizes   1: reached an inclusive prefix   2: met an unpublished frame
    int taken = 0;
#pragma unroll
    for (int k = 0; k < LB_ROWS; k++) {
        if (result == 0) {
            const uint32_t flag = (uint32_t)(v[k] >> 62);
            const uint32_t mP = __ballot_sync(FULLMASK, flag == 2u), mN = __ballot_sync(FULLMASK, flag == 0u);
            const int firstP = mP ? (__ffs((int)mP) - 1) : 32, firstN = mN ? (__ffs((int)mN) - 1) : 32;
            const int stop = min(firstP, firstN);
            if (lane < stop || (lane == stop && firstP < firstN)) sum += v[k] & VMASK;
            if (stop < 32) { result = firstP < firstN ? 1 : 2; taken = 32 * k + stop; }
        }
    }
    lb.exclusive += warp_sum_u64(sum);
    if (result == 1) return true;
    if (result == 2) { lb.pos -= taken; lb.blocked = true; }   // the frame that has not published its size yet
    else lb.pos -= 32 * LB_ROWS;
    return false;
}

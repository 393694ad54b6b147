
// ==========================================================================================
// per-subframe search: CONSTANT / VERBATIM / best FIXED order / LPC, each Rice-coded candidate costed from its partition
// sums, the winner's residuals left in the buffer, its exact size counted
// ==========================================================================================
#define BUF_PCM 0
#define BUF_RES 1
#define BUF_JUNK 2
struct BufState { int kind, vch, type, order; };

__device__ __forceinline__ void shift_wasted(uint32_t *pcm, int nsamp, int wasted, int lane) {
    const int gend = min(FRX_MAX_BLOCK, ((nsamp + BLK - 1) / BLK) * BLK);
    for (int g = lane; g < gend; g += 32) pcm[PIDX(g)] = (uint32_t)((int32_t)pcm[PIDX(g)] >> wasted);
    __syncwarp();
}
__device__ __forceinline__ void save_warmup(WarpSh &ws, const uint32_t *pcm, int order, int lane) {
    if (lane < order) ws.warm[lane] = (int32_t)pcm[PIDX(lane)];
    __syncwarp();
}

// residuals of the recorded decision into the buffer (the buffer holds the PCM of d.vch, wasted bits already shifted out)
template <bool WIDE, int H>
__device__ __forceinline__ bool residuals_inplace(WarpSh &ws, uint32_t *buf, const SubDec &d, int nsamp, unsigned long long *bins, int psf, int lane,
                                                  uint64_t &sumA, uint64_t &sumB) {
    uint32_t *pcm = buf + SLACK_W;
    const int type = d.type, order = d.order, shift = d.shift, narrow = d.narrow;
    const bool fast = d.fast != 0;
    save_warmup(ws, pcm, order, lane);
    bool ok = true;
    sumA = 0; sumB = 0;
    if (fast) {
        if (type == 3) lpc_inplace_full_any<H, WIDE>(ws, pcm, shift, narrow, lane, sumA, sumB);
        else fixed_inplace_full(pcm, order, lane);
        // warm-up samples are not coded: take their |r| out of the first half's sum and zero them
        uint32_t wv = 0;
        if (lane < order) { wv = unzig_abs(pcm[lane]); pcm[lane] = 0u; }
        const uint64_t wsum = warp_sum_u64(wv);
        if (lane == 0 && type == 3) sumA -= wsum;
        __syncwarp();
    } else {
        ok = resid_inplace_generic(ws, pcm, bins, nsamp, type == 3, order, shift, psf, lane);
    }
    return ok;
}

template <bool WIDE, int H>
__device__ __noinline__ int analyse_channel(const EncArgs &a, WarpSh &ws, uint32_t *buf, int vch, int ci, uint32_t startbits, uint32_t &phase,
                                             BufState &bs, int lane) {
    SubDec &d = ws.dec[ci];
    const FrameDesc &fd = ws.fd;
    uint32_t *pcm = buf + SLACK_W;
    const int nsamp = fd.nsamp;
    const bool full = nsamp == FRX_MAX_BLOCK;
    const long long t_0 = DBG_CLK();
    constexpr bool ALIGNED = !WIDE;
    int nbar = 0;                     // group barriers passed (phase alignment of the frame warps, see the kernel)
    const uint2 st = load_channel(a, ws, buf, vch, phase, lane);
    const uint32_t orv = st.x, ora = st.y;
    DBG_ADD(5, DBG_CLK() - t_0); DBG_ADD(15, 1);
    const int wasted = orv ? (__ffs((int)orv) - 1) : 0;
    if (wasted) shift_wasted(pcm, nsamp, wasted, lane);
    const int magbits = 32 - __clz((int)(ora >> wasted));          // |sample| <= 2^magbits
    const int sbps = a.bps + (vch == 9 ? 1 : 0) - wasted;
    const bool fast = full && magbits <= 23;
    const uint32_t kmax = a.bps > 16 ? 30u : 14u;                    // RICE2 only for bps > 16 (libFLAC)
    const int po_max = fd.po_max;
    const int max_lpc = min((int)fd.max_lpc, H);
    const uint64_t verb_bits = 8ull + (uint64_t)wasted + (uint64_t)nsamp * (uint64_t)sbps;
    bs.kind = BUF_PCM; bs.vch = vch; bs.type = 1; bs.order = 0;
    if (lane == 0) {
        d.type = 1; d.order = 0; d.wasted = (uint8_t)wasted; d.po = 0; d.rice2 = 0; d.precision = 0; d.sbps = (uint8_t)sbps; d.vch = (uint8_t)vch;
        d.fast = fast ? 1 : 0; d.narrow = 1; d.runs = 0; d.shift = 0; d.bits = (uint32_t)verb_bits;
    }
    __syncwarp();

    ALIGN_BAR(nbar);
    // ---- K2 + K3 ----
    const long long t_1 = DBG_CLK();
    uint64_t tot[5];
    uint64_t fsl[5][2];
    uint32_t badmask = 0;
    float autoc[H + 1];
    if (fast) {
        float acl[H + 1];
        pass1_full<H>(pcm, lane, fsl, acl);
        if (lane == 0) {   // warm-up rule: sample g counts for order k only when g >= k
            int32_t prev = 0, d1p = 0, d2p = 0, d3p = 0;
#pragma unroll
            for (int g = 0; g < 4; g++) {
                const int32_t x = (int32_t)pcm[g];
                const int32_t d1 = x - prev, d2 = d1 - d1p, d3 = d2 - d2p, d4 = d3 - d3p;
                prev = x; d1p = d1; d2p = d2; d3p = d3;
                if (g < 1) fsl[1][0] -= (uint32_t)abs(d1);
                if (g < 2) fsl[2][0] -= (uint32_t)abs(d2);
                if (g < 3) fsl[3][0] -= (uint32_t)abs(d3);
                fsl[4][0] -= (uint32_t)abs(d4);
            }
        }
#pragma unroll
        for (int o = 0; o < 5; o++) tot[o] = warp_sum_u64(fsl[o][0] + fsl[o][1]);
#pragma unroll
        for (int k = 0; k <= H; k++) autoc[k] = warp_sum_f32(acl[k]);
    } else {
#pragma unroll
        for (int o = 0; o < 5; o++) { fsl[o][0] = 0; fsl[o][1] = 0; }
        pass1_generic<H>(ws, pcm, nsamp, max_lpc, lane);
#pragma unroll
        for (int o = 0; o < 5; o++) tot[o] = ws.g_tot[o];
#pragma unroll
        for (int k = 0; k <= H; k++) autoc[k] = ws.g_ac[k];
        badmask = ws.g_bad;
    }
    DBG_ADD(6, DBG_CLK() - t_1);
    if (tot[1] == 0 && !(badmask & 2u)) {     // sum |s[g] - s[g-1]| == 0  <=>  all samples equal
        if (lane == 0) { d.type = 0; d.bits = 8u + (uint32_t)wasted + (uint32_t)sbps; }
        __syncwarp();
        return nbar;
    }
    ALIGN_BAR(nbar);
    int fixed_order = -1;
    if (nsamp > 4) {
        uint64_t bestv = ~0ull;
#pragma unroll
        for (int o = 0; o < 5; o++)
            if (!((badmask >> o) & 1u) && tot[o] < bestv) { bestv = tot[o]; fixed_order = o; }
    }
    // ---- K4 ----
    LpcChoice lc;
    lc.ok = 0; lc.order = 0; lc.shift = 0; lc.precision = 0; lc.narrow = 1;
    const long long t_2 = DBG_CLK();
    if (max_lpc > 0 && nsamp > 4) lc = lpc_design<WIDE, H>(ws, autoc, max_lpc, nsamp, sbps, a.qlp_precision, magbits, lane);
    DBG_ADD(7, DBG_CLK() - t_2);
    const long long t_3 = DBG_CLK();

    unsigned long long *bins = reinterpret_cast<unsigned long long *>(buf);   // generic path: 32 sums in the slack words
    const int Fg = po_max > 5 ? 5 : po_max;
    const int psf = nsamp >> Fg;
    uint64_t best_bits = verb_bits;
    int type = 1, order = 0, po = 0, cand = 0;
    bool rice2 = false;
    // ---- fixed candidate: cost from its partition sums ----
    if (fixed_order >= 0) {
        RiceChoice rf;
        if (fast) {
            uint64_t sA = 0, sB = 0;
#pragma unroll
            for (int o = 0; o < 5; o++) if (o == fixed_order) { sA = fsl[o][0]; sB = fsl[o][1]; }
            rf = analyse_partitions(ws.kparam[0], sA, sB, nsamp, fixed_order, po_max, 5, kmax);
        } else {
            bins[lane] = 0ull;
            __syncwarp();
            fixed_sums_generic(pcm, bins, nsamp, fixed_order, psf, lane);
            const uint64_t sA = lane < (1 << Fg) ? bins[lane] : 0ull;
            __syncwarp();
            rf = analyse_partitions(ws.kparam[0], sA, 0ull, nsamp, fixed_order, Fg, Fg, kmax);
        }
        const uint64_t b = 8ull + wasted + (uint64_t)fixed_order * sbps + rf.bits;
        if (b < best_bits) { best_bits = b; type = 2; order = fixed_order; po = rf.po; rice2 = rf.rice2; cand = 0; }
    }
    DBG_ADD(8, DBG_CLK() - t_3);
    const long long t_4 = DBG_CLK();
    ALIGN_BAR(nbar);
    // ---- LPC candidate: residuals in place (the usual winner), cost from their partition sums ----
    if (lc.ok) {
        if (lane == 0) { d.type = 3; d.order = (uint8_t)lc.order; d.shift = lc.shift; d.narrow = (uint8_t)lc.narrow; d.precision = (uint8_t)lc.precision; }
        __syncwarp();
        uint64_t sA = 0, sB = 0;
        if (!fast) { bins[lane] = 0ull; __syncwarp(); }
        const bool ok = residuals_inplace<WIDE, H>(ws, buf, d, nsamp, bins, psf, lane, sA, sB);
        bs.kind = ok ? BUF_RES : BUF_JUNK; bs.type = 3; bs.order = lc.order;
        if (ok) {
            RiceChoice rl;
            if (fast) rl = analyse_partitions(ws.kparam[1], sA, sB, nsamp, lc.order, po_max, 5, kmax);
            else {
                sA = lane < (1 << Fg) ? bins[lane] : 0ull;
                __syncwarp();
                rl = analyse_partitions(ws.kparam[1], sA, 0ull, nsamp, lc.order, Fg, Fg, kmax);
            }
            const uint64_t b = 8ull + wasted + 9ull + (uint64_t)lc.order * (uint64_t)(lc.precision + sbps) + rl.bits;
            if (b < best_bits) { best_bits = b; type = 3; order = lc.order; po = rl.po; rice2 = rl.rice2; cand = 1; }
        }
    }
    DBG_ADD(9, DBG_CLK() - t_4);
    ALIGN_BAR(nbar);
    if (type == 1) {
        if (lane == 0) { d.type = 1; d.order = 0; }
        __syncwarp();
        return nbar;
    }
    const long long t_5 = DBG_CLK();
    // ---- record the winner; make the buffer hold its residuals ----
    if (lane == 0) {
        d.type = (uint8_t)type; d.order = (uint8_t)order; d.po = (uint8_t)po; d.rice2 = rice2 ? 1 : 0;
        if (type == 2) { d.precision = 0; d.shift = 0; }
    }
    if (lane < FRX_MAX_LPC) d.qlp[lane] = type == 3 ? (int16_t)ws.qlp[lane] : (int16_t)0;
    d.k[lane] = ws.kparam[cand][lane]; d.k[lane + 32] = ws.kparam[cand][lane + 32];
    __syncwarp();
    if (type == 2) {
        if (bs.kind != BUF_PCM) {
            (void)load_channel(a, ws, buf, vch, phase, lane);
            if (wasted) shift_wasted(pcm, nsamp, wasted, lane);
        }
        uint64_t x0, x1;
        (void)residuals_inplace<WIDE, H>(ws, buf, d, nsamp, nullptr, psf, lane, x0, x1);
        bs.kind = BUF_RES; bs.type = 2; bs.order = order;
    }
    DBG_ADD(10, DBG_CLK() - t_5);
    const long long t_6 = DBG_CLK();
    // ---- exact size ----
    const uint32_t pb = rice2 ? 5u : 4u;
    const uint32_t hb = 8u + (uint32_t)wasted + (uint32_t)order * (uint32_t)sbps + (type == 3 ? 9u + (uint32_t)order * (uint32_t)lc.precision : 0u) + 6u;
    uint32_t total;
    bool inplace_ok, runs_ok = false;
    if (fast) {
        const int psh = 12 - po;                                   // log2(samples per partition)
        const int g0 = lane * BLK;
        const uint32_t k0 = d.k[g0 >> psh], k1 = d.k[(g0 + BLK / 2) >> psh];
        bool lane_ok;
        uint32_t c = count_block(pcm, lane, k0, k1, (g0 & ((1 << psh) - 1)) == 0 ? pb : 0u, po == 6 ? pb : 0u, lane_ok);
        if (lane == 0) c -= (uint32_t)order * (k0 + 1u);           // warm-up samples carry no code (their slots hold 0)
        if (lane == 0) c += hb;
        uint32_t incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t n = __shfl_up_sync(FULLMASK, incl, o);
            if (lane >= o) incl += n;
        }
        total = __shfl_sync(FULLMASK, incl, 31);
        inplace_ok = __all_sync(FULLMASK, startbits + incl <= 32u * (SLACK_W + BLK_W * ((uint32_t)lane + 1u) - 4u));
        runs_ok = __all_sync(FULLMASK, lane_ok);
    } else {
        total = hb + count_generic(pcm, d.k, nsamp, order, nsamp >> po, pb, startbits + hb, inplace_ok, lane);
    }
    if ((uint64_t)total >= verb_bits || !inplace_ok) {
        // the Rice-coded subframe would not beat VERBATIM (or cannot be packed in place): write it verbatim
        if (lane == 0) { d.type = 1; d.order = 0; d.bits = (uint32_t)verb_bits; }
    } else if (lane == 0) { d.bits = total; d.runs = runs_ok ? 1 : 0; }
    __syncwarp();
    DBG_ADD(11, DBG_CLK() - t_6);
    return nbar;
}

// make the buffer hold what packing subframe d needs: its PCM (CONSTANT / VERBATIM) or its residuals
template <bool WIDE, int H>
__device__ __noinline__ void materialise(const EncArgs &a, WarpSh &ws, uint32_t *buf, const SubDec &d, uint32_t &phase, BufState &bs, int lane) {
    const bool want_res = d.type >= 2;
    if (bs.vch == d.vch) {
        if (want_res && bs.kind == BUF_RES && bs.type == d.type && bs.order == d.order) return;
        if (!want_res && bs.kind == BUF_PCM) return;
    }
    uint32_t *pcm = buf + SLACK_W;
    (void)load_channel(a, ws, buf, d.vch, phase, lane);
    if (d.wasted) shift_wasted(pcm, ws.fd.nsamp, d.wasted, lane);
    bs.kind = BUF_PCM; bs.vch = d.vch;
    if (want_res) {
        if (lane < FRX_MAX_LPC) ws.qlp[lane] = d.qlp[lane];
        __syncwarp();
        uint64_t x0, x1;
        (void)residuals_inplace<WIDE, H>(ws, buf, d, ws.fd.nsamp, nullptr, 1, lane, x0, x1);
        bs.kind = BUF_RES; bs.type = d.type; bs.order = d.order;
    }
}

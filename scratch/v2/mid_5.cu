
// ---- K5: LPC residuals of the lane's block with all H taps (coefficients beyond the order are zero), zig-zag folded and
//      stored in place of the samples; |r| sums per half.  MODE 0: 32-bit accumulate, 1: 64-bit accumulate ----
template <int H, int MODE, bool WIDE>
struct P2 {
    static constexpr int W = H + 4;
    int32_t s[W];
    int32_t q[H];
    int shift;
    uint32_t acc32;
    uint64_t acc64;
    uint64_t cur, first;
    template <int P>
    __device__ __forceinline__ void step(uint32_t *p) {
        const uint4 v = *reinterpret_cast<const uint4 *>(p);
        s[(4 * P + H) % W] = (int32_t)v.x; s[(4 * P + H + 1) % W] = (int32_t)v.y; s[(4 * P + H + 2) % W] = (int32_t)v.z; s[(4 * P + H + 3) % W] = (int32_t)v.w;
        uint32_t u[4];
#pragma unroll
        for (int i = 0; i < 4; i++) {
            int32_t r;
            if (MODE == 0) {
                int32_t a = 0;
#pragma unroll
                for (int j = 0; j < H; j++) a += q[j] * s[(4 * P + H + i - 1 - j) % W];
                r = s[(4 * P + H + i) % W] - (a >> shift);
            } else {
                long long a = 0;
#pragma unroll
                for (int j = 0; j < H; j++) a += (long long)q[j] * (long long)s[(4 * P + H + i - 1 - j) % W];
                r = (int32_t)((long long)s[(4 * P + H + i) % W] - (a >> shift));
            }
            u[i] = zigzag(r);
            if (WIDE) acc64 += (uint32_t)abs(r); else acc32 = __sad(r, 0, acc32);
        }
        *reinterpret_cast<uint4 *>(p) = make_uint4(u[0], u[1], u[2], u[3]);
    }
    __device__ __forceinline__ void fold() {
        if (WIDE) { cur += acc64; acc64 = 0; } else { cur += acc32; acc32 = 0; }
    }
    __device__ __forceinline__ void end_first_half() { first = cur; cur = 0; }
    template <int R>
    __device__ __forceinline__ void realign() {
        int32_t t[W];
#pragma unroll
        for (int k = 0; k < W; k++) t[k] = s[(4 * R + k) % W];
#pragma unroll
        for (int k = 0; k < W; k++) s[k] = t[k];
    }
};
template <int H, int MODE, bool WIDE>
__device__ __forceinline__ void lpc_inplace_full(WarpSh &ws, uint32_t *pcm, int shift, int lane, uint64_t &sumA, uint64_t &sumB) {
    uint32_t *blk = pcm + lane * BLK_W;
    P2<H, MODE, WIDE> st;
    int32_t hs[H];
    load_history<H>(blk, lane, hs);
    __syncwarp();                      // every lane holds its history before any lane overwrites samples
#pragma unroll
    for (int k = 0; k < H; k++) st.s[k] = hs[k];
#pragma unroll
    for (int j = 0; j < H; j++) st.q[j] = j < FRX_MAX_LPC ? ws.qlp[j] : 0;
    st.shift = shift; st.acc32 = 0; st.acc64 = 0; st.cur = 0; st.first = 0;
    walk_block<H>(st, blk);
    sumA = st.first; sumB = st.cur;
    __syncwarp();
}
template <int H, bool WIDE>
__device__ __noinline__ void lpc_inplace_full_any(WarpSh &ws, uint32_t *pcm, int shift, int narrow, int lane, uint64_t &sumA, uint64_t &sumB) {
    if (!WIDE || narrow) lpc_inplace_full<H, 0, WIDE>(ws, pcm, shift, lane, sumA, sumB);
    else lpc_inplace_full<H, 1, WIDE>(ws, pcm, shift, lane, sumA, sumB);
}

// fixed predictor of order `order` (0..4), residuals zig-zag folded in place (32-bit arithmetic: lean data)
__device__ __noinline__ void fixed_inplace_full(uint32_t *pcm, int order, int lane) {
    uint32_t *blk = pcm + lane * BLK_W;
    int32_t hs[4];
    load_history<4>(blk, lane, hs);
    __syncwarp();
    int32_t prev = hs[3], d1p = hs[3] - hs[2];
    const int32_t dm2 = hs[2] - hs[1], dm3 = hs[1] - hs[0];
    int32_t d2p = d1p - dm2, d3p = d2p - (dm2 - dm3);
#pragma unroll 2
    for (int it = 0; it < BLK / 4; it++) {
        const uint4 v = *reinterpret_cast<const uint4 *>(blk + 4 * it);
        const int32_t x[4] = { (int32_t)v.x, (int32_t)v.y, (int32_t)v.z, (int32_t)v.w };
        uint32_t u[4];
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const int32_t d1 = x[i] - prev;
            const int32_t d2 = d1 - d1p, d3 = d2 - d2p, d4 = d3 - d3p;
            prev = x[i]; d1p = d1; d2p = d2; d3p = d3;
            const int32_t r = order == 0 ? x[i] : (order == 1 ? d1 : (order == 2 ? d2 : (order == 3 ? d3 : d4)));
            u[i] = zigzag(r);
        }
        *reinterpret_cast<uint4 *>(blk + 4 * it) = make_uint4(u[0], u[1], u[2], u[3]);
    }
    __syncwarp();
}

// Exact Rice bits of the lane's block in stream order (k0 / k1 = parameters of its two halves, pb0 / pb1 = parameter bits in
// front of them) and the lane-local condition of pack_runs: a lane that writes its run over its own block, starting four
// words in front of it, never overtakes its reads when the bits through every group of four samples stay within
// 128 t + 224.
__device__ __forceinline__ uint32_t count_block(const uint32_t *pcm, int lane, uint32_t k0, uint32_t k1, uint32_t pb0, uint32_t pb1, bool &ok) {
    const uint32_t *blk = pcm + lane * BLK_W;
    uint32_t c = pb0, lim = 224u, bad = 0;
    const uint32_t s0 = 4u * (k0 + 1u), s1 = 4u * (k1 + 1u);
#pragma unroll 4
    for (int it = 0; it < BLK / 8; it++) {
        const uint4 v = *reinterpret_cast<const uint4 *>(blk + 4 * it);
        c += s0 + ((v.x >> k0) + (v.y >> k0)) + ((v.z >> k0) + (v.w >> k0));
        bad |= (c > lim) ? 1u : 0u;
        lim += 128u;
    }
    c += pb1;
#pragma unroll 4
    for (int it = BLK / 8; it < BLK / 4; it++) {
        const uint4 v = *reinterpret_cast<const uint4 *>(blk + 4 * it);
        c += s1 + ((v.x >> k1) + (v.y >> k1)) + ((v.z >> k1) + (v.w >> k1));
        bad |= (c > lim) ? 1u : 0u;
        lim += 128u;
    }
    ok = bad == 0u;
    return c;
}

// ==========================================================================================
// GENERIC PATH: short frames and fat data (|sample| > 2^23): 64-bit arithmetic with range checks, samples striped over
// the lanes (g = 32 i + lane), in-place passes walk the frame from its end so that history is still intact.
// ==========================================================================================
__device__ __forceinline__ long long pcm_at(const uint32_t *pcm, int g) { return g >= 0 ? (long long)(int32_t)pcm[PIDX(g)] : 0ll; }

__device__ __forceinline__ long long fixed_resid64(const uint32_t *pcm, int g, int order) {
    const long long A = pcm_at(pcm, g);
    if (order == 0) return A;
    const long long B = pcm_at(pcm, g - 1);
    if (order == 1) return A - B;
    const long long C = pcm_at(pcm, g - 2);
    if (order == 2) return A - 2 * B + C;
    const long long D = pcm_at(pcm, g - 3);
    if (order == 3) return A - 3 * B + 3 * C - D;
    return A - 4 * B + 6 * C - 4 * D + pcm_at(pcm, g - 4);
}

// results go through shared memory (ws.g_*): reference parameters of an out-of-line function would pin the caller's arrays to local memory
template <int MAXO>
__device__ __noinline__ void pass1_generic(WarpSh &ws, const uint32_t *pcm, int nsamp, int max_lpc, int lane) {
    uint64_t tot[5];
    float ac[MAXO + 1];
    uint32_t badmask;
#pragma unroll
    for (int o = 0; o < 5; o++) tot[o] = 0;
#pragma unroll
    for (int k = 0; k <= MAXO; k++) ac[k] = 0.f;
    uint32_t bad = 0;
    float xw_prev = 0.f;
    const int niter = (nsamp + 31) >> 5;
    for (int it = 0; it < niter; it++) {
        const int g = 32 * it + lane;
        const bool valid = g < nsamp;
        const long long A = valid ? pcm_at(pcm, g) : 0, B = pcm_at(pcm, g - 1), C = pcm_at(pcm, g - 2), D = pcm_at(pcm, g - 3), E = pcm_at(pcm, g - 4);
        const long long e[5] = { A, A - B, A - 2 * B + C, A - 3 * B + 3 * C - D, A - 4 * B + 6 * C - 4 * D + E };
        if (valid) {
#pragma unroll
            for (int o = 0; o < 5; o++) {
                if (g >= o) {
                    const unsigned long long av = (unsigned long long)llabs(e[o]);
                    tot[o] += av;
                    if (av > 0x7FFFFFFFull) bad |= 1u << o;
                }
            }
        }
        if (max_lpc > 0) {
            const float xw = valid ? (float)(int32_t)A * tukey_w(g, nsamp) : 0.f;
#pragma unroll
            for (int lag = 0; lag <= MAXO; lag++) {
                const float same = __shfl_up_sync(FULLMASK, xw, lag);                        // lane - lag of this row
                const float prev = __shfl_sync(FULLMASK, xw_prev, (32 + lane - lag) & 31);   // tail of the previous row
                const float partner = lag == 0 ? xw : (lane >= lag ? same : prev);
                if (lag <= max_lpc) ac[lag] = fmaf(xw, partner, ac[lag]);
            }
            xw_prev = xw;
        }
    }
#pragma unroll
    for (int o = 0; o < 5; o++) tot[o] = warp_sum_u64(tot[o]);
#pragma unroll
    for (int k = 0; k <= MAXO; k++) ac[k] = warp_sum_f32(ac[k]);
    badmask = __reduce_or_sync(FULLMASK, bad);
    if (lane == 0) {
#pragma unroll
        for (int o = 0; o < 5; o++) ws.g_tot[o] = tot[o];
#pragma unroll
        for (int k = 0; k <= MAXO; k++) ws.g_ac[k] = ac[k];
        ws.g_bad = badmask;
    }
    __syncwarp();
}

// |r| sums of the fixed predictor into the partition bins, nothing stored (the PCM stays intact for the LPC trial)
__device__ __noinline__ void fixed_sums_generic(const uint32_t *pcm, unsigned long long *bins, int nsamp, int order, int psf, int lane) {
    const int niter = (nsamp + 31) >> 5;
    for (int it = 0; it < niter; it++) {
        const int g = 32 * it + lane;
        if (g < nsamp && g >= order) atomicAdd(&bins[g / psf], (unsigned long long)llabs(fixed_resid64(pcm, g, order)));
    }
    __syncwarp();
}

// residuals of the predictor (fixed: order 0..4, lpc == false; LPC: coefficients ws.qlp, shift) folded in place, warm-up
// samples zeroed; |r| sums added to the 64-bit bins of the finest partition level (psf = samples per finest partition).
// Returns false when a residual does not fit (INT32_MIN, INT32_MAX]: the buffer is then garbage and must be reloaded.
__device__ __noinline__ bool resid_inplace_generic(WarpSh &ws, uint32_t *pcm, unsigned long long *bins, int nsamp, bool lpc, int order, int shift,
                                                   int psf, int lane) {
    bool bad = false;
    const int niter = (nsamp + 31) >> 5;
    for (int it = niter - 1; it >= 0; it--) {
        const int g = 32 * it + lane;
        const bool valid = g < nsamp;
        long long r = 0;
        if (valid && g >= order) {
            if (lpc) {
                long long acc = 0;
                for (int j = 0; j < order; j++) acc += (long long)ws.qlp[j] * pcm_at(pcm, g - 1 - j);
                r = pcm_at(pcm, g) - (acc >> shift);
            } else r = fixed_resid64(pcm, g, order);
            if (r <= (long long)INT32_MIN || r > (long long)INT32_MAX) bad = true;
        }
        __syncwarp();
        if (valid) {
            pcm[PIDX(g)] = g >= order ? zigzag((int32_t)r) : 0u;
            if (bins && g >= order) atomicAdd(&bins[g / psf], (unsigned long long)llabs(r));
        }
        __syncwarp();
    }
    return !__any_sync(FULLMASK, bad);
}

// exact Rice bits of the whole subframe body (codes + parameters) and the in-place check: the bits of chunk c (128 samples)
// must end before the first word of chunk c + 1.  kk = the chosen order's parameters, psize = samples per partition.
__device__ __noinline__ uint32_t count_generic(const uint32_t *pcm, const uint8_t *kk, int nsamp, int order, int psize, uint32_t pb, uint32_t headbits,
                                               bool &inplace_ok, int lane) {
    uint32_t cnt = 0, run = headbits;
    bool ok = true;
    const int niter = (nsamp + 31) >> 5;
    for (int it = 0; it < niter; it++) {
        const int g = 32 * it + lane;
        uint32_t c = 0;
        if (g < nsamp) {
            const int pi = g / psize;
            const uint32_t k = kk[pi];
            if (g >= order) c = (pcm[PIDX(g)] >> k) + k + 1u;
            if (g == pi * psize) c += pb;
        }
        cnt += c;
        if ((it & 3) == 3 || it == niter - 1) {
            run += __reduce_add_sync(FULLMASK, cnt);
            cnt = 0;
            const uint32_t chunk = (uint32_t)it >> 2;
            if (run > 32u * (SLACK_W + BLK_W * (chunk + 1u))) ok = false;
        }
    }
    inplace_ok = ok;
    return run - headbits;
}

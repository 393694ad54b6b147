// ------------------------------------------------------------------------------------------
// per-thread pieces of the subframe search
// ------------------------------------------------------------------------------------------
// K2 fixed-order trial, 32-bit: sum |e_k| over the thread's 16 samples for k = 0..4 (history in s[HISTN-4..HISTN-1])
template <int HISTN>
__device__ __forceinline__ void fixed_trial32(const int32_t *s, uint32_t *fa) {
    int32_t dm1 = s[HISTN - 1] - s[HISTN - 2], dm2 = s[HISTN - 2] - s[HISTN - 3], dm3 = s[HISTN - 3] - s[HISTN - 4];
    int32_t d1p = dm1, d2p = dm1 - dm2, d3p = (dm1 - dm2) - (dm2 - dm3);
    uint32_t f0 = 0, f1 = 0, f2 = 0, f3 = 0, f4 = 0;
#pragma unroll
    for (int i = 0; i < FRX_SPT; i++) {
        const int32_t x = s[HISTN + i];
        const int32_t d1 = x - s[HISTN + i - 1];
        const int32_t d2 = d1 - d1p, d3 = d2 - d2p, d4 = d3 - d3p;
        d1p = d1; d2p = d2; d3p = d3;
        f0 += (uint32_t)abs(x); f1 += (uint32_t)abs(d1); f2 += (uint32_t)abs(d2); f3 += (uint32_t)abs(d3); f4 += (uint32_t)abs(d4);
    }
    fa[0] = f0; fa[1] = f1; fa[2] = f2; fa[3] = f3; fa[4] = f4;
}
// masked form for thread 0 (warm-up samples do not count: sample g contributes to order k only when g >= k) and for
// the partial thread of a short frame
template <int HISTN>
__device__ __forceinline__ void fixed_trial32_masked(const int32_t *s, int g0, int nv, uint32_t *fa) {
    int32_t dm1 = s[HISTN - 1] - s[HISTN - 2], dm2 = s[HISTN - 2] - s[HISTN - 3], dm3 = s[HISTN - 3] - s[HISTN - 4];
    int32_t d1p = dm1, d2p = dm1 - dm2, d3p = (dm1 - dm2) - (dm2 - dm3);
#pragma unroll
    for (int o = 0; o < 5; o++) fa[o] = 0;
#pragma unroll
    for (int i = 0; i < FRX_SPT; i++) {
        const int32_t x = s[HISTN + i];
        const int32_t d1 = x - s[HISTN + i - 1];
        const int32_t d2 = d1 - d1p, d3 = d2 - d2p, d4 = d3 - d3p;
        d1p = d1; d2p = d2; d3p = d3;
        if (i < nv) {
            const int g = g0 + i;
            fa[0] += (uint32_t)abs(x);
            if (g >= 1) fa[1] += (uint32_t)abs(d1);
            if (g >= 2) fa[2] += (uint32_t)abs(d2);
            if (g >= 3) fa[3] += (uint32_t)abs(d3);
            if (g >= 4) fa[4] += (uint32_t)abs(d4);
        }
    }
}

// tukey(0.5) window value of sample g in a short frame of n samples (libFLAC's definition).  Out of line on purpose:
// scalars in, scalar out, so the compiler cannot speculate the cosine into the full-frame path (which reads a table)
// and no register array escapes into the call.
__device__ __noinline__ float tukey_w(int g, int n) {
    if (g < 0 || g >= n) return 0.f;
    const int Np = (int)(0.25f * (float)n) - 1;
    if (Np <= 0) return 1.0f;
    if (g <= Np) return 0.5f - 0.5f * cospif((float)g / (float)Np);
    if (g >= n - Np - 1) return 0.5f - 0.5f * cospif((float)(g - (n - Np - 1) + Np) / (float)Np);
    return 1.0f;
}

// K4 on one warp, float32 (MAXO = largest order of the compression level): every lane runs the Levinson-Durbin recursion
// in registers; lane o-1 prices order o (expected bits), lane j owns coefficient j for the row store, the |max| and the
// quantised value, so the serial part is the recursion and the error-feedback rounding only (libFLAC's scheme).
// The decisions are ours; the bitstream stays exact because the residual is computed from the quantised integers.
template <bool WIDE, int MAXO>
__device__ __noinline__ void lpc_design(EncSmall &sh, int max_lpc, int nsamp, int wasted, int sbps, int qlp_precision, int magbits, bool lean,
                                        int lane) {
    if (lane == 0) sh.lpc_ok = 0;
    const float a0 = sh.autoc[0];
    if (!(a0 > 0.f)) return;
    const float inv0 = 1.0f / a0;
    float an[MAXO + 1], lpc[MAXO];
#pragma unroll
    for (int i = 0; i <= MAXO; i++) an[i] = sh.autoc[i] * inv0;   // normalised autocorrelation
    float err = 1.0f, myerr = 0.f;
    int mo = max_lpc;
#pragma unroll
    for (int i = 0; i < MAXO; i++) {
        if (i < mo) {
            float r = -an[i + 1];
#pragma unroll
            for (int j = 0; j < i; j++) r -= lpc[j] * an[i - j];
            r = __fdividef(r, err);
            lpc[i] = r;
#pragma unroll
            for (int j = 0; j < (i >> 1); j++) {
                const float tmp = lpc[j];
                lpc[j] += r * lpc[i - 1 - j];
                lpc[i - 1 - j] += r * tmp;
            }
            if (i & 1) lpc[i >> 1] += lpc[i >> 1] * r;
            err *= (1.0f - r * r);
            if (lane == 0) {                       // (a per-lane select of lpc[j] would turn lpc[] into a local-memory array)
#pragma unroll
                for (int j = 0; j <= i; j++) sh.lpc_rows[i][j] = -lpc[j];
            }
            if (lane == i) myerr = err;
            if (!(err > 0.f)) mo = i + 1;
        }
    }
    __syncwarp();
    // expected bits per order; errs are relative to autoc[0] (taken before the wasted-bit shift)
    const float lscale = __log2f(a0) + __log2f(0.5f / (float)nsamp) - 2.0f * (float)wasted;
    uint32_t key = 0xFFFFFFFFu;
    if (lane < mo) {
        const int o = lane + 1;
        const float bpsr = myerr > 0.f ? fmaxf(0.f, 0.5f * (__log2f(myerr) + lscale)) : 0.f;
        const float b = bpsr * (float)(nsamp - o) + (float)(o * (sbps + qlp_precision));
        key = __float_as_uint(b);                  // b >= 0: the bit pattern orders like the value
        if (!(b >= 0.f)) key = 0xFFFFFFFEu;
    }
    const uint32_t kmin = __reduce_min_sync(FULLMASK, key);
    const int order = __ffs((int)__ballot_sync(FULLMASK, key == kmin));   // lowest order among ties
    int precision = qlp_precision;
    if (!WIDE) {
        const int lim = 32 - sbps - (31 - __clz(order));
        if (lim < precision) precision = lim;
    }
    if (precision < 5) return;
    const float rv = lane < order ? sh.lpc_rows[order - 1][lane] : 0.f;
    const float cmax = __uint_as_float(__reduce_max_sync(FULLMASK, __float_as_uint(fabsf(rv))));
    if (!(cmax > 0.f) || !(cmax < 3.0e38f)) return;
    int e2;
    (void)frexpf(cmax, &e2);
    int shift = (precision - 1) - (e2 - 1) - 1;
    if (shift > 15) shift = 15;
    if (shift < 0) return;
    const int qmax = (1 << (precision - 1)) - 1, qmin = -(1 << (precision - 1));
    float error = 0.f;
    long long sumabs = 0;
    const float scale = (float)(1 << shift);
    int myq = 0;
#pragma unroll
    for (int j = 0; j < MAXO; j++) {
        const float c = __shfl_sync(FULLMASK, rv, j);
        if (j < order) {
            error += c * scale;
            int qq = __float2int_rn(error);
            if (qq > qmax) qq = qmax; else if (qq < qmin) qq = qmin;
            error -= (float)qq;
            if (lane == j) myq = qq;
            sumabs += qq < 0 ? -qq : qq;
        }
    }
    if (lane < FRX_MAX_LPC) sh.qlp[lane] = myq;
    int ok = 1, narrow = 1;
    if (WIDE) {
        // |prediction sum| <= sumabs * 2^mag; the residual bound decides which arithmetic is safe
        const int mag = magbits - wasted;
        const long long M = 1ll << (mag < 0 ? 0 : mag);
        narrow = (sumabs * M < (1ll << 31)) ? 1 : 0;
        if (lean) {
            const long long Bnd = M + ((sumabs * M) >> shift) + 1;
            if (Bnd >= (1ll << 27)) ok = 0;   // keeps 32-bit residual sums exact; such wild predictors never win anyway
        }
    }
    if (lane == 0) { sh.lpc_order = order; sh.shift = shift; sh.precision = precision; sh.lpc_narrow = narrow; sh.lpc_ok = ok; }
    __syncwarp();
}

// OR an n-bit field (1..32 bits, v < 2^n) into the zeroed frame buffer at bit position p
__device__ __forceinline__ void put_field(uint32_t *buf, uint32_t p, uint32_t v, uint32_t n) {
    const uint64_t x = (uint64_t)v << (64u - (p & 31u) - n);
    const uint32_t hi = (uint32_t)(x >> 32), lo = (uint32_t)x;
    if (hi) atomicOr(&buf[p >> 5], hi);
    if (lo) atomicOr(&buf[(p >> 5) + 1u], lo);
}

// K7b: Rice-code the thread's residuals straight into the (zeroed) frame buffer.  The unary zeros need no store at all;
// every sample is one (k+1)-bit field -- stop bit plus low bits -- OR-ed at its bit position, straddling at most two
// words.  The residuals come from the shared residual buffer (sample-major, conflict-free).  `skip` leading samples are warm-up (thread 0 only), `nv` samples are valid, `plain` = neither applies.
__device__ __forceinline__ void pack_rice(uint32_t *buf, uint32_t p, const int32_t *res, int tid, uint32_t kk, int skip, int nv, bool plain) {
    const uint32_t kmask = (1u << kk) - 1u;   // kk <= 30
    const uint32_t len = kk + 1u;
#pragma unroll
    for (int i = 0; i < FRX_SPT; i++) {
        if (plain || (i >= skip && i < nv)) {
            const uint32_t u = zigzag(res[RES_IDX(i, tid)]);
            p += u >> kk;
            const uint32_t code = (1u << kk) | (u & kmask);
            const uint64_t v = (uint64_t)code << (64u - (p & 31u) - len);
            atomicOr(&buf[p >> 5], (uint32_t)(v >> 32));
            if ((uint32_t)v) atomicOr(&buf[(p >> 5) + 1u], (uint32_t)v);
            p += len;
        }
    }
}

// Fat data (|s| > 2^23: lossless int32 / float bit patterns): 64-bit fixed trial and range-checked LPC residuals, the two
// Rice searches, and the decisions -- a complete, slower replacement of the lean K2/K5/K6 stages.  All compute threads call it.
template <int HISTN>
__device__ __noinline__ void fat_search(EncSmall &sh, const int32_t *pcm, uint64_t *tabs, int tid, int nv, int nsamp, int wasted,
                                        int po_max, uint32_t kmax) {
    const int lane = tid & 31, warp = tid >> 5, g0 = tid * FRX_SPT;
    int32_t s[HISTN + FRX_SPT];
    read_pcm<HISTN>(pcm, tid, wasted, s);
    const bool lpc_try = sh.lpc_ok != 0;
    const int lpc_order = sh.lpc_order;
    uint64_t acc[5] = { 0, 0, 0, 0, 0 };
    uint32_t fbad = 0;
#pragma unroll
    for (int i = 0; i < FRX_SPT; i++) {
        const int g = g0 + i;
        if (i < nv) {
            const int32_t *p = s + HISTN + i;
            const long long A = p[0], B = p[-1], C = p[-2], D = p[-3], E = p[-4];
            const long long e[5] = { A, A - B, A - 2 * B + C, A - 3 * B + 3 * C - D, A - 4 * B + 6 * C - 4 * D + E };
#pragma unroll
            for (int o = 0; o < 5; o++) {
                if (g >= o) {
                    const unsigned long long av = (unsigned long long)llabs(e[o]);
                    acc[o] += av;
                    if (av > 0x7FFFFFFFull) fbad |= 1u << o;
                }
            }
        }
    }
#pragma unroll
    for (int o = 0; o < 5; o++) tabs[o * FRX_ENC_THREADS + tid] = acc[o];
#pragma unroll
    for (int o = 0; o < 5; o++) acc[o] = warp_sum_u64(acc[o]);
    fbad = __reduce_or_sync(FULLMASK, fbad);
    bool bad = false;
    uint64_t ab = 0;
    if (lpc_try) {
        int32_t q[HISTN], r[FRX_SPT];
#pragma unroll
        for (int j = 0; j < HISTN; j++) q[j] = sh.qlp[j];
        const uint32_t badm = lpc_residuals_dispatch<2, HISTN>(lpc_order, s, q, sh.shift, r);
#pragma unroll
        for (int i = 0; i < FRX_SPT; i++) {
            if (i < nv && g0 + i >= lpc_order) {
                if ((badm >> i) & 1u) bad = true;
                ab += (uint32_t)abs(r[i]);
            }
        }
    }
    tabs[5 * FRX_ENC_THREADS + tid] = ab;
    const unsigned anybad = __ballot_sync(FULLMASK, bad);
    if (lane == 0) {
#pragma unroll
        for (int o = 0; o < 5; o++) sh.wfix[warp][o] = acc[o];
        sh.wbad[warp][0] = fbad; sh.wbad[warp][1] = anybad ? 1u : 0u;
    }
    compute_sync();
    if (warp == 0) {
        uint32_t fb = 0;
        for (int w2 = 0; w2 < NWARPS; w2++) fb |= sh.wbad[w2][0];
        uint64_t bestv = ~0ull, f1 = 0;
        int besto = -1;
        for (int o = 0; o < 5; o++) {
            uint64_t tot = 0;
            for (int w2 = 0; w2 < NWARPS; w2++) tot += sh.wfix[w2][o];
            if (o == 1) f1 = tot;
            if (!((fb >> o) & 1u) && tot < bestv) { bestv = tot; besto = o; }
        }
        const int fo = (nsamp > 4) ? besto : -1;
        if (lane == 0) { sh.is_const = (f1 == 0 && !(fb & 2u)) ? 1 : 0; sh.fixed_order = fo; }
        if (fo >= 0) analyse_partitions(sh, tabs + fo * FRX_ENC_THREADS, 0, nsamp, fo, po_max, kmax, 0);
    } else if (warp == 1) {
        uint32_t b = 0;
        for (int w2 = 0; w2 < NWARPS; w2++) b |= sh.wbad[w2][1];
        const bool valid = lpc_try && !b;
        if (lane == 0) sh.lpc_ok = valid ? 1 : 0;
        if (valid) analyse_partitions(sh, tabs + 5 * FRX_ENC_THREADS, 1, nsamp, lpc_order, po_max, kmax, 0);
    }
    compute_sync();
}


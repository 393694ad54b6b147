
// ------------------------------------------------------------------------------------------
// K4: LPC design on one warp from the (warp-uniform) autocorrelation.  Every lane runs the Levinson-Durbin recursion in
// registers; lane o-1 prices order o; the recursion is repeated up to the chosen order to get its coefficients, lane j owns
// coefficient j for the |max| and the quantised value (libFLAC's error-feedback rounding).  The decisions are ours; the
// bitstream stays exact because the residual is computed from the quantised integers.
// ------------------------------------------------------------------------------------------
struct LpcChoice { int ok, order, shift, precision, narrow; };
template <int MAXO>
__device__ __forceinline__ void levinson(const float *an, int upto, float *lpc, float &myerr, int &mo, int lane) {
    float err = 1.0f;
#pragma unroll
    for (int i = 0; i < MAXO; i++) {
        if (i < upto && i < mo) {
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
            if (lane == i) myerr = err;
            if (!(err > 0.f)) mo = i + 1;
        }
    }
}
template <bool WIDE, int MAXO>
__device__ __forceinline__ LpcChoice lpc_design(WarpSh &ws, const float *autoc, int max_lpc, int nsamp, int sbps, int qlp_precision, int magbits, int lane) {
    LpcChoice out;
    out.ok = 0; out.order = 0; out.shift = 0; out.precision = 0; out.narrow = 1;
    const float a0 = autoc[0];
    if (!(a0 > 0.f)) return out;
    const float inv0 = 1.0f / a0;
    float an[MAXO + 1], lpc[MAXO];
#pragma unroll
    for (int i = 0; i <= MAXO; i++) an[i] = autoc[i] * inv0;   // normalised autocorrelation
#pragma unroll
    for (int i = 0; i < MAXO; i++) lpc[i] = 0.f;
    float myerr = 0.f;
    int mo = max_lpc;
    levinson<MAXO>(an, MAXO, lpc, myerr, mo, lane);
    // expected bits per order; errs are relative to autoc[0]
    const float lscale = __log2f(a0) + __log2f(0.5f / (float)nsamp);
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
    if (order < 1 || kmin == 0xFFFFFFFFu) return out;
    int precision = qlp_precision;
    if (!WIDE) {
        const int lim = 32 - sbps - (31 - __clz(order));
        if (lim < precision) precision = lim;
    }
    if (precision < 5) return out;
    if (order < MAXO) {                            // coefficients of the chosen order: repeat the recursion up to it
#pragma unroll
        for (int i = 0; i < MAXO; i++) lpc[i] = 0.f;
        float e2 = 0.f; int mo2 = max_lpc;
        levinson<MAXO>(an, order, lpc, e2, mo2, lane);
    }
    float rv = 0.f;
#pragma unroll
    for (int j = 0; j < MAXO; j++) if (lane == j && j < order) rv = -lpc[j];
    const float cmax = __uint_as_float(__reduce_max_sync(FULLMASK, __float_as_uint(fabsf(rv))));
    if (!(cmax > 0.f) || !(cmax < 3.0e38f)) return out;
    int e2;
    (void)frexpf(cmax, &e2);
    int shift = (precision - 1) - (e2 - 1) - 1;
    if (shift > 15) shift = 15;
    if (shift < 0) return out;
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
    if (lane < FRX_MAX_LPC) ws.qlp[lane] = myq;
    int ok = 1, narrow = 1;
    if (WIDE) {
        // |prediction sum| <= sumabs * 2^mag; the residual bound decides which arithmetic is safe
        const long long M = 1ll << (magbits < 0 ? 0 : magbits);
        narrow = (sumabs * M < (1ll << 31)) ? 1 : 0;
        const long long Bnd = M + ((sumabs * M) >> shift) + 1;
        if (Bnd >= (1ll << 31)) ok = 0;            // such wild predictors never win anyway
    }
    __syncwarp();
    out.ok = ok; out.order = order; out.shift = shift; out.precision = precision; out.narrow = narrow;
    return out;
}

// tukey(0.5) window value of sample g in a frame of n samples (libFLAC's definition), for the generic path
__device__ __noinline__ float tukey_w(int g, int n) {
    if (g < 0 || g >= n) return 0.f;
    const int Np = (int)(0.25f * (float)n) - 1;
    if (Np <= 0) return 1.0f;
    if (g <= Np) return 0.5f - 0.5f * cospif((float)g / (float)Np);
    if (g >= n - Np - 1) return 0.5f - 0.5f * cospif((float)(g - (n - Np - 1) + Np) / (float)Np);
    return 1.0f;
}

// ==========================================================================================
// FAST PATH: full 4096-sample frames of lean data (|sample| <= 2^23).  Lane L owns samples [128 L, 128 L + 128).
// Sliding windows live in registers: in phase P of a rotation, slot (4 P + j) % W holds logical element j
// (0..H-1 = history, oldest first; H..H+3 = the four new samples).  NP = W / 4 steps bring the window back to phase 0.
// ==========================================================================================
template <int P, int NP, class S>
struct RunPh {
    static __device__ __forceinline__ void go(S &s, uint32_t *base, int it) {
        s.template step<P>(base + 4 * (it + P));
        if constexpr (P + 1 < NP) RunPh<P + 1, NP, S>::go(s, base, it);
    }
};
// the lane's block in two halves of 16 steps (the level-6 Rice partitions); fold() after every rotation
template <int H, class S>
__device__ __forceinline__ void walk_block(S &s, uint32_t *blk) {
    constexpr int NP = (H + 4) / 4, STEPS_H = (BLK / 2) / 4, MAINH = (STEPS_H / NP) * NP, REMH = STEPS_H - MAINH;
#pragma unroll 1
    for (int half = 0; half < 2; half++) {
        uint32_t *hb = blk + (BLK / 2) * half;
#pragma unroll 1
        for (int it = 0; it < MAINH; it += NP) { RunPh<0, NP, S>::go(s, hb, it); s.fold(); }
        if constexpr (REMH > 0) { RunPh<0, REMH, S>::go(s, hb, MAINH); s.fold(); s.template realign<REMH>(); }
        if (half == 0) s.end_first_half();
    }
}
template <int H>
__device__ __forceinline__ void load_history(const uint32_t *blk, int lane, int32_t (&hs)[H]) {
    if (lane > 0) {
#pragma unroll
        for (int k = 0; k < H / 4; k++) {
            const uint4 v = *reinterpret_cast<const uint4 *>(blk - BLK_W + BLK - H + 4 * k);
            hs[4 * k] = (int32_t)v.x; hs[4 * k + 1] = (int32_t)v.y; hs[4 * k + 2] = (int32_t)v.z; hs[4 * k + 3] = (int32_t)v.w;
        }
    } else {
#pragma unroll
        for (int k = 0; k < H; k++) hs[k] = 0;
    }
}

// ---- K2 + K3: |e_k| sums of the fixed predictors per 64-sample half, windowed autocorrelation partials ----
template <int H>
struct P1 {
    static constexpr int W = H + 4;
    float xw[W];
    float ac[H + 1];
    float c0, c1, twoc, wa, wb;   // window recurrence: c_{n+1} = twoc c_n - c_{n-1}, w = wa - wb c;  c1 = cos(phase of the next sample)
    int32_t prev, d1p, d2p, d3p;
    uint32_t f[5];                // sums of the running rotation (<= 16 samples, |e4| < 2^27)
    uint64_t cur[5], first[5];    // sums of the running half / of the first half (no dynamically indexed arrays: registers)
    template <int P>
    __device__ __forceinline__ void step(uint32_t *p) {
        const uint4 v = *reinterpret_cast<const uint4 *>(p);
        const int32_t x[4] = { (int32_t)v.x, (int32_t)v.y, (int32_t)v.z, (int32_t)v.w };
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const int32_t d1 = x[i] - prev;
            const int32_t d2 = d1 - d1p, d3 = d2 - d2p, d4 = d3 - d3p;
            prev = x[i]; d1p = d1; d2p = d2; d3p = d3;
            f[0] = __sad(x[i], 0, f[0]); f[1] = __sad(d1, 0, f[1]); f[2] = __sad(d2, 0, f[2]);      // |e| + f in one VABSDIFF each
            f[3] = __sad(d3, 0, f[3]); f[4] = __sad(d4, 0, f[4]);
            const float cn = fmaf(twoc, c1, -c0);
            const float wv = fmaf(-wb, c1, wa);
            c0 = c1; c1 = cn;
            const float xv = (float)x[i] * wv;
            xw[(4 * P + H + i) % W] = xv;
#pragma unroll
            for (int lag = 0; lag <= H; lag++) ac[lag] = fmaf(xv, xw[(4 * P + H + i - lag) % W], ac[lag]);
        }
    }
    __device__ __forceinline__ void fold() {
#pragma unroll
        for (int o = 0; o < 5; o++) { cur[o] += f[o]; f[o] = 0; }
    }
    __device__ __forceinline__ void end_first_half() {
#pragma unroll
        for (int o = 0; o < 5; o++) { first[o] = cur[o]; cur[o] = 0; }
    }
    template <int R>
    __device__ __forceinline__ void realign() {
        float t[W];
#pragma unroll
        for (int k = 0; k < W; k++) t[k] = xw[(4 * R + k) % W];
#pragma unroll
        for (int k = 0; k < W; k++) xw[k] = t[k];
    }
};
template <int H>
__device__ __forceinline__ void pass1_full(uint32_t *pcm, int lane, uint64_t (&fs)[5][2], float (&ac)[H + 1]) {
    uint32_t *blk = pcm + lane * BLK_W;
    P1<H> st;
    int32_t hs[H];
    load_history<H>(blk, lane, hs);
    // window: rising taper on lanes 0..7 (g <= 1023), falling on lanes 24..31 (g >= 3072), flat between (Np = 1023)
    {
        const int g0 = lane * BLK - H;
        const float inv = 1.0f / 1023.0f;
        const int ph = lane < 8 ? g0 : g0 - 2049;
        st.c0 = cospif((float)(ph - 1) * inv);
        st.c1 = cospif((float)ph * inv);
        st.twoc = 2.0f * cospif(inv);
        const bool taper = lane < 8 || lane >= 24;
        st.wa = taper ? 0.5f : 1.0f;
        st.wb = taper ? 0.5f : 0.0f;
    }
#pragma unroll
    for (int k = 0; k < H; k++) {
        const float cn = fmaf(st.twoc, st.c1, -st.c0);
        const float wv = fmaf(-st.wb, st.c1, st.wa);
        st.c0 = st.c1; st.c1 = cn;
        st.xw[k] = (float)hs[k] * wv;
    }
#pragma unroll
    for (int k = 0; k <= H; k++) st.ac[k] = 0.f;
    st.prev = hs[H - 1];
    st.d1p = hs[H - 1] - hs[H - 2];
    {
        const int32_t dm2 = hs[H - 2] - hs[H - 3], dm3 = hs[H - 3] - hs[H - 4];
        st.d2p = st.d1p - dm2;
        st.d3p = st.d2p - (dm2 - dm3);
    }
#pragma unroll
    for (int o = 0; o < 5; o++) { st.f[o] = 0; st.cur[o] = 0; st.first[o] = 0; }
    walk_block<H>(st, blk);
#pragma unroll
    for (int o = 0; o < 5; o++) { fs[o][0] = st.first[o]; fs[o][1] = st.cur[o]; }
#pragma unroll
    for (int k = 0; k <= H; k++) ac[k] = st.ac[k];
}

// residual of one sample, lean paths (no overflow possible by construction)
template <int ORD>
__device__ __forceinline__ int32_t lpc_res32(const int32_t *s, int i, const int32_t *q, int shift) {
    int32_t acc = 0;
#pragma unroll
    for (int j = 0; j < ORD; j++) acc += q[j] * s[i - 1 - j];
    return s[i] - (acc >> shift);
}
template <int ORD>
__device__ __forceinline__ long long lpc_res64(const int32_t *s, int i, const int32_t *q, int shift) {
    long long acc = 0;
#pragma unroll
    for (int j = 0; j < ORD; j++) acc += (long long)q[j] * (long long)s[i - 1 - j];
    return (long long)s[i] - (acc >> shift);
}

// All 16 residuals of this thread for LPC order ORD.  MODE 0: 32-bit accumulate, 1: 64-bit accumulate (result
// known to fit), 2: 64-bit with range check (fat data).  Returns a bitmask of samples whose residual does not fit.
template <int ORD, int MODE, int HISTN>
__device__ __forceinline__ uint32_t lpc_residuals(const int32_t *s, const int32_t *q, int shift, int32_t *r) {
    uint32_t bad = 0;
#pragma unroll
    for (int i = 0; i < FRX_SPT; i++) {
        if (MODE == 0) r[i] = lpc_res32<ORD>(s, HISTN + i, q, shift);
        else {
            long long rr = lpc_res64<ORD>(s, HISTN + i, q, shift);
            if (MODE == 2 && (rr <= (long long)INT32_MIN || rr > (long long)INT32_MAX)) bad |= 1u << i;
            r[i] = (int32_t)rr;
        }
    }
    return bad;
}
// Lean K5: residuals of the thread's 16 samples with all HISTN taps (coefficients beyond the order are zero), parked in the
// shared residual buffer as they are produced; returns the thread's sum of |r| over the samples that will be coded.
template <int MODE, int HISTN>
__device__ __forceinline__ uint32_t lpc_residuals_park(const int32_t *s, const int32_t *q, int shift, int32_t *res, int tid, bool plain, int nv,
                                                       int first_coded) {
    uint32_t ab = 0;
#pragma unroll
    for (int i = 0; i < FRX_SPT; i++) {
        const int32_t rr = MODE == 0 ? lpc_res32<HISTN>(s, HISTN + i, q, shift) : (int32_t)lpc_res64<HISTN>(s, HISTN + i, q, shift);
        res[RES_IDX(i, tid)] = rr;
        if (plain || (i < nv && i >= first_coded)) ab += (uint32_t)abs(rr);
    }
    return ab;
}
template <int MODE, int HISTN>
__device__ __forceinline__ uint32_t lpc_residuals_dispatch(int order, const int32_t *s, const int32_t *q, int shift, int32_t *r) {
    switch (order) {
        case 1: return lpc_residuals<1, MODE, HISTN>(s, q, shift, r);
        case 2: return lpc_residuals<2, MODE, HISTN>(s, q, shift, r);
        case 3: return lpc_residuals<3, MODE, HISTN>(s, q, shift, r);
        case 4: return lpc_residuals<4, MODE, HISTN>(s, q, shift, r);
        default: break;
    }
    if (HISTN >= 8) {
        switch (order) {
            case 5: return lpc_residuals<5, MODE, HISTN>(s, q, shift, r);
            case 6: return lpc_residuals<6, MODE, HISTN>(s, q, shift, r);
            case 7: return lpc_residuals<7, MODE, HISTN>(s, q, shift, r);
            case 8: return lpc_residuals<8, MODE, HISTN>(s, q, shift, r);
            default: break;
        }
    }
    if (HISTN >= 12) {
        switch (order) {
            case 9: return lpc_residuals<9, MODE, HISTN>(s, q, shift, r);
            case 10: return lpc_residuals<10, MODE, HISTN>(s, q, shift, r);
            case 11: return lpc_residuals<11, MODE, HISTN>(s, q, shift, r);
            default: return lpc_residuals<12, MODE, HISTN>(s, q, shift, r);
        }
    }
    return 0;
}

// fixed-order residuals; WIDE64: 64-bit with range check
template <bool WIDE64, int HISTN>
__device__ __forceinline__ uint32_t fixed_residuals(int order, const int32_t *s, int32_t *r) {
    uint32_t bad = 0;
#pragma unroll
    for (int i = 0; i < FRX_SPT; i++) {
        const int32_t *p = s + HISTN + i;
        if (WIDE64) {
            long long A = p[0], B = p[-1], C = p[-2], D = p[-3], E = p[-4], rr;
            switch (order) {
                case 0: rr = A; break;
                case 1: rr = A - B; break;
                case 2: rr = A - 2 * B + C; break;
                case 3: rr = A - 3 * B + 3 * C - D; break;
                default: rr = A - 4 * B + 6 * C - 4 * D + E; break;
            }
            if (rr <= (long long)INT32_MIN || rr > (long long)INT32_MAX) bad |= 1u << i;
            r[i] = (int32_t)rr;
        } else {
            int32_t A = p[0], B = p[-1], C = p[-2], D = p[-3], E = p[-4], rr;
            switch (order) {
                case 0: rr = A; break;
                case 1: rr = A - B; break;
                case 2: rr = A - 2 * B + C; break;
                case 3: rr = A - 3 * B + 3 * C - D; break;
                default: rr = A - 4 * B + 6 * C - 4 * D + E; break;
            }
            r[i] = rr;
        }
    }
    return bad;
}


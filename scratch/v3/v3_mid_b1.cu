// ------------------------------------------------------------------------------------------
// K1: raw staging buffer -> PCM in shared memory.  Thread t owns samples [16t, 16t+16).
// ------------------------------------------------------------------------------------------
template <int NWORDS>
__device__ __forceinline__ void lds_words16(const unsigned char *p, uint32_t *w) {   // p is 16-byte aligned (or 8/4 for short rows)
    if (NWORDS % 4 == 0) {
#pragma unroll
        for (int k = 0; k < NWORDS / 4; k++) {
            const uint4 v = *reinterpret_cast<const uint4 *>(p + 16 * k);
            w[4 * k] = v.x; w[4 * k + 1] = v.y; w[4 * k + 2] = v.z; w[4 * k + 3] = v.w;
        }
    } else {
#pragma unroll
        for (int k = 0; k < NWORDS; k++) w[k] = *reinterpret_cast<const uint32_t *>(p + 4 * k);
    }
}
template <int DT>
__device__ __forceinline__ typename FrxElem<DT>::T elem_of(const uint32_t *w, int i) {
    typedef typename FrxElem<DT>::T T;
    if (sizeof(T) == 4) {
        if (DT == FRX_F32) return (T)__uint_as_float(w[i]);
        return (T)w[i];
    } else if (sizeof(T) == 2) {
        return (T)((w[i >> 1] >> ((i & 1) * 16)) & 0xFFFFu);
    } else {
        return (T)((w[i >> 2] >> ((i & 3) * 8)) & 0xFFu);
    }
}

// the thread's 16 raw elements -> 16 PCM samples (registers)
template <int DT>
__device__ __forceinline__ void quantize_own(const unsigned char *stage, const FrxQuant &qz, int mode, int tid, int32_t *x) {
    typedef typename FrxElem<DT>::T T;
    constexpr int ESZ = (int)sizeof(T);
    constexpr int OW = FRX_SPT * ESZ / 4;
    uint32_t w[OW];
    lds_words16<OW>(stage + (size_t)(FRX_SPT * tid) * ESZ, w);
    if (mode == FRX_PCM_REFERENCE) {
#pragma unroll
        for (int i = 0; i < FRX_SPT; i++) x[i] = frx_quant_ref<DT>(elem_of<DT>(w, i), qz);
    } else if (mode == FRX_PCM_LOSSLESS) {
#pragma unroll
        for (int i = 0; i < FRX_SPT; i++) x[i] = frx_quant_lossless<DT>(elem_of<DT>(w, i));
    } else {
#pragma unroll
        for (int i = 0; i < FRX_SPT; i++) x[i] = frx_quant_spatial<DT>(elem_of<DT>(w, i));
    }
}
// K1 for one thread: quantise, zero what lies behind a short frame, store to the padded PCM buffer;
// returns (OR of the samples, OR of their magnitudes) for the wasted-bits / magnitude decisions
template <int DT, bool WIDE>
__device__ __forceinline__ uint2 k1_quantize(const unsigned char *stage, const FrxQuant &qz, int mode, int tid, int nv, int32_t *pcm) {
    int32_t x[FRX_SPT];
    quantize_own<DT>(stage, qz, mode, tid, x);
    uint32_t orv = 0, ora = 0;
#pragma unroll
    for (int i = 0; i < FRX_SPT; i++) {
        if (i >= nv) x[i] = 0;
        orv |= (uint32_t)x[i];
        if (WIDE) ora |= (uint32_t)(x[i] ^ (x[i] >> 31));
    }
#pragma unroll
    for (int k = 0; k < FRX_SPT / 4; k++)
        *reinterpret_cast<int4 *>(pcm + PCM_IDX(FRX_SPT * tid + 4 * k)) = make_int4(x[4 * k], x[4 * k + 1], x[4 * k + 2], x[4 * k + 3]);
    return make_uint2(orv, ora);
}
template <bool WIDE>
__device__ __noinline__ uint2 k1_quantize_any(int dtype, const unsigned char *stage, const FrxQuant &qz, int mode, int tid, int nv, int32_t *pcm) {
    switch (dtype) {   // the dtypes the benchmarks do not use share one out-of-line copy
        case FRX_U8: return k1_quantize<FRX_U8, WIDE>(stage, qz, mode, tid, nv, pcm);
        case FRX_I16: return k1_quantize<FRX_I16, WIDE>(stage, qz, mode, tid, nv, pcm);
        case FRX_I32: return k1_quantize<FRX_I32, WIDE>(stage, qz, mode, tid, nv, pcm);
        default: return k1_quantize<FRX_U32, WIDE>(stage, qz, mode, tid, nv, pcm);
    }
}

// PCM shared memory -> registers: HISTN samples of history + the thread's 16 samples, arithmetic shift by `sh`
template <int HISTN>
__device__ __forceinline__ void read_pcm(const int32_t *pcm, int tid, int sh, int32_t *s) {
    const int base = FRX_SPT * tid - HISTN;   // multiple of 4
#pragma unroll
    for (int k = 0; k < (HISTN + FRX_SPT) / 4; k++) {
        const int4 v = *reinterpret_cast<const int4 *>(pcm + PCM_IDX(base + 4 * k));
        s[4 * k] = v.x >> sh; s[4 * k + 1] = v.y >> sh; s[4 * k + 2] = v.z >> sh; s[4 * k + 3] = v.w >> sh;
    }
}

// generic gather (edge tiles whose rows are not 16-byte multiples): coalesced element loads into the raw stage
template <int DT>
__device__ __forceinline__ void gather_raw_t(const EncArgs &a, const frx_tile &t, int ch, uint32_t s0, int nsamp, unsigned char *stage, int tid) {
    typedef typename FrxElem<DT>::T T;
    const T *band = (const T *)a.raster + (long long)(t.first_band + ch) * a.band_stride + (long long)t.row_off * a.row_stride + t.col_off;
    T *dst = reinterpret_cast<T *>(stage);
    const uint32_t w = (uint32_t)t.w;
    uint32_t sidx = s0 + (uint32_t)tid;
    uint32_t row = sidx / w, col = sidx - row * w;
    for (int i = tid; i < nsamp; i += FRX_ENC_THREADS) {
        dst[i] = __ldg(band + (long long)row * a.row_stride + col);
        col += FRX_ENC_THREADS;
        if (col >= w) { const uint32_t q = col / w; row += q; col -= q * w; }
    }
}
__device__ __noinline__ void gather_raw(const EncArgs &a, const frx_tile &t, int ch, uint32_t s0, int nsamp, unsigned char *stage, int tid) {
    switch (frx_dtype_size(a.dtype)) {   // a pure move: only the element size matters
        case 1: gather_raw_t<FRX_U8>(a, t, ch, s0, nsamp, stage, tid); break;
        case 2: gather_raw_t<FRX_U16>(a, t, ch, s0, nsamp, stage, tid); break;
        default: gather_raw_t<FRX_U32>(a, t, ch, s0, nsamp, stage, tid); break;
    }
}


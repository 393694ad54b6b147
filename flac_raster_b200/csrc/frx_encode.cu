// frx_encode.cu -- raster -> FLAC frames on sm_100a.
//
// Kernels (SURVEY.md section 8a ids):
//   K0  frx_minmax_kernel        per-stream min/max                      (converter.py:67-68, 152-153)
//   K1..K8 frx_encode_kernel     one CTA per 4096-sample frame, persistent grid, ticket-ordered:
//        K1 tile cut + band->channel gather + PCM quantise            (converter.py:56-86, 185-194; cli.py:691-699)
//        K2 fixed-order 0..4 trial            K3 windowed autocorrelation
//        K4 Levinson-Durbin + order pick + coefficient quantisation    K5 residual
//        K6 Rice partition sums, parameter + partition-order choice    K7 bit packing + CRC-8/16
//        K8 decoupled look-back over frame sizes -> final byte offset, frame copied straight to its place
//   frx_tile_index_kernel        per-tile offset/size/minmax from the per-frame table
//
// What libFLAC does per frame (process_subframe_; restated in oracle/flac_oracle.c) is kept in spirit:
// candidates are CONSTANT, VERBATIM, the best FIXED order by sum|e|, and the LPC order picked from the
// Levinson error curve; each candidate is costed with the partition-sum Rice estimate and the cheapest
// wins.  The search is our own (the bitstream, not the decisions, is the contract).
#include <math_constants.h>

#include "frx_internal.cuh"

#define HIST FRX_MAX_LPC
#define NWARPS (FRX_ENC_THREADS / 32)
#define PCM_PAD(i) ((i) + ((i) >> 4))


// ------------------------------------------------------------------------------------------
// K0: min/max per stream.  unit u -> (tile, virtual-row range); virtual row = (channel, row).
// ------------------------------------------------------------------------------------------
struct MinMaxRaw {
    long long imin, imax;
    unsigned int fmin_o, fmax_o;  // ordered-uint encodings
    int has_nan, pad;
};

__global__ void frx_minmax_init_kernel(MinMaxRaw *mm, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        mm[i].imin = LLONG_MAX; mm[i].imax = LLONG_MIN;
        mm[i].fmin_o = 0xFFFFFFFFu; mm[i].fmax_o = 0u; mm[i].has_nan = 0; mm[i].pad = 0;
    }
}

template <int DT>
__global__ void __launch_bounds__(256) frx_minmax_kernel(const void *raster, long long band_stride, long long row_stride,
                                                         const frx_tile *tiles, const uint32_t *unit_prefix, int n_tiles,
                                                         int rows_per_unit_elems, MinMaxRaw *mm) {
    typedef typename FrxElem<DT>::T T;
    const T *base = (const T *)raster;
    // locate tile
    uint32_t u = blockIdx.x;
    int lo = 0, hi = n_tiles;
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (unit_prefix[mid] <= u) lo = mid; else hi = mid;
    }
    const frx_tile t = tiles[lo];
    uint32_t lu = u - unit_prefix[lo];
    int rpu = max(1, rows_per_unit_elems / max(1, t.w));
    long long vrow0 = (long long)lu * rpu;
    long long vrows = (long long)t.n_channels * t.h;
    long long vrow1 = min(vrows, vrow0 + rpu);
    long long imin = LLONG_MAX, imax = LLONG_MIN;
    float fmin = CUDART_INF_F, fmax = -CUDART_INF_F;
    int nan = 0;
    for (long long vr = vrow0; vr < vrow1; vr++) {
        int c = (int)(vr / t.h);
        int r = (int)(vr - (long long)c * t.h);
        const T *row = base + (long long)(t.first_band + c) * band_stride + (long long)(t.row_off + r) * row_stride + t.col_off;
        for (int x = threadIdx.x; x < t.w; x += blockDim.x) {
            T v = row[x];
            if (DT == FRX_F32) {
                float f = (float)v;
                if (f != f) nan = 1;
                fmin = fminf(fmin, f); fmax = fmaxf(fmax, f);
            } else {
                long long iv = (long long)v;
                imin = min(imin, iv); imax = max(imax, iv);
            }
        }
    }
    // warp reduce
    for (int o = 16; o > 0; o >>= 1) {
        if (DT == FRX_F32) {
            fmin = fminf(fmin, __shfl_xor_sync(0xFFFFFFFFu, fmin, o));
            fmax = fmaxf(fmax, __shfl_xor_sync(0xFFFFFFFFu, fmax, o));
            nan |= __shfl_xor_sync(0xFFFFFFFFu, nan, o);
        } else {
            imin = min(imin, __shfl_xor_sync(0xFFFFFFFFu, imin, o));
            imax = max(imax, __shfl_xor_sync(0xFFFFFFFFu, imax, o));
        }
    }
    if ((threadIdx.x & 31) == 0) {
        if (DT == FRX_F32) {
            if (fmin <= fmax) {
                atomicMin(&mm[lo].fmin_o, frx_float_ordered(fmin));
                atomicMax(&mm[lo].fmax_o, frx_float_ordered(fmax));
            }
            if (nan) atomicOr(&mm[lo].has_nan, 1);
        } else if (imin <= imax) {
            atomicMin(&mm[lo].imin, imin);
            atomicMax(&mm[lo].imax, imax);
        }
    }
}

// raw min/max -> FrxQuant (device-side, so no host round trip between K0 and K1)
__global__ void frx_quant_setup_kernel(const MinMaxRaw *mm, FrxQuant *q, int n, int dtype) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    FrxQuant o;
    o.imin = mm[i].imin; o.imax = mm[i].imax;
    o.fmin = 0.f; o.fmax = 0.f; o.range = 1.0; o.degenerate = 0; o.pad = 0;
    if (dtype == FRX_F32) {
        if (mm[i].has_nan) { o.fmin = CUDART_NAN_F; o.fmax = CUDART_NAN_F; }
        else { o.fmin = frx_float_unordered(mm[i].fmin_o); o.fmax = frx_float_unordered(mm[i].fmax_o); }
    } else {
        o.degenerate = !(o.imax > o.imin);
        switch (dtype) {
            case FRX_U8: o.range = (double)(uint8_t)((uint8_t)o.imax - (uint8_t)o.imin); break;
            case FRX_I16: o.range = (double)(int16_t)((int16_t)o.imax - (int16_t)o.imin); break;
            case FRX_U16: o.range = (double)(uint16_t)((uint16_t)o.imax - (uint16_t)o.imin); break;
            case FRX_I32: o.range = (double)(int32_t)((uint32_t)(int32_t)o.imax - (uint32_t)(int32_t)o.imin); break;
            case FRX_U32: o.range = (double)(uint32_t)((uint32_t)o.imax - (uint32_t)o.imin); break;
        }
    }
    q[i] = o;
}

__global__ void frx_tile_index_kernel(const uint32_t *frame_prefix, const uint32_t *frame_size, const uint64_t *frame_offset,
                                      const FrxQuant *q, int dtype, int n_tiles, uint64_t *tile_offset, uint64_t *tile_size,
                                      double *tile_minmax) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_tiles) return;
    uint32_t f0 = frame_prefix[i], f1 = frame_prefix[i + 1];
    uint64_t off = frame_offset[f0];
    if (tile_offset) tile_offset[i] = off;
    if (tile_size) tile_size[i] = frame_offset[f1 - 1] + frame_size[f1 - 1] - off;
    if (tile_minmax) {
        if (dtype == FRX_F32) { tile_minmax[2 * i] = (double)q[i].fmin; tile_minmax[2 * i + 1] = (double)q[i].fmax; }
        else { tile_minmax[2 * i] = (double)q[i].imin; tile_minmax[2 * i + 1] = (double)q[i].imax; }
    }
}

// ------------------------------------------------------------------------------------------
// bit writer into the zero-initialised shared frame buffer (big-endian 32-bit words).  A thread's first and
// last words may be shared with its neighbours (atomicOr); the words in between are its own (plain stores).
// ------------------------------------------------------------------------------------------
struct BitW {
    uint32_t *buf;
    uint32_t idx;
    uint64_t acc;
    int nb;
    bool first;
    __device__ __forceinline__ void init(uint32_t *b, uint32_t bitpos) {
        buf = b; idx = bitpos >> 5; nb = (int)(bitpos & 31u); acc = 0; first = true;
    }
    __device__ __forceinline__ void flushword() {
        uint32_t w = (uint32_t)(acc >> 32);
        if (first) { if (w) atomicOr(&buf[idx], w); first = false; }
        else buf[idx] = w;
        idx++; acc <<= 32; nb -= 32;
    }
    // n in 1..32, v < 2^n
    __device__ __forceinline__ void put(uint32_t v, int n) {
        acc |= (uint64_t)v << (64 - nb - n);
        nb += n;
        if (nb >= 32) flushword();
    }
    __device__ __forceinline__ void skip(uint32_t q) {  // q zero bits
        uint32_t t = (uint32_t)nb + q;
        if (t >= 32u) {
            uint32_t w = (uint32_t)(acc >> 32);
            if (first) { if (w) atomicOr(&buf[idx], w); first = false; }
            else if (w) buf[idx] = w;
            idx += t >> 5; nb = (int)(t & 31u); acc = 0;
        } else nb = (int)t;
    }
    __device__ __forceinline__ void finish() {
        if (nb > 0) {
            uint32_t w = (uint32_t)(acc >> 32);
            if (w) atomicOr(&buf[idx], w);
        }
    }
};

// ------------------------------------------------------------------------------------------
// warp helpers
// ------------------------------------------------------------------------------------------
#define FULLMASK 0xFFFFFFFFu
__device__ __forceinline__ uint64_t warp_sum_u64(uint64_t v) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULLMASK, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum_f64(double v) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULLMASK, v, o);
    return v;
}
// sum of 32 u32 values as u64 with two redux.sync (low / high halves cannot overflow)
__device__ __forceinline__ uint64_t warp_sum_u32_wide(uint32_t v) {
    uint32_t lo = __reduce_add_sync(FULLMASK, v & 0xFFFFu);
    uint32_t hi = __reduce_add_sync(FULLMASK, v >> 16);
    return (uint64_t)lo + ((uint64_t)hi << 16);
}
__device__ __forceinline__ uint32_t zigzag(int32_t r) { return ((uint32_t)r << 1) ^ (uint32_t)(r >> 31); }

// libFLAC-style Rice cost estimate from the partition sum (oracle/flac_oracle.c rice_bits_estimate)
__device__ __forceinline__ uint64_t rice_estimate(uint32_t k, uint32_t n, uint64_t s) {
    return 4ull + (uint64_t)(1 + k) * n + (k ? (s >> (k - 1)) : (s << 1)) - (n >> 1);
}
// smallest k with (n << k) >= s, capped at kmax  (libFLAC: rice parameter from the partition mean)
__device__ __forceinline__ uint32_t rice_param(uint32_t n, uint64_t s, uint32_t kmax) {
    if (n == 0 || s <= (uint64_t)n) return 0;
    int k = (64 - __clzll((long long)s)) - (32 - __clz((int)n)) - 1;  // floor(log2 s) - floor(log2 n) - 1
    if (k < 0) k = 0;
    while (((uint64_t)n << k) < s) k++;  // at most 2 iterations
    return (uint32_t)k > kmax ? kmax : (uint32_t)k;
}

// ------------------------------------------------------------------------------------------
// shared state of one encoder CTA
// ------------------------------------------------------------------------------------------
struct EncSmall {
    uint32_t idx[2];         // ticket double buffer
    uint32_t bitpos;         // running bit position in the frame buffer
    uint32_t tailw;          // the frame's last partial word (CRC tail)
    uint64_t out_offset;
    // per-subframe decisions
    int wasted, magbits, fixed_order, is_const, lean;
    int lpc_ok, lpc_order, shift, precision, lpc_narrow, lpc_valid;
    int32_t qlp[FRX_MAX_LPC];
    int32_t warm[FRX_MAX_LPC];
    uint32_t kparam[2][64];  // candidate 0 = fixed, 1 = lpc
    int cand_po[2], cand_rice2[2];
    uint64_t cand_bits[2];
    uint64_t fsum[5];
    // reductions
    uint64_t red64[NWARPS][5];
    uint32_t red32[NWARPS][4];
    double autoc[FRX_MAX_LPC + 1];
    uint32_t scan[NWARPS];
    uint32_t crc_part[2];
    uint16_t crc_tab[4][256];
};
#define ENC_UNION_BYTES 17408  // pcm[4096 + 255 pad + 1] int32  ==  acpart[13][256] float + tabs[2][256] u64
#define ENC_SMALL_BYTES ((sizeof(EncSmall) + 15) & ~(size_t)15)
#define CRC_THREADS 64

// One warp analyses one candidate: partition sums from the per-thread |r| sums, Rice parameter per partition,
// estimated bits per partition order, best order.  Lane p owns finest-level partition p (2p, 2p+1 when po_max == 6);
// coarser levels are built with shuffles, so partition q of level L lives in lane q << (F - L).
__device__ void analyse_partitions(EncSmall &sh, const uint64_t *tabs, int cand, int nsamp, int pred_order, int po_max, uint32_t kmax) {
    const int lane = threadIdx.x & 31;
    const int F = po_max > 5 ? 5 : po_max;               // finest level held one-per-lane
    const int nthr = (nsamp + FRX_SPT - 1) / FRX_SPT;
    uint64_t sum = 0, sumB = 0;                          // sumB: second half when po_max == 6
    uint64_t best = ~0ull;
    int best_po = 0;
    uint32_t kbest0 = 0, kbest1 = 0;
    if (po_max == 6) {
        const int tpp = (nsamp >> 6) / FRX_SPT;          // threads per level-6 partition
        for (int t = 0; t < tpp; t++) { sum += tabs[(2 * lane) * tpp + t]; sumB += tabs[(2 * lane + 1) * tpp + t]; }
        uint32_t n0 = (uint32_t)(nsamp >> 6) - (lane == 0 ? (uint32_t)pred_order : 0u), n1 = (uint32_t)(nsamp >> 6);
        uint32_t k0 = rice_param(n0, sum, kmax), k1 = rice_param(n1, sumB, kmax);
        uint64_t b = rice_estimate(k0, n0, sum) + rice_estimate(k1, n1, sumB);
        b = warp_sum_u64(b) + 6;
        best = b; best_po = 6; kbest0 = k0; kbest1 = k1;
        sum += sumB;
    } else if (F == 0) {
        for (int t = lane; t < nthr; t += 32) sum += tabs[t];
        sum = warp_sum_u64(sum);
    } else {
        const int tpp = (nsamp >> F) / FRX_SPT;          // threads per finest partition (>= 1)
        if (lane < (1 << F))
            for (int t = 0; t < tpp; t++) sum += tabs[lane * tpp + t];
    }
    // levels F .. 0
    for (int L = F; L >= 0; L--) {
        const int step = 1 << (F - L);                   // lanes that are multiples of step hold level-L partitions
        const bool mine = (lane & (step - 1)) == 0 && lane < (1 << F);
        const int p = lane >> (F - L);
        uint32_t n = (uint32_t)(nsamp >> L) - (p == 0 ? (uint32_t)pred_order : 0u);
        uint32_t k = 0;
        uint64_t b = 0;
        if (mine) { k = rice_param(n, sum, kmax); b = rice_estimate(k, n, sum); }
        b = warp_sum_u64(b) + 6;
        if (b < best) { best = b; best_po = L; kbest0 = k; }
        if (L > 0) {
            uint64_t o = __shfl_down_sync(FULLMASK, sum, step);
            sum += o;                                    // valid in lanes that are multiples of 2*step
        }
    }
    // publish parameters of the best level
    if (best_po == 6) { sh.kparam[cand][2 * lane] = kbest0; sh.kparam[cand][2 * lane + 1] = kbest1; }
    else {
        const int step = 1 << (F - best_po);
        if ((lane & (step - 1)) == 0 && lane < (1 << F)) sh.kparam[cand][lane >> (F - best_po)] = kbest0;
    }
    uint32_t big = __ballot_sync(FULLMASK, kbest0 >= 15u || (best_po == 6 && kbest1 >= 15u));
    if (lane == 0) { sh.cand_po[cand] = best_po; sh.cand_bits[cand] = best; sh.cand_rice2[cand] = big ? 1 : 0; }
    __syncwarp();
}

// ------------------------------------------------------------------------------------------
// K1: tile cut + band -> channel gather + PCM quantise, coalesced (thread t takes samples t, t+256, ...)
// ------------------------------------------------------------------------------------------
template <int DT>
__device__ __forceinline__ void enc_load(const EncArgs &a, const frx_tile &t, const FrxQuant &qz, int ch, uint32_t s0, int nsamp,
                                         int32_t *pcm, int tid) {
    typedef typename FrxElem<DT>::T T;
    const T *band = (const T *)a.raster + (long long)(t.first_band + ch) * a.band_stride + (long long)t.row_off * a.row_stride + t.col_off;
    const uint32_t w = (uint32_t)t.w;
    uint32_t sidx = s0 + (uint32_t)tid;
    uint32_t row = sidx / w, col = sidx - row * w;
    T raw[FRX_SPT];
#pragma unroll
    for (int k = 0; k < FRX_SPT; k++) {
        const int i = tid + k * FRX_ENC_THREADS;
        raw[k] = (i < nsamp) ? __ldg(band + (long long)row * a.row_stride + col) : (T)0;
        col += FRX_ENC_THREADS;
        if (col >= w) {
            if (w >= FRX_ENC_THREADS) { col -= w; row++; }
            else { uint32_t q = col / w; row += q; col -= q * w; }
        }
    }
#pragma unroll
    for (int k = 0; k < FRX_SPT; k++) {
        const int i = tid + k * FRX_ENC_THREADS;
        pcm[PCM_PAD(i)] = (i < nsamp) ? frx_quant<DT>(raw[k], qz, a.pcm_mode) : 0;
    }
}

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

// ------------------------------------------------------------------------------------------
// The frame encoder.  WIDE: PCM wider than 16 bits (64-bit LPC accumulate, RICE2).  HISTN: history samples a thread
// keeps = the largest LPC order of the compression level (4 when the level has no LPC stage).
// ------------------------------------------------------------------------------------------
template <bool WIDE, int HISTN>
__global__ void __launch_bounds__(FRX_ENC_THREADS, 2) frx_encode_kernel(const EncArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    EncSmall &sh = *reinterpret_cast<EncSmall *>(smem_raw);
    unsigned char *ureg = smem_raw + ENC_SMALL_BYTES;
    int32_t *pcm = reinterpret_cast<int32_t *>(ureg);
    float *acpart = reinterpret_cast<float *>(ureg);                          // [13][256]
    uint64_t *tabs = reinterpret_cast<uint64_t *>(ureg + 13 * FRX_ENC_THREADS * 4);  // [2][256]
    uint32_t *bits = reinterpret_cast<uint32_t *>(ureg + ENC_UNION_BYTES);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t kmax = a.bps > 16 ? 30u : 14u;  // RICE2 only for bps > 16 (libFLAC)
    const int g0 = tid * FRX_SPT;                  // first sample index of this thread

    for (int i = tid; i < a.frame_words; i += FRX_ENC_THREADS) bits[i] = 0u;
    for (int i = tid; i < 1024; i += FRX_ENC_THREADS) (&sh.crc_tab[0][0])[i] = a.crc_tab[i];
    if (tid == 0) sh.idx[0] = atomicAdd(a.ticket, 1u);
    int par = 0;

    for (;;) {
        __syncthreads();  // S0: previous frame left the buffers clean; ticket visible
        const uint32_t f = sh.idx[par];
        if (f >= a.total_frames) break;
        if (tid == 0) sh.idx[par ^ 1] = atomicAdd(a.ticket, 1u);  // prefetch the next ticket (tickets stay ordered per CTA)
        par ^= 1;

        // ---- locate tile / frame ----
        int lo = 0, hi = a.n_tiles;
        while (hi - lo > 1) {
            int mid = (lo + hi) >> 1;
            if (__ldg(&a.frame_prefix[mid]) <= f) lo = mid; else hi = mid;
        }
        const frx_tile t = a.tiles[lo];
        const FrxQuant qz = a.quant[lo];
        const uint32_t fj = f - a.frame_prefix[lo];
        const long long tile_n = (long long)t.h * t.w;
        const uint32_t s0 = fj * (uint32_t)a.blocksize;
        const int nsamp = (int)min((long long)a.blocksize, tile_n - (long long)s0);
        const int nthr = (nsamp + FRX_SPT - 1) / FRX_SPT;
        const int nv = min(FRX_SPT, max(0, nsamp - g0));  // valid samples of this thread
        const bool full = (nsamp == a.blocksize);

        // ---- frame header (thread 0) ----
        if (tid == 0) {
            uint8_t hb[16];
            int n = 0;
            int bs_code, bs_extra = 0;
            switch (nsamp) {
                case 192: bs_code = 1; break; case 576: bs_code = 2; break; case 1152: bs_code = 3; break;
                case 2304: bs_code = 4; break; case 4608: bs_code = 5; break; case 256: bs_code = 8; break;
                case 512: bs_code = 9; break; case 1024: bs_code = 10; break; case 2048: bs_code = 11; break;
                case 4096: bs_code = 12; break; case 8192: bs_code = 13; break; case 16384: bs_code = 14; break;
                case 32768: bs_code = 15; break;
                default: if (nsamp <= 256) { bs_code = 6; bs_extra = 1; } else { bs_code = 7; bs_extra = 2; }
            }
            int sz_code = a.bps == 16 ? 4 : (a.bps == 32 ? 7 : (a.bps == 8 ? 1 : (a.bps == 24 ? 6 : 0)));
            hb[n++] = 0xFF; hb[n++] = 0xF8;
            hb[n++] = (uint8_t)((bs_code << 4) | a.sr_code);
            hb[n++] = (uint8_t)(((t.n_channels - 1) << 4) | (sz_code << 1));
            uint32_t v = fj;
            if (v < 0x80u) hb[n++] = (uint8_t)v;
            else if (v < 0x800u) { hb[n++] = (uint8_t)(0xC0 | (v >> 6)); hb[n++] = (uint8_t)(0x80 | (v & 0x3F)); }
            else if (v < 0x10000u) { hb[n++] = (uint8_t)(0xE0 | (v >> 12)); hb[n++] = (uint8_t)(0x80 | ((v >> 6) & 0x3F)); hb[n++] = (uint8_t)(0x80 | (v & 0x3F)); }
            else if (v < 0x200000u) { hb[n++] = (uint8_t)(0xF0 | (v >> 18)); hb[n++] = (uint8_t)(0x80 | ((v >> 12) & 0x3F)); hb[n++] = (uint8_t)(0x80 | ((v >> 6) & 0x3F)); hb[n++] = (uint8_t)(0x80 | (v & 0x3F)); }
            else if (v < 0x4000000u) { hb[n++] = (uint8_t)(0xF8 | (v >> 24)); hb[n++] = (uint8_t)(0x80 | ((v >> 18) & 0x3F)); hb[n++] = (uint8_t)(0x80 | ((v >> 12) & 0x3F)); hb[n++] = (uint8_t)(0x80 | ((v >> 6) & 0x3F)); hb[n++] = (uint8_t)(0x80 | (v & 0x3F)); }
            else { hb[n++] = (uint8_t)(0xFC | (v >> 30)); hb[n++] = (uint8_t)(0x80 | ((v >> 24) & 0x3F)); hb[n++] = (uint8_t)(0x80 | ((v >> 18) & 0x3F)); hb[n++] = (uint8_t)(0x80 | ((v >> 12) & 0x3F)); hb[n++] = (uint8_t)(0x80 | ((v >> 6) & 0x3F)); hb[n++] = (uint8_t)(0x80 | (v & 0x3F)); }
            if (bs_extra == 1) hb[n++] = (uint8_t)(nsamp - 1);
            else if (bs_extra == 2) { hb[n++] = (uint8_t)((nsamp - 1) >> 8); hb[n++] = (uint8_t)(nsamp - 1); }
            if (a.sr_code == 13) { hb[n++] = (uint8_t)(a.sample_rate >> 8); hb[n++] = (uint8_t)a.sample_rate; }
            uint32_t c = 0;
            for (int i = 0; i < n; i++) c = frx_crc8_byte(c, hb[i]);
            hb[n++] = (uint8_t)c;
            for (int i = n; i < 16; i++) hb[i] = 0;
            for (int wv = 0; wv < 4; wv++)
                bits[wv] = ((uint32_t)hb[4 * wv] << 24) | ((uint32_t)hb[4 * wv + 1] << 16) | ((uint32_t)hb[4 * wv + 2] << 8) | hb[4 * wv + 3];
            sh.bitpos = (uint32_t)n * 8u;
        }

        // ================= per channel =================
        for (int ch = 0; ch < t.n_channels; ch++) {
          bool force_verbatim = false;
          for (;;) {  // second pass only when the Rice-coded subframe would not beat VERBATIM (then it is redone as VERBATIM)
            __syncthreads();  // S2: union region free, bitpos visible
            switch (a.dtype) {
                case FRX_U8: enc_load<FRX_U8>(a, t, qz, ch, s0, nsamp, pcm, tid); break;
                case FRX_I16: enc_load<FRX_I16>(a, t, qz, ch, s0, nsamp, pcm, tid); break;
                case FRX_U16: enc_load<FRX_U16>(a, t, qz, ch, s0, nsamp, pcm, tid); break;
                case FRX_I32: enc_load<FRX_I32>(a, t, qz, ch, s0, nsamp, pcm, tid); break;
                case FRX_U32: enc_load<FRX_U32>(a, t, qz, ch, s0, nsamp, pcm, tid); break;
                default: enc_load<FRX_F32>(a, t, qz, ch, s0, nsamp, pcm, tid); break;
            }
            __syncthreads();  // S3
            // ---- registers: HISTN history + 16 samples ----
            int32_t s[HISTN + FRX_SPT];
#pragma unroll
            for (int i = 0; i < HISTN + FRX_SPT; i++) {
                int g = g0 - HISTN + i;
                s[i] = (g >= 0) ? pcm[PCM_PAD(g)] : 0;
            }
            uint32_t orv = 0, ora = 0;
#pragma unroll
            for (int i = 0; i < FRX_SPT; i++) { orv |= (uint32_t)s[HISTN + i]; ora |= (uint32_t)abs(s[HISTN + i]); }
            orv = __reduce_or_sync(FULLMASK, orv);
            ora = __reduce_or_sync(FULLMASK, ora);
            // ---- K2 (lean): fixed-order trial in 32-bit, sum |e_k| over samples g >= k ----
            uint32_t fa0 = 0, fa1 = 0, fa2 = 0, fa3 = 0, fa4 = 0;
            {
                const bool first = (tid == 0);
                int32_t dm1 = s[HISTN - 1] - s[HISTN - 2], dm2 = s[HISTN - 2] - s[HISTN - 3], dm3 = s[HISTN - 3] - s[HISTN - 4];
                int32_t d1p = dm1, d2p = dm1 - dm2, d3p = (dm1 - dm2) - (dm2 - dm3);
#pragma unroll
                for (int i = 0; i < FRX_SPT; i++) {
                    int32_t x = s[HISTN + i];
                    int32_t d1 = x - s[HISTN + i - 1];
                    int32_t d2 = d1 - d1p, d3 = d2 - d2p, d4 = d3 - d3p;
                    d1p = d1; d2p = d2; d3p = d3;
                    if (i < nv) {
                        fa0 += (uint32_t)abs(x);
                        if (i >= 1 || !first) fa1 += (uint32_t)abs(d1);
                        if (i >= 2 || !first) fa2 += (uint32_t)abs(d2);
                        if (i >= 3 || !first) fa3 += (uint32_t)abs(d3);
                        if (i >= 4 || !first) fa4 += (uint32_t)abs(d4);
                    }
                }
            }
            {
                uint64_t w0 = warp_sum_u32_wide(fa0), w1 = warp_sum_u32_wide(fa1), w2 = warp_sum_u32_wide(fa2),
                         w3 = warp_sum_u32_wide(fa3), w4 = warp_sum_u32_wide(fa4);
                if (lane == 0) {
                    sh.red64[warp][0] = w0; sh.red64[warp][1] = w1; sh.red64[warp][2] = w2; sh.red64[warp][3] = w3; sh.red64[warp][4] = w4;
                    sh.red32[warp][0] = orv; sh.red32[warp][1] = ora;
                }
            }
            __syncthreads();  // S4: every thread holds its samples; pcm may be overwritten by acpart
            // ---- K3: windowed autocorrelation partials (float within the thread, double across threads) ----
            const int max_lpc = min(min(a.max_lpc_order, HISTN), nsamp - 1);
            if (max_lpc > 0) {
                float xw[HISTN + FRX_SPT];
                if (full) {
                    const float4 *w4 = reinterpret_cast<const float4 *>(a.window + g0);
#pragma unroll
                    for (int v = 0; v < FRX_SPT / 4; v++) {
                        float4 wv = __ldg(w4 + v);
                        xw[HISTN + 4 * v] = (float)s[HISTN + 4 * v] * wv.x; xw[HISTN + 4 * v + 1] = (float)s[HISTN + 4 * v + 1] * wv.y;
                        xw[HISTN + 4 * v + 2] = (float)s[HISTN + 4 * v + 2] * wv.z; xw[HISTN + 4 * v + 3] = (float)s[HISTN + 4 * v + 3] * wv.w;
                    }
                    if (tid > 0) {
                        const float4 *h4 = reinterpret_cast<const float4 *>(a.window + g0 - HISTN);
#pragma unroll
                        for (int v = 0; v < HISTN / 4; v++) {
                            float4 wv = __ldg(h4 + v);
                            xw[4 * v] = (float)s[4 * v] * wv.x; xw[4 * v + 1] = (float)s[4 * v + 1] * wv.y;
                            xw[4 * v + 2] = (float)s[4 * v + 2] * wv.z; xw[4 * v + 3] = (float)s[4 * v + 3] * wv.w;
                        }
                    } else {
#pragma unroll
                        for (int i = 0; i < HISTN; i++) xw[i] = 0.f;
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < HISTN + FRX_SPT; i++) {
                        int g = g0 - HISTN + i;
                        float wv = 0.f;
                        if (g >= 0 && g < nsamp) {
                            int Np = (int)(0.25f * (float)nsamp) - 1;  // tukey(0.5) of length nsamp
                            wv = 1.0f;
                            if (Np > 0) {
                                if (g <= Np) wv = 0.5f - 0.5f * cospif((float)g / (float)Np);
                                else if (g >= nsamp - Np - 1) wv = 0.5f - 0.5f * cospif((float)(g - (nsamp - Np - 1) + Np) / (float)Np);
                            }
                        }
                        xw[i] = (float)s[i] * wv;
                    }
                }
                float ac[HISTN + 1];
#pragma unroll
                for (int lag = 0; lag <= HISTN; lag++) ac[lag] = 0.f;
#pragma unroll
                for (int i = 0; i < FRX_SPT; i++) {
#pragma unroll
                    for (int lag = 0; lag <= HISTN; lag++) ac[lag] = fmaf(xw[HISTN + i], xw[HISTN + i - lag], ac[lag]);
                }
#pragma unroll
                for (int lag = 0; lag <= HISTN; lag++) acpart[lag * FRX_ENC_THREADS + tid] = ac[lag];
            }
            __syncthreads();  // S5
            // warp w reduces lags w, w+8 (double accumulation across threads)
            for (int lag = warp; lag <= max_lpc; lag += NWARPS) {
                double d = 0.0;
#pragma unroll
                for (int j = 0; j < FRX_ENC_THREADS / 32; j++) d += (double)acpart[lag * FRX_ENC_THREADS + lane + 32 * j];
                d = warp_sum_f64(d);
                if (lane == 0) sh.autoc[lag] = d;
            }
            // warp 7: block totals of the fixed trial, wasted bits, magnitude, order decision
            if (warp == NWARPS - 1) {
                uint64_t tot = 0;
                uint32_t o32 = 0;
                if (lane < 5) { for (int w2 = 0; w2 < NWARPS; w2++) tot += sh.red64[w2][lane]; }
                else if (lane < 7) { for (int w2 = 0; w2 < NWARPS; w2++) o32 |= sh.red32[w2][lane - 5]; }
                const uint32_t all_or = __shfl_sync(FULLMASK, o32, 5), all_abs = __shfl_sync(FULLMASK, o32, 6);
                const int wasted = all_or ? (__ffs((int)all_or) - 1) : 0;
                const int magbits = 32 - __clz((int)all_abs);  // bits of the largest magnitude (0 for silence)
                const bool lean = !WIDE || magbits <= 23;
                const uint64_t f1 = __shfl_sync(FULLMASK, tot, 1);
                // best fixed order among lanes 0..4 (lowest order wins ties)
                uint64_t bestv = (lane < 5) ? tot : ~0ull;
                int besto = lane;
                for (int o = 1; o < 8; o <<= 1) {
                    uint64_t ov = __shfl_xor_sync(FULLMASK, bestv, o);
                    int oo = __shfl_xor_sync(FULLMASK, besto, o);
                    if (ov < bestv || (ov == bestv && oo < besto)) { bestv = ov; besto = oo; }
                }
                if (lane < 5) sh.fsum[lane] = tot;
                if (lane == 0) {
                    sh.wasted = wasted; sh.magbits = magbits; sh.lean = lean ? 1 : 0;
                    sh.is_const = (f1 == 0) ? 1 : 0;   // sum |s[g]-s[g-1]| == 0  <=>  all samples equal
                    sh.fixed_order = (nsamp > 4) ? besto : -1;
                }
            }
            __syncthreads();  // S6
            const int wasted = sh.wasted;
            const bool lean = sh.lean != 0;
            const int sbps = a.bps - wasted;
            if (wasted) {
#pragma unroll
                for (int i = 0; i < HISTN + FRX_SPT; i++) s[i] >>= wasted;
            }
            // ---- fat data (|s| >= 2^23: lossless int32/float bit patterns): redo the fixed trial in 64-bit ----
            if (WIDE && !lean) {
                uint64_t acc[5] = { 0, 0, 0, 0, 0 };
                uint32_t fbad = 0;
#pragma unroll
                for (int i = 0; i < FRX_SPT; i++) {
                    int g = g0 + i;
                    if (i < nv) {
                        const int32_t *p = s + HISTN + i;
                        long long A = p[0], B = p[-1], C = p[-2], D = p[-3], E = p[-4];
                        long long e[5] = { A, A - B, A - 2 * B + C, A - 3 * B + 3 * C - D, A - 4 * B + 6 * C - 4 * D + E };
#pragma unroll
                        for (int o = 0; o < 5; o++) {
                            if (g >= o) {
                                unsigned long long av = (unsigned long long)llabs(e[o]);
                                acc[o] += av;
                                if (av > 0x7FFFFFFFull) fbad |= 1u << o;
                            }
                        }
                    }
                }
#pragma unroll
                for (int o = 0; o < 5; o++) acc[o] = warp_sum_u64(acc[o]);
                fbad = __reduce_or_sync(FULLMASK, fbad);
                if (lane == 0) {
#pragma unroll
                    for (int o = 0; o < 5; o++) sh.red64[warp][o] = acc[o];
                    sh.red32[warp][2] = fbad;
                }
                __syncthreads();
                if (tid == 0) {
                    uint64_t bestv = ~0ull;
                    int besto = -1;
                    uint32_t fb = 0;
                    for (int w2 = 0; w2 < NWARPS; w2++) fb |= sh.red32[w2][2];
                    for (int o = 0; o < 5; o++) {
                        uint64_t tot = 0;
                        for (int w2 = 0; w2 < NWARPS; w2++) tot += sh.red64[w2][o];
                        sh.fsum[o] = tot;
                        if (!((fb >> o) & 1u) && tot < bestv) { bestv = tot; besto = o; }
                    }
                    sh.is_const = (sh.fsum[1] == 0 && !(fb & 2u)) ? 1 : 0;
                    sh.fixed_order = (nsamp > 4) ? besto : -1;
                }
                // (published by the S7 barrier below)
            }
            // ---- K4: Levinson-Durbin, order pick, coefficient quantisation (thread 0) ----
            if (tid == 0) {
                sh.lpc_ok = 0;
                const bool is_const0 = sh.is_const != 0;  // lean: warp 7 before S6; fat: this thread, just above
                const double a0 = sh.autoc[0];
                if (!is_const0 && max_lpc > 0 && nsamp > 4 && a0 != 0.0) {
                    double lpc[FRX_MAX_LPC], err = a0;
                    double errs[FRX_MAX_LPC];
                    double coefs[FRX_MAX_LPC][FRX_MAX_LPC];
                    int mo = max_lpc;
                    for (int i = 0; i < mo; i++) {
                        double r = -sh.autoc[i + 1];
                        for (int j = 0; j < i; j++) r -= lpc[j] * sh.autoc[i - j];
                        r /= err;
                        lpc[i] = r;
                        int j;
                        for (j = 0; j < (i >> 1); j++) {
                            double tmp = lpc[j];
                            lpc[j] += r * lpc[i - 1 - j];
                            lpc[i - 1 - j] += r * tmp;
                        }
                        if (i & 1) lpc[j] += lpc[j] * r;
                        err *= (1.0 - r * r);
                        for (j = 0; j <= i; j++) coefs[i][j] = -lpc[j];
                        errs[i] = err;
                        if (!(err > 0.0)) { mo = i + 1; break; }
                    }
                    // expected bits per order (the autocorrelation was taken before the wasted-bit shift)
                    double best_bits = 1e300;
                    int best = 1;
                    const double scale = 0.5 / (double)nsamp * (wasted ? ldexp(1.0, -2 * wasted) : 1.0);
                    for (int o = 1; o <= mo; o++) {
                        double e = errs[o - 1];
                        double bpsr = e > 0.0 ? fmax(0.0, 0.5 * log2(scale * e)) : 0.0;
                        double b = bpsr * (double)(nsamp - o) + (double)(o * (sbps + a.qlp_precision));
                        if (b < best_bits) { best_bits = b; best = o; }
                    }
                    int order = best;
                    int precision = a.qlp_precision;
                    if (!WIDE) {
                        int lim = 32 - sbps - (31 - __clz(order));
                        if (lim < precision) precision = lim;
                    }
                    if (precision >= 5) {
                        double cmax = 0.0;
                        for (int j = 0; j < order; j++) cmax = fmax(cmax, fabs(coefs[order - 1][j]));
                        if (cmax > 0.0) {
                            int e2;
                            (void)frexp(cmax, &e2);
                            int shift = (precision - 1) - (e2 - 1) - 1;
                            if (shift > 15) shift = 15;
                            if (shift >= 0) {
                                const int qmax = (1 << (precision - 1)) - 1, qmin = -(1 << (precision - 1));
                                double error = 0.0;
                                long long sumabs = 0;
                                for (int j = 0; j < order; j++) {
                                    error += coefs[order - 1][j] * (double)(1 << shift);
                                    long long qq = llrint(error);
                                    if (qq > qmax) qq = qmax; else if (qq < qmin) qq = qmin;
                                    error -= (double)qq;
                                    sh.qlp[j] = (int32_t)qq;
                                    sumabs += qq < 0 ? -qq : qq;
                                }
                                for (int j = order; j < FRX_MAX_LPC; j++) sh.qlp[j] = 0;
                                sh.lpc_order = order; sh.shift = shift; sh.precision = precision;
                                int ok = 1, narrow = 1;
                                if (WIDE) {
                                    // |prediction sum| <= sumabs * 2^mag; the residual bound decides which arithmetic is safe
                                    const int mag = sh.magbits - wasted;
                                    const long long M = 1ll << (mag < 0 ? 0 : mag);
                                    narrow = (sumabs * M < (1ll << 31)) ? 1 : 0;
                                    if (lean) {
                                        long long Bnd = M + ((sumabs * M) >> shift) + 1;
                                        if (Bnd >= (1ll << 27)) ok = 0;  // keeps 32-bit residual sums exact; such wild predictors never win anyway
                                    }
                                }
                                sh.lpc_narrow = narrow;
                                sh.lpc_ok = ok;
                            }
                        }
                    }
                }
            }
            __syncthreads();  // S7
            // ---- K5: per-thread |r| sums of the two candidates (tabs lives where acpart was: S7 passed) ----
            const int fixed_order = sh.fixed_order;
            const bool is_const = sh.is_const != 0;
            const bool lpc_try = sh.lpc_ok != 0;
            const int lpc_order = sh.lpc_order;
            const int lshift = sh.shift;
            int32_t q[HISTN];
#pragma unroll
            for (int j = 0; j < HISTN; j++) q[j] = sh.qlp[j];
            const int lmode = !WIDE ? 0 : (lean ? (sh.lpc_narrow ? 0 : 1) : 2);
            {
                uint64_t ab = 0;
                bool bad = false;
                if (lpc_try) {
                    int32_t r[FRX_SPT];
                    uint32_t badm = 0;
                    if (lmode == 0) badm = lpc_residuals_dispatch<0, HISTN>(lpc_order, s, q, lshift, r);
                    else if (WIDE && lmode == 1) badm = lpc_residuals_dispatch<1, HISTN>(lpc_order, s, q, lshift, r);
                    else if (WIDE) badm = lpc_residuals_dispatch<2, HISTN>(lpc_order, s, q, lshift, r);
                    uint32_t ab32 = 0;
#pragma unroll
                    for (int i = 0; i < FRX_SPT; i++) {
                        const bool valid = (i < nv) && (g0 + i >= lpc_order);
                        if (valid) {
                            if ((badm >> i) & 1u) bad = true;
                            if (WIDE && !lean) ab += (uint32_t)abs(r[i]); else ab32 += (uint32_t)abs(r[i]);
                        }
                    }
                    if (!(WIDE && !lean)) ab = ab32;
                }
                tabs[FRX_ENC_THREADS + tid] = ab;
                if (WIDE && !lean) {
                    unsigned anybad = __ballot_sync(FULLMASK, bad);
                    if (lane == 0) sh.red32[warp][3] = anybad ? 1u : 0u;
                }
                uint64_t fb = 0;
                if (fixed_order >= 0) {
                    if (WIDE && !lean) {
                        int32_t rf[FRX_SPT];
                        (void)fixed_residuals<true, HISTN>(fixed_order, s, rf);
#pragma unroll
                        for (int i = 0; i < FRX_SPT; i++)
                            if (i < nv && g0 + i >= fixed_order) fb += (uint32_t)abs(rf[i]);
                    } else {
                        uint32_t sel = fixed_order == 0 ? fa0 : fixed_order == 1 ? fa1 : fixed_order == 2 ? fa2 : fixed_order == 3 ? fa3 : fa4;
                        fb = sel >> wasted;  // all samples are multiples of 2^wasted, so the sums scale exactly
                    }
                }
                tabs[tid] = fb;
            }
            __syncthreads();  // S8
            // ---- K6: Rice partition analysis, warp 0 = fixed, warp 1 = lpc ----
            int po_max = 0;
            {
                int tz = __ffs(nsamp) - 1;               // trailing zeros of the block size
                po_max = min(a.max_po, max(0, tz - 4));  // partitions stay multiples of 16 samples
                while (po_max > 0 && (nsamp >> po_max) <= FRX_MAX_LPC) po_max--;
            }
            bool lpc_valid = lpc_try;
            if (WIDE && !lean && lpc_try) {
                uint32_t b = 0;
#pragma unroll
                for (int w2 = 0; w2 < NWARPS; w2++) b |= sh.red32[w2][3];
                if (b) lpc_valid = false;
            }
            if (warp == 0 && fixed_order >= 0) analyse_partitions(sh, tabs, 0, nsamp, fixed_order, po_max, kmax);
            if (warp == 1 && lpc_valid) analyse_partitions(sh, tabs + FRX_ENC_THREADS, 1, nsamp, lpc_order, po_max, kmax);
            __syncthreads();  // S9

            // ---- choose ----
            const uint64_t verb_bits = 8ull + (uint64_t)wasted + (uint64_t)nsamp * (uint64_t)sbps;
            int type = 1, order = 0, cand = 0;
            uint64_t best_bits = verb_bits;
            if (is_const) { type = 0; best_bits = 8ull + wasted + sbps; }
            else if (!force_verbatim) {
                if (fixed_order >= 0) {
                    uint64_t b = 8ull + wasted + (uint64_t)fixed_order * sbps + sh.cand_bits[0];
                    if (b < best_bits) { best_bits = b; type = 2; order = fixed_order; cand = 0; }
                }
                if (lpc_valid) {
                    uint64_t b = 8ull + wasted + 9ull + (uint64_t)lpc_order * (uint64_t)(sh.precision + sbps) + sh.cand_bits[1];
                    if (b < best_bits) { best_bits = b; type = 3; order = lpc_order; cand = 1; }
                }
            }

            // ---- K7a: winner's residuals (zig-zag folded), exact bit count per thread, block scan ----
            const uint32_t sub_start = sh.bitpos;
            uint32_t u[FRX_SPT];
            int po = 0;
            uint32_t kk = 0;
            int psize = nsamp;
            bool rice2 = false;
            uint32_t mybits = 0;
            if (type >= 2) {
                po = sh.cand_po[cand];
                psize = nsamp >> po;
                kk = sh.kparam[cand][min(g0 / psize, (1 << po) - 1)];
                rice2 = sh.cand_rice2[cand] != 0;
                int32_t r[FRX_SPT];
                if (type == 3) {
                    if (lmode == 0) (void)lpc_residuals_dispatch<0, HISTN>(order, s, q, lshift, r);
                    else if (WIDE) (void)lpc_residuals_dispatch<1, HISTN>(order, s, q, lshift, r);
                } else {
                    if (WIDE && !lean) (void)fixed_residuals<true, HISTN>(order, s, r);
                    else (void)fixed_residuals<false, HISTN>(order, s, r);
                }
#pragma unroll
                for (int i = 0; i < FRX_SPT; i++) {
                    u[i] = zigzag(r[i]);
                    if (i < nv && g0 + i >= order) mybits += (u[i] >> kk) + 1u + kk;
                }
                if (tid < nthr && g0 % psize == 0) mybits += rice2 ? 5u : 4u;
                if (tid == 0) {
                    mybits += 8u + (uint32_t)wasted + (uint32_t)order * (uint32_t)sbps + 6u;
                    if (type == 3) mybits += 9u + (uint32_t)order * (uint32_t)sh.precision;
#pragma unroll
                    for (int i = 0; i < HISTN; i++) sh.warm[i] = s[HISTN + i];
                }
            } else if (type == 1) {
                const uint32_t smask = sbps >= 32 ? 0xFFFFFFFFu : ((1u << sbps) - 1u);
#pragma unroll
                for (int i = 0; i < FRX_SPT; i++) u[i] = (uint32_t)s[HISTN + i] & smask;
                mybits = (uint32_t)nv * (uint32_t)sbps;
                if (tid == 0) mybits += 8u + (uint32_t)wasted;
            } else {
                u[0] = (uint32_t)s[HISTN] & (sbps >= 32 ? 0xFFFFFFFFu : ((1u << sbps) - 1u));
                if (tid == 0) mybits = 8u + (uint32_t)wasted + (uint32_t)sbps;
            }
            uint32_t incl = mybits;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                uint32_t n = __shfl_up_sync(FULLMASK, incl, o);
                if (lane >= o) incl += n;
            }
            if (lane == 31) sh.scan[warp] = incl;
            __syncthreads();  // S10
            uint32_t wbase = 0, total = 0;
#pragma unroll
            for (int w2 = 0; w2 < NWARPS; w2++) {
                uint32_t v = sh.scan[w2];
                if (w2 < warp) wbase += v;
                total += v;
            }
            uint32_t start = sub_start + wbase + incl - mybits;
            // Rice-coded size must beat verbatim, else redo the subframe as VERBATIM (keeps the frame inside its buffer)
            if (type >= 2 && (uint64_t)total >= verb_bits) { force_verbatim = true; continue; }
            if (tid == 0) sh.bitpos = sub_start + total;  // next read of bitpos is behind the next S2

            // ---- K7b: pack ----
            if (mybits) {
                BitW bw;
                bw.init(bits, start);
                if (tid == 0) {
                    // subframe header: 0 | type(6) | wasted flag
                    uint32_t tcode = type == 0 ? 0u : type == 1 ? 1u : type == 2 ? (8u | (uint32_t)order) : (32u | (uint32_t)(order - 1));
                    bw.put((tcode << 1) | (wasted ? 1u : 0u), 8);
                    if (wasted) { if (wasted > 1) bw.skip((uint32_t)wasted - 1u); bw.put(1u, 1); }
                }
                const uint32_t smask = sbps >= 32 ? 0xFFFFFFFFu : ((1u << sbps) - 1u);
                if (type == 0) {
                    if (tid == 0) bw.put(u[0], sbps);
                } else if (type == 1) {
#pragma unroll
                    for (int i = 0; i < FRX_SPT; i++)
                        if (i < nv) bw.put(u[i], sbps);
                } else {
                    if (tid == 0) {
                        for (int i = 0; i < order; i++) bw.put((uint32_t)sh.warm[i] & smask, sbps);
                        if (type == 3) {
                            bw.put((uint32_t)(sh.precision - 1), 4);
                            bw.put((uint32_t)sh.shift & 31u, 5);
                            const uint32_t pmask = (1u << sh.precision) - 1u;
                            for (int j = 0; j < order; j++) bw.put((uint32_t)sh.qlp[j] & pmask, sh.precision);
                        }
                        bw.put(((rice2 ? 1u : 0u) << 4) | (uint32_t)po, 6);
                    }
                    if (g0 % psize == 0) bw.put(kk, rice2 ? 5 : 4);
                    const uint32_t kmask = kk ? ((1u << kk) - 1u) : 0u;
#pragma unroll
                    for (int i = 0; i < FRX_SPT; i++) {
                        if (i < nv && g0 + i >= order) {
                            uint32_t uq = u[i] >> kk;
                            uint32_t code = (1u << kk) | (u[i] & kmask);
                            if (uq + 1u + kk <= 32u) bw.put(code, (int)(uq + 1u + kk));
                            else { bw.skip(uq); bw.put(code, (int)(kk + 1u)); }
                        }
                    }
                }
                bw.finish();
            }
            break;
          }
        }  // channels
        __syncthreads();  // Sa: every subframe is packed

        // ---- frame footer: pad to byte, CRC-16 over the body (warps 6,7), byte offset by look-back (warp 0) ----
        const uint32_t body_bytes = (sh.bitpos + 7u) >> 3;
        const uint32_t nbytes = body_bytes + 2u;
        if (warp >= NWARPS - 2) {
            const int ct = tid - (FRX_ENC_THREADS - CRC_THREADS);   // 0..63
            const uint32_t nw = body_bytes >> 2;                   // whole words
            uint32_t per = (nw + CRC_THREADS - 1) / CRC_THREADS;
            int lg = per <= 1 ? 0 : 32 - __clz((int)per - 1);      // Lw = 2^lg words per CRC thread
            const uint32_t Lw = 1u << lg;
            const int padw = (int)(CRC_THREADS * Lw) - (int)nw;    // virtual leading zero words
            int w0 = ct * (int)Lw - padw, w1 = w0 + (int)Lw;
            if (w0 < 0) w0 = 0;
            uint32_t c = 0;
            for (int wi = w0; wi < w1; wi++) {
                uint32_t x = bits[wi] ^ (c << 16);
                c = (uint32_t)sh.crc_tab[3][x >> 24] ^ sh.crc_tab[2][(x >> 16) & 0xFFu] ^ sh.crc_tab[1][(x >> 8) & 0xFFu] ^ sh.crc_tab[0][x & 0xFFu];
            }
            if (c) c = frx_gf16_mulmod(c, (uint32_t)a.crc_xpow[lg * CRC_THREADS + (CRC_THREADS - 1 - ct)]);
            c = __reduce_xor_sync(FULLMASK, c);
            if (lane == 0) sh.crc_part[warp - (NWARPS - 2)] = c;
            if (ct == 0) sh.tailw = bits[nw];
        }
        // ---- K8: decoupled look-back for the byte offset (warp 0) ----
        if (warp == 0) {
            const uint64_t FLAG_A = 1ull << 62, FLAG_P = 2ull << 62, VMASK = (1ull << 62) - 1;
            uint64_t exclusive = 0;
            if (f == 0) {
                if (lane == 0) frx_st_release_u64(&a.desc[0], FLAG_P | (uint64_t)nbytes);
            } else {
                if (lane == 0) frx_st_release_u64(&a.desc[f], FLAG_A | (uint64_t)nbytes);
                long long pos = (long long)f - 1;
                for (;;) {
                    long long idx = pos - lane;
                    uint64_t v = idx >= 0 ? frx_ld_acquire_u64(&a.desc[idx]) : FLAG_P;
                    uint32_t flag = (uint32_t)(v >> 62);
                    uint32_t notready = __ballot_sync(FULLMASK, flag == 0u);
                    uint32_t isp = __ballot_sync(FULLMASK, flag == 2u);
                    int firstp = isp ? (__ffs((int)isp) - 1) : 32;
                    uint32_t need = firstp < 32 ? (firstp == 31 ? 0xFFFFFFFFu : ((2u << firstp) - 1u)) : 0xFFFFFFFFu;
                    if (notready & need) { __nanosleep(40); continue; }
                    uint64_t contrib = (lane <= firstp) ? (v & VMASK) : 0ull;
                    exclusive += warp_sum_u64(contrib);
                    if (firstp < 32) break;
                    pos -= 32;
                }
                if (lane == 0) frx_st_release_u64(&a.desc[f], FLAG_P | (exclusive + (uint64_t)nbytes));
            }
            if (lane == 0) {
                sh.out_offset = exclusive;
                if (a.frame_size) a.frame_size[f] = nbytes;
                if (a.frame_offset) a.frame_offset[f] = exclusive;
                if (f == a.total_frames - 1 && a.total_bytes) *a.total_bytes = exclusive + nbytes;
            }
        }
        __syncthreads();  // Sb
        // ---- copy the frame to its final place ----
        {
            const uint64_t off = sh.out_offset;
            if (off + nbytes > a.capacity) {
                if (tid == 0) atomicExch(a.status, FRX_E_OVERFLOW);
            } else {
                uint8_t *dst = a.bytes + off;
                if (tid == FRX_ENC_THREADS - 1) {
                    // finish the CRC over the <= 3 tail bytes and store it big-endian behind the body
                    uint32_t c = sh.crc_part[0] ^ sh.crc_part[1];
                    const uint32_t tw = sh.tailw;
                    for (uint32_t b = 0; b < (body_bytes & 3u); b++) {
                        uint32_t byte = (tw >> (24 - 8 * b)) & 0xFFu;
                        c = ((c << 8) ^ sh.crc_tab[0][((c >> 8) ^ byte) & 0xFFu]) & 0xFFFFu;
                    }
                    dst[body_bytes] = (uint8_t)(c >> 8);
                    dst[body_bytes + 1] = (uint8_t)c;
                }
                const uint32_t head = min(body_bytes, (uint32_t)((4u - (uint32_t)((uintptr_t)dst & 3u)) & 3u));
                if (tid < (int)head) dst[tid] = (uint8_t)((bits[tid >> 2] >> (24 - 8 * (tid & 3))) & 0xFFu);
                const uint32_t nwords = (body_bytes - head) >> 2;
                uint32_t *dw = reinterpret_cast<uint32_t *>(dst + head);
                const uint32_t sh8 = head * 8u;
                for (uint32_t j = tid; j < nwords; j += FRX_ENC_THREADS) {
                    // frame bytes [head+4j, head+4j+4): head < 4 so they straddle words j and j+1
                    uint32_t w0 = bits[j], w1 = bits[j + 1];
                    uint32_t v = head ? __funnelshift_l(w1, w0, sh8) : w0;
                    dw[j] = frx_bswap32(v);
                }
                const uint32_t tail0 = head + (nwords << 2);
                if (tid < (int)(body_bytes - tail0)) {
                    uint32_t b = tail0 + (uint32_t)tid;
                    dst[b] = (uint8_t)((bits[b >> 2] >> (24 - 8 * (b & 3))) & 0xFFu);
                }
            }
        }
        __syncthreads();  // Sz: all reads of the frame buffer done
        for (uint32_t j = tid; j <= (body_bytes >> 2); j += FRX_ENC_THREADS) bits[j] = 0u;
    }
}

// ------------------------------------------------------------------------------------------
// K0+K1 alone (frx_quantize): interleaved int32 PCM of one tile
// ------------------------------------------------------------------------------------------
template <int DT>
__global__ void frx_quantize_kernel(const void *raster, long long band_stride, long long row_stride, frx_tile t,
                                    const FrxQuant *q, int mode, int32_t *pcm) {
    typedef typename FrxElem<DT>::T T;
    const T *base = (const T *)raster;
    const FrxQuant qz = q[0];
    long long n = (long long)t.h * t.w * t.n_channels;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        int c = (int)(i % t.n_channels);
        long long px = i / t.n_channels;
        int row = (int)(px / t.w), col = (int)(px - (long long)row * t.w);
        T d = base[(long long)(t.first_band + c) * band_stride + (long long)(t.row_off + row) * row_stride + t.col_off + col];
        pcm[i] = frx_quant<DT>(d, qz, mode);
    }
}

__global__ void frx_minmax_out_kernel(const FrxQuant *q, int dtype, double *mm) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        if (dtype == FRX_F32) { mm[0] = (double)q[0].fmin; mm[1] = (double)q[0].fmax; }
        else { mm[0] = (double)q[0].imin; mm[1] = (double)q[0].imax; }
    }
}

// ------------------------------------------------------------------------------------------
// host launchers (called from frx_api.cu)
// ------------------------------------------------------------------------------------------
size_t frx_enc_smem_bytes(int frame_words) {
    return ENC_SMALL_BYTES + ENC_UNION_BYTES + sizeof(uint32_t) * (size_t)frame_words;
}

size_t frx_minmax_raw_size() { return sizeof(MinMaxRaw); }

template <int DT>
static cudaError_t launch_minmax_t(const frx_raster &r, const frx_tile *tiles_dev, const uint32_t *unit_prefix_dev, int n_tiles,
                                   uint32_t total_units, int unit_elems, void *mm_raw, FrxQuant *quant, cudaStream_t st) {
    frx_minmax_init_kernel<<<(n_tiles + 255) / 256, 256, 0, st>>>((MinMaxRaw *)mm_raw, n_tiles);
    if (total_units)
        frx_minmax_kernel<DT><<<total_units, 256, 0, st>>>(r.data, r.band_stride, r.row_stride, tiles_dev, unit_prefix_dev, n_tiles,
                                                          unit_elems, (MinMaxRaw *)mm_raw);
    frx_quant_setup_kernel<<<(n_tiles + 255) / 256, 256, 0, st>>>((const MinMaxRaw *)mm_raw, quant, n_tiles, r.dtype);
    return cudaGetLastError();
}

cudaError_t frx_launch_minmax(const frx_raster &r, const frx_tile *tiles_dev, const uint32_t *unit_prefix_dev, int n_tiles,
                              uint32_t total_units, int unit_elems, void *mm_raw, FrxQuant *quant, cudaStream_t st) {
    switch (r.dtype) {
        case FRX_U8: return launch_minmax_t<FRX_U8>(r, tiles_dev, unit_prefix_dev, n_tiles, total_units, unit_elems, mm_raw, quant, st);
        case FRX_I16: return launch_minmax_t<FRX_I16>(r, tiles_dev, unit_prefix_dev, n_tiles, total_units, unit_elems, mm_raw, quant, st);
        case FRX_U16: return launch_minmax_t<FRX_U16>(r, tiles_dev, unit_prefix_dev, n_tiles, total_units, unit_elems, mm_raw, quant, st);
        case FRX_I32: return launch_minmax_t<FRX_I32>(r, tiles_dev, unit_prefix_dev, n_tiles, total_units, unit_elems, mm_raw, quant, st);
        case FRX_U32: return launch_minmax_t<FRX_U32>(r, tiles_dev, unit_prefix_dev, n_tiles, total_units, unit_elems, mm_raw, quant, st);
        case FRX_F32: return launch_minmax_t<FRX_F32>(r, tiles_dev, unit_prefix_dev, n_tiles, total_units, unit_elems, mm_raw, quant, st);
    }
    return cudaErrorInvalidValue;
}

template <bool WIDE, int HISTN>
static cudaError_t launch_encode_t(const EncArgs &a, int sm_count, cudaStream_t st) {
    size_t smem = frx_enc_smem_bytes(a.frame_words);
    cudaError_t e = cudaFuncSetAttribute(frx_encode_kernel<WIDE, HISTN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(frx_encode_kernel<WIDE, HISTN>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if (e != cudaSuccess) return e;
    int occ = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, frx_encode_kernel<WIDE, HISTN>, FRX_ENC_THREADS, smem);
    if (e != cudaSuccess) return e;
    if (occ < 1) return cudaErrorLaunchOutOfResources;
    long long grid = (long long)sm_count * occ;   // persistent: every CTA resident, tickets hand out the frames in order
    if (grid > (long long)a.total_frames) grid = a.total_frames;
    if (grid < 1) grid = 1;
    frx_encode_kernel<WIDE, HISTN><<<(unsigned)grid, FRX_ENC_THREADS, smem, st>>>(a);
    return cudaGetLastError();
}

cudaError_t frx_launch_encode(const EncArgs &a, int dtype, int sm_count, cudaStream_t st) {
    (void)dtype;
    const bool wide = a.bps > 16;
    const int hist = a.max_lpc_order <= 4 ? 4 : (a.max_lpc_order <= 8 ? 8 : 12);
    if (wide) {
        if (hist == 4) return launch_encode_t<true, 4>(a, sm_count, st);
        if (hist == 8) return launch_encode_t<true, 8>(a, sm_count, st);
        return launch_encode_t<true, 12>(a, sm_count, st);
    }
    if (hist == 4) return launch_encode_t<false, 4>(a, sm_count, st);
    if (hist == 8) return launch_encode_t<false, 8>(a, sm_count, st);
    return launch_encode_t<false, 12>(a, sm_count, st);
}

cudaError_t frx_launch_tile_index(const uint32_t *frame_prefix, const uint32_t *frame_size, const uint64_t *frame_offset,
                                  const FrxQuant *q, int dtype, int n_tiles, uint64_t *tile_offset, uint64_t *tile_size,
                                  double *tile_minmax, cudaStream_t st) {
    frx_tile_index_kernel<<<(n_tiles + 255) / 256, 256, 0, st>>>(frame_prefix, frame_size, frame_offset, q, dtype, n_tiles, tile_offset,
                                                                tile_size, tile_minmax);
    return cudaGetLastError();
}

cudaError_t frx_launch_quantize(const frx_raster &r, const frx_tile &t, const FrxQuant *q, int mode, int32_t *pcm, double *minmax,
                                cudaStream_t st) {
    long long n = (long long)t.h * t.w * t.n_channels;
    int grid = (int)min((long long)148 * 8, (n + 255) / 256);
    if (grid < 1) grid = 1;
    switch (r.dtype) {
        case FRX_U8: frx_quantize_kernel<FRX_U8><<<grid, 256, 0, st>>>(r.data, r.band_stride, r.row_stride, t, q, mode, pcm); break;
        case FRX_I16: frx_quantize_kernel<FRX_I16><<<grid, 256, 0, st>>>(r.data, r.band_stride, r.row_stride, t, q, mode, pcm); break;
        case FRX_U16: frx_quantize_kernel<FRX_U16><<<grid, 256, 0, st>>>(r.data, r.band_stride, r.row_stride, t, q, mode, pcm); break;
        case FRX_I32: frx_quantize_kernel<FRX_I32><<<grid, 256, 0, st>>>(r.data, r.band_stride, r.row_stride, t, q, mode, pcm); break;
        case FRX_U32: frx_quantize_kernel<FRX_U32><<<grid, 256, 0, st>>>(r.data, r.band_stride, r.row_stride, t, q, mode, pcm); break;
        case FRX_F32: frx_quantize_kernel<FRX_F32><<<grid, 256, 0, st>>>(r.data, r.band_stride, r.row_stride, t, q, mode, pcm); break;
        default: return cudaErrorInvalidValue;
    }
    if (minmax) frx_minmax_out_kernel<<<1, 32, 0, st>>>(q, r.dtype, minmax);
    return cudaGetLastError();
}

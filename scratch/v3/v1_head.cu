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
    // 16-byte loads when the tile rows allow it (every thread keeps several in flight), element loads otherwise
    constexpr int VEC = 16 / (int)sizeof(T);
    const bool vec_ok = (((uintptr_t)base & 15u) == 0) && (band_stride % VEC) == 0 && (row_stride % VEC) == 0 && (t.col_off % VEC) == 0 && (t.w % VEC) == 0;
    auto take = [&](T v) {
        if (DT == FRX_F32) {
            const float f = (float)v;
            if (f != f) nan = 1;
            fmin = fminf(fmin, f); fmax = fmaxf(fmax, f);
        } else {
            const long long iv = (long long)v;
            imin = min(imin, iv); imax = max(imax, iv);
        }
    };
    auto row_ptr = [&](long long vr) {
        const int c = (int)(vr / t.h);
        const int r = (int)(vr - (long long)c * t.h);
        return base + (long long)(t.first_band + c) * band_stride + (long long)(t.row_off + r) * row_stride + t.col_off;
    };
    if (vec_ok) {
        const int n4 = t.w / VEC;
        for (long long vr = vrow0; vr < vrow1; vr += 4) {     // four rows at a time: four independent loads per thread
            const uint4 *rp[4];
#pragma unroll
            for (int k = 0; k < 4; k++) rp[k] = reinterpret_cast<const uint4 *>(row_ptr(min(vr + k, vrow1 - 1)));
            for (int x = threadIdx.x; x < n4; x += blockDim.x) {
                uint4 q[4];
#pragma unroll
                for (int k = 0; k < 4; k++) q[k] = __ldg(rp[k] + x);
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    T e[VEC];
                    memcpy(e, &q[k], 16);
#pragma unroll
                    for (int j = 0; j < VEC; j++) take(e[j]);
                }
            }
        }
    } else {
        for (long long vr = vrow0; vr < vrow1; vr++) {
            const T *row = row_ptr(vr);
            for (int x = threadIdx.x; x < t.w; x += blockDim.x) take(row[x]);
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
    __device__ __forceinline__ uint32_t bitpos() const { return idx * 32u + (uint32_t)nb; }
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


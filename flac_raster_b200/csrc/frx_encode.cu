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
// bit writer into the zero-initialised shared frame buffer (big-endian 32-bit words)
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
// block-wide helpers (256 threads)
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t warp_sum_u32(uint32_t v) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
    return v;
}
__device__ __forceinline__ uint64_t warp_sum_u64(uint64_t v) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum_f64(double v) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
    return v;
}

struct EncShared {
    // per-frame
    uint32_t frame_idx;
    uint32_t bitpos;        // running bit position in the frame buffer
    uint64_t out_offset;
    // per-subframe decisions
    int wasted;
    int type;               // 0 const 1 verbatim 2 fixed 3 lpc
    int order;
    int po;                 // partition order
    int rice2;
    int shift, precision;
    int lpc_ok;
    int lpc_order;
    int32_t qlp[FRX_MAX_LPC];
    uint32_t kparam[2][64];  // candidate 0 = fixed, 1 = lpc
    int cand_po[2];
    uint64_t cand_bits[2];
    // reductions
    uint64_t red64[NWARPS][8];
    uint32_t red32[NWARPS][8];
    double autoc[FRX_MAX_LPC + 1];
    uint64_t psum[2][128];
    uint64_t tabs[2][FRX_ENC_THREADS];  // per-thread sum|r| for the two candidates
    uint32_t scan[NWARPS];
    uint32_t crc_part[NWARPS];
    float acpart[FRX_MAX_LPC + 1][FRX_ENC_THREADS];
};

// libFLAC-style Rice cost estimate from the partition sum (oracle/flac_oracle.c rice_bits_estimate)
__device__ __forceinline__ uint64_t rice_estimate(uint32_t k, uint32_t n, uint64_t s) {
    return 4ull + (uint64_t)(1 + k) * n + (k ? (s >> (k - 1)) : (s << 1)) - (n >> 1);
}

// One warp analyses one candidate: partition sums from per-thread sums, Rice parameter per
// partition, bits per partition order, best order.  tpb = threads per block-partition at po_max.
__device__ void analyse_partitions(EncShared &sh, int cand, int nsamp, int pred_order, int po_max, uint32_t k_limit) {
    const int lane = threadIdx.x & 31;
    uint64_t *ps = sh.psum[cand];
    const uint64_t *tabs = sh.tabs[cand];
    int np = 1 << po_max;
    int tpp = (nsamp >> po_max) / FRX_SPT;  // threads per partition (>=1 unless po_max == 0)
    if (po_max == 0) {
        int nthr = (nsamp + FRX_SPT - 1) / FRX_SPT;
        uint64_t s = 0;
        for (int t = lane; t < nthr; t += 32) s += tabs[t];
        s = warp_sum_u64(s);
        if (lane == 0) ps[0] = s;
    } else {
        for (int p = lane; p < np; p += 32) {
            uint64_t s = 0;
            for (int t = 0; t < tpp; t++) s += tabs[p * tpp + t];
            ps[p] = s;
        }
    }
    __syncwarp();
    // merge down: level po stored at offset (sum of 2^j for j in (po, po_max])
    int off = 0;
    for (int po = po_max - 1; po >= 0; po--) {
        int n2 = 1 << po;
        for (int p = lane; p < n2; p += 32) ps[off + np + p] = ps[off + 2 * p] + ps[off + 2 * p + 1];
        off += np;
        np = n2;
        __syncwarp();
    }
    // cost per level
    uint64_t best = ~0ull;
    int best_po = 0;
    off = 0;
    for (int po = po_max; po >= 0; po--) {
        int n2 = 1 << po;
        uint64_t bits = 0;
        for (int p = lane; p < n2; p += 32) {
            uint32_t n = (uint32_t)(nsamp >> po) - (p == 0 ? (uint32_t)pred_order : 0u);
            uint64_t s = ps[off + p];
            uint32_t k = 0;
            while (k < k_limit - 1 && ((uint64_t)n << k) < s) k++;
            bits += rice_estimate(k, n, s);
        }
        bits = warp_sum_u64(bits) + 6;
        if (bits < best) { best = bits; best_po = po; }
        off += n2;
    }
    // store params of the best level
    off = 0;
    for (int po = po_max; po > best_po; po--) off += 1 << po;
    for (int p = lane; p < (1 << best_po); p += 32) {
        uint32_t n = (uint32_t)(nsamp >> best_po) - (p == 0 ? (uint32_t)pred_order : 0u);
        uint64_t s = ps[off + p];
        uint32_t k = 0;
        while (k < k_limit - 1 && ((uint64_t)n << k) < s) k++;
        sh.kparam[cand][p] = k;
    }
    if (lane == 0) { sh.cand_po[cand] = best_po; sh.cand_bits[cand] = best; }
    __syncwarp();
}

// LPC residual of this thread's samples; ORD taps.  s[HIST + i] is sample i of the thread.
template <int ORD, bool WIDE>
__device__ __forceinline__ void lpc_residual_thread(const int32_t *s, const int32_t *q, int shift, int32_t *r, uint32_t &bad) {
#pragma unroll
    for (int i = 0; i < FRX_SPT; i++) {
        if (WIDE) {
            long long acc = 0;
#pragma unroll
            for (int j = 0; j < ORD; j++) acc += (long long)q[j] * (long long)s[HIST + i - 1 - j];
            long long rr = (long long)s[HIST + i] - (acc >> shift);
            if (rr <= (long long)INT32_MIN || rr > (long long)INT32_MAX) bad |= 1u << i;
            r[i] = (int32_t)rr;
        } else {
            int32_t acc = 0;
#pragma unroll
            for (int j = 0; j < ORD; j++) acc += q[j] * s[HIST + i - 1 - j];
            r[i] = s[HIST + i] - (acc >> shift);
        }
    }
}

template <bool WIDE>
__device__ __forceinline__ void lpc_residual_dispatch(int order, const int32_t *s, const int32_t *q, int shift, int32_t *r, uint32_t &bad) {
    switch (order) {
        case 1: lpc_residual_thread<1, WIDE>(s, q, shift, r, bad); break;
        case 2: lpc_residual_thread<2, WIDE>(s, q, shift, r, bad); break;
        case 3: lpc_residual_thread<3, WIDE>(s, q, shift, r, bad); break;
        case 4: lpc_residual_thread<4, WIDE>(s, q, shift, r, bad); break;
        case 5: lpc_residual_thread<5, WIDE>(s, q, shift, r, bad); break;
        case 6: lpc_residual_thread<6, WIDE>(s, q, shift, r, bad); break;
        case 7: lpc_residual_thread<7, WIDE>(s, q, shift, r, bad); break;
        case 8: lpc_residual_thread<8, WIDE>(s, q, shift, r, bad); break;
        case 9: lpc_residual_thread<9, WIDE>(s, q, shift, r, bad); break;
        case 10: lpc_residual_thread<10, WIDE>(s, q, shift, r, bad); break;
        case 11: lpc_residual_thread<11, WIDE>(s, q, shift, r, bad); break;
        default: lpc_residual_thread<12, WIDE>(s, q, shift, r, bad); break;
    }
}

template <bool WIDE>
__device__ __forceinline__ void fixed_residual_thread(int order, const int32_t *s, int32_t *r, uint32_t &bad) {
#pragma unroll
    for (int i = 0; i < FRX_SPT; i++) {
        const int32_t *p = s + HIST + i;
        if (WIDE) {
            long long a = p[0], b = p[-1], c = p[-2], d = p[-3], e = p[-4], rr;
            switch (order) {
                case 0: rr = a; break;
                case 1: rr = a - b; break;
                case 2: rr = a - 2 * b + c; break;
                case 3: rr = a - 3 * b + 3 * c - d; break;
                default: rr = a - 4 * b + 6 * c - 4 * d + e; break;
            }
            if (rr <= (long long)INT32_MIN || rr > (long long)INT32_MAX) bad |= 1u << i;
            r[i] = (int32_t)rr;
        } else {
            int32_t a = p[0], b = p[-1], c = p[-2], d = p[-3], e = p[-4], rr;
            switch (order) {
                case 0: rr = a; break;
                case 1: rr = a - b; break;
                case 2: rr = a - 2 * b + c; break;
                case 3: rr = a - 3 * b + 3 * c - d; break;
                default: rr = a - 4 * b + 6 * c - 4 * d + e; break;
            }
            r[i] = rr;
        }
    }
}

__device__ __forceinline__ uint32_t zigzag(int32_t r) { return ((uint32_t)r << 1) ^ (uint32_t)(r >> 31); }

// ------------------------------------------------------------------------------------------
// The frame encoder
// ------------------------------------------------------------------------------------------
template <int DT, bool WIDE>
__global__ void __launch_bounds__(FRX_ENC_THREADS) frx_encode_kernel(const EncArgs a) {
    typedef typename FrxElem<DT>::T T;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    EncShared &sh = *reinterpret_cast<EncShared *>(smem_raw);
    int32_t *pcm = reinterpret_cast<int32_t *>(smem_raw + ((sizeof(EncShared) + 15) & ~(size_t)15));
    uint32_t *bits = reinterpret_cast<uint32_t *>(pcm + (FRX_MAX_BLOCK + (FRX_MAX_BLOCK >> 4) + 32));

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const T *base = (const T *)a.raster;
    const uint32_t k_limit = a.bps > 16 ? 31u : 15u;  // RICE2 only for bps > 16 (libFLAC)

    for (;;) {
        __syncthreads();
        if (tid == 0) sh.frame_idx = atomicAdd(a.ticket, 1u);
        // zero the frame buffer
        for (int i = tid; i < a.frame_words; i += FRX_ENC_THREADS) bits[i] = 0u;
        __syncthreads();
        const uint32_t f = sh.frame_idx;
        if (f >= a.total_frames) break;

        // ---- locate tile / frame ----
        int lo = 0, hi = a.n_tiles;
        while (hi - lo > 1) {
            int mid = (lo + hi) >> 1;
            if (a.frame_prefix[mid] <= f) lo = mid; else hi = mid;
        }
        const frx_tile t = a.tiles[lo];
        const FrxQuant qz = a.quant[lo];
        const uint32_t fj = f - a.frame_prefix[lo];
        const long long tile_n = (long long)t.h * t.w;
        const long long s0 = (long long)fj * a.blocksize;
        const int nsamp = (int)min((long long)a.blocksize, tile_n - s0);
        const int nthr = (nsamp + FRX_SPT - 1) / FRX_SPT;

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
            __syncthreads();
            // ---- K1: gather + quantise (coalesced: thread t takes samples t, t+256, ...) ----
            {
                const T *band = base + (long long)(t.first_band + ch) * a.band_stride + (long long)t.row_off * a.row_stride + t.col_off;
                long long sidx = s0 + tid;
                int row = (int)(sidx / t.w);
                int col = (int)(sidx - (long long)row * t.w);
                for (int i = tid; i < FRX_MAX_BLOCK; i += FRX_ENC_THREADS) {
                    int32_t v = 0;
                    if (i < nsamp) {
                        T d = band[(long long)row * a.row_stride + col];
                        v = frx_quant<DT>(d, qz, a.pcm_mode);
                    }
                    pcm[PCM_PAD(i)] = v;
                    col += FRX_ENC_THREADS;
                    while (col >= t.w) { col -= t.w; row++; }
                }
            }
            __syncthreads();
            // ---- registers: HIST history + 16 samples ----
            int32_t s[HIST + FRX_SPT];
            const int g0 = tid * FRX_SPT;  // first sample index of this thread
#pragma unroll
            for (int i = 0; i < HIST + FRX_SPT; i++) {
                int g = g0 - HIST + i;
                s[i] = (g >= 0 && g < FRX_MAX_BLOCK) ? pcm[PCM_PAD(g)] : 0;
            }
            // ---- wasted bits ----
            uint32_t orv = 0;
#pragma unroll
            for (int i = 0; i < FRX_SPT; i++) orv |= (uint32_t)s[HIST + i];
            for (int o = 16; o > 0; o >>= 1) orv |= __shfl_xor_sync(0xFFFFFFFFu, orv, o);
            if (lane == 0) sh.red32[warp][0] = orv;
            __syncthreads();
            int wasted = 0;
            {
                uint32_t all = 0;
#pragma unroll
                for (int w2 = 0; w2 < NWARPS; w2++) all |= sh.red32[w2][0];
                wasted = all ? (__ffs((int)all) - 1) : 0;
                if (wasted) {
#pragma unroll
                    for (int i = 0; i < HIST + FRX_SPT; i++) s[i] >>= wasted;
                }
            }
            const int sbps = a.bps - wasted;

            // ---- K2: fixed-order trial (sum |e_k| over samples g >= k) ----
            uint64_t fsum[5];
            uint32_t fbad = 0;
            {
                uint64_t acc0 = 0, acc1 = 0, acc2 = 0, acc3 = 0, acc4 = 0;
#pragma unroll
                for (int i = 0; i < FRX_SPT; i++) {
                    int g = g0 + i;
                    if (g < nsamp) {
                        const int32_t *p = s + HIST + i;
                        if (WIDE) {
                            long long A = p[0], B = p[-1], C = p[-2], D = p[-3], E = p[-4];
                            long long e0 = A, e1 = A - B, e2 = A - 2 * B + C, e3 = A - 3 * B + 3 * C - D, e4 = A - 4 * B + 6 * C - 4 * D + E;
                            unsigned long long a0 = (unsigned long long)llabs(e0), a1 = (unsigned long long)llabs(e1), a2 = (unsigned long long)llabs(e2),
                                               a3 = (unsigned long long)llabs(e3), a4 = (unsigned long long)llabs(e4);
                            acc0 += a0;
                            if (a0 > 0x7FFFFFFFull) fbad |= 1u;
                            if (g >= 1) { acc1 += a1; if (a1 > 0x7FFFFFFFull) fbad |= 2u; }
                            if (g >= 2) { acc2 += a2; if (a2 > 0x7FFFFFFFull) fbad |= 4u; }
                            if (g >= 3) { acc3 += a3; if (a3 > 0x7FFFFFFFull) fbad |= 8u; }
                            if (g >= 4) { acc4 += a4; if (a4 > 0x7FFFFFFFull) fbad |= 16u; }
                        } else {
                            int32_t A = p[0], B = p[-1], C = p[-2], D = p[-3], E = p[-4];
                            int32_t e0 = A, e1 = A - B, e2 = e1 - (B - C), e3 = e2 - (B - 2 * C + D), e4 = e3 - (B - 3 * C + 3 * D - E);
                            acc0 += (uint32_t)abs(e0);
                            if (g >= 1) acc1 += (uint32_t)abs(e1);
                            if (g >= 2) acc2 += (uint32_t)abs(e2);
                            if (g >= 3) acc3 += (uint32_t)abs(e3);
                            if (g >= 4) acc4 += (uint32_t)abs(e4);
                        }
                    }
                }
                acc0 = warp_sum_u64(acc0); acc1 = warp_sum_u64(acc1); acc2 = warp_sum_u64(acc2);
                acc3 = warp_sum_u64(acc3); acc4 = warp_sum_u64(acc4);
                for (int o = 16; o > 0; o >>= 1) fbad |= __shfl_xor_sync(0xFFFFFFFFu, fbad, o);
                if (lane == 0) {
                    sh.red64[warp][0] = acc0; sh.red64[warp][1] = acc1; sh.red64[warp][2] = acc2;
                    sh.red64[warp][3] = acc3; sh.red64[warp][4] = acc4; sh.red32[warp][1] = fbad;
                }
            }
            // ---- K3: windowed autocorrelation partials ----
            const int max_lpc = min(a.max_lpc_order, nsamp - 1);
            if (max_lpc > 0) {
                float xw[HIST + FRX_SPT];
                const bool table = (nsamp == a.blocksize);
#pragma unroll
                for (int i = 0; i < HIST + FRX_SPT; i++) {
                    int g = g0 - HIST + i;
                    float wv = 0.f;
                    if (g >= 0 && g < nsamp) {
                        if (table) wv = a.window[g];
                        else {
                            // tukey(0.5) of length nsamp
                            int Np = (int)(0.25f * (float)nsamp) - 1;
                            wv = 1.0f;
                            if (Np > 0) {
                                if (g <= Np) wv = 0.5f - 0.5f * cospif((float)g / (float)Np);
                                else if (g >= nsamp - Np - 1) wv = 0.5f - 0.5f * cospif((float)(g - (nsamp - Np - 1) + Np) / (float)Np);
                            }
                        }
                    }
                    xw[i] = (float)s[i] * wv;
                }
#pragma unroll
                for (int lag = 0; lag <= FRX_MAX_LPC; lag++) {
                    float acc = 0.f;
                    if (lag <= max_lpc) {
#pragma unroll
                        for (int i = 0; i < FRX_SPT; i++) acc = fmaf(xw[HIST + i], xw[HIST + i - lag], acc);
                    }
                    sh.acpart[lag][tid] = acc;
                }
            }
            __syncthreads();
            // warp w reduces lags w, w+8 (double accumulation across threads)
            for (int lag = warp; lag <= max_lpc; lag += NWARPS) {
                double d = 0.0;
#pragma unroll
                for (int j = 0; j < FRX_ENC_THREADS / 32; j++) d += (double)sh.acpart[lag][lane + 32 * j];
                d = warp_sum_f64(d);
                if (lane == 0) sh.autoc[lag] = d;
            }
            // fixed totals
            for (int o5 = 0; o5 < 5; o5++) {
                uint64_t tot = 0;
#pragma unroll
                for (int w2 = 0; w2 < NWARPS; w2++) tot += sh.red64[w2][o5];
                fsum[o5] = tot;
            }
            {
                uint32_t fb = 0;
#pragma unroll
                for (int w2 = 0; w2 < NWARPS; w2++) fb |= sh.red32[w2][1];
                fbad = fb;
            }
            __syncthreads();

            // ---- constant? (sum |e1| over g>=1 is zero) ----
            const bool is_const = (nsamp >= 1) && (fsum[1] == 0) && !(fbad & 2u);
            int fixed_order = -1;
            if (!is_const && nsamp > 4) {
                uint64_t bestv = ~0ull;
                for (int o5 = 0; o5 < 5; o5++) {
                    // sums for order k skip k samples; compare per-sample averages fairly by adding nothing (negligible)
                    if (!((fbad >> o5) & 1u) && fsum[o5] < bestv) { bestv = fsum[o5]; fixed_order = o5; }
                }
            }

            // ---- K4: Levinson-Durbin, order pick, coefficient quantisation (thread 0) ----
            if (tid == 0) {
                sh.lpc_ok = 0;
                if (!is_const && max_lpc > 0 && nsamp > 4 && sh.autoc[0] != 0.0) {
                    double lpc[FRX_MAX_LPC], err = sh.autoc[0];
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
                    // expected bits per order
                    double best_bits = 1e300;
                    int best = 1;
                    const double scale = 0.5 / (double)nsamp;
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
                                for (int j = 0; j < order; j++) {
                                    error += coefs[order - 1][j] * (double)(1 << shift);
                                    long long qq = llrint(error);
                                    if (qq > qmax) qq = qmax; else if (qq < qmin) qq = qmin;
                                    error -= (double)qq;
                                    sh.qlp[j] = (int32_t)qq;
                                }
                                for (int j = order; j < FRX_MAX_LPC; j++) sh.qlp[j] = 0;
                                sh.lpc_order = order; sh.shift = shift; sh.precision = precision; sh.lpc_ok = 1;
                            }
                        }
                    }
                }
            }
            __syncthreads();

            // ---- K5: residuals of both candidates, per-thread |r| sums ----
            int32_t r_lpc[FRX_SPT];
            const bool lpc_try = sh.lpc_ok != 0;
            const int lpc_order = sh.lpc_order;
            {
                uint32_t badm = 0;
                bool bad = false;
                uint64_t ab = 0;
                if (lpc_try) {
                    int32_t q[FRX_MAX_LPC];
#pragma unroll
                    for (int j = 0; j < FRX_MAX_LPC; j++) q[j] = sh.qlp[j];
                    lpc_residual_dispatch<WIDE>(lpc_order, s, q, sh.shift, r_lpc, badm);
#pragma unroll
                    for (int i = 0; i < FRX_SPT; i++) {
                        int g = g0 + i;
                        if (g < lpc_order || g >= nsamp) r_lpc[i] = 0;
                        else if ((badm >> i) & 1u) bad = true;
                        ab += (uint32_t)abs(r_lpc[i]);
                    }
                }
                sh.tabs[1][tid] = ab;
                unsigned anybad = __ballot_sync(0xFFFFFFFFu, bad);
                if (lane == 0) sh.red32[warp][2] = anybad ? 1u : 0u;
                // fixed candidate sums
                uint64_t fb = 0;
                if (fixed_order >= 0) {
                    int32_t rf[FRX_SPT];
                    uint32_t badf = 0;
                    fixed_residual_thread<WIDE>(fixed_order, s, rf, badf);
#pragma unroll
                    for (int i = 0; i < FRX_SPT; i++) {
                        int g = g0 + i;
                        if (g >= fixed_order && g < nsamp) fb += (uint32_t)abs(rf[i]);
                    }
                }
                sh.tabs[0][tid] = fb;
            }
            __syncthreads();
            // ---- K6: Rice partition analysis, warp 0 = fixed, warp 1 = lpc ----
            int po_max = 0;
            {
                int tz = __ffs(nsamp) - 1;          // trailing zeros of the block size
                po_max = min(a.max_po, max(0, tz - 4));  // partitions stay multiples of 16 samples
                while (po_max > 0 && (nsamp >> po_max) <= FRX_MAX_LPC) po_max--;
            }
            bool lpc_valid = lpc_try;
            {
                uint32_t b = 0;
#pragma unroll
                for (int w2 = 0; w2 < NWARPS; w2++) b |= sh.red32[w2][2];
                if (b) lpc_valid = false;
            }
            if (warp == 0 && fixed_order >= 0) analyse_partitions(sh, 0, nsamp, fixed_order, po_max, k_limit);
            if (warp == 1 && lpc_valid) analyse_partitions(sh, 1, nsamp, lpc_order, po_max, k_limit);
            __syncthreads();

            // ---- choose ----
            const uint64_t verb_bits = 8ull + (uint64_t)wasted + (uint64_t)nsamp * (uint64_t)sbps;
            int type = 1, order = 0, cand = 0;
            uint64_t best_bits = verb_bits;
            if (is_const) { type = 0; best_bits = 8ull + wasted + sbps; }
            else {
                if (fixed_order >= 0) {
                    uint64_t b = 8ull + wasted + (uint64_t)fixed_order * sbps + sh.cand_bits[0];
                    if (b < best_bits) { best_bits = b; type = 2; order = fixed_order; cand = 0; }
                }
                if (lpc_valid) {
                    uint64_t b = 8ull + wasted + 9ull + (uint64_t)lpc_order * (uint64_t)(sh.precision + sbps) + sh.cand_bits[1];
                    if (b < best_bits) { best_bits = b; type = 3; order = lpc_order; cand = 1; }
                }
            }

            // ---- K7: exact bit count per thread, block scan, pack ----
            const uint32_t sub_start = sh.bitpos;
            int32_t r[FRX_SPT];
            int po = 0;
            uint32_t kk = 0;
            int psize = nsamp;
            bool rice2 = false;
            if (type >= 2) {
                po = sh.cand_po[cand];
                psize = nsamp >> po;
                kk = sh.kparam[cand][min(g0 / psize, (1 << po) - 1)];
                // RICE2 iff any parameter needs 5 bits
                bool big = false;
                for (int p = 0; p < (1 << po); p++) big |= (sh.kparam[cand][p] >= 15u);
                rice2 = big;
                if (type == 3) {
#pragma unroll
                    for (int i = 0; i < FRX_SPT; i++) r[i] = r_lpc[i];
                } else {
                    uint32_t badf = 0;
                    fixed_residual_thread<WIDE>(order, s, r, badf);
                }
            }
            uint32_t mybits = 0;
            if (type >= 2) {
                if (tid < nthr) {
#pragma unroll
                    for (int i = 0; i < FRX_SPT; i++) {
                        int g = g0 + i;
                        if (g >= order && g < nsamp) mybits += (zigzag(r[i]) >> kk) + 1u + kk;
                    }
                    if (g0 % psize == 0) mybits += rice2 ? 5u : 4u;
                    if (tid == 0) {
                        mybits += 8u + (wasted ? (uint32_t)wasted : 0u) + (uint32_t)order * (uint32_t)sbps + 6u;
                        if (type == 3) mybits += 9u + (uint32_t)order * (uint32_t)sh.precision;
                    }
                }
            } else if (type == 1) {
                if (tid < nthr) mybits = (uint32_t)min(FRX_SPT, nsamp - g0) * (uint32_t)sbps;
                if (tid == 0) mybits += 8u + (uint32_t)wasted;
            } else {
                if (tid == 0) mybits = 8u + (uint32_t)wasted + (uint32_t)sbps;
            }
            // block exclusive scan of mybits
            uint32_t incl = mybits;
            for (int o = 1; o < 32; o <<= 1) {
                uint32_t n = __shfl_up_sync(0xFFFFFFFFu, incl, o);
                if (lane >= o) incl += n;
            }
            if (lane == 31) sh.scan[warp] = incl;
            __syncthreads();
            uint32_t wbase = 0, total = 0;
#pragma unroll
            for (int w2 = 0; w2 < NWARPS; w2++) {
                uint32_t v = sh.scan[w2];
                if (w2 < warp) wbase += v;
                total += v;
            }
            uint32_t start = sub_start + wbase + incl - mybits;
            // Rice-coded size must beat verbatim, else fall back (keeps the frame inside its buffer)
            if (type >= 2 && (uint64_t)total >= verb_bits) {
                type = 1;
                mybits = (tid < nthr) ? (uint32_t)min(FRX_SPT, nsamp - g0) * (uint32_t)sbps : 0u;
                start = sub_start + 8u + (uint32_t)wasted + (uint32_t)min(g0, nsamp) * (uint32_t)sbps;
                if (tid == 0) { mybits += 8u + (uint32_t)wasted; start = sub_start; }
                total = (uint32_t)verb_bits;
            }
            __syncthreads();  // all reads of sh.scan / sh.bitpos done
            if (tid == 0) sh.bitpos = sub_start + total;

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
                    if (tid == 0) bw.put((uint32_t)s[HIST] & smask, sbps);
                } else if (type == 1) {
#pragma unroll
                    for (int i = 0; i < FRX_SPT; i++)
                        if (g0 + i < nsamp) bw.put((uint32_t)s[HIST + i] & smask, sbps);
                } else {
                    if (tid == 0) {
                        for (int i = 0; i < order; i++) bw.put((uint32_t)s[HIST + i] & smask, sbps);
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
                        int g = g0 + i;
                        if (g >= order && g < nsamp) {
                            uint32_t u = zigzag(r[i]);
                            uint32_t q = u >> kk;
                            uint32_t code = (1u << kk) | (u & kmask);
                            if (q + 1u + kk <= 32u) bw.put(code, (int)(q + 1u + kk));
                            else { bw.skip(q); bw.put(code, (int)(kk + 1u)); }
                        }
                    }
                }
                bw.finish();
            }
        }  // channels
        __syncthreads();

        // ---- frame footer: pad to byte, CRC-16 ----
        const uint32_t body_bits = (sh.bitpos + 7u) & ~7u;
        const uint32_t body_bytes = body_bits >> 3;
        {
            // left-padded chunking: virtual length 256*L, thread t owns virtual bytes [tL, tL+L)
            const int L = a.crc_chunk;
            const int padv = FRX_ENC_THREADS * L - (int)body_bytes;
            uint32_t c = 0;
            int v0 = tid * L, v1 = v0 + L;
            int b0 = max(0, v0 - padv), b1 = v1 - padv;
            for (int b = b0; b < b1; b++) {
                uint32_t byte = (bits[b >> 2] >> (24 - 8 * (b & 3))) & 0xFFu;
                c ^= byte << 8;
#pragma unroll
                for (int i = 0; i < 8; i++) c = (c & 0x8000u) ? ((c << 1) ^ 0x8005u) & 0xFFFFu : (c << 1) & 0xFFFFu;
            }
            // shift by (255 - t) chunks
            if (c) c = frx_gf16_mulmod(c, (uint32_t)a.crc_xpow[FRX_ENC_THREADS - 1 - tid]);
            for (int o = 16; o > 0; o >>= 1) c ^= __shfl_xor_sync(0xFFFFFFFFu, c, o);
            if (lane == 0) sh.crc_part[warp] = c;
        }
        __syncthreads();
        const uint32_t nbytes = body_bytes + 2u;
        if (tid == 0) {
            uint32_t c = 0;
            for (int w2 = 0; w2 < NWARPS; w2++) c ^= sh.crc_part[w2];
            // append big-endian CRC-16 at byte body_bytes
            for (int k2 = 0; k2 < 2; k2++) {
                uint32_t byte = k2 == 0 ? (c >> 8) : (c & 0xFFu);
                uint32_t b = body_bytes + (uint32_t)k2;
                bits[b >> 2] |= byte << (24 - 8 * (b & 3));
            }
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
                    uint32_t notready = __ballot_sync(0xFFFFFFFFu, flag == 0u);
                    uint32_t isp = __ballot_sync(0xFFFFFFFFu, flag == 2u);
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
        __syncthreads();
        // ---- copy the frame to its final place ----
        {
            const uint64_t off = sh.out_offset;
            if (off + nbytes > a.capacity) {
                if (tid == 0) atomicExch(a.status, FRX_E_OVERFLOW);
            } else {
                uint8_t *dst = a.bytes + off;
                const uint32_t head = min(nbytes, (uint32_t)((4u - (uint32_t)((uintptr_t)dst & 3u)) & 3u));
                // head bytes
                if (tid < (int)head) dst[tid] = (uint8_t)((bits[tid >> 2] >> (24 - 8 * (tid & 3))) & 0xFFu);
                const uint32_t nwords = (nbytes - head) >> 2;
                uint32_t *dw = reinterpret_cast<uint32_t *>(dst + head);
                const uint32_t sh8 = head * 8u;
                for (uint32_t j = tid; j < nwords; j += FRX_ENC_THREADS) {
                    // frame bytes [head+4j, head+4j+4): head < 4 so they straddle words j and j+1
                    uint32_t w0 = bits[j], w1 = bits[j + 1];
                    uint32_t v = head ? __funnelshift_l(w1, w0, sh8) : w0;
                    dw[j] = frx_bswap32(v);
                }
                const uint32_t tail0 = head + (nwords << 2);
                if (tid < (int)(nbytes - tail0)) {
                    uint32_t b = tail0 + (uint32_t)tid;
                    dst[b] = (uint8_t)((bits[b >> 2] >> (24 - 8 * (b & 3))) & 0xFFu);
                }
            }
        }
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
    return ((sizeof(EncShared) + 15) & ~(size_t)15) + sizeof(int32_t) * (FRX_MAX_BLOCK + (FRX_MAX_BLOCK >> 4) + 32) +
           sizeof(uint32_t) * (size_t)frame_words;
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

template <int DT, bool WIDE>
static cudaError_t launch_encode_t(const EncArgs &a, int sm_count, cudaStream_t st) {
    size_t smem = frx_enc_smem_bytes(a.frame_words);
    cudaError_t e = cudaFuncSetAttribute(frx_encode_kernel<DT, WIDE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int occ = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, frx_encode_kernel<DT, WIDE>, FRX_ENC_THREADS, smem);
    if (e != cudaSuccess) return e;
    if (occ < 1) return cudaErrorLaunchOutOfResources;
    long long grid = (long long)sm_count * occ;
    if (grid > (long long)a.total_frames) grid = a.total_frames;
    if (grid < 1) grid = 1;
    frx_encode_kernel<DT, WIDE><<<(unsigned)grid, FRX_ENC_THREADS, smem, st>>>(a);
    return cudaGetLastError();
}

cudaError_t frx_launch_encode(const EncArgs &a, int dtype, int sm_count, cudaStream_t st) {
    const bool wide = a.bps > 16;
    switch (dtype) {
        case FRX_U8: return wide ? launch_encode_t<FRX_U8, true>(a, sm_count, st) : launch_encode_t<FRX_U8, false>(a, sm_count, st);
        case FRX_I16: return wide ? launch_encode_t<FRX_I16, true>(a, sm_count, st) : launch_encode_t<FRX_I16, false>(a, sm_count, st);
        case FRX_U16: return wide ? launch_encode_t<FRX_U16, true>(a, sm_count, st) : launch_encode_t<FRX_U16, false>(a, sm_count, st);
        case FRX_I32: return launch_encode_t<FRX_I32, true>(a, sm_count, st);
        case FRX_U32: return launch_encode_t<FRX_U32, true>(a, sm_count, st);
        case FRX_F32: return launch_encode_t<FRX_F32, true>(a, sm_count, st);
    }
    return cudaErrorInvalidValue;
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

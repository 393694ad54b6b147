// frx_encode.cu -- raster -> FLAC frames on sm_100a.
//
// Kernels (SURVEY.md section 8a ids):
//   K0  frx_minmax_kernel        per-stream min/max                      (converter.py:67-68, 152-153)
//   K1..K8 frx_encode_kernel     one CTA per 4096-sample frame, persistent grid, ticket-ordered; 8 compute warps plus one
//        control warp that takes tickets, stages the next raw samples with TMA bulk copies (mbarrier pipeline), builds the
//        frame header and runs the look-back while the compute warps search, pack and checksum:
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
#define NWARPS (FRX_ENC_THREADS / 32)   // compute warps


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
    o.fmin = 0.f; o.fmax = 0.f; o.range = 1.0; o.degenerate = 0; o.flags = 0;
    if (dtype == FRX_F32) {
        if (mm[i].has_nan) { o.fmin = CUDART_NAN_F; o.fmax = CUDART_NAN_F; }
        else {
            o.fmin = frx_float_unordered(mm[i].fmin_o); o.fmax = frx_float_unordered(mm[i].fmax_o);
            const double m = fmax(fabs((double)o.fmin), fabs((double)o.fmax));
            if (m * 8388607.0 < 2147483000.0) o.flags |= FRX_QF_F32_SAFE;
        }
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

#ifdef FRX_DBG
__device__ unsigned long long g_frx_dbg[16];
#define DBG_ADD(i, v) atomicAdd(&g_frx_dbg[i], (unsigned long long)(v))
#define DBG_CLK() clock64()
#define DBG_PH(i) do { if (tid == 32 * DBG_PH_WARP) { const long long n_ = clock64(); atomicAdd(&g_frx_dbg[i], (unsigned long long)(n_ - dbg_ph)); dbg_ph = n_; } } while (0)
#define DBG_PH_DECL long long dbg_ph = clock64()
#ifndef DBG_PH_WARP
#define DBG_PH_WARP 2
#endif
__device__ unsigned long long *g_frx_trace;   // 8 timestamps per frame
__device__ __forceinline__ unsigned long long dbg_now() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define DBG_TRACE(f, k) do { if (g_frx_trace) g_frx_trace[(size_t)(f) * 8 + (k)] = dbg_now(); } while (0)
#define DBG_TRACEV(f, k, v) do { if (g_frx_trace) g_frx_trace[(size_t)(f) * 8 + (k)] = (v); } while (0)
#else
#define DBG_ADD(i, v) do {} while (0)
#define DBG_CLK() 0ll
#define DBG_PH(i) do {} while (0)
#define DBG_PH_DECL
#define DBG_TRACE(f, k) do {} while (0)
#define DBG_TRACEV(f, k, v) do {} while (0)
#endif
// ------------------------------------------------------------------------------------------
// shared state of one encoder CTA
// ------------------------------------------------------------------------------------------
#define FD_SLOTS 4
#define STOP_FRAME 0xFFFFFFFFu
#define TABS_BYTES (6 * FRX_ENC_THREADS * 8)     // |e| sums per thread: fixed orders 0..4 + the LPC candidate
// PCM of the current subframe in shared memory, padded so that a thread's 16-byte chunks at a 64-byte stride fall into
// distinct bank groups: sample i lives at word PCM_IDX(i); 32 zero samples sit in front (history of the first thread).
#define PCM_IDX(i) (((i) + 32) + 4 * (((i) + 32) >> 5))
#define PCM_WORDS 4672
#define PCM_BYTES (PCM_WORDS * 4)

struct FrameDesc {                               // written by the control warp, read by the compute warps
    uint32_t f, fj, s0;
    int32_t nsamp;
    frx_tile t;
    uint32_t hdr[4];
    uint32_t hdr_bits;
    int32_t tma;                                 // 1: raw samples staged by TMA bulk copies; 0: compute warps gather them
    int32_t tile_idx;
    int16_t po_max, max_lpc;                     // largest Rice partition order / LPC order this frame can use
    FrxQuant q;
};

struct EncSmall {
    uint64_t full[1], empty[1];                  // raw stage hand-over (mbarriers): one stage, refilled as soon as K1 has read it
    uint64_t sized[FD_SLOTS], placed[FD_SLOTS];  // frame size known -> look-back -> byte offset known
    FrameDesc fd[FD_SLOTS];
    uint32_t frame_bytes[FD_SLOTS];
    uint64_t out_offset[FD_SLOTS];
    // per-subframe decisions
    int wasted, magbits, fixed_order, is_const, lean;
    int lpc_ok, lpc_order, shift, precision, lpc_narrow;
    int32_t qlp[FRX_MAX_LPC];
    int32_t warm[FRX_MAX_LPC];
    uint32_t kparam[2][128];                     // candidate 0 = fixed, 1 = lpc; level L's parameters start at (1 << L) - 1
    uint64_t lvl_bits[2][8];                     // Rice search: estimated bits of partition order L (~0: not evaluated)
    uint32_t lvl_big[2][8];                      // ... and whether it needs 5-bit parameters (RICE2)
    // reductions
    uint64_t wfix[NWARPS][5];
    uint32_t wor[NWARPS][2];
    uint32_t wbad[NWARPS][2];
    float wac[NWARPS][FRX_MAX_LPC + 1];
    float autoc[FRX_MAX_LPC + 1];
    float lpc_rows[FRX_MAX_LPC][FRX_MAX_LPC];    // predictor coefficients of every order (Levinson-Durbin by-product)
    uint32_t scan[NWARPS], scan2[NWARPS];
    uint32_t crc_part[2][NWARPS];                // per frame buffer
    uint16_t crc_tab[4][256];
};
#define RES_BYTES (FRX_MAX_BLOCK * 4)             // the candidate's residuals, sample-major: res[i * 256 + tid]
#define RES_IDX(i, tid) ((i) * FRX_ENC_THREADS + (tid))
#define ENC_SMALL_BYTES ((sizeof(EncSmall) + 127) & ~(size_t)127)

// ------------------------------------------------------------------------------------------
// mbarrier / TMA (bulk async copy) / named barrier primitives
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *b, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *b) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *b, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(b)), "r"(parity) : "memory");
    return ok != 0;
}
// non-blocking probe: try_wait may park the thread for a system-defined time, which a loop polling two barriers cannot afford
__device__ __forceinline__ bool mbar_test(uint64_t *b, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(b)), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t *b, uint32_t parity) {
    while (!mbar_try_wait(b, parity)) {}
}
// TMA 1-D bulk copy global -> shared, completion counted in bytes on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void compute_sync() { asm volatile("bar.sync 1, %0;" ::"n"(FRX_ENC_THREADS) : "memory"); }

// One warp analyses one candidate: partition sums from the per-thread |r| sums, Rice parameter per partition,
// estimated bits per partition order, best order.  Lane p owns finest-level partition p (2p, 2p+1 when po_max == 6);
// coarser levels are built with shuffles, so partition q of level L lives in lane q << (F - L).
__device__ __noinline__ void analyse_partitions(EncSmall &sh, const uint64_t *tabs, int cand, int nsamp, int pred_order, int po_max, uint32_t kmax, int rsh) {
    const int lane = threadIdx.x & 31;
    const int F = po_max > 5 ? 5 : po_max;               // finest level held one-per-lane
    const int nthr = (nsamp + FRX_SPT - 1) / FRX_SPT;
    uint64_t sum = 0, sumB = 0;                          // sumB: second half when po_max == 6
    uint64_t best = ~0ull;
    int best_po = 0;
    uint32_t kbest0 = 0, kbest1 = 0;
    if (po_max == 6) {
        const int tpp = (nsamp >> 6) / FRX_SPT;          // threads per level-6 partition
        for (int t = 0; t < tpp; t++) { sum += tabs[(2 * lane) * tpp + t] >> rsh; sumB += tabs[(2 * lane + 1) * tpp + t] >> rsh; }
        uint32_t n0 = (uint32_t)(nsamp >> 6) - (lane == 0 ? (uint32_t)pred_order : 0u), n1 = (uint32_t)(nsamp >> 6);
        uint32_t k0 = rice_param(n0, sum, kmax), k1 = rice_param(n1, sumB, kmax);
        uint64_t b = rice_estimate(k0, n0, sum) + rice_estimate(k1, n1, sumB);
        b = warp_sum_u64(b) + 6;
        best = b; best_po = 6; kbest0 = k0; kbest1 = k1;
        sum += sumB;
    } else if (F == 0) {
        for (int t = lane; t < nthr; t += 32) sum += tabs[t] >> rsh;
        sum = warp_sum_u64(sum);
    } else {
        const int tpp = (nsamp >> F) / FRX_SPT;          // threads per finest partition (>= 1)
        if (lane < (1 << F))
            for (int t = 0; t < tpp; t++) sum += tabs[lane * tpp + t] >> rsh;
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
    // publish in the per-level format pick_level() reads: only the winning level carries a finite cost
    uint32_t *kp = &sh.kparam[cand][(1 << best_po) - 1];
    if (best_po == 6) { kp[2 * lane] = kbest0; kp[2 * lane + 1] = kbest1; }
    else {
        const int step = 1 << (F - best_po);
        if ((lane & (step - 1)) == 0 && lane < (1 << F)) kp[lane >> (F - best_po)] = kbest0;
    }
    uint32_t big = __ballot_sync(FULLMASK, kbest0 >= 15u || (best_po == 6 && kbest1 >= 15u));
    if (lane < 8) { sh.lvl_bits[cand][lane] = (lane == best_po) ? best : ~0ull; sh.lvl_big[cand][lane] = big ? 1u : 0u; }
    __syncwarp();
}

// The same search for ONE partition order of a full 4096-sample frame (2^L partitions of 256 >> L threads each): the compute
// warps run it side by side, warp L on level L.  The per-thread |r| sums must fit 32 bits (the lean paths guarantee it).
__device__ __noinline__ void analyse_level(EncSmall &sh, const uint64_t *tabs, int cand, int L, int pred_order, uint32_t kmax, int rsh) {
    const int lane = threadIdx.x & 31;
    const uint32_t nfull = (uint32_t)FRX_MAX_BLOCK >> L;
    uint32_t *kp = &sh.kparam[cand][(1 << L) - 1];
    uint64_t bits = 0;
    uint32_t big = 0;
    if (L <= 3) {
        // a row of 32 threads never straddles a partition: reduce the eight rows (lane t keeps row t), then groups of rows
        uint64_t mine = 0;
#pragma unroll
        for (int t = 0; t < 8; t++) {
            const uint64_t s = warp_sum_u32_wide((uint32_t)(tabs[32 * t + lane] >> rsh));
            if (lane == t) mine = s;
        }
        const int g = 8 >> L;
        for (int o = 1; o < g; o <<= 1) mine += __shfl_xor_sync(FULLMASK, mine, o);
        if (lane < 8 && (lane & (g - 1)) == 0) {
            const int p = lane >> (3 - L);
            const uint32_t n = nfull - (p == 0 ? (uint32_t)pred_order : 0u);
            const uint32_t k = rice_param(n, mine, kmax);
            bits = rice_estimate(k, n, mine);
            kp[p] = k;
            big = k >= 15u ? 1u : 0u;
        }
    } else {
        const int T = FRX_ENC_THREADS >> L;            // threads per partition: 16, 8, 4
        for (int p = lane; p < (1 << L); p += 32) {
            uint64_t s = 0;
            for (int i = 0; i < T; i++) s += tabs[p * T + i] >> rsh;
            const uint32_t n = nfull - (p == 0 ? (uint32_t)pred_order : 0u);
            const uint32_t k = rice_param(n, s, kmax);
            bits += rice_estimate(k, n, s);
            kp[p] = k;
            big |= k >= 15u ? 1u : 0u;
        }
    }
    bits = warp_sum_u64(bits) + 6;
    const uint32_t anybig = __ballot_sync(FULLMASK, big != 0u);
    if (lane == 0) { sh.lvl_bits[cand][L] = bits; sh.lvl_big[cand][L] = anybig ? 1u : 0u; }
}

// best partition order of a candidate: highest order first, a lower one only when strictly cheaper (libFLAC's loop).
// Called by whole warps: lane L looks at order L, one redux picks the minimum, the highest lane holding it wins.
__device__ __forceinline__ void pick_level(const EncSmall &sh, int cand, int po_max, int &po, uint64_t &bits, bool &rice2) {
    const int lane = threadIdx.x & 31;
    uint32_t key = 0xFFFFFFFFu, big = 0;
    if (lane <= po_max) {
        const uint64_t b = sh.lvl_bits[cand][lane];
        key = b < 0xFFFFFFFEull ? (uint32_t)b : 0xFFFFFFFEu;
        big = sh.lvl_big[cand][lane];
    }
    const uint32_t m = __reduce_min_sync(FULLMASK, key);
    const int bp = 31 - __clz((int)(__ballot_sync(FULLMASK, key == m) & 0xFFu));
    po = bp; bits = m; rice2 = __shfl_sync(FULLMASK, big, bp) != 0u;
}
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

// ------------------------------------------------------------------------------------------
// control warp pieces
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ int frame_header_bytes(uint8_t *hb, int nsamp, int sr_code, int sample_rate, int channels, int bps, uint32_t fj) {
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
    hb[n++] = (uint8_t)(((channels - 1) << 4) | (sz_code << 1));
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

// control warp: everything the compute warps need to know about frame f.  The lanes probe 32 tile boundaries at once
// and fetch the tile and its quantiser words side by side, so the whole descriptor costs two memory round trips.
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
        const uint32_t s0 = fj * (uint32_t)a.blocksize;
        const int nsamp = (int)min((long long)a.blocksize, tile_n - (long long)s0);
        d.f = f; d.fj = fj; d.s0 = s0; d.nsamp = nsamp; d.tile_idx = ti;
        {
            const int tz = __ffs(nsamp) - 1;               // trailing zeros of the block size
            int po_max = min(a.max_po, max(0, tz - 4));    // partitions stay multiples of 16 samples
            while (po_max > 0 && (nsamp >> po_max) <= FRX_MAX_LPC) po_max--;
            d.po_max = (int16_t)po_max;
            d.max_lpc = (int16_t)min(a.max_lpc_order, nsamp - 1);
        }
        const int epv = 16 / frx_dtype_size(a.dtype);            // elements per 16 bytes
        d.tma = (a.tma_ok && (t.w % epv) == 0 && (t.col_off % epv) == 0) ? 1 : 0;
        uint8_t hb[16];
        const int n = frame_header_bytes(hb, nsamp, a.sr_code, a.sample_rate, t.n_channels, a.bps, fj);
#pragma unroll
        for (int k = 0; k < 4; k++)
            d.hdr[k] = ((uint32_t)hb[4 * k] << 24) | ((uint32_t)hb[4 * k + 1] << 16) | ((uint32_t)hb[4 * k + 2] << 8) | hb[4 * k + 3];
        d.hdr_bits = (uint32_t)n * 8u;
    }
    __syncwarp();
}

// K8: decoupled look-back over the frame sizes (whole control warp).  A descriptor word carries flag and value
// together (A: the frame's size, P: the inclusive byte prefix), so relaxed device-scope accesses are enough.  The compute
// warps publish A themselves as soon as the size is final; here 512 descriptors are examined per round trip (16
// consecutive ones per lane, nearest first).  When the walk meets a frame that has not published yet, only that single
// word is polled -- re-reading the window would hammer the L2 slice the publishers store to -- and the walk is resumable,
// so the control warp keeps feeding the raw-sample pipeline while it waits.
#define LB_PER_LANE 16
struct LookBack {
    long long pos;        // nearest descriptor not yet accounted for
    uint64_t exclusive;   // bytes summed so far
    bool blocked;         // desc[pos] had not been published at the last visit
};
// one step; true when the exclusive prefix is complete
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
    if (lane == 0) DBG_ADD(3, 1);
    uint64_t v[LB_PER_LANE];
    const long long base = lb.pos - (long long)lane * LB_PER_LANE;
#pragma unroll
    for (int k = 0; k < LB_PER_LANE; k++) {
        const long long idx = base - k;
        v[k] = idx >= 0 ? frx_ld_relaxed_u64(&a.desc[idx]) : FLAG_P;
    }
    // per lane: walk its 16 entries nearest-first up to the first inclusive prefix (P) or the first one not ready
    uint64_t sum = 0;
    int state = 0;                     // 0: all 16 ready, no P   1: stopped at a P   2: stopped at a not-ready entry
    int taken = 0;
#pragma unroll
    for (int k = 0; k < LB_PER_LANE; k++) {
        if (state == 0) {
            const uint32_t flag = (uint32_t)(v[k] >> 62);
            if (flag == 0u) state = 2;
            else { sum += v[k] & VMASK; taken++; if (flag == 2u) state = 1; }
        }
    }
    const uint32_t mP = __ballot_sync(FULLMASK, state == 1), mN = __ballot_sync(FULLMASK, state == 2);
    const int firstP = mP ? (__ffs((int)mP) - 1) : 32, firstN = mN ? (__ffs((int)mN) - 1) : 32;
    // lanes before the first stopper contribute everything; the stopper contributes what it took
    const int stop = min(firstP, firstN);
    lb.exclusive += warp_sum_u64((lane <= stop) ? sum : 0ull);
    if (firstP < firstN) return true;  // reached an inclusive prefix
    if (firstN < 32) {
        const int tk = __shfl_sync(FULLMASK, taken, firstN);
        lb.pos -= (long long)firstN * LB_PER_LANE + tk;      // the frame that has not published its size yet
        lb.blocked = true;
        if (lane == 0) DBG_ADD(4, 1);
    } else lb.pos -= 32 * LB_PER_LANE;
    return false;
}

// The control warp's event loop.  A CTA works on one frame and has at most the next one staged; the next ticket is
// taken only when the current frame's size has been published, because a ticket that is reserved but not yet sized
// holds up the look-back of every later frame (and convoys of such tickets serialise the whole grid).
__device__ __noinline__ void control_warp(const EncArgs &a, EncSmall &sh, unsigned char *raw, int lane) {
    const int esz = frx_dtype_size(a.dtype);
    uint32_t unit_iss = 0, fseq_iss = 0, fseq_lb = 0;
    int ch_iss = 0, nch_cur = 0, ti = 0;
    bool have_frame = false, want_frame = true, stop_issued = false, lb_active = false;
    LookBack lb;
    lb.pos = -1; lb.exclusive = 0; lb.blocked = false;
    for (;;) {
        bool progress = false;
        // ---- 1. a frame's size became known: get the next frame going, then start this frame's look-back ----
        if (!lb_active && fseq_lb < fseq_iss) {
            const int slot = (int)(fseq_lb & (FD_SLOTS - 1));
            if (mbar_test(&sh.sized[slot], (fseq_lb / FD_SLOTS) & 1u)) {
                if (!stop_issued) want_frame = true;
                lb.pos = (long long)sh.fd[slot].f - 1;
                lb.exclusive = 0; lb.blocked = false;
                lb_active = true;
                if (lane == 0) DBG_TRACE(sh.fd[slot].f, 2);
                progress = true;
            }
        }
        // ---- 2. next unit (frame, channel) into the raw stage ring ----
        if (have_frame || want_frame) {
            const int stage = 0;
            if (mbar_test(&sh.empty[stage], (unit_iss & 1u) ^ 1u)) {
                progress = true;
                if (!have_frame) {
                    const int slot = (int)(fseq_iss & (FD_SLOTS - 1));
                    uint32_t ticket = 0;
                    if (lane == 0) ticket = atomicAdd(a.ticket, 1u);
                    ticket = __shfl_sync(FULLMASK, ticket, 0);
                    want_frame = false;
                    if (ticket >= a.total_frames) {
                        if (lane == 0) { sh.fd[slot].f = STOP_FRAME; mbar_arrive(&sh.full[stage]); }
                        __syncwarp();
                        stop_issued = true;
                    } else {
                        build_frame_desc(a, sh.fd[slot], ticket, ti, lane);
                        if (lane == 0) { DBG_TRACE(ticket, 0); DBG_TRACEV(ticket, 6, blockIdx.x); }
                        nch_cur = sh.fd[slot].t.n_channels;
                        have_frame = true;
                        ch_iss = 0;
                    }
                }
                if (have_frame) {
                    const FrameDesc &d = sh.fd[fseq_iss & (FD_SLOTS - 1)];
                    unsigned char *dst = raw;
                    if (d.tma) {
                        const long long w = d.t.w;
                        const long long s0 = d.s0, s1 = s0 + d.nsamp;
                        const long long row0 = s0 / w, row1 = (s1 - 1) / w;
                        const unsigned char *band = (const unsigned char *)a.raster +
                            ((long long)(d.t.first_band + ch_iss) * a.band_stride + (long long)d.t.row_off * a.row_stride + d.t.col_off) * esz;
                        if (lane == 0) mbar_arrive_expect_tx(&sh.full[stage], (uint32_t)d.nsamp * (uint32_t)esz);
                        __syncwarp();
                        for (long long r = row0 + lane; r <= row1; r += 32) {
                            const long long b0 = max(s0, r * w), b1 = min(s1, (r + 1) * w);
                            bulk_g2s(dst + (b0 - s0) * esz, band + (r * a.row_stride + (b0 - r * w)) * esz, (uint32_t)((b1 - b0) * esz), &sh.full[stage]);
                        }
                    } else if (lane == 0) {
                        mbar_arrive(&sh.full[stage]);
                    }
                    __syncwarp();
                    unit_iss++;
                    if (++ch_iss == nch_cur) { have_frame = false; fseq_iss++; }
                }
            }
        }
        // ---- 3. the look-back in flight ----
        if (lb_active) {
            if (lookback_step(a, lb, lane)) {
                const int slot = (int)(fseq_lb & (FD_SLOTS - 1));
                if (lane == 0) {
                    const uint32_t f = sh.fd[slot].f;
                    const uint32_t nbytes = sh.frame_bytes[slot];
                    const uint64_t off = lb.exclusive;
                    frx_st_relaxed_u64(&a.desc[f], (2ull << 62) | (off + (uint64_t)nbytes));
                    sh.out_offset[slot] = off;
                    if (a.frame_size) a.frame_size[f] = nbytes;
                    if (a.frame_offset) a.frame_offset[f] = off;
                    if (f == a.total_frames - 1 && a.total_bytes) *a.total_bytes = off + nbytes;
                    DBG_TRACE(f, 3);
                    DBG_ADD(6, 1);
                    mbar_arrive(&sh.placed[slot]);
                }
                __syncwarp();
                lb_active = false;
                fseq_lb++;
                progress = true;
            } else if (!lb.blocked) progress = true;
        }
        if (stop_issued && !lb_active && fseq_lb == fseq_iss) break;
        if (!progress) __nanosleep(200);
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

// copy a finished frame from its shared-memory buffer to its final place and append the CRC-16
__device__ __forceinline__ void copy_out_frame(const EncArgs &a, EncSmall &sh, const uint32_t *bits, const uint32_t *crc_part, uint64_t off,
                                               uint32_t body_bytes, int tid) {
    const uint32_t nbytes = body_bytes + 2u;
    if (off + nbytes > a.capacity) {
        if (tid == 0) atomicExch(a.status, FRX_E_OVERFLOW);
        return;
    }
    uint8_t *dst = a.bytes + off;
    if (tid == FRX_ENC_THREADS - 1) {
        // undo the zero-word padding of the last quad, finish over the <= 3 tail bytes, store big-endian
        uint32_t c = 0;
#pragma unroll
        for (int w2 = 0; w2 < NWARPS; w2++) c ^= crc_part[w2];
        const uint32_t nw = body_bytes >> 2;
        const uint32_t z = ((nw + 3u) & ~3u) - nw;
        if (z && c) c = frx_gf16_mulmod(c, (uint32_t)a.crc_xinv[z]);
        const uint32_t tw = bits[nw];
        for (uint32_t b = 0; b < (body_bytes & 3u); b++) {
            const uint32_t byte = (tw >> (24 - 8 * b)) & 0xFFu;
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
        const uint32_t w0 = bits[j], w1 = bits[j + 1];
        const uint32_t v = head ? __funnelshift_l(w1, w0, sh8) : w0;
        dw[j] = frx_bswap32(v);
    }
    const uint32_t tail0 = head + (nwords << 2);
    if (tid < (int)(body_bytes - tail0)) {
        const uint32_t b = tail0 + (uint32_t)tid;
        dst[b] = (uint8_t)((bits[b >> 2] >> (24 - 8 * (b & 3))) & 0xFFu);
    }
}

// CRC-16 of the frame body's whole words: every compute thread takes a chunk of Lq quads, the chunk CRCs are shifted to
// their place with x^(128 Lq j) mod P and xor-ed (per warp here, across warps by copy_out_frame)
// shift factor of this thread's chunk; loaded ahead of the barrier in front of crc_frame so its latency is hidden
__device__ __forceinline__ uint32_t crc_shift_factor(const EncArgs &a, uint32_t body_bytes, int tid) {
    const uint32_t nq = ((body_bytes >> 2) + 3u) >> 2;
    const uint32_t Lq = (nq + FRX_ENC_THREADS - 1) / FRX_ENC_THREADS;
    return (uint32_t)__ldg(&a.crc_xpow[(max(Lq, 1u) - 1) * FRX_ENC_THREADS + (FRX_ENC_THREADS - 1 - tid)]);
}
__device__ __forceinline__ void crc_frame(const EncArgs &a, EncSmall &sh, const uint32_t *bits, uint32_t *crc_part, uint32_t body_bytes,
                                          int tid, uint32_t shift_factor) {
    const uint32_t nw = body_bytes >> 2;                       // whole words
    const uint32_t nq = (nw + 3u) >> 2;                        // quads; the buffer is zero behind the frame
    const uint32_t Lq = (nq + FRX_ENC_THREADS - 1) / FRX_ENC_THREADS;
    const int padq = (int)(FRX_ENC_THREADS * Lq) - (int)nq;    // virtual leading zero quads
    int q0 = tid * (int)Lq - padq;
    const int q1 = q0 + (int)Lq;
    if (q0 < 0) q0 = 0;
    uint32_t c = 0;
    for (int qi = q0; qi < q1; qi++) {
        uint4 wv = *reinterpret_cast<const uint4 *>(bits + 4 * qi);
        if (qi == (int)nq - 1) {   // the word behind the last whole word holds the <= 3 tail bytes: not part of this pass
            const uint32_t k = nw - 4u * (uint32_t)qi;   // whole words in the last quad (1..4)
            if (k < 4u) wv.w = 0u;
            if (k < 3u) wv.z = 0u;
            if (k < 2u) wv.y = 0u;
        }
        uint32_t x = wv.x ^ (c << 16);
        c = (uint32_t)sh.crc_tab[3][x >> 24] ^ sh.crc_tab[2][(x >> 16) & 0xFFu] ^ sh.crc_tab[1][(x >> 8) & 0xFFu] ^ sh.crc_tab[0][x & 0xFFu];
        x = wv.y ^ (c << 16);
        c = (uint32_t)sh.crc_tab[3][x >> 24] ^ sh.crc_tab[2][(x >> 16) & 0xFFu] ^ sh.crc_tab[1][(x >> 8) & 0xFFu] ^ sh.crc_tab[0][x & 0xFFu];
        x = wv.z ^ (c << 16);
        c = (uint32_t)sh.crc_tab[3][x >> 24] ^ sh.crc_tab[2][(x >> 16) & 0xFFu] ^ sh.crc_tab[1][(x >> 8) & 0xFFu] ^ sh.crc_tab[0][x & 0xFFu];
        x = wv.w ^ (c << 16);
        c = (uint32_t)sh.crc_tab[3][x >> 24] ^ sh.crc_tab[2][(x >> 16) & 0xFFu] ^ sh.crc_tab[1][(x >> 8) & 0xFFu] ^ sh.crc_tab[0][x & 0xFFu];
    }
    if (c) c = frx_gf16_mulmod(c, shift_factor);
    c = __reduce_xor_sync(FULLMASK, c);
    if ((tid & 31) == 0) crc_part[tid >> 5] = c;
}

// ------------------------------------------------------------------------------------------
// The frame encoder.  WIDE: PCM wider than 16 bits (64-bit LPC accumulate, RICE2).  HISTN: history samples a thread
// reads = the largest LPC order of the compression level (4 when the level has no LPC stage).
// 8 compute warps (one frame at a time, 16 samples per thread) + 1 control warp.  The PCM of the current subframe
// lives in padded shared memory and every stage re-reads it, so almost no register state crosses a barrier.
// ------------------------------------------------------------------------------------------
template <bool WIDE, int HISTN>
__global__ void __launch_bounds__(FRX_ENC_CTA, 2) frx_encode_kernel(const EncArgs a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    EncSmall &sh = *reinterpret_cast<EncSmall *>(smem_raw);
    const uint32_t stage_stride = 128u + (uint32_t)FRX_MAX_BLOCK * (uint32_t)frx_dtype_size(a.dtype);
    unsigned char *raw = smem_raw + ENC_SMALL_BYTES;
    int32_t *pcm = reinterpret_cast<int32_t *>(raw + stage_stride);
    int32_t *res = reinterpret_cast<int32_t *>(reinterpret_cast<unsigned char *>(pcm) + PCM_BYTES);
    uint64_t *tabs = reinterpret_cast<uint64_t *>(reinterpret_cast<unsigned char *>(res) + RES_BYTES);
    uint32_t *bits0 = reinterpret_cast<uint32_t *>(reinterpret_cast<unsigned char *>(tabs) + TABS_BYTES);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < 2 * a.frame_words; i += FRX_ENC_CTA) bits0[i] = 0u;
    for (int i = tid; i < PCM_WORDS; i += FRX_ENC_CTA) pcm[i] = 0;
    for (int i = tid; i < 1024; i += FRX_ENC_CTA) (&sh.crc_tab[0][0])[i] = a.crc_tab[i];
    if (tid == 0) {
        mbar_init(&sh.full[0], 1); mbar_init(&sh.empty[0], NWARPS);
        for (int i = 0; i < FD_SLOTS; i++) { mbar_init(&sh.sized[i], 1); mbar_init(&sh.placed[i], 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();
    if (warp == NWARPS) { control_warp(a, sh, raw, lane); return; }

    const uint32_t kmax = a.bps > 16 ? 30u : 14u;  // RICE2 only for bps > 16 (libFLAC)
    const int g0 = tid * FRX_SPT;                  // first sample index of this thread
    uint32_t unit = 0, fseq = 0;
    // copy-out runs one frame behind (double-buffered frame buffer), so the look-back never stalls the compute warps
    bool pend = false;
    uint32_t pend_body = 0;

    const long long ck0 = DBG_CLK();
    DBG_PH_DECL;
    for (;;) {
        // ---- next frame: its descriptor is valid once the first raw stage of the frame is full ----
        const int slot = (int)(fseq & (FD_SLOTS - 1));
        const int buf = (int)(fseq & 1u);
        uint32_t *bits = bits0 + (size_t)buf * a.frame_words;
        mbar_wait(&sh.full[0], unit & 1u);
        const FrameDesc &fd = sh.fd[slot];
        const uint32_t f = fd.f;
        if (f == STOP_FRAME) { if (tid == 0) DBG_ADD(7, DBG_CLK() - ck0); break; }
        if (tid == 0) DBG_TRACE(f, 4);
        DBG_PH(15);
        const int n_channels = fd.t.n_channels;
        const uint32_t s0 = fd.s0;
        const int nsamp = fd.nsamp;
        const bool use_tma = fd.tma != 0;
        const int nv = min(FRX_SPT, max(0, nsamp - g0));  // valid samples of this thread
        const bool full = (nsamp == FRX_MAX_BLOCK);
        const bool plain = full && tid != 0;              // no per-sample masks needed for this thread
        uint32_t bitpos = fd.hdr_bits;
        if (tid < 4) bits[tid] = fd.hdr[tid];
        const int po_max = fd.po_max;
        const int max_lpc = min((int)fd.max_lpc, HISTN);

        // ================= per channel =================
        for (int ch = 0; ch < n_channels; ch++) {
            unsigned char *stg = raw;
            if (ch > 0) mbar_wait(&sh.full[0], unit & 1u);
            if (!use_tma) {
                gather_raw(a, fd.t, ch, s0, nsamp, stg, tid);
                compute_sync();
            }
            // ---- K1: the thread's 16 samples -> PCM -> shared memory ----
            uint32_t orv = 0, ora = 0;
            {
                uint2 oo;
                if (a.dtype == FRX_F32) oo = k1_quantize<FRX_F32, WIDE>(stg, fd.q, a.pcm_mode, tid, nv, pcm);
                else if (a.dtype == FRX_U16) oo = k1_quantize<FRX_U16, WIDE>(stg, fd.q, a.pcm_mode, tid, nv, pcm);
                else oo = k1_quantize_any<WIDE>(a.dtype, stg, fd.q, a.pcm_mode, tid, nv, pcm);
                orv = oo.x; ora = oo.y;
                // the raw stage may be refilled as soon as every warp has read its samples
                if (!use_tma) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                __syncwarp();
                if (lane == 0) mbar_arrive(&sh.empty[0]);
                unit++;
            }
            orv = __reduce_or_sync(FULLMASK, orv);
            if (WIDE) ora = __reduce_or_sync(FULLMASK, ora);
            compute_sync();  // S0: PCM of the subframe is in shared memory
            DBG_PH(9);
            // ---- K2 (lean, 32-bit) fixed-order trial + K3 windowed autocorrelation partials ----
            {
                int32_t s[HISTN + FRX_SPT];
                read_pcm<HISTN>(pcm, tid, 0, s);
                uint32_t fa[5];
                if (plain) fixed_trial32<HISTN>(s, fa);
                else fixed_trial32_masked<HISTN>(s, g0, nv, fa);
#pragma unroll
                for (int o = 0; o < 5; o++) tabs[o * FRX_ENC_THREADS + tid] = fa[o];
                {
                    uint64_t w5[5];
#pragma unroll
                    for (int o = 0; o < 5; o++) w5[o] = WIDE ? warp_sum_u32_wide(fa[o]) : (uint64_t)__reduce_add_sync(FULLMASK, fa[o]);
                    if (lane == 0) {
#pragma unroll
                        for (int o = 0; o < 5; o++) sh.wfix[warp][o] = w5[o];
                        sh.wor[warp][0] = orv; sh.wor[warp][1] = ora;
                    }
                }
                if (max_lpc > 0) {
                    float xw[HISTN + FRX_SPT];
                    if (full) {
                        // window table: zero history of thread 0 comes from the zero samples in front of the PCM buffer
                        const float4 *w4 = reinterpret_cast<const float4 *>(a.window + g0 - HISTN);
#pragma unroll
                        for (int v = 0; v < (HISTN + FRX_SPT) / 4; v++) {
                            float4 wv = make_float4(0.f, 0.f, 0.f, 0.f);
                            if (tid > 0 || v >= HISTN / 4) wv = __ldg(w4 + v);
                            xw[4 * v] = (float)s[4 * v] * wv.x; xw[4 * v + 1] = (float)s[4 * v + 1] * wv.y;
                            xw[4 * v + 2] = (float)s[4 * v + 2] * wv.z; xw[4 * v + 3] = (float)s[4 * v + 3] * wv.w;
                        }
                    } else {
#pragma unroll
                        for (int i = 0; i < HISTN + FRX_SPT; i++) xw[i] = (float)s[i] * tukey_w(g0 - HISTN + i, nsamp);
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
                    for (int lag = 0; lag <= HISTN; lag++) {
                        float v = ac[lag];
#pragma unroll
                        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULLMASK, v, o);
                        if (lane == 0) sh.wac[warp][lag] = v;
                    }
                }
            }
            compute_sync();  // SA: per-warp partials and tabs[0..4] are in shared memory
            DBG_PH(10);
            // ---- every warp: fixed-order decision (cheap, redundant).  Warps 0..6: Rice search of the fixed candidate, one
            //      partition order each; warp 7: LPC design ----
            {
                uint64_t tot = 0;
                uint32_t o32 = 0;
                if (lane < 5) { for (int w2 = 0; w2 < NWARPS; w2++) tot += sh.wfix[w2][lane]; }
                else if (lane < 7) { for (int w2 = 0; w2 < NWARPS; w2++) o32 |= sh.wor[w2][lane - 5]; }
                const uint32_t all_or = __shfl_sync(FULLMASK, o32, 5), all_abs = __shfl_sync(FULLMASK, o32, 6);
                const int wasted = all_or ? (__ffs((int)all_or) - 1) : 0;
                const int magbits = WIDE ? 32 - __clz((int)all_abs) : 15;  // bits of the largest magnitude
                const bool lean = !WIDE || magbits <= 23;
                if (warp == NWARPS - 1) {
                    float acs = 0.f;
                    if (lane <= max_lpc) { for (int w2 = 0; w2 < NWARPS; w2++) acs += sh.wac[w2][lane]; }
                    if (lane <= FRX_MAX_LPC) sh.autoc[lane] = acs;
                    if (lane == 0) { sh.magbits = magbits; sh.lean = lean ? 1 : 0; sh.lpc_ok = 0; }
                    __syncwarp();
                    const long long cl0 = DBG_CLK();
                    if (max_lpc > 0 && nsamp > 4)
                        lpc_design<WIDE, HISTN>(sh, max_lpc, nsamp, wasted, a.bps - wasted, a.qlp_precision, magbits, lean, lane);
                    if (lane == 0) DBG_ADD(1, DBG_CLK() - cl0);
                } else {
                    const uint64_t f1 = __shfl_sync(FULLMASK, tot, 1);
                    uint64_t bestv = (lane < 5) ? tot : ~0ull;   // best fixed order among lanes 0..4 (lowest order wins ties)
                    int besto = lane;
                    for (int o = 1; o < 8; o <<= 1) {
                        const uint64_t ov = __shfl_xor_sync(FULLMASK, bestv, o);
                        const int oo = __shfl_xor_sync(FULLMASK, besto, o);
                        if (ov < bestv || (ov == bestv && oo < besto)) { bestv = ov; besto = oo; }
                    }
                    besto = __shfl_sync(FULLMASK, besto, 0);
                    const int fixed_order = (nsamp > 4) ? besto : -1;
                    if (warp == 0 && lane == 0) {
                        sh.wasted = wasted;
                        sh.is_const = (f1 == 0) ? 1 : 0;   // sum |s[g]-s[g-1]| == 0  <=>  all samples equal
                        sh.fixed_order = fixed_order;
                    }
                    if (fixed_order >= 0 && lean) {          // fat data: fat_search redoes the fixed candidate in 64 bits
                        const long long ca0 = DBG_CLK();
                        if (full) {
                            if (warp <= po_max) analyse_level(sh, tabs + fixed_order * FRX_ENC_THREADS, 0, warp, fixed_order, kmax, wasted);
                            if (lane == 0 && warp == 6) DBG_ADD(0, DBG_CLK() - ca0);
                            if (lane == 0 && warp == 0) DBG_ADD(2, DBG_CLK() - ca0);
                        } else if (warp == 0) analyse_partitions(sh, tabs + fixed_order * FRX_ENC_THREADS, 0, nsamp, fixed_order, po_max, kmax, wasted);
                    }
                }
            }
            compute_sync();  // SB
            DBG_PH(11);
            const int wasted = sh.wasted;
            const bool lean = sh.lean != 0;
            const int sbps = a.bps - wasted;
            const int lpc_order = sh.lpc_order;
            const int lshift = sh.shift;
            if (WIDE && !lean) {
                fat_search<HISTN>(sh, pcm, tabs, tid, nv, nsamp, wasted, po_max, kmax);
            } else {
                // ---- K5: LPC residuals (kept in registers) and their |r| sums ----
                const bool lpc_try0 = sh.lpc_ok != 0;
                if (lpc_try0) {
                    int32_t s[HISTN + FRX_SPT], q[HISTN];
                    read_pcm<HISTN>(pcm, tid, wasted, s);
#pragma unroll
                    for (int j = 0; j < HISTN; j++) q[j] = sh.qlp[j];
                    uint32_t ab;
                    if (!WIDE || sh.lpc_narrow) ab = lpc_residuals_park<0, HISTN>(s, q, lshift, res, tid, plain, nv, lpc_order - g0);
                    else ab = lpc_residuals_park<1, HISTN>(s, q, lshift, res, tid, plain, nv, lpc_order - g0);
                    tabs[5 * FRX_ENC_THREADS + tid] = ab;
                }
                compute_sync();  // SC
                DBG_PH(8);
                if (lpc_try0) {
                    if (full) { if (warp <= po_max) analyse_level(sh, tabs + 5 * FRX_ENC_THREADS, 1, warp, lpc_order, kmax, 0); }
                    else if (warp == 1) analyse_partitions(sh, tabs + 5 * FRX_ENC_THREADS, 1, nsamp, lpc_order, po_max, kmax, 0);
                }
                compute_sync();  // SD
                DBG_PH(12);
            }
            const int fixed_order = sh.fixed_order;
            const bool is_const = sh.is_const != 0;
            const bool lpc_try = sh.lpc_ok != 0;

            // ---- choose ----
            const uint64_t verb_bits = 8ull + (uint64_t)wasted + (uint64_t)nsamp * (uint64_t)sbps;
            int type = 1, order = 0, cand = 0, po = 0;
            bool rice2 = false;
            uint64_t best_bits = verb_bits;
            if (is_const) { type = 0; best_bits = 8ull + wasted + sbps; }
            else {
                if (fixed_order >= 0) {
                    int lp; uint64_t lb; bool l2;
                    pick_level(sh, 0, po_max, lp, lb, l2);
                    const uint64_t b = 8ull + wasted + (uint64_t)fixed_order * sbps + lb;
                    if (b < best_bits) { best_bits = b; type = 2; order = fixed_order; cand = 0; po = lp; rice2 = l2; }
                }
                if (lpc_try) {
                    int lp; uint64_t lb; bool l2;
                    pick_level(sh, 1, po_max, lp, lb, l2);
                    const uint64_t b = 8ull + wasted + 9ull + (uint64_t)lpc_order * (uint64_t)(sh.precision + sbps) + lb;
                    if (b < best_bits) { best_bits = b; type = 3; order = lpc_order; cand = 1; po = lp; rice2 = l2; }
                }
            }
            // ---- K7a: winner's residuals (zig-zag folded), exact bit count per thread, block scan ----
            const uint32_t sub_start = bitpos;
            const uint32_t smask = sbps >= 32 ? 0xFFFFFFFFu : ((1u << sbps) - 1u);
            int psize = nsamp;
            uint32_t kk = 0;
            uint32_t mybits = 0;
            const int skip = (tid == 0) ? order : 0;
            if (type >= 2) {
                psize = nsamp >> po;
                kk = sh.kparam[cand][(1 << po) - 1 + min(g0 / psize, (1 << po) - 1)];
                int32_t r[FRX_SPT];
                if (type == 2 || (WIDE && !lean)) {
                    // fixed winner (4 ops per sample) or a fat-path LPC winner: recompute from the PCM, park in the residual buffer
                    int32_t s[HISTN + FRX_SPT];
                    read_pcm<HISTN>(pcm, tid, wasted, s);
                    if (type == 2) {
                        if (WIDE && !lean) (void)fixed_residuals<true, HISTN>(order, s, r);
                        else (void)fixed_residuals<false, HISTN>(order, s, r);
                    } else if (WIDE) {
                        int32_t q[HISTN];
#pragma unroll
                        for (int j = 0; j < HISTN; j++) q[j] = sh.qlp[j];
                        (void)lpc_residuals_dispatch<1, HISTN>(order, s, q, lshift, r);
                    }
#pragma unroll
                    for (int i = 0; i < FRX_SPT; i++) res[RES_IDX(i, tid)] = r[i];
                } else {
                    // lean LPC winner: K5 left its residuals in the buffer
#pragma unroll
                    for (int i = 0; i < FRX_SPT; i++) r[i] = res[RES_IDX(i, tid)];
                }
                if (plain) {
                    uint32_t qs = 0;
#pragma unroll
                    for (int i = 0; i < FRX_SPT; i++) qs += zigzag(r[i]) >> kk;
                    mybits = qs + (uint32_t)FRX_SPT * (1u + kk);
                } else {
#pragma unroll
                    for (int i = 0; i < FRX_SPT; i++)
                        if (i >= skip && i < nv) mybits += (zigzag(r[i]) >> kk) + 1u + kk;
                }
                if (nv > 0 && g0 % psize == 0) mybits += rice2 ? 5u : 4u;
                if (tid == 0) {
                    mybits += 8u + (uint32_t)wasted + (uint32_t)order * (uint32_t)sbps + 6u;
                    if (type == 3) mybits += 9u + (uint32_t)order * (uint32_t)sh.precision;
                }
            } else if (type == 1) {
                mybits = (uint32_t)nv * (uint32_t)sbps;
                if (tid == 0) mybits += 8u + (uint32_t)wasted;
            } else {
                if (tid == 0) mybits = 8u + (uint32_t)wasted + (uint32_t)sbps;
            }
            uint32_t incl = mybits;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t n = __shfl_up_sync(FULLMASK, incl, o);
                if (lane >= o) incl += n;
            }
            if (lane == 31) sh.scan[warp] = incl;
            compute_sync();  // SE
            DBG_PH(13);
            uint32_t wbase = 0, total = 0;
#pragma unroll
            for (int w2 = 0; w2 < NWARPS; w2++) {
                const uint32_t v = sh.scan[w2];
                if (w2 < warp) wbase += v;
                total += v;
            }
            if (type >= 2 && (uint64_t)total >= verb_bits) {
                // the Rice-coded subframe would not beat VERBATIM: write it verbatim (keeps the frame inside its buffer)
                type = 1;
                mybits = (uint32_t)nv * (uint32_t)sbps;
                if (tid == 0) mybits += 8u + (uint32_t)wasted;
                incl = mybits;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t n = __shfl_up_sync(FULLMASK, incl, o);
                    if (lane >= o) incl += n;
                }
                if (lane == 31) sh.scan2[warp] = incl;
                compute_sync();
                wbase = 0; total = 0;
#pragma unroll
                for (int w2 = 0; w2 < NWARPS; w2++) {
                    const uint32_t v = sh.scan2[w2];
                    if (w2 < warp) wbase += v;
                    total += v;
                }
            }
            const uint32_t start = sub_start + wbase + incl - mybits;
            bitpos = sub_start + total;
            if (ch == n_channels - 1 && tid == 0) {
                // the frame's size is final: hand it to the control warp so the look-back overlaps packing and the CRC
                const uint32_t nb = ((bitpos + 7u) >> 3) + 2u;
                if (f != 0) frx_st_relaxed_u64(&a.desc[f], (1ull << 62) | (uint64_t)nb);
                DBG_TRACE(f, 1);   // flag A: size known, prefix not yet
                sh.frame_bytes[slot] = nb;
                mbar_arrive(&sh.sized[slot]);
            }
            // ---- K7b: pack ----
            if (type >= 2) {
                // every header field sits at a position known in closed form, so the lanes of warp 0 write them side by side
                const uint32_t precision = (uint32_t)sh.precision;
                const uint32_t p_warm = sub_start + 8u + (uint32_t)wasted;
                const uint32_t p_lpc = p_warm + (uint32_t)order * (uint32_t)sbps;
                const uint32_t p_rice = p_lpc + (type == 3 ? 9u + (uint32_t)order * precision : 0u);
                if (tid == 0) {
                    // subframe header: 0 | type(6) | wasted flag, then the wasted-bits count in unary
                    const uint32_t tcode = type == 2 ? (8u | (uint32_t)order) : (32u | (uint32_t)(order - 1));
                    put_field(bits, sub_start, (tcode << 1) | (wasted ? 1u : 0u), 8u);
                    if (wasted) put_field(bits, sub_start + 8u + (uint32_t)wasted - 1u, 1u, 1u);
                    if (type == 3) put_field(bits, p_lpc, ((precision - 1u) << 5) | ((uint32_t)sh.shift & 31u), 9u);
                    put_field(bits, p_rice, ((rice2 ? 1u : 0u) << 4) | (uint32_t)po, 6u);
                }
                if (tid < order) {
                    put_field(bits, p_warm + (uint32_t)tid * (uint32_t)sbps, (uint32_t)(pcm[PCM_IDX(tid)] >> wasted) & smask, (uint32_t)sbps);   // warm-up
                    if (type == 3) put_field(bits, p_lpc + 9u + (uint32_t)tid * precision, (uint32_t)sh.qlp[tid] & ((1u << precision) - 1u), precision);
                }
                if (mybits) {
                    uint32_t p = tid == 0 ? p_rice + 6u : start;
                    if (g0 % psize == 0) {
                        const uint32_t pb = rice2 ? 5u : 4u;
                        if (kk) put_field(bits, p, kk, pb);
                        p += pb;
                    }
                    pack_rice(bits, p, res, tid, kk, skip, nv, plain);
                }
            } else if (mybits) {
                BitW bw;
                bw.init(bits, start);
                if (tid == 0) {
                    bw.put((type == 0 ? 0u : 2u) | (wasted ? 1u : 0u), 8);
                    if (wasted) { if (wasted > 1) bw.skip((uint32_t)wasted - 1u); bw.put(1u, 1); }
                }
                if (type == 0) {
                    if (tid == 0) bw.put((uint32_t)(pcm[PCM_IDX(0)] >> wasted) & smask, sbps);
                } else {
                    for (int i = 0; i < nv; i++) bw.put((uint32_t)(pcm[PCM_IDX(g0 + i)] >> wasted) & smask, sbps);
                }
                bw.finish();
            }
            // the PCM buffer is rewritten by the next channel's K1: every reader of it (verbatim / warm-up packing) must be done
            if (ch + 1 < n_channels) compute_sync();
        }  // channels
        const uint32_t crc_xp = crc_shift_factor(a, (bitpos + 7u) >> 3, tid);
        compute_sync();  // Sa: every subframe is packed
        DBG_PH(14);

        // ---- frame footer: CRC-16 of this frame; copy-out of the PREVIOUS frame (its byte offset has had a frame time to arrive) ----
        const uint32_t body_bytes = (bitpos + 7u) >> 3;
        const long long cc0 = DBG_CLK();
        crc_frame(a, sh, bits, sh.crc_part[buf], body_bytes, tid, crc_xp);

        if (pend) {
            const int pslot = (int)((fseq - 1u) & (FD_SLOTS - 1));
            uint32_t *pbits = bits0 + (size_t)(buf ^ 1) * a.frame_words;
            const long long cp0 = DBG_CLK();
            mbar_wait(&sh.placed[pslot], ((fseq - 1u) / FD_SLOTS) & 1u);
            if (tid == 0) { DBG_TRACE(f, 7); }
            copy_out_frame(a, sh, pbits, sh.crc_part[buf ^ 1], sh.out_offset[pslot], pend_body, tid);
            compute_sync();  // all reads of the previous frame's buffer are done
            for (uint32_t j = tid; j <= (pend_body >> 2); j += FRX_ENC_THREADS) pbits[j] = 0u;
        }
        pend = true;
        pend_body = body_bytes;
        fseq++;
        compute_sync();  // crc_part / cleared buffer visible; the next frame's header words go into a clean buffer
        DBG_PH(5);
    }
    if (pend) {   // flush the last frame
        const int pslot = (int)((fseq - 1u) & (FD_SLOTS - 1));
        const int pbuf = (int)((fseq - 1u) & 1u);
        mbar_wait(&sh.placed[pslot], ((fseq - 1u) / FD_SLOTS) & 1u);
        copy_out_frame(a, sh, bits0 + (size_t)pbuf * a.frame_words, sh.crc_part[pbuf], sh.out_offset[pslot], pend_body, tid);
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
size_t frx_enc_smem_bytes(int frame_words, int dtype) {
    return ENC_SMALL_BYTES + (128 + (size_t)FRX_MAX_BLOCK * frx_dtype_size(dtype)) + PCM_BYTES + RES_BYTES + TABS_BYTES + 2 * sizeof(uint32_t) * (size_t)frame_words;
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
    size_t smem = frx_enc_smem_bytes(a.frame_words, a.dtype);
    cudaError_t e = cudaFuncSetAttribute(frx_encode_kernel<WIDE, HISTN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(frx_encode_kernel<WIDE, HISTN>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if (e != cudaSuccess) return e;
    int occ = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, frx_encode_kernel<WIDE, HISTN>, FRX_ENC_CTA, smem);
    if (e != cudaSuccess) return e;
    if (occ < 1) return cudaErrorLaunchOutOfResources;
    long long grid = (long long)sm_count * occ;   // persistent: every CTA resident, tickets hand out the frames in order
    if (grid > (long long)a.total_frames) grid = a.total_frames;
    if (grid < 1) grid = 1;
    frx_encode_kernel<WIDE, HISTN><<<(unsigned)grid, FRX_ENC_CTA, smem, st>>>(a);
    return cudaGetLastError();
}

cudaError_t frx_launch_encode(const EncArgs &a, int dtype, int sm_count, cudaStream_t st) {
    (void)dtype;
    const bool wide = a.bps > 16;
    const int hist = a.max_lpc_order <= 4 ? 4 : (a.max_lpc_order <= 8 ? 8 : 12);
#ifdef FRX_DEV_L5   // development builds (scratch/dev/build.sh): only the level-5 instantiations, 20 s instead of 5 min
    if (hist != 8) return cudaErrorInvalidValue;
    return wide ? launch_encode_t<true, 8>(a, sm_count, st) : launch_encode_t<false, 8>(a, sm_count, st);
#else
    if (wide) {
        if (hist == 4) return launch_encode_t<true, 4>(a, sm_count, st);
        if (hist == 8) return launch_encode_t<true, 8>(a, sm_count, st);
        return launch_encode_t<true, 12>(a, sm_count, st);
    }
    if (hist == 4) return launch_encode_t<false, 4>(a, sm_count, st);
    if (hist == 8) return launch_encode_t<false, 8>(a, sm_count, st);
    return launch_encode_t<false, 12>(a, sm_count, st);
#endif
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

#ifdef FRX_DBG
extern "C" void frx_debug_trace(unsigned long long *dev_ptr) { cudaMemcpyToSymbol(g_frx_trace, &dev_ptr, sizeof(dev_ptr)); }
extern "C" void frx_debug_read(unsigned long long *out, int reset) {
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out, g_frx_dbg, sizeof(unsigned long long) * 16);
    if (reset) { unsigned long long z[16] = {0}; cudaMemcpyToSymbol(g_frx_dbg, z, sizeof(z)); }
}
#endif

// frx_encode.cu -- raster -> FLAC frames on sm_100a.
//
// Kernels (SURVEY.md section 8a ids):
//   K0  frx_minmax_kernel        per-stream min/max                      (converter.py:67-68, 152-153)
//   K1..K8 frx_encode_kernel     one WARP per 4096-sample frame, twelve frame warps per SM, persistent grid,
//        ticket-ordered:
//        K1 tile cut + band->channel gather (TMA bulk copies) + PCM quantise   (converter.py:56-86, 185-194; cli.py:691-699)
//        K2 fixed-order 0..4 trial            K3 windowed autocorrelation
//        K4 Levinson-Durbin + order pick + coefficient quantisation    K5 residual (in place)
//        K6 Rice partition sums, parameter + partition-order choice    K7 bit packing (in place) + CRC-8/16
//        K8 decoupled look-back over frame sizes -> final byte offset; frame copied to its place, at once or from its slot
//   frx_tile_index_kernel        per-tile offset/size/minmax from the per-frame table
//
// What libFLAC does per frame (process_subframe_; restated in oracle/flac_oracle.c) is kept in spirit:
// candidates are CONSTANT, VERBATIM, the best FIXED order by sum|e|, and the LPC order picked from the
// Levinson error curve; each candidate is costed with the partition-sum Rice estimate and the cheapest
// wins; two-channel streams also try mid and side (left/side, side/right, mid/side).  The search is our own
// (the bitstream, not the decisions, is the contract).
#include <math_constants.h>

#include "frx_internal.cuh"



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
    // integer dtypes accumulate in 32 bits (one min and one max instruction per element; 64-bit min/max made K0 compute-bound on
    // 16-bit rasters): biased to unsigned order for uint32, widened to int64 once per thread at the end
    uint32_t amin = 0xFFFFFFFFu, amax = 0u;
    float fmin = CUDART_INF_F, fmax = -CUDART_INF_F;
    int nan = 0;
    // 16-byte loads when the tile rows allow it (every thread keeps several in flight), element loads otherwise
    constexpr int VEC = 16 / (int)sizeof(T);
    const bool vec_ok = (((uintptr_t)base & 15u) == 0) && (band_stride % VEC) == 0 && (row_stride % VEC) == 0 && (t.col_off % VEC) == 0 && (t.w % VEC) == 0 &&
                        vrows < 0x7FFFFFFFLL;     // (the vector path indexes the tile's virtual rows in 32 bits)
    auto take = [&](T v) {
        if (DT == FRX_F32) {
            const float f = (float)v;
            if (f != f) nan = 1;
            fmin = fminf(fmin, f); fmax = fmaxf(fmax, f);
        } else {
            const uint32_t iv = DT == FRX_U32 ? (uint32_t)v : ((uint32_t)(int32_t)v ^ 0x80000000u);     // order-preserving
            amin = min(amin, iv); amax = max(amax, iv);
        }
    };
    auto row_ptr = [&](long long vr) {
        const int c = (int)(vr / t.h);
        const int r = (int)(vr - (long long)c * t.h);
        return base + (long long)(t.first_band + c) * band_stride + (long long)(t.row_off + r) * row_stride + t.col_off;
    };
    if (vec_ok) {
        // the unit's 16-byte vectors, rows x (w / VEC), spread over the whole CTA with eight loads in flight per thread (narrow
        // tiles: with one row per pass only w / VEC of the 256 threads had work -- C4's 256-wide tiles ran at 0.7 TB/s)
        const int n4 = t.w / VEC;
        const int items = (int)(vrow1 - vrow0) * n4;                 // <= rows_per_unit_elems / VEC
        const int th = t.h;
        if (n4 >= (int)blockDim.x) {
            // wide rows: every thread has work in every row; four rows at a time, row pointers computed once per pass
            for (long long vr = vrow0; vr < vrow1; vr += 4) {
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
        } else
        for (int i0 = threadIdx.x; i0 < items; i0 += 8 * (int)blockDim.x) {
            uint4 q[8];
            bool ok[8];
#pragma unroll
            for (int k = 0; k < 8; k++) {
                const int i = i0 + k * (int)blockDim.x;
                ok[k] = i < items;
                if (ok[k]) {
                    const int rr = i / n4, x = i - rr * n4;
                    const int vr = (int)vrow0 + rr, c = vr / th, r = vr - c * th;
                    q[k] = __ldg(reinterpret_cast<const uint4 *>(base + (long long)(t.first_band + c) * band_stride +
                                                                 (long long)(t.row_off + r) * row_stride + t.col_off) + x);
                }
            }
#pragma unroll
            for (int k = 0; k < 8; k++) {
                if (ok[k]) {
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
    if (DT != FRX_F32 && amin <= amax) {
        imin = DT == FRX_U32 ? (long long)amin : (long long)(int32_t)(amin ^ 0x80000000u);
        imax = DT == FRX_U32 ? (long long)amax : (long long)(int32_t)(amax ^ 0x80000000u);
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
    o.rcp = o.range != 0.0 ? 1.0 / o.range : 0.0;   // correctly rounded (IEEE division)
    o.irange = 0; o.rcpf = 0.f;
    if ((dtype == FRX_U8 || dtype == FRX_U16 || dtype == FRX_I16) && !o.degenerate && o.imax - o.imin <= (dtype == FRX_I16 ? 32767 : 65535)) {
        // neither max - min nor any x - min wraps in the raster dtype: the integer form of the quantiser applies
        o.irange = (int)(o.imax - o.imin); o.rcpf = __fdiv_rz(1.0f, (float)o.irange); o.flags |= FRX_QF_INT16;
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

// ==========================================================================================
// Frame encoder, round 2: ONE WARP PER FRAME, twelve frame warps per SM.
//
// A warp owns a 4096-sample frame from ticket to copy-out.  A lane owns a contiguous block of 128 samples, so every search
// stage is a register sliding window over shared memory that is read with 16-byte loads, the level-5 Rice partitions ARE
// the lanes, and all reductions are shuffles.  Per warp the shared memory is one 16.9 KB buffer that holds, in turn and in
// place: the raw samples as TMA delivered them, the PCM, the zig-zag residuals of the winning predictor, and finally the
// packed frame (the packer reads a chunk of 128 residuals into registers before it writes the bits of that chunk, and the
// bits never overtake the reader: checked per subframe from the exact bit counts, VERBATIM otherwise).  Frame sizes go
// through the decoupled look-back as soon as the exact bit count is known; a frame whose offset is not there when it is
// packed is parked in a slot in global memory instead of waiting (see the kernel).  The warps of the wide (32-bit PCM)
// kernels run free; those of the narrow kernels walk the phases of their frames in step, in groups of ALIGN_GROUP, because
// twelve warps at twelve places of this kernel overrun the instruction cache (see the kernel).
// ==========================================================================================
#define FULLMASK 0xFFFFFFFFu
#ifdef FRX_DBG
// development counters (cycles, events) summed over all frame warps; read and cleared by frx_debug_read
__device__ unsigned long long g_frx_dbg[16];
#define DBG_ADD(i, v) do { if (lane == 0) atomicAdd(&g_frx_dbg[i], (unsigned long long)(v)); } while (0)
#define DBG_CLK() clock64()
#else
#define DBG_ADD(i, v) do { } while (0)
#define DBG_CLK() 0ll
#endif
#ifndef ENC_WARPS
#define ENC_WARPS 12
#endif
#define ENC_CTA (ENC_WARPS * 32)
#define SLACK_W 64                                   // words in front of the PCM area: the packed stream runs ahead by its headers
#define BLK 128                                      // samples per lane block of a full frame = samples per pack chunk
#define BLK_W (BLK + 4)                              // words per block: 4 words of padding keep 16-byte lane-strided loads conflict-free
#define PIDX(g) ((g) + 4 * ((g) >> 7))               // word of sample g inside the PCM area
#define PCM_W (FRX_MAX_BLOCK + 4 * (FRX_MAX_BLOCK / BLK))
#define WBUF_W (SLACK_W + PCM_W + 12)
// Phase alignment (kernels with ALIGNED = true, see frx_encode_kernel): the frame warps of a CTA are split into groups of
// ALIGN_GROUP warps that take their frames together and meet again at the phase boundaries of the frame (one named barrier
// per group).  Every warp of a group passes exactly EncArgs.bars_per_frame barriers per frame, whatever path its frame takes.
#ifndef ALIGN_GROUP
#define ALIGN_GROUP 6
#endif
static_assert(ENC_WARPS % ALIGN_GROUP == 0, "groups must tile the CTA");
__device__ __forceinline__ void group_bar() {
    asm volatile("bar.sync %0, %1;" :: "r"(1 + (int)(threadIdx.x >> 5) / ALIGN_GROUP), "r"(ALIGN_GROUP * 32) : "memory");
}
__device__ __forceinline__ bool group_or(bool pred) {
    int r;
    asm volatile("{ .reg .pred p, q; setp.ne.s32 q, %3, 0; bar.red.or.pred p, %1, %2, q; selp.s32 %0, 1, 0, p; }"
                 : "=r"(r) : "r"(1 + (int)(threadIdx.x >> 5) / ALIGN_GROUP), "r"(ALIGN_GROUP * 32), "r"((int)pred) : "memory");
    return r != 0;
}
#define ALIGN_BAR(n) do { if (ALIGNED) { group_bar(); (n)++; } } while (0)

__device__ __forceinline__ uint64_t warp_sum_u64(uint64_t v) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULLMASK, v, o);
    return v;
}
__device__ __forceinline__ float warp_sum_f32(float v) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULLMASK, v, o);
    return v;
}
__device__ __forceinline__ uint32_t zigzag(int32_t r) { return ((uint32_t)r << 1) ^ (uint32_t)(r >> 31); }
__device__ __forceinline__ uint32_t unzig_abs(uint32_t u) { return (u >> 1) + (u & 1u); }   // |r| of a zig-zag folded residual

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
// mbarrier / TMA (bulk async copy) primitives
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *b, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count) : "memory");
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
__device__ __forceinline__ void mbar_wait(uint64_t *b, uint32_t parity) {
    while (!mbar_try_wait(b, parity)) {}
}
// TMA 1-D bulk copy global -> shared, completion counted in bytes on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// ------------------------------------------------------------------------------------------
// per-warp shared state
// ------------------------------------------------------------------------------------------
struct FrameDesc {
    uint32_t f, fj;
    long long s0;                                // first sample of the frame inside its tile (row-major), 64-bit: tiles may hold >= 2^32 samples
    int32_t nsamp;
    frx_tile t;
    uint32_t hdr[4];
    uint32_t hdr_bits;
    int32_t tma;                                 // 1: rows are 16-byte multiples, raw samples staged by TMA bulk copies
    int32_t tile_idx;
    int16_t po_max, max_lpc;                     // largest Rice partition order / LPC order this frame can use
    FrxQuant q;
};
struct SubDec {                                  // decision record of one subframe
    uint8_t type, order, wasted, po, rice2, precision, sbps, vch;   // type: 0 CONSTANT 1 VERBATIM 2 FIXED 3 LPC; vch: virtual channel
    uint8_t fast, narrow, runs, pad1;            // fast: full frame of lean data (blocked 32-bit passes); narrow: 32-bit LPC accumulate;
                                                 // runs: every lane may pack its own block as one run (pack_runs)
    int32_t shift;
    uint32_t bits;                               // exact size of the subframe
    int16_t qlp[FRX_MAX_LPC];
    uint8_t k[64];
};
struct WarpSh {
    uint64_t mbar;
    FrameDesc fd;
    int32_t qlp[FRX_MAX_LPC];
    int32_t warm[FRX_MAX_LPC];
    uint8_t kparam[2][64];                       // candidate 0 = fixed, 1 = lpc; parameters of the candidate's best partition order
    uint64_t g_tot[5];                           // generic path: results of pass1_generic
    float g_ac[FRX_MAX_LPC + 1];
    uint32_t g_bad;
    SubDec dec[FRX_MAX_CH];
};
#define WARP_SH_BYTES ((sizeof(WarpSh) + 127) & ~(size_t)127)
#define WARP_BYTES (WARP_SH_BYTES + ((WBUF_W * 4 + 127) & ~(size_t)127))

// ------------------------------------------------------------------------------------------
// frame header, frame descriptor
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ int frame_header_bytes(uint8_t *hb, int nsamp, int sr_code, int sample_rate, int chan_code, int bps, uint32_t fj) {
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
    hb[n++] = (uint8_t)((chan_code << 4) | (sz_code << 1));
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
// header words of the frame for a channel-assignment code (0..7 independent, 8 left/side, 9 side/right, 10 mid/side)
__device__ __noinline__ void make_frame_header(const EncArgs &a, FrameDesc &d, int chan_code) {
    uint8_t hb[16];
    const int n = frame_header_bytes(hb, d.nsamp, a.sr_code, a.sample_rate, chan_code, a.bps, d.fj);
#pragma unroll
    for (int k = 0; k < 4; k++)
        d.hdr[k] = ((uint32_t)hb[4 * k] << 24) | ((uint32_t)hb[4 * k + 1] << 16) | ((uint32_t)hb[4 * k + 2] << 8) | hb[4 * k + 3];
    d.hdr_bits = (uint32_t)n * 8u;
}

// Everything the warp needs to know about frame f.  The lanes probe 32 tile boundaries at once and fetch the tile and its
// quantiser words side by side, so the whole descriptor costs two memory round trips.
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
        const long long s0 = (long long)fj * (long long)a.blocksize;
        const int nsamp = (int)min((long long)a.blocksize, tile_n - s0);
        d.f = f; d.fj = fj; d.s0 = s0; d.nsamp = nsamp; d.tile_idx = ti;
        {
            const int tz = __ffs(nsamp) - 1;               // trailing zeros of the block size
            int po_max = min(a.max_po, max(0, tz - 4));    // partitions stay multiples of 16 samples
            while (po_max > 0 && (nsamp >> po_max) <= FRX_MAX_LPC) po_max--;
            if (nsamp != FRX_MAX_BLOCK && po_max > 5) po_max = 5;   // short frames: the partition bins hold 32 sums
            d.po_max = (int16_t)po_max;
            d.max_lpc = (int16_t)min(a.max_lpc_order, nsamp - 1);
        }
        const int epv = 16 / frx_dtype_size(a.dtype);            // elements per 16 bytes
        d.tma = (a.tma_ok && (t.w % epv) == 0 && (t.col_off % epv) == 0 && (a.blocksize % epv) == 0) ? 1 : 0;
        make_frame_header(a, d, t.n_channels - 1);
    }
    __syncwarp();
}

// K8: decoupled look-back over the frame sizes (whole warp).  A descriptor word carries flag and value together (A: the
// frame's size, P: the inclusive byte prefix), so relaxed device-scope accesses are enough.  512 descriptors are examined
// per round trip, as 16 rows of 32 consecutive words (one coalesced 256-byte load per row, all 16 in flight together); the
// rows are then walked nearest-first with two ballots each.  When the walk meets a frame that has not published yet, only
// that single word is polled.
#define LB_ROWS 16
struct LookBack {
    long long pos;        // nearest descriptor not yet accounted for
    uint64_t exclusive;   // bytes summed so far
    bool blocked;         // desc[pos] had not been published at the last visit
};
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
    uint64_t v[LB_ROWS];
    const long long base = lb.pos - lane;
#pragma unroll
    for (int k = 0; k < LB_ROWS; k++) {
        const long long idx = base - 32 * k;
        v[k] = idx >= 0 ? frx_ld_relaxed_u64(&a.desc[idx]) : FLAG_P;      // in front of frame 0: an inclusive prefix of 0
    }
    uint64_t sum = 0;
    int result = 0;                    // 0: all 512 were sizes   1: reached an inclusive prefix   2: met an unpublished frame
    int taken = 0;
#pragma unroll
    for (int k = 0; k < LB_ROWS; k++) {
        if (result == 0) {
            const uint32_t flag = (uint32_t)(v[k] >> 62);
            const uint32_t mP = __ballot_sync(FULLMASK, flag == 2u), mN = __ballot_sync(FULLMASK, flag == 0u);
            const int firstP = mP ? (__ffs((int)mP) - 1) : 32, firstN = mN ? (__ffs((int)mN) - 1) : 32;
            const int stop = min(firstP, firstN);
            if (lane < stop || (lane == stop && firstP < firstN)) sum += v[k] & VMASK;
            if (stop < 32) { result = firstP < firstN ? 1 : 2; taken = 32 * k + stop; }
        }
    }
    lb.exclusive += warp_sum_u64(sum);
    if (result == 1) return true;
    if (result == 2) { lb.pos -= taken; lb.blocked = true; }   // the frame that has not published its size yet
    else lb.pos -= 32 * LB_ROWS;
    return false;
}

// ------------------------------------------------------------------------------------------
// K1: raw samples -> PCM in the warp buffer.  Virtual channel vch: 0..7 = band first_band + vch; for two-channel stereo
// trials 8 = mid ((L+R)>>1), 9 = side (L-R).
// ------------------------------------------------------------------------------------------
template <int DT>
__device__ __forceinline__ int32_t quant_one(typename FrxElem<DT>::T v, const FrxQuant &qz, int mode) {
    if (mode == FRX_PCM_REFERENCE) return frx_quant_ref<DT>(v, qz);
    if (mode == FRX_PCM_LOSSLESS) return frx_quant_lossless<DT>(v);
    return frx_quant_spatial<DT>(v);
}

// in-place expansion of one block: raw elements sit end-aligned in the block's first 512 bytes + 16, PCM words grow from
// its start (a 16-byte read is always ahead of the words written so far).  FULL: all 128 samples are valid.
// F32SAFE: float32 reference mode with the stream-wide guarantee that x * 8388607 fits int32 (no per-sample range check).
// INT16: 8/16-bit reference mode of a tile under FRX_QF_INT16 (frx_common.cuh).
template <int DT, bool FULL, bool F32SAFE, bool INT16 = false>
__device__ __forceinline__ void quantize_block_t(uint32_t *blk, const FrxQuant &qz, int mode, int nvalid, uint32_t &orv, uint32_t &ora) {
    typedef typename FrxElem<DT>::T T;
    constexpr int ESZ = (int)sizeof(T);
    constexpr int EPV = 16 / ESZ;                           // elements per 16-byte load
    const unsigned char *raw = reinterpret_cast<const unsigned char *>(blk) + (BLK_W * 4 - 16 - BLK * ESZ);
#pragma unroll 2
    for (int t = 0; t < BLK / EPV; t++) {
        const uint4 v = *reinterpret_cast<const uint4 *>(raw + 16 * t);
        T e[EPV];
        memcpy(e, &v, 16);
        int32_t x[EPV];
        if (INT16) {
            // integer form of the 16-bit quantiser for all the elements of the load, one (rare) branch for those whose
            // quotient is exact and needs the float64 chain
            bool exact = false;
#pragma unroll
            for (int i = 0; i < EPV; i++) x[i] = frx_quant16_int<DT>(e[i], qz, exact);
            if (exact) {
#pragma unroll
                for (int i = 0; i < EPV; i++) x[i] = frx_quant_ref_chain<DT>(e[i], qz);
            }
        }
#pragma unroll
        for (int i = 0; i < EPV; i++) {
            if (INT16) { }
            else if (F32SAFE) x[i] = __float2int_rz(__fmul_rn((float)e[i], 8388607.0f));
            else x[i] = quant_one<DT>(e[i], qz, mode);
            if (!FULL && t * EPV + i >= nvalid) x[i] = 0;
            orv |= (uint32_t)x[i];
            ora |= (uint32_t)(x[i] ^ (x[i] >> 31));
        }
#pragma unroll
        for (int i = 0; i < EPV / 4; i++)
            *reinterpret_cast<uint4 *>(blk + t * EPV + 4 * i) = make_uint4((uint32_t)x[4 * i], (uint32_t)x[4 * i + 1], (uint32_t)x[4 * i + 2], (uint32_t)x[4 * i + 3]);   // same type as the read: ordered
    }
}
template <int DT>
__device__ __forceinline__ void quantize_block(uint32_t *blk, const FrxQuant &qz, int mode, int nvalid, uint32_t &orv, uint32_t &ora) {
    if (nvalid == BLK) {
        if (DT == FRX_F32 && mode == FRX_PCM_REFERENCE && (qz.flags & FRX_QF_F32_SAFE)) quantize_block_t<DT, true, DT == FRX_F32>(blk, qz, mode, nvalid, orv, ora);
        else if ((DT == FRX_U8 || DT == FRX_I16 || DT == FRX_U16) && mode == FRX_PCM_REFERENCE && (qz.flags & FRX_QF_INT16))
            quantize_block_t<DT, true, false, DT == FRX_U8 || DT == FRX_I16 || DT == FRX_U16>(blk, qz, mode, nvalid, orv, ora);
        else quantize_block_t<DT, true, false>(blk, qz, mode, nvalid, orv, ora);
    } else quantize_block_t<DT, false, false>(blk, qz, mode, nvalid, orv, ora);
}
__device__ __noinline__ void quantize_block_any(int dtype, uint32_t *blk, const FrxQuant &qz, int mode, int nvalid, uint32_t &orv, uint32_t &ora) {
    switch (dtype) {   // the dtypes the benchmarks do not use share one out-of-line copy
        case FRX_U8: quantize_block<FRX_U8>(blk, qz, mode, nvalid, orv, ora); break;
        case FRX_I16: quantize_block<FRX_I16>(blk, qz, mode, nvalid, orv, ora); break;
        case FRX_I32: quantize_block<FRX_I32>(blk, qz, mode, nvalid, orv, ora); break;
        default: quantize_block<FRX_U32>(blk, qz, mode, nvalid, orv, ora); break;
    }
}

template <int DT>
__device__ __forceinline__ int32_t gather_one(const EncArgs &a, const FrameDesc &d, int band, long long t) {
    typedef typename FrxElem<DT>::T T;
    const long long row = t / d.t.w, col = t - row * d.t.w;
    const T *p = (const T *)a.raster + (long long)band * a.band_stride + ((long long)d.t.row_off + row) * a.row_stride + d.t.col_off + col;
    return quant_one<DT>(__ldg(p), d.q, a.pcm_mode);
}
__device__ __noinline__ int32_t gather_any(const EncArgs &a, const FrameDesc &d, int band, long long t) {
    switch (a.dtype) {
        case FRX_U8: return gather_one<FRX_U8>(a, d, band, t);
        case FRX_I16: return gather_one<FRX_I16>(a, d, band, t);
        case FRX_U16: return gather_one<FRX_U16>(a, d, band, t);
        case FRX_I32: return gather_one<FRX_I32>(a, d, band, t);
        case FRX_U32: return gather_one<FRX_U32>(a, d, band, t);
        default: return gather_one<FRX_F32>(a, d, band, t);
    }
}

// Mid / side of a two-channel frame, aligned rows: every lane reads the 128 samples of its block from both band planes with
// 16-byte loads, quantises both and stores mid or side as PCM (no raw staging: the buffer holds one channel at a time)
template <int DT>
__device__ __forceinline__ void load_ms_block(const EncArgs &a, const FrameDesc &d, uint32_t *blk, int lane, int vch, int nvalid, uint32_t &orv, uint32_t &ora) {
    typedef typename FrxElem<DT>::T T;
    constexpr int EPV = 16 / (int)sizeof(T);
    const long long w = d.t.w;
    const long long t0 = d.s0 + (long long)lane * BLK;
    long long row = t0 / w, col = t0 - row * w;
    const T *L = (const T *)a.raster + (long long)d.t.first_band * a.band_stride + (long long)d.t.row_off * a.row_stride + d.t.col_off;
    const T *R = L + a.band_stride;
#pragma unroll 1
    for (int i = 0; i < BLK; i += EPV) {
        int32_t x[EPV];
        if (i < nvalid) {
            const long long e = row * a.row_stride + col;
            const uint4 vl = __ldg(reinterpret_cast<const uint4 *>(L + e)), vr = __ldg(reinterpret_cast<const uint4 *>(R + e));
            T el[EPV], er[EPV];
            memcpy(el, &vl, 16); memcpy(er, &vr, 16);
            int32_t ql[EPV], qr[EPV];
            if ((DT == FRX_U8 || DT == FRX_I16 || DT == FRX_U16) && a.pcm_mode == FRX_PCM_REFERENCE && (d.q.flags & FRX_QF_INT16)) {
                bool exact = false;
#pragma unroll
                for (int j = 0; j < EPV; j++) { ql[j] = frx_quant16_int<DT>(el[j], d.q, exact); qr[j] = frx_quant16_int<DT>(er[j], d.q, exact); }
                if (exact) {
#pragma unroll
                    for (int j = 0; j < EPV; j++) { ql[j] = frx_quant_ref_chain<DT>(el[j], d.q); qr[j] = frx_quant_ref_chain<DT>(er[j], d.q); }
                }
            } else {
#pragma unroll
                for (int j = 0; j < EPV; j++) { ql[j] = quant_one<DT>(el[j], d.q, a.pcm_mode); qr[j] = quant_one<DT>(er[j], d.q, a.pcm_mode); }
            }
#pragma unroll
            for (int j = 0; j < EPV; j++) {
                const int32_t l = ql[j], r = qr[j];
                x[j] = vch == 8 ? ((l + r) >> 1) : (l - r);
                if (i + j >= nvalid) x[j] = 0;
                orv |= (uint32_t)x[j];
                ora |= (uint32_t)(x[j] ^ (x[j] >> 31));
            }
            col += EPV;
            if (col >= w) { col = 0; row++; }
        } else {
#pragma unroll
            for (int j = 0; j < EPV; j++) x[j] = 0;
        }
#pragma unroll
        for (int j = 0; j < EPV / 4; j++)
            *reinterpret_cast<uint4 *>(blk + i + 4 * j) = make_uint4((uint32_t)x[4 * j], (uint32_t)x[4 * j + 1], (uint32_t)x[4 * j + 2], (uint32_t)x[4 * j + 3]);
    }
}
__device__ __noinline__ void load_ms_block_any(const EncArgs &a, const FrameDesc &d, uint32_t *blk, int lane, int vch, int nvalid, uint32_t &orv, uint32_t &ora) {
    switch (a.dtype) {
        case FRX_U8: load_ms_block<FRX_U8>(a, d, blk, lane, vch, nvalid, orv, ora); break;
        case FRX_I16: load_ms_block<FRX_I16>(a, d, blk, lane, vch, nvalid, orv, ora); break;
        case FRX_U16: load_ms_block<FRX_U16>(a, d, blk, lane, vch, nvalid, orv, ora); break;
        case FRX_I32: load_ms_block<FRX_I32>(a, d, blk, lane, vch, nvalid, orv, ora); break;
        case FRX_U32: load_ms_block<FRX_U32>(a, d, blk, lane, vch, nvalid, orv, ora); break;
        default: load_ms_block<FRX_F32>(a, d, blk, lane, vch, nvalid, orv, ora); break;
    }
}

// Stage virtual channel `vch` of the frame as PCM in buf[SLACK_W + PIDX(g)].  Returns (OR of the samples, OR of their
// magnitudes), warp-uniform.  The samples behind a short frame are zero up to the next block boundary + one block.
__device__ __noinline__ uint2 load_channel(const EncArgs &a, WarpSh &ws, uint32_t *buf, int vch, uint32_t &phase, int lane) {
    const FrameDesc &d = ws.fd;
    const int nsamp = d.nsamp;
    uint32_t *pcm = buf + SLACK_W;
    uint32_t orv = 0, ora = 0;
    const int nblk = (nsamp + BLK - 1) / BLK;
    if (d.tma && vch < 8) {
        const int esz = frx_dtype_size(a.dtype);
        const long long w = d.t.w;
        const unsigned char *band = (const unsigned char *)a.raster +
            ((long long)(d.t.first_band + vch) * a.band_stride + (long long)d.t.row_off * a.row_stride + d.t.col_off) * esz;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // earlier generic accesses to the buffer before the async writes
        __syncwarp();
        if (lane == 0) mbar_arrive_expect_tx(&ws.mbar, (uint32_t)nsamp * (uint32_t)esz);
        __syncwarp();
        if (lane < nblk) {
            const long long g0 = (long long)lane * BLK, g1 = min((long long)nsamp, g0 + BLK);
            long long t0 = d.s0 + g0;
            const long long t1 = d.s0 + g1;
            long long row = t0 / w;
            unsigned char *dst0 = reinterpret_cast<unsigned char *>(pcm + lane * BLK_W) + (BLK_W * 4 - 16 - BLK * esz);
            while (t0 < t1) {
                const long long rend = min(t1, (row + 1) * w);
                bulk_g2s(dst0 + (t0 - d.s0 - g0) * esz, band + (row * a.row_stride + (t0 - row * w)) * esz, (uint32_t)((rend - t0) * esz), &ws.mbar);
                t0 = rend; row++;
            }
        }
        mbar_wait(&ws.mbar, phase);
        phase ^= 1u;
        if (lane < nblk) {
            const int nvalid = min(BLK, nsamp - lane * BLK);
            uint32_t *blk = pcm + lane * BLK_W;
            if (a.dtype == FRX_F32) quantize_block<FRX_F32>(blk, d.q, a.pcm_mode, nvalid, orv, ora);
            else if (a.dtype == FRX_U16) quantize_block<FRX_U16>(blk, d.q, a.pcm_mode, nvalid, orv, ora);
            else quantize_block_any(a.dtype, blk, d.q, a.pcm_mode, nvalid, orv, ora);
        } else if (lane == nblk) {
            uint32_t *blk = pcm + lane * BLK_W;
            for (int i = 0; i < BLK; i += 4) *reinterpret_cast<int4 *>(blk + i) = make_int4(0, 0, 0, 0);
        }
    } else if (d.tma) {
        // mid / side (vch >= 8) of aligned rows
        uint32_t o1 = 0, o2 = 0;
        if (lane <= nblk && lane < 32) load_ms_block_any(a, d, pcm + lane * BLK_W, lane, vch, max(0, min(BLK, nsamp - lane * BLK)), o1, o2);
        orv = o1; ora = o2;
    } else {
        const int b0 = d.t.first_band;
        for (int g = lane; g < nsamp; g += 32) {
            int32_t x;
            if (vch < 8) x = gather_any(a, d, b0 + vch, d.s0 + g);
            else {
                const int32_t l = gather_any(a, d, b0, d.s0 + g), r = gather_any(a, d, b0 + 1, d.s0 + g);
                x = vch == 8 ? ((l + r) >> 1) : (l - r);
            }
            pcm[PIDX(g)] = (uint32_t)x;
            orv |= (uint32_t)x;
            ora |= (uint32_t)(x ^ (x >> 31));
        }
        const int gend = min(FRX_MAX_BLOCK, nblk * BLK + BLK);
        for (int g = nsamp + lane; g < gend; g += 32) pcm[PIDX(g)] = 0u;
    }
    orv = __reduce_or_sync(FULLMASK, orv);
    ora = __reduce_or_sync(FULLMASK, ora);
    __syncwarp();
    return make_uint2(orv, ora);
}

// ------------------------------------------------------------------------------------------
// Rice search of one candidate from its partition sums.  F = the level whose partitions the lanes hold one each (5 for
// full frames): lane p holds the sum of level-F partition p in sumA + sumB; for po_max == 6 sumA / sumB are the two
// level-6 partitions 2p, 2p+1.  Coarser levels are built with shuffles; levels above po_max are not evaluated.  Writes the
// best order's parameters to kparam[] and returns (estimated bits incl. the 6-bit method/order field, best order, RICE2).
// ------------------------------------------------------------------------------------------
struct RiceChoice { uint64_t bits; int po; bool rice2; };
// sum of per-lane bit estimates: one redux when they are small (always, unless a partition is hopeless), exact otherwise
__device__ __forceinline__ uint64_t warp_sum_bits(uint64_t b) {
    if (__all_sync(FULLMASK, b < (1ull << 26))) return (uint64_t)__reduce_add_sync(FULLMASK, (uint32_t)b);
    return warp_sum_u64(b);
}
__device__ __noinline__ RiceChoice analyse_partitions(uint8_t *kparam, uint64_t sumA, uint64_t sumB, int nsamp, int pred_order, int po_max,
                                                      int F, uint32_t kmax) {
    const int lane = threadIdx.x & 31;
    uint64_t sum = sumA + sumB;
    uint64_t best = ~0ull;
    int best_po = 0;
    uint32_t kbest0 = 0, kbest1 = 0;
    if (po_max == 6) {
        const uint32_t n0 = (uint32_t)(nsamp >> 6) - (lane == 0 ? (uint32_t)pred_order : 0u), n1 = (uint32_t)(nsamp >> 6);
        const uint32_t k0 = rice_param(n0, sumA, kmax), k1 = rice_param(n1, sumB, kmax);
        uint64_t b = rice_estimate(k0, n0, sumA) + rice_estimate(k1, n1, sumB);
        b = warp_sum_bits(b) + 6;
        best = b; best_po = 6; kbest0 = k0; kbest1 = k1;
    }
    for (int L = F; L >= 0; L--) {
        const int step = 1 << (F - L);                   // lanes that are multiples of step hold level-L partitions
        const bool mine = (lane & (step - 1)) == 0 && lane < (1 << F);
        const int p = lane >> (F - L);
        const uint32_t n = (uint32_t)(nsamp >> L) - (p == 0 ? (uint32_t)pred_order : 0u);
        uint32_t k = 0;
        uint64_t b = 0;
        if (L <= po_max) {
            if (mine) { k = rice_param(n, sum, kmax); b = rice_estimate(k, n, sum); }
            b = warp_sum_bits(b) + 6;
            if (b < best) { best = b; best_po = L; kbest0 = k; }
        }
        if (L > 0) {
            const uint64_t o = __shfl_down_sync(FULLMASK, sum, step);
            sum += o;                                    // valid in lanes that are multiples of 2*step
        }
    }
    if (best_po == 6) { kparam[2 * lane] = (uint8_t)kbest0; kparam[2 * lane + 1] = (uint8_t)kbest1; }
    else {
        const int step = 1 << (F - best_po);
        if ((lane & (step - 1)) == 0 && lane < (1 << F)) kparam[lane >> (F - best_po)] = (uint8_t)kbest0;
    }
    const uint32_t big = __ballot_sync(FULLMASK, kbest0 >= 15u || (best_po == 6 && kbest1 >= 15u));
    __syncwarp();
    RiceChoice rc;
    rc.bits = best; rc.po = best_po; rc.rice2 = big != 0u;
    return rc;
}

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

// ==========================================================================================
// K7: bit packing.  Every lane appends a run of fields to the stream in lane order; a field is `zer` zero bits followed by
// the low `len` bits of `val` (len <= 32).  Lanes accumulate their bits in registers and store whole words; the words two
// lanes share are merged with a segmented OR-scan, so shared memory sees plain stores only.
// ==========================================================================================
struct Emitter {
    uint32_t P;        // bit position of the end of the stream inside the warp buffer
    uint32_t carry;    // bits of the incomplete last word (top-aligned)
};
struct LaneBits {
    uint32_t hi, lo, firstw;
    int nb;
    uint32_t *wp;
    bool flushed;
    __device__ __forceinline__ void init(uint32_t *out, uint32_t bitpos) {
        wp = out + (bitpos >> 5); nb = (int)(bitpos & 31u); hi = 0; lo = 0; firstw = 0; flushed = false;
    }
    __device__ __forceinline__ void flush() {
        if (!flushed) { firstw = hi; flushed = true; } else *wp = hi;
        wp++; hi = lo; lo = 0; nb -= 32;
    }
    __device__ __forceinline__ void put(uint32_t c, int n) {   // n in 1..32, c < 2^n
        const uint32_t ct = c << (32 - n);
        hi |= ct >> nb;
        lo = __funnelshift_r(0u, ct, (uint32_t)nb);            // ct << (32 - nb), 0 when nb == 0
        nb += n;
        if (nb >= 32) flush();
    }
    __device__ __forceinline__ void zeros(uint32_t q) {
        nb += (int)q;
        while (nb >= 32) flush();
    }
};
template <int NF>
__device__ __forceinline__ void emit_fields(uint32_t *out, Emitter &em, const uint32_t (&val)[NF], const uint32_t (&len)[NF], const uint32_t (&zer)[NF], int lane) {
    uint32_t T = 0;
#pragma unroll
    for (int i = 0; i < NF; i++) T += len[i] + zer[i];
    uint32_t incl = T;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t n = __shfl_up_sync(FULLMASK, incl, o);
        if (lane >= o) incl += n;
    }
    const uint32_t total = __shfl_sync(FULLMASK, incl, 31);
    const uint32_t start = em.P + incl - T;
    LaneBits lb;
    lb.init(out, start);
#pragma unroll
    for (int i = 0; i < NF; i++) {
        const uint32_t n = len[i] + zer[i];
        if (n) {
            if (n <= 32u) lb.put(val[i], (int)n);
            else { lb.zeros(zer[i]); if (len[i]) lb.put(val[i], (int)len[i]); }
        }
    }
    // the incomplete last word of every lane flows to the lanes behind it until one of them completes the word
    uint32_t v = lb.hi;
    uint32_t fl = lb.flushed ? 1u : 0u;
    if (lane == 0 && !lb.flushed) v |= em.carry;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t pv = __shfl_up_sync(FULLMASK, v, o);
        const uint32_t pf = __shfl_up_sync(FULLMASK, fl, o);
        if (lane >= o && !fl) { v |= pv; fl = pf; }
    }
    uint32_t cin = __shfl_up_sync(FULLMASK, v, 1);
    if (lane == 0) cin = em.carry;
    if (lb.flushed) out[start >> 5] = lb.firstw | cin;
    em.carry = __shfl_sync(FULLMASK, v, 31);
    em.P += total;
    __syncwarp();
}

// subframe header: 0 | type(6) | wasted flag, wasted-bits count in unary, warm-up samples, LPC precision / shift /
// coefficients, Rice method + partition order -- one field per lane, in stream order
__device__ __noinline__ Emitter pack_subframe_header(WarpSh &ws, uint32_t *out, Emitter em, const SubDec &d, const uint32_t *pcm, int lane) {
    const int type = d.type, order = d.order, wasted = d.wasted, sbps = d.sbps;
    uint32_t val[2] = { 0, 0 }, len[2] = { 0, 0 }, zer[2] = { 0, 0 };
    const uint32_t smask = sbps >= 32 ? 0xFFFFFFFFu : ((1u << sbps) - 1u);
    if (lane == 0) {
        const uint32_t tcode = type == 0 ? 0u : (type == 1 ? 1u : (type == 2 ? (8u | (uint32_t)order) : (32u | (uint32_t)(order - 1))));
        val[0] = (tcode << 1) | (wasted ? 1u : 0u); len[0] = 8;
        if (wasted) { val[1] = 1u; len[1] = 1; zer[1] = (uint32_t)wasted - 1u; }
    }
    if (type == 0) {
        if (lane == 1) { val[0] = pcm[PIDX(0)] & smask; len[0] = (uint32_t)sbps; }
    } else if (type >= 2) {
        if (lane >= 1 && lane <= order) { val[0] = (uint32_t)ws.warm[lane - 1] & smask; len[0] = (uint32_t)sbps; }
        int nxt = order + 1;
        if (type == 3) {
            const uint32_t precision = d.precision;
            if (lane == nxt) { val[0] = ((precision - 1u) << 5) | ((uint32_t)d.shift & 31u); len[0] = 9; }
            if (lane > nxt && lane <= nxt + order) { val[0] = (uint32_t)(int32_t)d.qlp[lane - nxt - 1] & ((1u << precision) - 1u); len[0] = precision; }
            nxt += order + 1;
        }
        if (lane == nxt) { val[0] = ((d.rice2 ? 1u : 0u) << 4) | (uint32_t)d.po; len[0] = 6; }
    }
    emit_fields<2>(out, em, val, len, zer, lane);
    return em;
}

// Rice-coded residual section, chunk by chunk: 128 residuals into registers, then their bits into the words behind them.
// Fast path (whole warp): the lane's four codes (and a partition parameter in front of them) are shift-appended into one
// 128-bit string, the string is moved so that its END sits on the word grid of the stream, and the words come out right to
// left: X0 = the incomplete last word (handed to the lanes behind), X1..Xn = the words this lane completes.  A chunk in which
// some code is longer than 32 bits, or a lane's bits do not fit the string, takes emit_fields.
__device__ __forceinline__ void append128(uint32_t &w3, uint32_t &w2, uint32_t &w1, uint32_t &w0, uint32_t v, uint32_t n) {   // n <= 32
    w3 = __funnelshift_lc(w2, w3, n); w2 = __funnelshift_lc(w1, w2, n); w1 = __funnelshift_lc(w0, w1, n);
    w0 = __funnelshift_lc(0u, w0, n) | v;
}
__device__ __noinline__ Emitter pack_rice(uint32_t *out, Emitter em, const SubDec &d, uint32_t *pcm, int nsamp, int lane) {
    const int order = d.order, po = d.po;
    const int psize = nsamp >> po;
    const uint32_t pb = d.rice2 ? 5u : 4u;
    const bool pow2 = (psize & (psize - 1)) == 0;
    const int psh = 31 - __clz(psize);
    const int nchunks = (nsamp + BLK - 1) / BLK;
    for (int c = 0; c < nchunks; c++) {
        const int g = BLK * c + 4 * lane;
        const uint4 v = *reinterpret_cast<const uint4 *>(pcm + PIDX(g));
        __syncwarp();
        const uint32_t u[4] = { v.x, v.y, v.z, v.w };
        const bool allcoded = (c > 0 || order == 0) && BLK * (c + 1) <= nsamp;      // warp-uniform
        int pi;
        if (pow2) pi = g >> psh;
        else pi = g < nsamp ? g / psize : 0;
        const bool inside = allcoded || g < nsamp;
        const uint32_t k = inside ? d.k[pi] : 0u;
        const bool pstart = inside && g == pi * psize;
        const uint32_t kmask = (1u << k) - 1u, kbit = 1u << k, k1 = k + 1u;
        uint32_t q[4], cv[4], cl[4];
        if (allcoded) {
#pragma unroll
            for (int i = 0; i < 4; i++) { q[i] = u[i] >> k; cv[i] = (u[i] & kmask) | kbit; cl[i] = q[i] + k1; }
        } else {
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const bool coded = (g + i) < nsamp && (g + i) >= order;
                q[i] = u[i] >> k;
                cv[i] = coded ? ((u[i] & kmask) | kbit) : 0u;
                cl[i] = coded ? q[i] + k1 : 0u;                   // code length: zeros, stop bit, k low bits
            }
        }
        const uint32_t p0 = pstart ? pb : 0u;
        const uint32_t T = p0 + (cl[0] + cl[1]) + (cl[2] + cl[3]);
        uint32_t incl = T;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t n = __shfl_up_sync(FULLMASK, incl, o);
            if (lane >= o) incl += n;
        }
        const uint32_t start = em.P + incl - T, sh = start & 31u;
        const bool fits = ((cl[0] | cl[1] | cl[2] | cl[3]) <= 32u) && (sh + T <= 128u);   // (the OR bounds the largest length from above)
        if (__all_sync(FULLMASK, fits)) {
            const uint32_t total = __shfl_sync(FULLMASK, incl, 31);
            uint32_t w3 = 0, w2 = 0, w1 = 0, w0 = pstart ? k : 0u;
            append128(w3, w2, w1, w0, cv[0], cl[0]);
            append128(w3, w2, w1, w0, cv[1], cl[1]);
            append128(w3, w2, w1, w0, cv[2], cl[2]);
            append128(w3, w2, w1, w0, cv[3], cl[3]);
            const uint32_t endb = sh + T;                          // bits from the lane's first word boundary to the end of its string
            const uint32_t a = 32u - (endb & 31u);                 // 1..32: move the end of the string onto the word grid
            const uint32_t X4 = __funnelshift_lc(w3, 0u, a), X3 = __funnelshift_lc(w2, w3, a), X2 = __funnelshift_lc(w1, w2, a),
                           X1 = __funnelshift_lc(w0, w1, a), X0 = __funnelshift_lc(0u, w0, a);
            const uint32_t nfull = endb >> 5;                      // words this lane completes (0..4): X_nfull (first) .. X1 (last)
            uint32_t vv = X0, fl = nfull ? 1u : 0u;
            if (lane == 0 && !fl) vv |= em.carry;
            if (!__all_sync(FULLMASK, fl)) {                      // some lane completes no word: its bits flow on (segmented OR scan)
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t pv = __shfl_up_sync(FULLMASK, vv, o);
                    const uint32_t pf = __shfl_up_sync(FULLMASK, fl, o);
                    if (lane >= o && !fl) { vv |= pv; fl = pf; }
                }
            }
            uint32_t cin = __shfl_up_sync(FULLMASK, vv, 1);
            if (lane == 0) cin = em.carry;
            uint32_t *wp = out + (start >> 5) + nfull;             // one past the last word this lane completes
            const uint32_t first = nfull == 1u ? X1 : (nfull == 2u ? X2 : (nfull == 3u ? X3 : X4));
            if (nfull >= 2u) wp[-1] = X1;
            if (nfull >= 3u) wp[-2] = X2;
            if (nfull >= 4u) wp[-3] = X3;
            if (nfull >= 1u) out[start >> 5] = first | cin;
            em.carry = __shfl_sync(FULLMASK, vv, 31);
            em.P += total;
            __syncwarp();
        } else {
            uint32_t val[5], len[5], zer[5];
            val[0] = k; len[0] = p0; zer[0] = 0;
#pragma unroll
            for (int i = 0; i < 4; i++) { val[1 + i] = cv[i]; len[1 + i] = cl[i] ? k1 : 0u; zer[1 + i] = cl[i] ? q[i] : 0u; }
            emit_fields<5>(out, em, val, len, zer, lane);
        }
    }
    return em;
}
// Rice-coded residual section of a full frame, one RUN per lane: lane L codes the 128 residuals of its own block with a serial
// bit writer (64-bit accumulator, no shuffles) and stores the words over the block it is reading, starting four words in front
// of it (lane 0: at the end of the stream); count_block has checked that a run never overtakes its reads.  The words of a run
// lie on the word grid of the stream and hold the lane's own bits only, so the runs are then moved down to their place one
// after the other (the global in-place condition keeps a destination below every unread source) and the word two runs share
// is OR-ed.
struct RunWriter {
    uint32_t hi, lo, nb;                                                 // 64-bit accumulator, nb bits used (< 64 between samples)
    uint32_t wa;                                                         // shared-memory byte address of the next word
    // store the completed word when there is one: predicated, no branch (lanes complete their words at different samples)
    __device__ __forceinline__ void word_if_full() {
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ge.u32 p, %2, 32;\n\t@p st.shared.u32 [%0], %1;\n\t}" ::"r"(wa), "r"(hi), "r"(nb) : "memory");
        const bool f = nb >= 32u;
        wa += f ? 4u : 0u; hi = f ? lo : hi; lo = f ? 0u : lo; nb -= f ? 32u : 0u;
    }
    // append n <= 32 bits; on entry nb + n <= 64
    __device__ __forceinline__ void bits(uint32_t v, uint32_t n) {
        const uint64_t x = (uint64_t)v << (64u - n - nb);
        hi |= (uint32_t)(x >> 32); lo |= (uint32_t)x;
        nb += n;
    }
    __device__ __forceinline__ void put(uint32_t v, uint32_t n) { word_if_full(); bits(v, n); }     // nb < 64 on entry
    // one Rice code: at most one word leaves per sample; a run of zeros that would not fit takes the loop (rare)
    __device__ __forceinline__ void code(uint32_t u, uint32_t k, uint32_t kmask, uint32_t kbit, uint32_t k1) {
        word_if_full();                                                 // nb < 32
        nb += u >> k;                                                   // the unary zeros
        if (nb + k1 > 63u) { while (nb >= 32u) word_if_full(); }
        bits((u & kmask) | kbit, k1);                                   // nb <= 63
    }
    __device__ __forceinline__ void finish() { word_if_full(); }        // nb < 32: the incomplete last word is in hi
};
__device__ __noinline__ Emitter pack_runs(uint32_t *out, Emitter em, const SubDec &d, uint32_t *pcm, int lane) {
    const int order = d.order, po = d.po;
    const uint32_t pb = d.rice2 ? 5u : 4u;
    const int psh = 12 - po, g0 = lane * BLK;
    uint32_t *blk = pcm + lane * BLK_W;
    const uint32_t k0 = d.k[g0 >> psh], k1 = d.k[(g0 + BLK / 2) >> psh];
    const uint32_t pb0 = (g0 & ((1 << psh) - 1)) == 0 ? pb : 0u, pb1 = po == 6 ? pb : 0u;
    bool ok;
    uint32_t B = count_block(pcm, lane, k0, k1, pb0, pb1, ok);
    if (lane == 0) B -= (uint32_t)order * (k0 + 1u);
    uint32_t incl = B;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t n = __shfl_up_sync(FULLMASK, incl, o);
        if (lane >= o) incl += n;
    }
    const uint32_t total = __shfl_sync(FULLMASK, incl, 31);
    const uint32_t start = em.P + incl - B;
    const uint32_t gw = start >> 5;
    uint32_t *sp = lane == 0 ? out + gw : blk - 4;
    RunWriter w;
    w.hi = 0u; w.lo = 0u; w.nb = start & 31u; w.wa = smem_u32(sp);
    {
        uint32_t k = k0, kmask = (1u << k0) - 1u, kbit = 1u << k0, kk1 = k0 + 1u;
        if (pb0) w.put(k0, pb);
        // lane 0 skips the warm-up samples (their slots hold 0, they carry no code): only the first groups need the test
        const int skip = lane == 0 ? order : 0;
#pragma unroll 1
        for (int it = 0; it < BLK / 4; it++) {
            if (it == BLK / 8) {
                k = k1; kmask = (1u << k1) - 1u; kbit = 1u << k1; kk1 = k1 + 1u;
                if (pb1) w.put(k1, pb);
            }
            const uint4 v = *reinterpret_cast<const uint4 *>(blk + 4 * it);
            const uint32_t u[4] = { v.x, v.y, v.z, v.w };
            if (4 * it >= skip) {
#pragma unroll
                for (int i = 0; i < 4; i++) w.code(u[i], k, kmask, kbit, kk1);
            } else {
#pragma unroll
                for (int i = 0; i < 4; i++)
                    if (4 * it + i >= skip) w.code(u[i], k, kmask, kbit, kk1);
            }
        }
        w.finish();
    }
    const uint32_t nwhole = (w.wa - smem_u32(sp)) >> 2;
    if (w.nb) sp[nwhole] = w.hi;                             // the incomplete last word, own bits only
    const uint32_t m = nwhole + (w.nb ? 1u : 0u);
    __syncwarp();
    if (lane == 0 && (em.P & 31u)) out[gw] |= em.carry;      // the bits in front of the first run
    __syncwarp();
    // move runs 1..31 down to their place
#pragma unroll 1
    for (int L = 1; L < 32; L++) {
        const uint32_t gwL = __shfl_sync(FULLMASK, gw, L), mL = __shfl_sync(FULLMASK, m, L), shared = __shfl_sync(FULLMASK, start & 31u, L);
        const uint32_t *src = pcm + L * BLK_W - 4;
        uint32_t *dst = out + gwL;
        // (a run of a full block has 4..132 words: two batches of 32 cover the usual ones)
        uint32_t x0 = lane < mL ? src[lane] : 0u, x1 = lane + 32u < mL ? src[lane + 32u] : 0u;
        if (lane == 0 && shared) x0 |= dst[0];
        __syncwarp();
        if (lane < mL) dst[lane] = x0;
        if (lane + 32u < mL) dst[lane + 32u] = x1;
        for (uint32_t t = lane + 64u; t < mL; t += 32) {
            const uint32_t x = src[t];
            __syncwarp(__activemask());
            dst[t] = x;
        }
        __syncwarp();
    }
    em.P += total;
    em.carry = (em.P & 31u) ? out[em.P >> 5] : 0u;
    __syncwarp();
    return em;
}

// VERBATIM body: the samples as sbps-bit fields
__device__ __noinline__ Emitter pack_verbatim(uint32_t *out, Emitter em, const SubDec &d, uint32_t *pcm, int nsamp, int lane) {
    const uint32_t sbps = d.sbps;
    const int nchunks = (nsamp + BLK - 1) / BLK;
    for (int c = 0; c < nchunks; c++) {
        const int g = BLK * c + 4 * lane;
        const uint4 v = *reinterpret_cast<const uint4 *>(pcm + PIDX(g));
        __syncwarp();
        const uint32_t u[4] = { v.x, v.y, v.z, v.w };
        const uint32_t smask = sbps >= 32u ? 0xFFFFFFFFu : ((1u << sbps) - 1u);   // sbps <= 32: a side channel is only tried for 16-bit PCM
        uint32_t val[4], len[4], zer[4];
#pragma unroll
        for (int i = 0; i < 4; i++) { val[i] = u[i] & smask; len[i] = (g + i) < nsamp ? sbps : 0u; zer[i] = 0; }
        emit_fields<4>(out, em, val, len, zer, lane);
    }
    return em;
}

// ------------------------------------------------------------------------------------------
// CRC-16 of `nw` whole words of the warp buffer (big-endian words): every lane takes a run of Lq 16-byte quads through the
// 4x256 table, the runs are combined with x^(128 Lq j) mod P.  Warp-uniform result.
// ------------------------------------------------------------------------------------------
__device__ __noinline__ uint32_t crc_words(const EncArgs &a, const uint16_t (*tab)[256], const uint32_t *bits, uint32_t nw, int lane) {
    if (nw == 0) return 0;
    const uint32_t nq = (nw + 3u) >> 2;
    const uint32_t Lq = (nq + 31u) / 32u;
    const uint32_t shift_factor = (uint32_t)__ldg(&a.crc_xpow[(Lq - 1) * 256 + (31 - lane)]);
    const int padq = (int)(32u * Lq) - (int)nq;                // virtual leading zero quads
    int q0 = lane * (int)Lq - padq;
    const int q1 = q0 + (int)Lq;
    if (q0 < 0) q0 = 0;
    uint32_t c = 0;
    for (int qi = q0; qi < q1; qi++) {
        uint4 wv = *reinterpret_cast<const uint4 *>(bits + 4 * qi);
        if (qi == (int)nq - 1) {
            const uint32_t k = nw - 4u * (uint32_t)qi;         // whole words in the last quad (1..4)
            if (k < 4u) wv.w = 0u;
            if (k < 3u) wv.z = 0u;
            if (k < 2u) wv.y = 0u;
        }
        uint32_t x = wv.x ^ (c << 16);
        c = (uint32_t)tab[3][x >> 24] ^ tab[2][(x >> 16) & 0xFFu] ^ tab[1][(x >> 8) & 0xFFu] ^ tab[0][x & 0xFFu];
        x = wv.y ^ (c << 16);
        c = (uint32_t)tab[3][x >> 24] ^ tab[2][(x >> 16) & 0xFFu] ^ tab[1][(x >> 8) & 0xFFu] ^ tab[0][x & 0xFFu];
        x = wv.z ^ (c << 16);
        c = (uint32_t)tab[3][x >> 24] ^ tab[2][(x >> 16) & 0xFFu] ^ tab[1][(x >> 8) & 0xFFu] ^ tab[0][x & 0xFFu];
        x = wv.w ^ (c << 16);
        c = (uint32_t)tab[3][x >> 24] ^ tab[2][(x >> 16) & 0xFFu] ^ tab[1][(x >> 8) & 0xFFu] ^ tab[0][x & 0xFFu];
    }
    if (c) c = frx_gf16_mulmod(c, shift_factor);
    c = __reduce_xor_sync(FULLMASK, c);
    const uint32_t z = 4u * nq - nw;                           // zero words appended to the last quad
    if (z && c) c = frx_gf16_mulmod(c, (uint32_t)__ldg(&a.crc_xinv[z]));
    return c;
}
// x^(32 n) mod P, n < 32768, from the CRC chunk tables: n = 128 j1 + 4 j0 + b, x^(32 n) = x^(128 * 32 * j1) * x^(128 * j0) * x^(32 b)
__device__ __noinline__ uint32_t gf16_xpow32(const EncArgs &a, uint32_t n) {
    const uint32_t j1 = (n >> 7) & 255u, j0 = (n >> 2) & 31u, b = n & 3u;
    uint32_t r = (uint32_t)__ldg(&a.crc_xpow[31 * 256 + j1]);                    // row Lq = 32
    if (j0) r = frx_gf16_mulmod(r, (uint32_t)__ldg(&a.crc_xpow[j0]));           // row Lq = 1
    if (b) r = frx_gf16_mulmod(r, b == 1u ? 0x8017u : (b == 2u ? 0x8113u : 0x9543u));   // x^32, x^64, x^96
    return r;
}
// words [0, nw) of the warp buffer -> dst (any byte alignment), big-endian byte order
__device__ __forceinline__ void copy_words_out(uint8_t *dst, const uint32_t *bits, uint32_t nw, int lane) {
    const uint32_t nbytes = nw * 4u;
    const uint32_t head = min(nbytes, (uint32_t)((4u - (uint32_t)((uintptr_t)dst & 3u)) & 3u));
    if (lane < (int)head) dst[lane] = (uint8_t)((bits[lane >> 2] >> (24 - 8 * (lane & 3))) & 0xFFu);
    const uint32_t nwords = (nbytes - head) >> 2;
    uint32_t *dw = reinterpret_cast<uint32_t *>(dst + head);
    const uint32_t sh8 = head * 8u;
    for (uint32_t j = lane; j < nwords; j += 32) {
        const uint32_t w0 = bits[j], w1 = bits[j + 1];
        const uint32_t v = head ? __funnelshift_l(w1, w0, sh8) : w0;
        dw[j] = frx_bswap32(v);
    }
    const uint32_t tail0 = head + (nwords << 2);
    if (lane < (int)(nbytes - tail0)) {
        const uint32_t b = tail0 + (uint32_t)lane;
        dst[b] = (uint8_t)((bits[b >> 2] >> (24 - 8 * (b & 3))) & 0xFFu);
    }
}
// a finished frame from the warp's slot (16-byte aligned, global memory, written by this warp) to its final place (any byte
// alignment): aligned 32-bit stores, every destination word funnel-shifted out of two slot words
__device__ __noinline__ void copy_slot_out(uint8_t *dst, const uint8_t *slot, uint32_t nbytes, int lane) {
    __syncwarp();
    const uint32_t head = min(nbytes, (uint32_t)((4u - (uint32_t)((uintptr_t)dst & 3u)) & 3u));
    if (lane < (int)head) dst[lane] = __ldcg(slot + lane);
    const uint32_t nwords = (nbytes - head) >> 2;
    uint32_t *dw = reinterpret_cast<uint32_t *>(dst + head);
    const uint32_t *sw = reinterpret_cast<const uint32_t *>(slot);
    const uint32_t sh8 = head * 8u;
    uint32_t j = lane;
    for (; j + 7u * 32u < nwords; j += 8u * 32u) {            // eight independent L2 round trips in flight per lane
        uint32_t w0[8], w1[8];
#pragma unroll
        for (int k = 0; k < 8; k++) { w0[k] = __ldcg(sw + j + 32 * k); w1[k] = head ? __ldcg(sw + j + 32 * k + 1) : 0u; }
#pragma unroll
        for (int k = 0; k < 8; k++) dw[j + 32 * k] = head ? __funnelshift_r(w0[k], w1[k], sh8) : w0[k];
    }
    for (; j < nwords; j += 32) {
        const uint32_t w0 = __ldcg(sw + j), w1 = head ? __ldcg(sw + j + 1) : 0u;
        dw[j] = head ? __funnelshift_r(w0, w1, sh8) : w0;      // little-endian words: bytes [head + 4j, head + 4j + 4)
    }
    const uint32_t tail0 = head + (nwords << 2);
    if (lane < (int)(nbytes - tail0)) dst[tail0 + lane] = __ldcg(slot + tail0 + lane);
}

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

// ==========================================================================================
// The frame encoder.  WIDE: PCM wider than 16 bits (RICE2 parameters, 64-bit LPC accumulate when needed).  H: history
// samples of the register windows = the largest LPC order of the compression level (4 when the level has no LPC stage).
// ==========================================================================================
template <bool WIDE, int H>
__global__ void __launch_bounds__(ENC_CTA, 1) frx_encode_kernel(const EncArgs a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    uint16_t (*crc_tab)[256] = reinterpret_cast<uint16_t (*)[256]>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned char *wbase = smem_raw + 2048 + (size_t)warp * WARP_BYTES;
    WarpSh &ws = *reinterpret_cast<WarpSh *>(wbase);
    uint32_t *buf = reinterpret_cast<uint32_t *>(wbase + WARP_SH_BYTES);
    uint32_t *pcm = buf + SLACK_W;
    for (int i = tid; i < 1024; i += ENC_CTA) (&crc_tab[0][0])[i] = a.crc_tab[i];
    if (lane == 0) {
        mbar_init(&ws.mbar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    uint32_t phase = 0;
    int ti = 0;
    const bool stereo = a.stereo != 0;
    // Phase alignment for the narrow kernels: with the frame warps of an SM each in a phase of its own, the 8/16-bit
    // workloads ran 1.6-2.3x slower (instruction fetch: 18 % of the stall samples were no_instruction, all phases stretched,
    // and the frames drifted so far apart that nearly every look-back had to wait); in step they share the instruction cache
    // and publish their sizes together.  The wide kernels (float32 / 32-bit) lose a few per cent to the waiting at the
    // barriers instead (their load phase is exposed), so they run free (profiles/r2_encode_experiments.md).
    constexpr bool ALIGNED = !WIDE;

    // Where the bytes go.  A frame's place in the output is the sum of the sizes of all frames in front of it (decoupled
    // look-back over the descriptors), and a frame whose predecessors have not all published their sizes must not hold up the
    // warp: it is PARKED in one of the warp's two slots in global memory (L2-resident) and the warp goes on to its next frame;
    // the parked frame is sent to its final place as soon as a (non-blocking) look-back completes, at the latest when the next
    // frame is finished -- a whole frame of slack, so that the blocking wait is practically never taken.
    // Multi-channel frames are STREAMED: a channel is packed right after its analysis (the buffer holds one channel at a time,
    // packing later would mean staging and predicting it a second time) and the complete words go to the current slot; the
    // frame's size is published when its last channel has been analysed.  Single-channel frames and stereo frames (whose
    // channel assignment needs all four analyses) run two sweeps: analyse, publish the size, pack.
    uint8_t *slots2 = a.slots + (size_t)(blockIdx.x * ENC_WARPS + warp) * 2u * a.slot_stride;
    int cur = 0;                          // the slot the next streamed / parked frame uses
    bool pend = false;                    // a parked frame waits in pend_slot
    uint32_t pend_f = 0, pend_nbytes = 0;
    uint8_t *pend_slot = slots2;
    LookBack plb;                         // its look-back, resumed where it stopped
    plb.pos = -1; plb.exclusive = 0; plb.blocked = false;
    auto lb_run = [&](LookBack &lb, bool block) -> bool {      // true: lb.exclusive is the byte offset
        for (;;) {
            if (lookback_step(a, lb, lane)) return true;
            if (lb.blocked) {
                if (!block) return false;
                DBG_ADD(1, 1);
                __nanosleep(100);
            }
        }
    };
    auto lb_commit = [&](uint32_t ff, uint32_t nb, uint64_t off) -> bool {     // publishes the inclusive prefix; true on overflow
        if (lane == 0) {
            frx_st_relaxed_u64(&a.desc[ff], (2ull << 62) | (off + (uint64_t)nb));
            if (a.frame_offset) a.frame_offset[ff] = off;
            if (ff == a.total_frames - 1 && a.total_bytes) *a.total_bytes = off + nb;
        }
        const bool ovf = off + nb > a.capacity;
        if (ovf && lane == 0) atomicExch(a.status, FRX_E_OVERFLOW);
        return ovf;
    };
    auto finish_pending = [&](bool block) {
        const long long tl0 = DBG_CLK();
        const bool ok = lb_run(plb, block);
        if (block) DBG_ADD(0, DBG_CLK() - tl0);
        if (!ok) return;
        const bool ovf = lb_commit(pend_f, pend_nbytes, plb.exclusive);
        const long long ts0 = DBG_CLK();
        if (!ovf) copy_slot_out(a.bytes + plb.exclusive, pend_slot, pend_nbytes, lane);
        DBG_ADD(2, DBG_CLK() - ts0);
        pend = false;
    };

    for (;;) {
        uint32_t f = 0;
        if (lane == 0) f = atomicAdd(a.ticket, 1u);
        f = __shfl_sync(FULLMASK, f, 0);
        if (ALIGNED) {
            if (!group_or(f < a.total_frames)) break;          // the group's frames start together
            if (f >= a.total_frames) {
                for (int k = 0; k < a.bars_per_frame; k++) group_bar();
                continue;
            }
        } else if (f >= a.total_frames) break;
        const long long tf0 = DBG_CLK();
        build_frame_desc(a, ws.fd, f, ti, lane);
        const int nsamp = ws.fd.nsamp;
        const int n_ch = ws.fd.t.n_channels;
        const bool two_sweeps = n_ch == 1 || (n_ch == 2 && stereo);
        BufState bs;
        bs.kind = BUF_JUNK; bs.vch = -1; bs.type = 0; bs.order = 0;
        int nbar = 0;
        Emitter em;
        em.P = 0; em.carry = 0;
        uint8_t *const slot = slots2 + (size_t)cur * a.slot_stride;
        uint32_t done = 0;          // bytes of the frame already in the slot
        uint32_t crc_run = 0;
        uint32_t nbytes = 0;
        auto emit_frame_header = [&]() {
            uint32_t val[1], len[1], zer[1] = { 0 };
            const uint32_t hbits = ws.fd.hdr_bits;
            val[0] = lane < 4 ? ws.fd.hdr[lane] : 0u;
            len[0] = (uint32_t)max(0, min(32, (int)hbits - 32 * lane));
            if (len[0] < 32u && len[0] > 0u) val[0] >>= (32u - len[0]);
            emit_fields<1>(buf, em, val, len, zer, lane);
        };
        auto publish_size = [&](uint32_t frame_bits) {
            nbytes = ((frame_bits + 7u) >> 3) + 2u;
            if (lane == 0) {
                // the frame's size is final: the look-back of the frames behind this one can proceed
                if (f != 0) frx_st_relaxed_u64(&a.desc[f], (1ull << 62) | (uint64_t)nbytes);
                if (a.frame_size) a.frame_size[f] = nbytes;
            }
        };
        auto pack_subframe = [&](const SubDec &d) {
            em = pack_subframe_header(ws, buf, em, d, pcm, lane);
            if (d.type >= 2) em = d.runs ? pack_runs(buf, em, d, pcm, lane) : pack_rice(buf, em, d, pcm, nsamp, lane);
            else if (d.type == 1) em = pack_verbatim(buf, em, d, pcm, nsamp, lane);
            bs.kind = BUF_JUNK;                     // the packed bits now lie over the samples
            ALIGN_BAR(nbar);
        };
        auto flush_words = [&]() {                   // the next channel needs the whole buffer: the complete words go to the slot
            const uint32_t nw = em.P >> 5;
            __syncwarp();
            const uint32_t cs = crc_words(a, crc_tab, buf, nw, lane);
            crc_run = (done ? frx_gf16_mulmod(crc_run, gf16_xpow32(a, nw)) : 0u) ^ cs;
            copy_words_out(slot + done, buf, nw, lane);
            done += nw * 4u;
            em.P &= 31u;
            __syncwarp();
        };

        if (two_sweeps) {
            // ---- sweep 1: search every subframe, exact sizes ----
            int sub0 = 0, sub1 = 1;     // decision records of the two subframes (the stereo choice permutes them)
            if (n_ch == 2) {
                // left, right, mid, side; the cheapest of L/R, L/S, S/R, M/S (libFLAC's exhaustive stereo decision)
                nbar += analyse_channel<WIDE, H>(a, ws, buf, 0, 0, ws.fd.hdr_bits, phase, bs, lane);
                nbar += analyse_channel<WIDE, H>(a, ws, buf, 1, 1, ws.fd.hdr_bits, phase, bs, lane);
                if (pend) finish_pending(false);
                nbar += analyse_channel<WIDE, H>(a, ws, buf, 8, 2, ws.fd.hdr_bits, phase, bs, lane);
                nbar += analyse_channel<WIDE, H>(a, ws, buf, 9, 3, ws.fd.hdr_bits, phase, bs, lane);
                const uint32_t bL = ws.dec[0].bits, bR = ws.dec[1].bits, bM = ws.dec[2].bits, bS = ws.dec[3].bits;
                uint32_t best = bL + bR; int ca = 1; sub0 = 0; sub1 = 1;
                if (bL + bS < best) { best = bL + bS; ca = 8; sub0 = 0; sub1 = 3; }
                if (bR + bS < best) { best = bR + bS; ca = 9; sub0 = 3; sub1 = 1; }
                if (bM + bS < best) { best = bM + bS; ca = 10; sub0 = 2; sub1 = 3; }
                if (lane == 0) make_frame_header(a, ws.fd, ca);
                __syncwarp();
                publish_size(ws.fd.hdr_bits + ws.dec[sub0].bits + ws.dec[sub1].bits);
            } else {
                nbar += analyse_channel<WIDE, H>(a, ws, buf, 0, 0, ws.fd.hdr_bits, phase, bs, lane);
                publish_size(ws.fd.hdr_bits + ws.dec[0].bits);
            }
            if (pend) finish_pending(false);
            // ---- sweep 2: pack ----
            emit_frame_header();
            for (int c = 0; c < n_ch; c++) {
                const SubDec &d = ws.dec[c == 0 ? sub0 : sub1];
                materialise<WIDE, H>(a, ws, buf, d, phase, bs, lane);
                pack_subframe(d);
                if (c + 1 < n_ch) flush_words();
            }
        } else {
            // ---- streamed: analyse, pack and send each channel on its way ----
            for (int c = 0; c < n_ch; c++) {
                nbar += analyse_channel<WIDE, H>(a, ws, buf, c, 0, c == 0 ? ws.fd.hdr_bits : (em.P & 31u), phase, bs, lane);
                const SubDec &d = ws.dec[0];
                if (c == 0) emit_frame_header();
                if (c + 1 == n_ch) publish_size(done * 8u + em.P + d.bits);
                if (pend) finish_pending(false);
                materialise<WIDE, H>(a, ws, buf, d, phase, bs, lane);      // (a no-op unless the subframe fell back to VERBATIM)
                pack_subframe(d);
                if (c + 1 < n_ch) flush_words();
            }
        }
        // ---- frame footer: zero padding to a byte boundary, CRC-16 ----
        {
            const uint32_t p0 = em.P;
            em.P = (em.P + 7u) & ~7u;
            if ((p0 & 31u) && !(em.P & 31u)) {       // the padding completes the last word: it is still in the carry
                if (lane == 0) buf[p0 >> 5] = em.carry;
                em.carry = 0;
            }
        }
        const uint32_t nw = em.P >> 5, tb = (em.P & 31u) >> 3;
        __syncwarp();
        uint32_t crc = crc_words(a, crc_tab, buf, nw, lane);
        if (done) crc ^= frx_gf16_mulmod(crc_run, gf16_xpow32(a, nw));
        for (uint32_t b = 0; b < tb; b++) {
            const uint32_t byte = (em.carry >> (24 - 8 * b)) & 0xFFu;
            crc = ((crc << 8) ^ crc_tab[0][((crc >> 8) ^ byte) & 0xFFu]) & 0xFFFFu;
        }
        // ---- out: to its final place when the offset is there, else parked ----
        if (pend) finish_pending(true);              // (the frame before this one: it has had this whole frame's time)
        LookBack lb;
        lb.pos = (long long)f - 1; lb.exclusive = 0; lb.blocked = false;
        const bool have = lb_run(lb, false);
        bool skip = false;
        uint8_t *dst = slot + done;
        if (have) {
            skip = lb_commit(f, nbytes, lb.exclusive);
            if (!skip && done) copy_slot_out(a.bytes + lb.exclusive, slot, done, lane);
            dst = a.bytes + lb.exclusive + done;
        }
        if (!skip) {
            copy_words_out(dst, buf, nw, lane);
            if (lane < (int)tb) dst[nw * 4u + lane] = (uint8_t)((em.carry >> (24 - 8 * lane)) & 0xFFu);
            if (lane == 4) dst[nw * 4u + tb] = (uint8_t)(crc >> 8);
            if (lane == 5) dst[nw * 4u + tb + 1u] = (uint8_t)crc;
        }
        if (!have) {
            pend = true; pend_f = f; pend_nbytes = nbytes; pend_slot = slot; plb = lb;
            cur ^= 1;
            DBG_ADD(12, 1);
        }
        __syncwarp();
        if (ALIGNED) { while (nbar < a.bars_per_frame) ALIGN_BAR(nbar); }      // the same number of barriers for every frame
        DBG_ADD(3, DBG_CLK() - tf0); DBG_ADD(4, 1);
    }
    if (pend) finish_pending(true);
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
    (void)frame_words; (void)dtype;
    return 2048 + (size_t)ENC_WARPS * WARP_BYTES;
}
int frx_enc_warps_per_cta() { return ENC_WARPS; }

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
    const size_t smem = frx_enc_smem_bytes(0, 0);
    cudaError_t e = cudaFuncSetAttribute(frx_encode_kernel<WIDE, HISTN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(frx_encode_kernel<WIDE, HISTN>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if (e != cudaSuccess) return e;
    // persistent: one CTA of twelve frame warps per SM, tickets hand out the frames in order
    long long grid = sm_count;
    const long long need = ((long long)a.total_frames + ENC_WARPS - 1) / ENC_WARPS;
    if (grid > need) grid = need;
    if (grid < 1) grid = 1;
    frx_encode_kernel<WIDE, HISTN><<<(unsigned)grid, ENC_CTA, smem, st>>>(a);
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


// ------------------------------------------------------------------------------------------
// MD5 of a stream's PCM (STREAMINFO's signature: the channel-interleaved samples as little-endian bytes of the effective
// sample size).  One thread per stream: MD5 is a serial chain over 64-byte blocks, so this is off the hot path -- the
// reference leaves the field zero and so does create-streaming unless asked.
// ------------------------------------------------------------------------------------------
__constant__ uint32_t c_md5_k[64] = {
    0xd76aa478, 0xe8c7b756, 0x242070db, 0xc1bdceee, 0xf57c0faf, 0x4787c62a, 0xa8304613, 0xfd469501, 0x698098d8, 0x8b44f7af, 0xffff5bb1,
    0x895cd7be, 0x6b901122, 0xfd987193, 0xa679438e, 0x49b40821, 0xf61e2562, 0xc040b340, 0x265e5a51, 0xe9b6c7aa, 0xd62f105d, 0x02441453,
    0xd8a1e681, 0xe7d3fbc8, 0x21e1cde6, 0xc33707d6, 0xf4d50d87, 0x455a14ed, 0xa9e3e905, 0xfcefa3f8, 0x676f02d9, 0x8d2a4c8a, 0xfffa3942,
    0x8771f681, 0x6d9d6122, 0xfde5380c, 0xa4beea44, 0x4bdecfa9, 0xf6bb4b60, 0xbebfbc70, 0x289b7ec6, 0xeaa127fa, 0xd4ef3085, 0x04881d05,
    0xd9d4d039, 0xe6db99e5, 0x1fa27cf8, 0xc4ac5665, 0xf4292244, 0x432aff97, 0xab9423a7, 0xfc93a039, 0x655b59c3, 0x8f0ccc92, 0xffeff47d,
    0x85845dd1, 0x6fa87e4f, 0xfe2ce6e0, 0xa3014314, 0x4e0811a1, 0xf7537e82, 0xbd3af235, 0x2ad7d2bb, 0xeb86d391 };
__constant__ uint8_t c_md5_s[64] = { 7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22, 5, 9, 14, 20, 5, 9, 14, 20, 5, 9, 14, 20, 5, 9, 14, 20,
                                     4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23, 6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21 };
__device__ void md5_block(uint32_t (&st)[4], const uint32_t (&w)[16]) {
    uint32_t a = st[0], b = st[1], c = st[2], d = st[3];
    for (int i = 0; i < 64; i++) {
        uint32_t f;
        int g;
        if (i < 16) { f = (b & c) | (~b & d); g = i; }
        else if (i < 32) { f = (d & b) | (~d & c); g = (5 * i + 1) & 15; }
        else if (i < 48) { f = b ^ c ^ d; g = (3 * i + 5) & 15; }
        else { f = c ^ (b | ~d); g = (7 * i) & 15; }
        f += a + c_md5_k[i] + w[g];
        a = d; d = c; c = b;
        b += (f << c_md5_s[i]) | (f >> (32 - c_md5_s[i]));
    }
    st[0] += a; st[1] += b; st[2] += c; st[3] += d;
}
__device__ int32_t md5_sample(const void *raster, long long band_stride, long long row_stride, const frx_tile &t, const FrxQuant &q, int dtype, int mode,
                              long long i) {
    const int ch = (int)(i % t.n_channels);
    const long long px = i / t.n_channels;
    const long long row = px / t.w, col = px - row * t.w;
    const long long e = (long long)(t.first_band + ch) * band_stride + (t.row_off + row) * row_stride + t.col_off + col;
    switch (dtype) {
        case FRX_U8: return frx_quant<FRX_U8>(((const uint8_t *)raster)[e], q, mode);
        case FRX_I16: return frx_quant<FRX_I16>(((const int16_t *)raster)[e], q, mode);
        case FRX_U16: return frx_quant<FRX_U16>(((const uint16_t *)raster)[e], q, mode);
        case FRX_I32: return frx_quant<FRX_I32>(((const int32_t *)raster)[e], q, mode);
        case FRX_U32: return frx_quant<FRX_U32>(((const uint32_t *)raster)[e], q, mode);
        default: return frx_quant<FRX_F32>(((const float *)raster)[e], q, mode);
    }
}
__global__ void frx_md5_kernel(const void *raster, long long band_stride, long long row_stride, const frx_tile *tiles, const FrxQuant *quant,
                               int n_tiles, int dtype, int mode, int bps, uint8_t *out) {
    const int ti = blockIdx.x * blockDim.x + threadIdx.x;
    if (ti >= n_tiles) return;
    const frx_tile t = tiles[ti];
    const FrxQuant q = quant[ti];
    const long long n = (long long)t.h * t.w * t.n_channels;      // samples
    const int spw = bps == 16 ? 2 : 1;                            // samples per 32-bit word
    const long long nbytes = n * (bps / 8);
    uint32_t st[4] = { 0x67452301u, 0xefcdab89u, 0x98badcfeu, 0x10325476u };
    uint32_t w[16];
    long long i = 0;
    const long long nwords = (nbytes + 3) / 4;                    // data words (the last may be half full at bps 16)
    long long wi = 0;
    bool padded = false, done = false;
    while (!done) {
        for (int k = 0; k < 16; k++) {
            uint32_t v = 0;
            if (wi < nwords) {
                if (spw == 1) v = (uint32_t)md5_sample(raster, band_stride, row_stride, t, q, dtype, mode, i++);
                else {
                    v = (uint32_t)md5_sample(raster, band_stride, row_stride, t, q, dtype, mode, i++) & 0xFFFFu;
                    if (i < n) v |= (uint32_t)md5_sample(raster, band_stride, row_stride, t, q, dtype, mode, i++) << 16;
                    else { v |= 0x80u << 16; padded = true; }    // odd number of 16-bit samples: the pad byte follows them
                }
            } else if (!padded) { v = 0x80u; padded = true; }
            w[k] = v;
            wi++;
        }
        // the length goes into the last two words of the first block that has room behind the pad byte
        if (padded && wi - 16 + 14 >= nwords + ((nbytes & 3) ? 0 : 1)) {
            w[14] = (uint32_t)((unsigned long long)nbytes << 3);
            w[15] = (uint32_t)(((unsigned long long)nbytes << 3) >> 32);
            done = true;
        }
        md5_block(st, w);
    }
    for (int k = 0; k < 4; k++)
        for (int b = 0; b < 4; b++) out[(size_t)ti * 16 + 4 * k + b] = (uint8_t)(st[k] >> (8 * b));
}
cudaError_t frx_launch_md5(const frx_raster &r, const frx_tile *tiles_dev, const FrxQuant *quant, int n_tiles, int mode, int bps, uint8_t *out,
                           cudaStream_t st) {
    frx_md5_kernel<<<(n_tiles + 31) / 32, 32, 0, st>>>(r.data, r.band_stride, r.row_stride, tiles_dev, quant, n_tiles, r.dtype, mode, bps, out);
    return cudaGetLastError();
}

#ifdef FRX_DBG
extern "C" void frx_debug_read(unsigned long long *out16, int reset) {
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out16, g_frx_dbg, sizeof(unsigned long long) * 16);
    if (reset) { unsigned long long z[16] = { 0 }; cudaMemcpyToSymbol(g_frx_dbg, z, sizeof(z)); }
}
#endif

// frx_common.cuh -- shared device helpers for libfrx (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/frx.h"

#define FRX_MAX_CH 8
#define FRX_MAX_LPC 12          // libFLAC levels 0..8 never exceed order 12 (dump:6926-6934)
#define FRX_MAX_BLOCK 4096      // largest frame: one warp per frame, 128 samples per lane

struct FrxCtxDev;  // opaque

// ------------------------------------------------------------------------------------------
// dtype helpers
// ------------------------------------------------------------------------------------------
__host__ __device__ inline int frx_dtype_size(int dt) {
    switch (dt) {
        case FRX_U8: return 1;
        case FRX_I16: case FRX_U16: return 2;
        default: return 4;
    }
}

// Effective PCM bits per sample (pyFLAC: array itemsize * 8, docs/sonos-pyflac.txt:1986-1992).
__host__ __device__ inline int frx_bps_for(int dt, int mode) {
    if (mode == FRX_PCM_REFERENCE_SPATIAL) return 32;  // float32 array -> 4 bytes -> 32
    if (dt == FRX_U8 || dt == FRX_I16 || dt == FRX_U16) return 16;
    return 32;
}

// x86 cvttsd2si / cvttss2si semantics, which is what numpy's astype(int32) compiles to:
// NaN and out-of-range give INT_MIN ("integer indefinite").  PTX cvt.rzi saturates instead.
__device__ __forceinline__ int32_t frx_trunc_i32(double v) {
    if (!(v > -2147483649.0 && v < 2147483648.0)) return INT32_MIN;
    return __double2int_rz(v);
}
__device__ __forceinline__ int32_t frx_trunc_i32f(float v) {
    if (!(v >= -2147483648.0f && v < 2147483648.0f)) return INT32_MIN;
    return __float2int_rz(v);
}

// Per-stream quantisation constants (one per tile), produced by the min/max kernel.
struct FrxQuant {
    long long imin, imax;   // integer dtypes: min/max as int64
    float fmin, fmax;       // float32: min/max (NaN propagates like np.min)
    double range;           // (double)(T)(dmax - dmin), wrapped in the raster dtype (converter.py:72)
    double rcp;             // RN(1 / range): the 16-bit dtypes divide by multiplying (frx_div16)
    int degenerate;         // dmax == dmin  -> zeros (converter.py:73-75)
    int flags;              // FRX_QF_*
    int irange;             // FRX_QF_INT16: max - min as an integer, 1..65535
    float rcpf;             // FRX_QF_INT16: RZ(1 / irange) in float32
};
#define FRX_QF_INT16 2      // 8/16-bit dtypes whose max - min and x - min do not wrap in the raster dtype: 0 <= x - min <= range <=
                            // 65535, the case the integer form of the quantiser is proven for (oracle/quant16_check.c)
#define FRX_QF_F32_SAFE 1   // float32 reference mode: no NaN and |x| * 8388607 < 2^31 for the whole stream, so the
                            // x86 "integer indefinite" check of frx_trunc_i32f can never fire

// raw element -> int64 (integers) ; floats handled separately
template <int DT> struct FrxElem;
template <> struct FrxElem<FRX_U8>  { typedef uint8_t  T; };
template <> struct FrxElem<FRX_I16> { typedef int16_t  T; };
template <> struct FrxElem<FRX_U16> { typedef uint16_t T; };
template <> struct FrxElem<FRX_I32> { typedef int32_t  T; };
template <> struct FrxElem<FRX_U32> { typedef uint32_t T; };
template <> struct FrxElem<FRX_F32> { typedef float    T; };

// (double)(T)(d - dmin): the subtraction is done in the raster's own dtype and wraps
// (numpy array - numpy scalar of the same dtype).
template <int DT>
__device__ __forceinline__ double frx_wrap_sub(typename FrxElem<DT>::T d, long long dmin) {
    typedef typename FrxElem<DT>::T T;
    return (double)(T)((T)d - (T)dmin);
}

// RN(a / b) for integer-valued |a| <= 131070, 1 <= |b| <= 65535 without the division sequence: y0 = RN(a rcp), r = a - y0 b
// (exact, FMA), RN(y0 + r rcp).  Markstein's correction step; checked exhaustively over the whole operand range against the
// IEEE quotient by oracle/div_check.c (8.6e9 pairs, no mismatch), so the quantiser stays bit-exact.
__device__ __forceinline__ double frx_div16(double a, double b, double rcp) {
    const double y0 = __dmul_rn(a, rcp);
    const double r = __fma_rn(-y0, b, a);
    return __fma_rn(r, rcp, y0);
}

// Integer form of the 16-bit float64 chain of frx_quant_ref (only under FRX_QF_INT16): with v = x - min and R = max - min the
// chain's value is N / R for N = (2 v - R) * 32767 (|N| < 2^31) up to rounding errors far smaller than 1 / R, so truncating
// the exact quotient gives the same integer unless N / R is itself an integer -- then (quotients 0 and +-32767 apart) the
// chain decides, and `exact` is set.  oracle/quant16_check.c proves both statements over all 2^31 (R, v) pairs, and that
// the float32 product below is floor(|N| / R) or one less.
template <int DT>
__device__ __forceinline__ int32_t frx_quant16_int(typename FrxElem<DT>::T d, const FrxQuant &q, bool &exact) {
    typedef typename FrxElem<DT>::T T;
    const int iv = (int)(T)((T)d - (T)q.imin);
    const int N = (2 * iv - q.irange) * 32767;
    const uint32_t M = (uint32_t)abs(N), R = (uint32_t)q.irange;
    uint32_t qh = __float2uint_rz(__fmul_rz(__uint2float_rz(M), q.rcpf));
    uint32_t rem = M - qh * R;
    if (rem >= R) { qh++; rem -= R; }
    exact |= rem == 0u && qh - 1u < 32766u;
    return N < 0 ? -(int32_t)qh : (int32_t)qh;
}

// K1 quantiser, FRX_PCM_REFERENCE: bit-exact restatement of
// RasterFLACConverter._normalize_to_audio (converter.py:56-86) followed by pyFLAC's
// astype(int32) (dump:1994).  Every floating-point op is a separately rounded IEEE op.
template <int DT>
__device__ __forceinline__ int32_t frx_quant_ref_chain(typename FrxElem<DT>::T d, const FrxQuant &q) {
    if (DT == FRX_F32) {
        // float data "as-is"; float32 array * python int 8388607 stays float32 (NEP 50)
        float y = __fmul_rn((float)d, 8388607.0f);
        if (q.flags & FRX_QF_F32_SAFE) return __float2int_rz(y);
        return frx_trunc_i32f(y);
    } else {
        if (q.degenerate) return 0;
        double v = frx_wrap_sub<DT>(d, q.imin);
        if (DT == FRX_U8 || DT == FRX_I16 || DT == FRX_U16) {
            const double norm = __dsub_rn(frx_div16(__dmul_rn(2.0, v), q.range, q.rcp), 1.0);
            return (int32_t)(int16_t)frx_trunc_i32(__dmul_rn(norm, 32767.0));  // astype(int16)
        } else {
            const double norm = __dsub_rn(__ddiv_rn(__dmul_rn(2.0, v), q.range), 1.0);
            return frx_trunc_i32(__dmul_rn(norm, 8388607.0));
        }
    }
}
template <int DT>
__device__ __forceinline__ int32_t frx_quant_ref(typename FrxElem<DT>::T d, const FrxQuant &q) {
    if ((DT == FRX_U8 || DT == FRX_I16 || DT == FRX_U16) && (q.flags & FRX_QF_INT16)) {
        bool exact = false;
        const int32_t r = frx_quant16_int<DT>(d, q, exact);
        if (!exact) return r;
    }
    return frx_quant_ref_chain<DT>(d, q);
}

// FRX_PCM_LOSSLESS: identity mapping into a signed PCM of the effective width.
template <int DT>
__device__ __forceinline__ int32_t frx_quant_lossless(typename FrxElem<DT>::T d) {
    if (DT == FRX_U8) return (int32_t)d;
    if (DT == FRX_I16) return (int32_t)d;
    if (DT == FRX_U16) return (int32_t)d - 32768;
    if (DT == FRX_I32) return (int32_t)d;
    if (DT == FRX_U32) return (int32_t)((uint32_t)d ^ 0x80000000u);
    // float32: order-preserving bit map (sign-magnitude -> two's complement)
    int32_t b = __float_as_int((float)d);
    return b >= 0 ? b : (int32_t)(0x80000000u - (uint32_t)b);
}

// FRX_PCM_REFERENCE_SPATIAL: SpatialFLACEncoder._normalize_to_audio (spatial_encoder.py:229-248)
// then pyFLAC's float32 -> int32 truncation (dump:1994)  => {-1, 0, 1}.
template <int DT>
__device__ __forceinline__ int32_t frx_quant_spatial(typename FrxElem<DT>::T d) {
    float f;
    if (DT == FRX_U8) f = __fdiv_rn(__fsub_rn((float)d, 127.5f), 127.5f);
    else if (DT == FRX_U16) f = __fdiv_rn(__fsub_rn((float)d, 32767.5f), 32767.5f);
    else if (DT == FRX_I16) f = __fdiv_rn((float)d, 32767.0f);
    else if (DT == FRX_I32) f = __fdiv_rn((float)d, 2147483647.0f);
    else if (DT == FRX_F32) {
        float x = (float)d;  // np.clip propagates NaN
        f = (x != x) ? x : fminf(fmaxf(x, -1.0f), 1.0f);
    } else f = 0.0f;
    return frx_trunc_i32f(f);
}

template <int DT>
__device__ __forceinline__ int32_t frx_quant(typename FrxElem<DT>::T d, const FrxQuant &q, int mode) {
    if (mode == FRX_PCM_REFERENCE) return frx_quant_ref<DT>(d, q);
    if (mode == FRX_PCM_LOSSLESS) return frx_quant_lossless<DT>(d);
    return frx_quant_spatial<DT>(d);
}

// K12 dequantiser constants
struct FrxDequant {
    float range_f, dmin_f;  // float32(dmax - dmin), float32(dmin)  (NEP 50 weak python scalars)
};

// FRX_PCM_REFERENCE decode: pyFLAC FileDecoder (int16 narrowing, PCM_16 WAV, /32768;
// dump:1809-1854, 1621-1640) then _denormalize_from_audio (converter.py:99-110) in float32.
template <int DT>
__device__ __forceinline__ typename FrxElem<DT>::T frx_dequant_ref(int32_t s, int bps, const FrxDequant &q) {
    typedef typename FrxElem<DT>::T T;
    int16_t s16 = (bps == 32) ? (int16_t)(s >> 16) : (int16_t)s;
    float x = (float)s16 * 3.0517578125e-05f;  // exact: int16 / 32768
    if (DT == FRX_F32) return (T)x;
    // (x + 1) / 2 as a multiplication by 0.5: the same float for every x here (x + 1 is 0 or lies in [2^-15, 2), halving is exact),
    // and one instruction instead of the IEEE division sequence with its range check and slow-path call
    float den = __fadd_rn(__fmul_rn(__fmul_rn(__fadd_rn(x, 1.0f), 0.5f), q.range_f), q.dmin_f);
    float r = rintf(den);  // np.round: half to even
    if (DT == FRX_U32) return (T)(uint32_t)__float2ll_rz(r);
    return (T)frx_trunc_i32f(r);
}

template <int DT>
__device__ __forceinline__ typename FrxElem<DT>::T frx_dequant_lossless(int32_t s) {
    typedef typename FrxElem<DT>::T T;
    if (DT == FRX_U8) return (T)s;
    if (DT == FRX_I16) return (T)s;
    if (DT == FRX_U16) return (T)(s + 32768);
    if (DT == FRX_I32) return (T)s;
    if (DT == FRX_U32) return (T)((uint32_t)s ^ 0x80000000u);
    int32_t b = s >= 0 ? s : (int32_t)(0x80000000u - (uint32_t)s);
    return (T)__int_as_float(b);
}

template <int DT>
__device__ __forceinline__ typename FrxElem<DT>::T frx_dequant(int32_t s, int bps, const FrxDequant &q, int mode) {
    if (mode == FRX_PCM_LOSSLESS) return frx_dequant_lossless<DT>(s);
    return frx_dequant_ref<DT>(s, bps, q);  // REFERENCE and REFERENCE_SPATIAL share the decode side
}

// ------------------------------------------------------------------------------------------
// CRC (FLAC: CRC-8 poly 0x07 over the frame header, CRC-16 poly 0x8005 over the frame; MSB first)
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t frx_crc8_byte(uint32_t crc, uint32_t byte) {
    crc ^= byte;
#pragma unroll
    for (int i = 0; i < 8; i++) crc = (crc & 0x80u) ? ((crc << 1) ^ 0x07u) & 0xFFu : (crc << 1) & 0xFFu;
    return crc;
}

// multiply a 16-bit polynomial by x^8 k times... generic GF(2) multiply modulo 0x18005
__host__ __device__ inline uint32_t frx_gf16_mulmod(uint32_t a, uint32_t b) {
    uint32_t r = 0;
    for (int i = 15; i >= 0; i--) {
        r <<= 1;
        if (r & 0x10000u) r ^= 0x18005u;
        if ((b >> i) & 1u) r ^= a;
    }
    return r & 0xFFFFu;
}

// ------------------------------------------------------------------------------------------
// misc
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t frx_bswap32(uint32_t v) { return __byte_perm(v, 0, 0x0123); }

__device__ __forceinline__ uint64_t frx_ld_acquire_u64(const uint64_t *p) {
    uint64_t v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
// relaxed device-scope accesses: enough when the 64-bit word itself is the whole message
__device__ __forceinline__ uint64_t frx_ld_relaxed_u64(const uint64_t *p) {
    uint64_t v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void frx_st_relaxed_u64(uint64_t *p, uint64_t v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void frx_st_release_u64(uint64_t *p, uint64_t v) {
    asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// ordered-uint mapping for float min/max via integer atomics
__device__ __forceinline__ uint32_t frx_float_ordered(float f) {
    uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float frx_float_unordered(uint32_t u) {
    uint32_t b = (u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u;
    return __uint_as_float(b);
}

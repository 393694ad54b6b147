// frx_ingest.cu -- uncompressed TIFF strips/tiles (as they lie in the file) -> band-planar device raster.
//
// Replaces the host side of rasterio's read(): GDAL copies every strip/tile into a (bands,h,w) array on the CPU
// (converter.py:137-141; spatial_encoder.py:154-183; cli.py:699) and the caller then ships that array to the GPU.  Here
// the file image goes to the device once and frx_tiff_unpack_kernel de-interleaves, un-tiles, clips and byte-swaps there.
//
// Work unit = one row segment of a wide chunk, or a run of whole rows of a narrow one (<= FRX_ING_SEG_BYTES of source; frx_ingest_plan).  The segment is staged in shared memory with
// 16-byte loads (source and staging buffer share their alignment phase, so the bulk is aligned on both sides), then
// written band by band with 16-byte stores where the destination allows.  HBM-bound: 2 x raster bytes per call.
#include "frx_internal.cuh"

#define ING_THREADS 256

template <int ESZ> struct IngElem;
template <> struct IngElem<1> { typedef uint8_t T; };
template <> struct IngElem<2> { typedef uint16_t T; };
template <> struct IngElem<4> { typedef uint32_t T; };

template <int ESZ>
__device__ __forceinline__ typename IngElem<ESZ>::T ing_swap(typename IngElem<ESZ>::T v, int swap) {
    if (ESZ == 2) return swap ? (typename IngElem<ESZ>::T)(((uint32_t)v >> 8) | (((uint32_t)v & 0xFFu) << 8)) : v;
    if (ESZ == 4) return swap ? (typename IngElem<ESZ>::T)__byte_perm((uint32_t)v, 0, 0x0123) : v;
    return v;
}

// One destination row of one band: `n` samples, sample i of the row at srow + (i * nb + b) * ESZ in the staging buffer.
// Called by a group of `wsize` threads (a warp, or the whole CTA), `wid` = index inside the group.  Full 16-byte stores
// in the middle; single-band rows assemble them from aligned 32-bit shared-memory words with funnel shifts (the source's
// alignment phase is whatever the file gave it), interleaved rows gather sample by sample.
template <int ESZ>
__device__ __forceinline__ void ing_write_row(typename IngElem<ESZ>::T *dp, const uint8_t *srow, int n, int nb, int b, int swap, int wid,
                                              int wsize) {
    typedef typename IngElem<ESZ>::T T;
    constexpr int VEC = 16 / ESZ;
    const uint8_t *s0 = srow + (size_t)b * ESZ;
    const bool el_ok = ((uintptr_t)s0 % ESZ) == 0 && (((size_t)nb * ESZ) % ESZ) == 0;   // staged samples naturally aligned
    auto elem = [&](int i) -> T {
        const uint8_t *e = s0 + (size_t)i * nb * ESZ;
        T v;
        if (el_ok) v = *reinterpret_cast<const T *>(e);
        else { v = 0; for (int k = 0; k < ESZ; k++) v |= (T)((T)e[k] << (8 * k)); }
        return ing_swap<ESZ>(v, swap);
    };
    const int dhead = min(n, (int)(((16u - (uint32_t)((uintptr_t)dp & 15u)) & 15u) / ESZ));
    for (int i = wid; i < dhead; i += wsize) dp[i] = elem(i);
    const int nvec = (n - dhead) / VEC;
    if (nb == 1) {
        const uint8_t *a0 = s0 + (size_t)dhead * ESZ;
        const uint32_t sh = (uint32_t)((uintptr_t)a0 & 3u) * 8u;
        const uint32_t *w0 = reinterpret_cast<const uint32_t *>((uintptr_t)a0 & ~(uintptr_t)3);
        if (((uintptr_t)a0 & 15u) == 0 && !swap) {                   // source and destination share their phase: straight copy
            const uint4 *q0 = reinterpret_cast<const uint4 *>(a0);
            for (int v = wid; v < nvec; v += wsize) *reinterpret_cast<uint4 *>(dp + dhead + v * VEC) = q0[v];
        } else
        for (int v = wid; v < nvec; v += wsize) {
            const uint32_t *wp = w0 + 4 * v;
            uint32_t w[5], o[4];
#pragma unroll
            for (int k = 0; k < 5; k++) w[k] = wp[k];                 // the staging buffer has slack behind the data
#pragma unroll
            for (int k = 0; k < 4; k++) {
                o[k] = __funnelshift_r(w[k], w[k + 1], sh);              // sh == 0: w[k]
                if (swap) o[k] = ESZ == 2 ? __byte_perm(o[k], 0, 0x2301) : (ESZ == 4 ? __byte_perm(o[k], 0, 0x0123) : o[k]);
            }
            *reinterpret_cast<uint4 *>(dp + dhead + v * VEC) = make_uint4(o[0], o[1], o[2], o[3]);
        }
    } else {
        for (int v = wid; v < nvec; v += wsize) {
            T tmp[VEC];
#pragma unroll
            for (int k = 0; k < VEC; k++) tmp[k] = elem(dhead + v * VEC + k);
            uint4 q;
            memcpy(&q, tmp, 16);
            *reinterpret_cast<uint4 *>(dp + dhead + v * VEC) = q;
        }
    }
    const int t0 = dhead + nvec * VEC;
    for (int i = t0 + wid; i < n; i += wsize) dp[i] = elem(i);
}

// Pixel-interleaved rows, all bands in one pass: a thread takes 16 / ESZ pixels = NB x 16 contiguous bytes of the staged row
// (NB 16-byte shared-memory loads, conflict-free at this stride), separates the bands in registers and writes one 16-byte
// vector per band plane, so a warp writes 512 contiguous bytes per band.  (Band by band, every output vector cost 16 / ESZ
// element loads from shared memory at a stride of NB elements -- a 4-way bank conflict each for three 16-bit bands.)
// Needs the vector part of the row 16-byte aligned in the staging buffer and the band planes congruent mod 16; the
// caller falls back to ing_write_row otherwise.
template <int ESZ, int NB>
__device__ __forceinline__ void ing_write_pixels_t(typename IngElem<ESZ>::T *dp0, long long band_stride, const uint8_t *srow, int n, int swap,
                                                   int dhead, int wid, int wsize) {
    typedef typename IngElem<ESZ>::T T;
    constexpr int VEC = 16 / ESZ;
    auto elem = [&](int p, int b) -> T {
        const uint8_t *e = srow + ((size_t)p * NB + b) * ESZ;
        T v = 0;
        for (int k = 0; k < ESZ; k++) v |= (T)((T)e[k] << (8 * k));
        return ing_swap<ESZ>(v, swap);
    };
    const int nvec = (n - dhead) / VEC;
    for (int i = wid; i < dhead * NB; i += wsize) { const int pp = i / NB, b = i - pp * NB; dp0[(long long)b * band_stride + pp] = elem(pp, b); }
    const int t0 = dhead + nvec * VEC;
    for (int i = wid; i < (n - t0) * NB; i += wsize) { const int pp = t0 + i / NB, b = i % NB; dp0[(long long)b * band_stride + pp] = elem(pp, b); }
    const uint4 *v0 = reinterpret_cast<const uint4 *>(srow + (size_t)dhead * NB * ESZ);
    for (int v = wid; v < nvec; v += wsize) {
        uint4 q[NB];
#pragma unroll
        for (int k = 0; k < NB; k++) q[k] = v0[(size_t)v * NB + k];
        T e[NB * VEC];
        memcpy(e, q, 16 * NB);
#pragma unroll
        for (int b = 0; b < NB; b++) {
            T o[VEC];
#pragma unroll
            for (int pp = 0; pp < VEC; pp++) o[pp] = ing_swap<ESZ>(e[pp * NB + b], swap);
            uint4 w;
            memcpy(&w, o, 16);
            *reinterpret_cast<uint4 *>(dp0 + (long long)b * band_stride + dhead + v * VEC) = w;
        }
    }
}
template <int ESZ>
__device__ __forceinline__ bool ing_write_pixels(typename IngElem<ESZ>::T *dp0, long long band_stride, const uint8_t *srow, int n, int nb, int swap,
                                                 int wid, int wsize) {
    if (nb < 2 || nb > 4 || (((unsigned long long)band_stride * ESZ) & 15ull)) return false;
    const int dhead = min(n, (int)(((16u - (uint32_t)((uintptr_t)dp0 & 15u)) & 15u) / ESZ));
    if (((uintptr_t)(srow + (size_t)dhead * nb * ESZ) & 15u) != 0) return false;
    if (nb == 2) ing_write_pixels_t<ESZ, 2>(dp0, band_stride, srow, n, swap, dhead, wid, wsize);
    else if (nb == 3) ing_write_pixels_t<ESZ, 3>(dp0, band_stride, srow, n, swap, dhead, wid, wsize);
    else ing_write_pixels_t<ESZ, 4>(dp0, band_stride, srow, n, swap, dhead, wid, wsize);
    return true;
}
template <int ESZ>
// six CTAs per SM (40 registers): with four, only half of 64 KB in flight per SM was loading at any time (C2 layout 63 % -> 76 % of the
// copy peak, interleaved C3 layout 51 % -> 57 %; eight CTAs spill in the interleaved path)
__global__ void __launch_bounds__(ING_THREADS, 6) frx_tiff_unpack_kernel(const uint8_t *__restrict__ src, const frx_tiff_chunk *__restrict__ chunks,
                                                                      const uint32_t *__restrict__ unit_prefix, int n_chunks, int spp,
                                                                      int swap, uint8_t *__restrict__ dst, long long band_stride,
                                                                      long long row_stride) {
    typedef typename IngElem<ESZ>::T T;
    __shared__ __align__(16) uint8_t stage[FRX_ING_SEG_BYTES + 48];
    const uint32_t u = blockIdx.x;
    // chunk of this unit: chunks of one file nearly always hold the same number of units, so the proportional guess is
    // right (two independent loads); the bisection -- log2(n_chunks) dependent L2 round trips per CTA -- is the fallback
    int lo = (int)(((unsigned long long)u * (unsigned long long)n_chunks) / gridDim.x);
    if (!(__ldg(&unit_prefix[lo]) <= u && u < __ldg(&unit_prefix[lo + 1]))) {
        int hi = n_chunks;
        lo = 0;
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (__ldg(&unit_prefix[mid]) <= u) lo = mid; else hi = mid;
        }
    }
    const frx_tiff_chunk ch = chunks[lo];
    const uint32_t lu = u - unit_prefix[lo];
    const int nb = ch.band < 0 ? spp : 1;                    // samples per pixel in this chunk
    int rows_per_unit, seg_px;
    (void)frx_ingest_plan(ESZ, nb, ch.rows, ch.cols, ch.pitch, &rows_per_unit, &seg_px);
    // unit -> (first row, row count, first column, column count)
    int r0, nr, c0, n;
    if (rows_per_unit > 1) {
        r0 = (int)lu * rows_per_unit; nr = min(rows_per_unit, ch.rows - r0); c0 = 0; n = ch.cols;
    } else {
        const uint32_t segs = ((uint32_t)ch.cols + (uint32_t)seg_px - 1u) / (uint32_t)seg_px;
        r0 = (int)(lu / segs); nr = 1; c0 = (int)(lu - (uint32_t)r0 * segs) * seg_px; n = min(seg_px, ch.cols - c0);
    }
    const size_t row_bytes = (size_t)ch.pitch * nb * ESZ;
    const size_t nbytes = (size_t)(nr - 1) * row_bytes + (size_t)n * nb * ESZ;
    const uint8_t *sp = src + ch.src_offset + ((size_t)r0 * ch.pitch + (size_t)c0) * nb * ESZ;
    // ---- stage: head bytes up to the first 16-byte boundary, aligned body, tail (sb and sp are congruent mod 16) ----
    const uint32_t phase = (uint32_t)((uintptr_t)sp & 15u);
    uint8_t *sb = stage + phase;
    const size_t head = min(nbytes, (size_t)((16u - phase) & 15u));
    if (threadIdx.x < head) sb[threadIdx.x] = sp[threadIdx.x];
    const size_t body = (nbytes - head) & ~(size_t)15;
    {
        const uint4 *s4 = reinterpret_cast<const uint4 *>(sp + head);
        uint4 *d4 = reinterpret_cast<uint4 *>(sb + head);
        for (size_t i = threadIdx.x; i < body / 16; i += ING_THREADS) d4[i] = __ldg(s4 + i);
    }
    const size_t tail0 = head + body;
    if (threadIdx.x < nbytes - tail0) sb[tail0 + threadIdx.x] = sp[tail0 + threadIdx.x];
    __syncthreads();
    // ---- write out: (row, band) pairs, one warp each when the unit holds several rows, the whole CTA otherwise ----
    T *base = reinterpret_cast<T *>(dst) + (long long)(ch.row0 + r0) * row_stride + ch.col0 + c0;
    if (ch.band < 0 && nb >= 2 && nb <= 4) {
        // interleaved pixels of a few bands: all bands of a row in one pass (uniform decision per row; falls through when the
        // alignment does not allow it)
        bool all = true;
        if (nr > 1) {
            const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
            for (int rr = warp; rr < nr && all; rr += ING_THREADS / 32)
                all = ing_write_pixels<ESZ>(base + (long long)rr * row_stride, band_stride, sb + (size_t)rr * row_bytes, n, nb, swap, lane, 32);
            // (a row that cannot take the fast path is rewritten below together with the others: same values)
            all = __syncthreads_and(all);
        } else {
            all = ing_write_pixels<ESZ>(base, band_stride, sb, n, nb, swap, threadIdx.x, ING_THREADS);
        }
        if (all) return;
    }
    if (nr > 1 || nb >= ING_THREADS / 32) {
        const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
        for (int p = warp; p < nr * nb; p += ING_THREADS / 32) {
            const int rr = p / nb, b = p - rr * nb;
            const int band = ch.band < 0 ? b : ch.band;
            ing_write_row<ESZ>(base + (long long)band * band_stride + (long long)rr * row_stride, sb + (size_t)rr * row_bytes, n, nb, b, swap,
                               lane, 32);
        }
    } else {
        for (int b = 0; b < nb; b++) {
            const int band = ch.band < 0 ? b : ch.band;
            ing_write_row<ESZ>(base + (long long)band * band_stride, sb, n, nb, b, swap, threadIdx.x, ING_THREADS);
        }
    }
}

cudaError_t frx_launch_tiff_unpack(const uint8_t *src, const frx_tiff_chunk *chunks_dev, const uint32_t *unit_prefix_dev, int n_chunks,
                                   uint32_t total_units, int spp, int swap, const frx_raster &dst, cudaStream_t st) {
    if (!total_units) return cudaSuccess;
    const int esz = frx_dtype_size(dst.dtype);
    uint8_t *out = (uint8_t *)const_cast<void *>(dst.data);
    switch (esz) {
        case 1: frx_tiff_unpack_kernel<1><<<total_units, ING_THREADS, 0, st>>>(src, chunks_dev, unit_prefix_dev, n_chunks, spp, swap, out, dst.band_stride, dst.row_stride); break;
        case 2: frx_tiff_unpack_kernel<2><<<total_units, ING_THREADS, 0, st>>>(src, chunks_dev, unit_prefix_dev, n_chunks, spp, swap, out, dst.band_stride, dst.row_stride); break;
        case 4: frx_tiff_unpack_kernel<4><<<total_units, ING_THREADS, 0, st>>>(src, chunks_dev, unit_prefix_dev, n_chunks, spp, swap, out, dst.band_stride, dst.row_stride); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

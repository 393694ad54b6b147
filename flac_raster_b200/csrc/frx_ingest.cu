// frx_ingest.cu -- uncompressed TIFF strips/tiles (as they lie in the file) -> band-planar device raster.
//
// Replaces the host side of rasterio's read(): GDAL copies every strip/tile into a (bands,h,w) array on the CPU
// (converter.py:137-141; spatial_encoder.py:154-183; cli.py:699) and the caller then ships that array to the GPU.  Here
// the file image goes to the device once and frx_tiff_unpack_kernel de-interleaves, un-tiles, clips and byte-swaps there.
//
// Work unit = one row segment of one chunk (<= SEG_BYTES of source).  The segment is staged in shared memory with
// 16-byte loads (source and staging buffer share their alignment phase, so the bulk is aligned on both sides), then
// written band by band with 16-byte stores where the destination allows.  HBM-bound: 2 x raster bytes per call.
#include "frx_internal.cuh"

#define ING_THREADS 256
#define SEG_BYTES 16384

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

template <int ESZ>
__global__ void __launch_bounds__(ING_THREADS) frx_tiff_unpack_kernel(const uint8_t *__restrict__ src, const frx_tiff_chunk *__restrict__ chunks,
                                                                      const uint32_t *__restrict__ unit_prefix, int n_chunks, int spp, int seg_px,
                                                                      int swap, uint8_t *__restrict__ dst, long long band_stride,
                                                                      long long row_stride) {
    typedef typename IngElem<ESZ>::T T;
    __shared__ __align__(16) uint8_t stage[SEG_BYTES + 32];
    const uint32_t u = blockIdx.x;
    int lo = 0, hi = n_chunks;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (unit_prefix[mid] <= u) lo = mid; else hi = mid;
    }
    const frx_tiff_chunk ch = chunks[lo];
    const uint32_t lu = u - unit_prefix[lo];
    const uint32_t segs = ((uint32_t)ch.cols + (uint32_t)seg_px - 1u) / (uint32_t)seg_px;
    const uint32_t r = lu / segs, sg = lu - r * segs;
    const int c0 = (int)sg * seg_px;
    const int n = min(seg_px, ch.cols - c0);                 // pixels of this segment
    const int nb = ch.band < 0 ? spp : 1;                    // samples per pixel in this chunk
    const size_t nbytes = (size_t)n * nb * ESZ;
    const uint8_t *sp = src + ch.src_offset + ((size_t)r * ch.pitch + (size_t)c0) * nb * ESZ;
    // ---- stage: head bytes up to the first 16-byte boundary, aligned body, tail ----
    const uint32_t phase = (uint32_t)((uintptr_t)sp & 15u);
    uint8_t *sb = stage + phase;                             // sb and sp are congruent mod 16
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
    // ---- write out, band by band ----
    const bool el_ok = (phase % ESZ) == 0;                   // staged elements are naturally aligned
    for (int b = 0; b < nb; b++) {
        const int band = ch.band < 0 ? b : ch.band;
        T *dp = reinterpret_cast<T *>(dst) + (long long)band * band_stride + (long long)(ch.row0 + (int)r) * row_stride + ch.col0 + c0;
        auto elem = [&](int i) -> T {
            const uint8_t *e = sb + ((size_t)i * nb + b) * ESZ;
            T v;
            if (el_ok) v = *reinterpret_cast<const T *>(e);
            else { v = 0; for (int k = 0; k < ESZ; k++) v |= (T)((T)e[k] << (8 * k)); }
            return ing_swap<ESZ>(v, swap);
        };
        constexpr int VEC = 16 / ESZ;
        const int dhead = min(n, (int)(((16u - (uint32_t)((uintptr_t)dp & 15u)) & 15u) / ESZ));
        if ((int)threadIdx.x < dhead) dp[threadIdx.x] = elem(threadIdx.x);
        const int nvec = (n - dhead) / VEC;
        for (int v = threadIdx.x; v < nvec; v += ING_THREADS) {
            T tmp[VEC];
#pragma unroll
            for (int k = 0; k < VEC; k++) tmp[k] = elem(dhead + v * VEC + k);
            uint4 q;
            memcpy(&q, tmp, 16);
            *reinterpret_cast<uint4 *>(dp + dhead + v * VEC) = q;
        }
        const int t0 = dhead + nvec * VEC;
        if ((int)threadIdx.x < n - t0) dp[t0 + threadIdx.x] = elem(t0 + threadIdx.x);
    }
}

int frx_ingest_seg_pixels(int esz, int spp) {
    int px = SEG_BYTES / (esz * spp);
    px &= ~15;                                               // keeps segment starts 16-byte aligned in the destination
    return px < 16 ? 16 : px;
}

cudaError_t frx_launch_tiff_unpack(const uint8_t *src, const frx_tiff_chunk *chunks_dev, const uint32_t *unit_prefix_dev, int n_chunks,
                                   uint32_t total_units, int spp, int swap, const frx_raster &dst, cudaStream_t st) {
    if (!total_units) return cudaSuccess;
    const int esz = frx_dtype_size(dst.dtype);
    const int seg_px = frx_ingest_seg_pixels(esz, spp);
    uint8_t *out = (uint8_t *)const_cast<void *>(dst.data);
    switch (esz) {
        case 1: frx_tiff_unpack_kernel<1><<<total_units, ING_THREADS, 0, st>>>(src, chunks_dev, unit_prefix_dev, n_chunks, spp, seg_px, swap, out, dst.band_stride, dst.row_stride); break;
        case 2: frx_tiff_unpack_kernel<2><<<total_units, ING_THREADS, 0, st>>>(src, chunks_dev, unit_prefix_dev, n_chunks, spp, seg_px, swap, out, dst.band_stride, dst.row_stride); break;
        case 4: frx_tiff_unpack_kernel<4><<<total_units, ING_THREADS, 0, st>>>(src, chunks_dev, unit_prefix_dev, n_chunks, spp, seg_px, swap, out, dst.band_stride, dst.row_stride); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

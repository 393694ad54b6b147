// frx_decode.cu -- FLAC frames -> raster on sm_100a (batched byte-range decode).
//
// Kernels (SURVEY.md section 8a ids):
//   K9a frx_frame_pos_kernel     frame byte positions from a frame-size table (our own files)
//   K9b frx_frame_discover_kernel sync-code scan + header/CRC-8 validation (reference-produced files carry
//                                no seek table and STREAMINFO frame sizes are 0)
//   K10-K12 frx_decode_kernel    one thread per frame: header parse, Rice/Rice2/escape unpack, FIXED/LPC
//                                restore, wasted-bit shift, CRC-16 check, PCM -> raster dtype
//                                (pyflac.FileDecoder.process + converter.py:88-110, 265-273)
// Rice decoding and the LPC recurrence are serial inside a subframe; the parallelism is across frames
// (config 5: 10^4 queries x 256 frames), so one thread owns one frame and lanes never diverge on data
// except at escape partitions.
#include "frx_internal.cuh"


#define FRX_NOPOS 0xFFFFFFFFFFFFFFFFull

__device__ __forceinline__ int frx_size_code_bps(uint32_t c) {
    switch (c) { case 1: return 8; case 2: return 12; case 4: return 16; case 5: return 20; case 6: return 24; case 7: return 32; case 0: return 0; }
    return -1;
}

__device__ __forceinline__ int dec_find_tile(const uint32_t *prefix, int n, uint32_t f) {
    int lo = 0, hi = n;
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (prefix[mid] <= f) lo = mid; else hi = mid;
    }
    return lo;
}

// ---- K9a: one warp per tile, chunked shuffle scan over the tile's frame sizes ----
__global__ void frx_frame_pos_kernel(DecArgs a) {
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= a.n_tiles) return;
    const frx_dtile t = a.tiles[warp];
    uint32_t f0 = a.dprefix[warp], f1 = a.dprefix[warp + 1];
    if (t.first_frame == 0xFFFFFFFFu) {
        for (uint32_t f = f0 + lane; f < f1; f += 32) a.frame_pos[f] = FRX_NOPOS;
        return;
    }
    uint64_t base = t.byte_offset;
    for (uint32_t c = 0; c < f1 - f0; c += 32) {
        uint32_t j = c + lane;
        uint64_t sz = (j < f1 - f0) ? a.frame_size[t.first_frame + j] : 0;
        uint64_t incl = sz;
        for (int o = 1; o < 32; o <<= 1) {
            uint64_t n = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= o) incl += n;
        }
        if (j < f1 - f0) a.frame_pos[f0 + j] = base + incl - sz;
        base += __shfl_sync(0xFFFFFFFFu, incl, 31);
    }
}

// ---- K9b: sync scan.  grid.x = units of 4096 bytes over tiles that need discovery ----
__global__ void frx_frame_discover_kernel(DecArgs a, const uint32_t *unit_prefix, const int32_t *unit_tiles, int n_scan_tiles) {
    // locate scan tile
    uint32_t u = blockIdx.x;
    int k = dec_find_tile(unit_prefix, n_scan_tiles, u);
    int ti = unit_tiles[k];
    const frx_dtile t = a.tiles[ti];
    uint64_t lo = (uint64_t)(u - unit_prefix[k]) * 4096ull;
    uint32_t f0 = a.dprefix[ti];
    const uint8_t *p = a.bytes + t.byte_offset;
    for (uint32_t i = threadIdx.x; i < 4096u; i += blockDim.x) {
        uint64_t pos = lo + i;
        if (pos + 6 > t.byte_size) break;
        if (p[pos] != 0xFF || p[pos + 1] != 0xF8) continue;
        uint32_t b2 = p[pos + 2], b3 = p[pos + 3];
        uint32_t bs_code = b2 >> 4, sr_code = b2 & 15u, ch_code = b3 >> 4, sz_code = (b3 >> 1) & 7u;
        if (b3 & 1u) continue;
        if (bs_code == 0 || sr_code == 15u) continue;
        if (ch_code > 10u) continue;
        int channels = ch_code < 8u ? (int)ch_code + 1 : 2;
        if (channels != t.channels) continue;
        int bps = frx_size_code_bps(sz_code);
        if (bps < 0 || (bps != 0 && bps != t.bps)) continue;
        // utf-8 number
        uint32_t n = 4;
        uint32_t x = p[pos + n++];
        uint64_t number;
        int extra;
        if (!(x & 0x80u)) { number = x; extra = 0; }
        else if ((x & 0xE0u) == 0xC0u) { number = x & 0x1Fu; extra = 1; }
        else if ((x & 0xF0u) == 0xE0u) { number = x & 0x0Fu; extra = 2; }
        else if ((x & 0xF8u) == 0xF0u) { number = x & 0x07u; extra = 3; }
        else if ((x & 0xFCu) == 0xF8u) { number = x & 0x03u; extra = 4; }
        else if ((x & 0xFEu) == 0xFCu) { number = x & 0x01u; extra = 5; }
        else continue;
        if (pos + n + extra + 3 > t.byte_size) continue;
        bool ok = true;
        for (int e = 0; e < extra; e++) {
            uint32_t y = p[pos + n++];
            if ((y & 0xC0u) != 0x80u) { ok = false; break; }
            number = (number << 6) | (y & 0x3Fu);
        }
        if (!ok || number >= t.n_frames) continue;
        uint32_t blocksize;
        if (bs_code == 1) blocksize = 192;
        else if (bs_code <= 5) blocksize = 576u << (bs_code - 2);
        else if (bs_code == 6) { blocksize = (uint32_t)p[pos + n] + 1; n += 1; }
        else if (bs_code == 7) { blocksize = (((uint32_t)p[pos + n] << 8) | p[pos + n + 1]) + 1; n += 2; }
        else blocksize = 256u << (bs_code - 8);
        if (sr_code == 12u) n += 1; else if (sr_code == 13u || sr_code == 14u) n += 2;
        if (pos + n + 1 > t.byte_size) continue;
        long long total = (long long)t.h * t.w;
        long long expect = (number + 1 < t.n_frames) ? a.blocksize : total - (long long)number * a.blocksize;
        if ((long long)blocksize != expect) continue;
        uint32_t c = 0;
        for (uint32_t q = 0; q < n; q++) c = frx_crc8_byte(c, p[pos + q]);
        if (c != p[pos + n]) continue;
        atomicMin((unsigned long long *)&a.frame_pos[f0 + (uint32_t)number], (unsigned long long)(t.byte_offset + pos));
    }
}


// frx_decode.cu -- FLAC frames -> raster on sm_100a (batched byte-range decode).
//
// Kernels (SURVEY.md section 8a ids):
//   K9a frx_frame_pos_kernel     frame byte positions from a frame-size table (our own files)
//   K9b frx_frame_discover_kernel sync-code scan + header/CRC-8 validation (reference-produced files carry
//                                no seek table and STREAMINFO frame sizes are 0)
//   K10-K12 frx_decode_kernel    one lane per frame, 32 frames of a warp in lockstep: header parse, Rice/Rice2/escape
//                                unpack, FIXED/LPC restore, wasted-bit shift, PCM -> raster dtype, stores transposed
//                                through shared memory into full lines (pyflac.FileDecoder.process +
//                                converter.py:88-110, 265-273)
//   K10c frx_crc_check_kernel    CRC-16 of every frame, one warp per frame, coalesced
// Rice decoding and the LPC recurrence are serial inside a subframe; the parallelism is across frames
// (config 5: 10^4 queries x 256 frames).
#include "frx_internal.cuh"

#define FRX_DEC_CTAS 20
#define FRX_DEC_RINGW 32
#define FRX_DEC_NS dec_w20
#include "frx_decode_frame.cuh"

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


// ---- CRC-16 of every decoded frame: one warp per frame, coalesced reads.  Lane l takes a contiguous run of Lq 16-byte
// quads (the run of lane 0 may be short: virtual leading zeros do not change a CRC); the 32 chunk CRCs are moved to
// their place with x^(128 Lq j) mod P -- the encoder's table -- and xor-ed.  A frame including its stored CRC gives 0.
__global__ void __launch_bounds__(128) frx_crc_check_kernel(const DecArgs a, const uint16_t *xpow) {
    __shared__ uint16_t crc_tab[4][256];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) (&crc_tab[0][0])[i] = a.crc_tab[i];
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t f = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; f < a.total_frames; f += nwarps) {
        const uint32_t fsize = a.frame_len[f];
        if (fsize == 0) continue;                       // the decoder already flagged it
        const uint8_t *p = a.bytes + a.frame_pos[f];
        uint32_t crc = 0;                               // running CRC over the passes (frames above 32 * 48 quads)
        uint32_t done = 0;
        while (done < fsize) {
            const uint32_t pass = min(fsize - done, 32u * 16u * (uint32_t)FRX_CRC_MAXLQ);
            const uint32_t Lq = (pass + 511u) / 512u, C = Lq * 16u;
            // lane l covers pass bytes [pass - (32 - l) C, pass - (31 - l) C), clipped at 0
            const long long b1 = (long long)pass - (long long)(31 - lane) * C;
            long long b0 = b1 - C;
            if (b0 < 0) b0 = 0;
            uint32_t c = 0;
            if (b1 > 0) {
                const uint8_t *q = p + done + b0;
                long long n = b1 - b0;
                while (n > 0 && ((uintptr_t)q & 3u)) { c = ((c << 8) ^ crc_tab[0][((c >> 8) ^ *q) & 0xFFu]) & 0xFFFFu; q++; n--; }
                for (; n >= 4; n -= 4, q += 4) {
                    const uint32_t x = frx_bswap32(__ldg(reinterpret_cast<const uint32_t *>(q))) ^ (c << 16);
                    c = (uint32_t)crc_tab[3][x >> 24] ^ crc_tab[2][(x >> 16) & 0xFFu] ^ crc_tab[1][(x >> 8) & 0xFFu] ^ crc_tab[0][x & 0xFFu];
                }
                for (; n > 0; n--, q++) c = ((c << 8) ^ crc_tab[0][((c >> 8) ^ *q) & 0xFFu]) & 0xFFFFu;
            }
            if (c && lane != 31) c = frx_gf16_mulmod(c, (uint32_t)__ldg(&xpow[(Lq - 1) * 256 + (31 - lane)]));
            c = __reduce_xor_sync(0xFFFFFFFFu, c);
            // crc <- crc * x^(8 pass) + c ; x^(8 pass) = x^(128 Lq 32) when the pass is full, and a full pass is the only one followed by another
            if (done) {
                // previous passes were full: shift the running CRC by this pass's length, 8 bits at a time is too slow -> square-and-multiply
                uint32_t e = pass * 8u, base = 2u, acc = 1u;
                while (e) { if (e & 1u) acc = frx_gf16_mulmod(acc, base); base = frx_gf16_mulmod(base, base); e >>= 1; }
                crc = frx_gf16_mulmod(crc, acc) ^ c;
            } else crc = c;
            done += pass;
        }
        if (crc != 0 && lane == 0) {
            const int ti = dec_find_tile(a.dprefix, a.n_tiles, f);
            atomicExch(&a.tile_status[ti], FRX_E_BITSTREAM);
        }
    }
}

// K12 alone
template <int DT>
__global__ void frx_dequantize_kernel(const int32_t *pcm, int h, int w, int channels, int bps, int mode, FrxDequant dq, void *out) {
    typedef typename FrxElem<DT>::T T;
    long long n = (long long)h * w * channels;
    long long plane = (long long)h * w;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        int c = (int)(i % channels);
        long long px = i / channels;
        ((T *)out)[(long long)c * plane + px] = frx_dequant<DT>(pcm[i], bps, dq, mode);
    }
}

// ------------------------------------------------------------------------------------------
// host launchers
// ------------------------------------------------------------------------------------------
void frx_make_crc16_tables(uint16_t *tab /* [4][256] */) {
    for (int i = 0; i < 256; i++) {
        uint32_t c = (uint32_t)i << 8;
        for (int b = 0; b < 8; b++) c = (c & 0x8000u) ? ((c << 1) ^ 0x8005u) & 0xFFFFu : (c << 1) & 0xFFFFu;
        tab[i] = (uint16_t)c;
    }
    for (int k = 1; k < 4; k++)
        for (int i = 0; i < 256; i++) {
            uint32_t c = tab[(k - 1) * 256 + i];
            tab[k * 256 + i] = (uint16_t)(((c << 8) ^ tab[(c >> 8) & 0xFFu]) & 0xFFFFu);
        }
}

cudaError_t frx_launch_frame_pos(const DecArgs &a, cudaStream_t st) {
    int threads = 128, warps_per_block = threads / 32;
    frx_frame_pos_kernel<<<(a.n_tiles + warps_per_block - 1) / warps_per_block, threads, 0, st>>>(a);
    return cudaGetLastError();
}

cudaError_t frx_launch_discover(const DecArgs &a, const uint32_t *unit_prefix, const int32_t *unit_tiles, int n_scan_tiles,
                                uint32_t total_units, cudaStream_t st) {
    if (total_units) frx_frame_discover_kernel<<<total_units, 256, 0, st>>>(a, unit_prefix, unit_tiles, n_scan_tiles);
    return cudaGetLastError();
}

cudaError_t frx_launch_crc_check(const DecArgs &a, const uint16_t *xpow, int sm_count, cudaStream_t st) {
    if (!a.total_frames) return cudaSuccess;
    unsigned want = (a.total_frames + 3u) / 4u, cap = (unsigned)sm_count * 16u;
    frx_crc_check_kernel<<<want < cap ? want : cap, 128, 0, st>>>(a, xpow);
    return cudaGetLastError();
}

cudaError_t frx_launch_decode_w22(const DecArgs &a, int out_dtype, cudaStream_t st);   // frx_decode_w22.cu

// Which copy: a frame takes milliseconds on its lane, so what matters is whether the call fits ONE wave.  The 22-per-SM copy
// exactly when it turns two waves into one (C2: 20.6 warps per SM); otherwise the 96-register copy (C3: 63 vs 68 ms).
// FRX_DECODE_COPY=20|22 forces one (A/B runs, tests of both).
cudaError_t frx_launch_decode(const DecArgs &a, int out_dtype, int sm_count, cudaStream_t st) {
    const uint32_t warps = (a.total_frames + 31u) / 32u;
    static const int forced = getenv("FRX_DECODE_COPY") ? atoi(getenv("FRX_DECODE_COPY")) : 0;
    if (forced == 22) return frx_launch_decode_w22(a, out_dtype, st);
    if (forced == 20) return dec_w20::launch(a, out_dtype, st);
    if (warps > 20u * (uint32_t)sm_count && warps <= 22u * (uint32_t)sm_count) return frx_launch_decode_w22(a, out_dtype, st);
    return dec_w20::launch(a, out_dtype, st);
}

cudaError_t frx_launch_dequantize(const int32_t *pcm, int h, int w, int channels, int bps, int out_dtype, int mode, double dmin,
                                  double dmax, void *out, cudaStream_t st) {
    FrxDequant dq;
    dq.range_f = (float)(dmax - dmin);
    dq.dmin_f = (float)dmin;
    long long n = (long long)h * w * channels;
    int grid = (int)((n + 255) / 256 < 148 * 8 ? (n + 255) / 256 : 148 * 8);
    if (grid < 1) grid = 1;
    switch (out_dtype) {
        case FRX_U8: frx_dequantize_kernel<FRX_U8><<<grid, 256, 0, st>>>(pcm, h, w, channels, bps, mode, dq, out); break;
        case FRX_I16: frx_dequantize_kernel<FRX_I16><<<grid, 256, 0, st>>>(pcm, h, w, channels, bps, mode, dq, out); break;
        case FRX_U16: frx_dequantize_kernel<FRX_U16><<<grid, 256, 0, st>>>(pcm, h, w, channels, bps, mode, dq, out); break;
        case FRX_I32: frx_dequantize_kernel<FRX_I32><<<grid, 256, 0, st>>>(pcm, h, w, channels, bps, mode, dq, out); break;
        case FRX_U32: frx_dequantize_kernel<FRX_U32><<<grid, 256, 0, st>>>(pcm, h, w, channels, bps, mode, dq, out); break;
        case FRX_F32: frx_dequantize_kernel<FRX_F32><<<grid, 256, 0, st>>>(pcm, h, w, channels, bps, mode, dq, out); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

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

// ------------------------------------------------------------------------------------------
// bit reader: 64-bit window refilled with aligned, bounds-clamped 32-bit words
// ------------------------------------------------------------------------------------------
struct BitR {
    const uint8_t *lo, *hi;     // valid byte range of the buffer
    const uint8_t *wp;          // next 4-byte aligned word to load
    uint64_t acc;  // left aligned
    int cnt;       // valid bits in acc
    uint64_t consumed;  // bits consumed since init
    __device__ __forceinline__ uint32_t load_word() {
        uint32_t w;
        if (wp >= lo && wp + 4 <= hi) w = frx_bswap32(__ldg((const uint32_t *)wp));
        else {
            w = 0;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const uint8_t *q = wp + i;
                uint32_t b = (q >= lo && q < hi) ? (uint32_t)__ldg(q) : 0u;
                w |= b << (24 - 8 * i);
            }
        }
        wp += 4;
        return w;
    }
    __device__ __forceinline__ void init(const uint8_t *base, uint64_t n_bytes, uint64_t pos) {
        lo = base; hi = base + n_bytes;
        uintptr_t addr = (uintptr_t)(base + pos);
        int lead = (int)(addr & 3u);
        wp = (const uint8_t *)(addr - lead);
        consumed = 0;
        uint32_t w = load_word();
        acc = (uint64_t)w << (32 + 8 * lead);
        cnt = 32 - 8 * lead;
        refill();
    }
    __device__ __forceinline__ void refill() {
        if (cnt <= 32) {
            uint32_t w = load_word();
            acc |= (uint64_t)w << (32 - cnt);
            cnt += 32;
        }
    }
    // n in 0..32
    __device__ __forceinline__ uint32_t get(int n) {
        refill();
        uint32_t v = n ? (uint32_t)(acc >> (64 - n)) : 0u;
        acc = n >= 64 ? 0 : (acc << n);
        cnt -= n; consumed += (uint64_t)n;
        return v;
    }
    __device__ __forceinline__ int32_t sget(int n) {  // signed, n in 1..32
        uint32_t v = get(n);
        return n >= 32 ? (int32_t)v : ((int32_t)(v << (32 - n)) >> (32 - n));
    }
    __device__ __forceinline__ uint32_t unary() {
        uint32_t q = 0;
        for (;;) {
            refill();
            if (acc == 0) {
                q += (uint32_t)cnt; consumed += (uint64_t)cnt; cnt = 0;
                if (wp > hi + 8) return q;  // ran off the buffer: malformed
                continue;
            }
            int lz = __clzll((long long)acc);
            q += (uint32_t)lz;
            acc = lz == 63 ? 0ull : (acc << (lz + 1));  // a set bit exists within cnt bits, so lz < cnt
            cnt -= lz + 1; consumed += (uint64_t)lz + 1u;
            return q;
        }
    }
};

// ------------------------------------------------------------------------------------------
// raster output cursor: walks the tile row-major, one element per put()
// ------------------------------------------------------------------------------------------
template <int DT>
struct OutW {
    typedef typename FrxElem<DT>::T T;
    T *base;           // element pointer to (band c,row 0,col 0) of the tile
    long long row_stride;
    int w, row, col;
    __device__ __forceinline__ void init(T *b, long long rs, int width, long long px) {
        base = b; row_stride = rs; w = width;
        row = (int)(px / width); col = (int)(px - (long long)row * width);
    }
    __device__ __forceinline__ void put(T v) {
        base[(long long)row * row_stride + col] = v;
        if (++col == w) { col = 0; row++; }
    }
};

// ------------------------------------------------------------------------------------------
// residual + predictor restore for one subframe, ORD-tap ring held in registers
// ------------------------------------------------------------------------------------------
struct RiceState {
    uint32_t k;        // current parameter
    uint32_t left;     // samples left in the current partition
    int pbits;         // 4 or 5
    uint32_t esc;      // 15 or 31
    int rawbits;       // >= 0: escape partition with this many raw bits
    uint32_t psize;    // samples per partition
};

__device__ __forceinline__ int32_t rice_next(BitR &br, RiceState &rs) {
    if (rs.left == 0) {
        rs.k = br.get(rs.pbits);
        rs.left = rs.psize;
        rs.rawbits = -1;
        if (rs.k == rs.esc) rs.rawbits = (int)br.get(5);
    }
    rs.left--;
    if (rs.rawbits >= 0) return rs.rawbits ? br.sget(rs.rawbits) : 0;
    uint32_t q = br.unary();
    uint32_t u = (q << rs.k) | br.get((int)rs.k);
    return (int32_t)(u >> 1) ^ -(int32_t)(u & 1u);
}

template <int DT, int ORD, bool WIDE>
__device__ __forceinline__ void restore_subframe(BitR &br, RiceState &rs, OutW<DT> &ow, const int32_t *coef, int order, int shift,
                                                 const int32_t *warm, int nsamp, int wasted, int bps, const FrxDequant &dq, int mode) {
    int32_t h[ORD];
#pragma unroll
    for (int i = 0; i < ORD; i++) h[i] = i < order ? warm[i] : 0;
    for (int i = 0; i < order; i++) ow.put(frx_dequant<DT>((int32_t)((uint32_t)warm[i] << wasted), bps, dq, mode));
    for (int n0 = 0; n0 < nsamp; n0 += ORD) {
#pragma unroll
        for (int m = 0; m < ORD; m++) {
            int n = n0 + m;
            if (n >= order && n < nsamp) {
                int32_t r = rice_next(br, rs);
                int32_t v;
                if (WIDE) {
                    long long acc = 0;
#pragma unroll
                    for (int j = 0; j < ORD; j++) acc += (long long)coef[j] * (long long)h[(m - 1 - j + 2 * ORD) % ORD];
                    v = (int32_t)((long long)r + (acc >> shift));
                } else {
                    int32_t acc = 0;
#pragma unroll
                    for (int j = 0; j < ORD; j++) acc += coef[j] * h[(m - 1 - j + 2 * ORD) % ORD];
                    v = r + (acc >> shift);
                }
                h[m] = v;
                ow.put(frx_dequant<DT>((int32_t)((uint32_t)v << wasted), bps, dq, mode));
            }
        }
    }
}

// generic order (13..32): history ring in local memory
template <int DT>
__device__ void restore_subframe_generic(BitR &br, RiceState &rs, OutW<DT> &ow, const int32_t *coef, int order, int shift,
                                         const int32_t *warm, int nsamp, int wasted, int bps, const FrxDequant &dq, int mode) {
    int32_t h[32];
    for (int i = 0; i < 32; i++) h[i] = i < order ? warm[i] : 0;
    for (int i = 0; i < order; i++) ow.put(frx_dequant<DT>((int32_t)((uint32_t)warm[i] << wasted), bps, dq, mode));
    for (int n = order; n < nsamp; n++) {
        int32_t r = rice_next(br, rs);
        long long acc = 0;
        for (int j = 0; j < order; j++) acc += (long long)coef[j] * (long long)h[(n - 1 - j) & 31];
        int32_t v = (int32_t)((long long)r + (acc >> shift));
        h[n & 31] = v;
        ow.put(frx_dequant<DT>((int32_t)((uint32_t)v << wasted), bps, dq, mode));
    }
}

__device__ __forceinline__ int ilog2_u32(uint32_t v) { return 31 - __clz((int)v); }

template <int DT>
__global__ void __launch_bounds__(128) frx_decode_kernel(const DecArgs a) {
    typedef typename FrxElem<DT>::T T;
    __shared__ uint16_t crc_tab[4][256];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) (&crc_tab[0][0])[i] = a.crc_tab[i];
    __syncthreads();
    uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= a.total_frames) return;
    const int ti = dec_find_tile(a.dprefix, a.n_tiles, f);
    const frx_dtile t = a.tiles[ti];
    const uint32_t fj = f - a.dprefix[ti];
    const uint64_t pos = a.frame_pos[f];
    if (pos == FRX_NOPOS || pos < t.byte_offset || pos >= t.byte_offset + t.byte_size) { atomicExch(&a.tile_status[ti], FRX_E_BITSTREAM); return; }
    uint64_t end = t.byte_offset + t.byte_size;
    if (fj + 1 < t.n_frames) {
        uint64_t nx = a.frame_pos[f + 1];
        if (nx != FRX_NOPOS && nx > pos && nx <= end) end = nx;
    }
    const long long tile_n = (long long)t.h * t.w;
    const long long s0 = (long long)fj * a.blocksize;
    const int expect = (int)min((long long)a.blocksize, tile_n - s0);
    FrxDequant dq;
    dq.range_f = (float)(t.dmax - t.dmin);
    dq.dmin_f = (float)t.dmin;

    BitR br;
    br.init(a.bytes, a.n_bytes, pos);
    bool bad = false;
    // ---- header ----
    uint32_t sync = br.get(16);
    if (sync != 0xFFF8u) bad = true;
    uint32_t bs_code = br.get(4), sr_code = br.get(4), ch_code = br.get(4), sz_code = br.get(3);
    br.get(1);
    {
        uint32_t x = br.get(8);
        int extra = !(x & 0x80u) ? 0 : (x & 0xE0u) == 0xC0u ? 1 : (x & 0xF0u) == 0xE0u ? 2 : (x & 0xF8u) == 0xF0u ? 3 : (x & 0xFCu) == 0xF8u ? 4 : 5;
        for (int e = 0; e < extra; e++) br.get(8);
    }
    int nsamp;
    if (bs_code == 1) nsamp = 192;
    else if (bs_code >= 2 && bs_code <= 5) nsamp = 576 << (bs_code - 2);
    else if (bs_code == 6) nsamp = (int)br.get(8) + 1;
    else if (bs_code == 7) nsamp = (int)br.get(16) + 1;
    else if (bs_code >= 8) nsamp = 256 << (bs_code - 8);
    else { nsamp = 0; bad = true; }
    if (sr_code == 12u) br.get(8); else if (sr_code == 13u || sr_code == 14u) br.get(16);
    br.get(8);  // CRC-8 (validated by discovery for foreign files; CRC-16 below covers the header too)
    int bps = sz_code ? frx_size_code_bps(sz_code) : t.bps;
    int channels = ch_code < 8u ? (int)ch_code + 1 : 2;
    if (nsamp != expect || bps != t.bps || channels != t.channels || ch_code >= 8u) bad = true;  // stereo decorrelation: see frx_api (unsupported)
    if (bad) { atomicExch(&a.tile_status[ti], ch_code >= 8u && ch_code <= 10u ? FRX_E_UNSUPPORTED : FRX_E_BITSTREAM); return; }

    T *obase = (T *)a.out + t.out_offset;
    for (int c = 0; c < channels && !bad; c++) {
        OutW<DT> ow;
        ow.init(obase + (long long)c * t.out_band_stride, t.out_row_stride, t.w, s0);
        uint32_t hdr = br.get(8);
        if (hdr & 0x80u) { bad = true; break; }
        uint32_t type = (hdr >> 1) & 0x3Fu;
        int wasted = 0;
        if (hdr & 1u) wasted = (int)br.unary() + 1;
        int sbps = bps - wasted;
        if (sbps <= 0) { bad = true; break; }
        if (type == 0) {
            int32_t v = br.sget(sbps);
            T o = frx_dequant<DT>((int32_t)((uint32_t)v << wasted), bps, dq, a.pcm_mode);
            for (int i = 0; i < nsamp; i++) ow.put(o);
        } else if (type == 1) {
            for (int i = 0; i < nsamp; i++) {
                int32_t v = br.sget(sbps);
                ow.put(frx_dequant<DT>((int32_t)((uint32_t)v << wasted), bps, dq, a.pcm_mode));
            }
        } else {
            int order, shift = 0, prec = 0;
            int32_t coef[32];
            int32_t warm[32];
            bool lpc = false;
            if (type >= 8 && type <= 12) order = (int)type - 8;
            else if (type >= 32) { order = (int)(type & 31u) + 1; lpc = true; }
            else { bad = true; break; }
            if (order > nsamp) { bad = true; break; }
            for (int i = 0; i < order; i++) warm[i] = br.sget(sbps);
            for (int j = 0; j < 32; j++) coef[j] = 0;
            if (lpc) {
                prec = (int)br.get(4) + 1;
                shift = br.sget(5);
                if (prec == 16 || shift < 0) { bad = true; break; }
                for (int j = 0; j < order; j++) coef[j] = br.sget(prec);
            } else {
                if (order == 1) { coef[0] = 1; }
                else if (order == 2) { coef[0] = 2; coef[1] = -1; }
                else if (order == 3) { coef[0] = 3; coef[1] = -3; coef[2] = 1; }
                else if (order == 4) { coef[0] = 4; coef[1] = -6; coef[2] = 4; coef[3] = -1; }
                prec = 3;
            }
            // residual header
            uint32_t method = br.get(2);
            uint32_t po = br.get(4);
            if (method > 1u || ((uint32_t)nsamp >> po) << po != (uint32_t)nsamp && po > 0) { bad = true; break; }
            if (((uint32_t)nsamp >> po) < (uint32_t)order) { bad = true; break; }
            RiceState rs;
            rs.pbits = method ? 5 : 4; rs.esc = method ? 31u : 15u; rs.psize = (uint32_t)nsamp >> po;
            rs.left = 0; rs.k = 0; rs.rawbits = -1;
            // the first partition is short by `order`: prime it here
            rs.k = br.get(rs.pbits);
            rs.left = rs.psize - (uint32_t)order;
            if (rs.k == rs.esc) rs.rawbits = (int)br.get(5);
            const bool wide = order > 0 && (sbps + prec + ilog2_u32((uint32_t)order) > 32);
            if (order <= 4) {
                if (wide) restore_subframe<DT, 4, true>(br, rs, ow, coef, order, shift, warm, nsamp, wasted, bps, dq, a.pcm_mode);
                else restore_subframe<DT, 4, false>(br, rs, ow, coef, order, shift, warm, nsamp, wasted, bps, dq, a.pcm_mode);
            } else if (order <= 8) {
                if (wide) restore_subframe<DT, 8, true>(br, rs, ow, coef, order, shift, warm, nsamp, wasted, bps, dq, a.pcm_mode);
                else restore_subframe<DT, 8, false>(br, rs, ow, coef, order, shift, warm, nsamp, wasted, bps, dq, a.pcm_mode);
            } else if (order <= 12) {
                if (wide) restore_subframe<DT, 12, true>(br, rs, ow, coef, order, shift, warm, nsamp, wasted, bps, dq, a.pcm_mode);
                else restore_subframe<DT, 12, false>(br, rs, ow, coef, order, shift, warm, nsamp, wasted, bps, dq, a.pcm_mode);
            } else {
                restore_subframe_generic<DT>(br, rs, ow, coef, order, shift, warm, nsamp, wasted, bps, dq, a.pcm_mode);
            }
        }
    }
    // ---- footer: byte align + CRC-16 ----
    uint64_t body_bytes = (br.consumed + 7u) >> 3;
    uint64_t fsize = body_bytes + 2u;
    if (!bad && pos + fsize > end) bad = true;
    if (!bad && fj + 1 < t.n_frames && a.frame_pos[f + 1] != FRX_NOPOS && pos + fsize != a.frame_pos[f + 1]) bad = true;
    if (!bad && a.verify_crc) {
        // CRC over the whole frame including the stored CRC must be 0
        const uint8_t *p = a.bytes + pos;
        uint32_t c = 0;
        uint64_t i = 0;
        while (i < fsize && ((uintptr_t)(p + i) & 3u)) { c = ((c << 8) ^ crc_tab[0][((c >> 8) ^ p[i]) & 0xFFu]) & 0xFFFFu; i++; }
        for (; i + 4 <= fsize; i += 4) {
            uint32_t w = frx_bswap32(__ldg((const uint32_t *)(p + i)));
            uint32_t x = w ^ (c << 16);
            c = (uint32_t)crc_tab[3][x >> 24] ^ crc_tab[2][(x >> 16) & 0xFFu] ^ crc_tab[1][(x >> 8) & 0xFFu] ^ crc_tab[0][x & 0xFFu];
        }
        for (; i < fsize; i++) c = ((c << 8) ^ crc_tab[0][((c >> 8) ^ p[i]) & 0xFFu]) & 0xFFFFu;
        if (c != 0) bad = true;
    }
    if (bad) atomicExch(&a.tile_status[ti], FRX_E_BITSTREAM);
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

cudaError_t frx_launch_decode(const DecArgs &a, int out_dtype, cudaStream_t st) {
    unsigned grid = (a.total_frames + 127u) / 128u;
    if (!grid) return cudaSuccess;
    switch (out_dtype) {
        case FRX_U8: frx_decode_kernel<FRX_U8><<<grid, 128, 0, st>>>(a); break;
        case FRX_I16: frx_decode_kernel<FRX_I16><<<grid, 128, 0, st>>>(a); break;
        case FRX_U16: frx_decode_kernel<FRX_U16><<<grid, 128, 0, st>>>(a); break;
        case FRX_I32: frx_decode_kernel<FRX_I32><<<grid, 128, 0, st>>>(a); break;
        case FRX_U32: frx_decode_kernel<FRX_U32><<<grid, 128, 0, st>>>(a); break;
        case FRX_F32: frx_decode_kernel<FRX_F32><<<grid, 128, 0, st>>>(a); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
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

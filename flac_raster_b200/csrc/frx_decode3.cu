// frx_decode3.cu -- the phased frame decoder (K10-K12 as four kernels), used for small multi-channel calls (frx_api.cu decode_core).
//
// Rice codes and the predictor recurrence are both serial inside a subframe, but they are serial in different ways, and
// the single-kernel decoder (frx_decode_frame.cuh: one lane walks one frame from its first bit to its last sample) pays for
// both with one lane: a frame takes ~3 ms on its lane, which is also the latency of the smallest call.  Here the work is
// cut where the dependencies allow it:
//   A  scan     lane = frame.  Parses the headers and walks the residual codes WITHOUT decoding them (find the stop bit,
//               skip k bits), noting the bit position and Rice state at every 128th sample.
//   B  residual lane = one block of 128 samples, warp = subframe.  Decodes the block's residuals from the noted position:
//               3 million independent lanes on C2 instead of 98 thousand.
//   C  restore  lane = subframe (32 subframes of a warp in lockstep).  Only the recurrence is left on the serial path:
//               residual + prediction, a dozen multiply-adds per sample; residuals and samples travel through a 32x33
//               shared-memory transpose so that global memory only sees full lines.
//   D  output   thread = sample.  Stereo recombination, dequantisation, raster geometry, coalesced stores.
// Residuals / PCM live in an int32 scratch row per subframe between the phases (EncArgs-like D3Args, owned by the context).
// The bitstream semantics are those of frx_decode_frame.cuh (RFC 9639; reference: pyflac's FileDecoder, dump:1571-1640,
// then converter.py:99-110 for the dequantiser) and the two decoders are tested against each other and the oracle.
#include "frx_decode_common.cuh"

namespace {

#define D3_BLK 128               // samples per residual block
#define D3_MAXBLK 32             // blocks per subframe: blocksize <= 4096
#define D3_RING 32               // words per lane of the scan kernel's input ring

struct D3Sub {                   // one per (frame, channel)
    uint8_t type;                // 0 constant, 1 verbatim, 2 predictive, 3 nothing to do (absent channel / rejected frame)
    uint8_t order, shift, wasted;
    uint8_t sbps, wide, st_mode; // st_mode: 0 plain, 1 first channel of a decorrelated pair, 2/3/4 second channel of L/S, S/R, M/S
    uint8_t pbits;               // 4 or 5
    int32_t cval;
    uint32_t nsamp;
    uint32_t psize;              // samples per Rice partition
    uint32_t tile;
    uint32_t bps;
    long long s0;                // first sample of the frame inside its tile
    unsigned long long pos;      // byte position of the frame
    int16_t coef[32];
    int32_t warm[32];
};
struct D3Block { uint32_t bitoff; uint32_t state; };   // state: k | rawflag << 5 | rawbits << 6 | left << 12

struct RiceState {
    uint32_t k;        // current parameter
    uint32_t left;     // samples left in the current partition
    int pbits;         // 4 or 5
    uint32_t esc;      // 15 or 31
    int rawbits;       // >= 0: escape partition with this many raw bits
    uint32_t psize;    // samples per partition
};

// ---- bit reader of the scan kernel: per-lane ring in shared memory, topped up for all lanes at once (see frx_decode_frame.cuh) ----
__device__ __noinline__ uint4 d3_load16_clamped(const uint8_t *np, const uint8_t *lo, const uint8_t *hi) {
    uint32_t w[4];
    for (int k = 0; k < 4; k++) {
        w[k] = 0;
        for (int i = 0; i < 4; i++) {
            const uint8_t *p = np + 4 * k + i;
            const uint32_t b = (p >= lo && p < hi) ? (uint32_t)__ldg(p) : 0u;
            w[k] |= b << (8 * i);
        }
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
}
struct RingR {
    const uint8_t *lo, *hi, *np;
    uint32_t *ring;
    uint32_t rd, wr;
    uint32_t nx;                // ring word `rd`, fetched ahead so that a refill has no shared-memory load on its critical path
    uint64_t acc;
    int cnt;
    uint32_t start_bits;
    __device__ __forceinline__ uint4 load16() const {
        if (np >= lo && np + 16 <= hi) return __ldg(reinterpret_cast<const uint4 *>(np));
        return d3_load16_clamped(np, lo, hi);
    }
    __device__ __forceinline__ void push16(const uint4 v) {
        ring[((wr + 0u) % D3_RING) * 32] = frx_bswap32(v.x);
        ring[((wr + 1u) % D3_RING) * 32] = frx_bswap32(v.y);
        ring[((wr + 2u) % D3_RING) * 32] = frx_bswap32(v.z);
        ring[((wr + 3u) % D3_RING) * 32] = frx_bswap32(v.w);
        wr += 4; np += 16;
    }
    __device__ __forceinline__ void topup(bool want) {
        for (;;) {
            const uint32_t free_blocks = want ? (D3_RING - (wr - rd)) >> 2 : 0u;
            if (!__any_sync(0xFFFFFFFFu, free_blocks != 0u)) break;
            uint4 v[4];
#pragma unroll
            for (int k = 0; k < 4; k++) if ((uint32_t)k < free_blocks) { const uint8_t *p = np; np += 16 * k; v[k] = load16(); np = p; }
#pragma unroll
            for (int k = 0; k < 4; k++) if ((uint32_t)k < free_blocks) push16(v[k]);
            if (!__any_sync(0xFFFFFFFFu, free_blocks > 4u)) break;
        }
        nx = ring[(rd % D3_RING) * 32];
    }
    __device__ __forceinline__ uint32_t pop() {
        if (rd == wr) { push16(load16()); nx = ring[(rd % D3_RING) * 32]; }     // ring ran dry inside a chunk: fetch on demand
        const uint32_t w = nx;
        rd++;
        nx = ring[(rd % D3_RING) * 32];      // (stale when rd == wr: reloaded when the words arrive)
        return w;
    }
    __device__ __forceinline__ void init(const uint8_t *base, uint64_t n_bytes, uint64_t pos, uint32_t *ring_col, bool live) {
        lo = base; hi = base + n_bytes;
        const uintptr_t addr = (uintptr_t)(base + pos);
        const int lead = (int)(addr & 15u);
        np = (const uint8_t *)(addr - lead);
        ring = ring_col; rd = 0; wr = 0; nx = 0; start_bits = (uint32_t)lead * 8u; acc = 0; cnt = 0;
        topup(live);
        if (live) {
            rd = (uint32_t)(lead >> 2);
            nx = ring[(rd % D3_RING) * 32];
            const int lb = 8 * (lead & 3);
            acc = (uint64_t)pop() << (32 + lb);
            cnt = 32 - lb;
            refill();
        }
    }
    __device__ __forceinline__ uint32_t consumed() const { return rd * 32u - (uint32_t)cnt - start_bits; }
    __device__ __forceinline__ void refill() {
        if (cnt <= 32) {
            acc |= (uint64_t)pop() << (32 - cnt);
            cnt += 32;
        }
    }
    __device__ __forceinline__ bool past_end() const { return np > hi + 256; }
};
// ---- bit reader of the residual kernel: lane-private, straight from global memory (neighbouring lanes read neighbouring
// blocks of the same subframe, so the lines they touch are shared through L1) ----
struct GlobR {
    const uint8_t *lo, *hi, *np;
    uint64_t acc;
    int cnt;
    __device__ __forceinline__ uint32_t pop() {
        uint32_t w;
        if (np >= lo && np + 4 <= hi) w = __ldg(reinterpret_cast<const uint32_t *>(np));
        else {
            w = 0;
            for (int i = 0; i < 4; i++) { const uint8_t *p = np + i; if (p >= lo && p < hi) w |= (uint32_t)__ldg(p) << (8 * i); }
        }
        np += 4;
        return frx_bswap32(w);
    }
    __device__ __forceinline__ void init(const uint8_t *base, uint64_t n_bytes, uint64_t bitpos) {
        lo = base; hi = base + n_bytes;
        const uintptr_t addr = (uintptr_t)(base + (bitpos >> 3));
        const int lead = (int)(addr & 3u);
        np = (const uint8_t *)(addr - lead);
        const int sh = lead * 8 + (int)(bitpos & 7u);            // 0..31
        acc = (uint64_t)pop() << 32;
        acc |= (uint64_t)pop();
        acc <<= sh;
        cnt = 64 - sh;
    }
    __device__ __forceinline__ void refill() {
        if (cnt <= 32) {
            acc |= (uint64_t)pop() << (32 - cnt);
            cnt += 32;
        }
    }
    __device__ __forceinline__ bool past_end() const { return np > hi + 256; }
};
// n in 0..32
template <class R> __device__ __forceinline__ uint32_t rd_get(R &br, int n) {
    br.refill();
    const uint32_t v = n ? (uint32_t)(br.acc >> (64 - n)) : 0u;
    br.acc <<= n;
    br.cnt -= n;
    return v;
}
template <class R> __device__ __forceinline__ int32_t rd_sget(R &br, int n) {  // signed, n in 1..32
    const uint32_t v = rd_get(br, n);
    return n >= 32 ? (int32_t)v : ((int32_t)(v << (32 - n)) >> (32 - n));
}
template <class R> __device__ __forceinline__ uint32_t rd_unary(R &br) {
    uint32_t z = 0;
    for (;;) {
        br.refill();
        if (br.acc == 0) {
            z += (uint32_t)br.cnt; br.cnt = 0;
            if (br.past_end()) return z;      // ran off the buffer: malformed
            continue;
        }
        const int lz = __clzll((long long)br.acc);
        z += (uint32_t)lz;
        br.acc = lz == 63 ? 0ull : (br.acc << (lz + 1));
        br.cnt -= lz + 1;
        return z;
    }
}
template <class R> __device__ __forceinline__ void rice_partition(R &br, RiceState &rs) {
    rs.k = rd_get(br, rs.pbits);
    rs.left = rs.psize;
    rs.rawbits = -1;
    if (rs.k == rs.esc) rs.rawbits = (int)rd_get(br, 5);
}
// one residual; the common case (zeros, stop bit and k low bits inside the upper half of the window) takes no loop
template <class R> __device__ __forceinline__ int32_t rice_next(R &br, RiceState &rs) {
    if (rs.left == 0) rice_partition(br, rs);
    rs.left--;
    if (rs.rawbits >= 0) return rs.rawbits ? rd_sget(br, rs.rawbits) : 0;
    br.refill();
    const int k = (int)rs.k;
    const uint32_t hi = (uint32_t)(br.acc >> 32);
    if (hi != 0u) {
        const int lz = __clz((int)hi);
        if (lz + k < 32) {
            const uint32_t t = (hi << lz) << 1;
            const uint32_t low = k ? (t >> (32 - k)) : 0u;
            const int used = lz + 1 + k;
            br.acc = (br.acc << (used - 1)) << 1;
            br.cnt -= used;
            const uint32_t u = ((uint32_t)lz << k) | low;
            return (int32_t)(u >> 1) ^ -(int32_t)(u & 1u);
        }
    }
    const uint32_t z = rd_unary(br);
    const uint32_t u = (z << rs.k) | rd_get(br, (int)rs.k);
    return (int32_t)(u >> 1) ^ -(int32_t)(u & 1u);
}
// the same walk without the value
template <class R> __device__ __forceinline__ void rice_skip(R &br, RiceState &rs) {
    if (rs.left == 0) rice_partition(br, rs);
    rs.left--;
    if (rs.rawbits >= 0) { (void)rd_get(br, rs.rawbits); return; }
    br.refill();
    const int k = (int)rs.k;
    const uint32_t hi = (uint32_t)(br.acc >> 32);
    if (hi != 0u) {
        const int used = __clz((int)hi) + 1 + k;
        if (used <= 32) {
            br.acc = (br.acc << (used - 1)) << 1;
            br.cnt -= used;
            return;
        }
    }
    (void)rd_unary(br);
    (void)rd_get(br, (int)rs.k);
}
// the scan's walk over one code with the ring reader: the common case (inside a Rice partition, window topped up from the word
// fetched ahead, code within 32 bits) is straight-line code with selects; everything else goes through rice_skip
__device__ __forceinline__ void rice_skip_ring(RingR &br, RiceState &rs) {
    const bool need = br.cnt <= 32;
    if (rs.left == 0 || rs.rawbits >= 0 || (need && br.rd == br.wr)) { rice_skip(br, rs); return; }
    rs.left--;
    const uint64_t add = (uint64_t)br.nx << ((32 - br.cnt) & 31);
    br.acc |= need ? add : 0ull;
    br.cnt += need ? 32 : 0;
    if (need) { br.rd++; br.nx = br.ring[(br.rd % D3_RING) * 32]; }
    const uint32_t hi = (uint32_t)(br.acc >> 32);
    const int used = __clz((int)hi) + 1 + (int)rs.k;          // hi == 0: clz = 32, used > 32
    if (used <= 32) {
        br.acc = (br.acc << (used - 1)) << 1;
        br.cnt -= used;
        return;
    }
    (void)rd_unary(br);
    (void)rd_get(br, (int)rs.k);
}
__device__ __forceinline__ uint32_t pack_state(const RiceState &rs) {
    return rs.k | (rs.rawbits >= 0 ? 1u << 5 : 0u) | ((uint32_t)(rs.rawbits >= 0 ? rs.rawbits : 0) << 6) | (min(rs.left, 0xFFFFFu) << 12);
}
__device__ __forceinline__ int ilog2_u32(uint32_t v) { return 31 - __clz((int)v); }

struct D3Args {
    D3Sub *sub;          // [frame - f0][chmax]
    D3Block *blk;        // [subframe][D3_MAXBLK]
    int32_t *res;        // [subframe][row] residuals, then PCM
    uint32_t f0, nf;     // frames of this pass
    int chmax;           // channels of the widest tile of the call
    int row;             // samples per scratch row (blocksize rounded up to a multiple of D3_BLK)
};

// ------------------------------------------------------------------------------------------
// A: scan.  Lane = frame; the 32 frames of a warp walk their codes in lockstep, 32 samples between ring top-ups.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(32) d3_scan_kernel(const DecArgs a, const D3Args x) {
    __shared__ uint32_t ring[D3_RING][32];
    const int lane = threadIdx.x;
    const uint32_t lf = blockIdx.x * 32u + (uint32_t)lane;
    const uint32_t f = x.f0 + lf;
    bool bad = false, live = lf < x.nf;
    int status = FRX_E_BITSTREAM;
    int ti = 0;
    frx_dtile t;
    memset(&t, 0, sizeof(t));
    uint32_t fj = 0;
    uint64_t pos = 0, end = 0;
    long long s0 = 0;
    int nsamp = 0, channels = 0, bps = 0;
    uint32_t stereo = 0;
    RingR br;
    if (live) {
        ti = dec_find_tile(a.dprefix, a.n_tiles, f);
        t = a.tiles[ti];
        fj = f - a.dprefix[ti];
        pos = a.frame_pos[f];
        if (pos == FRX_NOPOS || pos < t.byte_offset || pos >= t.byte_offset + t.byte_size) bad = true;
        end = t.byte_offset + t.byte_size;
        if (!bad && fj + 1 < t.n_frames) {
            const uint64_t nx = a.frame_pos[f + 1];
            if (nx != FRX_NOPOS && nx > pos && nx <= end) end = nx;
        }
    }
    br.init(a.bytes, a.n_bytes, pos, &ring[0][lane], live && !bad);
    if (live && !bad) {
        const long long tile_n = (long long)t.h * t.w;
        s0 = (long long)fj * a.blocksize;
        const int expect = (int)min((long long)a.blocksize, tile_n - s0);
        // ---- frame header ----
        const uint32_t sync = rd_get(br, 16);
        if (sync != 0xFFF8u) bad = true;
        const uint32_t bs_code = rd_get(br, 4), sr_code = rd_get(br, 4), ch_code = rd_get(br, 4), sz_code = rd_get(br, 3);
        rd_get(br, 1);
        {
            const uint32_t v = rd_get(br, 8);
            const int extra = !(v & 0x80u) ? 0 : (v & 0xE0u) == 0xC0u ? 1 : (v & 0xF0u) == 0xE0u ? 2 : (v & 0xF8u) == 0xF0u ? 3 : (v & 0xFCu) == 0xF8u ? 4 : 5;
            for (int e = 0; e < extra; e++) rd_get(br, 8);
        }
        if (bs_code == 1) nsamp = 192;
        else if (bs_code >= 2 && bs_code <= 5) nsamp = 576 << (bs_code - 2);
        else if (bs_code == 6) nsamp = (int)rd_get(br, 8) + 1;
        else if (bs_code == 7) nsamp = (int)rd_get(br, 16) + 1;
        else if (bs_code >= 8) nsamp = 256 << (bs_code - 8);
        else { nsamp = 0; bad = true; }
        if (sr_code == 12u) rd_get(br, 8); else if (sr_code == 13u || sr_code == 14u) rd_get(br, 16);
        rd_get(br, 8);  // CRC-8 (validated by discovery for foreign files; the CRC-16 pass covers the header too)
        bps = sz_code ? frx_size_code_bps(sz_code) : t.bps;
        channels = ch_code < 8u ? (int)ch_code + 1 : 2;
        if (nsamp != expect || bps != t.bps || channels != t.channels || ch_code > 10u) bad = true;
        if (ch_code >= 8u && !bad) {
            stereo = ch_code - 7u;                       // 1 left/side, 2 side/right, 3 mid/side
            if (bps >= 32) { bad = true; status = FRX_E_UNSUPPORTED; }
        }
    }
    int my_n = (live && !bad) ? nsamp : 0;
    const int my_ch = (live && !bad) ? channels : 0;
    const int nmax = __reduce_max_sync(0xFFFFFFFFu, my_n), chmax = __reduce_max_sync(0xFFFFFFFFu, my_ch);

    for (int c = 0; c < chmax; c++) {
        RiceState rs;
        rs.k = 0; rs.left = 0; rs.pbits = 4; rs.esc = 15u; rs.rawbits = -1; rs.psize = 1;
        int type = 3, order = 0;
        D3Sub *sd = live ? &x.sub[(size_t)lf * x.chmax + c] : nullptr;
        D3Block *bk = live ? &x.blk[((size_t)lf * x.chmax + c) * D3_MAXBLK] : nullptr;
        if (c < my_ch && !bad) {
            const uint32_t hdr = rd_get(br, 8);
            const uint32_t tcode = (hdr >> 1) & 0x3Fu;
            int wasted = 0, shift = 0, prec = 3;
            int32_t cval = 0;
            bool wide = false;
            if (hdr & 0x80u) bad = true;
            if (hdr & 1u) wasted = (int)rd_unary(br) + 1;
            const int side = ((stereo == 1u && c == 1) || (stereo == 2u && c == 0) || (stereo == 3u && c == 1)) ? 1 : 0;
            const int sbps = bps + side - wasted;
            if (sbps <= 0 || sbps > 32) bad = true;
            if (bad) {}
            else if (tcode == 0) { type = 0; cval = rd_sget(br, sbps); }
            else if (tcode == 1) {
                type = 1;
                rs.rawbits = sbps; rs.left = 0xFFFFFu; rs.psize = 0xFFFFFu;      // "one escaped partition" of sbps-bit samples
            } else {
                bool lpc = false;
                if (tcode >= 8 && tcode <= 12) order = (int)tcode - 8;
                else if (tcode >= 32) { order = (int)(tcode & 31u) + 1; lpc = true; }
                else bad = true;
                if (order > nsamp) bad = true;
                if (!bad) {
                    for (int i = 0; i < order; i++) sd->warm[i] = rd_sget(br, sbps);
                    if (lpc) {
                        prec = (int)rd_get(br, 4) + 1;
                        shift = rd_sget(br, 5);
                        if (prec == 16 || shift < 0) bad = true;
                        for (int j = 0; j < order; j++) sd->coef[j] = (int16_t)rd_sget(br, prec);
                    } else {
                        const int16_t fc[5][4] = { { 0, 0, 0, 0 }, { 1, 0, 0, 0 }, { 2, -1, 0, 0 }, { 3, -3, 1, 0 }, { 4, -6, 4, -1 } };
                        for (int j = 0; j < order; j++) sd->coef[j] = fc[order][j];
                    }
                    const uint32_t method = rd_get(br, 2), po = rd_get(br, 4);
                    if (method > 1u || (po > 0 && ((uint32_t)nsamp >> po) << po != (uint32_t)nsamp) || ((uint32_t)nsamp >> po) < (uint32_t)order) bad = true;
                    if (!bad) {
                        type = 2;
                        rs.pbits = method ? 5 : 4; rs.esc = method ? 31u : 15u; rs.psize = (uint32_t)nsamp >> po;
                        // the first partition is short by `order`
                        rs.k = rd_get(br, rs.pbits);
                        rs.left = rs.psize - (uint32_t)order;
                        rs.rawbits = -1;
                        if (rs.k == rs.esc) rs.rawbits = (int)rd_get(br, 5);
                        wide = order > 0 && (sbps + prec + ilog2_u32((uint32_t)order) > 32);
                    }
                }
            }
            if (!bad) {
                sd->type = (uint8_t)type; sd->order = (uint8_t)order; sd->shift = (uint8_t)shift; sd->wasted = (uint8_t)wasted;
                sd->sbps = (uint8_t)sbps; sd->wide = wide ? 1 : 0; sd->pbits = (uint8_t)rs.pbits;
                sd->st_mode = (uint8_t)(stereo ? (c == 0 ? 1u : 1u + stereo) : 0u);
                sd->cval = cval; sd->nsamp = (uint32_t)nsamp; sd->psize = rs.psize; sd->tile = (uint32_t)ti; sd->bps = (uint32_t)bps;
                sd->s0 = s0; sd->pos = pos;
            }
        }
        // ---- walk the codes: the block table ----
        const bool walk = !bad && (type == 1 || type == 2);
        const int first = type == 2 ? order : 0;
        for (int n0 = 0; n0 < nmax; n0 += 32) {
            br.topup(walk && n0 < my_n);
            // chunks in which every lane walks 32 codes need no per-sample bounds; a block boundary can only be their first sample
            const bool steady = __all_sync(0xFFFFFFFFu, walk && n0 >= first && n0 > 0 && n0 + 32 <= my_n);
            if (steady) {
                if ((n0 & (D3_BLK - 1)) == 0) {
                    D3Block e;
                    e.bitoff = br.consumed(); e.state = pack_state(rs);
                    bk[n0 >> 7] = e;
                }
#pragma unroll 8
                for (int i = 0; i < 32; i++) rice_skip_ring(br, rs);
            } else if (walk) {
#pragma unroll 4
                for (int i = 0; i < 32; i++) {
                    const int n = n0 + i;
                    if (n >= first && n < my_n) {
                        if ((n & (D3_BLK - 1)) == 0 || n == first) {
                            D3Block e;
                            e.bitoff = br.consumed(); e.state = pack_state(rs);
                            bk[n >> 7] = e;
                        }
                        rice_skip(br, rs);
                    }
                }
            }
        }
        if (bad) my_n = 0;
    }
    if (!live) return;
    // ---- footer: byte align; the frame must end where the next one starts (CRC-16: frx_crc_check_kernel) ----
    const uint64_t fsize = (uint64_t)((br.consumed() + 7u) >> 3) + 2u;
    if (!bad && pos + fsize > end) bad = true;
    if (!bad && fj + 1 < t.n_frames && a.frame_pos[f + 1] != FRX_NOPOS && pos + fsize != a.frame_pos[f + 1]) bad = true;
    if (a.frame_len) a.frame_len[f] = bad ? 0u : (uint32_t)fsize;
    for (int c = 0; c < x.chmax; c++)
        if (bad || c >= channels) x.sub[(size_t)lf * x.chmax + c].type = 3;
    if (bad) atomicExch(&a.tile_status[ti], status);
}

// ------------------------------------------------------------------------------------------
// B: residuals.  Warp = subframe, lane = block of 128 samples.  Every lane decodes its block 32 residuals at a time into
// a padded 32x33 stage; the warp then writes the stage row by row, so the scratch row only sees full 128-byte lines.
// (Verbatim subframes come through here too: the scan describes them as one escaped partition of sbps-bit samples.)
// ------------------------------------------------------------------------------------------
#define D3B_WARPS 4
__global__ void __launch_bounds__(32 * D3B_WARPS) d3_residual_kernel(const DecArgs a, const D3Args x) {
    __shared__ uint32_t stage_all[D3B_WARPS][32][33];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t (*stage)[33] = stage_all[warp];
    const size_t s = (size_t)blockIdx.x * D3B_WARPS + warp;
    if (s >= (size_t)x.nf * x.chmax) return;
    const D3Sub *sd = &x.sub[s];
    const int type = sd->type;
    if (type != 1 && type != 2) return;
    const int nsamp = (int)sd->nsamp, first = type == 2 ? sd->order : 0;
    const int nblk = (nsamp + D3_BLK - 1) / D3_BLK;
    const int b0 = lane * D3_BLK;
    const int lo = max(first, b0), hi = min(nsamp, b0 + D3_BLK);      // this lane's samples [lo, hi)
    const bool mine = lane < nblk && lo < hi;
    GlobR br;
    RiceState rs;
    rs.pbits = sd->pbits; rs.esc = rs.pbits == 5 ? 31u : 15u; rs.psize = sd->psize; rs.k = 0; rs.left = 0; rs.rawbits = -1;
    if (mine) {
        const D3Block e = x.blk[s * D3_MAXBLK + lane];
        rs.k = e.state & 31u;
        rs.rawbits = (e.state >> 5) & 1u ? (int)((e.state >> 6) & 63u) : -1;
        rs.left = e.state >> 12;
        br.init(a.bytes, a.n_bytes, sd->pos * 8ull + e.bitoff);
    } else { br.lo = br.hi = br.np = a.bytes; br.acc = 0; br.cnt = 0; }
    int32_t *row = x.res + s * (size_t)x.row;
    for (int q = 0; q < D3_BLK; q += 32) {
        if (mine) {
#pragma unroll 4
            for (int i = 0; i < 32; i++) {
                const int n = b0 + q + i;
                int32_t r = 0;
                if (n >= lo && n < hi) r = rice_next(br, rs);
                stage[lane][i] = (uint32_t)r;
            }
        }
        __syncwarp();
        for (int j = 0; j < nblk; j++) {
            const int n = j * D3_BLK + q + lane;
            if (n < nsamp) row[n] = (int32_t)stage[j][lane];
        }
        __syncwarp();
    }
}

// ------------------------------------------------------------------------------------------
// C: restore.  Lane = subframe; the 32 subframes of a warp restore 32 samples at a time, in place in the stage: residuals
// in (row by row, coalesced), samples out (the same rows).  ORD = history length (a multiple of 4 dividing 32),
// WIDE = 64-bit accumulate; both warp-uniform.  What is stored is the PCM with its wasted bits shifted back in.
// ------------------------------------------------------------------------------------------
template <int ORD, bool WIDE>
__device__ __forceinline__ void d3_restore_rows(uint32_t (*stage)[33], const int32_t *coef, int32_t *h, int lane, int type, int order, int shift,
                                               int wasted, int32_t cval, int my_n, int nmax, int32_t *myrow, int32_t *const *rows, const int *ns) {
    // the residual rows of a chunk are fetched one chunk ahead (32 independent loads per lane in flight during the recurrence of
    // the chunk before: fetched on demand, row after row, they were 32 dependent round trips per chunk and 2.3 of the 4.4 ms of a
    // single three-band tile)
    int32_t pre[32];
#pragma unroll
    for (int j = 0; j < 32; j++) pre[j] = lane < ns[j] ? rows[j][lane] : 0;
    for (int n0 = 0; n0 < nmax; n0 += 32) {
        // residual rows in
#pragma unroll
        for (int j = 0; j < 32; j++) stage[j][lane] = (uint32_t)pre[j];
        __syncwarp();
        if (n0 + 32 < nmax) {
#pragma unroll
            for (int j = 0; j < 32; j++) { const int n = n0 + 32 + lane; pre[j] = n < ns[j] ? rows[j][n] : 0; }
        }
        if (n0 < my_n) {
            for (int mb = 0; mb < 32; mb += ORD) {
#pragma unroll
                for (int mi = 0; mi < ORD; mi++) {
                    const int n = n0 + mb + mi;
                    int32_t v = cval;
                    if (type == 2) {
                        if (n >= order) {
                            const int32_t r = (int32_t)stage[lane][mb + mi];
                            if (WIDE) {
                                long long acc = 0;
#pragma unroll
                                for (int j = 0; j < ORD; j++) acc += (long long)coef[j] * (long long)h[(mi - 1 - j + 2 * ORD) % ORD];
                                v = (int32_t)((long long)r + (acc >> shift));
                            } else {
                                int32_t acc = 0;
#pragma unroll
                                for (int j = 0; j < ORD; j++) acc += coef[j] * h[(mi - 1 - j + 2 * ORD) % ORD];
                                v = r + (acc >> shift);
                            }
                            h[mi] = v;
                        } else v = h[mi];          // warm-up sample (n < order <= ORD, so n == mi)
                    } else if (type == 1) v = (int32_t)stage[lane][mb + mi];
                    stage[lane][mb + mi] = (uint32_t)v << wasted;
                }
            }
        }
        __syncwarp();
        // sample rows out
        for (int j = 0; j < 32; j++) {
            const int n = n0 + lane;
            if (n < ns[j]) rows[j][n] = (int32_t)stage[j][lane];
        }
        __syncwarp();
    }
}

#define D3C_WARPS 2
__global__ void __launch_bounds__(32 * D3C_WARPS) d3_restore_kernel(const D3Args x) {
    __shared__ uint32_t stage_all[D3C_WARPS][32][33];
    __shared__ int32_t *rows_all[D3C_WARPS][32];
    __shared__ int ns_all[D3C_WARPS][32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t s = ((size_t)blockIdx.x * D3C_WARPS + warp) * 32u + (size_t)lane;
    const bool live = s < (size_t)x.nf * x.chmax;
    int type = 3, order = 0, shift = 0, wasted = 0, my_n = 0;
    int32_t cval = 0;
    bool wide = false;
    int32_t coef[32], h[32];
#pragma unroll
    for (int j = 0; j < 32; j++) { coef[j] = 0; h[j] = 0; }
    if (live) {
        const D3Sub *sd = &x.sub[s];
        type = sd->type;
        if (type != 3) {
            order = sd->order; shift = sd->shift; wasted = sd->wasted; cval = sd->cval; my_n = (int)sd->nsamp; wide = sd->wide != 0;
            if (type == 2) {
#pragma unroll
                for (int j = 0; j < 32; j++) if (j < order) { coef[j] = sd->coef[j]; h[j] = sd->warm[j]; }
            }
        }
    }
    rows_all[warp][lane] = x.res + (live ? s : 0) * (size_t)x.row;
    ns_all[warp][lane] = my_n;
    __syncwarp();
    const int nmax = __reduce_max_sync(0xFFFFFFFFu, my_n);
    const int omax = __reduce_max_sync(0xFFFFFFFFu, (uint32_t)(type == 2 ? order : 0));
    const bool anywide = __any_sync(0xFFFFFFFFu, wide && my_n);
    uint32_t (*stage)[33] = stage_all[warp];
    int32_t *const *rows = rows_all[warp];
    const int *ns = ns_all[warp];
    int32_t *myrow = rows_all[warp][lane];
#define D3_RUN(O, W) d3_restore_rows<O, W>(stage, coef, h, lane, type, order, shift, wasted, cval, my_n, nmax, myrow, rows, ns)
    if (omax <= 4) { if (anywide) D3_RUN(4, true); else D3_RUN(4, false); }
    else if (omax <= 8) { if (anywide) D3_RUN(8, true); else D3_RUN(8, false); }
    else if (omax <= 16) { if (anywide) D3_RUN(16, true); else D3_RUN(16, false); }
    else { if (anywide) D3_RUN(32, true); else D3_RUN(32, false); }
#undef D3_RUN
}

// ------------------------------------------------------------------------------------------
// D: output.  Block = frame, thread = sample (all channels): stereo recombination (RFC 9639 section 4.2), dequantisation,
// raster geometry.  Consecutive threads write consecutive pixels of a raster row.
// ------------------------------------------------------------------------------------------
template <int DT>
__global__ void __launch_bounds__(256) d3_output_kernel(const DecArgs a, const D3Args x) {
    typedef typename FrxElem<DT>::T T;
    const uint32_t lf = blockIdx.x;
    const D3Sub *sd0 = &x.sub[(size_t)lf * x.chmax];
    // the frame's tile and place (taken from the first subframe the scan accepted)
    int cfirst = -1;
    for (int c = 0; c < x.chmax; c++) if (sd0[c].type != 3) { cfirst = c; break; }
    if (cfirst < 0) return;
    const frx_dtile t = a.tiles[sd0[cfirst].tile];
    const long long s0 = sd0[cfirst].s0;
    const int nsamp = (int)sd0[cfirst].nsamp, bps = (int)sd0[cfirst].bps;
    FrxDequant dq;
    dq.range_f = (float)(t.dmax - t.dmin);
    dq.dmin_f = (float)t.dmin;
    const uint32_t w = (uint32_t)t.w;
    const long long row0 = s0 / t.w;
    const uint32_t col0 = (uint32_t)(s0 - row0 * t.w);
    T *out = (T *)a.out + t.out_offset;
    for (int c = 0; c < t.channels && c < x.chmax; c++) {
        const D3Sub *sd = &sd0[c];
        if (sd->type == 3) continue;
        const uint32_t st = sd->st_mode;
        if (st == 1u) continue;                                   // first channel of a decorrelated pair: written with the second
        const int32_t *r1 = x.res + ((size_t)lf * x.chmax + c) * (size_t)x.row;
        const int32_t *r0 = st >= 2u ? r1 - x.row : r1;
        T *band = out + (long long)c * t.out_band_stride;
        for (int n = threadIdx.x; n < nsamp; n += 256) {
            const uint32_t rel = col0 + (uint32_t)n;
            const uint32_t dr = rel / w, col = rel - dr * w;
            T *p = band + (row0 + dr) * t.out_row_stride + col;
            const int32_t s1 = r1[n];
            if (st == 0u) { *p = frx_dequant<DT>(s1, bps, dq, a.pcm_mode); continue; }
            const int32_t sA = r0[n];
            int32_t L, R;
            if (st == 2u) { L = sA; R = sA - s1; }                  // left, side
            else if (st == 3u) { R = s1; L = sA + s1; }             // side, right
            else { const int32_t mid = (int32_t)(((uint32_t)sA << 1) | ((uint32_t)s1 & 1u)); L = (mid + s1) >> 1; R = (mid - s1) >> 1; }
            *(p - t.out_band_stride) = frx_dequant<DT>(L, bps, dq, a.pcm_mode);
            *p = frx_dequant<DT>(R, bps, dq, a.pcm_mode);
        }
    }
}

template <int DT>
static cudaError_t launch_output(const DecArgs &a, const D3Args &x, cudaStream_t st) {
    d3_output_kernel<DT><<<x.nf, 256, 0, st>>>(a, x);
    return cudaGetLastError();
}

}  // namespace

// bytes of scratch the phased decoder needs for a pass over `frames` frames of up to `chmax` channels
void frx_decode3_scratch(uint32_t frames, int chmax, int blocksize, size_t *sub_bytes, size_t *blk_bytes, size_t *res_bytes) {
    const size_t subs = (size_t)frames * (size_t)chmax;
    const size_t row = (size_t)((blocksize + D3_BLK - 1) / D3_BLK) * D3_BLK;
    *sub_bytes = subs * sizeof(D3Sub);
    *blk_bytes = subs * D3_MAXBLK * sizeof(D3Block);
    *res_bytes = subs * row * sizeof(int32_t);
}
bool frx_decode3_supports(int blocksize) { return blocksize <= D3_BLK * D3_MAXBLK; }

// the four phases over the frames [f0, f0 + nf) of the call
cudaError_t frx_launch_decode3(const DecArgs &a, void *sub, void *blk, void *res, uint32_t f0, uint32_t nf, int chmax, int out_dtype,
                               cudaStream_t st) {
    if (!nf) return cudaSuccess;
    D3Args x;
    x.sub = (D3Sub *)sub; x.blk = (D3Block *)blk; x.res = (int32_t *)res; x.f0 = f0; x.nf = nf; x.chmax = chmax;
    x.row = ((a.blocksize + D3_BLK - 1) / D3_BLK) * D3_BLK;
    const size_t subs = (size_t)nf * (size_t)chmax;
    d3_scan_kernel<<<(nf + 31u) / 32u, 32, 0, st>>>(a, x);
    d3_residual_kernel<<<(unsigned)((subs + D3B_WARPS - 1) / D3B_WARPS), 32 * D3B_WARPS, 0, st>>>(a, x);
    d3_restore_kernel<<<(unsigned)((subs + 32 * D3C_WARPS - 1) / (32 * D3C_WARPS)), 32 * D3C_WARPS, 0, st>>>(x);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    switch (out_dtype) {
        case FRX_U8: return launch_output<FRX_U8>(a, x, st);
        case FRX_I16: return launch_output<FRX_I16>(a, x, st);
        case FRX_U16: return launch_output<FRX_U16>(a, x, st);
        case FRX_I32: return launch_output<FRX_I32>(a, x, st);
        case FRX_U32: return launch_output<FRX_U32>(a, x, st);
        case FRX_F32: return launch_output<FRX_F32>(a, x, st);
        default: return cudaErrorInvalidValue;
    }
}

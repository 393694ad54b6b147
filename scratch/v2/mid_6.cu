
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

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



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

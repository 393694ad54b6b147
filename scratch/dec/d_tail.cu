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

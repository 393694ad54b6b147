// ------------------------------------------------------------------------------------------
// K0+K1 alone (frx_quantize): interleaved int32 PCM of one tile
// ------------------------------------------------------------------------------------------
template <int DT>
__global__ void frx_quantize_kernel(const void *raster, long long band_stride, long long row_stride, frx_tile t,
                                    const FrxQuant *q, int mode, int32_t *pcm) {
    typedef typename FrxElem<DT>::T T;
    const T *base = (const T *)raster;
    const FrxQuant qz = q[0];
    long long n = (long long)t.h * t.w * t.n_channels;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        int c = (int)(i % t.n_channels);
        long long px = i / t.n_channels;
        int row = (int)(px / t.w), col = (int)(px - (long long)row * t.w);
        T d = base[(long long)(t.first_band + c) * band_stride + (long long)(t.row_off + row) * row_stride + t.col_off + col];
        pcm[i] = frx_quant<DT>(d, qz, mode);
    }
}

__global__ void frx_minmax_out_kernel(const FrxQuant *q, int dtype, double *mm) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        if (dtype == FRX_F32) { mm[0] = (double)q[0].fmin; mm[1] = (double)q[0].fmax; }
        else { mm[0] = (double)q[0].imin; mm[1] = (double)q[0].imax; }
    }
}

// ------------------------------------------------------------------------------------------
// host launchers (called from frx_api.cu)
// ------------------------------------------------------------------------------------------
size_t frx_enc_smem_bytes(int frame_words) {
    return ENC_SMALL_BYTES + ENC_UNION_BYTES + sizeof(uint32_t) * (size_t)frame_words;
}

size_t frx_minmax_raw_size() { return sizeof(MinMaxRaw); }

template <int DT>
static cudaError_t launch_minmax_t(const frx_raster &r, const frx_tile *tiles_dev, const uint32_t *unit_prefix_dev, int n_tiles,
                                   uint32_t total_units, int unit_elems, void *mm_raw, FrxQuant *quant, cudaStream_t st) {
    frx_minmax_init_kernel<<<(n_tiles + 255) / 256, 256, 0, st>>>((MinMaxRaw *)mm_raw, n_tiles);
    if (total_units)
        frx_minmax_kernel<DT><<<total_units, 256, 0, st>>>(r.data, r.band_stride, r.row_stride, tiles_dev, unit_prefix_dev, n_tiles,
                                                          unit_elems, (MinMaxRaw *)mm_raw);
    frx_quant_setup_kernel<<<(n_tiles + 255) / 256, 256, 0, st>>>((const MinMaxRaw *)mm_raw, quant, n_tiles, r.dtype);
    return cudaGetLastError();
}

cudaError_t frx_launch_minmax(const frx_raster &r, const frx_tile *tiles_dev, const uint32_t *unit_prefix_dev, int n_tiles,
                              uint32_t total_units, int unit_elems, void *mm_raw, FrxQuant *quant, cudaStream_t st) {
    switch (r.dtype) {
        case FRX_U8: return launch_minmax_t<FRX_U8>(r, tiles_dev, unit_prefix_dev, n_tiles, total_units, unit_elems, mm_raw, quant, st);
        case FRX_I16: return launch_minmax_t<FRX_I16>(r, tiles_dev, unit_prefix_dev, n_tiles, total_units, unit_elems, mm_raw, quant, st);
        case FRX_U16: return launch_minmax_t<FRX_U16>(r, tiles_dev, unit_prefix_dev, n_tiles, total_units, unit_elems, mm_raw, quant, st);
        case FRX_I32: return launch_minmax_t<FRX_I32>(r, tiles_dev, unit_prefix_dev, n_tiles, total_units, unit_elems, mm_raw, quant, st);
        case FRX_U32: return launch_minmax_t<FRX_U32>(r, tiles_dev, unit_prefix_dev, n_tiles, total_units, unit_elems, mm_raw, quant, st);
        case FRX_F32: return launch_minmax_t<FRX_F32>(r, tiles_dev, unit_prefix_dev, n_tiles, total_units, unit_elems, mm_raw, quant, st);
    }
    return cudaErrorInvalidValue;
}

template <bool WIDE, int HISTN>
static cudaError_t launch_encode_t(const EncArgs &a, int sm_count, cudaStream_t st) {
    size_t smem = frx_enc_smem_bytes(a.frame_words);
    cudaError_t e = cudaFuncSetAttribute(frx_encode_kernel<WIDE, HISTN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(frx_encode_kernel<WIDE, HISTN>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if (e != cudaSuccess) return e;
    int occ = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, frx_encode_kernel<WIDE, HISTN>, FRX_ENC_THREADS, smem);
    if (e != cudaSuccess) return e;
    if (occ < 1) return cudaErrorLaunchOutOfResources;
    long long grid = (long long)sm_count * occ;   // persistent: every CTA resident, tickets hand out the frames in order
    if (grid > (long long)a.total_frames) grid = a.total_frames;
    if (grid < 1) grid = 1;
    frx_encode_kernel<WIDE, HISTN><<<(unsigned)grid, FRX_ENC_THREADS, smem, st>>>(a);
    return cudaGetLastError();
}

cudaError_t frx_launch_encode(const EncArgs &a, int dtype, int sm_count, cudaStream_t st) {
    (void)dtype;
    const bool wide = a.bps > 16;
    const int hist = a.max_lpc_order <= 4 ? 4 : (a.max_lpc_order <= 8 ? 8 : 12);
    if (wide) {
        if (hist == 4) return launch_encode_t<true, 4>(a, sm_count, st);
        if (hist == 8) return launch_encode_t<true, 8>(a, sm_count, st);
        return launch_encode_t<true, 12>(a, sm_count, st);
    }
    if (hist == 4) return launch_encode_t<false, 4>(a, sm_count, st);
    if (hist == 8) return launch_encode_t<false, 8>(a, sm_count, st);
    return launch_encode_t<false, 12>(a, sm_count, st);
}

cudaError_t frx_launch_tile_index(const uint32_t *frame_prefix, const uint32_t *frame_size, const uint64_t *frame_offset,
                                  const FrxQuant *q, int dtype, int n_tiles, uint64_t *tile_offset, uint64_t *tile_size,
                                  double *tile_minmax, cudaStream_t st) {
    frx_tile_index_kernel<<<(n_tiles + 255) / 256, 256, 0, st>>>(frame_prefix, frame_size, frame_offset, q, dtype, n_tiles, tile_offset,
                                                                tile_size, tile_minmax);
    return cudaGetLastError();
}

cudaError_t frx_launch_quantize(const frx_raster &r, const frx_tile &t, const FrxQuant *q, int mode, int32_t *pcm, double *minmax,
                                cudaStream_t st) {
    long long n = (long long)t.h * t.w * t.n_channels;
    int grid = (int)min((long long)148 * 8, (n + 255) / 256);
    if (grid < 1) grid = 1;
    switch (r.dtype) {
        case FRX_U8: frx_quantize_kernel<FRX_U8><<<grid, 256, 0, st>>>(r.data, r.band_stride, r.row_stride, t, q, mode, pcm); break;
        case FRX_I16: frx_quantize_kernel<FRX_I16><<<grid, 256, 0, st>>>(r.data, r.band_stride, r.row_stride, t, q, mode, pcm); break;
        case FRX_U16: frx_quantize_kernel<FRX_U16><<<grid, 256, 0, st>>>(r.data, r.band_stride, r.row_stride, t, q, mode, pcm); break;
        case FRX_I32: frx_quantize_kernel<FRX_I32><<<grid, 256, 0, st>>>(r.data, r.band_stride, r.row_stride, t, q, mode, pcm); break;
        case FRX_U32: frx_quantize_kernel<FRX_U32><<<grid, 256, 0, st>>>(r.data, r.band_stride, r.row_stride, t, q, mode, pcm); break;
        case FRX_F32: frx_quantize_kernel<FRX_F32><<<grid, 256, 0, st>>>(r.data, r.band_stride, r.row_stride, t, q, mode, pcm); break;
        default: return cudaErrorInvalidValue;
    }
    if (minmax) frx_minmax_out_kernel<<<1, 32, 0, st>>>(q, r.dtype, minmax);
    return cudaGetLastError();
}

#ifdef FRX_DBG
extern "C" void frx_debug_trace(unsigned long long *dev_ptr) { cudaMemcpyToSymbol(g_frx_trace, &dev_ptr, sizeof(dev_ptr)); }
extern "C" void frx_debug_read(unsigned long long *out, int reset) {
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out, g_frx_dbg, sizeof(unsigned long long) * 16);
    if (reset) { unsigned long long z[16] = {0}; cudaMemcpyToSymbol(g_frx_dbg, z, sizeof(z)); }
}
#endif

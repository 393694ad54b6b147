// frx_internal.cuh -- argument blocks and launcher prototypes shared by the libfrx translation units.
#pragma once
#include "frx_common.cuh"

#define FRX_CRC_MAXLQ 48   // CRC chunk tables cover frames up to 48 * 256 quads = 192 KB

struct EncArgs {
    const void *raster;
    long long H, W, band_stride, row_stride;
    const frx_tile *tiles;
    const uint32_t *frame_prefix;
    const FrxQuant *quant;
    const uint16_t *crc_xpow;   // [FRX_CRC_MAXLQ][256]: x^(128 * Lq * j) mod P, j = quad-chunks that follow a CRC thread's chunk of Lq quads
    const uint16_t *crc_xinv;   // [4]: x^(-32 z) mod P, undoes z zero words appended to the last quad
    const uint16_t *crc_tab;    // [4][256] slicing-by-4 CRC-16 tables
    uint8_t *bytes;
    uint64_t capacity;
    uint8_t *slots;             // two slots of slot_stride bytes per resident frame warp: streamed channels and parked frames
    uint32_t slot_stride;
    int bars_per_frame;         // CTA barriers every frame warp passes per frame (phase alignment; padded to this count)
    uint32_t *frame_size;
    uint64_t *frame_offset;
    uint64_t *desc;
    uint32_t *ticket;
    int32_t *status;
    uint64_t *total_bytes;
    uint32_t total_frames;
    int n_tiles;
    int blocksize;
    int bps;
    int pcm_mode;
    int sr_code;
    int sample_rate;
    int max_lpc_order;
    int max_po;
    int qlp_precision;
    int frame_words;
    int dtype;
    int tma_ok;                 // raster base and strides are 16-byte multiples: rows can be staged with TMA bulk copies
    int stereo;                 // two-channel streams try mid/side decorrelation (libFLAC levels 1, 2, 4..8; 16-bit PCM only)
};
struct DecArgs {
    const uint8_t *bytes;
    uint64_t n_bytes;
    const frx_dtile *tiles;
    const uint32_t *dprefix;
    uint64_t *frame_pos;
    const uint32_t *frame_size;
    int32_t *stereo_scratch;        // one blocksize row of int32 per frame, only when two-channel tiles are decoded (else NULL)
    uint32_t *frame_len;            // decoded size of every frame (0: rejected), input of the CRC pass
    const uint16_t *crc_tab;
    void *out;
    int32_t *tile_status;
    uint32_t total_frames;
    int n_tiles;
    int blocksize;
    int pcm_mode;
    int verify_crc;
};
size_t frx_enc_smem_bytes(int frame_words, int dtype);
int frx_enc_warps_per_cta();
cudaError_t frx_launch_crc_check(const DecArgs &a, const uint16_t *xpow, int sm_count, cudaStream_t st);
size_t frx_minmax_raw_size();
cudaError_t frx_launch_minmax(const frx_raster &r, const frx_tile *tiles_dev, const uint32_t *unit_prefix_dev, int n_tiles,
                              uint32_t total_units, int unit_elems, void *mm_raw, FrxQuant *quant, cudaStream_t st);
cudaError_t frx_launch_encode(const EncArgs &a, int dtype, int sm_count, cudaStream_t st);
cudaError_t frx_launch_tile_index(const uint32_t *frame_prefix, const uint32_t *frame_size, const uint64_t *frame_offset,
                                  const FrxQuant *q, int dtype, int n_tiles, uint64_t *tile_offset, uint64_t *tile_size,
                                  double *tile_minmax, cudaStream_t st);
cudaError_t frx_launch_quantize(const frx_raster &r, const frx_tile &t, const FrxQuant *q, int mode, int32_t *pcm, double *minmax,
                                cudaStream_t st);
void frx_make_crc16_tables(uint16_t *tab);
cudaError_t frx_launch_md5(const frx_raster &r, const frx_tile *tiles_dev, const FrxQuant *quant, int n_tiles, int mode, int bps, uint8_t *out,
                           cudaStream_t st);
cudaError_t frx_launch_frame_pos(const DecArgs &a, cudaStream_t st);
cudaError_t frx_launch_discover(const DecArgs &a, const uint32_t *unit_prefix, const int32_t *unit_tiles, int n_scan_tiles,
                                uint32_t total_units, cudaStream_t st);
cudaError_t frx_launch_decode(const DecArgs &a, int out_dtype, int sm_count, cudaStream_t st);
// the phased decoder (frx_decode3.cu): scan, residuals, restore, output over the frames [f0, f0 + nf) of the call
void frx_decode3_scratch(uint32_t frames, int chmax, int blocksize, size_t *sub_bytes, size_t *blk_bytes, size_t *res_bytes);
bool frx_decode3_supports(int blocksize);
cudaError_t frx_launch_decode3(const DecArgs &a, void *sub, void *blk, void *res, uint32_t f0, uint32_t nf, int chmax, int out_dtype,
                               cudaStream_t st);
cudaError_t frx_launch_dequantize(const int32_t *pcm, int h, int w, int channels, int bps, int out_dtype, int mode, double dmin,
                                  double dmax, void *out, cudaStream_t st);
// frx_tiff_unpack work split, shared by the host (unit counts) and the kernel: a chunk whose rows are short is cut into runs of
// whole rows (one contiguous source block per unit), a chunk with long rows into row segments of seg_px pixels.
#define FRX_ING_SEG_BYTES 16384
__host__ __device__ inline unsigned long long frx_ingest_plan(int esz, int nb, int rows, int cols, int pitch, int *rows_per_unit, int *seg_px) {
    const long long row_bytes = (long long)pitch * nb * esz;
    if (row_bytes * 2 <= FRX_ING_SEG_BYTES) {
        *rows_per_unit = (int)(FRX_ING_SEG_BYTES / row_bytes);
        *seg_px = pitch;
        return ((unsigned long long)rows + *rows_per_unit - 1) / *rows_per_unit;
    }
    int px = (FRX_ING_SEG_BYTES / (esz * nb)) & ~15;      // multiples of 16 pixels keep segment starts 16-byte aligned
    if (px < 16) px = 16;
    *rows_per_unit = 1;
    *seg_px = px;
    return (unsigned long long)rows * (((unsigned long long)cols + px - 1) / px);
}
cudaError_t frx_launch_tiff_unpack(const uint8_t *src, const frx_tiff_chunk *chunks_dev, const uint32_t *unit_prefix_dev, int n_chunks,
                                   uint32_t total_units, int spp, int swap, const frx_raster &dst, cudaStream_t st);

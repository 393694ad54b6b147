/*
 * frx.h -- C-ABI of libfrx.so, the B200-native raster<->FLAC tile codec.
 *
 * This is the drop-in boundary for flac-raster's codec hot path.  In the reference the only
 * interface the code talks to is pyFLAC 3.0.0 (cffi over libFLAC 1.4.3):
 *
 *   encode:  pyflac.StreamEncoder(write_callback, sample_rate, compression_level, blocksize=4096)
 *            .process(ndarray) / .finish()
 *              reference call sites: src/flac_raster/converter.py:201-216,
 *                                    src/flac_raster/spatial_encoder.py:278-291
 *              pyFLAC side:          docs/sonos-pyflac.txt:1968-2014, 2175-2213
 *              C level underneath:   FLAC__stream_encoder_{new,set_*,init_stream,
 *                                    process_interleaved,finish}  docs/sonos-pyflac.txt:3205-3262
 *   decode:  pyflac.FileDecoder(path).process() -> (float64 ndarray, sample_rate)
 *              reference call sites: src/flac_raster/converter.py:241-242, cli.py:170-173, 469-472
 *              pyFLAC side:          docs/sonos-pyflac.txt:1584-1640, 1809-1854
 *   staging fused into the same kernels (numpy in the reference):
 *              band->channel interleave   converter.py:185-194, spatial_encoder.py:186-190
 *              min/max + PCM scaling      converter.py:56-86, 152-153; spatial_encoder.py:229-248
 *              tile cut                   cli.py:691-699, spatial_encoder.py:92-103, 182-183
 *              PCM -> raster dtype        converter.py:88-110, 265-273
 *
 * Conventions: plain C, no exceptions; every function returns 0 on success and a negative
 * FRX_E_* code on failure, with a human-readable reason in frx_last_error() (thread-local).
 * The caller owns every buffer passed in; the library keeps only per-context scratch.
 * "dev" pointers are CUDA device pointers on the context's device; `stream` is a
 * cudaStream_t (pass 0 for the legacy default stream).  Device entry points are
 * ordered on `stream` and return with their kernels still running; frx_encode_tiles,
 * frx_decode_tiles, frx_tile_md5 and frx_quantize synchronise `stream` ONCE on the way in
 * (the tile tables are staged from host vectors), so work queued on that stream earlier
 * has finished when they return.
 * There is NO CPU fallback: without a CUDA device every compute entry point fails.
 */
#ifndef FRX_H
#define FRX_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FRX_VERSION_MAJOR 0
#define FRX_VERSION_MINOR 1

/* error codes */
#define FRX_OK 0
#define FRX_E_INVALID (-1)    /* bad argument */
#define FRX_E_CUDA (-2)       /* CUDA runtime error (see frx_last_error) */
#define FRX_E_NOMEM (-3)
#define FRX_E_UNSUPPORTED (-4) /* e.g. > 8 channels per stream (FLAC__MAX_CHANNELS, dump:3501) */
#define FRX_E_OVERFLOW (-5)   /* output capacity too small */
#define FRX_E_BITSTREAM (-6)  /* decode: malformed stream / CRC mismatch */

/* raster dtypes (numpy names) */
enum frx_dtype { FRX_U8 = 0, FRX_I16 = 1, FRX_U16 = 2, FRX_I32 = 3, FRX_U32 = 4, FRX_F32 = 5 };

/* raster <-> PCM mapping */
enum frx_pcm_mode {
    FRX_PCM_REFERENCE = 0,         /* converter.py:56-86 (+ pyFLAC astype(int32), dump:1986-1994); decode =
                                      pyFLAC/soundfile int16/32768 + converter.py:99-110 */
    FRX_PCM_LOSSLESS = 1,          /* identity mapping (extension; SURVEY decision D1) */
    FRX_PCM_REFERENCE_SPATIAL = 2  /* spatial_encoder.py:229-248 followed by pyFLAC's int32 truncation */
};

/* One independent FLAC stream ("tile"): a window of the raster and a run of bands that
 * become its channels.  Mirrors one RasterFLACConverter.tiff_to_flac() call on a
 * window (cli.py:691-733) or one SpatialFLACEncoder._encode_tile_to_flac() call.      */
typedef struct frx_tile {
    int32_t row_off, col_off, h, w;
    int32_t first_band, n_channels; /* 1..8 */
} frx_tile;

typedef struct frx_encode_params {
    int32_t compression_level; /* 0..8, libFLAC preset semantics (dump:6926-6934) */
    int32_t blocksize;         /* samples per frame; the reference always passes 4096 */
    int32_t sample_rate;       /* written to frame headers (converter.py:44-51) */
    int32_t pcm_mode;          /* enum frx_pcm_mode */
} frx_encode_params;

/* Geometry of a band-planar raster in device (or host) memory, element strides. */
typedef struct frx_raster {
    const void *data;
    int32_t dtype; /* enum frx_dtype */
    int32_t bands;
    int64_t height, width;
    int64_t band_stride, row_stride; /* in elements; 0 => dense (height*width, width) */
} frx_raster;

/* Encoder outputs (device pointers unless noted).  Any pointer except `bytes` may be NULL. */
typedef struct frx_encode_out {
    uint8_t *bytes;          /* [capacity] FLAC frames of all tiles, tile after tile, frame after frame */
    uint64_t capacity;
    uint64_t *tile_offset;   /* [n_tiles] byte offset of the tile's first frame in `bytes` */
    uint64_t *tile_size;     /* [n_tiles] bytes of the tile's frames */
    double *tile_minmax;     /* [n_tiles][2] data_min, data_max over the tile's bands (converter.py:152-153) */
    uint32_t *frame_size;    /* [total_frames] bytes per frame (frame order = tile order, then time) */
    uint64_t *frame_offset;  /* [total_frames] byte offset of each frame in `bytes` */
    uint64_t *total_bytes;   /* [1] */
    int32_t *status;         /* [1] 0 ok, FRX_E_OVERFLOW if capacity was too small (nothing written out of bounds) */
} frx_encode_out;

/* One tile to decode: its frame region and where/how to write the raster block. */
typedef struct frx_dtile {
    uint64_t byte_offset, byte_size; /* FLAC frames (no stream header) inside `bytes` */
    uint64_t out_offset;             /* element offset of (band 0,row 0,col 0) in the output raster */
    int64_t out_band_stride, out_row_stride; /* elements */
    int32_t h, w, channels, bps;
    double dmin, dmax;               /* GEOSPATIAL_DATA_MIN/MAX of this stream */
    uint32_t first_frame;            /* index into frame_size[] of this tile's first frame, or
                                        0xFFFFFFFF: discover frame starts by sync-code scan */
    uint32_t n_frames;               /* frames in the tile (ceil(h*w/blocksize)) */
} frx_dtile;

typedef struct frx_ctx frx_ctx;

/* ---- library ------------------------------------------------------------------------ */
int frx_version(void);
const char *frx_last_error(void);
int frx_device_count(void);

int frx_ctx_create(int device, frx_ctx **out);
void frx_ctx_destroy(frx_ctx *ctx);

/* Number of FLAC frames the tile list produces, and a worst-case bound on the bytes. */
int frx_encode_bound(const frx_tile *tiles, int32_t n_tiles, int32_t dtype, const frx_encode_params *p,
                     uint64_t *total_frames, uint64_t *max_bytes);
/* Effective bits per sample of the PCM a raster dtype produces (pyFLAC semantics). */
int frx_pcm_bps(int32_t dtype, int32_t pcm_mode);

/* ---- hot path, device buffers ----------------------------------------------------------- */
/* K0 min/max -> K1 tile cut + interleave + quantise -> K2..K7 FLAC frame encode -> K8 offsets.
 * replaces: converter.py:152-216 (per tile), pyflac.StreamEncoder.process/finish.            */
int frx_encode_tiles(frx_ctx *ctx, const frx_raster *raster_dev, const frx_tile *tiles_host, int32_t n_tiles,
                     const frx_encode_params *p, const frx_encode_out *out_dev, void *stream);

/* K9 frame discovery -> K10 Rice unpack -> K11 predictor restore -> K12 PCM -> raster dtype.
 * replaces: pyflac.FileDecoder.process (converter.py:241-242) + converter.py:265-282.
 * frame_size_dev may be NULL when every tile has first_frame == 0xFFFFFFFF.
 * tile_status_dev: [n_tiles] 0 ok / FRX_E_BITSTREAM.                                          */
int frx_decode_tiles(frx_ctx *ctx, const uint8_t *bytes_dev, uint64_t n_bytes, const frx_dtile *tiles_host,
                     int32_t n_tiles, const uint32_t *frame_size_dev, int32_t blocksize, int32_t out_dtype,
                     int32_t pcm_mode, void *out_dev, int32_t *tile_status_dev, void *stream);

/* K0+K1 alone: PCM (int32, channel-interleaved like pyFLAC's input, dump:1994) of one tile. */
int frx_quantize(frx_ctx *ctx, const frx_raster *raster_dev, const frx_tile *tile_host, int32_t pcm_mode,
                 int32_t *pcm_dev, double *minmax_dev, void *stream);
/* STREAMINFO's MD5 signature of every tile's PCM (the channel-interleaved samples as little-endian bytes of the effective
 * sample size; layout docs/sonos-pyflac.txt:3938-3949): md5_dev[n_tiles][16].  The reference leaves the field zero
 * (libFLAC computes it only when the stream is finalised through seek callbacks, which converter.py:201-216 does not
 * supply), so this is opt-in; one thread per tile, MD5 being a serial chain -- not part of the timed path.             */
int frx_tile_md5(frx_ctx *ctx, const frx_raster *raster_dev, const frx_tile *tiles_host, int32_t n_tiles, int32_t pcm_mode,
                 uint8_t *md5_dev, void *stream);
/* K12 alone: int32 channel-interleaved PCM -> (channels,h,w) raster block. */
int frx_dequantize(frx_ctx *ctx, const int32_t *pcm_dev, int32_t h, int32_t w, int32_t channels, int32_t bps,
                   int32_t out_dtype, int32_t pcm_mode, double dmin, double dmax, void *out_dev, void *stream);

/* ---- hot path, host buffers (what a pyflac-style caller binds) ---------------------------- */
/* Same as frx_encode_tiles but `raster_host` and every output live in host memory; the call
 * stages H2D/D2H itself (chunked over internal streams) and returns when the outputs are
 * complete.  out_host->capacity bounds out_host->bytes.                                       */
int frx_encode_tiles_host(frx_ctx *ctx, const frx_raster *raster_host, const frx_tile *tiles_host,
                          int32_t n_tiles, const frx_encode_params *p, const frx_encode_out *out_host);
int frx_decode_tiles_host(frx_ctx *ctx, const uint8_t *bytes_host, uint64_t n_bytes, const frx_dtile *tiles_host,
                          int32_t n_tiles, const uint32_t *frame_size_host, int32_t blocksize, int32_t out_dtype,
                          int32_t pcm_mode, void *out_host, uint64_t out_elems, int32_t *tile_status_host);

/* ---- allocation-free decode ---------------------------------------------------------------- */
/* No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid allocation).  Same arguments, same kernels and same outputs as frx_decode_tiles, but the
 * call allocates nothing: no cudaMalloc/cudaFree, no allocating library routine.  Its scratch
 * (tile descriptors, frame positions and lengths, CRC tables; about 430 KB) is a file-scope
 * `static __device__` array inside libfrx, which every GPU that loads the library holds until
 * the process exits.  Limits: n_tiles <= FRX_NOALLOC_MAX_TILES, the tiles' frames together
 * <= FRX_NOALLOC_MAX_FRAMES, no two-channel tiles (their stereo-decorrelated frames need a scratch
 * row per frame) -- anything beyond returns FRX_E_UNSUPPORTED and says which limit was hit.
 * All calls on one device share the array, and it keeps the previous call's contents (the call
 * clears what must start at zero): two calls that overlap on the GPU -- different streams or
 * contexts of the same device -- race.  Calls on one stream, or ordered by events, are safe.   */
#define FRX_NOALLOC_MAX_TILES 128
#define FRX_NOALLOC_MAX_FRAMES 32768
int frx_decode_tiles_noalloc(frx_ctx *ctx, const uint8_t *bytes_dev, uint64_t n_bytes, const frx_dtile *tiles_host,
                             int32_t n_tiles, const uint32_t *frame_size_dev, int32_t blocksize, int32_t out_dtype,
                             int32_t pcm_mode, void *out_dev, int32_t *tile_status_dev, void *stream);

/* ---- raster ingestion -------------------------------------------------- */
/* One strip or tile of an uncompressed TIFF: where its bytes sit in the staged file image and
 * which rectangle of the raster they fill.  Chunk rows hold `pitch` pixels (the tile width; the
 * image width for strips), of which the first `cols` are stored; `rows`/`cols` are already
 * clipped to the image (TIFF pads edge tiles).  band < 0: pixel-interleaved chunk holding every
 * band (PlanarConfiguration 1); band >= 0: one band's chunk (PlanarConfiguration 2).           */
typedef struct frx_tiff_chunk {
    uint64_t src_offset;      /* byte offset of the chunk inside `file_dev` */
    int32_t row0, col0, rows, cols;
    int32_t pitch;
    int32_t band;
} frx_tiff_chunk;

/* File bytes (device) -> band-planar raster (device): de-interleave, un-tile, clip, byte-swap.
 * replaces: rasterio's read() into a (bands,h,w) ndarray followed by the H2D copy of that array
 * (converter.py:137-141, spatial_encoder.py:154-183, cli.py:699) -- the file image is copied to
 * the device once, as it lies on disk, and this kernel does the rearrangement there.
 * `dst_dev` describes the destination raster (dtype, bands, geometry; `data` is written).
 * samples_per_pixel: bands stored per pixel in interleaved chunks.  swap_bytes: file byte order
 * differs from the device's (big-endian "MM" files).                                            */
int frx_tiff_unpack(frx_ctx *ctx, const uint8_t *file_dev, uint64_t n_bytes, const frx_tiff_chunk *chunks_host,
                    int32_t n_chunks, int32_t samples_per_pixel, int32_t swap_bytes, const frx_raster *dst_dev,
                    void *stream);

/* ---- instrumentation ---------------------------------------------------------------------- */
/* Kernel launches issued by this context since creation (bench.py reports it). */
uint64_t frx_launch_count(const frx_ctx *ctx);
/* Device time (ms) of the dominant encode / decode kernel in the most recent call, measured
 * with CUDA events on the launching stream; valid after the stream has been synchronised.    */
int frx_last_kernel_ms(frx_ctx *ctx, float *encode_ms, float *decode_ms);

#ifdef __cplusplus
}
#endif
#endif /* FRX_H */

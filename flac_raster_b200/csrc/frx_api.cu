// frx_api.cu -- the C-ABI of libfrx.so (include/frx.h): context, scratch, launch sequencing.
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <pthread.h>
#include <string.h>

#include <vector>

#include "frx_internal.cuh"

// ---- error handling ----------------------------------------------------------------------
static thread_local char g_err[512] = "";
static int set_err(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}
#define CK(call)                                                                                              \
    do {                                                                                                      \
        cudaError_t _e = (call);                                                                              \
        if (_e != cudaSuccess) return set_err(FRX_E_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(_e), __FILE__, __LINE__); \
    } while (0)

struct Buf {
    void *p = nullptr;
    size_t cap = 0;
};

struct frx_ctx {
    int device = 0;
    int sm_count = 148;
    uint64_t launches = 0;
    Buf tiles, prefix, unit_prefix, mmraw, quant, desc, misc, xpow, crctab, fsize, foff, slots;
    Buf dtiles, dprefix, fpos, flen, stereo, scan_prefix, scan_tiles, ing_chunks, ing_prefix, d3sub, d3blk, d3res;
    Buf h_raster, h_bytes, h_small, h_out;  // device staging for the *_host entry points
    int xpow_L = -1;
    bool crctab_ready = false;
    cudaEvent_t ev[4] = { nullptr, nullptr, nullptr, nullptr };
    bool enc_timed = false, dec_timed = false;
    int verify_crc = 1;
    // *_host pipeline: copy-in / compute / copy-out streams, one event per tile group, pinned per-group totals
    cudaStream_t hs[3] = { nullptr, nullptr, nullptr };
    std::vector<cudaEvent_t> hev;
    void *pin_small = nullptr;
    size_t pin_small_cap = 0;
};

static int ensure(Buf &b, size_t bytes) {
    if (bytes <= b.cap && b.p) return FRX_OK;
    if (b.p) cudaFree(b.p);
    b.p = nullptr; b.cap = 0;
    size_t want = bytes < 256 ? 256 : bytes;
    cudaError_t e = cudaMalloc(&b.p, want);
    if (e != cudaSuccess) return set_err(FRX_E_NOMEM, "cudaMalloc(%zu) failed: %s", want, cudaGetErrorString(e));
    b.cap = want;
    return FRX_OK;
}
#define ENSURE(buf, bytes) do { int _r = ensure(buf, bytes); if (_r) return _r; } while (0)

static int ensure_crctab(frx_ctx *c) {
    if (!c->crctab_ready) {
        uint16_t tab[1024];
        frx_make_crc16_tables(tab);
        ENSURE(c->crctab, sizeof(tab));
        CK(cudaMemcpy(c->crctab.p, tab, sizeof(tab), cudaMemcpyHostToDevice));
        c->crctab_ready = true;
    }
    return FRX_OK;
}

// CRC chunk combine tables (frx_encode.cu, frx_crc_check_kernel): x^(128 * Lq * j) mod P for Lq = 1..FRX_CRC_MAXLQ, j = 0..255;
// x^(-32 z), z = 0..3
static int ensure_xpow(frx_ctx *c) {
    if (c->xpow_L != 2) {
        std::vector<uint16_t> xp(FRX_CRC_MAXLQ * 256 + 4);
        for (int lq = 1; lq <= FRX_CRC_MAXLQ; lq++) {
            uint32_t base = 2, e = 128u * lq, acc = 1;
            while (e) { if (e & 1u) acc = frx_gf16_mulmod(acc, base); base = frx_gf16_mulmod(base, base); e >>= 1; }
            uint32_t cur = 1;
            for (int j = 0; j < 256; j++) { xp[(lq - 1) * 256 + j] = (uint16_t)cur; cur = frx_gf16_mulmod(cur, acc); }
        }
        uint32_t inv32 = 0xC002u;                       // x^-1: x * (x^15 + x^14 + x) = P + 1
        for (int i = 0; i < 5; i++) inv32 = frx_gf16_mulmod(inv32, inv32);   // x^-32
        uint32_t cur = 1;
        for (int z = 0; z < 4; z++) { xp[FRX_CRC_MAXLQ * 256 + z] = (uint16_t)cur; cur = frx_gf16_mulmod(cur, inv32); }
        ENSURE(c->xpow, sizeof(uint16_t) * xp.size());
        CK(cudaMemcpy(c->xpow.p, xp.data(), sizeof(uint16_t) * xp.size(), cudaMemcpyHostToDevice));
        c->xpow_L = 2;
    }
    return FRX_OK;
}

extern "C" {

int frx_version(void) { return FRX_VERSION_MAJOR * 100 + FRX_VERSION_MINOR; }
const char *frx_last_error(void) { return g_err; }

int frx_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int frx_ctx_create(int device, frx_ctx **out) {
    if (!out) return set_err(FRX_E_INVALID, "out is NULL");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return set_err(FRX_E_CUDA, "no CUDA device available (%s); libfrx has no CPU fallback", cudaGetErrorString(e));
    if (device < 0 || device >= n) return set_err(FRX_E_INVALID, "device %d out of range (0..%d)", device, n - 1);
    CK(cudaSetDevice(device));
    frx_ctx *c = new frx_ctx();
    c->device = device;
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device));
    c->sm_count = prop.multiProcessorCount;
    for (int i = 0; i < 4; i++) CK(cudaEventCreate(&c->ev[i]));
    *out = c;
    return FRX_OK;
}

void frx_ctx_destroy(frx_ctx *c) {
    if (!c) return;
    cudaSetDevice(c->device);
    Buf *all[] = { &c->tiles, &c->prefix, &c->unit_prefix, &c->mmraw, &c->quant, &c->desc, &c->misc, &c->xpow,
                   &c->crctab, &c->fsize, &c->foff, &c->slots, &c->d3sub, &c->d3blk, &c->d3res, &c->dtiles, &c->dprefix, &c->fpos, &c->flen, &c->stereo, &c->scan_prefix, &c->scan_tiles, &c->ing_chunks, &c->ing_prefix,
                   &c->h_raster, &c->h_bytes, &c->h_small, &c->h_out };
    for (Buf *b : all) if (b->p) cudaFree(b->p);
    for (int i = 0; i < 4; i++) if (c->ev[i]) cudaEventDestroy(c->ev[i]);
    for (int i = 0; i < 3; i++) if (c->hs[i]) cudaStreamDestroy(c->hs[i]);
    for (cudaEvent_t e : c->hev) cudaEventDestroy(e);
    if (c->pin_small) cudaFreeHost(c->pin_small);
    delete c;
}

uint64_t frx_launch_count(const frx_ctx *c) { return c ? c->launches : 0; }

int frx_last_kernel_ms(frx_ctx *c, float *enc_ms, float *dec_ms) {
    if (!c) return set_err(FRX_E_INVALID, "ctx is NULL");
    CK(cudaSetDevice(c->device));
    if (enc_ms) { *enc_ms = 0.f; if (c->enc_timed) CK(cudaEventElapsedTime(enc_ms, c->ev[0], c->ev[1])); }
    if (dec_ms) { *dec_ms = 0.f; if (c->dec_timed) CK(cudaEventElapsedTime(dec_ms, c->ev[2], c->ev[3])); }
    return FRX_OK;
}

int frx_pcm_bps(int32_t dtype, int32_t pcm_mode) {
    if (dtype < FRX_U8 || dtype > FRX_F32) return set_err(FRX_E_INVALID, "bad dtype %d", dtype);
    return frx_bps_for(dtype, pcm_mode);
}

static int check_params(const frx_encode_params *p) {
    if (!p) return set_err(FRX_E_INVALID, "params is NULL");
    if (p->blocksize < 16 || p->blocksize > FRX_MAX_BLOCK)
        return set_err(FRX_E_UNSUPPORTED, "blocksize %d outside 16..%d", p->blocksize, FRX_MAX_BLOCK);
    if (p->compression_level < 0 || p->compression_level > 8) return set_err(FRX_E_INVALID, "compression_level %d outside 0..8", p->compression_level);
    if (p->pcm_mode < 0 || p->pcm_mode > 2) return set_err(FRX_E_INVALID, "bad pcm_mode %d", p->pcm_mode);
    if (p->sample_rate <= 0 || p->sample_rate > 655350) return set_err(FRX_E_INVALID, "bad sample_rate %d", p->sample_rate);
    return FRX_OK;
}

static int check_tiles(const frx_tile *tiles, int n, int bands, long long H, long long W) {
    if (!tiles || n <= 0) return set_err(FRX_E_INVALID, "no tiles");
    for (int i = 0; i < n; i++) {
        const frx_tile &t = tiles[i];
        if (t.n_channels < 1 || t.n_channels > FRX_MAX_CH)
            return set_err(FRX_E_UNSUPPORTED, "tile %d: %d channels; FLAC allows 1..8 per stream (EncoderInitException in the reference)", i, t.n_channels);
        if (t.h <= 0 || t.w <= 0 || t.row_off < 0 || t.col_off < 0 || (bands >= 0 && (t.first_band < 0 || t.first_band + t.n_channels > bands)) ||
            (H >= 0 && ((long long)t.row_off + t.h > H || (long long)t.col_off + t.w > W)))
            return set_err(FRX_E_INVALID, "tile %d: window/bands outside the raster", i);
    }
    return FRX_OK;
}

int frx_encode_bound(const frx_tile *tiles, int32_t n_tiles, int32_t dtype, const frx_encode_params *p, uint64_t *total_frames,
                     uint64_t *max_bytes) {
    int r = check_params(p);
    if (r) return r;
    r = check_tiles(tiles, n_tiles, -1, -1, -1);
    if (r) return r;
    if (dtype < FRX_U8 || dtype > FRX_F32) return set_err(FRX_E_INVALID, "bad dtype %d", dtype);
    int bps = frx_bps_for(dtype, p->pcm_mode);
    uint64_t nf = 0, nb = 0;
    for (int i = 0; i < n_tiles; i++) {
        uint64_t n = (uint64_t)tiles[i].h * tiles[i].w;
        uint64_t f = (n + p->blocksize - 1) / p->blocksize;
        nf += f;
        nb += f * (16ull + 3ull * tiles[i].n_channels + 2ull) + (n * tiles[i].n_channels * bps + 7) / 8 + 8ull * f;
    }
    if (total_frames) *total_frames = nf;
    if (max_bytes) *max_bytes = nb + 64;
    return FRX_OK;
}

static int sample_rate_code(int sr) {
    switch (sr) {
        case 88200: return 1; case 176400: return 2; case 192000: return 3; case 8000: return 4; case 16000: return 5;
        case 22050: return 6; case 24000: return 7; case 32000: return 8; case 44100: return 9; case 48000: return 10;
        case 96000: return 11;
    }
    return sr <= 65535 ? 13 : 0;
}

static int qlp_precision_for(int bps, int blocksize) {  // libFLAC's automatic precision
    if (bps < 16) { int p = 2 + bps / 2; return p < 5 ? 5 : p; }
    if (bps == 16) {
        if (blocksize <= 192) return 7; if (blocksize <= 384) return 8; if (blocksize <= 576) return 9;
        if (blocksize <= 1152) return 10; if (blocksize <= 2304) return 11; if (blocksize <= 4608) return 12;
        return 13;
    }
    if (blocksize <= 384) return 13;
    if (blocksize <= 1152) return 14;
    return 15;
}

// Everything the encode path needs on the device before the main kernel: tiles, prefixes, min/max.
struct EncPrep {
    uint32_t total_frames = 0;
    int bps = 0;
    int max_channels = 1;
};

static int encode_prepare(frx_ctx *c, const frx_raster &r, const frx_tile *tiles, int n_tiles, const frx_encode_params *p,
                          EncPrep &prep, cudaStream_t st) {
    std::vector<uint32_t> fprefix(n_tiles + 1), uprefix(n_tiles + 1);
    const int unit_elems = 65536;
    uint64_t nf = 0, nu = 0;
    int maxc = 1;
    for (int i = 0; i < n_tiles; i++) {
        fprefix[i] = (uint32_t)nf; uprefix[i] = (uint32_t)nu;
        uint64_t n = (uint64_t)tiles[i].h * tiles[i].w;
        nf += (n + p->blocksize - 1) / p->blocksize;
        int rpu = unit_elems / tiles[i].w; if (rpu < 1) rpu = 1;
        uint64_t vrows = (uint64_t)tiles[i].n_channels * tiles[i].h;
        nu += (vrows + rpu - 1) / rpu;
        if (tiles[i].n_channels > maxc) maxc = tiles[i].n_channels;
        if (nf > 0xFFFFFFF0ull || nu > 0x7FFFFFF0ull) return set_err(FRX_E_UNSUPPORTED, "too many frames in one call");
    }
    fprefix[n_tiles] = (uint32_t)nf; uprefix[n_tiles] = (uint32_t)nu;
    prep.total_frames = (uint32_t)nf;
    prep.bps = frx_bps_for(r.dtype, p->pcm_mode);
    prep.max_channels = maxc;
    ENSURE(c->tiles, sizeof(frx_tile) * n_tiles);
    ENSURE(c->prefix, sizeof(uint32_t) * (n_tiles + 1));
    ENSURE(c->unit_prefix, sizeof(uint32_t) * (n_tiles + 1));
    ENSURE(c->mmraw, frx_minmax_raw_size() * n_tiles);
    ENSURE(c->quant, sizeof(FrxQuant) * n_tiles);
    CK(cudaMemcpyAsync(c->tiles.p, tiles, sizeof(frx_tile) * n_tiles, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(c->prefix.p, fprefix.data(), sizeof(uint32_t) * (n_tiles + 1), cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(c->unit_prefix.p, uprefix.data(), sizeof(uint32_t) * (n_tiles + 1), cudaMemcpyHostToDevice, st));
    // the vectors die at return: pageable async copies are staged before the call returns, but be explicit
    CK(cudaStreamSynchronize(st));
    CK(frx_launch_minmax(r, (const frx_tile *)c->tiles.p, (const uint32_t *)c->unit_prefix.p, n_tiles, (uint32_t)nu, unit_elems,
                         c->mmraw.p, (FrxQuant *)c->quant.p, st));
    c->launches += 3;
    return FRX_OK;
}

static frx_raster normalise_raster(const frx_raster *in) {
    frx_raster r = *in;
    if (r.row_stride == 0) r.row_stride = r.width;
    if (r.band_stride == 0) r.band_stride = r.height * r.width;
    return r;
}

static int check_raster(const frx_raster *r) {
    if (!r || !r->data) return set_err(FRX_E_INVALID, "raster is NULL");
    if (r->dtype < FRX_U8 || r->dtype > FRX_F32) return set_err(FRX_E_INVALID, "bad raster dtype %d", r->dtype);
    if (r->bands < 1 || r->height < 1 || r->width < 1) return set_err(FRX_E_INVALID, "bad raster shape");
    return FRX_OK;
}

int frx_encode_tiles(frx_ctx *c, const frx_raster *raster_dev, const frx_tile *tiles, int32_t n_tiles, const frx_encode_params *p,
                     const frx_encode_out *out, void *stream) {
    if (!c) return set_err(FRX_E_INVALID, "ctx is NULL");
    int rc = check_params(p); if (rc) return rc;
    rc = check_raster(raster_dev); if (rc) return rc;
    if (!out || !out->bytes) return set_err(FRX_E_INVALID, "out->bytes is NULL");
    frx_raster r = normalise_raster(raster_dev);
    rc = check_tiles(tiles, n_tiles, r.bands, r.height, r.width); if (rc) return rc;
    if (p->pcm_mode == FRX_PCM_REFERENCE_SPATIAL && r.dtype == FRX_U32)
        return set_err(FRX_E_UNSUPPORTED, "reference --spatial mapping yields a float64 array for uint32 input, which pyFLAC rejects (bps 64)");
    CK(cudaSetDevice(c->device));
    cudaStream_t st = (cudaStream_t)stream;
    EncPrep prep;
    rc = encode_prepare(c, r, tiles, n_tiles, p, prep, st); if (rc) return rc;

    const int frame_words = 0;   // (round 1 sized per-CTA frame buffers with this; a frame warp now packs in place)
    { int rc2 = ensure_xpow(c); if (rc2) return rc2; }
    {
        int rc2 = ensure_crctab(c);
        if (rc2) return rc2;
    }
    ENSURE(c->desc, sizeof(uint64_t) * (size_t)prep.total_frames);
    ENSURE(c->misc, 64);
    uint32_t *frame_size = out->frame_size;
    uint64_t *frame_offset = out->frame_offset;
    if (!frame_size) { ENSURE(c->fsize, sizeof(uint32_t) * (size_t)prep.total_frames); frame_size = (uint32_t *)c->fsize.p; }
    if (!frame_offset) { ENSURE(c->foff, sizeof(uint64_t) * (size_t)prep.total_frames); frame_offset = (uint64_t *)c->foff.p; }
    CK(cudaMemsetAsync(c->desc.p, 0, sizeof(uint64_t) * (size_t)prep.total_frames, st));
    CK(cudaMemsetAsync(c->misc.p, 0, 64, st));
    if (out->status) CK(cudaMemsetAsync(out->status, 0, sizeof(int32_t), st));

    static const int lvl_lpc[9] = { 0, 0, 0, 6, 8, 8, 8, 12, 12 };
    static const int lvl_po[9] = { 3, 3, 3, 4, 4, 5, 6, 6, 6 };
    EncArgs a;
    memset(&a, 0, sizeof(a));
    a.raster = r.data; a.H = r.height; a.W = r.width; a.band_stride = r.band_stride; a.row_stride = r.row_stride;
    a.tiles = (const frx_tile *)c->tiles.p; a.frame_prefix = (const uint32_t *)c->prefix.p; a.quant = (const FrxQuant *)c->quant.p;
    a.crc_xpow = (const uint16_t *)c->xpow.p; a.crc_xinv = a.crc_xpow + FRX_CRC_MAXLQ * 256;
    a.bytes = out->bytes; a.capacity = out->capacity; a.frame_size = frame_size; a.frame_offset = frame_offset;
    a.desc = (uint64_t *)c->desc.p; a.ticket = (uint32_t *)c->misc.p;
    a.status = out->status ? out->status : (int32_t *)((char *)c->misc.p + 16);
    a.total_bytes = out->total_bytes ? out->total_bytes : (uint64_t *)((char *)c->misc.p + 32);
    a.total_frames = prep.total_frames; a.n_tiles = n_tiles; a.blocksize = p->blocksize; a.bps = prep.bps; a.pcm_mode = p->pcm_mode;
    a.sr_code = sample_rate_code(p->sample_rate); a.sample_rate = p->sample_rate;
    a.max_lpc_order = lvl_lpc[p->compression_level]; a.max_po = lvl_po[p->compression_level];
    a.qlp_precision = qlp_precision_for(prep.bps, p->blocksize);
    a.frame_words = frame_words; a.dtype = r.dtype; a.crc_tab = (const uint16_t *)c->crctab.p;
    {
        const long long esz = frx_dtype_size(r.dtype);
        a.tma_ok = (((uintptr_t)r.data & 15u) == 0 && (r.band_stride * esz) % 16 == 0 && (r.row_stride * esz) % 16 == 0) ? 1 : 0;
    }
    {
        static const int lvl_ms[9] = { 0, 1, 1, 0, 1, 1, 1, 1, 1 };   // libFLAC presets with mid-side coding (dump:6926-6934)
        a.stereo = (lvl_ms[p->compression_level] && prep.bps <= 16) ? 1 : 0;
    }
    {
        // two slots per resident frame warp (streamed and parked frames, frx_encode.cu): each holds the largest frame of this
        // call (all channels verbatim + headers), 16-byte multiples
        const uint32_t stride = (uint32_t)((((uint64_t)prep.max_channels * ((uint64_t)p->blocksize * 4u + 16u) + 64u) + 15u) & ~15ull);
        ENSURE(c->slots, 2u * (size_t)stride * (size_t)c->sm_count * (size_t)frx_enc_warps_per_cta());
        a.slots = (uint8_t *)c->slots.p; a.slot_stride = stride;
        // per channel: four barriers in the analysis, one after packing; a stereo frame analyses four virtual channels
        a.bars_per_frame = 5 * prep.max_channels;
        if (a.stereo && a.bars_per_frame < 18) a.bars_per_frame = 18;
    }
    CK(cudaEventRecord(c->ev[0], st));
    CK(frx_launch_encode(a, r.dtype, c->sm_count, st));
    CK(cudaEventRecord(c->ev[1], st));
    c->enc_timed = true;
    CK(frx_launch_tile_index((const uint32_t *)c->prefix.p, frame_size, frame_offset, (const FrxQuant *)c->quant.p, r.dtype, n_tiles,
                             out->tile_offset, out->tile_size, out->tile_minmax, st));
    c->launches += 2;
    return FRX_OK;
}

int frx_tiff_unpack(frx_ctx *c, const uint8_t *file_dev, uint64_t n_bytes, const frx_tiff_chunk *chunks, int32_t n_chunks,
                    int32_t samples_per_pixel, int32_t swap_bytes, const frx_raster *dst_dev, void *stream) {
    if (!c) return set_err(FRX_E_INVALID, "ctx is NULL");
    if (!file_dev) return set_err(FRX_E_INVALID, "file_dev is NULL");
    int rc = check_raster(dst_dev); if (rc) return rc;
    if (n_chunks < 0 || (n_chunks > 0 && !chunks)) return set_err(FRX_E_INVALID, "bad chunk list");
    if (samples_per_pixel < 1 || samples_per_pixel > 64) return set_err(FRX_E_INVALID, "samples_per_pixel %d out of range", samples_per_pixel);
    frx_raster r = normalise_raster(dst_dev);
    const int esz = frx_dtype_size(r.dtype);
    std::vector<uint32_t> prefix((size_t)n_chunks + 1);
    uint64_t nu = 0;
    for (int i = 0; i < n_chunks; i++) {
        const frx_tiff_chunk &k = chunks[i];
        prefix[i] = (uint32_t)nu;
        if (k.rows < 0 || k.cols < 0 || k.row0 < 0 || k.col0 < 0 || k.pitch < k.cols || (int64_t)k.row0 + k.rows > r.height ||
            (int64_t)k.col0 + k.cols > r.width || k.band >= r.bands || (k.band < 0 && samples_per_pixel > r.bands))
            return set_err(FRX_E_INVALID, "chunk %d lies outside the raster", i);
        if (k.rows == 0 || k.cols == 0) continue;
        const uint64_t nb = k.band < 0 ? (uint64_t)samples_per_pixel : 1u;
        const uint64_t span = ((uint64_t)(k.rows - 1) * (uint64_t)k.pitch + (uint64_t)k.cols) * nb * (uint64_t)esz;
        if (k.src_offset > n_bytes || span > n_bytes - k.src_offset) return set_err(FRX_E_INVALID, "chunk %d reads past the file image", i);
        { int rpu, spx; nu += frx_ingest_plan(esz, (int)nb, k.rows, k.cols, k.pitch, &rpu, &spx); }
        if (nu > 0x7FFFFFF0ull) return set_err(FRX_E_UNSUPPORTED, "too many row segments in one call");
    }
    prefix[n_chunks] = (uint32_t)nu;
    if (!nu) return FRX_OK;
    CK(cudaSetDevice(c->device));
    cudaStream_t st = (cudaStream_t)stream;
    ENSURE(c->ing_chunks, sizeof(frx_tiff_chunk) * (size_t)n_chunks);
    ENSURE(c->ing_prefix, sizeof(uint32_t) * ((size_t)n_chunks + 1));
    CK(cudaMemcpyAsync(c->ing_chunks.p, chunks, sizeof(frx_tiff_chunk) * (size_t)n_chunks, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(c->ing_prefix.p, prefix.data(), sizeof(uint32_t) * ((size_t)n_chunks + 1), cudaMemcpyHostToDevice, st));
    CK(cudaStreamSynchronize(st));   // `prefix` dies at return
    CK(cudaEventRecord(c->ev[0], st));
    CK(frx_launch_tiff_unpack(file_dev, (const frx_tiff_chunk *)c->ing_chunks.p, (const uint32_t *)c->ing_prefix.p, n_chunks, (uint32_t)nu,
                              samples_per_pixel, swap_bytes ? 1 : 0, r, st));
    CK(cudaEventRecord(c->ev[1], st));
    c->enc_timed = true;   // frx_last_kernel_ms reports this kernel until the next encode call
    c->launches += 1;
    return FRX_OK;
}

int frx_quantize(frx_ctx *c, const frx_raster *raster_dev, const frx_tile *tile, int32_t pcm_mode, int32_t *pcm_dev, double *minmax_dev,
                 void *stream) {
    if (!c) return set_err(FRX_E_INVALID, "ctx is NULL");
    int rc = check_raster(raster_dev); if (rc) return rc;
    if (!pcm_dev) return set_err(FRX_E_INVALID, "pcm_dev is NULL");
    frx_raster r = normalise_raster(raster_dev);
    rc = check_tiles(tile, 1, r.bands, r.height, r.width); if (rc) return rc;
    CK(cudaSetDevice(c->device));
    cudaStream_t st = (cudaStream_t)stream;
    frx_encode_params p = { 5, 4096, 44100, pcm_mode };
    rc = check_params(&p); if (rc) return rc;
    EncPrep prep;
    rc = encode_prepare(c, r, tile, 1, &p, prep, st); if (rc) return rc;
    CK(frx_launch_quantize(r, *tile, (const FrxQuant *)c->quant.p, pcm_mode, pcm_dev, minmax_dev, st));
    c->launches += minmax_dev ? 2 : 1;
    return FRX_OK;
}


static int decode_core(frx_ctx *c, const uint8_t *bytes_dev, uint64_t n_bytes, const frx_dtile *tiles, int32_t n_tiles,
                       const uint32_t *frame_size_dev, int32_t blocksize, int32_t out_dtype, int32_t pcm_mode, void *out_dev,
                       int32_t *tile_status_dev, cudaStream_t st, uint64_t out_elems = 0) {
    if (!bytes_dev || !tiles || n_tiles <= 0 || !out_dev || !tile_status_dev) return set_err(FRX_E_INVALID, "NULL argument");
    if (blocksize < 16 || blocksize > 65535) return set_err(FRX_E_INVALID, "bad blocksize %d", blocksize);
    if (out_dtype < FRX_U8 || out_dtype > FRX_F32) return set_err(FRX_E_INVALID, "bad out_dtype %d", out_dtype);
    std::vector<uint32_t> dprefix(n_tiles + 1), sprefix;
    std::vector<int32_t> stiles;
    uint64_t nf = 0, nu = 0;
    for (int i = 0; i < n_tiles; i++) {
        const frx_dtile &t = tiles[i];
        if (t.channels < 1 || t.channels > FRX_MAX_CH || t.h <= 0 || t.w <= 0 || (t.bps != 16 && t.bps != 32 && t.bps != 8 && t.bps != 24 && t.bps != 12 && t.bps != 20))
            return set_err(FRX_E_INVALID, "tile %d: bad geometry/bps", i);
        if (t.byte_offset > n_bytes || t.byte_size > n_bytes - t.byte_offset) return set_err(FRX_E_INVALID, "tile %d: byte range outside the buffer", i);
        if (out_elems) {   // the caller told us the size of the output: every element the tile writes must lie inside it
            if (t.out_band_stride < 0 || t.out_row_stride < 0) return set_err(FRX_E_INVALID, "tile %d: negative output stride", i);
            const unsigned __int128 last = (unsigned __int128)t.out_offset + (unsigned __int128)(t.channels - 1) * (uint64_t)t.out_band_stride +
                                           (unsigned __int128)(t.h - 1) * (uint64_t)t.out_row_stride + (uint64_t)t.w;
            if (last > (unsigned __int128)out_elems) return set_err(FRX_E_INVALID, "tile %d: output block ends at element %llu, beyond the %llu the output holds", i,
                                                                     (unsigned long long)last, (unsigned long long)out_elems);
        }
        uint64_t expect = ((uint64_t)t.h * t.w + blocksize - 1) / blocksize;
        if (t.n_frames != expect) return set_err(FRX_E_INVALID, "tile %d: n_frames %u != ceil(h*w/blocksize) %llu", i, t.n_frames, (unsigned long long)expect);
        if (t.first_frame != 0xFFFFFFFFu && !frame_size_dev) return set_err(FRX_E_INVALID, "tile %d needs frame_size", i);
        dprefix[i] = (uint32_t)nf;
        nf += t.n_frames;
        if (t.first_frame == 0xFFFFFFFFu) {
            sprefix.push_back((uint32_t)nu);
            stiles.push_back(i);
            nu += (t.byte_size + 4095) / 4096;
        }
        if (nf > 0xFFFFFFF0ull || nu > 0x7FFFFFF0ull) return set_err(FRX_E_UNSUPPORTED, "too many frames in one call");
    }
    dprefix[n_tiles] = (uint32_t)nf;
    sprefix.push_back((uint32_t)nu);
    ENSURE(c->dtiles, sizeof(frx_dtile) * n_tiles);
    ENSURE(c->dprefix, sizeof(uint32_t) * (n_tiles + 1));
    ENSURE(c->fpos, sizeof(uint64_t) * (size_t)(nf + 1));
    ENSURE(c->flen, sizeof(uint32_t) * (size_t)(nf + 1));
    bool two_ch = false, two_ch_32 = false;
    for (int i = 0; i < n_tiles; i++) { two_ch = two_ch || tiles[i].channels == 2; two_ch_32 = two_ch_32 || (tiles[i].channels == 2 && tiles[i].bps == 32); }
    // scratch rows for stereo decorrelation: int32 per sample, int64 when a 33-bit side channel can occur
    if (two_ch) ENSURE(c->stereo, (two_ch_32 ? sizeof(int64_t) : sizeof(int32_t)) * (size_t)nf * (size_t)blocksize);
    ENSURE(c->scan_prefix, sizeof(uint32_t) * sprefix.size());
    ENSURE(c->scan_tiles, sizeof(int32_t) * (stiles.size() + 1));
    { int rc2 = ensure_crctab(c); if (rc2) return rc2; }
    { int rc2 = ensure_xpow(c); if (rc2) return rc2; }
    CK(cudaMemcpyAsync(c->dtiles.p, tiles, sizeof(frx_dtile) * n_tiles, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(c->dprefix.p, dprefix.data(), sizeof(uint32_t) * (n_tiles + 1), cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(c->scan_prefix.p, sprefix.data(), sizeof(uint32_t) * sprefix.size(), cudaMemcpyHostToDevice, st));
    if (!stiles.empty()) CK(cudaMemcpyAsync(c->scan_tiles.p, stiles.data(), sizeof(int32_t) * stiles.size(), cudaMemcpyHostToDevice, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaMemsetAsync(tile_status_dev, 0, sizeof(int32_t) * n_tiles, st));
    DecArgs a;
    memset(&a, 0, sizeof(a));
    a.bytes = bytes_dev; a.n_bytes = n_bytes; a.tiles = (const frx_dtile *)c->dtiles.p; a.dprefix = (const uint32_t *)c->dprefix.p;
    a.frame_pos = (uint64_t *)c->fpos.p; a.frame_len = (uint32_t *)c->flen.p; a.stereo_scratch = two_ch ? (int32_t *)c->stereo.p : nullptr; a.frame_size = frame_size_dev; a.crc_tab = (const uint16_t *)c->crctab.p;
    a.out = out_dev; a.tile_status = tile_status_dev; a.total_frames = (uint32_t)nf; a.n_tiles = n_tiles; a.blocksize = blocksize;
    a.pcm_mode = pcm_mode; a.verify_crc = c->verify_crc;
    CK(frx_launch_frame_pos(a, st));
    c->launches += 1;
    if (!stiles.empty()) {
        CK(frx_launch_discover(a, (const uint32_t *)c->scan_prefix.p, (const int32_t *)c->scan_tiles.p, (int)stiles.size(), (uint32_t)nu, st));
        c->launches += 1;
    }
    CK(cudaEventRecord(c->ev[2], st));
    // Small calls are latency-bound: one lane of the single-kernel decoder walks a whole frame, channel after channel (1.7 ms for
    // one 1024 x 1024 float32 tile, 5.5 ms for a three-band uint16 tile, however idle the GPU is).  They take the phased decoder
    // (frx_decode3.cu: codes scanned, then the residuals of all channels decoded 128 samples per lane, the recurrence restored per
    // subframe, output written -- four short kernels: 1.4 and 3.0 ms for those tiles).  Its throughput is lower, so it is used up to
    // 8192 frames per channel of the widest tile (the measured crossovers: 32 one-band tiles, 128 three-band tiles of 256 frames;
    // profiles/r2_decode_experiments.md) and 2 GB of scratch rows.  FRX_D3_MAX_FRAMES=n forces it for every call of up to n frames
    // (0: never).
    static const long long d3_env = getenv("FRX_D3_MAX_FRAMES") ? atoll(getenv("FRX_D3_MAX_FRAMES")) : -1;
    int chmax = 1;
    for (int i = 0; i < n_tiles; i++) chmax = tiles[i].channels > chmax ? tiles[i].channels : chmax;
    bool phased = frx_decode3_supports(blocksize) && !two_ch_32;      // (33-bit side channels: single-kernel decoder only)
    if (phased) {
        size_t sb, bb, rb;
        frx_decode3_scratch((uint32_t)((nf + 31ull) / 32ull * 32ull), chmax, blocksize, &sb, &bb, &rb);
        phased = d3_env >= 0 ? (long long)nf <= d3_env : (nf <= 8192ull * (uint64_t)chmax && rb <= (2ull << 30));
    }
    if (phased) {
        size_t sb, bb, rb;
        frx_decode3_scratch((uint32_t)((nf + 31ull) / 32ull * 32ull), chmax, blocksize, &sb, &bb, &rb);
        ENSURE(c->d3sub, sb);
        ENSURE(c->d3blk, bb);
        ENSURE(c->d3res, rb);
        CK(frx_launch_decode3(a, c->d3sub.p, c->d3blk.p, c->d3res.p, 0u, (uint32_t)nf, chmax, out_dtype, st));
        c->launches += 3;
    } else {
        CK(frx_launch_decode(a, out_dtype, c->sm_count, st));
    }
    if (c->verify_crc) { CK(frx_launch_crc_check(a, (const uint16_t *)c->xpow.p, c->sm_count, st)); c->launches += 1; }
    CK(cudaEventRecord(c->ev[3], st));
    c->dec_timed = true;
    c->launches += 1;
    return FRX_OK;
}

int frx_decode_tiles(frx_ctx *c, const uint8_t *bytes_dev, uint64_t n_bytes, const frx_dtile *tiles, int32_t n_tiles,
                     const uint32_t *frame_size_dev, int32_t blocksize, int32_t out_dtype, int32_t pcm_mode, void *out_dev,
                     int32_t *tile_status_dev, void *stream) {
    if (!c) return set_err(FRX_E_INVALID, "ctx is NULL");
    CK(cudaSetDevice(c->device));
    return decode_core(c, bytes_dev, n_bytes, tiles, n_tiles, frame_size_dev, blocksize, out_dtype, pcm_mode, out_dev, tile_status_dev,
                       (cudaStream_t)stream);
}

int frx_tile_md5(frx_ctx *c, const frx_raster *raster_dev, const frx_tile *tiles, int32_t n_tiles, int32_t pcm_mode, uint8_t *md5_dev, void *stream) {
    if (!c) return set_err(FRX_E_INVALID, "ctx is NULL");
    int rc = check_raster(raster_dev); if (rc) return rc;
    if (!md5_dev) return set_err(FRX_E_INVALID, "md5_dev is NULL");
    frx_raster r = normalise_raster(raster_dev);
    rc = check_tiles(tiles, n_tiles, r.bands, r.height, r.width); if (rc) return rc;
    CK(cudaSetDevice(c->device));
    cudaStream_t st = (cudaStream_t)stream;
    frx_encode_params p = { 5, 4096, 44100, pcm_mode };
    rc = check_params(&p); if (rc) return rc;
    EncPrep prep;
    rc = encode_prepare(c, r, tiles, n_tiles, &p, prep, st); if (rc) return rc;   // K0: the quantiser needs each stream's min/max
    CK(frx_launch_md5(r, (const frx_tile *)c->tiles.p, (const FrxQuant *)c->quant.p, n_tiles, pcm_mode, prep.bps, md5_dev, st));
    c->launches += 1;
    return FRX_OK;
}

int frx_dequantize(frx_ctx *c, const int32_t *pcm_dev, int32_t h, int32_t w, int32_t channels, int32_t bps, int32_t out_dtype,
                   int32_t pcm_mode, double dmin, double dmax, void *out_dev, void *stream) {
    if (!c) return set_err(FRX_E_INVALID, "ctx is NULL");
    if (!pcm_dev || !out_dev || h <= 0 || w <= 0 || channels < 1) return set_err(FRX_E_INVALID, "bad argument");
    if (out_dtype < FRX_U8 || out_dtype > FRX_F32) return set_err(FRX_E_INVALID, "bad out_dtype %d", out_dtype);
    CK(cudaSetDevice(c->device));
    CK(frx_launch_dequantize(pcm_dev, h, w, channels, bps, out_dtype, pcm_mode, dmin, dmax, out_dev, (cudaStream_t)stream));
    c->launches += 1;
    return FRX_OK;
}

// ---- host-buffer entry points: stage, run, fetch ---------------------------------------------
// Host buffers in, host buffers out.  The tiles are cut into groups of consecutive tiles (a group = the tiles that
// share a band of raster rows, at least FRX_HOST_GROUP_BYTES of samples -- the environment variable of that name overrides it); while group k is encoded, the rows of the
// groups behind it are still arriving over PCIe and the bytes of group k-1 are already leaving, each on its own stream.
// Every group is one frx_encode_tiles() call into its own slice of the device output (slices sized by the worst-case
// bound), so the result is byte-identical to a single call: the group streams are simply concatenated on the host side
// and the per-group offsets rebased.
#define FRX_HOST_GROUP_BYTES (32ull << 20)
#define FRX_HOST_ROW_BLOCK_BYTES (64ull << 20)   // H2D copies below ~32 MB lose 8 % of the PCIe rate to per-copy overhead (C2: 36.6 -> 33.6 ms)
int frx_encode_tiles_host(frx_ctx *c, const frx_raster *raster_host, const frx_tile *tiles, int32_t n_tiles,
                          const frx_encode_params *p, const frx_encode_out *out) {
    if (!c) return set_err(FRX_E_INVALID, "ctx is NULL");
    int rc = check_params(p); if (rc) return rc;
    rc = check_raster(raster_host); if (rc) return rc;
    if (!out || !out->bytes) return set_err(FRX_E_INVALID, "out->bytes is NULL");
    frx_raster r = normalise_raster(raster_host);
    rc = check_tiles(tiles, n_tiles, r.bands, r.height, r.width); if (rc) return rc;
    CK(cudaSetDevice(c->device));
    uint64_t total_frames = 0, max_bytes = 0;
    rc = frx_encode_bound(tiles, n_tiles, r.dtype, p, &total_frames, &max_bytes); if (rc) return rc;
    const size_t esz = (size_t)frx_dtype_size(r.dtype);
    for (int i = 0; i < 3; i++) if (!c->hs[i]) CK(cudaStreamCreateWithFlags(&c->hs[i], cudaStreamNonBlocking));
    cudaStream_t s_in = c->hs[0], s_cmp = c->hs[1], s_out = c->hs[2];
    // the band range and the row range the tiles touch: the device copy is the dense covering box
    long long r0 = r.height, r1 = 0, b0 = r.bands, b1 = 0;
    for (int i = 0; i < n_tiles; i++) {
        if (tiles[i].row_off < r0) r0 = tiles[i].row_off;
        if (tiles[i].row_off + tiles[i].h > r1) r1 = tiles[i].row_off + tiles[i].h;
        if (tiles[i].first_band < b0) b0 = tiles[i].first_band;
        if (tiles[i].first_band + tiles[i].n_channels > b1) b1 = tiles[i].first_band + tiles[i].n_channels;
    }
    const long long rows = r1 - r0, nb = b1 - b0;
    const size_t plane = (size_t)rows * (size_t)r.width;
    ENSURE(c->h_raster, plane * (size_t)nb * esz);
    // ---- tile groups ----
    struct Group { int t0, t1; long long gr0, gr1; uint64_t frames, bound, byte_base, frame_base; };
    std::vector<Group> groups;
    {
        uint64_t group_bytes = FRX_HOST_GROUP_BYTES;
        if (const char *e = getenv("FRX_HOST_GROUP_BYTES")) { const long long v = atoll(e); if (v > 0) group_bytes = (uint64_t)v; }
        int t0 = 0;
        while (t0 < n_tiles) {
            Group g;
            g.t0 = t0; g.gr0 = tiles[t0].row_off; g.gr1 = tiles[t0].row_off + tiles[t0].h;
            uint64_t bytes = 0;
            int t = t0;
            while (t < n_tiles) {
                const bool same_rows = tiles[t].row_off >= g.gr0 && tiles[t].row_off + tiles[t].h <= g.gr1;
                if (!same_rows && bytes >= group_bytes) break;
                if (tiles[t].row_off < g.gr0) g.gr0 = tiles[t].row_off;
                if (tiles[t].row_off + tiles[t].h > g.gr1) g.gr1 = tiles[t].row_off + tiles[t].h;
                bytes += (uint64_t)tiles[t].h * tiles[t].w * tiles[t].n_channels * esz;
                t++;
            }
            g.t1 = t;
            groups.push_back(g);
            t0 = t;
        }
        uint64_t bb = 0, fb = 0;
        for (Group &g : groups) {
            rc = frx_encode_bound(tiles + g.t0, g.t1 - g.t0, r.dtype, p, &g.frames, &g.bound); if (rc) return rc;
            g.byte_base = bb; g.frame_base = fb;
            bb += (g.bound + 15) & ~15ull;      // 16-byte aligned slices
            fb += g.frames;
        }
        max_bytes = bb;
    }
    const int ng = (int)groups.size();
    while ((int)c->hev.size() < 2 * ng) { cudaEvent_t e; CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); c->hev.push_back(e); }
    if (c->pin_small_cap < (size_t)ng * 16) {
        if (c->pin_small) cudaFreeHost(c->pin_small);
        c->pin_small = nullptr; c->pin_small_cap = 0;
        CK(cudaHostAlloc(&c->pin_small, (size_t)ng * 16 + 64, cudaHostAllocDefault));
        c->pin_small_cap = (size_t)ng * 16 + 64;
    }
    struct Tot { uint64_t total; int32_t status, pad; };
    Tot *tot = (Tot *)c->pin_small;
    // ---- device outputs ----
    ENSURE(c->h_bytes, max_bytes + 16);
    size_t small = sizeof(uint64_t) * 2 * (size_t)n_tiles + sizeof(double) * 2 * (size_t)n_tiles + sizeof(uint32_t) * (size_t)total_frames +
                   sizeof(uint64_t) * (size_t)total_frames + 16 * (size_t)ng + 64;
    ENSURE(c->h_small, small);
    char *sp = (char *)c->h_small.p;
    Tot *tot_dev = (Tot *)sp; sp += 16 * (size_t)ng;
    uint64_t *d_tile_offset = (uint64_t *)sp; sp += sizeof(uint64_t) * n_tiles;
    uint64_t *d_tile_size = (uint64_t *)sp; sp += sizeof(uint64_t) * n_tiles;
    double *d_tile_minmax = (double *)sp; sp += sizeof(double) * 2 * n_tiles;
    uint64_t *d_frame_offset = (uint64_t *)sp; sp += sizeof(uint64_t) * total_frames;
    uint32_t *d_frame_size = (uint32_t *)sp;
    // ---- copy-in: row blocks in the order the groups need them, all queued up front on their own stream ----
    {
        uint64_t row_block_bytes = FRX_HOST_ROW_BLOCK_BYTES;
        if (const char *e = getenv("FRX_HOST_ROW_BLOCK_BYTES")) { const long long v = atoll(e); if (v > 0) row_block_bytes = (uint64_t)v; }
        long long rb = (long long)(row_block_bytes / ((size_t)r.width * esz * (size_t)nb));
        if (rb < 1) rb = 1;
        const long long nblk = (rows + rb - 1) / rb;
        std::vector<char> have((size_t)nblk, 0);
        for (int k = 0; k < ng; k++) {
            const long long k0 = (groups[k].gr0 - r0) / rb, k1 = (groups[k].gr1 - r0 + rb - 1) / rb;
            for (long long blk = k0; blk < k1; blk++) {
                if (have[(size_t)blk]) continue;
                have[(size_t)blk] = 1;
                const long long y0 = blk * rb, y1 = (y0 + rb < rows) ? y0 + rb : rows;
                for (long long b = 0; b < nb; b++) {
                    const char *src = (const char *)r.data + ((size_t)(b0 + b) * (size_t)r.band_stride + (size_t)(r0 + y0) * (size_t)r.row_stride) * esz;
                    char *dst = (char *)c->h_raster.p + ((size_t)b * plane + (size_t)y0 * (size_t)r.width) * esz;
                    if (r.row_stride == r.width)      // contiguous rows: one linear copy (the 2-D path of the copy engine is slower)
                        CK(cudaMemcpyAsync(dst, src, (size_t)r.width * esz * (size_t)(y1 - y0), cudaMemcpyHostToDevice, s_in));
                    else
                        CK(cudaMemcpy2DAsync(dst, (size_t)r.width * esz, src, (size_t)r.row_stride * esz, (size_t)r.width * esz, (size_t)(y1 - y0),
                                             cudaMemcpyHostToDevice, s_in));
                }
            }
            CK(cudaEventRecord(c->hev[k], s_in));
        }
    }
    frx_raster rd = r;
    rd.data = c->h_raster.p; rd.bands = (int32_t)nb; rd.height = rows; rd.row_stride = r.width; rd.band_stride = (long long)plane;
    // ---- encode group k while group k-1 goes home ----
    std::vector<frx_tile> lt;
    std::vector<uint64_t> host_base((size_t)ng, 0);
    uint64_t host_off = 0;
    int failed = 0;
    auto copy_out_group = [&](int k) -> int {
        // the group's encode and its totals copy have completed (the caller synchronised the compute stream)
        const Group &g = groups[k];
        host_base[k] = host_off;
        if (tot[k].status) { failed = tot[k].status; host_off += tot[k].total; return FRX_OK; }
        if (host_off + tot[k].total > out->capacity) { failed = FRX_E_OVERFLOW; host_off += tot[k].total; return FRX_OK; }
        CK(cudaMemcpyAsync(out->bytes + host_off, (const uint8_t *)c->h_bytes.p + g.byte_base, tot[k].total, cudaMemcpyDeviceToHost, s_out));
        host_off += tot[k].total;
        return FRX_OK;
    };
    for (int k = 0; k < ng; k++) {
        const Group &g = groups[k];
        lt.assign(tiles + g.t0, tiles + g.t1);
        for (auto &t : lt) { t.row_off -= (int32_t)r0; t.first_band -= (int32_t)b0; }
        frx_encode_out od;
        memset(&od, 0, sizeof(od));
        od.bytes = (uint8_t *)c->h_bytes.p + g.byte_base; od.capacity = g.bound;
        od.total_bytes = &tot_dev[k].total; od.status = &tot_dev[k].status;
        od.tile_offset = d_tile_offset + g.t0; od.tile_size = d_tile_size + g.t0; od.tile_minmax = d_tile_minmax + 2 * (size_t)g.t0;
        od.frame_offset = d_frame_offset + g.frame_base; od.frame_size = d_frame_size + g.frame_base;
        // group k-1 goes home as soon as IT is done -- not when the rows of group k have arrived, which is what the stream
        // synchronisation inside frx_encode_tiles below waits for (it sits behind the wait on hev[k]): one group earlier,
        // so that only the last group's bytes are still to travel when the last rows have come in
        if (k > 0) { CK(cudaEventSynchronize(c->hev[ng + k - 1])); rc = copy_out_group(k - 1); if (rc) return rc; }
        CK(cudaStreamWaitEvent(s_cmp, c->hev[k], 0));
        rc = frx_encode_tiles(c, &rd, lt.data(), g.t1 - g.t0, p, &od, s_cmp); if (rc) return rc;
        CK(cudaMemcpyAsync(&tot[k], &tot_dev[k], sizeof(Tot), cudaMemcpyDeviceToHost, s_cmp));
        CK(cudaEventRecord(c->hev[ng + k], s_cmp));
    }
    CK(cudaStreamSynchronize(s_cmp));
    rc = copy_out_group(ng - 1); if (rc) return rc;
    if (out->total_bytes) *out->total_bytes = host_off;
    if (out->status) *out->status = failed;
    if (failed) {
        cudaStreamSynchronize(s_out);
        return set_err(failed, "output capacity %llu too small (need %llu)", (unsigned long long)out->capacity, (unsigned long long)host_off);
    }
    if (out->tile_offset) CK(cudaMemcpyAsync(out->tile_offset, d_tile_offset, sizeof(uint64_t) * n_tiles, cudaMemcpyDeviceToHost, s_out));
    if (out->tile_size) CK(cudaMemcpyAsync(out->tile_size, d_tile_size, sizeof(uint64_t) * n_tiles, cudaMemcpyDeviceToHost, s_out));
    if (out->tile_minmax) CK(cudaMemcpyAsync(out->tile_minmax, d_tile_minmax, sizeof(double) * 2 * n_tiles, cudaMemcpyDeviceToHost, s_out));
    if (out->frame_offset) CK(cudaMemcpyAsync(out->frame_offset, d_frame_offset, sizeof(uint64_t) * total_frames, cudaMemcpyDeviceToHost, s_out));
    if (out->frame_size) CK(cudaMemcpyAsync(out->frame_size, d_frame_size, sizeof(uint32_t) * total_frames, cudaMemcpyDeviceToHost, s_out));
    CK(cudaStreamSynchronize(s_out));
    // offsets inside a group start at 0: rebase them onto the concatenated stream
    for (int k = 1; k < ng; k++) {
        const Group &g = groups[k];
        if (out->tile_offset) for (int t = g.t0; t < g.t1; t++) out->tile_offset[t] += host_base[k];
        if (out->frame_offset) for (uint64_t f = g.frame_base; f < g.frame_base + g.frames; f++) out->frame_offset[f] += host_base[k];
    }
    return FRX_OK;
}

// Host buffers in, host buffers out, pipelined like the encode side: the tiles are cut into groups of consecutive tiles (at
// least FRX_HOST_GROUP_BYTES of compressed bytes, extended while the next tile's output block interleaves with the group's);
// while group k is decoded, the bytes of the groups behind it are still arriving and the rasters of group k-1 are already
// leaving, each on its own stream.  Output ranges of different groups may overlap (tiles of one raster): copies run in group
// order, so every element's last copy is made after the group that owns it has finished.
int frx_decode_tiles_host(frx_ctx *c, const uint8_t *bytes_host, uint64_t n_bytes, const frx_dtile *tiles, int32_t n_tiles,
                          const uint32_t *frame_size_host, int32_t blocksize, int32_t out_dtype, int32_t pcm_mode, void *out_host,
                          uint64_t out_elems, int32_t *tile_status_host) {
    if (!c) return set_err(FRX_E_INVALID, "ctx is NULL");
    if (!bytes_host || !out_host || !tiles || n_tiles <= 0) return set_err(FRX_E_INVALID, "NULL argument");
    if (out_dtype < FRX_U8 || out_dtype > FRX_F32) return set_err(FRX_E_INVALID, "bad out_dtype %d", out_dtype);
    CK(cudaSetDevice(c->device));
    for (int i = 0; i < 3; i++) if (!c->hs[i]) CK(cudaStreamCreateWithFlags(&c->hs[i], cudaStreamNonBlocking));
    cudaStream_t s_in = c->hs[0], s_cmp = c->hs[1], s_out = c->hs[2];
    const size_t esz = (size_t)frx_dtype_size(out_dtype);
    uint64_t nfs = 0;
    for (int i = 0; i < n_tiles; i++) {
        const frx_dtile &t = tiles[i];
        if (t.first_frame != 0xFFFFFFFFu && (uint64_t)t.first_frame + t.n_frames > nfs) nfs = (uint64_t)t.first_frame + t.n_frames;
        if (t.byte_offset > n_bytes || t.byte_size > n_bytes - t.byte_offset) return set_err(FRX_E_INVALID, "tile %d: byte range outside the buffer", i);
        if (t.channels < 1 || t.h <= 0 || t.w <= 0 || t.out_band_stride < 0 || t.out_row_stride < 0) return set_err(FRX_E_INVALID, "tile %d: bad geometry", i);
    }
    if (nfs && !frame_size_host) return set_err(FRX_E_INVALID, "frame_size_host is NULL");
    ENSURE(c->h_bytes, n_bytes + 16);
    ENSURE(c->h_out, out_elems * esz + 16);
    const size_t st_bytes = ((size_t)n_tiles * 4 + 63) & ~(size_t)63;
    ENSURE(c->h_small, st_bytes + sizeof(uint32_t) * (size_t)nfs + 64);
    int32_t *status_dev = (int32_t *)c->h_small.p;
    uint32_t *fs_dev = nfs ? (uint32_t *)((char *)c->h_small.p + st_bytes) : nullptr;
    // ---- groups ----
    struct Group { int t0, t1; uint64_t b0, b1, o0, o1; };
    auto out_last = [&](const frx_dtile &t) {
        return t.out_offset + (uint64_t)(t.channels - 1) * (uint64_t)t.out_band_stride + (uint64_t)(t.h - 1) * (uint64_t)t.out_row_stride + (uint64_t)t.w;
    };
    std::vector<Group> groups;
    {
        // a decode launch costs a frame time (~3 ms: a frame is decoded by one lane) however few frames it holds, so the groups are
        // large: a handful per call is enough to overlap the copies, which are what bounds this entry point
        uint64_t group_bytes = 192ull << 20;
        if (const char *e = getenv("FRX_HOST_GROUP_BYTES")) { const long long v = atoll(e); if (v > 0) group_bytes = (uint64_t)v; }
        int t0 = 0;
        while (t0 < n_tiles) {
            Group g = { t0, t0, tiles[t0].byte_offset, tiles[t0].byte_offset + tiles[t0].byte_size, tiles[t0].out_offset, out_last(tiles[t0]) };
            uint64_t bytes = 0;
            int t = t0;
            while (t < n_tiles) {
                const uint64_t o0 = tiles[t].out_offset, o1 = out_last(tiles[t]);
                const bool interleaved = o0 < g.o1 && o1 > g.o0;
                if (t > t0 && bytes >= group_bytes && !interleaved) break;
                if (tiles[t].byte_offset < g.b0) g.b0 = tiles[t].byte_offset;
                if (tiles[t].byte_offset + tiles[t].byte_size > g.b1) g.b1 = tiles[t].byte_offset + tiles[t].byte_size;
                if (o0 < g.o0) g.o0 = o0;
                if (o1 > g.o1) g.o1 = o1;
                bytes += tiles[t].byte_size;
                t++;
            }
            g.t1 = t;
            if (g.o1 > out_elems) return set_err(FRX_E_INVALID, "tile output blocks end at element %llu, beyond the %llu the output holds",
                                                  (unsigned long long)g.o1, (unsigned long long)out_elems);
            groups.push_back(g);
            t0 = t;
        }
    }
    const int ng = (int)groups.size();
    while ((int)c->hev.size() < 2 * ng) { cudaEvent_t e; CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); c->hev.push_back(e); }
    if (nfs) CK(cudaMemcpyAsync(fs_dev, frame_size_host, sizeof(uint32_t) * (size_t)nfs, cudaMemcpyHostToDevice, s_in));
    for (int k = 0; k < ng; k++) {      // copy-in of every group's byte range, queued up front on its own stream
        const Group &g = groups[k];
        CK(cudaMemcpyAsync((uint8_t *)c->h_bytes.p + g.b0, bytes_host + g.b0, g.b1 - g.b0, cudaMemcpyHostToDevice, s_in));
        CK(cudaEventRecord(c->hev[k], s_in));
    }
    for (int k = 0; k < ng; k++) {
        const Group &g = groups[k];
        CK(cudaStreamWaitEvent(s_cmp, c->hev[k], 0));
        // (decode_core synchronises its stream once while staging the tile tables: group k-1 is finished after it)
        int rc = decode_core(c, (const uint8_t *)c->h_bytes.p, n_bytes, tiles + g.t0, g.t1 - g.t0, fs_dev, blocksize, out_dtype, pcm_mode, c->h_out.p,
                             status_dev + g.t0, s_cmp, out_elems);
        if (rc) { cudaStreamSynchronize(s_in); cudaStreamSynchronize(s_out); return rc; }
        CK(cudaEventRecord(c->hev[ng + k], s_cmp));
        CK(cudaStreamWaitEvent(s_out, c->hev[ng + k], 0));
        CK(cudaMemcpyAsync((char *)out_host + g.o0 * esz, (const char *)c->h_out.p + g.o0 * esz, (g.o1 - g.o0) * esz, cudaMemcpyDeviceToHost, s_out));
    }
    std::vector<int32_t> status(n_tiles);
    CK(cudaMemcpyAsync(status.data(), status_dev, sizeof(int32_t) * n_tiles, cudaMemcpyDeviceToHost, s_out));
    CK(cudaStreamSynchronize(s_out));
    CK(cudaStreamSynchronize(s_in));
    int worst = 0;
    for (int i = 0; i < n_tiles; i++) {
        if (tile_status_host) tile_status_host[i] = status[i];
        if (status[i] && !worst) worst = status[i];
    }
    if (worst) return set_err(worst, "decode failed for at least one tile (status %d)", worst);
    return FRX_OK;
}

}  // extern "C"

// ---- allocation-free decode (include/frx.h: frx_decode_tiles_noalloc) -----------------------------------------
// The scratch the decode launch sequence needs, laid out in one file-scope device array.  Every GPU that loads libfrx
// holds its own instance until the process exits.  In host code the array's NAME is only a host-side stand-in: the
// device address comes from cudaGetSymbolAddress, taken once per device with that device current (noalloc_setup).
struct NoallocStage {                                   // host-built tables, copied to the head of the scratch in one go
    frx_dtile tiles[FRX_NOALLOC_MAX_TILES];
    uint32_t dprefix[FRX_NOALLOC_MAX_TILES + 1];
    uint32_t scan_prefix[FRX_NOALLOC_MAX_TILES + 1];
    int32_t scan_tiles[FRX_NOALLOC_MAX_TILES + 1];
};
#define NA_ALIGN(x) (((x) + 255) & ~(size_t)255)
#define NA_OFF_STAGE 0
#define NA_OFF_CRCTAB NA_ALIGN(sizeof(NoallocStage))
#define NA_XPOW_ELEMS (FRX_CRC_MAXLQ * 256 + 4)
#define NA_OFF_XPOW (NA_OFF_CRCTAB + NA_ALIGN(1024 * sizeof(uint16_t)))
#define NA_OFF_FPOS (NA_OFF_XPOW + NA_ALIGN(NA_XPOW_ELEMS * sizeof(uint16_t)))
#define NA_OFF_FLEN (NA_OFF_FPOS + NA_ALIGN(sizeof(uint64_t) * (FRX_NOALLOC_MAX_FRAMES + 1)))
#define NA_SCRATCH_BYTES (NA_OFF_FLEN + NA_ALIGN(sizeof(uint32_t) * (FRX_NOALLOC_MAX_FRAMES + 1)))
static __device__ __align__(256) unsigned char g_frx_noalloc_scratch[NA_SCRATCH_BYTES];

#define NA_MAX_DEVICES 64
static unsigned char *g_na_ptr[NA_MAX_DEVICES];         // device address of the array on each device, 0 until set up
static pthread_mutex_t g_na_lock = PTHREAD_MUTEX_INITIALIZER;

// once per device, with that device current: the array's device address, and the constant tables into it
static int noalloc_setup(int device, unsigned char **out) {
    if (device < 0 || device >= NA_MAX_DEVICES) return set_err(FRX_E_UNSUPPORTED, "device %d beyond the %d the allocation-free path tracks", device, NA_MAX_DEVICES);
    pthread_mutex_lock(&g_na_lock);
    unsigned char *p = g_na_ptr[device];
    int rc = FRX_OK;
    if (!p) {
        cudaError_t e = cudaGetSymbolAddress((void **)&p, g_frx_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            rc = set_err(FRX_E_CUDA, "cudaGetSymbolAddress(g_frx_noalloc_scratch) failed on device %d: %s", device, cudaGetErrorString(e));
            p = nullptr;
        } else {
            static uint16_t tab[1024], xp[NA_XPOW_ELEMS];   // host tables (static: 27 KB is too much for a caller's stack)
            frx_make_crc16_tables(tab);
            for (int lq = 1; lq <= FRX_CRC_MAXLQ; lq++) {   // x^(128 lq j) mod P, as ensure_xpow builds them
                uint32_t base = 2, ex = 128u * lq, acc = 1;
                while (ex) { if (ex & 1u) acc = frx_gf16_mulmod(acc, base); base = frx_gf16_mulmod(base, base); ex >>= 1; }
                uint32_t cur = 1;
                for (int j = 0; j < 256; j++) { xp[(lq - 1) * 256 + j] = (uint16_t)cur; cur = frx_gf16_mulmod(cur, acc); }
            }
            uint32_t inv32 = 0xC002u;
            for (int i = 0; i < 5; i++) inv32 = frx_gf16_mulmod(inv32, inv32);
            uint32_t cur = 1;
            for (int z = 0; z < 4; z++) { xp[FRX_CRC_MAXLQ * 256 + z] = (uint16_t)cur; cur = frx_gf16_mulmod(cur, inv32); }
            e = cudaMemcpy(p + NA_OFF_CRCTAB, tab, sizeof(tab), cudaMemcpyHostToDevice);
            if (e == cudaSuccess) e = cudaMemcpy(p + NA_OFF_XPOW, xp, sizeof(xp), cudaMemcpyHostToDevice);
            if (e != cudaSuccess) { rc = set_err(FRX_E_CUDA, "copying the CRC tables into the static scratch failed: %s", cudaGetErrorString(e)); p = nullptr; }
            else g_na_ptr[device] = p;
        }
    }
    pthread_mutex_unlock(&g_na_lock);
    *out = p;
    return rc;
}

extern "C" int frx_decode_tiles_noalloc(frx_ctx *c, const uint8_t *bytes_dev, uint64_t n_bytes, const frx_dtile *tiles, int32_t n_tiles,
                                        const uint32_t *frame_size_dev, int32_t blocksize, int32_t out_dtype, int32_t pcm_mode,
                                        void *out_dev, int32_t *tile_status_dev, void *stream) {
    if (!c) return set_err(FRX_E_INVALID, "ctx is NULL");
    if (!bytes_dev || !tiles || n_tiles <= 0 || !out_dev || !tile_status_dev) return set_err(FRX_E_INVALID, "NULL argument");
    if (blocksize < 16 || blocksize > 65535) return set_err(FRX_E_INVALID, "bad blocksize %d", blocksize);
    if (out_dtype < FRX_U8 || out_dtype > FRX_F32) return set_err(FRX_E_INVALID, "bad out_dtype %d", out_dtype);
    if (n_tiles > FRX_NOALLOC_MAX_TILES)
        return set_err(FRX_E_UNSUPPORTED, "%d tiles in one call; the allocation-free decode holds scratch for at most FRX_NOALLOC_MAX_TILES = %d (use frx_decode_tiles or split the batch)",
                       n_tiles, FRX_NOALLOC_MAX_TILES);
    NoallocStage hs;                                    // 11.8 KB on the stack: no heap either
    uint64_t nf = 0, nu = 0;
    int n_scan = 0;
    for (int i = 0; i < n_tiles; i++) {
        const frx_dtile &t = tiles[i];
        if (t.channels < 1 || t.channels > FRX_MAX_CH || t.h <= 0 || t.w <= 0 || (t.bps != 16 && t.bps != 32 && t.bps != 8 && t.bps != 24 && t.bps != 12 && t.bps != 20))
            return set_err(FRX_E_INVALID, "tile %d: bad geometry/bps", i);
        if (t.channels == 2)
            return set_err(FRX_E_UNSUPPORTED, "tile %d has two channels: stereo-decorrelated frames need a scratch row per frame, which the fixed scratch of the allocation-free decode does not hold (use frx_decode_tiles)", i);
        if (t.byte_offset > n_bytes || t.byte_size > n_bytes - t.byte_offset) return set_err(FRX_E_INVALID, "tile %d: byte range outside the buffer", i);
        const uint64_t expect = ((uint64_t)t.h * t.w + blocksize - 1) / blocksize;
        if (t.n_frames != expect) return set_err(FRX_E_INVALID, "tile %d: n_frames %u != ceil(h*w/blocksize) %llu", i, t.n_frames, (unsigned long long)expect);
        if (t.first_frame != 0xFFFFFFFFu && !frame_size_dev) return set_err(FRX_E_INVALID, "tile %d needs frame_size", i);
        hs.tiles[i] = t;
        hs.dprefix[i] = (uint32_t)nf;
        nf += t.n_frames;
        if (nf > FRX_NOALLOC_MAX_FRAMES)
            return set_err(FRX_E_UNSUPPORTED, "the tiles hold more than FRX_NOALLOC_MAX_FRAMES = %d frames (reached at tile %d); the allocation-free decode has no scratch for more (use frx_decode_tiles or split the batch)",
                           FRX_NOALLOC_MAX_FRAMES, i);
        if (t.first_frame == 0xFFFFFFFFu) {
            hs.scan_prefix[n_scan] = (uint32_t)nu;
            hs.scan_tiles[n_scan] = i;
            n_scan++;
            nu += (t.byte_size + 4095) / 4096;
            if (nu > 0x7FFFFFF0ull) return set_err(FRX_E_UNSUPPORTED, "too many bytes to scan for frame starts in one call");
        }
    }
    hs.dprefix[n_tiles] = (uint32_t)nf;
    hs.scan_prefix[n_scan] = (uint32_t)nu;
    CK(cudaSetDevice(c->device));
    unsigned char *base = nullptr;
    { int rc = noalloc_setup(c->device, &base); if (rc) return rc; }
    cudaStream_t st = (cudaStream_t)stream;
    // pageable source: the runtime copies `hs` to its own staging area before this returns, so the stack copy may die
    CK(cudaMemcpyAsync(base + NA_OFF_STAGE, &hs, sizeof(hs), cudaMemcpyHostToDevice, st));
    // what must start at zero: the tile status, and the frame lengths the CRC pass reads (the array keeps the last call's)
    CK(cudaMemsetAsync(tile_status_dev, 0, sizeof(int32_t) * n_tiles, st));
    CK(cudaMemsetAsync(base + NA_OFF_FLEN, 0, sizeof(uint32_t) * (size_t)(nf + 1), st));
    NoallocStage *ds = (NoallocStage *)(base + NA_OFF_STAGE);
    DecArgs a;
    memset(&a, 0, sizeof(a));
    a.bytes = bytes_dev; a.n_bytes = n_bytes; a.tiles = ds->tiles; a.dprefix = ds->dprefix;
    a.frame_pos = (uint64_t *)(base + NA_OFF_FPOS); a.frame_len = (uint32_t *)(base + NA_OFF_FLEN); a.stereo_scratch = nullptr;
    a.frame_size = frame_size_dev; a.crc_tab = (const uint16_t *)(base + NA_OFF_CRCTAB);
    a.out = out_dev; a.tile_status = tile_status_dev; a.total_frames = (uint32_t)nf; a.n_tiles = n_tiles; a.blocksize = blocksize;
    a.pcm_mode = pcm_mode; a.verify_crc = c->verify_crc;
    CK(frx_launch_frame_pos(a, st));
    c->launches += 1;
    if (n_scan) {
        CK(frx_launch_discover(a, ds->scan_prefix, ds->scan_tiles, n_scan, (uint32_t)nu, st));
        c->launches += 1;
    }
    CK(cudaEventRecord(c->ev[2], st));
    CK(frx_launch_decode(a, out_dtype, c->sm_count, st));
    if (c->verify_crc) { CK(frx_launch_crc_check(a, (const uint16_t *)(base + NA_OFF_XPOW), c->sm_count, st)); c->launches += 1; }
    CK(cudaEventRecord(c->ev[3], st));
    c->dec_timed = true;
    c->launches += 1;
    return FRX_OK;
}

/*
 * flac_oracle.c -- CPU ORACLE (test infrastructure, NOT the product).
 *
 * A plain-C restatement of the codec arithmetic that flac-raster reaches through
 * pyFLAC 3.0.0 -> libFLAC 1.4.3 (third-party; neither is vendored under
 * /root/reference and neither is installable here):
 *
 *   - ofl_decode():  an independent RFC 9639 FLAC stream decoder.  It follows the
 *     bitstream description restated in SURVEY.md Appendix A and replaces
 *     pyflac.FileDecoder's libFLAC hop (reference call sites converter.py:241-242,
 *     cli.py:170-173; pyFLAC dump docs/sonos-pyflac.txt:1584-1640, 1809-1854).
 *   - ofl_encode():  a restatement of libFLAC's compression-level search
 *     (preset table docs/sonos-pyflac.txt:6926-6934, limits :3491-3546) as reached
 *     by pyflac.StreamEncoder.process()/finish() (converter.py:201-216,
 *     spatial_encoder.py:278-291; dump:1968-2014, 2175-2213).  libFLAC's .c files
 *     are NOT in the reference tree; this follows its published algorithm
 *     (tukey(0.5) window, autocorrelation, Levinson-Durbin, best-order estimate,
 *     coefficient quantisation with error feedback, fixed-order 0-4 trial, Rice
 *     partition-order search by partition sums) and is PINNED against the
 *     libFLAC-1.4.3-produced fixtures test_data/sample_rgb.flac and
 *     test_data/sample_dem.flac by tests/test_oracle_golden.py.
 *   - ofl_ref_encode_tiles() / ofl_ref_decode_tiles(): the reference's tile loop
 *     (cli.py:691-763 create-streaming body; cli.py:1016-1023 + converter.py:234-282
 *     extract-streaming body) restated end to end in C with an OpenMP loop over
 *     tiles, used ONLY as the timed CPU baseline of bench.py and as the checker in
 *     tests/.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this library.  The product (flac_raster_b200) never does.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#ifndef M_LN2
#define M_LN2 0.69314718055994530942
#endif
#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

#define OFL_MAX_CHANNELS 8
#define OFL_MAX_LPC_ORDER 32
#define OFL_MAX_BLOCKSIZE 65535

/* ------------------------------------------------------------------------- */
/* CRC-8 (poly 0x07) and CRC-16 (poly 0x8005), MSB first, init 0.              */
/* ------------------------------------------------------------------------- */
static uint8_t crc8_tab[256];
static uint16_t crc16_tab[256];
static int crc_ready = 0;

static void crc_init(void) {
    if (crc_ready) return;
    for (int i = 0; i < 256; i++) {
        uint8_t c = (uint8_t)i;
        for (int b = 0; b < 8; b++) c = (uint8_t)((c & 0x80) ? ((c << 1) ^ 0x07) : (c << 1));
        crc8_tab[i] = c;
        uint16_t d = (uint16_t)(i << 8);
        for (int b = 0; b < 8; b++) d = (uint16_t)((d & 0x8000) ? ((d << 1) ^ 0x8005) : (d << 1));
        crc16_tab[i] = d;
    }
    crc_ready = 1;
}
static uint8_t crc8(const uint8_t *p, size_t n) {
    uint8_t c = 0;
    for (size_t i = 0; i < n; i++) c = crc8_tab[c ^ p[i]];
    return c;
}
static uint16_t crc16(const uint8_t *p, size_t n) {
    uint16_t c = 0;
    for (size_t i = 0; i < n; i++) c = (uint16_t)((c << 8) ^ crc16_tab[(c >> 8) ^ p[i]]);
    return c;
}
uint32_t ofl_crc8(const uint8_t *p, size_t n) { crc_init(); return crc8(p, n); }
uint32_t ofl_crc16(const uint8_t *p, size_t n) { crc_init(); return crc16(p, n); }

/* ------------------------------------------------------------------------- */
/* Bit reader (MSB first)                                                      */
/* ------------------------------------------------------------------------- */
typedef struct {
    const uint8_t *p;
    size_t nbytes;
    size_t bitpos;
    int err;
} br_t;

static inline uint32_t br_bit(br_t *b) {
    if ((b->bitpos >> 3) >= b->nbytes) { b->err = 1; return 0; }
    uint32_t v = (b->p[b->bitpos >> 3] >> (7 - (b->bitpos & 7))) & 1u;
    b->bitpos++;
    return v;
}
static inline uint64_t br_bits(br_t *b, int n) { /* n in 0..64 */
    uint64_t v = 0;
    while (n > 0) {
        size_t byte = b->bitpos >> 3;
        if (byte >= b->nbytes) { b->err = 1; return 0; }
        int avail = 8 - (int)(b->bitpos & 7);
        int take = n < avail ? n : avail;
        uint32_t cur = (b->p[byte] >> (avail - take)) & ((1u << take) - 1u);
        v = (v << take) | cur;
        b->bitpos += (size_t)take;
        n -= take;
    }
    return v;
}
static inline int64_t br_sbits(br_t *b, int n) {
    if (n == 0) return 0;
    uint64_t v = br_bits(b, n);
    if (n < 64 && (v >> (n - 1)) & 1) v |= ~((((uint64_t)1) << n) - 1);
    return (int64_t)v;
}
static inline uint32_t br_unary(br_t *b) { /* count zeros until a one */
    uint32_t q = 0;
    for (;;) {
        size_t byte = b->bitpos >> 3;
        if (byte >= b->nbytes) { b->err = 1; return q; }
        int off = (int)(b->bitpos & 7);
        uint32_t cur = (uint32_t)(b->p[byte] << off) & 0xFFu;
        if (cur) {
            int lz = __builtin_clz(cur) - 24;
            q += (uint32_t)lz;
            b->bitpos += (size_t)lz + 1;
            return q;
        }
        q += (uint32_t)(8 - off);
        b->bitpos += (size_t)(8 - off);
    }
}

/* ------------------------------------------------------------------------- */
/* Public structs (mirrored by ctypes in oracle/__init__.py)                   */
/* ------------------------------------------------------------------------- */
typedef struct {
    uint32_t min_blocksize, max_blocksize, min_framesize, max_framesize;
    uint32_t sample_rate, channels, bps, pad;
    uint64_t total_samples;
    uint8_t md5[16];
} ofl_streaminfo;

typedef struct {
    uint64_t byte_offset; /* from start of buffer */
    uint64_t number;      /* coded frame/sample number */
    uint32_t byte_size;
    uint32_t blocksize;
    uint32_t sample_rate;
    uint8_t channels, bps, chan_assign, crc8_ok, crc16_ok, pad[3];
    uint8_t sf_type[OFL_MAX_CHANNELS];   /* 0 const, 1 verbatim, 2 fixed, 3 lpc */
    uint8_t sf_order[OFL_MAX_CHANNELS];
    uint8_t sf_wasted[OFL_MAX_CHANNELS];
    uint8_t sf_part_order[OFL_MAX_CHANNELS];
    uint8_t sf_rice_method[OFL_MAX_CHANNELS];
    uint8_t sf_precision[OFL_MAX_CHANNELS];
    int8_t sf_shift[OFL_MAX_CHANNELS];
    uint8_t sf_escapes[OFL_MAX_CHANNELS];
} ofl_frameinfo;

/* ------------------------------------------------------------------------- */
/* Decoder                                                                     */
/* ------------------------------------------------------------------------- */
static int decode_residual(br_t *b, int64_t *out, uint32_t blocksize, uint32_t pred_order,
                           ofl_frameinfo *fi, int ch) {
    uint32_t method = (uint32_t)br_bits(b, 2);
    if (method > 1) return -1;
    uint32_t porder = (uint32_t)br_bits(b, 4);
    uint32_t nparts = 1u << porder;
    int pbits = method ? 5 : 4;
    uint32_t esc = method ? 31u : 15u;
    if (fi) { fi->sf_part_order[ch] = (uint8_t)porder; fi->sf_rice_method[ch] = (uint8_t)method; }
    if ((blocksize >> porder) << porder != blocksize && porder > 0) return -1;
    uint32_t i = pred_order;
    for (uint32_t p = 0; p < nparts; p++) {
        uint32_t n = blocksize >> porder;
        if (p == 0) {
            if (n < pred_order) return -1;
            n -= pred_order;
        }
        uint32_t k = (uint32_t)br_bits(b, pbits);
        if (k == esc) {
            uint32_t raw = (uint32_t)br_bits(b, 5);
            if (fi && fi->sf_escapes[ch] < 255) fi->sf_escapes[ch]++;
            for (uint32_t j = 0; j < n; j++) out[i++] = br_sbits(b, (int)raw);
        } else {
            for (uint32_t j = 0; j < n; j++) {
                uint32_t q = br_unary(b);
                uint64_t u = ((uint64_t)q << k) | br_bits(b, (int)k);
                out[i++] = (int64_t)(u >> 1) ^ -(int64_t)(u & 1);
            }
        }
        if (b->err) return -1;
    }
    return 0;
}

static int decode_subframe(br_t *b, int64_t *out, uint32_t blocksize, int bps,
                           ofl_frameinfo *fi, int ch) {
    if (br_bit(b)) return -1; /* zero pad */
    uint32_t type = (uint32_t)br_bits(b, 6);
    uint32_t wasted = 0;
    if (br_bit(b)) { wasted = br_unary(b) + 1; }
    if ((int)wasted >= bps) return -1;
    bps -= (int)wasted;
    if (fi) { fi->sf_wasted[ch] = (uint8_t)wasted; fi->sf_precision[ch] = 0; fi->sf_shift[ch] = 0;
              fi->sf_part_order[ch] = 0; fi->sf_rice_method[ch] = 0; fi->sf_escapes[ch] = 0; }
    if (type == 0) { /* CONSTANT */
        int64_t v = br_sbits(b, bps);
        for (uint32_t i = 0; i < blocksize; i++) out[i] = v;
        if (fi) { fi->sf_type[ch] = 0; fi->sf_order[ch] = 0; }
    } else if (type == 1) { /* VERBATIM */
        for (uint32_t i = 0; i < blocksize; i++) out[i] = br_sbits(b, bps);
        if (fi) { fi->sf_type[ch] = 1; fi->sf_order[ch] = 0; }
    } else if (type >= 8 && type <= 12) { /* FIXED */
        uint32_t order = type - 8;
        if (order > blocksize) return -1;
        if (fi) { fi->sf_type[ch] = 2; fi->sf_order[ch] = (uint8_t)order; }
        for (uint32_t i = 0; i < order; i++) out[i] = br_sbits(b, bps);
        if (decode_residual(b, out, blocksize, order, fi, ch)) return -1;
        for (uint32_t i = order; i < blocksize; i++) {
            int64_t pred = 0;
            switch (order) {
                case 0: pred = 0; break;
                case 1: pred = out[i - 1]; break;
                case 2: pred = 2 * out[i - 1] - out[i - 2]; break;
                case 3: pred = 3 * out[i - 1] - 3 * out[i - 2] + out[i - 3]; break;
                case 4: pred = 4 * out[i - 1] - 6 * out[i - 2] + 4 * out[i - 3] - out[i - 4]; break;
            }
            out[i] += pred;
        }
    } else if (type >= 32) { /* LPC */
        uint32_t order = (type & 31u) + 1;
        if (order > blocksize) return -1;
        if (fi) { fi->sf_type[ch] = 3; fi->sf_order[ch] = (uint8_t)order; }
        for (uint32_t i = 0; i < order; i++) out[i] = br_sbits(b, bps);
        uint32_t prec = (uint32_t)br_bits(b, 4) + 1;
        if (prec == 16) return -1;
        int shift = (int)br_sbits(b, 5);
        if (shift < 0) return -1;
        int32_t coef[OFL_MAX_LPC_ORDER];
        for (uint32_t j = 0; j < order; j++) coef[j] = (int32_t)br_sbits(b, (int)prec);
        if (fi) { fi->sf_precision[ch] = (uint8_t)prec; fi->sf_shift[ch] = (int8_t)shift; }
        if (decode_residual(b, out, blocksize, order, fi, ch)) return -1;
        for (uint32_t i = order; i < blocksize; i++) {
            int64_t sum = 0;
            for (uint32_t j = 0; j < order; j++) sum += (int64_t)coef[j] * out[i - 1 - j];
            out[i] += sum >> shift;
        }
    } else {
        return -1;
    }
    if (b->err) return -1;
    if (wasted) for (uint32_t i = 0; i < blocksize; i++) out[i] = (int64_t)((uint64_t)out[i] << wasted);
    return 0;
}

/* Decode ONE frame starting at p.  si supplies defaults.  Returns frame size in bytes,
 * or <0 on error.  out: channels buffers of int64 [blocksize].                        */
static long decode_frame(const uint8_t *p, size_t n, const ofl_streaminfo *si, int64_t **chbuf,
                         ofl_frameinfo *fi) {
    crc_init();
    if (n < 6) return -1;
    if (p[0] != 0xFF || (p[1] & 0xFE) != 0xF8) return -1;
    br_t b = { p, n, 0, 0 };
    br_bits(&b, 14);
    if (br_bit(&b)) return -1;
    uint32_t variable = br_bit(&b);
    uint32_t bs_code = (uint32_t)br_bits(&b, 4);
    uint32_t sr_code = (uint32_t)br_bits(&b, 4);
    uint32_t ch_code = (uint32_t)br_bits(&b, 4);
    uint32_t sz_code = (uint32_t)br_bits(&b, 3);
    if (br_bit(&b)) return -1;
    /* UTF-8 style number */
    uint64_t number;
    {
        uint32_t x = (uint32_t)br_bits(&b, 8);
        int extra;
        if (!(x & 0x80)) { number = x; extra = 0; }
        else if ((x & 0xE0) == 0xC0) { number = x & 0x1F; extra = 1; }
        else if ((x & 0xF0) == 0xE0) { number = x & 0x0F; extra = 2; }
        else if ((x & 0xF8) == 0xF0) { number = x & 0x07; extra = 3; }
        else if ((x & 0xFC) == 0xF8) { number = x & 0x03; extra = 4; }
        else if ((x & 0xFE) == 0xFC) { number = x & 0x01; extra = 5; }
        else if (x == 0xFE) { number = 0; extra = 6; }
        else return -1;
        for (int i = 0; i < extra; i++) {
            uint32_t y = (uint32_t)br_bits(&b, 8);
            if ((y & 0xC0) != 0x80) return -1;
            number = (number << 6) | (y & 0x3F);
        }
    }
    (void)variable;
    uint32_t blocksize;
    if (bs_code == 0) return -1;
    else if (bs_code == 1) blocksize = 192;
    else if (bs_code <= 5) blocksize = 576u << (bs_code - 2);
    else if (bs_code == 6) blocksize = (uint32_t)br_bits(&b, 8) + 1;
    else if (bs_code == 7) blocksize = (uint32_t)br_bits(&b, 16) + 1;
    else blocksize = 256u << (bs_code - 8);
    uint32_t sample_rate;
    static const uint32_t sr_tab[12] = { 0, 88200, 176400, 192000, 8000, 16000, 22050, 24000, 32000, 44100, 48000, 96000 };
    if (sr_code == 0) sample_rate = si ? si->sample_rate : 0;
    else if (sr_code < 12) sample_rate = sr_tab[sr_code];
    else if (sr_code == 12) sample_rate = (uint32_t)br_bits(&b, 8) * 1000u;
    else if (sr_code == 13) sample_rate = (uint32_t)br_bits(&b, 16);
    else if (sr_code == 14) sample_rate = (uint32_t)br_bits(&b, 16) * 10u;
    else return -1;
    int bps;
    static const int sz_tab[8] = { 0, 8, 12, -1, 16, 20, 24, 32 };
    if (sz_code == 0) bps = si ? (int)si->bps : 0; else bps = sz_tab[sz_code];
    if (bps <= 0) return -1;
    uint32_t channels;
    if (ch_code < 8) channels = ch_code + 1;
    else if (ch_code <= 10) channels = 2;
    else return -1;
    if (b.err) return -1;
    size_t hdr_bytes = b.bitpos >> 3;
    if (hdr_bytes + 1 > n) return -1;
    uint8_t c8 = crc8(p, hdr_bytes);
    uint8_t got8 = (uint8_t)br_bits(&b, 8);
    if (fi) {
        memset(fi, 0, sizeof(*fi));
        fi->number = number; fi->blocksize = blocksize; fi->sample_rate = sample_rate;
        fi->channels = (uint8_t)channels; fi->bps = (uint8_t)bps; fi->chan_assign = (uint8_t)ch_code;
        fi->crc8_ok = (uint8_t)(c8 == got8);
    }
    if (c8 != got8) return -2;
    for (uint32_t c = 0; c < channels; c++) {
        int sbps = bps;
        if ((ch_code == 8 && c == 1) || (ch_code == 9 && c == 0) || (ch_code == 10 && c == 1)) sbps++;
        if (decode_subframe(&b, chbuf[c], blocksize, sbps, fi, (int)c)) return -3;
    }
    /* byte align */
    if (b.bitpos & 7) {
        uint32_t padbits = 8 - (uint32_t)(b.bitpos & 7);
        if (br_bits(&b, (int)padbits) != 0) return -4;
    }
    size_t body = b.bitpos >> 3;
    if (body + 2 > n) return -1;
    uint16_t c16 = crc16(p, body);
    uint16_t got16 = (uint16_t)br_bits(&b, 16);
    if (fi) fi->crc16_ok = (uint8_t)(c16 == got16);
    if (c16 != got16) return -5;
    /* stereo undo */
    if (ch_code == 8) {
        for (uint32_t i = 0; i < blocksize; i++) chbuf[1][i] = chbuf[0][i] - chbuf[1][i];
    } else if (ch_code == 9) {
        for (uint32_t i = 0; i < blocksize; i++) chbuf[0][i] = chbuf[0][i] + chbuf[1][i];
    } else if (ch_code == 10) {
        for (uint32_t i = 0; i < blocksize; i++) {
            int64_t m = chbuf[0][i], s = chbuf[1][i];
            m = (int64_t)(((uint64_t)m << 1) | (uint64_t)(s & 1));
            chbuf[0][i] = (m + s) >> 1;
            chbuf[1][i] = (m - s) >> 1;
        }
    }
    if (fi) fi->byte_size = (uint32_t)(body + 2);
    return (long)(body + 2);
}

/* Parse "fLaC" + metadata blocks.  Returns offset of first frame or <0.
 * vc_off/vc_len: location of the VORBIS_COMMENT payload (0 if none).        */
long ofl_parse_header(const uint8_t *buf, size_t len, ofl_streaminfo *si, size_t *vc_off, size_t *vc_len) {
    if (len < 42 || memcmp(buf, "fLaC", 4) != 0) return -1;
    size_t pos = 4;
    int last = 0, first = 1;
    if (vc_off) *vc_off = 0;
    if (vc_len) *vc_len = 0;
    while (!last) {
        if (pos + 4 > len) return -1;
        last = buf[pos] >> 7;
        int type = buf[pos] & 0x7F;
        size_t blen = ((size_t)buf[pos + 1] << 16) | ((size_t)buf[pos + 2] << 8) | buf[pos + 3];
        pos += 4;
        if (pos + blen > len) return -1;
        if (first) {
            if (type != 0 || blen != 34) return -1;
            const uint8_t *q = buf + pos;
            si->min_blocksize = ((uint32_t)q[0] << 8) | q[1];
            si->max_blocksize = ((uint32_t)q[2] << 8) | q[3];
            si->min_framesize = ((uint32_t)q[4] << 16) | ((uint32_t)q[5] << 8) | q[6];
            si->max_framesize = ((uint32_t)q[7] << 16) | ((uint32_t)q[8] << 8) | q[9];
            si->sample_rate = ((uint32_t)q[10] << 12) | ((uint32_t)q[11] << 4) | (q[12] >> 4);
            si->channels = ((q[12] >> 1) & 7u) + 1;
            si->bps = (((uint32_t)(q[12] & 1) << 4) | (q[13] >> 4)) + 1;
            si->total_samples = ((uint64_t)(q[13] & 0xF) << 32) | ((uint64_t)q[14] << 24) | ((uint64_t)q[15] << 16) | ((uint64_t)q[16] << 8) | q[17];
            memcpy(si->md5, q + 18, 16);
            si->pad = 0;
            first = 0;
        } else if (type == 4) {
            if (vc_off) *vc_off = pos;
            if (vc_len) *vc_len = blen;
        }
        pos += blen;
    }
    return (long)pos;
}

/* Decode a complete stream located at buf[0..len).  Decoding stops at the end of
 * the buffer or at the first position that is not a valid frame (e.g. the next
 * concatenated "fLaC" stream).  out is channel-interleaved int32.
 * Returns 0 on success (>=0 frames decoded, stopped cleanly), <0 on a hard error. */
int ofl_decode(const uint8_t *buf, size_t len, ofl_streaminfo *si, int32_t *out,
               size_t out_cap_frames /* samples per channel */, size_t *n_samples,
               ofl_frameinfo *frames, size_t frames_cap, size_t *n_frames,
               size_t *bytes_consumed, size_t *first_frame_offset) {
    crc_init();
    long hp = ofl_parse_header(buf, len, si, NULL, NULL);
    if (hp < 0) return -1;
    size_t pos = (size_t)hp;
    if (first_frame_offset) *first_frame_offset = pos;
    int64_t *chbuf[OFL_MAX_CHANNELS];
    for (int c = 0; c < OFL_MAX_CHANNELS; c++) chbuf[c] = (int64_t *)malloc(sizeof(int64_t) * (OFL_MAX_BLOCKSIZE + 1));
    size_t ns = 0, nf = 0;
    int rc = 0;
    while (pos < len) {
        if (len - pos >= 4 && memcmp(buf + pos, "fLaC", 4) == 0) break;
        ofl_frameinfo fi;
        long fs = decode_frame(buf + pos, len - pos, si, chbuf, &fi);
        if (fs < 0) { rc = (int)fs * 10; break; }
        if (fi.channels != si->channels) { rc = -100; break; }
        fi.byte_offset = pos;
        if (frames && nf < frames_cap) frames[nf] = fi;
        if (out) {
            if (ns + fi.blocksize > out_cap_frames) { rc = -101; break; }
            for (uint32_t c = 0; c < fi.channels; c++)
                for (uint32_t i = 0; i < fi.blocksize; i++)
                    out[(ns + i) * fi.channels + c] = (int32_t)chbuf[c][i];
        }
        ns += fi.blocksize;
        nf++;
        pos += (size_t)fs;
    }
    for (int c = 0; c < OFL_MAX_CHANNELS; c++) free(chbuf[c]);
    if (n_samples) *n_samples = ns;
    if (n_frames) *n_frames = nf;
    if (bytes_consumed) *bytes_consumed = pos;
    return rc;
}

/* Decode a bare run of frames (no "fLaC" header); si gives channels/bps/sample_rate
 * defaults.  Used to check the product's frame regions directly.                   */
int ofl_decode_frames(const uint8_t *buf, size_t len, const ofl_streaminfo *si, int32_t *out,
                      size_t out_cap_frames, size_t *n_samples, ofl_frameinfo *frames,
                      size_t frames_cap, size_t *n_frames) {
    crc_init();
    int64_t *chbuf[OFL_MAX_CHANNELS];
    for (int c = 0; c < OFL_MAX_CHANNELS; c++) chbuf[c] = (int64_t *)malloc(sizeof(int64_t) * (OFL_MAX_BLOCKSIZE + 1));
    size_t ns = 0, nf = 0, pos = 0;
    int rc = 0;
    while (pos < len) {
        ofl_frameinfo fi;
        long fs = decode_frame(buf + pos, len - pos, si, chbuf, &fi);
        if (fs < 0) { rc = (int)fs * 10; break; }
        fi.byte_offset = pos;
        if (frames && nf < frames_cap) frames[nf] = fi;
        if (out) {
            if (ns + fi.blocksize > out_cap_frames) { rc = -101; break; }
            for (uint32_t c = 0; c < fi.channels; c++)
                for (uint32_t i = 0; i < fi.blocksize; i++)
                    out[(ns + i) * fi.channels + c] = (int32_t)chbuf[c][i];
        }
        ns += fi.blocksize;
        nf++;
        pos += (size_t)fs;
    }
    for (int c = 0; c < OFL_MAX_CHANNELS; c++) free(chbuf[c]);
    if (n_samples) *n_samples = ns;
    if (n_frames) *n_frames = nf;
    return rc;
}

/* ------------------------------------------------------------------------- */
/* Bit writer (MSB first)                                                      */
/* ------------------------------------------------------------------------- */
typedef struct {
    uint8_t *p;
    size_t cap;
    size_t nbytes;
    uint64_t acc;
    int nacc; /* bits in acc */
    int err;
} bw_t;

static inline void bw_flush_bytes(bw_t *w) {
    while (w->nacc >= 8) {
        if (w->nbytes >= w->cap) { w->err = 1; w->nacc -= 8; continue; }
        w->p[w->nbytes++] = (uint8_t)(w->acc >> (w->nacc - 8));
        w->nacc -= 8;
    }
}
static inline void bw_bits(bw_t *w, uint64_t v, int n) { /* n <= 32 */
    if (n == 0) return;
    if (n < 64) v &= ((((uint64_t)1) << n) - 1);
    w->acc = (w->acc << n) | v;
    w->nacc += n;
    bw_flush_bytes(w);
}
static inline void bw_bits64(bw_t *w, uint64_t v, int n) {
    if (n > 32) { bw_bits(w, v >> 32, n - 32); bw_bits(w, v & 0xFFFFFFFFu, 32); }
    else bw_bits(w, v, n);
}
static inline void bw_unary(bw_t *w, uint32_t q) {
    while (q >= 32) { bw_bits(w, 0, 32); q -= 32; }
    bw_bits(w, 1, (int)q + 1);
}
static inline void bw_align(bw_t *w) {
    if (w->nacc & 7) bw_bits(w, 0, 8 - (w->nacc & 7));
}

/* ------------------------------------------------------------------------- */
/* Encoder: libFLAC compression-level search restated                          */
/* ------------------------------------------------------------------------- */
typedef struct {
    int do_mid_side, loose_mid_side;
    int max_lpc_order;
    int min_part_order, max_part_order;
} preset_t;

/* docs/sonos-pyflac.txt:6926-6934.  Levels 6-8 use subdivide_tukey(n) apodizations in
 * libFLAC; this oracle keeps tukey(0.5) for them (documented approximation -- the
 * reference never uses a level other than 5 on the create-streaming path, cli.py:733). */
static const preset_t presets[9] = {
    { 0, 0, 0, 0, 3 }, { 1, 1, 0, 0, 3 }, { 1, 0, 0, 0, 3 }, { 0, 0, 6, 0, 4 }, { 1, 1, 8, 0, 4 },
    { 1, 0, 8, 0, 5 }, { 1, 0, 8, 0, 6 }, { 1, 0, 12, 0, 6 }, { 1, 0, 12, 0, 6 },
};

typedef struct {
    int type;  /* 0 const 1 verbatim 2 fixed 3 lpc */
    int order;
    int wasted;
    int bps;   /* subframe bps after wasted-bit removal (incl. side bit) */
    int qlp_precision, qlp_shift;
    int32_t qlp[OFL_MAX_LPC_ORDER];
    int part_order;
    int rice2;
    uint32_t rice_param[1 << 8];
    int32_t *residual; /* [blocksize] (index from 0 = sample `order`) */
    uint64_t bits;     /* estimate */
} subframe_t;

typedef struct {
    int blocksize_max;
    float *window;       /* tukey(0.5) of current length */
    int window_len;
    float *windowed;
    int64_t *sig[4];     /* L, R, M, S working signals (after wasted-bit shift) */
    int32_t *res[2];     /* two residual scratch buffers */
    int32_t *best_res[4];
    uint64_t *psums;     /* partition sums scratch */
} enc_ws_t;

static int ilog2u(uint32_t v) { int l = 0; while (v >>= 1) l++; return l; }

static void make_tukey(float *w, int L, float p) {
    for (int n = 0; n < L; n++) w[n] = 1.0f;
    if (p <= 0.0f) return;
    if (p >= 1.0f) { /* hann */
        for (int n = 0; n < L; n++) w[n] = (float)(0.5f - 0.5f * cosf(2.0f * (float)M_PI * n / (L - 1)));
        return;
    }
    int Np = (int)(p / 2.0f * L) - 1;
    if (Np > 0) {
        for (int n = 0; n <= Np; n++) {
            w[n] = (float)(0.5f - 0.5f * cosf((float)(M_PI * n / Np)));
            w[L - Np - 1 + n] = (float)(0.5f - 0.5f * cosf((float)(M_PI * (n + Np) / Np)));
        }
    }
}

static uint64_t rice_bits_estimate(uint32_t k, uint32_t n, uint64_t abs_sum) {
    uint64_t v = 4 + (uint64_t)(1 + k) * n + (k ? (abs_sum >> (k - 1)) : (abs_sum << 1)) - (n >> 1);
    return v > 0xFFFFFFFFull ? 0xFFFFFFFFull : v;
}

/* Partition-order search by partition sums (libFLAC find_best_partition_order_ with
 * rice_parameter_search_dist 0, no escape coding).  Returns estimated residual bits. */
static uint64_t find_best_partition_order(const int32_t *residual, uint32_t blocksize, uint32_t pred_order,
                                          int min_po, int max_po, uint32_t rice_limit, uint64_t *psums,
                                          int *best_po, uint32_t *best_params) {
    /* limit the order */
    int mpo = 0;
    { uint32_t b = blocksize; while (!(b & 1) && mpo < 15) { mpo++; b >>= 1; } }
    if (max_po > mpo) max_po = mpo;
    while (max_po > 0 && (blocksize >> max_po) <= pred_order) max_po--;
    if (min_po > max_po) min_po = max_po;
    /* sums at max order */
    uint32_t nparts = 1u << max_po;
    uint32_t psize = blocksize >> max_po;
    {
        uint32_t ri = 0, end = psize - pred_order;
        for (uint32_t p = 0; p < nparts; p++) {
            uint64_t s = 0;
            for (; ri < end; ri++) { int64_t r = residual[ri]; s += (uint64_t)(r < 0 ? -r : r); }
            psums[p] = s;
            end += psize;
        }
    }
    /* merge down */
    {
        uint32_t from = 0, to = nparts;
        for (int po = max_po - 1; po >= min_po; po--) {
            uint32_t np = 1u << po;
            for (uint32_t i = 0; i < np; i++) psums[to + i] = psums[from + 2 * i] + psums[from + 2 * i + 1];
            from = to;
            to += np;
        }
    }
    uint64_t best = UINT64_MAX;
    uint32_t params[1 << 8];
    uint32_t off = 0;
    uint32_t residual_samples = blocksize - pred_order;
    for (int po = max_po; po >= min_po; po--) {
        uint32_t np = 1u << po;
        uint64_t bits = 6;
        for (uint32_t p = 0; p < np; p++) {
            uint32_t n = (residual_samples + pred_order) >> po;
            if (p == 0) n -= pred_order;
            uint64_t mean = psums[off + p];
            uint32_t k = 0;
            uint64_t k2 = n;
            while (k2 * 8 < mean) { k += 4; k2 <<= 4; }
            while (k2 < mean) { k++; k2 <<= 1; }
            if (k >= rice_limit) k = rice_limit - 1;
            params[p] = k;
            bits += rice_bits_estimate(k, n, mean);
        }
        off += np;
        if (bits < best) {
            best = bits;
            *best_po = po;
            memcpy(best_params, params, sizeof(uint32_t) * np);
        }
    }
    return best;
}

/* fixed-order residual; returns 0 if every residual fits (INT32_MIN, INT32_MAX] */
static int fixed_residual(const int64_t *s, uint32_t n, int order, int32_t *res) {
    for (uint32_t i = (uint32_t)order; i < n; i++) {
        int64_t r;
        switch (order) {
            case 0: r = s[i]; break;
            case 1: r = s[i] - s[i - 1]; break;
            case 2: r = s[i] - 2 * s[i - 1] + s[i - 2]; break;
            case 3: r = s[i] - 3 * s[i - 1] + 3 * s[i - 2] - s[i - 3]; break;
            default: r = s[i] - 4 * s[i - 1] + 6 * s[i - 2] - 4 * s[i - 3] + s[i - 4]; break;
        }
        if (r <= INT32_MIN || r > INT32_MAX) return -1;
        res[i - (uint32_t)order] = (int32_t)r;
    }
    return 0;
}

static int lpc_residual(const int64_t *s, uint32_t n, int order, const int32_t *q, int shift, int32_t *res) {
    for (uint32_t i = (uint32_t)order; i < n; i++) {
        int64_t sum = 0;
        for (int j = 0; j < order; j++) sum += (int64_t)q[j] * s[i - 1 - (uint32_t)j];
        int64_t r = s[i] - (sum >> shift);
        if (r <= INT32_MIN || r > INT32_MAX) return -1;
        res[i - (uint32_t)order] = (int32_t)r;
    }
    return 0;
}

static int quantize_coeffs(const float *lp, int order, int precision, int32_t *q, int *shift_out) {
    precision--;
    int32_t qmax = 1 << precision, qmin = -qmax;
    qmax--;
    double cmax = 0.0;
    for (int i = 0; i < order; i++) { double d = fabs((double)lp[i]); if (d > cmax) cmax = d; }
    if (cmax <= 0.0) return 2;
    int log2cmax;
    (void)frexp(cmax, &log2cmax);
    log2cmax--;
    int shift = precision - log2cmax - 1;
    if (shift > 15) shift = 15;
    else if (shift < -16) return 1;
    if (shift >= 0) {
        double error = 0.0;
        for (int i = 0; i < order; i++) {
            error += (double)lp[i] * (double)(1 << shift);
            long qq = lround(error);
            if (qq > qmax) qq = qmax; else if (qq < qmin) qq = qmin;
            error -= (double)qq;
            q[i] = (int32_t)qq;
        }
    } else {
        int nshift = -shift;
        double error = 0.0;
        for (int i = 0; i < order; i++) {
            error += (double)lp[i] / (double)(1 << nshift);
            long qq = lround(error);
            if (qq > qmax) qq = qmax; else if (qq < qmin) qq = qmin;
            error -= (double)qq;
            q[i] = (int32_t)qq;
        }
        shift = 0;
    }
    *shift_out = shift;
    return 0;
}

static double expected_bits_per_residual_sample(double lpc_error, double error_scale) {
    if (lpc_error > 0.0) {
        double bps = 0.5 * log(error_scale * lpc_error) / M_LN2;
        return bps >= 0.0 ? bps : 0.0;
    } else if (lpc_error < 0.0) return 1e32;
    return 0.0;
}

/* Evaluate one channel signal (after wasted-bit removal); fills sf with the best
 * candidate.  res_a/res_b are scratch; the winning residual is copied to sf->residual. */
static void process_subframe(enc_ws_t *ws, const preset_t *ps, const int64_t *sig, uint32_t blocksize,
                             int subframe_bps, int stream_bps, int qlp_precision_cfg, subframe_t *sf) {
    const uint32_t rice_limit = stream_bps > 16 ? 31u : 15u;
    sf->bps = subframe_bps;
    /* verbatim baseline */
    sf->type = 1; sf->order = 0;
    sf->bits = 8 + (uint64_t)sf->wasted + (uint64_t)blocksize * (uint64_t)subframe_bps;
    if (blocksize <= 4) return;

    /* fixed predictor trial */
    uint64_t tot[5] = { 0, 0, 0, 0, 0 };
    int guess;
    float fixed_bps[5];
    if (subframe_bps < 28) {
        /* libFLAC FLAC__fixed_compute_best_predictor[_wide]: errors over data[4..], ties prefer
         * the lower order. */
        for (uint32_t i = 4; i < blocksize; i++) {
            int64_t e0 = sig[i];
            int64_t e1 = e0 - sig[i - 1];
            int64_t e2 = e1 - (sig[i - 1] - sig[i - 2]);
            int64_t e3 = e2 - (sig[i - 1] - 2 * sig[i - 2] + sig[i - 3]);
            int64_t e4 = e3 - (sig[i - 1] - 3 * sig[i - 2] + 3 * sig[i - 3] - sig[i - 4]);
            int64_t e[5] = { e0, e1, e2, e3, e4 };
            for (int o = 0; o < 5; o++) tot[o] += (uint64_t)(e[o] < 0 ? -e[o] : e[o]);
        }
        uint64_t m1234 = tot[1] < tot[2] ? tot[1] : tot[2]; if (tot[3] < m1234) m1234 = tot[3]; if (tot[4] < m1234) m1234 = tot[4];
        uint64_t m234 = tot[2] < tot[3] ? tot[2] : tot[3]; if (tot[4] < m234) m234 = tot[4];
        uint64_t m34 = tot[3] < tot[4] ? tot[3] : tot[4];
        if (tot[0] <= m1234) guess = 0;
        else if (tot[1] <= m234) guess = 1;
        else if (tot[2] <= m34) guess = 2;
        else if (tot[3] <= tot[4]) guess = 3;
        else guess = 4;
        for (int o = 0; o < 5; o++)
            fixed_bps[o] = (float)((tot[o] > 0) ? log(M_LN2 * (double)tot[o] / (double)(blocksize - 4)) / M_LN2 : 0.0);
    } else {
        /* libFLAC FLAC__fixed_compute_best_predictor_limit_residual (subframe_bps >= 28): the
         * order-k error runs over samples k.., an order is valid only if every residual is in
         * [-INT32_MAX, INT32_MAX], the smallest valid total wins with a strict '<' walked from
         * order 0 up, and every non-selected order reports 34.0 bits -- which is why an all-zero
         * bps-32 block is coded FIXED-0 and not CONSTANT (fixture sample_dem.flac). */
        int valid[5] = { 1, 1, 1, 1, 1 };
        for (uint32_t i = 0; i < blocksize; i++) {
            int64_t e[5];
            e[0] = sig[i];
            e[1] = i >= 1 ? sig[i] - sig[i - 1] : 0;
            e[2] = i >= 2 ? sig[i] - 2 * sig[i - 1] + sig[i - 2] : 0;
            e[3] = i >= 3 ? sig[i] - 3 * sig[i - 1] + 3 * sig[i - 2] - sig[i - 3] : 0;
            e[4] = i >= 4 ? sig[i] - 4 * sig[i - 1] + 6 * sig[i - 2] - 4 * sig[i - 3] + sig[i - 4] : 0;
            for (int o = 0; o < 5; o++) {
                uint64_t a = (uint64_t)(e[o] < 0 ? -e[o] : e[o]);
                tot[o] += a;
                if (a > (uint64_t)INT32_MAX) valid[o] = 0;
            }
        }
        uint64_t smallest = UINT64_MAX;
        guess = 0;
        for (int o = 0; o < 5; o++) {
            if (valid[o] && tot[o] < smallest) {
                guess = o; smallest = tot[o];
                fixed_bps[o] = (float)((tot[o] > 0) ? log(M_LN2 * (double)tot[o] / (double)(blocksize - 4)) / M_LN2 : 0.0);
            } else fixed_bps[o] = 34.0f;
        }
    }
    /* constant? */
    if (fixed_bps[1] == 0.0f) {
        int constant = 1;
        for (uint32_t i = 1; i < blocksize; i++) if (sig[i] != sig[0]) { constant = 0; break; }
        if (constant) {
            uint64_t bits = 8 + (uint64_t)sf->wasted + (uint64_t)subframe_bps;
            if (bits < sf->bits) { sf->type = 0; sf->order = 0; sf->bits = bits; }
            return;
        }
    }
    int32_t *cand = ws->res[0];
    uint32_t params[1 << 8];
    /* fixed */
    {
        int fo = guess;
        if ((uint32_t)fo >= blocksize) fo = (int)blocksize - 1;
        if (!(fixed_bps[fo] >= (float)subframe_bps)) {
            if (fixed_residual(sig, blocksize, fo, cand) == 0) {
                int po = 0;
                uint64_t rb = find_best_partition_order(cand, blocksize, (uint32_t)fo, ps->min_part_order, ps->max_part_order,
                                                        rice_limit, ws->psums, &po, params);
                uint64_t bits = 8 + (uint64_t)sf->wasted + (uint64_t)fo * (uint64_t)subframe_bps + rb;
                if (bits < sf->bits) {
                    sf->type = 2; sf->order = fo; sf->bits = bits; sf->part_order = po;
                    memcpy(sf->rice_param, params, sizeof(uint32_t) << po);
                    memcpy(sf->residual, cand, sizeof(int32_t) * (blocksize - (uint32_t)fo));
                }
            }
        }
    }
    /* lpc */
    int max_lpc_order = ps->max_lpc_order;
    if ((uint32_t)max_lpc_order >= blocksize) max_lpc_order = (int)blocksize - 1;
    if (max_lpc_order > 0) {
        if (ws->window_len != (int)blocksize) { make_tukey(ws->window, (int)blocksize, 0.5f); ws->window_len = (int)blocksize; }
        float *wd = ws->windowed;
        if (subframe_bps <= 32) for (uint32_t i = 0; i < blocksize; i++) wd[i] = (float)(int32_t)sig[i] * ws->window[i];
        else for (uint32_t i = 0; i < blocksize; i++) wd[i] = (float)sig[i] * ws->window[i];
        double autoc[OFL_MAX_LPC_ORDER + 1];
        int lag = max_lpc_order + 1;
        for (int c = 0; c < lag; c++) autoc[c] = 0.0;
        {
            uint32_t limit = blocksize - (uint32_t)lag, s;
            for (s = 0; s <= limit; s++) {
                double d = wd[s];
                for (int c = 0; c < lag; c++) autoc[c] += d * (double)wd[s + (uint32_t)c];
            }
            for (; s < blocksize; s++) {
                double d = wd[s];
                for (uint32_t c = 0; c < blocksize - s; c++) autoc[c] += d * (double)wd[s + c];
            }
        }
        if (autoc[0] != 0.0) {
            float lp_coeff[OFL_MAX_LPC_ORDER][OFL_MAX_LPC_ORDER];
            double lpc_error[OFL_MAX_LPC_ORDER];
            double lpc[OFL_MAX_LPC_ORDER];
            int mo = max_lpc_order;
            double err = autoc[0];
            for (int i = 0; i < mo; i++) {
                double r = -autoc[i + 1];
                for (int j = 0; j < i; j++) r -= lpc[j] * autoc[i - j];
                r /= err;
                lpc[i] = r;
                int j;
                for (j = 0; j < (i >> 1); j++) {
                    double tmp = lpc[j];
                    lpc[j] += r * lpc[i - 1 - j];
                    lpc[i - 1 - j] += r * tmp;
                }
                if (i & 1) lpc[j] += lpc[j] * r;
                err *= (1.0 - r * r);
                for (j = 0; j <= i; j++) lp_coeff[i][j] = (float)(-lpc[j]);
                lpc_error[i] = err;
                if (err == 0.0) { mo = i + 1; break; }
            }
            /* best order by expected bits */
            int best_order;
            {
                double error_scale = 0.5 / (double)blocksize;
                uint32_t overhead = (uint32_t)(subframe_bps + qlp_precision_cfg);
                double best_bits = (double)(uint32_t)(-1);
                int best_index = 0;
                for (int idx = 0, order = 1; idx < mo; idx++, order++) {
                    double bits = expected_bits_per_residual_sample(lpc_error[idx], error_scale) * (double)(blocksize - (uint32_t)order) + (double)((uint32_t)order * overhead);
                    if (bits < best_bits) { best_index = idx; best_bits = bits; }
                }
                best_order = best_index + 1;
            }
            {
                int order = best_order;
                double rbps = expected_bits_per_residual_sample(lpc_error[order - 1], 0.5 / (double)(blocksize - (uint32_t)order));
                if (!(rbps >= (double)subframe_bps)) {
                    int precision = qlp_precision_cfg;
                    if (subframe_bps <= 17) {
                        int lim = 32 - subframe_bps - ilog2u((uint32_t)order);
                        if (lim < precision) precision = lim;
                    }
                    int32_t q[OFL_MAX_LPC_ORDER];
                    int shift = 0;
                    if (precision >= 5 && quantize_coeffs(lp_coeff[order - 1], order, precision, q, &shift) == 0) {
                        cand = ws->res[1];
                        if (lpc_residual(sig, blocksize, order, q, shift, cand) == 0) {
                            int po = 0;
                            uint64_t rb = find_best_partition_order(cand, blocksize, (uint32_t)order, ps->min_part_order, ps->max_part_order,
                                                                    rice_limit, ws->psums, &po, params);
                            uint64_t bits = 8 + (uint64_t)sf->wasted + 4 + 5 + (uint64_t)order * (uint64_t)(precision + subframe_bps) + rb;
                            if (bits < sf->bits) {
                                sf->type = 3; sf->order = order; sf->bits = bits; sf->part_order = po;
                                sf->qlp_precision = precision; sf->qlp_shift = shift;
                                memcpy(sf->qlp, q, sizeof(int32_t) * (size_t)order);
                                memcpy(sf->rice_param, params, sizeof(uint32_t) << po);
                                memcpy(sf->residual, cand, sizeof(int32_t) * (blocksize - (uint32_t)order));
                            }
                        }
                    }
                }
            }
        }
    }
}

/* Test-only switches (tests/test_gpu_decode_cases.py): streams a FLAC decoder must accept although libFLAC's presets
 * never write them -- LPC orders up to 32 (non-subset) and escaped (raw binary) Rice partitions (RFC 9639 9.2.7.1).
 * Both default to off, which leaves ofl_encode() the libFLAC restatement pinned by the golden fixtures. */
static int g_test_max_lpc_order = 0, g_test_escape_mod = 0;
void ofl_set_test_options(int max_lpc_order, int escape_mod) {
    g_test_max_lpc_order = max_lpc_order < 0 ? 0 : (max_lpc_order > OFL_MAX_LPC_ORDER ? OFL_MAX_LPC_ORDER : max_lpc_order);
    g_test_escape_mod = escape_mod < 0 ? 0 : escape_mod;
}

static void write_subframe(bw_t *w, const subframe_t *sf, const int64_t *sig, uint32_t blocksize) {
    int bps = sf->bps;
    uint32_t typecode = sf->type == 0 ? 0u : sf->type == 1 ? 1u : sf->type == 2 ? (8u | (uint32_t)sf->order) : (32u | (uint32_t)(sf->order - 1));
    bw_bits(w, 0, 1);
    bw_bits(w, typecode, 6);
    if (sf->wasted) { bw_bits(w, 1, 1); bw_unary(w, (uint32_t)sf->wasted - 1); }
    else bw_bits(w, 0, 1);
    if (sf->type == 0) { bw_bits64(w, (uint64_t)sig[0], bps); return; }
    if (sf->type == 1) { for (uint32_t i = 0; i < blocksize; i++) bw_bits64(w, (uint64_t)sig[i], bps); return; }
    for (int i = 0; i < sf->order; i++) bw_bits64(w, (uint64_t)sig[i], bps);
    if (sf->type == 3) {
        bw_bits(w, (uint64_t)(sf->qlp_precision - 1), 4);
        bw_bits(w, (uint64_t)sf->qlp_shift & 31u, 5);
        for (int i = 0; i < sf->order; i++) bw_bits(w, (uint64_t)(int64_t)sf->qlp[i], sf->qlp_precision);
    }
    /* residual */
    int rice2 = 0;
    uint32_t np = 1u << sf->part_order;
    for (uint32_t p = 0; p < np; p++) if (sf->rice_param[p] >= 15) { rice2 = 1; break; }
    bw_bits(w, (uint64_t)rice2, 2);
    bw_bits(w, (uint64_t)sf->part_order, 4);
    uint32_t ri = 0;
    for (uint32_t p = 0; p < np; p++) {
        uint32_t n = blocksize >> sf->part_order;
        if (p == 0) n -= (uint32_t)sf->order;
        uint32_t k = sf->rice_param[p];
        if (g_test_escape_mod > 0 && (p % (uint32_t)g_test_escape_mod) == 0) {
            /* escape code: parameter all ones, 5-bit raw width, residuals as two's-complement fields */
            int nb = 0;
            for (uint32_t j = 0; j < n; j++) {
                int64_t r = sf->residual[ri + j];
                int need = 1;
                while (need < 32 && (r < -((int64_t)1 << (need - 1)) || r > ((int64_t)1 << (need - 1)) - 1)) need++;
                if (need > nb) nb = need;
            }
            if (n == 0) nb = 0;
            int all_zero = 1;
            for (uint32_t j = 0; j < n; j++) if (sf->residual[ri + j]) { all_zero = 0; break; }
            if (all_zero) nb = 0;
            if (nb <= 31) {
                bw_bits(w, rice2 ? 31 : 15, rice2 ? 5 : 4);
                bw_bits(w, (uint64_t)nb, 5);
                for (uint32_t j = 0; j < n; j++) { int32_t r = sf->residual[ri++]; if (nb) bw_bits64(w, (uint64_t)(int64_t)r, nb); }
                continue;
            }
        }
        bw_bits(w, k, rice2 ? 5 : 4);
        for (uint32_t j = 0; j < n; j++) {
            int32_t r = sf->residual[ri++];
            uint32_t u = ((uint32_t)r << 1) ^ (uint32_t)(r >> 31);
            bw_unary(w, u >> k);
            if (k) bw_bits(w, u & ((1u << k) - 1u), (int)k);
        }
    }
}

static int sample_rate_code(uint32_t sr) {
    switch (sr) {
        case 88200: return 1; case 176400: return 2; case 192000: return 3; case 8000: return 4;
        case 16000: return 5; case 22050: return 6; case 24000: return 7; case 32000: return 8;
        case 44100: return 9; case 48000: return 10; case 96000: return 11; default: return 0;
    }
}
static int bps_code(int bps) {
    switch (bps) { case 8: return 1; case 12: return 2; case 16: return 4; case 20: return 5; case 24: return 6; case 32: return 7; default: return 0; }
}

static void write_frame_header(bw_t *w, uint32_t blocksize, uint32_t sample_rate, int chan_code, int bps, uint64_t frame_number) {
    size_t start = w->nbytes;
    int bs_code, bs_extra = 0;
    switch (blocksize) {
        case 192: bs_code = 1; break; case 576: bs_code = 2; break; case 1152: bs_code = 3; break;
        case 2304: bs_code = 4; break; case 4608: bs_code = 5; break; case 256: bs_code = 8; break;
        case 512: bs_code = 9; break; case 1024: bs_code = 10; break; case 2048: bs_code = 11; break;
        case 4096: bs_code = 12; break; case 8192: bs_code = 13; break; case 16384: bs_code = 14; break;
        case 32768: bs_code = 15; break;
        default: if (blocksize <= 256) { bs_code = 6; bs_extra = 8; } else { bs_code = 7; bs_extra = 16; }
    }
    bw_bits(w, 0x3FFE, 14);
    bw_bits(w, 0, 1);
    bw_bits(w, 0, 1); /* fixed blocksize */
    bw_bits(w, (uint64_t)bs_code, 4);
    bw_bits(w, (uint64_t)sample_rate_code(sample_rate), 4);
    bw_bits(w, (uint64_t)chan_code, 4);
    bw_bits(w, (uint64_t)bps_code(bps), 3);
    bw_bits(w, 0, 1);
    /* utf-8 coded frame number */
    {
        uint64_t v = frame_number;
        if (v < 0x80) bw_bits(w, v, 8);
        else if (v < 0x800) { bw_bits(w, 0xC0 | (v >> 6), 8); bw_bits(w, 0x80 | (v & 0x3F), 8); }
        else if (v < 0x10000) { bw_bits(w, 0xE0 | (v >> 12), 8); bw_bits(w, 0x80 | ((v >> 6) & 0x3F), 8); bw_bits(w, 0x80 | (v & 0x3F), 8); }
        else if (v < 0x200000) { bw_bits(w, 0xF0 | (v >> 18), 8); bw_bits(w, 0x80 | ((v >> 12) & 0x3F), 8); bw_bits(w, 0x80 | ((v >> 6) & 0x3F), 8); bw_bits(w, 0x80 | (v & 0x3F), 8); }
        else if (v < 0x4000000) { bw_bits(w, 0xF8 | (v >> 24), 8); bw_bits(w, 0x80 | ((v >> 18) & 0x3F), 8); bw_bits(w, 0x80 | ((v >> 12) & 0x3F), 8); bw_bits(w, 0x80 | ((v >> 6) & 0x3F), 8); bw_bits(w, 0x80 | (v & 0x3F), 8); }
        else { bw_bits(w, 0xFC | (v >> 30), 8); bw_bits(w, 0x80 | ((v >> 24) & 0x3F), 8); bw_bits(w, 0x80 | ((v >> 18) & 0x3F), 8); bw_bits(w, 0x80 | ((v >> 12) & 0x3F), 8); bw_bits(w, 0x80 | ((v >> 6) & 0x3F), 8); bw_bits(w, 0x80 | (v & 0x3F), 8); }
    }
    if (bs_extra) bw_bits(w, blocksize - 1, bs_extra);
    if (sample_rate_code(sample_rate) == 0) { /* not representable: fall back to "from STREAMINFO" */ }
    bw_bits(w, crc8(w->p + start, w->nbytes - start), 8);
}

static int get_wasted_bits(int64_t *s, uint32_t n) {
    int64_t x = 0;
    for (uint32_t i = 0; i < n && !(x & 1); i++) x |= s[i];
    if (x == 0) return 0;
    int shift = 0;
    while (!(x & 1)) { x >>= 1; shift++; }
    if (shift > 0) for (uint32_t i = 0; i < n; i++) s[i] >>= shift;
    return shift;
}

static enc_ws_t *ws_new(int blocksize) {
    enc_ws_t *ws = (enc_ws_t *)calloc(1, sizeof(enc_ws_t));
    ws->blocksize_max = blocksize;
    ws->window = (float *)malloc(sizeof(float) * (size_t)blocksize);
    ws->windowed = (float *)malloc(sizeof(float) * (size_t)blocksize);
    ws->window_len = -1;
    for (int i = 0; i < 4; i++) ws->sig[i] = (int64_t *)malloc(sizeof(int64_t) * (size_t)blocksize);
    for (int i = 0; i < 2; i++) ws->res[i] = (int32_t *)malloc(sizeof(int32_t) * (size_t)blocksize);
    for (int i = 0; i < 4; i++) ws->best_res[i] = (int32_t *)malloc(sizeof(int32_t) * (size_t)blocksize);
    ws->psums = (uint64_t *)malloc(sizeof(uint64_t) * 1024);
    return ws;
}
static void ws_free(enc_ws_t *ws) {
    free(ws->window); free(ws->windowed);
    for (int i = 0; i < 4; i++) free(ws->sig[i]);
    for (int i = 0; i < 2; i++) free(ws->res[i]);
    for (int i = 0; i < 4; i++) free(ws->best_res[i]);
    free(ws->psums); free(ws);
}

static int qlp_precision_for(int bps, int blocksize) {
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

/* Encode one frame of `n` samples (n <= stream blocksize).  pcm is channel-interleaved
 * int32 with stride `channels`.  Appends to w.                                        */
static void encode_frame(enc_ws_t *ws, const preset_t *ps, bw_t *w, const int32_t *pcm, uint32_t n, int channels,
                         int bps, uint32_t sample_rate, int stream_blocksize, uint64_t frame_number) {
    size_t start = w->nbytes;
    int qlp_prec = qlp_precision_for(bps, stream_blocksize);
    subframe_t sf[4];
    memset(sf, 0, sizeof(sf));
    if (channels == 2 && ps->do_mid_side) {
        for (uint32_t i = 0; i < n; i++) {
            int64_t l = pcm[2 * i], r = pcm[2 * i + 1];
            ws->sig[0][i] = l; ws->sig[1][i] = r;
            ws->sig[2][i] = (l + r) >> 1; ws->sig[3][i] = l - r;
        }
        for (int c = 0; c < 4; c++) {
            sf[c].residual = ws->best_res[c];
            sf[c].wasted = get_wasted_bits(ws->sig[c], n);
            int sbps = bps + (c == 3 ? 1 : 0) - sf[c].wasted;
            process_subframe(ws, ps, ws->sig[c], n, sbps, bps, qlp_prec, &sf[c]);
        }
        uint64_t bits[4] = { sf[0].bits + sf[1].bits, sf[0].bits + sf[3].bits, sf[1].bits + sf[3].bits, sf[2].bits + sf[3].bits };
        int ca = 0;
        for (int i = 1; i < 4; i++) if (bits[i] < bits[ca]) ca = i;
        static const int code[4] = { 1, 8, 9, 10 };
        static const int first[4] = { 0, 0, 3, 2 }, second[4] = { 1, 3, 1, 3 };
        write_frame_header(w, n, sample_rate, code[ca], bps, frame_number);
        write_subframe(w, &sf[first[ca]], ws->sig[first[ca]], n);
        write_subframe(w, &sf[second[ca]], ws->sig[second[ca]], n);
    } else {
        write_frame_header(w, n, sample_rate, channels - 1, bps, frame_number);
        for (int c = 0; c < channels; c++) {
            for (uint32_t i = 0; i < n; i++) ws->sig[0][i] = pcm[(size_t)i * (size_t)channels + (size_t)c];
            sf[0].residual = ws->best_res[0];
            sf[0].wasted = get_wasted_bits(ws->sig[0], n);
            process_subframe(ws, ps, ws->sig[0], n, bps - sf[0].wasted, bps, qlp_prec, &sf[0]);
            write_subframe(w, &sf[0], ws->sig[0], n);
        }
    }
    bw_align(w);
    uint16_t c16 = crc16(w->p + start, w->nbytes - start);
    bw_bits(w, c16, 16);
}

static const char OFL_VENDOR[] = "flac-raster-b200 oracle (libFLAC 1.4.3 level search restated)";

static void write_stream_header(bw_t *w, int channels, int bps, uint32_t sample_rate, uint32_t blocksize) {
    const uint8_t magic[4] = { 'f', 'L', 'a', 'C' };
    for (int i = 0; i < 4; i++) bw_bits(w, magic[i], 8);
    bw_bits(w, 0x00, 8); bw_bits(w, 34, 24);
    bw_bits(w, blocksize, 16); bw_bits(w, blocksize, 16);
    bw_bits(w, 0, 24); bw_bits(w, 0, 24);
    bw_bits(w, sample_rate, 20); bw_bits(w, (uint64_t)(channels - 1), 3); bw_bits(w, (uint64_t)(bps - 1), 5);
    bw_bits(w, 0, 4); bw_bits(w, 0, 32); /* total samples: unknown, like the reference */
    for (int i = 0; i < 16; i++) bw_bits(w, 0, 8);
    uint32_t vl = (uint32_t)strlen(OFL_VENDOR);
    bw_bits(w, 0x84, 8); bw_bits(w, 8 + vl, 24);
    bw_bits(w, vl & 0xFF, 8); bw_bits(w, (vl >> 8) & 0xFF, 8); bw_bits(w, (vl >> 16) & 0xFF, 8); bw_bits(w, (vl >> 24) & 0xFF, 8);
    for (uint32_t i = 0; i < vl; i++) bw_bits(w, (uint8_t)OFL_VENDOR[i], 8);
    bw_bits(w, 0, 32);
}

/* pcm: interleaved int32 (n per channel).  Returns 0 or <0.                         */
int ofl_encode(const int32_t *pcm, size_t n, int channels, int bps, int sample_rate, int blocksize, int level,
               int write_header, uint8_t *out, size_t out_cap, size_t *out_len, uint32_t *frame_sizes,
               size_t frame_sizes_cap, size_t *n_frames) {
    crc_init();
    if (channels < 1 || channels > 8) return -2; /* FLAC__MAX_CHANNELS, dump:3501 */
    if (bps < 4 || bps > 32) return -3;
    if (blocksize < 16 || blocksize > 65535) return -4;
    if (level < 0) level = 0; if (level > 8) level = 8;
    preset_t pst = presets[level];
    if (g_test_max_lpc_order > 0) pst.max_lpc_order = g_test_max_lpc_order;
    const preset_t *ps = &pst;
    bw_t w = { out, out_cap, 0, 0, 0, 0 };
    if (write_header) write_stream_header(&w, channels, bps, (uint32_t)sample_rate, (uint32_t)blocksize);
    enc_ws_t *ws = ws_new(blocksize);
    size_t nf = 0;
    for (size_t pos = 0; pos < n; pos += (size_t)blocksize) {
        uint32_t cur = (uint32_t)((n - pos) < (size_t)blocksize ? (n - pos) : (size_t)blocksize);
        size_t before = w.nbytes;
        encode_frame(ws, ps, &w, pcm + pos * (size_t)channels, cur, channels, bps, (uint32_t)sample_rate, blocksize, nf);
        if (frame_sizes && nf < frame_sizes_cap) frame_sizes[nf] = (uint32_t)(w.nbytes - before);
        nf++;
    }
    ws_free(ws);
    if (out_len) *out_len = w.nbytes;
    if (n_frames) *n_frames = nf;
    return w.err ? -5 : 0;
}

size_t ofl_encode_bound(size_t n, int channels, int bps, int blocksize) {
    size_t frames = (n + (size_t)blocksize - 1) / (size_t)blocksize;
    return 256 + frames * (32 + (size_t)channels * 8) + (n * (size_t)channels * (size_t)bps + 7) / 8 + 64;
}

/* ------------------------------------------------------------------------- */
/* The reference's tile loop, restated end to end (timed CPU baseline).        */
/*                                                                             */
/* Encode (cli.py:691-763 -> converter.py:152-216 per tile):                   */
/*   per tile: min/max over the tile (converter.py:67-68), float64 stretch      */
/*   (converter.py:72), *32767 -> int16 | *8388607 -> int32 truncation         */
/*   (converter.py:79-81), then pyFLAC (array dtype -> bps 16/32, dump:1986-1994)*/
/*   -> libFLAC level `level`, blocksize 4096.                                  */
/* dtype codes: 0 u8, 1 i16, 2 u16, 3 i32, 4 u32, 5 f32                         */
/* ------------------------------------------------------------------------- */
typedef struct {
    int32_t row_off, col_off, h, w;
} ofl_tile;

static inline int64_t load_int(const void *base, int dtype, size_t idx) {
    switch (dtype) {
        case 0: return ((const uint8_t *)base)[idx];
        case 1: return ((const int16_t *)base)[idx];
        case 2: return ((const uint16_t *)base)[idx];
        case 3: return ((const int32_t *)base)[idx];
        case 4: return ((const uint32_t *)base)[idx];
    }
    return 0;
}

/* numpy `d - dmin` is done in the raster's own dtype (wraps), converter.py:72 / SURVEY 0.7 */
static inline double wrap_sub(int64_t d, int64_t dmin, int dtype) {
    switch (dtype) {
        case 0: return (double)(uint8_t)(d - dmin);
        case 1: return (double)(int16_t)(d - dmin);
        case 2: return (double)(uint16_t)(d - dmin);
        case 3: return (double)(int32_t)(d - dmin);
        case 4: return (double)(uint32_t)(d - dmin);
    }
    return 0.0;
}

static inline int32_t trunc_i32(double v) { /* x86 cvttsd2si semantics: out of range / NaN -> INT_MIN */
    if (!(v > -2147483649.0 && v < 2147483648.0)) return INT32_MIN;
    return (int32_t)v;
}
static inline int16_t trunc_i16(double v) { /* numpy float64->int16: via int32 then wrap on x86 */
    return (int16_t)trunc_i32(v);
}

/* Quantise one tile into interleaved int32 PCM; returns effective bps (16 or 32).   */
int ofl_ref_quantize_tile(const void *raster, int dtype, int bands, int64_t H, int64_t W, const ofl_tile *t,
                          int first_band, int nch, int32_t *pcm, double *dmin_out, double *dmax_out) {
    size_t plane = (size_t)H * (size_t)W;
    int label16 = (dtype == 0 || dtype == 1 || dtype == 2);
    (void)bands;
    if (dtype == 5) {
        const float *f = (const float *)raster;
        float mn = INFINITY, mx = -INFINITY;
        int nan = 0;
        for (int c = 0; c < nch; c++)
            for (int r = 0; r < t->h; r++)
                for (int x = 0; x < t->w; x++) {
                    float v = f[(size_t)(first_band + c) * plane + (size_t)(t->row_off + r) * (size_t)W + (size_t)(t->col_off + x)];
                    if (v != v) nan = 1;
                    if (v < mn) mn = v;
                    if (v > mx) mx = v;
                    /* float data used "as-is" (converter.py:61-64); float32 * 8388607 in float32 (NEP 50) */
                    float p = v * 8388607.0f;
                    pcm[((size_t)r * (size_t)t->w + (size_t)x) * (size_t)nch + (size_t)c] = trunc_i32((double)p);
                }
        *dmin_out = nan ? NAN : (double)mn;
        *dmax_out = nan ? NAN : (double)mx;
        return 32;
    }
    int64_t mn = INT64_MAX, mx = INT64_MIN;
    for (int c = 0; c < nch; c++)
        for (int r = 0; r < t->h; r++)
            for (int x = 0; x < t->w; x++) {
                int64_t v = load_int(raster, dtype, (size_t)(first_band + c) * plane + (size_t)(t->row_off + r) * (size_t)W + (size_t)(t->col_off + x));
                if (v < mn) mn = v;
                if (v > mx) mx = v;
            }
    *dmin_out = (double)mn;
    *dmax_out = (double)mx;
    double range = wrap_sub(mx, mn, dtype);
    for (int c = 0; c < nch; c++)
        for (int r = 0; r < t->h; r++)
            for (int x = 0; x < t->w; x++) {
                int64_t v = load_int(raster, dtype, (size_t)(first_band + c) * plane + (size_t)(t->row_off + r) * (size_t)W + (size_t)(t->col_off + x));
                int32_t q;
                if (mx > mn) {
                    double norm = 2.0 * wrap_sub(v, mn, dtype) / range - 1.0;
                    q = label16 ? (int32_t)trunc_i16(norm * 32767.0) : trunc_i32(norm * 8388607.0);
                } else q = 0;
                pcm[((size_t)r * (size_t)t->w + (size_t)x) * (size_t)nch + (size_t)c] = q;
            }
    return label16 ? 16 : 32;
}

/* Encode n_tiles tiles; tile i's FLAC frames (no stream header) go to out + i*slot.
 * sizes[i] = bytes.  Returns 0 or <0.                                                */
int ofl_ref_encode_tiles(const void *raster, int dtype, int bands, int64_t H, int64_t W, const ofl_tile *tiles,
                         int n_tiles, int first_band, int nch, int level, int sample_rate, uint8_t *out,
                         size_t slot, uint64_t *sizes, double *dmins, double *dmaxs, int n_threads) {
    crc_init();
    int rc = 0;
#ifdef _OPENMP
    if (n_threads > 0) omp_set_num_threads(n_threads);
#endif
#pragma omp parallel for schedule(dynamic, 1)
    for (int i = 0; i < n_tiles; i++) {
        const ofl_tile *t = &tiles[i];
        size_t n = (size_t)t->h * (size_t)t->w;
        int32_t *pcm = (int32_t *)malloc(sizeof(int32_t) * n * (size_t)nch);
        int bps = ofl_ref_quantize_tile(raster, dtype, bands, H, W, t, first_band, nch, pcm, &dmins[i], &dmaxs[i]);
        size_t len = 0, nf = 0;
        int r = ofl_encode(pcm, n, nch, bps, sample_rate, 4096, level, 0, out + (size_t)i * slot, slot, &len, NULL, 0, &nf);
        if (r) {
#pragma omp critical
            rc = r;
        }
        sizes[i] = len;
        free(pcm);
    }
    return rc;
}

/* Decode (cli.py:1016-1023 -> converter.py:241-282): libFLAC decode, pyFLAC int16
 * narrowing (dump:1809-1854), PCM_16 WAV through soundfile -> float64 = int16/32768
 * (dump:1621-1640; bps-32 streams keep their top 16 bits [inferred, SURVEY 0.6]),
 * then _denormalize_from_audio in float32 (converter.py:99-110).
 * Output band-planar into out + i*out_slot (raster dtype).                           */
static inline float rint_f32(float v) { return nearbyintf(v); }

int ofl_ref_decode_tiles(const uint8_t *bytes, const uint64_t *offsets, const uint64_t *sizes, int n_tiles,
                         int nch, int bps, int sample_rate, const int32_t *hw /* h,w pairs */, int dtype,
                         const double *dmins, const double *dmaxs, void *out, size_t out_slot, int n_threads) {
    crc_init();
    int rc = 0;
#ifdef _OPENMP
    if (n_threads > 0) omp_set_num_threads(n_threads);
#endif
#pragma omp parallel for schedule(dynamic, 1)
    for (int i = 0; i < n_tiles; i++) {
        size_t n = (size_t)hw[2 * i] * (size_t)hw[2 * i + 1];
        int32_t *pcm = (int32_t *)malloc(sizeof(int32_t) * n * (size_t)nch);
        ofl_streaminfo si;
        memset(&si, 0, sizeof(si));
        si.channels = (uint32_t)nch; si.bps = (uint32_t)bps; si.sample_rate = (uint32_t)sample_rate;
        size_t ns = 0, nf = 0;
        int r = ofl_decode_frames(bytes + offsets[i], sizes[i], &si, pcm, n, &ns, NULL, 0, &nf);
        if (r || ns != n) {
#pragma omp critical
            rc = r ? r : -200;
        }
        float dmin = (float)dmins[i];
        float range = (float)(dmaxs[i] - dmins[i]);
        uint8_t *o = (uint8_t *)out + (size_t)i * out_slot;
        for (int c = 0; c < nch; c++)
            for (size_t k = 0; k < n; k++) {
                int32_t s = pcm[k * (size_t)nch + (size_t)c];
                int16_t s16 = bps == 32 ? (int16_t)(s >> 16) : (int16_t)s;
                float x = (float)((double)s16 / 32768.0);
                size_t oi = (size_t)c * n + k;
                if (dtype == 5) { ((float *)o)[oi] = x; continue; }
                float d = rint_f32((x + 1.0f) / 2.0f * range + dmin);
                switch (dtype) { /* numpy float32 -> int casts (in-range values only are defined) */
                    case 0: ((uint8_t *)o)[oi] = (uint8_t)(int32_t)d; break;
                    case 1: ((int16_t *)o)[oi] = (int16_t)(int32_t)d; break;
                    case 2: ((uint16_t *)o)[oi] = (uint16_t)(int32_t)d; break;
                    case 3: ((int32_t *)o)[oi] = trunc_i32((double)d); break;
                    case 4: ((uint32_t *)o)[oi] = (uint32_t)(int64_t)d; break;
                }
            }
        free(pcm);
    }
    return rc;
}

int ofl_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

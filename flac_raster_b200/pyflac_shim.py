"""pyFLAC-shaped front end of libfrx: ``StreamEncoder`` / ``FileDecoder`` with pyFLAC 3.0.0's semantics
(docs/sonos-pyflac.txt:1947-2213 encoder, 1571-1640 + 1809-1854 decoder), so code written against
``pyflac`` -- the reference's converter.py:201-216, 241-242 and spatial_encoder.py:278-291 -- can run on
the GPU codec by ``sys.modules['pyflac'] = flac_raster_b200.pyflac_shim``.

Semantics kept: channels = samples.shape[1], bits per sample = itemsize * 8, ``astype(int32)`` truncation
of float input (dump:1986-1994); header chunks arrive with num_samples == 0, then one chunk per frame
(dump:2311-2332); > 8 channels or an unsupported sample size raise EncoderInitException (dump:2209-2210);
FileDecoder.process() returns (float64 (N, ch) = int16 / 32768, sample_rate), bps-32 streams keeping their
top 16 bits (PCM_16 WAV in pyFLAC, dump:1621-1640).  Difference: frames are produced in ``finish()``
(one batched GPU call over everything passed to ``process``), not incrementally.
"""
from __future__ import annotations

from typing import Callable, Optional

import numpy as np

from . import _lib as L
from . import flacmeta
from .codec import make_dtiles
from .converter import get_codec


class EncoderInitException(Exception):
    pass


class EncoderProcessException(Exception):
    pass


class DecoderInitException(Exception):
    pass


class DecoderProcessException(Exception):
    pass


class StreamEncoder:
    def __init__(self, sample_rate: int, write_callback: Callable[[bytes, int, int, int], None], seek_callback=None,
                 tell_callback=None, metadata_callback=None, compression_level: int = 5, blocksize: int = 0,
                 streamable_subset: bool = True, verify: bool = False, limit_min_bitrate: bool = False, device: int = 0):
        self.write_callback = write_callback
        self._sample_rate = int(sample_rate)
        self._blocksize = int(blocksize) or 4096  # libFLAC picks 4096 at levels >= 3 when 0 is passed
        self._compression_level = int(compression_level)
        self._channels = 0
        self._bits_per_sample = 0
        self._initialised = False
        self._chunks = []
        self._device = device

    def _init(self):
        if not (1 <= self._channels <= 8):
            raise EncoderInitException(f"FLAC__STREAM_ENCODER_INIT_STATUS_INVALID_NUMBER_OF_CHANNELS ({self._channels})")
        if self._bits_per_sample not in (16, 32):
            raise EncoderInitException(f"FLAC__STREAM_ENCODER_INIT_STATUS_INVALID_BITS_PER_SAMPLE ({self._bits_per_sample}); "
                                       "libfrx encodes 16- and 32-bit PCM")
        if not (16 <= self._blocksize <= 4096):
            raise EncoderInitException(f"FLAC__STREAM_ENCODER_INIT_STATUS_INVALID_BLOCK_SIZE ({self._blocksize})")
        si = flacmeta.StreamInfo(self._blocksize, self._blocksize, 0, 0, self._sample_rate, self._channels, self._bits_per_sample)
        head = b"fLaC" + bytes([flacmeta.BLOCK_STREAMINFO]) + (34).to_bytes(3, "big") + si.pack()
        vc = flacmeta.pack_vorbis_comment([])
        vcb = bytes([0x80 | flacmeta.BLOCK_VORBIS_COMMENT]) + len(vc).to_bytes(3, "big") + vc
        for chunk in (head, vcb):
            self.write_callback(chunk, len(chunk), 0, 0)
        self._initialised = True

    def process(self, samples: np.ndarray):
        if not isinstance(samples, np.ndarray):
            raise TypeError("Processing only supports numpy arrays")
        if not self._initialised:
            try:
                self._channels = samples.shape[1]
            except IndexError:
                self._channels = 1
            self._bits_per_sample = samples.dtype.itemsize * 8
            self._init()
        with np.errstate(all="ignore"):
            s = np.ascontiguousarray(samples).astype(np.int32)  # dump:1994 (float input truncates toward zero)
        self._chunks.append(s.reshape(len(s), -1))

    def finish(self) -> bool:
        if not self._initialised or not self._chunks:
            return True
        pcm = np.concatenate(self._chunks, axis=0)
        self._chunks = []
        n, ch = pcm.shape
        dtype = np.int16 if self._bits_per_sample == 16 else np.int32
        planar = np.ascontiguousarray(pcm.T.astype(dtype)).reshape(ch, 1, n)
        res = get_codec(self._device).encode_tiles_host(planar, [(0, 0, 1, n, 0, ch)], level=self._compression_level,
                                                        blocksize=self._blocksize, sample_rate=self._sample_rate,
                                                        pcm_mode="lossless")
        if int(res.status[0]) != 0:
            raise EncoderProcessException("FLAC__STREAM_ENCODER_CLIENT_ERROR")
        data = res.bytes[:int(res.total_bytes[0])]
        pos = 0
        for i, sz in enumerate(res.frame_size):
            sz = int(sz)
            ns = min(self._blocksize, n - i * self._blocksize)
            self.write_callback(data[pos:pos + sz].tobytes(), sz, ns, i)
            pos += sz
        self._initialised = False
        return True


class FileDecoder:
    def __init__(self, input_file, output_file=None, device: int = 0):
        self._input = str(input_file)
        self._device = device

    def process(self):
        blob = np.fromfile(self._input, dtype=np.uint8)
        try:
            hdr = flacmeta.parse_header(blob[:1 << 24].tobytes())
        except ValueError as e:
            raise DecoderInitException(str(e)) from e
        si = hdr.streaminfo
        if si.bps not in (16, 32):
            raise DecoderProcessException(f"Unsupported bit depth: {si.bps}")  # dump:1829-1830
        frames = blob[hdr.first_frame_offset:]
        if si.total_samples:
            n = si.total_samples
        else:  # libFLAC-written streams of the reference leave it 0: count frames by their headers
            n = _count_samples(frames, si.max_blocksize)
        bs = si.max_blocksize
        dtype = np.int16 if si.bps == 16 else np.int32
        out = np.empty((si.channels, 1, n), dtype=dtype)
        dt = make_dtiles([dict(byte_offset=0, byte_size=len(frames), out_offset=0, out_band_stride=n, out_row_stride=n, h=1, w=n,
                               channels=si.channels, bps=si.bps, dmin=0.0, dmax=0.0, first_frame=None,
                               n_frames=(n + bs - 1) // bs)])
        try:
            get_codec(self._device).decode_tiles_host(frames, dt, out, blocksize=bs, pcm_mode="lossless")
        except L.FrxError as e:
            raise DecoderProcessException(str(e)) from e
        pcm = out.reshape(si.channels, n).T
        if si.bps == 32:
            pcm = (pcm >> 16).astype(np.int16)  # libsndfile int32 -> PCM_16 keeps the top 16 bits
        return pcm.astype(np.float64) / 32768.0, si.sample_rate


def _count_samples(frames: np.ndarray, blocksize: int) -> int:
    """Total samples of a fixed-blocksize stream whose STREAMINFO says 'unknown': every frame but the last
    holds `blocksize`; the last frame's header carries its own size.  Finds the LAST valid frame header."""
    b = frames
    cand = np.nonzero((b[:-1] == 0xFF) & (b[1:] == 0xF8))[0]
    for pos in cand[::-1]:
        h = _parse_frame_header(b, int(pos))
        if h is not None:
            number, bsz = h
            return number * blocksize + bsz
    raise DecoderProcessException("no FLAC frame found")


def _crc8(data) -> int:
    c = 0
    for x in data:
        c ^= int(x)
        for _ in range(8):
            c = ((c << 1) ^ 0x07) & 0xFF if c & 0x80 else (c << 1) & 0xFF
    return c


def _parse_frame_header(b, pos):
    if pos + 6 > len(b):
        return None
    b2, b3 = int(b[pos + 2]), int(b[pos + 3])
    bs_code, sr_code = b2 >> 4, b2 & 15
    if bs_code == 0 or sr_code == 15 or (b3 & 1) or (b3 >> 4) > 10:
        return None
    p = pos + 4
    x = int(b[p]); p += 1
    if x < 0x80:
        num, extra = x, 0
    elif x & 0xE0 == 0xC0:
        num, extra = x & 0x1F, 1
    elif x & 0xF0 == 0xE0:
        num, extra = x & 0x0F, 2
    elif x & 0xF8 == 0xF0:
        num, extra = x & 0x07, 3
    elif x & 0xFC == 0xF8:
        num, extra = x & 0x03, 4
    elif x & 0xFE == 0xFC:
        num, extra = x & 0x01, 5
    else:
        return None
    if p + extra + 3 > len(b):
        return None
    for _ in range(extra):
        y = int(b[p]); p += 1
        if y & 0xC0 != 0x80:
            return None
        num = (num << 6) | (y & 0x3F)
    if bs_code == 1:
        bsz = 192
    elif bs_code <= 5:
        bsz = 576 << (bs_code - 2)
    elif bs_code == 6:
        bsz = int(b[p]) + 1; p += 1
    elif bs_code == 7:
        bsz = ((int(b[p]) << 8) | int(b[p + 1])) + 1; p += 2
    else:
        bsz = 256 << (bs_code - 8)
    if sr_code == 12:
        p += 1
    elif sr_code in (13, 14):
        p += 2
    if p >= len(b) or _crc8(b[pos:p]) != int(b[p]):
        return None
    return num, bsz

"""TEST INFRASTRUCTURE: an oracle-backed stand-in for flac_raster_b200.codec.Codec's host entry points, so the
host-side logic (containers, indexes, sharding, CLI plumbing) is exercised by the CPU suite without a GPU.
Never imported by the product; the GPU suite runs the same scenarios through libfrx."""
from types import SimpleNamespace

import numpy as np

import oracle
from oracle import quant


def _lossless_fwd(block):
    dt = block.dtype
    if dt == np.uint16:
        return block.astype(np.int32) - 32768, 16
    if dt in (np.uint8, np.int16):
        return block.astype(np.int32), 16
    if dt == np.int32:
        return block.astype(np.int32), 32
    if dt == np.uint32:
        return (block ^ np.uint32(0x80000000)).view(np.int32), 32
    raise NotImplementedError(dt)


def _lossless_inv(pcm, dtype):
    dtype = np.dtype(dtype)
    if dtype == np.uint16:
        return (pcm + 32768).astype(np.uint16)
    if dtype in (np.uint8, np.int16, np.int32):
        return pcm.astype(dtype)
    if dtype == np.uint32:
        return pcm.astype(np.int32).view(np.uint32) ^ np.uint32(0x80000000)
    raise NotImplementedError(dtype)


class OracleCodec:
    device = 0

    def close(self):
        pass

    def read_tiff(self, src, row0=0, row1=None, bands=None):
        """Codec.read_tiff through the numpy restatement of frx_tiff_unpack (oracle/tiff_unpack.py)."""
        import numpy as np
        from oracle import tiff_unpack
        row1 = src.height if row1 is None else row1
        if not hasattr(src, "chunk_table"):
            from flac_raster_b200.rio import Window
            out = src.read(window=Window(0, row0, src.width, row1 - row0))
        else:
            chunks = src.chunk_table(row0, row1)
            with open(src.name, "rb") as f:
                data = f.read()
            out = tiff_unpack.unpack(data, chunks, src.dtypes[0], src.count, row1 - row0, src.width, src.count,
                                     swap_bytes=(src._bo == ">"))
        return out if bands is None else np.ascontiguousarray(out[list(bands)])

    def tile_md5(self, raster, tiles, pcm_mode="reference"):
        """hashlib over the oracle's PCM (what libFLAC's MD5 covers: interleaved little-endian samples)."""
        import hashlib
        import numpy as np
        from oracle import quant
        out = []
        for (r0, c0, h, w, b0, nch) in tiles:
            assert pcm_mode == "reference"
            pcm, bps, *_ = quant.reference_pcm(np.asarray(raster)[b0:b0 + nch, r0:r0 + h, c0:c0 + w])
            out.append(np.frombuffer(hashlib.md5(pcm.astype("<i2" if bps == 16 else "<i4").tobytes()).digest(), dtype=np.uint8))
        return np.stack(out)

    def encode_tiles_any(self, raster, tiles, **kw):
        return self.encode_tiles_host(raster, tiles, **kw)

    def encode_tiles_host(self, raster, tiles, level=5, blocksize=4096, sample_rate=44100, pcm_mode="reference", out=None):
        chunks, fsizes, mm, toff, tsz, fpt = [], [], [], [], [], []
        pos = 0
        bps = None
        for (r0, c0, h, w, b0, nch) in tiles:
            if nch > 8:
                from flac_raster_b200._lib import E_UNSUPPORTED, FrxUnsupported
                raise FrxUnsupported(E_UNSUPPORTED, "more than 8 channels")
            block = raster[b0:b0 + nch, r0:r0 + h, c0:c0 + w]
            if pcm_mode == "reference":
                pcm, bps, dmin, dmax, _ = quant.reference_pcm(block)
            elif pcm_mode == "lossless":
                p, bps = _lossless_fwd(block)
                pcm = p.reshape(nch, h * w).T
                dmin, dmax = float(block.min()), float(block.max())
            else:
                pcm, bps = quant.spatial_reference_pcm(block)[:2]
                dmin, dmax = float(block.min()), float(block.max())
            enc, fs = oracle.encode_stream(np.ascontiguousarray(pcm), bps, sample_rate, blocksize, level, write_header=False)
            chunks.append(enc)
            fsizes.extend(int(x) for x in fs)
            fpt.append(len(fs))
            toff.append(pos)
            tsz.append(len(enc))
            pos += len(enc)
            mm.append((dmin, dmax))
        data = np.frombuffer(b"".join(chunks), dtype=np.uint8)
        return SimpleNamespace(bytes=data, tile_offset=np.asarray(toff, np.int64), tile_size=np.asarray(tsz, np.int64),
                               tile_minmax=np.asarray(mm, np.float64).reshape(-1, 2), frame_size=np.asarray(fsizes, np.int32),
                               frame_offset=np.concatenate([[0], np.cumsum(fsizes)[:-1]]).astype(np.int64),
                               total_bytes=np.asarray([pos], np.int64), status=np.zeros(1, np.int32),
                               frames_per_tile=np.asarray(fpt, np.int64), bps=bps, blocksize=blocksize)

    def decode_tiles_host(self, data, dtiles, out, frame_size=None, blocksize=4096, pcm_mode="reference"):
        flat = out.reshape(-1)
        for d in dtiles:
            raw = bytes(data[d.byte_offset:d.byte_offset + d.byte_size])
            pcm, frames, rc = oracle.decode_frames(raw, d.channels, d.bps, max_samples=d.h * d.w + 16)
            assert len(pcm) == d.h * d.w, (len(pcm), d.h * d.w, rc)
            if pcm_mode == "lossless":
                blk = _lossless_inv(pcm.T.reshape(d.channels, d.h, d.w), out.dtype)
            else:
                blk = quant.reference_decode_raster(pcm, d.bps, d.h, d.w, d.channels, out.dtype, d.dmin, d.dmax).reshape(
                    d.channels, d.h, d.w)
            for c in range(d.channels):
                for r in range(d.h):
                    o = d.out_offset + c * d.out_band_stride + r * d.out_row_stride
                    flat[o:o + d.w] = blk[c, r]
        return np.zeros(len(dtiles), np.int32)

"""Spatial ("raw frames") FLAC: one independent FLAC stream per tile, concatenated, with an embedded
bbox index -- the reference's spatial_encoder.py (SpatialFrame :21-45, SpatialIndex :48-78,
SpatialFLACEncoder :81-373, SpatialFLACStreamer :376-506) with every tile encoded by ONE batched
libfrx call instead of a Python loop of pyflac.StreamEncoder objects.

Deliberate, documented differences from the reference (SURVEY.md decisions D1/D2, section 0.2/0.3/0.9):
  * pcm_mode defaults to "lossless": the reference hands float32 in [-1,1] to pyFLAC, whose astype(int32)
    turns every sample into -1/0/1.  ``pcm_mode="reference_spatial"`` reproduces that bit-for-bit.
  * byte offsets in the index are the real file offsets (the reference's go stale when mutagen grows
    the first header).
  * more than 8 bands (FLAC's channel limit; the reference raises EncoderInitException) are split into
    consecutive groups of <= 8 channels, each its own stream, listed under "streams" in the tile entry.
"""
from __future__ import annotations

import base64
import gzip
import json
import logging
from pathlib import Path
from typing import Dict, List, Optional, Tuple

import numpy as np

from . import flacmeta, rio
from .codec import make_dtiles, np_dtype_of, pcm_bps
from .converter import BLOCKSIZE, ENCODER_TAG, get_codec, stream_header
from .rio import CRS, Affine, Window

MAX_CHANNELS = 8  # FLAC__MAX_CHANNELS (docs/sonos-pyflac.txt:3501)


class SpatialFrame:
    """spatial_encoder.py:21-45"""

    def __init__(self, frame_id: int, bbox: Tuple[float, float, float, float], window: Window, byte_offset: int = 0,
                 byte_size: int = 0, streams: Optional[List[Dict]] = None):
        self.frame_id = frame_id
        self.bbox = bbox  # (xmin, ymin, xmax, ymax)
        self.window = window
        self.byte_offset = byte_offset
        self.byte_size = byte_size
        self.streams = streams  # extension D2: [{"bands":[first,count],"byte_offset":..,"byte_size":..}]

    def to_dict(self) -> Dict:
        d = {"frame_id": self.frame_id, "bbox": self.bbox,
             "window": {"row_off": self.window.row_off, "col_off": self.window.col_off, "height": self.window.height,
                        "width": self.window.width},
             "byte_offset": self.byte_offset, "byte_size": self.byte_size}
        if self.streams:
            d["streams"] = self.streams
        return d


class SpatialIndex:
    """spatial_encoder.py:48-78"""

    def __init__(self, frames: List[SpatialFrame], crs, transform):
        self.frames = frames
        self.crs = crs
        self.transform = transform
        self.total_bytes = sum(frame.byte_size for frame in frames)

    def query_bbox(self, bbox) -> List[SpatialFrame]:
        """Strict-inequality intersection, linear scan, index order (spatial_encoder.py:57-70)."""
        xmin, ymin, xmax, ymax = bbox
        return [f for f in self.frames
                if xmin < f.bbox[2] and xmax > f.bbox[0] and ymin < f.bbox[3] and ymax > f.bbox[1]]

    def query_bbox_batch(self, bboxes: np.ndarray) -> List[np.ndarray]:
        """Vectorised form for many queries: list of frame-position arrays, same order as query_bbox."""
        fb = np.asarray([f.bbox for f in self.frames], dtype=np.float64).reshape(-1, 4)
        q = np.asarray(bboxes, dtype=np.float64).reshape(-1, 4)
        hit = ((q[:, None, 0] < fb[None, :, 2]) & (q[:, None, 2] > fb[None, :, 0]) &
               (q[:, None, 1] < fb[None, :, 3]) & (q[:, None, 3] > fb[None, :, 1]))
        return [np.nonzero(r)[0] for r in hit]

    def to_dict(self) -> Dict:
        return {"crs": str(self.crs), "transform": list(self.transform), "frames": [f.to_dict() for f in self.frames]}


def calculate_tiles(height: int, width: int, tile_size: int):
    """spatial_encoder.py:92-103: row-major (row_off, col_off, h, w), edge tiles clipped."""
    return [(r, c, min(tile_size, height - r), min(tile_size, width - c))
            for r in range(0, height, tile_size) for c in range(0, width, tile_size)]


def tile_to_bbox(row_off, col_off, height, width, transform):
    """spatial_encoder.py:105-112"""
    xmin, ymax = transform * (col_off, row_off)
    xmax, ymin = transform * (col_off + width, row_off + height)
    return (xmin, ymin, xmax, ymax)


def band_groups(count: int):
    return [(b, min(MAX_CHANNELS, count - b)) for b in range(0, count, MAX_CHANNELS)]


class SpatialFLACEncoder:
    """Drop-in for spatial_encoder.SpatialFLACEncoder (:81)."""

    def __init__(self, tile_size: int = 512, device: int = 0, pcm_mode: str = "lossless"):
        self.tile_size = tile_size
        self.logger = logging.getLogger("flac_raster.spatial_encoder")
        self.frames: List[SpatialFrame] = []
        self.device = device
        if pcm_mode not in ("lossless", "reference_spatial"):
            # converter.py's min/max stretch ("reference") would be applied per tile here, while the file keeps one global
            # GEOSPATIAL_DATA_MIN/MAX: integer rasters would come back wrong.  The reference's --spatial path never uses it.
            raise ValueError(f"pcm_mode {pcm_mode!r} is not available for spatial files: use 'lossless' (default) or "
                             "'reference_spatial' (the reference's own --spatial mapping)")
        self.pcm_mode = pcm_mode

    def _calculate_tiles(self, height, width):
        tiles = calculate_tiles(height, width, self.tile_size)
        self.logger.info(f"Created {len(tiles)} tiles of size {self.tile_size}x{self.tile_size}")
        return tiles

    _tile_to_bbox = staticmethod(tile_to_bbox)

    def _calculate_audio_params(self, audio_data, dtype):
        """spatial_encoder.py:250-262: always 44 100 Hz; label 16 for <=16-bit ints else 24."""
        dtype = np.dtype(dtype)
        return 44100, (16 if dtype in (np.dtype(np.uint8), np.dtype(np.uint16), np.dtype(np.int16)) else 24)

    def encode_array(self, raster: np.ndarray, compression_level: int = 5):
        """(bands,h,w) host array -> (tiles, groups, EncodeResult).  Stream order: tile-major, band group minor."""
        count, height, width = raster.shape
        tiles = self._calculate_tiles(height, width)
        groups = band_groups(count)
        if self.pcm_mode == "reference_spatial" and len(groups) > 1:
            # the reference hands all bands to one libFLAC encoder, which refuses > 8 channels
            from ._lib import E_UNSUPPORTED, FrxUnsupported
            raise FrxUnsupported(E_UNSUPPORTED, f"{count} channels: FLAC allows at most {MAX_CHANNELS} per stream")
        units = [(r, c, h, w, b0, nb) for (r, c, h, w) in tiles for (b0, nb) in groups]
        if isinstance(raster, np.ndarray):
            raster = np.ascontiguousarray(raster)
        res = get_codec(self.device).encode_tiles_any(raster, units, level=compression_level, blocksize=BLOCKSIZE,
                                                      sample_rate=44100, pcm_mode=self.pcm_mode)
        return tiles, groups, res

    def encode_spatial_flac(self, tiff_path: Path, flac_path: Path, compression_level: int = 5,
                            enable_streaming: bool = True) -> SpatialIndex:
        """spatial_encoder.py:136-227"""
        tiff_path, flac_path = Path(tiff_path), Path(flac_path)
        self.logger.info(f"Starting spatial FLAC encoding: {tiff_path} -> {flac_path}")
        with rio.open(tiff_path) as src:
            raster = get_codec(self.device).read_tiff(src)    # file bytes -> device -> frx_tiff_unpack
            transform, crs = src.transform, src.crs
        count, height, width = raster.shape
        tiles, groups, res = self.encode_array(raster, compression_level)
        blob = res.bytes[:int(res.total_bytes[0])]
        bps = res.bps
        ng = len(groups)
        # per-stream byte strings: minimal header (STREAMINFO + frame table) + frames
        streams, fpos = [], 0
        for u in range(len(tiles) * ng):
            r, c, h, w = tiles[u // ng]
            nb = groups[u % ng][1]
            nf = int(res.frames_per_tile[u])
            fs = res.frame_size[fpos:fpos + nf]
            fpos += nf
            off, size = int(res.tile_offset[u]), int(res.tile_size[u])
            streams.append((stream_header(nb, bps, 44100, h * w, fs, None), blob[off:off + size]))
        sizes = [len(hd) + len(body) for hd, body in streams]
        dmin, dmax = float(res.tile_minmax[:, 0].min()), float(res.tile_minmax[:, 1].max())

        def build(first_header_len: int):
            frames, pos = [], 0
            for i, (r, c, h, w) in enumerate(tiles):
                tsz = sum(sizes[i * ng:(i + 1) * ng]) + (first_header_len if i == 0 else 0)
                sl = None
                if ng > 1:
                    sl, p = [], pos + (first_header_len if i == 0 else 0)
                    for g, (b0, nb) in enumerate(groups):
                        extra = first_header_len if (i == 0 and g == 0) else 0
                        sl.append({"bands": [b0, nb], "byte_offset": p - extra, "byte_size": sizes[i * ng + g] + extra})
                        p += sizes[i * ng + g]
                frames.append(SpatialFrame(i, tile_to_bbox(r, c, h, w, transform), Window(c, r, w, h), pos, tsz, sl))
                pos += tsz
            return frames

        # The first stream's header carries the comments + index, whose length depends on the offsets it
        # lists: reserve a fixed size, absorb the slack with a PADDING block (the reference's files carry one too).
        u0 = tiles[0]
        fs0 = res.frame_size[:int(res.frames_per_tile[0])]
        base_len = len(streams[0][0])
        reserve = 0
        for _ in range(8):
            frames = build(reserve)
            index = SpatialIndex(frames, crs, transform)
            comments = self._comments(index, crs, transform, height, width, count, np_dtype_of(raster), dmin, dmax, len(tiles))
            h0 = stream_header(groups[0][1], bps, 44100, u0[2] * u0[3], fs0, comments)
            need = len(h0) - base_len
            if reserve >= need and (reserve == need or reserve - need >= 4):
                pad = reserve - need
                if pad:
                    h0 = stream_header(groups[0][1], bps, 44100, u0[2] * u0[3], fs0, comments, padding=pad - 4)
                break
            reserve = ((need + 4 + 255) // 256) * 256 + 256
        else:  # pragma: no cover
            raise RuntimeError("could not settle the spatial header size")
        assert len(h0) == base_len + reserve
        with open(flac_path, "wb") as f:
            f.write(h0)
            f.write(streams[0][1].tobytes())
            for hd, body in streams[1:]:
                f.write(hd)
                f.write(body.tobytes())
        self.frames = frames
        self.logger.info(f"Encoded {len(tiles)} spatial tiles to FLAC: {flac_path}")
        return index

    def _comments(self, index, crs, transform, height, width, count, dtype, dmin, dmax, n_tiles):
        """spatial_encoder.py:313-347: same keys and order."""
        bounds = [index.frames[0].bbox[0], index.frames[0].bbox[1], index.frames[-1].bbox[2], index.frames[-1].bbox[3]]
        packed = base64.b64encode(gzip.compress(json.dumps(index.to_dict(), separators=(",", ":")).encode("utf-8"),
                                                mtime=0)).decode("ascii")
        c = [("TITLE", "Geospatial Raster Data"),
             ("DESCRIPTION", f"TIFF raster converted to spatial FLAC with {n_tiles} tiles"),
             ("ENCODER", ENCODER_TAG), ("DATE", str(np.datetime64("now", "D"))),
             ("GEOSPATIAL_CRS", str(crs)), ("GEOSPATIAL_WIDTH", str(width)), ("GEOSPATIAL_HEIGHT", str(height)),
             ("GEOSPATIAL_COUNT", str(count)), ("GEOSPATIAL_DTYPE", str(np.dtype(dtype))),
             ("GEOSPATIAL_DATA_MIN", str(dmin)), ("GEOSPATIAL_DATA_MAX", str(dmax)),
             ("GEOSPATIAL_TRANSFORM", json.dumps(list(transform))), ("GEOSPATIAL_BOUNDS", json.dumps(bounds)),
             ("GEOSPATIAL_SPATIAL_TILING", "true"), ("GEOSPATIAL_TILE_SIZE", str(self.tile_size)),
             ("GEOSPATIAL_NUM_TILES", str(n_tiles)), ("GEOSPATIAL_SPATIAL_INDEX", packed)]
        if self.pcm_mode != "reference":
            c.append(("GEOSPATIAL_PCM_MODE", self.pcm_mode))
        return c


def index_from_dict(data: Dict) -> SpatialIndex:
    """spatial_encoder.py:441-462"""
    frames = []
    for fd in data["frames"]:
        wd = fd["window"]
        frames.append(SpatialFrame(fd["frame_id"], tuple(fd["bbox"]), Window(wd["col_off"], wd["row_off"], wd["width"], wd["height"]),
                                   fd["byte_offset"], fd["byte_size"], fd.get("streams")))
    return SpatialIndex(frames, CRS.from_string(data["crs"]), Affine(*data["transform"][:6]))


class SpatialFLACStreamer:
    """Drop-in for spatial_encoder.SpatialFLACStreamer (:376): bbox -> byte ranges -> bytes."""

    def __init__(self, flac_path):
        self.flac_path = flac_path
        self.is_url = isinstance(flac_path, str) and flac_path.startswith(("http://", "https://"))
        self.logger = logging.getLogger("flac_raster.spatial_streamer")
        self.header: Optional[flacmeta.FlacHeader] = None
        self.spatial_index = self._load_spatial_index()

    def _read_head(self) -> bytes:
        if self.is_url:
            import requests
            r = requests.get(self.flac_path, headers={"Range": "bytes=0-1048575"}, stream=True)  # spatial_encoder.py:398
            r.raise_for_status()
            return r.content
        with open(self.flac_path, "rb") as f:
            return f.read(1 << 24)

    def _load_spatial_index(self) -> SpatialIndex:
        """spatial_encoder.py:385-462: embedded GEOSPATIAL_SPATIAL_INDEX first, `.spatial.json` sidecar second."""
        try:
            self.header = flacmeta.parse_header(self._read_head())
            enc = self.header.get("GEOSPATIAL_SPATIAL_INDEX")
            if enc is None:
                raise ValueError("No embedded spatial index found")
            data = json.loads(gzip.decompress(base64.b64decode(enc.encode("ascii"))).decode("utf-8"))
        except Exception as e:  # noqa: BLE001
            self.logger.warning(f"Failed to read embedded metadata: {e}")
            index_path = Path(self.flac_path).with_suffix(".spatial.json")
            if not index_path.exists():
                raise FileNotFoundError(f"Spatial index not found in FLAC metadata or sidecar file: {index_path}")
            with open(index_path) as f:
                data = json.load(f)
        return index_from_dict(data)

    def get_byte_ranges_for_bbox(self, bbox) -> List[Tuple[int, int]]:
        """spatial_encoder.py:464-484: inclusive (start,end), sorted, merged when start <= prev_end + 1."""
        ranges = sorted((f.byte_offset, f.byte_offset + f.byte_size - 1) for f in self.spatial_index.query_bbox(bbox)
                        if f.byte_size > 0)
        merged: List[Tuple[int, int]] = []
        for start, end in ranges:
            if merged and start <= merged[-1][1] + 1:
                merged[-1] = (merged[-1][0], max(merged[-1][1], end))
            else:
                merged.append((start, end))
        self.logger.info(f"Found {len(merged)} byte ranges for bbox {bbox}")
        return merged

    def _fetch(self, start: int, end: int) -> bytes:
        if self.is_url:
            import requests
            r = requests.get(self.flac_path, headers={"Range": f"bytes={start}-{end}"})
            r.raise_for_status()
            return r.content
        with open(self.flac_path, "rb") as f:
            f.seek(start)
            return f.read(end - start + 1)

    def stream_bbox_data(self, bbox) -> bytes:
        """spatial_encoder.py:486-506"""
        return b"".join(self._fetch(s, e) for s, e in self.get_byte_ranges_for_bbox(bbox))

    # ---- extension: decode the intersecting tiles in one batched GPU call -----------
    def decode_frames(self, frames: List[SpatialFrame], device: int = 0):
        """[(frame, (count,h,w) array)] for tiles of a file written by SpatialFLACEncoder."""
        if self.header is None:
            raise ValueError("tile decode needs the embedded metadata of a libfrx-written spatial file")
        hd = self.header
        count, dtype = int(hd.get("GEOSPATIAL_COUNT")), np.dtype(hd.get("GEOSPATIAL_DTYPE"))
        mode = hd.get("GEOSPATIAL_PCM_MODE", "reference_spatial")
        dmin, dmax = float(hd.get("GEOSPATIAL_DATA_MIN", "0")), float(hd.get("GEOSPATIAL_DATA_MAX", "0"))
        chunks, entries, tables, pos, ff, oe = [], [], [], 0, 0, 0
        for fr in frames:
            h, w = int(fr.window.height), int(fr.window.width)
            for st in (fr.streams or [{"bands": [0, count], "byte_offset": fr.byte_offset, "byte_size": fr.byte_size}]):
                raw = self._fetch(st["byte_offset"], st["byte_offset"] + st["byte_size"] - 1)
                sh = flacmeta.parse_header(raw)
                body = raw[sh.first_frame_offset:]
                bs = sh.streaminfo.max_blocksize or BLOCKSIZE
                nfr = (h * w + bs - 1) // bs
                first = None
                app = sh.applications.get(flacmeta.FRAME_TABLE_APP_ID)
                if app is not None:
                    tbs, sizes = flacmeta.unpack_frame_table(app)
                    if tbs == bs and len(sizes) == nfr:
                        tables.extend(sizes)
                        first, ff = ff, ff + nfr
                entries.append(dict(byte_offset=pos, byte_size=len(body), out_offset=oe + st["bands"][0] * h * w,
                                    out_band_stride=h * w, out_row_stride=w, h=h, w=w, channels=st["bands"][1],
                                    bps=sh.streaminfo.bps, dmin=dmin, dmax=dmax, first_frame=first, n_frames=nfr))
                chunks.append(body)
                pos += len(body)
            oe += count * h * w
        out = np.empty(oe, dtype=dtype)
        data = np.frombuffer(b"".join(chunks), dtype=np.uint8)
        get_codec(device).decode_tiles_host(data, make_dtiles(entries), out,
                                            frame_size=np.asarray(tables, dtype=np.uint32) if tables else None,
                                            blocksize=BLOCKSIZE, pcm_mode=mode)
        res, p = [], 0
        for fr in frames:
            n = count * int(fr.window.height) * int(fr.window.width)
            res.append((fr, out[p:p + n].reshape(count, int(fr.window.height), int(fr.window.width))))
            p += n
        return res

    def decode_bbox(self, bbox, device: int = 0):
        return self.decode_frames(self.spatial_index.query_bbox(bbox), device)


def decode_spatial_file(flac_path, device: int = 0) -> np.ndarray:
    """Whole raster (count,H,W) back from a spatial file: every tile in one batched decode."""
    s = SpatialFLACStreamer(Path(flac_path))
    hd = s.header
    count, H, W = int(hd.get("GEOSPATIAL_COUNT")), int(hd.get("GEOSPATIAL_HEIGHT")), int(hd.get("GEOSPATIAL_WIDTH"))
    out = np.empty((count, H, W), dtype=np.dtype(hd.get("GEOSPATIAL_DTYPE")))
    for fr, arr in s.decode_frames(s.spatial_index.frames, device):
        wd = fr.window
        out[:, wd.row_off:wd.row_off + wd.height, wd.col_off:wd.col_off + wd.width] = arr
    return out

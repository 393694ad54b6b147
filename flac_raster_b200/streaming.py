"""Streaming format: ``[u32 BE index length][JSON index][complete FLAC tile]...``
(reference: cli.py:620-804 create-streaming, cli.py:875-1039 extract-streaming).

The reference builds it with a serial loop -- window read, temp GeoTIFF, RasterFLACConverter.tiff_to_flac
(level 5), temp FLAC read back -- once per tile.  Here every tile of the raster (or of this rank's
contiguous tile range) is encoded by ONE libfrx call; the host only composes the per-tile stream
headers and the JSON index around the bytes the GPU produced.

Multi-GPU: tiles are independent, so rank r encodes the contiguous tile range
``shard_range(n_tiles, world, r)`` and the ranks' (bytes, sizes) are concatenated on the host in
rank order; byte offsets are an exclusive prefix sum over all tile sizes.  No device collective.
"""
from __future__ import annotations

import json
import struct
from pathlib import Path
from typing import Dict, Iterable, List, Optional, Sequence, Tuple

import numpy as np

from . import flacmeta, rio
from .codec import make_dtiles, np_dtype_of, pcm_bps
from .converter import (BLOCKSIZE, decode_stream, get_codec, metadata_comments, metadata_from_comments, raster_metadata,
                        sample_rate_for, stream_header)
from .geotiff import BoundingBox
from .spatial_encoder import MAX_CHANNELS, calculate_tiles

STREAMING_LEVEL = 5  # cli.py:733 hard-wires compression_level=5


# ---- tile sharding ---------------------------------------------------------------------------------
def shard_range(n_tiles: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous range of the row-major tile list owned by `rank`: [r*ceil(n/w), (r+1)*ceil(n/w))."""
    per = (n_tiles + world - 1) // world
    lo = min(n_tiles, rank * per)
    return lo, min(n_tiles, lo + per)


def merge_shards(shards: Sequence[Tuple[Sequence[bytes], Sequence[int]]]):
    """Host-side offset concatenation: shards = per-rank (tile byte strings, tile sizes) in rank order ->
    (all tile byte strings, global byte offsets, sizes).  The only cross-rank step of the path."""
    chunks: List[bytes] = []
    sizes: List[int] = []
    for blobs, sz in shards:
        chunks.extend(blobs)
        sizes.extend(int(s) for s in sz)
    offsets = np.concatenate([[0], np.cumsum(sizes)[:-1]]).astype(np.int64) if sizes else np.zeros(0, np.int64)
    return chunks, offsets, sizes


# ---- encode ----------------------------------------------------------------------------------------
def tile_window_transform(transform, col_off: int, row_off: int):
    return transform * rio.Affine.translation(col_off, row_off)


def tile_bbox(transform, col_off, row_off, tw, th):
    """cli.py:702-706"""
    tt = tile_window_transform(transform, col_off, row_off)
    xmin, ymax = tt.c, tt.f
    return [xmin, ymax + th * tt.e, xmin + tw * tt.a, ymax], tt


def encode_streaming_tiles(raster: np.ndarray, tiles, crs, transform, nodata=None, bands: str = "first",
                           pcm_mode: str = "reference", device: int = 0, row_base: int = 0, codec=None, md5: bool = False) -> Tuple[List[bytes], List[int]]:
    """Encode `tiles` [(row_off,col_off,h,w)] of the host raster (bands,H,W) into complete per-tile FLAC files.
    `row_base`: raster row 0 is row `row_base` of the full image (a rank holding only its own rows).
    Each tile is what RasterFLACConverter.tiff_to_flac(level 5) makes of the tile's own GeoTIFF
    (cli.py:711-737): per-tile min/max, count = encoded bands, tile transform and bounds."""
    count = raster.shape[0]
    nch = 1 if bands == "first" else count  # cli.py:699 reads band 1 only; "all" is extension D3
    if nch > MAX_CHANNELS:
        from ._lib import E_UNSUPPORTED, FrxUnsupported
        raise FrxUnsupported(E_UNSUPPORTED, f"{nch} channels: FLAC allows at most {MAX_CHANNELS} per stream")
    if not tiles:
        return [], []
    # the per-tile tiff_to_flac call sees a (nch, th, tw) array: sample rate keyed on nch*th (always 44100 for th<=1024)
    rates = {sample_rate_for(nch, t[2]) for t in tiles}
    blobs: List[Optional[bytes]] = [None] * len(tiles)
    dtype = np_dtype_of(raster)
    for sr in sorted(rates):
        sel = [i for i, t in enumerate(tiles) if sample_rate_for(nch, t[2]) == sr]
        units = [(tiles[i][0] - row_base, tiles[i][1], tiles[i][2], tiles[i][3], 0, nch) for i in sel]
        res = (codec or get_codec(device)).encode_tiles_any(raster, units, level=STREAMING_LEVEL, blocksize=BLOCKSIZE, sample_rate=sr,
                                                            pcm_mode=pcm_mode)
        data = res.bytes[:int(res.total_bytes[0])]
        sigs = (codec or get_codec(device)).tile_md5(raster, units, pcm_mode) if md5 else None
        fpos = 0
        for k, i in enumerate(sel):
            r, c, th, tw = tiles[i]
            nf = int(res.frames_per_tile[k])
            fs = res.frame_size[fpos:fpos + nf]
            fpos += nf
            bbox, tt = tile_bbox(transform, c, r, tw, th)
            md = raster_metadata(tw, th, nch, dtype, crs, tt, BoundingBox(bbox[0], bbox[1], bbox[2], bbox[3]),
                                 float(res.tile_minmax[k, 0]), float(res.tile_minmax[k, 1]), nodata)
            hdr = stream_header(nch, res.bps, sr, th * tw, fs, metadata_comments(md, pcm_mode), md5=None if sigs is None else sigs[k].tobytes())
            off, size = int(res.tile_offset[k]), int(res.tile_size[k])
            blobs[i] = hdr + data[off:off + size].tobytes()
    return blobs, [len(b) for b in blobs]


def build_index(crs, transform, width: int, height: int, tile_size: int, tiles, offsets, sizes) -> Dict:
    """cli.py:679-686, 748-759 (same key order)."""
    frames = []
    for i, (r, c, th, tw) in enumerate(tiles):
        bbox, _ = tile_bbox(transform, c, r, tw, th)
        frames.append({"frame_id": i, "bbox": bbox, "window": {"col_off": c, "row_off": r, "width": tw, "height": th},
                       "byte_offset": int(offsets[i]), "byte_size": int(sizes[i])})
    return {"crs": str(crs), "transform": list(transform), "width": width, "height": height, "tile_size": tile_size,
            "frames": frames}


def write_streaming_file(path, index: Dict, chunks: Iterable[bytes]) -> int:
    """cli.py:769-780"""
    index_json = json.dumps(index, separators=(",", ":")).encode("utf-8")
    with open(path, "wb") as f:
        f.write(len(index_json).to_bytes(4, "big"))
        f.write(index_json)
        for ch in chunks:
            f.write(ch)
    return len(index_json)


_make_codec = None   # tests substitute the oracle-backed fake here


def _new_codec(dev):
    from .codec import Codec
    return Codec(dev)


def _encode_shard(input_file, dev, own_codec, tiles, lo, hi, crs, transform, nodata, bands, pcm_mode, md5=False):
    """One device's part of create-streaming: its own file handle and (when several shards run at once) its own libfrx context;
    only the raster rows the tile range touches -- and, in band-1 mode, only band 1 -- are ingested."""
    if lo == hi:
        return [], []
    codec = (_make_codec or _new_codec)(dev) if own_codec else get_codec(dev)
    try:
        with rio.open(input_file) as src:
            r0 = min(t[0] for t in tiles[lo:hi])
            r1 = max(t[0] + t[2] for t in tiles[lo:hi])
            raster = codec.read_tiff(src, r0, r1, bands=[0] if bands == "first" else None)
        return encode_streaming_tiles(raster, tiles[lo:hi], crs, transform, nodata, bands, pcm_mode, dev, row_base=r0, codec=codec, md5=md5)
    finally:
        if own_codec:
            codec.close()


def create_streaming(input_file, output_file, tile_size: int = 1024, bands: str = "first", pcm_mode: str = "reference",
                     devices: Sequence[int] = (0,), md5: bool = False) -> Dict:
    """create-streaming (cli.py:620-804).  `devices`: GPUs of this process; contiguous tile ranges are sharded over them and
    encoded CONCURRENTLY, one thread and one libfrx context per device (the C-ABI allows one context per thread), then merged on
    the host by offset concatenation."""
    with rio.open(input_file) as src:
        crs, transform, nodata = src.crs, src.transform, getattr(src, "nodata", None)
        width, height = src.width, src.height
    tiles = calculate_tiles(height, width, tile_size)
    ranges = [shard_range(len(tiles), len(devices), k) for k in range(len(devices))]
    if len(devices) == 1:
        shards = [_encode_shard(input_file, devices[0], False, tiles, ranges[0][0], ranges[0][1], crs, transform, nodata, bands, pcm_mode, md5)]
    else:
        from concurrent.futures import ThreadPoolExecutor
        with ThreadPoolExecutor(max_workers=len(devices)) as ex:
            futs = [ex.submit(_encode_shard, input_file, dev, True, tiles, lo, hi, crs, transform, nodata, bands, pcm_mode, md5)
                    for dev, (lo, hi) in zip(devices, ranges)]
            shards = [f.result() for f in futs]
    chunks, offsets, sizes = merge_shards(shards)
    index = build_index(crs, transform, width, height, tile_size, tiles, offsets, sizes)
    write_streaming_file(output_file, index, chunks)
    return index


# ---- decode ----------------------------------------------------------------------------------------
def read_streaming_index(path_or_url) -> Tuple[int, Dict]:
    """cli.py:899-919 -> (index_size, index)."""
    s = str(path_or_url)
    if s.startswith(("http://", "https://")):
        import requests
        r = requests.get(s, headers={"Range": "bytes=0-3"})
        if r.status_code != 206:
            raise ValueError(f"Server doesn't support range requests: {r.status_code}")
        index_size = struct.unpack(">I", r.content)[0]
        r = requests.get(s, headers={"Range": f"bytes=4-{3 + index_size}"})
        if r.status_code != 206:
            raise ValueError(f"Failed to get spatial index: {r.status_code}")
        return index_size, json.loads(r.content.decode("utf-8"))
    with open(s, "rb") as f:
        index_size = struct.unpack(">I", f.read(4))[0]
        return index_size, json.loads(f.read(index_size).decode("utf-8"))


def select_frame(index: Dict, tile_id: Optional[int] = None, last: bool = False, center: bool = False,
                 bbox: Optional[Sequence[float]] = None) -> Optional[Dict]:
    """cli.py:926-990: by id | max id | nearest to the centre | FIRST intersecting frame (strict < >)."""
    frames = index["frames"]
    if tile_id is not None:
        for fr in frames:
            if fr["frame_id"] == tile_id:
                return fr
        return None
    if last:
        return max(frames, key=lambda f: f["frame_id"])
    if center:
        cx = (min(f["bbox"][0] for f in frames) + max(f["bbox"][2] for f in frames)) / 2
        cy = (min(f["bbox"][1] for f in frames) + max(f["bbox"][3] for f in frames)) / 2
        best, best_d = None, float("inf")
        for fr in frames:
            fx, fy = (fr["bbox"][0] + fr["bbox"][2]) / 2, (fr["bbox"][1] + fr["bbox"][3]) / 2
            d = ((fx - cx) ** 2 + (fy - cy) ** 2) ** 0.5
            if d < best_d:
                best, best_d = fr, d
        return best
    if bbox is not None:
        for fr in frames:
            fb = fr["bbox"]
            if bbox[0] < fb[2] and bbox[2] > fb[0] and bbox[1] < fb[3] and bbox[3] > fb[1]:
                return fr
        return None
    raise ValueError("Must specify tile_id, bbox, center or last")


def select_frames_batch(index: Dict, queries: Sequence[Dict]) -> np.ndarray:
    """Vectorised select_frame for many queries ({"tile_id":..} or {"bbox":[..]}): positions into index["frames"],
    -1 when nothing matches.  Same answers as the reference loop (first intersecting frame in index order)."""
    frames = index["frames"]
    ids = np.asarray([f["frame_id"] for f in frames])
    fb = np.asarray([f["bbox"] for f in frames], dtype=np.float64).reshape(-1, 4)
    pos_of_id = {int(v): i for i, v in reversed(list(enumerate(ids)))}
    out = np.full(len(queries), -1, dtype=np.int64)
    bq = [(i, q["bbox"]) for i, q in enumerate(queries) if q.get("tile_id") is None and q.get("bbox") is not None]
    for i, q in enumerate(queries):
        if q.get("tile_id") is not None:
            out[i] = pos_of_id.get(int(q["tile_id"]), -1)
    if bq:
        qi = np.asarray([i for i, _ in bq])
        qb = np.asarray([b for _, b in bq], dtype=np.float64)
        for s in range(0, len(qb), 4096):  # bounded (queries x frames) boolean blocks
            b = qb[s:s + 4096]
            hit = ((b[:, None, 0] < fb[None, :, 2]) & (b[:, None, 2] > fb[None, :, 0]) &
                   (b[:, None, 1] < fb[None, :, 3]) & (b[:, None, 3] > fb[None, :, 1]))
            first = hit.argmax(axis=1)
            out[qi[s:s + 4096]] = np.where(hit.any(axis=1), first, -1)
    return out


def fetch_tile(path_or_url, index_size: int, frame: Dict) -> bytes:
    """cli.py:997-1011"""
    start = 4 + index_size + frame["byte_offset"]
    s = str(path_or_url)
    if s.startswith(("http://", "https://")):
        import requests
        r = requests.get(s, headers={"Range": f"bytes={start}-{start + frame['byte_size'] - 1}"})
        if r.status_code != 206:
            raise ValueError(f"Failed to download tile: {r.status_code}")
        return r.content
    with open(s, "rb") as f:
        f.seek(start)
        return f.read(frame["byte_size"])


def tile_metadata(tile_bytes: bytes) -> Tuple[flacmeta.FlacHeader, Dict]:
    hdr = flacmeta.parse_header(tile_bytes[:1 << 20] if len(tile_bytes) > (1 << 20) else tile_bytes)
    md = metadata_from_comments(hdr)
    if md is None:
        raise ValueError("No metadata found in FLAC tile")
    return hdr, md


def write_tile_tiff(raster: np.ndarray, md: Dict, output) -> None:
    meta = {"driver": "GTiff", "width": md["width"], "height": md["height"], "count": md["count"],
            "dtype": np.dtype(md["dtype"]), "nodata": md.get("nodata")}
    if md.get("crs"):
        meta["crs"] = rio.CRS.from_string(md["crs"])
    if md.get("transform"):
        t = md["transform"]
        meta["transform"] = rio.Affine(t[0], t[1], t[2], t[3], t[4], t[5])
    with rio.open(output, "w", **meta) as dst:
        if md["count"] == 1:
            dst.write(raster[0], 1)
        else:
            dst.write(raster)


def extract_streaming(path_or_url, output, tile_id=None, bbox=None, center=False, last=False, device: int = 0) -> Dict:
    """extract-streaming (cli.py:875-1039): one query -> one tile -> GeoTIFF.  Returns the chosen frame."""
    index_size, index = read_streaming_index(path_or_url)
    frame = select_frame(index, tile_id, last, center, bbox)
    if frame is None:
        raise LookupError("no tile matches the query")
    data = fetch_tile(path_or_url, index_size, frame)
    _, md = tile_metadata(data)
    write_tile_tiff(decode_stream(np.frombuffer(data, dtype=np.uint8), md, device), md, output)
    return frame


class StreamingFile:
    """A streaming file opened for many queries (config C5): index parsed once, tile headers parsed once per
    distinct tile, compressed bytes staged once; each batch of queries is ONE libfrx decode call."""

    def __init__(self, path, device: int = 0):
        self.path = str(path)
        self.device = device
        self.index_size, self.index = read_streaming_index(self.path)
        self.data = np.fromfile(self.path, dtype=np.uint8)
        self.base = 4 + self.index_size
        self._tiles: Dict[int, Dict] = {}

    def _tile(self, pos: int) -> Dict:
        t = self._tiles.get(pos)
        if t is None:
            fr = self.index["frames"][pos]
            start = self.base + fr["byte_offset"]
            raw = self.data[start:start + fr["byte_size"]]
            hdr, md = tile_metadata(raw[:1 << 20].tobytes())
            bs = hdr.streaminfo.max_blocksize or BLOCKSIZE
            h, w = md["height"], md["width"]
            nfr = (h * w + bs - 1) // bs
            sizes = None
            app = hdr.applications.get(flacmeta.FRAME_TABLE_APP_ID)
            if app is not None:
                tbs, sz = flacmeta.unpack_frame_table(app)
                if tbs == bs and len(sz) == nfr:
                    sizes = np.asarray(sz, dtype=np.uint32)
            t = self._tiles[pos] = dict(frame=fr, md=md, bps=hdr.streaminfo.bps, channels=hdr.streaminfo.channels, blocksize=bs,
                                        frames_start=start + hdr.first_frame_offset, frames_bytes=fr["byte_size"] - hdr.first_frame_offset,
                                        n_frames=nfr, sizes=sizes)
        return t

    def plan(self, positions: Sequence[int]):
        """DTile list for a batch of frame positions (duplicates allowed): byte ranges point into the whole
        file image (`self.data`), outputs are packed back to back.  -> (dtiles, frame_size table, out elems, metas)"""
        entries, tables, metas, ff, oe = [], [], [], 0, 0
        table_of: Dict[int, int] = {}
        for p in positions:
            t = self._tile(int(p))
            md = t["md"]
            first = None
            if t["sizes"] is not None:
                if int(p) not in table_of:
                    table_of[int(p)] = ff
                    tables.append(t["sizes"])
                    ff += t["n_frames"]
                first = table_of[int(p)]
            h, w, c = md["height"], md["width"], md["count"]
            entries.append(dict(byte_offset=t["frames_start"], byte_size=t["frames_bytes"], out_offset=oe, out_band_stride=h * w,
                                out_row_stride=w, h=h, w=w, channels=c, bps=t["bps"], dmin=md["data_min"], dmax=md["data_max"],
                                first_frame=first, n_frames=t["n_frames"]))
            metas.append(md)
            oe += c * h * w
        table = np.concatenate(tables).astype(np.uint32) if tables else None
        return make_dtiles(entries), table, oe, metas

    def device_image(self):
        """The file image as a CUDA uint8 tensor, uploaded once (torch is device-memory plumbing only)."""
        if getattr(self, "_dev", None) is None:
            import torch
            self._dev = torch.from_numpy(self.data).to(f"cuda:{self.device}")
        return self._dev

    def decode_device(self, positions: Sequence[int], out=None):
        """Batched decode with the file image resident in HBM: -> (flat CUDA tensor of all tiles back to back,
        metas, status tensor).  This is the timed body of the C5 bench."""
        import torch
        from .codec import _tdtype
        dt, table, oe, metas = self.plan(positions)
        dev = self.device_image()
        if out is None:
            out = torch.empty(oe, dtype=_tdtype(np.dtype(metas[0]["dtype"]).name), device=dev.device)
        tab = None
        if table is not None:
            tab = torch.from_numpy(table.view(np.int32)).to(dev.device)
        status = get_codec(self.device).decode_tiles(dev, dt, out, out_dtype=np.dtype(metas[0]["dtype"]).name, frame_size=tab,
                                                     blocksize=BLOCKSIZE, pcm_mode=metas[0].get("pcm_mode", "reference"))
        return out, metas, status

    def decode(self, positions: Sequence[int]) -> List[np.ndarray]:
        """Batched decode -> one (count,h,w) host array per query."""
        if len(positions) == 0:
            return []
        from . import _lib as L
        from .synth import to_numpy
        out, metas, status = self.decode_device(positions)
        st = status.cpu().numpy()
        if st.any():
            raise L.FrxError(int(st[st != 0][0]), f"decode failed for {int((st != 0).sum())} tile(s)")
        flat = to_numpy(out)
        res, p = [], 0
        for md in metas:
            n = md["count"] * md["height"] * md["width"]
            res.append(flat[p:p + n].reshape(md["count"], md["height"], md["width"]))
            p += n
        return res

    def query(self, queries: Sequence[Dict]) -> Tuple[np.ndarray, List[Optional[np.ndarray]]]:
        pos = select_frames_batch(self.index, queries)
        ok = pos >= 0
        arrays = iter(self.decode(pos[ok]))
        return pos, [next(arrays) if k else None for k in ok]

"""RasterFLACConverter -- same entry points as the reference's converter (converter.py:17-437),
with the numpy staging + pyFLAC/libFLAC codec replaced by one libfrx call per direction.

    tiff_to_flac : converter.py:112-232  ->  Codec.encode_tiles_host (K0..K8)
    flac_to_tiff : converter.py:234-313  ->  Codec.decode_tiles_host (K9..K12)

The stream header (fLaC, STREAMINFO, VORBIS_COMMENT with the GEOSPATIAL_* keys of
converter.py:315-373) is composed on the host in front of the frames the GPU produced, so
nothing is rewritten afterwards (the reference has mutagen rewrite the file).
"""
from __future__ import annotations

import json
import logging
from pathlib import Path
from typing import Dict, Optional, Tuple

import numpy as np

from . import _lib as L
from . import flacmeta, rio
from .codec import Codec, make_dtiles, np_dtype_of, pcm_bps

BLOCKSIZE = 4096  # converter.py:205
ENCODER_TAG = "FLAC-Raster v0.1.0"  # converter.py:330 (kept so reference readers see the same key set)

_codecs: Dict[int, Codec] = {}


def get_codec(device: int = 0) -> Codec:
    """One libfrx context per device, created on first use (raises when there is no GPU)."""
    c = _codecs.get(device)
    if c is None:
        c = _codecs[device] = Codec(device)
    return c


def sample_rate_for(bands: int, height: int) -> int:
    """converter.py:41-51: keyed on shape[0]*shape[1] of the (bands,h,w) array."""
    total = bands * height
    if total < 1_000_000:
        return 44100
    if total < 10_000_000:
        return 48000
    if total < 100_000_000:
        return 96000
    return 192000


def bits_label_for(dtype) -> int:
    """converter.py:29-37: the label the reference prints (16 or 24); the effective FLAC sample
    size is pcm_bps() (pyFLAC overrides it from the array itemsize)."""
    dtype = np.dtype(dtype)
    return 16 if dtype in (np.dtype(np.uint8), np.dtype(np.uint16), np.dtype(np.int16)) else 24


def metadata_comments(metadata: Dict, pcm_mode: str = "reference"):
    """The VORBIS_COMMENT list of converter.py:327-348, same keys, order and string forms."""
    c = [("TITLE", "Geospatial Raster Data"),
         ("DESCRIPTION", "TIFF raster converted to FLAC with geospatial metadata"),
         ("ENCODER", ENCODER_TAG),
         ("GEOSPATIAL_CRS", str(metadata.get("crs", ""))),
         ("GEOSPATIAL_WIDTH", str(metadata.get("width", 0))),
         ("GEOSPATIAL_HEIGHT", str(metadata.get("height", 0))),
         ("GEOSPATIAL_COUNT", str(metadata.get("count", 1))),
         ("GEOSPATIAL_DTYPE", str(metadata.get("dtype", ""))),
         ("GEOSPATIAL_NODATA", str(metadata.get("nodata", ""))),
         ("GEOSPATIAL_DATA_MIN", str(metadata.get("data_min", ""))),
         ("GEOSPATIAL_DATA_MAX", str(metadata.get("data_max", ""))),
         ("GEOSPATIAL_TRANSFORM", json.dumps(metadata.get("transform", []))),
         ("GEOSPATIAL_BOUNDS", json.dumps(metadata.get("bounds", []))),
         ("GEOSPATIAL_SPATIAL_TILING", str(metadata.get("spatial_tiling", False)))]
    if pcm_mode != "reference":
        c.append(("GEOSPATIAL_PCM_MODE", pcm_mode))  # additive key; reference readers ignore it
    return c


def metadata_from_comments(hdr: flacmeta.FlacHeader) -> Optional[Dict]:
    """converter.py:375-416: requires GEOSPATIAL_CRS, converts the typed fields."""
    if "GEOSPATIAL_CRS" not in hdr:
        return None
    md: Dict = {}
    for field in ("CRS", "WIDTH", "HEIGHT", "COUNT", "DTYPE", "NODATA", "DATA_MIN", "DATA_MAX", "TRANSFORM", "BOUNDS",
                  "SPATIAL_TILING", "PCM_MODE"):
        value = hdr.get("GEOSPATIAL_" + field)
        if value is None:
            continue
        key = field.lower()
        if key in ("width", "height", "count"):
            md[key] = int(value) if value else 0
        elif key in ("data_min", "data_max"):
            md[key] = float(value) if value else 0.0
        elif key in ("transform", "bounds"):
            md[key] = json.loads(value) if value else []
        elif key == "spatial_tiling":
            md[key] = value.lower() == "true"
        elif key == "nodata":
            md[key] = None if value == "None" else float(value) if value else None
        else:
            md[key] = value
    return md


def raster_metadata(width, height, count, dtype, crs, transform, bounds, data_min, data_max, nodata=None,
                    driver="GTiff") -> Dict:
    """converter.py:157-174"""
    return {"width": int(width), "height": int(height), "count": int(count), "dtype": str(dtype),
            "crs": (crs.to_string() if hasattr(crs, "to_string") else str(crs)) if crs else None,
            "transform": list(transform) if transform is not None else None,
            "bounds": {"left": bounds.left, "bottom": bounds.bottom, "right": bounds.right, "top": bounds.top},
            "data_min": data_min, "data_max": data_max, "nodata": nodata, "driver": driver}


def stream_header(channels: int, bps: int, sample_rate: int, total_samples: int, frame_sizes, comments,
                  blocksize: int = BLOCKSIZE, frame_table: bool = True, padding=None, md5: Optional[bytes] = None) -> bytes:
    """fLaC + STREAMINFO (finalised: total samples and min/max frame size, which the reference leaves 0,
    SURVEY 0.10; the MD5 signature of the PCM when the caller computed one, Codec.tile_md5) + optional frxT frame-size
    table + VORBIS_COMMENT."""
    fs = np.asarray(frame_sizes)
    si = flacmeta.StreamInfo(min_blocksize=blocksize, max_blocksize=blocksize,
                             min_framesize=int(fs.min()) if fs.size else 0, max_framesize=int(fs.max()) if fs.size else 0,
                             sample_rate=sample_rate, channels=channels, bps=bps, total_samples=int(total_samples),
                             md5=bytes(md5) if md5 is not None else b"\0" * 16)
    apps = {flacmeta.FRAME_TABLE_APP_ID: flacmeta.pack_frame_table(blocksize, fs)} if frame_table else None
    return flacmeta.build_header(si, comments, padding=padding, applications=apps)


def encode_array(data, compression_level: int = 5, sample_rate: Optional[int] = None,
                 pcm_mode: str = "reference", device: int = 0):
    """(bands,h,w) host array or CUDA tensor (Codec.read_tiff) -> (frame bytes, frame_sizes, data_min, data_max, bps): the hot-path part of
    tiff_to_flac (converter.py:152-153, 180-216) as ONE libfrx call (whole raster = one stream)."""
    if data.ndim == 2:
        data = data[None]
    bands, h, w = data.shape
    codec = get_codec(device)
    sr = sample_rate or sample_rate_for(bands, h)
    if isinstance(data, np.ndarray):
        data = np.ascontiguousarray(data)
    res = codec.encode_tiles_any(data, [(0, 0, h, w, 0, bands)], level=compression_level, blocksize=BLOCKSIZE, sample_rate=sr,
                                 pcm_mode=pcm_mode)
    total = int(res.total_bytes[0])
    return (res.bytes[:total], res.frame_size.astype(np.int64), float(res.tile_minmax[0, 0]), float(res.tile_minmax[0, 1]),
            res.bps)


def decode_stream(blob, metadata: Dict, device: int = 0, offset: int = 0) -> np.ndarray:
    """Complete FLAC stream (header + frames) at blob[offset:] + its raster metadata -> (count,h,w) array.
    pyflac.FileDecoder.process + reshape + _denormalize_from_audio (converter.py:241-282) in one libfrx call."""
    blob = np.frombuffer(blob, dtype=np.uint8) if not isinstance(blob, np.ndarray) else blob
    hdr = flacmeta.parse_header(blob[offset:offset + min(len(blob) - offset, 1 << 24)].tobytes())
    si = hdr.streaminfo
    h, w, count = int(metadata["height"]), int(metadata["width"]), int(metadata["count"])
    if si.channels != count:
        raise ValueError(f"stream has {si.channels} channels but metadata says {count} band(s)")
    dtype = np.dtype(metadata["dtype"])
    if dtype.name not in L.DTYPE_CODE:
        raise L.FrxUnsupported(L.E_UNSUPPORTED, f"raster dtype {dtype} is not supported by libfrx")
    blocksize = si.max_blocksize or BLOCKSIZE
    n_frames = (h * w + blocksize - 1) // blocksize
    frames = blob[offset + hdr.first_frame_offset:]
    table = None
    first_frame = None
    app = hdr.applications.get(flacmeta.FRAME_TABLE_APP_ID)
    if app is not None:
        bs, sizes = flacmeta.unpack_frame_table(app)
        if bs == blocksize and len(sizes) == n_frames:
            table, first_frame = np.asarray(sizes, dtype=np.uint32), 0
    nbytes = int(table.sum()) if table is not None else len(frames)
    out = np.empty((count, h, w), dtype=dtype)
    dt = make_dtiles([dict(byte_offset=0, byte_size=nbytes, out_offset=0, out_band_stride=h * w, out_row_stride=w, h=h, w=w,
                           channels=count, bps=si.bps, dmin=metadata["data_min"], dmax=metadata["data_max"],
                           first_frame=first_frame, n_frames=n_frames)])
    get_codec(device).decode_tiles_host(frames[:nbytes], dt, out, frame_size=table, blocksize=blocksize,
                                        pcm_mode=metadata.get("pcm_mode", "reference"))
    return out


class RasterFLACConverter:
    """Drop-in for flac_raster.RasterFLACConverter (converter.py:17)."""

    def __init__(self, device: int = 0, pcm_mode: str = "reference", md5: bool = False):
        self.metadata_key = "RASTER_METADATA"
        self.logger = logging.getLogger("flac_raster.converter")
        self.device = device
        self.pcm_mode = pcm_mode
        self.md5 = md5   # fill STREAMINFO's MD5 signature (the reference's files carry zeros)

    def _calculate_audio_params(self, raster_data: np.ndarray, dtype) -> Tuple[int, int]:
        """converter.py:25-54"""
        return sample_rate_for(raster_data.shape[0], raster_data.shape[1]), bits_label_for(dtype)

    def tiff_to_flac(self, tiff_path: Path, flac_path: Path, compression_level: int = 5, spatial_tiling: bool = False,
                     tile_size: int = 512):
        """converter.py:112-232.  Returns the SpatialIndex when spatial_tiling, else None."""
        tiff_path, flac_path = Path(tiff_path), Path(flac_path)
        self.logger.info(f"Starting TIFF to FLAC conversion: {tiff_path} -> {flac_path}")
        if spatial_tiling:
            from .spatial_encoder import SpatialFLACEncoder
            return SpatialFLACEncoder(tile_size=tile_size, device=self.device).encode_spatial_flac(
                tiff_path, flac_path, compression_level)
        with rio.open(tiff_path) as src:
            data = get_codec(self.device).read_tiff(src)      # file bytes -> device -> frx_tiff_unpack (no host rearrangement)
            meta = dict(src.meta)
            bounds, crs, transform = src.bounds, src.crs, src.transform
        count, h, w = data.shape
        dtype = np_dtype_of(data)
        sample_rate, label = self._calculate_audio_params(data, dtype)
        self.logger.info(f"Raster {w}x{h}, {count} band(s), {dtype}; sample rate {sample_rate} Hz, label {label}-bit, "
                         f"effective {pcm_bps(dtype.name, self.pcm_mode)}-bit")
        frames, frame_sizes, dmin, dmax, bps = encode_array(data, compression_level, sample_rate, self.pcm_mode, self.device)
        md = raster_metadata(meta["width"], meta["height"], meta["count"], meta["dtype"], crs, transform, bounds, dmin, dmax,
                             meta.get("nodata"), meta.get("driver", "GTiff"))
        sig = get_codec(self.device).tile_md5(data, [(0, 0, h, w, 0, count)], self.pcm_mode)[0].tobytes() if self.md5 else None
        header = stream_header(count, bps, sample_rate, h * w, frame_sizes, metadata_comments(md, self.pcm_mode), md5=sig)
        with open(flac_path, "wb") as f:
            f.write(header)
            f.write(frames.tobytes())
        out_size, in_size = flac_path.stat().st_size, tiff_path.stat().st_size
        self.logger.info(f"Conversion complete: {out_size / 1048576:.2f} MB (compression {(1 - out_size / in_size) * 100:.1f}%)")
        return None

    def flac_to_tiff(self, flac_path: Path, tiff_path: Path):
        """converter.py:234-313"""
        flac_path, tiff_path = Path(flac_path), Path(tiff_path)
        self.logger.info(f"Starting FLAC to TIFF conversion: {flac_path} -> {tiff_path}")
        blob = np.fromfile(flac_path, dtype=np.uint8)
        metadata = self._read_embedded_metadata(flac_path, blob)
        if not metadata:
            self.logger.error("No metadata found in FLAC file or sidecar file")
            raise ValueError("No metadata found in FLAC file or sidecar file")
        if metadata.get("spatial_tiling") and flacmeta.parse_header(blob[:1 << 24].tobytes()).get("GEOSPATIAL_SPATIAL_INDEX"):
            from .spatial_encoder import decode_spatial_file
            raster = decode_spatial_file(flac_path, device=self.device)
        else:
            raster = decode_stream(blob, metadata, self.device)
        meta = {"driver": "GTiff", "width": metadata["width"], "height": metadata["height"], "count": metadata["count"],
                "dtype": np.dtype(metadata["dtype"]), "nodata": metadata.get("nodata")}
        if metadata.get("crs"):
            meta["crs"] = rio.CRS.from_string(metadata["crs"])
        if metadata.get("transform"):
            t = metadata["transform"]
            meta["transform"] = rio.Affine(t[0], t[1], t[2], t[3], t[4], t[5])
        with rio.open(tiff_path, "w", **meta) as dst:
            if metadata["count"] == 1:
                dst.write(raster[0], 1)
            else:
                dst.write(raster)
        self.logger.info(f"TIFF written successfully: {tiff_path.stat().st_size / 1048576:.2f} MB")

    def _read_embedded_metadata(self, flac_path: Path, blob: Optional[np.ndarray] = None) -> Optional[Dict]:
        """converter.py:375-427 (embedded VORBIS_COMMENT first, `.json` sidecar second)."""
        try:
            if blob is None:
                with open(flac_path, "rb") as f:
                    head = f.read(1 << 24)
            else:
                head = blob[:1 << 24].tobytes()
            md = metadata_from_comments(flacmeta.parse_header(head))
            if md is None:
                raise ValueError("No embedded metadata found")
            return md
        except Exception as e:  # noqa: BLE001 - same breadth as the reference
            self.logger.warning(f"Failed to read embedded metadata: {e}")
            sidecar = Path(flac_path).with_suffix(".json")
            if sidecar.exists():
                with open(sidecar) as f:
                    return json.load(f)
        return None

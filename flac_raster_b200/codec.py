"""Host-side driver of the libfrx hot path.  PyTorch is used for device memory and streams
only; every byte of codec work happens in the CUDA kernels behind the C-ABI (include/frx.h).

``Codec.encode_tiles``  replaces the per-tile body of create-streaming / tiff_to_flac
    (cli.py:691-763, converter.py:152-216): min/max, PCM scaling, band->channel
    interleave, libFLAC encode.
``Codec.decode_tiles``  replaces the per-tile body of extract-streaming / flac_to_tiff
    (cli.py:1016-1023, converter.py:241-282).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import _lib as L

_TORCH_NAMES = {"torch.uint8": "uint8", "torch.int16": "int16", "torch.uint16": "uint16", "torch.int32": "int32",
                "torch.uint32": "uint32", "torch.float32": "float32"}


def _torch():
    import torch
    return torch


def _tdtype(name: str):
    torch = _torch()
    return {"uint8": torch.uint8, "int16": torch.int16, "uint16": torch.uint16, "int32": torch.int32,
            "uint32": torch.uint32, "float32": torch.float32}[name]


def np_dtype_of(raster) -> np.dtype:
    """numpy dtype of a numpy array or a torch tensor"""
    return raster.dtype if isinstance(raster, np.ndarray) else np.dtype(dtype_name(raster))


def dtype_name(t) -> str:
    name = _TORCH_NAMES.get(str(t.dtype)) if not isinstance(t, np.ndarray) else str(t.dtype)
    if name not in L.DTYPE_CODE:
        raise L.FrxUnsupported(L.E_UNSUPPORTED, f"raster dtype {t.dtype} is not supported")
    return name


def pcm_bps(dtype: str, pcm_mode: str = "reference") -> int:
    """Effective FLAC bits/sample for a raster dtype (pyFLAC itemsize rule)."""
    return L.check(L.lib().frx_pcm_bps(L.DTYPE_CODE[dtype], L.PCM_MODE[pcm_mode]))


def tile_grid(height: int, width: int, tile_size: int, first_band: int = 0, n_channels: int = 1):
    """Row-major tile list, edge tiles clipped (cli.py:691-698, spatial_encoder.py:92-103)."""
    out = []
    for r in range(0, height, tile_size):
        for c in range(0, width, tile_size):
            out.append((r, c, min(tile_size, height - r), min(tile_size, width - c), first_band, n_channels))
    return out


@dataclass
class EncodeResult:
    """Device tensors (or numpy arrays for the host entry point)."""
    bytes: object          # uint8 [capacity]; valid prefix = total_bytes
    tile_offset: object    # uint64-as-int64 [n_tiles]
    tile_size: object      # int64 [n_tiles]
    tile_minmax: object    # float64 [n_tiles, 2]
    frame_size: object     # int32 [total_frames]
    frame_offset: object   # int64 [total_frames]
    total_bytes: object    # int64 [1]
    status: object         # int32 [1]
    frames_per_tile: np.ndarray
    bps: int
    blocksize: int


class Codec:
    """One libfrx context on one CUDA device."""

    def __init__(self, device: int = 0):
        self._lib = L.lib()
        h = C.c_void_p()
        L.check(self._lib.frx_ctx_create(int(device), C.byref(h)))
        self._h = h
        self.device = int(device)

    def close(self):
        if getattr(self, "_h", None):
            self._lib.frx_ctx_destroy(self._h)
            self._h = None

    __del__ = close

    # ---- helpers ---------------------------------------------------------------------------
    def _stream(self, stream=None):
        if stream is not None:
            return C.c_void_p(int(stream))
        torch = _torch()
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    @staticmethod
    def _params(level, blocksize, sample_rate, pcm_mode):
        return L.EncodeParams(int(level), int(blocksize), int(sample_rate), L.PCM_MODE[pcm_mode])

    @staticmethod
    def _raster_desc(ptr, dtype, bands, H, W, band_stride=0, row_stride=0):
        return L.Raster(C.c_void_p(ptr), L.DTYPE_CODE[dtype], bands, H, W, band_stride, row_stride)

    @staticmethod
    def prepare_tiles(tiles):
        """The frx_tile array of a tile list, built once: every entry point that takes `tiles` also takes this array
        (a caller that encodes the same grid repeatedly skips the per-call conversion)."""
        return L.make_tiles(tiles)

    def encode_bound(self, tiles, dtype: str, level=5, blocksize=4096, sample_rate=44100, pcm_mode="reference"):
        arr = L.make_tiles(tiles)
        p = self._params(level, blocksize, sample_rate, pcm_mode)
        nf, nb = C.c_uint64(), C.c_uint64()
        L.check(self._lib.frx_encode_bound(arr, len(arr), L.DTYPE_CODE[dtype], C.byref(p), C.byref(nf), C.byref(nb)))
        return nf.value, nb.value

    @property
    def launch_count(self) -> int:
        return int(self._lib.frx_launch_count(self._h))

    def last_kernel_ms(self):
        e, d = C.c_float(), C.c_float()
        L.check(self._lib.frx_last_kernel_ms(self._h, C.byref(e), C.byref(d)))
        return e.value, d.value

    # ---- raster ingestion ----------------------------------------------------
    def tiff_unpack(self, file_dev, chunks, dtype: str, bands: int, height: int, width: int, samples_per_pixel: int,
                    swap_bytes: bool = False, out=None, stream=None):
        """file_dev: CUDA uint8 tensor holding (a span of) an uncompressed TIFF as it lies on disk.
        chunks: [(src_offset, row0, col0, rows, cols, pitch, band)] (see include/frx.h frx_tiff_chunk).
        Returns the band-planar CUDA tensor (bands, height, width) that rasterio's read() would return."""
        torch = _torch()
        assert file_dev.is_cuda and file_dev.dtype == torch.uint8 and file_dev.is_contiguous()
        if out is None:
            out = torch.empty((bands, height, width), dtype=_tdtype(dtype), device=file_dev.device)
        assert out.is_cuda and out.is_contiguous() and tuple(out.shape) == (bands, height, width)
        arr = L.make_tiff_chunks(chunks)
        r = self._raster_desc(out.data_ptr(), dtype, bands, height, width)
        L.check(self._lib.frx_tiff_unpack(self._h, C.c_void_p(file_dev.data_ptr()), file_dev.numel(), arr, len(chunks),
                                          int(samples_per_pixel), 1 if swap_bytes else 0, C.byref(r), self._stream(stream)))
        return out

    def read_tiff(self, src, row0: int = 0, row1: int | None = None, bands=None):
        """Rows [row0, row1) of an open raster as a band-planar CUDA tensor (only the 0-based `bands` when given).  With the
        built-in GeoTIFF reader the file bytes go to the device as they are and frx_tiff_unpack rearranges them there; any
        other dataset object (rasterio) is read on the host and uploaded."""
        if hasattr(src, "read_device") and src.dtypes[0] in L.DTYPE_CODE:
            return src.read_device(self, row0, row1, bands)
        from .rio import Window
        from .synth import from_numpy
        row1 = src.height if row1 is None else row1
        idx = None if bands is None else [b + 1 for b in bands]
        return from_numpy(src.read(indexes=idx, window=Window(0, row0, src.width, row1 - row0)), device=f"cuda:{self.device}")

    def encode_tiles_any(self, raster, tiles, level: int = 5, blocksize: int = 4096, sample_rate: int = 44100,
                         pcm_mode: str = "reference") -> EncodeResult:
        """encode_tiles_host for a numpy raster; for a CUDA tensor (read_tiff) the device entry point, results copied
        back as numpy arrays of the same layout."""
        if isinstance(raster, np.ndarray):
            return self.encode_tiles_host(raster, tiles, level=level, blocksize=blocksize, sample_rate=sample_rate, pcm_mode=pcm_mode)
        torch = _torch()
        res = self.encode_tiles(raster, tiles, level=level, blocksize=blocksize, sample_rate=sample_rate, pcm_mode=pcm_mode)
        torch.cuda.synchronize(raster.device)
        if int(res.status.item()) != 0:
            raise L.FrxError(int(res.status.item()), "encode output buffer too small")
        total = int(res.total_bytes.item())
        return EncodeResult(bytes=res.bytes[:total].cpu().numpy(), tile_offset=res.tile_offset.cpu().numpy(),
                            tile_size=res.tile_size.cpu().numpy(), tile_minmax=res.tile_minmax.cpu().numpy(),
                            frame_size=res.frame_size.cpu().numpy(), frame_offset=res.frame_offset.cpu().numpy(),
                            total_bytes=np.array([total], dtype=np.int64), status=np.zeros(1, dtype=np.int32),
                            frames_per_tile=res.frames_per_tile, bps=res.bps, blocksize=res.blocksize)

    # ---- device path -----------------------------------------------------------------------
    def encode_tiles(self, raster, tiles, level: int = 5, blocksize: int = 4096, sample_rate: int = 44100,
                     pcm_mode: str = "reference", capacity: int | None = None, out: EncodeResult | None = None,
                     dtype: str | None = None, stream=None) -> EncodeResult:
        """raster: CUDA tensor (bands,H,W), contiguous.  tiles: [(row_off,col_off,h,w,first_band,n_ch)]."""
        torch = _torch()
        assert raster.is_cuda and raster.dim() == 3 and raster.is_contiguous()
        dtype = dtype or dtype_name(raster)
        bands, H, W = raster.shape
        arr = L.make_tiles(tiles)
        n = len(arr)
        p = self._params(level, blocksize, sample_rate, pcm_mode)
        nf, nb = C.c_uint64(), C.c_uint64()
        L.check(self._lib.frx_encode_bound(arr, n, L.DTYPE_CODE[dtype], C.byref(p), C.byref(nf), C.byref(nb)))
        nf, nb = nf.value, nb.value
        dev = raster.device
        if out is None:
            cap = int(capacity) if capacity is not None else nb
            out = EncodeResult(
                bytes=torch.empty(cap + 16, dtype=torch.uint8, device=dev),
                tile_offset=torch.empty(n, dtype=torch.int64, device=dev),
                tile_size=torch.empty(n, dtype=torch.int64, device=dev),
                tile_minmax=torch.empty((n, 2), dtype=torch.float64, device=dev),
                frame_size=torch.empty(nf, dtype=torch.int32, device=dev),
                frame_offset=torch.empty(nf, dtype=torch.int64, device=dev),
                total_bytes=torch.zeros(1, dtype=torch.int64, device=dev),
                status=torch.zeros(1, dtype=torch.int32, device=dev),
                frames_per_tile=np.array([(t.h * t.w + blocksize - 1) // blocksize for t in arr], dtype=np.int64),
                bps=pcm_bps(dtype, pcm_mode), blocksize=blocksize)
        cap = out.bytes.numel() - 16
        o = L.EncodeOut(out.bytes.data_ptr(), cap, out.tile_offset.data_ptr(), out.tile_size.data_ptr(),
                        out.tile_minmax.data_ptr(), out.frame_size.data_ptr(), out.frame_offset.data_ptr(),
                        out.total_bytes.data_ptr(), out.status.data_ptr())
        r = self._raster_desc(raster.data_ptr(), dtype, bands, H, W)
        L.check(self._lib.frx_encode_tiles(self._h, C.byref(r), arr, n, C.byref(p), C.byref(o), self._stream(stream)))
        return out

    def decode_tiles(self, data, dtiles, out, out_dtype: str | None = None, frame_size=None, blocksize: int = 4096,
                     pcm_mode: str = "reference", status=None, n_bytes: int | None = None, stream=None):
        """data: CUDA uint8 tensor; dtiles: ctypes array of _lib.DTile; out: CUDA tensor that receives the
        raster blocks.  Returns the per-tile status tensor (int32, 0 = ok)."""
        torch = _torch()
        out_dtype = out_dtype or dtype_name(out)
        n = len(dtiles)
        if status is None:
            status = torch.zeros(n, dtype=torch.int32, device=data.device)
        fs = C.c_void_p(frame_size.data_ptr()) if frame_size is not None else None
        nb = int(n_bytes) if n_bytes is not None else data.numel()
        L.check(self._lib.frx_decode_tiles(self._h, C.c_void_p(data.data_ptr()), nb, dtiles, n, fs, blocksize,
                                           L.DTYPE_CODE[out_dtype], L.PCM_MODE[pcm_mode], C.c_void_p(out.data_ptr()),
                                           C.c_void_p(status.data_ptr()), self._stream(stream)))
        return status

    def decode_tiles_noalloc(self, data, dtiles, out, status, out_dtype: str | None = None, frame_size=None, blocksize: int = 4096,
                             pcm_mode: str = "reference", n_bytes: int | None = None, stream=None):
        """decode_tiles through frx_decode_tiles_noalloc: nothing is allocated, here or in the library (the caller passes
        `status`; scratch is a static device array inside libfrx).  At most _lib.NOALLOC_MAX_TILES tiles and
        _lib.NOALLOC_MAX_FRAMES frames per call, no two-channel tiles; calls that overlap on one GPU race (include/frx.h)."""
        out_dtype = out_dtype or dtype_name(out)
        fs = C.c_void_p(frame_size.data_ptr()) if frame_size is not None else None
        nb = int(n_bytes) if n_bytes is not None else data.numel()
        L.check(self._lib.frx_decode_tiles_noalloc(self._h, C.c_void_p(data.data_ptr()), nb, dtiles, len(dtiles), fs, blocksize,
                                                   L.DTYPE_CODE[out_dtype], L.PCM_MODE[pcm_mode], C.c_void_p(out.data_ptr()),
                                                   C.c_void_p(status.data_ptr()), self._stream(stream)))
        return status

    def quantize(self, raster, tile, pcm_mode: str = "reference", dtype: str | None = None):
        """K0+K1 only: (pcm int32 (h*w, n_ch) tensor, minmax float64[2] tensor)."""
        torch = _torch()
        dtype = dtype or dtype_name(raster)
        bands, H, W = raster.shape
        arr = L.make_tiles([tile])
        t = arr[0]
        pcm = torch.empty((t.h * t.w, t.n_channels), dtype=torch.int32, device=raster.device)
        mm = torch.empty(2, dtype=torch.float64, device=raster.device)
        r = self._raster_desc(raster.data_ptr(), dtype, bands, H, W)
        L.check(self._lib.frx_quantize(self._h, C.byref(r), arr, L.PCM_MODE[pcm_mode], C.c_void_p(pcm.data_ptr()),
                                       C.c_void_p(mm.data_ptr()), self._stream()))
        return pcm, mm

    def tile_md5(self, raster, tiles, pcm_mode: str = "reference", dtype: str | None = None) -> np.ndarray:
        """MD5 signatures (n_tiles, 16) uint8 of the tiles' PCM, as STREAMINFO wants them.  raster: CUDA tensor or numpy array."""
        torch = _torch()
        if isinstance(raster, np.ndarray):
            from .synth import from_numpy
            raster = from_numpy(raster, device=f"cuda:{self.device}")
        dtype = dtype or dtype_name(raster)
        bands, H, W = raster.shape
        arr = L.make_tiles(tiles)
        out = torch.zeros((len(arr), 16), dtype=torch.uint8, device=raster.device)
        r = self._raster_desc(raster.data_ptr(), dtype, bands, H, W)
        L.check(self._lib.frx_tile_md5(self._h, C.byref(r), arr, len(arr), L.PCM_MODE[pcm_mode], C.c_void_p(out.data_ptr()), self._stream()))
        torch.cuda.synchronize(raster.device)
        return out.cpu().numpy()

    def dequantize(self, pcm, h, w, channels, bps, out_dtype: str, dmin: float, dmax: float,
                   pcm_mode: str = "reference"):
        """K12 only: int32 PCM (h*w, channels) -> (channels,h,w) raster tensor."""
        torch = _torch()
        out = torch.empty((channels, h, w), dtype=_tdtype(out_dtype), device=pcm.device)
        L.check(self._lib.frx_dequantize(self._h, C.c_void_p(pcm.data_ptr()), h, w, channels, bps,
                                         L.DTYPE_CODE[out_dtype], L.PCM_MODE[pcm_mode], float(dmin), float(dmax),
                                         C.c_void_p(out.data_ptr()), self._stream()))
        return out

    # ---- host path (what a pyflac-style caller binds) -----------------------------------------
    def encode_tiles_host(self, raster: np.ndarray, tiles, level: int = 5, blocksize: int = 4096,
                          sample_rate: int = 44100, pcm_mode: str = "reference",
                          out: EncodeResult | None = None) -> EncodeResult:
        """Host buffers in, host buffers out (numpy; pass views of pinned torch tensors for full PCIe speed).
        `out` lets the caller reuse (pinned) output arrays across calls."""
        raster = np.ascontiguousarray(raster)
        dtype = str(raster.dtype)
        bands, H, W = raster.shape
        arr = L.make_tiles(tiles)
        n = len(arr)
        p = self._params(level, blocksize, sample_rate, pcm_mode)
        nf, nb = self.encode_bound(tiles, dtype, level, blocksize, sample_rate, pcm_mode)
        res = out if out is not None else EncodeResult(bytes=np.empty(nb, dtype=np.uint8), tile_offset=np.zeros(n, dtype=np.int64),
                           tile_size=np.zeros(n, dtype=np.int64), tile_minmax=np.zeros((n, 2), dtype=np.float64),
                           frame_size=np.zeros(nf, dtype=np.int32), frame_offset=np.zeros(nf, dtype=np.int64),
                           total_bytes=np.zeros(1, dtype=np.int64), status=np.zeros(1, dtype=np.int32),
                           frames_per_tile=np.array([(t.h * t.w + blocksize - 1) // blocksize for t in arr],
                                                    dtype=np.int64),
                           bps=pcm_bps(dtype, pcm_mode), blocksize=blocksize)
        nb = int(res.bytes.size)
        o = L.EncodeOut(res.bytes.ctypes.data, nb, res.tile_offset.ctypes.data, res.tile_size.ctypes.data,
                        res.tile_minmax.ctypes.data, res.frame_size.ctypes.data, res.frame_offset.ctypes.data,
                        res.total_bytes.ctypes.data, res.status.ctypes.data)
        r = self._raster_desc(raster.ctypes.data, dtype, bands, H, W)
        L.check(self._lib.frx_encode_tiles_host(self._h, C.byref(r), arr, n, C.byref(p), C.byref(o)))
        return res

    def decode_tiles_host(self, data: np.ndarray, dtiles, out: np.ndarray, frame_size: np.ndarray | None = None,
                          blocksize: int = 4096, pcm_mode: str = "reference"):
        data = np.ascontiguousarray(data, dtype=np.uint8)
        assert out.flags.c_contiguous
        n = len(dtiles)
        status = np.zeros(n, dtype=np.int32)
        fs = None
        if frame_size is not None:
            frame_size = np.ascontiguousarray(frame_size, dtype=np.uint32)
            fs = C.c_void_p(frame_size.ctypes.data)
        L.check(self._lib.frx_decode_tiles_host(self._h, C.c_void_p(data.ctypes.data), data.size, dtiles, n, fs,
                                                blocksize, L.DTYPE_CODE[str(out.dtype)], L.PCM_MODE[pcm_mode],
                                                C.c_void_p(out.ctypes.data), out.size,
                                                C.c_void_p(status.ctypes.data)))
        return status


def make_dtiles(entries):
    """entries: iterable of dicts with byte_offset, byte_size, out_offset, out_band_stride, out_row_stride, h, w,
    channels, bps, dmin, dmax, first_frame (or None), n_frames."""
    entries = list(entries)
    n = len(entries)
    rec = np.zeros(n, dtype=_DTILE_DTYPE)           # filled column by column, copied into the ctypes array in one go
    for name in ("byte_offset", "byte_size", "out_offset", "out_band_stride", "out_row_stride", "h", "w", "channels", "bps", "n_frames"):
        rec[name] = [int(e[name]) for e in entries]
    for name in ("dmin", "dmax"):
        rec[name] = [float(e[name]) for e in entries]
    rec["first_frame"] = [L.NO_FRAME_TABLE if e.get("first_frame") is None else int(e["first_frame"]) for e in entries]
    return (L.DTile * n).from_buffer_copy(rec.tobytes()) if n else (L.DTile * 0)()


_DTILE_DTYPE = np.dtype([("byte_offset", "<u8"), ("byte_size", "<u8"), ("out_offset", "<u8"), ("out_band_stride", "<i8"),
                         ("out_row_stride", "<i8"), ("h", "<i4"), ("w", "<i4"), ("channels", "<i4"), ("bps", "<i4"),
                         ("dmin", "<f8"), ("dmax", "<f8"), ("first_frame", "<u4"), ("n_frames", "<u4")])
assert _DTILE_DTYPE.itemsize == C.sizeof(L.DTile)

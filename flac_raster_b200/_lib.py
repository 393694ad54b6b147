"""ctypes binding of libfrx.so (include/frx.h).  There is no CPU fallback: if the shared
library is missing, or no CUDA device is present when a context is created, this raises."""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

_PKG = Path(__file__).resolve().parent
LIB_PATH = _PKG / "libfrx.so"

FRX_U8, FRX_I16, FRX_U16, FRX_I32, FRX_U32, FRX_F32 = range(6)
DTYPE_CODE = {"uint8": FRX_U8, "int16": FRX_I16, "uint16": FRX_U16, "int32": FRX_I32, "uint32": FRX_U32,
              "float32": FRX_F32}
CODE_DTYPE = {v: k for k, v in DTYPE_CODE.items()}
PCM_REFERENCE, PCM_LOSSLESS, PCM_REFERENCE_SPATIAL = 0, 1, 2
PCM_MODE = {"reference": PCM_REFERENCE, "lossless": PCM_LOSSLESS, "reference_spatial": PCM_REFERENCE_SPATIAL}
NO_FRAME_TABLE = 0xFFFFFFFF

E_INVALID, E_CUDA, E_NOMEM, E_UNSUPPORTED, E_OVERFLOW, E_BITSTREAM = -1, -2, -3, -4, -5, -6


class FrxError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libfrx error {code}: {msg}")
        self.code = code


class FrxUnsupported(FrxError):
    """e.g. > 8 channels per stream: the reference raises pyflac.EncoderInitException here."""


class Tile(C.Structure):
    _fields_ = [("row_off", C.c_int32), ("col_off", C.c_int32), ("h", C.c_int32), ("w", C.c_int32),
                ("first_band", C.c_int32), ("n_channels", C.c_int32)]


class EncodeParams(C.Structure):
    _fields_ = [("compression_level", C.c_int32), ("blocksize", C.c_int32), ("sample_rate", C.c_int32),
                ("pcm_mode", C.c_int32)]


class Raster(C.Structure):
    _fields_ = [("data", C.c_void_p), ("dtype", C.c_int32), ("bands", C.c_int32), ("height", C.c_int64),
                ("width", C.c_int64), ("band_stride", C.c_int64), ("row_stride", C.c_int64)]


class EncodeOut(C.Structure):
    _fields_ = [("bytes", C.c_void_p), ("capacity", C.c_uint64), ("tile_offset", C.c_void_p),
                ("tile_size", C.c_void_p), ("tile_minmax", C.c_void_p), ("frame_size", C.c_void_p),
                ("frame_offset", C.c_void_p), ("total_bytes", C.c_void_p), ("status", C.c_void_p)]


class TiffChunk(C.Structure):
    _fields_ = [("src_offset", C.c_uint64), ("row0", C.c_int32), ("col0", C.c_int32), ("rows", C.c_int32),
                ("cols", C.c_int32), ("pitch", C.c_int32), ("band", C.c_int32)]


class DTile(C.Structure):
    _fields_ = [("byte_offset", C.c_uint64), ("byte_size", C.c_uint64), ("out_offset", C.c_uint64),
                ("out_band_stride", C.c_int64), ("out_row_stride", C.c_int64),
                ("h", C.c_int32), ("w", C.c_int32), ("channels", C.c_int32), ("bps", C.c_int32),
                ("dmin", C.c_double), ("dmax", C.c_double), ("first_frame", C.c_uint32), ("n_frames", C.c_uint32)]


EXPORTS = ["frx_version", "frx_last_error", "frx_device_count", "frx_ctx_create", "frx_ctx_destroy",
           "frx_encode_bound", "frx_pcm_bps", "frx_encode_tiles", "frx_decode_tiles", "frx_quantize",
           "frx_dequantize", "frx_encode_tiles_host", "frx_decode_tiles_host", "frx_launch_count",
           "frx_last_kernel_ms", "frx_tiff_unpack", "frx_decode_tiles_noalloc", "frx_tile_md5"]
NOALLOC_MAX_TILES, NOALLOC_MAX_FRAMES = 128, 32768   # include/frx.h FRX_NOALLOC_MAX_*

_lib = None


def build(force: bool = False) -> Path:
    """Compile libfrx.so for sm_100a with nvcc (csrc/Makefile)."""
    src = _PKG / "csrc"
    newest = max(p.stat().st_mtime for p in list(src.glob("*.cu")) + list(src.glob("*.cuh")) +
                 [_PKG.parent / "include" / "frx.h"])
    if force or not LIB_PATH.exists() or LIB_PATH.stat().st_mtime < newest:
        subprocess.run(["make", "-C", str(src), "-j4"], check=True, capture_output=True)
    return LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise FrxError(E_CUDA, f"{LIB_PATH} is missing: build it with flac_raster_b200._lib.build() "
                                   "(nvcc, sm_100a). There is no CPU fallback.")
        L = C.CDLL(str(LIB_PATH))
        L.frx_version.restype = C.c_int
        L.frx_last_error.restype = C.c_char_p
        L.frx_device_count.restype = C.c_int
        L.frx_ctx_create.restype = C.c_int
        L.frx_ctx_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
        L.frx_ctx_destroy.restype = None
        L.frx_ctx_destroy.argtypes = [C.c_void_p]
        L.frx_encode_bound.restype = C.c_int
        L.frx_encode_bound.argtypes = [C.POINTER(Tile), C.c_int32, C.c_int32, C.POINTER(EncodeParams),
                                       C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.frx_pcm_bps.restype = C.c_int
        L.frx_pcm_bps.argtypes = [C.c_int32, C.c_int32]
        L.frx_encode_tiles.restype = C.c_int
        L.frx_encode_tiles.argtypes = [C.c_void_p, C.POINTER(Raster), C.POINTER(Tile), C.c_int32,
                                       C.POINTER(EncodeParams), C.POINTER(EncodeOut), C.c_void_p]
        L.frx_decode_tiles.restype = C.c_int
        L.frx_decode_tiles.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(DTile), C.c_int32, C.c_void_p,
                                       C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]
        L.frx_decode_tiles_noalloc.restype = C.c_int
        L.frx_decode_tiles_noalloc.argtypes = L.frx_decode_tiles.argtypes
        L.frx_tile_md5.restype = C.c_int
        L.frx_tile_md5.argtypes = [C.c_void_p, C.POINTER(Raster), C.POINTER(Tile), C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]
        L.frx_quantize.restype = C.c_int
        L.frx_quantize.argtypes = [C.c_void_p, C.POINTER(Raster), C.POINTER(Tile), C.c_int32, C.c_void_p, C.c_void_p,
                                   C.c_void_p]
        L.frx_dequantize.restype = C.c_int
        L.frx_dequantize.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                     C.c_int32, C.c_double, C.c_double, C.c_void_p, C.c_void_p]
        L.frx_encode_tiles_host.restype = C.c_int
        L.frx_encode_tiles_host.argtypes = [C.c_void_p, C.POINTER(Raster), C.POINTER(Tile), C.c_int32,
                                            C.POINTER(EncodeParams), C.POINTER(EncodeOut)]
        L.frx_decode_tiles_host.restype = C.c_int
        L.frx_decode_tiles_host.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(DTile), C.c_int32,
                                            C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_uint64,
                                            C.c_void_p]
        L.frx_tiff_unpack.restype = C.c_int
        L.frx_tiff_unpack.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(TiffChunk), C.c_int32, C.c_int32,
                                      C.c_int32, C.POINTER(Raster), C.c_void_p]
        L.frx_launch_count.restype = C.c_uint64
        L.frx_launch_count.argtypes = [C.c_void_p]
        L.frx_last_kernel_ms.restype = C.c_int
        L.frx_last_kernel_ms.argtypes = [C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_float)]
        _lib = L
    return _lib


def check(rc: int):
    if rc < 0:
        msg = lib().frx_last_error().decode("utf-8", "replace")
        if rc == E_UNSUPPORTED:
            raise FrxUnsupported(rc, msg)
        raise FrxError(rc, msg)
    return rc


def make_tiles(tiles) -> "C.Array[Tile]":
    """tiles: iterable of (row_off, col_off, h, w, first_band, n_channels)."""
    if isinstance(tiles, C.Array) and tiles._type_ is Tile:
        return tiles                      # already prepared (hot loops build the array once)
    import numpy as np
    a = np.ascontiguousarray(np.asarray(list(tiles), dtype=np.int64).reshape(-1, 6).astype(np.int32))
    # one memcpy instead of six Python attribute stores per tile (400 tiles: 1 ms -> 20 us)
    return (Tile * a.shape[0]).from_buffer_copy(a.tobytes()) if a.shape[0] else (Tile * 0)()


def make_tiff_chunks(chunks) -> "C.Array[TiffChunk]":
    """chunks: iterable of (src_offset, row0, col0, rows, cols, pitch, band) -> frx_tiff_chunk array (at least one slot)."""
    import numpy as np
    chunks = list(chunks)
    rec = np.zeros(max(1, len(chunks)), dtype=np.dtype([("src_offset", "<u8"), ("row0", "<i4"), ("col0", "<i4"), ("rows", "<i4"),
                                                         ("cols", "<i4"), ("pitch", "<i4"), ("band", "<i4")]))
    if chunks:
        a = np.asarray(chunks, dtype=np.int64).reshape(-1, 7)
        for k, name in enumerate(rec.dtype.names):
            rec[name][:len(chunks)] = a[:, k]
    assert rec.dtype.itemsize == C.sizeof(TiffChunk)
    return (TiffChunk * rec.shape[0]).from_buffer_copy(rec.tobytes())

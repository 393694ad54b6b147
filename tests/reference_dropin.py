"""SURVEY 8(b) "strongest drop-in test": the UNMODIFIED reference package (/root/reference/src/flac_raster) runs on top of
`flac_raster_b200.pyflac_shim` injected as `sys.modules['pyflac']`, with rasterio replaced by the built-in GeoTIFF adapter
(rasterio/GDAL is not installable on the image; mutagen is absent too, so the reference takes its own JSON-sidecar fallback).
Shared by the CPU suite (oracle-backed fake codec) and the GPU suite (libfrx)."""
import importlib
import sys
import types
from pathlib import Path

import numpy as np

REF_SRC = Path("/root/reference/src")
GOLDEN = Path(__file__).resolve().parent / "golden"


def available() -> bool:
    return (REF_SRC / "flac_raster" / "converter.py").exists()


def import_reference():
    """import flac_raster (the reference) with pyflac -> our shim and rasterio -> flac_raster_b200.geotiff"""
    from flac_raster_b200 import geotiff, pyflac_shim
    ras = types.ModuleType("rasterio")
    ras.open = geotiff.open
    crs = types.ModuleType("rasterio.crs"); crs.CRS = geotiff.CRS
    tr = types.ModuleType("rasterio.transform"); tr.Affine = geotiff.Affine; tr.from_origin = geotiff.from_origin
    win = types.ModuleType("rasterio.windows"); win.Window = geotiff.Window
    ras.crs, ras.transform, ras.windows = crs, tr, win
    saved = {k: sys.modules.get(k) for k in ("rasterio", "rasterio.crs", "rasterio.transform", "rasterio.windows", "pyflac", "flac_raster",
                                              "flac_raster.converter", "flac_raster.compare", "flac_raster.spatial_encoder")}
    sys.modules.update({"rasterio": ras, "rasterio.crs": crs, "rasterio.transform": tr, "rasterio.windows": win, "pyflac": pyflac_shim})
    for k in list(sys.modules):
        if k == "flac_raster" or k.startswith("flac_raster."):
            del sys.modules[k]
    sys.path.insert(0, str(REF_SRC))
    try:
        mod = importlib.import_module("flac_raster.converter")
    finally:
        sys.path.remove(str(REF_SRC))
    return mod, saved


def restore(saved):
    for k in list(sys.modules):
        if k == "flac_raster" or k.startswith("flac_raster."):
            del sys.modules[k]
    for k, v in saved.items():
        if v is None:
            sys.modules.pop(k, None)
        else:
            sys.modules[k] = v


def reference_converter_roundtrip(tmp):
    """reference tiff_to_flac -> reference flac_to_tiff, every codec call served by the shim; the result must be the
    reference round trip the oracle restates, the stream must be valid FLAC whose PCM is the reference quantiser's, and our
    own converter must read the file the reference code wrote."""
    import oracle
    from oracle import quant
    from flac_raster_b200 import rio
    from flac_raster_b200.converter import RasterFLACConverter
    mod, saved = import_reference()
    try:
        assert Path(mod.__file__).resolve().is_relative_to(REF_SRC), mod.__file__
        conv = mod.RasterFLACConverter()
        for name in ("sample_dem.tif", "sample_rgb.tif"):
            src = GOLDEN / name
            flac, back = tmp / (name + ".flac"), tmp / (name + ".back.tif")
            conv.tiff_to_flac(src, flac)                    # unmodified reference code path (converter.py:112-232)
            assert flac.with_suffix(".json").exists()       # mutagen is absent: the reference's own sidecar fallback
            conv.flac_to_tiff(flac, back)                   # converter.py:234-313
            with rio.open(src) as s:
                arr = s.read()
            with rio.open(back) as s:
                got = s.read()
            pcm_ref, bps, dmin, dmax, _ = quant.reference_pcm(arr)
            want = quant.reference_decode_raster(pcm_ref, bps, arr.shape[1], arr.shape[2], arr.shape[0], arr.dtype, dmin, dmax)
            assert np.array_equal(got, want.reshape(got.shape))
            pcm, sid, frames, used, first, rc = oracle.decode_stream(flac.read_bytes())
            assert rc == 0 and all(f["crc8_ok"] and f["crc16_ok"] for f in frames)
            assert (sid["channels"], sid["bps"]) == (arr.shape[0], bps) and np.array_equal(pcm, pcm_ref)
            ours = tmp / (name + ".ours.tif")
            RasterFLACConverter().flac_to_tiff(flac, ours)  # our reader on the reference-written file + sidecar
            with rio.open(ours) as s:
                assert np.array_equal(s.read(), got)
    finally:
        restore(saved)

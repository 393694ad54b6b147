"""Raster I/O adapter: rasterio when it is importable, the built-in GeoTIFF subset otherwise.

The reference reads and writes rasters through rasterio (converter.py:137-143, 285-309;
spatial_encoder.py:154-183; cli.py:668-729).  rasterio/GDAL is not installable on the B200
image, so everything in this package imports ``open``, ``Window``, ``CRS`` and ``Affine``
from here; both back-ends expose the same attribute names.
"""
from __future__ import annotations

try:  # pragma: no cover - rasterio is absent on the build and GPU boxes
    import rasterio as _rio
    from rasterio.crs import CRS
    from rasterio.transform import Affine, from_origin
    from rasterio.windows import Window

    open = _rio.open  # noqa: A001
    BACKEND = "rasterio"
except ImportError:
    from .geotiff import CRS, Affine, Window, from_origin, open  # noqa: A001,F401

    BACKEND = "builtin-geotiff"

__all__ = ["open", "Window", "CRS", "Affine", "from_origin", "BACKEND"]

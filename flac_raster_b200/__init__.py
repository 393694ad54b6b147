"""flac_raster_b200 -- B200-native raster<->FLAC tile codec behind flac-raster's API surface.

Hot path: hand-written sm_100a CUDA kernels in ``csrc/`` behind the
C-ABI of ``include/frx.h`` (``libfrx.so``, loaded with ctypes by ``_lib``).  Host side mirrors
the reference's entry points: ``RasterFLACConverter`` (converter.py), ``SpatialFLACEncoder``,
``SpatialFLACStreamer``, ``SpatialIndex``, ``SpatialFrame`` (spatial_encoder.py), the streaming
format (streaming.py), the CLI commands (cli.py) and a pyFLAC-shaped shim (pyflac_shim.py).
"""
__version__ = "0.1.0"
__all__ = ["RasterFLACConverter", "SpatialFLACEncoder", "SpatialFLACStreamer", "SpatialIndex", "SpatialFrame",
           "compare_tiffs", "display_comparison_table"]


def __getattr__(name):  # lazy: importing the package must not need numpy/torch/libfrx
    if name == "RasterFLACConverter":
        from .converter import RasterFLACConverter
        return RasterFLACConverter
    if name in ("SpatialFLACEncoder", "SpatialFLACStreamer", "SpatialIndex", "SpatialFrame"):
        from . import spatial_encoder
        return getattr(spatial_encoder, name)
    if name in ("compare_tiffs", "display_comparison_table"):
        from . import compare
        return getattr(compare, name)
    raise AttributeError(name)

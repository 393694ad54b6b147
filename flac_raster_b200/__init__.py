"""flac_raster_b200 -- B200-native raster<->FLAC tile codec behind flac-raster's API surface.

Hot path: hand-written sm_100a CUDA kernels in ``csrc/`` behind the
C-ABI of ``include/frx.h`` (``libfrx.so``, loaded with ctypes by ``_lib``).  Host side mirrors
the reference's entry points: ``RasterFLACConverter``, ``SpatialFLACEncoder``,
``SpatialFLACStreamer``, ``SpatialIndex``, ``SpatialFrame`` and the CLI commands.
"""
__version__ = "0.1.0"

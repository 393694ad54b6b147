"""TEST INFRASTRUCTURE (never imported by the product): numpy restatement of what rasterio's read() does with an
uncompressed TIFF -- the checker for frx_tiff_unpack (include/frx.h).

Reference call sites: converter.py:137-141 (src.read()), spatial_encoder.py:154-183 (read(window=...)), cli.py:699.
rasterio/GDAL is a third-party dependency that is not installed here; what is restated is the TIFF 6.0 layout rule GDAL
follows: a chunk (strip or tile) stores its rows top to bottom, `pitch` pixels per row, samples of a pixel adjacent
(PlanarConfiguration 1) or one band per chunk (PlanarConfiguration 2), in the file's byte order; edge tiles are padded
and the padding is dropped.  Pinned by tests/test_oracle_golden.py against the reference's own fixtures
(test_data/*.tif decoded by PIL/libtiff) and against the built-in reader."""
import numpy as np


def unpack(file_bytes, chunks, dtype, bands, height, width, samples_per_pixel, swap_bytes=False):
    """file_bytes: bytes-like; chunks: [(src_offset, row0, col0, rows, cols, pitch, band)] -> (bands, height, width)."""
    dt = np.dtype(dtype)
    src_dt = dt.newbyteorder("S") if swap_bytes else dt
    buf = np.frombuffer(file_bytes, dtype=np.uint8)
    out = np.zeros((bands, height, width), dtype=dt)
    for (off, row0, col0, rows, cols, pitch, band) in chunks:
        if rows == 0 or cols == 0:
            continue
        nb = samples_per_pixel if band < 0 else 1
        n = ((rows - 1) * pitch + cols) * nb
        flat = np.frombuffer(buf[off:off + n * dt.itemsize].tobytes(), dtype=src_dt).astype(dt)
        full = np.zeros(rows * pitch * nb, dtype=dt)
        full[:n] = flat
        blk = full.reshape(rows, pitch, nb)[:, :cols, :]
        if band < 0:
            out[:nb, row0:row0 + rows, col0:col0 + cols] = blk.transpose(2, 0, 1)
        else:
            out[band, row0:row0 + rows, col0:col0 + cols] = blk[:, :, 0]
    return out

"""TIFF strips/tiles -> band-planar raster.  CPU: the chunk table + numpy restatement
(oracle/tiff_unpack.py) against the host reader and against PIL/libtiff on the reference's own fixtures.
GPU: frx_tiff_unpack (through the C-ABI) bit-exact against the same."""
import itertools
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(__file__))
from helpers import write_tiff_variant  # noqa: E402

from flac_raster_b200 import geotiff  # noqa: E402
from oracle import tiff_unpack  # noqa: E402

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")

LAYOUTS = [dict(rows_per_strip=7), dict(rows_per_strip=None), dict(tile=(32, 16)), dict(tile=(64, 64))]
CASES = [("uint8", 3), ("uint16", 3), ("int16", 1), ("float32", 2), ("int32", 1), ("uint16", 12)]


def _raster(dtype, bands, h=83, w=117, seed=0):
    rng = np.random.default_rng(seed)
    dt = np.dtype(dtype)
    if dt.kind == "f":
        return rng.standard_normal((bands, h, w)).astype(dt)
    info = np.iinfo(dt)
    return rng.integers(info.min, info.max, size=(bands, h, w), endpoint=True, dtype=np.int64).astype(dt)


def _variants():
    for (dtype, bands), layout, planar, be, pad in itertools.product(CASES, LAYOUTS, (1, 2), (False, True), (0, 3)):
        if pad and (be or planar == 2) and dtype != "uint16":
            continue   # keep the matrix small: unaligned chunks are covered once per layout
        yield dtype, bands, layout, planar, be, pad


@pytest.mark.parametrize("dtype,bands,layout,planar,be,pad", list(_variants()))
def test_chunk_table_and_restatement_match_host_reader(tmp_path, dtype, bands, layout, planar, be, pad):
    a = _raster(dtype, bands)
    p = tmp_path / "v.tif"
    write_tiff_variant(p, a, planar=planar, big_endian=be, pad=pad, **layout)
    with geotiff.open(p) as src:
        assert np.array_equal(src.read(), a)
        data = open(p, "rb").read()
        for (r0, r1) in ((0, None), (5, 60), (16, 17)):
            rr1 = src.height if r1 is None else r1
            got = tiff_unpack.unpack(data, src.chunk_table(r0, r1), dtype, bands, rr1 - r0, src.width, bands, swap_bytes=be)
            assert np.array_equal(got, a[:, r0:rr1]), (r0, r1)


@pytest.mark.parametrize("name", ["sample_dem.tif", "sample_rgb.tif", "sample_multispectral.tif"])
def test_restatement_matches_libtiff_on_reference_fixtures(name):
    """Pins the restatement on the reference's own files, decoded independently by PIL/libtiff."""
    PIL = pytest.importorskip("PIL.Image")
    path = os.path.join(GOLDEN, name)
    if not os.path.exists(path):
        pytest.skip(f"{name} not among the committed fixtures")
    with geotiff.open(path) as src:
        got = tiff_unpack.unpack(open(path, "rb").read(), src.chunk_table(), src.dtypes[0], src.count, src.height, src.width,
                                 src.count, swap_bytes=(src._bo == ">"))
        assert np.array_equal(got, src.read())
    if got.shape[0] in (1, 3):
        ref = np.asarray(PIL.open(path))
        ref = ref[None] if ref.ndim == 2 else ref.transpose(2, 0, 1)
        assert np.array_equal(got, ref)


@pytest.mark.gpu
@pytest.mark.parametrize("dtype,bands,layout,planar,be,pad", list(_variants()))
def test_gpu_unpack_bit_exact(tmp_path, dtype, bands, layout, planar, be, pad):
    import torch
    from flac_raster_b200.codec import Codec
    a = _raster(dtype, bands, seed=1)
    p = tmp_path / "v.tif"
    write_tiff_variant(p, a, planar=planar, big_endian=be, pad=pad, **layout)
    codec = Codec(0)
    with geotiff.open(p) as src:
        for (r0, r1) in ((0, None), (5, 60)):
            rr1 = src.height if r1 is None else r1
            dev = src.read_device(codec, r0, r1)
            torch.cuda.synchronize()
            got = dev.cpu().numpy() if dtype not in ("uint16", "uint32") else dev.view(torch.int16 if dtype == "uint16" else torch.int32).cpu().numpy().view(dtype)
            want = tiff_unpack.unpack(open(p, "rb").read(), src.chunk_table(r0, r1), dtype, bands, rr1 - r0, src.width, bands, swap_bytes=be)
            assert np.array_equal(want, a[:, r0:rr1])
            assert got.dtype == a.dtype and np.array_equal(got, want), (r0, r1)


@pytest.mark.gpu
def test_gpu_unpack_large_and_rejects_bad_chunks(tmp_path):
    import torch
    from flac_raster_b200._lib import FrxError
    from flac_raster_b200.codec import Codec
    a = _raster("uint16", 3, h=1500, w=4100, seed=2)     # rows longer than one staging segment
    p = tmp_path / "big.tif"
    write_tiff_variant(p, a, rows_per_strip=64)
    codec = Codec(0)
    with geotiff.open(p) as src:
        dev = src.read_device(codec)
        torch.cuda.synchronize()
        assert np.array_equal(dev.view(torch.int16).cpu().numpy().view(np.uint16), a)
    file_dev = torch.zeros(1024, dtype=torch.uint8, device="cuda")
    with pytest.raises(FrxError):   # reads past the file image
        codec.tiff_unpack(file_dev, [(512, 0, 0, 4, 64, 64, -1)], "uint16", 3, 4, 64, 3)
    with pytest.raises(FrxError):   # rectangle outside the raster
        codec.tiff_unpack(file_dev, [(0, 2, 0, 4, 8, 8, 0)], "uint8", 1, 4, 8, 1)

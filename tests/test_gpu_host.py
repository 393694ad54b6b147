"""GPU suite for the reference-facing host layer: the scenarios of tests/host_scenarios.py through libfrx
(C-ABI host entry points, device-resident batched decode), checked against the oracle and the golden fixtures."""
import numpy as np
import pytest

import host_scenarios as hs

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _need_gpu(codec):  # noqa: ARG001 - skips cleanly when there is no device
    yield


def test_convert_rgb_fixture_roundtrip(tmp_path):
    hs.convert_rgb_fixture_roundtrip(tmp_path)


def test_convert_writes_md5_when_asked(tmp_path):
    hs.convert_writes_md5_when_asked(tmp_path)


def test_convert_reference_fixture_decodes(tmp_path):
    hs.convert_reference_fixture_decodes(tmp_path)


def test_convert_dem_c1_roundtrip(tmp_path):
    hs.convert_dem_c1_roundtrip(tmp_path)


@pytest.mark.parametrize("dtype,bands,mode", [("uint16", 3, "first"), ("uint16", 3, "all"), ("float32", 1, "first"),
                                              ("uint8", 2, "all"), ("int16", 1, "first")])
def test_streaming_roundtrip(tmp_path, dtype, bands, mode):
    hs.streaming_roundtrip(tmp_path, dtype, bands, mode)


def test_streaming_two_shards_equal_one(tmp_path):
    (tmp_path / "a").mkdir(); (tmp_path / "b").mkdir()
    one, _, _ = hs.streaming_roundtrip(tmp_path / "a", "uint16", 3, "first", devices=(0,))
    two, _, _ = hs.streaming_roundtrip(tmp_path / "b", "uint16", 3, "first", devices=(0, 0))
    assert one.read_bytes() == two.read_bytes()


def test_streaming_batch_queries(tmp_path):
    hs.streaming_batch_queries(tmp_path)


@pytest.mark.parametrize("bands", [12, 3])
def test_spatial_roundtrip(tmp_path, bands):
    hs.spatial_roundtrip(tmp_path, bands)


def test_spatial_rejects_per_tile_stretch(tmp_path):
    hs.spatial_rejects_per_tile_stretch(tmp_path)


def test_spatial_reference_mode(tmp_path):
    hs.spatial_reference_mode_reproduces_fixture(tmp_path)


def test_pyflac_shim(tmp_path):
    hs.pyflac_shim_semantics(tmp_path)


def test_cli(tmp_path):
    hs.cli_commands(tmp_path)


def test_unmodified_reference_converter_over_the_shim(tmp_path):
    """SURVEY 8(b), GPU side: runs where /root/reference is mounted (not on the GPU box)."""
    import reference_dropin as rd
    if not rd.available():
        pytest.skip("/root/reference is not mounted on this machine")
    rd.reference_converter_roundtrip(tmp_path)


def test_streaming_band1_of_planar_file_uploads_one_band(tmp_path, monkeypatch):
    """create-streaming in the reference's band-1 mode on a band-planar TIFF (PlanarConfiguration 2): only band 1's strips go
    to the device, and the file written equals the one made from the pixel-interleaved layout of the same raster."""
    from helpers import write_tiff_variant
    from flac_raster_b200 import geotiff, streaming
    arr = hs.synthetic(3, 300, 420, "uint16")
    write_tiff_variant(tmp_path / "planar.tif", arr, rows_per_strip=7, planar=2)
    write_tiff_variant(tmp_path / "chunky.tif", arr, rows_per_strip=7, planar=1)
    sent = []
    real = geotiff.DatasetReader._upload_and_unpack

    def spy(self, codec, chunks, count, spp, row0, row1):
        sent.append((count, sum(c[3] * c[4] for c in chunks)))
        return real(self, codec, chunks, count, spp, row0, row1)
    monkeypatch.setattr(geotiff.DatasetReader, "_upload_and_unpack", spy)
    streaming.create_streaming(tmp_path / "planar.tif", tmp_path / "p.flac", tile_size=128, bands="first")
    assert sent == [(1, 300 * 420)]                       # one band's pixels, not three
    streaming.create_streaming(tmp_path / "chunky.tif", tmp_path / "c.flac", tile_size=128, bands="first")
    assert (tmp_path / "p.flac").read_bytes() == (tmp_path / "c.flac").read_bytes()

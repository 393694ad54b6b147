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


def test_spatial_reference_mode(tmp_path):
    hs.spatial_reference_mode_reproduces_fixture(tmp_path)


def test_pyflac_shim(tmp_path):
    hs.pyflac_shim_semantics(tmp_path)


def test_cli(tmp_path):
    hs.cli_commands(tmp_path)

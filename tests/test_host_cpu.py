"""CPU suite for the reference-facing host layer: tests/host_scenarios.py on top of the oracle-backed fake codec
(tests/fake_codec.py), plus the pure-host pieces (index selection, sharding, world_size-2 gloo merge)."""
import json
import os
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

import host_scenarios as hs
from fake_codec import OracleCodec

ROOT = Path(__file__).resolve().parents[1]


@pytest.fixture()
def fake(monkeypatch):
    import flac_raster_b200.converter as conv
    import flac_raster_b200.pyflac_shim as shim
    import flac_raster_b200.spatial_encoder as sp
    import flac_raster_b200.streaming as st
    c = OracleCodec()
    for m in (conv, sp, st, shim):
        monkeypatch.setattr(m, "get_codec", lambda device=0: c)
    monkeypatch.setattr(st, "_make_codec", lambda dev: OracleCodec())
    monkeypatch.setattr(st.StreamingFile, "decode", _host_decode)
    return c


def _host_decode(self, positions):
    """StreamingFile.decode without torch/CUDA: same plan, host entry point of the (fake) codec."""
    from flac_raster_b200.converter import BLOCKSIZE
    from flac_raster_b200 import streaming
    if len(positions) == 0:
        return []
    dt, table, oe, metas = self.plan(positions)
    out = np.empty(oe, dtype=np.dtype(metas[0]["dtype"]))
    streaming.get_codec(self.device).decode_tiles_host(self.data, dt, out, frame_size=table, blocksize=BLOCKSIZE,
                                                       pcm_mode=metas[0].get("pcm_mode", "reference"))
    res, p = [], 0
    for md in metas:
        n = md["count"] * md["height"] * md["width"]
        res.append(out[p:p + n].reshape(md["count"], md["height"], md["width"]))
        p += n
    return res


def test_convert_rgb_fixture_roundtrip(fake, tmp_path):
    hs.convert_rgb_fixture_roundtrip(tmp_path)


def test_convert_writes_md5_when_asked(fake, tmp_path):
    hs.convert_writes_md5_when_asked(tmp_path)


def test_convert_reference_fixture_decodes(fake, tmp_path):
    hs.convert_reference_fixture_decodes(tmp_path)


def test_convert_dem_c1_roundtrip(fake, tmp_path):
    hs.convert_dem_c1_roundtrip(tmp_path)


@pytest.mark.parametrize("dtype,bands,mode", [("uint16", 3, "first"), ("uint16", 3, "all"), ("float32", 1, "first")])
def test_streaming_roundtrip(fake, tmp_path, dtype, bands, mode):
    hs.streaming_roundtrip(tmp_path, dtype, bands, mode)


def test_streaming_two_devices_equals_one(fake, tmp_path):
    (tmp_path / "a").mkdir(); (tmp_path / "b").mkdir()
    one, _, _ = hs.streaming_roundtrip(tmp_path / "a", "uint16", 3, "first", devices=(0,))
    two, _, _ = hs.streaming_roundtrip(tmp_path / "b", "uint16", 3, "first", devices=(0, 0))
    assert one.read_bytes() == two.read_bytes()


def test_streaming_batch_queries(fake, tmp_path):
    hs.streaming_batch_queries(tmp_path)


@pytest.mark.parametrize("bands", [12, 3])
def test_spatial_roundtrip(fake, tmp_path, bands):
    hs.spatial_roundtrip(tmp_path, bands)


def test_spatial_rejects_per_tile_stretch(fake, tmp_path):
    hs.spatial_rejects_per_tile_stretch(tmp_path)


def test_spatial_reference_mode(fake, tmp_path):
    hs.spatial_reference_mode_reproduces_fixture(tmp_path)


def test_pyflac_shim(fake, tmp_path):
    hs.pyflac_shim_semantics(tmp_path)


def test_cli(fake, tmp_path):
    hs.cli_commands(tmp_path)


def test_unmodified_reference_converter_over_the_shim(fake, tmp_path):
    """the reference's own converter.py, unmodified, on top of the pyflac-shaped shim."""
    import reference_dropin as rd
    if not rd.available():
        pytest.skip("/root/reference is not mounted on this machine")
    rd.reference_converter_roundtrip(tmp_path)


def test_shard_ranges_cover_and_are_contiguous():
    from flac_raster_b200.streaming import shard_range
    for n in (0, 1, 7, 400, 1600, 1601):
        for w in (1, 2, 3, 4, 8):
            rs = [shard_range(n, w, r) for r in range(w)]
            assert rs[0][0] == 0 and rs[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(rs, rs[1:]))
            assert max(hi - lo for lo, hi in rs) <= (n + w - 1) // w


def test_select_frames_batch_matches_reference_loop():
    from flac_raster_b200 import streaming
    from flac_raster_b200.rio import from_origin
    from flac_raster_b200.spatial_encoder import calculate_tiles
    tiles = calculate_tiles(1000, 1300, 256)
    idx = streaming.build_index("EPSG:4326", from_origin(-120.0, 37.0, 1e-4, 1e-4), 1300, 1000, 256, tiles,
                                np.arange(len(tiles)) * 10, [10] * len(tiles))
    rng = np.random.default_rng(0)
    qs = [{"tile_id": int(rng.integers(0, 25))} for _ in range(50)]
    for _ in range(200):
        x, y = -120.0 + rng.uniform(-0.01, 0.14), 37.0 - rng.uniform(-0.01, 0.11)
        qs.append({"bbox": [x, y - rng.uniform(0, 0.03), x + rng.uniform(0, 0.03), y]})
    pos = streaming.select_frames_batch(idx, qs)
    for q, p in zip(qs, pos):
        ref = streaming.select_frame(idx, q.get("tile_id"), False, False, q.get("bbox"))
        assert (ref is None and p == -1) or idx["frames"][p] is ref


_GLOO = r'''
import os, sys, pickle
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, os.path.join(sys.argv[1], "tests"))
import numpy as np, torch.distributed as dist
import flac_raster_b200.streaming as st
from fake_codec import OracleCodec
from host_scenarios import synthetic
from flac_raster_b200.rio import from_origin
from flac_raster_b200.spatial_encoder import calculate_tiles
c = OracleCodec(); st.get_codec = lambda device=0: c
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
arr = synthetic(2, 200, 260, "uint16")
tiles = calculate_tiles(200, 260, 64)
tr = from_origin(-120.0, 37.0, 1e-4, 1e-4)
lo, hi = st.shard_range(len(tiles), world, rank)
# each rank holds only the raster rows its tile range touches
r0 = min(t[0] for t in tiles[lo:hi]); r1 = max(t[0] + t[2] for t in tiles[lo:hi])
mine = st.encode_streaming_tiles(np.ascontiguousarray(arr[:, r0:r1]), tiles[lo:hi], "EPSG:4326", tr, row_base=r0)
gathered = [None] * world
dist.gather_object(mine, gathered if rank == 0 else None, dst=0)   # host-side concatenation; no device collective
if rank == 0:
    chunks, offsets, sizes = st.merge_shards(gathered)
    whole = st.encode_streaming_tiles(arr, tiles, "EPSG:4326", tr)
    assert chunks == whole[0] and sizes == whole[1]
    assert list(offsets) == [sum(sizes[:i]) for i in range(len(sizes))]
    print("GLOO_OK", len(chunks))
dist.destroy_process_group()
'''


def test_two_rank_gloo_sharding_equals_single_rank(tmp_path):
    """contiguous tile ranges per rank, host-side offset concatenation, world_size 2 on gloo."""
    script = tmp_path / "gloo_shard.py"
    script.write_text(_GLOO)
    env = dict(os.environ, OMP_NUM_THREADS="2")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", "29653", str(script), str(ROOT)], capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0 and "GLOO_OK 20" in r.stdout, r.stdout[-2000:] + r.stderr[-3000:]

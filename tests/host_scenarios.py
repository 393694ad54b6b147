"""Scenarios of the reference-facing host layer (converter / spatial encoder / streaming / CLI / pyflac shim).
The same functions run twice: in the CPU suite on top of tests/fake_codec.OracleCodec (host logic only) and in
the GPU suite on top of libfrx (tests/test_gpu_host.py).  Expected values come from the golden fixtures and
from oracle/ (the checker)."""
import json
import struct
from pathlib import Path

import numpy as np

import oracle
from oracle import quant
from flac_raster_b200 import flacmeta, rio, streaming
from flac_raster_b200.converter import RasterFLACConverter
from flac_raster_b200.spatial_encoder import SpatialFLACEncoder, SpatialFLACStreamer, decode_spatial_file

GOLDEN = Path(__file__).resolve().parent / "golden"


def read_tif(path):
    with rio.open(path) as src:
        return src.read(), src


def write_tif(path, arr, transform=None, crs="EPSG:4326"):
    transform = transform or rio.from_origin(-120.0, 37.0, 1e-4, 1e-4)
    with rio.open(path, "w", driver="GTiff", width=arr.shape[2], height=arr.shape[1], count=arr.shape[0], dtype=arr.dtype,
                  crs=rio.CRS.from_string(crs), transform=transform) as dst:
        dst.write(arr)


def synthetic(bands, h, w, dtype, seed=3):
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w]
    out = []
    for b in range(bands):
        z = 1200 + 500 * np.sin(xx / (11.0 + b)) * np.cos(yy / (17.0 + 2 * b)) + rng.normal(0, 5, (h, w))
        out.append(z)
    a = np.stack(out)
    if np.dtype(dtype) == np.float32:
        return (a / 8192.0).astype(np.float32)
    if np.dtype(dtype) == np.uint8:
        return np.clip(a / 8, 0, 255).astype(np.uint8)
    return a.astype(dtype)


# ---- convert (config C1 and fixture G1-G3) --------------------------------------------------------
def convert_rgb_fixture_roundtrip(tmp):
    """sample_rgb.tif -> FLAC -> TIFF must reproduce sample_rgb_reconstructed.tif's pixels (G3), the embedded
    metadata must carry the reference's key set, and the stream must be libFLAC-grade in size (G1)."""
    flac, tif = tmp / "rgb.flac", tmp / "rgb_back.tif"
    conv = RasterFLACConverter()
    assert conv.tiff_to_flac(GOLDEN / "sample_rgb.tif", flac) is None
    blob = flac.read_bytes()
    hdr = flacmeta.parse_header(blob)
    keys = [k for k, _ in hdr.comments]
    assert keys[:14] == ["TITLE", "DESCRIPTION", "ENCODER", "GEOSPATIAL_CRS", "GEOSPATIAL_WIDTH", "GEOSPATIAL_HEIGHT",
                         "GEOSPATIAL_COUNT", "GEOSPATIAL_DTYPE", "GEOSPATIAL_NODATA", "GEOSPATIAL_DATA_MIN",
                         "GEOSPATIAL_DATA_MAX", "GEOSPATIAL_TRANSFORM", "GEOSPATIAL_BOUNDS", "GEOSPATIAL_SPATIAL_TILING"]
    assert hdr.get("GEOSPATIAL_DATA_MIN") == "1.0" and hdr.get("GEOSPATIAL_DATA_MAX") == "255.0"
    assert hdr.get("GEOSPATIAL_NODATA") == "None" and hdr.get("GEOSPATIAL_SPATIAL_TILING") == "False"
    assert len(json.loads(hdr.get("GEOSPATIAL_TRANSFORM"))) == 9
    si = hdr.streaminfo
    assert (si.channels, si.bps, si.sample_rate, si.min_blocksize, si.total_samples) == (3, 16, 44100, 4096, 65536)
    # independent decoder: whole file is one valid stream whose PCM is the fixture's PCM (G1/G2)
    pcm, sid, frames, used, first, rc = oracle.decode_stream(blob)
    assert rc == 0 and used == len(blob) and all(f["crc8_ok"] and f["crc16_ok"] for f in frames)
    ref_pcm = oracle.decode_stream((GOLDEN / "sample_rgb.flac").read_bytes())[0]
    assert np.array_equal(pcm, ref_pcm)
    frame_bytes = len(blob) - first
    assert frame_bytes <= 1.02 * 178857, frame_bytes
    conv.flac_to_tiff(flac, tif)
    got, _ = read_tif(tif)
    want, src = read_tif(GOLDEN / "sample_rgb_reconstructed.tif")
    assert got.dtype == want.dtype and np.array_equal(got, want)
    back, bsrc = read_tif(tif)
    assert tuple(bsrc.transform)[:6] == tuple(src.transform)[:6] and str(bsrc.crs) == str(src.crs)


def convert_writes_md5_when_asked(tmp):
    """STREAMINFO's MD5: opt-in; equals the MD5 of the reference PCM's little-endian bytes -- for sample_rgb.tif
    that is the fingerprint of the libFLAC-written fixture's own PCM."""
    import hashlib
    flac = tmp / "rgb_md5.flac"
    RasterFLACConverter(md5=True).tiff_to_flac(GOLDEN / "sample_rgb.tif", flac)
    si = flacmeta.parse_header(flac.read_bytes()).streaminfo
    assert si.md5.hex() == "4bf5483c1452f8c591c813b52db99491"
    pcm = oracle.decode_stream(flac.read_bytes())[0]
    assert hashlib.md5(pcm.astype("<i2").tobytes()).digest() == si.md5
    RasterFLACConverter().tiff_to_flac(GOLDEN / "sample_rgb.tif", tmp / "rgb_nomd5.flac")
    assert flacmeta.parse_header((tmp / "rgb_nomd5.flac").read_bytes()).streaminfo.md5 == b"\0" * 16     # the reference's default


def convert_reference_fixture_decodes(tmp):
    """flac_to_tiff on the reference-written sample_rgb.flac (+ its .json sidecar, no embedded comments, no
    frame table -> frame discovery) reproduces sample_rgb_reconstructed.tif."""
    import shutil
    shutil.copy(GOLDEN / "sample_rgb.flac", tmp / "sample_rgb.flac")
    shutil.copy(GOLDEN / "sample_rgb.json", tmp / "sample_rgb.json")
    RasterFLACConverter().flac_to_tiff(tmp / "sample_rgb.flac", tmp / "out.tif")
    got, _ = read_tif(tmp / "out.tif")
    want, _ = read_tif(GOLDEN / "sample_rgb_reconstructed.tif")
    assert np.array_equal(got, want)


def convert_dem_c1_roundtrip(tmp):
    """Config C1: sample_dem.tif round trip == the reference's own round trip (oracle restatement)."""
    src_arr, _ = read_tif(GOLDEN / "sample_dem.tif")
    conv = RasterFLACConverter()
    conv.tiff_to_flac(GOLDEN / "sample_dem.tif", tmp / "dem.flac")
    conv.flac_to_tiff(tmp / "dem.flac", tmp / "dem_back.tif")
    got, _ = read_tif(tmp / "dem_back.tif")
    pcm, bps, dmin, dmax, _ = quant.reference_pcm(src_arr)
    want = quant.reference_decode_raster(pcm, bps, src_arr.shape[1], src_arr.shape[2], 1, src_arr.dtype, dmin, dmax)
    assert np.array_equal(got, want.reshape(got.shape))


# ---- streaming format (configs C2/C3/C5) ----------------------------------------------------------
def streaming_roundtrip(tmp, dtype="uint16", bands=3, mode_bands="first", devices=(0,)):
    arr = synthetic(bands, 300, 420, dtype)
    write_tif(tmp / "in.tif", arr)
    out = tmp / "s.flac"
    index = streaming.create_streaming(tmp / "in.tif", out, tile_size=128, bands=mode_bands, devices=devices)
    raw = out.read_bytes()
    n = struct.unpack(">I", raw[:4])[0]
    idx = json.loads(raw[4:4 + n].decode())
    assert idx == json.loads(json.dumps(index))
    assert list(idx.keys()) == ["crs", "transform", "width", "height", "tile_size", "frames"]
    assert list(idx["frames"][0].keys()) == ["frame_id", "bbox", "window", "byte_offset", "byte_size"]
    assert list(idx["frames"][0]["window"].keys()) == ["col_off", "row_off", "width", "height"]
    assert len(idx["frames"]) == 3 * 4 and idx["frames"][-1]["window"] == {"col_off": 384, "row_off": 256, "width": 36, "height": 44}
    nch = 1 if mode_bands == "first" else bands
    pos = 0
    for fr in idx["frames"]:
        assert fr["byte_offset"] == pos
        pos += fr["byte_size"]
        wd = fr["window"]
        tile = raw[4 + n + fr["byte_offset"]: 4 + n + fr["byte_offset"] + fr["byte_size"]]
        assert tile[:4] == b"fLaC"
        block = arr[:nch, wd["row_off"]:wd["row_off"] + wd["height"], wd["col_off"]:wd["col_off"] + wd["width"]]
        pcm_ref, bps, dmin, dmax, _ = quant.reference_pcm(block)
        pcm, sid, frames, used, first, rc = oracle.decode_stream(tile)
        assert rc == 0 and used == len(tile) and np.array_equal(pcm, pcm_ref)
        hdr = flacmeta.parse_header(tile)
        assert float(hdr.get("GEOSPATIAL_DATA_MIN")) == dmin and float(hdr.get("GEOSPATIAL_DATA_MAX")) == dmax
        assert int(hdr.get("GEOSPATIAL_WIDTH")) == wd["width"] and int(hdr.get("GEOSPATIAL_COUNT")) == nch
    assert pos == len(raw) - 4 - n
    # extract: single query per call (reference shape), every selector
    want_of = {}
    for fr in idx["frames"]:
        wd = fr["window"]
        block = arr[:nch, wd["row_off"]:wd["row_off"] + wd["height"], wd["col_off"]:wd["col_off"] + wd["width"]]
        pcm_ref, bps, dmin, dmax, _ = quant.reference_pcm(block)
        want_of[fr["frame_id"]] = quant.reference_decode_raster(pcm_ref, bps, wd["height"], wd["width"], nch, arr.dtype, dmin,
                                                                dmax).reshape(nch, wd["height"], wd["width"])
    for kw, fid in (({"tile_id": 5}, 5), ({"last": True}, 11), ({"center": True}, None),
                    ({"bbox": [-119.99, 36.98, -119.985, 36.99]}, None)):
        fr = streaming.extract_streaming(out, tmp / "t.tif", **kw)
        if fid is not None:
            assert fr["frame_id"] == fid
        got, gsrc = read_tif(tmp / "t.tif")
        assert np.array_equal(got, want_of[fr["frame_id"]])
        assert abs(gsrc.transform.c - fr["bbox"][0]) < 1e-12 and abs(gsrc.transform.f - fr["bbox"][3]) < 1e-12
    return out, idx, want_of


def streaming_batch_queries(tmp):
    """C5 shape: many tile-id / bbox queries in one batched decode == per-query answers of the reference loop."""
    out, idx, want_of = streaming_roundtrip(tmp, "uint16", 3, "first")
    rng = np.random.default_rng(5)
    queries = []
    for k in range(40):
        if k % 2 == 0:
            queries.append({"tile_id": int(rng.integers(0, 12))})
        else:
            px, py = rng.integers(0, 410), rng.integers(0, 290)
            x0, y1 = -120.0 + px * 1e-4, 37.0 - py * 1e-4
            queries.append({"bbox": [x0, y1 - 10e-4, x0 + 10e-4, y1]})
    sf = streaming.StreamingFile(out)
    pos, arrays = sf.query(queries)
    for q, p, a in zip(queries, pos, arrays):
        ref = streaming.select_frame(idx, q.get("tile_id"), False, False, q.get("bbox"))
        assert ref is not None and idx["frames"][p] == ref
        assert np.array_equal(a, want_of[ref["frame_id"]])


# ---- --spatial raw-frames format (config C4) ------------------------------------------------------
def spatial_roundtrip(tmp, bands=12, dtype="int16"):
    arr = synthetic(bands, 150, 200, dtype, seed=9)
    write_tif(tmp / "ms.tif", arr)
    out = tmp / "ms.flac"
    index = SpatialFLACEncoder(tile_size=64).encode_spatial_flac(tmp / "ms.tif", out)
    raw = out.read_bytes()
    assert len(index.frames) == 3 * 4
    s = SpatialFLACStreamer(out)
    assert json.loads(json.dumps([f.to_dict() for f in s.spatial_index.frames])) == json.loads(json.dumps([f.to_dict() for f in index.frames]))
    # offsets are real: every tile (and every band-group stream) starts with a FLAC marker
    end = 0
    for fr in s.spatial_index.frames:
        assert raw[fr.byte_offset:fr.byte_offset + 4] == b"fLaC" and fr.byte_offset == end
        end = fr.byte_offset + fr.byte_size
        for st in (fr.streams or []):
            assert raw[st["byte_offset"]:st["byte_offset"] + 4] == b"fLaC"
    assert end == len(raw)
    if bands > 8:
        assert [st["bands"] for st in s.spatial_index.frames[0].streams] == [[0, 8], [8, bands - 8]]
    hdr = s.header
    assert hdr.get("GEOSPATIAL_SPATIAL_TILING") == "true" and hdr.get("GEOSPATIAL_NUM_TILES") == "12"
    assert hdr.get("GEOSPATIAL_PCM_MODE") == "lossless"
    # byte ranges: same as the reference algorithm on the same index
    fb = [f.bbox for f in s.spatial_index.frames]
    offs = [f.byte_offset for f in s.spatial_index.frames]
    szs = [f.byte_size for f in s.spatial_index.frames]
    for bbox in ([-120.0, 36.99, -119.99, 37.0], [-119.995, 36.985, -119.985, 36.995], [-121, 30, -119, 38], [0, 0, 1, 1]):
        assert s.get_byte_ranges_for_bbox(bbox) == quant.byte_ranges_for_bbox(fb, offs, szs, bbox)
        data = s.stream_bbox_data(bbox)
        assert len(data) == sum(e - st + 1 for st, e in s.get_byte_ranges_for_bbox(bbox))
    # lossless: whole raster comes back bit-exact, and so does a bbox subset
    assert np.array_equal(decode_spatial_file(out), arr)
    for fr, a in s.decode_bbox([-119.995, 36.99, -119.99, 36.995]):
        wd = fr.window
        assert np.array_equal(a, arr[:, wd.row_off:wd.row_off + wd.height, wd.col_off:wd.col_off + wd.width])
    RasterFLACConverter().flac_to_tiff(out, tmp / "ms_back.tif")
    assert np.array_equal(read_tif(tmp / "ms_back.tif")[0], arr)


def spatial_rejects_per_tile_stretch(tmp):
    """converter.py's min/max stretch would be applied per tile while the file keeps one global DATA_MIN/MAX:
    the spatial encoder refuses it, from the API and from the CLI."""
    import pytest
    from typer.testing import CliRunner
    from flac_raster_b200.cli import app
    with pytest.raises(ValueError, match="not available for spatial files"):
        SpatialFLACEncoder(tile_size=64, pcm_mode="reference")
    import shutil
    shutil.copy(GOLDEN / "sample_dem.tif", tmp / "dem.tif")
    r = CliRunner().invoke(app, ["convert", str(tmp / "dem.tif"), "-o", str(tmp / "sp.flac"), "--spatial", "--pcm-mode", "reference"])
    assert r.exit_code == 1 and not (tmp / "sp.flac").exists()


def spatial_reference_mode_reproduces_fixture(tmp):
    """pcm_mode=reference_spatial on sample_dem.tif: the reference's effective PCM is all zeros at bps 32 (SURVEY 0.2,
    fixture G4).  Our streams must decode to exactly the fixture's PCM and never be larger than the fixture's frames
    (libFLAC 1.4.3 spends 523 B/frame on them; a CONSTANT subframe is legal and smaller)."""
    enc = SpatialFLACEncoder(tile_size=256, pcm_mode="reference_spatial")
    arr, _ = read_tif(GOLDEN / "sample_dem.tif")
    tiles, groups, res = enc.encode_array(arr)
    assert len(tiles) == 4 and res.bps == 32
    fixture = (GOLDEN / "sample_dem.flac").read_bytes()
    want = oracle.decode_stream(fixture)[0]
    assert want.shape == (65536, 1) and not want.any()
    for i in range(4):
        off, size = int(res.tile_offset[i]), int(res.tile_size[i])
        assert size <= 16 * 523
        pcm, frames, rc = oracle.decode_frames(bytes(res.bytes[off:off + size]), 1, 32, max_samples=65536 + 16)
        assert rc == 0 and len(frames) == 16 and np.array_equal(pcm, want)


# ---- pyflac-shaped shim -----------------------------------------------------------------------------
def pyflac_shim_semantics(tmp):
    from flac_raster_b200 import pyflac_shim as pf
    rng = np.random.default_rng(1)
    pcm = (rng.normal(0, 900, (10000, 3)).cumsum(axis=0) % 20000 - 10000).astype(np.int16)
    chunks = []
    enc = pf.StreamEncoder(write_callback=lambda b, n, s, f: chunks.append((bytes(b), n, s, f)), sample_rate=44100,
                           compression_level=5, blocksize=4096)
    enc._channels, enc._bits_per_sample = 9, 24  # the reference sets these; process() overrides them
    enc.process(pcm)
    assert enc.finish()
    assert [c[2] for c in chunks[:2]] == [0, 0] and chunks[0][0][:4] == b"fLaC"
    assert [c[2] for c in chunks[2:]] == [4096, 4096, 1808] and [c[3] for c in chunks[2:]] == [0, 1, 2]
    assert all(len(c[0]) == c[1] for c in chunks)
    blob = b"".join(c[0] for c in chunks)
    got, sid, frames, used, first, rc = oracle.decode_stream(blob)
    assert rc == 0 and (sid["channels"], sid["bps"]) == (3, 16) and np.array_equal(got, pcm.astype(np.int32))
    (tmp / "a.flac").write_bytes(blob)
    audio, sr = pf.FileDecoder(str(tmp / "a.flac")).process()
    assert sr == 44100 and audio.dtype == np.float64 and np.array_equal(audio, pcm.astype(np.float64) / 32768.0)
    # float32 input truncates toward zero and becomes a bps-32 stream (what --spatial does in the reference)
    chunks.clear()
    enc = pf.StreamEncoder(write_callback=lambda b, n, s, f: chunks.append(bytes(b)), sample_rate=44100, blocksize=4096)
    enc.process(np.linspace(-1.0, 1.0, 5000, dtype=np.float32).reshape(-1, 1))
    enc.finish()
    got, sid = oracle.decode_stream(b"".join(chunks))[:2]
    assert sid["bps"] == 32 and set(np.unique(got)) <= {-1, 0, 1} and got[0, 0] == -1 and got[-1, 0] == 1
    # the reference's 12-band case: libFLAC refuses > 8 channels at init
    enc = pf.StreamEncoder(write_callback=lambda *a: None, sample_rate=44100, blocksize=4096)
    try:
        enc.process(np.zeros((100, 12), dtype=np.int16))
        raise AssertionError("expected EncoderInitException")
    except pf.EncoderInitException:
        pass
    try:
        enc.process([1, 2, 3])
        raise AssertionError("expected TypeError")
    except TypeError:
        pass
    # reference-written file through the shim decoder == the oracle's restatement of FileDecoder
    audio, sr = pf.FileDecoder(str(GOLDEN / "sample_rgb.flac")).process()
    ref = oracle.decode_stream((GOLDEN / "sample_rgb.flac").read_bytes())[0]
    assert np.array_equal(audio, quant.filedecoder_output(ref, 16))


# ---- CLI -----------------------------------------------------------------------------------------
def cli_commands(tmp):
    from typer.testing import CliRunner
    from flac_raster_b200.cli import app
    run = CliRunner().invoke
    import shutil
    shutil.copy(GOLDEN / "sample_rgb.tif", tmp / "rgb.tif")
    r = run(app, ["convert", str(tmp / "rgb.tif")])
    assert r.exit_code == 0 and (tmp / "rgb.flac").exists(), r.output
    r = run(app, ["convert", str(tmp / "rgb.tif")])
    assert r.exit_code == 1 and "already exists" in r.output                       # overwrite guard, cli.py:74-77
    r = run(app, ["convert", str(tmp / "rgb.tif"), "-f", "-c", "8"])
    assert r.exit_code == 0
    r = run(app, ["convert", str(tmp / "rgb.flac"), "-o", str(tmp / "back.tif")])
    assert r.exit_code == 0
    assert np.array_equal(read_tif(tmp / "back.tif")[0], read_tif(GOLDEN / "sample_rgb_reconstructed.tif")[0])
    r = run(app, ["convert", str(tmp / "nope.tif")])
    assert r.exit_code == 1
    (tmp / "x.txt").write_text("x")
    assert run(app, ["convert", str(tmp / "x.txt")]).exit_code == 1                 # unsupported suffix, cli.py:64-67
    r = run(app, ["create-streaming", str(tmp / "rgb.tif"), "--tile-size", "128"])
    assert r.exit_code == 0 and (tmp / "rgb_streaming.flac").exists(), r.output
    r = run(app, ["extract-streaming", str(tmp / "rgb_streaming.flac"), "--tile-id", "3", "-o", str(tmp / "t3.tif")])
    assert r.exit_code == 0, r.output
    assert read_tif(tmp / "t3.tif")[0].shape == (1, 128, 128)
    assert run(app, ["extract-streaming", str(tmp / "rgb_streaming.flac"), "--tile-id", "99", "-o", str(tmp / "t.tif")]).exit_code == 1
    assert run(app, ["extract-streaming", str(tmp / "rgb_streaming.flac"), "-o", str(tmp / "t.tif")]).exit_code == 1
    r = run(app, ["convert", str(tmp / "rgb.tif"), "-o", str(tmp / "sp.flac"), "--spatial", "--tile-size", "128"])
    assert r.exit_code == 0 and "4 tiles" in r.output, r.output
    r = run(app, ["query", str(tmp / "sp.flac"), "--bbox", "-180,-90,180,90", "-o", str(tmp / "ranges.json")])
    assert r.exit_code == 0, r.output
    rj = json.loads((tmp / "ranges.json").read_text())
    assert rj["total_ranges"] == 1 and rj["total_bytes"] == (tmp / "sp.flac").stat().st_size
    r = run(app, ["query", str(tmp / "sp.flac"), "--bbox", "-180,-90,180,90", "-f", "data", "-o", str(tmp / "d.bin")])
    assert r.exit_code == 0 and (tmp / "d.bin").read_bytes() == (tmp / "sp.flac").read_bytes()
    assert run(app, ["query", str(tmp / "sp.flac"), "--bbox", "1,2,3"]).exit_code == 1
    assert run(app, ["query", str(tmp / "rgb.flac"), "--bbox", "0,0,1,1"]).exit_code == 1   # no spatial index
    # inspection commands around the path (cli.py:96-290, 409-617, 807-874)
    r = run(app, ["info", str(tmp / "rgb.tif")])
    assert r.exit_code == 0 and "Dimensions: 256 x 256" in r.output and "Bands: 3" in r.output, r.output
    r = run(app, ["info", str(tmp / "rgb.flac")])
    assert r.exit_code == 0 and "Channels: 3" in r.output and "(65536, 3)" in r.output and "Original dtype: uint8" in r.output, r.output
    assert run(app, ["info", str(tmp / "nope.flac")]).exit_code == 1
    assert run(app, ["info", str(tmp / "x.txt")]).exit_code == 1
    r = run(app, ["compare", str(tmp / "rgb.tif"), str(tmp / "back.tif"), "-e", str(tmp / "cmp.json")])
    assert r.exit_code == 0, r.output
    cj = json.loads((tmp / "cmp.json").read_text())
    assert cj["shape_match"] and cj["arrays_equal"] and cj["max_difference"] == 0.0 and len(cj["bands"]) == 3
    assert run(app, ["compare", str(tmp / "rgb.tif"), str(tmp / "rgb.flac")]).exit_code == 1       # not a TIFF, cli.py:267-273
    r = run(app, ["spatial-info", str(tmp / "sp.flac")])
    assert r.exit_code == 0 and "Total frames/tiles: 4" in r.output, r.output
    assert run(app, ["spatial-info", str(tmp / "rgb.flac")]).exit_code == 1                          # no spatial index
    from flac_raster_b200.spatial_encoder import SpatialFLACStreamer
    fr = SpatialFLACStreamer(tmp / "sp.flac").spatial_index.frames[1]
    x0, y0, x1, y1 = fr.bbox
    inner = f"{x0 + (x1 - x0) * 0.25},{y0 + (y1 - y0) * 0.25},{x0 + (x1 - x0) * 0.5},{y0 + (y1 - y0) * 0.5}"
    r = run(app, ["extract-tile", str(tmp / "sp.flac"), "--bbox", inner, "-o", str(tmp / "tile1.tif")])
    assert r.exit_code == 0 and "Using frame 1" in r.output, r.output
    w = fr.window
    want = read_tif(GOLDEN / "sample_rgb.tif")[0][0, w.row_off:w.row_off + w.height, w.col_off:w.col_off + w.width]
    assert np.array_equal(read_tif(tmp / "tile1.tif")[0][0], want)            # --spatial defaults to the lossless mapping
    assert run(app, ["extract-tile", str(tmp / "sp.flac"), "--bbox", "1000,1000,1001,1001", "-o", str(tmp / "t.tif")]).exit_code == 1

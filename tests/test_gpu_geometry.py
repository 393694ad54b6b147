"""Parity at the BASELINE configs' REAL tile geometry ("what's weak" 1): the rasters are crops of the
benchmark generators that keep every tile shape the full-size configs contain, so the paths the toy cases never
reach are covered -- 4096-sample frames that are exactly 4 rows of 1024 (C2/C3 interior), 7.53 rows of 544 (C2
edge), 64- and 16-wide edge tiles (C3/C4), 12 bands in groups of 8 + 4 (C4, decision D2), frames shorter than one
raster row (`convert` of a wide raster, C1-style), batched tile-id / bbox queries (C5).

Every case runs: CUDA encode -> oracle decoder (CRCs, frame numbers, frame sizes) -> PCM == the reference quantiser
-> size <= 1.02 x the oracle's libFLAC level-5 restatement -> CUDA decode -> raster == the reference round trip.
"""
import json
import struct

import numpy as np
import pytest

import oracle
from oracle import quant
from tests.helpers import check_encode_against_oracle, lossless_pcm, spatial_pcm, tiles_for

pytestmark = pytest.mark.gpu


def _np(t):
    from flac_raster_b200.synth import to_numpy
    return to_numpy(t)


def _encode_dev(codec, dev_raster, tiles, **kw):
    import torch
    res = codec.encode_tiles(dev_raster, tiles, **kw)
    torch.cuda.synchronize()
    assert int(res.status.item()) == 0
    total = int(res.total_bytes.item())
    return res, (res.bytes[:total].cpu().numpy(), res.tile_offset.cpu().numpy(), res.tile_size.cpu().numpy(),
                 res.tile_minmax.cpu().numpy(), res.frame_size.cpu().numpy())


def _decode_dev(codec, res, tiles, dtype, pcm_mode="reference"):
    """Every tile decoded by one batched call into a packed output; returns [(nch,h,w) arrays]."""
    import torch
    from flac_raster_b200.codec import _tdtype, make_dtiles
    toff, tsize, mm = res.tile_offset.cpu().numpy(), res.tile_size.cpu().numpy(), res.tile_minmax.cpu().numpy()
    entries, ff, oe = [], 0, 0
    for i, (r0, c0, h, w, b0, nch) in enumerate(tiles):
        nfr = (h * w + 4095) // 4096
        entries.append(dict(byte_offset=toff[i], byte_size=tsize[i], out_offset=oe, out_band_stride=h * w, out_row_stride=w,
                            h=h, w=w, channels=nch, bps=res.bps, dmin=mm[i, 0], dmax=mm[i, 1], first_frame=ff, n_frames=nfr))
        ff += nfr
        oe += nch * h * w
    out = torch.zeros(oe, dtype=_tdtype(dtype), device="cuda")
    status = codec.decode_tiles(res.bytes, make_dtiles(entries), out, out_dtype=dtype, frame_size=res.frame_size,
                                pcm_mode=pcm_mode, n_bytes=int(res.total_bytes.item()))
    torch.cuda.synchronize()
    assert not status.cpu().numpy().any(), status.cpu().numpy()
    flat = _np(out)
    return [flat[e["out_offset"]:e["out_offset"] + e["channels"] * e["h"] * e["w"]].reshape(e["channels"], e["h"], e["w"])
            for e in entries]


def _reference_roundtrip(block):
    pcm, bps, dmin, dmax, _ = quant.reference_pcm(block)
    nch, h, w = block.shape
    return quant.reference_decode_raster(pcm, bps, h, w, nch, block.dtype, dmin, dmax).reshape(nch, h, w)


def test_c2_geometry_f32_tile1024(codec):
    """C2 crop: float32 DEM 2592^2 at tile 1024 -> tiles of 1024^2, 1024x544, 544x1024, 544^2 (bps 32)."""
    from flac_raster_b200 import synth
    dev = synth.dem_f32(2592, 2592, seed=20000)
    r = _np(dev)
    tiles = tiles_for(r, 1024, "all")
    assert sorted({(t[2], t[3]) for t in tiles}) == [(544, 544), (544, 1024), (1024, 544), (1024, 1024)]
    res, outs = _encode_dev(codec, dev, tiles)
    ours, ref = check_encode_against_oracle(r, tiles, *outs)
    assert ours <= 1.02 * ref
    for (r0, c0, h, w, b0, nch), got in zip(tiles, _decode_dev(codec, res, tiles, "float32")):
        assert np.array_equal(got, _reference_roundtrip(r[b0:b0 + nch, r0:r0 + h, c0:c0 + w]))


@pytest.mark.parametrize("bands", ["first", "all"])
def test_c3_geometry_u16_tile1024(codec, bands):
    """C3 crop: uint16 x 3, 2112^2 at tile 1024 -> 1024^2, 1024x64, 64x1024 and 64^2 tiles; band 1 only (the
    reference, cli.py:699) and all bands as channels (decision D3)."""
    from flac_raster_b200 import synth
    dev = synth.rgb_u16(2112, 2112, seed=40000)
    r = _np(dev)
    tiles = tiles_for(r, 1024, bands)
    assert sorted({(t[2], t[3]) for t in tiles}) == [(64, 64), (64, 1024), (1024, 64), (1024, 1024)]
    res, outs = _encode_dev(codec, dev, tiles)
    check_encode_against_oracle(r, tiles, *outs)
    for (r0, c0, h, w, b0, nch), got in zip(tiles, _decode_dev(codec, res, tiles, "uint16")):
        assert np.array_equal(got, _reference_roundtrip(r[b0:b0 + nch, r0:r0 + h, c0:c0 + w]))


def test_c4_geometry_i16_12band_tile256_lossless(codec):
    """C4 crop: int16 x 12, 528^2 at tile 256 -> 256^2, 256x16, 16x256 and 16^2 tiles, each in band groups 8 + 4
    (decision D2), pcm_mode=lossless (decision D1): bit-exact back."""
    from flac_raster_b200 import synth
    from flac_raster_b200.spatial_encoder import band_groups, calculate_tiles
    dev = synth.multispectral_i16(528, 528, 12, seed=10000)
    r = _np(dev)
    groups = band_groups(12)
    assert groups == [(0, 8), (8, 4)]
    units = [(rr, cc, h, w, b0, nb) for (rr, cc, h, w) in calculate_tiles(528, 528, 256) for (b0, nb) in groups]
    assert sorted({(t[2], t[3]) for t in units}) == [(16, 16), (16, 256), (256, 16), (256, 256)]
    res, outs = _encode_dev(codec, dev, units, pcm_mode="lossless")
    check_encode_against_oracle(r, units, *outs, pcm_fn=lossless_pcm)
    for (r0, c0, h, w, b0, nch), got in zip(units, _decode_dev(codec, res, units, "int16", "lossless")):
        assert np.array_equal(got, r[b0:b0 + nch, r0:r0 + h, c0:c0 + w])


def test_c4_geometry_reference_spatial_8band(codec):
    """The reference's own `--spatial` mapping (float32 in [-1,1] truncated by pyFLAC to -1/0/1, bps 32) on the
    8 bands libFLAC accepts, tile 256 with 16-wide edges."""
    from flac_raster_b200 import synth
    dev = synth.multispectral_i16(272, 528, 8, seed=10001)
    r = _np(dev)
    units = tiles_for(r, 256, "all")
    res, outs = _encode_dev(codec, dev, units, pcm_mode="reference_spatial")
    assert res.bps == 32
    check_encode_against_oracle(r, units, *outs, pcm_fn=spatial_pcm)


def test_c5_queries_against_c3_geometry_file(codec, tmp_path):
    """C5 shape on the C3 crop: a streaming file at tile 1024, 200 seeded queries (half tile ids, half 10x10-pixel
    bboxes) answered by batched decodes; every answer == the reference loop's tile and round trip."""
    from flac_raster_b200 import rio, streaming, synth
    from flac_raster_b200.spatial_encoder import calculate_tiles
    H = W = 2112
    r = _np(synth.rgb_u16(H, W, seed=40000))
    tiles = calculate_tiles(H, W, 1024)
    tr = rio.from_origin(-120.0, 37.0, 1e-4, 1e-4)
    blobs, sizes = streaming.encode_streaming_tiles(r, tiles, "EPSG:4326", tr, bands="first")
    chunks, offsets, sizes = streaming.merge_shards([(blobs, sizes)])
    index = streaming.build_index("EPSG:4326", tr, W, H, 1024, tiles, offsets, sizes)
    path = tmp_path / "c3.flac"
    streaming.write_streaming_file(path, index, chunks)
    raw = path.read_bytes()
    n = struct.unpack(">I", raw[:4])[0]
    assert json.loads(raw[4:4 + n]) == json.loads(json.dumps(index))
    want = {}
    for i, (r0, c0, h, w) in enumerate(tiles):
        want[i] = _reference_roundtrip(r[0:1, r0:r0 + h, c0:c0 + w])
    rng = np.random.default_rng(5)
    queries = []
    for k in range(200):
        if k % 2 == 0:
            queries.append({"tile_id": int(rng.integers(0, len(tiles)))})
        else:
            px, py = int(rng.integers(0, W - 10)), int(rng.integers(0, H - 10))
            x0, y1 = -120.0 + px * 1e-4, 37.0 - py * 1e-4
            queries.append({"bbox": [x0, y1 - 10e-4, x0 + 10e-4, y1]})
    sf = streaming.StreamingFile(path)
    for s in range(0, len(queries), 50):
        pos, arrays = sf.query(queries[s:s + 50])
        for q, p, a in zip(queries[s:s + 50], pos, arrays):
            ref = streaming.select_frame(index, q.get("tile_id"), False, False, q.get("bbox"))
            assert ref is not None and index["frames"][p] == ref
            assert np.array_equal(a, want[ref["frame_id"]])


def test_c1_wide_convert_frame_shorter_than_row(codec, tmp_path):
    """`convert` of a 1-band int16 raster 8200 px wide: the whole raster is one stream, so a 4096-sample frame is
    shorter than a raster row and frames start mid-row."""
    from host_scenarios import read_tif, write_tif
    from flac_raster_b200 import flacmeta
    from flac_raster_b200.converter import RasterFLACConverter
    rng = np.random.default_rng(8)
    xx = np.arange(8200)[None, :]
    yy = np.arange(5)[:, None]
    arr = (900 + 400 * np.sin(xx / 37.0) * np.cos(yy / 3.0) + rng.normal(0, 5, (5, 8200))).astype(np.int16)[None]
    write_tif(tmp_path / "wide.tif", arr)
    conv = RasterFLACConverter()
    conv.tiff_to_flac(tmp_path / "wide.tif", tmp_path / "wide.flac")
    blob = (tmp_path / "wide.flac").read_bytes()
    pcm, sid, frames, used, first, rc = oracle.decode_stream(blob)
    pcm_ref, bps, dmin, dmax, _ = quant.reference_pcm(arr)
    assert rc == 0 and used == len(blob) and all(f["crc8_ok"] and f["crc16_ok"] for f in frames)
    assert len(frames) == (5 * 8200 + 4095) // 4096 and np.array_equal(pcm, pcm_ref)
    enc, _ = oracle.encode_stream(pcm_ref, bps, 44100, 4096, 5, write_header=False)
    assert len(blob) - first <= 1.02 * len(enc) + 16
    hdr = flacmeta.parse_header(blob)
    assert hdr.streaminfo.total_samples == 5 * 8200
    conv.flac_to_tiff(tmp_path / "wide.flac", tmp_path / "wide_back.tif")
    got, _ = read_tif(tmp_path / "wide_back.tif")
    assert np.array_equal(got, _reference_roundtrip(arr))


@pytest.mark.parametrize("kind,nch,tile", [("f32", 1, 1024), ("u16", 3, 1024), ("u16", 2, 512), ("u16", 1, 200)])
def test_encode_is_deterministic_across_calls(codec, kind, nch, tile):
    """The frame warps take tickets, park frames and (narrow kernels) meet at barriers in an order that differs from run to run;
    the bytes must not.  Twenty encodes of the same raster, compared byte for byte (sizes, offsets, stream)."""
    import torch
    from flac_raster_b200 import synth
    from flac_raster_b200.codec import tile_grid
    H = 3000
    r = synth.dem_f32(H, H, seed=5) if kind == "f32" else synth.rgb_u16(H, H, seed=6)
    tiles = tile_grid(H, H, tile, 0, nch)
    res = codec.encode_tiles(r, tiles)
    torch.cuda.synchronize()
    assert int(res.status.item()) == 0
    total = int(res.total_bytes.item())
    ref_bytes, ref_fs, ref_fo = res.bytes[:total].clone(), res.frame_size.clone(), res.frame_offset.clone()
    for _ in range(20):
        res.bytes.zero_()
        codec.encode_tiles(r, tiles, out=res)
        torch.cuda.synchronize()
        assert int(res.total_bytes.item()) == total
        assert torch.equal(res.frame_size, ref_fs) and torch.equal(res.frame_offset, ref_fo)
        assert torch.equal(res.bytes[:total], ref_bytes)

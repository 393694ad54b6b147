"""GPU parity tests: the CUDA path (through the C-ABI) against the CPU oracle."""
import numpy as np
import pytest

import oracle
from oracle import quant
from tests.helpers import check_encode_against_oracle, tiles_for

pytestmark = pytest.mark.gpu


def _torch_raster(a):
    import torch
    tmap = {"uint16": torch.uint16, "uint32": torch.uint32}
    if str(a.dtype) in tmap:
        signed = a.view({"uint16": np.int16, "uint32": np.int32}[str(a.dtype)])
        return torch.from_numpy(signed.copy()).cuda().view(tmap[str(a.dtype)])
    return torch.from_numpy(a.copy()).cuda()


def _encode(codec, raster_np, tiles, **kw):
    import torch
    t = _torch_raster(np.ascontiguousarray(raster_np))
    res = codec.encode_tiles(t, tiles, **kw)
    torch.cuda.synchronize()
    assert int(res.status.item()) == 0
    total = int(res.total_bytes.item())
    return (res, res.bytes[:total].cpu().numpy(), res.tile_offset.cpu().numpy(), res.tile_size.cpu().numpy(),
            res.tile_minmax.cpu().numpy(), res.frame_size.cpu().numpy())


@pytest.mark.parametrize("name", ["rgb", "dem", "ms", "u8_small", "u8_narrow", "u16_12bit", "u16_full", "i16_dem",
                                  "i16_wrap", "i16_neg", "i32_mid", "i32_wide", "u32_mid", "const_u16", "f32_norm",
                                  "f32_dem", "f32_overflow"])
def test_quantize_matches_reference(codec, golden, name):
    """P1: frx_quantize == RasterFLACConverter._normalize_to_audio + pyFLAC astype(int32) (golden from the reference)."""
    import torch
    r = golden[f"in_{name}"]
    bands, h, w = r.shape
    pcm, mm = codec.quantize(_torch_raster(r), (0, 0, h, w, 0, bands))
    torch.cuda.synchronize()
    want = golden[f"audio_{name}"].astype(np.int32).reshape(-1, bands)
    assert np.array_equal(pcm.cpu().numpy(), want)
    mmw = golden[f"minmax_{name}"]
    got = mm.cpu().numpy()
    assert np.array_equal(got, mmw, equal_nan=True)


@pytest.mark.parametrize("name", ["rgb", "dem", "ms", "u8_small", "u16_12bit", "u16_full", "i16_dem", "i16_wrap",
                                  "i32_mid", "u32_mid", "const_u16", "f32_norm", "f32_dem"])
def test_dequantize_matches_reference(codec, golden, name):
    """P3: frx_dequantize == FileDecoder (/32768, top 16 bits) + _denormalize_from_audio (golden from the reference)."""
    import torch
    r = golden[f"in_{name}"]
    bands, h, w = r.shape
    audio = golden[f"audio_{name}"]
    bps = audio.dtype.itemsize * 8
    pcm = torch.from_numpy(audio.astype(np.int32).reshape(-1, bands).copy()).cuda()
    mm = golden[f"minmax_{name}"]
    out = codec.dequantize(pcm, h, w, bands, bps, str(r.dtype), float(mm[0]), float(mm[1]))
    torch.cuda.synchronize()
    got = out.cpu().view({"uint16": torch.int16, "uint32": torch.int32}.get(str(r.dtype), out.dtype)).numpy().view(r.dtype)
    assert np.array_equal(got, golden[f"denorm_{name}"], equal_nan=True)


@pytest.mark.parametrize("name,tile", [("rgb", 256), ("rgb", 100), ("dem", 512), ("dem", 256), ("ms", 200), ("ms", 64),
                                       ("u16_12bit", 64), ("i16_dem", 96), ("i32_mid", 50), ("f32_dem", 48),
                                       ("f32_norm", 64), ("const_u16", 16), ("u8_narrow", 40), ("i16_wrap", 8),
                                       ("u32_mid", 20), ("f32_overflow", 12)])
def test_encode_streams_decode_to_reference_pcm(codec, golden, name, tile):
    """P1+P2 and the <= 2 % size gate on the fixtures and seeded cases, all bands as channels."""
    r = golden[f"in_{name}"]
    tiles = tiles_for(r, tile, "all")
    _, data, toff, tsize, tmm, fsize = _encode(codec, r, tiles)
    check_encode_against_oracle(r, tiles, data, toff, tsize, tmm, fsize)


def test_rgb_fixture_size_gate(codec, golden):
    """frames for the G1 PCM must be <= 1.02 x 178 857 B (libFLAC 1.4.3 level 5)."""
    r = golden["in_rgb"]
    tiles = [(0, 0, 256, 256, 0, 3)]
    _, data, toff, tsize, tmm, fsize = _encode(codec, r, tiles)
    assert int(tsize[0]) <= 182434, int(tsize[0])


@pytest.mark.parametrize("level", [0, 3, 8])
def test_levels(codec, golden, level):
    r = golden["in_dem"]
    tiles = tiles_for(r, 256, "all")
    _, data, toff, tsize, tmm, fsize = _encode(codec, r, tiles, level=level)
    check_encode_against_oracle(r, tiles, data, toff, tsize, tmm, fsize, level=level, size_margin=1.03)


def _roundtrip_decode(codec, r, tiles, res, pcm_mode="reference"):
    import torch
    from flac_raster_b200.codec import make_dtiles
    bands, H, W = r.shape
    entries = []
    out_elems = 0
    toff = res.tile_offset.cpu().numpy(); tsize = res.tile_size.cpu().numpy(); mm = res.tile_minmax.cpu().numpy()
    ff = 0
    for i, (r0, c0, h, w, b0, nch) in enumerate(tiles):
        nfr = (h * w + 4095) // 4096
        entries.append(dict(byte_offset=toff[i], byte_size=tsize[i], out_offset=out_elems, out_band_stride=h * w,
                            out_row_stride=w, h=h, w=w, channels=nch, bps=res.bps, dmin=mm[i, 0], dmax=mm[i, 1],
                            first_frame=ff, n_frames=nfr))
        ff += nfr
        out_elems += nch * h * w
    from flac_raster_b200.codec import _tdtype
    out = torch.zeros(out_elems, dtype=_tdtype(str(r.dtype)), device="cuda")
    status = codec.decode_tiles(res.bytes, make_dtiles(entries), out, out_dtype=str(r.dtype), frame_size=res.frame_size,
                                pcm_mode=pcm_mode, n_bytes=int(res.total_bytes.item()))
    torch.cuda.synchronize()
    assert not status.cpu().numpy().any(), status.cpu().numpy()
    flat = out.cpu().view({"uint16": torch.int16, "uint32": torch.int32}.get(str(r.dtype), out.dtype)).numpy().view(r.dtype)
    return flat, entries


@pytest.mark.parametrize("name,tile", [("rgb", 256), ("rgb", 100), ("dem", 256), ("ms", 64), ("u16_12bit", 64),
                                       ("i32_mid", 50), ("f32_dem", 48), ("u8_narrow", 40), ("const_u16", 16)])
def test_gpu_decode_matches_reference_roundtrip(codec, golden, name, tile):
    """P3 end to end: GPU decode of our own streams == reference flac_to_tiff output on the same PCM."""
    r = golden[f"in_{name}"]
    tiles = tiles_for(r, tile, "all")
    res, *_ = _encode(codec, r, tiles)
    flat, entries = _roundtrip_decode(codec, r, tiles, res)
    for (r0, c0, h, w, b0, nch), e in zip(tiles, entries):
        block = r[b0:b0 + nch, r0:r0 + h, c0:c0 + w]
        pcm_ref, bps, dmin, dmax, _ = quant.reference_pcm(block)
        want = quant.reference_decode_raster(pcm_ref, bps, h, w, nch, r.dtype, dmin, dmax)
        got = flat[e["out_offset"]: e["out_offset"] + nch * h * w].reshape(nch, h, w)
        assert np.array_equal(got, want.reshape(nch, h, w), equal_nan=True)


@pytest.mark.parametrize("name", ["u16_full", "i16_dem", "u8_small", "i32_wide", "f32_overflow", "u32_mid"])
def test_lossless_mode_roundtrip(codec, golden, name):
    r = golden[f"in_{name}"]
    tiles = tiles_for(r, 64, "all")
    res, *_ = _encode(codec, r, tiles, pcm_mode="lossless")
    flat, entries = _roundtrip_decode(codec, r, tiles, res, pcm_mode="lossless")
    for (r0, c0, h, w, b0, nch), e in zip(tiles, entries):
        block = r[b0:b0 + nch, r0:r0 + h, c0:c0 + w]
        got = flat[e["out_offset"]: e["out_offset"] + nch * h * w].reshape(nch, h, w)
        assert np.array_equal(got.view(np.uint8 if r.dtype.itemsize == 1 else got.dtype).tobytes(), np.ascontiguousarray(block).tobytes())


def test_decode_reference_fixture_rgb(codec, golden):
    """Our decoder on a libFLAC-1.4.3-produced file (no frame table -> sync discovery): G1 -> G3."""
    import torch
    from pathlib import Path
    from flac_raster_b200 import flacmeta
    from flac_raster_b200.codec import make_dtiles
    data = (Path(__file__).parent / "golden" / "sample_rgb.flac").read_bytes()
    hdr = flacmeta.parse_header(data)
    si = hdr.streaminfo
    n = len(data) - hdr.first_frame_offset
    e = dict(byte_offset=hdr.first_frame_offset, byte_size=n, out_offset=0, out_band_stride=256 * 256, out_row_stride=256,
             h=256, w=256, channels=si.channels, bps=si.bps, dmin=1.0, dmax=255.0, first_frame=None, n_frames=16)
    buf = torch.from_numpy(np.frombuffer(data, dtype=np.uint8).copy()).cuda()
    out = torch.zeros(3 * 256 * 256, dtype=torch.uint8, device="cuda")
    status = codec.decode_tiles(buf, make_dtiles([e]), out, out_dtype="uint8")
    torch.cuda.synchronize()
    assert int(status[0].item()) == 0
    assert np.array_equal(out.cpu().numpy().reshape(3, 256, 256), golden["fixture_rgb"])


def test_decode_reference_fixture_dem_spatial(codec):
    """G4: four concatenated bps-32 streams, all-zero PCM."""
    import torch
    from pathlib import Path
    from flac_raster_b200 import flacmeta
    from flac_raster_b200.codec import make_dtiles
    data = (Path(__file__).parent / "golden" / "sample_dem.flac").read_bytes()
    starts = [0, 10426, 18880, 27334]
    entries = []
    for i, s in enumerate(starts):
        hdr = flacmeta.parse_header(data, s)
        end = starts[i + 1] if i + 1 < 4 else len(data)
        entries.append(dict(byte_offset=hdr.first_frame_offset, byte_size=end - hdr.first_frame_offset, out_offset=i * 65536,
                            out_band_stride=65536, out_row_stride=256, h=256, w=256, channels=1, bps=32, dmin=0.0,
                            dmax=0.0, first_frame=None, n_frames=16))
    buf = torch.from_numpy(np.frombuffer(data, dtype=np.uint8).copy()).cuda()
    out = torch.full((4 * 65536,), 7, dtype=torch.int32, device="cuda")
    status = codec.decode_tiles(buf, make_dtiles(entries), out, out_dtype="int32", pcm_mode="lossless")
    torch.cuda.synchronize()
    assert not status.cpu().numpy().any()
    assert int(out.abs().max().item()) == 0


def test_decode_stereo_decorrelation(codec):
    """Two-channel streams as libFLAC writes them at level 5: left/side, side/right and mid/side frames (RFC 9639 4.2).
    The oracle's encoder restates libFLAC's mid-side search; its decoder is the reference for the PCM.  Both the
    frame-table path and sync discovery are covered, with tile widths that make chunks straddle raster rows."""
    import torch
    import oracle
    from flac_raster_b200 import flacmeta
    from flac_raster_b200.codec import make_dtiles
    rng = np.random.default_rng(11)
    h, w = 96, 200
    n = h * w
    base = np.cumsum(rng.integers(-40, 41, n)).astype(np.int64)
    cases = {
        "correlated": np.stack([base + rng.integers(-3, 4, n), base + rng.integers(-3, 4, n)], axis=1),
        "left_quiet": np.stack([rng.integers(-5, 6, n), base], axis=1),
        "right_quiet": np.stack([base, rng.integers(-5, 6, n)], axis=1),
    }
    seen = set()
    for name, pcm in cases.items():
        pcm = np.clip(pcm, -30000, 30000).astype(np.int16)
        data, fsizes = oracle.encode_stream(pcm, 16, level=5)
        ref, _si, frames = oracle.decode_stream(data)[:3]
        seen.update(int(f["chan_assign"]) for f in frames)
        assert np.array_equal(ref.reshape(-1, 2), pcm)
        hdr = flacmeta.parse_header(data)
        nfr = (n + 4095) // 4096
        buf = torch.from_numpy(np.frombuffer(data, dtype=np.uint8).copy()).cuda()
        for table in (True, False):
            e = dict(byte_offset=hdr.first_frame_offset, byte_size=len(data) - hdr.first_frame_offset, out_offset=0, out_band_stride=n,
                     out_row_stride=w, h=h, w=w, channels=2, bps=16, dmin=0.0, dmax=0.0, first_frame=0 if table else None, n_frames=nfr)
            out = torch.zeros(2 * n, dtype=torch.int16, device="cuda")
            fs = torch.from_numpy(fsizes.astype(np.int32)).cuda() if table else None
            status = codec.decode_tiles(buf, make_dtiles([e]), out, out_dtype="int16", pcm_mode="lossless", frame_size=fs)
            torch.cuda.synchronize()
            assert int(status[0].item()) == 0, (name, table)
            got = out.cpu().numpy().reshape(2, n)
            assert np.array_equal(got[0], pcm[:, 0]) and np.array_equal(got[1], pcm[:, 1]), (name, table)
    assert seen & {8, 9, 10}, f"the oracle encoder never chose a decorrelated assignment: {seen}"


def test_host_entry_points_roundtrip(codec, golden):
    from flac_raster_b200.codec import make_dtiles
    r = golden["in_dem"]
    tiles = tiles_for(r, 256, "all")
    res = codec.encode_tiles_host(r, tiles)
    total = int(res.total_bytes[0])
    check_encode_against_oracle(r, tiles, res.bytes[:total], res.tile_offset, res.tile_size, res.tile_minmax, res.frame_size)
    entries, ff, oe = [], 0, 0
    for i, (r0, c0, h, w, b0, nch) in enumerate(tiles):
        nfr = (h * w + 4095) // 4096
        entries.append(dict(byte_offset=res.tile_offset[i], byte_size=res.tile_size[i], out_offset=oe, out_band_stride=h * w,
                            out_row_stride=w, h=h, w=w, channels=nch, bps=res.bps, dmin=res.tile_minmax[i, 0],
                            dmax=res.tile_minmax[i, 1], first_frame=ff, n_frames=nfr))
        ff += nfr; oe += nch * h * w
    out = np.zeros(oe, dtype=r.dtype)
    status = codec.decode_tiles_host(res.bytes[:total], make_dtiles(entries), out, frame_size=res.frame_size.astype(np.uint32))
    assert not status.any()
    for (r0, c0, h, w, b0, nch), e in zip(tiles, entries):
        block = r[b0:b0 + nch, r0:r0 + h, c0:c0 + w]
        pcm_ref, bps, dmin, dmax, _ = quant.reference_pcm(block)
        want = quant.reference_decode_raster(pcm_ref, bps, h, w, nch, r.dtype, dmin, dmax)
        assert np.array_equal(out[e["out_offset"]: e["out_offset"] + nch * h * w].reshape(nch, h, w), want.reshape(nch, h, w))


def test_host_pipeline_groups_equal_single_call(codec, monkeypatch):
    """frx_encode_tiles_host cuts the tiles into groups that overlap copy-in, encode and copy-out; however the groups fall,
    the stream and every index array must equal the one-call device path."""
    import torch
    from flac_raster_b200 import synth
    r = synth.dem_f32(1100, 900, seed=5).cpu().numpy()
    tiles = tiles_for(r, 256, "all")
    dev = codec.encode_tiles(torch.from_numpy(r).cuda(), tiles)
    total = int(dev.total_bytes.cpu()[0])
    want = dev.bytes[:total].cpu().numpy()
    for group_bytes in ("1", "300000", "1000000000"):
        monkeypatch.setenv("FRX_HOST_GROUP_BYTES", group_bytes)
        res = codec.encode_tiles_host(r, tiles)
        assert int(res.total_bytes[0]) == total
        assert np.array_equal(res.bytes[:total], want)
        assert np.array_equal(res.tile_offset, dev.tile_offset.cpu().numpy())
        assert np.array_equal(res.tile_size, dev.tile_size.cpu().numpy())
        assert np.array_equal(res.frame_offset, dev.frame_offset.cpu().numpy())
        assert np.array_equal(res.frame_size, dev.frame_size.cpu().numpy())
        assert np.array_equal(res.tile_minmax, dev.tile_minmax.cpu().numpy())
    # a capacity that is too small is reported, not overrun
    from flac_raster_b200._lib import FrxError
    small = codec.encode_tiles_host(r, tiles)
    small.bytes = np.empty(total // 2, dtype=np.uint8)
    with pytest.raises(FrxError):
        codec.encode_tiles_host(r, tiles, out=small)


def test_more_than_8_channels_is_rejected(codec):
    import torch
    from flac_raster_b200._lib import FrxUnsupported
    t = torch.zeros((12, 64, 64), dtype=torch.int16, device="cuda")
    with pytest.raises(FrxUnsupported):
        codec.encode_tiles(t, [(0, 0, 64, 64, 0, 12)])


def test_host_decode_rejects_output_blocks_outside_the_buffer(codec, golden):
    """frx_decode_tiles_host knows the size of the output: a descriptor whose block does not fit is refused before any launch."""
    from flac_raster_b200._lib import FrxError
    from flac_raster_b200.codec import make_dtiles
    r = golden["in_dem"]
    tiles = tiles_for(r, 256, "all")
    res = codec.encode_tiles_host(r, tiles)
    total = int(res.total_bytes[0])
    e = dict(byte_offset=res.tile_offset[0], byte_size=res.tile_size[0], out_offset=0, out_band_stride=256 * 256, out_row_stride=256, h=256, w=256,
             channels=1, bps=res.bps, dmin=res.tile_minmax[0, 0], dmax=res.tile_minmax[0, 1], first_frame=0, n_frames=16)
    out = np.zeros(256 * 256, dtype=r.dtype)
    fs = res.frame_size.astype(np.uint32)
    assert not codec.decode_tiles_host(res.bytes[:total], make_dtiles([e]), out, frame_size=fs).any()
    for bad in (dict(e, out_offset=1), dict(e, out_row_stride=257), dict(e, byte_offset=total - 10, byte_size=11)):
        with pytest.raises(FrxError):
            codec.decode_tiles_host(res.bytes[:total], make_dtiles([bad]), out, frame_size=fs)


@pytest.mark.parametrize("name,tile,mode", [("rgb", 256, "reference"), ("rgb", 100, "reference"), ("f32_dem", 48, "reference"),
                                            ("i16_dem", 37, "lossless"), ("u8_narrow", 5, "reference"), ("i32_mid", 50, "lossless")])
def test_tile_md5_matches_hashlib(codec, golden, name, tile, mode):
    """frx_tile_md5 == MD5 of the stream's PCM as little-endian bytes (what libFLAC stores in STREAMINFO); tile sizes with odd
    sample counts put the pad byte in the middle of a word."""
    import hashlib
    from tests.helpers import lossless_pcm
    r = golden[f"in_{name}"]
    tiles = tiles_for(r, tile, "all")
    got = codec.tile_md5(_torch_raster(r), tiles, pcm_mode=mode)
    for (r0, c0, h, w, b0, nch), sig in zip(tiles, got):
        block = r[b0:b0 + nch, r0:r0 + h, c0:c0 + w]
        pcm, bps, *_ = (quant.reference_pcm(block) if mode == "reference" else lossless_pcm(block))
        want = hashlib.md5(pcm.astype("<i2" if bps == 16 else "<i4").tobytes()).digest()
        assert bytes(sig) == want, (name, (r0, c0, h, w))


def test_host_decode_pipeline_groups_equal_single_call(codec, monkeypatch):
    """frx_decode_tiles_host cuts the tiles into groups that overlap copy-in, decode and copy-out; however the groups fall --
    with packed outputs or with tiles written into one raster, where the output ranges of neighbouring groups interleave --
    the rasters must equal the device path's."""
    import torch
    from flac_raster_b200 import synth
    from flac_raster_b200.codec import make_dtiles
    H, W = 700, 900
    dev = synth.dem_f32(H, W, seed=6)
    tiles = tiles_for(synth.to_numpy(dev), 256, "all")
    res, data, toff, tsize, mm, fsize = _encode(codec, synth.to_numpy(dev), tiles)
    packed, mosaic, ff, oe = [], [], 0, 0
    for i, (r0, c0, h, w, b0, nch) in enumerate(tiles):
        nfr = (h * w + 4095) // 4096
        base = dict(byte_offset=toff[i], byte_size=tsize[i], h=h, w=w, channels=1, bps=res.bps, dmin=mm[i, 0], dmax=mm[i, 1], first_frame=ff, n_frames=nfr)
        packed.append(dict(base, out_offset=oe, out_band_stride=h * w, out_row_stride=w))
        mosaic.append(dict(base, out_offset=r0 * W + c0, out_band_stride=H * W, out_row_stride=W))
        ff += nfr; oe += h * w
    want_m = torch.zeros(H * W, dtype=torch.float32, device="cuda")
    st = codec.decode_tiles(res.bytes, make_dtiles(mosaic), want_m, out_dtype="float32", frame_size=res.frame_size, n_bytes=len(data))
    torch.cuda.synchronize()
    assert not st.cpu().numpy().any()
    want_m = want_m.cpu().numpy()
    fs = fsize.astype(np.uint32)
    for group_bytes in ("1", "200000", "1000000000"):
        monkeypatch.setenv("FRX_HOST_GROUP_BYTES", group_bytes)
        out = np.full(H * W, np.nan, dtype=np.float32)
        assert not codec.decode_tiles_host(data, make_dtiles(mosaic), out, frame_size=fs).any()
        assert np.array_equal(out, want_m)
        outp = np.full(oe, np.nan, dtype=np.float32)
        assert not codec.decode_tiles_host(data, make_dtiles(packed), outp, frame_size=fs).any()
        for e, m in zip(packed, mosaic):
            r0, c0 = divmod(m["out_offset"], W)
            assert np.array_equal(outp[e["out_offset"]:e["out_offset"] + e["h"] * e["w"]].reshape(e["h"], e["w"]),
                                  want_m.reshape(H, W)[r0:r0 + e["h"], c0:c0 + e["w"]])


@pytest.mark.parametrize("dtype,lo,rng_", [("uint16", 100, 32767), ("uint16", 0, 65534), ("uint16", 7, 4340), ("uint16", 1000, 453),
                                           ("uint16", 0, 65535), ("uint8", 0, 255), ("uint8", 3, 217), ("int16", -16000, 32767),
                                           ("int16", -5, 31 * 151), ("int16", -20000, 40000)])
def test_integer_quantiser_on_ranges_full_of_exact_quotients(codec, dtype, lo, rng_):
    """The 8/16-bit quantiser's integer form (frx_common.cuh frx_quant16_int) against the reference's float64 chain on
    tiles that hold EVERY value of their range, with ranges sharing factors with 32767 = 7 * 31 * 151 so that many quotients
    are exact integers (the samples that take the float64 branch), and an int16 range that wraps (the chain only).  Both
    kernels that quantise are covered: frx_quantize (per sample) and the encoder's block loader (per 16-byte load)."""
    import torch
    n_val = rng_ + 1
    w = 512
    h = (n_val + w - 1) // w + 3
    vals = (lo + (np.arange(h * w, dtype=np.int64) % n_val))
    rs = np.random.default_rng(rng_)
    vals[n_val:] = lo + rs.integers(0, n_val, h * w - n_val)              # every value once, then a random refill
    r = vals.astype(np.int64).astype(dtype).reshape(1, h, w)
    from oracle import quant
    want, bps, dmin, dmax, _ = quant.reference_pcm(r)
    pcm, mm = codec.quantize(_torch_raster(r), (0, 0, h, w, 0, 1))
    torch.cuda.synchronize()
    assert np.array_equal(pcm.cpu().numpy().reshape(-1), want.reshape(-1))
    tiles = [(0, 0, h, w, 0, 1)]
    _, data, toff, tsize, tmm, fsize = _encode(codec, r, tiles)
    check_encode_against_oracle(r, tiles, data, toff, tsize, tmm, fsize)

"""frx_decode_tiles_noalloc (include/frx.h): the allocation-free variant of the batched decode, whose scratch is a static
device array inside libfrx.  Called through the C-ABI and compared bit for bit (integer/byte work: the tests' tolerance
is equality) with the existing frx_decode_tiles on the same inputs, on calls in a row whose inputs differ, so that
whatever one call leaves in the shared scratch cannot leak into the next.  This file is collected last on purpose."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _case(codec, raster_dev, tile, bands, table=True):
    """Encode with the existing entry point; -> (EncodeResult, dtile entries, output elements)."""
    import torch
    from tests.helpers import tiles_for
    from flac_raster_b200.synth import to_numpy
    r = to_numpy(raster_dev)
    tiles = tiles_for(r, tile, bands)
    res = codec.encode_tiles(raster_dev, tiles)
    torch.cuda.synchronize()
    assert int(res.status.item()) == 0
    toff, tsize, mm = res.tile_offset.cpu().numpy(), res.tile_size.cpu().numpy(), res.tile_minmax.cpu().numpy()
    entries, ff, oe = [], 0, 0
    for i, (r0, c0, h, w, b0, nch) in enumerate(tiles):
        nfr = (h * w + 4095) // 4096
        entries.append(dict(byte_offset=toff[i], byte_size=tsize[i], out_offset=oe, out_band_stride=h * w, out_row_stride=w, h=h, w=w,
                            channels=nch, bps=res.bps, dmin=mm[i, 0], dmax=mm[i, 1], first_frame=ff if table else None, n_frames=nfr))
        ff += nfr
        oe += nch * h * w
    return res, entries, oe


def _both(codec, res, entries, oe, dtype, check_free=False):
    """Decode with frx_decode_tiles and with frx_decode_tiles_noalloc; the outputs must be identical."""
    import torch
    from flac_raster_b200.codec import _tdtype, make_dtiles
    dt = make_dtiles(entries)
    nb = int(res.total_bytes.item())
    want = torch.zeros(oe, dtype=_tdtype(dtype), device="cuda")
    st0 = codec.decode_tiles(res.bytes, dt, want, out_dtype=dtype, frame_size=res.frame_size, n_bytes=nb)
    got = torch.empty_like(want)
    got.view(torch.uint8).fill_(7)
    status = torch.full((len(entries),), -99, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    if check_free:
        # the driver loads a kernel's image on its first launch (lazy module loading), which takes device memory that is not ours:
        # one unmeasured call first, then the measured one
        codec.decode_tiles_noalloc(res.bytes, dt, got, status, out_dtype=dtype, frame_size=res.frame_size, n_bytes=nb)
        torch.cuda.synchronize()
        got.view(torch.uint8).fill_(7)
        status.fill_(-99)
        torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    codec.decode_tiles_noalloc(res.bytes, dt, got, status, out_dtype=dtype, frame_size=res.frame_size, n_bytes=nb)
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    assert not st0.cpu().numpy().any() and not status.cpu().numpy().any(), (st0.cpu().numpy(), status.cpu().numpy())
    assert torch.equal(got.view(torch.uint8), want.view(torch.uint8))
    if check_free:
        assert free1 >= free0, f"device memory shrank by {free0 - free1} B across an allocation-free call"
    return want


def test_noalloc_decode_equals_decode_tiles_on_calls_in_a_row(codec):
    import torch
    from flac_raster_b200 import synth
    from oracle import quant
    a = synth.dem_f32(600, 700, seed=31)                      # 9 float32 tiles (bps 32), frame table
    b = synth.rgb_u16(300, 420, seed=32)                      # 12 uint16 tiles x 3 channels (bps 16), sync discovery
    ca = _case(codec, a, 256, "all", table=True)
    cb = _case(codec, b, 128, "all", table=False)
    cc = _case(codec, b, 200, "first", table=True)            # fewer, larger tiles than the call before
    wa = _both(codec, *ca, "float32")
    _both(codec, *cb, "uint16", check_free=True)              # second call in a row: different dtype, tiles, discovery path
    _both(codec, *cc, "uint16", check_free=True)
    wa2 = _both(codec, *ca, "float32", check_free=True)       # and the first inputs again
    assert torch.equal(wa.view(torch.uint8), wa2.view(torch.uint8))
    # the shared expectation is itself the reference round trip (spot check of the first tile)
    r = synth.to_numpy(a)
    e = ca[1][0]
    pcm, bps, dmin, dmax, _ = quant.reference_pcm(r[:, :e["h"], :e["w"]])
    ref = quant.reference_decode_raster(pcm, bps, e["h"], e["w"], 1, r.dtype, dmin, dmax)
    assert np.array_equal(synth.to_numpy(wa[:e["h"] * e["w"]]).reshape(1, e["h"], e["w"]), ref.reshape(1, e["h"], e["w"]))


def test_noalloc_decode_rejects_what_its_scratch_cannot_hold(codec):
    import torch
    from flac_raster_b200 import _lib as L, synth
    from flac_raster_b200.codec import make_dtiles
    res, entries, oe = _case(codec, synth.rgb_u16(128, 256, seed=33), 64, "first")
    out = torch.zeros(oe, dtype=torch.uint16, device="cuda")
    status = torch.zeros(L.NOALLOC_MAX_TILES + 8, dtype=torch.int32, device="cuda")
    nb = int(res.total_bytes.item())
    too_many = [entries[i % len(entries)] for i in range(L.NOALLOC_MAX_TILES + 1)]
    with pytest.raises(L.FrxUnsupported, match="FRX_NOALLOC_MAX_TILES"):
        codec.decode_tiles_noalloc(res.bytes, make_dtiles(too_many), out, status, frame_size=res.frame_size, n_bytes=nb)
    big = [dict(e, h=4096, w=4096, n_frames=4096, out_band_stride=4096 * 4096, out_row_stride=4096) for e in entries[:8]] + \
          [dict(entries[0], h=64, w=64, n_frames=1)]
    assert sum(e["n_frames"] for e in big) > L.NOALLOC_MAX_FRAMES
    with pytest.raises(L.FrxUnsupported, match="FRX_NOALLOC_MAX_FRAMES"):
        codec.decode_tiles_noalloc(res.bytes, make_dtiles(big), out, status, frame_size=res.frame_size, n_bytes=nb)
    with pytest.raises(L.FrxUnsupported, match="two channels"):
        codec.decode_tiles_noalloc(res.bytes, make_dtiles([dict(entries[0], channels=2)]), out, status, frame_size=res.frame_size, n_bytes=nb)
    # the limits themselves are accepted: exactly FRX_NOALLOC_MAX_TILES tiles in one call
    full = [dict(entries[i % len(entries)], out_offset=i * 4096) for i in range(L.NOALLOC_MAX_TILES)]
    out = torch.zeros(L.NOALLOC_MAX_TILES * 4096, dtype=torch.uint16, device="cuda")
    codec.decode_tiles_noalloc(res.bytes, make_dtiles(full), out, status, frame_size=res.frame_size, n_bytes=nb)
    torch.cuda.synchronize()
    assert not status[:L.NOALLOC_MAX_TILES].cpu().numpy().any()

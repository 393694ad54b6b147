"""GPU decoder on streams our own encoder never writes but a reference-produced (or any legal) FLAC file may hold
("what's weak" 4): libFLAC levels 0/3/8, escaped Rice partitions, LPC orders 13..32, each with the frame
table and through sync-code discovery.  Streams come from the oracle's encoder (test switches in oracle/__init__.py),
the expected PCM from the oracle's decoder."""
import numpy as np
import pytest

import oracle

pytestmark = pytest.mark.gpu


def _gpu_decode(codec, data, fsizes, h, w, channels, bps, table):
    import torch
    from flac_raster_b200 import flacmeta
    from flac_raster_b200.codec import make_dtiles
    hdr = flacmeta.parse_header(data)
    n = h * w
    nfr = (n + 4095) // 4096
    buf = torch.from_numpy(np.frombuffer(data, dtype=np.uint8).copy()).cuda()
    e = dict(byte_offset=hdr.first_frame_offset, byte_size=len(data) - hdr.first_frame_offset, out_offset=0, out_band_stride=n,
             out_row_stride=w, h=h, w=w, channels=channels, bps=bps, dmin=0.0, dmax=0.0, first_frame=0 if table else None,
             n_frames=nfr)
    dt = torch.int16 if bps == 16 else torch.int32
    out = torch.zeros(channels * n, dtype=dt, device="cuda")
    fs = torch.from_numpy(fsizes.astype(np.int32)).cuda() if table else None
    status = codec.decode_tiles(buf, make_dtiles([e]), out, out_dtype="int16" if bps == 16 else "int32", pcm_mode="lossless",
                                frame_size=fs)
    torch.cuda.synchronize()
    assert int(status[0].item()) == 0
    return out.cpu().numpy().reshape(channels, n)


def _signal(n, seed, tones=14, noise=2.0):
    rng = np.random.default_rng(seed)
    t = np.arange(n)
    sig = sum(3000 * np.sin(t * (0.05 + 0.09 * k) + k) for k in range(tones)) / 4 + rng.normal(0, noise, n)
    return sig


@pytest.fixture(autouse=True)
def _reset_oracle_switches():
    yield
    oracle.set_test_options(0, 0)


@pytest.mark.parametrize("level", [0, 3, 8])
@pytest.mark.parametrize("channels", [1, 2, 3])
def test_decode_oracle_levels(codec, level, channels):
    """Streams as libFLAC's presets 0, 3 and 8 write them (fixed-only, LPC <= 6 with partition order <= 4, LPC <= 12 with
    partition order <= 6; 2-channel streams with the presets' stereo decisions)."""
    h, w = 70, 300
    n = h * w
    rng = np.random.default_rng(level * 10 + channels)
    base = _signal(n, level, tones=5, noise=6.0)
    pcm = np.stack([base * (0.6 + 0.2 * c) + rng.normal(0, 3 + c, n) for c in range(channels)], axis=1)
    pcm = np.clip(pcm, -32000, 32000).astype(np.int16)
    data, fsizes = oracle.encode_stream(pcm, 16, level=level)
    ref, _si, frames = oracle.decode_stream(data)[:3]
    assert np.array_equal(ref, pcm)
    if level == 8:
        assert max(sf["order"] for f in frames for sf in f["subframes"] if sf["type"] == "LPC") > 8
    for table in (True, False):
        got = _gpu_decode(codec, data, fsizes, h, w, channels, 16, table)
        for c in range(channels):
            assert np.array_equal(got[c], pcm[:, c]), (level, channels, table, c)


@pytest.mark.parametrize("bps", [16, 32])
def test_decode_high_lpc_orders(codec, bps):
    """LPC orders 13..32 (legal outside the streamable subset; the out-of-line path of the frame decoder)."""
    h, w = 50, 400
    n = h * w
    sig = _signal(n, 3)
    if bps == 32:
        pcm = (sig * 4000.0).astype(np.int64).clip(-(1 << 30), (1 << 30) - 1).astype(np.int32)
    else:
        pcm = np.clip(sig, -32000, 32000).astype(np.int16)
    seen = set()
    for max_order in (16, 20, 32):
        oracle.set_test_options(max_order, 0)
        data, fsizes = oracle.encode_stream(pcm, bps, level=8)
        ref, _si, frames = oracle.decode_stream(data)[:3]
        assert np.array_equal(ref[:, 0], pcm)
        seen.update(sf["order"] for f in frames for sf in f["subframes"] if sf["type"] == "LPC")
        for table in (True, False):
            got = _gpu_decode(codec, data, fsizes, h, w, 1, bps, table)
            assert np.array_equal(got[0], pcm), (bps, max_order, table)
    assert any(o > 12 for o in seen) and any(o > 16 for o in seen), seen


@pytest.mark.parametrize("bps", [16, 32])
def test_decode_escaped_partitions(codec, bps):
    """Escaped (raw binary) Rice partitions mixed with ordinary ones, including all-zero partitions of width 0."""
    h, w = 40, 512
    n = h * w
    rng = np.random.default_rng(9)
    scale = np.repeat(rng.integers(1, 400, n // 256), 256).astype(np.float64)    # variance changes -> high partition orders
    sig = np.cumsum(rng.normal(0, 1, n) * scale) * 0.05 + rng.normal(0, 1, n) * scale
    sig[4096 * 2:4096 * 2 + 2048] = 17                                            # flat stretch -> zero residual partitions
    if bps == 32:
        pcm = (sig * 3000.0).astype(np.int64).clip(-(1 << 30), (1 << 30) - 1).astype(np.int32)
    else:
        pcm = np.clip(sig, -32000, 32000).astype(np.int16)
    total_esc = 0
    for mod in (1, 2, 3):
        oracle.set_test_options(0, mod)
        data, fsizes = oracle.encode_stream(pcm, bps, level=5)
        ref, _si, frames = oracle.decode_stream(data)[:3]
        assert np.array_equal(ref[:, 0], pcm)
        total_esc += sum(sf["escapes"] for f in frames for sf in f["subframes"])
        for table in (True, False):
            got = _gpu_decode(codec, data, fsizes, h, w, 1, bps, table)
            assert np.array_equal(got[0], pcm), (bps, mod, table)
    assert total_esc > 20, total_esc


@pytest.mark.parametrize("table", [True, False])
def test_decode_stereo_at_32_bits(codec, table):
    """Two-channel streams at 32 bits per sample as libFLAC 1.4 writes them at level 5: left/side, side/right and mid/side
    frames whose side channel is a 33-bit signal (the oracle's encoder restates that search).  Includes full-scale samples
    of opposite sign (side = 2^32 - 1), a constant and a verbatim side channel."""
    h, w = 30, 400
    n = h * w
    rng = np.random.default_rng(12)
    t = np.arange(n)
    base = (np.sin(t / 50.0) * 1.5e9).astype(np.int64)
    L = np.clip(base + rng.integers(-1000, 1000, n), -2 ** 31, 2 ** 31 - 1)
    R = np.clip(base // 3 * 2 + rng.integers(-1000, 1000, n), -2 ** 31, 2 ** 31 - 1)
    # frame 1: extremes of opposite sign; frame 2 (tail): noise that no predictor helps
    L[4096:8192] = np.where(t[4096:8192] % 2 == 0, 2 ** 31 - 1, -2 ** 31)
    R[4096:8192] = -L[4096:8192] - 1
    L[8192:] = rng.integers(-2 ** 31, 2 ** 31 - 1, n - 8192)
    R[8192:] = rng.integers(-2 ** 31, 2 ** 31 - 1, n - 8192)
    pcm = np.stack([L, R], axis=1).astype(np.int32)
    data, fsizes = oracle.encode_stream(pcm, 32, level=5)
    ref, _si, frames = oracle.decode_stream(data)[:3]
    assert np.array_equal(ref, pcm)
    assert any(f["chan_assign"] >= 8 for f in frames), [f["chan_assign"] for f in frames]
    got = _gpu_decode(codec, data, fsizes, h, w, 2, 32, table)
    assert np.array_equal(got[0], pcm[:, 0]) and np.array_equal(got[1], pcm[:, 1])


@pytest.mark.parametrize("warps_per_sm", [20.5, 21.5])
def test_decode_copies_for_calls_that_just_miss_one_wave(codec, warps_per_sm):
    """frx_launch_decode picks the 22-CTA copy of the frame decoder when a call needs between 20 and 22 warps per SM (it then fits
    one wave); other calls take the 20-CTA copy.  Calls inside that window must produce exactly what the same tiles give when
    they are decoded a few at a time."""
    import torch
    from flac_raster_b200 import synth
    from flac_raster_b200.codec import make_dtiles, tile_grid
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    n_tiles = int(warps_per_sm * sms * 32) // 256               # 256 frames per 1024 x 1024 tile
    cols = 20
    rows = (n_tiles + cols - 1) // cols
    n_tiles = rows * cols
    warps = n_tiles * 256 / 32 / sms
    if not (int(warps_per_sm) < warps <= int(warps_per_sm) + 1):
        pytest.skip(f"no 20-column tile grid lands in the window on {sms} SMs ({warps:.2f} warps per SM)")
    H, W = rows * 1024, cols * 1024
    r = synth.rgb_u16(H, W, seed=77)[:1].contiguous()
    tiles = tile_grid(H, W, 1024, 0, 1)
    res = codec.encode_tiles(r, tiles)
    torch.cuda.synchronize()
    assert int(res.status.item()) == 0
    total = int(res.total_bytes.item())
    toff, tsize, mm = res.tile_offset.cpu().numpy(), res.tile_size.cpu().numpy(), res.tile_minmax.cpu().numpy()
    entries = []
    for i, (r0, c0, h, w, b0, nch) in enumerate(tiles):
        entries.append(dict(byte_offset=toff[i], byte_size=tsize[i], out_offset=r0 * W + c0, out_band_stride=H * W, out_row_stride=W, h=h, w=w,
                            channels=1, bps=16, dmin=mm[i, 0], dmax=mm[i, 1], first_frame=i * 256, n_frames=256))
    big = torch.zeros((1, H, W), dtype=torch.uint16, device="cuda")
    st = codec.decode_tiles(res.bytes, make_dtiles(entries), big, out_dtype="uint16", frame_size=res.frame_size, n_bytes=total)
    torch.cuda.synchronize()
    assert not st.cpu().numpy().any()
    small = torch.zeros_like(big)
    for i0 in range(0, len(entries), 64):
        st = codec.decode_tiles(res.bytes, make_dtiles(entries[i0:i0 + 64]), small, out_dtype="uint16", frame_size=res.frame_size, n_bytes=total)
        torch.cuda.synchronize()
        assert not st.cpu().numpy().any()
    assert torch.equal(big.view(torch.int16), small.view(torch.int16))

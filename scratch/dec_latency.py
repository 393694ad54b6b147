"""Decode latency of small calls (1..N tiles of 1024 x 1024): wall time of frx_decode_tiles incl. its launches, per decoder.
dec_latency.py [f32|u16x3]   env FRX_D3_MAX_FRAMES selects the phased decoder for calls up to that many frames"""
import sys, time, torch, numpy as np
sys.path.insert(0, '/root/repo')
from flac_raster_b200 import synth
from flac_raster_b200.codec import Codec, tile_grid, make_dtiles
kind = sys.argv[1] if len(sys.argv) > 1 else "f32"
codec = Codec(0)
H = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
if kind == "f32":
    r, nch, dt_name, bps = synth.dem_f32(H, H, seed=3), 1, "float32", 32
else:
    r, nch, dt_name, bps = synth.rgb_u16(H, H, seed=3), 3, "uint16", 16
tiles = tile_grid(H, H, 1024, 0, nch)
res = codec.encode_tiles(r, tiles); torch.cuda.synchronize()
total = int(res.total_bytes.item())
toff, tsize, mm = res.tile_offset.cpu().numpy(), res.tile_size.cpu().numpy(), res.tile_minmax.cpu().numpy()
entries = []
for i, (r0, c0, h, w, b0, n) in enumerate(tiles):
    entries.append(dict(byte_offset=toff[i], byte_size=tsize[i], out_offset=r0 * H + c0, out_band_stride=H * H, out_row_stride=H, h=h, w=w,
                        channels=n, bps=bps, dmin=mm[i, 0], dmax=mm[i, 1], first_frame=i * 256, n_frames=256))
out = torch.zeros_like(r)
ref = None
line = []
for nt in [int(x) for x in (sys.argv[3].split(",") if len(sys.argv) > 3 else "1,4,16,32,64".split(","))]:
    dt = make_dtiles(entries[:nt])
    st = torch.zeros(nt, dtype=torch.int32, device='cuda')
    ms = []
    for it in range(6):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        codec.decode_tiles(res.bytes, dt, out, out_dtype=dt_name, frame_size=res.frame_size, status=st, n_bytes=total)
        torch.cuda.synchronize(); ms.append((time.perf_counter() - t0) * 1e3)
    assert not st.cpu().numpy().any()
    line.append(f"{nt} tiles: {min(ms[1:]):.2f} ms (kernel {codec.last_kernel_ms()[1]:.2f})")
print(kind, "|", " | ".join(line), "| checksum", float(out.double().sum().item()))

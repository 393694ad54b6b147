import sys, torch
sys.path.insert(0, '/root/repo')
from flac_raster_b200 import synth
from flac_raster_b200.codec import Codec, tile_grid, make_dtiles, pcm_bps
codec = Codec(0)
H = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
r = synth.dem_f32(H, H, seed=1)
tiles = tile_grid(H, H, 1024, 0, 1)
res = codec.encode_tiles(r, tiles)
torch.cuda.synchronize()
total = int(res.total_bytes.cpu()[0])
toff = res.tile_offset.cpu().numpy(); tsize = res.tile_size.cpu().numpy(); mm = res.tile_minmax.cpu().numpy()
entries, ff = [], 0
for i, (r0, c0, h, w, b0, nch) in enumerate(tiles):
    nfr = (h * w + 4095) // 4096
    entries.append(dict(byte_offset=toff[i], byte_size=tsize[i], out_offset=b0 * H * H + r0 * H + c0, out_band_stride=H * H, out_row_stride=H,
                        h=h, w=w, channels=nch, bps=pcm_bps('float32', 'reference'), dmin=mm[i, 0], dmax=mm[i, 1], first_frame=ff, n_frames=nfr))
    ff += nfr
dt = make_dtiles(entries)
out = torch.empty_like(r)
status = torch.zeros(len(tiles), dtype=torch.int32, device='cuda')
for _ in range(3):
    codec.decode_tiles(res.bytes, dt, out, out_dtype='float32', frame_size=res.frame_size, status=status, n_bytes=total)
    torch.cuda.synchronize()
    print('decode kernel ms', codec.last_kernel_ms()[1], 'status', int(status.abs().sum()))

import sys, torch, numpy as np
sys.path.insert(0, '/root/repo')
from flac_raster_b200 import synth
from flac_raster_b200.codec import Codec, tile_grid, make_dtiles
codec = Codec(0)
H = 2048
r = synth.rgb_u16(H, H, seed=3)
tiles = tile_grid(H, H, 1024, 0, 3)
res = codec.encode_tiles(r, tiles); torch.cuda.synchronize()
total = int(res.total_bytes.item())
toff, tsize, mm = res.tile_offset.cpu().numpy(), res.tile_size.cpu().numpy(), res.tile_minmax.cpu().numpy()
e = [dict(byte_offset=toff[0], byte_size=tsize[0], out_offset=0, out_band_stride=H * H, out_row_stride=H, h=1024, w=1024, channels=3, bps=16,
          dmin=mm[0, 0], dmax=mm[0, 1], first_frame=0, n_frames=256)]
out = torch.zeros_like(r)
for _ in range(3):
    codec.decode_tiles(res.bytes, make_dtiles(e), out, out_dtype="uint16", frame_size=res.frame_size, n_bytes=total); torch.cuda.synchronize()
print("ok")

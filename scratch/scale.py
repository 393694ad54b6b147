import sys, torch
sys.path.insert(0, '/root/repo')
from flac_raster_b200 import synth
from flac_raster_b200.codec import Codec, tile_grid
codec = Codec(0)
for H in (1024, 2048, 4096, 8192):
    r = synth.dem_f32(H, H, seed=1)
    tiles = tile_grid(H, H, 1024, 0, 1)
    res = codec.encode_tiles(r, tiles)
    torch.cuda.synchronize()
    ts = []
    for _ in range(3):
        codec.encode_tiles(r, tiles, out=res); torch.cuda.synchronize(); ts.append(codec.last_kernel_ms()[0])
    nf = res.frame_size.numel()
    print(H, 'frames', nf, 'kernel ms', [round(t, 3) for t in ts], 'us/frame/cta', round(min(ts) * 1e3 / max(1, nf / 296), 2))

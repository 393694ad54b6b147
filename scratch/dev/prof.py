import sys, torch
sys.path.insert(0, '/root/repo')
from pathlib import Path
from flac_raster_b200 import _lib
tag = sys.argv[1]
if tag != 'prod': _lib.LIB_PATH = Path(f'/root/repo/scratch/dev/libfrx_{tag}.so')
from flac_raster_b200 import synth
from flac_raster_b200.codec import Codec, tile_grid
codec = Codec(0)
H = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
r = synth.dem_f32(H, H, seed=1)
tiles = tile_grid(H, H, 1024, 0, 1)
res = codec.encode_tiles(r, tiles)
for _ in range(2):
    codec.encode_tiles(r, tiles, out=res)
torch.cuda.synchronize()
print('kernel ms', codec.last_kernel_ms())

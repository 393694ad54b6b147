import sys, torch
sys.path.insert(0, '/root/repo')
from flac_raster_b200 import synth
from flac_raster_b200.codec import Codec, tile_grid
codec = Codec(0)
H = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
nch = int(sys.argv[2]) if len(sys.argv) > 2 else 3
r = synth.rgb_u16(H, H, seed=40000)
tiles = tile_grid(H, H, 1024, 0, nch)
res = codec.encode_tiles(r, tiles)
for _ in range(2):
    codec.encode_tiles(r, tiles, out=res)
torch.cuda.synchronize()
ms = codec.last_kernel_ms()[0]
nsub = H * H * nch / 4096
print(f'u16 {H}^2 x{nch}ch: kernel ms {ms:.3f}  ns/subframe {ms*1e6/nsub:.1f}  bytes {int(res.total_bytes.item())}')

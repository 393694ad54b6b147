import sys, torch
sys.path.insert(0, '/root/repo')
from flac_raster_b200 import synth
from flac_raster_b200.codec import Codec, tile_grid
codec = Codec(0)
H = 16384
for mode in ("reference", "lossless"):
    for nch in (1, 3):
        r = synth.rgb_u16(H, H, seed=40000)
        tiles = tile_grid(H, H, 1024, 0, nch)
        res = codec.encode_tiles(r, tiles, pcm_mode=mode)
        for _ in range(2):
            codec.encode_tiles(r, tiles, pcm_mode=mode, out=res)
        torch.cuda.synchronize()
        ms = codec.last_kernel_ms()[0]
        print(f'u16 {mode} x{nch}ch: kernel ms {ms:.3f}  ns/subframe {ms*1e6/(H*H*nch/4096):.1f}  bytes {int(res.total_bytes.item())}')

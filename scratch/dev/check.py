# byte-level regression + timing of a dev build: python scratch/dev/check.py <tag|prod>
import sys, hashlib, torch, numpy as np
sys.path.insert(0, '/root/repo')
from pathlib import Path
from flac_raster_b200 import _lib
tag = sys.argv[1]
if tag != 'prod':
    _lib.LIB_PATH = Path(f'/root/repo/scratch/dev/libfrx_{tag}.so')
from flac_raster_b200 import synth
from flac_raster_b200.codec import Codec, tile_grid
codec = Codec(0)
def run(name, r, tiles, reps=3):
    res = codec.encode_tiles(r, tiles)
    torch.cuda.synchronize()
    assert int(res.status.item()) == 0, name
    ms = []
    for _ in range(reps):
        codec.encode_tiles(r, tiles, out=res); torch.cuda.synchronize()
        ms.append(codec.last_kernel_ms()[0])
    n = int(res.total_bytes.item())
    h = hashlib.sha256(res.bytes[:n].cpu().numpy().tobytes()).hexdigest()[:16]
    h2 = hashlib.sha256(res.frame_size.cpu().numpy().tobytes()).hexdigest()[:8]
    print(f'{tag:8s} {name:28s} bytes {n:12d} sha {h} fs {h2} kernel ms {min(ms):.3f}')
g = torch.Generator(device='cuda').manual_seed(3)
run('dem_f32 8192 t1024', synth.dem_f32(8192, 8192, seed=1), tile_grid(8192, 8192, 1024, 0, 1))
r = synth.rgb_u16(4096, 4096, seed=2)
run('rgb_u16 4096 t1024 3ch', r, tile_grid(4096, 4096, 1024, 0, 3))
run('rgb_u16 4096 t1024 band1', r, tile_grid(4096, 4096, 1024, 0, 1))
x = (torch.randn(2, 1000, 1301, generator=g, device='cuda') * 300).cumsum(2).to(torch.int16)
run('i16 odd 1000x1301 t256 2ch', x, tile_grid(1000, 1301, 256, 0, 2))
y = torch.randint(-2**31, 2**31 - 1, (1, 700, 900), generator=g, device='cuda', dtype=torch.int64).to(torch.int32)
run('i32 noise fat', y, tile_grid(700, 900, 512, 0, 1))
run('dem_f32 20000 t1024 (C2)', synth.dem_f32(20000, 20000, seed=20000), tile_grid(20000, 20000, 1024, 0, 1), reps=5)

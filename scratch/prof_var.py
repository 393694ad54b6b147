"""Kernel time of the encode kernel for a given build of the library: prof_var.py <lib path>"""
import sys, torch
sys.path.insert(0, '/root/repo')
from pathlib import Path
import flac_raster_b200._lib as L
L.LIB_PATH = Path(sys.argv[1])
from flac_raster_b200 import synth
from flac_raster_b200.codec import Codec, tile_grid
codec = Codec(0)
out = []
for name, r, H, nch in (("f32 20000^2", synth.dem_f32(20000, 20000, seed=1), 20000, 1), ("u16x1", synth.rgb_u16(16384, 16384, seed=40000), 16384, 1),
                        ("u16x3", None, 16384, 3), ("u16x2 stereo", None, 16384, 2)):
    if r is None:
        r = prev
    prev = r
    tiles = tile_grid(H, H, 1024, 0, nch)
    res = codec.encode_tiles(r, tiles)
    ms = []
    for _ in range(4):
        codec.encode_tiles(r, tiles, out=res)
        torch.cuda.synchronize()
        ms.append(codec.last_kernel_ms()[0])
    out.append(f"{name}: {min(ms):.3f} ms (bytes {int(res.total_bytes.item())})")
print(Path(sys.argv[1]).name, " | ".join(out))

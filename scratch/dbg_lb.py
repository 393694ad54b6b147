import sys, ctypes, torch
sys.path.insert(0, '/root/repo')
import flac_raster_b200._lib as L
from pathlib import Path
L.LIB_PATH = Path('/root/repo/flac_raster_b200/libfrx_dbg.so')
from flac_raster_b200 import synth
from flac_raster_b200.codec import Codec, tile_grid
codec = Codec(0)
lib = L.lib()
H = 16384
def rd(reset=1):
    out = (ctypes.c_ulonglong * 16)()
    lib.frx_debug_read(out, reset)
    return list(out)
for name, r, nch in (("f32", synth.dem_f32(H, H, seed=1), 1), ("u16", synth.rgb_u16(H, H, seed=40000), 1), ("u16x3", synth.rgb_u16(H, H, seed=40000), 3)):
    tiles = tile_grid(H, H, 1024, 0, nch)
    res = codec.encode_tiles(r, tiles)
    torch.cuda.synchronize(); rd()
    codec.encode_tiles(r, tiles, out=res)
    torch.cuda.synchronize()
    d = rd()
    nf = H * H // 4096
    print(f"{name}: kernel {codec.last_kernel_ms()[0]:.2f} ms; per frame: polls {d[0]/nf:.1f} steps {d[1]/nf:.2f} lookback cycles {d[2]/max(1,d[3]):.0f} (frames {d[3]}), warp total cycles/frame {d[4]/nf:.0f}; analysis cycles: all {d[7]/nf:.0f}, fixed-winner frames {d[5]/max(1,d[6]):.0f} ({d[6]} frames), others {(d[7]-d[5])/max(1,nf-d[6]):.0f}; channel loads/frame {d[8]/nf:.2f}, 16-bit overflows {d[9]}")

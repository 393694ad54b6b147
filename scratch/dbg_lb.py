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
names = {0: "lookback", 2: "slot copy", 3: "frame", 5: "load", 6: "pass1", 7: "lpc_design", 8: "fixed cand", 9: "lpc resid+cost", 10: "winner fixed redo", 11: "count", 12: "materialise"}
cases = [("f32", synth.dem_f32(H, H, seed=1), 1), ("u16x1", synth.rgb_u16(H, H, seed=40000), 1), ("u16x3", synth.rgb_u16(H, H, seed=40000), 3)]
for name, r, nch in cases:
    tiles = tile_grid(H, H, 1024, 0, nch)
    res = codec.encode_tiles(r, tiles)
    torch.cuda.synchronize(); rd()
    codec.encode_tiles(r, tiles, out=res)
    torch.cuda.synchronize()
    d = rd()
    nf = max(1, d[4]); ns = max(1, d[15])
    print(f"{name}: kernel {codec.last_kernel_ms()[0]:.2f} ms; frames {d[4]} analyses {d[15]}; frames parked {d[12]/nf:.3f}; blocked polls/frame {d[1]/nf:.2f}")
    print("   cycles per frame: " + ", ".join(f"{n} {d[i]/nf:.0f}" for i, n in names.items() if i in (0, 2, 3)))
    print("   cycles per analysis: " + ", ".join(f"{n} {d[i]/ns:.0f}" for i, n in names.items() if i not in (0, 2, 3, 12)))

# phase cycle counters of a -DFRX_DBG dev build: python scratch/dev/phases.py <tag>
import sys, ctypes, torch
sys.path.insert(0, '/root/repo')
from pathlib import Path
from flac_raster_b200 import _lib
tag = sys.argv[1]
_lib.LIB_PATH = Path(f'/root/repo/scratch/dev/libfrx_{tag}.so')
from flac_raster_b200 import synth
from flac_raster_b200.codec import Codec, tile_grid
lib = ctypes.CDLL(str(_lib.LIB_PATH))
codec = Codec(0)
names = ['analyse L6 (w6)', 'lpc_design (w7)', 'analyse L0 (w0)', 'lb_rounds', 'lb_notready', 'PH footer(crc+copy)', 'frames', 'compute_total', 'PH lpcres ->SC', 'PH K1 ->S0', 'PH trial+autoc ->SA', 'PH decide/design ->SB', 'PH analyse ->SD', 'PH choose/scan ->SE', 'PH pack ->Sa', 'PH waitfull']
H = 8192
r = synth.dem_f32(H, H, seed=1)
tiles = tile_grid(H, H, 1024, 0, 1)
res = codec.encode_tiles(r, tiles)
buf = (ctypes.c_ulonglong * 16)()
lib.frx_debug_read(buf, 1)
codec.encode_tiles(r, tiles, out=res); torch.cuda.synchronize()
ms = codec.last_kernel_ms()[0]
lib.frx_debug_read(buf, 1)
nf = buf[6]
print(H, 'ms', round(ms, 3), 'frames', nf)
for i, n in enumerate(names):
    print('   ', n, buf[i], 'per frame', round(buf[i] / max(1, nf), 1))

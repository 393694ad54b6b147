"""Determinism stress: encode the same raster repeatedly, compare frame sizes and bytes with the first run.
stress_det.py <lib> [runs]"""
import sys, torch
sys.path.insert(0, '/root/repo')
from pathlib import Path
import flac_raster_b200._lib as L
L.LIB_PATH = Path(sys.argv[1])
runs = int(sys.argv[2]) if len(sys.argv) > 2 else 20
from flac_raster_b200 import synth
from flac_raster_b200.codec import Codec, tile_grid
codec = Codec(0)
cases = [("f32", synth.dem_f32(20000, 20000, seed=1), 20000, 1, 1024, {}), ("u16x3", synth.rgb_u16(8192, 8192, seed=40000), 8192, 3, 1024, {}),
         ("u16x1", synth.rgb_u16(8192, 8192, seed=7), 8192, 1, 1024, {}), ("u16x2", synth.rgb_u16(8192, 8192, seed=8), 8192, 2, 1024, {}),
         ("u16x3_t200", synth.rgb_u16(4000, 4000, seed=9), 4000, 3, 200, {}), ("u16x3_lossless_t256", synth.rgb_u16(4096, 4096, seed=10), 4096, 3, 256, {"pcm_mode": "lossless"})]
if len(sys.argv) > 3:
    cases = [c for c in cases if c[0] == sys.argv[3]]
for name, r, H, nch, tile, kw in cases:
    tiles = tile_grid(H, H, tile, 0, nch)
    res = codec.encode_tiles(r, tiles, **kw)
    torch.cuda.synchronize()
    fs0 = res.frame_size.clone(); tb0 = int(res.total_bytes.item()); b0 = res.bytes[:tb0].clone()
    bad = 0
    for i in range(runs):
        res.bytes.zero_()
        codec.encode_tiles(r, tiles, out=res, **kw)
        torch.cuda.synchronize()
        tb = int(res.total_bytes.item())
        d = (res.frame_size != fs0).nonzero().flatten()
        same = tb == tb0 and d.numel() == 0 and torch.equal(res.bytes[:tb0], b0)
        if not same:
            bad += 1
            nd = int((res.bytes[:min(tb, tb0)] != b0[:min(tb, tb0)]).sum().item())
            print(f"  {name} run {i}: total {tb} vs {tb0}; frames with another size {d[:6].tolist()} ({d.numel()}); sizes {res.frame_size[d[:3]].tolist()} vs {fs0[d[:3]].tolist()}; differing bytes {nd}")
    print(f"{Path(sys.argv[1]).name} {name}: {bad} of {runs} runs differ")

# interleaved A/B timing of dev builds on C2: python scratch/dev/ab.py prod a2 a3 ...   (each tag in its own subprocess, 3 rounds)
import sys, subprocess
if sys.argv[1] == '--one':
    import torch
    sys.path.insert(0, '/root/repo')
    from pathlib import Path
    from flac_raster_b200 import _lib
    tag = sys.argv[2]
    if tag != 'prod': _lib.LIB_PATH = Path(f'/root/repo/scratch/dev/libfrx_{tag}.so')
    from flac_raster_b200 import synth
    from flac_raster_b200.codec import Codec, tile_grid
    codec = Codec(0)
    H = int(sys.argv[3])
    r = synth.dem_f32(H, H, seed=20000); tiles = tile_grid(H, H, 1024, 0, 1)
    res = codec.encode_tiles(r, tiles); torch.cuda.synchronize()
    ms = []
    for _ in range(8):
        codec.encode_tiles(r, tiles, out=res); torch.cuda.synchronize(); ms.append(codec.last_kernel_ms()[0])
    ms.sort()
    clk = subprocess.run(['nvidia-smi', '--query-gpu=clocks.sm,temperature.gpu,power.draw', '--format=csv,noheader'], capture_output=True, text=True).stdout.strip()
    print(f'{tag:6s} min {ms[0]:.3f} med {ms[4]:.3f} max {ms[-1]:.3f}  [{clk}]')
else:
    tags = sys.argv[1:]
    for rnd in range(3):
        for t in tags:
            subprocess.run([sys.executable, __file__, '--one', t, '20000'])

import sys, time, torch, numpy as np
sys.path.insert(0, '/root/repo')
from flac_raster_b200 import synth
from flac_raster_b200.codec import Codec, EncodeResult, tile_grid
codec = Codec(0)
H = 20000
r = synth.dem_f32(H, H, seed=20000)
tiles = tile_grid(H, H, 1024, 0, 1)
res = codec.encode_tiles(r, tiles); torch.cuda.synchronize()
total = int(res.total_bytes.item()); nf = res.frame_size.numel()
host_t = torch.empty(r.shape, dtype=r.dtype, pin_memory=True); host_t.copy_(r); torch.cuda.synchronize()
host = host_t.numpy()
pin = lambda shape, dt: torch.empty(shape, dtype=dt, pin_memory=True)
hres = EncodeResult(bytes=pin(int(total * 1.05) + (1 << 20), torch.uint8).numpy(), tile_offset=pin(len(tiles), torch.int64).numpy(),
                    tile_size=pin(len(tiles), torch.int64).numpy(), tile_minmax=pin((len(tiles), 2), torch.float64).numpy(),
                    frame_size=pin(nf, torch.int32).numpy(), frame_offset=pin(nf, torch.int64).numpy(), total_bytes=pin(1, torch.int64).numpy(),
                    status=pin(1, torch.int32).numpy(), frames_per_tile=res.frames_per_tile, bps=32, blocksize=4096)
del r, res; torch.cuda.empty_cache()
for _ in range(2): codec.encode_tiles_host(host, tiles, out=hres)
ts = []
for _ in range(6):
    t0 = time.perf_counter(); codec.encode_tiles_host(host, tiles, out=hres); torch.cuda.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
print(sys.argv[1:], 'e2e ms min %.2f med %.2f' % (min(ts), sorted(ts)[3]))

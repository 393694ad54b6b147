# C5: 10 000 random bbox queries against a streaming file (reduced C3: 16384^2 x 3 uint16, tile 1024, --bands all)
import sys, time, numpy as np, torch
sys.path.insert(0, '/root/repo')
from flac_raster_b200 import synth, streaming
H = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
NQ = int(sys.argv[2]) if len(sys.argv) > 2 else 10000
r = synth.rgb_u16(H, H, seed=40000).cpu().numpy()
tiles = streaming.calculate_tiles(H, H, 1024)
from flac_raster_b200 import rio
transform = rio.Affine(1.0, 0.0, 0.0, 0.0, -1.0, float(H))
t0 = time.perf_counter()
chunks, sizes = streaming.encode_streaming_tiles(r, tiles, "EPSG:3857", transform, None, "all", "reference", 0)
offsets = np.concatenate([[0], np.cumsum(sizes)[:-1]])
index = streaming.build_index("EPSG:3857", transform, H, H, 1024, tiles, offsets, sizes)
streaming.write_streaming_file('/tmp/c5.flacs', index, chunks)
print('create-streaming s', round(time.perf_counter() - t0, 2), 'tiles', len(tiles), 'bytes', int(sum(sizes)))
sf = streaming.StreamingFile('/tmp/c5.flacs', 0)
rng = np.random.default_rng(5)
xs = rng.uniform(0, H - 64, NQ); ys = rng.uniform(0, H - 64, NQ)
queries = [{"bbox": [float(x), float(y), float(x + 50), float(y + 50)]} for x, y in zip(xs, ys)]
t0 = time.perf_counter()
pos = streaming.select_frames_batch(sf.index, queries)
t_sel = time.perf_counter() - t0
assert (pos >= 0).all()
sf.device_image(); torch.cuda.synchronize()
B = 1000
out = None
for warm in (True, False):
    t0 = time.perf_counter(); kms = 0.0
    for i in range(0, NQ, B):
        out, metas, status = sf.decode_device(pos[i:i + B], out=out)
        torch.cuda.synchronize()
        kms += streaming.get_codec(0).last_kernel_ms()[1]
        assert not bool(status.any().item())
    dt = time.perf_counter() - t0
    if not warm:
        print('select s', round(t_sel, 4), 'decode wall s', round(dt, 3), 'kernel ms', round(kms, 1), 'tiles/s wall', round(NQ / dt, 1), 'tiles/s kernel', round(NQ / (kms * 1e-3), 1))

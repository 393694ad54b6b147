# throughput of frx_tiff_unpack on file images built in device memory (no disk): python scratch/ingest_bench.py
import sys, torch
sys.path.insert(0, '/root/repo')
from flac_raster_b200.codec import Codec
codec = Codec(0)
def bench(name, dtype, tdt, bands, H, W, layout, planar):
    esz = torch.empty(0, dtype=tdt).element_size()
    chunks = []
    if layout[0] == 'strip':
        rps = layout[1]
        off = 8
        planes = [-1] if planar == 1 else list(range(bands))
        for b in planes:
            for r0 in range(0, H, rps):
                rows = min(rps, H - r0)
                chunks.append((off, r0, 0, rows, W, W, b))
                off += rows * W * (bands if b < 0 else 1) * esz
    else:
        tw, th = layout[1], layout[2]
        off = 8
        planes = [-1] if planar == 1 else list(range(bands))
        for b in planes:
            for r0 in range(0, H, th):
                for c0 in range(0, W, tw):
                    chunks.append((off, r0, c0, min(th, H - r0), min(tw, W - c0), tw, b))
                    off += th * tw * (bands if b < 0 else 1) * esz
    img = torch.randint(0, 255, (off,), dtype=torch.uint8, device='cuda')
    out = torch.empty((bands, H, W), dtype=tdt, device='cuda')
    for _ in range(3):
        codec.tiff_unpack(img, chunks, dtype, bands, H, W, bands, out=out)
    torch.cuda.synchronize()
    ms = []
    for _ in range(10):
        codec.tiff_unpack(img, chunks, dtype, bands, H, W, bands, out=out); torch.cuda.synchronize()
        ms.append(codec.last_kernel_ms()[0])
    t = min(ms)
    nbytes = 2 * bands * H * W * esz
    print(f'{name:44s} {t:8.3f} ms  {nbytes / t / 1e6:8.1f} GB/s (read + write), {len(chunks)} chunks')
bench('C2 f32 1 band 20000^2 strips(8)', 'float32', torch.float32, 1, 20000, 20000, ('strip', 8), 1)
bench('C2 f32 1 band 20000^2 tiles 512', 'float32', torch.float32, 1, 20000, 20000, ('tile', 512, 512), 1)
bench('C3 u16 3 band interleaved 20000^2 strips(1)', 'uint16', torch.uint16, 3, 20000, 20000, ('strip', 1), 1)
bench('C3 u16 3 band interleaved 20000^2 tiles 256', 'uint16', torch.uint16, 3, 20000, 20000, ('tile', 256, 256), 1)
bench('C3 u16 3 band planar 20000^2 strips(16)', 'uint16', torch.uint16, 3, 20000, 20000, ('strip', 16), 2)
bench('C4 i16 12 band interleaved 10000^2 strips(1)', 'int16', torch.int16, 12, 10000, 10000, ('strip', 1), 1)
bench('u8 3 band interleaved 16384^2 tiles 512', 'uint8', torch.uint8, 3, 16384, 16384, ('tile', 512, 512), 1)

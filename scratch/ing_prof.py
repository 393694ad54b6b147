# one frx_tiff_unpack launch on the C3 layout (3-band uint16 pixel-interleaved strips, 1 row each) for ncu
import sys, torch
sys.path.insert(0, '/root/repo')
from flac_raster_b200.codec import Codec
codec = Codec(0)
H = W = 20000; bands = 3; esz = 2
img = torch.randint(0, 255, (8 + H * W * bands * esz,), dtype=torch.uint8, device='cuda')
strips = [(8 + r * W * bands * esz, r, 0, 1, W, W, -1) for r in range(H)]
out = torch.empty((bands, H, W), dtype=torch.uint16, device='cuda')
for _ in range(3):
    codec.tiff_unpack(img, strips, 'uint16', bands, H, W, bands, out=out)
torch.cuda.synchronize()
print('unpack ms', codec.last_kernel_ms()[0])

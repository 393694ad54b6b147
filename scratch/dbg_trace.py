import sys, ctypes, torch, numpy as np
sys.path.insert(0, '/root/repo')
from pathlib import Path
from flac_raster_b200 import _lib
_lib.LIB_PATH = Path('/root/repo/scratch/dev/libfrx_%s.so' % sys.argv[2])
lib = ctypes.CDLL(str(_lib.LIB_PATH))
from flac_raster_b200 import synth
from flac_raster_b200.codec import Codec, tile_grid
lib.frx_debug_trace.argtypes = [ctypes.c_void_p]
codec = Codec(0)
H = int(sys.argv[1])
r = synth.dem_f32(H, H, seed=1)
tiles = tile_grid(H, H, 1024, 0, 1)
res = codec.encode_tiles(r, tiles)
nf = res.frame_size.numel()
tr = torch.zeros(nf * 8, dtype=torch.int64, device='cuda')
lib.frx_debug_trace(tr.data_ptr())
codec.encode_tiles(r, tiles, out=res); torch.cuda.synchronize()
lib.frx_debug_trace(None)
t = tr.cpu().numpy().reshape(nf, 8).astype(np.int64)
t0 = t[:, 0].min()
res_t, A, lb0, lb1, start, blk, cta, copy = [t[:, k] for k in range(8)]
us = lambda x: (x - t0) / 1e3
print('kernel ms', codec.last_kernel_ms()[0], 'span us', us(lb1.max()))
print('reserve->start  mean %.1f p50 %.1f p99 %.1f max %.1f' % tuple(np.percentile((start - res_t) / 1e3, q) if q else ((start - res_t) / 1e3).mean() for q in (0, 50, 99, 100)))
print('start->A        mean %.1f p50 %.1f p99 %.1f max %.1f' % tuple(np.percentile((A - start) / 1e3, q) if q else ((A - start) / 1e3).mean() for q in (0, 50, 99, 100)))
print('A->lbstart      mean %.1f p50 %.1f p99 %.1f max %.1f' % tuple(np.percentile((lb0 - A) / 1e3, q) if q else ((lb0 - A) / 1e3).mean() for q in (0, 50, 99, 100)))
print('lb duration     mean %.1f p50 %.1f p99 %.1f max %.1f' % tuple(np.percentile((lb1 - lb0) / 1e3, q) if q else ((lb1 - lb0) / 1e3).mean() for q in (0, 50, 99, 100)))
# A time vs frame index: how out-of-order?
Aus = us(A)
order = np.argsort(Aus)
print('A-time inversion: max over f of (max A-time of frames < f) - A(f):')
runmax = np.maximum.accumulate(Aus)
lag = runmax - Aus
print('   mean %.1f p50 %.1f p99 %.1f max %.1f' % (lag.mean(), np.percentile(lag, 50), np.percentile(lag, 99), lag.max()))
# blocking distance
m = blk > 0
d = (np.arange(nf)[m] - blk[m])
if m.sum(): print('blocked look-backs', m.sum(), 'distance mean %.1f p50 %.1f p99 %.1f max %d' % (d.mean(), np.percentile(d, 50), np.percentile(d, 99), d.max()))
# the frames that were waited on: reserve->A of those vs all
bl = np.unique(blk[m])
if m.sum(): print('blockers', len(bl), 'their reserve->A mean %.1f vs all %.1f' % (((A - res_t)[bl] / 1e3).mean(), ((A - res_t) / 1e3).mean()))
# per-CTA frame counts
c = np.bincount(cta.astype(int), minlength=296)
print('frames per CTA min %d max %d' % (c.min(), c.max()))
# timeline sample: for CTA 0, print its frames
for cid in (0, 17):
    idx = np.where(cta == cid)[0]
    print('CTA', cid)
    for f in idx[:12]:
        print('   f %6d reserve %8.1f start %8.1f A %8.1f lb0 %8.1f lb1 %8.1f copy %8.1f blk %d' % (f, us(res_t[f]), us(start[f]), us(A[f]), us(lb0[f]), us(lb1[f]), us(copy[f]) if copy[f] else -1, blk[f]))
np.save('/root/repo/gpurun_out/trace.npy', t)

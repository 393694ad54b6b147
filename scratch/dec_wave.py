import sys, torch, numpy as np
sys.path.insert(0, '/root/repo')
from flac_raster_b200 import synth
from flac_raster_b200.codec import Codec, tile_grid, make_dtiles
codec = Codec(0)
H = 20000
r = synth.dem_f32(H, H, seed=20000)
tiles = tile_grid(H, H, 1024, 0, 1)
res = codec.encode_tiles(r, tiles); torch.cuda.synchronize()
total = int(res.total_bytes.item())
toff, tsize, mm = res.tile_offset.cpu().numpy(), res.tile_size.cpu().numpy(), res.tile_minmax.cpu().numpy()
def run(sel, name):
    entries = []
    ffs = np.concatenate([[0], np.cumsum(res.frames_per_tile)])
    for i in sel:
        r0, c0, h, w, b0, nch = tiles[i]
        entries.append(dict(byte_offset=toff[i], byte_size=tsize[i], out_offset=r0 * H + c0, out_band_stride=H * H, out_row_stride=H, h=h, w=w,
                            channels=1, bps=32, dmin=mm[i, 0], dmax=mm[i, 1], first_frame=int(ffs[i]), n_frames=int(res.frames_per_tile[i])))
    dt = make_dtiles(entries)
    out = torch.empty_like(r)
    st = torch.zeros(len(sel), dtype=torch.int32, device='cuda')
    ms = []
    for _ in range(5):
        codec.decode_tiles(res.bytes, dt, out, out_dtype='float32', frame_size=res.frame_size, status=st, n_bytes=total); torch.cuda.synchronize()
        ms.append(codec.last_kernel_ms()[1])
    nf = sum(e['n_frames'] for e in entries)
    print(f'{name}: tiles {len(sel)} frames {nf} warps {(nf + 31) // 32} kernel ms {min(ms):.3f}')
full = [i for i, t in enumerate(tiles) if t[2] == 1024 and t[3] == 1024]
run(list(range(len(tiles))), 'all')
run(full, 'full tiles (361)')
run(full[:180], 'half')
run(full[:90], 'quarter')

import sys, numpy as np, torch
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import oracle
from oracle import quant
from flac_raster_b200.codec import Codec, tile_grid
from flac_raster_b200 import synth
codec = Codec(0)
def run(name, raster, tile, nch, level=5):
    H, W = raster.shape[1:]
    tiles = tile_grid(H, W, tile, 0, nch)
    dev = synth.from_numpy(raster)
    res = codec.encode_tiles(dev, tiles, level=level)
    torch.cuda.synchronize()
    total = int(res.total_bytes.item())
    data = res.bytes[:total].cpu().numpy()
    toff, tsize = res.tile_offset.cpu().numpy(), res.tile_size.cpu().numpy()
    fs = res.frame_size.cpu().numpy()
    print(name, 'status', int(res.status.item()), 'total', total, 'frames', len(fs), 'fs[:6]', fs[:6])
    bad = 0
    for i, (r0, c0, h, w, b0, nc) in enumerate(tiles):
        block = raster[b0:b0+nc, r0:r0+h, c0:c0+w]
        pcm_ref, bps, dmin, dmax, _ = quant.reference_pcm(block)
        pcm, frames, rc = oracle.decode_frames(bytes(data[toff[i]:toff[i]+tsize[i]]), nc, bps, max_samples=h*w+16)
        ok8 = all(f['crc8_ok'] for f in frames); ok16 = all(f['crc16_ok'] for f in frames)
        same = pcm.shape == pcm_ref.shape and np.array_equal(pcm, pcm_ref)
        if not (rc == 0 and ok8 and ok16 and same):
            bad += 1
            if bad <= 3:
                print('  tile', i, 'rc', rc, 'nframes', len(frames), 'crc8', ok8, 'crc16', [f['crc16_ok'] for f in frames][:8], 'pcm', same, pcm.shape, pcm_ref.shape)
                if frames: print('   sf0', frames[0]['subframes'], 'size', frames[0]['byte_size'])
                if pcm.shape == pcm_ref.shape and not same:
                    d = np.nonzero((pcm != pcm_ref).any(axis=1))[0]; print('   first diff at', d[:5], pcm[d[0]], pcm_ref[d[0]])
    ref = sum(len(oracle.encode_stream(quant.reference_pcm(raster[b0:b0+nc, r0:r0+h, c0:c0+w])[0], res.bps, 44100, 4096, level, write_header=False)[0]) for (r0,c0,h,w,b0,nc) in tiles[:4])
    print('  bad tiles', bad, 'of', len(tiles), ' size ours(first4)', int(tsize[:4].sum()), 'oracle', ref)
rng = np.random.default_rng(0)
yy, xx = np.mgrid[0:256, 0:256]
dem = ((1000 + 300*np.sin(xx/20.)*np.cos(yy/30.) + rng.random((256,256))*50)/8192).astype(np.float32)[None]
run('f32 256 tile128', dem, 128, 1)
i16 = np.stack([(1200 + 400*np.sin(xx/17.)*np.cos(yy/23.) + rng.normal(0,4,xx.shape)), (900+300*np.cos(xx/11.)+rng.normal(0,6,xx.shape)), (500+100*np.sin(yy/7.)+rng.normal(0,3,xx.shape))]).astype(np.int16)
run('i16x3 tile128', i16, 128, 3)
run('i16x1 tile100', i16, 100, 1)
run('u8 tile 64', (i16[:2] // 8).astype(np.uint8), 64, 2)
run('f32 level0', dem, 128, 1, level=0)
run('f32 level8', dem, 128, 1, level=8)
print('---- byte dump')
def dump(raster, tile, nch):
    H, W = raster.shape[1:]
    tiles = tile_grid(H, W, tile, 0, nch)
    res = codec.encode_tiles(synth.from_numpy(raster), tiles)
    torch.cuda.synchronize()
    data = res.bytes[:int(res.total_bytes.item())].cpu().numpy()
    fo = res.frame_offset.cpu().numpy(); fs = res.frame_size.cpu().numpy()
    for k in range(min(6, len(fo))):
        b = data[fo[k]:fo[k]+fs[k]]
        print(k, int(fo[k]), int(fs[k]), bytes(b[:12]).hex(), '...', bytes(b[-6:]).hex())
dump(dem, 128, 1)
dump(dem, 128, 1)

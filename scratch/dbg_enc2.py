"""Debug driver for the encode kernel: encodes small seeded cases and reports, frame by frame, what the oracle's
decoder makes of them (first failing frame, its decisions, first PCM mismatch)."""
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import oracle  # noqa: E402
from oracle import quant  # noqa: E402
from flac_raster_b200.codec import Codec  # noqa: E402

codec = Codec(0)


def run(name, raster, tiles, **kw):
    t = torch.from_numpy(raster.view(np.int16) if raster.dtype == np.uint16 else raster).cuda()
    if raster.dtype == np.uint16:
        t = t.view(torch.uint16)
    res = codec.encode_tiles(t, tiles, **kw)
    torch.cuda.synchronize()
    total = int(res.total_bytes.item())
    data = res.bytes[:total].cpu().numpy()
    fsz = res.frame_size.cpu().numpy()
    foff = res.frame_offset.cpu().numpy()
    toff, tsz = res.tile_offset.cpu().numpy(), res.tile_size.cpu().numpy()
    print(f"== {name}: status {int(res.status.item())} total {total} frames {len(fsz)} sizes {fsz[:6]} offs {foff[:4]}")
    fpos = 0
    bad = 0
    for i, (r0, c0, h, w, b0, nch) in enumerate(tiles):
        block = raster[b0:b0 + nch, r0:r0 + h, c0:c0 + w]
        mode = kw.get("pcm_mode", "reference")
        if mode == "reference":
            pcm_ref, bps, *_ = quant.reference_pcm(block)
        else:
            from tests.helpers import lossless_pcm
            pcm_ref, bps, *_ = lossless_pcm(block)
        nfr = (h * w + 4095) // 4096
        for j in range(nfr):
            o, s = int(foff[fpos + j]), int(fsz[fpos + j])
            pcm, frames, rc = oracle.decode_frames(bytes(data[o:o + s]), nch, bps, max_samples=8192)
            want = pcm_ref[j * 4096:(j + 1) * 4096]
            ok = rc == 0 and len(frames) == 1 and frames[0]["crc16_ok"] and pcm.shape == want.shape and np.array_equal(pcm, want)
            if not ok:
                bad += 1
                if bad <= 4:
                    info = frames[0] if frames else None
                    print(f"  tile {i} frame {j}: rc {rc} size {s} hdr {bytes(data[o:o + 8]).hex()} info {info and (info['crc8_ok'], info['crc16_ok'], info['chan_assign'], info['subframes'])}")
                    if pcm.shape == want.shape and pcm.size:
                        d = np.argwhere(pcm != want)
                        print("   first mismatch", d[:3].tolist(), pcm[tuple(d[0])] if len(d) else None, want[tuple(d[0])] if len(d) else None)
                    enc, fs = oracle.encode_stream(want, bps, 44100, 4096, 5, write_header=False)
                    _, fr2, _ = oracle.decode_frames(enc, nch, bps, max_samples=8192)
                    print("   oracle would write", len(enc), fr2[0]["subframes"] if fr2 else None)
                    # the oracle numbers its frame 0: compare from the byte after the header
                    ours = bytes(data[o:o + s])
                    ea = np.frombuffer(enc, dtype=np.uint8)
                    oa = np.frombuffer(ours, dtype=np.uint8)
                    hl = 6 if j < 128 else 7
                    if len(ea) == len(oa):
                        dd = np.nonzero(ea[hl:-2] != oa[hl:-2])[0]
                        print("   differing bytes vs oracle:", len(dd), "first", (dd[:8] + hl).tolist(), "last", (dd[-3:] + hl).tolist() if len(dd) else None)
                        if len(dd):
                            k0 = int(dd[0]) + hl
                            print("   ours  ", ours[k0 - 4:k0 + 12].hex(), "\n   oracle", enc[k0 - 4:k0 + 12].hex())
        fpos += nfr
    print(f"   bad frames: {bad}")
    return bad


rng = np.random.default_rng(1)
yy, xx = np.mgrid[0:128, 0:256]
base = (1200 + 400 * np.sin(xx / 17.0) * np.cos(yy / 23.0) + rng.normal(0, 4, xx.shape))
which = sys.argv[1:] or ["i16_1", "i16_3", "u8_3", "f32", "short", "stereo", "lossless32"]
if "i16_1" in which:
    run("int16 1ch 128x256 (8 full frames)", base.astype(np.int16)[None], [(0, 0, 128, 256, 0, 1)])
if "i16_3" in which:
    r = np.stack([base, base * 0.5 + 30, base[::-1]]).astype(np.int16)
    run("int16 3ch", r, [(0, 0, 128, 256, 0, 3)])
if "u8_3" in which:
    r = np.stack([base / 8, base / 9 + 3, base[::-1] / 10]).astype(np.uint8)
    run("uint8 3ch", r, [(0, 0, 128, 256, 0, 3)])
if "f32" in which:
    run("float32 1ch", (base / 8192).astype(np.float32)[None], [(0, 0, 128, 256, 0, 1)])
if "short" in which:
    run("int16 1ch short frames 50x100 + 30x37", base.astype(np.int16)[None], [(0, 0, 50, 100, 0, 1), (60, 8, 30, 37, 0, 1)])
if "stereo" in which:
    r = np.stack([base, base + rng.integers(-3, 4, base.shape)]).astype(np.int16)
    run("int16 2ch stereo", r, [(0, 0, 128, 256, 0, 2)])
if "lossless32" in which:
    r = (base * 100000).astype(np.int32)[None]
    run("int32 lossless fat", r, [(0, 0, 128, 256, 0, 1)], pcm_mode="lossless")

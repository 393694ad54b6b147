import sys, collections
import numpy as np, torch
sys.path.insert(0, '/root/repo')
import oracle
from flac_raster_b200 import synth
from flac_raster_b200.codec import Codec, tile_grid
codec = Codec(0)
for name, r, bps in (("u16", synth.rgb_u16(2048, 2048, seed=40000), 16), ("f32", synth.dem_f32(2048, 2048, seed=20000), 32)):
    tiles = tile_grid(2048, 2048, 1024, 0, 1)
    res = codec.encode_tiles(r, tiles)
    torch.cuda.synchronize()
    data = res.bytes[:int(res.total_bytes.item())].cpu().numpy()
    pcm, frames, rc = oracle.decode_frames(bytes(data), 1, bps, max_samples=4 * 1024 * 1024 + 16)
    c = collections.Counter((f["subframes"][0]["type"], f["subframes"][0]["order"], f["subframes"][0]["part_order"]) for f in frames)
    print(name, rc, len(frames), c.most_common(8))

"""Shared checkers for the parity tests: everything the CUDA path emits is checked against oracle/."""
import numpy as np

import oracle
from oracle import quant


def tiles_for(raster, tile, bands="all"):
    b, H, W = raster.shape
    nch = b if bands == "all" else 1
    out = []
    for r in range(0, H, tile):
        for c in range(0, W, tile):
            out.append((r, c, min(tile, H - r), min(tile, W - c), 0, nch))
    return out


def check_encode_against_oracle(raster_np, tiles, res_bytes, tile_offset, tile_size, tile_minmax, frame_size,
                                blocksize=4096, level=5, size_margin=1.02, check_size=True, sample_rate=44100):
    """P1 (PCM == reference quantiser), P2 (independent decoder accepts the stream and returns the PCM),
    size gate vs the libFLAC-level restatement.  Returns (ours_total, oracle_total)."""
    ours_total = oracle_total = 0
    fpos = 0
    for i, (r0, c0, h, w, b0, nch) in enumerate(tiles):
        block = raster_np[b0:b0 + nch, r0:r0 + h, c0:c0 + w]
        pcm_ref, bps, dmin, dmax, _sr = quant.reference_pcm(block)
        off, size = int(tile_offset[i]), int(tile_size[i])
        data = bytes(res_bytes[off:off + size])
        pcm, frames, rc = oracle.decode_frames(data, nch, bps, sample_rate, max_samples=h * w + 16)
        assert rc == 0, f"tile {i}: oracle decoder rejected the stream rc={rc}"
        assert all(f["crc8_ok"] and f["crc16_ok"] for f in frames)
        assert pcm.shape == pcm_ref.shape, (pcm.shape, pcm_ref.shape)
        assert np.array_equal(pcm, pcm_ref), f"tile {i}: PCM mismatch"
        nfr = (h * w + blocksize - 1) // blocksize
        assert len(frames) == nfr
        assert [f["number"] for f in frames] == list(range(nfr))
        assert [f["byte_size"] for f in frames] == [int(x) for x in frame_size[fpos:fpos + nfr]]
        fpos += nfr
        mm = tile_minmax[i]
        assert (mm[0] == dmin or (np.isnan(mm[0]) and np.isnan(dmin))) and (mm[1] == dmax or (np.isnan(mm[1]) and np.isnan(dmax)))
        ours_total += size
        if check_size:
            enc, _ = oracle.encode_stream(pcm_ref, bps, sample_rate, blocksize, level, write_header=False)
            oracle_total += len(enc)
    if check_size:
        assert ours_total <= size_margin * oracle_total + 16, (ours_total, oracle_total)
    return ours_total, oracle_total

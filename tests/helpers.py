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


def lossless_pcm(block):
    """pcm_mode=lossless (SURVEY decision D1): identity mapping into a signed PCM of the effective width.
    -> (int32 PCM (h*w, bands), bps, dmin, dmax, sample_rate) like quant.reference_pcm."""
    bands = block.shape[0]
    a = block.transpose(1, 2, 0).reshape(-1, bands)
    dt = block.dtype
    if dt == np.uint16:
        pcm, bps = a.astype(np.int32) - 32768, 16
    elif dt == np.uint8 or dt == np.int16:
        pcm, bps = a.astype(np.int32), 16
    elif dt == np.int32:
        pcm, bps = a.astype(np.int32), 32
    elif dt == np.uint32:
        pcm, bps = (a ^ np.uint32(0x80000000)).view(np.int32), 32
    else:
        b = np.ascontiguousarray(a).view(np.int32).astype(np.int64)
        pcm, bps = np.where(b >= 0, b, 0x80000000 - (b & 0xFFFFFFFF)).astype(np.int64).astype(np.int32), 32
    return pcm, bps, float(np.min(block)), float(np.max(block)), 44100


def spatial_pcm(block):
    """pcm_mode=reference_spatial: spatial_encoder.py:229-248 + pyFLAC truncation (oracle/quant.py)."""
    pcm, bps = quant.spatial_reference_pcm(block)
    return pcm, bps, float(np.min(block)), float(np.max(block)), 44100


def check_encode_against_oracle(raster_np, tiles, res_bytes, tile_offset, tile_size, tile_minmax, frame_size,
                                blocksize=4096, level=5, size_margin=1.02, check_size=True, sample_rate=44100,
                                pcm_fn=None):
    """P1 (PCM == reference quantiser), P2 (independent decoder accepts the stream and returns the PCM),
    size gate vs the libFLAC-level restatement.  Returns (ours_total, oracle_total).
    pcm_fn: block -> (pcm, bps, dmin, dmax, sr); default = the reference quantiser (pcm_mode=reference)."""
    pcm_fn = pcm_fn or quant.reference_pcm
    ours_total = oracle_total = 0
    fpos = 0
    for i, (r0, c0, h, w, b0, nch) in enumerate(tiles):
        block = raster_np[b0:b0 + nch, r0:r0 + h, c0:c0 + w]
        pcm_ref, bps, dmin, dmax, _sr = pcm_fn(block)
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


def write_tiff_variant(path, arr, tile=None, rows_per_strip=None, planar=1, big_endian=False, pad=0):
    """Test-only TIFF writer for the layouts the ingestion kernel must handle: `arr` (C,H,W); strips (rows_per_strip) or
    tiles (tile=(tw,th), multiples of 16, edge tiles zero-padded); PlanarConfiguration 1 or 2; either byte order;
    `pad` stray bytes in front of every chunk (odd values make the chunks unaligned)."""
    import struct
    arr = np.asarray(arr)
    C, H, W = arr.shape
    bo = ">" if big_endian else "<"
    dt = arr.dtype.newbyteorder(bo)
    chunks = []
    if tile:
        tw, th = tile
        tx, ty = (W + tw - 1) // tw, (H + th - 1) // th
        padded = np.zeros((C, ty * th, tx * tw), dtype=arr.dtype)
        padded[:, :H, :W] = arr
        planes = [None] if planar == 1 else list(range(C))
        for c in planes:
            for j in range(ty):
                for i in range(tx):
                    blk = padded[:, j * th:(j + 1) * th, i * tw:(i + 1) * tw]
                    chunks.append(blk.transpose(1, 2, 0).astype(dt).tobytes() if c is None else blk[c].astype(dt).tobytes())
    else:
        rps = rows_per_strip or H
        planes = [None] if planar == 1 else list(range(C))
        for c in planes:
            for r0 in range(0, H, rps):
                blk = arr[:, r0:r0 + rps, :]
                chunks.append(blk.transpose(1, 2, 0).astype(dt).tobytes() if c is None else blk[c].astype(dt).tobytes())
    body = bytearray(b"\0" * 8)
    offsets, counts = [], []
    for ch in chunks:
        body += b"\xAA" * pad
        offsets.append(len(body))
        counts.append(len(ch))
        body += ch
    if len(body) & 1:
        body += b"\0"
    sf = {"u": 1, "i": 2, "f": 3}[arr.dtype.kind]
    entries = [(256, 4, [W]), (257, 4, [H]), (258, 3, [arr.dtype.itemsize * 8] * C), (259, 3, [1]), (262, 3, [1]),
               (277, 3, [C]), (284, 3, [planar]), (339, 3, [sf] * C)]
    if tile:
        entries += [(322, 4, [tile[0]]), (323, 4, [tile[1]]), (324, 4, offsets), (325, 4, counts)]
    else:
        entries += [(273, 4, offsets), (278, 4, [rows_per_strip or H]), (279, 4, counts)]
    entries.sort()
    extra = bytearray()
    ifd_off = len(body)
    extra_base = ifd_off + 2 + 12 * len(entries) + 4
    ifd = struct.pack(bo + "H", len(entries))
    for tag, typ, vals in entries:
        fmt = {3: "H", 4: "I"}[typ]
        raw = struct.pack(bo + fmt * len(vals), *vals)
        if len(raw) <= 4:
            ifd += struct.pack(bo + "HHI", tag, typ, len(vals)) + raw.ljust(4, b"\0")
        else:
            ifd += struct.pack(bo + "HHII", tag, typ, len(vals), extra_base + len(extra))
            extra += raw
            if len(extra) & 1:
                extra += b"\0"
    ifd += struct.pack(bo + "I", 0)
    body[0:8] = (b"MM" if big_endian else b"II") + struct.pack(bo + "HI", 42, ifd_off)
    with open(path, "wb") as f:
        f.write(bytes(body) + ifd + bytes(extra))

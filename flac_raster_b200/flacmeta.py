"""FLAC container blocks (fLaC marker, STREAMINFO, VORBIS_COMMENT, PADDING) -- host side.

Replaces the reference's use of ``mutagen.flac.FLAC`` (converter.py:315-427,
spatial_encoder.py:296-373, 385-462).  The reference lets libFLAC write the stream header
and then has mutagen REWRITE the whole file to grow the comment block (which is what makes
its --spatial byte offsets stale); here the header is composed once, in front of
the frames the GPU produced, so offsets are correct by construction.
"""
from __future__ import annotations

import struct
from dataclasses import dataclass, field

VENDOR = "flac-raster-b200 libfrx 0.1"
BLOCK_STREAMINFO, BLOCK_PADDING, BLOCK_APPLICATION, BLOCK_SEEKTABLE, BLOCK_VORBIS_COMMENT = 0, 1, 2, 3, 4


@dataclass
class StreamInfo:
    min_blocksize: int = 4096
    max_blocksize: int = 4096
    min_framesize: int = 0
    max_framesize: int = 0
    sample_rate: int = 44100
    channels: int = 1
    bps: int = 16
    total_samples: int = 0
    md5: bytes = b"\0" * 16

    def pack(self) -> bytes:
        v = (self.sample_rate << 44) | ((self.channels - 1) << 41) | ((self.bps - 1) << 36) | (self.total_samples & ((1 << 36) - 1))
        return (struct.pack(">HH", self.min_blocksize, self.max_blocksize) + self.min_framesize.to_bytes(3, "big") +
                self.max_framesize.to_bytes(3, "big") + v.to_bytes(8, "big") + self.md5)

    @classmethod
    def unpack(cls, b: bytes) -> "StreamInfo":
        mnb, mxb = struct.unpack(">HH", b[:4])
        v = int.from_bytes(b[10:18], "big")
        return cls(mnb, mxb, int.from_bytes(b[4:7], "big"), int.from_bytes(b[7:10], "big"), v >> 44, ((v >> 41) & 7) + 1,
                   ((v >> 36) & 31) + 1, v & ((1 << 36) - 1), bytes(b[18:34]))


@dataclass
class FlacHeader:
    streaminfo: StreamInfo
    vendor: str = ""
    comments: list = field(default_factory=list)  # [(KEY, value)] in file order
    first_frame_offset: int = 0
    blocks: list = field(default_factory=list)    # [(type, offset, length)]
    applications: dict = field(default_factory=dict)  # {4-byte id: payload}

    def get(self, key: str, default=None):
        key = key.upper()
        for k, v in self.comments:
            if k.upper() == key:
                return v
        return default

    def __contains__(self, key: str) -> bool:
        return self.get(key) is not None


def _block(btype: int, payload: bytes, last: bool) -> bytes:
    if len(payload) >= 1 << 24:
        raise ValueError("FLAC metadata block larger than 16 MiB")
    return bytes([(0x80 if last else 0) | btype]) + len(payload).to_bytes(3, "big") + payload


def pack_vorbis_comment(comments, vendor: str = VENDOR) -> bytes:
    v = vendor.encode("utf-8")
    out = [struct.pack("<I", len(v)), v, struct.pack("<I", len(comments))]
    for k, val in comments:
        e = f"{k}={val}".encode("utf-8")
        out.append(struct.pack("<I", len(e)))
        out.append(e)
    return b"".join(out)


FRAME_TABLE_APP_ID = b"frxT"


def pack_frame_table(blocksize: int, frame_sizes) -> bytes:
    """APPLICATION 'frxT' payload: u8 version, u32 blocksize, u32 n, u32 sizes[n] (big-endian).
    Lets the batched decoder skip frame discovery; other readers skip the block."""
    sizes = [int(x) for x in frame_sizes]
    return struct.pack(">BII", 1, int(blocksize), len(sizes)) + struct.pack(f">{len(sizes)}I", *sizes)


def unpack_frame_table(payload: bytes):
    ver, blocksize, n = struct.unpack(">BII", payload[:9])
    if ver != 1 or len(payload) < 9 + 4 * n:
        raise ValueError("bad frxT frame table")
    return blocksize, list(struct.unpack(f">{n}I", payload[9:9 + 4 * n]))


def build_header(si: StreamInfo, comments=None, padding=None, vendor: str = VENDOR, applications=None) -> bytes:
    """fLaC + STREAMINFO [+ APPLICATION...] [+ VORBIS_COMMENT] [+ PADDING]."""
    blocks = [(BLOCK_STREAMINFO, si.pack())]
    for app_id, payload in (applications or {}).items():
        assert len(app_id) == 4
        blocks.append((BLOCK_APPLICATION, bytes(app_id) + payload))
    if comments is not None:
        blocks.append((BLOCK_VORBIS_COMMENT, pack_vorbis_comment(comments, vendor)))
    if padding is not None:  # a zero-length PADDING block is legal
        blocks.append((BLOCK_PADDING, b"\0" * padding))
    out = [b"fLaC"]
    for i, (t, p) in enumerate(blocks):
        out.append(_block(t, p, i == len(blocks) - 1))
    return b"".join(out)


def parse_header(data: bytes, offset: int = 0) -> FlacHeader:
    """Parse the metadata blocks of the stream that starts at data[offset:]."""
    if data[offset:offset + 4] != b"fLaC":
        raise ValueError("not a FLAC stream (missing fLaC marker)")
    pos = offset + 4
    hdr = None
    last = False
    blocks = []
    apps = {}
    vendor, comments = "", []
    while not last:
        if pos + 4 > len(data):
            raise ValueError("truncated FLAC metadata")
        last = bool(data[pos] & 0x80)
        btype = data[pos] & 0x7F
        blen = int.from_bytes(data[pos + 1:pos + 4], "big")
        pos += 4
        if pos + blen > len(data):
            raise ValueError("truncated FLAC metadata block")
        blocks.append((btype, pos, blen))
        if hdr is None:
            if btype != BLOCK_STREAMINFO or blen != 34:
                raise ValueError("first metadata block is not STREAMINFO")
            hdr = StreamInfo.unpack(data[pos:pos + 34])
        elif btype == BLOCK_APPLICATION and blen >= 4:
            apps[bytes(data[pos:pos + 4])] = bytes(data[pos + 4:pos + blen])
        elif btype == BLOCK_VORBIS_COMMENT:
            q = pos
            vl = struct.unpack("<I", data[q:q + 4])[0]; q += 4
            vendor = data[q:q + vl].decode("utf-8", "replace"); q += vl
            n = struct.unpack("<I", data[q:q + 4])[0]; q += 4
            for _ in range(n):
                cl = struct.unpack("<I", data[q:q + 4])[0]; q += 4
                e = data[q:q + cl].decode("utf-8", "replace"); q += cl
                k, _, v = e.partition("=")
                comments.append((k, v))
        pos += blen
    return FlacHeader(hdr, vendor, comments, pos, blocks, apps)

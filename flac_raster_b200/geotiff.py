"""Minimal GeoTIFF reader/writer with a rasterio-shaped surface.

The reference does its raster I/O through rasterio/GDAL (converter.py:137-143, 285-309;
spatial_encoder.py:154-183; cli.py:668-729).  rasterio is not installable on the B200
boxes, so this module supplies the small subset the hot path's callers use:
``open(path)`` -> dataset with ``read()``, ``read(band, window=...)``, ``width``,
``height``, ``count``, ``dtypes``, ``crs``, ``transform``, ``bounds``, ``nodata``,
``meta``, ``window_transform``; and ``open(path, 'w', **meta)`` -> ``write()``.

Supported files: classic TIFF and BigTIFF, either byte order, uncompressed, strips or
tiles, PlanarConfiguration 1 (pixel interleaved) or 2 (band planar), SampleFormat
uint/int/float at 8/16/32/64 bits; GeoTIFF tags 33550/33922/34264/34735/34736/34737 and
GDAL_NODATA 42113.  That covers every fixture the reference ships (test_data/*.tif).
"""
from __future__ import annotations

import io
import struct
from collections import namedtuple
from pathlib import Path

import numpy as np

BoundingBox = namedtuple("BoundingBox", "left bottom right top")


class Affine(tuple):
    """(a, b, c, d, e, f) with x = a*col + b*row + c, y = d*col + e*row + f.
    Iterates as 9 numbers like ``affine.Affine`` (converter.py:163, cli.py:681)."""

    def __new__(cls, a, b, c, d, e, f, *rest):
        return super().__new__(cls, (float(a), float(b), float(c), float(d), float(e), float(f), 0.0, 0.0, 1.0))

    a = property(lambda s: s[0]); b = property(lambda s: s[1]); c = property(lambda s: s[2])
    d = property(lambda s: s[3]); e = property(lambda s: s[4]); f = property(lambda s: s[5])

    def __mul__(self, other):
        if isinstance(other, Affine):
            oa, ob, oc, od, oe, of = other[:6]
            return Affine(self.a * oa + self.b * od, self.a * ob + self.b * oe, self.a * oc + self.b * of + self.c,
                          self.d * oa + self.e * od, self.d * ob + self.e * oe, self.d * oc + self.e * of + self.f)
        x, y = other
        return (x * self.a + y * self.b + self.c, x * self.d + y * self.e + self.f)

    @classmethod
    def translation(cls, xoff, yoff):
        return cls(1.0, 0.0, xoff, 0.0, 1.0, yoff)

    def __repr__(self):
        return "Affine(%r, %r, %r,\n       %r, %r, %r)" % tuple(self[:6])

    __str__ = lambda self: "|% .2f,% .2f,% .2f|\n|% .2f,% .2f,% .2f|\n| 0.00, 0.00, 1.00|" % tuple(self[:6])


def from_origin(west, north, xsize, ysize):
    """rasterio.transform.from_origin"""
    return Affine(xsize, 0.0, west, 0.0, -ysize, north)


class Window:
    """rasterio.windows.Window(col_off, row_off, width, height)"""

    def __init__(self, col_off, row_off, width, height):
        self.col_off, self.row_off, self.width, self.height = col_off, row_off, width, height

    def __repr__(self):
        return f"Window(col_off={self.col_off}, row_off={self.row_off}, width={self.width}, height={self.height})"

    def __eq__(self, o):
        return (self.col_off, self.row_off, self.width, self.height) == (o.col_off, o.row_off, o.width, o.height)


class CRS:
    """Just enough of rasterio.crs.CRS: string round trip ("EPSG:4326")."""

    def __init__(self, s):
        self._s = str(s) if s is not None else ""

    @classmethod
    def from_string(cls, s):
        return cls(s)

    @classmethod
    def from_epsg(cls, code):
        return cls(f"EPSG:{int(code)}")

    def to_string(self):
        return self._s

    def to_epsg(self):
        if self._s.upper().startswith("EPSG:"):
            try:
                return int(self._s.split(":", 1)[1])
            except ValueError:
                return None
        return None

    def __str__(self):
        return self._s

    def __repr__(self):
        return f"CRS.from_string({self._s!r})"

    def __bool__(self):
        return bool(self._s)

    def __eq__(self, o):
        return str(self) == str(o)

    def __hash__(self):
        return hash(self._s)


_TYPE_FMT = {1: "B", 2: "c", 3: "H", 4: "I", 5: "II", 6: "b", 7: "B", 8: "h", 9: "i", 10: "ii", 11: "f", 12: "d",
             16: "Q", 17: "q", 18: "Q"}
_TYPE_SIZE = {1: 1, 2: 1, 3: 2, 4: 4, 5: 8, 6: 1, 7: 1, 8: 2, 9: 4, 10: 8, 11: 4, 12: 8, 16: 8, 17: 8, 18: 8}


def _np_dtype(sample_format, bits, bo):
    kind = {1: "u", 2: "i", 3: "f"}.get(sample_format)
    if kind is None or bits not in (8, 16, 32, 64) or (kind == "f" and bits < 32):
        raise ValueError(f"unsupported TIFF sample format {sample_format}/{bits} bits")
    return np.dtype(f"{bo}{kind}{bits // 8}")


class DatasetReader:
    def __init__(self, path):
        self.name = str(path)
        self._f = io.open(path, "rb")
        self._parse()

    # ---- parsing -----------------------------------------------------------------
    def _parse(self):
        f = self._f
        hdr = f.read(16)
        if hdr[:2] == b"II":
            bo = "<"
        elif hdr[:2] == b"MM":
            bo = ">"
        else:
            raise ValueError("not a TIFF file")
        self._bo = bo
        magic = struct.unpack(bo + "H", hdr[2:4])[0]
        if magic == 42:
            big = False
            ifd = struct.unpack(bo + "I", hdr[4:8])[0]
        elif magic == 43:
            big = True
            ifd = struct.unpack(bo + "Q", hdr[8:16])[0]
        else:
            raise ValueError("not a TIFF file")
        f.seek(ifd)
        if big:
            n = struct.unpack(bo + "Q", f.read(8))[0]
            esz, cfmt, vsz = 20, "HHQ", 8
        else:
            n = struct.unpack(bo + "H", f.read(2))[0]
            esz, cfmt, vsz = 12, "HHI", 4
        raw = f.read(esz * n)
        tags = {}
        for i in range(n):
            e = raw[i * esz:(i + 1) * esz]
            tag, typ, cnt = struct.unpack(bo + cfmt, e[: esz - vsz])
            if typ not in _TYPE_SIZE:
                continue
            size = _TYPE_SIZE[typ] * cnt
            if size <= vsz:
                data = e[esz - vsz: esz - vsz + size]
            else:
                off = struct.unpack(bo + ("Q" if big else "I"), e[esz - vsz:])[0]
                pos = f.tell()
                f.seek(off)
                data = f.read(size)
                f.seek(pos)
            if typ == 2:
                val = data.split(b"\0")[0].decode("latin-1")
            elif typ in (5, 10):
                v = struct.unpack(bo + _TYPE_FMT[typ][0] * (2 * cnt), data)
                val = tuple(v[2 * j] / v[2 * j + 1] if v[2 * j + 1] else 0.0 for j in range(cnt))
            else:
                val = struct.unpack(bo + _TYPE_FMT[typ] * cnt, data)
            tags[tag] = val
        self.tags = tags
        g = lambda t, d=None: tags.get(t, d)
        self.width = int(g(256)[0])
        self.height = int(g(257)[0])
        self.count = int(g(277, (1,))[0])
        bits = g(258, (1,))
        if len(set(bits)) != 1:
            raise ValueError("mixed bit depths unsupported")
        if int(g(259, (1,))[0]) != 1:
            raise ValueError("compressed TIFFs are not supported by the built-in reader")
        sf = g(339, (1,))
        self._dtype = _np_dtype(int(sf[0]), int(bits[0]), bo)
        self.dtypes = tuple(str(self._dtype.newbyteorder("=")) for _ in range(self.count))
        self._planar = int(g(284, (1,))[0])
        self._tiled = 322 in tags
        if self._tiled:
            self._tw, self._th = int(g(322)[0]), int(g(323)[0])
            self._offsets, self._counts = g(324), g(325)
        else:
            self._rps = min(int(g(278, (self.height,))[0]), self.height)
            self._offsets, self._counts = g(273), g(279)
        # georeferencing
        self.transform = Affine(1.0, 0.0, 0.0, 0.0, 1.0, 0.0)
        if 34264 in tags:
            m = tags[34264]
            self.transform = Affine(m[0], m[1], m[3], m[4], m[5], m[7])
        elif 33550 in tags and 33922 in tags:
            sx, sy = tags[33550][0], tags[33550][1]
            tp = tags[33922]
            self.transform = Affine(sx, 0.0, tp[3] - tp[0] * sx, 0.0, -sy, tp[4] + tp[1] * sy)
        self.crs = None
        gk = tags.get(34735)
        if gk:
            keys = {gk[4 + 4 * i]: gk[4 + 4 * i + 1: 4 + 4 * i + 4] for i in range(gk[3])}
            for kid in (3072, 2048):
                if kid in keys and keys[kid][0] == 0 and keys[kid][2] not in (0, 32767):
                    self.crs = CRS.from_epsg(keys[kid][2])
                    break
        nd = tags.get(42113)
        self.nodata = None
        if nd:
            try:
                self.nodata = float(nd)
            except ValueError:
                self.nodata = None

    # ---- rasterio-like surface ----------------------------------------------------
    @property
    def bounds(self):
        t = self.transform
        x0, y0 = t * (0, 0)
        x1, y1 = t * (self.width, self.height)
        return BoundingBox(min(x0, x1), min(y0, y1), max(x0, x1), max(y0, y1))

    @property
    def meta(self):
        return {"driver": "GTiff", "dtype": self.dtypes[0], "nodata": self.nodata, "width": self.width,
                "height": self.height, "count": self.count, "crs": self.crs, "transform": self.transform}

    def window_transform(self, window):
        return self.transform * Affine.translation(window.col_off, window.row_off)

    def _read_rows(self, r0, r1):
        """(count, r1-r0, width) for rows [r0, r1)"""
        f, dt, W, C = self._f, self._dtype, self.width, self.count
        out = np.empty((C, r1 - r0, W), dtype=dt.newbyteorder("="))
        if not self._tiled:
            rps = self._rps
            spp = (self.height + rps - 1) // rps  # strips per plane
            for s in range(r0 // rps, (r1 - 1) // rps + 1):
                sr0 = s * rps
                nrows = min(rps, self.height - sr0)
                a, b = max(r0, sr0), min(r1, sr0 + nrows)
                if self._planar == 1:
                    f.seek(self._offsets[s] + (a - sr0) * W * C * dt.itemsize)
                    blk = np.frombuffer(f.read((b - a) * W * C * dt.itemsize), dtype=dt).reshape(b - a, W, C)
                    out[:, a - r0: b - r0, :] = blk.transpose(2, 0, 1)
                else:
                    for c in range(C):
                        f.seek(self._offsets[c * spp + s] + (a - sr0) * W * dt.itemsize)
                        blk = np.frombuffer(f.read((b - a) * W * dt.itemsize), dtype=dt).reshape(b - a, W)
                        out[c, a - r0: b - r0, :] = blk
        else:
            tw, th = self._tw, self._th
            tx, ty = (W + tw - 1) // tw, (self.height + th - 1) // th
            for j in range(r0 // th, (r1 - 1) // th + 1):
                a, b = max(r0, j * th), min(r1, (j + 1) * th)
                for i in range(tx):
                    cw = min(tw, W - i * tw)
                    if self._planar == 1:
                        f.seek(self._offsets[j * tx + i])
                        blk = np.frombuffer(f.read(tw * th * C * dt.itemsize), dtype=dt).reshape(th, tw, C)
                        out[:, a - r0: b - r0, i * tw: i * tw + cw] = blk[a - j * th: b - j * th, :cw, :].transpose(2, 0, 1)
                    else:
                        for c in range(C):
                            f.seek(self._offsets[c * tx * ty + j * tx + i])
                            blk = np.frombuffer(f.read(tw * th * dt.itemsize), dtype=dt).reshape(th, tw)
                            out[c, a - r0: b - r0, i * tw: i * tw + cw] = blk[a - j * th: b - j * th, :cw]
        return out

    # ---- device ingestion (include/frx.h frx_tiff_unpack) ---------------------------------
    def chunk_table(self, row0: int = 0, row1: int | None = None):
        """Strips/tiles intersecting rows [row0, row1) as frx_tiff_chunk tuples
        (file_offset, dst_row0, dst_col0, rows, cols, pitch, band), destination rows relative to `row0`.
        The same geometry _read_rows() walks on the host."""
        H, W, C = self.height, self.width, self.count
        row1 = H if row1 is None else row1
        esz = self._dtype.itemsize
        out = []
        if not self._tiled:
            rps = self._rps
            spp = (H + rps - 1) // rps
            for s in range(row0 // rps, (row1 - 1) // rps + 1):
                sr0 = s * rps
                a, b = max(row0, sr0), min(row1, sr0 + min(rps, H - sr0))
                if self._planar == 1:
                    out.append((self._offsets[s] + (a - sr0) * W * C * esz, a - row0, 0, b - a, W, W, -1))
                else:
                    for c in range(C):
                        out.append((self._offsets[c * spp + s] + (a - sr0) * W * esz, a - row0, 0, b - a, W, W, c))
        else:
            tw, th = self._tw, self._th
            tx, ty = (W + tw - 1) // tw, (H + th - 1) // th
            for j in range(row0 // th, (row1 - 1) // th + 1):
                a, b = max(row0, j * th), min(row1, min((j + 1) * th, H))
                for i in range(tx):
                    cw = min(tw, W - i * tw)
                    if self._planar == 1:
                        out.append((self._offsets[j * tx + i] + (a - j * th) * tw * C * esz, a - row0, i * tw, b - a, cw, tw, -1))
                    else:
                        for c in range(C):
                            out.append((self._offsets[c * tx * ty + j * tx + i] + (a - j * th) * tw * esz, a - row0, i * tw,
                                        b - a, cw, tw, c))
        return out

    def read_device(self, codec, row0: int = 0, row1: int | None = None, bands=None):
        """read() without the host-side rearrangement: the byte span of the file that holds the chunks goes to the
        device as it is (pinned staging buffer, one H2D copy) and frx_tiff_unpack lays it out as (count, rows, width).
        bands: 0-based band indices wanted; a band-planar file (PlanarConfiguration 2) then uploads only those bands' chunks
        (pixel-interleaved files hold every band in every byte run, so all of it travels and the unwanted planes are dropped
        on the device)."""
        import torch
        row1 = self.height if row1 is None else row1
        chunks = self.chunk_table(row0, row1)
        if bands is not None and all(c[6] >= 0 for c in chunks):
            pos = {int(b): i for i, b in enumerate(bands)}
            chunks = [tuple(c[:6]) + (pos[c[6]],) for c in chunks if c[6] in pos]
            return self._upload_and_unpack(codec, chunks, len(pos), 1, row0, row1)
        out = self._upload_and_unpack(codec, chunks, self.count, self.count, row0, row1)
        return out if bands is None else torch.cat([out[b:b + 1] for b in bands])   # (no fancy indexing for unsigned dtypes)

    def _upload_and_unpack(self, codec, chunks, count, spp, row0, row1):
        """chunks -> (count, row1 - row0, width) on the device; spp = bands stored per pixel in interleaved chunks."""
        import torch
        esz = self._dtype.itemsize
        lo = min(c[0] for c in chunks)
        hi = max(c[0] + ((c[3] - 1) * c[5] + c[4]) * (spp if c[6] < 0 else 1) * esz for c in chunks)
        lo_al = lo & ~255                       # keeps the file's own alignment phase on the device
        try:
            host = torch.empty(hi - lo_al, dtype=torch.uint8, pin_memory=True)
        except RuntimeError:                    # not enough lockable memory for the whole span: a pageable copy still works
            host = torch.empty(hi - lo_al, dtype=torch.uint8)
        self._f.seek(lo_al)
        got = self._f.readinto(memoryview(host.numpy()))
        if got != hi - lo_al:
            raise ValueError(f"{self.name}: truncated TIFF (wanted {hi - lo_al} bytes at {lo_al}, got {got})")
        dev = host.to(f"cuda:{codec.device}", non_blocking=True)
        rel = [(c[0] - lo_al,) + tuple(c[1:]) for c in chunks]
        swap = (self._bo == ">") != (__import__("sys").byteorder == "big")
        return codec.tiff_unpack(dev, rel, self.dtypes[0], count, row1 - row0, self.width, spp, swap_bytes=swap)

    def read(self, indexes=None, window=None):
        if window is None:
            r0, r1, c0, c1 = 0, self.height, 0, self.width
        else:
            r0, r1 = int(window.row_off), int(window.row_off + window.height)
            c0, c1 = int(window.col_off), int(window.col_off + window.width)
        data = self._read_rows(r0, r1)[:, :, c0:c1]
        if indexes is None:
            return np.ascontiguousarray(data)
        if isinstance(indexes, int):
            return np.ascontiguousarray(data[indexes - 1])
        return np.ascontiguousarray(data[[i - 1 for i in indexes]])

    def close(self):
        self._f.close()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


class DatasetWriter:
    """Writes an uncompressed, strip-organised, pixel-interleaved GeoTIFF (BigTIFF when > 3.9 GB)."""

    def __init__(self, path, driver="GTiff", width=None, height=None, count=1, dtype=None, crs=None, transform=None,
                 nodata=None, **_ignored):
        self.name = str(path)
        self.width, self.height, self.count = int(width), int(height), int(count)
        self._dtype = np.dtype(dtype)
        self.crs = CRS(crs) if (crs is not None and not isinstance(crs, CRS)) else crs
        self.transform = transform
        self.nodata = nodata
        self._data = np.zeros((self.count, self.height, self.width), dtype=self._dtype)

    def write(self, arr, indexes=None):
        arr = np.asarray(arr)
        if indexes is None:
            if arr.ndim == 2:
                arr = arr[None]
            self._data[...] = arr.astype(self._dtype, copy=False)
        else:
            self._data[int(indexes) - 1] = arr.astype(self._dtype, copy=False)

    def close(self):
        C, H, W = self._data.shape
        dt = self._dtype
        pix = np.ascontiguousarray(self._data.transpose(1, 2, 0)).astype(dt.newbyteorder("<"), copy=False)
        nbytes = pix.nbytes
        big = nbytes > 3_900_000_000
        rps = max(1, min(H, 8192 // max(1, W * C * dt.itemsize)))
        nstrips = (H + rps - 1) // rps
        strip_bytes = [min(rps, H - s * rps) * W * C * dt.itemsize for s in range(nstrips)]
        sf = {"u": 1, "i": 2, "f": 3}[dt.kind]
        entries = []  # (tag, type, values)
        entries.append((256, 4 if big else 3, (W,)) if W > 65535 else (256, 3, (W,)))
        entries.append((257, 4, (H,)) if H > 65535 else (257, 3, (H,)))
        entries.append((258, 3, (dt.itemsize * 8,) * C))
        entries.append((259, 3, (1,)))
        entries.append((262, 3, (2 if (C == 3 and dt == np.uint8) else 1,)))
        entries.append((273, 16 if big else 4, None))  # patched below
        entries.append((277, 3, (C,)))
        entries.append((278, 4 if rps > 65535 else 3, (rps,)))
        entries.append((279, 16 if big else 4, tuple(strip_bytes)))
        entries.append((284, 3, (1,)))
        if C > 3 or (C > 1 and not (C == 3 and dt == np.uint8)):
            entries.append((338, 3, (0,) * (C - 1)))
        entries.append((339, 3, (sf,) * C))
        if self.transform is not None:
            t = self.transform
            if t[1] == 0.0 and t[3] == 0.0:
                entries.append((33550, 12, (t[0], -t[4], 0.0)))
                entries.append((33922, 12, (0.0, 0.0, 0.0, t[2], t[5], 0.0)))
            else:
                entries.append((34264, 12, (t[0], t[1], 0.0, t[2], t[3], t[4], 0.0, t[5], 0, 0, 0, 0, 0, 0, 0, 1.0)))
        epsg = self.crs.to_epsg() if self.crs else None
        if epsg:
            if epsg == 4326 or 4000 <= epsg < 5000:
                gk = (1, 1, 0, 3, 1024, 0, 1, 2, 1025, 0, 1, 1, 2048, 0, 1, epsg)
            else:
                gk = (1, 1, 0, 3, 1024, 0, 1, 1, 1025, 0, 1, 1, 3072, 0, 1, epsg)
            entries.append((34735, 3, gk))
        if self.nodata is not None:
            entries.append((42113, 2, repr(float(self.nodata)).encode() + b"\0"))
        entries.sort(key=lambda e: e[0])
        # layout: header | IFD | out-of-line values | pixel data
        osz = 8 if big else 4
        hdr_len = 16 if big else 8
        ifd_len = (8 + 20 * len(entries) + 8) if big else (2 + 12 * len(entries) + 4)
        ool = bytearray()
        ool_base = hdr_len + ifd_len

        def pack_vals(typ, vals):
            if typ == 2:
                return bytes(vals)
            return struct.pack("<" + _TYPE_FMT[typ] * len(vals), *vals)

        # first pass to find size of out-of-line area (strip offsets depend on it)
        sizes = []
        for tag, typ, vals in entries:
            cnt = nstrips if vals is None else len(vals)
            sizes.append(_TYPE_SIZE[typ] * cnt)
        ool_total = sum((s + 1) & ~1 for s in sizes if s > osz)
        data_off = (ool_base + ool_total + 15) & ~15
        offs, p = [], data_off
        for sb in strip_bytes:
            offs.append(p)
            p += sb
        ifd = bytearray()
        for (tag, typ, vals), size in zip(entries, sizes):
            if vals is None:
                vals = tuple(offs)
            cnt = len(vals)
            payload = pack_vals(typ, vals)
            if size <= osz:
                field = payload.ljust(osz, b"\0")
            else:
                field = struct.pack("<Q" if big else "<I", ool_base + len(ool))
                ool += payload
                if len(ool) & 1:
                    ool += b"\0"
            ifd += struct.pack("<HHQ" if big else "<HHI", tag, typ, cnt) + field
        with io.open(self.name, "wb") as f:
            if big:
                f.write(b"II" + struct.pack("<HHHQ", 43, 8, 0, 16))
                f.write(struct.pack("<Q", len(entries)) + ifd + struct.pack("<Q", 0))
            else:
                f.write(b"II" + struct.pack("<HI", 42, 8))
                f.write(struct.pack("<H", len(entries)) + ifd + struct.pack("<I", 0))
            f.write(ool)
            f.write(b"\0" * (data_off - ool_base - len(ool)))
            f.write(pix.tobytes())

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


def open(path, mode="r", **kwargs):  # noqa: A001 - mirrors rasterio.open
    if mode == "r":
        return DatasetReader(path)
    if mode == "w":
        return DatasetWriter(path, **kwargs)
    raise ValueError(f"unsupported mode {mode!r}")

"""Generate tests/golden/* by running the REFERENCE's own Python (run in the build container only).

The reference (/root/reference) is importable here once ``rasterio`` and ``pyflac`` are
stubbed (they are only used inside functions we do not call); its numpy staging functions
then run unmodified.  The GPU box has no /root/reference, so the outputs are committed:

  tests/golden/quant_golden.npz   inputs + outputs of the reference's
      RasterFLACConverter._normalize_to_audio / _denormalize_from_audio /
      _calculate_audio_params, SpatialFLACEncoder._normalize_to_audio / _calculate_tiles /
      _tile_to_bbox, SpatialIndex.query_bbox, SpatialFLACStreamer.get_byte_ranges_for_bbox
  tests/golden/sample_*.{tif,flac,json}   the reference's shipped fixtures (test_data/), copied
      verbatim as DATA (they are the libFLAC-1.4.3-produced golden vectors G1/G4/G5).

Usage:  python oracle/gen_golden.py
"""
from __future__ import annotations

import json
import shutil
import sys
import types
import warnings
from pathlib import Path

import numpy as np

REF = Path("/root/reference")
OUT = Path(__file__).resolve().parents[1] / "tests" / "golden"


def _install_stubs():
    class Affine(tuple):
        def __new__(cls, a, b, c, d, e, f, *r):
            return super().__new__(cls, (a, b, c, d, e, f, 0.0, 0.0, 1.0))

        def __mul__(self, pt):
            x, y = pt
            return (x * self[0] + y * self[1] + self[2], x * self[3] + y * self[4] + self[5])

    class CRS:
        def __init__(self, s):
            self.s = s

        @classmethod
        def from_string(cls, s):
            return cls(s)

        def __str__(self):
            return self.s

    class Window:
        def __init__(self, col_off, row_off, width, height):
            self.col_off, self.row_off, self.width, self.height = col_off, row_off, width, height

    ras = types.ModuleType("rasterio")
    crs = types.ModuleType("rasterio.crs"); crs.CRS = CRS
    tr = types.ModuleType("rasterio.transform"); tr.Affine = Affine
    win = types.ModuleType("rasterio.windows"); win.Window = Window
    ras.crs, ras.transform, ras.windows = crs, tr, win
    sys.modules.update({"rasterio": ras, "rasterio.crs": crs, "rasterio.transform": tr, "rasterio.windows": win,
                        "pyflac": types.ModuleType("pyflac")})
    return Affine, CRS, Window


def read_tiff_pixels(path):
    """raw strip reader for the shipped fixtures (classic LE TIFF, uncompressed, PlanarConfig 1)."""
    import struct
    d = Path(path).read_bytes()
    assert d[:2] == b"II"
    off = struct.unpack("<I", d[4:8])[0]
    n = struct.unpack("<H", d[off:off + 2])[0]
    tags = {}
    for i in range(n):
        tag, typ, cnt, val = struct.unpack("<HHII", d[off + 2 + 12 * i: off + 14 + 12 * i])
        size = {1: 1, 2: 1, 3: 2, 4: 4, 12: 8}[typ] * cnt
        raw = d[off + 10 + 12 * i: off + 14 + 12 * i] if size <= 4 else d[val: val + size]
        fmt = {1: "B", 2: "c", 3: "H", 4: "I", 12: "d"}[typ]
        tags[tag] = struct.unpack("<" + fmt * cnt, raw[:size])
    W, H, C = tags[256][0], tags[257][0], tags.get(277, (1,))[0]
    bits, sf = tags[258][0], tags.get(339, (1,))[0]
    dt = np.dtype({(8, 1): "u1", (16, 1): "<u2", (16, 2): "<i2", (32, 3): "<f4", (32, 2): "<i4"}[(bits, sf)])
    buf = b"".join(d[o: o + c] for o, c in zip(tags[273], tags[279]))
    return np.frombuffer(buf, dtype=dt).reshape(H, W, C).transpose(2, 0, 1).copy()


def main():
    OUT.mkdir(parents=True, exist_ok=True)
    Affine, CRS, Window = _install_stubs()
    sys.path.insert(0, str(REF / "src"))
    import logging
    logging.disable(logging.CRITICAL)
    from flac_raster.converter import RasterFLACConverter
    from flac_raster.spatial_encoder import SpatialFLACEncoder, SpatialFrame, SpatialIndex, SpatialFLACStreamer

    conv = RasterFLACConverter()
    g = {}
    rng = np.random.default_rng(20251104)

    # ---- fixtures (data, not source) -------------------------------------------------
    for name in ["sample_rgb.tif", "sample_rgb.flac", "sample_rgb.json", "sample_rgb_reconstructed.tif",
                 "sample_dem.tif", "sample_dem.flac", "sample_multispectral.tif"]:
        shutil.copyfile(REF / "test_data" / name, OUT / name)
    rgb = read_tiff_pixels(REF / "test_data/sample_rgb.tif")
    dem = read_tiff_pixels(REF / "test_data/sample_dem.tif")
    ms = read_tiff_pixels(REF / "test_data/sample_multispectral.tif")
    g["fixture_rgb"], g["fixture_dem"], g["fixture_ms"] = rgb, dem, ms

    # ---- _normalize_to_audio on the fixtures (whole raster = one convert call) ---------
    cases = {"rgb": rgb, "dem": dem, "ms": ms}
    # seeded synthetic cases for every dtype the converter labels, incl. edge cases
    cases["u8_small"] = rng.integers(0, 256, size=(2, 17, 33), dtype=np.uint8)
    cases["u8_narrow"] = rng.integers(100, 104, size=(1, 40, 40), dtype=np.uint8)
    cases["u16_12bit"] = rng.integers(0, 4096, size=(3, 64, 64)).astype(np.uint16)
    cases["u16_full"] = rng.integers(0, 65536, size=(1, 128, 64)).astype(np.uint16)
    cases["i16_dem"] = (1000 + 300 * np.sin(np.arange(96 * 96).reshape(1, 96, 96) * 0.01)
                        + rng.normal(0, 5, (1, 96, 96))).astype(np.int16)
    cases["i16_wrap"] = np.array([[[-32768, 0, 32767, 100, -100, 12345, -12345, 1]]], dtype=np.int16)
    cases["i16_neg"] = rng.integers(-3000, -1000, size=(2, 31, 7)).astype(np.int16)
    cases["i32_mid"] = rng.integers(-1_000_000, 1_000_000, size=(1, 50, 50)).astype(np.int32)
    cases["i32_wide"] = np.array([[[-2147483648, 2147483647, 0, 5, -7, 1 << 30, -(1 << 30)]]], dtype=np.int32)
    cases["u32_mid"] = rng.integers(0, 3_000_000_000, size=(1, 20, 20)).astype(np.uint32)
    cases["const_u16"] = np.full((1, 16, 16), 777, dtype=np.uint16)
    cases["f32_norm"] = (rng.random((1, 64, 64), dtype=np.float32) * 2 - 1).astype(np.float32)
    cases["f32_dem"] = ((1000 + 300 * rng.random((1, 48, 48))) / 8192).astype(np.float32)
    cases["f32_overflow"] = np.array([[[0.5, -0.25, 255.9, 256.0, 300.0, -256.0, -257.0, np.nan, np.inf, -np.inf,
                                        1e-9, -1e-9]]], dtype=np.float32)
    names = sorted(cases)
    g["case_names"] = np.array(names)
    with warnings.catch_warnings(), np.errstate(all="ignore"):
        warnings.simplefilter("ignore")
        for nm in names:
            r = cases[nm]
            bands = r.shape[0]
            sr, label = conv._calculate_audio_params(r, r.dtype)
            reshaped = r.transpose(1, 2, 0).reshape(-1, bands)
            audio = conv._normalize_to_audio(reshaped, label)
            g[f"in_{nm}"] = r
            g[f"audio_{nm}"] = audio
            g[f"params_{nm}"] = np.array([sr, label])
            g[f"minmax_{nm}"] = np.array([float(np.min(r)), float(np.max(r))])
            # decode side: what flac_to_tiff would hand back for this PCM (see oracle/quant.py for
            # the pyFLAC/soundfile glue, which is NOT reference code and is restated there)
            eff = audio.astype(np.int32)
            bps = audio.dtype.itemsize * 8
            s16 = eff.astype(np.int16) if bps == 16 else (eff >> 16).astype(np.int16)
            f64 = s16.astype(np.float64) / 32768.0
            if bands > 1:
                rr = f64.reshape(r.shape[1], r.shape[2], bands).transpose(2, 0, 1)
            else:
                rr = f64.reshape(r.shape[1], r.shape[2])
            den = conv._denormalize_from_audio(rr, r.dtype, float(np.min(r)), float(np.max(r)))
            g[f"denorm_{nm}"] = den.reshape(r.shape)
        # sample-rate thresholds (bands x height)
        g["sr_shapes"] = np.array([[1, 999_999], [1, 1_000_000], [3, 3_333_334], [1, 10_000_000], [2, 50_000_000],
                                   [12, 10_000]])
        g["sr_values"] = np.array([conv._calculate_audio_params(np.empty((a, b, 0)), np.uint16)[0]
                                   for a, b in g["sr_shapes"]])

        # ---- spatial path ---------------------------------------------------------------
        se = SpatialFLACEncoder(tile_size=256)
        for nm in ["dem", "u8_small", "u16_12bit", "i32_mid", "f32_overflow", "rgb"]:
            r = cases[nm]
            bands, h, w = r.shape
            a = r.reshape(bands, h * w).T
            g[f"spatial_audio_{nm}"] = np.asarray(se._normalize_to_audio(a, r.dtype))
        g["tiles_512_256"] = np.array(se._calculate_tiles(512, 512))
        se2 = SpatialFLACEncoder(tile_size=1024)
        g["tiles_2500x3000_1024"] = np.array(se2._calculate_tiles(2500, 3000))
        T = Affine(0.001, 0.0, -105.5, 0.0, -0.001, 40.5)
        g["bbox_512_256"] = np.array([se._tile_to_bbox(*t, T) for t in se._calculate_tiles(512, 512)])

        # ---- index queries ----------------------------------------------------------------
        frames = []
        off = 0
        for i, t in enumerate(se._calculate_tiles(1000, 1300)):
            bb = se._tile_to_bbox(*t, T)
            size = 1000 + 37 * i if i % 5 else 0
            frames.append(SpatialFrame(i, bb, Window(t[1], t[0], t[3], t[2]), off, size))
            off += size
        idx = SpatialIndex(frames, CRS("EPSG:4326"), T)
        streamer = SpatialFLACStreamer.__new__(SpatialFLACStreamer)
        streamer.spatial_index = idx
        streamer.logger = logging.getLogger("x")
        qs = []
        for _ in range(40):
            x0 = -105.5 + rng.random() * 1.3
            y1 = 40.5 - rng.random() * 1.0
            qs.append((x0, y1 - rng.random() * 0.4, x0 + rng.random() * 0.5, y1))
        qs.append((-105.5, 40.5 - 0.256, -105.5 + 0.256, 40.5))  # exactly one tile (touching edges)
        qs.append((-200.0, 0.0, -190.0, 1.0))  # nothing
        g["q_bboxes"] = np.array(qs)
        g["q_frame_bboxes"] = np.array([f.bbox for f in frames])
        g["q_offsets"] = np.array([f.byte_offset for f in frames])
        g["q_sizes"] = np.array([f.byte_size for f in frames])
        hits = [[f.frame_id for f in idx.query_bbox(q)] for q in qs]
        g["q_hits"] = np.array([json.dumps(h) for h in hits])
        g["q_ranges"] = np.array([json.dumps(streamer.get_byte_ranges_for_bbox(q)) for q in qs])
        g["index_to_dict"] = np.array(json.dumps(idx.to_dict(), separators=(",", ":")))

    np.savez_compressed(OUT / "quant_golden.npz", **g)
    print("wrote", OUT / "quant_golden.npz", len(g), "arrays")


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""bench.py -- encode MPixel/s and byte-range decode tiles/s of the raster<->FLAC tile codec on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c3|c4] [--small]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

One "step" = one create-streaming pass of the hot path over one synthetic raster: per-tile min/max, PCM scaling, tile cut,
FLAC encode of every 4096-sample frame, frame/tile byte offsets.
Workloads (BASELINE.json configs):
  c2  20 000^2 float32 DEM, tile 1024 (configs[1]): the N=1 default, the configuration the metric is quoted on.
  c3  40 000^2 x 3 uint16, tile 1024, bands as channels (configs[2]): the default when N > 1 -- ONE raster whose 1 600 tiles
      are cut into contiguous ranges over the ranks (strong scaling); a rank holds only the raster rows its tiles touch; the
      per-tile sizes are gathered and rank 0 makes the global byte offsets (host-side offset concatenation, no data-path
      collective).
  c4  10 000^2 x 12 int16, `convert --spatial` lossless, tile 256, band groups 8 + 4 (configs[3]).
  c5  (the `decode` block of every line) 10 000 seeded queries -- half tile ids, half 10x10-pixel bboxes -- against the C3
      streaming file held in HBM, decoded in batches of 1 000 tiles (configs[4]); with N ranks the queries are sorted by byte
      offset and cut into N chunks of equal compressed bytes.
`value`   : device-resident raster -> device-resident FLAC bytes + index arrays, CUDA events, max over ranks.
`e2e`     : the same pass through the C-ABI host entry point (frx_encode_tiles_host) with pinned HOST buffers, H2D of the
            raster and D2H of the bytes + index inside the timed region; `floor_ms` = the same bytes copied and nothing else.
`roofline`: dominant kernel (frx_encode_kernel), algorithmic bytes (raster read once + FLAC bytes written once) over its
            CUDA-event duration (recorded by libfrx on the launching stream) vs the measured HBM copy peak.
`parity_check` / `size_vs_oracle`: what the timed loop wrote is decoded by the CPU oracle and compared with the reference
            quantiser, the GPU decode of it with the reference round trip, its size with the oracle's libFLAC level-5
            restatement (first 16 tiles); a mismatch or a ratio above 1.02 ends the run with a non-zero exit code.
`cpu_baseline` / `--impl reference`: the CPU oracle's restatement of the reference tile loop (reference quantiser +
            libFLAC-level-5 search; oracle/flac_oracle.c) on a bounded sample of the same tiles.  pyFLAC/libFLAC themselves
            are not installable on this image.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

WORKLOADS = {
    "c2": dict(gen="dem_f32", H=20000, W=20000, bands=1, dtype="float32", tile=1024, seed=20000, itemsize=4, pcm_mode="reference",
               desc="create-streaming, synthetic 20000x20000 float32 DEM, tile 1024 (BASELINE configs[1])"),
    "c3": dict(gen="rgb_u16", H=40000, W=40000, bands=3, dtype="uint16", tile=1024, seed=40000, itemsize=2, pcm_mode="reference",
               desc="create-streaming --bands all, synthetic 40000x40000x3 uint16, tile 1024 (BASELINE configs[2])"),
    "c4": dict(gen="multispectral_i16", H=10000, W=10000, bands=12, dtype="int16", tile=256, seed=10000, itemsize=2, pcm_mode="lossless",
               desc="convert --spatial (lossless, band groups 8+4), synthetic 10000x10000x12 int16, tile 256 (BASELINE configs[3])"),
}
SMALL = {"c2": (4096, 4096), "c3": (4096, 4096), "c4": (2048, 2048)}
C5_QUERIES, C5_BATCH, C5_SEED = 10000, 1000, 5
FALLBACK_HBM_GBS = 6650.0
PARITY_TILES = 16


def ncu_traffic(workload: str, kernel: str):
    """dram bytes (read + write) per launch of `kernel` from the committed ncu --set full capture of this workload and the
    file it came from; (None, None) when no capture exists for it."""
    for name in ("r2_traffic.json", "r1_traffic.json"):
        try:
            t = json.loads((ROOT / "profiles" / name).read_text())[workload][kernel]
            return int(t["dram_read_bytes"]) + int(t["dram_write_bytes"]), f"profiles/{name} (ncu --set full capture, not measured in this run)"
        except Exception:
            continue
    return None, None


class ClockSampler(threading.Thread):
    """Samples SM clock + throttle reasons of one GPU through NVML while the timed region runs."""

    def __init__(self, index: int, period: float = 0.05):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._halt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.ok = True
        except Exception:  # NVML missing: report nulls, the driver's own sampler still sees the run
            self.ok = False

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
            getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake",
        }
        while not self._halt.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                try:
                    mask = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                except Exception:
                    mask = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                for bit, name in names.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._halt.wait(self.period)

    def finish(self):
        self._halt.set()
        if self.is_alive():
            self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def hbm_peak():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md 6.65 TB/s)"


def workload(name, small):
    wl = dict(WORKLOADS[name])
    wl["name"] = name
    if small:
        wl["H"], wl["W"] = SMALL[name]
        wl["desc"] += f" [REDUCED to {wl['H']}x{wl['W']} by --small: not a benchmark number]"
    return wl


def units_for(wl, lo=None, hi=None, row_base=0):
    """(tiles [(r,c,h,w)], encode units [(r,c,h,w,first_band,n_channels)]) of the workload; units = tile x band group.
    lo/hi: tile range of a shard; row_base: raster row 0 of the shard is row `row_base` of the full image."""
    from flac_raster_b200.spatial_encoder import band_groups, calculate_tiles
    tiles = calculate_tiles(wl["H"], wl["W"], wl["tile"])
    if lo is not None:
        tiles = tiles[lo:hi]
    groups = band_groups(wl["bands"])
    units = [(r - row_base, c, h, w, b0, nb) for (r, c, h, w) in tiles for (b0, nb) in groups]
    return tiles, units


# ----------------------------------------------------------------------------------------------------------
# CPU arm: the oracle's restatement of the reference tile loop on a bounded sample
# ----------------------------------------------------------------------------------------------------------
def run_cpu(wl, strip, units, threads=0):
    """Oracle tile loop over `units` of the host strip; lossless workloads (C4) feed the oracle encoder the identity PCM
    group by group.  -> (MPixel/s counted in pixels of whole tiles, seconds, bytes)."""
    import numpy as np
    import oracle
    t0 = time.perf_counter()
    nbytes = 0
    if wl["pcm_mode"] == "reference":
        by_group = {}
        for u in units:
            by_group.setdefault((u[4], u[5]), []).append(u[:4])
        for (b0, nb), ts in by_group.items():
            streams, _, _ = oracle.ref_encode_tiles(strip, ts, first_band=b0, nch=nb, level=5, sample_rate=44100, threads=threads)
            nbytes += sum(len(s) for s in streams)
    else:
        from concurrent.futures import ThreadPoolExecutor

        def one(u):
            r, c, h, w, b0, nb = u
            pcm = np.ascontiguousarray(strip[b0:b0 + nb, r:r + h, c:c + w].transpose(1, 2, 0).reshape(-1, nb)).astype(np.int32)
            return len(oracle.encode_stream(pcm, 16, 44100, 4096, 5, write_header=False)[0])   # ctypes call: releases the GIL
        with ThreadPoolExecutor(max_workers=max(1, threads or oracle.num_threads())) as ex:
            nbytes = sum(ex.map(one, units))
    dt = time.perf_counter() - t0
    groups = max(1, len({(u[4], u[5]) for u in units}))
    px = sum(u[2] * u[3] for u in units) / groups
    return px / dt / 1e6, dt, nbytes


def host_strip_generator(wl):
    """CPU-side generation of the first rows of the workload (used by --impl reference, which must not need a GPU).  Same
    formulas as flac_raster_b200/synth.py; the noise is numpy's, not torch's (the CPU arm only needs the same statistics)."""
    import numpy as np

    def gen(r0, r1):
        rng = np.random.default_rng(wl["seed"])
        H, W = wl["H"], wl["W"]
        rows = np.arange(r0, r1, dtype=np.float32)
        if wl["gen"] == "dem_f32":
            x = np.linspace(0.0, 20.0 * (W / 512.0), W, dtype=np.float32)[None, :]
            y = (rows * np.float32(20.0 * (H / 512.0) / max(1, H - 1)))[:, None]
            z = 1000.0 + 300.0 * np.sin(0.5 * x) * np.cos(0.3 * y) + 150.0 * np.sin(1.2 * x) * np.sin(1.1 * y)
            z = z + 50.0 * rng.random((r1 - r0, W), dtype=np.float32)
            return (z / 8192.0).astype(np.float32)[None]
        if wl["gen"] == "rgb_u16":
            x = np.linspace(0.0, 2.0 * np.pi * (W / 256.0), W, dtype=np.float32)[None, :]
            y = (rows * np.float32(2.0 * np.pi * (H / 256.0) / max(1, H - 1)))[:, None]
            out = np.empty((wl["bands"], r1 - r0, W), dtype=np.uint16)
            for k in range(wl["bands"]):
                f = np.cos(np.sqrt(x * x + y * y)) if k % 3 == 0 else (np.broadcast_to(np.sin(x), (r1 - r0, W)) if k % 3 == 1
                                                                       else np.broadcast_to(np.sin(y), (r1 - r0, W)))
                v = 2048.0 + 1800.0 * f + 12.0 * rng.standard_normal((r1 - r0, W), dtype=np.float32)
                out[k] = np.clip(v, 0.0, 4095.0).astype(np.uint16)
            return out
        x = np.linspace(0.0, 4.0 * np.pi * (W / 200.0), W, dtype=np.float32)[None, :]
        y = (rows * np.float32(4.0 * np.pi * (H / 200.0) / max(1, H - 1)))[:, None]
        out = np.empty((wl["bands"], r1 - r0, W), dtype=np.int16)
        for i in range(wl["bands"]):
            fi, ph = 0.5 + 0.3 * i, i * np.pi / 3.0
            v = 1000.0 * (1.28 + np.sin(x * fi + ph) * np.cos(y * fi)) + 8.0 * rng.standard_normal((r1 - r0, W), dtype=np.float32)
            out[i] = np.clip(v, -32768.0, 32767.0).astype(np.int16)
        return out
    return gen


def cpu_sample_units(wl, threads):
    """The bounded CPU sample: the first tiles of the first tile row, every band group of them."""
    tiles, units = units_for(wl)
    per_row = max(1, (wl["W"] + wl["tile"] - 1) // wl["tile"])
    groups = len(units) // len(tiles)
    n_tiles = min(len(tiles), per_row, max(8, threads) * (4 if wl["tile"] <= 256 else 1))
    return tiles[:n_tiles], units[:n_tiles * groups]


def main_reference(args):
    """--impl reference: the reference's CPU implementation of the path (oracle port; libFLAC is not installable here), all
    host threads, each step a bounded sample of the workload's tiles.  Rank 0 alone works."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    world = int(os.environ.get("WORLD_SIZE", "1"))
    wl = workload(args.workload or ("c2" if world == 1 else "c3"), args.small)
    # every host core this process may run on, set explicitly: torchrun exports OMP_NUM_THREADS=1 to its workers
    try:
        threads = len(os.sched_getaffinity(0))
    except AttributeError:
        threads = os.cpu_count() or 1
    tiles, units = cpu_sample_units(wl, threads)
    strip = host_strip_generator(wl)(0, max(t[0] + t[2] for t in tiles))
    run_cpu(wl, strip, units, threads)
    dts = []
    for _ in range(args.steps):
        _, dt, _ = run_cpu(wl, strip, units, threads)
        dts.append(dt)
    px = sum(t[2] * t[3] for t in tiles)
    value = px * args.steps / sum(dts) / 1e6
    sample = f"first {len(tiles)} tiles ({tiles[0][3]}x{tiles[0][2]}, {wl['bands']} band(s)) of the workload per step, one tile per thread"
    line = {
        "impl": "reference", "metric": "encode_mpixel_per_s", "value": round(value, 3), "unit": "MPixel/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(1e3 * sum(dts) / args.steps, 3),
        "higher_is_better": True, "scaling": "strong" if (wl["name"] == "c3" and world > 1) else "weak", "vs_baseline": None,
        "dtype": "int64" if wl["dtype"] == "float32" else "int32", "data": "synthetic",
        "config": {"workload": wl["desc"], "tile_size": wl["tile"], "compression_level": 5, "blocksize": 4096,
                   "bands_as_channels": wl["bands"]},
        "cpu_baseline": {"value": round(value, 3), "unit": "MPixel/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": round(value, 3), "unit": "MPixel/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


# ----------------------------------------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------------------------------------
class Env:
    """torch.distributed plumbing: one process per GPU, NCCL for the barrier, the max over ranks and the size gather."""

    def __init__(self):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device; the codec has no CPU fallback (use --impl reference for the CPU arm)")
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            # NCCL prints its version banner on stdout when the communicator comes up: keep stdout for the JSON line only
            sys.stdout.flush()
            saved = os.dup(1)
            os.dup2(2, 1)
            try:
                dist.init_process_group("nccl", device_id=self.dev)
                dist.barrier()
                torch.cuda.synchronize()
            finally:
                sys.stdout.flush()
                try:
                    import ctypes
                    ctypes.CDLL(None).fflush(None)   # the banner sits in C stdio's buffer when stdout is a pipe
                except Exception:
                    pass
                os.dup2(saved, 1)
                os.close(saved)

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def reduce(self, x: float, op="max") -> float:
        if self.world == 1:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX if op == "max" else self.dist.ReduceOp.SUM)
        return float(t.item())

    def gather_sizes(self, sizes, n_total, lo):
        """Per-tile byte sizes of every rank -> the full array on every rank (the one cross-rank step of the path)."""
        torch = self.torch
        full = torch.zeros(n_total, dtype=torch.int64, device=self.dev)
        full[lo:lo + sizes.numel()] = sizes
        if self.world > 1:
            self.dist.all_reduce(full, op=self.dist.ReduceOp.SUM)   # disjoint ranges: a sum is a gather (<= 1 600 int64)
        return full

    def close(self):
        if self.world > 1:
            self.dist.destroy_process_group()


def _np_view(t):
    """numpy view of a CPU tensor (keeps unsigned dtypes)."""
    import numpy as np
    import torch
    if t.dtype == torch.uint16:
        return t.view(torch.int16).numpy().view(np.uint16)
    if t.dtype == torch.uint32:
        return t.view(torch.int32).numpy().view(np.uint32)
    return t.numpy()


def encode_timed(env, codec, raster, units, wl, steps, warmup):
    """W warm-up + K timed device-resident encodes.  -> (EncodeResult, ms/step on this rank, kernel ms, launches, clocks)"""
    torch = env.torch
    arr = codec.prepare_tiles(units)                        # the frx_tile array, built once (static metadata of the workload)
    res = codec.encode_tiles(raster, arr, pcm_mode=wl["pcm_mode"])
    torch.cuda.synchronize()
    assert int(res.status.item()) == 0, "encode overflowed its output buffer"
    for _ in range(warmup):
        codec.encode_tiles(raster, arr, pcm_mode=wl["pcm_mode"], out=res)
    env.barrier()
    sampler = ClockSampler(env.local)
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = codec.launch_count
    kms = []
    ev0.record()
    for _ in range(steps):
        codec.encode_tiles(raster, arr, pcm_mode=wl["pcm_mode"], out=res)
        torch.cuda.current_stream().synchronize()
        kms.append(codec.last_kernel_ms()[0])
    ev1.record()
    env.barrier()
    clocks = sampler.finish()
    return res, ev0.elapsed_time(ev1) / steps, sum(kms) / len(kms), codec.launch_count - l0, clocks, arr


def decode_entries(units, res_np, bps, plane_of=None):
    """DTile dicts for `units` (tile-major, band group minor): outputs packed back to back."""
    toff, tsize, mm, fpt = res_np
    entries, ff, oe = [], 0, 0
    for i, (r0, c0, h, w, b0, nch) in enumerate(units):
        nfr = int(fpt[i])
        entries.append(dict(byte_offset=int(toff[i]), byte_size=int(tsize[i]), out_offset=oe, out_band_stride=h * w, out_row_stride=w,
                            h=h, w=w, channels=nch, bps=bps, dmin=float(mm[i, 0]), dmax=float(mm[i, 1]), first_frame=ff, n_frames=nfr))
        ff += nfr
        oe += nch * h * w
    return entries, oe


def parity_check(env, codec, wl, raster, units, res, n_units):
    """What the timed loop wrote, checked on the first `n_units` units: the CPU oracle decodes the frames (CRCs, frame count)
    and its PCM must equal the reference quantiser's; the GPU decode of them must equal the reference round trip (lossless
    workloads: the raster itself); the bytes are compared with the oracle's libFLAC level-5 restatement of the same PCM."""
    import numpy as np
    import oracle
    from oracle import quant
    from flac_raster_b200.codec import _tdtype, make_dtiles
    from flac_raster_b200.synth import to_numpy
    torch = env.torch
    n = min(n_units, len(units))
    toff = res.tile_offset[:n].cpu().numpy(); tsize = res.tile_size[:n].cpu().numpy(); mm = res.tile_minmax[:n].cpu().numpy()
    end = int(toff[n - 1] + tsize[n - 1])
    data = res.bytes[:end].cpu().numpy()
    r1 = max(u[0] + u[2] for u in units[:n])
    host = to_numpy(raster[:, :r1])
    entries, oe = decode_entries(units[:n], (toff, tsize, mm, res.frames_per_tile[:n]), res.bps)
    out = torch.zeros(oe, dtype=_tdtype(wl["dtype"]), device=env.dev)
    status = codec.decode_tiles(res.bytes, make_dtiles(entries), out, out_dtype=wl["dtype"], frame_size=res.frame_size, pcm_mode=wl["pcm_mode"],
                                n_bytes=int(res.total_bytes.item()))
    torch.cuda.synchronize()
    flat = to_numpy(out)
    pcm_ok = raster_ok = not bool(status.cpu().numpy().any())
    ours = ref = 0
    for i, (r0, c0, h, w, b0, nch) in enumerate(units[:n]):
        block = host[b0:b0 + nch, r0:r0 + h, c0:c0 + w]
        if wl["pcm_mode"] == "reference":
            pcm_ref, bps, dmin, dmax, _ = quant.reference_pcm(block)
            want = quant.reference_decode_raster(pcm_ref, bps, h, w, nch, block.dtype, dmin, dmax).reshape(nch, h, w)
        else:
            pcm_ref, bps = block.transpose(1, 2, 0).reshape(-1, nch).astype(np.int32), 16
            want = block
        pcm, frames, rc = oracle.decode_frames(bytes(data[toff[i]:toff[i] + tsize[i]]), nch, bps, max_samples=h * w + 16)
        pcm_ok = pcm_ok and rc == 0 and len(frames) == (h * w + 4095) // 4096 and all(f["crc8_ok"] and f["crc16_ok"] for f in frames) \
            and pcm.shape == pcm_ref.shape and bool(np.array_equal(pcm, pcm_ref))
        e = entries[i]
        got = flat[e["out_offset"]:e["out_offset"] + nch * h * w].reshape(nch, h, w)
        raster_ok = raster_ok and bool(np.array_equal(got, want, equal_nan=True))
        ours += int(tsize[i])
        ref += len(oracle.encode_stream(pcm_ref, bps, 44100, 4096, 5, write_header=False)[0])
    return {"tiles": n, "pcm_equal": bool(pcm_ok), "raster_equal": bool(raster_ok)}, (ours / ref if ref else None)


def c5_queries(wl, tiles):
    """10 000 queries, half `tile_id ~ U{0..n-1}`, half a 10x10-pixel bbox around a uniform random pixel, mapped
    through the file's transform from_origin(-120, 37, 1e-4, 1e-4)."""
    import numpy as np
    rng = np.random.default_rng(C5_SEED)
    qs = []
    for k in range(C5_QUERIES):
        if k % 2 == 0:
            qs.append({"tile_id": int(rng.integers(0, len(tiles)))})
        else:
            px, py = int(rng.integers(0, max(1, wl["W"] - 10))), int(rng.integers(0, max(1, wl["H"] - 10)))
            x0, y1 = -120.0 + px * 1e-4, 37.0 - py * 1e-4
            qs.append({"bbox": [x0, y1 - 10e-4, x0 + 10e-4, y1]})
    return qs


def c5_decode_leg(env, codec, wl, tiles_all, lo, hi, res, sizes_full, peak, steps):
    """Config C5 on this rank's part of the C3 file (compressed image resident in HBM): index selection on the host (timed
    apart), then batched decodes of 1 000 tiles each.  With N ranks the selected tiles, sorted by byte offset, are cut into N
    chunks of equal compressed bytes; a tile-sharded file lets each rank answer the chunk that lies in its own shard, so the
    chunk boundaries follow the shards (the byte split of SURVEY 8e, restricted to what a rank holds)."""
    import numpy as np
    from flac_raster_b200 import rio, streaming
    from flac_raster_b200.codec import _tdtype, make_dtiles
    torch = env.torch
    offs = np.concatenate([[0], np.cumsum(sizes_full.cpu().numpy())[:-1]])
    tr = rio.from_origin(-120.0, 37.0, 1e-4, 1e-4)
    index = streaming.build_index("EPSG:4326", tr, wl["W"], wl["H"], wl["tile"], tiles_all, offs, sizes_full.cpu().numpy())
    queries = c5_queries(wl, tiles_all)
    t0 = time.perf_counter()
    pos = streaming.select_frames_batch(index, queries)
    select_ms = (time.perf_counter() - t0) * 1e3
    assert (pos >= 0).all()
    pos = np.sort(pos)                                       # byte-offset order (offsets grow with the tile position)
    mine = pos[(pos >= lo) & (pos < hi)] - lo                # the queries whose tile this rank holds
    nb = wl["bands"]
    groups = len(res.frames_per_tile) // (hi - lo)
    assert groups == 1, "C5 runs against a streaming file (one stream per tile)"
    toff = res.tile_offset.cpu().numpy(); tsize = res.tile_size.cpu().numpy(); mm = res.tile_minmax.cpu().numpy()
    fpt = res.frames_per_tile
    ffirst = np.concatenate([[0], np.cumsum(fpt)[:-1]])
    tiles = tiles_all[lo:hi]
    batches = []
    for s in range(0, len(mine), C5_BATCH):
        entries, oe = [], 0
        for p in mine[s:s + C5_BATCH]:
            r0, c0, h, w = tiles[p]
            entries.append(dict(byte_offset=int(toff[p]), byte_size=int(tsize[p]), out_offset=oe, out_band_stride=h * w, out_row_stride=w,
                                h=h, w=w, channels=nb, bps=res.bps, dmin=float(mm[p, 0]), dmax=float(mm[p, 1]), first_frame=int(ffirst[p]),
                                n_frames=int(fpt[p])))
            oe += nb * h * w
        batches.append((make_dtiles(entries), oe, len(entries), sum(e["byte_size"] for e in entries)))
    cap = max([b[1] for b in batches] + [1])
    out = torch.empty(cap, dtype=_tdtype(wl["dtype"]), device=env.dev)
    status = torch.zeros(C5_BATCH, dtype=torch.int32, device=env.dev)
    total_bytes = int(res.total_bytes.item())

    def run_all():
        kms = 0.0
        for dt, oe, n, _ in batches:
            codec.decode_tiles(res.bytes, dt, out, out_dtype=wl["dtype"], frame_size=res.frame_size, status=status[:n], n_bytes=total_bytes)
            torch.cuda.current_stream().synchronize()
            kms += codec.last_kernel_ms()[1]
        return kms
    run_all()
    assert not bool(status.any().item()), "decode reported a bitstream error"
    env.barrier()
    d0, d1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = max(1, min(steps, 3))
    d0.record()
    kms = sum(run_all() for _ in range(reps)) / reps
    d1.record()
    env.barrier()
    ms = env.reduce(d0.elapsed_time(d1) / reps)
    kms = env.reduce(kms)
    n_q = int(env.reduce(float(len(mine)), "sum"))
    comp = sum(b[3] for b in batches)
    raw = sum(b[1] for b in batches) * wl["itemsize"]
    alg = env.reduce(float(comp + raw), "sum")
    ach = alg / (kms * 1e-3) / 1e9 / env.world
    return {"metric": "decode_tiles_per_s", "value": round(n_q / (ms * 1e-3), 1), "unit": "tiles/s", "ms_per_step": round(ms, 3),
            "workload": f"C5: {C5_QUERIES} seeded queries (half tile ids, half 10x10-pixel bboxes) against the C3 streaming file "
                        f"({wl['H']}x{wl['W']}x{nb} {wl['dtype']}, tile {wl['tile']}), image resident in HBM, batches of {C5_BATCH} tiles"
                        + (f", queries sorted by byte offset and answered by the rank that holds the tile ({env.world} ranks)" if env.world > 1 else ""),
            "queries": n_q, "tiles_per_step": n_q, "batches_per_rank": len(batches), "index_select_ms": round(select_ms, 2),
            "mpixel_per_s": round(n_q * wl["tile"] * wl["tile"] / (ms * 1e-3) / 1e6, 1),
            "roofline": {"bound": "hbm", "achieved": round(ach, 2), "peak": peak, "unit": "GB/s", "frac": round(ach / peak, 4),
                         "traffic": None, "kernel": "frx_decode_kernel", "kernel_ms": round(kms, 4), "algorithmic_bytes": int(alg),
                         "note": "B_dec = compressed bytes read + raster bytes written, per GPU"}}


def e2e_encode(env, codec, wl, raster, arr, res, steps, warmup):
    """Host buffers through frx_encode_tiles_host (pinned memory), H2D + D2H timed; and the floor: the same bytes copied
    over the same links, both directions at once, with no kernel in between."""
    from flac_raster_b200.codec import EncodeResult
    torch = env.torch
    n, nf = len(arr), int(res.frame_size.numel())
    total_bytes = int(res.total_bytes.item())
    host_t = torch.empty(raster.shape, dtype=raster.dtype, pin_memory=True)
    host_t.copy_(raster)
    torch.cuda.synchronize()
    host = _np_view(host_t)

    def pinned(shape, dt):
        return torch.empty(shape, dtype=dt, pin_memory=True)
    hb = pinned(int(total_bytes * 1.05) + (1 << 20), torch.uint8)
    hres = EncodeResult(bytes=hb.numpy(), tile_offset=pinned(n, torch.int64).numpy(), tile_size=pinned(n, torch.int64).numpy(),
                        tile_minmax=pinned((n, 2), torch.float64).numpy(), frame_size=pinned(nf, torch.int32).numpy(),
                        frame_offset=pinned(nf, torch.int64).numpy(), total_bytes=pinned(1, torch.int64).numpy(),
                        status=pinned(1, torch.int32).numpy(), frames_per_tile=res.frames_per_tile, bps=res.bps, blocksize=4096)
    for _ in range(max(1, min(warmup, 2))):
        codec.encode_tiles_host(host, arr, pcm_mode=wl["pcm_mode"], out=hres)
    env.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        codec.encode_tiles_host(host, arr, pcm_mode=wl["pcm_mode"], out=hres)
    torch.cuda.synchronize()
    ms = env.reduce((time.perf_counter() - t0) * 1e3 / steps)
    env.barrier()
    assert int(hres.total_bytes[0]) == total_bytes
    # floor: H2D of the raster and D2H of the stream, on two streams, nothing else
    dbuf = res.bytes[:total_bytes]                          # (the stream itself: the pinned buffer keeps valid contents)
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    fl = []
    for _ in range(3):
        env.barrier()
        t0 = time.perf_counter()
        with torch.cuda.stream(s1):
            raster.copy_(host_t, non_blocking=True)
        with torch.cuda.stream(s2):
            hb[:total_bytes].copy_(dbuf, non_blocking=True)
        torch.cuda.synchronize()
        fl.append((time.perf_counter() - t0) * 1e3)
    floor = env.reduce(min(fl))
    raster_bytes = raster.numel() * raster.element_size()
    d2h = total_bytes + n * 32 + nf * 12 + 12
    return ms, floor, raster_bytes, d2h, host_t, hb, hres


def e2e_decode(env, codec, wl, units, res, hb, steps):
    """frx_decode_tiles_host: compressed bytes from pinned host memory in, rasters to pinned host memory out."""
    import numpy as np
    from flac_raster_b200.codec import make_dtiles
    torch = env.torch
    total = int(res.total_bytes.item())
    res_np = (res.tile_offset.cpu().numpy(), res.tile_size.cpu().numpy(), res.tile_minmax.cpu().numpy(), res.frames_per_tile)
    entries, oe = decode_entries(units, res_np, res.bps)
    dt = make_dtiles(entries)
    out_t = torch.empty(oe, dtype={"float32": torch.float32, "uint16": torch.int16, "int16": torch.int16}[wl["dtype"]], pin_memory=True)
    out = out_t.numpy().view(np.dtype(wl["dtype"]))
    fs = res.frame_size.cpu().numpy().astype(np.uint32)
    data = hb.numpy()[:total]
    codec.decode_tiles_host(data, dt, out, frame_size=fs, pcm_mode=wl["pcm_mode"])
    env.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        st = codec.decode_tiles_host(data, dt, out, frame_size=fs, pcm_mode=wl["pcm_mode"])
    ms = env.reduce((time.perf_counter() - t0) * 1e3 / steps)
    assert not st.any()
    return {"metric": "decode_tiles_per_s", "value": round(env.world * len(units) / (ms * 1e-3), 1), "unit": "tiles/s", "ms_per_step": round(ms, 3),
            "h2d_bytes_per_step": total + fs.nbytes, "d2h_bytes_per_step": oe * wl["itemsize"], "api": "frx_decode_tiles_host, pinned host buffers"}


def device_decode_leg(env, codec, wl, units, res, steps, peak):
    """The whole stream the timed loop wrote, decoded device to device in one frx_decode_tiles call (K9-K12)."""
    import numpy as np
    from flac_raster_b200.codec import _tdtype, make_dtiles
    torch = env.torch
    total = int(res.total_bytes.item())
    res_np = (res.tile_offset.cpu().numpy(), res.tile_size.cpu().numpy(), res.tile_minmax.cpu().numpy(), res.frames_per_tile)
    entries, oe = decode_entries(units, res_np, res.bps)
    dt = make_dtiles(entries)
    out = torch.empty(oe, dtype=_tdtype(wl["dtype"]), device=env.dev)
    status = torch.zeros(len(units), dtype=torch.int32, device=env.dev)
    ms, kms = [], []
    for i in range(steps + 1):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        codec.decode_tiles(res.bytes, dt, out, out_dtype=wl["dtype"], frame_size=res.frame_size, pcm_mode=wl["pcm_mode"], status=status, n_bytes=total)
        e1.record()
        torch.cuda.synchronize()
        if i:                                              # first call: warm-up
            ms.append(e0.elapsed_time(e1)); kms.append(codec.last_kernel_ms()[1])
    assert not status.cpu().numpy().any()
    step_ms, kernel_ms = float(np.mean(ms)), float(np.mean(kms))
    alg = total + oe * wl["itemsize"]
    return {"metric": "decode_tiles_per_s", "value": round(len(units) / (step_ms * 1e-3), 1), "unit": "tiles/s", "ms_per_step": round(step_ms, 3),
            "msample_per_s": round(oe / (step_ms * 1e-3) / 1e6, 1), "api": "frx_decode_tiles, device buffers, one call",
            "roofline": {"bound": "hbm", "achieved": round(alg / (kernel_ms * 1e-3) / 1e9, 2), "peak": peak, "unit": "GB/s",
                         "frac": round(alg / (kernel_ms * 1e-3) / 1e9 / peak, 4), "kernel": "frx_decode_kernel", "kernel_ms": round(kernel_ms, 4),
                         "algorithmic_bytes": alg, "note": "B_dec = compressed bytes read + raster bytes written"}}


def ingest_leg(env, codec, wl, raster, steps, peak):
    """the raster as a pixel-interleaved, one-row-per-strip TIFF image in device memory (GDAL's default layout for
    rasters this wide) -> band-planar raster; checked against the raster it came from."""
    torch = env.torch
    bands, H, W = raster.shape
    esz = wl["itemsize"]
    image = torch.empty(8 + H * W * bands * esz, dtype=torch.uint8, device=env.dev)
    signed = {torch.uint16: torch.int16, torch.uint32: torch.int32}.get(raster.dtype, raster.dtype)   # same bits, full op coverage
    image[8:].view(signed).view(H, W, bands).copy_(raster.view(signed).permute(1, 2, 0))
    strips = [(8 + r * W * bands * esz, r, 0, 1, W, W, -1 if bands > 1 else 0) for r in range(H)]
    out = torch.empty_like(raster)
    codec.tiff_unpack(image, strips, wl["dtype"], bands, H, W, bands, out=out)
    torch.cuda.synchronize()
    assert torch.equal(out.view(torch.uint8), raster.view(torch.uint8)), "tiff_unpack differs from the source raster"
    ms = []
    for _ in range(steps):
        codec.tiff_unpack(image, strips, wl["dtype"], bands, H, W, bands, out=out)
        torch.cuda.current_stream().synchronize()
        ms.append(codec.last_kernel_ms()[0])
    k = env.reduce(sum(ms) / len(ms))
    rb = raster.numel() * raster.element_size()
    return {"metric": "tiff_unpack_gb_per_s", "value": round(env.world * 2 * rb / (k * 1e-3) / 1e9, 1), "unit": "GB/s",
            "kernel": "frx_tiff_unpack_kernel", "kernel_ms": round(k, 4), "algorithmic_bytes": 2 * rb,
            "layout": "pixel-interleaved strips, 1 row per strip", "roofline": {"bound": "hbm", "peak": peak, "unit": "GB/s",
                                                                                "frac": round(2 * rb / (k * 1e-3) / 1e9 / peak, 4)}}


def roofline_block(wl, kernel_ms, alg_bytes, peak, peak_src, small):
    ach = alg_bytes / (kernel_ms * 1e-3) / 1e9
    traffic, tsrc = (None, None) if small else ncu_traffic(wl["name"], "frx_encode_kernel")
    return {"bound": "hbm", "achieved": round(ach, 2), "peak": peak, "unit": "GB/s", "frac": round(ach / peak, 4), "traffic": traffic,
            "traffic_source": tsrc, "kernel": "frx_encode_kernel", "kernel_ms": round(kernel_ms, 4), "algorithmic_bytes": int(alg_bytes),
            "peak_source": peak_src}


def side_workload(env, codec, name, small, steps, peak, peak_src, want_c5):
    """An extra configuration measured in the same run at N=1 (device-resident encode, parity on 4 tiles; C3 also feeds C5)."""
    from flac_raster_b200 import synth
    torch = env.torch
    wl = workload(name, small)
    tiles, units = units_for(wl)
    raster = getattr(synth, wl["gen"])(wl["H"], wl["W"], seed=wl["seed"], device=env.dev)
    res, ms, kms, launches, clocks, arr = encode_timed(env, codec, raster, units, wl, max(2, min(steps, 5)), 3)
    total = int(res.total_bytes.item())
    rb = raster.numel() * raster.element_size()
    groups = len(units) // len(tiles)
    par, ratio = parity_check(env, codec, wl, raster, units, res, 4 * groups)
    blk = {"workload": wl["desc"], "metric": "encode_mpixel_per_s", "value": round(wl["H"] * wl["W"] / (ms * 1e-3) / 1e6, 1), "unit": "MPixel/s",
           "ms_per_step": round(ms, 3), "msample_per_s": round(wl["H"] * wl["W"] * wl["bands"] / (ms * 1e-3) / 1e6, 1),
           "compressed_bytes": total, "compression_ratio": round(rb / total, 4), "streams": len(units),
           "roofline": roofline_block(wl, kms, rb + total, peak, peak_src, small), "parity_check": par,
           "size_vs_oracle": None if ratio is None else round(ratio, 5)}
    dec = None
    if want_c5:
        sizes = res.tile_size.clone()
        dec = c5_decode_leg(env, codec, wl, tiles, 0, len(tiles), res, sizes, peak, steps)
    ok = par["pcm_equal"] and par["raster_equal"] and (ratio is None or ratio <= 1.02)
    del raster, res
    torch.cuda.empty_cache()
    return blk, dec, ok


def main_ours(args):
    import numpy as np
    from flac_raster_b200 import synth
    from flac_raster_b200.codec import Codec
    from flac_raster_b200.streaming import shard_range
    env = Env()
    torch = env.torch
    world, rank = env.world, env.rank
    wl = workload(args.workload or ("c2" if world == 1 else "c3"), args.small)
    sharded = wl["name"] == "c3" and world > 1              # one raster, tile ranges over the ranks (strong scaling)
    peak, peak_src = hbm_peak()
    codec = Codec(env.local)
    tiles_all, _ = units_for(wl)
    H, W, bands = wl["H"], wl["W"], wl["bands"]
    if sharded:
        lo, hi = shard_range(len(tiles_all), world, rank)
        r0 = min(t[0] for t in tiles_all[lo:hi]); r1 = max(t[0] + t[2] for t in tiles_all[lo:hi])
        tiles, units = units_for(wl, lo, hi, row_base=r0)
        raster = getattr(synth, wl["gen"])(H, W, seed=wl["seed"], device=env.dev, rows=(r0, r1))
    else:
        lo, hi, r0 = 0, len(tiles_all), 0
        tiles, units = units_for(wl)
        raster = getattr(synth, wl["gen"])(H, W, seed=wl["seed"] + (rank if world > 1 else 0), device=env.dev)
    torch.cuda.synchronize()
    res, ms_local, kms, launches, clocks, arr = encode_timed(env, codec, raster, units, wl, args.steps, args.warmup)
    ms_step = env.reduce(ms_local)
    k_ms = env.reduce(kms)
    total_bytes = int(res.total_bytes.item())
    raster_bytes = raster.numel() * raster.element_size()
    px_step = H * W if sharded or world == 1 else world * H * W
    value = px_step / (ms_step * 1e-3) / 1e6
    alg_local = raster_bytes + total_bytes                   # raster read once + compressed bytes written once
    # ---- the cross-rank step: per-tile sizes -> global byte offsets (sharded) ----
    groups = len(units) // len(tiles)
    tile_sizes = res.tile_size.view(len(tiles), groups).sum(dim=1)
    merged = None
    if sharded:
        t0 = time.perf_counter()
        sizes_full = env.gather_sizes(tile_sizes, len(tiles_all), lo)
        offsets = torch.cumsum(sizes_full, 0) - sizes_full
        torch.cuda.synchronize()
        merged = {"tiles": len(tiles_all), "total_bytes": int(sizes_full.sum().item()), "first_offset_of_rank": int(offsets[lo].item()),
                  "merge_ms": round((time.perf_counter() - t0) * 1e3, 3),
                  "how": "per-tile sizes all-reduced over disjoint ranges (<= 1 600 int64), exclusive prefix sum; bytes stay on their GPU"}
    else:
        sizes_full = tile_sizes
    # ---- parity of what was timed (rank 0) ----
    par, ratio = ({"tiles": 0, "pcm_equal": None, "raster_equal": None}, None)
    if rank == 0:
        par, ratio = parity_check(env, codec, wl, raster, units, res, PARITY_TILES * groups)
    ok = rank != 0 or (par["pcm_equal"] and par["raster_equal"] and (ratio is None or ratio <= 1.02))
    # ---- C5 decode leg ----
    decode = None
    if wl["name"] == "c3":
        decode = c5_decode_leg(env, codec, wl, tiles_all, lo, hi, res, sizes_full, peak, args.steps)
    extra = {}
    # ---- ingest, e2e, CPU baseline on the headline workload ----
    ingest = ingest_leg(env, codec, wl, raster, max(2, min(args.steps, 5)), peak) if raster_bytes < 12e9 else None
    e2e_ms, floor_ms, h2d, d2h, host_t, hb, hres = e2e_encode(env, codec, wl, raster, arr, res, max(2, min(args.steps, 10)), args.warmup)
    e2e = {"value": round(px_step / (e2e_ms * 1e-3) / 1e6, 1), "unit": "MPixel/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
           "ms_per_step": round(e2e_ms, 3), "floor_ms": round(floor_ms, 3), "over_floor": round(e2e_ms / floor_ms, 3),
           "api": "frx_encode_tiles_host, pinned host buffers; floor = the same H2D + D2H bytes copied concurrently, no kernels"}
    decode_own = None
    if wl["name"] == "c2" and world == 1:
        e2e["decode"] = e2e_decode(env, codec, wl, units, res, hb, max(2, min(args.steps, 5)))
        decode_own = device_decode_leg(env, codec, wl, units, res, max(2, min(args.steps, 5)), peak)
    cpu = None
    if world == 1 and not args.no_cpu:
        import oracle
        threads = oracle.num_threads()
        stiles, sunits = cpu_sample_units(wl, threads)
        rr = max(t[0] + t[2] for t in stiles)
        strip = np.ascontiguousarray(_np_view(host_t)[:, :rr, :])
        run_cpu(wl, strip, sunits[:max(1, len(sunits) // 4)])
        _, dt, _ = run_cpu(wl, strip, sunits)
        reps = 1
        while dt < 10.0 and reps < 400:   # aim for ~10 s of CPU work in total
            dt += run_cpu(wl, strip, sunits)[1]
            reps += 1
        px = sum(t[2] * t[3] for t in stiles) * reps
        cpu = {"value": round(px / dt / 1e6, 3), "unit": "MPixel/s", "cores": threads, "kind": "port",
               "sample": f"first {len(stiles)} tiles of the same raster x {reps} passes ({dt:.1f} s), oracle tile loop, one tile per thread"}
    del host_t, hb, hres, raster, res
    torch.cuda.empty_cache()
    # ---- the other configurations, N=1 default run only ----
    if world == 1 and wl["name"] == "c2" and not args.only:
        blk, dec, ok3 = side_workload(env, codec, "c3", args.small, args.steps, peak, peak_src, True)
        extra["c3"] = blk
        decode = dec
        blk4, _, ok4 = side_workload(env, codec, "c4", args.small, args.steps, peak, peak_src, False)
        extra["c4"] = blk4
        ok = ok and ok3 and ok4
    launches_all = int(env.reduce(float(launches), "sum"))
    if rank == 0:
        line = {
            "metric": "encode_mpixel_per_s", "value": round(value, 1), "unit": "MPixel/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms_step, 3), "higher_is_better": True, "scaling": "strong" if sharded else "weak",
            "vs_baseline": None, "dtype": "int64" if res_bps(wl) == 32 else "int32", "data": "synthetic",
            "config": {"workload": wl["desc"], "tile_size": wl["tile"], "compression_level": 5, "blocksize": 4096,
                       "bands_as_channels": bands, "pcm_mode": wl["pcm_mode"], "pcm_bps": res_bps(wl), "tiles_per_gpu": len(tiles),
                       "streams_per_gpu": len(units), "compressed_bytes_rank0": total_bytes,
                       "compression_ratio": round(raster_bytes / total_bytes, 4),
                       "l2": f"inputs larger than L2 ({raster_bytes / 1e6:.0f} MB raster per GPU per step vs 126 MB), no flush needed",
                       "per_rank": ("contiguous tile range shard_range(n_tiles, N, rank) of ONE raster, only its rows resident; sizes gathered, "
                                    "rank 0 prefix-sums the byte offsets; no data-path collective") if sharded else
                                   ("each rank encodes its own raster of this shape (seed + rank); no collective" if world > 1 else "single GPU")},
            "roofline": roofline_block(wl, k_ms, alg_local, peak, peak_src, args.small),
            "parity_check": par, "size_vs_oracle": None if ratio is None else round(ratio, 5),
            "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches_all, "clocks": clocks, "decode": decode, "ingest": ingest,
        }
        if decode_own:
            line["decode_own_stream"] = decode_own
        if merged:
            line["merged_index"] = merged
        line.update(extra)
        print(json.dumps(line))
        if not ok:
            print("bench.py: PARITY OR SIZE GATE FAILED: " + json.dumps({"parity_check": par, "size_vs_oracle": ratio,
                  "c3": extra.get("c3", {}).get("parity_check"), "c4": extra.get("c4", {}).get("parity_check")}), file=sys.stderr)
    codec.close()
    env.close()
    return 0 if ok else 1


def res_bps(wl):
    return 32 if wl["dtype"] in ("float32", "int32", "uint32") else 16


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--workload", choices=sorted(WORKLOADS), default=None, help="default: c2 on one GPU, c3 (tiles sharded) on several")
    ap.add_argument("--small", action="store_true", help="reduced rasters for a quick functional check (not a benchmark)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--only", action="store_true", help="headline workload only (skip the C3/C4/C5 blocks of the default run)")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours" and not args.small:
        args.warmup = 3
    if args.impl == "reference":
        return main_reference(args)
    return main_ours(args)


if __name__ == "__main__":
    sys.exit(main())

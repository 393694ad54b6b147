"""Seeded synthetic rasters of the benchmark configs, generated on the device.

They follow the reference's own sample generators (examples/create_test_data.py:15-25, 53-64,
96-106) scaled to the BASELINE.json shapes.  Generation is chunked by rows so the float
temporaries stay small next to the raster itself.
"""
from __future__ import annotations

import math


def _torch():
    import torch
    return torch


ROWS_PER_CHUNK = 1024   # the noise of rows [1024 k, 1024 k + 1024) comes from its own generator (seed + k): any range of whole
                        # chunks of a raster can be generated on its own, identical to the same rows of the full raster


def _row_chunks(height, rows):
    r0, r1 = (0, height) if rows is None else rows
    assert r0 % ROWS_PER_CHUNK == 0 and 0 <= r0 <= r1 <= height
    return r0, r1, [(c0, min(r1, c0 + ROWS_PER_CHUNK)) for c0 in range(r0, r1, ROWS_PER_CHUNK)]


def _gen(device, seed, chunk_row0):
    return _torch().Generator(device=device).manual_seed(int(seed) * 100003 + chunk_row0 // ROWS_PER_CHUNK)


def dem_f32(height: int, width: int, seed: int = 20000, device="cuda", rows=None):
    """C2: float32 DEM, z = 1000 + 300 sin(.5X) cos(.3Y) + 150 sin(1.2X) sin(1.1Y) + 50 U[0,1), stored as z/8192
    (the reference uses floats as-is, so they must already lie in [-1,1]; converter.py:61-64).  -> (1,H,W), or only the
    rows [rows[0], rows[1]) of it."""
    torch = _torch()
    r0, r1, chunks = _row_chunks(height, rows)
    out = torch.empty((1, r1 - r0, width), dtype=torch.float32, device=device)
    span = 20.0 * (height / 512.0)
    x = torch.linspace(0.0, 20.0 * (width / 512.0), width, device=device, dtype=torch.float32)[None, :]
    for c0, c1 in chunks:
        y = (torch.arange(c0, c1, device=device, dtype=torch.float32) * (span / max(1, height - 1)))[:, None]
        z = 1000.0 + 300.0 * torch.sin(0.5 * x) * torch.cos(0.3 * y) + 150.0 * torch.sin(1.2 * x) * torch.sin(1.1 * y)
        z += 50.0 * torch.rand((c1 - c0, width), generator=_gen(device, seed, c0), device=device, dtype=torch.float32)
        out[0, c0 - r0:c1 - r0] = z / 8192.0
    return out


def rgb_u16(height: int, width: int, seed: int = 40000, device="cuda", bands: int = 3, rows=None):
    """C3: uint16 (12-bit depth), b_k = 2048 + 1800 f_k(X,Y) + N(0,12) clipped to [0,4095],
    f = {cos(sqrt(X^2+Y^2)), sin X, sin Y}, X,Y = linspace(0, 2 pi (size/256)).  -> (3,H,W) torch.uint16, or only the rows
    [rows[0], rows[1]) of it."""
    torch = _torch()
    r0, r1, chunks = _row_chunks(height, rows)
    out = torch.empty((bands, r1 - r0, width), dtype=torch.int16, device=device)
    x = torch.linspace(0.0, 2.0 * math.pi * (width / 256.0), width, device=device, dtype=torch.float32)[None, :]
    span = 2.0 * math.pi * (height / 256.0)
    for c0, c1 in chunks:
        g = _gen(device, seed, c0)
        y = (torch.arange(c0, c1, device=device, dtype=torch.float32) * (span / max(1, height - 1)))[:, None]
        for k in range(bands):
            if k % 3 == 0:
                f = torch.cos(torch.sqrt(x * x + y * y))
            elif k % 3 == 1:
                f = torch.sin(x).expand(c1 - c0, width)
            else:
                f = torch.sin(y).expand(c1 - c0, width)
            v = 2048.0 + 1800.0 * f + 12.0 * torch.randn((c1 - c0, width), generator=g, device=device, dtype=torch.float32)
            out[k, c0 - r0:c1 - r0] = v.clamp_(0.0, 4095.0).to(torch.int16)
    return out.view(torch.uint16)


def multispectral_i16(height: int, width: int, bands: int = 12, seed: int = 10000, device="cuda", rows=None):
    """C4: int16, 1000 (1.28 + sin(X f_i + phi_i) cos(Y f_i)) + N(0,8), f_i = .5 + .3 i, phi_i = i pi/3."""
    torch = _torch()
    r0, r1, chunks = _row_chunks(height, rows)
    out = torch.empty((bands, r1 - r0, width), dtype=torch.int16, device=device)
    x = torch.linspace(0.0, 4.0 * math.pi * (width / 200.0), width, device=device, dtype=torch.float32)[None, :]
    span = 4.0 * math.pi * (height / 200.0)
    for c0, c1 in chunks:
        g = _gen(device, seed, c0)
        y = (torch.arange(c0, c1, device=device, dtype=torch.float32) * (span / max(1, height - 1)))[:, None]
        for i in range(bands):
            fi, ph = 0.5 + 0.3 * i, i * math.pi / 3.0
            v = 1000.0 * (1.28 + torch.sin(x * fi + ph) * torch.cos(y * fi))
            v = v + 8.0 * torch.randn((c1 - c0, width), generator=g, device=device, dtype=torch.float32)
            out[i, c0 - r0:c1 - r0] = v.clamp_(-32768.0, 32767.0).to(torch.int16)
    return out


def to_numpy(t):
    """CUDA/CPU tensor -> numpy array of the same dtype (torch.uint16/uint32 go through a signed view)."""
    import numpy as np
    torch = _torch()
    c = t.detach().cpu()
    if c.dtype == torch.uint16:
        return c.view(torch.int16).numpy().view(np.uint16)
    if c.dtype == torch.uint32:
        return c.view(torch.int32).numpy().view(np.uint32)
    return c.numpy()


def from_numpy(a, device="cuda"):
    """numpy array -> tensor on `device`, keeping unsigned dtypes."""
    import numpy as np
    torch = _torch()
    a = np.ascontiguousarray(a)
    if a.dtype == np.uint16:
        return torch.from_numpy(a.view(np.int16).copy()).to(device).view(torch.uint16)
    if a.dtype == np.uint32:
        return torch.from_numpy(a.view(np.int32).copy()).to(device).view(torch.uint32)
    return torch.from_numpy(a.copy()).to(device)

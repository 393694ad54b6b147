"""``compare_tiffs`` / ``display_comparison_table``: the reference's comparison helpers (compare.py:15-156, re-exported
from the package, __init__.py:4-8) on top of this package's raster adapter.  Same result keys and table layout, so
scripts that read the ``--export`` JSON keep working.  Host-side reporting only; no codec work."""
from __future__ import annotations

import logging
from pathlib import Path
from typing import Dict

import numpy as np
from rich.console import Console
from rich.table import Table

from . import rio

console = Console()
logger = logging.getLogger("flac_raster.compare")


def _load(path: Path):
    with rio.open(path) as src:
        return src.read(), src.meta.get("crs")


def _diff_stats(a: np.ndarray, b: np.ndarray):
    """(equal, max |a-b|, mean |a-b|) with the difference taken in the rasters' own dtype, as the reference does
    (compare.py:56-58): unsigned rasters wrap, which is part of the numbers the reference prints."""
    d = np.abs(a - b)
    return bool(np.array_equal(a, b)), float(d.max()), float(d.mean())


def compare_tiffs(file1_path: Path, file2_path: Path, show_bands: bool = True) -> Dict:
    file1_path, file2_path = Path(file1_path), Path(file2_path)
    logger.info(f"Comparing {file1_path} and {file2_path}")
    (a, crs_a), (b, crs_b) = _load(file1_path), _load(file2_path)
    res = {"file1": file1_path.name, "file2": file2_path.name}
    res.update(shape_match=a.shape == b.shape, dtype_match=a.dtype == b.dtype, crs_match=crs_a == crs_b,
               file1_shape=a.shape, file2_shape=b.shape, file1_dtype=str(a.dtype), file2_dtype=str(b.dtype),
               file1_crs=str(crs_a if crs_a is not None else "None"), file2_crs=str(crs_b if crs_b is not None else "None"))
    if not res["shape_match"]:
        return res
    eq, mx, mean = _diff_stats(a, b)
    res.update(arrays_equal=eq, max_difference=mx, mean_difference=mean, rmse=float(np.sqrt(np.mean((a - b) ** 2))))
    for tag, arr in (("file1", a), ("file2", b)):
        res[f"{tag}_min"], res[f"{tag}_max"] = float(arr.min()), float(arr.max())
    if show_bands and a.ndim == 3:
        res["bands"] = []
        for k, (x, y) in enumerate(zip(a, b), start=1):
            eq, mx, mean = _diff_stats(x, y)
            res["bands"].append({"band": k, "equal": eq, "max_diff": mx, "mean_diff": mean,
                                 "file1_range": [float(x.min()), float(x.max())],
                                 "file2_range": [float(y.min()), float(y.max())]})
    return res


def _table(title, columns, rows):
    t = Table(title=title, show_header=True)
    for name, style in columns:
        t.add_column(name, style=style)
    for r in rows:
        t.add_row(*r)
    console.print(t)


def display_comparison_table(results: Dict):
    yn = lambda v: "YES" if v else "NO"
    f1, f2 = results["file1"], results["file2"]
    _table("TIFF Comparison Results", [("Property", "cyan"), (f1, "green"), (f2, "yellow"), ("Match", "bold")],
           [("Shape", str(results["file1_shape"]), str(results["file2_shape"]), yn(results["shape_match"])),
            ("Data Type", results["file1_dtype"], results["file2_dtype"], yn(results["dtype_match"])),
            ("CRS", results["file1_crs"], results["file2_crs"], yn(results["crs_match"]))])
    if not results.get("shape_match"):
        console.print("[red]Cannot compute detailed statistics - shapes don't match![/red]")
        return
    _table("Statistical Comparison", [("Metric", "cyan"), ("Value", "bold")],
           [("Arrays Equal", yn(results["arrays_equal"])), ("Max Difference", f"{results['max_difference']:.6f}"),
            ("Mean Difference", f"{results['mean_difference']:.6f}"), ("RMSE", f"{results['rmse']:.6f}")])
    _table("Data Ranges", [("File", "cyan"), ("Min", "blue"), ("Max", "red")],
           [(f1, f"{results['file1_min']:.2f}", f"{results['file1_max']:.2f}"),
            (f2, f"{results['file2_min']:.2f}", f"{results['file2_max']:.2f}")])
    if "bands" in results:
        _table("Per-Band Statistics", [("Band", "cyan"), ("Equal", "bold"), ("Max Diff", "yellow"), ("Mean Diff", "yellow"),
                                       (f"{f1} Range", "green"), (f"{f2} Range", "blue")],
               [(str(b["band"]), yn(b["equal"]), f"{b['max_diff']:.3f}", f"{b['mean_diff']:.6f}",
                 f"[{b['file1_range'][0]:.1f}, {b['file1_range'][1]:.1f}]",
                 f"[{b['file2_range'][0]:.1f}, {b['file2_range'][1]:.1f}]") for b in results["bands"]])

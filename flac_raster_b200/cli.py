"""``flac-raster`` command line, same commands and flags as the reference: the four on the hot path
(convert cli.py:32-93, query :293-406, create-streaming :620-804, extract-streaming :875-1039) and the
inspection commands around it (info :96-244, compare :247-290, extract-tile :409-617, spatial-info :807-874).
Additive flags (``--device/--gpus``, ``--bands``, ``--pcm-mode``) default to the reference's behaviour.
Errors follow the reference: message on the console, exit code 1.
"""
from __future__ import annotations

import json
import logging
from pathlib import Path
from typing import Optional

import typer
from rich.console import Console
from rich.logging import RichHandler
from rich.table import Table

app = typer.Typer(name="flac-raster", help="Raster <-> FLAC tile codec on B200 (drop-in for flac-raster's encode/decode path)",
                  add_completion=False)
console = Console()
logging.basicConfig(level=logging.INFO, format="%(message)s", datefmt="[%X]",
                    handlers=[RichHandler(console=console, rich_tracebacks=True)])
logger = logging.getLogger("flac_raster")


def _fail(msg: str):
    console.print(f"[red]{msg}[/red]")
    raise typer.Exit(1)


def _parse_bbox(bbox: str):
    try:
        coords = [float(x.strip()) for x in bbox.split(",")]
        if len(coords) != 4:
            raise ValueError("Bbox must have 4 coordinates")
        return coords
    except (ValueError, IndexError) as e:
        _fail(f"Error: Invalid bbox format. Use 'xmin,ymin,xmax,ymax': {e}")


@app.command()
def convert(
    input_file: Path = typer.Argument(..., help="Input file (TIFF or FLAC)"),
    output_file: Optional[Path] = typer.Option(None, "--output", "-o", help="Output file path"),
    compression_level: int = typer.Option(5, "--compression", "-c", min=0, max=8, help="FLAC compression level (0-8)"),
    force: bool = typer.Option(False, "--force", "-f", help="Overwrite existing output file"),
    verbose: bool = typer.Option(False, "--verbose", "-v", help="Enable verbose logging"),
    spatial_tiling: bool = typer.Option(False, "--spatial", "-s", help="Enable spatial tiling for HTTP range streaming"),
    tile_size: int = typer.Option(512, "--tile-size", help="Size of spatial tiles (default: 512x512)"),
    device: int = typer.Option(0, "--device", help="CUDA device index"),
    pcm_mode: Optional[str] = typer.Option(None, "--pcm-mode", help="reference | lossless | reference_spatial "
                                           "(default: reference; lossless with --spatial)"),
):
    """Convert between TIFF and FLAC formats"""
    if verbose:
        logging.getLogger("flac_raster").setLevel(logging.DEBUG)
    if not input_file.exists():
        _fail(f"Error: Input file does not exist: {input_file}")
    suffix = input_file.suffix.lower()
    if suffix in (".tif", ".tiff"):
        direction, default_suffix = "tiff_to_flac", ".flac"
    elif suffix == ".flac":
        direction, default_suffix = "flac_to_tiff", ".tif"
    else:
        console.print(f"[red]Error: Unsupported file format: {suffix}[/red]")
        console.print("[yellow]Supported formats: .tif, .tiff, .flac[/yellow]")
        raise typer.Exit(1)
    if output_file is None:
        output_file = input_file.with_suffix(default_suffix)
    if output_file.exists() and not force:
        console.print(f"[red]Error: Output file already exists: {output_file}[/red]")
        console.print("[yellow]Use --force to overwrite[/yellow]")
        raise typer.Exit(1)
    try:
        from .converter import RasterFLACConverter
        if direction == "tiff_to_flac":
            if spatial_tiling:
                from .spatial_encoder import SpatialFLACEncoder
                result = SpatialFLACEncoder(tile_size, device, pcm_mode or "lossless").encode_spatial_flac(
                    input_file, output_file, compression_level)
                console.print(f"[green]Spatial index created with {len(result.frames)} tiles[/green]")
            else:
                RasterFLACConverter(device, pcm_mode or "reference").tiff_to_flac(input_file, output_file, compression_level)
            console.print(f"[green]SUCCESS: Converted to FLAC: {output_file}[/green]")
        else:
            RasterFLACConverter(device).flac_to_tiff(input_file, output_file)
            console.print(f"[green]SUCCESS: Converted to TIFF: {output_file}[/green]")
    except typer.Exit:
        raise
    except Exception as e:  # noqa: BLE001 - cli.py:90-93
        logger.exception("Conversion failed")
        _fail(f"Error during conversion: {e}")


@app.command()
def query(
    flac_file: str = typer.Argument(..., help="Spatial FLAC file to query (local path or HTTP URL)"),
    bbox: str = typer.Option(..., "--bbox", "-b", help="Bounding box as 'xmin,ymin,xmax,ymax'"),
    output: Optional[Path] = typer.Option(None, "--output", "-o", help="Output file for extracted data"),
    format: str = typer.Option("ranges", "--format", "-f", help="Output format: 'ranges' or 'data'"),
):
    """Query spatial FLAC file by bounding box for HTTP range streaming"""
    is_url = flac_file.startswith(("http://", "https://"))
    target = flac_file if is_url else Path(flac_file)
    if not is_url and not target.exists():
        _fail(f"Error: FLAC file does not exist: {flac_file}")
    coords = _parse_bbox(bbox)
    try:
        from .spatial_encoder import SpatialFLACStreamer
        streamer = SpatialFLACStreamer(target)
        if format == "ranges":
            ranges = streamer.get_byte_ranges_for_bbox(tuple(coords))
            console.print(f"[green]Found {len(ranges)} byte ranges for bbox {bbox}[/green]")
            table = Table(title=f"HTTP Byte Ranges for {flac_file.split('/')[-1] if is_url else target.name}")
            for col, style in (("Range #", "cyan"), ("Start Byte", "green"), ("End Byte", "yellow"), ("Size (bytes)", "blue"),
                               ("HTTP Range Header", "magenta")):
                table.add_column(col, style=style)
            total = 0
            for i, (start, end) in enumerate(ranges, 1):
                total += end - start + 1
                table.add_row(str(i), f"{start:,}", f"{end:,}", f"{end - start + 1:,}", f"bytes={start}-{end}")
            console.print(table)
            console.print(f"[bold]Total data to fetch: {total:,} bytes[/bold]")
            if output:
                with open(output, "w") as f:
                    json.dump({"bbox": coords, "total_ranges": len(ranges), "total_bytes": total,
                               "ranges": [{"start": s, "end": e, "size": e - s + 1} for s, e in ranges],
                               "http_headers": [f"bytes={s}-{e}" for s, e in ranges]}, f, indent=2)
                console.print(f"[green]Ranges saved to: {output}[/green]")
        elif format == "data":
            data = streamer.stream_bbox_data(tuple(coords))
            console.print(f"[green]Extracted {len(data):,} bytes of FLAC data[/green]")
            if output:
                with open(output, "wb") as f:
                    f.write(data)
                console.print(f"[green]Data saved to: {output}[/green]")
            else:
                console.print("[yellow]Use --output to save extracted data to file[/yellow]")
        else:
            _fail(f"Error: Unknown format '{format}'. Use 'ranges' or 'data'")
    except typer.Exit:
        raise
    except FileNotFoundError as e:
        _fail(f"Error: Spatial index not found. File may not have spatial tiling enabled: {e}")
    except Exception as e:  # noqa: BLE001
        logger.exception("Query failed")
        _fail(f"Error during query: {e}")


def _suffix_of(path_or_url: str):
    if path_or_url.startswith(("http://", "https://")):
        from urllib.parse import urlparse
        return Path(urlparse(path_or_url).path).suffix.lower(), True
    return Path(path_or_url).suffix.lower(), False


@app.command()
def info(file_path: str = typer.Argument(..., help="FLAC or TIFF file to inspect (local path or HTTP URL)")):
    """Display information about a FLAC or TIFF file"""
    suffix, is_url = _suffix_of(file_path)
    if not is_url and not Path(file_path).exists():
        _fail(f"Error: File does not exist: {file_path}")
    mb = lambda n: f"{n / 1024 / 1024:.2f} MB"
    if suffix in (".tif", ".tiff"):
        if is_url:
            _fail("Error: URL-based TIFF inspection not supported yet")
        from . import rio
        with rio.open(file_path) as src:
            console.print("[cyan]TIFF Information:[/cyan]")
            for label, value in (("Dimensions", f"{src.width} x {src.height}"), ("Bands", src.count), ("Data type", src.dtypes[0]),
                                 ("CRS", src.crs), ("Bounds", src.bounds), ("File size", mb(Path(file_path).stat().st_size))):
                console.print(f"  {label}: {value}")
        return
    if suffix != ".flac":
        _fail(f"Error: Unsupported file format: {suffix}")
    console.print("[cyan]FLAC Information:[/cyan]")
    if is_url:   # cli.py:141-165: size from a HEAD request, index through ranged reads only
        try:
            import requests
            from .spatial_encoder import SpatialFLACStreamer
            console.print(f"  Remote file size: {mb(int(requests.head(file_path).headers.get('content-length', 0)))}")
            console.print("  [cyan]Loading metadata (lazy loading)...[/cyan]")
            idx = SpatialFLACStreamer(file_path).spatial_index
            console.print(f"  Spatial tiles: {len(idx.frames)}")
            console.print(f"  Total indexed data: {idx.total_bytes:,} bytes")
            console.print("\n[green]Spatial FLAC Metadata:[/green]")
            console.print("  Spatial format: Yes")
            console.print(f"  Total frames/tiles: {len(idx.frames)}")
            console.print(f"  CRS: {idx.crs}")
            console.print(f"  Transform: {idx.transform}")
        except Exception as e:  # noqa: BLE001
            console.print(f"  [red]Error reading remote FLAC file: {e}[/red]")
        return
    size = mb(Path(file_path).stat().st_size)
    try:   # cli.py:170-180 decodes the whole file to report its shape; here that is one batched GPU decode
        from .pyflac_shim import FileDecoder
        audio, rate = FileDecoder(str(file_path)).process()
        console.print(f"  Sample rate: {rate} Hz")
        console.print(f"  Channels: {audio.shape[1] if audio.ndim > 1 else 1}")
        console.print(f"  Audio shape: {audio.shape}")
        console.print(f"  Data type: {audio.dtype}")
    except Exception as e:  # noqa: BLE001
        console.print(f"  [red]Error reading FLAC file: {e}[/red]")
    console.print(f"  File size: {size}")
    shown = False
    try:
        from . import flacmeta
        with open(file_path, "rb") as f:
            hd = flacmeta.parse_header(f.read(1 << 24))
        if "GEOSPATIAL_CRS" in hd:
            g = lambda k: hd.get("GEOSPATIAL_" + k, "N/A")
            console.print("\n[green]Embedded Geospatial Metadata:[/green]")
            console.print(f"  Original dimensions: {g('WIDTH')} x {g('HEIGHT')}")
            console.print(f"  Original bands: {g('COUNT')}")
            console.print(f"  Original dtype: {g('DTYPE')}")
            console.print(f"  CRS: {g('CRS')}")
            console.print(f"  Data range: [{g('DATA_MIN')}, {g('DATA_MAX')}]")
            console.print(f"  Spatial tiling: {'Yes' if str(hd.get('GEOSPATIAL_SPATIAL_TILING', 'false')).lower() == 'true' else 'No'}")
            raw = hd.get("GEOSPATIAL_BOUNDS")
            if raw:
                b = json.loads(raw)
                b = [b[k] for k in ("left", "bottom", "right", "top")] if isinstance(b, dict) else b
                console.print(f"  Bounds: ({b[0]:.6f}, {b[1]:.6f}, {b[2]:.6f}, {b[3]:.6f})")
            shown = True
    except Exception:  # noqa: BLE001 - fall through to the sidecar like the reference
        shown = False
    if not shown:
        side = Path(file_path).with_suffix(".json")
        if side.exists():
            md = json.loads(side.read_text())
            console.print(f"\n[yellow]Raster Metadata (from {side.name}):[/yellow]")
            console.print(f"  Original dimensions: {md['width']} x {md['height']}")
            console.print(f"  Original bands: {md['count']}")
            console.print(f"  Original dtype: {md['dtype']}")
            console.print(f"  CRS: {md.get('crs', 'None')}")
            if md.get("bounds"):
                b = md["bounds"]
                console.print(f"  Bounds: ({b['left']}, {b['bottom']}, {b['right']}, {b['top']})")
            shown = True
    if not shown:
        console.print("\n[yellow]No embedded or sidecar metadata found[/yellow]")


@app.command()
def compare(
    file1: Path = typer.Argument(..., help="First TIFF file"),
    file2: Path = typer.Argument(..., help="Second TIFF file"),
    show_bands: bool = typer.Option(True, "--show-bands/--no-bands", help="Show per-band statistics"),
    export_json: Optional[Path] = typer.Option(None, "--export", "-e", help="Export comparison results to JSON"),
):
    """Compare two TIFF files and display statistics"""
    for f in (file1, file2):
        if not f.exists():
            _fail(f"Error: File does not exist: {f}")
    for f in (file1, file2):
        if f.suffix.lower() not in (".tif", ".tiff"):
            _fail(f"Error: {f} is not a TIFF file")
    try:
        from .compare import compare_tiffs, display_comparison_table
        results = compare_tiffs(file1, file2, show_bands)
        display_comparison_table(results)
        if export_json:
            export_json.write_text(json.dumps(results, indent=2))
            console.print(f"\n[green]SUCCESS: Comparison results exported to: {export_json}[/green]")
    except Exception as e:  # noqa: BLE001 - cli.py:287-290
        logger.exception("Comparison failed")
        _fail(f"Error during comparison: {e}")


@app.command("extract-tile")
def extract_tile(
    flac_file: str = typer.Argument(..., help="Spatial FLAC file (local path or HTTP URL)"),
    bbox: str = typer.Option(..., "--bbox", "-b", help="Bounding box as 'xmin,ymin,xmax,ymax'"),
    output: Path = typer.Option(..., "--output", "-o", help="Output TIFF file path"),
    device: int = typer.Option(0, "--device", help="CUDA device index"),
):
    """Extract a specific tile from spatial FLAC as TIFF file"""
    is_url = flac_file.startswith(("http://", "https://"))
    if not is_url and not Path(flac_file).exists():
        _fail(f"Error: FLAC file does not exist: {flac_file}")
    box = tuple(_parse_bbox(bbox))
    try:
        from . import rio
        from .spatial_encoder import SpatialFLACStreamer
        console.print(f"[cyan]Loading spatial index from: {flac_file}[/cyan]")
        streamer = SpatialFLACStreamer(flac_file if is_url else Path(flac_file))
        console.print(f"[cyan]Finding tiles for bbox: {bbox}[/cyan]")
        ranges = streamer.get_byte_ranges_for_bbox(box)
        if not ranges:
            _fail(f"Error: No tiles found for bbox {bbox}")
        console.print(f"[green]Found {len(ranges)} tile(s) for extraction[/green]")
        # cli.py:474-480, 547-553: the tile written is the FIRST frame whose bbox contains the query box
        inside = [fr for fr in streamer.spatial_index.frames
                  if box[0] >= fr.bbox[0] and box[1] >= fr.bbox[1] and box[2] <= fr.bbox[2] and box[3] <= fr.bbox[3]]
        if not inside:
            _fail("Error: Could not find exact frame for bbox")
        target = inside[0]
        console.print(f"[green]Using frame {target.frame_id} with bbox {target.bbox}[/green]")
        # the reference decodes the whole file (or trusts stale offsets); here only this tile's byte range is
        # fetched and decoded, in one batched GPU call (SpatialFLACStreamer.decode_frames)
        (_, block), = streamer.decode_frames([target], device)
        tile = block[0]                                   # band 1, like src.read(1, window=...) (cli.py:571)
        w = target.window
        tt = streamer.spatial_index.transform
        tt = rio.Affine(*list(tt)[:6]) * rio.Affine.translation(w.col_off, w.row_off)
        console.print(f"[cyan]Writing TIFF to: {output}[/cyan]")
        with rio.open(output, "w", driver="GTiff", height=tile.shape[0], width=tile.shape[1], count=1, dtype=tile.dtype,
                      crs=streamer.spatial_index.crs, transform=tt) as dst:
            dst.write(tile, 1)
        console.print(f"[green]SUCCESS: Tile extracted to: {output}[/green]")
        with rio.open(output) as src:
            console.print("[cyan]Verification:[/cyan]")
            console.print(f"  Dimensions: {src.width} x {src.height}")
            console.print(f"  CRS: {src.crs}")
            console.print(f"  Bounds: {src.bounds}")
            console.print(f"  Data type: {src.dtypes[0]}")
            data = src.read(1)
            console.print(f"  Data range: [{data.min()}, {data.max()}]")
    except typer.Exit:
        raise
    except Exception as e:  # noqa: BLE001 - cli.py:614-617
        logger.exception("Tile extraction failed")
        _fail(f"Error during tile extraction: {e}")


@app.command("spatial-info")
def spatial_info(flac_file: str = typer.Argument(..., help="Spatial FLAC file to analyze (local path or HTTP URL)")):
    """Show spatial index information for FLAC file"""
    is_url = flac_file.startswith(("http://", "https://"))
    if not is_url and not Path(flac_file).exists():
        _fail(f"Error: FLAC file does not exist: {flac_file}")
    try:
        from .spatial_encoder import SpatialFLACStreamer
        idx = SpatialFLACStreamer(flac_file if is_url else Path(flac_file)).spatial_index
        console.print(f"[green]Spatial FLAC File: {flac_file if is_url else Path(flac_file).name}[/green]")
        console.print(f"CRS: {idx.crs}")
        console.print(f"Transform: {idx.transform}")
        console.print(f"Total frames/tiles: {len(idx.frames)}")
        table = Table(title="Spatial Frames")
        for name, style in (("Frame ID", "cyan"), ("Bbox (xmin, ymin, xmax, ymax)", "green"),
                            ("Window (col, row, width, height)", "yellow"), ("Byte Range", "blue"), ("Size", "magenta")):
            table.add_column(name, style=style)
        shown = idx.frames[:10]                            # cli.py:846: the first ten frames, and their bytes only
        for fr in shown:
            w = fr.window
            table.add_row(str(fr.frame_id), "(" + ", ".join(f"{v:.6f}" for v in fr.bbox) + ")",
                          f"({w.col_off}, {w.row_off}, {w.width}, {w.height})",
                          f"{fr.byte_offset}-{fr.byte_offset + fr.byte_size - 1}", f"{fr.byte_size:,} bytes")
        console.print(table)
        if len(idx.frames) > 10:
            console.print(f"[yellow]... and {len(idx.frames) - 10} more frames[/yellow]")
        console.print(f"[bold]Total indexed data: {sum(fr.byte_size for fr in shown):,} bytes[/bold]")
    except FileNotFoundError as e:
        _fail(f"Error: Spatial index not found. File may not have spatial tiling enabled: {e}")
    except typer.Exit:
        raise
    except Exception as e:  # noqa: BLE001 - cli.py:871-874
        logger.exception("Spatial info failed")
        _fail(f"Error reading spatial info: {e}")


@app.command("create-streaming")
def create_streaming(
    input_file: Path = typer.Argument(..., help="Input TIFF file to convert"),
    output_file: Optional[Path] = typer.Option(None, "--output", "-o", help="Output streaming FLAC file"),
    tile_size: int = typer.Option(1024, "--tile-size", help="Size of streaming tiles (default: 1024x1024)"),
    force: bool = typer.Option(False, "--force", "-f", help="Overwrite existing output file"),
    bands: str = typer.Option("first", "--bands", help="first (reference: band 1 only) | all (bands as channels)"),
    gpus: int = typer.Option(1, "--gpus", help="Shard contiguous tile ranges over this many GPUs"),
    pcm_mode: str = typer.Option("reference", "--pcm-mode", help="reference | lossless"),
):
    """Create Netflix-style streaming FLAC with self-contained tiles"""
    if not input_file.exists():
        _fail(f"Error: Input file does not exist: {input_file}")
    if input_file.suffix.lower() not in (".tif", ".tiff"):
        _fail(f"Error: Input must be a TIFF file, got: {input_file.suffix}")
    if output_file is None:
        output_file = input_file.with_suffix(".flac").with_name(input_file.stem + "_streaming.flac")
    if output_file.exists() and not force:
        console.print(f"[red]Error: Output file already exists: {output_file}[/red]")
        console.print("[yellow]Use --force to overwrite[/yellow]")
        raise typer.Exit(1)
    try:
        from . import streaming
        index = streaming.create_streaming(input_file, output_file, tile_size, bands, pcm_mode, tuple(range(gpus)))
        sizes = [f["byte_size"] for f in index["frames"]]
        console.print("[green]SUCCESS: Netflix-style streaming FLAC created![/green]")
        console.print(f"  File: {output_file}")
        console.print(f"  Size: {output_file.stat().st_size / 1048576:.2f} MB")
        console.print(f"  Tiles: {len(sizes)}")
        console.print(f"  Avg tile size: {sum(sizes) / max(1, len(sizes)) / 1024:.1f} KB")
    except typer.Exit:
        raise
    except Exception as e:  # noqa: BLE001 - cli.py:801-804
        logger.exception("Streaming FLAC creation failed")
        _fail(f"Error creating streaming FLAC: {e}")


@app.command("extract-streaming")
def extract_streaming(
    flac_url: str = typer.Argument(..., help="Streaming FLAC file (local path or HTTP URL)"),
    bbox: Optional[str] = typer.Option(None, "--bbox", "-b", help="Bounding box as 'xmin,ymin,xmax,ymax'"),
    tile_id: Optional[int] = typer.Option(None, "--tile-id", help="Extract specific tile by ID"),
    output: Path = typer.Option(..., "--output", "-o", help="Output TIFF file path"),
    center: bool = typer.Option(False, "--center", help="Extract center tile"),
    last: bool = typer.Option(False, "--last", help="Extract last tile"),
    device: int = typer.Option(0, "--device", help="CUDA device index"),
):
    """Extract tiles from Netflix-style streaming FLAC files"""
    try:
        from . import streaming
        coords = _parse_bbox(bbox) if bbox else None
        if tile_id is None and not last and not center and coords is None:
            _fail("Error: Must specify --tile-id, --bbox, --center, or --last")
        index_size, index = streaming.read_streaming_index(flac_url)
        console.print(f"[green]Loaded metadata with {len(index['frames'])} tiles[/green]")
        frame = streaming.select_frame(index, tile_id, last, center, coords)
        if frame is None:
            _fail(f"Error: Tile ID {tile_id} not found" if tile_id is not None else f"Error: No tiles intersect with bbox {bbox}")
        data = streaming.fetch_tile(flac_url, index_size, frame)
        import numpy as np
        from .converter import decode_stream
        _, md = streaming.tile_metadata(data)
        streaming.write_tile_tiff(decode_stream(np.frombuffer(data, dtype=np.uint8), md, device), md, output)
        console.print(f"[bold green]SUCCESS: Extracted tile to {output}[/bold green]")
        console.print(f"[green]Geographic bounds: {frame['bbox']}[/green]")
        total = sum(f["byte_size"] for f in index["frames"])
        console.print(f"[blue]Bandwidth savings: {(1 - frame['byte_size'] / total) * 100:.1f}% "
                      f"({frame['byte_size'] / 1048576:.1f} MB vs {total / 1048576:.1f} MB)[/blue]")
    except typer.Exit:
        raise
    except Exception as e:  # noqa: BLE001 - cli.py:1036-1039
        logger.exception("Streaming extraction failed")
        _fail(f"Error during streaming extraction: {e}")


if __name__ == "__main__":
    app()

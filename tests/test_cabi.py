"""CPU tests of the drop-in boundary: libfrx.so loads, exports every symbol include/frx.h declares, its host-only
entry points behave, and the compute entry points fail loudly without a CUDA device (no CPU fallback)."""
import ctypes as C
import re
from pathlib import Path

import pytest

from flac_raster_b200 import _lib as L

ROOT = Path(__file__).resolve().parents[1]


def _declared_functions():
    text = (ROOT / "include" / "frx.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(frx_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree():
    decl = _declared_functions()
    assert decl, "no functions parsed from include/frx.h"
    assert sorted(L.EXPORTS) == decl


def test_library_exports_every_declared_symbol():
    lib = C.CDLL(str(L.LIB_PATH))
    for name in _declared_functions():
        assert hasattr(lib, name), f"libfrx.so does not export {name}"


def test_version_and_bps_rule():
    lib = L.lib()
    assert lib.frx_version() == 1
    # pyFLAC: bits per sample = array itemsize * 8: never 24
    assert lib.frx_pcm_bps(L.FRX_U8, L.PCM_REFERENCE) == 16
    assert lib.frx_pcm_bps(L.FRX_I16, L.PCM_REFERENCE) == 16
    assert lib.frx_pcm_bps(L.FRX_U16, L.PCM_REFERENCE) == 16
    assert lib.frx_pcm_bps(L.FRX_I32, L.PCM_REFERENCE) == 32
    assert lib.frx_pcm_bps(L.FRX_F32, L.PCM_REFERENCE) == 32
    assert lib.frx_pcm_bps(L.FRX_U8, L.PCM_REFERENCE_SPATIAL) == 32
    assert lib.frx_pcm_bps(99, 0) == L.E_INVALID


def test_encode_bound_counts_frames():
    lib = L.lib()
    tiles = L.make_tiles([(0, 0, 1024, 1024, 0, 1), (0, 1024, 1024, 544, 0, 1), (1024, 0, 544, 544, 0, 3)])
    p = L.EncodeParams(5, 4096, 44100, L.PCM_REFERENCE)
    nf, nb = C.c_uint64(), C.c_uint64()
    assert lib.frx_encode_bound(tiles, 3, L.FRX_F32, C.byref(p), C.byref(nf), C.byref(nb)) == 0
    assert nf.value == 256 + 136 + 73  # SURVEY section 8 derived sizes for C2's edge tiles
    assert nb.value >= (1024 * 1024 + 1024 * 544 + 3 * 544 * 544) * 4


def test_more_than_8_channels_is_unsupported_like_the_reference():
    """FLAC__MAX_CHANNELS = 8: the reference raises EncoderInitException."""
    lib = L.lib()
    tiles = L.make_tiles([(0, 0, 256, 256, 0, 12)])
    p = L.EncodeParams(5, 4096, 44100, L.PCM_REFERENCE)
    nf, nb = C.c_uint64(), C.c_uint64()
    rc = lib.frx_encode_bound(tiles, 1, L.FRX_I16, C.byref(p), C.byref(nf), C.byref(nb))
    assert rc == L.E_UNSUPPORTED
    assert b"8" in lib.frx_last_error()
    with pytest.raises(L.FrxUnsupported):
        L.check(rc)


def test_bad_params_are_rejected():
    lib = L.lib()
    tiles = L.make_tiles([(0, 0, 16, 16, 0, 1)])
    nf, nb = C.c_uint64(), C.c_uint64()
    for p in [L.EncodeParams(9, 4096, 44100, 0), L.EncodeParams(5, 8, 44100, 0), L.EncodeParams(5, 4096, 0, 0),
              L.EncodeParams(5, 4096, 44100, 7)]:
        assert lib.frx_encode_bound(tiles, 1, L.FRX_I16, C.byref(p), C.byref(nf), C.byref(nb)) < 0


def test_no_cpu_fallback():
    """Without a CUDA device the context cannot be created and says so; with one, this test is a no-op."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA device present")
    lib = L.lib()
    assert lib.frx_device_count() == 0
    h = C.c_void_p()
    rc = lib.frx_ctx_create(0, C.byref(h))
    assert rc == L.E_CUDA and b"no CPU fallback" in lib.frx_last_error()
    from flac_raster_b200.codec import Codec
    with pytest.raises(L.FrxError):
        Codec(0)


def test_product_never_imports_the_oracle():
    """The oracle is test infrastructure: nothing under flac_raster_b200/ may import or load it."""
    for p in (ROOT / "flac_raster_b200").rglob("*"):
        if p.suffix in (".py", ".cu", ".cuh", ".h") and p.is_file():
            txt = p.read_text()
            assert "import oracle" not in txt and "from oracle" not in txt and "libflac_oracle" not in txt, p

"""The exhaustive proof behind the integer form of the 8/16-bit reference quantiser (frx_common.cuh frx_quant16_int):
oracle/quant16_check.c walks all 2.1e9 (range, value) pairs and must report no pair on which the integer quotient and the
reference's float64 chain (converter.py:56-86, astype(int16)) disagree outside the exact-quotient branch."""
import os
import shutil
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


@pytest.mark.skipif(shutil.which("gcc") is None, reason="gcc not available")
def test_integer_quantiser_proof(tmp_path):
    exe = tmp_path / "quant16_check"
    subprocess.run(["gcc", "-O2", "-fopenmp", "-frounding-math", "-ffp-contract=off", str(ROOT / "oracle" / "quant16_check.c"),
                    "-o", str(exe), "-lm"], check=True)
    env = dict(os.environ)
    env.setdefault("OMP_NUM_THREADS", str(os.cpu_count() or 4))
    out = subprocess.run([str(exe)], capture_output=True, text=True, timeout=1500, env=env)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "mismatches 0;" in out.stdout and "mismatching 0;" in out.stdout and "{floor - 1, floor}: 0" in out.stdout, out.stdout

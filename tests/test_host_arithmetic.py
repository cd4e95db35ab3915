"""Exactness of the reciprocal-based 64-by-32-bit remainder the kernels use when several draws share one divisor
(csrc/common.cuh mod_u64_u32_rcp): restated in C with the rounding modes the device intrinsics name, against x % d."""
import shutil
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


@pytest.mark.skipif(shutil.which("gcc") is None, reason="needs gcc")
def test_reciprocal_remainder_is_exact(tmp_path):
    exe = tmp_path / "mod_rcp_check"
    subprocess.run(["gcc", "-O2", "-frounding-math", "-o", str(exe), str(ROOT / "tests" / "cpp" / "mod_rcp_check.c"), "-lm"], check=True)
    out = subprocess.run([str(exe)], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "0 bad" in out.stdout, out.stdout

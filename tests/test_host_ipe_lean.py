"""The IPE inversion of csrc/ipe_lean.cuh, run on the host (its functions are __host__ __device__): bracketed
thresholds + tightening + plain summation must agree with plain summation alone for every uniform - no GPU needed."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(shutil.which("nvcc") is None, reason="needs nvcc to compile the host harness")
def test_lean_inversion_equals_plain_summation(tmp_path):
    exe = str(tmp_path / "ipe_lean_check")
    src = os.path.join(ROOT, "tests", "host", "ipe_lean_check.cu")
    inc = ["-I" + os.path.join(ROOT, "sq-learn_b200", "csrc"), "-I" + os.path.join(ROOT, "include")]
    subprocess.run(["nvcc", "-O2", "-std=c++17", "-x", "cu", src, "-o", exe] + inc, check=True, capture_output=True)
    out = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout[-2000:]
    assert "TOTAL disagreements 0" in out.stdout

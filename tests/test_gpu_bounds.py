"""Memcheck stand-in (compute-sanitizer is closed on the GPU pool, profiles/README_r2.md): tools/sanitize_smoke.py -- every launch
mode at smoke size: inline, wavefront, chunk stealing, work units, banded read-back, gather pack / unpack, small and general
builds, ray lists -- run against libradiance_b200_bounds.so, the same sources compiled with -DRD_BOUNDS_CHECK=1: every queue /
order / record index is checked against the size of what it indexes before the access. No violation may be counted, the
device status flags stay clear, and five renderings of the wavefront frame are byte-identical (no race reaches the result)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_bounds_checked_build_sees_no_violation():
    lib = os.path.join(ROOT, "radiance-ray-tracing_b200", "libradiance_b200_bounds.so")
    if not os.path.exists(lib):
        pytest.skip("libradiance_b200_bounds.so not built (make -C radiance-ray-tracing_b200/csrc bounds)")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "sanitize_smoke.py")], env=dict(os.environ, RADIANCE_B200_LIB=lib),
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:]
    assert "bounds-checked build: 1, violations: 0, wavefront frames identical over 5 runs: True, device status (0, 0)" in r.stdout, r.stdout[-2000:]
    assert "sanitize smoke done" in r.stdout

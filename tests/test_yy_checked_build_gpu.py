"""The -DRC_CHECKED build of librc (csrc/Makefile: variants/librc_checked.so) renders the frames of tests/checked_scenario.py
without hitting one of its bounds traps.  Stands in for compute-sanitizer's memcheck, which is closed on the GPU pool."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_checked_build_renders_without_a_trap():
    lib = os.path.join(ROOT, "raytracing_test_b200", "variants", "librc_checked.so")
    if not os.path.exists(lib):
        pytest.skip("variants/librc_checked.so not built (make -C raytracing_test_b200/csrc)")
    env = dict(os.environ, RC_LIBRC=lib)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "checked_scenario.py")], capture_output=True, text=True, timeout=900,
                       env=env)
    assert r.returncode == 0 and "checked build ok" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]

"""The free-flight tail of the primary kernel (raytracing_test_b200/csrc/free_flight.h: closed-form leap + walked rest +
step counts recovered from the growth of deltaPos) compiled for the host and held to the plain step-by-step loop of
nextIntersectDDA (src/shaders/low_res.frag:446-491) over millions of adversarial rays -- ties, power-of-two and half-way
increments, axis-parallel directions (absDelta = inf, NaN deltaPos), binade crossings.  The same header is what
primary_kernel compiles; tests/test_parity_gpu.py::test_free_flight_changes_nothing is the device-side counterpart."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def fuzz_binary(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("ff") / "free_flight_fuzz")
    subprocess.run(["g++", "-std=gnu++17", "-O2", "-ffp-contract=off", "-fno-fast-math", "-I",
                    os.path.join(ROOT, "raytracing_test_b200", "csrc"), os.path.join(ROOT, "tests", "free_flight_fuzz.cpp"), "-o", out],
                   check=True)
    return out


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_free_flight_equals_the_walk(fuzz_binary, seed):
    r = subprocess.run([fuzz_binary, "1500000", str(seed)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:]
    last = r.stdout.strip().splitlines()[-1].split()
    rays, leaps, bad = int(last[1]), int(last[3]), int(last[5])
    assert bad == 0 and rays == 1500000
    assert leaps > 0.6 * rays  # the closed form really ran

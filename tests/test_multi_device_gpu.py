"""Several GPUs behind the C-ABI from ONE process and one thread (rc_init_multi): the reference's host is a single
process with a single render loop (src/main.cpp:173-232).  The library deals the 16-row bands of every frame to its devices;
the frame the caller reads back must be, byte for byte, the frame one device renders -- hit records too -- through
rc_read_image, the pipelined rc_read_image_async, after incremental world edits, with a device-generated world, and in the
peer-store frame on device 0.  With fewer than two GPUs on the box one device plays both (RC_ALLOW_DUPLICATE_DEVICES): every
code path of the multi-device entry points still runs.  The headless C++ host does the same at BASELINE config 5 against the
oracle's golden hash."""
import hashlib
import json
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _devices(n=2):
    import torch

    have = torch.cuda.device_count()
    if have >= n:
        return list(range(n))
    os.environ["RC_ALLOW_DUPLICATE_DEVICES"] = "1"
    return [0] * n


def test_multi_device_frames_equal_single_device_frames():
    """tests/multi_device_scenario.py in a process of its own (librc's state is process-global)."""
    import sys

    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "multi_device_scenario.py")], capture_output=True, text=True,
                       timeout=900)
    assert r.returncode == 0 and "multi device ok" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


def test_headless_cpp_host_on_several_gpus_renders_the_config5_golden(tmp_path):
    """host/rc_headless --gpus N: the reference's main loop against rc.h, C++, no Python in the process.  The 8K frame of
    BASELINE config 5 it reads back equals the oracle's golden hash, and equals the frame of --gpus 1 byte for byte; a run with
    clicks (rc_update_world on every device) writes the same PPMs on one and on two devices."""
    exe = os.path.join(ROOT, "host", "rc_headless")
    if not os.path.exists(exe):
        pytest.skip("host/rc_headless not built")
    n = 2
    devs = ",".join(str(v) for v in _devices(n))  # "0,1", or "0,0" on a one-GPU box (RC_ALLOW_DUPLICATE_DEVICES)
    env = dict(os.environ)
    with open(os.path.join(ROOT, "tests", "golden", "frames_big.json")) as f:
        want = json.load(f)["cases"]["config5_8k_ao64"]["rgba8_sha256"]
    got = {}
    for gpus in (1, n):
        out = str(tmp_path / f"frame_{gpus}.rgba")
        r = subprocess.run([exe] + (["--devices", devs] if gpus > 1 else []) + ["--frames", "2", "--size", "7680x4320", "--ao", "64", "--still", "--dump-last", out],
                           capture_output=True, text=True, timeout=600, env=env)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        got[gpus] = hashlib.sha256(open(out, "rb").read()).hexdigest()
        os.remove(out)
    assert got[1] == want, "single-device headless frame differs from the oracle's config-5 golden"
    assert got[n] == want, "multi-device headless frame differs from the oracle's config-5 golden"
    ppm = {}
    for gpus in (1, n):
        prefix = str(tmp_path / f"edit{gpus}_")
        r = subprocess.run([exe] + (["--devices", devs] if gpus > 1 else []) + ["--frames", "24", "--size", "640x360", "--ao", "16", "--edit", "--out", prefix],
                           capture_output=True, text=True, timeout=600, env=env)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        assert "click at frame" in r.stdout
        ppm[gpus] = [hashlib.sha256(open(f"{prefix}{k:04d}.ppm", "rb").read()).hexdigest() for k in range(24)]
    assert ppm[1] == ppm[n]
    assert len(set(ppm[1])) > 12  # the camera turns: the frames really differ

"""Generates tests/golden/frames_big.json + tests/golden/samples_big.npz: the ORACLE's output at the
full-size BASELINE configs 4 and 5 (too slow for the test suites: ~10-30 s of CPU each), so that the
GPU box can hold the CUDA path to them at full size by hash + sampled records, and bench.py has the
reference algorithm's byte counts (roofline) for exactly those frames.

  config4_4k_ao64_world2048 : 3840x2160, shadows + AO 64, world = genWorld's formula over 2048x2048
                              columns in a depth-7 tree (built by libvoxelhost, the host stand-in that
                              tests/test_host_cpu.py holds to the reference's own code)
  config5_8k_ao64           : 7680x4320, shadows + AO 64, default world

Run in the build container:  python tools/make_golden_big.py
"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import bindings as ob  # noqa: E402
from raytracing_test_b200 import voxel_host as vh  # noqa: E402

CASES = [
    # name, w, h, world kwargs, depth
    ("config4_4k_ao64_world2048", 3840, 2160, dict(width=2048, length=2048, max_depth=7), 7),
    ("config5_8k_ao64", 7680, 4320, dict(), 6),
]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    pos, d, sun = vh.default_camera()
    out = {"cases": {}}
    samples = {}
    rng = np.random.default_rng(4096)
    tab = np.zeros((64, 3), dtype=np.float32)
    ob.oracle_lib().oracle_hemisphere(64, tab.ctypes.data)
    ob.oracle_lib().oracle_set_hemisphere(64, tab.ctypes.data)
    for name, w, h, wkw, depth in CASES:
        host = vh.World.default(**wkw)
        nodes, arrays = host.pools()
        world = ob.make_world(nodes, arrays, depth)
        look, _, _ = host.pick(pos, d, 30)
        look = tuple(int(v) for v in look)
        r = ob.render(world, ob.make_frame(w, h, pos, d, sun, shadows=True, ao_samples=64, looking_at=look))
        c = r["counters"].as_dict()
        idx = np.sort(rng.choice(w * h, size=1024, replace=False))
        samples[name + "_idx"] = idx.astype(np.int64)
        samples[name + "_hits"] = r["hits"][idx]
        samples[name + "_rgba8"] = r["rgba8"].reshape(-1, 4)[idx]
        out["cases"][name] = {"width": w, "height": h, "world": wkw, "max_depth": depth, "looking_at": list(look),
                              "kwargs": dict(shadows=True, ao_samples=64), "hits_sha256": sha(r["hits"]),
                              "rgba8_sha256": sha(r["rgba8"]), "counters": c}
        print(name, c["rays"], c["algorithmic_bytes"], out["cases"][name]["hits_sha256"][:12], flush=True)
        del r, world, nodes, arrays
        host.close()
    with open(os.path.join(ROOT, "tests", "golden", "frames_big.json"), "w") as f:
        json.dump(out, f, indent=1)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "samples_big.npz"), **samples)


if __name__ == "__main__":
    main()

"""Generates tests/golden/frames.json + tests/golden/samples.npz.

The reference holds no golden frames: these vectors are the ORACLE's output
(oracle/oracle_render.cpp) over the world built by the reference's own unmodified CPU code
(oracle/_ref), at the BASELINE configs 1-3 and a few small cameras.  They pin the oracle
against regressions and let the GPU box check the CUDA path without /root/reference.

Run in the build container:  python tools/make_golden.py
"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import camera_dir  # noqa: E402
from oracle import bindings as ob  # noqa: E402

CASES = [
    # name, w, h, origin, (yaw,pitch) or None = default dir, kwargs
    ("config1_720p_primary", 1280, 720, None, None, dict(shadows=False)),
    ("config2_1080p_shadow", 1920, 1080, None, None, dict(shadows=True)),
    ("config3_1080p_ao64", 1920, 1080, None, None, dict(shadows=True, ao_samples=64)),
    ("small_default_ao16", 320, 180, None, None, dict(shadows=True, ao_samples=16)),
    ("small_mirror", 320, 180, (8.5, 101.5, 4.5), (15.0, 5.0), dict(shadows=True, ao_samples=20, delta_time=1.25)),
    ("small_refractor", 320, 180, (21.5, 12.5, 210.5), (180.0, 10.0), dict(shadows=True, ao_samples=20, delta_time=1.25)),
    ("small_lake", 320, 180, (150.5, 30.5, 140.5), (45.0, 45.0), dict(shadows=True, ao_samples=20, delta_time=1.25)),
    ("small_underwater", 320, 180, (162.5, 19.5, 128.5), (30.0, 30.0), dict(shadows=True, ao_samples=20, delta_time=0.5)),
    ("small_negative_origin", 320, 180, (-20.5, 45.5, -20.5), (45.0, 10.0), dict(shadows=True, ao_samples=20)),
    ("small_down", 320, 180, (100.5, 60.25, 100.5), (0.0, 60.0), dict(shadows=True, ao_samples=20, mouse_pos=(160.0, 90.0))),
]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    rw = ob.RefWorld()
    world = ob.make_world(rw.nodes, rw.arrays, rw.max_depth)
    pos0, dir0, sun = rw.defaults()
    out = {"world": {"node_bytes_used": int(rw.node_bytes_used), "array_bytes_used": int(rw.array_bytes_used),
                     "nodes_sha256": sha(rw.nodes[: rw.node_bytes_used // 8]),
                     "arrays_sha256": sha(rw.arrays[: rw.array_bytes_used // 4])},
           "cases": {}}
    samples = {}
    rng = np.random.default_rng(2024)
    for name, w, h, origin, yp, kw in CASES:
        pos = pos0 if origin is None else np.array(origin, dtype=np.float32)
        d = dir0 if yp is None else camera_dir(*yp)
        ao = kw.get("ao_samples", 0)
        if ao:
            tab = np.zeros((ao, 3), dtype=np.float32)
            ob.oracle_lib().oracle_hemisphere(ao, tab.ctypes.data)
            ob.oracle_lib().oracle_set_hemisphere(ao, tab.ctypes.data)
        r = ob.render(world, ob.make_frame(w, h, pos, d, sun, **kw))
        c = r["counters"].as_dict()
        idx = np.sort(rng.choice(w * h, size=256, replace=False))
        samples[name + "_idx"] = idx.astype(np.int64)
        samples[name + "_hits"] = r["hits"][idx]
        samples[name + "_rgba8"] = r["rgba8"].reshape(-1, 4)[idx]
        out["cases"][name] = {
            "width": w, "height": h, "origin": [float(v) for v in pos], "camera_dir_bits": [int(v) for v in d.view(np.uint32)],
            "yaw_pitch": yp, "kwargs": kw, "hits_sha256": sha(r["hits"]), "rgba8_sha256": sha(r["rgba8"]), "counters": c,
        }
        print(name, c["rays"], c["algorithmic_bytes"], out["cases"][name]["hits_sha256"][:12])
    with open(os.path.join(ROOT, "tests", "golden", "frames.json"), "w") as f:
        json.dump(out, f, indent=1)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "samples.npz"), **samples)


if __name__ == "__main__":
    main()

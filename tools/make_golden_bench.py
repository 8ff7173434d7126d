"""Generates tests/golden/frames_bench.json: the ORACLE's output for BASELINE configs 1-3 rendered with exactly the frame
parameters bench.py times (default world from the host stand-in, default camera, lookingAtBlock = the host's picking ray,
src/main.cpp:81) -- tests/golden/frames.json holds the same configs WITHOUT a looked-at voxel, and the highlight
(low_res.frag:347-350) changes a handful of pixels.  bench.py hashes the frames it has just timed against these
(`parity_check`), outside every timed region; configs 4 and 5 are in frames_big.json already (same parameters).

Run in the build container:  python tools/make_golden_bench.py
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
    ("config1_720p_primary", 1280, 720, dict(shadows=False)),
    ("config2_1080p_shadow", 1920, 1080, dict(shadows=True)),
    ("config3_1080p_ao64", 1920, 1080, dict(shadows=True, ao_samples=64)),
]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    host = vh.World.default()
    nodes, arrays = host.pools()
    world = ob.make_world(nodes, arrays, 6)
    pos, d, sun = vh.default_camera()
    look = tuple(int(v) for v in host.pick(pos, d, 30)[0])
    tab = np.zeros((64, 3), dtype=np.float32)
    ob.oracle_lib().oracle_hemisphere(64, tab.ctypes.data)
    ob.oracle_lib().oracle_set_hemisphere(64, tab.ctypes.data)
    out = {"cases": {}}
    for name, w, h, kw in CASES:
        r = ob.render(world, ob.make_frame(w, h, pos, d, sun, looking_at=look, **kw))
        c = r["counters"].as_dict()
        out["cases"][name] = {"width": w, "height": h, "looking_at": list(look), "kwargs": kw, "hits_sha256": sha(r["hits"]),
                              "rgba8_sha256": sha(r["rgba8"]), "counters": c,
                              "highlight_pixels": int((r["hits"]["kind"] == ob.KIND_HIGHLIGHT).sum())}
        print(name, c["rays"], out["cases"][name]["rgba8_sha256"][:12], out["cases"][name]["highlight_pixels"], flush=True)
    with open(os.path.join(ROOT, "tests", "golden", "frames_bench.json"), "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()

"""World generation: the host stand-in's putBlock loop (one core, what the reference does:
src/world_gen.cpp:13-42) + flatten + upload, against rc_generate_world on the device.

    python tools/worldgen_time.py [out.json]

Prints / writes one record per footprint."""
import json
import sys
import time

import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from raytracing_test_b200 import rc, voxel_host as vh  # noqa: E402


def main():
    caster = rc.RayCaster(0)
    out = []
    for (w, l, d) in [(200, 200, 6), (1024, 1024, 6), (2048, 2048, 7), (4096, 4096, 7)]:
        rec = {"width": w, "length": l, "max_depth": d}
        if w <= 2048:
            t0 = time.perf_counter()
            host = vh.World.default(w, l, d)
            t1 = time.perf_counter()
            info = caster.upload_world(host.describe())
            t2 = time.perf_counter()
            host.close()
            rec.update(host_genworld_ms=(t1 - t0) * 1e3, host_flatten_upload_ms=(t2 - t1) * 1e3, flatten_ms=info.flatten_ms,
                       upload_ms=info.upload_ms, packed_nodes=info.packed_nodes, packed_voxels=info.packed_voxels)
        best = None
        for _ in range(5):
            gi = caster.generate_world(w, l, d)
            if best is None or gi.total_ms < best.total_ms:
                best = gi
        rec.update(device=best.as_dict())
        if "host_genworld_ms" in rec:
            rec["speedup_vs_host_path"] = (rec["host_genworld_ms"] + rec["host_flatten_upload_ms"]) / best.total_ms
        print(json.dumps(rec), flush=True)
        out.append(rec)
    caster.close()
    if len(sys.argv) > 1:
        with open(sys.argv[1], "w") as f:
            json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()

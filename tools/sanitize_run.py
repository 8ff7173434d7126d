"""Developer probe: every device entry point of librc.so once, at small sizes (a quick whole-ABI smoke; written for
compute-sanitizer, which is closed on this GPU pool -- where it is open: compute-sanitizer --tool racecheck python
tools/sanitize_run.py).
Covers: upload (flatten + H2D), primary + shadow + AO kernels (hits, RGBA8, RGBA32F), sharded bands, the frame-target
peer-store path on one device, async read-back, device pick, incremental update (append + patch), device world
generation (radix sort, scans, linking), lookup, blob save / load."""
import os
import sys
import tempfile

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from raytracing_test_b200 import rc, voxel_host as vh

pos, d, sun = vh.default_camera()
host = vh.World.default()
with rc.RayCaster(0) as c:
    c.upload_world(host.describe())
    for (w, h, ao) in [(320, 180, 16), (257, 129, 64), (33, 17, 8)]:
        p = rc.frame_params(w, h, pos, d, sun, shadows=True, ao_samples=ao, hits=True, rgba32f=True, mouse_pos=(w / 2, h / 2),
                            delta_time=0.5)
        c.render_frame(p)
        img, hits = c.read_image(), c.read_hits()
        f32 = c.read_image(fmt=rc.RC_RGBA32F)
        print(f"{w}x{h} ao {ao}: {int((hits['kind'] == 1).sum())} hit pixels, checksum {int(img.astype(np.uint64).sum())}")
    # sharded bands + async read-back
    c.set_shard(1, 3)
    p = rc.frame_params(320, 180, pos, d, sun, shadows=True, ao_samples=16)
    buf = np.zeros((180, 320, 4), dtype=np.uint8)
    c.render_frame(p)
    t = c.read_image_async(buf)
    c.wait_image(t)
    c.set_shard(0, 1)
    # frame target on the same device (the peer-store path of store_pixel)
    c.frame_target_create(320, 180, 2)
    c.render_frame(p)
    c.sync()
    c.frame_target_close()
    # pick + incremental update
    print("pick", c.pick(pos, d, 30))
    c.update_world(host.describe(), None, None)  # first call = full flatten with tracking (the baseline to diff against)
    for k in range(6):
        host.put_block(40 + k, 60 + k, 40, color=123456789, flags=0, metadata=0.0)
    info = c.update_world(host.describe(), None, None)
    print("update", info.as_dict() if hasattr(info, "as_dict") else info)
    c.render_frame(p)
    c.sync()
    pts = np.random.default_rng(0).integers(0, 200, size=(5000, 3)).astype(np.int32)
    words, _ = c.lookup_device(pts)
    print("lookup solid", int((words != 0).sum()))
    # device world generation, blob round trip
    gi = c.generate_world(96, 64, 6)
    print("generated", gi.as_dict())
    c.render_frame(p)
    c.sync()
    with tempfile.TemporaryDirectory() as td:
        path = os.path.join(td, "w.rcw")
        c.save_world(path)
        c.load_world(path)
        c.render_frame(p)
        c.write_ppm(os.path.join(td, "f.ppm"))
    gi = c.generate_world(300, 300, 7)
    c.render_frame(rc.frame_params(200, 120, pos, d, sun, shadows=True, ao_samples=20))
    c.sync()
    print("done")
host.close()

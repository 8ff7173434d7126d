"""Runs in a process of its own with RC_LIBRC = the -DRC_CHECKED build of librc (bounds traps on the AO window index, the
seed-list cursor, the pixel and peer-store addresses; csrc/Makefile): frames of every kind the traps guard -- AO through the
window kernels in bit and byte mode, every schedule, ragged sizes, sharded bands storing into a frame target -- must come out
without a trap (a trap kills the context: every later call fails) and equal to the golden / to the unsharded frame."""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from raytracing_test_b200 import rc, voxel_host as vh  # noqa: E402


def main():
    assert "checked" in os.path.basename(rc.LIB_PATH), rc.LIB_PATH
    host = vh.World.default()
    pos, d, sun = vh.default_camera()
    look = tuple(int(v) for v in host.pick(pos, d, 30)[0])
    with open(os.path.join(ROOT, "tests", "golden", "frames_bench.json")) as f:
        want = json.load(f)["cases"]["config3_1080p_ao64"]["rgba8_sha256"]
    frames = 0
    with rc.RayCaster(0) as c:
        c.upload_world(host.describe())
        p = rc.frame_params(1920, 1080, pos, d, sun, shadows=True, ao_samples=64, looking_at=look)
        for sched in range(5):
            c.set_option("schedule", sched)
            c.render_frame(p)
            img = c.read_image()
            assert hashlib.sha256(img.tobytes()).hexdigest() == want, f"schedule {sched}"
            frames += 1
        c.set_option("schedule", "auto")
        # cameras with many distinct windows per warp (bit mode), liquid, ragged sizes, other sample counts
        for (w, h, ao, o, yaw_pitch) in ((1001, 563, 20, (150.5, 30.5, 140.5), (45.0, 45.0)), (333, 777, 256, (100.5, 60.25, 100.5), (0.0, 60.0)),
                                         (640, 360, 7, (8.5, 101.5, 4.5), (15.0, 5.0)), (1920, 1080, 64, (60.2, 25.9, 140.8), (120.0, 35.0))):
            sys.path.insert(0, os.path.join(ROOT, "tests"))
            from conftest import camera_dir

            q = rc.frame_params(w, h, o, camera_dir(*yaw_pitch), sun, shadows=True, ao_samples=ao, hits=True)
            c.render_frame(q)
            full, hits = c.read_image().copy(), c.read_hits().copy()
            frames += 1
            # the same frame as 3 ranks' bands, all storing into one frame target as well
            import torch

            from raytracing_test_b200 import frame_gather as fg

            c.frame_target_create(w, h, 1)
            try:
                tgt = fg.device_bytes(c.frame_target_ptr(0), h * w * 4, torch.device("cuda:0")).view(h, w, 4)
                got = np.zeros_like(full)
                for r in range(3):
                    c.set_shard(r, 3)
                    c.render_frame(q)
                    c.read_image(got)
                    frames += 1
                c.sync()
                assert np.array_equal(got, full) and np.array_equal(tgt.cpu().numpy(), full), (w, h, ao)
            finally:
                c.set_shard(0, 1)
                c.frame_target_close()
        c.render_frame(p)
        st = c.stats()
    print(f"checked build ok: {frames} frames, no trap (last frame: {st.hit_pixels} hit pixels)")


if __name__ == "__main__":
    main()

"""Developer probe: primary kernel time vs max_steps (slope = cost per step, intercept = everything else)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from raytracing_test_b200 import rc, voxel_host as vh

host = vh.World.default()
pos, d, sun = vh.default_camera()
with rc.RayCaster(0) as c:
    c.upload_world(host.describe())
    for ms in (0, 1, 25, 50, 100, 150, 200, 300, 600):
        p = rc.frame_params(1920, 1080, pos, d, sun, shadows=False, ao_samples=0, max_steps=ms)
        best = 1e9
        for _ in range(8):
            c.render_frame(p)
            st = c.stats()
            best = min(best, st.primary_ms)
        p2 = rc.frame_params(1920, 1080, pos, d, sun, shadows=False, ao_samples=0, max_steps=ms, hits=True)
        c.render_frame(p2)
        st = c.stats()
        print(f"max_steps {ms:4d}: primary {best:.3f} ms, steps {st.primary_steps / 1e6:8.1f} M, hit pixels {st.hit_pixels}")

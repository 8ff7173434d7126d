"""Developer probe (not the bench): times rc_render_frame for the BASELINE configs on cuda:0."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import bindings as ob  # world only (dev tool)
from raytracing_test_b200 import rc

rw = ob.RefWorld()
pos, d, sun = rw.defaults()
nodes, arrays = rw.block_ptrs()
desc = rc.make_world_desc(nodes, arrays, rw.node_bytes_used, rw.array_bytes_used)
c = rc.RayCaster(0)
info = c.upload_world(desc)
print("world", info.as_dict())
for (w, h, sh, ao) in [(1280, 720, False, 0), (1920, 1080, True, 0), (1920, 1080, True, 64), (3840, 2160, True, 64)]:
    p = rc.frame_params(w, h, pos, d, sun, shadows=sh, ao_samples=ao)
    for _ in range(3):
        c.render_frame(p)
    c.sync()
    best = None
    for _ in range(10):
        c.render_frame(p)
        st = c.stats()
        if best is None or st.total_ms < best.total_ms:
            best = st
    rays = best.primary_rays + best.shadow_rays + best.ao_rays
    print(f"{w}x{h} shadows={sh} ao={ao}: primary {best.primary_ms:.3f} ms, ao {best.ao_ms:.3f} ms, total {best.total_ms:.3f} ms; "
          f"{rays/1e6:.1f} Mrays -> {rays/best.total_ms/1e3:.1f} Mrays/s; primary-only {best.primary_rays/best.primary_ms/1e3:.1f} Mrays/s")
    t = time.time()
    for _ in range(10):
        c.render_frame(p)
        img = c.read_image()
    print(f"   e2e (render+read_image pageable): {(time.time()-t)/10*1e3:.3f} ms/frame")
c.close()

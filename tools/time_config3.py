"""Developer probe: kernel times of config 3 (1080p, shadows, AO 64) for the librc named by RC_LIBRC (default: the built one)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from raytracing_test_b200 import rc, voxel_host as vh

host = vh.World.default()
pos, d, sun = vh.default_camera()
c = rc.RayCaster(0)
c.upload_world(host.describe())
w, h = (int(v) for v in (sys.argv[1] if len(sys.argv) > 1 else "1920x1080").split("x"))
ao_steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5  # 32 = the stress row of SURVEY 8d (config 3)
p = rc.frame_params(w, h, pos, d, sun, shadows=True, ao_samples=64, ao_steps=ao_steps, kernel_times=True)
for _ in range(5):
    c.render_frame(p)
c.sync()
pr, ao = [], []
for _ in range(30):
    c.render_frame(p)
    st = c.stats()
    pr.append(st.primary_ms)
    ao.append(st.ao_ms)
pr.sort(), ao.sort()
st = c.stats()
rays = st.primary_rays + st.shadow_rays + st.ao_rays
print(f"{w}x{h} ao_steps {ao_steps}: {rays / 1e6:.1f} Mrays, {rays / ((pr[15] + ao[15]) * 1e-3) / 1e9:.1f} Grays/s (median kernel times)")
print(f"{os.environ.get('RC_LIBRC', 'librc.so')}: primary min {pr[0]:.4f} med {pr[15]:.4f}  ao min {ao[0]:.4f} med {ao[15]:.4f} ms")
c.close()

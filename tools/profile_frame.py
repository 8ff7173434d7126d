"""Developer probe for ncu: renders a few frames of one BASELINE config on cuda:0 and exits.
python tools/profile_frame.py [width height shadows ao_samples frames max_steps world_columns]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from raytracing_test_b200 import rc, voxel_host as vh

a = sys.argv[1:]
w, h = (int(a[0]), int(a[1])) if len(a) >= 2 else (1920, 1080)
sh = bool(int(a[2])) if len(a) >= 3 else True
ao = int(a[3]) if len(a) >= 4 else 64
frames = int(a[4]) if len(a) >= 5 else 4
max_steps = int(a[5]) if len(a) >= 6 else 300
world = int(a[6]) if len(a) >= 7 else 200  # columns per side; 2048 = config 4's world (depth 7), generated on the device
pos, d, sun = vh.default_camera()
with rc.RayCaster(0) as c:
    if world == 200:
        c.upload_world(vh.World.default().describe())
    else:
        c.generate_world(world, world, 7)
    if "RC_SCHEDULE" in os.environ:  # A/B captures: rc_debug_set_option
        c.set_option("schedule", int(os.environ["RC_SCHEDULE"]))
    if "RC_AO_OPTS" in os.environ:
        c.set_option("ao_opts", int(os.environ["RC_AO_OPTS"]))
    p = rc.frame_params(w, h, pos, d, sun, shadows=sh, ao_samples=ao, max_steps=max_steps,
                        kernel_times=bool(int(os.environ.get("RC_KERNEL_TIMES", "0"))))
    for _ in range(frames):
        c.render_frame(p)
        c.sync()
    st = c.stats()
    print(f"{w}x{h} shadows={sh} ao={ao} schedule {st.schedule}: total {st.total_ms:.3f} ms (primary {st.primary_ms:.3f} ms ao {st.ao_ms:.3f} ms)")

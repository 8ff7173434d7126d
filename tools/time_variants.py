"""Developer probe: frame time of a workload under every schedule / AO option (rc_debug_set_option), for the librc named by
RC_LIBRC (default: the built one).  L2 is flushed between frames (a 256 MB cudaMemset through torch is avoided: plain
rc frames of another resolution would not flush, so the flush is a device buffer written by torch when available).
    python tools/time_variants.py [WxH] [ao_samples] [frames]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from raytracing_test_b200 import rc, voxel_host as vh

w, h = (int(v) for v in (sys.argv[1] if len(sys.argv) > 1 else "1920x1080").split("x"))
ao = int(sys.argv[2]) if len(sys.argv) > 2 else 64
frames = int(sys.argv[3]) if len(sys.argv) > 3 else 40
try:
    import torch

    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
except Exception:
    flush = None

host = vh.World.default()
pos, d, sun = vh.default_camera()
c = rc.RayCaster(0)
c.upload_world(host.describe())
look, _, _ = host.pick(pos, d, 30)


def run(sched, opts, **kw):
    c.set_option("schedule", sched)
    c.set_option("ao_opts", opts)
    p = rc.frame_params(w, h, pos, d, sun, shadows=True, ao_samples=ao, looking_at=tuple(int(v) for v in look), **kw)
    for _ in range(5):
        c.render_frame(p)
    c.sync()
    tot, pr, aos = [], [], []
    for _ in range(frames):
        if flush is not None:
            flush.zero_()
            torch.cuda.synchronize()
        c.render_frame(p)
        st = c.stats()
        tot.append(st.total_ms)
        pr.append(st.primary_ms)
        aos.append(st.ao_ms)
    rays = st.primary_rays + st.shadow_rays + st.ao_rays
    return float(np.median(tot)), float(np.min(tot)), float(np.median(pr)), float(np.median(aos)), rays, st.schedule


print(f"{os.environ.get('RC_LIBRC', 'librc.so')}  {w}x{h} ao {ao}, {frames} frames each, L2 flush {'on' if flush is not None else 'OFF'}")
names = {v: k for k, v in rc.SCHEDULES.items()}
for sched in range(5):
    for opts in ((0, 1) if ao else (1,)):
        med, mn, _, _, rays, used = run(sched, opts)
        line = f"schedule {names[sched]:18s} ao_opts {opts}: total median {med:.4f} min {mn:.4f} ms  {rays / med / 1e6:.1f} Grays/s"
        if sched < 2:
            _, _, pr, aos, _, _ = run(sched, opts, kernel_times=True)
            line += f"   [kernel_times: primary {pr:.4f} + ao {aos:.4f}]"
        if used != sched:
            line += f"  (ran as {names[used]})"
        print(line, flush=True)
for smem in (0, 12 * 1024, 25 * 1024):
    c.set_option("primary_smem", smem)
    _, _, pr, aos, _, _ = run(rc.SCHED_TWO_KERNELS, 1, kernel_times=True)
    print(f"two kernels, primary kernel launched with {smem} B of (unused) dynamic shared memory: primary {pr:.4f} ms")
c.set_option("primary_smem", 0)
_, _, pr, aos, _, _ = run(rc.SCHED_TWO_KERNELS, 1, kernel_times=True, column_regions=False)
print(f"two kernels, column regions off: primary {pr:.4f} ms")
_, _, pr, aos, _, _ = run(rc.SCHED_TWO_KERNELS, 1, kernel_times=True, column_regions=True)
print(f"two kernels, column regions on : primary {pr:.4f} ms")
for tt in (0, 1):
    c.set_option("top_table", tt)
    med, _, pr, aos, _, _ = run(rc.SCHED_TWO_KERNELS, 1, kernel_times=True)
    print(f"two kernels, top table {'on ' if tt else 'off'}: primary {pr:.4f} + ao {aos:.4f} ms")
for dg in (False, True):
    _, _, pr, aos, _, _ = run(rc.SCHED_TWO_KERNELS, 1, kernel_times=True, drop_gates=dg)
    print(f"two kernels, drop gates {'on ' if dg else 'off'}: primary {pr:.4f} ms")
med, mn, _, _, rays, _ = run(rc.SCHED_FUSED, 1, always_tail=True)
print(f"fused, free-flight tail forced: {med:.4f} ms")
med, mn, _, _, rays, _ = run(rc.SCHED_FUSED, 1, hits=True)
print(f"fused, hit records: {med:.4f} ms")
c.close()

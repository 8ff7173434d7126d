import os, sys
import numpy as np
sys.path.insert(0, os.getcwd())
from raytracing_test_b200 import rc, voxel_host as vh
import torch
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
host = vh.World.default(); pos, d, sun = vh.default_camera()
c = rc.RayCaster(0); c.upload_world(host.describe())
c.set_option("schedule", 0)
for (w,h) in ((1920,1080),(3840,2160)):
    p = rc.frame_params(w, h, pos, d, sun, shadows=True, ao_samples=64, kernel_times=True)
    for _ in range(5): c.render_frame(p)
    pr=[]; ao=[]
    for _ in range(40):
        flush.zero_(); torch.cuda.synchronize(); c.render_frame(p); st=c.stats(); pr.append(st.primary_ms); ao.append(st.ao_ms)
    print(f"{os.environ.get('RC_LIBRC','librc.so').split('/')[-1]:22s} {w}x{h}: primary median {np.median(pr):.4f} min {np.min(pr):.4f}  ao median {np.median(ao):.4f}")
c.close()

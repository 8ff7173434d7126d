"""Developer probe: where the primary kernel's time goes over the image.  Needs a librc built with -DRC_COST_MAP
(tools/build_variants.sh cost "-DRC_COST_MAP"): the alpha channel of the float frame then holds the cycles each pixel's warp
spent in shade_pixel.  Prints the share of warp-cycles per pixel class and per band of rows.
    RC_LIBRC=raytracing_test_b200/variants/librc_cost.so python tools/cost_map.py"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from raytracing_test_b200 import rc, voxel_host as vh

host = vh.World.default()
pos, d, sun = vh.default_camera()
c = rc.RayCaster(0)
c.upload_world(host.describe())
w, h = 1920, 1080
c.set_option("schedule", 0)
p = rc.frame_params(w, h, pos, d, sun, shadows=True, ao_samples=0, hits=True, rgba32f=True)
for _ in range(3):
    c.render_frame(p)
hits = c.read_hits().reshape(h, w)
if len(sys.argv) > 1 and sys.argv[1] == "notail":  # the product frame: no hit records, free-flight rays skip their tail
    p = rc.frame_params(w, h, pos, d, sun, shadows=True, ao_samples=0, rgba32f=True)
    for _ in range(3):
        c.render_frame(p)
f4 = c.read_image(fmt=rc.RC_RGBA32F).reshape(h, w, 4).astype(np.float64)
f = f4[:, :, 3]
# one value per warp tile (8 wide, 4 high)
tile = f.reshape(h // 4, 4, w // 8, 8).max(axis=(1, 3))
kind = hits["kind"].reshape(h // 4, 4, w // 8, 8)
steps = hits["steps"].reshape(h // 4, 4, w // 8, 8).astype(np.int64)
sky = (kind == 0).sum(axis=(1, 3))
tot = tile.sum()
print(f"warps {tile.size}, mean cycles {tile.mean():.0f}, primary {c.stats().primary_ms:.3f} ms")
for name, m in (("all sky", sky == 32), ("all terrain", sky == 0), ("mixed", (sky > 0) & (sky < 32))):
    print(f"{name:12s}: {m.mean() * 100:5.1f}% of warps, {tile[m].sum() / tot * 100:5.1f}% of warp-cycles, mean {tile[m].mean():.0f} cycles, "
          f"mean max-steps {steps.max(axis=(1, 3))[m].mean():.0f}")
# sections: per warp the lane maxima (x: set-up + loop A + free flight, y: ratios + replay, z: loop B + shading, rest: shadow ray)
sec = f4.reshape(h // 4, 4, w // 8, 8, 4).max(axis=(1, 3))
for name, m in (("all sky", sky == 32), ("all terrain", sky == 0)):
    t = sec[m]
    print(f"{name:12s}: loop A {t[:, 0].mean():.0f}  replay {t[:, 1].mean():.0f}  loop B + shading {t[:, 2].mean():.0f}  "
          f"shadow + rest {(t[:, 3] - t[:, 0] - t[:, 1] - t[:, 2]).mean():.0f}  total {t[:, 3].mean():.0f} cycles (lane maxima, they overlap)")
rows = tile.sum(axis=1)
nb = 10
for b in range(nb):
    lo, hi = b * len(rows) // nb, (b + 1) * len(rows) // nb
    print(f"rows {lo * 4:4d}-{hi * 4 - 1:4d} (bottom-up): {rows[lo:hi].sum() / tot * 100:5.1f}% of warp-cycles, sky pixels {sky[lo:hi].mean() / 32 * 100:5.1f}%")
# the most expensive all-sky warps: how far up the image are they?
m = sky == 32
if m.any():
    q = np.quantile(tile[m], [0.5, 0.9, 0.99])
    print("all-sky warps: median / p90 / p99 cycles", q)
    ys, xs = np.nonzero(m & (tile > q[1]))
    print("rows of the p90+ all-sky warps (pixel rows):", np.quantile(ys * 4, [0, 0.25, 0.5, 0.75, 1.0]))

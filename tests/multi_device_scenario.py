"""Own-process scenario of tests/test_multi_device_gpu.py (librc's state is process-global: rc_init_multi would take the
session fixture's context away).  python tests/multi_device_scenario.py -> prints "multi device ok"."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import camera_dir  # noqa: E402
from raytracing_test_b200 import rc as rcmod  # noqa: E402


def _devices(n=2):
    import torch

    have = torch.cuda.device_count()
    if have >= n:
        return list(range(n))
    os.environ["RC_ALLOW_DUPLICATE_DEVICES"] = "1"
    return [0] * n


def expect_refused(call, code=-8):
    try:
        call()
    except rcmod.RcError as e:
        assert e.code == code, (e.code, code)
        return
    raise AssertionError("call was not refused")


def main():
    import torch

    from raytracing_test_b200 import frame_gather as fg, voxel_host as vh

    world = vh.World.default()
    world_desc = world.describe()
    pos, d, sun = vh.default_camera()
    views = [(1920, 1080, pos, d, 64), (641, 357, (100.5, 60.25, 100.5), camera_dir(0.0, 60.0), 64),
             (1000, 562, (60.2, 25.9, 140.8), camera_dir(120.0, 35.0), 16), (320, 20, pos, d, 64), (64, 8, pos, d, 0)]
    single = {}
    with rcmod.RayCaster(0) as c:
        c.upload_world(world_desc)
        for i, (w, h, o, dd, ao) in enumerate(views):
            c.render_frame(rcmod.frame_params(w, h, o, dd, sun, shadows=True, ao_samples=ao, hits=True, rgba32f=True))
            single[i] = (c.read_image().copy(), c.read_hits().copy(), c.read_image(fmt=rcmod.RC_RGBA32F).copy(), c.stats())
    for n in (2, 3):
        devs = _devices(n)
        with rcmod.RayCaster(devs) as c:
            assert c.lib.rc_device_count() == n
            c.upload_world(world_desc)
            for i, (w, h, o, dd, ao) in enumerate(views):
                p = rcmod.frame_params(w, h, o, dd, sun, shadows=True, ao_samples=ao, hits=True, rgba32f=True)
                c.render_frame(p)
                img, hits, f32, st = c.read_image(), c.read_hits(), c.read_image(fmt=rcmod.RC_RGBA32F), c.stats()
                assert np.array_equal(img, single[i][0]), (n, i)
                assert hits.tobytes() == single[i][1].tobytes(), (n, i)
                assert np.array_equal(f32.view(np.uint32), single[i][2].view(np.uint32)), (n, i)
                for k in ("primary_rays", "shadow_rays", "ao_rays", "hit_pixels", "primary_steps"):
                    assert getattr(st, k) == getattr(single[i][3], k), (n, i, k)
            # pipelined read-back into page-locked memory: two frames in flight, each lands intact
            w, h, o, dd, ao = views[0]
            pinned = [torch.empty((h, w, 4), dtype=torch.uint8, pin_memory=True).numpy() for _ in range(2)]
            prev = None
            for k in range(5):
                c.render_frame(rcmod.frame_params(w, h, o, dd, sun, shadows=True, ao_samples=ao))
                t = c.read_image_async(pinned[k & 1])
                if prev is not None:
                    c.wait_image(prev)
                    assert np.array_equal(pinned[(k - 1) & 1], single[0][0]), (n, k)
                prev = t
            c.wait_image(prev)
            assert np.array_equal(pinned[0], single[0][0])
            # one frame on device 0, written by every device's kernels (peer stores): complete in device 0's stream order
            c.frame_target_create(w, h, 2)
            for k in range(3):
                c.render_frame(rcmod.frame_params(w, h, o, dd, sun, shadows=True, ao_samples=ao))
                slot = c.frame_target_last_slot()
                assert slot == k % 2
                c.sync()
                frame = fg.device_bytes(c.frame_target_ptr(slot), h * w * 4, torch.device("cuda", devs[0])).view(h, w, 4).cpu().numpy()
                assert np.array_equal(frame, single[0][0]), (n, k)
                assert np.array_equal(c.read_image(), single[0][0])
            c.frame_target_close()
            # the calls of the one-process-per-GPU deployment are refused, not silently misapplied
            for call in (lambda: c.set_shard(0, 2), lambda: c.set_stream(0), lambda: c.frame_target_open(w, h, 1, bytes(64))):
                expect_refused(call)
            # a world generated on every device, and an incremental edit pushed to every device
            c.generate_world()
            c.render_frame(rcmod.frame_params(w, h, o, dd, sun, shadows=True, ao_samples=ao, hits=True))
            assert np.array_equal(c.read_image(), single[0][0]) and c.read_hits().tobytes() == single[0][1].tobytes()
            host = vh.World.default()
            c.update_world(host.describe(), [], [])   # first call: full upload with tracking
            host.put_block(60, 60, 60, 0xFF0000, 0, 0.0)
            host.put_block(61, 60, 60, 0x00FF00, 2, 0.0)
            info = c.update_world(host.describe(), None, None)  # RC_ALL_BLOCKS: diff everything
            assert not info.full_rebuild and info.changed_ref_nodes > 0
            c.render_frame(rcmod.frame_params(640, 360, (50.5, 62.5, 50.5), camera_dir(45.0, 10.0), sun, shadows=True, ao_samples=16, hits=True))
            multi_edit = (c.read_image().copy(), c.read_hits().copy())
            host_desc = host.describe()
            with_single = None
        with rcmod.RayCaster(0) as c1:
            c1.upload_world(host_desc)
            c1.render_frame(rcmod.frame_params(640, 360, (50.5, 62.5, 50.5), camera_dir(45.0, 10.0), sun, shadows=True, ao_samples=16, hits=True))
            with_single = (c1.read_image().copy(), c1.read_hits().copy())
        assert np.array_equal(multi_edit[0], with_single[0]) and multi_edit[1].tobytes() == with_single[1].tobytes(), n
        assert int((with_single[1]["x"] == 60).sum()) > 0  # the placed block is in view
        host.close()
    world.close()
    print("multi device ok")


if __name__ == "__main__":
    main()

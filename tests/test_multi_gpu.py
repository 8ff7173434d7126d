"""GPU test of the image-band sharding: every rank's bands rendered by librc (rc_set_shard), put
together by rc_compose_frame, must equal the single-GPU frame byte for byte.
One GPU is enough: the ranks are rendered one after the other into the layout the NCCL gather
produces (the gather itself is exercised over gloo in test_multi_gpu_cpu.py and by bench.py --gpus N)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("world,w,h", [(2, 640, 360), (3, 333, 200), (4, 1920, 1080), (8, 1280, 720), (8, 50, 40)])
def test_sharded_frame_equals_single_gpu_frame(rcmod, caster, defaults, world, w, h):
    import torch

    from raytracing_test_b200 import frame_gather as fg

    pos, d, sun = defaults
    p = rcmod.frame_params(w, h, pos, d, sun, shadows=True, ao_samples=16, hits=True)
    caster.set_shard(0, 1)
    caster.render_frame(p)
    full_img, full_hits = caster.read_image(), caster.read_hits()
    dev = torch.device("cuda", 0)
    pr = fg.padded_rows(h, world)
    g_img = torch.zeros((world, pr * w * 4), dtype=torch.uint8, device=dev)
    g_hit = torch.zeros((world, pr * w * 20), dtype=torch.uint8, device=dev)
    rays = 0
    try:
        for r in range(world):
            caster.set_shard(r, world)
            caster.render_frame(p)
            caster.sync()
            g_img[r].copy_(fg.device_bytes(caster.device_image_ptr(), pr * w * 4, dev))
            g_hit[r].copy_(fg.device_bytes(caster.device_hits_ptr(), pr * w * 20, dev))
            # rc_read_image of a shard: this rank's rows in place, the others untouched
            part = np.full((h, w, 4), 7, dtype=np.uint8)
            caster.read_image(part)
            mine = rcmod.band_rows_of_rank(h, r, world)
            other = sorted(set(range(h)) - set(mine))
            assert np.array_equal(part[mine], full_img[mine]) and (part[other] == 7).all()
            rays += caster.stats().primary_rays
        torch.cuda.synchronize()
        img = torch.empty((h, w, 4), dtype=torch.uint8, device=dev)
        hit = torch.empty((h * w * 20,), dtype=torch.uint8, device=dev)
        caster.compose_frame(g_img.data_ptr(), img.data_ptr(), w, h, world, 4)
        caster.compose_frame(g_hit.data_ptr(), hit.data_ptr(), w, h, world, 20)
        caster.sync()
        assert rays == w * h
        assert np.array_equal(img.cpu().numpy(), full_img)
        assert hit.cpu().numpy().tobytes() == full_hits.tobytes()
        # the torch expression used on the gloo path agrees with the kernel
        ref = fg.compose_reference(g_img.view(world, pr, w, 4), w, h, world)
        assert torch.equal(ref, img)
    finally:
        caster.set_shard(0, 1)

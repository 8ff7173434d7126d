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


@pytest.mark.parametrize("world,w,h,slots", [(2, 640, 360, 1), (3, 333, 200, 2), (8, 1280, 720, 1), (8, 50, 40, 2)])
def test_direct_frame_target_equals_single_gpu_frame(rcmod, caster, defaults, world, w, h, slots):
    """rc_frame_target_*: every rank's kernels store their pixels straight into ONE whole frame (on a real box:
    rank 0's, mapped by the peers over CUDA IPC / NVLink; here one process plays all ranks into its own
    target).  The result must be the single-GPU frame byte for byte, with no gather and no compose."""
    import torch

    from raytracing_test_b200 import frame_gather as fg

    pos, d, sun = defaults
    p = rcmod.frame_params(w, h, pos, d, sun, shadows=True, ao_samples=16)
    caster.set_shard(0, 1)
    caster.render_frame(p)
    full_img = caster.read_image()
    dev = torch.device("cuda", 0)
    try:
        handles = caster.frame_target_create(w, h, slots)
        assert len(handles) == 64 * slots and any(handles)
        frames = [fg.device_bytes(caster.frame_target_ptr(s), h * w * 4, dev).view(h, w, 4) for s in range(slots)]
        assert caster.frame_target_ptr(slots) is None
        for rounds in range(2):  # with 2 slots: the two rounds land in different frames only if world is odd
            used = set()
            for r in range(world):
                caster.set_shard(r, world)
                caster.render_frame(p)
                used.add(caster.frame_target_last_slot())
            caster.sync()
            got = torch.zeros((h, w, 4), dtype=torch.uint8, device=dev)
            for s in used:  # one process playing all ranks advances the slot per call: merge what each slot received
                got |= frames[s]
            assert np.array_equal(got.cpu().numpy(), full_img)
            for s in range(slots):
                frames[s].zero_()
        # another resolution does not touch the target; the compact per-rank buffer still works beside it
        caster.set_shard(0, 1)
        caster.render_frame(rcmod.frame_params(w + 16, h, pos, d, sun, shadows=False))
        assert caster.frame_target_last_slot() == -1
        caster.render_frame(p)
        assert caster.frame_target_last_slot() >= 0 and np.array_equal(caster.read_image(), full_img)
    finally:
        caster.set_shard(0, 1)
        caster.frame_target_close()
    caster.render_frame(p)
    assert caster.frame_target_last_slot() == -1 and np.array_equal(caster.read_image(), full_img)


def test_direct_frame_across_processes():
    """tests/multi_gpu_direct.py under torchrun, one rank per GPU (needs >= 2 GPUs: `gpurun --gpus 2`)."""
    import os
    import subprocess
    import sys

    import torch

    n = min(torch.cuda.device_count(), 4)
    if n < 2:
        pytest.skip("one GPU: the cross-process IPC path needs two (run with gpurun --gpus 2)")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr",
                        "127.0.0.1", "--master-port", "29533", os.path.join(root, "tests", "multi_gpu_direct.py")],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "direct frame ok" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]

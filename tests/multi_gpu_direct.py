"""Multi-process check of the direct multi-GPU frame (run under torchrun, one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
        tests/multi_gpu_direct.py

Every rank renders its bands; rank 0's frame, written by all ranks' kernels through CUDA IPC peer memory
(frame_gather.ShardedRayCaster mode "direct"), must equal (a) the frame the NCCL gather + compose path
produces and (b) the frame one GPU renders alone, byte for byte -- for several frames in a row (the two
frame slots alternate), after a change of resolution, and through the pipelined read-back; and so must the
frame that assembles itself in shared page-locked HOST memory when every rank reads back only its own bands
(frame_gather.SharedHostFrames: no device-side gather at all).
Prints "direct frame ok" on rank 0."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import camera_dir  # noqa: E402
from raytracing_test_b200 import frame_gather as fg, rc, voxel_host as vh  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    import datetime

    dist.init_process_group("nccl", device_id=torch.device("cuda", local), timeout=datetime.timedelta(seconds=120))
    host = vh.World.default()
    pos, _, sun = vh.default_camera()
    direct = fg.ShardedRayCaster(local, rank, world, mode="direct")
    direct.upload_world(host.describe())
    views = [(1280, 720, 45.0), (1280, 720, 60.0), (1280, 720, 75.0), (640, 360, 30.0), (640, 360, 100.0), (333, 200, 45.0)]

    # (b) the whole frame on one GPU: the same library, unsharded (rank 0 only)
    alone = {}
    if rank == 0:
        direct.caster.set_shard(0, 1)
        for (w, h, yaw) in views:
            direct.caster.render_frame(rc.frame_params(w, h, pos, camera_dir(yaw, 15.0), sun, shadows=True, ao_samples=16))
            alone[(w, h, yaw)] = direct.caster.read_image().copy()
        direct.caster.set_shard(rank, world)

    got_direct = {}
    for (w, h, yaw) in views:
        direct.render(rc.frame_params(w, h, pos, camera_dir(yaw, 15.0), sun, shadows=True, ao_samples=16))
        torch.cuda.synchronize()
        if rank == 0:
            got_direct[(w, h, yaw)] = direct.frame().cpu().numpy().copy()
    # pipelined read-back: frames in flight, each must arrive intact
    w, h = 1280, 720
    pinned = [torch.empty((h, w, 4), dtype=torch.uint8, pin_memory=True) for _ in range(2)]
    yaws = [45.0, 60.0, 75.0, 45.0, 60.0, 75.0, 45.0]
    landed, prev = [], None
    for i, yaw in enumerate(yaws):
        direct.render(rc.frame_params(w, h, pos, camera_dir(yaw, 15.0), sun, shadows=True, ao_samples=16))
        ev = direct.read_frame_async(pinned[i & 1])
        if prev is not None:
            if rank == 0:
                prev[0].synchronize()
                landed.append((prev[2], prev[1].numpy().copy()))
        prev = (ev, pinned[i & 1], yaw)
    if rank == 0:
        prev[0].synchronize()
        landed.append((prev[2], prev[1].numpy().copy()))
    torch.cuda.synchronize()
    # the frames above were signalled complete by flags in rank 0's memory (rc_frame_sync_*), no collective per frame; no
    # wait gave up
    assert direct.signal == "flags", "flag block could not be mapped: the all-reduce fall-back ran instead"
    if rank == 0:
        assert direct.caster.frame_sync_timeouts() == 0

    # host delivery with no device-side gather: every rank reads its own bands back into ONE shared page-locked
    # host frame; rank 0 sees each frame complete and intact
    sink = fg.SharedHostFrames(w, h, rank, world, slots=2, pin=direct.caster)
    host_ok, prev = 0, None

    def finish(ticket, i, yaw):
        nonlocal host_ok
        direct.caster.wait_image(ticket)
        sink.mark_landed(i)
        if rank == 0:
            frame = sink.acquire(i)
            assert np.array_equal(frame, alone[(1280, 720, yaw)]), f"shared host frame {i} (yaw {yaw}) differs"
            host_ok += 1
            sink.release(i)

    for i, yaw in enumerate(yaws):
        direct.caster.render_frame(rc.frame_params(w, h, pos, camera_dir(yaw, 15.0), sun, shadows=True, ao_samples=16))
        sink.wait_slot_free(i)
        t = direct.caster.read_image_async(sink.frames[i % 2])
        if prev is not None:
            finish(*prev)
        prev = (t, i, yaw)
    finish(*prev)
    assert rank != 0 or host_ok == len(yaws)
    sink.close()
    direct.close()

    # the completion signal this replaced (one-element NCCL all-reduce per frame): same frames, pipelined read-back included
    ar = fg.ShardedRayCaster(local, rank, world, mode="direct", signal="allreduce")
    ar.upload_world(host.describe())
    w, h = 1280, 720
    prev = None
    for i, yaw in enumerate(yaws):
        ar.render(rc.frame_params(w, h, pos, camera_dir(yaw, 15.0), sun, shadows=True, ao_samples=16))
        ev = ar.read_frame_async(pinned[i & 1])
        if prev is not None and rank == 0:
            prev[0].synchronize()
            assert np.array_equal(prev[1].numpy(), alone[(1280, 720, prev[2])]), f"all-reduce signal: pipelined frame (yaw {prev[2]}) differs"
        prev = (ev, pinned[i & 1], yaw)
    if rank == 0:
        prev[0].synchronize()
        assert np.array_equal(prev[1].numpy(), alone[(1280, 720, prev[2])])
    torch.cuda.synchronize()
    assert ar.signal == "allreduce"
    ar.close()

    gather = fg.ShardedRayCaster(local, rank, world, mode="gather")
    gather.upload_world(host.describe())
    for (w, h, yaw) in views:
        gather.render(rc.frame_params(w, h, pos, camera_dir(yaw, 15.0), sun, shadows=True, ao_samples=16))
        torch.cuda.synchronize()
        if rank == 0:
            g = gather.frame().cpu().numpy()
            assert np.array_equal(g, alone[(w, h, yaw)]), f"gather path differs from the single-GPU frame at {(w, h, yaw)}"
            assert np.array_equal(got_direct[(w, h, yaw)], g), f"direct frame differs from the gathered frame at {(w, h, yaw)}"
    if rank == 0:
        assert len(landed) == len(yaws)
        for yaw, img in landed:
            assert np.array_equal(img, alone[(1280, 720, yaw)]), f"pipelined frame (yaw {yaw}) arrived damaged"
    gather.close()
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print(f"direct frame ok: {len(views)} views x {world} ranks, {len(yaws)} pipelined frames, {host_ok} shared host frames")


if __name__ == "__main__":
    main()

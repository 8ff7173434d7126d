"""Host logic of the multi-GPU path on CPU: band sharding layout, gather over gloo (world_size 2
and 3), frame composition.  The per-rank bands come from the oracle's band sampling (the GPU
version of the same test lives in test_multi_gpu.py and renders them with librc)."""
import os
import socket
import sys
import time

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_layout_partitions_the_rows(rcmod):
    for h in (1, 15, 16, 17, 100, 1080, 2160, 4320):
        for world in (1, 2, 3, 4, 8):
            seen = []
            pads = set()
            for r in range(world):
                rows, padded = rcmod.shard_layout(h, r, world)
                mine = rcmod.band_rows_of_rank(h, r, world)
                assert rows == len(mine) <= padded and padded % rcmod.BAND_ROWS == 0
                pads.add(padded)
                seen += mine
            assert sorted(seen) == list(range(h)) and len(pads) == 1
    lib = rcmod.load_library()
    assert lib.rc_shard_layout(100, 2, 2, None, None) == -2
    assert lib.rc_shard_layout(0, 0, 1, None, None) == -2


def test_compose_reference_inverts_numpy_bands(rcmod):
    from raytracing_test_b200 import frame_gather as fg

    rng = np.random.default_rng(0)
    for (h, w, world) in [(100, 7, 3), (1080, 16, 8), (33, 5, 2), (16, 4, 4), (5, 3, 4)]:
        frame = rng.integers(0, 255, size=(h, w, 4), dtype=np.uint8)
        gathered = np.stack([fg.numpy_bands(frame, r, world) for r in range(world)])
        out = fg.compose_reference(torch.from_numpy(gathered), w, h, world).numpy()
        assert np.array_equal(out, frame)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, w, h, q):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    os.environ["OMP_NUM_THREADS"] = "2"
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import bindings as ob
        from raytracing_test_b200 import frame_gather as fg
        from raytracing_test_b200 import rc, voxel_host as vh

        world_host = vh.World.default()  # replicated: every rank builds the same world
        nodes, arrays = world_host.pools()
        ow = ob.make_world(nodes, arrays, 6)
        pos, d, sun = vh.default_camera()
        # this rank's bands only, then packed the way librc stores them
        part = ob.render(ow, ob.make_frame(w, h, pos, d, sun, shadows=True, bands=(rc.BAND_ROWS, world, rank), threads=2))
        rows = rc.band_rows_of_rank(h, rank, world)
        assert part["counters"].rays[0] == len(rows) * w
        local = torch.from_numpy(fg.numpy_bands(part["rgba8"], rank, world))
        gathered = fg.gather_bands(local, dst=0)
        if rank == 0:
            frame = fg.compose_reference(gathered, w, h, world).numpy()
            full = ob.render(ow, ob.make_frame(w, h, pos, d, sun, shadows=True, threads=2))
            q.put(bool(np.array_equal(frame, full["rgba8"]) and frame.any()))
        else:
            assert gathered is None
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,w,h", [(2, 160, 90), (3, 96, 100)])
def test_gloo_sharded_frame_equals_single_frame(world, w, h):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, w, h, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=180)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True


def _host_frames_worker(rank, world, port, w, h, n_frames, q):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from raytracing_test_b200 import frame_gather as fg
        from raytracing_test_b200 import rc

        sink = fg.SharedHostFrames(w, h, rank, world, slots=2, pin=None)
        rows = rc.band_rows_of_rank(h, rank, world)
        ok = True
        for i in range(n_frames):
            # frame i: every pixel = i + 1 in the rows this rank owns; what rc_read_image_async does with a sharded
            # frame (this rank's rows to their place, the others untouched), here by a numpy store
            sink.wait_slot_free(i)
            if rank == 1 and i % 3 == 0:
                time.sleep(0.01)  # a slow producer: the consumer must wait for it, the fast ones must not lap it
            sink.frames[i % 2][rows] = (i + 1) & 0xFF
            sink.mark_landed(i)
            if rank == 0:
                frame = sink.acquire(i)
                ok = ok and bool((frame == ((i + 1) & 0xFF)).all())
                sink.release(i)
        sink.close()
        if rank == 0:
            q.put(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,w,h", [(2, 64, 100), (4, 33, 70)])
def test_shared_host_frames_assemble_without_a_gather(world, w, h):
    """SharedHostFrames: every rank delivers its bands into one shared-memory frame; the consumer sees each
    frame complete and no producer overwrites a frame that has not been released (2 slots, 40 frames)."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_host_frames_worker, args=(r, world, port, w, h, 40, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=180)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True

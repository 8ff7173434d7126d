"""Image-tile sharding across the GPUs of one box: one process per GPU (torch.distributed), every
rank holds the whole flattened world and renders the 16-row bands with band % world_size == rank
(librc: rc_set_shard).  Two ways for the bands to become one frame on rank 0:

  direct (default on CUDA): rank 0 owns the whole frame (rc_frame_target_create), the other ranks
      map it through CUDA IPC (rc_frame_target_open) and every rank's kernels store each finished
      pixel straight into it over NVLink peer memory -- the transfer is fused into the tracing
      kernels, there is no gather and no compose pass.  "Frame complete" crosses the processes as
      flags in rank 0's memory (rc_frame_sync_*: one store behind each peer's kernels, one polling
      warp on rank 0's stream, and the other way round for "this frame slot may be written again");
      no collective per frame (signal="allreduce": the one-element NCCL all-reduce this replaces,
      kept for A/B and as the fall-back when the flag block cannot be mapped).
  gather: the compact band buffers are gathered to rank 0 over NCCL and de-interleaved by one
      librc kernel (rc_compose_frame).  Kept as the comparison baseline and for hosts without
      peer access.

torch is plumbing here: process group, handle exchange, streams.  The gather / compose logic also
runs on the gloo backend with CPU tensors (tests/test_multi_gpu_cpu.py), where the per-rank bands
come from the test's renderer instead of librc.
"""
import os

import numpy as np
import torch
import torch.distributed as dist

from . import rc


class _DevicePtr:
    """Exposes a raw device pointer through __cuda_array_interface__ so torch can view it."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3,
                                         "strides": None}


def device_bytes(ptr, nbytes, device):
    """uint8 torch view (no copy) of nbytes of device memory at ptr."""
    if not ptr:
        raise RuntimeError("null device pointer (no frame rendered?)")
    return torch.as_tensor(_DevicePtr(ptr, nbytes), device=device)


def padded_rows(height, world_size):
    return rc.shard_layout(height, 0, world_size)[1]


def compose_reference(gathered, width, height, world_size):
    """The de-interleave rc_compose_frame does, as a torch expression (CPU/gloo path and tests).
    gathered: [world_size, padded_rows, width, C] -> [height, width, C]."""
    pr = gathered.shape[1]
    bands = pr // rc.BAND_ROWS
    g = gathered.reshape(world_size, bands, rc.BAND_ROWS, width, -1)
    return g.permute(1, 0, 2, 3, 4).reshape(bands * world_size * rc.BAND_ROWS, width, -1)[:height]


def gather_bands(local, dst=0, group=None):
    """Gathers equal-size per-rank band buffers to rank dst. Returns [world, ...] on dst, else None.
    NCCL: grouped send/recv under dist.gather; one message per peer (16.6 MB at 8K / 8 GPUs)."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    if world == 1:
        return local.unsqueeze(0)
    out = None
    if rank == dst:
        out = torch.empty((world,) + tuple(local.shape), dtype=local.dtype, device=local.device)
        dist.gather(local, list(out.unbind(0)), dst=dst, group=group)
    else:
        dist.gather(local, None, dst=dst, group=group)
    return out


class ShardedRayCaster:
    """rc.RayCaster for one rank of a one-process-per-GPU job. All ranks call render(); rank 0's
    frame() returns the composed RGBA8 frame as a device tensor [height, width, 4]."""

    def __init__(self, device_index, rank=None, world_size=None, group=None, mode="direct", slots=2, signal="flags"):
        if mode not in ("direct", "gather"):
            raise ValueError("mode must be 'direct' or 'gather'")
        if signal not in ("flags", "allreduce"):
            raise ValueError("signal must be 'flags' or 'allreduce'")
        self.mode = mode
        self.signal = signal
        self._sync_open = False
        self._allowed_by_copy = set()  # frames whose slot is released by a read-back in flight (rank 0, signal "flags")
        self.slots = slots
        self.group = group
        self.rank = dist.get_rank(group) if rank is None else rank
        self.world = dist.get_world_size(group) if world_size is None else world_size
        self.device = torch.device("cuda", device_index)
        self.caster = rc.RayCaster(device_index)
        self.caster.set_shard(self.rank, self.world)
        # kernels, the gather and the compose are all ordered on ONE non-default torch stream (torch's
        # default stream has handle 0, which rc_set_stream reads as "use the library's own stream")
        self.stream = torch.cuda.Stream(self.device)
        self.caster.set_stream(self.stream.cuda_stream)
        self._frame = None
        self._gathered = None
        self._dims = None
        self._token = None
        self._frames = None
        self._target_open = False
        self._copy_stream = None
        self._pending = {}  # frame slot -> event of the read-back still reading it (rank 0)
        self.completion_signal = "not sharded"

    def close(self):
        if self._target_open:
            self._close_target()
        self.caster.close()

    def _meet(self):
        """All ranks' streams have drained (device-side all-reduce + host sync): safe to map / unmap."""
        dist.all_reduce(self._token, group=self.group)
        torch.cuda.synchronize(self.device)

    def _close_target(self):
        # nobody may unmap / free while a peer's kernels can still write
        self._meet()
        if self._sync_open:
            self.caster.frame_sync_close()
            self._sync_open = False
        self.caster.frame_target_close()
        self._frames = None
        self._target_open = False
        self._pending.clear()
        self._allowed_by_copy.clear()

    def _open_sync(self):
        """Rank 0 allocates the completion flags, everyone else maps them; all ranks end up with the same answer."""
        box = [None]
        if self.rank == 0:
            try:
                box[0] = self.caster.frame_sync_create()
            except rc.RcError as e:
                box[0] = e
        dist.broadcast_object_list(box, src=0, group=self.group)
        ok = torch.ones(1, dtype=torch.int32, device=self.device)
        if isinstance(box[0], Exception):
            ok.zero_()
        elif self.rank != 0:
            try:
                self.caster.frame_sync_open(box[0])
            except rc.RcError:
                ok.zero_()
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=self.group)
        torch.cuda.synchronize(self.device)
        if int(ok.item()) == 0:
            try:
                self.caster.frame_sync_close()
            except rc.RcError:
                pass
            return False
        self._sync_open = True
        return True

    def _open_target(self, w, h):
        """Rank 0 allocates the frame(s), everyone else maps them (CUDA IPC handles travel as bytes)."""
        if self._token is None:
            self._token = torch.zeros(1, dtype=torch.int32, device=self.device)
        if self._target_open:
            self._close_target()
        box = [None]
        if self.rank == 0:
            try:
                box[0] = self.caster.frame_target_create(w, h, self.slots)
            except rc.RcError as e:
                box[0] = e  # every rank must learn of it, or the others wait in the broadcast for ever
        dist.broadcast_object_list(box, src=0, group=self.group)
        ok = torch.ones(1, dtype=torch.int32, device=self.device)
        if isinstance(box[0], Exception):
            ok.zero_()
        elif self.rank == 0:
            self._frames = [device_bytes(self.caster.frame_target_ptr(s), h * w * 4, self.device).view(h, w, 4)
                            for s in range(self.slots)]
        else:
            try:
                self.caster.frame_target_open(w, h, self.slots, box[0])
            except rc.RcError as e:  # no peer access to rank 0's GPU from this one
                self._open_error = e
                ok.zero_()
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=self.group)
        torch.cuda.synchronize(self.device)
        if int(ok.item()) == 0:
            # some rank cannot map the frame: all ranks fall back to the gather path together
            self.caster.frame_target_close()
            self._frames = None
            self.mode = "gather"
            self.fallback_reason = str(getattr(self, "_open_error", box[0] if isinstance(box[0], Exception) else "a peer rank"))
            if self.rank == 0:
                import sys

                print(f"frame_gather: direct frame unavailable ({self.fallback_reason}); using the NCCL gather path", file=sys.stderr)
            return False
        self._target_open = True
        if self.signal == "flags" and not self._open_sync():
            self.signal = "allreduce"
        self.completion_signal = (
            "frame completion by per-rank flags in rank 0's memory (rc_frame_sync_*: a store behind each peer's kernels, one "
            "polling warp on rank 0's stream), no collective per frame" if self.signal == "flags" else
            "one-element NCCL all-reduce behind the kernels as the completion signal")
        self._meet()  # every peer has the frame (and the flags) mapped before anyone renders into it
        return True

    def upload_world(self, desc):
        return self.caster.upload_world(desc)  # replicated: every rank uploads the same host world

    def render(self, params):
        """Enqueue one frame on self.stream (asynchronous): direct = kernels (storing into rank 0's frame) +
        completion all-reduce; gather = kernels -> NCCL gather to rank 0 -> compose."""
        with torch.cuda.stream(self.stream):
            self._render(params)

    def _render(self, params):
        w, h = params.width, params.height
        direct = self.world > 1 and self.mode == "direct"
        if direct and self._dims != (w, h):
            direct = self._open_target(w, h)
            self._dims = (w, h) if direct else None
        flags = direct and self._sync_open
        if flags:
            # frame f of this sync block goes into slot f % slots (target and flags were created together)
            f = self.caster.frame_sync_frames()
            if self.rank == 0:
                # rank 0's own kernels overwrite that slot too: the read-back that last used it must be done
                ev = self._pending.pop(f % self.slots, None)
                if ev is not None:
                    torch.cuda.current_stream(self.device).wait_event(ev)
                # the peers may start frame f + 1 as soon as ITS slot is free: now, unless a read-back still uses it (then
                # the copy stream says so when that copy is done, read_frame_async)
                if (f + 1) in self._allowed_by_copy:
                    self._allowed_by_copy.discard(f + 1)
                else:
                    self.caster.frame_sync_allow(f + 1)
            else:
                self.caster.frame_sync_begin()
        self.caster.render_frame(params)
        pr = padded_rows(h, self.world)
        local = device_bytes(self.caster.device_image_ptr(rc.RC_RGBA8), pr * w * 4, self.device)
        if self.world == 1:
            self._frame = local[: h * w * 4].view(h, w, 4)
            return
        if flags:
            slot = self.caster.frame_target_last_slot()
            self.caster.frame_sync_end()
            if self.rank == 0:
                self._frame = self._frames[slot]
                self._slot = slot
                self._frame_index = f
            return
        if direct:
            # frame complete on rank 0 <=> every rank's kernels have finished: a one-element all-reduce queued
            # behind them on each rank's stream (peer stores are visible when the storing kernel has completed)
            slot = self.caster.frame_target_last_slot()
            if self.rank == 0:
                # once this all-reduce completes the peers may start the NEXT frame, which overwrites the other
                # slot: a read-back still copying out of it must have finished before rank 0 lets them go
                nxt = (slot + 1) % self.slots
                ev = self._pending.pop(nxt, None)
                if ev is not None:
                    torch.cuda.current_stream(self.device).wait_event(ev)
            dist.all_reduce(self._token, group=self.group)
            if self.rank == 0:
                self._frame = self._frames[slot]
                self._slot = slot
            return
        if self._dims != (w, h):
            self._dims = (w, h)
            if self.rank == 0:
                self._gathered = torch.empty((self.world, pr * w * 4), dtype=torch.uint8, device=self.device)
                self._frame = torch.empty((h, w, 4), dtype=torch.uint8, device=self.device)
        if self.rank == 0:
            dist.gather(local, list(self._gathered.unbind(0)), dst=0, group=self.group)
            self.caster.compose_frame(self._gathered.data_ptr(), self._frame.data_ptr(), w, h, self.world, 4)
        else:
            dist.gather(local, None, dst=0, group=self.group)

    def frame(self):
        return self._frame if self.rank == 0 else None

    def read_frame_async(self, pinned):
        """Rank 0, direct mode: queues the copy of the last frame into `pinned` ([h, w, 4] uint8, pinned) on a
        copy stream and returns an event; the next frame renders (on every rank) meanwhile into the other slot."""
        if self.rank != 0:
            return None
        if self.world == 1 or self.mode != "direct":
            raise RuntimeError("read_frame_async is the multi-GPU direct-mode read-back; world 1: RayCaster.read_image_async")
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(self.device)
        ready = torch.cuda.Event()
        ready.record(self.stream)
        self._copy_stream.wait_event(ready)
        with torch.cuda.stream(self._copy_stream):
            pinned.copy_(self._frame, non_blocking=True)
        done = torch.cuda.Event()
        done.record(self._copy_stream)
        self._pending[self._slot] = done
        if self._sync_open:
            # this slot is free again once the copy has left it: the copy stream tells the peers so
            nxt = self._frame_index + self.slots
            self.caster.frame_sync_allow(nxt, self._copy_stream.cuda_stream)
            self._allowed_by_copy.add(nxt)
        return done

    def stats(self):
        return self.caster.stats()


def numpy_bands(full_frame, rank, world_size):
    """Test helper: the compact band buffer (padded) a rank would hold for a given full frame."""
    h, w = full_frame.shape[:2]
    pr = padded_rows(h, world_size)
    out = np.zeros((pr, w) + full_frame.shape[2:], dtype=full_frame.dtype)
    rows = rc.band_rows_of_rank(h, rank, world_size)
    out[: len(rows)] = full_frame[rows]
    return out


class SharedHostFrames:
    """Host delivery without any device-side gather: `slots` whole RGBA8 frames in POSIX shared memory, mapped
    by every rank of the box and page-locked in each (rc_host_register), plus two small counter arrays.  Every
    rank reads its own bands back into the shared frame (rc_read_image_async scatters a sharded frame's rows to
    their place), so a frame assembles itself in host memory over all the GPUs' PCIe links at once.

    Protocol (single writer per counter; a full fence -- rc_host_fence -- orders each counter store after the frame bytes
    that were waited for, and each counter load before the frame bytes read afterwards, so hosts weaker than x86 TSO see
    them in order too; waiting loops yield the core between polls):
      landed[r]  = number of frames rank r has completely delivered        (written by rank r)
      consumed   = number of frames the consumer (rank 0) has released      (written by rank 0)
    Rank r may start delivering frame i (into slot i % slots) once consumed >= i - slots + 1; the consumer owns
    frame i once min(landed) >= i + 1.

    pin=None: no page-locking (CPU tests of the protocol)."""

    COUNTER_BYTES = 4096
    SPIN_TIMEOUT_S = 30.0  # a peer that died must not leave the others spinning for ever

    def __init__(self, width, height, rank, world_size, group=None, slots=2, pin=None):
        from multiprocessing import shared_memory

        self.rank, self.world, self.slots, self.group = rank, world_size, slots, group
        self.frame_bytes = width * height * 4
        size = slots * self.frame_bytes + self.COUNTER_BYTES
        box = [None]
        if rank == 0:
            try:
                self.shm = shared_memory.SharedMemory(create=True, size=size)
                self.shm.buf[size - self.COUNTER_BYTES:size] = bytes(self.COUNTER_BYTES)
                box[0] = self.shm.name
            except OSError as e:  # /dev/shm too small for two frames
                box[0] = e
        dist.broadcast_object_list(box, src=0, group=group)
        if isinstance(box[0], Exception):
            raise RuntimeError(f"SharedHostFrames: rank 0 could not create the shared segment: {box[0]}")
        if rank != 0:
            self.shm = shared_memory.SharedMemory(name=box[0])
            # Python < 3.13 registers attached segments with this process's resource tracker too, which would
            # unlink the creator's segment (and warn) when this process exits: rank 0 alone owns the name
            try:
                from multiprocessing import resource_tracker

                resource_tracker.unregister(self.shm._name, "shared_memory")
            except Exception:
                pass
        self.whole = np.ndarray((size,), dtype=np.uint8, buffer=self.shm.buf)
        self.frames = [self.whole[s * self.frame_bytes:(s + 1) * self.frame_bytes].reshape(height, width, 4) for s in range(slots)]
        counters = self.whole[slots * self.frame_bytes:].view(np.int64)
        self.landed = counters[:world_size]
        self.consumed = counters[64:65]
        self._pin = pin
        if pin is not None:
            pin.host_register(self.whole)
        dist.barrier(group=group)

    def _spin(self, ready, what):
        import time

        if ready():
            return
        deadline = time.monotonic() + self.SPIN_TIMEOUT_S
        polls = 0
        while not ready():
            polls += 1
            if polls & 15 == 0:
                os.sched_yield()  # the cores are shared with the other ranks' read-backs
            if time.monotonic() > deadline:
                raise TimeoutError(f"SharedHostFrames rank {self.rank}: {what} (landed {self.landed.tolist()}, "
                                   f"consumed {int(self.consumed[0])})")

    def wait_slot_free(self, i):
        """Producer side, before delivering frame i."""
        need = i - self.slots + 1
        self._spin(lambda: int(self.consumed[0]) >= need, f"frame {i}: its slot was never released")
        rc.host_fence()  # acquire: the consumer's reads of the old frame are done before this rank overwrites the slot

    def mark_landed(self, i):
        """Producer side, after frame i is complete in host memory (rc_wait_image returned)."""
        rc.host_fence()  # release: the frame bytes (device DMA, waited for by rc_wait_image) before the counter
        self.landed[self.rank] = i + 1

    def acquire(self, i):
        """Consumer (rank 0): blocks until frame i is complete; returns it (valid until release(i))."""
        self._spin(lambda: int(self.landed.min()) >= i + 1, f"frame {i} never completed")
        rc.host_fence()  # acquire: the counters before the frame bytes
        return self.frames[i % self.slots]

    def release(self, i):
        rc.host_fence()
        self.consumed[0] = i + 1

    def close(self):
        dist.barrier(group=self.group)
        if self._pin is not None:
            self._pin.host_unregister(self.whole)
        self.frames = self.landed = self.consumed = self.whole = None
        self.shm.close()
        dist.barrier(group=self.group)
        if self.rank == 0:
            self.shm.unlink()

"""Image-tile sharding across the GPUs of one box: one process per GPU (torch.distributed), every
rank holds the whole flattened world, renders the 16-row bands with band % world_size == rank
(librc: rc_set_shard) into compact device buffers, and the finished bands are gathered to rank 0
over NCCL / NVLink and de-interleaved into the frame by one librc kernel (rc_compose_frame).

There is no data-path collective other than that gather.  torch is plumbing here:
process group, device memory for the gathered bands, streams.  The same code runs on the gloo
backend with CPU tensors (tests/test_multi_gpu_cpu.py), where the per-rank bands come from the
test's renderer instead of librc.
"""
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

    def __init__(self, device_index, rank=None, world_size=None, group=None):
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

    def close(self):
        self.caster.close()

    def upload_world(self, desc):
        return self.caster.upload_world(desc)  # replicated: every rank uploads the same host world

    def render(self, params):
        """Enqueue: local bands -> gather to rank 0 -> compose. Asynchronous on self.stream."""
        with torch.cuda.stream(self.stream):
            self._render(params)

    def _render(self, params):
        w, h = params.width, params.height
        self.caster.render_frame(params)
        pr = padded_rows(h, self.world)
        local = device_bytes(self.caster.device_image_ptr(rc.RC_RGBA8), pr * w * 4, self.device)
        if self.world == 1:
            self._frame = local[: h * w * 4].view(h, w, 4)
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

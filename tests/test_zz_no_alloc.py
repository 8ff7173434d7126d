"""rc_render_frame_static (include/rc.h): the frame entry point that allocates nothing -- output in the caller's device
buffer, scratch in a fixed static __device__ array of librc -- against rc_render_frame, through the C-ABI.

Collected last on purpose.  Passes alone in a fresh process and after the other tests:
it re-uploads the default world and resets shard / schedule before it starts."""
import numpy as np
import pytest

from conftest import CAMERAS, camera_dir

pytestmark = pytest.mark.gpu


def _regular(rcmod, caster, p):
    caster.render_frame(p)
    return caster.read_image().copy()


def test_static_entry_point_renders_the_same_frames(rcmod, caster, world_desc, defaults):
    import torch

    caster.upload_world(world_desc)
    caster.set_shard(0, 1)
    caster.set_option("schedule", "auto")
    pos, d, sun = defaults
    origin, yaw, pitch = CAMERAS[2]  # steep look down: most pixels hit, most cast shadow + AO rays
    frames = [
        rcmod.frame_params(640, 360, pos, d, sun, shadows=True, ao_samples=64),
        rcmod.frame_params(500, 281, origin, camera_dir(yaw, pitch), sun, shadows=True, ao_samples=20, max_steps=200),  # ragged size,
        # other AO table, other gate table, fewer pixels than the call before: stale seeds / counters would show
        rcmod.frame_params(640, 360, pos, d, sun, shadows=True, ao_samples=64),
        rcmod.frame_params(512, 512, pos, d, sun, shadows=False, ao_samples=0),  # exactly RC_STATIC_MAX_PIXELS, no AO
    ]
    assert frames[3].width * frames[3].height == rcmod.RC_STATIC_MAX_PIXELS
    want = [_regular(rcmod, caster, p) for p in frames]

    dev = torch.device("cuda:0")
    outs = [torch.full((p.height, p.width, 4), 0xAB, dtype=torch.uint8, device=dev) for p in frames]
    torch.cuda.synchronize()
    # two calls in a row (and two more), nothing between them: the calls share the scratch and are ordered by the stream
    for p, o in zip(frames, outs):
        caster.render_frame_static(p, o.data_ptr(), o.numel())
    caster.sync()
    for k, (p, o) in enumerate(zip(frames, outs)):
        got = o.cpu().numpy()
        assert np.array_equal(got, want[k]), (k, int((got != want[k]).any(axis=2).sum()))

    # ... and alternating with the allocating entry point (they share the __constant__ AO tables): still the same frames
    for k in (1, 0):
        outs[k].fill_(0xCD)
        torch.cuda.synchronize()
        assert np.array_equal(_regular(rcmod, caster, frames[k]), want[k])
        caster.render_frame_static(frames[k], outs[k].data_ptr(), outs[k].numel())
        caster.sync()
        assert np.array_equal(outs[k].cpu().numpy(), want[k]), k
    # the last rc_render_frame is still what the read-back calls describe
    assert np.array_equal(caster.read_image(), want[0])

    # a world change reaches the static entry point's tables too
    from raytracing_test_b200 import voxel_host as vh

    host = vh.World.default()
    for y in range(40, 120):  # a pillar in view, far above the terrain: other pixels, other column tops / sky gates
        host.put_block(60, y, 60, (1 << 62) | (1 << 20), 0, 0.0)
    try:
        caster.upload_world(host.describe())
        edited = _regular(rcmod, caster, frames[0])
        assert not np.array_equal(edited, want[0])
        outs[0].fill_(0)
        caster.render_frame_static(frames[0], outs[0].data_ptr(), outs[0].numel())
        caster.sync()
        assert np.array_equal(outs[0].cpu().numpy(), edited)
    finally:
        caster.upload_world(world_desc)


def test_static_entry_point_refuses_what_it_cannot_hold(rcmod, caster, world_desc, defaults):
    import torch

    caster.upload_world(world_desc)
    caster.set_shard(0, 1)
    pos, d, sun = defaults
    buf = torch.zeros(1920 * 1080 * 4, dtype=torch.uint8, device="cuda:0")
    with pytest.raises(rcmod.RcError) as e:  # beyond the fixed scratch: a clear error, nothing rendered
        caster.render_frame_static(rcmod.frame_params(1920, 1080, pos, d, sun, ao_samples=64), buf.data_ptr(), buf.numel())
    assert e.value.code == -2 and "RC_STATIC_MAX_PIXELS" in str(e.value) and "1920x1080" in str(e.value)
    with pytest.raises(rcmod.RcError) as e:
        caster.render_frame_static(rcmod.frame_params(513, 512, pos, d, sun), buf.data_ptr(), buf.numel())
    assert e.value.code == -2
    with pytest.raises(rcmod.RcError) as e:  # frame larger than the caller's buffer
        caster.render_frame_static(rcmod.frame_params(640, 360, pos, d, sun), buf.data_ptr(), 640 * 360 * 4 - 1)
    assert e.value.code == -7
    with pytest.raises(rcmod.RcError) as e:  # only the RGBA8 frame has a caller-owned output
        caster.render_frame_static(rcmod.frame_params(640, 360, pos, d, sun, hits=True), buf.data_ptr(), buf.numel())
    assert e.value.code == -8
    with pytest.raises(rcmod.RcError) as e:
        caster.render_frame_static(rcmod.frame_params(640, 360, pos, d, sun), None, buf.numel())
    assert e.value.code == -2
    torch.cuda.synchronize()
    assert int(buf.max()) == 0

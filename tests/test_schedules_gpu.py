"""Every way librc can schedule a frame on the device renders the same frame, bit for bit.

rc_debug_set_option("schedule", ...): two kernels (primary + AO over a seed list), persistent AO grid behind a programmatic
dependent launch, one fused kernel (the default), persistent fused kernel over a global atomic tile queue, the same with the
top tree levels staged in shared memory.  rc_debug_set_option("ao_opts", ...): AO sample octants that cannot meet an occupied
voxel skipped or cast.  Hit records (ao_occluded included) and RGBA8 must not depend on any of it, and one setting
of each frame is held to the oracle."""
import numpy as np
import pytest

from conftest import CAMERAS, camera_dir
from test_parity_gpu import assert_parity, build_zoo_world, render_both

pytestmark = pytest.mark.gpu


def _render(rcmod, caster, w, h, pos, d, sun, **kw):
    caster.render_frame(rcmod.frame_params(w, h, pos, d, sun, hits=True, **kw))
    return caster.read_image().copy(), caster.read_hits().copy(), caster.stats()


@pytest.fixture()
def restore_options(rcmod, caster):
    yield
    caster.set_option("schedule", rcmod.SCHED_AUTO)
    caster.set_option("ao_opts", rcmod.AO_OPT_OCTANT)


def test_schedules_and_ao_options_render_the_same_frame(ob, rcmod, caster, oracle_world, defaults, restore_options):
    pos, d, sun = defaults
    views = [(1920, 1080, pos, d), (1280, 720, (100.5, 60.25, 100.5), camera_dir(0.0, 60.0)),
             (1000, 562, (60.2, 25.9, 140.8), camera_dir(120.0, 35.0)), (640, 360, (150.5, 30.5, 140.5), camera_dir(45.0, 45.0)),
             (641, 357, (199.5, 35.5, 199.5), camera_dir(45.0, -10.0))]
    for (w, h, o, dd) in views:
        caster.set_option("schedule", rcmod.SCHED_TWO_KERNELS)
        caster.set_option("ao_opts", 0)
        ref, got = render_both(ob, rcmod, caster, oracle_world, w, h, o, dd, sun, shadows=True, ao_samples=64)
        assert_parity(ref, got, f"two kernels, plain AO, {w}x{h}")
        base = (got["rgba8"], got["hits"])
        seen = set()
        for sched in range(5):
            for opts in (0, 1):
                caster.set_option("schedule", sched)
                caster.set_option("ao_opts", opts)
                img, hits, st = _render(rcmod, caster, w, h, o, dd, sun, shadows=True, ao_samples=64)
                assert st.schedule == sched, (sched, st.schedule)
                seen.add(st.schedule)
                assert hits.tobytes() == base[1].tobytes(), (w, h, sched, opts)
                assert np.array_equal(img, base[0]), (w, h, sched, opts)
                assert st.ao_rays == ref["counters"].rays[2] and st.primary_rays == w * h
                assert st.kernel_launches == (2 if sched < 2 else 1)
        assert seen == set(range(5))
    # frames the fused schedules cannot serve fall back to two kernels: more AO steps than the window code covers, fewer
    caster.set_option("ao_opts", 1)
    for steps in (3, 9):
        caster.set_option("schedule", rcmod.SCHED_TWO_KERNELS)
        ref, got = render_both(ob, rcmod, caster, oracle_world, 480, 270, pos, d, sun, shadows=True, ao_samples=16, ao_steps=steps)
        assert_parity(ref, got, f"ao_steps {steps}")
        for sched in range(1, 5):
            caster.set_option("schedule", sched)
            img, hits, st = _render(rcmod, caster, 480, 270, pos, d, sun, shadows=True, ao_samples=16, ao_steps=steps)
            assert st.schedule == rcmod.SCHED_TWO_KERNELS
            assert hits.tobytes() == got["hits"].tobytes() and np.array_equal(img, got["rgba8"])
    # no AO at all: every schedule renders the primary pass alone
    for sched in range(5):
        caster.set_option("schedule", sched)
        img, hits, st = _render(rcmod, caster, 800, 450, pos, d, sun, shadows=True)
        if sched == 0:
            base = (img, hits)
        assert hits.tobytes() == base[1].tobytes() and np.array_equal(img, base[0]), sched


def test_schedules_on_the_material_zoo(ob, rcmod, caster, world_desc, defaults, restore_options):
    """Liquid next to the hit face, mirrors, refractors, uniform cells: the AO windows' general paths (no forced step) and
    loop B under every schedule, sharded rendering included (bands of a rank under the persistent tile queue)."""
    from raytracing_test_b200 import voxel_host as vh

    rng = np.random.default_rng(4242)
    _, _, sun = defaults
    host = build_zoo_world(vh, 6, rng)
    try:
        nodes, arrays = host.pools()
        oworld = ob.make_world(nodes, arrays, 6)
        caster.upload_world(host.describe())
        for i, (origin, yaw, pitch) in enumerate([((96.5, 40.5, 96.5), 30.0, 25.0), ((20.5, 100.5, 30.5), 60.0, 40.0),
                                                  ((64.5, 5.5, 64.5), 135.0, -35.0)]):
            caster.set_option("schedule", rcmod.SCHED_TWO_KERNELS)
            caster.set_option("ao_opts", 0)
            ref, got = render_both(ob, rcmod, caster, oworld, 480, 270, origin, camera_dir(yaw, pitch), sun, shadows=True,
                                   ao_samples=64, delta_time=0.25)
            assert_parity(ref, got, f"zoo view {i}")
            for sched in range(1, 5):
                for opts in (0, 1):
                    caster.set_option("schedule", sched)
                    caster.set_option("ao_opts", opts)
                    img, hits, _ = _render(rcmod, caster, 480, 270, origin, camera_dir(yaw, pitch), sun, shadows=True,
                                           ao_samples=64, delta_time=0.25)
                    assert hits.tobytes() == got["hits"].tobytes(), (i, sched, opts)
                    assert np.array_equal(img, got["rgba8"]), (i, sched, opts)
        # sharded: rank 1 of 3 renders its bands under every schedule; its rows equal the whole frame's
        origin, dd = (96.5, 40.5, 96.5), camera_dir(30.0, 25.0)
        caster.set_option("schedule", rcmod.SCHED_TWO_KERNELS)
        full_img, full_hits, _ = _render(rcmod, caster, 480, 270, origin, dd, sun, shadows=True, ao_samples=64)
        rows = rcmod.band_rows_of_rank(270, 1, 3)
        try:
            caster.set_shard(1, 3)
            for sched in range(5):
                caster.set_option("schedule", sched)
                img, hits, _ = _render(rcmod, caster, 480, 270, origin, dd, sun, shadows=True, ao_samples=64)
                assert np.array_equal(img[rows], full_img[rows]), sched
                assert hits.reshape(270, 480)[rows].tobytes() == full_hits.reshape(270, 480)[rows].tobytes(), sched
        finally:
            caster.set_shard(0, 1)
    finally:
        caster.upload_world(world_desc)
        host.close()

"""Parity of the CUDA path (through the C-ABI) against the CPU oracle.

Bar (BASELINE.json north_star): hit voxel, face, step counts, flags bit-exact;
RGBA8 within 1/255 per channel (in practice identical: the deterministic sin/exp
make the shading bit-exact too, which the tests report)."""
import os

import numpy as np
import pytest

from conftest import CAMERAS, camera_dir

pytestmark = pytest.mark.gpu

HIT_FIELDS = ["x", "y", "z", "steps", "shadow_steps", "face", "kind", "flags", "ao_occluded"]


def render_both(ob, rcmod, caster, oracle_world, w, h, pos, d, sun, **kw):
    ao = kw.get("ao_samples", 0)
    if ao:
        tab = caster.hemisphere_table(ao)
        ob.oracle_lib().oracle_set_hemisphere(ao, tab.ctypes.data)
    okw = dict(kw)
    fo = ob.make_frame(w, h, pos, d, sun, **okw)
    ref = ob.render(oracle_world, fo, want_f32=True)
    fp = rcmod.frame_params(w, h, pos, d, sun, hits=True, rgba32f=True, **kw)
    caster.render_frame(fp)
    got = {"rgba8": caster.read_image(), "rgba32f": caster.read_image(fmt=rcmod.RC_RGBA32F), "hits": caster.read_hits(),
           "stats": caster.stats()}
    return ref, got


def assert_parity(ref, got, label=""):
    for f in HIT_FIELDS:
        a, b = ref["hits"][f], got["hits"][f]
        if not np.array_equal(a, b):
            bad = np.flatnonzero(a != b)
            i = bad[0]
            raise AssertionError(f"{label}: hit field {f} differs at {len(bad)} pixels; first idx {i}: "
                                 f"oracle {ref['hits'][i]} cuda {got['hits'][i]}")
    diff = np.abs(ref["rgba8"].astype(np.int16) - got["rgba8"].astype(np.int16))
    assert diff.max() <= 1, f"{label}: RGBA8 differs by {diff.max()} (> 1/255) at {np.argwhere(diff > 1)[:4]}"
    return int((diff != 0).sum())


def test_config1_720p_primary_only(ob, rcmod, caster, oracle_world, defaults):
    pos, d, sun = defaults
    ref, got = render_both(ob, rcmod, caster, oracle_world, 1280, 720, pos, d, sun, shadows=False)
    n_off = assert_parity(ref, got, "config1")
    c = ref["counters"]
    st = got["stats"]
    assert st.primary_rays == 1280 * 720 and st.shadow_rays == 0
    assert st.primary_steps == c.dda_steps[0]
    assert st.hit_pixels == c.kind_pixels[ob.KIND_HIT]
    assert n_off == 0  # stronger than required: shading is bit-exact too
    assert np.array_equal(ref["rgba32f"].view(np.uint32), got["rgba32f"].view(np.uint32))


def test_config2_1080p_shadow(ob, rcmod, caster, oracle_world, defaults):
    pos, d, sun = defaults
    ref, got = render_both(ob, rcmod, caster, oracle_world, 1920, 1080, pos, d, sun, shadows=True)
    assert_parity(ref, got, "config2")
    assert got["stats"].shadow_rays == ref["counters"].rays[1]
    assert np.array_equal(ref["rgba8"], got["rgba8"])


def test_config3_1080p_ao64(ob, rcmod, caster, oracle_world, defaults):
    pos, d, sun = defaults
    ref, got = render_both(ob, rcmod, caster, oracle_world, 1920, 1080, pos, d, sun, shadows=True, ao_samples=64)
    assert_parity(ref, got, "config3")
    assert got["stats"].ao_rays == ref["counters"].rays[2]
    assert np.array_equal(ref["rgba8"], got["rgba8"])


def test_free_flight_changes_nothing(rcmod, caster, defaults):
    """Primary rays that rise above every column they can still reach finish their step budget in free flight
    (kernels.cu, loop A; FrameArgs::sky_gate): most of the remaining steps in closed form (leap_axis), the last few one by
    one.  Same frame, same hit records -- the final voxel and last face of a sky ray after its 300 steps included -- with
    free flight off, with the leap off (every step walked), and with both on; and both do run for most sky rays of the
    default view."""
    pos, d, sun = defaults
    # ... and with the column regions of loop A (the whole 16 x 16 column above the terrain as one empty region, found
    # in the tops table instead of the tree) off, alone and together with free flight: "plain" is the tree-only walk
    modes = {"on": {}, "walk": {"leap": False}, "nocol": {"column_regions": False}, "nodrop": {"drop_gates": False},
             "plain": {"free_flight": False, "column_regions": False}, "off": {"free_flight": False}}
    frames = {}
    for name, kw in modes.items():
        caster.render_frame(rcmod.frame_params(1920, 1080, pos, d, sun, shadows=True, ao_samples=16, hits=True, rgba32f=True, **kw))
        frames[name] = (caster.read_image().copy(), caster.read_image(fmt=rcmod.RC_RGBA32F).copy(), caster.read_hits().copy(),
                        caster.stats())
    b = frames["plain"]
    for name in ("on", "walk", "nocol", "nodrop", "off"):
        a = frames[name]
        assert a[2].tobytes() == b[2].tobytes(), name
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1].view(np.uint32), b[1].view(np.uint32)), name
        assert a[3].primary_steps == b[3].primary_steps
        sky = int((a[2]["kind"] == 0).sum())  # RC_KIND_SKY
        if name != "off":
            assert a[3].free_flight_rays > 0.5 * sky, (name, a[3].free_flight_rays, sky)
    assert b[3].free_flight_rays == 0 and frames["off"][3].free_flight_rays == 0
    # rays that step DOWN but not as far as anything they can still reach (the grazing rays under the horizon, 300 steps
    # each before) are set free too: sky_drop_distance
    assert frames["on"][3].free_flight_rays > frames["nodrop"][3].free_flight_rays + 80_000
    # Without hit records (and without the position debug output) nobody can see where a free ray ends, and the tail is not
    # worked out at all -- unless the looked-at voxel lies exactly max_steps steps from the camera's, the one case in which
    # the highlight test (low_res.frag:347-350) could fire on a ray that ran out of steps.  Same frame in every case.
    sky_px = np.flatnonzero(frames["on"][2]["kind"] == 0)
    end = frames["on"][2][sky_px[len(sky_px) // 2]]           # where some sky ray ends: 300 steps from the camera's voxel
    for la in ((1 << 30, 1 << 30, 1 << 30), (int(end["x"]), int(end["y"]), int(end["z"])), (int(end["x"]) + 1, int(end["y"]), int(end["z"]))):
        imgs = []
        for kw in ({"hits": True}, {}, {"always_tail": True}, {"free_flight": False}):
            caster.render_frame(rcmod.frame_params(1920, 1080, pos, d, sun, shadows=True, rgba32f=True, looking_at=la, **kw))
            imgs.append((caster.read_image().copy(), caster.read_image(fmt=rcmod.RC_RGBA32F).copy()))
        for im in imgs[1:]:
            assert np.array_equal(im[0], imgs[0][0]) and np.array_equal(im[1].view(np.uint32), imgs[0][1].view(np.uint32)), la
    # (the middle one really is highlighted)
    caster.render_frame(rcmod.frame_params(1920, 1080, pos, d, sun, shadows=True, hits=True,
                                           looking_at=(int(end["x"]), int(end["y"]), int(end["z"]))))
    assert int((caster.read_hits()["kind"] == 4).sum()) > 0  # RC_KIND_HIGHLIGHT
    # other step budgets (the gate table is rebuilt per budget), cameras looking down / high above the terrain / along an
    # axis (absDelta = inf), all yaw / pitch combinations of the parity cameras
    cams = [(64, pos, d), (1000, pos, d), (1024, pos, d), (300, (100.5, 120.25, 100.5), camera_dir(30.0, -50.0)),
            (300, (60.5, 90.5, 20.5), camera_dir(-120.0, 20.0)), (300, (35.0, 50.0, 35.0), (0.0, 1.0, 0.0)),
            (300, (35.5, 80.5, 35.5), (1.0, 0.0, 0.0)), (300, (100.0, 75.0, 100.0), (0.0, 0.0, 1.0)),
            (500, (120.5, 60.5, 90.5), camera_dir(200.0, 5.0))]
    cams += [(300, c[0], camera_dir(c[1], c[2])) for c in CAMERAS]
    for steps, o, dd in cams:
        got = []
        for kw in modes.values():
            caster.render_frame(rcmod.frame_params(640, 360, o, dd, sun, shadows=True, hits=True, max_steps=steps, **kw))
            got.append((caster.read_image().copy(), caster.read_hits().copy()))
        for g in got[:-1]:
            assert g[1].tobytes() == got[-1][1].tobytes(), (steps, o, dd)
            assert np.array_equal(g[0], got[-1][0])


def test_sky_gates_follow_the_world(ob, rcmod, caster, world_desc, defaults):
    """The free-flight gates are rebuilt from the tree after every world change.  Worlds made for them: empty (every ray that
    steps up is free at once), a few blocks floating high above bare ground, then edits through rc_update_world that put a
    block into empty sky and take it away again -- every frame against the oracle (hit records bit-exact, shadows included:
    shadow rays use the gates too), and identical with free flight off."""
    from raytracing_test_b200 import voxel_host as vh

    _, _, sun = defaults
    cams = [((60.5, 40.5, 60.5), camera_dir(45.0, -35.0)), ((60.5, 40.5, 60.5), camera_dir(-100.0, -60.0)),  # pitch < 0: up
            ((100.5, 130.5, 90.5), camera_dir(200.0, 40.0)), ((20.5, 5.5, 20.5), camera_dir(45.0, -10.0))]

    def check(host, label):
        nodes, arrays = host.pools()
        ow = ob.make_world(nodes, arrays, 6)
        n_hit = 0
        for i, (o, dd) in enumerate(cams):
            ref = ob.render(ow, ob.make_frame(480, 270, o, dd, sun, shadows=True))
            frames = []
            for ff in (True, False):
                caster.render_frame(rcmod.frame_params(480, 270, o, dd, sun, shadows=True, hits=True, free_flight=ff))
                frames.append((caster.read_image().copy(), caster.read_hits().copy()))
            assert frames[0][1].tobytes() == ref["hits"].tobytes(), (label, i)
            assert frames[1][1].tobytes() == ref["hits"].tobytes(), (label, i)
            assert np.array_equal(frames[0][0], frames[1][0]) and np.array_equal(frames[0][0], ref["rgba8"])
            n_hit += int((ref["hits"]["kind"] == 1).sum())
        return n_hit

    host = vh.World(6)
    try:
        caster.upload_world(host.describe())
        assert check(host, "empty world") == 0
        for x in range(0, 160, 2):  # bare ground: a strip of 1-voxel columns, so that shadow rays have something to start from
            for z in range(40, 120, 2):
                host.put_block(x, 3, z, color=0x3FF00000000, flags=0)
        host.put_block(90, 100, 90, color=0x1FFFFF, flags=0)
        host.put_block(64, 120, 64, color=0x7FE00000, flags=0, level=5)   # a 4^3 cell
        host.put_block(30, 200, 100, color=0x1FFFFF, flags=0)
        caster.upload_world(host.describe())
        before = check(host, "floating blocks")
        assert before > 0
        host.put_block(78, 58, 78, color=0x1FFFFF, flags=0, level=5)       # into the empty sky, in the middle of the first view
        info = caster.update_world(host.describe(), None, None)
        after = check(host, "after put")
        assert after > before
        host.put_block(95, 75, 95, color=0x1FFFFF, flags=0)
        caster.update_world(host.describe(), None, None)
        check(host, "after second put")
    finally:
        host.close()
        caster.upload_world(world_desc)


def _spike_world(vh, depth=6, off=0):
    """Bare ground with a forest of one-voxel-wide towers of pseudo-random heights: rays that rise between them graze
    tower tops at every distance and slope -- what the sky gates' slope / ring arithmetic has to get right."""
    host = vh.World(depth)
    for x in range(0, 400, 3):
        for z in range(0, 400, 3):
            host.put_block(off + x, 2, off + z, color=0x3FF00000000, flags=0)
    state = 12345
    for x in range(4, 400, 7):
        for z in range(5, 400, 7):
            state = (state * 1103515245 + 12345) & 0x7FFFFFFF
            top = 4 + (state >> 8) % 140
            if (state >> 3) % 5 == 0:
                top = 4 + (state >> 12) % 20      # many low ones, so that the gates near the camera stay low
            for y in range(max(3, top - 2), top + 1):   # the top three voxels are enough for rays from below / the side
                host.put_block(off + x, y, off + z, color=0x1FFFFF, flags=0)
            for y in range(3, top - 2, 5):              # and a sparse stem
                host.put_block(off + x, y, off + z, color=0x7FE00000, flags=0)
    return host


def test_sky_gates_on_a_forest_of_towers(ob, rcmod, caster, world_desc, defaults):
    """Adversarial world for free flight (primary and shadow rays): hit records bit-exact against the oracle and identical
    with free flight off, from cameras low between the towers looking up at many pitches, and with a low sun."""
    from raytracing_test_b200 import voxel_host as vh

    host = _spike_world(vh)
    try:
        caster.upload_world(host.describe())
        nodes, arrays = host.pools()
        ow = ob.make_world(nodes, arrays, 6)
        f = np.float32
        suns = [defaults[2], tuple(float(v) for v in (np.array([3.0, 0.35, 1.0], dtype=f) / f(np.sqrt(f(10.1225)))))]
        cams = [((150.5, 6.5, 150.5), camera_dir(45.0, -8.0)), ((150.5, 6.5, 150.5), camera_dir(20.0, -25.0)),
                ((201.5, 30.5, 180.5), camera_dir(-60.0, -3.0)), ((120.5, 12.5, 260.5), camera_dir(135.0, -15.0)),
                ((180.5, 60.5, 170.5), camera_dir(80.0, -45.0)), ((160.5, 4.5, 140.5), camera_dir(5.0, -1.5))]
        free = 0
        for i, (o, dd) in enumerate(cams):
            sun = suns[i % 2]
            ref = ob.render(ow, ob.make_frame(480, 270, o, dd, sun, shadows=True))
            frames = []
            for ff in (True, False):
                caster.render_frame(rcmod.frame_params(480, 270, o, dd, sun, shadows=True, hits=True, free_flight=ff))
                frames.append((caster.read_image().copy(), caster.read_hits().copy()))
                if ff:
                    free += caster.stats().free_flight_rays
            assert frames[0][1].tobytes() == ref["hits"].tobytes(), i
            assert frames[1][1].tobytes() == ref["hits"].tobytes(), i
            assert np.array_equal(frames[0][0], ref["rgba8"]) and np.array_equal(frames[1][0], ref["rgba8"])
        assert free > 0
        # AO on tower faces (thin columns: every window holds several faces' worth of edges), both sample counts
        for ao, (o, dd) in ((64, cams[1]), (20, cams[4]), (64, cams[5])):
            tab = caster.hemisphere_table(ao)
            ob.oracle_lib().oracle_set_hemisphere(ao, tab.ctypes.data)
            ref = ob.render(ow, ob.make_frame(320, 180, o, dd, suns[0], shadows=True, ao_samples=ao))
            caster.render_frame(rcmod.frame_params(320, 180, o, dd, suns[0], shadows=True, ao_samples=ao, hits=True))
            assert caster.read_hits().tobytes() == ref["hits"].tobytes(), (ao, o)
            assert np.array_equal(caster.read_image(), ref["rgba8"])
        # other step budgets: the gate tables (reach in blocks, rings) are rebuilt for each
        for steps in (40, 120, 700, 1024):
            for o, dd in (cams[0], cams[3], cams[5]):
                ref = ob.render(ow, ob.make_frame(320, 180, o, dd, suns[1], shadows=True, max_steps=steps, shadow_steps=min(steps, 75)))
                caster.render_frame(rcmod.frame_params(320, 180, o, dd, suns[1], shadows=True, hits=True, max_steps=steps,
                                                       shadow_steps=min(steps, 75)))
                assert caster.read_hits().tobytes() == ref["hits"].tobytes(), (steps, o)
    finally:
        host.close()
        caster.upload_world(world_desc)


def test_sky_gates_on_a_forest_of_towers_depth7(ob, rcmod, caster, world_desc, defaults):
    """The same forest in a depth-7 tree (4096^3 torus, column blocks of 64 x 64), planted beyond x, z = 1024."""
    from raytracing_test_b200 import voxel_host as vh

    host = _spike_world(vh, 7, 1100)
    try:
        caster.upload_world(host.describe())
        nodes, arrays = host.pools()
        ow = ob.make_world(nodes, arrays, 7)
        sun = defaults[2]
        for o, dd in (((1250.5, 6.5, 1250.5), camera_dir(45.0, -8.0)), ((1301.5, 30.5, 1280.5), camera_dir(-60.0, -3.0)),
                      ((1280.5, 60.5, 1270.5), camera_dir(80.0, -45.0)), ((1260.5, 4.5, 1240.5), camera_dir(5.0, -1.5))):
            ref = ob.render(ow, ob.make_frame(480, 270, o, dd, sun, shadows=True))
            for ff in (True, False):
                caster.render_frame(rcmod.frame_params(480, 270, o, dd, sun, shadows=True, hits=True, free_flight=ff))
                assert caster.read_hits().tobytes() == ref["hits"].tobytes(), (o, ff)
                assert np.array_equal(caster.read_image(), ref["rgba8"])
    finally:
        host.close()
        caster.upload_world(world_desc)


def test_cuda_frame_equals_the_reference_shader_frame(ob, rcmod, caster, ref_world, defaults):
    """No oracle in between: the CUDA frame against src/shaders/low_res.frag itself, compiled as C++
    (oracle/_ref/libref_shader.so, see tests/test_ref_shader_cpu.py) -- FragColor float for float at config 2's
    1920x1080, hit voxels / faces against the state the shader's globals are left in."""
    try:
        shader = ob.RefShader(ref_world.nodes, ref_world.arrays, oracle_math=True)
    except FileNotFoundError:
        pytest.skip("oracle/_ref/libref_shader.so not built (needs /root/reference once)")
    pos, d, sun = defaults
    w, h = 1920, 1080
    look, _, _ = ref_world.pick(pos, d, 30)
    look = tuple(int(v) for v in look)
    ref = shader.render(w, h, pos, d, sun, looking_at=look, mouse_pos=(960.0, 540.0), delta_time=0.5)
    caster.render_frame(rcmod.frame_params(w, h, pos, d, sun, hits=True, rgba32f=True, shadows=True, looking_at=look,
                                           mouse_pos=(960.0, 540.0), delta_time=0.5))
    img = caster.read_image(fmt=rcmod.RC_RGBA32F)
    hits = caster.read_hits().reshape(h, w)
    assert np.array_equal(ref["rgba32f"].view(np.uint32), img.view(np.uint32))
    primary_only = (hits["shadow_steps"] == 0) & np.isin(hits["kind"], [0, 1, 4])
    assert primary_only.sum() > w * h // 2
    for k, n in enumerate("xyz"):
        assert np.array_equal(ref["state"]["round"][..., k][primary_only], hits[n][primary_only])
    stepped = primary_only & (hits["steps"] > 0)
    assert np.array_equal(ref["state"]["last_hit"][stepped], hits["face"][stepped])
    # and the debug outputs: pos.exact (one rounded product per DDA step) and pos.round of EVERY pixel
    for mode in (1, 2):
        ref = shader.render(w, h, pos, d, sun, render_pos_data=mode, want_state=False)
        caster.render_frame(rcmod.frame_params(w, h, pos, d, sun, rgba32f=True, shadows=True, render_pos_data=mode))
        assert np.array_equal(ref["rgba32f"].view(np.uint32), caster.read_image(fmt=rcmod.RC_RGBA32F).view(np.uint32))


@pytest.mark.parametrize("cam", range(len(CAMERAS)))
def test_camera_sweep(ob, rcmod, caster, oracle_world, defaults, cam):
    origin, yaw, pitch = CAMERAS[cam]
    _, _, sun = defaults
    d = camera_dir(yaw, pitch)
    ref, got = render_both(ob, rcmod, caster, oracle_world, 640, 360, origin, d, sun, shadows=True, ao_samples=20,
                           delta_time=1.25, mouse_pos=(320.0, 180.0))
    assert_parity(ref, got, f"camera {cam}")


def test_looking_at_highlight_and_debug_modes(ob, rcmod, caster, oracle_world, defaults, ref_world):
    pos, d, sun = defaults
    # the block under a pixel a little below the screen centre (the host would pass castRayFromCam(30).pos)
    probe = ob.render(oracle_world, ob.make_frame(320, 200, pos, d, sun))["hits"][60 * 320 + 160]
    assert probe["kind"] == ob.KIND_HIT
    look = (int(probe["x"]), int(probe["y"]), int(probe["z"]))
    ref, got = render_both(ob, rcmod, caster, oracle_world, 320, 200, pos, d, sun, looking_at=look)
    assert_parity(ref, got, "highlight")
    assert (got["hits"]["kind"] == ob.KIND_HIGHLIGHT).sum() > 0
    # and the voxel the reference's own CPU picker reports (src/main.cpp:81,89), hit or miss
    look, _, _ = ref_world.pick(pos, d, 30)
    ref, got = render_both(ob, rcmod, caster, oracle_world, 320, 200, pos, d, sun, looking_at=tuple(int(v) for v in look))
    assert_parity(ref, got, "highlight(castRayFromCam)")
    for mode in (1, 2):
        ref, got = render_both(ob, rcmod, caster, oracle_world, 320, 200, pos, d, sun, render_pos_data=mode)
        assert_parity(ref, got, f"debug mode {mode}")
        assert np.array_equal(ref["rgba32f"].view(np.uint32), got["rgba32f"].view(np.uint32))


def test_ragged_resolutions(ob, rcmod, caster, oracle_world, defaults):
    pos, d, sun = defaults
    for (w, h) in [(1, 1), (7, 3), (33, 17), (130, 75), (257, 129)]:
        ref, got = render_both(ob, rcmod, caster, oracle_world, w, h, pos, d, sun, shadows=True, ao_samples=8)
        assert_parity(ref, got, f"{w}x{h}")


def test_step_limits(ob, rcmod, caster, oracle_world, defaults):
    pos, d, sun = defaults
    for (ms, ss, aos) in [(0, 0, 0), (1, 1, 1), (17, 3, 2), (300, 75, 32), (1000, 200, 9)]:
        ref, got = render_both(ob, rcmod, caster, oracle_world, 200, 120, pos, d, sun, max_steps=ms, shadow_steps=ss,
                               ao_samples=16, ao_steps=aos)
        assert_parity(ref, got, f"limits {ms}/{ss}/{aos}")


def test_device_lookup_matches_reference_getblock(ob, rcmod, caster, ref_world):
    rng = np.random.default_rng(7)
    xyz = np.concatenate([
        rng.integers(0, 256, size=(200000, 3)),
        rng.integers(0, 768, size=(50000, 3)),
        np.array([[35, 20, 35], [10, 100, 10], [21, 9, 201], [24, 9, 201], [35, 50, 35]]),
    ]).astype(np.int32)
    words, _ = caster.lookup_device(xyz)
    refb = ref_world.get_blocks(xyz)
    solid = words != 0
    assert np.array_equal(solid, refb[:, 0] != np.uint64(0xFFFFFFFFFFFFFFFF))
    hw, _, leaves = rcmod.lookup_host(xyz, which=0)
    assert np.array_equal(hw, words)
    leaf = leaves.view(np.uint64).reshape(-1, 2)
    assert np.array_equal(leaf[solid], refb[solid])


def test_device_detmath_matches_oracle(ob, caster):
    L = ob.oracle_lib()
    x = np.concatenate([np.linspace(-100, 100, 20001), np.linspace(-3000, 3000, 20001),
                        np.array([0.0, -0.0, 1e-30, 88.5, -87.5, 1e7, np.inf, -np.inf])]).astype(np.float32)
    s, e = caster.detmath(x)
    so = np.array([L.oracle_sinf(float(v)) for v in x], dtype=np.float32)
    eo = np.array([L.oracle_expf(float(v)) for v in x], dtype=np.float32)
    assert np.array_equal(s.view(np.uint32), so.view(np.uint32))
    assert np.array_equal(e.view(np.uint32), eo.view(np.uint32))


def test_gpu_pick_matches_reference_castRayFromCam(rcmod, caster, ref_world, defaults):
    """rc_pick against the reference's own RAY_CASTER::castRayFromCam (src/ray_caster.cpp:54-87,
    compiled unmodified into oracle/_ref): pos, lastPos and steps left, hit or miss."""
    pos, d, _ = defaults
    cases = [(pos, d, 30)]
    rng = np.random.default_rng(11)
    for (origin, yaw, pitch) in CAMERAS:
        cases.append((np.array(origin, dtype=np.float32), camera_dir(yaw, pitch), 30))
    for _ in range(200):
        o = np.array([rng.uniform(1, 199), rng.uniform(15, 70), rng.uniform(1, 199)], dtype=np.float32)
        cases.append((o, camera_dir(rng.uniform(-180, 180), rng.uniform(-80, 80)), int(rng.integers(1, 120))))
    hits = 0
    for (o, dd, n) in cases:
        rp, rl, rs = ref_world.pick(o, dd, n)
        gp, gl, gs = caster.pick(o, dd, n)
        assert (tuple(gp), tuple(gl), gs) == (tuple(rp), tuple(rl), rs), f"pick from {o} dir {dd} steps {n}"
        hits += rs > 0
    assert hits > 20


def test_incremental_world_update_edit_scenario_on_device():
    """rc_update_world end to end (tests/edit_scenario.py gpu, own process: it edits the reference's global
    tree): after every round of the reference's putBlock / deleteBlock + updateSsboData(), device lookups
    equal the shader's getBlock over the SSBO view and a frame equals the oracle's, record for record."""
    import os
    import subprocess
    import sys

    here = os.path.dirname(os.path.abspath(__file__))
    r = subprocess.run([sys.executable, os.path.join(here, "edit_scenario.py"), "gpu"], capture_output=True, text=True,
                       timeout=600)
    assert r.returncode == 0 and "edit scenario ok" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


def test_world_blob_and_ppm_on_device(rcmod, caster, world_desc, defaults, tmp_path):
    """rc_save_world -> rc_load_world renders the same frame byte for byte (no host pools, no flatten);
    rc_write_ppm holds the RGBA8 frame top row first."""
    pos, d, sun = defaults
    p = rcmod.frame_params(640, 360, pos, d, sun, shadows=True, ao_samples=16, hits=True)
    caster.render_frame(p)
    img, hits = caster.read_image(), caster.read_hits()
    path = str(tmp_path / "default.rcpw")
    caster.save_world(path)
    try:
        info = caster.load_world(path)
        assert info.packed_bytes == os.path.getsize(path) - 64 and info.ref_bytes == 0
        caster.render_frame(p)
        assert np.array_equal(caster.read_image(), img) and caster.read_hits().tobytes() == hits.tobytes()
        ppm = str(tmp_path / "frame.ppm")
        caster.write_ppm(ppm)
        raw = open(ppm, "rb").read()
        header = b"P6\n640 360\n255\n"
        assert raw.startswith(header) and len(raw) == len(header) + 640 * 360 * 3
        pix = np.frombuffer(raw[len(header):], dtype=np.uint8).reshape(360, 640, 3)
        assert np.array_equal(pix, img[::-1, :, :3])
    finally:
        caster.upload_world(world_desc)


def test_headless_cpp_host_drives_the_c_abi(rcmod, caster, defaults, ref_world, tmp_path):
    """host/rc_headless (the reference's main loop against include/rc.h, C++, no Python): runs, clicks
    (rc_update_world), picks on the device, writes PPM frames; its first frame equals the same frame asked
    for through the ctypes binding."""
    import subprocess

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "host", "rc_headless")
    assert os.path.exists(exe), "host/rc_headless not built (make -C raytracing_test_b200/csrc)"
    r = subprocess.run([exe, "--frames", "14", "--size", "640x360", "--ao", "16", "--out", str(tmp_path) + "/f", "--edit",
                        "--gpu-pick"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "click at frame 4" in r.stdout and "render time:" in r.stdout
    pos, d, sun = defaults
    look, _, _ = ref_world.pick(pos, d, 30)
    caster.render_frame(rcmod.frame_params(640, 360, pos, d, sun, shadows=True, ao_samples=16, mouse_pos=(320.0, 180.0),
                                           looking_at=tuple(int(v) for v in look)))
    img = caster.read_image()
    raw = open(str(tmp_path) + "/f0000.ppm", "rb").read()
    header = b"P6\n640 360\n255\n"
    pix = np.frombuffer(raw[len(header):], dtype=np.uint8).reshape(360, 640, 3)
    assert np.array_equal(pix, img[::-1, :, :3])
    assert len([f for f in os.listdir(tmp_path) if f.endswith(".ppm")]) == 14


def test_headless_host_with_device_generated_world(tmp_path):
    """--gen-device: rc_generate_world instead of the host's genWorld + upload; every frame of the run
    (edits included: the host tree is built at the first click) is byte-identical to the host-built run."""
    import subprocess

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "host", "rc_headless")
    assert os.path.exists(exe), "host/rc_headless not built (make -C raytracing_test_b200/csrc)"
    common = ["--frames", "14", "--size", "480x270", "--ao", "16", "--edit", "--gpu-pick"]
    for tag, extra in (("h", []), ("d", ["--gen-device"])):
        r = subprocess.run([exe] + common + ["--out", f"{tmp_path}/{tag}"] + extra, capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        assert "click at frame 4" in r.stdout
        assert ("generated chunk on the device!" in r.stdout) == bool(extra)
    for f in range(14):
        a = open(f"{tmp_path}/h{f:04d}.ppm", "rb").read()
        b = open(f"{tmp_path}/d{f:04d}.ppm", "rb").read()
        assert a == b, f"frame {f} differs between the host-built and the device-generated world"


def test_pipelined_readback_delivers_every_frame_intact(rcmod, caster, defaults):
    """rc_read_image_async / rc_wait_image: frames N and N+1 are in flight together (two device frame
    buffers, separate copy stream); every delivered frame is byte-identical to the synchronous read of the
    same frame, in order, also when the resolution changes mid-stream and when reads are skipped."""
    import torch

    pos, d, sun = defaults
    cams = [(pos, camera_dir(yaw, 20.0)) for yaw in (45.0, 60.0, 75.0, 90.0, 105.0, 120.0, 135.0)]
    sizes = [(640, 360)] * 4 + [(800, 450)] * 3
    want = []
    for (o, dd), (w, h) in zip(cams, sizes):
        caster.render_frame(rcmod.frame_params(w, h, o, dd, sun, shadows=True, ao_samples=16))
        want.append(caster.read_image().copy())
    bufs = [torch.empty((450, 800, 4), dtype=torch.uint8, pin_memory=True).numpy() for _ in range(2)]
    got, prev = [], None
    for i, ((o, dd), (w, h)) in enumerate(zip(cams, sizes)):
        caster.render_frame(rcmod.frame_params(w, h, o, dd, sun, shadows=True, ao_samples=16))
        view = bufs[i % 2].reshape(-1)[: w * h * 4].reshape(h, w, 4)
        t = caster.read_image_async(view)
        with pytest.raises(rcmod.RcError):  # one queued read-back per frame
            caster.read_image_async(view)
        if prev is not None:
            caster.wait_image(prev[0])
            got.append(prev[1].copy())
        prev = (t, view)
    caster.wait_image(prev[0])
    got.append(prev[1].copy())
    assert len(got) == len(want)
    for i, (a, b) in enumerate(zip(want, got)):
        assert np.array_equal(a, b), f"pipelined frame {i} differs from its synchronous read"
    # a frame rendered but never read asynchronously does not disturb the next ones; the synchronous read still works
    caster.render_frame(rcmod.frame_params(640, 360, cams[0][0], cams[0][1], sun, shadows=True, ao_samples=16))
    assert np.array_equal(caster.read_image(), want[0])
    with pytest.raises(rcmod.RcError):
        caster.wait_image(7)


def build_zoo_world(vh, depth, rng):
    """A world the terrain never produces: every material flag combination (src/globals.hpp:68-74) as single
    voxels, as 4^3 / 16^3 / 64^3 uniform cells, uniform cells split by later single-voxel puts, liquid next to
    opaque, mirrors facing mirrors, non-zero metadata.  Built with the host stand-in's putBlock."""
    host = vh.World(depth)
    voxel = depth
    flag_sets = [0x0, 0x2, 0x4, 0x10, 0x14, 0x6, 0x12, 0x16]
    col = lambda: int(vh.load_library().vh_rgb(int(rng.integers(0, 256)), int(rng.integers(0, 256)), int(rng.integers(0, 256))))
    for _ in range(2):  # 64^3
        x, y, z = (int(v) for v in rng.integers(0, 3, 3) * 64)
        host.put_block(x, y + 64, z, col(), int(rng.choice([0x0, 0x4, 0x14])), 0.0, voxel - 3)
    for _ in range(14):  # 16^3
        x, y, z = (int(v) for v in rng.integers(0, 12, 3) * 16)
        host.put_block(x, y, z, col(), int(rng.choice(flag_sets)), float(rng.choice([0.0, 0.5])), voxel - 2)
    for _ in range(150):  # 4^3
        x, y, z = (int(v) for v in rng.integers(0, 48, 3) * 4)
        host.put_block(x, y, z, col(), int(rng.choice(flag_sets)), 0.0, voxel - 1)
    for _ in range(9000):  # single voxels, many of them inside the uniform cells above (splits them)
        x, y, z = (int(v) for v in rng.integers(0, 192, 3))
        host.put_block(x, y, z, col(), int(rng.choice(flag_sets, p=[0.6, 0.08, 0.08, 0.04, 0.1, 0.04, 0.03, 0.03])), 0.0, voxel)
    # a floor so that most rays end on something and cast shadow + AO rays
    for x in range(0, 192, 4):
        for z in range(0, 192, 4):
            host.put_block(x, 0, z, col(), 0x0, 0.0, voxel - 1)
    return host


@pytest.mark.parametrize("depth", [6, 7])
def test_material_zoo_world(ob, rcmod, caster, world_desc, defaults, depth):
    from raytracing_test_b200 import voxel_host as vh

    rng = np.random.default_rng(900 + depth)
    _, _, sun = defaults
    host = build_zoo_world(vh, depth, rng)
    try:
        nodes, arrays = host.pools()
        oworld = ob.make_world(nodes, arrays, depth)
        caster.upload_world(host.describe())
        pts = rng.integers(0, 200, size=(50000, 3)).astype(np.int32)
        words, _ = caster.lookup_device(pts)
        refb = ob.get_blocks(oworld, pts)
        assert np.array_equal(words != 0, refb[:, 0] != np.uint64(0xFFFFFFFFFFFFFFFF))
        seen = {"refl": 0, "refr": 0, "shadow": 0, "inside": 0, "ao": 0}
        views = [((96.5, 40.5, 96.5), 30.0, 25.0), ((20.5, 100.5, 30.5), 60.0, 40.0), ((180.5, 8.5, 10.5), -45.0, 5.0),
                 ((-10.5, 60.5, 96.5), 90.0, 15.0), ((96.5, 190.5, 96.5), 10.0, 80.0), ((100.2, 70.7, 33.3), 200.0, -20.0),
                 ((64.5, 5.5, 64.5), 135.0, -35.0), ((8.5, 72.5, 8.5), 45.0, 0.0)]
        # plus cameras placed inside solid / liquid voxels picked from the world itself
        solid = pts[words != 0]
        for k in range(4):
            p = solid[k * 7]
            views.append(((float(p[0]) + 0.5, float(p[1]) + 0.5, float(p[2]) + 0.5), 33.0 * k, 10.0))
        for i, (origin, yaw, pitch) in enumerate(views):
            ref, got = render_both(ob, rcmod, caster, oworld, 480, 270, origin, camera_dir(yaw, pitch), sun, shadows=True,
                                   ao_samples=24, delta_time=0.75)
            assert_parity(ref, got, f"zoo depth {depth} view {i}")
            assert np.array_equal(ref["rgba8"], got["rgba8"]), f"zoo depth {depth} view {i}: shading not bit-exact"
            fl = got["hits"]["flags"]
            seen["refl"] += int(((fl & 0x02) != 0).sum())
            seen["refr"] += int(((fl & 0x04) != 0).sum())
            seen["shadow"] += int(((fl & 0x01) != 0).sum())
            seen["inside"] += int((got["hits"]["kind"] == ob.KIND_CAMERA_INSIDE).sum())
            seen["ao"] += int(got["hits"]["ao_occluded"].astype(np.int64).sum())
        # the views really exercise what the world was built for
        assert seen["refl"] > 1000 and seen["refr"] > 1000 and seen["shadow"] > 1000 and seen["ao"] > 10000, seen
        # more AO steps than the window kernel is specialised for, and ragged step limits, on this world too
        for (ms, ss, aos) in [(300, 75, 9), (40, 10, 3)]:
            ref, got = render_both(ob, rcmod, caster, oworld, 320, 180, views[0][0], camera_dir(30.0, 25.0), sun, shadows=True,
                                   ao_samples=16, ao_steps=aos, max_steps=ms, shadow_steps=ss)
            assert_parity(ref, got, f"zoo depth {depth} limits {ms}/{ss}/{aos}")
    finally:
        caster.upload_world(world_desc)
        host.close()

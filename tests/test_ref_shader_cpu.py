"""The oracle held to the reference's OWN fragment shader.

oracle/_ref/libref_shader.so is src/shaders/low_res.frag itself -- the only implementation of the render path the
reference has -- compiled as C++ (oracle/ref_shader.cpp, oracle/shim/glsl/; three declaration-level rewrites, no
expression touched) and run one instance per pixel on the host cores, over the pools the reference's own
putBlock / genWorld built.  The restated oracle (oracle/oracle_render.cpp), which every GPU parity test and golden
fixture hangs on, has to reproduce it FLOAT FOR FLOAT: FragColor of every pixel, the debug outputs (pos.exact after
every DDA step of the primary loops, pos.round), and the state the shader's globals are left in.

What stays outside this pin (no GL driver on this box): how a GLSL compiler would contract / reorder the float
arithmetic and its sin / exp / normalize implementations -- the shim uses IEEE binary32 in source order and, for
bit-for-bit colour comparisons, the oracle's sin / exp (the libm variant is held to 1/255 below)."""
import os

import numpy as np
import pytest

from conftest import CAMERAS, camera_dir

FAR = (1 << 30, 1 << 30, 1 << 30)


@pytest.fixture(scope="module")
def shader(ob, ref_world):
    try:
        return ob.RefShader(ref_world.nodes, ref_world.arrays, oracle_math=True)
    except FileNotFoundError:
        pytest.skip("oracle/_ref/libref_shader.so not built (needs /root/reference once)")


def both(ob, shader, oworld, w, h, pos, d, sun, **kw):
    """(shader result, oracle result) for one frame; the oracle with shadows on and AO off = the live shader."""
    ref = shader.render(w, h, pos, d, sun, **kw)
    okw = dict(kw)
    frame = ob.make_frame(w, h, pos, d, sun, shadows=True, ao_samples=0, **okw)
    orc = ob.render(oworld, frame, want_f32=True)
    return ref, orc


def assert_same_frame(ref, orc, label, h, w):
    a, b = ref["rgba32f"].view(np.uint32), orc["rgba32f"].view(np.uint32)
    if not np.array_equal(a, b):
        bad = np.argwhere((a != b).any(axis=2))
        y, x = bad[0]
        raise AssertionError(f"{label}: FragColor differs at {len(bad)} pixels; first (x={x}, y={y}): shader "
                             f"{ref['rgba32f'][y, x]} oracle {orc['rgba32f'][y, x]} hit {orc['hits'][y * w + x]}")
    hits, st = orc["hits"].reshape(h, w), ref["state"]
    # pixels whose shader run ended before the shadow ray: the globals still hold the primary result
    primary_only = (hits["shadow_steps"] == 0) & np.isin(hits["kind"], [0, 1, 4])
    for k, n in enumerate("xyz"):
        assert np.array_equal(st["round"][..., k][primary_only], hits[n][primary_only]), f"{label}: pos.round.{n}"
    stepped = primary_only & (hits["steps"] > 0)
    assert np.array_equal(st["last_hit"][stepped], hits["face"][stepped]), f"{label}: lastHit"
    assert np.array_equal(st["reflected"][primary_only] != 0, (hits["flags"][primary_only] & 2) != 0), f"{label}: rayReflected"
    sky = hits["kind"] == 0
    assert (st["block_empty"][sky & (hits["shadow_steps"] == 0)] == 1).all()
    return int(primary_only.sum())


@pytest.mark.parametrize("w,h", [(1280, 720), (1920, 1080)])
def test_oracle_reproduces_the_reference_shader_on_the_baseline_frames(ob, shader, oracle_world, defaults, w, h):
    pos, d, sun = defaults
    ref, orc = both(ob, shader, oracle_world, w, h, pos, d, sun)
    assert ref["fragments"] == w * h
    n = assert_same_frame(ref, orc, f"{w}x{h}", h, w)
    assert n > w * h // 2
    kinds = np.bincount(orc["hits"]["kind"], minlength=6)
    assert kinds[0] > 0 and kinds[1] > 0 and (orc["hits"]["flags"] & 1).sum() > 1000  # sky, terrain, shadowed pixels


def test_debug_outputs_carry_every_dda_step(ob, shader, oracle_world, defaults):
    """renderPosData 1 / 2 (low_res.frag:335-341): FragColor = pos.exact / pos.round when the primary loops exit.
    pos.exact accumulates one rounded fp32 product per DDA step, so equality here holds the whole step sequence."""
    pos, d, sun = defaults
    for mode in (1, 2):
        ref, orc = both(ob, shader, oracle_world, 640, 360, pos, d, sun, render_pos_data=mode)
        assert np.array_equal(ref["rgba32f"].view(np.uint32), orc["rgba32f"].view(np.uint32)), f"debug mode {mode}"
    hits = orc["hits"].reshape(360, 640)
    vox = ref["rgba32f"][..., :3]
    traced = hits["kind"] == 5  # KIND_DEBUG: every pixel that went through the loops
    assert traced.sum() > 600 * 300
    for k, n in enumerate("xyz"):
        assert np.array_equal(vox[..., k][traced].astype(np.int32), hits[n][traced])


@pytest.mark.parametrize("cam", range(len(CAMERAS)))
def test_camera_sweep_against_the_reference_shader(ob, shader, oracle_world, ref_world, defaults, cam):
    origin, yaw, pitch = CAMERAS[cam]
    _, _, sun = defaults
    d = camera_dir(yaw, pitch)
    look, _, _ = ref_world.pick(origin, d, 30)  # what main.cpp:81,89 uploads as lookingAtBlock, hit or miss
    ref, orc = both(ob, shader, oracle_world, 480, 270, origin, d, sun, delta_time=1.25, mouse_pos=(240.0, 135.0),
                    looking_at=tuple(int(v) for v in look))
    assert_same_frame(ref, orc, f"camera {cam}", 270, 480)
    assert (orc["hits"]["kind"] == 3).sum() > 0  # the crosshair disc


def test_material_zoo_against_the_reference_shader(ob, defaults):
    """Mirrors, refractors, liquid (sin-perturbed normals, non-zero deltaTime), uniform cells, cameras inside solid
    and refractive voxels: everything reflectRay / refractRay / the camera-voxel test do (low_res.frag:178-240,299-333)."""
    from raytracing_test_b200 import voxel_host as vh
    from test_parity_gpu import build_zoo_world

    rng = np.random.default_rng(906)
    _, _, sun = defaults
    host = build_zoo_world(vh, 6, rng)
    try:
        nodes, arrays = host.pools()
        oworld = ob.make_world(nodes, arrays, 6)
        try:
            sh = ob.RefShader(nodes, arrays, oracle_math=True)
        except FileNotFoundError:
            pytest.skip("oracle/_ref/libref_shader.so not built")
        pts = rng.integers(0, 200, size=(20000, 3)).astype(np.int32)
        blocks = ob.get_blocks(oworld, pts)
        views = [((96.5, 40.5, 96.5), 30.0, 25.0), ((20.5, 100.5, 30.5), 60.0, 40.0), ((180.5, 8.5, 10.5), -45.0, 5.0),
                 ((-10.5, 60.5, 96.5), 90.0, 15.0), ((96.5, 190.5, 96.5), 10.0, 80.0), ((100.2, 70.7, 33.3), 200.0, -20.0),
                 ((64.5, 5.5, 64.5), 135.0, -35.0), ((8.5, 72.5, 8.5), 45.0, 0.0)]
        solid = pts[blocks[:, 0] != np.uint64(0xFFFFFFFFFFFFFFFF)]
        refr = pts[(blocks[:, 0] != np.uint64(0xFFFFFFFFFFFFFFFF)) & ((blocks[:, 1] & np.uint64(4)) != 0)]
        for k in range(3):
            p = solid[k * 7]
            views.append(((float(p[0]) + 0.5, float(p[1]) + 0.5, float(p[2]) + 0.5), 33.0 * k, 10.0))
        for k in range(3):
            p = refr[k * 5]
            views.append(((float(p[0]) + 0.3, float(p[1]) + 0.6, float(p[2]) + 0.4), 77.0 * k, -12.0))
        seen = np.zeros(8, dtype=np.int64)
        for i, (origin, yaw, pitch) in enumerate(views):
            ref, orc = both(ob, sh, oworld, 320, 180, origin, camera_dir(yaw, pitch), sun, delta_time=0.75)
            assert_same_frame(ref, orc, f"zoo view {i}", 180, 320)
            fl = orc["hits"]["flags"]
            seen[0] += ((fl & 2) != 0).sum()
            seen[1] += ((fl & 4) != 0).sum()
            seen[2] += (orc["hits"]["kind"] == 2).sum()
        assert seen[0] > 1000 and seen[1] > 1000 and seen[2] > 0, seen  # reflected, refracted, camera-inside pixels
    finally:
        host.close()


def test_libm_sin_exp_stay_within_one_255th(ob, shader, oracle_world, defaults):
    """With the C library's sin / exp in the shader instead of the oracle's deterministic ones, the 8-bit frame moves by
    at most one count (sky pixels only: exp in genSkyBox) -- the tolerance BASELINE.json allows for colour."""
    pos, d, sun = defaults
    shader.use_oracle_math(False)
    try:
        ref = shader.render(640, 360, pos, d, sun, want_state=False)
    finally:
        shader.use_oracle_math(True)
    orc = ob.render(oracle_world, ob.make_frame(640, 360, pos, d, sun, shadows=True), want_f32=True)
    q = lambda img: np.rint(np.clip(img, 0.0, 1.0) * 255.0).astype(np.int16)
    diff = np.abs(q(ref["rgba32f"]) - q(orc["rgba32f"]))
    assert diff.max() <= 1
    hit = (orc["hits"]["kind"] == 1).reshape(360, 640)
    assert np.array_equal(ref["rgba32f"][hit].view(np.uint32), orc["rgba32f"][hit].view(np.uint32))  # no sin/exp on terrain


def test_strided_sample_is_a_subset_of_the_full_frame(shader, defaults):
    pos, d, sun = defaults
    full = shader.render(320, 180, pos, d, sun, want_state=False)
    part = shader.render(320, 180, pos, d, sun, col_stride=3, row_stride=5, rows=(10, 170), want_state=False)
    assert part["fragments"] == len(range(10, 170, 5)) * len(range(0, 320, 3))
    assert np.array_equal(part["rgba32f"][10:170:5, ::3], full["rgba32f"][10:170:5, ::3])
    untouched = np.ones((180, 320), dtype=bool)
    untouched[10:170:5, ::3] = False
    assert (part["rgba32f"][untouched] == 0).all()


@pytest.mark.skipif(not os.path.isdir("/root/reference/src/shaders"), reason="needs the reference tree (build container only)")
def test_build_recipe_rewrites_only_declarations():
    """The sed script between the reference's file and g++ changes seven lines, none of them an expression."""
    import subprocess

    here = os.path.dirname(os.path.abspath(__file__))
    script = os.path.join(here, "..", "oracle", "shim", "glsl", "frag_to_cpp.sed")
    src = "/root/reference/src/shaders/low_res.frag"
    out = subprocess.run(["sed", "-E", "-f", script, src], check=True, capture_output=True, text=True).stdout.split("\n")
    orig = open(src).read().split("\n")
    assert len(out) == len(orig)
    changed = [(a, b) for a, b in zip(orig, out) if a != b]
    assert len(changed) == 7
    for a, b in changed:
        a = a.strip()
        assert (a.startswith("#version") or a.startswith("#extension") or a == "uint[64] children_array[];"
                or a in ("void getBlock();", "void nextIntersect(int step);", "void nextIntersectDDA();")), a
        assert b.strip() in ("", "uint children_array[][64];")

"""The CUDA source itself on the CPU.  raytracing_test_b200/csrc/kernels.cu -- not a restatement of it -- is compiled by g++
through tests/kernel_host/cuda_on_host.h (one host thread plays one CUDA thread; every IEEE intrinsic is the corresponding C
operation under -ffp-contract=off) and run for whole frames: the table kernels as rc_api.cu launches them, then
shade_pixel() for every pixel -- camera ray, loop A over tree regions / column regions / the top table, sky gates and drop
gates, free flight with the closed-form leap, the replay of pos.exact, loop B (mirrors, refraction), shading, shadow rays.
Hit records, RGBA8 and the float frame must equal the oracle's bit for bit, on the BASELINE configs without AO at full
size, the parity cameras, every A/B mode, ragged sizes, step budgets, the debug outputs, the material-zoo and tower-forest
worlds and a sample of the random worlds of the GPU suite.  What g++ sees is kernels.cu up to its first host-side launch
helper with its five inline-PTX lines rewritten (three blanked, the two shared-memory byte loads as the same load in C;
kernel launches further down need nvcc); a test checks exactly that.

Kernel mode (second half of this file) goes further: the __global__ functions themselves -- primary_kernel<false / true>,
ao_window_kernel, ao_window_persistent_kernel, ao_cached_kernel, frame_persistent_kernel<false / true> -- run CTA by CTA as
launch_frame schedules them, every warp as 32 coroutines that meet at each *_sync intrinsic (KhWarp in cuda_on_host.h), with
the CTA's shared memory as a plain array: warp votes, shuffles, match_any, the AO windows with their denormal-float indices,
the seed list, the atomic tile / chunk queues, the staged top levels.  BASELINE config 3's frame (64-sample AO) under all
five schedules, the parity cameras with AO, sample / step counts from 5 x 0 to 255 x 5 and 12 x 32: bit for bit the oracle's,
ao_occluded of every pixel included.

This is the parity gate a reader WITHOUT a GPU can run.  Not covered here: what only exists on the device -- real
concurrency (races between warps, memory ordering, the peer stores and completion flags between GPUs), nvcc's code
generation -- which the `-m gpu` suite holds."""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import CAMERAS, camera_dir

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "raytracing_test_b200", "csrc")
MARKER = "// Kernels with dynamic shared memory above 48 KB need the opt-in once per device"
LDS = re.compile(r'^(\s*)asm volatile\("ld\.shared\.u8 %0, \[%1\];" : "=r"\((\w+)\) : "r"\((.+)\) : "memory"\);$')


def device_part_of_kernels_cu():
    """kernels.cu as g++ gets it: (text, lines changed).  Everything before the first host-side launch helper (the device code
    and upload_hemisphere, which builds the AO tables), verbatim but for the inline-PTX statements (one line each: the two
    shared-memory byte loads become the same load in C, the other three are blanked), plus the brace that closes namespace rc."""
    out, changed = [], []
    # KHOST_KERNELS=<file>: another version of kernels.cu (a kernel change under development, held to the oracle before it
    # ever reaches a GPU)
    with open(os.environ.get("KHOST_KERNELS") or os.path.join(CSRC, "kernels.cu")) as f:
        for n, line in enumerate(f.read().split("\n"), 1):
            if line.startswith(MARKER):
                break
            m = LDS.match(line)
            if m:  # the shared-memory byte load of the AO window code: the same load, in C
                changed.append((n, line.strip()))
                out.append(f"{m.group(1)}{m.group(2)} = kh_lds_u8({m.group(3)});")
            elif "asm volatile(" in line:
                changed.append((n, line.strip()))
                out.append("// (inline PTX blanked for the host build)")
            else:
                out.append(line)
        else:
            raise AssertionError("kernels.cu: the marker before the host-side launch helpers is gone")
    out.append("}  // namespace rc")
    return "\n".join(out) + "\n", changed


class KernelOnHost:
    def __init__(self, lib_path, rcmod):
        self.rc = rcmod
        lib = ctypes.CDLL(lib_path)
        lib.khost_world_create.restype = ctypes.c_void_p
        lib.khost_world_create.argtypes = [ctypes.POINTER(rcmod.WorldDesc)]
        lib.khost_world_destroy.argtypes = [ctypes.c_void_p]
        lib.khost_render.restype = ctypes.c_int
        lib.khost_render.argtypes = [ctypes.c_void_p, ctypes.POINTER(rcmod.FrameParams)] + [ctypes.c_void_p] * 4
        lib.khost_lookup.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p]
        lib.khost_render_kernels.restype = ctypes.c_int
        lib.khost_render_kernels.argtypes = [ctypes.c_void_p, ctypes.POINTER(rcmod.FrameParams), ctypes.c_uint32] + [ctypes.c_void_p] * 5
        lib.khost_last_error.restype = ctypes.c_char_p
        self.lib = lib

    def world(self, desc):
        w = self.lib.khost_world_create(ctypes.byref(desc))
        assert w, self.lib.khost_last_error()
        return w

    def free(self, world):
        self.lib.khost_world_destroy(world)

    def render(self, world, params):
        n = params.width * params.height
        hits = np.zeros(n, dtype=self.rc.HIT_DTYPE)
        img = np.zeros((params.height, params.width, 4), dtype=np.uint8)
        f32 = np.zeros((params.height, params.width, 4), dtype=np.float32)
        stats = np.zeros(3, dtype=np.uint64)
        r = self.lib.khost_render(world, ctypes.byref(params), hits.ctypes.data, img.ctypes.data, f32.ctypes.data, stats.ctypes.data)
        assert r == 0, self.lib.khost_last_error()
        return {"hits": hits, "rgba8": img, "rgba32f": f32, "free_flight_rays": int(stats[0]), "hit_pixels": int(stats[1]),
                "shadow_rays": int(stats[2])}

    def render_kernels(self, world, params, schedule=0, octant_skip=True, lane_order=0):
        """The __global__ functions themselves, scheduled as launch_frame schedules them, every warp as 32 coroutines."""
        n = params.width * params.height
        hits = np.zeros(n, dtype=self.rc.HIT_DTYPE)
        img = np.zeros((params.height, params.width, 4), dtype=np.uint8)
        f32 = np.zeros((params.height, params.width, 4), dtype=np.float32)
        stats = np.zeros(4, dtype=np.uint64)
        tab = None
        if params.ao_samples:
            tab = np.zeros((params.ao_samples, 3), dtype=np.float32)
            assert self.rc.load_library().rc_hemisphere_table(params.ao_samples, tab.ctypes.data) == 0  # host code of librc, no device
        r = self.lib.khost_render_kernels(world, ctypes.byref(params), schedule | (0 if octant_skip else 0x100) | (lane_order << 9),
                                          tab.ctypes.data if tab is not None else None, hits.ctypes.data, img.ctypes.data,
                                          f32.ctypes.data, stats.ctypes.data)
        assert r == 0, self.lib.khost_last_error()
        return {"hits": hits, "rgba8": img, "rgba32f": f32, "free_flight_rays": int(stats[0]), "hit_pixels": int(stats[1]),
                "shadow_rays": int(stats[2]), "ao_seeds": int(stats[3])}

    def lookup(self, world, xyz):
        xyz = np.ascontiguousarray(xyz, dtype=np.int32).reshape(-1, 3)
        words = np.zeros(len(xyz), dtype=np.uint32)
        shifts = np.zeros(len(xyz), dtype=np.int32)
        self.lib.khost_lookup(world, xyz.ctypes.data, len(xyz), words.ctypes.data, shifts.ctypes.data)
        return words, shifts


@pytest.fixture(scope="module")
def khost(tmp_path_factory, rcmod):
    d = tmp_path_factory.mktemp("kernel_on_host")
    text, _ = device_part_of_kernels_cu()
    inc = str(d / "kernels_device_part.inc")
    with open(inc, "w") as f:
        f.write(text)
    out = str(d / "libkernel_on_host.so")
    cuda_inc = os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "include")
    if not os.path.exists(os.path.join(cuda_inc, "cuda_runtime.h")):
        pytest.skip("no CUDA toolkit headers here (vector types of cuda_runtime.h are needed to compile kernels.cu for the host)")
    # KHOST_DEFINES="-DRC_MUTATE_DROP": the deliberately wrong variants of tools/mutation_check.sh -- this file then fails
    extra = os.environ.get("KHOST_DEFINES", "").split()
    subprocess.run(["g++", "-std=gnu++17", "-O2", "-ffp-contract=off", "-fno-fast-math", "-fPIC", "-shared", "-w", *extra, "-I", cuda_inc,
                    "-I", CSRC, "-I", os.path.join(ROOT, "tests", "kernel_host"), f'-DKERNEL_SOURCE="{inc}"',
                    os.path.join(ROOT, "tests", "kernel_host", "kernel_on_host.cpp"), os.path.join(CSRC, "flatten.cpp"), "-o", out],
                   check=True)
    return KernelOnHost(out, rcmod)


@pytest.fixture(scope="module")
def default_world(khost, ob):
    from raytracing_test_b200 import voxel_host as vh

    host = vh.World.default()
    nodes, arrays = host.pools()
    world = khost.world(host.describe())
    yield {"host": host, "kw": world, "oracle": ob.make_world(nodes, arrays, 6), "camera": vh.default_camera()}
    khost.free(world)
    host.close()


def both(khost, ob, rcmod, kw, oworld, w, h, pos, d, sun, *, hits=True, modes=None, **kw_frame):
    """(oracle frame, kernel-on-host frame) for the same parameters; modes = frame_params switches of the kernel only."""
    ref = ob.render(oworld, ob.make_frame(w, h, pos, d, sun, **kw_frame), want_f32=True)
    got = khost.render(kw, rcmod.frame_params(w, h, pos, d, sun, hits=hits, **kw_frame, **(modes or {})))
    return ref, got


def assert_identical(ref, got, label, hits=True):
    if hits and got["hits"].tobytes() != ref["hits"].tobytes():
        a, b = got["hits"].view(np.uint8).reshape(-1, 20), ref["hits"].view(np.uint8).reshape(-1, 20)
        bad = np.flatnonzero((a != b).any(axis=1))
        raise AssertionError(f"{label}: {len(bad)} hit records differ, first pixel {bad[0]}: kernel {got['hits'][bad[0]]} "
                             f"oracle {ref['hits'][bad[0]]}")
    assert np.array_equal(got["rgba8"], ref["rgba8"]), f"{label}: RGBA8 differs"
    assert np.array_equal(got["rgba32f"].view(np.uint32), ref["rgba32f"].view(np.uint32)), f"{label}: float frame differs"


def test_what_reaches_gpp_is_kernels_cu_but_for_the_inline_ptx():
    text, changed = device_part_of_kernels_cu()
    src = open(os.environ.get("KHOST_KERNELS") or os.path.join(CSRC, "kernels.cu")).read()
    cut = src.index(MARKER)
    # five single-line PTX statements: the optimisation fence of the empty-cell loop and the two griddepcontrol statements of
    # the kernel prologues (blanked: no effect on results), the two shared-memory byte loads of the AO window code (the same
    # load in C: kh_lds_u8)
    assert len(changed) == 5, changed
    assert text.count("= kh_lds_u8(") == 2 and text.count("// (inline PTX blanked for the host build)") == 3
    assert sum("ld.shared.u8" in c for _, c in changed) == 2 and sum("griddepcontrol" in c for _, c in changed) == 2
    assert sum(c.startswith('asm volatile("" :') for _, c in changed) == 1
    before, after = src[:cut].split("\n")[:-1], text.split("\n")
    assert len(after) == len(before) + 2 and after[len(before)] == "}  // namespace rc"   # + the closing brace, + the final newline
    assert [n for n, (x, y) in enumerate(zip(before, after), 1) if x != y] == [n for n, _ in changed]
    # nothing of the device code lies behind the cut: only launch helpers (kernel<<<...>>>)
    tail = src[cut:]
    assert "__global__" not in tail and "__device__" not in tail and "<<<" in tail


def test_baseline_configs_1_and_2_at_full_size(khost, ob, rcmod, default_world):
    pos, d, sun = default_world["camera"]
    kw, ow = default_world["kw"], default_world["oracle"]
    ref, got = both(khost, ob, rcmod, kw, ow, 1280, 720, pos, d, sun, shadows=False)           # BASELINE configs[0]
    assert_identical(ref, got, "config 1")
    assert got["free_flight_rays"] > 400_000 and got["shadow_rays"] == 0
    ref, got = both(khost, ob, rcmod, kw, ow, 1920, 1080, pos, d, sun, shadows=True)           # BASELINE configs[1]
    assert_identical(ref, got, "config 2")
    sky = int((got["hits"]["kind"] == 0).sum())
    assert got["free_flight_rays"] > 0.95 * sky and got["shadow_rays"] > 100_000
    # the frame as bench.py times it: no hit records, so free rays skip their tail; same pixels
    _, fast = both(khost, ob, rcmod, kw, ow, 1920, 1080, pos, d, sun, hits=False, shadows=True)
    assert_identical(ref, fast, "config 2 without hit records", hits=False)


@pytest.mark.parametrize("cam", range(len(CAMERAS)))
def test_camera_sweep(khost, ob, rcmod, default_world, cam):
    _, _, sun = default_world["camera"]
    origin, yaw, pitch = CAMERAS[cam]
    ref, got = both(khost, ob, rcmod, default_world["kw"], default_world["oracle"], 640, 360, origin, camera_dir(yaw, pitch), sun,
                    shadows=True, delta_time=0.75, mouse_pos=(300.5, 200.5))
    assert_identical(ref, got, f"camera {cam}")


def test_every_mode_renders_the_same_frame(khost, ob, rcmod, default_world):
    """Free flight, the leap, column regions, drop gates off one at a time and all together: same hit records (final voxel,
    face and step count of every sky ray included), same pixels; and each mechanism really runs when it is on."""
    pos, d, sun = default_world["camera"]
    kw, ow = default_world["kw"], default_world["oracle"]
    modes = {"on": {}, "walk": dict(leap=False), "no_columns": dict(column_regions=False), "no_drop": dict(drop_gates=False),
             "no_free_flight": dict(free_flight=False), "plain": dict(free_flight=False, column_regions=False),
             "always_tail": dict(always_tail=True)}
    views = [(300, pos, d), (64, pos, d), (1024, pos, d), (2000, pos, d), (300, (100.5, 120.25, 100.5), camera_dir(30.0, -50.0)),
             (300, (35.0, 50.0, 35.0), (0.0, 1.0, 0.0)), (300, (35.5, 80.5, 35.5), (1.0, 0.0, 0.0)),
             (500, (120.5, 60.5, 90.5), camera_dir(200.0, 5.0)), (300, (35.0, 50.0, 35.0), camera_dir(45.0, 4.0))]
    free = {}
    for steps, o, dd in views:
        ref = None
        for name, m in modes.items():
            r, got = both(khost, ob, rcmod, kw, ow, 320, 180, o, dd, sun, shadows=True, max_steps=steps, modes=m)
            assert_identical(r, got, f"mode {name}, {steps} steps from {o}")
            free[name] = free.get(name, 0) + got["free_flight_rays"]
    assert free["plain"] == 0 and free["no_free_flight"] == 0
    assert free["on"] == free["walk"] == free["always_tail"] > 130_000
    assert free["on"] > free["no_drop"] + 8_000   # rays that step down are set free too


def test_ragged_sizes_step_limits_and_debug_outputs(khost, ob, rcmod, default_world):
    pos, d, sun = default_world["camera"]
    kw, ow, host = default_world["kw"], default_world["oracle"], default_world["host"]
    for (w, h) in [(1, 1), (7, 5), (33, 17), (250, 141), (641, 3)]:
        ref, got = both(khost, ob, rcmod, kw, ow, w, h, pos, d, sun, shadows=True)
        assert_identical(ref, got, f"{w}x{h}")
    for (ms, ss) in [(0, 0), (1, 1), (40, 10), (300, 0), (700, 300), (1024, 1024), (1500, 75)]:
        ref, got = both(khost, ob, rcmod, kw, ow, 200, 120, pos, d, sun, shadows=True, max_steps=ms, shadow_steps=ss)
        assert_identical(ref, got, f"steps {ms}/{ss}")
    look, _, _ = host.pick(pos, d, 30)
    for la in (tuple(int(v) for v in look), (1 << 30, 1 << 30, 1 << 30)):
        ref, got = both(khost, ob, rcmod, kw, ow, 320, 200, pos, d, sun, shadows=True, looking_at=la)
        assert_identical(ref, got, f"looking at {la}")
    probe = ob.render(ow, ob.make_frame(320, 200, pos, d, sun))["hits"][60 * 320 + 160]  # a block a little below the centre
    assert probe["kind"] == ob.KIND_HIT
    ref, got = both(khost, ob, rcmod, kw, ow, 320, 200, pos, d, sun, shadows=True,
                    looking_at=(int(probe["x"]), int(probe["y"]), int(probe["z"])))
    assert_identical(ref, got, "highlight")
    assert int((got["hits"]["kind"] == ob.KIND_HIGHLIGHT).sum()) > 0
    for mode in (1, 2):   # renderPosData: pos.exact (every rounded product of every DDA step) / pos.round as the colour
        ref, got = both(khost, ob, rcmod, kw, ow, 320, 200, pos, d, sun, shadows=True, render_pos_data=mode)
        assert_identical(ref, got, f"render_pos_data {mode}")


def test_lookups_through_the_device_routine(khost, ob, default_world):
    """lookup(): top table, cursor, popcount addressing -- against the oracle's getBlock on the reference-format pools."""
    rng = np.random.default_rng(5)
    pts = np.concatenate([rng.integers(0, 208, size=(60000, 3)), rng.integers(-40, 1064, size=(20000, 3)),
                          np.array([[10, 100, 10], [1000, 1000, 1000], [0, 0, 0], [1023, 1023, 1023], [767, 50, 767]])]).astype(np.int32)
    inside = ((pts >= 768) & (pts <= 1023)).all(axis=1)  # the aliased root cell
    pts = pts[~inside]
    words, shifts = khost.lookup(default_world["kw"], pts)
    ref = ob.get_blocks(default_world["oracle"], pts)
    assert np.array_equal(words != 0, ref[:, 0] != np.uint64(0xFFFFFFFFFFFFFFFF))
    assert (words != 0).sum() > 5000 and set(np.unique(shifts[words == 0])) >= {2, 4, 6, 8}


@pytest.mark.parametrize("depth", [6, 7])
def test_material_zoo_world(khost, ob, rcmod, default_world, depth):
    """Every material flag combination as voxels and as uniform 4^3 / 16^3 / 64^3 cells, mirrors facing mirrors, liquid with
    animated normals, cameras inside solid and refractive voxels (the world of the GPU suite's test of the same name):
    loop B, reflectRay, refractRay, det_sin."""
    from raytracing_test_b200 import voxel_host as vh
    from test_parity_gpu import build_zoo_world

    rng = np.random.default_rng(900 + depth)
    _, _, sun = default_world["camera"]
    host = build_zoo_world(vh, depth, rng)
    kw = khost.world(host.describe())
    try:
        nodes, arrays = host.pools()
        ow = ob.make_world(nodes, arrays, depth)
        pts = rng.integers(0, 200, size=(50000, 3)).astype(np.int32)
        words, _ = khost.lookup(kw, pts)
        assert np.array_equal(words != 0, ob.get_blocks(ow, pts)[:, 0] != np.uint64(0xFFFFFFFFFFFFFFFF))
        views = [((96.5, 40.5, 96.5), 30.0, 25.0), ((20.5, 100.5, 30.5), 60.0, 40.0), ((180.5, 8.5, 10.5), -45.0, 5.0),
                 ((-10.5, 60.5, 96.5), 90.0, 15.0), ((96.5, 190.5, 96.5), 10.0, 80.0), ((100.2, 70.7, 33.3), 200.0, -20.0),
                 ((64.5, 5.5, 64.5), 135.0, -35.0), ((8.5, 72.5, 8.5), 45.0, 0.0)]
        solid = pts[words != 0]
        for k in range(4):
            p = solid[k * 7]
            views.append(((float(p[0]) + 0.5, float(p[1]) + 0.5, float(p[2]) + 0.5), 33.0 * k, 10.0))
        seen = {"refl": 0, "refr": 0, "shadow": 0, "inside": 0}
        for i, (origin, yaw, pitch) in enumerate(views):
            ref, got = both(khost, ob, rcmod, kw, ow, 480, 270, origin, camera_dir(yaw, pitch), sun, shadows=True, delta_time=0.75)
            assert_identical(ref, got, f"zoo depth {depth} view {i}")
            fl = got["hits"]["flags"]
            seen["refl"] += int(((fl & 0x02) != 0).sum())
            seen["refr"] += int(((fl & 0x04) != 0).sum())
            seen["shadow"] += int(((fl & 0x01) != 0).sum())
            seen["inside"] += int((got["hits"]["kind"] == ob.KIND_CAMERA_INSIDE).sum())
        assert seen["refl"] > 1000 and seen["refr"] > 1000 and seen["shadow"] > 1000 and seen["inside"] > 0, seen
    finally:
        khost.free(kw)
        host.close()


def test_tower_forests_and_empty_space(khost, ob, rcmod, default_world):
    """Adversarial for the gates: bare ground under 3,200 one-voxel-wide towers of pseudo-random heights, cameras low between
    them looking up at 1.5 to 45 degrees and grazing down, a low sun; the same forest beyond x, z = 1024 in a depth-7 tree; an
    empty world."""
    from raytracing_test_b200 import voxel_host as vh
    from test_parity_gpu import _spike_world

    suns = [np.array(v, dtype=np.float32) / np.float32(np.linalg.norm(v)) for v in ([2.0, 1.0, 4.0], [3.0, 0.4, -1.0])]
    for depth, off in ((6, 0), (7, 1500)):
        host = _spike_world(vh, depth, off)
        kw = khost.world(host.describe())
        try:
            nodes, arrays = host.pools()
            ow = ob.make_world(nodes, arrays, depth)
            free = 0
            for (o, yaw, pitch) in [((50.5, 4.5, 50.5), 45.0, -1.5), ((50.5, 4.5, 50.5), 45.0, -8.0), ((120.5, 6.5, 80.5), 200.0, -20.0),
                                    ((200.5, 10.5, 200.5), -60.0, -45.0), ((30.5, 60.5, 30.5), 45.0, 3.0), ((30.5, 150.5, 30.5), 45.0, 12.0)]:
                origin = (o[0] + off, o[1], o[2] + off)
                for sun, steps in ((suns[0], 300), (suns[1], 1024)):
                    ref, got = both(khost, ob, rcmod, kw, ow, 320, 180, origin, camera_dir(yaw, pitch), sun, shadows=True,
                                    max_steps=steps, shadow_steps=min(steps, 300))
                    assert_identical(ref, got, f"towers depth {depth} from {origin} yaw {yaw} pitch {pitch} steps {steps}")
                    free += got["free_flight_rays"]
            assert free > 5_000
        finally:
            khost.free(kw)
            host.close()
    host = vh.World(6)
    kw = khost.world(host.describe())
    try:
        nodes, arrays = host.pools()
        pos, d, sun = default_world["camera"]
        ref, got = both(khost, ob, rcmod, kw, ob.make_world(nodes, arrays, 6), 320, 180, pos, d, sun, shadows=True)
        assert_identical(ref, got, "empty world")
        assert got["free_flight_rays"] > 0.95 * 320 * 180  # all but the steepest rays down, which would leave the torus
    finally:
        khost.free(kw)
        host.close()


def test_random_worlds(khost, ob, rcmod):
    """A sample of the randomised cases of tests/test_sky_gates_random_gpu.py (same generators, other seeds): sparse worlds with
    columns, floating voxels / 4^3 / 16^3 cells, cells that straddle column-block borders and the torus seam, depths 6 and 7;
    cameras looking up and down, grazing, axis-parallel, near the seam; step budgets 40..1024; suns down to the horizon and
    below.  Every case: kernel source on the host == oracle, with all shortcuts on and with all of them off."""
    from raytracing_test_b200 import voxel_host as vh
    from test_sky_gates_random_gpu import random_case, random_world

    cases = free = 0
    for wi in range(12):
        seed = 8_100_000 + wi
        rng = np.random.default_rng(seed)
        depth = 6 if wi % 3 else 7
        host, base, span, edge = random_world(vh, rng, depth)
        kw = khost.world(host.describe())
        try:
            nodes, arrays = host.pools()
            ow = ob.make_world(nodes, arrays, depth)
            for ci in range(10):
                o, d, sun, budget = random_case(rng, base, span, edge)
                shadow_steps = int(rng.choice([75, min(budget, 75), budget]))
                where = f"seed {seed} (depth {depth}) case {ci}: origin {o} dir {tuple(float(v) for v in d)} sun {sun} steps {budget}/{shadow_steps}"
                for m in ({}, dict(free_flight=False, column_regions=False)):
                    ref, got = both(khost, ob, rcmod, kw, ow, 160, 90, o, d, sun, shadows=True, max_steps=budget,
                                    shadow_steps=shadow_steps, modes=m)
                    assert_identical(ref, got, where + f" modes {m}")
                    free += got["free_flight_rays"]
                cases += 1
        finally:
            khost.free(kw)
            host.close()
    assert cases == 120 and free > 100 * cases


# ---- kernel mode: the __global__ functions themselves, every warp as 32 coroutines ------------------------------------------
# primary_kernel<false / true>, ao_window_kernel, ao_window_persistent_kernel, ao_cached_kernel, frame_persistent_kernel<false /
# true> run CTA by CTA as launch_frame schedules them: warp votes, shuffles, match_any, the shared-memory AO windows with
# their denormal-float indices, the seed list, the atomic tile / chunk queues, the staged top levels -- everything but real
# concurrency.


def both_kernels(khost, ob, rcmod, kw, oworld, w, h, pos, d, sun, *, schedule=0, octant_skip=True, lane_order=0, **kw_frame):
    ao = kw_frame.get("ao_samples", 0)
    if ao:
        tab = np.zeros((ao, 3), dtype=np.float32)
        assert rcmod.load_library().rc_hemisphere_table(ao, tab.ctypes.data) == 0
        ob.oracle_lib().oracle_set_hemisphere(ao, tab.ctypes.data)
    ref = ob.render(oworld, ob.make_frame(w, h, pos, d, sun, **kw_frame), want_f32=True)
    got = khost.render_kernels(kw, rcmod.frame_params(w, h, pos, d, sun, hits=True, rgba32f=True, **kw_frame), schedule, octant_skip,
                               lane_order)
    return ref, got


def test_config3_kernels_under_every_schedule(khost, ob, rcmod, default_world):
    """BASELINE config 3's frame (default view, shadows, 64-sample AO) at a quarter of its size: two kernels, persistent AO grid,
    fused, persistent fused over the atomic tile queue, the same with the top levels staged in shared memory; octant skip on
    and off.  ao_occluded of every pixel, hit records, RGBA8 and the float frame against the oracle."""
    pos, d, sun = default_world["camera"]
    kw, ow = default_world["kw"], default_world["oracle"]
    for schedule in range(5):
        ref, got = both_kernels(khost, ob, rcmod, kw, ow, 480, 270, pos, d, sun, schedule=schedule, shadows=True, ao_samples=64)
        assert_identical(ref, got, f"config 3, schedule {schedule}")
        assert got["ao_seeds"] == got["hit_pixels"] > 50_000 and int(got["hits"]["ao_occluded"].astype(np.int64).sum()) > 500_000
    ref, got = both_kernels(khost, ob, rcmod, kw, ow, 480, 270, pos, d, sun, octant_skip=False, shadows=True, ao_samples=64)
    assert_identical(ref, got, "config 3, octant skip off")


@pytest.mark.parametrize("cam", range(len(CAMERAS)))
def test_camera_sweep_with_ao(khost, ob, rcmod, default_world, cam):
    """The parity cameras (mirror, refractor, lake, under water, negative origin, near-vertical ...) with 20 AO samples: pixels
    with LIQUID in their window and pixels off their hit face take ao_ray_general, the rest the window code in byte and bit mode."""
    _, _, sun = default_world["camera"]
    origin, yaw, pitch = CAMERAS[cam]
    ref, got = both_kernels(khost, ob, rcmod, default_world["kw"], default_world["oracle"], 192, 108, origin, camera_dir(yaw, pitch), sun,
                            schedule=(0, 2, 1, 3)[cam % 4], shadows=True, ao_samples=20, delta_time=0.75)
    assert_identical(ref, got, f"camera {cam} with AO")


def test_ao_sample_and_step_counts_ragged_sizes(khost, ob, rcmod, default_world):
    pos, d, sun = default_world["camera"]
    kw, ow = default_world["kw"], default_world["oracle"]
    o2, d2 = (100.5, 60.25, 100.5), camera_dir(0.0, 60.0)
    for (w, h, ao, steps) in [(33, 17, 8, 5), (250, 141, 7, 5), (1, 1, 64, 5), (200, 120, 255, 5), (200, 120, 16, 9), (200, 120, 16, 3),
                              (200, 120, 12, 32), (200, 120, 5, 0)]:
        for (o, dd) in ((pos, d), (o2, d2)):
            ref, got = both_kernels(khost, ob, rcmod, kw, ow, w, h, o, dd, sun, shadows=True, ao_samples=ao, ao_steps=steps,
                                    ao_strength=0.7)
            assert_identical(ref, got, f"{w}x{h} ao {ao} x {steps} steps from {o}")


def test_zoo_world_kernels_with_ao(khost, ob, rcmod, default_world):
    from raytracing_test_b200 import voxel_host as vh
    from test_parity_gpu import build_zoo_world

    rng = np.random.default_rng(906)
    _, _, sun = default_world["camera"]
    host = build_zoo_world(vh, 6, rng)
    kw = khost.world(host.describe())
    try:
        nodes, arrays = host.pools()
        ow = ob.make_world(nodes, arrays, 6)
        occluded = 0
        for i, (origin, yaw, pitch) in enumerate([((96.5, 40.5, 96.5), 30.0, 25.0), ((20.5, 100.5, 30.5), 60.0, 40.0),
                                                  ((64.5, 5.5, 64.5), 135.0, -35.0), ((100.2, 70.7, 33.3), 200.0, -20.0)]):
            ref, got = both_kernels(khost, ob, rcmod, kw, ow, 256, 144, origin, camera_dir(yaw, pitch), sun, schedule=(0, 2)[i % 2],
                                    shadows=True, ao_samples=24, delta_time=0.75)
            assert_identical(ref, got, f"zoo view {i} with AO")
            occluded += int(got["hits"]["ao_occluded"].astype(np.int64).sum())
        assert occluded > 5_000
    finally:
        khost.free(kw)
        host.close()


def test_results_do_not_depend_on_the_order_of_the_lanes(khost, ob, rcmod, default_world):
    """A race check compute-sanitizer would do on the device (it is closed on the GPU pool): between two rendezvous the
    emulator runs the lanes of a warp one after the other -- ascending by default; here descending and shuffled anew in
    every pass.  If a lane read shared or global memory that another lane of its warp writes without a *_sync between them
    (the cooperative window fill with its atomicOr, the liquid word, the seed-list append, the counters), some order would
    show it.  Every schedule, cameras with many distinct windows per warp (bit mode) and with liquid.  (Teeth: without the
    __syncwarp behind the window fill, the lake camera's frame has 1 / 313 / 15 wrong pixels under the three orders.)"""
    pos, d, sun = default_world["camera"]
    kw, ow = default_world["kw"], default_world["oracle"]
    views = [(pos, d), (CAMERAS[2][0], camera_dir(*CAMERAS[2][1:])), (CAMERAS[7][0], camera_dir(*CAMERAS[7][1:]))]
    for order in (1, 2):
        for schedule in range(5):
            o, dd = views[schedule % len(views)]
            ref, got = both_kernels(khost, ob, rcmod, kw, ow, 208, 117, o, dd, sun, schedule=schedule, lane_order=order, shadows=True,
                                    ao_samples=16, delta_time=0.75)
            assert_identical(ref, got, f"lane order {order}, schedule {schedule}")

"""CPU tests of the ORACLE: the restatement of low_res.frag is pinned against everything the
reference lets us pin: its own world builder and getBlock compiled unmodified
(oracle/_ref), the test.cpp DDA scenario, castRayFromCam, the packed-colour and Morton KATs,
the hemisphere-table KAT, and the committed golden frames (tests/golden, made by
tools/make_golden.py)."""
import ctypes
import hashlib
import json
import math
import os

import numpy as np
import pytest

from conftest import camera_dir

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
EMPTY = np.uint64(0xFFFFFFFFFFFFFFFF)


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.fixture(scope="module")
def golden():
    with open(os.path.join(GOLDEN_DIR, "frames.json")) as f:
        return json.load(f), np.load(os.path.join(GOLDEN_DIR, "samples.npz"))


# ---- the world the reference builds (SURVEY 0 probe numbers, 4 KATs) ---------------------------

def test_reference_world_sizes_and_root(ref_world):
    assert ref_world.max_depth == 6 and ref_world.block_bytes == 1 << 22
    assert ref_world.node_bytes_used == 18_753_888  # 1,172,118 nodes
    assert ref_world.array_bytes_used == 5_305_344  # 20,724 arrays
    assert ref_world.lib.refworld_num_node_blocks() == 5 and ref_world.lib.refworld_num_array_blocks() == 2
    bitmap, word1 = int(ref_world.nodes[0]), int(ref_world.nodes[1])
    assert bitmap == 0x8000000000000001
    assert word1 & 0xFFFFFFFF == 0 and word1 >> 32 == 1  # flags 0, children array 1


def test_reference_world_matches_golden_hash(ref_world, golden):
    g = golden[0]["world"]
    assert g["node_bytes_used"] == ref_world.node_bytes_used
    assert sha(ref_world.nodes[: ref_world.node_bytes_used // 8]) == g["nodes_sha256"]
    assert sha(ref_world.arrays[: ref_world.array_bytes_used // 4]) == g["arrays_sha256"]


def test_reference_getblock_kats(ref_world):
    # SURVEY appendix A.14 / A.15
    assert ref_world.get_block(35, 20, 35)[:2] == (0x1090908484842424, 1)           # stone
    assert ref_world.get_block(10, 100, 10)[:2] == (0, 3)                           # black mirror
    assert ref_world.get_block(21, 9, 201)[:2] == (0, 5)                            # 4^3 refractor
    assert ref_world.get_block(24, 9, 201)[0] == int(EMPTY)
    assert ref_world.get_block(1001, 1001, 1001)[0] == int(EMPTY)
    assert ref_world.get_block(35, 50, 35)[0] == int(EMPTY)                         # the camera voxel
    L = ref_world.lib
    assert L.refworld_rgb_to_u64(33, 33, 33) == 0x1090908484842424
    assert L.refworld_rgb_to_u64(45, 18, 0) == 0x1696944848400000
    assert L.refworld_rgb_to_u64(0, 150, 10) == 0x0000025A5A414141


def test_surface_heights(ref_world):
    ys = np.arange(0, 64)
    cols = [(x, z) for x in range(0, 200, 9) for z in range(0, 200, 11)]
    xyz = np.array([[x, y, z] for (x, z) in cols for y in ys], dtype=np.int32)
    solid = (ref_world.get_blocks(xyz)[:, 0] != EMPTY).reshape(len(cols), len(ys))
    top = np.array([ys[row][-1] if row.any() else -1 for row in solid])
    assert top.min() >= 20 and top.max() <= 50


# ---- oracle getBlock == reference getBlock -------------------------------------------------------

def test_oracle_getblock_matches_reference(ob, ref_world, oracle_world):
    rng = np.random.default_rng(11)
    box = np.stack(np.meshgrid(np.arange(0, 204, 3), np.arange(0, 108, 2), np.arange(0, 204, 3), indexing="ij"), -1)
    xyz = np.concatenate([box.reshape(-1, 3), rng.integers(0, 768, size=(100_000, 3)),
                          np.array([[35, 20, 35], [10, 100, 10], [21, 9, 201], [24, 9, 201], [20, 8, 200], [23, 11, 203]])])
    xyz = xyz.astype(np.int32)
    a, b = ob.get_blocks(oracle_world, xyz), ref_world.get_blocks(xyz)
    assert np.array_equal(a[:, 0], b[:, 0])
    solid = b[:, 0] != EMPTY
    assert np.array_equal(a[solid, 1], b[solid, 1])
    assert solid.sum() > 50_000


def test_oracle_getblock_wraps_like_the_shader(ob, oracle_world):
    # low_res.frag:454-459: bits 0-9 of a coordinate survive the bit spread => 1024-torus for x >= 0
    base = np.array([[35, 20, 35], [10, 100, 10], [21, 9, 201], [199, 30, 199], [35, 50, 35]], dtype=np.int32)
    a = ob.get_blocks(oracle_world, base)
    for off in ([1024, 0, 0], [0, 2048, 0], [3072, 1024, 5120]):
        assert np.array_equal(ob.get_blocks(oracle_world, base + np.array(off, dtype=np.int32)), a)
    # ... and the mask 0x030000FF also keeps bits 24-25 of the raw value: a negative coordinate lands
    # in 768 | (x & 255), whatever its bits 8-9 are
    neg = np.array([[-1, 30, 5], [-256, 30, 5], [-257, 30, 5], [-700, 30, 5], [5, 30, -1000]], dtype=np.int32)
    img = np.array([[1023, 30, 5], [768, 30, 5], [1023, 30, 5], [768 | (-700 & 255), 30, 5], [5, 30, 768 | (-1000 & 255)]],
                   dtype=np.int32)
    assert np.array_equal(ob.get_blocks(oracle_world, neg), ob.get_blocks(oracle_world, img))
    L = ob.oracle_lib()
    assert L.oracle_morton(-257, 0, 0, 6) == L.oracle_morton(1023, 0, 0, 6)
    assert L.oracle_morton(-257, 0, 0, 7) == L.oracle_morton(4096 - 257, 0, 0, 7)  # extension: plain torus


def test_f11_region_is_empty_in_the_oracle(ob, oracle_world):
    # root cell 63 holds child index 0 (aliasing bug): documented divergence => empty
    xyz = np.array([[1000, 1000, 1000], [1002, 1002, 1003], [800, 900, 1023], [768, 768, 768]], dtype=np.int32)
    assert (ob.get_blocks(oracle_world, xyz)[:, 0] == EMPTY).all()


# ---- Morton / colour / DDA known answers ---------------------------------------------------------

def test_morton_matches_reference_masks(ob):
    L = ob.oracle_lib()
    rng = np.random.default_rng(3)
    for x, y, z in rng.integers(-2000, 3000, size=(2000, 3)):
        x, y, z = int(x), int(y), int(z)
        assert L.oracle_morton(x, y, z, 6) == L.oracle_morton_ref10(x, y, z)
    # getMortonPos (main.cpp:45-66): 6-bit groups z<<4 | y<<2 | x, least significant group = lowest 2 bits
    assert L.oracle_morton(1, 0, 0, 6) == 1 and L.oracle_morton(0, 1, 0, 6) == 4 and L.oracle_morton(0, 0, 1, 6) == 16
    assert L.oracle_morton(4, 0, 0, 6) == 64 and L.oracle_morton(1023, 1023, 1023, 6) == (1 << 30) - 1
    assert L.oracle_morton(4095, 0, 0, 7) == sum(3 << (6 * g) for g in range(6))


def test_dda_test_cpp_scenario(ob):
    """test.cpp:78-134: camDir = normalize(11,20,3), camPos = (10.5,12.1,14.7), 500 double-precision
    steps => lastHit = 1, pos.round = (172, 306, 58)."""
    f = np.float32
    v = np.array([11, 20, 3], dtype=np.float32)
    inv = f(1.0) / np.sqrt(f(f(f(v[0] * v[0]) + f(v[1] * v[1])) + f(v[2] * v[2])))
    d = (v * inv).astype(np.float32)
    assert [float(x).hex() for x in d] == ["0x1.e946e20000000p-2", "0x1.bccc140000000p-1", "0x1.0ae0d80000000p-3"]
    o = np.array([10.5, 12.1, 14.7], dtype=np.float32)
    out = np.zeros(4, dtype=np.int32)
    ob.oracle_lib().oracle_dda_walk(o.ctypes.data, d.ctypes.data, 500, 1, out.ctypes.data)
    assert out.tolist() == [172, 306, 58, 1]
    # the fp32 variant of the shader visits the same voxel here (no near-tie in 500 steps)
    ob.oracle_lib().oracle_dda_walk(o.ctypes.data, d.ctypes.data, 500, 0, out.ctypes.data)
    assert out.tolist() == [172, 306, 58, 1]


def test_double_dda_matches_cast_ray_from_cam(ob, ref_world):
    """oracle_dda_walk(use_double) against RAY_CASTER::castRayFromCam on rays that hit nothing
    (so the reference's loop runs all its steps): same voxel after n steps."""
    o = np.array([35.0, 120.5, 35.0], dtype=np.float32)  # above everything
    out = np.zeros(4, dtype=np.int32)
    for yaw, pitch, n in [(45.0, -10.0, 30), (10.0, -40.0, 57), (-100.0, -5.0, 200), (170.0, -60.0, 99)]:
        d = camera_dir(yaw, pitch)
        pos, last, left = ref_world.pick(o, d, n)
        assert left == 0
        ob.oracle_lib().oracle_dda_walk(o.ctypes.data, d.ctypes.data, n, 1, out.ctypes.data)
        assert out[:3].tolist() == pos.tolist()


def test_center_pixel_agrees_with_reference_picker(ob, ref_world, oracle_world, defaults):
    """The reference's CPU ray (double deltas) and the oracle's fp32 shader ray through the screen
    centre reach the same first solid voxel where no near-tie separates them."""
    _, _, sun = defaults
    for origin, yaw, pitch in [((100.5, 60.25, 100.5), 0.0, 60.0), ((100.5, 60.25, 100.5), 180.0, 45.0),
                               ((60.2, 55.9, 140.8), 120.0, 35.0), ((35.0, 50.0, 35.0), 45.0, 30.0)]:
        d = camera_dir(yaw, pitch)
        pos, _, left = ref_world.pick(np.array(origin, dtype=np.float32), d, 300)
        assert left > 0
        # an odd-sized frame has a pixel whose centre is exactly uv = (0.5, 0.5): the camera axis
        r = ob.render(oracle_world, ob.make_frame(3, 3, origin, d, sun, shadows=False))
        h = r["hits"][4]
        assert h["kind"] == ob.KIND_HIT
        assert (int(h["x"]), int(h["y"]), int(h["z"])) == tuple(int(v) for v in pos)
        assert int(h["steps"]) == 300 - left


# ---- deterministic transcendentals and the hemisphere table ----------------------------------------

def test_det_sin_exp_close_to_libm(ob):
    L = ob.oracle_lib()
    xs = np.concatenate([np.linspace(-50, 50, 4001), np.linspace(-700, 700, 2001)]).astype(np.float32)
    s = np.array([L.oracle_sinf(float(v)) for v in xs])
    assert np.abs(s - np.sin(xs.astype(np.float64))).max() < 2e-6
    xe = np.linspace(-20, 20, 4001).astype(np.float32)
    e = np.array([L.oracle_expf(float(v)) for v in xe])
    ref = np.exp(xe.astype(np.float64))
    assert (np.abs(e - ref) / ref).max() < 3e-7
    assert L.oracle_expf(0.0) == 1.0 and L.oracle_sinf(0.0) == 0.0


def test_hemisphere_table_kat(ob, rcmod):
    # gen_hemisphare_distrib.py with num_pts = 20 prints (x, z, y); ALL 20 entries as pasted into
    # light_scattering.frag:133-154 (tests/golden/hemisphere20.json, parsed from the shader by
    # tools/make_golden_hemisphere.py): the oracle's generator and librc's (rc_hemisphere_table, what the AO kernels use)
    # must give exactly those numbers after rounding to fp32
    import json

    with open(os.path.join(os.path.dirname(__file__), "golden", "hemisphere20.json")) as f:
        lit = json.load(f)["literals_x_up_y"]
    assert len(lit) == 20
    want20 = np.array([[np.float32(float(x)), np.float32(float(y)), np.float32(float(up))] for x, up, y in lit], dtype=np.float32)
    tab = np.zeros((20, 3), dtype=np.float32)
    ob.oracle_lib().oracle_hemisphere(20, tab.ctypes.data)
    assert np.array_equal(tab.view(np.uint32), want20.view(np.uint32)), np.argwhere(tab != want20)
    tab2 = np.zeros((20, 3), dtype=np.float32)
    assert rcmod.load_library().rc_hemisphere_table(20, tab2.ctypes.data) == 0
    assert np.array_equal(tab2.view(np.uint32), want20.view(np.uint32)), np.argwhere(tab2 != want20)
    x, y, up = tab[0]
    assert x == np.float32(0.07430756460710652) and up == np.float32(0.97875) and y == np.float32(-0.19111991874778692)
    # the script's formula, evaluated independently in numpy
    for n in (20, 64):
        i = np.arange(n) + 0.5
        phi = np.arccos(1 - i * 0.85 / n)
        theta = np.pi * (1 + 5 ** 0.5) * i
        want = np.stack([np.cos(theta) * np.sin(phi), np.sin(theta) * np.sin(phi), np.cos(phi)], -1)
        tab = np.zeros((n, 3), dtype=np.float32)
        ob.oracle_lib().oracle_hemisphere(n, tab.ctypes.data)
        assert np.abs(tab - want).max() < 1e-6
        assert (tab[:, 2] > 0).all() and np.allclose(np.linalg.norm(tab.astype(np.float64), axis=1), 1, atol=1e-6)


# ---- frames ----------------------------------------------------------------------------------------

def test_primary_ray_statistics_of_the_default_view(ob, oracle_world, defaults):
    """169.76 steps, 178.23 node reads, 7.47 array reads per primary ray, hit fraction 0.469."""
    pos, d, sun = defaults
    r = ob.render(oracle_world, ob.make_frame(1280, 720, pos, d, sun, shadows=False), want_rgba8=False)
    c = r["counters"]
    n = 1280 * 720
    assert c.rays[0] == n
    assert abs(c.dda_steps[0] / n - 169.76) < 0.05
    assert abs(c.node_reads[0] / n - 178.23) < 0.05
    assert abs(c.array_reads[0] / n - 7.47) < 0.05
    assert abs(c.kind_pixels[ob.KIND_HIT] / n - 0.469) < 0.002
    h = r["hits"]
    sky = h["kind"] == ob.KIND_SKY
    assert (h["steps"][sky] == 300).all()
    # step count is the L1 distance between start and end voxel
    l1 = np.abs(h["x"] - 35) + np.abs(h["y"] - 50) + np.abs(h["z"] - 35)
    assert np.array_equal(l1.astype(np.int64), h["steps"].astype(np.int64))


def _render_case(ob, oracle_world, sun, case, rows=None):
    kw = dict(case["kwargs"])
    if "mouse_pos" in kw:
        kw["mouse_pos"] = tuple(kw["mouse_pos"])
    ao = kw.get("ao_samples", 0)
    if ao:
        tab = np.zeros((ao, 3), dtype=np.float32)
        ob.oracle_lib().oracle_hemisphere(ao, tab.ctypes.data)
        ob.oracle_lib().oracle_set_hemisphere(ao, tab.ctypes.data)
    d = np.array(case["camera_dir_bits"], dtype=np.uint32).view(np.float32)
    return ob.render(oracle_world, ob.make_frame(case["width"], case["height"], case["origin"], d, sun, rows=rows, **kw))


def test_oracle_reproduces_golden_frames(ob, oracle_world, defaults, golden):
    meta, samples = golden
    _, _, sun = defaults
    for name, case in meta["cases"].items():
        if case["width"] * case["height"] > 1280 * 720 and case["kwargs"].get("ao_samples", 0):
            continue  # the 1080p AO frame is covered by its sampled rows below
        r = _render_case(ob, oracle_world, sun, case)
        assert sha(r["hits"]) == case["hits_sha256"], name
        assert sha(r["rgba8"]) == case["rgba8_sha256"], name
        idx = samples[name + "_idx"]
        assert np.array_equal(r["hits"][idx], samples[name + "_hits"])
        c = r["counters"].as_dict()
        assert c["rays"] == case["counters"]["rays"] and c["algorithmic_bytes"] == case["counters"]["algorithmic_bytes"]


def test_golden_ao_frame_sampled_rows(ob, oracle_world, defaults, golden):
    meta, samples = golden
    _, _, sun = defaults
    case = meta["cases"]["config3_1080p_ao64"]
    w = case["width"]
    idx = samples["config3_1080p_ao64_idx"]
    rows = np.unique(idx // w)[:24]
    for row in rows:
        r = _render_case(ob, oracle_world, sun, case, rows=(int(row), int(row) + 1))
        sel = idx[idx // w == row]
        k = np.flatnonzero(np.isin(idx, sel))
        assert np.array_equal(r["hits"][sel], samples["config3_1080p_ao64_hits"][k])
        assert np.array_equal(r["rgba8"].reshape(-1, 4)[sel], samples["config3_1080p_ao64_rgba8"][k])


def test_band_sampling_is_a_subset_of_the_full_frame(ob, oracle_world, defaults):
    pos, d, sun = defaults
    full = ob.render(oracle_world, ob.make_frame(200, 100, pos, d, sun, ao_samples=0))
    part = ob.render(oracle_world, ob.make_frame(200, 100, pos, d, sun, ao_samples=0, bands=(16, 3, 1)))
    rows = np.arange(100)
    mine = (rows // 16) % 3 == 1
    assert np.array_equal(part["rgba8"][mine], full["rgba8"][mine])
    assert (part["rgba8"][~mine] == 0).all()
    assert part["counters"].rays[0] == mine.sum() * 200


def test_pixel_zero_is_bottom_left_and_looks_down(ob, oracle_world, defaults):
    pos, d, sun = defaults
    r = ob.render(oracle_world, ob.make_frame(64, 36, pos, d, sun))
    kinds = r["hits"]["kind"].reshape(36, 64)
    assert (kinds[0] == ob.KIND_HIT).all()      # row 0 = bottom of the screen: terrain
    assert (kinds[-1] == ob.KIND_SKY).all()     # top row: sky

"""rc_generate_world: initTetraHexaTree() + genWorld() on the device, straight into the
flattened form.  Held to
  * the reference's own world (oracle/_ref: its unmodified putBlock / genWorld), voxel for voxel through the
    shader's getBlock, and NODE FOR NODE against rc_upload_world's flattening of that world (same breadth-first
    order, same child runs, same voxel runs; materials compared through the two palettes);
  * frames: a frame over the generated world == the frame over the uploaded world, and config 4's full-size
    frame over a generated 2048 x 2048 depth-7 world == the oracle's golden for the host-built one;
  * the host stand-in (libvoxelhost, itself held to the reference in tests/test_host_cpu.py) for other
    seeds / footprints / depths."""
import json
import os

import numpy as np
import pytest

from test_fullsize_gpu import GOLDEN_DIR, HIT_FIELDS, check_against_golden

pytestmark = pytest.mark.gpu

EMPTY = np.uint64(0xFFFFFFFFFFFFFFFF)
NODE = np.dtype([("mask", "<u8"), ("base", "<u4"), ("kind", "<u4")])
PAL = np.dtype([("color", "<u8"), ("flags", "<u4"), ("meta", "<u4")])
PK_INNER, PK_BRICK, PK_LEAF = 0, 1, 2


def read_blob(path):
    """The rc_save_world file (csrc/world_blob.cpp): 64-byte header, nodes, voxels, palette."""
    raw = np.fromfile(path, dtype=np.uint8)
    assert raw[:8].tobytes() == b"RCPWORLD"
    version, depth = np.frombuffer(raw[8:16].tobytes(), dtype="<u4")
    n_nodes, n_voxels, n_pal = (int(v) for v in np.frombuffer(raw[16:40].tobytes(), dtype="<u8"))
    o = 64
    nodes = np.frombuffer(raw[o:o + 16 * n_nodes].tobytes(), dtype=NODE)
    o += 16 * n_nodes
    voxels = np.frombuffer(raw[o:o + 4 * n_voxels].tobytes(), dtype="<u4")
    o += 4 * n_voxels
    pal = np.frombuffer(raw[o:o + 16 * n_pal].tobytes(), dtype=PAL)
    assert o + 16 * n_pal == raw.size and version == 1
    return {"depth": int(depth), "nodes": nodes, "voxels": voxels, "palette": pal}


def assert_same_packed_world(a, b):
    """Same tree, node for node; materials may be numbered differently."""
    assert a["depth"] == b["depth"]
    assert a["nodes"].size == b["nodes"].size and a["voxels"].size == b["voxels"].size
    na, nb = a["nodes"], b["nodes"]
    assert np.array_equal(na["kind"], nb["kind"]), "node kinds (inner / brick / uniform cell, liquid flag) differ"
    assert np.array_equal(na["mask"], nb["mask"]), "child masks differ"
    leaf = (na["kind"] & 0xFF) == PK_LEAF
    assert np.array_equal(na["base"][~leaf], nb["base"][~leaf]), "child / voxel run bases differ"

    def materials(w, words):
        p = w["palette"][words & 0xFFFFFF]
        return np.stack([p["color"], p["flags"].astype(np.uint64) | (p["meta"].astype(np.uint64) << np.uint64(32)),
                         (words >> 24).astype(np.uint64)], axis=1)

    assert np.array_equal(materials(a, a["voxels"]), materials(b, b["voxels"])), "voxel materials differ"
    assert np.array_equal(materials(a, na["base"][leaf]), materials(b, nb["base"][leaf])), "uniform-cell materials differ"
    pa = {(int(e["color"]), int(e["flags"]), int(e["meta"])) for e in a["palette"]}
    pb = {(int(e["color"]), int(e["flags"]), int(e["meta"])) for e in b["palette"]}
    # b (generated) lists every material genWorld / the start-up blocks name, a only those that survived
    assert pa <= pb, "the generated palette lacks a material of the flattened world"


def check_lookups(caster, rcmod, pts, refb):
    words, _ = caster.lookup_device(pts)
    hwords, _, leaves = rcmod.lookup_host(pts, which=0)  # the (lazily fetched) host copy of the generated world
    assert np.array_equal(words, hwords)
    solid = words != 0
    assert np.array_equal(solid, refb[:, 0] != EMPTY), "occupancy differs"
    leaf = leaves.view(np.uint64).reshape(-1, 2)
    assert np.array_equal(leaf[solid], refb[solid]), "materials differ"
    assert np.array_equal((words[solid] >> 24).astype(np.uint64), refb[solid][:, 1] & np.uint64(0xFF))


def probe_points(rng, width, length, top=80):
    box = np.stack(np.meshgrid(np.arange(0, min(width + 8, 72)), np.arange(0, top), np.arange(0, min(length + 8, 72)),
                               indexing="ij"), -1).reshape(-1, 3)
    rnd = np.stack([rng.integers(0, width + 16, 60000), rng.integers(0, 110, 60000), rng.integers(0, length + 16, 60000)], 1)
    startup = [(10, 100, 10), (100, 10, 100), (21, 9, 201), (2, 10, 10), (1000, 1000, 1000)]
    near = [np.stack(np.meshgrid(np.arange(x - 3, x + 7), np.arange(y - 3, y + 7), np.arange(z - 3, z + 7), indexing="ij"),
                     -1).reshape(-1, 3) for (x, y, z) in startup]
    p = np.concatenate([box, rnd] + near).astype(np.int32)
    return p[(p >= 0).all(axis=1)]


def test_generated_default_world_is_the_reference_world(ob, rcmod, caster, ref_world, oracle_world, world_desc, defaults, tmp_path):
    pos, d, sun = defaults
    rng = np.random.default_rng(11)
    p = rcmod.frame_params(960, 540, pos, d, sun, shadows=True, ao_samples=16, hits=True)
    caster.upload_world(world_desc)
    up_info = caster.world_info()
    caster.save_world(tmp_path / "uploaded.rcw")
    caster.render_frame(p)
    ref_img, ref_hits = caster.read_image(), caster.read_hits()
    try:
        gi = caster.generate_world()  # the reference's constants: 200 x 200, seeds 42 / 64 / 100, depth 6
        wi = caster.world_info()
        assert gi.columns == 200 * 200 and gi.min_height >= 0 and gi.max_height < 100
        assert (wi.packed_nodes, wi.packed_voxels) == (up_info.packed_nodes, up_info.packed_voxels)
        assert (gi.nodes, gi.voxels) == (wi.packed_nodes, wi.packed_voxels) and wi.max_depth == 6
        # frames first: the host copy has not been fetched yet, the kernels see only what the generator wrote
        caster.render_frame(p)
        got_hits = caster.read_hits()
        for f in HIT_FIELDS:
            assert np.array_equal(ref_hits[f], got_hits[f]), f"hit field {f} differs from the uploaded world's frame"
        assert np.array_equal(ref_img, caster.read_image())
        # voxel for voxel against the reference's own tree
        pts = probe_points(rng, 200, 200)
        check_lookups(caster, rcmod, pts, ob.get_blocks(oracle_world, pts))
        # node for node against the flattening of the reference's pools
        caster.save_world(tmp_path / "generated.rcw")
        assert_same_packed_world(read_blob(tmp_path / "uploaded.rcw"), read_blob(tmp_path / "generated.rcw"))
        # picking works on it (N2) and an update after it is a full upload (no baseline), like a loaded world
        assert [list(v) for v in caster.pick(pos, d, 30)[:2]] == [list(v) for v in ref_world.pick(pos, d, 30)[:2]]
        info = caster.update_world(world_desc, [], [])
        assert info.full_rebuild == 1
    finally:
        caster.upload_world(world_desc)


def test_generated_2048_world_renders_config4_golden(rcmod, caster, world_desc, defaults):
    from raytracing_test_b200 import voxel_host as vh

    with open(os.path.join(GOLDEN_DIR, "frames_big.json")) as f:
        case = json.load(f)["cases"]["config4_4k_ao64_world2048"]
    samples = np.load(os.path.join(GOLDEN_DIR, "samples_big.npz"))
    pos, d, sun = defaults
    try:
        gi = caster.generate_world(**case["world"])
        assert gi.columns == 2048 * 2048 and caster.world_info().max_depth == 7
        p = rcmod.frame_params(case["width"], case["height"], pos, d, sun, shadows=True, ao_samples=64, hits=True,
                               looking_at=tuple(case["looking_at"]))
        caster.render_frame(p)
        check_against_golden("config4_4k_ao64_world2048", case, samples, caster.read_image(), caster.read_hits())
        # the generator must beat one host core's putBlock loop by a wide margin (it is ~100 ms vs ~10 s)
        assert gi.total_ms < 2000.0, gi.as_dict()
        # sampled voxels against the host-built world
        host = vh.World.default(**case["world"])
        try:
            rng = np.random.default_rng(12)
            pts = np.stack([rng.integers(0, 2060, 400000), rng.integers(0, 90, 400000), rng.integers(0, 2060, 400000)],
                           1).astype(np.int32)
            refb = host.get_blocks(pts)
            words, _ = caster.lookup_device(pts)
            solid = words != 0
            assert np.array_equal(solid, refb[:, 0] != EMPTY)
            assert np.array_equal((words[solid] >> 24).astype(np.uint64), refb[solid][:, 1] & np.uint64(0xFF))
            hi = caster.upload_world(host.describe())
            assert (hi.packed_nodes, hi.packed_voxels) == (gi.nodes, gi.voxels)
        finally:
            host.close()
    finally:
        caster.upload_world(world_desc)


@pytest.mark.parametrize("width,length,depth,seeds,startup", [
    (37, 53, 6, (1, 2, 3), False),      # footprint not a multiple of the brick edge
    (64, 64, 5, (7, 7, 7), True),       # 256^3 torus: the start-up blocks' coordinates keep their low bits only
    (130, 40, 7, (42, 64, 100), True),
    (1, 1, 6, (5, 6, 7), False),        # one column
])
def test_generated_world_equals_host_built_world(rcmod, caster, world_desc, tmp_path, width, length, depth, seeds, startup):
    from raytracing_test_b200 import voxel_host as vh

    host = vh.World(depth)
    if startup:
        host.add_startup_blocks()
    host.generate_terrain(width, length, seeds)
    rng = np.random.default_rng(width * 1000 + length)
    try:
        caster.upload_world(host.describe())
        caster.save_world(tmp_path / "host.rcw")
        caster.generate_world(width, length, depth, seeds, startup)
        caster.save_world(tmp_path / "gen.rcw")
        assert_same_packed_world(read_blob(tmp_path / "host.rcw"), read_blob(tmp_path / "gen.rcw"))
        edge = 1 << (2 * (depth - 1))
        pts = probe_points(rng, width, length)
        pts = pts[(pts < edge).all(axis=1)]
        check_lookups(caster, rcmod, pts, host.get_blocks(pts))
    finally:
        host.close()
        caster.upload_world(world_desc)


def test_generate_world_refuses_what_it_cannot_match(rcmod, caster, world_desc):
    before = caster.world_info()
    for kwargs in [dict(width=0), dict(length=0), dict(max_depth=2), dict(max_depth=9), dict(width=1025),
                   dict(width=64, length=64, max_depth=3)]:  # 16^3 torus: the terrain would wrap around it
        with pytest.raises(rcmod.RcError):
            caster.generate_world(**kwargs)
    # a refused call leaves the installed world alone
    after = caster.world_info()
    assert (before.packed_nodes, before.packed_voxels) == (after.packed_nodes, after.packed_voxels)
    pts = np.array([[100, 30, 100], [10, 100, 10]], dtype=np.int32)
    assert (caster.lookup_device(pts)[0] != 0).any()

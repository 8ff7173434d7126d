"""BASELINE configs 4 and 5 at FULL size on the GPU, held to the oracle's output for exactly those frames
(tests/golden/frames_big.json + samples_big.npz, made by tools/make_golden_big.py in the build container
-- the oracle needs 10-30 s of CPU per frame, too slow to run inside the suite), plus the properties that
do not depend on the size: a frame rendered twice is identical, band-sharded == unsharded.

config 4 is the declared extension of SURVEY 8d: genWorld's formula over 2048 x 2048 columns in a depth-7
tree (4096^3 torus), built by libvoxelhost; a small frame over the same world is also compared with the
oracle directly."""
import hashlib
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
HIT_FIELDS = ["x", "y", "z", "steps", "shadow_steps", "face", "kind", "flags", "ao_occluded"]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.fixture(scope="module")
def big_golden():
    with open(os.path.join(GOLDEN_DIR, "frames_big.json")) as f:
        meta = json.load(f)
    return meta["cases"], np.load(os.path.join(GOLDEN_DIR, "samples_big.npz"))


def check_against_golden(name, case, samples, img, hits):
    idx = samples[name + "_idx"]
    got_h, ref_h = hits[idx], samples[name + "_hits"]
    for f in HIT_FIELDS:  # a readable failure before the hash comparison
        assert np.array_equal(got_h[f], ref_h[f]), f"{name}: sampled hit field {f} differs"
    assert np.array_equal(img.reshape(-1, 4)[idx], samples[name + "_rgba8"]), f"{name}: sampled pixels differ"
    assert sha(hits) == case["hits_sha256"], f"{name}: hit buffer hash differs from the oracle's"
    assert sha(img) == case["rgba8_sha256"], f"{name}: RGBA8 hash differs from the oracle's"


def test_config5_8k_frame_matches_oracle_golden(rcmod, caster, defaults, big_golden):
    cases, samples = big_golden
    name = "config5_8k_ao64"
    case = cases[name]
    pos, d, sun = defaults
    p = rcmod.frame_params(case["width"], case["height"], pos, d, sun, shadows=True, ao_samples=64, hits=True,
                           looking_at=tuple(case["looking_at"]))
    caster.render_frame(p)
    img, hits, st = caster.read_image(), caster.read_hits(), caster.stats()
    check_against_golden(name, case, samples, img, hits)
    c = case["counters"]
    assert [st.primary_rays, st.shadow_rays, st.ao_rays] == c["rays"]
    assert st.primary_steps == c["dda_steps"][0]
    # idempotence: the same frame again, byte for byte
    caster.render_frame(p)
    assert np.array_equal(caster.read_image(), img)


def test_config4_4k_frame_over_2048_world_matches_oracle_golden(ob, rcmod, caster, world_desc, defaults, big_golden):
    from raytracing_test_b200 import voxel_host as vh

    cases, samples = big_golden
    name = "config4_4k_ao64_world2048"
    case = cases[name]
    pos, d, sun = defaults
    host = vh.World.default(**case["world"])
    try:
        info = caster.upload_world(host.describe())
        assert info.max_depth == 7 and info.packed_bytes < info.ref_bytes
        p = rcmod.frame_params(case["width"], case["height"], pos, d, sun, shadows=True, ao_samples=64, hits=True,
                               looking_at=tuple(case["looking_at"]))
        caster.render_frame(p)
        img, hits, st = caster.read_image(), caster.read_hits(), caster.stats()
        check_against_golden(name, case, samples, img, hits)
        assert [st.primary_rays, st.shadow_rays, st.ao_rays] == case["counters"]["rays"]

        # the same world against the oracle itself, at a size it renders in a second: a camera well inside
        # the 2048^2 terrain looking across it, and one outside the reference's 1024^3 address space
        nodes, arrays = host.pools()
        oworld = ob.make_world(nodes, arrays, 7)
        tab = caster.hemisphere_table(16)
        ob.oracle_lib().oracle_set_hemisphere(16, tab.ctypes.data)
        for origin in [(1000.5, 60.5, 1000.5), (1900.25, 45.5, 1500.75), (2047.5, 70.0, 2047.5)]:
            fo = ob.make_frame(480, 270, origin, d, sun, shadows=True, ao_samples=16)
            ref = ob.render(oworld, fo)
            caster.render_frame(rcmod.frame_params(480, 270, origin, d, sun, shadows=True, ao_samples=16, hits=True))
            got = caster.read_hits()
            for f in HIT_FIELDS:
                assert np.array_equal(ref["hits"][f], got[f]), f"depth-7 world at {origin}: hit field {f} differs"
            assert np.array_equal(ref["rgba8"], caster.read_image())

        # band-sharded == unsharded at full size (what config 5's gather composes), here on the big world
        full = img
        part = np.zeros_like(full)
        for r in range(4):
            caster.set_shard(r, 4)
            caster.render_frame(p)
            caster.read_image(part)  # writes this rank's rows in place
        caster.set_shard(0, 1)
        assert np.array_equal(part, full)
    finally:
        caster.set_shard(0, 1)
        caster.upload_world(world_desc)
        host.close()

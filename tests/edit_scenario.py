"""Edit scenario for rc_update_world, run as a SUBPROCESS by the tests (it mutates the reference's
process-global tree, so it must not share a process with the session fixtures):

    python tests/edit_scenario.py cpu|gpu

The world is the reference's own (oracle/_ref: initTetraHexaTree + genWorld); the edits are the
reference's own putBlock / deleteBlock (what src/input.cpp:141-160 does on a click); the dirty-block
lists are what the reference's own updateSsboData() pushes through glBufferSubData
(src/voxel_data/voxel_allocator.hpp:38-78, captured by the GL shim).  After every round the patched
packed world must answer like the reference's getBlock, and (gpu) a frame rendered over it must equal
the oracle's frame over the edited pools, hit record for hit record.
Prints one line "edit scenario ok ..." and exits 0."""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import bindings as ob  # noqa: E402  (checker)
from raytracing_test_b200 import rc  # noqa: E402

HIT_FIELDS = ["x", "y", "z", "steps", "shadow_steps", "face", "kind", "flags", "ao_occluded"]
EMPTY = np.uint64(0xFFFFFFFFFFFFFFFF)


def dirty_blocks(rw):
    """Runs the reference's updateSsboData(); returns (dirty node blocks, dirty array blocks)."""
    L = rw.lib
    n = L.refworld_run_update_ssbo()
    nodes, arrays = [], []
    rec = (ctypes.c_uint64 * 4)()
    for i in range(n):
        L.refworld_upload_record(i, rec)
        assert rec[2] == rw.block_bytes and rec[1] % rw.block_bytes == 0
        (nodes if rec[3] == 1 else arrays).append(int(rec[1] // rw.block_bytes))
    return nodes, arrays


def make_desc(rw):
    rw.refresh()
    nodes, arrays = rw.block_ptrs()
    return rc.make_world_desc(nodes, arrays, rw.node_bytes_used, rw.array_bytes_used, block_bytes=rw.block_bytes,
                              max_depth=rw.max_depth)


def probe_points(rng, extra):
    pts = [rng.integers(0, 200, size=(20000, 3)), rng.integers(0, 64, size=(20000, 3)) + np.array([0, 16, 0])]
    for (x, y, z) in extra:  # everything around each edit
        g = np.stack(np.meshgrid(np.arange(x - 5, x + 6), np.arange(y - 5, y + 6), np.arange(z - 5, z + 6), indexing="ij"), -1)
        pts.append(g.reshape(-1, 3))
    p = np.concatenate(pts).astype(np.int32)
    return p[(p >= 0).all(axis=1) & (p < 768).all(axis=1)]


class SsboView:
    """What the reference's shader would see: the two SSBOs as updateSsboData() fills them -- only blocks
    whose `modified` flag was set are pushed (voxel_allocator.hpp:50-56,69-75), everything else stays as
    last uploaded.  The oracle renders THIS, not the host's live pools."""

    def __init__(self, rw):
        rw.refresh()
        self.nodes, self.arrays = rw.nodes.copy(), rw.arrays.copy()
        self.bb = rw.block_bytes

    def push(self, rw, dirty_nodes, dirty_arrays):
        rw.refresh()
        for (view, live, dirty, item) in ((self, rw.nodes, dirty_nodes, 8), (None, rw.arrays, dirty_arrays, 4)):
            cur = self.nodes if item == 8 else self.arrays
            if live.size > cur.size:  # glBufferData on growth (voxel_allocator.hpp:45-48): new space, old content kept
                grown = np.zeros(live.size, dtype=cur.dtype)
                grown[: cur.size] = cur
                cur = grown
            per = self.bb // item
            for b in (range(live.size // per) if dirty is None else dirty):
                cur[b * per:(b + 1) * per] = live[b * per:(b + 1) * per]
            if item == 8:
                self.nodes = cur
            else:
                self.arrays = cur

    def world(self, rw):
        return ob.make_world(self.nodes, self.arrays, rw.max_depth)


def check_lookup(oworld, pts, lookup):
    refb = ob.get_blocks(oworld, pts)  # the shader's getBlock (low_res.frag:493-532) over the SSBO view
    words, leaves = lookup(pts)
    solid = words != 0
    assert np.array_equal(solid, refb[:, 0] != EMPTY), "occupancy differs from the shader's getBlock over the SSBOs"
    leaf = leaves.view(np.uint64).reshape(-1, 2)
    assert np.array_equal(leaf[solid], refb[solid]), "materials differ"
    # traversal flags travel in the word's top byte
    assert np.array_equal((words[solid] >> 24).astype(np.uint64), refb[solid][:, 1] & np.uint64(0xFF))


def rounds(rw, rng):
    """Yields (label, list of edited voxels) after applying each batch of edits to the reference tree."""
    L = rw.lib
    pos, _, _ = rw.defaults()
    down = np.array([1.0, -0.6, 1.0], dtype=np.float32)
    down /= np.float32(np.sqrt((down * down).sum()))
    hit, last, left = rw.pick(pos, down, 300)  # the block under a crosshair aimed at the ground
    assert left > 0
    hit, last = [int(v) for v in hit], [int(v) for v in last]
    red = L.refworld_rgb_to_u64(255, 30, 30)
    # The two click handlers of src/input.cpp:141-160.  Deletes use level maxDepth-1, the level at which
    # deleteBlock frees ONE voxel: the handler's own `deleteBlock(pos, 6)` overruns `Ptr stack[maxDepth]`
    # (tetrahexa_tree.cpp:300,343: stack[6]) and `bitmap ^= 1 << index` (:352) is an int shift, wrong for
    # child indices >= 32 -- reference defects; the scenario stays on the well-defined side (z & 3 < 2).
    vox = rw.max_depth - 1
    assert hit[2] & 3 < 2
    L.refworld_delete_block(*hit, vox)
    yield "delete the picked voxel", [hit]
    L.refworld_put_block(*last, 0x2, red, 0.0, rw.max_depth)  # a mirror where the ray came from
    yield "place a reflective voxel", [last]
    L.refworld_put_block(60, 72, 60, 0x0, L.refworld_rgb_to_u64(10, 200, 10), 0.0, rw.max_depth - 1)  # 4^3 uniform block
    yield "place a 4^3 block in the air", [(60, 72, 60)]
    L.refworld_delete_block(61, 73, 61, vox)  # splits the uniform leaf into 64 voxels
    yield "carve one voxel out of it", [(61, 73, 61)]
    L.refworld_put_block(40, 52, 40, 0x1 | 0x4 | 0x10, L.refworld_rgb_to_u64(20, 40, 220), 0.0, rw.max_depth)  # water in sight
    L.refworld_put_block(300, 300, 300, 0x0, red, 0.0, rw.max_depth)  # far away: new branches at every level
    yield "water + a voxel in an empty 256^3 cell", [(40, 52, 40), (300, 300, 300)]
    edits = []
    for _ in range(300):
        x, z = int(rng.integers(20, 90)), int(rng.integers(20, 90))
        y = int(rng.integers(25, 58))
        if rng.random() < 0.5:
            z &= ~2
            L.refworld_delete_block(x, y, z, vox)
        else:
            L.refworld_put_block(x, y, z, 0x0, L.refworld_rgb_to_u64(int(rng.integers(0, 256)), 99, 7), 0.0, rw.max_depth)
        edits.append((x, y, z))
    yield "300 random edits in view", edits


def main():
    mode = sys.argv[1] if len(sys.argv) > 1 else "cpu"
    rng = np.random.default_rng(5)
    rw = ob.RefWorld()
    dirty_blocks(rw)  # the start-up upload (src/main.cpp:212) clears every modified flag
    desc = make_desc(rw)
    caster = None
    if mode == "gpu":
        caster = rc.RayCaster(0)
        info = caster.update_world(desc, [], [])
        assert info.full_rebuild == 1  # first call: nothing to diff against

        def lookup(pts):
            w, _ = caster.lookup_device(pts)
            _, _, leaves = rc.lookup_host(pts, which=0)
            return w, leaves
    else:
        rc.flatten_host_tracked(desc)

        def lookup(pts):
            w, _, leaves = rc.lookup_host(pts, which=1)
            return w, leaves

    ssbo = SsboView(rw)
    check_lookup(ssbo.world(rw), probe_points(rng, []), lookup)
    pos, d, sun = rw.defaults()
    total = {"rounds": 0, "reflattened": 0, "bytes": 0, "stale_rounds": 0}

    def verify(label, edited):
        oworld = ssbo.world(rw)
        check_lookup(oworld, probe_points(rng, edited), lookup)
        if caster:
            tab = caster.hemisphere_table(16)
            ob.oracle_lib().oracle_set_hemisphere(16, tab.ctypes.data)
            ref = ob.render(oworld, ob.make_frame(480, 270, pos, d, sun, shadows=True, ao_samples=16))
            caster.render_frame(rc.frame_params(480, 270, pos, d, sun, shadows=True, ao_samples=16, hits=True))
            got = caster.read_hits()
            for f in HIT_FIELDS:
                assert np.array_equal(ref["hits"][f], got[f]), f"{label}: hit field {f} differs from the oracle"
            assert np.array_equal(ref["rgba8"], caster.read_image()), label

    def update(dn, da):
        return caster.update_world(desc, dn, da) if caster else rc.update_host(desc, dn, da)

    for label, edited in rounds(rw, rng):
        dn, da = dirty_blocks(rw)
        desc = make_desc(rw)
        if not dn and not da:
            # deleteBlock's final branch (tetrahexa_tree.cpp:348-353) sets no `modified` flag: the reference
            # pushes nothing and its shader keeps seeing the old world.  Same here ...
            info = update(dn, da)
            assert info.full_rebuild == 0 and info.reflattened == 0
            verify(label + " (not pushed: stale on both sides)", edited)
            total["stale_rounds"] += 1
            # ... until the host asks for a diff of every block
            dn, da = None, None
        ssbo.push(rw, dn, da)
        info = update(dn, da)
        assert info.full_rebuild == 0, f"{label}: fell back to a full rebuild"
        assert info.changed_ref_nodes + info.changed_ref_arrays > 0 and info.reflattened > 0, label
        # a click touches a handful of packed subtrees, not the world
        assert info.reflattened <= 4 * len(edited) + 4, (label, info.as_dict())
        if caster:
            assert info.uploaded_bytes < 2_000_000, (label, info.as_dict())
        verify(label, edited)
        total["rounds"] += 1
        total["reflattened"] += info.reflattened
        total["bytes"] += info.uploaded_bytes
    assert total["stale_rounds"] >= 1
    # bring everything in line (the random round's deletions are only partly pushed), then: an update with
    # nothing modified is a no-op; a dirty list naming unchanged blocks finds no change
    desc = make_desc(rw)
    ssbo.push(rw, None, None)
    info = update(None, None)
    assert info.full_rebuild == 0
    verify("diff of every block", [])
    info = update([], [])
    assert info.full_rebuild == 0 and info.reflattened == 0
    info = update([0, 1], [0])
    assert info.full_rebuild == 0 and info.changed_ref_nodes == 0 and info.reflattened == 0
    # and the patched world equals a fresh flatten of the same pools, voxel for voxel
    pts = probe_points(rng, [(60, 72, 60), (40, 52, 40)])
    patched_words, patched_leaves = lookup(pts)
    rc.flatten_host(desc)
    fresh_words, _, fresh_leaves = rc.lookup_host(pts, which=1)
    assert np.array_equal(patched_words != 0, fresh_words != 0) and np.array_equal(patched_leaves, fresh_leaves)
    if caster:
        caster.close()
    print(f"edit scenario ok ({mode}): {total}")


if __name__ == "__main__":
    main()

"""CPU tests of the product's HOST side: the C-ABI surface of librc.so / libvoxelhost.so, the
flattener (reference pools -> packed 64-tree) and the host stand-in (world generation, block
store, picking ray).  No compute call needs a GPU here."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EMPTY = np.uint64(0xFFFFFFFFFFFFFFFF)


def declared_symbols(header, prefix):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(%s[a-z0-9_]+)\s*\(" % prefix, text)))


# ---- ABI surface -----------------------------------------------------------------------------------

def test_librc_exports_every_declared_symbol(rcmod):
    lib = rcmod.load_library()
    names = declared_symbols("rc.h", "rc_") + declared_symbols("rc_debug.h", "rc_")
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"librc.so does not export {n}"
        assert n in rcmod.ABI, f"rc.py does not bind {n}"
    assert lib.rc_abi_version() == 1


def test_libvoxelhost_exports_every_declared_symbol():
    from raytracing_test_b200 import voxel_host as vh

    lib = vh.load_library()
    for n in declared_symbols("voxel_host.h", "vh_"):
        assert hasattr(lib, n), f"libvoxelhost.so does not export {n}"
        assert n in vh.ABI


def test_struct_layouts_match_the_header(rcmod):
    sizes = np.zeros(5, dtype=np.uint32)
    assert rcmod.load_library().rc_debug_struct_sizes(sizes.ctypes.data) == 0
    mine = [ctypes.sizeof(rcmod.WorldDesc), ctypes.sizeof(rcmod.WorldInfo), ctypes.sizeof(rcmod.FrameParams),
            rcmod.HIT_DTYPE.itemsize, ctypes.sizeof(rcmod.FrameStats)]
    assert sizes.tolist() == mine == [56, 72, 104, 20, 64]


def test_calls_before_init_fail_loudly(rcmod):
    lib = rcmod.load_library()
    lib.rc_shutdown()
    p = rcmod.frame_params(8, 8, (0, 0, 0), (0, 0, 1), (0, 1, 0))
    assert lib.rc_render_frame(ctypes.byref(p)) == -1  # RC_ERR_NOT_INITIALIZED
    assert b"rc_init" in lib.rc_last_error()
    assert lib.rc_sync() == -1
    buf = np.zeros(256, dtype=np.uint8)
    assert lib.rc_read_image(buf.ctypes.data, buf.nbytes, 0) == -1


def test_no_cpu_fallback_without_a_device(rcmod):
    """On a machine without a CUDA device rc_init must fail (RC_ERR_CUDA), never fall back."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    lib = rcmod.load_library()
    assert lib.rc_init(0) == -3
    assert b"no CPU path" in lib.rc_last_error()
    with pytest.raises(rcmod.RcError):
        rcmod.RayCaster(0)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "raytracing_test_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f
                assert not re.search(r'#include\s+"[^"]*oracle', text), f


# ---- flattener ---------------------------------------------------------------------------------------

def test_flatten_default_world_matches_reference_getblock(rcmod, ref_world, world_desc):
    info = rcmod.flatten_host(world_desc)
    assert info.ref_bytes == ref_world.node_bytes_used + ref_world.array_bytes_used
    assert info.max_depth == 6 and info.null_children == 1  # root cell 63
    assert info.packed_bytes < info.ref_bytes // 4
    # every voxel of the terrain box + random probes (outside [768,1023]^3 where the reference exits)
    box = np.stack(np.meshgrid(np.arange(0, 208), np.arange(0, 72), np.arange(0, 208), indexing="ij"), -1).reshape(-1, 3)
    rng = np.random.default_rng(5)
    xyz = np.concatenate([box, rng.integers(0, 768, size=(200_000, 3)), [[10, 100, 10], [21, 9, 201], [1000, 10, 10]]])
    xyz = xyz.astype(np.int32)
    words, shifts, leaves = rcmod.lookup_host(xyz, which=1)
    ref = ref_world.get_blocks(xyz)
    solid = ref[:, 0] != EMPTY
    assert np.array_equal(words != 0, solid)
    assert np.array_equal(leaves.view(np.uint64).reshape(-1, 2)[solid], ref[solid])
    # the material word carries the low 8 flag bits (traversal decisions need no palette load)
    assert np.array_equal((words[solid] >> 24).astype(np.uint64), ref[solid, 1] & np.uint64(0xFF))
    assert solid.sum() > 1_000_000


def test_flatten_uniform_cells_are_really_uniform(rcmod, ref_world, world_desc):
    """cell_shift contract of the lookup: an empty answer at shift s is empty for the whole aligned
    4^s... cell (that is what lets the kernel skip lookups while it steps inside the cell)."""
    rcmod.flatten_host(world_desc)
    rng = np.random.default_rng(9)
    xyz = rng.integers(0, 260, size=(3000, 3)).astype(np.int32)
    words, shifts, _ = rcmod.lookup_host(xyz, which=1)
    for (x, y, z), w, s in zip(xyz[:400], words[:400], shifts[:400]):
        if w != 0 or s == 0:
            continue
        size = 1 << int(s)
        base = np.array([x, y, z]) & ~(size - 1)
        probe = base + rng.integers(0, size, size=(64, 3))
        corners = base + np.array([[0, 0, 0], [size - 1] * 3, [size - 1, 0, 0], [0, size - 1, size - 1]])
        pts = np.concatenate([probe, corners]).astype(np.int32)
        assert (ref_world.get_blocks(pts)[:, 0] == EMPTY).all(), (x, y, z, s)
    assert (shifts[words == 0] >= 0).all() and shifts.max() <= 8


def test_flatten_follows_the_morton_quirk_for_negative_coordinates(rcmod, world_desc):
    rcmod.flatten_host(world_desc)
    neg = np.array([[-1, 30, 5], [-257, 30, 5], [-700, 30, 5], [5, 30, -1000], [35 + 1024, 20, 35 + 2048]], dtype=np.int32)
    img = np.array([[1023, 30, 5], [1023, 30, 5], [768 | (-700 & 255), 30, 5], [5, 30, 768 | (-1000 & 255)], [35, 20, 35]],
                   dtype=np.int32)
    a, b = rcmod.lookup_host(neg, which=1), rcmod.lookup_host(img, which=1)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    assert b[0][-1] != 0


def test_flatten_rejects_malformed_worlds(rcmod, ref_world):
    nodes, arrays = ref_world.nodes.copy(), ref_world.arrays.copy()

    def flatten(n, a, **kw):
        d = rcmod.world_desc_from_flat(n, a, ref_world.node_bytes_used, ref_world.array_bytes_used, **kw)
        return rcmod.flatten_host(d)

    flatten(nodes, arrays)  # sanity: the copy is fine
    with pytest.raises(rcmod.RcError) as e:
        flatten(nodes, arrays, max_depth=1)
    assert e.value.code in (-2, -4)
    bad = nodes.copy()
    bad[1] = (np.uint64(0x00FFFFFF) << np.uint64(32))  # root's child array index far outside the pool
    with pytest.raises(rcmod.RcError) as e:
        flatten(bad, arrays)
    assert e.value.code == -4
    bad_arrays = arrays.copy()
    root_array = int(nodes[1] >> np.uint64(32))
    bad_arrays[64 * root_array + 0] = 0x7FFFFFF0  # child node index outside the node pool
    # an out-of-pool child is "no child" (empty), exactly as the oracle treats it (divergence (1))
    assert flatten(nodes, bad_arrays).null_children == 2
    cyc = arrays.copy()
    cyc[64 * root_array + 0] = 0  # ... and so is child index 0, the root itself
    assert flatten(nodes, cyc).null_children == 2
    lib = rcmod.load_library()
    assert lib.rc_debug_flatten_host(None, None) == -2


def test_flatten_tiny_worlds(rcmod):
    from raytracing_test_b200 import voxel_host as vh

    # empty world: root with no children
    w = vh.World(6)
    info = rcmod.flatten_host(w.describe())
    assert info.packed_nodes == 1 and info.packed_voxels == 0
    words, shifts, _ = rcmod.lookup_host(np.array([[0, 0, 0], [500, 500, 500]], dtype=np.int32))
    assert (words == 0).all() and (shifts == 8).all()
    # one voxel, one 4^3 block, one 16^3 block
    w.put_block(5, 6, 7, color=123, flags=vh.REFLECTIVE)
    w.put_block(64, 64, 64, color=77, level=5)
    w.put_block(128, 0, 0, color=99, level=4)
    rcmod.flatten_host(w.describe())
    pts = np.array([[5, 6, 7], [4, 6, 7], [64, 64, 64], [67, 67, 67], [68, 64, 64], [128, 0, 0], [143, 15, 15], [144, 0, 0]],
                   dtype=np.int32)
    words, shifts, leaves = rcmod.lookup_host(pts)
    assert (words != 0).tolist() == [True, False, True, True, False, True, True, False]
    assert leaves[0, 0] == 123 and leaves[0, 2] == (vh.REFLECTIVE | 1)
    assert leaves[2, 0] == 77 and leaves[5, 0] == 99
    assert shifts[5] == 4 and shifts[6] == 4  # the 16^3 block stays ONE uniform cell in the packed form


# ---- host stand-in (libvoxelhost) vs the reference's own host code ------------------------------------

@pytest.fixture(scope="module")
def vh_world():
    from raytracing_test_b200 import voxel_host as vh

    return vh.World.default()


def test_simplex_noise_bit_exact(ref_world):
    from raytracing_test_b200 import voxel_host as vh

    for seed, scale in [(42, 0.005), (64, 0.05), (100, 0.1), (7, 0.37)]:
        a = vh.noise2_grid(seed, scale, 128, 96)
        b = np.zeros((128, 96))
        ref_world.lib.refworld_noise2_grid(seed, scale, 128, 96, b.ctypes.data)
        assert np.array_equal(a.view(np.uint64), b.view(np.uint64)), seed
    assert vh.noise2(42, -3.25, 1e3) == ref_world.lib.refworld_noise2(42, -3.25, 1e3)


def test_generated_world_equals_reference_world(rcmod, ref_world, vh_world, world_desc):
    box = np.stack(np.meshgrid(np.arange(0, 208), np.arange(0, 112), np.arange(0, 208), indexing="ij"), -1).reshape(-1, 3)
    rng = np.random.default_rng(1)
    xyz = np.concatenate([box, rng.integers(0, 768, size=(100_000, 3))]).astype(np.int32)
    assert np.array_equal(vh_world.get_blocks(xyz), ref_world.get_blocks(xyz))
    # and the flattened forms are identical node for node
    i_ref = rcmod.flatten_host(world_desc).as_dict()
    probe = xyz[::5]
    w_ref = rcmod.lookup_host(probe)
    i_vh = rcmod.flatten_host(vh_world.describe()).as_dict()
    w_vh = rcmod.lookup_host(probe)
    for k in ("ref_nodes", "packed_nodes", "packed_voxels", "packed_materials", "packed_bytes"):
        assert i_ref[k] == i_vh[k], k
    assert i_vh["null_children"] == 0  # the stand-in does not reproduce the root-array aliasing (F11)
    assert all(np.array_equal(a, b) for a, b in zip(w_ref, w_vh))


def test_defaults_colours_and_picker_match_reference(ref_world, vh_world):
    from raytracing_test_b200 import voxel_host as vh
    from conftest import CAMERAS, camera_dir

    for a, b in zip(vh.default_camera(), ref_world.defaults()):
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    lib = vh.load_library()
    for rgb in [(33, 33, 33), (45, 18, 0), (0, 150, 10), (255, 255, 255), (1, 2, 3)]:
        assert lib.vh_rgb(*rgb) == ref_world.lib.refworld_rgb_to_u64(*rgb)
    for origin, yaw, pitch in CAMERAS:
        if min(origin) < 0:
            continue  # the reference's getBlock is only defined in range
        d = camera_dir(yaw, pitch)
        for steps in (30, 120):
            pa, la, sa = vh_world.pick(origin, d, steps)
            pb, lb, sb = ref_world.pick(origin, d, steps)
            assert pa.tolist() == pb.tolist() and la.tolist() == lb.tolist() and sa == sb


def test_put_block_overwrite_and_split(vh_world):
    from raytracing_test_b200 import voxel_host as vh

    w = vh.World(6)
    w.put_block(16, 16, 16, color=5, level=4)               # a 16^3 block
    assert w.get_block(31, 31, 31)[0] == 5 and w.get_block(32, 16, 16) is None
    w.put_block(17, 18, 19, color=9, flags=vh.LIQUID)       # carve one voxel: the block splits
    assert w.get_block(17, 18, 19)[:2] == (9, vh.LIQUID | 1)
    assert w.get_block(16, 16, 16)[0] == 5 and w.get_block(31, 31, 31)[0] == 5 and w.get_block(17, 18, 20)[0] == 5
    w.put_block(16, 16, 16, color=6, level=4)               # overwrite the whole block again
    assert w.get_block(17, 18, 19)[0] == 6
    with pytest.raises(ValueError):
        w.put_block(0, 0, 0, level=9)


def test_deep_world_extension(rcmod):
    """max_depth 7 (4096^3 address space: BASELINE config 4's declared extension)."""
    from raytracing_test_b200 import voxel_host as vh

    w = vh.World.default(width=300, length=260, max_depth=7)
    info = rcmod.flatten_host(w.describe())
    assert info.max_depth == 7
    rng = np.random.default_rng(2)
    xyz = np.concatenate([rng.integers(0, 300, size=(50_000, 3)), rng.integers(0, 4096, size=(20_000, 3)),
                          [[299, 25, 259], [1500 + 4096, 30, 20]]]).astype(np.int32)
    words, _, leaves = rcmod.lookup_host(xyz)
    ref = w.get_blocks(xyz & 4095)
    solid = ref[:, 0] != EMPTY
    assert np.array_equal(words != 0, solid)
    assert np.array_equal(leaves.view(np.uint64).reshape(-1, 2)[solid], ref[solid])
    # same terrain as the depth-6 world where they overlap
    w6 = vh.World.default(width=300, length=260, max_depth=6)
    probe = rng.integers(0, 260, size=(50_000, 3)).astype(np.int32)
    a, b = w.get_blocks(probe), w6.get_blocks(probe)
    same = ~((probe[:, 0] >= 20) & (probe[:, 0] < 36) & (probe[:, 2] >= 200) & (probe[:, 2] < 216))  # start-up block level differs
    assert np.array_equal(a[same], b[same])


def test_incremental_update_host_side_edit_scenario():
    """rc_update_world's host half (diff of the modified blocks, re-emission of the changed runs) over the
    reference's own putBlock / deleteBlock and updateSsboData(): tests/edit_scenario.py, in a subprocess
    because it edits the reference's process-global tree."""
    import subprocess
    import sys

    here = os.path.dirname(os.path.abspath(__file__))
    if not os.path.exists(os.path.join(here, "..", "oracle", "_ref", "libref_world.so")):
        pytest.skip("oracle/_ref/libref_world.so not built")
    r = subprocess.run([sys.executable, os.path.join(here, "edit_scenario.py"), "cpu"], capture_output=True, text=True,
                       timeout=300)
    assert r.returncode == 0 and "edit scenario ok" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


def test_world_blob_round_trip_and_validation(rcmod, world_desc, tmp_path):
    """rc_save_world / rc_load_world host halves: a saved flattened world loads back voxel for voxel;
    truncated, bit-flipped and foreign files are refused (RC_ERR_MALFORMED_WORLD), never installed."""
    info = rcmod.flatten_host(world_desc)
    rng = np.random.default_rng(3)
    pts = np.concatenate([rng.integers(0, 256, size=(50000, 3)), rng.integers(-50, 1100, size=(5000, 3))]).astype(np.int32)
    w0, s0, l0 = rcmod.lookup_host(pts, which=1)
    path = str(tmp_path / "world.rcpw")
    rcmod.save_host(path)
    blob = open(path, "rb").read()
    assert blob[:8] == b"RCPWORLD" and len(blob) == 64 + info.packed_bytes
    rcmod.flatten_host(rcmod.world_desc_from_flat(np.zeros(2, dtype=np.uint64), np.zeros(64, dtype=np.uint32)))  # forget it
    back = rcmod.load_host(path)
    assert (back.packed_nodes, back.packed_voxels, back.packed_materials, back.max_depth) == (
        info.packed_nodes, info.packed_voxels, info.packed_materials, info.max_depth)
    w1, s1, l1 = rcmod.lookup_host(pts, which=1)
    assert np.array_equal(w0, w1) and np.array_equal(s0, s1) and np.array_equal(l0, l1)
    for name, data in [("short", blob[:-5]), ("long", blob + b"x"), ("magic", b"NOTAWRLD" + blob[8:]),
                       ("flip", blob[:5000] + bytes([blob[5000] ^ 0x10]) + blob[5001:]), ("empty", b"")]:
        bad = str(tmp_path / f"bad_{name}.rcpw")
        open(bad, "wb").write(data)
        with pytest.raises(rcmod.RcError) as e:
            rcmod.load_host(bad)
        assert e.value.code == -4, name
    with pytest.raises(rcmod.RcError):
        rcmod.load_host(str(tmp_path / "does_not_exist.rcpw"))


def test_world_blob_with_a_valid_checksum_but_inconsistent_levels_is_refused(rcmod, tmp_path):
    """The kernels pick the array a child run is read from by tree LEVEL (fetch_region, world_tops_kernel) or by node kind
    (lookup): a blob whose checksum is right but whose kinds contradict the levels -- a brick above the brick level, an inner
    node at it, a header depth that shifts every level -- would send them out of bounds.  load walks the tree from the
    root and refuses those."""
    import struct

    from raytracing_test_b200 import voxel_host as vh

    host = vh.World(6)
    for (x, y, z) in [(5, 5, 5), (70, 3, 9), (300, 40, 2), (301, 40, 2)]:
        host.put_block(x, y, z, 0x123456, 0, 0.0)
    host.put_block(64, 64, 64, 0x5, 0, 0.0, 4)  # a 16^3 uniform cell: a LEAF child
    rcmod.flatten_host(host.describe())
    path = str(tmp_path / "small.rcpw")
    rcmod.save_host(path)
    blob = bytearray(open(path, "rb").read())
    n_nodes, n_voxels, n_palette = struct.unpack_from("<QQQ", blob, 16)
    assert rcmod.load_host(path).packed_nodes == n_nodes

    def fnv(data):
        h = 0xCBF29CE484222325
        for b in data:
            h = ((h ^ b) * 0x100000001B3) & 0xFFFFFFFFFFFFFFFF
        return h

    def variant(name, edit):
        b = bytearray(blob)
        edit(b)
        struct.pack_into("<Q", b, 56, fnv(bytes(b[64:])))  # the checksum covers the payload: make it right again
        out = str(tmp_path / f"{name}.rcpw")
        open(out, "wb").write(b)
        return out

    kinds = [struct.unpack_from("<I", blob, 64 + 16 * i + 12)[0] & 0xFF for i in range(n_nodes)]
    inner = [i for i, k in enumerate(kinds) if k == 0]
    brick = [i for i, k in enumerate(kinds) if k == 1]
    assert inner and brick and kinds[0] == 0

    def set_kind(i, k):
        return lambda b: struct.pack_into("<I", b, 64 + 16 * i + 12, k)

    cases = {
        "brick_above_brick_level": set_kind(inner[-1], 1),
        "inner_at_brick_level": set_kind(brick[0], 0),
        "leaf_root": set_kind(0, 2),
        "unknown_kind": set_kind(brick[-1], 7),
        "shallower_header": lambda b: struct.pack_into("<I", b, 12, 5),
        "deeper_header": lambda b: struct.pack_into("<I", b, 12, 7),
        "child_run_outside": lambda b: struct.pack_into("<I", b, 64 + 16 * inner[-1] + 8, n_nodes),
    }
    for name, edit in cases.items():
        with pytest.raises(rcmod.RcError) as e:
            rcmod.load_host(variant(name, edit))
        assert e.value.code == -4, name
    # the untouched blob still loads (the variants did not poison anything)
    assert rcmod.load_host(variant("same", lambda b: None)).packed_nodes == n_nodes
    host.close()

"""oracle/bindings.py -- TEST INFRASTRUCTURE ONLY.

ctypes bindings for the two checker libraries:

* ``oracle/_build/liboracle.so``  -- CPU restatement of src/shaders/low_res.frag
  (oracle_render.cpp)
* ``oracle/_ref/libref_world.so`` -- the reference's own unmodified CPU sources
  (world builder, getBlock/putBlock, CPU picking ray), compiled by oracle/Makefile
* ``oracle/_ref/libref_shader.so`` -- the reference's fragment shader src/shaders/low_res.frag compiled as C++
  (oracle/ref_shader.cpp + oracle/shim/glsl/), one instance per pixel on the host cores

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.  The product package never does.
"""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "_build", "liboracle.so")
REF_SO = os.path.join(HERE, "_ref", "libref_world.so")
REF_SHADER_SO = os.path.join(HERE, "_ref", "libref_shader.so")

F_CROSSHAIR = 1
F_SHADOWS = 2

KIND_SKY, KIND_HIT, KIND_CAMERA_INSIDE, KIND_CROSSHAIR, KIND_HIGHLIGHT, KIND_DEBUG = range(6)
HF_SHADOWED, HF_REFLECTED, HF_REFRACTED, HF_FACING_SUN = 1, 2, 4, 8

# Byte layout shared with rc_hit (include/rc.h)
HIT_DTYPE = np.dtype(
    [
        ("x", "<i4"), ("y", "<i4"), ("z", "<i4"),
        ("steps", "<u2"), ("shadow_steps", "<u2"),
        ("face", "u1"), ("kind", "u1"), ("flags", "u1"), ("ao_occluded", "u1"),
    ]
)
assert HIT_DTYPE.itemsize == 20


class OracleWorld(ctypes.Structure):
    _fields_ = [
        ("nodes", ctypes.c_void_p), ("n_nodes", ctypes.c_uint64),
        ("arrays", ctypes.c_void_p), ("n_arrays", ctypes.c_uint64),
        ("max_depth", ctypes.c_int32), ("_pad", ctypes.c_int32),
    ]


class OracleFrame(ctypes.Structure):
    _fields_ = [
        ("origin", ctypes.c_float * 3), ("camera_dir", ctypes.c_float * 3), ("sun_dir", ctypes.c_float * 3),
        ("mouse_pos", ctypes.c_float * 2), ("proj_plane", ctypes.c_float * 2),
        ("delta_time", ctypes.c_float), ("ao_strength", ctypes.c_float),
        ("looking_at", ctypes.c_int32 * 3), ("render_pos_data", ctypes.c_int32),
        ("width", ctypes.c_uint32), ("height", ctypes.c_uint32),
        ("max_steps", ctypes.c_uint32), ("shadow_steps", ctypes.c_uint32),
        ("ao_samples", ctypes.c_uint32), ("ao_steps", ctypes.c_uint32),
        ("flags", ctypes.c_uint32), ("row_begin", ctypes.c_uint32), ("row_end", ctypes.c_uint32),
        ("threads", ctypes.c_int32),
        ("band_rows", ctypes.c_uint32), ("band_stride", ctypes.c_uint32), ("band_offset", ctypes.c_uint32),
    ]


class OracleCounters(ctypes.Structure):
    _fields_ = [
        ("node_reads", ctypes.c_uint64 * 3), ("array_reads", ctypes.c_uint64 * 3),
        ("dda_steps", ctypes.c_uint64 * 3), ("rays", ctypes.c_uint64 * 3),
        ("kind_pixels", ctypes.c_uint64 * 6),
    ]

    def as_dict(self):
        return {
            "node_reads": list(self.node_reads), "array_reads": list(self.array_reads),
            "dda_steps": list(self.dda_steps), "rays": list(self.rays),
            "kind_pixels": list(self.kind_pixels),
            "algorithmic_bytes": [16 * self.node_reads[i] + 4 * self.array_reads[i] for i in range(3)],
        }


def build(which=("oracle", "ref")):
    """Runs oracle/Makefile (the ref target is a no-op when /root/reference is absent)."""
    for target in which:
        subprocess.run(["make", "-C", HERE, target], check=True, stdout=subprocess.DEVNULL)


_oracle = None
_ref = None


def oracle_lib():
    global _oracle
    if _oracle is None:
        if not os.path.exists(ORACLE_SO):
            build(("oracle",))
        L = ctypes.CDLL(ORACLE_SO)
        L.oracle_render.restype = ctypes.c_int
        L.oracle_render.argtypes = [ctypes.c_void_p] * 6
        L.oracle_get_block.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, ctypes.c_void_p]
        L.oracle_get_block_batch.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p]
        L.oracle_morton.restype = ctypes.c_uint64
        L.oracle_morton.argtypes = [ctypes.c_int32] * 4
        L.oracle_morton_ref10.restype = ctypes.c_uint32
        L.oracle_morton_ref10.argtypes = [ctypes.c_int32] * 3
        L.oracle_dda_walk.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_int, ctypes.c_void_p]
        L.oracle_sinf.restype = ctypes.c_float
        L.oracle_sinf.argtypes = [ctypes.c_float]
        L.oracle_expf.restype = ctypes.c_float
        L.oracle_expf.argtypes = [ctypes.c_float]
        L.oracle_hemisphere.argtypes = [ctypes.c_uint32, ctypes.c_void_p]
        L.oracle_set_hemisphere.argtypes = [ctypes.c_uint32, ctypes.c_void_p]
        L.oracle_last_threads.restype = ctypes.c_int
        _oracle = L
    return _oracle


def ref_lib():
    """The reference's own code. Raises FileNotFoundError if it was never built here."""
    global _ref
    if _ref is None:
        if not os.path.exists(REF_SO):
            if os.path.isdir("/root/reference/src"):
                build(("ref",))
            else:
                raise FileNotFoundError(REF_SO)
        L = ctypes.CDLL(REF_SO)
        L.refworld_node_block.restype = ctypes.c_void_p
        L.refworld_array_block.restype = ctypes.c_void_p
        L.refworld_node_block.argtypes = [ctypes.c_uint32]
        L.refworld_array_block.argtypes = [ctypes.c_uint32]
        L.refworld_rgb_to_u64.restype = ctypes.c_uint64
        L.refworld_noise2.restype = ctypes.c_double
        L.refworld_noise2.argtypes = [ctypes.c_int64, ctypes.c_double, ctypes.c_double]
        L.refworld_noise2_grid.argtypes = [ctypes.c_int64, ctypes.c_double, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
        L.refworld_get_block.argtypes = [ctypes.c_int] * 3 + [ctypes.c_void_p] * 3
        L.refworld_get_block_batch.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p]
        L.refworld_put_block.argtypes = [ctypes.c_int] * 3 + [ctypes.c_uint32, ctypes.c_uint64, ctypes.c_float, ctypes.c_int]
        L.refworld_delete_block.argtypes = [ctypes.c_int] * 4
        L.refworld_pick.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
        L.refworld_pick_batch.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_int, ctypes.c_void_p]
        L.refworld_copy_nodes.argtypes = [ctypes.c_void_p]
        L.refworld_copy_arrays.argtypes = [ctypes.c_void_p]
        L.refworld_defaults.argtypes = [ctypes.c_void_p] * 3
        L.refworld_upload_record.argtypes = [ctypes.c_uint32, ctypes.c_void_p]
        L.refworld_upload_src.restype = ctypes.c_void_p
        L.refworld_upload_src.argtypes = [ctypes.c_uint32]
        _ref = L
    return _ref


class RefWorld:
    """The default world exactly as the reference host builds it (src/main.cpp:181-190)."""

    def __init__(self, gen_world=True):
        L = ref_lib()
        L.refworld_build(1 if gen_world else 0)
        self.lib = L
        self.block_bytes = L.refworld_block_bytes()
        self.refresh()

    def refresh(self):
        L = self.lib
        self.node_bytes_used = L.refworld_node_bytes_used()
        self.array_bytes_used = L.refworld_array_bytes_used()
        nb, ab = L.refworld_num_node_blocks(), L.refworld_num_array_blocks()
        self.nodes = np.zeros(nb * self.block_bytes // 8, dtype=np.uint64)
        self.arrays = np.zeros(ab * self.block_bytes // 4, dtype=np.uint32)
        L.refworld_copy_nodes(self.nodes.ctypes.data)
        L.refworld_copy_arrays(self.arrays.ctypes.data)
        self.max_depth = L.refworld_max_depth()

    def block_ptrs(self):
        L = self.lib
        nodes = [L.refworld_node_block(i) for i in range(L.refworld_num_node_blocks())]
        arrays = [L.refworld_array_block(i) for i in range(L.refworld_num_array_blocks())]
        return nodes, arrays

    def get_block(self, x, y, z):
        c, f, m = ctypes.c_uint64(), ctypes.c_uint32(), ctypes.c_float()
        self.lib.refworld_get_block(x, y, z, ctypes.byref(c), ctypes.byref(f), ctypes.byref(m))
        return c.value, f.value, m.value

    def get_blocks(self, xyz):
        """Reference getBlock for n voxels -> (n,2) u64 {color, flags | metadata<<32}."""
        xyz = np.ascontiguousarray(xyz, dtype=np.int32).reshape(-1, 3)
        out = np.zeros((len(xyz), 2), dtype=np.uint64)
        self.lib.refworld_get_block_batch(xyz.ctypes.data, len(xyz), out.ctypes.data)
        return out

    def defaults(self):
        pos, d, s = (ctypes.c_float * 3)(), (ctypes.c_float * 3)(), (ctypes.c_float * 3)()
        self.lib.refworld_defaults(pos, d, s)
        return (np.array(pos, dtype=np.float32), np.array(d, dtype=np.float32), np.array(s, dtype=np.float32))

    def pick(self, cam_pos, cam_dir, steps=30):
        p = np.asarray(cam_pos, dtype=np.float32)
        d = np.asarray(cam_dir, dtype=np.float32)
        out = np.zeros(7, dtype=np.int32)
        self.lib.refworld_pick(p.ctypes.data, d.ctypes.data, steps, out.ctypes.data)
        return out[0:3].copy(), out[3:6].copy(), int(out[6])


def make_world(nodes, arrays, max_depth=6):
    """OracleWorld over flat pool arrays (kept alive by the caller)."""
    w = OracleWorld()
    w.nodes = nodes.ctypes.data
    w.n_nodes = nodes.size // 2
    w.arrays = arrays.ctypes.data
    w.n_arrays = arrays.size // 64
    w.max_depth = max_depth
    return w


def make_frame(width, height, origin, camera_dir, sun_dir, *, shadows=True, ao_samples=0, ao_steps=5,
               ao_strength=1.0, mouse_pos=None, looking_at=(1 << 30, 1 << 30, 1 << 30), delta_time=0.0,
               render_pos_data=0, max_steps=300, shadow_steps=75, rows=None, threads=0, proj_plane=None,
               bands=None):
    f = OracleFrame()
    f.origin[:] = [float(v) for v in origin]
    f.camera_dir[:] = [float(v) for v in camera_dir]
    f.sun_dir[:] = [float(v) for v in sun_dir]
    flags = 0
    if mouse_pos is not None:
        f.mouse_pos[:] = [float(v) for v in mouse_pos]
        flags |= F_CROSSHAIR
    if shadows:
        flags |= F_SHADOWS
    if proj_plane is None:
        proj_plane = default_proj_plane(width, height)
    f.proj_plane[:] = [float(v) for v in proj_plane]
    f.delta_time = delta_time
    f.ao_strength = ao_strength
    f.looking_at[:] = list(looking_at)
    f.render_pos_data = render_pos_data
    f.width, f.height = width, height
    f.max_steps, f.shadow_steps = max_steps, shadow_steps
    f.ao_samples, f.ao_steps = ao_samples, ao_steps
    f.flags = flags
    f.row_begin, f.row_end = (0, height) if rows is None else rows
    f.threads = threads
    if bands is not None:  # (band_rows, stride, offset)
        f.band_rows, f.band_stride, f.band_offset = bands
    return f


def default_proj_plane(width, height):
    """src/main.cpp:94: evaluated in double on the host, passed as two floats."""
    import math

    t = math.tan(math.radians(45.0))
    # glm::tan(glm::radians(45.0)) * ((float)height)/width : double * float / int
    return (np.float32(t), np.float32(t * float(np.float32(height)) / width))


def render(world, frame, want_hits=True, want_rgba8=True, want_f32=False):
    L = oracle_lib()
    n = frame.width * frame.height
    hits = np.zeros(n, dtype=HIT_DTYPE) if want_hits else None
    rgba8 = np.zeros((frame.height, frame.width, 4), dtype=np.uint8) if want_rgba8 else None
    f32 = np.zeros((frame.height, frame.width, 4), dtype=np.float32) if want_f32 else None
    counters = OracleCounters()
    rc = L.oracle_render(
        ctypes.addressof(world), ctypes.addressof(frame),
        hits.ctypes.data if hits is not None else None,
        rgba8.ctypes.data if rgba8 is not None else None,
        f32.ctypes.data if f32 is not None else None,
        ctypes.addressof(counters),
    )
    if rc != 0:
        raise RuntimeError(f"oracle_render failed: {rc}")
    return {"hits": hits, "rgba8": rgba8, "rgba32f": f32, "counters": counters}


def get_blocks(world, xyz):
    """Oracle getBlock (shader semantics) for n voxels -> (n,2) u64 {block.x, block.y}."""
    xyz = np.ascontiguousarray(xyz, dtype=np.int32).reshape(-1, 3)
    out = np.zeros((len(xyz), 2), dtype=np.uint64)
    oracle_lib().oracle_get_block_batch(ctypes.addressof(world), xyz.ctypes.data, len(xyz), out.ctypes.data)
    return out


# ---------------------------------------------------------------------------------------------------------
# The reference's own fragment shader on the host cores (oracle/ref_shader.cpp)
class RefShaderFrame(ctypes.Structure):
    _fields_ = [
        ("origin", ctypes.c_float * 3), ("camera_dir", ctypes.c_float * 3), ("sun_dir", ctypes.c_float * 3),
        ("mouse_pos", ctypes.c_float * 2), ("proj_plane", ctypes.c_float * 2), ("delta_time", ctypes.c_float),
        ("looking_at", ctypes.c_int32 * 3), ("render_pos_data", ctypes.c_int32),
        ("width", ctypes.c_uint32), ("height", ctypes.c_uint32),
        ("row_begin", ctypes.c_uint32), ("row_end", ctypes.c_uint32),
        ("col_stride", ctypes.c_uint32), ("row_stride", ctypes.c_uint32),
        ("threads", ctypes.c_int32),
    ]


REF_SHADER_STATE = np.dtype(
    [
        ("round", "<i4", 3), ("exact", "<f4", 3), ("last_hit", "<u4"), ("reflected", "<u4"),
        ("block_empty", "<u4"), ("refractive_index", "<f4"), ("color_mod", "<f4", 3),
    ]
)

_ref_shader = None
_MATH_FN = ctypes.CFUNCTYPE(ctypes.c_float, ctypes.c_float)


def ref_shader_lib():
    """low_res.frag as compiled C++.  Raises FileNotFoundError if it was never built here."""
    global _ref_shader
    if _ref_shader is None:
        if not os.path.exists(REF_SHADER_SO):
            if os.path.isdir("/root/reference/src"):
                build(("ref",))
            else:
                raise FileNotFoundError(REF_SHADER_SO)
        L = ctypes.CDLL(REF_SHADER_SO)
        L.refshader_set_world.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.refshader_set_math.argtypes = [ctypes.c_void_p] * 3
        L.refshader_render.restype = ctypes.c_longlong
        L.refshader_render.argtypes = [ctypes.c_void_p] * 3
        L.refshader_last_threads.restype = ctypes.c_int
        assert L.refshader_sizeof_frame() == ctypes.sizeof(RefShaderFrame)
        assert L.refshader_sizeof_state() == REF_SHADER_STATE.itemsize
        _ref_shader = L
    return _ref_shader


class RefShader:
    """The reference's shader over flat copies of the host pools (what updateSsboData puts in the two SSBOs,
    src/voxel_data/voxel_allocator.hpp:38-91).  `nodes` / `arrays` are kept alive here."""

    def __init__(self, nodes, arrays, oracle_math=True):
        self.lib = ref_shader_lib()
        self.nodes = np.ascontiguousarray(nodes, dtype=np.uint64)
        self.arrays = np.ascontiguousarray(arrays, dtype=np.uint32)
        self.use_oracle_math(oracle_math)

    def use_oracle_math(self, on):
        """sin / exp of the shader: the oracle's deterministic ones (bit-for-bit colour comparisons) or libm."""
        if on:
            O = oracle_lib()
            addr = lambda f: ctypes.cast(f, ctypes.c_void_p)
            self.lib.refshader_set_math(addr(O.oracle_sinf), None, addr(O.oracle_expf))
        else:
            self.lib.refshader_set_math(None, None, None)

    def render(self, width, height, origin, camera_dir, sun_dir, *, mouse_pos=None,
               looking_at=(1 << 30, 1 << 30, 1 << 30), delta_time=0.0, render_pos_data=0, rows=None,
               col_stride=1, row_stride=1, threads=0, proj_plane=None, want_state=True):
        f = RefShaderFrame()
        f.origin[:] = [float(v) for v in origin]
        f.camera_dir[:] = [float(v) for v in camera_dir]
        f.sun_dir[:] = [float(v) for v in sun_dir]
        # the reference always evaluates the crosshair test; "disabled" = a mouse position far off the frame
        f.mouse_pos[:] = [float(v) for v in (mouse_pos if mouse_pos is not None else (-1.0e6, -1.0e6))]
        f.proj_plane[:] = [float(v) for v in (proj_plane or default_proj_plane(width, height))]
        f.delta_time = delta_time
        f.looking_at[:] = list(looking_at)
        f.render_pos_data = render_pos_data
        f.width, f.height = width, height
        f.row_begin, f.row_end = (0, height) if rows is None else rows
        f.col_stride, f.row_stride, f.threads = col_stride, row_stride, threads
        rgba = np.zeros((height, width, 4), dtype=np.float32)
        state = np.zeros((height, width), dtype=REF_SHADER_STATE) if want_state else None
        self.lib.refshader_set_world(self.nodes.ctypes.data, self.arrays.ctypes.data)
        n = self.lib.refshader_render(ctypes.addressof(f), rgba.ctypes.data,
                                      state.ctypes.data if state is not None else None)
        if n < 0:
            raise RuntimeError("refshader_render failed")
        return {"rgba32f": rgba, "state": state, "fragments": int(n), "threads": int(self.lib.refshader_last_threads())}

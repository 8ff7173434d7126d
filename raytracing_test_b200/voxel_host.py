"""ctypes binding of libvoxelhost.so (include/voxel_host.h): the stand-in for the host side the
reference keeps (64-tree block store, OpenSimplex terrain, CPU picking ray).

bench.py and the tests use it to produce worlds in the reference's pool wire format on
machines without /root/reference; rc_upload_world ingests them exactly as it would ingest
the real host's nodeBlocks[] / arrayBlocks[].
"""
import ctypes
import os

import numpy as np

from . import rc

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG_DIR, "libvoxelhost.so")

REFLECTIVE, REFRACTIVE, LUMINESCENT, LIQUID = 0x2, 0x4, 0x8, 0x10
DEFAULT_SEEDS = (42, 64, 100)  # src/world_gen.cpp:15-17


class Block(ctypes.Structure):
    _fields_ = [("color", ctypes.c_uint64), ("flags", ctypes.c_uint32), ("metadata", ctypes.c_float)]


ABI = {
    "vh_create": (ctypes.c_void_p, [ctypes.c_uint32]),
    "vh_destroy": (None, [ctypes.c_void_p]),
    "vh_put_block": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32,
                                    ctypes.POINTER(Block), ctypes.c_int]),
    "vh_get_block": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32,
                                    ctypes.POINTER(Block)]),
    "vh_get_block_batch": (None, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p]),
    "vh_add_startup_blocks": (None, [ctypes.c_void_p]),
    "vh_generate_terrain": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32, ctypes.c_void_p]),
    "vh_describe": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(rc.WorldDesc)]),
    "vh_rgb": (ctypes.c_uint64, [ctypes.c_int, ctypes.c_int, ctypes.c_int]),
    "vh_default_camera": (None, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]),
    "vh_pick": (None, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]),
    "vh_noise2": (ctypes.c_double, [ctypes.c_int64, ctypes.c_double, ctypes.c_double]),
    "vh_noise2_grid": (None, [ctypes.c_int64, ctypes.c_double, ctypes.c_int32, ctypes.c_int32, ctypes.c_void_p]),
    "vh_copy_pools": (None, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]),
}

_lib = None


def load_library(path=LIB_PATH):
    global _lib
    if _lib is None:
        if not os.path.exists(path):
            raise FileNotFoundError(f"{path} not built: run `make -C raytracing_test_b200/csrc`")
        lib = ctypes.CDLL(path)
        for name, (restype, argtypes) in ABI.items():
            fn = getattr(lib, name)
            fn.restype = restype
            fn.argtypes = argtypes
        _lib = lib
    return _lib


def default_camera():
    lib = load_library()
    pos, d, sun = (np.zeros(3, dtype=np.float32) for _ in range(3))
    lib.vh_default_camera(pos.ctypes.data, d.ctypes.data, sun.ctypes.data)
    return pos, d, sun


class World:
    """A 64-tree block store in the reference's wire format. max_depth 6 = the reference."""

    def __init__(self, max_depth=6):
        self.lib = load_library()
        self.handle = self.lib.vh_create(max_depth)
        if not self.handle:
            raise ValueError(f"vh_create({max_depth}) failed")
        self.max_depth = max_depth

    @classmethod
    def default(cls, width=200, length=200, max_depth=6, seeds=DEFAULT_SEEDS):
        """initTetraHexaTree() + genWorld() (src/main.cpp:181-190) for a width x length footprint."""
        w = cls(max_depth)
        w.add_startup_blocks()
        w.generate_terrain(width, length, seeds)
        return w

    def close(self):
        if self.handle:
            self.lib.vh_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def put_block(self, x, y, z, color=0, flags=0, metadata=0.0, level=None):
        b = Block(color, flags, metadata)
        r = self.lib.vh_put_block(self.handle, x, y, z, ctypes.byref(b), self.max_depth if level is None else level)
        if r != 0:
            raise ValueError("vh_put_block: bad arguments")

    def get_block(self, x, y, z):
        b = Block()
        if self.lib.vh_get_block(self.handle, x, y, z, ctypes.byref(b)):
            return b.color, b.flags, b.metadata
        return None

    def get_blocks(self, xyz):
        """(n,2) u64 {color, flags | metadata_bits << 32}; empty = (~0, 0) like the shader's getBlock."""
        xyz = np.ascontiguousarray(xyz, dtype=np.int32).reshape(-1, 3)
        out = np.zeros((len(xyz), 2), dtype=np.uint64)
        self.lib.vh_get_block_batch(self.handle, xyz.ctypes.data, len(xyz), out.ctypes.data)
        return out

    def add_startup_blocks(self):
        self.lib.vh_add_startup_blocks(self.handle)

    def generate_terrain(self, width, length, seeds=DEFAULT_SEEDS):
        s = np.asarray(seeds, dtype=np.int64)
        if self.lib.vh_generate_terrain(self.handle, width, length, s.ctypes.data) != 0:
            raise ValueError("vh_generate_terrain: bad arguments")

    def describe(self):
        """rc_world_desc pointing into the store's blocks (valid until the next mutation)."""
        d = rc.WorldDesc()
        if self.lib.vh_describe(self.handle, ctypes.byref(d)) != 0:
            raise ValueError("vh_describe failed")
        d._keep = (self,)
        return d

    def pools(self):
        """Flat copies of the two pools (the SSBO view): (nodes u64[2n], arrays u32[64m])."""
        d = self.describe()
        nodes = np.zeros(d.n_node_blocks * d.block_bytes // 8, dtype=np.uint64)
        arrays = np.zeros(max(d.n_array_blocks, 1) * d.block_bytes // 4, dtype=np.uint32)
        self.lib.vh_copy_pools(self.handle, nodes.ctypes.data, arrays.ctypes.data)
        return nodes, arrays

    def pick(self, cam_pos, cam_dir, steps=30):
        """RAY_CASTER::castRayFromCam: (pos, lastPos, steps_left)."""
        p = np.asarray(cam_pos, dtype=np.float32)
        d = np.asarray(cam_dir, dtype=np.float32)
        out = np.zeros(7, dtype=np.int32)
        self.lib.vh_pick(self.handle, p.ctypes.data, d.ctypes.data, steps, out.ctypes.data)
        return out[0:3].copy(), out[3:6].copy(), int(out[6])


def noise2(seed, x, y):
    return load_library().vh_noise2(seed, x, y)


def noise2_grid(seed, scale, w, l):
    out = np.zeros((w, l), dtype=np.float64)
    load_library().vh_noise2_grid(seed, scale, w, l, out.ctypes.data)
    return out

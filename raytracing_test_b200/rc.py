"""ctypes binding of librc.so (include/rc.h) -- plumbing for tests and bench.py.

The product is the shared library; this module only marshals numpy buffers into
the C-ABI the reference's C++ host would call.  It fails loudly when the library
is missing or when no CUDA device is usable: there is no CPU fallback.
"""
import ctypes
import math
import os

import numpy as np

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RC_LIBRC") or os.path.join(PKG_DIR, "librc.so")  # RC_LIBRC: a variant build (kernel tuning)

RC_F_CROSSHAIR, RC_F_SHADOWS, RC_F_HITS, RC_F_RGBA32F, RC_F_NO_FREE_FLIGHT, RC_F_NO_LEAP, RC_F_ALWAYS_TAIL = 1, 2, 4, 8, 16, 32, 64
RC_F_KERNEL_TIMES = 128
RC_F_NO_COLUMN_REGIONS = 256
RC_F_NO_DROP_GATES = 512
# rc_debug_set_option("schedule", ...): how a frame is scheduled on the device (csrc/kernels.h FrameSchedule)
SCHED_TWO_KERNELS, SCHED_AO_PERSISTENT, SCHED_FUSED, SCHED_PERSISTENT, SCHED_PERSISTENT_STAGED, SCHED_AUTO = range(6)
SCHEDULES = {"two_kernels": 0, "ao_persistent": 1, "fused": 2, "persistent": 3, "persistent_staged": 4, "auto": 5}
AO_OPT_OCTANT = 1
RC_RGBA8, RC_RGBA32F = 0, 1
RC_IPC_HANDLE_BYTES = 64
RC_ALL_BLOCKS = 0xFFFFFFFF
RC_STATIC_MAX_PIXELS = 262144
KIND_SKY, KIND_HIT, KIND_CAMERA_INSIDE, KIND_CROSSHAIR, KIND_HIGHLIGHT, KIND_DEBUG = range(6)
HF_SHADOWED, HF_REFLECTED, HF_REFRACTED, HF_FACING_SUN = 1, 2, 4, 8
BAND_ROWS = 16

HIT_DTYPE = np.dtype(
    [
        ("x", "<i4"), ("y", "<i4"), ("z", "<i4"),
        ("steps", "<u2"), ("shadow_steps", "<u2"),
        ("face", "u1"), ("kind", "u1"), ("flags", "u1"), ("ao_occluded", "u1"),
    ]
)


class RcError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"librc error {code}: {message}")
        self.code = code


class WorldDesc(ctypes.Structure):
    _fields_ = [
        ("node_blocks", ctypes.POINTER(ctypes.c_void_p)),
        ("array_blocks", ctypes.POINTER(ctypes.c_void_p)),
        ("n_node_blocks", ctypes.c_uint32), ("n_array_blocks", ctypes.c_uint32),
        ("block_bytes", ctypes.c_uint32), ("root_index", ctypes.c_uint32),
        ("node_bytes_used", ctypes.c_uint64), ("array_bytes_used", ctypes.c_uint64),
        ("max_depth", ctypes.c_uint32), ("reserved", ctypes.c_uint32),
    ]


class WorldInfo(ctypes.Structure):
    _fields_ = [
        ("ref_nodes", ctypes.c_uint64), ("ref_bytes", ctypes.c_uint64),
        ("packed_nodes", ctypes.c_uint64), ("packed_voxels", ctypes.c_uint64),
        ("packed_materials", ctypes.c_uint64), ("packed_bytes", ctypes.c_uint64),
        ("max_depth", ctypes.c_uint32), ("null_children", ctypes.c_uint32),
        ("flatten_ms", ctypes.c_double), ("upload_ms", ctypes.c_double),
    ]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class UpdateInfo(ctypes.Structure):
    _fields_ = [
        ("full_rebuild", ctypes.c_uint32), ("changed_ref_nodes", ctypes.c_uint32),
        ("changed_ref_arrays", ctypes.c_uint32), ("reflattened", ctypes.c_uint32),
        ("appended_nodes", ctypes.c_uint64), ("appended_voxels", ctypes.c_uint64),
        ("garbage_nodes", ctypes.c_uint64), ("garbage_voxels", ctypes.c_uint64),
        ("uploaded_bytes", ctypes.c_uint64), ("host_ms", ctypes.c_double), ("upload_ms", ctypes.c_double),
    ]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class GenParams(ctypes.Structure):
    """rc_gen_params (include/rc.h): what initTetraHexaTree() + genWorld() hard-code."""
    _fields_ = [("width", ctypes.c_uint32), ("length", ctypes.c_uint32), ("max_depth", ctypes.c_uint32),
                ("startup_blocks", ctypes.c_uint32), ("seeds", ctypes.c_int64 * 3)]


class GenInfo(ctypes.Structure):
    _fields_ = [("columns", ctypes.c_uint64), ("candidate_bricks", ctypes.c_uint64), ("bricks", ctypes.c_uint64),
                ("nodes", ctypes.c_uint64), ("voxels", ctypes.c_uint64), ("min_height", ctypes.c_int32),
                ("max_height", ctypes.c_int32), ("heights_ms", ctypes.c_float), ("build_ms", ctypes.c_float),
                ("total_ms", ctypes.c_float)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class FrameParams(ctypes.Structure):
    _fields_ = [
        ("origin", ctypes.c_float * 3), ("camera_dir", ctypes.c_float * 3), ("sun_dir", ctypes.c_float * 3),
        ("mouse_pos", ctypes.c_float * 2), ("proj_plane", ctypes.c_float * 2),
        ("delta_time", ctypes.c_float), ("ao_strength", ctypes.c_float),
        ("looking_at", ctypes.c_int32 * 3), ("render_pos_data", ctypes.c_int32),
        ("width", ctypes.c_uint32), ("height", ctypes.c_uint32),
        ("max_steps", ctypes.c_uint32), ("shadow_steps", ctypes.c_uint32),
        ("ao_samples", ctypes.c_uint32), ("ao_steps", ctypes.c_uint32),
        ("flags", ctypes.c_uint32),
    ]


class FrameStats(ctypes.Structure):
    _fields_ = [
        ("primary_rays", ctypes.c_uint64), ("shadow_rays", ctypes.c_uint64), ("ao_rays", ctypes.c_uint64),
        ("hit_pixels", ctypes.c_uint64), ("primary_steps", ctypes.c_uint64),
        ("kernel_launches", ctypes.c_uint32), ("free_flight_rays", ctypes.c_uint32),
        ("primary_ms", ctypes.c_float), ("ao_ms", ctypes.c_float), ("total_ms", ctypes.c_float),
        ("schedule", ctypes.c_uint32),
    ]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_ if not k.startswith("reserved")}


# name -> (restype, argtypes); every symbol include/rc.h and include/rc_debug.h declare
ABI = {
    "rc_init": (ctypes.c_int, [ctypes.c_int]),
    "rc_init_multi": (ctypes.c_int, [ctypes.c_int, ctypes.c_void_p]),
    "rc_device_count": (ctypes.c_int, []),
    "rc_build_id": (ctypes.c_char_p, []),
    "rc_shutdown": (None, []),
    "rc_last_error": (ctypes.c_char_p, []),
    "rc_abi_version": (ctypes.c_int, []),
    "rc_set_stream": (ctypes.c_int, [ctypes.c_void_p]),
    "rc_set_shard": (ctypes.c_int, [ctypes.c_uint32, ctypes.c_uint32]),
    "rc_shard_layout": (ctypes.c_int, [ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p]),
    "rc_compose_frame": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32,
                                         ctypes.c_uint32]),
    "rc_upload_world": (ctypes.c_int, [ctypes.POINTER(WorldDesc)]),
    "rc_world_info_get": (ctypes.c_int, [ctypes.POINTER(WorldInfo)]),
    "rc_update_world": (ctypes.c_int, [ctypes.POINTER(WorldDesc), ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_uint32,
                                   ctypes.POINTER(UpdateInfo)]),
    "rc_save_world": (ctypes.c_int, [ctypes.c_char_p]),
    "rc_load_world": (ctypes.c_int, [ctypes.c_char_p]),
    "rc_write_ppm": (ctypes.c_int, [ctypes.c_char_p]),
    "rc_generate_world": (ctypes.c_int, [ctypes.POINTER(GenParams), ctypes.POINTER(GenInfo)]),
    "rc_render_frame": (ctypes.c_int, [ctypes.POINTER(FrameParams)]),
    "rc_render_frame_static": (ctypes.c_int, [ctypes.POINTER(FrameParams), ctypes.c_void_p, ctypes.c_size_t]),
    "rc_read_image": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int]),
    "rc_read_image_async": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_size_t, ctypes.POINTER(ctypes.c_int)]),
    "rc_wait_image": (ctypes.c_int, [ctypes.c_int]),
    "rc_host_register": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_size_t]),
    "rc_host_unregister": (ctypes.c_int, [ctypes.c_void_p]),
    "rc_host_fence": (None, []),
    "rc_frame_target_create": (ctypes.c_int, [ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_void_p]),
    "rc_frame_target_open": (ctypes.c_int, [ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_void_p]),
    "rc_frame_target_close": (ctypes.c_int, []),
    "rc_frame_target_device": (ctypes.c_void_p, [ctypes.c_uint32]),
    "rc_frame_target_last_slot": (ctypes.c_int, []),
    "rc_frame_sync_create": (ctypes.c_int, [ctypes.c_void_p]),
    "rc_frame_sync_open": (ctypes.c_int, [ctypes.c_void_p]),
    "rc_frame_sync_close": (ctypes.c_int, []),
    "rc_frame_sync_begin": (ctypes.c_int, []),
    "rc_frame_sync_allow": (ctypes.c_int, [ctypes.c_uint32, ctypes.c_void_p]),
    "rc_frame_sync_end": (ctypes.c_int, []),
    "rc_frame_sync_frames": (ctypes.c_uint32, []),
    "rc_frame_sync_timeouts": (ctypes.c_int, [ctypes.c_void_p]),
    "rc_read_hits": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_size_t]),
    "rc_sync": (ctypes.c_int, []),
    "rc_frame_stats_get": (ctypes.c_int, [ctypes.POINTER(FrameStats)]),
    "rc_device_image": (ctypes.c_void_p, [ctypes.c_int]),
    "rc_device_hits": (ctypes.c_void_p, []),
    "rc_hemisphere_table": (ctypes.c_int, [ctypes.c_uint32, ctypes.c_void_p]),
    "rc_pick": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]),
    # rc_debug.h
    "rc_debug_flatten_host": (ctypes.c_int, [ctypes.POINTER(WorldDesc), ctypes.POINTER(WorldInfo)]),
    "rc_debug_flatten_host_tracked": (ctypes.c_int, [ctypes.POINTER(WorldDesc), ctypes.POINTER(WorldInfo)]),
    "rc_debug_update_host": (ctypes.c_int, [ctypes.POINTER(WorldDesc), ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_uint32,
                                   ctypes.POINTER(UpdateInfo)]),
    "rc_debug_save_host": (ctypes.c_int, [ctypes.c_char_p]),
    "rc_debug_load_host": (ctypes.c_int, [ctypes.c_char_p, ctypes.POINTER(WorldInfo)]),
    "rc_debug_lookup_host": (ctypes.c_int, [ctypes.c_int, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32,
                                             ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]),
    "rc_debug_lookup_host_batch": (ctypes.c_int, [ctypes.c_int, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p,
                                                   ctypes.c_void_p, ctypes.c_void_p]),
    "rc_debug_lookup_device": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p]),
    "rc_debug_palette": (ctypes.c_int, [ctypes.c_uint32, ctypes.c_void_p]),
    "rc_debug_struct_sizes": (ctypes.c_int, [ctypes.c_void_p]),
    "rc_debug_set_option": (ctypes.c_int, [ctypes.c_char_p, ctypes.c_int]),
    "rc_debug_reread_image": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_uint32, ctypes.c_void_p]),
    "rc_debug_detmath": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p]),
}

_lib = None


def load_library(path=LIB_PATH):
    """Loads librc.so and types every exported symbol. Raises if the library is absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(path):
            raise FileNotFoundError(
                f"{path} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a). There is no CPU fallback."
            )
        lib = ctypes.CDLL(path)
        for name, (restype, argtypes) in ABI.items():
            fn = getattr(lib, name)  # AttributeError if the symbol is missing
            fn.restype = restype
            fn.argtypes = argtypes
        _lib = lib
    return _lib


def _check(lib, code):
    if code != 0:
        raise RcError(code, lib.rc_last_error().decode("utf-8", "replace"))


def make_world_desc(node_block_ptrs, array_block_ptrs, node_bytes_used, array_bytes_used, *, block_bytes=1 << 22,
                    root_index=0, max_depth=6):
    """rc_world_desc over raw host block pointers (ints). Keeps the pointer arrays alive on the struct."""
    d = WorldDesc()
    nb = (ctypes.c_void_p * len(node_block_ptrs))(*node_block_ptrs)
    ab = (ctypes.c_void_p * len(array_block_ptrs))(*array_block_ptrs)
    d._keep = (nb, ab)
    d.node_blocks = ctypes.cast(nb, ctypes.POINTER(ctypes.c_void_p))
    d.array_blocks = ctypes.cast(ab, ctypes.POINTER(ctypes.c_void_p))
    d.n_node_blocks, d.n_array_blocks = len(node_block_ptrs), len(array_block_ptrs)
    d.block_bytes, d.root_index = block_bytes, root_index
    d.node_bytes_used, d.array_bytes_used = node_bytes_used, array_bytes_used
    d.max_depth = max_depth
    return d


def world_desc_from_flat(nodes_u64, arrays_u32, node_bytes_used=None, array_bytes_used=None, *, block_bytes=1 << 22,
                         max_depth=6):
    """rc_world_desc over flat numpy pools (the SSBO view) by slicing them into blocks."""
    nbytes, abytes = nodes_u64.nbytes, arrays_u32.nbytes
    nptr, aptr = nodes_u64.ctypes.data, arrays_u32.ctypes.data
    node_blocks = [nptr + o for o in range(0, max(nbytes, 1), block_bytes)]
    array_blocks = [aptr + o for o in range(0, max(abytes, 1), block_bytes)]
    d = make_world_desc(node_blocks, array_blocks,
                        nbytes if node_bytes_used is None else node_bytes_used,
                        abytes if array_bytes_used is None else array_bytes_used,
                        block_bytes=block_bytes, max_depth=max_depth)
    d._keep = d._keep + (nodes_u64, arrays_u32)
    return d


def default_proj_plane(width, height):
    """projPlaneSize as render() computes it (src/main.cpp:94): double maths, passed as floats."""
    t = math.tan(math.radians(45.0))
    return (float(np.float32(t)), float(np.float32(t * float(np.float32(height)) / width)))


def frame_params(width, height, origin, camera_dir, sun_dir, *, shadows=True, ao_samples=0, ao_steps=5,
                 ao_strength=1.0, mouse_pos=None, looking_at=(1 << 30, 1 << 30, 1 << 30), delta_time=0.0,
                 render_pos_data=0, max_steps=300, shadow_steps=75, hits=False, rgba32f=False, proj_plane=None, free_flight=True, leap=True,
                 always_tail=False, kernel_times=False, column_regions=True, drop_gates=True):
    p = FrameParams()
    p.origin[:] = [float(v) for v in origin]
    p.camera_dir[:] = [float(v) for v in camera_dir]
    p.sun_dir[:] = [float(v) for v in sun_dir]
    flags = 0
    if mouse_pos is not None:
        p.mouse_pos[:] = [float(v) for v in mouse_pos]
        flags |= RC_F_CROSSHAIR
    if shadows:
        flags |= RC_F_SHADOWS
    if hits:
        flags |= RC_F_HITS
    if rgba32f:
        flags |= RC_F_RGBA32F
    if not free_flight:
        flags |= RC_F_NO_FREE_FLIGHT
    if not leap:
        flags |= RC_F_NO_LEAP
    if always_tail:
        flags |= RC_F_ALWAYS_TAIL
    if kernel_times:
        flags |= RC_F_KERNEL_TIMES
    if not column_regions:
        flags |= RC_F_NO_COLUMN_REGIONS
    if not drop_gates:
        flags |= RC_F_NO_DROP_GATES
    pp = default_proj_plane(width, height) if proj_plane is None else proj_plane
    p.proj_plane[:] = [float(pp[0]), float(pp[1])]
    p.delta_time, p.ao_strength = delta_time, ao_strength
    p.looking_at[:] = list(looking_at)
    p.render_pos_data = render_pos_data
    p.width, p.height = width, height
    p.max_steps, p.shadow_steps = max_steps, shadow_steps
    p.ao_samples, p.ao_steps = ao_samples, ao_steps
    p.flags = flags
    return p


class RayCaster:
    """The reference host's view of the library: init -> upload_world -> render_frame -> read_image."""

    def __init__(self, device=0):
        """device: a CUDA device index (rc_init), or a list of them (rc_init_multi: one process, one thread, the library
        deals the bands of every frame to its devices)."""
        self.lib = load_library()
        if isinstance(device, (list, tuple)):
            devs = (ctypes.c_int * len(device))(*[int(d) for d in device])
            _check(self.lib, self.lib.rc_init_multi(len(device), devs))
            self.devices = [int(d) for d in device]
            self.device = self.devices[0]
        else:
            _check(self.lib, self.lib.rc_init(device))
            self.device = device
            self.devices = [device]
        self._last = None

    def build_id(self):
        """sha256 of the kernel sources librc.so was compiled from (rc_build_id)."""
        return self.lib.rc_build_id().decode()

    def close(self):
        if self.lib is not None:
            self.lib.rc_shutdown()
            self.lib = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def set_stream(self, cuda_stream):
        _check(self.lib, self.lib.rc_set_stream(cuda_stream))

    def set_option(self, name, value):
        """rc_debug_set_option: "schedule" (int or a key of SCHEDULES), "ao_opts" (AO_OPT_* bits)."""
        if isinstance(value, str):
            value = SCHEDULES[value]
        _check(self.lib, self.lib.rc_debug_set_option(name.encode(), int(value)))

    def set_shard(self, rank, world_size):
        _check(self.lib, self.lib.rc_set_shard(rank, world_size))

    def upload_world(self, desc):
        _check(self.lib, self.lib.rc_upload_world(ctypes.byref(desc)))
        return self.world_info()

    def world_info(self):
        info = WorldInfo()
        _check(self.lib, self.lib.rc_world_info_get(ctypes.byref(info)))
        return info

    def render_frame(self, params):
        _check(self.lib, self.lib.rc_render_frame(ctypes.byref(params)))
        self._last = (params.width, params.height)

    def render_frame_static(self, params, device_ptr, nbytes):
        """rc_render_frame_static: the RGBA8 frame into caller-owned DEVICE memory, scratch in the library's fixed array
        (no allocation; width * height <= RC_STATIC_MAX_PIXELS)."""
        _check(self.lib, self.lib.rc_render_frame_static(ctypes.byref(params), device_ptr, nbytes))

    def sync(self):
        _check(self.lib, self.lib.rc_sync())

    def read_image(self, out=None, fmt=RC_RGBA8):
        w, h = self._last
        if out is None:
            out = np.zeros((h, w, 4), dtype=np.uint8 if fmt == RC_RGBA8 else np.float32)
        _check(self.lib, self.lib.rc_read_image(out.ctypes.data, out.nbytes, fmt))
        return out

    def read_image_async(self, out):
        """rc_read_image_async into `out` (a uint8 array over PINNED memory, h x w x 4) -> ticket for wait_image."""
        t = ctypes.c_int(-1)
        _check(self.lib, self.lib.rc_read_image_async(out.ctypes.data, out.nbytes, ctypes.byref(t)))
        return t.value

    def wait_image(self, ticket):
        _check(self.lib, self.lib.rc_wait_image(ticket))

    def host_register(self, array):
        """rc_host_register: page-locks the memory under a numpy array (e.g. over POSIX shared memory)."""
        _check(self.lib, self.lib.rc_host_register(array.ctypes.data, array.nbytes))

    def host_unregister(self, array):
        _check(self.lib, self.lib.rc_host_unregister(array.ctypes.data))

    def frame_target_create(self, width, height, slots=2):
        """rc_frame_target_create -> the IPC handles (bytes) the other ranks pass to frame_target_open."""
        buf = ctypes.create_string_buffer(RC_IPC_HANDLE_BYTES * slots)
        _check(self.lib, self.lib.rc_frame_target_create(width, height, slots, buf))
        return buf.raw

    def frame_target_open(self, width, height, slots, handles):
        _check(self.lib, self.lib.rc_frame_target_open(width, height, slots, ctypes.create_string_buffer(handles, len(handles))))

    def frame_target_close(self):
        _check(self.lib, self.lib.rc_frame_target_close())

    def frame_target_ptr(self, slot):
        return self.lib.rc_frame_target_device(slot)

    def frame_sync_create(self):
        h = ctypes.create_string_buffer(RC_IPC_HANDLE_BYTES)
        _check(self.lib, self.lib.rc_frame_sync_create(h))
        return h.raw

    def frame_sync_open(self, handle):
        _check(self.lib, self.lib.rc_frame_sync_open(ctypes.create_string_buffer(bytes(handle), RC_IPC_HANDLE_BYTES)))

    def frame_sync_close(self):
        _check(self.lib, self.lib.rc_frame_sync_close())

    def frame_sync_begin(self):
        _check(self.lib, self.lib.rc_frame_sync_begin())

    def frame_sync_allow(self, frame, cuda_stream=None):
        _check(self.lib, self.lib.rc_frame_sync_allow(int(frame), cuda_stream))

    def frame_sync_end(self):
        _check(self.lib, self.lib.rc_frame_sync_end())

    def frame_sync_frames(self):
        return int(self.lib.rc_frame_sync_frames())

    def frame_sync_timeouts(self):
        n = ctypes.c_uint32(0)
        _check(self.lib, self.lib.rc_frame_sync_timeouts(ctypes.byref(n)))
        return int(n.value)

    def frame_target_last_slot(self):
        return self.lib.rc_frame_target_last_slot()

    def read_hits(self, out=None):
        w, h = self._last
        if out is None:
            out = np.zeros(w * h, dtype=HIT_DTYPE)
        _check(self.lib, self.lib.rc_read_hits(out.ctypes.data, out.size))
        return out

    def stats(self):
        s = FrameStats()
        _check(self.lib, self.lib.rc_frame_stats_get(ctypes.byref(s)))
        return s

    def hemisphere_table(self, n):
        out = np.zeros((n, 3), dtype=np.float32)
        _check(self.lib, self.lib.rc_hemisphere_table(n, out.ctypes.data))
        return out

    def update_world(self, desc, dirty_node_blocks, dirty_array_blocks):
        """rc_update_world: the blocks updateSsboData() would push (BlockPtr.modified) -> UpdateInfo."""
        dn, nn, da, na = _dirty_args(dirty_node_blocks, dirty_array_blocks)
        info = UpdateInfo()
        _check(self.lib, self.lib.rc_update_world(ctypes.byref(desc), dn.ctypes.data, nn, da.ctypes.data, na,
                                                  ctypes.byref(info)))
        return info

    def save_world(self, path):
        _check(self.lib, self.lib.rc_save_world(os.fsencode(path)))

    def load_world(self, path):
        _check(self.lib, self.lib.rc_load_world(os.fsencode(path)))
        info = WorldInfo()
        _check(self.lib, self.lib.rc_world_info_get(ctypes.byref(info)))
        return info

    def generate_world(self, width=200, length=200, max_depth=6, seeds=(42, 64, 100), startup_blocks=True):
        """rc_generate_world: initTetraHexaTree() + genWorld() (src/main.cpp:181-190) on the device -> GenInfo."""
        p = GenParams(int(width), int(length), int(max_depth), 1 if startup_blocks else 0, (ctypes.c_int64 * 3)(*seeds))
        info = GenInfo()
        _check(self.lib, self.lib.rc_generate_world(ctypes.byref(p), ctypes.byref(info)))
        return info

    def write_ppm(self, path):
        _check(self.lib, self.lib.rc_write_ppm(os.fsencode(path)))

    def pick(self, cam_pos, cam_dir, steps=30):
        """rc_pick = RAY_CASTER::castRayFromCam(steps) on the device -> (pos, lastPos, steps left)."""
        p = np.ascontiguousarray(cam_pos, dtype=np.float32)
        d = np.ascontiguousarray(cam_dir, dtype=np.float32)
        out = np.zeros(7, dtype=np.int32)
        _check(self.lib, self.lib.rc_pick(p.ctypes.data, d.ctypes.data, int(steps), out.ctypes.data))
        return out[0:3].copy(), out[3:6].copy(), int(out[6])

    def device_image_ptr(self, fmt=RC_RGBA8):
        return self.lib.rc_device_image(fmt)

    def device_hits_ptr(self):
        return self.lib.rc_device_hits()

    def compose_frame(self, gathered_ptr, dst_ptr, width, height, world_size, bytes_per_element=4):
        _check(self.lib, self.lib.rc_compose_frame(gathered_ptr, dst_ptr, width, height, world_size, bytes_per_element))

    # ---- rc_debug.h ----
    def lookup_device(self, xyz):
        xyz = np.ascontiguousarray(xyz, dtype=np.int32).reshape(-1, 3)
        words = np.zeros(len(xyz), dtype=np.uint32)
        shifts = np.zeros(len(xyz), dtype=np.int32)
        _check(self.lib, self.lib.rc_debug_lookup_device(xyz.ctypes.data, len(xyz), words.ctypes.data, shifts.ctypes.data))
        return words, shifts

    def reread_image_ms(self, out, repeats):
        """rc_debug_reread_image: wall-clock ms of `repeats` back-to-back copies of the last frame's rows into `out` (a uint8
        array over page-locked memory; None = a page-locked frame of the call's own)."""
        ms = ctypes.c_float(0.0)
        _check(self.lib, self.lib.rc_debug_reread_image(out.ctypes.data if out is not None else None, out.nbytes if out is not None else 0,
                                                        int(repeats), ctypes.byref(ms)))
        return float(ms.value)

    def detmath(self, x):
        x = np.ascontiguousarray(x, dtype=np.float32)
        s, e = np.zeros_like(x), np.zeros_like(x)
        _check(self.lib, self.lib.rc_debug_detmath(x.ctypes.data, x.size, s.ctypes.data, e.ctypes.data))
        return s, e


def host_fence():
    """rc_host_fence: a full memory fence (release before a counter store, acquire after a counter load). Needs no device."""
    load_library().rc_host_fence()


def shard_layout(height, rank, world_size):
    """(rows this rank renders, rows every rank's buffers hold) for a frame sharded in 16-row bands."""
    lib = load_library()
    a, b = ctypes.c_uint32(), ctypes.c_uint32()
    _check(lib, lib.rc_shard_layout(height, rank, world_size, ctypes.byref(a), ctypes.byref(b)))
    return a.value, b.value


def band_rows_of_rank(height, rank, world_size):
    """Global row indices a rank renders, in the order its compact buffers store them."""
    rows = []
    band = rank
    while band * BAND_ROWS < height:
        rows.extend(range(band * BAND_ROWS, min((band + 1) * BAND_ROWS, height)))
        band += world_size
    return rows


def flatten_host(desc):
    """CPU-only: the flattener half of rc_upload_world (no device needed)."""
    lib = load_library()
    info = WorldInfo()
    _check(lib, lib.rc_debug_flatten_host(ctypes.byref(desc), ctypes.byref(info)))
    return info


def _dirty_args(dirty_node_blocks, dirty_array_blocks):
    """None = RC_ALL_BLOCKS (diff every block)."""
    dn = np.ascontiguousarray([] if dirty_node_blocks is None else dirty_node_blocks, dtype=np.uint32)
    da = np.ascontiguousarray([] if dirty_array_blocks is None else dirty_array_blocks, dtype=np.uint32)
    return (dn, RC_ALL_BLOCKS if dirty_node_blocks is None else dn.size,
            da, RC_ALL_BLOCKS if dirty_array_blocks is None else da.size)


def save_host(path):
    """CPU-only: rc_save_world's writer on the debug world (which=1)."""
    lib = load_library()
    _check(lib, lib.rc_debug_save_host(os.fsencode(path)))


def load_host(path):
    """CPU-only: rc_load_world's reader + validation into the debug world (which=1)."""
    lib = load_library()
    info = WorldInfo()
    _check(lib, lib.rc_debug_load_host(os.fsencode(path), ctypes.byref(info)))
    return info


def flatten_host_tracked(desc):
    """CPU-only: full flatten keeping the maps rc_update_world needs (debug world, which=1)."""
    lib = load_library()
    info = WorldInfo()
    _check(lib, lib.rc_debug_flatten_host_tracked(ctypes.byref(desc), ctypes.byref(info)))
    return info


def update_host(desc, dirty_node_blocks, dirty_array_blocks):
    """CPU-only: the host half of rc_update_world on the tracked debug world."""
    lib = load_library()
    dn, nn, da, na = _dirty_args(dirty_node_blocks, dirty_array_blocks)
    info = UpdateInfo()
    _check(lib, lib.rc_debug_update_host(ctypes.byref(desc), dn.ctypes.data, nn, da.ctypes.data, na, ctypes.byref(info)))
    return info


def lookup_host(xyz, which=1):
    """Packed-form lookups on the host copy. Returns (words, shifts, leaves[n,4] u32)."""
    lib = load_library()
    xyz = np.ascontiguousarray(xyz, dtype=np.int32).reshape(-1, 3)
    words = np.zeros(len(xyz), dtype=np.uint32)
    shifts = np.zeros(len(xyz), dtype=np.int32)
    leaves = np.zeros((len(xyz), 4), dtype=np.uint32)
    _check(lib, lib.rc_debug_lookup_host_batch(which, xyz.ctypes.data, len(xyz), words.ctypes.data, shifts.ctypes.data,
                                               leaves.ctypes.data))
    return words, shifts, leaves

"""Shared fixtures. GPU tests are marked @pytest.mark.gpu and call through the C-ABI
(librc.so); everything else runs on CPU.  Nothing here reads /root/reference at run
time: the reference-built world comes from the prebuilt oracle/_ref/libref_world.so."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def ob():
    from oracle import bindings

    bindings.oracle_lib()
    return bindings


@pytest.fixture(scope="session")
def ref_world(ob):
    """Default world built by the reference's own unmodified code (initTetraHexaTree + genWorld)."""
    try:
        return ob.RefWorld()
    except FileNotFoundError:
        pytest.skip("oracle/_ref/libref_world.so not built (needs /root/reference once)")


@pytest.fixture(scope="session")
def oracle_world(ob, ref_world):
    return ob.make_world(ref_world.nodes, ref_world.arrays, ref_world.max_depth)


@pytest.fixture(scope="session")
def rcmod():
    from raytracing_test_b200 import rc

    rc.load_library()
    return rc


@pytest.fixture(scope="session")
def world_desc(rcmod, ref_world):
    nodes, arrays = ref_world.block_ptrs()
    return rcmod.make_world_desc(nodes, arrays, ref_world.node_bytes_used, ref_world.array_bytes_used,
                                 block_bytes=ref_world.block_bytes, max_depth=ref_world.max_depth)


@pytest.fixture(scope="session")
def caster(rcmod, world_desc):
    """librc initialised on cuda:0 with the default world uploaded."""
    c = rcmod.RayCaster(0)
    c.upload_world(world_desc)
    yield c
    c.close()


@pytest.fixture(scope="session")
def defaults(ref_world):
    """cameraPos, normalised cameraDir, sun of src/globals.cpp:20-23."""
    return ref_world.defaults()


def camera_dir(yaw_deg, pitch_deg):
    """rotateCamera's formula (src/input.cpp:25-28) in fp32: rotateX(0,0,1) by pitch, rotateY by yaw,
    normalising after each (glm pure-C++ definitions)."""
    f = np.float32
    p, y = f(np.radians(f(pitch_deg))), f(np.radians(f(yaw_deg)))

    def norm(v):
        d = f(f(f(v[0] * v[0]) + f(v[1] * v[1])) + f(v[2] * v[2]))
        return (v * f(f(1.0) / np.sqrt(d))).astype(np.float32)

    c, s = f(np.cos(p)), f(np.sin(p))
    v = np.array([0.0, f(f(0.0) * c - f(1.0) * s), f(f(0.0) * s + f(1.0) * c)], dtype=np.float32)  # rotateX
    v = norm(v)
    c, s = f(np.cos(y)), f(np.sin(y))
    v = np.array([f(f(v[0] * c) + f(v[2] * s)), v[1], f(f(-v[0] * s) + f(v[2] * c))], dtype=np.float32)  # rotateY
    return norm(v)


# (origin, yaw, pitch) poses for parity sweeps; what each one exercises was checked with the
# oracle (kinds / reflected / refracted / shadowed pixel counts).  None looks into
# [768,1023]^3.
CAMERAS = [
    ((35.0, 50.0, 35.0), 45.0, 0.0),        # the default view
    ((35.0, 50.0, 35.0), -135.0, 10.0),     # away from the terrain: all sky, 300 steps, wraps the torus
    ((100.5, 60.25, 100.5), 0.0, 60.0),     # steep look down: short rays, most cast shadow rays
    ((100.5, 60.25, 100.5), 180.0, 45.0),
    ((100.5, 45.5, 100.5), 90.0, -30.0),
    ((8.5, 101.5, 4.5), 15.0, 5.0),         # black 1^3 mirror at (10,100,10): ~6 % reflected pixels
    ((21.5, 12.5, 210.5), 180.0, 10.0),     # 4^3 refractor at (20..23,8..11,200..203) in front of the terrain edge
    ((150.5, 30.5, 140.5), 45.0, 45.0),     # down onto the lake: every ray refracts through animated water
    ((162.5, 19.5, 128.5), 30.0, -20.0),    # camera INSIDE water (10 deep) looking up: starts with n = 1.1
    ((162.5, 19.5, 128.5), 30.0, 30.0),     # camera inside water looking at the lake bed
    ((-20.5, 45.5, -20.5), 45.0, 10.0),     # starts at negative coordinates (= 1003 on the torus), enters the world
    ((2.5, 40.5, 2.5), -135.0, 20.0),       # leaves the world over the seam
    ((199.5, 35.5, 199.5), 45.0, -10.0),
    ((60.2, 25.9, 140.8), 120.0, 35.0),
    ((35.0, 50.0, 35.0), 45.0, -89.0),      # near-vertical rays: huge deltas on two axes
    ((35.0, 50.0, 35.0), 45.0, 89.0),
]

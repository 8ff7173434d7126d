"""raytracing_test_b200 -- B200-native voxel ray caster behind the reference's render boundary.

The product is ``librc.so`` (C-ABI in ``include/rc.h``, CUDA for sm_100a in ``csrc/``).
``rc`` is the ctypes binding used by tests and bench.py.
"""
from . import rc  # noqa: F401

__all__ = ["rc"]

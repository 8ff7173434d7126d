// oracle/shim/errhandlingapi.h -- TEST INFRASTRUCTURE ONLY.
// Empty stand-in so the reference's CPU sources (which include <windows.h>,
// src/voxel_data/voxel_allocator.hpp:3-5) compile unmodified on Linux.
#pragma once

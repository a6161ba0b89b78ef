// Stand-in for the CMake-generated ceres/internal/export.h (test only): lets the reference's own
// internal/ceres/block_structure.h compile outside its build system.
#pragma once
#define CERES_NO_EXPORT
#define CERES_EXPORT

// Forced-include compatibility header for building the UNMODIFIED reference
// (CILAB-MA/gpudrive_lab + vendored Madrona) with g++ instead of the
// clang/libc++ toolchain its CMake fetches from the network.
// Test infrastructure only (oracle/_ref); nothing here is arithmetic.
#pragma once
#ifdef __cplusplus
#include <cmath>
#include <cstring>
#include <cstdlib>
#include <cstdio>
#include <numeric>
#include <string>
#include <unordered_map>
#include <algorithm>
using std::isnan;     // external/madrona/src/physics/xpbd.cpp:57
using std::signbit;   // external/madrona/include/madrona/mesh_bvh.inl:615 calls it unqualified
#endif

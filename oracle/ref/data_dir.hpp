// DATA_DIR for src/mgr.cpp:273-279 resolves at run time so libgd_ref.so can
// find the two collision meshes on any box (GD_REF_ASSETS, else oracle/ref/assets
// next to the library). The meshes there are our own 8-vertex boxes with the same
// extents as assets/cube_collision.obj / agent_collision_simplified.obj.
#pragma once
#ifdef __cplusplus
#include <string>
std::string gd_ref_data_dir();
#endif

// loader.h -- host-side scene ingest (see loader.cpp).
#pragma once
#include <string>
#include <vector>

#include "../../include/ptx.h"

namespace ptx {
bool load_obj(const std::string& path, bool mirror_from_mtl, std::vector<ptx_primitive>& out, std::string& err);
bool load_pbrt(const std::string& path, bool mirror_ext, long long triangle_cap, std::vector<ptx_primitive>& out, std::string& err);
void split_quads(std::vector<ptx_primitive>& prims);
void subdivide(std::vector<ptx_primitive>& prims, int levels);
}  // namespace ptx

// loader.h -- host-side scene ingest (see loader.cpp).
#pragma once
#include <string>
#include <vector>

#include "../../include/ptx.h"

namespace ptx {
bool load_obj(const std::string& path, bool mirror_from_mtl, std::vector<ptx_primitive>& out, std::string& err);
bool load_pbrt(const std::string& path, bool mirror_ext, long long triangle_cap, std::vector<ptx_primitive>& out, std::string& err);
// Two-level form of a .pbrt scene (ptx_load_opts.keep_instances): the world's own shapes in world space, every instanced object ONCE in
// object space, and the instances in the order the reference's flattening visits them, so that the flattened primitive numbering
// is: world primitives, then for every instance the primitives of its object.  Returns false with `err` empty when the file
// cannot be kept two-level (nested instancing, no instances at all): the caller flattens instead.
struct InstancedScene {
    std::vector<ptx_primitive> world;
    std::vector<std::vector<ptx_primitive>> objects;     // object space; normal = NaN when the mesh carries no vertex normals
    struct Inst {
        int object;
        float m[12];                                       // object -> world: columns vx, vy, vz, then the translation
    };
    std::vector<Inst> instances;
};
bool load_pbrt_instanced(const std::string& path, bool mirror_ext, InstancedScene& out, std::string& err);
void split_quads(std::vector<ptx_primitive>& prims);
void subdivide(std::vector<ptx_primitive>& prims, int levels);
}  // namespace ptx

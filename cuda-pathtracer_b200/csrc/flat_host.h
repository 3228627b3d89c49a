// flat_host.h -- host-side construction of the flat unit list for tiny scenes (see flat.h).
#pragma once
#include <vector>

#include "bvh_common.h"
#include "flat.h"

namespace ptx {

struct FlatHost {
    std::vector<FlatUnit> units;       // <= kFlatUnits
    std::vector<f32x4> recs;           // kFlatRecQuads per unit
    float scale = 0.f;                 // max |coordinate| of the scene's vertices
};

// Groups the triangle soup into planar units: the two halves of a reference quad (consecutive records with one prim id)
// when coplanar, two coplanar triangles that share an edge, else a triangle on its own.  Returns false (and leaves `out`
// empty) when the scene needs more than kFlatUnits units or holds a degenerate (zero-area) triangle.
bool build_flat_units(const TriIn* tris, size_t n, FlatHost& out);

}  // namespace ptx

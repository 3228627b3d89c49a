// instancing_host.h -- host arithmetic of the two-level scene assembly (api.cu scene_from_instanced; tests/host_emu.cpp runs the
// same code on the CPU): inverse of an instance transform, the instance's world box, and the placeholder triangle the TLAS is built
// from (its bounding box is the instance's box, its prim id the instance index).
#pragma once
#include <algorithm>
#include <cmath>
#include <cstring>

#include "bvh_common.h"
#include "traverse_inst.h"

namespace ptx {

// 3 x 4 affine inverse in double precision (columns vx, vy, vz, translation)
inline bool invert_affine(const float* m, float* inv) {
    const double a = m[0], b = m[3], c = m[6], d = m[1], e = m[4], f = m[7], g = m[2], h = m[5], i = m[8];
    const double det = a * (e * i - f * h) - b * (d * i - f * g) + c * (d * h - e * g);
    if (!(std::fabs(det) > 1e-30)) return false;
    const double r[9] = {(e * i - f * h) / det, (c * h - b * i) / det, (b * f - c * e) / det,
                         (f * g - d * i) / det, (a * i - c * g) / det, (c * d - a * f) / det,
                         (d * h - e * g) / det, (b * g - a * h) / det, (a * e - b * d) / det};      // row-major inverse of the linear part
    // store as columns: inv[0..2] = first column = (r[0], r[3], r[6]) ...
    for (int col = 0; col < 3; col++) for (int row = 0; row < 3; row++) inv[3 * col + row] = (float)r[3 * row + col];
    for (int row = 0; row < 3; row++) inv[9 + row] = (float)(-(r[3 * row + 0] * m[9] + r[3 * row + 1] * m[10] + r[3 * row + 2] * m[11]));
    return true;
}


// `root` receives the OBJECT index here; the caller patches it to the node offset of that object's BLAS once the trees are laid out.
inline bool make_instance(const float* m, const aabb& obj_bounds, uint32_t object, uint32_t prim_base, uint32_t shade_base, uint32_t index,
                          InstanceRec& r, TriIn& placeholder) {
    memset(&r, 0, sizeof r);
    memcpy(r.m, m, sizeof r.m);
    if (!invert_affine(m, r.inv)) return false;
    r.root = object; r.prim_base = prim_base; r.shade_base = shade_base;
    aabb wb = aabb_empty();
    for (int k = 0; k < 8; k++)
        aabb_grow(wb, xfm_pt(m, mk3((k & 1) ? obj_bounds.hi.x : obj_bounds.lo.x, (k & 2) ? obj_bounds.hi.y : obj_bounds.lo.y, (k & 4) ? obj_bounds.hi.z : obj_bounds.lo.z)));
    const vec3 ext = wb.hi - wb.lo;                     // the corners were transformed in float: pad by a few ulps of the box
    const float big = std::max(std::max(std::fabs(wb.hi.x), std::fabs(wb.lo.x)), std::max(std::max(std::fabs(wb.hi.y), std::fabs(wb.lo.y)), std::max(std::fabs(wb.hi.z), std::fabs(wb.lo.z))));
    const float pad = 1e-5f * std::max(std::max(ext.x, ext.y), std::max(ext.z, 1e-3f)) + 1e-6f * big;
    st3(placeholder.v0, wb.lo - mk3(pad, pad, pad)); st3(placeholder.v1, wb.hi + mk3(pad, pad, pad)); st3(placeholder.v2, 0.5f * (wb.lo + wb.hi));
    placeholder.prim_id = index;
    return true;
}

}  // namespace ptx

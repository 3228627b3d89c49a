// scene_pack.cpp -- see scene_pack.h.  The reference keeps Kd / Ke / normal inside every 4 364-byte
// Primitive (rendering/triangle.h:98-105); here they become a 16-byte per-primitive record plus a
// small deduplicated material table, and each quad becomes the two triangles its intersect routine
// tests, (v00,v10,v11) and (v00,v11,v01) (rendering/quad.h:60-62, 88-90).
#include "scene_pack.h"

#include <cstring>
#include <map>

namespace ptx {

namespace {
struct MatKey {
    uint32_t w[7];
    bool operator<(const MatKey& o) const { return memcmp(w, o.w, sizeof w) < 0; }
};
}  // namespace

void make_quad_bits(const ptx_primitive* prims, size_t n, std::vector<uint32_t>& out) {
    out.clear();
    bool any = false;
    for (size_t i = 0; i < n && !any; i++) any = prims[i].type == 1;
    if (!any) return;
    out.assign((n + 31) / 32, 0u);
    for (size_t i = 0; i < n; i++)
        if (prims[i].type == 1) out[i >> 5] |= 1u << (i & 31);
}

void pack_primitives(const ptx_primitive* prims, size_t n, PackedScene& out) {
    out.tris.clear(); out.prim_shade.clear(); out.materials.clear();
    make_quad_bits(prims, n, out.quad_bits);
    out.tris.reserve(n); out.prim_shade.reserve(n);
    std::map<MatKey, uint32_t> table;
    for (size_t i = 0; i < n; i++) {
        const ptx_primitive& p = prims[i];
        MatKey k;
        memcpy(&k.w[0], p.bsdf, 12); memcpy(&k.w[3], p.Le, 12); k.w[6] = (uint32_t)p.material_type;
        auto it = table.find(k);
        uint32_t mid;
        if (it == table.end()) {
            mid = (uint32_t)table.size();
            table.emplace(k, mid);
            out.materials.push_back(f32x4{p.bsdf[0], p.bsdf[1], p.bsdf[2], u2f((uint32_t)p.material_type)});
            out.materials.push_back(f32x4{p.Le[0], p.Le[1], p.Le[2], 0.f});
        } else mid = it->second;
        out.prim_shade.push_back(f32x4{p.normal[0], p.normal[1], p.normal[2], u2f(mid)});
        TriIn t;
        memcpy(t.v0, p.v[0], 12); memcpy(t.v1, p.v[1], 12); memcpy(t.v2, p.v[2], 12); t.prim_id = (uint32_t)i;
        out.tris.push_back(t);
        if (p.type == 1) {
            memcpy(t.v1, p.v[2], 12); memcpy(t.v2, p.v[3], 12);
            out.tris.push_back(t);
        }
    }
}

void pack_triangles(const float* v, const float* normals, const uint32_t* mat_id, const ptx_material* mats, size_t n_tri,
                    size_t n_mat, PackedScene& out) {
    out.tris.resize(n_tri); out.prim_shade.resize(n_tri); out.materials.clear(); out.quad_bits.clear();
    for (size_t m = 0; m < n_mat; m++) {
        out.materials.push_back(f32x4{mats[m].bsdf[0], mats[m].bsdf[1], mats[m].bsdf[2], u2f((uint32_t)mats[m].type)});
        out.materials.push_back(f32x4{mats[m].Le[0], mats[m].Le[1], mats[m].Le[2], 0.f});
    }
    if (n_mat == 0) {   // default Material of the reference: grey 0.8, no emission (file_manager.h:24-27)
        out.materials.push_back(f32x4{0.8f, 0.8f, 0.8f, u2f(0u)});
        out.materials.push_back(f32x4{0.f, 0.f, 0.f, 0.f});
    }
    for (size_t i = 0; i < n_tri; i++) {
        TriIn& t = out.tris[i];
        memcpy(t.v0, v + 9 * i, 12); memcpy(t.v1, v + 9 * i + 3, 12); memcpy(t.v2, v + 9 * i + 6, 12);
        t.prim_id = (uint32_t)i;
        vec3 nn = normals ? ld3(normals + 3 * i) : unit(cross(ld3(t.v1) - ld3(t.v0), ld3(t.v2) - ld3(t.v0)));
        uint32_t mid = mat_id ? mat_id[i] : 0u;
        if (mid >= (n_mat ? n_mat : (size_t)1)) mid = 0;      // n_mat == 0: only the default material exists
        out.prim_shade[i] = f32x4{nn.x, nn.y, nn.z, u2f(mid)};
    }
}

}  // namespace ptx

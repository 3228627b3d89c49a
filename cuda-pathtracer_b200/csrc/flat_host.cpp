// flat_host.cpp -- see flat_host.h / flat.h.  Double precision throughout; the stored floats are within one rounding of it.
#include "flat_host.h"

#include <algorithm>
#include <cmath>
#include <cstring>

namespace ptx {

namespace {

struct D3 {
    double x, y, z;
};
D3 d3(const float* p) { return D3{p[0], p[1], p[2]}; }
D3 sub(D3 a, D3 b) { return D3{a.x - b.x, a.y - b.y, a.z - b.z}; }
D3 add(D3 a, D3 b) { return D3{a.x + b.x, a.y + b.y, a.z + b.z}; }
D3 mul(D3 a, double s) { return D3{a.x * s, a.y * s, a.z * s}; }
double dotd(D3 a, D3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
D3 crossd(D3 a, D3 b) { return D3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
double lend(D3 a) { return std::sqrt(dotd(a, a)); }

bool same_vertex(const float* a, const float* b) { return memcmp(a, b, 12) == 0; }
int shared_vertices(const TriIn& a, const TriIn& b) {
    const float* va[3] = {a.v0, a.v1, a.v2};
    const float* vb[3] = {b.v0, b.v1, b.v2};
    int n = 0;
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++)
            if (same_vertex(va[i], vb[j])) { n++; break; }
    return n;
}

}  // namespace

bool build_flat_units(const TriIn* tris, size_t n, FlatHost& out) {
    out.units.clear(); out.recs.clear(); out.scale = 0.f;
    if (n == 0 || n > 2 * (size_t)kFlatUnits) return false;
    double scale = 0.0;
    for (size_t i = 0; i < n; i++)
        for (int k = 0; k < 3; k++) scale = std::max({scale, std::fabs((double)tris[i].v0[k]), std::fabs((double)tris[i].v1[k]), std::fabs((double)tris[i].v2[k])});
    if (!(scale > 0.0) || !std::isfinite(scale)) return false;
    const double plane_tol = 1e-6 * scale;      // a tenth of the filter's margin m1 (flat.h)

    std::vector<D3> nrm(n);
    std::vector<double> off(n);
    for (size_t i = 0; i < n; i++) {
        D3 v0 = d3(tris[i].v0), N = crossd(sub(d3(tris[i].v1), v0), sub(d3(tris[i].v2), v0));
        double l = lend(N);
        if (!(l > 1e-12 * scale * scale)) return false;       // degenerate triangle: leave the scene to the BVH path
        nrm[i] = mul(N, 1.0 / l);
        off[i] = dotd(nrm[i], v0);
    }
    auto coplanar = [&](size_t a, size_t b) {
        const float* vb[3] = {tris[b].v0, tris[b].v1, tris[b].v2};
        for (int k = 0; k < 3; k++)
            if (std::fabs(dotd(nrm[a], d3(vb[k])) - off[a]) > plane_tol) return false;
        return true;
    };

    std::vector<int> partner(n, -1);
    std::vector<char> used(n, 0);
    for (size_t i = 0; i < n; i++) {
        if (used[i]) continue;
        used[i] = 1;
        int j = -1;
        if (i + 1 < n && !used[i + 1] && tris[i + 1].prim_id == tris[i].prim_id) {       // the two halves of a reference quad
            if (coplanar(i, i + 1)) j = (int)(i + 1);
        } else {
            for (size_t c = i + 1; c < n && j < 0; c++) {
                if (used[c] || tris[c].prim_id == tris[i].prim_id) continue;
                if (c + 1 < n && tris[c + 1].prim_id == tris[c].prim_id) continue;          // half of a quad: pairs with its other half
                if (c > 0 && tris[c - 1].prim_id == tris[c].prim_id) continue;
                if (shared_vertices(tris[i], tris[c]) >= 2 && coplanar(i, c)) j = (int)c;
            }
        }
        if (j >= 0) { used[j] = 1; partner[i] = j; }
        if (out.units.size() == (size_t)kFlatUnits) { out.units.clear(); out.recs.clear(); return false; }

        // bounding rectangle in the plane of triangle i: try every edge direction, keep the smallest rectangle
        std::vector<D3> pts = {d3(tris[i].v0), d3(tris[i].v1), d3(tris[i].v2)};
        if (j >= 0) { pts.push_back(d3(tris[j].v0)); pts.push_back(d3(tris[j].v1)); pts.push_back(d3(tris[j].v2)); }
        const D3 nn = nrm[i];
        const D3 base = pts[0];
        double best_area = -1.0; D3 ba1{}, ba2{}; double bmid1 = 0, bmid2 = 0, bh1 = 0, bh2 = 0;
        for (size_t a = 0; a < pts.size(); a++)
            for (size_t b = a + 1; b < pts.size(); b++) {
                D3 e = sub(pts[b], pts[a]);
                e = sub(e, mul(nn, dotd(e, nn)));                    // into the plane
                double l = lend(e);
                if (!(l > 1e-9 * scale)) continue;
                D3 a1 = mul(e, 1.0 / l), a2 = crossd(nn, a1);
                double lo1 = 1e300, hi1 = -1e300, lo2 = 1e300, hi2 = -1e300;
                for (const D3& p : pts) {
                    D3 r = sub(p, base);
                    double c1 = dotd(r, a1), c2 = dotd(r, a2);
                    lo1 = std::min(lo1, c1); hi1 = std::max(hi1, c1); lo2 = std::min(lo2, c2); hi2 = std::max(hi2, c2);
                }
                double area = (hi1 - lo1) * (hi2 - lo2);
                if (best_area < 0 || area < best_area) {
                    best_area = area; ba1 = a1; ba2 = a2;
                    bmid1 = 0.5 * (lo1 + hi1); bmid2 = 0.5 * (lo2 + hi2); bh1 = 0.5 * (hi1 - lo1); bh2 = 0.5 * (hi2 - lo2);
                }
            }
        if (best_area < 0) { out.units.clear(); out.recs.clear(); return false; }
        D3 ctr = add(base, add(mul(ba1, bmid1), mul(ba2, bmid2)));
        ctr = sub(ctr, mul(nn, dotd(ctr, nn) - off[i]));             // exactly on the plane
        FlatUnit u;
        u.nx = (float)nn.x; u.ny = (float)nn.y; u.nz = (float)nn.z; u.c = (float)off[i];
        u.cx = (float)ctr.x; u.cy = (float)ctr.y; u.cz = (float)ctr.z; u.h1 = (float)bh1;
        u.a1x = (float)ba1.x; u.a1y = (float)ba1.y; u.a1z = (float)ba1.z; u.h2 = (float)bh2;
        u.a2x = (float)ba2.x; u.a2y = (float)ba2.y; u.a2z = (float)ba2.z; u.pad = 0.f;
        out.units.push_back(u);
        f32x4 r[3];
        pack_triangle(tris[i], r);
        out.recs.insert(out.recs.end(), r, r + 3);
        if (j >= 0) pack_triangle(tris[j], r);
        else { r[0] = f32x4{0.f, 0.f, 0.f, u2f(kMiss)}; r[1] = r[2] = f32x4{0.f, 0.f, 0.f, 0.f}; }   // null record: a = 0, always rejected
        out.recs.insert(out.recs.end(), r, r + 3);
    }
    out.scale = (float)scale;
    return true;
}

}  // namespace ptx

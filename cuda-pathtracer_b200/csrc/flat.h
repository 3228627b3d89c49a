// flat.h -- closest hit for TINY scenes (at most kFlatUnits planar units, e.g. the 16 quads of the Cornell box) without a
// tree: every ray runs one straight-line pass over all units, then ONE exact test.
//
// Replaces, for such scenes, Scene::intersect -> intersect_bvh_optimized (rendering/scene.h:39-110) exactly like the BVH8
// traversal of traverse.h does: the accepted hit is decided by the very same intersect_tri() arithmetic
// (Triangle::intersect / Quad::intersect, triangle.h:64-96, quad.h:49-132), so distances and primitive ids are bit-identical
// to the BVH8 path (exact ties between two different primitives excepted, as everywhere else).  Only the CULLING differs.
//
// Why: on a 4-node tree the per-lane state machine of k_path_persist spends 28 % of its issue slots in a triangle loop that
// runs 5.1 rounds per warp iteration for 1.9 tests per lane (lane efficiency 0.38) and a ray needs 1.5 iterations, so only
// 21 of 32 lanes finish a segment per iteration and shading runs at 0.35 (profiles/r1ze_c2_*).  Here the per-unit filter is
// uniform work (all lanes, unit data as constant-bank operands straight from the kernel parameters, no loads, no stack), every
// lane finishes its segment in the iteration it started it, and exactly one unit (two records) is tested exactly.
//
// A unit = one planar primitive: a reference quad (both halves), or two coplanar triangles sharing an edge, or a single
// triangle.  Per unit the filter intersects the ray with the unit's plane and checks the hit point against the unit's
// bounding rectangle IN the plane (exact fit for rectangles / parallelograms' bounding box), with a margin that grows as
// the ray grazes the plane:      m = m0 + m1 / |n.d|.
// The filter is CONSERVATIVE: whatever intersect_tri() accepts lies inside the candidate set (the margins cover the
// rounding of both computations, see flat_margins()).  tlo = t - m is then a lower bound of the unit's exact distance; the
// candidate with the smallest tlo is tested exactly, and if its exact distance is below the second smallest tlo nothing
// else can be closer.  Otherwise (two candidates within their margins: rays into an edge between two walls; a grazing
// plane; a false candidate in front) the lane tests every candidate exactly -- rare (measured on the host emulation and
// reported by the GPU tests).
#pragma once
#include "layout.h"
#include "traverse.h"

namespace ptx {

constexpr int kFlatUnits = 20;      // the Cornell box: 15 planar quads + the skewed left wall as two triangles = 17
constexpr int kFlatRecQuads = 6;       // two 48-byte edge records per unit (a single triangle is followed by a null record)

struct FlatUnit {                      // 64 bytes, read as constant-bank operands
    float nx, ny, nz, c;               // unit normal, plane offset (n.x = c)
    float cx, cy, cz, h1;              // centre of the bounding rectangle (a point of the plane), half extent along a1
    float a1x, a1y, a1z, h2;           // in-plane unit axes a1, a2 = n x a1; half extent along a2
    float a2x, a2y, a2z, pad;
};

struct FlatScene {
    FlatUnit u[kFlatUnits];
    const f32x4* recs;                 // kFlatRecQuads float4 per unit: { v0, prim }, { e1 }, { e2 } twice
    uint32_t n_units;
    float m0, m1;                      // margins, see flat_margins()
    uint32_t pad;
};

// Margins for rays whose origins and the scene lie within `scale` of the coordinate origin.
//   error of t = num / den:  num = c - n.o cancels to ~ 4 ulp of `scale`, den to ~ 3 ulp of 1, |t| <= 2 * scale:
//                            |dt| <= (2.4e-7 + 2 * 1.8e-7 + rcp.approx 1.2e-7 * 2) * scale / |den|  < 1e-6 * scale / |den|
//   tolerance of intersect_tri at its borders (u, v computed as f * dot(s, h) with f = 1 / (e1 . (d x e2))): a point it accepts
//   lies within ~ 3e-7 * (|s| + |e|) / (sin(edge angle) * |den|) of the primitive:                           < 3e-6 * scale / |den|
//   storage rounding of the unit constants (host double -> float) and of the in-plane projections:         < 1e-6 * scale
// m0 = 1e-5 * scale and m1 = 1e-5 * scale leave a factor > 2 on top of their sum.
inline void flat_margins(float scale, float& m0, float& m1) { m0 = 1e-5f * scale; m1 = 1e-5f * scale; }

// 1 / x for the filter: one MUFU.RCP, denormals flushed (a denormal n.d becomes 0 -> 1 / 0 = inf -> infinite margin -> candidate).
// rcp.approx WITHOUT .ftz costs two range tests, two selects and two scalings around the MUFU.
PTX_HD float rcp_ftz(float x) {
#if defined(__CUDA_ARCH__)
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
#else
    return 1.0f / x;
#endif
}

// exact test of unit k (both records, the second replaces the first only if strictly closer: quad.h:76, 106)
PTX_HD void flat_test_unit(const f32x4* recs, uint32_t k, vec3 o, vec3 d, float t_min, float& closest, uint32_t& bestw) {
    const f32x4* tp = recs + (size_t)k * kFlatRecQuads;
    f32x4 a = ldq(tp + 0), b = ldq(tp + 1), c = ldq(tp + 2);
    f32x4 a2 = ldq(tp + 3), b2 = ldq(tp + 4), c2 = ldq(tp + 5);
    if (intersect_tri(mk3(a.x, a.y, a.z), mk3(b.x, b.y, b.z), mk3(c.x, c.y, c.z), o, d, t_min, closest)) bestw = f2u(a.w);
    if (intersect_tri(mk3(a2.x, a2.y, a2.z), mk3(b2.x, b2.y, b2.z), mk3(c2.x, c2.y, c2.z), o, d, t_min, closest)) bestw = f2u(a2.w);
}

struct FlatCounters {
    uint32_t rays, slow, unit_tests;
};

// Closest hit of (o, d) in [t_min, FLT_MAX).  closest / bestw (primitive id or kMiss) as Traversal leaves them.
template <bool COUNT>
PTX_HD void flat_closest(const FlatScene& fs, vec3 o, vec3 d, float t_min, float& closest, uint32_t& bestw, FlatCounters* ctr) {
    const float kInf = u2f(0x7F800000u);
    float tbest = kInf, t2 = kInf;
    uint32_t kbest = 0, mask = 0;
    const uint32_t n = fs.n_units;
#pragma unroll
    for (int k = 0; k < kFlatUnits; k++) {
        if ((uint32_t)k < n) {                                        // uniform
        const FlatUnit& u = fs.u[k];
        const float den = fmaf(u.nz, d.z, fmaf(u.ny, d.y, u.nx * d.x));
        const float num = fmaf(-u.nz, o.z, fmaf(-u.ny, o.y, fmaf(-u.nx, o.x, u.c)));
        const float rden = rcp_ftz(den);
        const float t = num * rden;
        const float m = fmaf(fabsf(rden), fs.m1, fs.m0);
        const float px = fmaf(t, d.x, o.x) - u.cx, py = fmaf(t, d.y, o.y) - u.cy, pz = fmaf(t, d.z, o.z) - u.cz;
        const float e1 = fabsf(fmaf(u.a1z, pz, fmaf(u.a1y, py, u.a1x * px))) - u.h1;
        const float e2 = fabsf(fmaf(u.a2z, pz, fmaf(u.a2y, py, u.a2x * px))) - u.h2;
        // negated comparisons: a NaN (den = 0, num = 0) or an infinite margin (den = 0) makes the unit a candidate
        const bool cand = !(fmaxf(e1, e2) > m) && !(t + m < t_min);
        const float tc = cand ? t - m : kInf;
        if (cand) mask |= 1u << k;
        t2 = fminf(t2, fmaxf(tc, tbest));                             // second smallest lower bound so far
        if (tc < tbest) kbest = (uint32_t)k;
        tbest = fminf(tbest, tc);
        }
    }
    closest = FLT_MAX; bestw = kMiss;
    if (COUNT) ctr->rays++;
    if (mask) {
        flat_test_unit(fs.recs, kbest, o, d, t_min, closest, bestw);
        if (COUNT) ctr->unit_tests++;
        if (!(bestw != kMiss && closest < t2)) {                      // not provably the closest: test the other candidates too
            if (COUNT && (mask & ~(1u << kbest))) ctr->slow++;
            mask &= ~(1u << kbest);
            while (mask) {
                const uint32_t k = hi_bit(mask);
                mask &= ~(1u << k);
                flat_test_unit(fs.recs, k, o, d, t_min, closest, bestw);
                if (COUNT) ctr->unit_tests++;
            }
        }
    }
}

}  // namespace ptx

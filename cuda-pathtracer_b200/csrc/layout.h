// layout.h -- HBM data layout of a scene (all records 16-byte aligned for 128-bit loads).
//
// Replaces the reference's AoS `Primitive[]` (4 364 B stride, rendering/primitive.h:21-29),
// `BVHNode[]` (36 B, rendering/bvh.h:63-72) and `int bvh_indices[]`.
//
//   nodes      : BVH8 wide nodes, 80 B = 5 x uint4, compressed as in Ylitie, Karras, Laine,
//                "Efficient Incoherent Ray Traversal on GPUs Through Compressed Wide BVHs" (HPG 2017):
//                  q0 = { p.x, p.y, p.z, ex | ey<<8 | ez<<16 | imask<<24 }
//                  q1 = { child_base, tri_base, meta[0..3], meta[4..7] }
//                  q2 = { qlo_x[0..3], qlo_x[4..7], qlo_y[0..3], qlo_y[4..7] }
//                  q3 = { qlo_z[0..3], qlo_z[4..7], qhi_x[0..3], qhi_x[4..7] }
//                  q4 = { qhi_y[0..3], qhi_y[4..7], qhi_z[0..3], qhi_z[4..7] }
//                child box = p + q * 2^(e-127) per axis (lo rounded down, hi rounded up: conservative);
//                meta[s] = 0 (empty) | 001:11sss (inner child in slot s) | uuu:ooooo (leaf: unary
//                triangle count 001/011/111, 5-bit offset from tri_base).  Children of a node are
//                contiguous from child_base in slot order of the set imask bits; so are its triangles.
//   tris       : precomputed edge records, 48 B = 3 x float4 in BVH leaf order:
//                  { v0.xyz, prim_id }, { e1.xyz, 0 }, { e2.xyz, 0 }   with e1 = v1-v0, e2 = v2-v0
//                (the exact float values Triangle::intersect recomputes per test, triangle.h:66-67).
//                A reference Quad contributes two records (v00,v10,v11) and (v00,v11,v01) (quad.h:60-62,
//                88-90) that share one prim_id.
//   prim_shade : 16 B per reference primitive: { n.xyz, material index } (stored flat normal).
//   materials  : 32 B each: { Kd.rgb, type }, { Ke.rgb, 0 }.
#pragma once
#include "common.h"

namespace ptx {

struct u32x4 {
    uint32_t x, y, z, w;
};
struct f32x4 {
    float x, y, z, w;
};

constexpr uint32_t kNodeQuads = 5;       // uint4 per node
constexpr uint32_t kTriQuads = 3;        // float4 per triangle
constexpr uint32_t kMiss = 0xFFFFFFFFu;

struct InstanceRec;

struct SceneView {
    const u32x4* nodes;
    const f32x4* tris;
    const f32x4* prim_shade;
    const f32x4* materials;
    uint32_t n_nodes, n_tris, n_prims, n_mats;
    // 0x3F800000 (1.0f), read by the kernels as a constant-bank operand of PRMT: see Traversal::step (traverse.h)
    uint32_t byte_magic = 0x3F800000u;   // PRMT source for the byte -> float trick in traverse.h (kept out of the immediate field)
    // grid-guided sampling (shade.h): kGuideFloats floats per reference primitive, or nullptr; Scene::mis_bsdf_fraction
    const float* guide = nullptr;
    float mis_bsdf_fraction = 0.5f;
    // one bit per reference primitive: 1 = Quad (its halves reject t == t_max in the any-hit query, trace_any); nullptr = no quads
    const uint32_t* quad_bits = nullptr;
    // two-level scenes (traverse_inst.h): instance records, or nullptr for an ordinary (flattened) scene
    const struct InstanceRec* inst = nullptr;
    uint32_t n_inst = 0;
};

// Camera in the reference's Sensor terms (rendering/sensor.h:92-95): what get_ray needs.
struct CameraView {
    float origin[3], llc[3], horizontal[3], vertical[3];
    int width, height;
};

// 128-bit read-only loads (LDG.E.128 through the non-coherent path on the device)
PTX_HD u32x4 ldq(const u32x4* p) {
#if defined(__CUDA_ARCH__)
    uint4 v = __ldg(reinterpret_cast<const uint4*>(p));
    u32x4 r; r.x = v.x; r.y = v.y; r.z = v.z; r.w = v.w; return r;
#else
    return *p;
#endif
}
PTX_HD f32x4 ldq(const f32x4* p) {
#if defined(__CUDA_ARCH__)
    float4 v = __ldg(reinterpret_cast<const float4*>(p));
    f32x4 r; r.x = v.x; r.y = v.y; r.z = v.z; r.w = v.w; return r;
#else
    return *p;
#endif
}
}  // namespace ptx

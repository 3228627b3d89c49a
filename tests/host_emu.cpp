// tests/host_emu.cpp -- TEST-ONLY: compiles the product's __host__ __device__ kernel cores
// (traverse.h, shade.h, rng.h) and host builders for the CPU so that `-m "not gpu"` tests can check
// BVH8 construction, traversal, RNG and shading logic against the oracle without a GPU.
// It is never linked into libptx.so; the product path always runs CUDA kernels.
#include <cstdint>
#include <cstring>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "../cuda-pathtracer_b200/csrc/bvh_host.h"
#include "../cuda-pathtracer_b200/csrc/cull.h"
#include "../cuda-pathtracer_b200/csrc/flat_host.h"
#include "../cuda-pathtracer_b200/csrc/instancing_host.h"
#include "../cuda-pathtracer_b200/csrc/loader.h"
#include "../cuda-pathtracer_b200/csrc/rng.h"
#include "../cuda-pathtracer_b200/csrc/scene_pack.h"
#include "../cuda-pathtracer_b200/csrc/shade.h"
#include "../cuda-pathtracer_b200/csrc/traverse.h"

using namespace ptx;

struct EmuScene {
    PackedScene packed;
    Bvh8Result bvh;
    SceneView view;
    FlatHost flat;          // tiny scenes: the unit list of flat.h (empty otherwise)
    FlatScene fs;
};

static void finish(EmuScene* s, int max_leaf) {
    Bvh2 b2;
    build_bvh2_sah(s->packed.tris.data(), s->packed.tris.size(), 16, max_leaf, b2);
    collapse_bvh8(b2, s->packed.tris.data(), s->packed.tris.size(), s->bvh);
    s->view.nodes = s->bvh.nodes.data();
    s->view.tris = s->bvh.tris.data();
    s->view.prim_shade = s->packed.prim_shade.data();
    s->view.materials = s->packed.materials.data();
    s->view.n_nodes = (uint32_t)(s->bvh.nodes.size() / kNodeQuads);
    s->view.n_tris = (uint32_t)s->bvh.tri_src.size();
    s->view.n_prims = (uint32_t)s->packed.prim_shade.size();
    s->view.n_mats = (uint32_t)(s->packed.materials.size() / 2);
    s->view.quad_bits = s->packed.quad_bits.empty() ? nullptr : s->packed.quad_bits.data();
    memset(&s->fs, 0, sizeof s->fs);
    if (build_flat_units(s->packed.tris.data(), s->packed.tris.size(), s->flat)) {
        for (size_t k = 0; k < s->flat.units.size(); k++) flat_set_unit(s->fs, (int)k, s->flat.units[k]);
        s->fs.recs = s->flat.recs.data();
        s->fs.n_units = (uint32_t)s->flat.units.size();
    }
}

extern "C" {

EmuScene* emu_scene_from_prims(const ptx_primitive* prims, int64_t n, int max_leaf) {
    EmuScene* s = new EmuScene();
    pack_primitives(prims, (size_t)n, s->packed);
    finish(s, max_leaf);
    return s;
}
EmuScene* emu_scene_load(const char* path, int convert_quads, int subdiv, int max_leaf) {
    std::vector<ptx_primitive> prims; std::string err;
    if (!load_obj(path, false, prims, err)) return nullptr;
    if (convert_quads) split_quads(prims);
    if (subdiv > 0) subdivide(prims, subdiv);
    return emu_scene_from_prims(prims.data(), (int64_t)prims.size(), max_leaf);
}
void emu_scene_free(EmuScene* s) { delete s; }
int64_t emu_scene_counts(const EmuScene* s, int64_t* n_tris, int64_t* n_nodes, int* depth, float* sah) {
    *n_tris = s->view.n_tris; *n_nodes = s->view.n_nodes; *depth = s->bvh.max_depth; *sah = s->bvh.sah_cost;
    return s->view.n_prims;
}
int64_t emu_parse_obj(const char* path, int convert_quads, int subdiv, ptx_primitive* out, int64_t cap) {
    std::vector<ptx_primitive> prims; std::string err;
    if (!load_obj(path, false, prims, err)) return -1;
    if (convert_quads) split_quads(prims);
    if (subdiv > 0) subdivide(prims, subdiv);
    for (size_t i = 0; i < prims.size() && (int64_t)i < cap; i++) out[i] = prims[i];
    return (int64_t)prims.size();
}

// closest hits for n rays (6 floats each; d normalised here like the Ray ctor)
void emu_trace(const EmuScene* s, const float* rays, int64_t n, float t_min, int32_t* prim, float* t, uint64_t* ctr) {
    for (int64_t i = 0; i < n; i++) {
        vec3 o = ld3(rays + 6 * i), d = unit(ld3(rays + 6 * i + 3));
        float th; TraceCounters c = {0, 0};
        uint32_t tri = trace_closest<32, true>(s->view, o, d, t_min, FLT_MAX, th, &c);
        prim[i] = tri == kMiss ? -1 : (int32_t)f2u(s->view.tris[(size_t)tri * kTriQuads].w);
        t[i] = tri == kMiss ? -1.0f : th;
        if (ctr) { ctr[0] += c.nodes; ctr[1] += c.tris; }
    }
}

void emu_primary_ids(const EmuScene* s, const CameraView* cam, int32_t* prim, float* t) {
    for (int y = 0; y < cam->height; y++)
        for (int x = 0; x < cam->width; x++) {
            float u = (x + 0.5f) / (float)cam->width, v = (y + 0.5f) / (float)cam->height;
            vec3 o, d; camera_ray(*cam, u, v, o, d);
            float th;
            uint32_t tri = trace_closest<32, false>(s->view, o, d, 1e-4f, FLT_MAX, th, nullptr);
            prim[y * cam->width + x] = tri == kMiss ? -1 : (int32_t)f2u(s->view.tris[(size_t)tri * kTriQuads].w);
            t[y * cam->width + x] = tri == kMiss ? -1.0f : th;
        }
}

// tiny scenes: the flat unit list (flat.h).  Returns the unit count (0 = the scene does not qualify).
int emu_flat_units(const EmuScene* s, float* units_out /* 16 floats per unit, may be NULL */) {
    if (units_out) memcpy(units_out, s->flat.units.data(), s->flat.units.size() * sizeof(FlatUnit));
    return (int)s->fs.n_units;
}
// closest hits through flat_closest; ctr3 = { rays, rays that needed the every-candidate path, exact unit tests }
void emu_trace_flat(const EmuScene* s, const float* rays, int64_t n, float t_min, float scale, int32_t* prim, float* t, uint64_t* ctr3) {
    FlatScene fs = s->fs;
    flat_margins(scale > 0.f ? scale : s->flat.scale, fs.m0, fs.m1);
    fs.key_mask = kFlatKeyMask;
    FlatCounters c = {0, 0, 0};
    for (int64_t i = 0; i < n; i++) {
        vec3 o = ld3(rays + 6 * i), d = unit(ld3(rays + 6 * i + 3));
        float th; uint32_t w;
        flat_closest<true>(fs, o, d, t_min, th, w, &c);
        prim[i] = w == kMiss ? -1 : (int32_t)w;
        t[i] = w == kMiss ? -1.0f : th;
    }
    if (ctr3) { ctr3[0] = c.rays; ctr3[1] = c.slow; ctr3[2] = c.unit_tests; }
}

// Scene::occluded(from, to) through trace_any (the arithmetic of k_occluded, kernels.cu)
void emu_occluded(const EmuScene* s, const float* segs, int64_t n, uint8_t* out) {
    for (int64_t i = 0; i < n; i++) {
        vec3 from = ld3(segs + 6 * i);
        vec3 dir = ld3(segs + 6 * i + 3) - from;
        float dist = length(dir);
        dir = div_scalar(dir, dist);
        vec3 o = from + 1e-4f * dir;
        vec3 d = unit(dir);
        out[i] = trace_any<32>(s->view, o, d, 1e-4f, dist - 1e-4f) ? 1 : 0;
    }
}

// ---- two-level scenes on the CPU: the assembly of api.cu scene_from_instanced with the HOST builder in place of the device builder ----
struct EmuInstScene {
    PackedScene packed;
    std::vector<u32x4> nodes;
    std::vector<f32x4> tris;
    std::vector<InstanceRec> inst;
    SceneView view;
    int64_t flattened_prims = 0;
    int depth_tlas = 0, depth_blas = 0;
};
EmuInstScene* emu_inst_load(const char* path, int mirror_ext) {
    InstancedScene is; std::string err;
    if (!load_pbrt_instanced(path, mirror_ext != 0, is, err)) return nullptr;
    EmuInstScene* s = new EmuInstScene();
    std::vector<size_t> obj_first;
    std::vector<ptx_primitive> all;
    const bool has_world = !is.world.empty();
    if (has_world) { obj_first.push_back(0); all.insert(all.end(), is.world.begin(), is.world.end()); }
    for (auto& o : is.objects) { obj_first.push_back(all.size()); all.insert(all.end(), o.begin(), o.end()); }
    obj_first.push_back(all.size());
    const size_t n_obj = obj_first.size() - 1;
    pack_primitives(all.data(), all.size(), s->packed);
    std::vector<Bvh8Result> blas(n_obj);
    for (size_t k = 0; k < n_obj; k++) {
        const size_t first = obj_first[k], n = obj_first[k + 1] - first;
        std::vector<TriIn> local(s->packed.tris.begin() + (ptrdiff_t)first, s->packed.tris.begin() + (ptrdiff_t)(first + n));
        for (auto& t : local) t.prim_id -= (uint32_t)first;
        Bvh2 b2;
        build_bvh2_sah(local.data(), n, 16, 3, b2);
        collapse_bvh8(b2, local.data(), n, blas[k]);
        s->depth_blas = std::max(s->depth_blas, blas[k].max_depth);
    }
    std::vector<TriIn> boxes;
    uint32_t prim_base = 0;
    auto add = [&](size_t obj, const float* m) {
        InstanceRec r; TriIn t;
        if (!make_instance(m, blas[obj].bounds, (uint32_t)obj, prim_base, (uint32_t)obj_first[obj], (uint32_t)s->inst.size(), r, t)) return false;
        prim_base += (uint32_t)(obj_first[obj + 1] - obj_first[obj]);
        boxes.push_back(t); s->inst.push_back(r);
        return true;
    };
    const float ident[12] = {1, 0, 0, 0, 1, 0, 0, 0, 1, 0, 0, 0};
    bool ok = true;
    if (has_world) ok = add(0, ident);
    for (const auto& in : is.instances) if (ok) ok = add((size_t)in.object + (has_world ? 1 : 0), in.m);
    if (!ok) { delete s; return nullptr; }
    Bvh2 t2; Bvh8Result tlas;
    build_bvh2_sah(boxes.data(), boxes.size(), 16, 1, t2);
    collapse_bvh8(t2, boxes.data(), boxes.size(), tlas);
    s->depth_tlas = tlas.max_depth;
    s->nodes = tlas.nodes; s->tris = tlas.tris;
    std::vector<uint32_t> node_off(n_obj);
    for (size_t k = 0; k < n_obj; k++) {
        const uint32_t noff = (uint32_t)(s->nodes.size() / kNodeQuads), toff = (uint32_t)(s->tris.size() / kTriQuads);
        node_off[k] = noff;
        for (size_t i = 0; i < blas[k].nodes.size(); i++) {
            u32x4 q = blas[k].nodes[i];
            if (i % kNodeQuads == 1) { q.x += noff; q.y += toff; }                  // k_offset_nodes
            s->nodes.push_back(q);
        }
        s->tris.insert(s->tris.end(), blas[k].tris.begin(), blas[k].tris.end());
    }
    for (auto& r : s->inst) r.root = node_off[r.root];
    s->flattened_prims = prim_base;
    s->view.nodes = s->nodes.data(); s->view.tris = s->tris.data();
    s->view.prim_shade = s->packed.prim_shade.data(); s->view.materials = s->packed.materials.data();
    s->view.n_nodes = (uint32_t)(s->nodes.size() / kNodeQuads); s->view.n_tris = (uint32_t)(s->tris.size() / kTriQuads);
    s->view.n_prims = prim_base; s->view.n_mats = (uint32_t)(s->packed.materials.size() / 2);
    s->view.inst = s->inst.data(); s->view.n_inst = (uint32_t)s->inst.size();
    return s;
}
void emu_inst_free(EmuInstScene* s) { delete s; }
int64_t emu_inst_counts(const EmuInstScene* s, int64_t* n_inst, int64_t* stored_tris, int* depth_tlas, int* depth_blas) {
    *n_inst = (int64_t)s->inst.size(); *stored_tris = (int64_t)s->packed.tris.size(); *depth_tlas = s->depth_tlas; *depth_blas = s->depth_blas;
    return s->flattened_prims;
}
// closest hits through the two-level traversal; prim = flattened numbering; normals_out (optional, 3 floats per ray) = the world-space
// shading normal resolve_instanced_hit computes
void emu_inst_occluded(const EmuInstScene* s, const float* segs, int64_t n, uint8_t* out) {      // k_occluded_inst on the host
    for (int64_t i = 0; i < n; i++) {
        vec3 from = ld3(segs + 6 * i);
        vec3 dir = ld3(segs + 6 * i + 3) - from;
        float dist = length(dir);
        dir = div_scalar(dir, dist);
        vec3 o = from + 1e-4f * dir;
        vec3 d = unit(dir);
        out[i] = trace_any_instanced<32>(s->view, o, d, 1e-4f, dist - 1e-4f) ? 1 : 0;
    }
}
void emu_inst_trace(const EmuInstScene* s, const float* rays, int64_t n, float t_min, int32_t* prim, float* t, float* normals_out) {
    for (int64_t i = 0; i < n; i++) {
        vec3 o = ld3(rays + 6 * i), d = unit(ld3(rays + 6 * i + 3));
        InstTraversal it; TravStack<32> stk;
        it.begin(o, d, FLT_MAX);
        while (!it.template step<32>(s->view, stk, o, d, t_min)) {}
        prim[i] = -1; t[i] = -1.0f;
        if (it.tr.best != kMiss) {
            uint32_t p, mat; vec3 nn;
            resolve_instanced_hit(s->view, it, p, nn, mat);
            prim[i] = (int32_t)p; t[i] = it.tr.closest;
            if (normals_out) st3(normals_out + 3 * i, nn);
        }
    }
}

void emu_jump_tables(uint32_t* out) { xorwow_build_jump_tables(out); }
void emu_xorwow_init(uint64_t seed, uint64_t subsequence, const uint32_t* tables, uint32_t* out6) {
    Xorwow s = xorwow_seed(seed);
    for (int b = 0; b < kXorwowJumps; b++)
        if ((subsequence >> b) & 1ull) xorwow_apply(s, tables + (size_t)b * kXorwowJumpWords);
    out6[0] = s.v0; out6[1] = s.v1; out6[2] = s.v2; out6[3] = s.v3; out6[4] = s.v4; out6[5] = s.d;
}
// whole-pixel culling (api.cu ptx_render): the stream skip and the screen rectangle
void emu_xorwow_skip(uint32_t* st6, uint64_t n) {               // T^n + the Weyl counter: what k_cull_pixels applies
    std::vector<uint32_t> m((size_t)kXorwowJumpWords);
    xorwow_build_skip_matrix(n, m.data());
    Xorwow s{st6[0], st6[1], st6[2], st6[3], st6[4], st6[5]};
    xorwow_apply(s, m.data());
    s.d += (uint32_t)(362437ull * n);
    st6[0] = s.v0; st6[1] = s.v1; st6[2] = s.v2; st6[3] = s.v3; st6[4] = s.v4; st6[5] = s.d;
}
void emu_xorwow_steps(uint32_t* st6, uint64_t n) {              // n draws one at a time
    Xorwow s{st6[0], st6[1], st6[2], st6[3], st6[4], st6[5]};
    for (uint64_t i = 0; i < n; i++) xorwow_next(s);
    st6[0] = s.v0; st6[1] = s.v1; st6[2] = s.v2; st6[3] = s.v3; st6[4] = s.v4; st6[5] = s.d;
}
int emu_cull_rect(const ptx_camera* cam, const float* bmin, const float* bmax, int w, int h, int* out4) {
    CullRect r{0, 0, w, h};
    bool ok = cull_rect(*cam, bmin, bmax, w, h, r);
    out4[0] = r.x0; out4[1] = r.y0; out4[2] = r.x1; out4[3] = r.y1;
    return ok ? 1 : 0;
}
void emu_philox(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t* out4) {
    Philox4 p = philox4x32_10(c0, c1, c2, c3, k0, k1);
    out4[0] = p.x; out4[1] = p.y; out4[2] = p.z; out4[3] = p.w;
}

// the render loop of integrator.h:371-408 driven through shade.h / traverse.h (compat RNG), serial
static TraceCounters g_render_ctr = {0, 0};     // node visits / triangle tests of the last emu_render (single-threaded test code)
void emu_render_counters(uint64_t* out2) { out2[0] = g_render_ctr.nodes; out2[1] = g_render_ctr.tris; }
int64_t emu_scene_nodes(const EmuScene* s, uint32_t* out, int64_t cap_nodes) {     // raw 80-byte nodes
    int64_t n = (int64_t)s->view.n_nodes;
    for (int64_t i = 0; i < n && i < cap_nodes; i++) memcpy(out + 20 * i, s->bvh.nodes.data() + kNodeQuads * i, 80);
    return n;
}
static void emu_render_impl(const EmuScene* s, const CameraView* cam, int spp, int max_depth, uint64_t seed_base, float* linear,
                            uint8_t* rgb8, uint64_t* segments, bool use_flat, uint64_t* flat_ctr3);
void emu_render(const EmuScene* s, const CameraView* cam, int spp, int max_depth, uint64_t seed_base, float* linear,
                uint8_t* rgb8, uint64_t* segments) {
    emu_render_impl(s, cam, spp, max_depth, seed_base, linear, rgb8, segments, false, nullptr);
}
// the same loop with flat_closest instead of the BVH8 traversal (tiny scenes); margins from the scene and camera extent like ptx_render
void emu_render_flat(const EmuScene* s, const CameraView* cam, int spp, int max_depth, uint64_t seed_base, float* linear,
                     uint8_t* rgb8, uint64_t* segments, uint64_t* flat_ctr3) {
    emu_render_impl(s, cam, spp, max_depth, seed_base, linear, rgb8, segments, true, flat_ctr3);
}
static void emu_render_impl(const EmuScene* s, const CameraView* cam, int spp, int max_depth, uint64_t seed_base, float* linear,
                            uint8_t* rgb8, uint64_t* segments, bool use_flat, uint64_t* flat_ctr3) {
    FlatScene fs = s->fs;
    {
        float scale = s->flat.scale;
        for (int k = 0; k < 3; k++) scale = fmaxf(scale, fabsf(cam->origin[k]));
        flat_margins(scale, fs.m0, fs.m1);
        fs.key_mask = kFlatKeyMask;
    }
    FlatCounters fc = {0, 0, 0};
    std::vector<uint32_t> tables((size_t)kXorwowJumps * kXorwowJumpWords);
    xorwow_build_jump_tables(tables.data());
    ShadeParams sp = {max_depth, 3, 0.95f, 0};
    uint64_t seg = 0;
    g_render_ctr = TraceCounters{0, 0};
    for (int y = 0; y < cam->height; y++)
        for (int x = 0; x < cam->width; x++) {
            int pix = y * cam->width + x;
            uint32_t st6[6];
            emu_xorwow_init(seed_base + (uint64_t)pix, (uint64_t)pix, tables.data(), st6);
            XorwowRng rng; rng.st = Xorwow{st6[0], st6[1], st6[2], st6[3], st6[4], st6[5]};
            vec3 color = mk3(0, 0, 0);
            for (int k = 0; k < spp; k++) {
                float u = (x + rng.next()) / (float)cam->width;
                float v = (y + rng.next()) / (float)cam->height;
                PathVertex pv; camera_ray(*cam, u, v, pv.o, pv.d);
                pv.beta = mk3(1, 1, 1); pv.depth = 0;
                vec3 L = mk3(0, 0, 0);
                for (;;) {
                    float th; seg++;
                    uint32_t prim;
                    if (use_flat) {
                        flat_closest<true>(fs, pv.o, pv.d, 1e-4f, th, prim, &fc);
                        if (getenv("EMU_FLAT_DEBUG")) {
                            float th2; uint32_t tri = trace_closest<32, false>(s->view, pv.o, pv.d, 1e-4f, FLT_MAX, th2, nullptr);
                            uint32_t prim2 = tri == kMiss ? kMiss : f2u(s->view.tris[(size_t)tri * kTriQuads].w);
                            if (prim2 != prim || (prim != kMiss && th != th2))
                                fprintf(stderr, "MISMATCH pix %d depth %d o %.9g %.9g %.9g d %.9g %.9g %.9g flat prim %d t %.9g bvh prim %d t %.9g beta %g %g %g\n", pix, pv.depth, pv.o.x, pv.o.y, pv.o.z, pv.d.x, pv.d.y, pv.d.z, (int)prim, th, (int)prim2, th2, pv.beta.x, pv.beta.y, pv.beta.z);
                        }
                    } else {
                        uint32_t tri = trace_closest<32, true>(s->view, pv.o, pv.d, 1e-4f, FLT_MAX, th, &g_render_ctr);
                        prim = tri == kMiss ? kMiss : f2u(s->view.tris[(size_t)tri * kTriQuads].w);
                    }
                    if (!shade_bounce(s->view, sp, pv, prim, th, rng, L)) break;
                }
                color = color + L;
            }
            color = div_scalar(color, (float)spp);
            if (linear) st3(linear + 3 * pix, color);
            if (rgb8) { rgb8[3 * pix] = tonemap_channel(color.x); rgb8[3 * pix + 1] = tonemap_channel(color.y); rgb8[3 * pix + 2] = tonemap_channel(color.z); }
        }
    if (segments) *segments = seg;
    if (flat_ctr3) { flat_ctr3[0] = fc.rays; flat_ctr3[1] = fc.slow; flat_ctr3[2] = fc.unit_tests; }
}

// grid-guided / MIS modes (shade.h guided_direction) on the CPU: guide records built by build_guide_record from
// n_prims x 256 x 3 radiosity floats, then the render loop above with shade_bounce<.., GUIDED = true>
void emu_build_guide(const float* rgb, int64_t n_prims, float* out) {
    for (int64_t p = 0; p < n_prims; p++) build_guide_record(rgb + p * 768, out + p * kGuideFloats);
}
void emu_render_guided(const EmuScene* s, const CameraView* cam, int spp, int max_depth, uint64_t seed_base, int mode, float mis_bsdf_fraction,
                       const float* guide, float* linear, uint32_t* rng_out, uint64_t* segments) {
    std::vector<uint32_t> tables((size_t)kXorwowJumps * kXorwowJumpWords);
    xorwow_build_jump_tables(tables.data());
    ShadeParams sp = {max_depth, 3, 0.95f, mode};
    SceneView view = s->view;
    view.guide = guide; view.mis_bsdf_fraction = mis_bsdf_fraction;
    uint64_t seg = 0;
    for (int y = 0; y < cam->height; y++)
        for (int x = 0; x < cam->width; x++) {
            int pix = y * cam->width + x;
            uint32_t st6[6];
            emu_xorwow_init(seed_base + (uint64_t)pix, (uint64_t)pix, tables.data(), st6);
            XorwowRng rng; rng.st = Xorwow{st6[0], st6[1], st6[2], st6[3], st6[4], st6[5]};
            vec3 color = mk3(0, 0, 0);
            for (int k = 0; k < spp; k++) {
                float u = (x + rng.next()) / (float)cam->width;
                float v = (y + rng.next()) / (float)cam->height;
                PathVertex pv; camera_ray(*cam, u, v, pv.o, pv.d);
                pv.beta = mk3(1, 1, 1); pv.depth = 0;
                vec3 L = mk3(0, 0, 0);
                for (;;) {
                    float th; seg++;
                    uint32_t tri = trace_closest<32, false>(view, pv.o, pv.d, 1e-4f, FLT_MAX, th, nullptr);
                    uint32_t prim = tri == kMiss ? kMiss : f2u(view.tris[(size_t)tri * kTriQuads].w);
                    if (!shade_bounce<XorwowRng, true>(view, sp, pv, prim, th, rng, L)) break;
                }
                color = color + L;
            }
            color = div_scalar(color, (float)spp);
            if (linear) st3(linear + 3 * pix, color);
            if (rng_out) { uint32_t* o = rng_out + 6 * (size_t)pix; o[0] = rng.st.v0; o[1] = rng.st.v1; o[2] = rng.st.v2; o[3] = rng.st.v3; o[4] = rng.st.v4; o[5] = rng.st.d; }
        }
    if (segments) *segments = seg;
}

}  // extern "C"

// api.cu -- implementation of the C ABI declared in include/ptx.h.
//
// Host-side orchestration only: scene packing / upload, film buffers, and the wavefront loop that
// launches the sm_100a kernels of kernels.cu.  There is no CPU rendering path in this library: every
// entry point that produces pixels or hits launches CUDA kernels and fails with PTX_ERR_NO_DEVICE /
// PTX_ERR_CUDA when that is impossible.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>          // types only: the functions are resolved with dlopen (a single-GPU host needs no NCCL)

#include <cub/device/device_radix_sort.cuh>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/ptx.h"
#include "bvh_gpu.h"
#include "bvh_host.h"
#include "flat_host.h"
#include "kernels.cuh"
#include "geometry_gpu.h"
#include "image_io.h"
#include "instancing_host.h"
#include "loader.h"
#include "rng.h"
#include "scene_pack.h"

using namespace ptx;

namespace {
thread_local std::string g_last_error;
constexpr int kTraversalStack = 24;      // entries of the traversal stack in the kernels (trace_closest<24>, TravStack<24>)
}

struct ptx_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    std::string err;
    cudaDeviceProp prop{};
    uint32_t* d_jump = nullptr;                 // XORWOW jump matrices J[0..63]
    // wavefront state (grown on demand, reused across renders)
    WaveState ws{};
    uint32_t* d_scratch = nullptr; size_t scratch_words = 0;
    void* d_flush = nullptr; int flush_value = 0;            // ptx_ctx_flush_l2
    uint32_t pixmap_capacity = 0, progress_capacity = 0;   // entries allocated in ws.slot_pixel / ws.progress
    unsigned long long* h_counters = nullptr;   // pinned
    int slot_key[5] = {-1, -1, -1, -1, -1};     // (w, h, shard_mode, rank, world) the slot map was built for
    // ws.slot_pixel points at d_slot_base (canonical order: what ptx_film_combine packs by) or at d_slot_sorted (cost order)
    uint32_t* d_slot_base = nullptr; uint32_t* d_slot_sorted = nullptr; uint32_t* d_cost_keys[2] = {nullptr, nullptr}; uint32_t* d_cost_vals = nullptr;
    void* d_cost_tmp = nullptr; size_t cost_tmp_bytes = 0;
    // whole-pixel culling (ptx_render): compacted list of the pixels that can see the scene, owned rows before each row, T^n
    uint32_t* d_slot_active = nullptr; uint32_t* d_cull_rows = nullptr; uint32_t cull_rows_capacity = 0;
    uint32_t* d_skip = nullptr; uint64_t skip_n = 0;
    std::vector<uint32_t> h_cull_rows;
    unsigned long long culled_paths = 0;        // paths (= segments) of the last render that were resolved by culling
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, tev0 = nullptr, tev1 = nullptr;
    std::vector<cudaEvent_t> stage_events;
    ptx_render_stats stats{};
    uint32_t* d_sort_keys[2] = {nullptr, nullptr}; void* d_sort_tmp = nullptr; size_t sort_capacity = 0, sort_tmp_bytes = 0;   // wavefront queue sorting
    bool stats_pending = false;                 // PTX_RENDER_ASYNC: the counters of the last render are still on their way
};

struct ptx_scene {
    ptx_ctx* ctx = nullptr;
    std::vector<ptx_primitive> prims;           // reference numbering (empty for triangle-soup scenes)
    PackedScene packed;                          // host copy kept for rebuilds
    u32x4* d_nodes = nullptr; f32x4* d_tris = nullptr; f32x4* d_prim_shade = nullptr; f32x4* d_materials = nullptr;
    SceneView view{};
    ptx_scene_stats stats{};
    float* d_guide = nullptr;                    // grid-guided sampling: kGuideFloats floats per primitive (shade.h)
    InstanceRec* d_inst = nullptr;               // two-level scenes (traverse_inst.h)
    uint32_t* d_quad_bits = nullptr;             // one bit per primitive: Quad (any-hit end-point rule, traverse.h trace_any)
    FlatHost flat;                               // tiny scenes: the planar-unit list of flat.h (empty when the scene does not qualify)
    f32x4* d_flat_recs = nullptr;
};

struct ptx_film {
    ptx_ctx* ctx = nullptr;
    int w = 0, h = 0;
    float4* d_accum = nullptr; uint4* d_rngA = nullptr; uint2* d_rngB = nullptr; uint8_t* d_rgb8 = nullptr; float* d_linear = nullptr;
    float total_spp = 0.f;
    float frame_spp = 0.f;                       // samples per pixel asked for over ALL ranks since the last clear (sample-range combine)
    uint32_t sample_base = 0;                    // global index of the next frame's first sample (progressive mode)
    unsigned long long seed_base = 2023ull;      // integrator.h:279
    FilmView view() const { FilmView f; f.accum = d_accum; f.rngA = d_rngA; f.rngB = d_rngB; f.width = w; f.height = h; return f; }
};

namespace {

int fail(ptx_ctx* ctx, int code, const std::string& msg) {
    g_last_error = msg;
    if (ctx) ctx->err = msg;
    return code;
}
int cuda_fail(ptx_ctx* ctx, cudaError_t e, const char* what) {
    return fail(ctx, PTX_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}
#define PTX_CUDA(ctx, call)                                              \
    do {                                                                 \
        cudaError_t e_ = (call);                                         \
        if (e_ != cudaSuccess) return cuda_fail((ctx), e_, #call);       \
    } while (0)

CameraView camera_view(const ptx_camera& c, int w, int h) {
    CameraView v;
    memcpy(v.origin, c.origin, 12); memcpy(v.llc, c.lower_left_corner, 12);
    memcpy(v.horizontal, c.horizontal, 12); memcpy(v.vertical, c.vertical, 12);
    v.width = w; v.height = h;
    return v;
}

}  // namespace

extern "C" {

const char* ptx_version(void) { return "ptx-b200 0.1 (sm_100a)"; }

const char* ptx_last_error(const ptx_ctx* ctx) { return ctx ? ctx->err.c_str() : g_last_error.c_str(); }

int ptx_ctx_create(int device, ptx_ctx** out) {
    if (!out) return fail(nullptr, PTX_ERR_INVALID, "ptx_ctx_create: out is NULL");
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return fail(nullptr, PTX_ERR_NO_DEVICE, std::string("no CUDA device available (this library has no CPU path): ") +
                                                     (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0"));
    if (device < 0 || device >= n) return fail(nullptr, PTX_ERR_INVALID, "ptx_ctx_create: bad device index");
    ptx_ctx* c = new ptx_ctx();
    c->device = device;
    std::vector<uint32_t> jump((size_t)kXorwowJumps * kXorwowJumpWords);
    xorwow_build_jump_tables(jump.data());
    cudaError_t ce = cudaSetDevice(device);
    if (ce == cudaSuccess) ce = cudaGetDeviceProperties(&c->prop, device);
    if (ce == cudaSuccess) ce = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
    if (ce == cudaSuccess) ce = cudaEventCreate(&c->ev0);
    if (ce == cudaSuccess) ce = cudaEventCreate(&c->ev1);
    if (ce == cudaSuccess) ce = cudaMallocHost(&c->h_counters, sizeof(unsigned long long) * CTR_COUNT * 2);
    if (ce == cudaSuccess) ce = cudaMalloc(&c->ws.counters, sizeof(unsigned long long) * CTR_COUNT);
    if (ce == cudaSuccess) ce = cudaMalloc(&c->d_jump, jump.size() * sizeof(uint32_t));
    if (ce == cudaSuccess) ce = cudaMemcpy(c->d_jump, jump.data(), jump.size() * sizeof(uint32_t), cudaMemcpyHostToDevice);
    if (ce != cudaSuccess) {                     // nothing half-built survives a failed creation
        int rc = cuda_fail(nullptr, ce, "ptx_ctx_create");
        ptx_ctx_destroy(c);
        return rc;
    }
    *out = c;
    return PTX_OK;
}

void ptx_ctx_destroy(ptx_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    cudaFree(c->ws.A); cudaFree(c->ws.B); cudaFree(c->ws.C); cudaFree(c->ws.D); cudaFree(c->ws.E);
    cudaFree(c->ws.queue[0]); cudaFree(c->ws.queue[1]); cudaFree(c->d_slot_base); cudaFree(c->d_slot_sorted); cudaFree(c->d_slot_active); cudaFree(c->d_cost_keys[0]); cudaFree(c->d_cost_keys[1]);
    cudaFree(c->d_cost_vals); cudaFree(c->d_cost_tmp); cudaFree(c->ws.progress); cudaFree(c->ws.counters); cudaFree(c->d_flush);
    cudaFree(c->d_scratch); cudaFree(c->d_jump); cudaFree(c->d_cull_rows); cudaFree(c->d_skip); cudaFree(c->d_sort_keys[0]); cudaFree(c->d_sort_keys[1]); cudaFree(c->d_sort_tmp);
    cudaFreeHost(c->h_counters);
    for (auto ev : c->stage_events) cudaEventDestroy(ev);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->tev0) { cudaEventDestroy(c->tev0); cudaEventDestroy(c->tev1); }
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

// PTX_RENDER_ASYNC: wait for the last render and complete its statistics from the counters it copied back
static int finalize_pending(ptx_ctx* c) {
    if (!c->stats_pending) return PTX_OK;
    c->stats_pending = false;
    PTX_CUDA(c, cudaSetDevice(c->device));
    PTX_CUDA(c, cudaStreamSynchronize(c->stream));
    const unsigned long long* hc = c->h_counters;
    cudaEventElapsedTime(&c->stats.render_ms, c->ev0, c->ev1);
    c->stats.shade_ms = c->stats.render_ms;
    c->stats.segments = hc[CTR_SEGMENTS] + c->culled_paths; c->stats.paths = hc[CTR_PATHS] + c->culled_paths;
    c->stats.culled_paths = c->culled_paths;
    return PTX_OK;
}

int ptx_sync(ptx_ctx* c) {
    if (!c) return fail(nullptr, PTX_ERR_INVALID, "ptx_sync: ctx is NULL");
    PTX_CUDA(c, cudaSetDevice(c->device));
    PTX_CUDA(c, cudaStreamSynchronize(c->stream));
    return finalize_pending(c);
}

int ptx_timer_start(ptx_ctx* c) {
    if (!c) return fail(nullptr, PTX_ERR_INVALID, "ptx_timer_start: ctx is NULL");
    PTX_CUDA(c, cudaSetDevice(c->device));
    if (!c->tev0) { PTX_CUDA(c, cudaEventCreate(&c->tev0)); PTX_CUDA(c, cudaEventCreate(&c->tev1)); }
    PTX_CUDA(c, cudaEventRecord(c->tev0, c->stream));
    return PTX_OK;
}

int ptx_timer_stop(ptx_ctx* c, float* ms) {
    if (!c || !ms || !c->tev0) return fail(c, PTX_ERR_INVALID, "ptx_timer_stop: timer not started");
    PTX_CUDA(c, cudaSetDevice(c->device));
    PTX_CUDA(c, cudaEventRecord(c->tev1, c->stream));
    PTX_CUDA(c, cudaEventSynchronize(c->tev1));
    PTX_CUDA(c, cudaEventElapsedTime(ms, c->tev0, c->tev1));
    return PTX_OK;
}

int ptx_ctx_stream(const ptx_ctx* c, void** stream) {
    if (!c || !stream) return fail(nullptr, PTX_ERR_INVALID, "ptx_ctx_stream: NULL argument");
    *stream = (void*)c->stream;
    return PTX_OK;
}

int ptx_ctx_device_info(const ptx_ctx* c, int* sm_count, int* cc_major, int* cc_minor, size_t* total_mem) {
    if (!c) return fail(nullptr, PTX_ERR_INVALID, "ptx_ctx_device_info: ctx is NULL");
    if (sm_count) *sm_count = c->prop.multiProcessorCount;
    if (cc_major) *cc_major = c->prop.major;
    if (cc_minor) *cc_minor = c->prop.minor;
    if (total_mem) *total_mem = c->prop.totalGlobalMem;
    return PTX_OK;
}


// ---------------------------------------------------------------------------------------------
// camera: host maths of the reference's Sensor (rendering/sensor.h:16-67), same operation order
// ---------------------------------------------------------------------------------------------
int ptx_camera_update(ptx_camera* cam) {
    if (!cam) return fail(nullptr, PTX_ERR_INVALID, "ptx_camera_update: cam is NULL");
    float theta = (float)((double)cam->vfov * 3.14159265358979323846 / (double)180.0f);   // `vfov * M_PI / 180.0f`
    float half_height = tanf(theta / 2.0f);
    float half_width = cam->aspect * half_height;
    vec3 origin = ld3(cam->origin);
    vec3 w = unit(origin - ld3(cam->lookat));
    vec3 u = unit(cross(ld3(cam->vup), w));
    vec3 v = cross(w, u);
    st3(cam->lower_left_corner, ((origin - half_width * u) - half_height * v) - w);
    st3(cam->horizontal, (2 * half_width) * u);
    st3(cam->vertical, (2 * half_height) * v);
    return PTX_OK;
}

int ptx_camera_set(ptx_camera* cam, const float eye[3], const float at[3], const float up[3], float vfov_deg, float aspect) {
    if (!cam || !eye || !at || !up) return fail(nullptr, PTX_ERR_INVALID, "ptx_camera_set: NULL argument");
    memcpy(cam->origin, eye, 12); memcpy(cam->lookat, at, 12); memcpy(cam->vup, up, 12);
    cam->vfov = vfov_deg; cam->aspect = aspect;
    cam->radius = length(ld3(eye) - ld3(at));
    cam->yaw = 90.0f; cam->pitch = 0.0f;
    if (cam->image_width <= 0) cam->image_width = 0;
    if (cam->image_height <= 0) cam->image_height = 0;
    return ptx_camera_update(cam);
}

int ptx_camera_orbit(ptx_camera* cam) {
    if (!cam) return fail(nullptr, PTX_ERR_INVALID, "ptx_camera_orbit: cam is NULL");
    float yawRad = (float)((double)cam->yaw * 3.14159265358979323846 / (double)180.0f);
    float pitchRad = (float)((double)cam->pitch * 3.14159265358979323846 / (double)180.0f);
    cam->origin[0] = cam->lookat[0] + cam->radius * cosf(pitchRad) * cosf(yawRad);
    cam->origin[1] = cam->lookat[1] + cam->radius * sinf(pitchRad);
    cam->origin[2] = cam->lookat[2] + cam->radius * cosf(pitchRad) * sinf(yawRad);
    return ptx_camera_update(cam);
}

int ptx_camera_default(ptx_camera* cam, int width, int height) {
    if (!cam || width <= 0 || height <= 0) return fail(nullptr, PTX_ERR_INVALID, "ptx_camera_default: bad argument");
    memset(cam, 0, sizeof *cam);
    const float eye[3] = {0.5f, 3.0f, 8.5f}, at[3] = {0.0f, 2.5f, 0.0f}, up[3] = {0.0f, 1.0f, 0.0f};
    cam->image_width = width; cam->image_height = height;
    ptx_camera_set(cam, eye, at, up, 40.0f, 1.0f);                 // RenderState ctor, application_state.h:88
    cam->aspect = (float)width / (float)height;                     // allocateBuffers, application_state.h:106-109
    ptx_camera_update(cam);
    return ptx_camera_orbit(cam);                                   // renderFrame, application.h:162
}


// ---------------------------------------------------------------------------------------------
// scene
// ---------------------------------------------------------------------------------------------
namespace {

void free_accel(ptx_scene* s) {
    cudaFree(s->d_nodes); cudaFree(s->d_tris); cudaFree(s->d_flat_recs);
    s->d_nodes = nullptr; s->d_tris = nullptr; s->d_flat_recs = nullptr;
    // nothing may traverse a scene whose rebuild failed half-way (ptx_render / ptx_trace_* check n_nodes)
    s->view.nodes = nullptr; s->view.tris = nullptr; s->view.n_nodes = 0; s->view.n_tris = 0;
    s->flat.units.clear(); s->flat.recs.clear();
}

int build_accel(ptx_scene* s, const ptx_build_opts* bo, const TriIn* d_tris_in = nullptr) {
    ptx_ctx* c = s->ctx;
    PTX_CUDA(c, cudaSetDevice(c->device));
    int builder = bo ? bo->builder : 0;
    int bins = bo && bo->sah_bins > 0 ? bo->sah_bins : 16;
    int max_leaf = bo && bo->max_leaf > 0 ? bo->max_leaf : 3;
    free_accel(s);
    size_t n = s->packed.tris.size();
    if (n > 0x7FFFFFF0ull) return fail(c, PTX_ERR_INVALID, "scene too large: triangle count must fit 31 bits");
    PTX_CUDA(c, cudaEventRecord(c->ev0, c->stream));
    size_t node_bytes = 0, tri_bytes = 0;
    float sah = 0.f; int depth = 0; aabb bounds;
    if (builder == 0) {
        // device builder: breadth-first binned SAH + BVH8 collapse, result stays in HBM (bvh_gpu.cu)
        GpuBvhResult g; std::string berr;
        cudaError_t e = d_tris_in ? build_bvh8_gpu(c->stream, d_tris_in, n, max_leaf, kTraversalStack, g, berr, true)   // device-resident soup (geometry_gpu.cu)
                                  : build_bvh8_gpu(c->stream, s->packed.tris.data(), n, max_leaf, kTraversalStack, g, berr);
        if (e != cudaSuccess) { cudaFree(g.d_nodes); cudaFree(g.d_tris); return fail(c, PTX_ERR_CUDA, "GPU BVH build failed: " + berr); }
        s->d_nodes = g.d_nodes; s->d_tris = g.d_tris;
        s->view.n_nodes = g.n_nodes; s->view.n_tris = g.n_tris;
        node_bytes = (size_t)g.n_nodes * kNodeQuads * sizeof(u32x4); tri_bytes = (size_t)g.n_tris * kTriQuads * sizeof(f32x4);
        sah = g.sah_cost; depth = g.max_depth; bounds = g.bounds;
        memcpy(s->stats.build_stage_ms, g.stage_ms, sizeof g.stage_ms);
    } else {
        Bvh8Result r;
        Bvh2 b2;
        build_bvh2_sah(s->packed.tris.data(), n, bins, max_leaf, b2);
        collapse_bvh8(b2, s->packed.tris.data(), n, r);
        if (r.max_depth > kTraversalStack) return fail(c, PTX_ERR_INVALID, "BVH8 deeper than the traversal stack");
        builder = 1;
        node_bytes = r.nodes.size() * sizeof(u32x4); tri_bytes = r.tris.size() * sizeof(f32x4);
        PTX_CUDA(c, cudaMalloc(&s->d_nodes, node_bytes));
        PTX_CUDA(c, cudaMalloc(&s->d_tris, std::max<size_t>(tri_bytes, sizeof(f32x4))));
        PTX_CUDA(c, cudaMemcpyAsync(s->d_nodes, r.nodes.data(), node_bytes, cudaMemcpyHostToDevice, c->stream));
        if (!r.tris.empty()) PTX_CUDA(c, cudaMemcpyAsync(s->d_tris, r.tris.data(), tri_bytes, cudaMemcpyHostToDevice, c->stream));
        PTX_CUDA(c, cudaStreamSynchronize(c->stream));      // r goes out of scope
        s->view.n_nodes = (uint32_t)(r.nodes.size() / kNodeQuads);
        s->view.n_tris = (uint32_t)r.tri_src.size();
        sah = r.sah_cost; depth = r.max_depth; bounds = r.bounds;
    }
    // tiny scenes also get the flat unit list (flat.h); the BVH8 stays for the queries and as the cross-check
    if (n <= 2 * (size_t)kFlatUnits && build_flat_units(s->packed.tris.data(), n, s->flat)) {
        const size_t bytes = s->flat.recs.size() * sizeof(f32x4);
        PTX_CUDA(c, cudaMalloc(&s->d_flat_recs, bytes));
        PTX_CUDA(c, cudaMemcpyAsync(s->d_flat_recs, s->flat.recs.data(), bytes, cudaMemcpyHostToDevice, c->stream));
    }
    PTX_CUDA(c, cudaEventRecord(c->ev1, c->stream));
    PTX_CUDA(c, cudaStreamSynchronize(c->stream));
    float ms = 0.f; cudaEventElapsedTime(&ms, c->ev0, c->ev1);
    s->view.nodes = s->d_nodes; s->view.tris = s->d_tris;
    s->stats.flat_units = (int32_t)s->flat.units.size();
    s->stats.n_nodes = s->view.n_nodes; s->stats.n_triangles = s->view.n_tris;
    s->stats.node_bytes = (int64_t)node_bytes; s->stats.triangle_bytes = (int64_t)tri_bytes;
    s->stats.sah_cost = sah; s->stats.max_depth = depth; s->stats.builder = builder; s->stats.build_ms = ms;
    st3(s->stats.bounds_min, bounds.lo); st3(s->stats.bounds_max, bounds.hi);
    return PTX_OK;
}

int finish_scene(ptx_ctx* c, ptx_scene* s, const ptx_build_opts* bo, ptx_scene** out, const TriIn* d_tris_in = nullptr) {
    // takes ownership of s: on any failure the scene (and whatever device memory it already holds) is destroyed
    size_t np = s->packed.prim_shade.size(), nm = s->packed.materials.size();
    cudaError_t e = cudaSetDevice(c->device);
    if (e == cudaSuccess) e = cudaMalloc(&s->d_prim_shade, std::max<size_t>(np, 1) * sizeof(f32x4));
    if (e == cudaSuccess) e = cudaMalloc(&s->d_materials, std::max<size_t>(nm, 2) * sizeof(f32x4));
    if (e == cudaSuccess && np) e = cudaMemcpy(s->d_prim_shade, s->packed.prim_shade.data(), np * sizeof(f32x4), cudaMemcpyHostToDevice);
    if (e == cudaSuccess && nm) e = cudaMemcpy(s->d_materials, s->packed.materials.data(), nm * sizeof(f32x4), cudaMemcpyHostToDevice);
    if (e == cudaSuccess && !s->packed.quad_bits.empty()) {
        e = cudaMalloc(&s->d_quad_bits, s->packed.quad_bits.size() * sizeof(uint32_t));
        if (e == cudaSuccess) e = cudaMemcpy(s->d_quad_bits, s->packed.quad_bits.data(), s->packed.quad_bits.size() * sizeof(uint32_t), cudaMemcpyHostToDevice);
    }
    if (e != cudaSuccess) { ptx_scene_destroy(s); return cuda_fail(c, e, "scene upload"); }
    s->view.prim_shade = s->d_prim_shade; s->view.materials = s->d_materials; s->view.quad_bits = s->d_quad_bits;
    s->view.n_prims = (uint32_t)np; s->view.n_mats = (uint32_t)(nm / 2);
    s->stats.n_primitives = (int64_t)np; s->stats.n_materials = (int64_t)(nm / 2);
    s->stats.shading_bytes = (int64_t)((np + nm) * sizeof(f32x4));
    int rc = build_accel(s, bo, d_tris_in);
    if (rc != PTX_OK) { ptx_scene_destroy(s); return rc; }
    *out = s;
    return PTX_OK;
}

// convert-quads / subdivision / packing as CUDA kernels (geometry_gpu.cu); the builder reads the triangle soup where the
// kernels left it.  Host mirrors (primitive list for ptx_scene_get_primitives, soup for rebuilds) are downloaded afterwards.
int scene_from_base_gpu(ptx_ctx* c, const std::vector<ptx_primitive>& base, bool convert_quads, int levels, ptx_scene** out) {
    PTX_CUDA(c, cudaSetDevice(c->device));
    PackedScene bp;
    pack_primitives(base.data(), base.size(), bp);             // material table + material index per parsed primitive
    std::vector<uint32_t> base_mat(base.size());
    for (size_t i = 0; i < base.size(); i++) base_mat[i] = f2u(bp.prim_shade[i].w);
    DevGeometry g; std::string gerr;
    cudaError_t e = preprocess_geometry_gpu(c->stream, base.data(), base_mat.data(), base.size(), convert_quads, levels, g, gerr);
    if (e != cudaSuccess) { g.release(); return fail(c, e == cudaErrorMemoryAllocation ? PTX_ERR_NOMEM : PTX_ERR_CUDA, "GPU geometry preprocessing failed: " + gerr); }
    ptx_scene* s = new ptx_scene();
    s->ctx = c;
    s->prims.resize(g.n_prims); s->packed.tris.resize(g.n_tris); s->packed.prim_shade.resize(g.n_prims);
    s->packed.materials = bp.materials;
    if (g.n_prims) {
        e = cudaMemcpyAsync(s->prims.data(), g.d_prims, sizeof(ptx_primitive) * g.n_prims, cudaMemcpyDeviceToHost, c->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(s->packed.tris.data(), g.d_tris, sizeof(TriIn) * g.n_tris, cudaMemcpyDeviceToHost, c->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(s->packed.prim_shade.data(), g.d_prim_shade, sizeof(f32x4) * g.n_prims, cudaMemcpyDeviceToHost, c->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
        if (e != cudaSuccess) { g.release(); delete s; return cuda_fail(c, e, "scene_from_base_gpu"); }
    }
    make_quad_bits(s->prims.data(), s->prims.size(), s->packed.quad_bits);
    s->stats.preprocess_ms = g.ms; s->stats.preprocess_on_device = 1;
    int rc = finish_scene(c, s, nullptr, out, g.d_tris);
    g.release();
    return rc;
}

// Two-level scene: one BLAS per object (the world's own shapes are object 0 under the identity), a TLAS over the instances' world
// boxes, all in ONE node array and ONE triangle array (TLAS first).  Every tree comes from the device builder; the TLAS is built
// from one placeholder triangle per instance whose bounding box is the instance's box and whose prim id is the instance index.
int scene_from_instanced(ptx_ctx* c, InstancedScene& is, ptx_scene** out) {
    PTX_CUDA(c, cudaSetDevice(c->device));
    ptx_scene* s = new ptx_scene();
    s->ctx = c;
    auto bail = [&](int rc) { ptx_scene_destroy(s); return rc; };
    // objects: 0 = world shapes (may be empty), then the instanced objects; one global material table / shading array
    std::vector<size_t> obj_first;           // first object-primitive of every object in the concatenated list
    std::vector<ptx_primitive> all;
    const bool has_world = !is.world.empty();
    if (has_world) { obj_first.push_back(0); all.insert(all.end(), is.world.begin(), is.world.end()); }
    for (auto& o : is.objects) { obj_first.push_back(all.size()); all.insert(all.end(), o.begin(), o.end()); o.clear(); o.shrink_to_fit(); }
    obj_first.push_back(all.size());
    const size_t n_obj = obj_first.size() - 1;
    pack_primitives(all.data(), all.size(), s->packed);
    all.clear(); all.shrink_to_fit();
    if (s->packed.tris.size() > 0x7FFFFFF0ull) return bail(fail(c, PTX_ERR_INVALID, "scene too large"));
    cudaError_t e = cudaMalloc(&s->d_prim_shade, std::max<size_t>(s->packed.prim_shade.size(), 1) * sizeof(f32x4));
    if (e == cudaSuccess) e = cudaMalloc(&s->d_materials, std::max<size_t>(s->packed.materials.size(), 2) * sizeof(f32x4));
    if (e == cudaSuccess) e = cudaMemcpy(s->d_prim_shade, s->packed.prim_shade.data(), s->packed.prim_shade.size() * sizeof(f32x4), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(s->d_materials, s->packed.materials.data(), s->packed.materials.size() * sizeof(f32x4), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) return bail(cuda_fail(c, e, "instanced scene upload"));
    PTX_CUDA(c, cudaEventRecord(c->ev0, c->stream));
    // BLAS per object (pbrt scenes hold triangles only: one record per primitive, so tris[k].prim_id == k)
    std::vector<GpuBvhResult> blas(n_obj);
    auto free_blas = [&]() { for (auto& b : blas) { cudaFree(b.d_nodes); cudaFree(b.d_tris); b.d_nodes = nullptr; b.d_tris = nullptr; } };
    int max_blas_depth = 0;
    for (size_t k = 0; k < n_obj; k++) {
        const size_t first = obj_first[k], n = obj_first[k + 1] - first;
        std::vector<TriIn> local(s->packed.tris.begin() + (ptrdiff_t)first, s->packed.tris.begin() + (ptrdiff_t)(first + n));
        for (auto& t : local) t.prim_id -= (uint32_t)first;             // index WITHIN the object
        std::string berr;
        e = build_bvh8_gpu(c->stream, local.data(), n, 3, kTraversalStack, blas[k], berr);
        if (e != cudaSuccess) { free_blas(); return bail(fail(c, PTX_ERR_CUDA, "GPU BVH build (object) failed: " + berr)); }
        max_blas_depth = std::max(max_blas_depth, blas[k].max_depth);
    }
    // instances: the world's own shapes first (identity), then the file's, in flattening order
    std::vector<InstanceRec> recs;
    std::vector<TriIn> boxes;
    uint32_t prim_base = 0;
    auto add_instance = [&](size_t obj, const float* m) -> bool {
        InstanceRec r; TriIn t;
        if (!make_instance(m, blas[obj].bounds, (uint32_t)obj, prim_base, (uint32_t)obj_first[obj], (uint32_t)recs.size(), r, t)) return false;
        prim_base += (uint32_t)(obj_first[obj + 1] - obj_first[obj]);
        boxes.push_back(t); recs.push_back(r);
        return true;
    };
    const float ident[12] = {1, 0, 0, 0, 1, 0, 0, 0, 1, 0, 0, 0};
    bool ok_inst = true;
    if (has_world) ok_inst = add_instance(0, ident);
    for (const auto& in : is.instances) if (ok_inst) ok_inst = add_instance((size_t)in.object + (has_world ? 1 : 0), in.m);
    if (!ok_inst) { free_blas(); return bail(fail(c, PTX_ERR_INVALID, "an instance transform is singular")); }
    GpuBvhResult tlas; std::string terr;
    e = build_bvh8_gpu(c->stream, boxes.data(), boxes.size(), 1, kTraversalStack, tlas, terr);
    if (e != cudaSuccess) { free_blas(); cudaFree(tlas.d_nodes); cudaFree(tlas.d_tris); return bail(fail(c, PTX_ERR_CUDA, "GPU BVH build (TLAS) failed: " + terr)); }
    // a TLAS level can park two stack entries: the rest of its group and its inner hits
    if (2 * tlas.max_depth + 1 + max_blas_depth > kTraversalStack) { free_blas(); cudaFree(tlas.d_nodes); cudaFree(tlas.d_tris); return bail(fail(c, PTX_ERR_INVALID, "two-level tree deeper than the traversal stack")); }
    // one node array, one triangle array: TLAS first, then every BLAS with its bases shifted
    size_t n_nodes = tlas.n_nodes, n_tris = tlas.n_tris;
    std::vector<uint32_t> node_off(n_obj), tri_off(n_obj);
    for (size_t k = 0; k < n_obj; k++) { node_off[k] = (uint32_t)n_nodes; tri_off[k] = (uint32_t)n_tris; n_nodes += blas[k].n_nodes; n_tris += blas[k].n_tris; }
    e = cudaMalloc(&s->d_nodes, n_nodes * kNodeQuads * sizeof(u32x4));
    if (e == cudaSuccess) e = cudaMalloc(&s->d_tris, std::max<size_t>(n_tris, 1) * kTriQuads * sizeof(f32x4));
    if (e == cudaSuccess) e = cudaMemcpyAsync(s->d_nodes, tlas.d_nodes, (size_t)tlas.n_nodes * kNodeQuads * sizeof(u32x4), cudaMemcpyDeviceToDevice, c->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(s->d_tris, tlas.d_tris, (size_t)tlas.n_tris * kTriQuads * sizeof(f32x4), cudaMemcpyDeviceToDevice, c->stream);
    for (size_t k = 0; k < n_obj && e == cudaSuccess; k++) {
        e = cudaMemcpyAsync(s->d_nodes + (size_t)node_off[k] * kNodeQuads, blas[k].d_nodes, (size_t)blas[k].n_nodes * kNodeQuads * sizeof(u32x4), cudaMemcpyDeviceToDevice, c->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(s->d_tris + (size_t)tri_off[k] * kTriQuads, blas[k].d_tris, (size_t)blas[k].n_tris * kTriQuads * sizeof(f32x4), cudaMemcpyDeviceToDevice, c->stream);
        if (e == cudaSuccess) { launch_offset_nodes(c->stream, s->d_nodes + (size_t)node_off[k] * kNodeQuads, blas[k].n_nodes, node_off[k], tri_off[k]); e = cudaGetLastError(); }
    }
    for (auto& r : recs) r.root = node_off[r.root];
    if (e == cudaSuccess) e = cudaMalloc(&s->d_inst, recs.size() * sizeof(InstanceRec));
    if (e == cudaSuccess) e = cudaMemcpyAsync(s->d_inst, recs.data(), recs.size() * sizeof(InstanceRec), cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) e = cudaEventRecord(c->ev1, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    free_blas(); cudaFree(tlas.d_nodes); cudaFree(tlas.d_tris);
    if (e != cudaSuccess) return bail(cuda_fail(c, e, "instanced scene assembly"));
    float ms = 0.f; cudaEventElapsedTime(&ms, c->ev0, c->ev1);
    s->view.nodes = s->d_nodes; s->view.tris = s->d_tris; s->view.prim_shade = s->d_prim_shade; s->view.materials = s->d_materials;
    s->view.n_nodes = (uint32_t)n_nodes; s->view.n_tris = (uint32_t)n_tris;
    s->view.n_prims = prim_base; s->view.n_mats = (uint32_t)(s->packed.materials.size() / 2);
    s->view.inst = s->d_inst; s->view.n_inst = (uint32_t)recs.size();
    s->stats.n_primitives = (int64_t)prim_base; s->stats.n_triangles = (int64_t)prim_base;       // what the flattened scene would hold
    s->stats.n_nodes = (int64_t)n_nodes; s->stats.n_materials = s->view.n_mats;
    s->stats.node_bytes = (int64_t)(n_nodes * kNodeQuads * sizeof(u32x4)); s->stats.triangle_bytes = (int64_t)(n_tris * kTriQuads * sizeof(f32x4));
    s->stats.shading_bytes = (int64_t)((s->packed.prim_shade.size() + s->packed.materials.size()) * sizeof(f32x4) + recs.size() * sizeof(InstanceRec));
    s->stats.sah_cost = tlas.sah_cost; s->stats.max_depth = tlas.max_depth + 1 + max_blas_depth; s->stats.builder = 0; s->stats.build_ms = ms;
    s->stats.n_instances = (int32_t)recs.size(); s->stats.stored_triangles = (int64_t)s->packed.tris.size();
    st3(s->stats.bounds_min, tlas.bounds.lo); st3(s->stats.bounds_max, tlas.bounds.hi);
    *out = s;
    return PTX_OK;
}

bool has_suffix(const std::string& s, const char* suf) {
    size_t n = strlen(suf);
    if (s.size() < n) return false;
    for (size_t i = 0; i < n; i++) if (tolower((unsigned char)s[s.size() - n + i]) != suf[i]) return false;
    return true;
}

int parse_file(const char* path, const ptx_load_opts* opts, std::vector<ptx_primitive>& prims, std::string& err) {
    std::string p = path ? path : "";
    if (has_suffix(p, ".obj")) {
        if (!load_obj(p, opts && opts->mirror_from_mtl, prims, err)) return PTX_ERR_IO;
    } else if (has_suffix(p, ".pbrt")) {
        if (!load_pbrt(p, opts && opts->mirror_from_mtl, opts ? (long long)opts->triangle_cap : 0ll, prims, err)) return PTX_ERR_IO;
    } else {
        err = "unsupported scene format (supported: .obj, .pbrt): " + p;      // loadScene, file_manager.h:291-315
        return PTX_ERR_UNSUPPORTED;
    }
    if (opts && opts->convert_quads) split_quads(prims);                // application_state.h:404-407
    if (opts && opts->subdivision > 0) subdivide(prims, opts->subdivision);   // application_state.h:409-413
    return PTX_OK;
}

}  // namespace

int ptx_parse_scene_file(const char* path, const ptx_load_opts* opts, ptx_primitive** prims_out, int64_t* n_out) {
    if (!path || !prims_out || !n_out) return fail(nullptr, PTX_ERR_INVALID, "ptx_parse_scene_file: NULL argument");
    std::vector<ptx_primitive> prims; std::string err;
    int rc = parse_file(path, opts, prims, err);
    if (rc != PTX_OK) return fail(nullptr, rc, err);
    ptx_primitive* buf = (ptx_primitive*)malloc(sizeof(ptx_primitive) * std::max<size_t>(prims.size(), 1));
    if (!buf) return fail(nullptr, PTX_ERR_NOMEM, "out of host memory");
    memcpy(buf, prims.data(), sizeof(ptx_primitive) * prims.size());
    *prims_out = buf; *n_out = (int64_t)prims.size();
    return PTX_OK;
}
void ptx_free_primitives(ptx_primitive* prims) { free(prims); }

int ptx_scene_from_primitives(ptx_ctx* c, const ptx_primitive* prims, int64_t n, const ptx_build_opts* bo, ptx_scene** out) {
    if (!c || !out || n < 0 || (n > 0 && !prims)) return fail(c, PTX_ERR_INVALID, "ptx_scene_from_primitives: bad argument");
    ptx_scene* s = new ptx_scene();
    s->ctx = c;
    s->prims.assign(prims, prims + n);
    pack_primitives(prims, (size_t)n, s->packed);
    return finish_scene(c, s, bo, out);
}

int ptx_scene_from_triangles(ptx_ctx* c, const float* v, const float* normals, const uint32_t* mat_id, const ptx_material* mats,
                             size_t n_tri, size_t n_mat, const ptx_build_opts* bo, ptx_scene** out) {
    if (!c || !out || (n_tri > 0 && !v) || (n_mat > 0 && !mats)) return fail(c, PTX_ERR_INVALID, "ptx_scene_from_triangles: bad argument");
    ptx_scene* s = new ptx_scene();
    s->ctx = c;
    pack_triangles(v, normals, mat_id, mats, n_tri, n_mat, s->packed);
    return finish_scene(c, s, bo, out);
}

int ptx_scene_load(ptx_ctx* c, const char* path, const ptx_load_opts* opts, ptx_scene** out) {
    if (!c || !path || !out) return fail(c, PTX_ERR_INVALID, "ptx_scene_load: NULL argument");
    std::vector<ptx_primitive> prims; std::string err;
    if (opts && opts->keep_instances && has_suffix(path, ".pbrt") && !opts->convert_quads && opts->subdivision <= 0 && opts->triangle_cap <= 0) {
        InstancedScene is;
        if (load_pbrt_instanced(path, opts->mirror_from_mtl != 0, is, err)) return scene_from_instanced(c, is, out);
        if (!err.empty()) return fail(c, PTX_ERR_IO, err);
        // no instances / nested instancing: fall through to the flattened load
    }
    const bool gpu_geometry = opts && (opts->convert_quads || opts->subdivision > 0) && !opts->host_preprocess;
    ptx_load_opts base_opts{};
    if (opts) base_opts = *opts;
    if (gpu_geometry) { base_opts.convert_quads = 0; base_opts.subdivision = 0; }      // the parser only parses
    int rc = parse_file(path, &base_opts, prims, err);
    if (rc != PTX_OK) return fail(c, rc, err);
    if (gpu_geometry) return scene_from_base_gpu(c, prims, opts->convert_quads != 0, opts->subdivision, out);
    return ptx_scene_from_primitives(c, prims.data(), (int64_t)prims.size(), nullptr, out);
}

int ptx_scene_build_accel(ptx_scene* s, const ptx_build_opts* opts) {
    if (!s) return fail(nullptr, PTX_ERR_INVALID, "ptx_scene_build_accel: scene is NULL");
    if (s->view.inst) return fail(s->ctx, PTX_ERR_UNSUPPORTED, "ptx_scene_build_accel: a two-level scene is built once at load time");
    return build_accel(s, opts);
}

int ptx_scene_stats_get(const ptx_scene* s, ptx_scene_stats* out) {
    if (!s || !out) return fail(nullptr, PTX_ERR_INVALID, "ptx_scene_stats_get: NULL argument");
    *out = s->stats;
    return PTX_OK;
}

int ptx_scene_get_primitives(const ptx_scene* s, ptx_primitive* out, int64_t capacity) {
    if (!s || (capacity > 0 && !out)) return fail(nullptr, PTX_ERR_INVALID, "ptx_scene_get_primitives: NULL argument");
    int64_t n = std::min<int64_t>(capacity, (int64_t)s->prims.size());
    if (n > 0) memcpy(out, s->prims.data(), sizeof(ptx_primitive) * (size_t)n);
    return (int)std::min<int64_t>((int64_t)s->prims.size(), 0x7FFFFFFF);
}

void ptx_scene_destroy(ptx_scene* s) {
    if (!s) return;
    cudaSetDevice(s->ctx->device);
    cudaStreamSynchronize(s->ctx->stream);
    free_accel(s);
    cudaFree(s->d_prim_shade); cudaFree(s->d_materials); cudaFree(s->d_guide); cudaFree(s->d_quad_bits); cudaFree(s->d_inst);
    delete s;
}

// SceneState::precomputeCDFs (application_state.h:492-585) + Scene::mis_bsdf_fraction (scene.h:215)
int ptx_scene_set_radiosity_grids(ptx_scene* s, const float* rgb_host, float mis_bsdf_fraction) {
    if (!s) return fail(nullptr, PTX_ERR_INVALID, "ptx_scene_set_radiosity_grids: scene is NULL");
    ptx_ctx* c = s->ctx;
    PTX_CUDA(c, cudaSetDevice(c->device));
    PTX_CUDA(c, cudaStreamSynchronize(c->stream));
    s->view.mis_bsdf_fraction = mis_bsdf_fraction;
    if (s->d_guide) { cudaFree(s->d_guide); s->d_guide = nullptr; s->view.guide = nullptr; }
    if (!rgb_host) return PTX_OK;
    const size_t n = s->view.n_prims;
    if (!n) return PTX_OK;
    float* d_rgb = nullptr;
    PTX_CUDA(c, cudaMalloc(&d_rgb, n * 768 * sizeof(float)));
    cudaError_t e = cudaMalloc(&s->d_guide, n * kGuideFloats * sizeof(float));
    if (e != cudaSuccess) { cudaFree(d_rgb); s->d_guide = nullptr; return cuda_fail(c, e, "cudaMalloc(guide)"); }
    e = cudaMemcpyAsync(d_rgb, rgb_host, n * 768 * sizeof(float), cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) { launch_build_guide(c->stream, d_rgb, n, s->d_guide); e = cudaGetLastError(); }
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(d_rgb);
    if (e != cudaSuccess) { cudaFree(s->d_guide); s->d_guide = nullptr; return cuda_fail(c, e, "guide build"); }
    s->view.guide = s->d_guide;
    return PTX_OK;
}
int ptx_scene_get_guide_records(const ptx_scene* s, int64_t first, int64_t n, float* out_host) {
    if (!s || !out_host || first < 0 || n < 0 || (uint64_t)(first + n) > s->view.n_prims) return fail(s ? s->ctx : nullptr, PTX_ERR_INVALID, "ptx_scene_get_guide_records: bad argument");
    if (!s->d_guide) return fail(s->ctx, PTX_ERR_INVALID, "ptx_scene_get_guide_records: the scene has no radiosity grids");
    ptx_ctx* c = s->ctx;
    PTX_CUDA(c, cudaSetDevice(c->device));
    PTX_CUDA(c, cudaStreamSynchronize(c->stream));
    PTX_CUDA(c, cudaMemcpy(out_host, s->d_guide + (size_t)first * kGuideFloats, (size_t)n * kGuideFloats * sizeof(float), cudaMemcpyDeviceToHost));
    return PTX_OK;
}


// ---------------------------------------------------------------------------------------------
// film
// ---------------------------------------------------------------------------------------------
int ptx_film_create(ptx_ctx* c, int width, int height, ptx_film** out) {
    if (!c || !out || width <= 0 || height <= 0) return fail(c, PTX_ERR_INVALID, "ptx_film_create: bad argument");
    if ((long long)width * height > 0x7FFFFFFFll || width > 65535 || height > 65535)      // div_exact_pos (common.h) covers coordinates up to 65536
        return fail(c, PTX_ERR_INVALID, "ptx_film_create: image too large (at most 65535 x 65535 and 2^31 - 1 pixels)");
    PTX_CUDA(c, cudaSetDevice(c->device));
    ptx_film* f = new ptx_film();
    f->ctx = c; f->w = width; f->h = height;
    size_t npix = (size_t)width * height;
    cudaError_t e;
    if ((e = cudaMalloc(&f->d_accum, npix * sizeof(float4))) != cudaSuccess || (e = cudaMalloc(&f->d_rngA, npix * sizeof(uint4))) != cudaSuccess ||
        (e = cudaMalloc(&f->d_rngB, npix * sizeof(uint2))) != cudaSuccess || (e = cudaMalloc(&f->d_rgb8, npix * 3)) != cudaSuccess ||
        (e = cudaMalloc(&f->d_linear, npix * 3 * sizeof(float))) != cudaSuccess) {
        ptx_film_destroy(f);
        return cuda_fail(c, e, "ptx_film_create: cudaMalloc");
    }
    *out = f;
    return ptx_film_reset(f, 2023ull);
}

int ptx_film_reset(ptx_film* f, uint64_t seed_base) {
    if (!f) return fail(nullptr, PTX_ERR_INVALID, "ptx_film_reset: film is NULL");
    ptx_ctx* c = f->ctx;
    PTX_CUDA(c, cudaSetDevice(c->device));
    f->seed_base = seed_base;
    launch_rng_init(c->stream, f->view(), seed_base, c->d_jump);
    PTX_CUDA(c, cudaGetLastError());
    return ptx_film_clear(f);
}

int ptx_film_clear(ptx_film* f) {
    if (!f) return fail(nullptr, PTX_ERR_INVALID, "ptx_film_clear: film is NULL");
    ptx_ctx* c = f->ctx;
    PTX_CUDA(c, cudaSetDevice(c->device));
    PTX_CUDA(c, cudaMemsetAsync(f->d_accum, 0, (size_t)f->w * f->h * sizeof(float4), c->stream));
    f->total_spp = 0.f; f->frame_spp = 0.f; f->sample_base = 0;
    return PTX_OK;
}

int ptx_film_resolve_rgb8(ptx_film* f, uint8_t* dst, int dst_is_device, int flip_y) {
    if (!f || (!dst && !dst_is_device)) return fail(nullptr, PTX_ERR_INVALID, "ptx_film_resolve_rgb8: NULL argument");
    if (!dst) dst = f->d_rgb8;                                         // device-resident result in the film's own buffer
    ptx_ctx* c = f->ctx;
    PTX_CUDA(c, cudaSetDevice(c->device));
    float inv = f->total_spp > 0.f ? 1.0f / f->total_spp : 0.f;      // `color /= float(spp)`: k = 1/spp, multiply
    size_t bytes = (size_t)f->w * f->h * 3;
    launch_resolve(c->stream, f->view(), inv, dst_is_device ? dst : f->d_rgb8, flip_y);
    PTX_CUDA(c, cudaGetLastError());
    if (!dst_is_device) PTX_CUDA(c, cudaMemcpyAsync(dst, f->d_rgb8, bytes, cudaMemcpyDeviceToHost, c->stream));
    PTX_CUDA(c, cudaStreamSynchronize(c->stream));
    return PTX_OK;
}

int ptx_film_read_linear(ptx_film* f, float* rgb_host) {
    if (!f || !rgb_host) return fail(nullptr, PTX_ERR_INVALID, "ptx_film_read_linear: NULL argument");
    ptx_ctx* c = f->ctx;
    PTX_CUDA(c, cudaSetDevice(c->device));
    float inv = f->total_spp > 0.f ? 1.0f / f->total_spp : 0.f;
    launch_film_scale(c->stream, f->view(), inv, f->d_linear);
    PTX_CUDA(c, cudaGetLastError());
    PTX_CUDA(c, cudaMemcpyAsync(rgb_host, f->d_linear, (size_t)f->w * f->h * 3 * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    PTX_CUDA(c, cudaStreamSynchronize(c->stream));
    return PTX_OK;
}

int ptx_film_write(ptx_film* f, const char* path) {
    if (!f || !path) return fail(nullptr, PTX_ERR_INVALID, "ptx_film_write: NULL argument");
    std::string p = path;
    size_t npix = (size_t)f->w * f->h;
    FILE* fp = nullptr;
    if (has_suffix(p, ".pfm")) {
        std::vector<float> lin(npix * 3);
        int rc = ptx_film_read_linear(f, lin.data());
        if (rc != PTX_OK) return rc;
        if (!(fp = fopen(path, "wb"))) return fail(f->ctx, PTX_ERR_IO, "cannot write " + p);
        fprintf(fp, "PF\n%d %d\n-1.0\n", f->w, f->h);                  // PFM rows run bottom to top, like the film
        fwrite(lin.data(), sizeof(float), lin.size(), fp);
    } else if (has_suffix(p, ".png")) {
        std::vector<uint8_t> img(npix * 3);
        int rc = ptx_film_resolve_rgb8(f, img.data(), 0, 0);             // film rows run bottom to top (the reference's h_image) ...
        if (rc != PTX_OK) return rc;
        if (!write_png_rgb8(path, f->w, f->h, img.data(), true))         // ... and the writer flips, like stbi_flip_vertically_on_write(1)
            return fail(f->ctx, PTX_ERR_IO, "cannot write " + p);
        return PTX_OK;
    } else {
        std::vector<uint8_t> img(npix * 3);
        int rc = ptx_film_resolve_rgb8(f, img.data(), 0, 1);             // vertical flip like the PNG export (ui_windows.h:205-207)
        if (rc != PTX_OK) return rc;
        if (!(fp = fopen(path, "wb"))) return fail(f->ctx, PTX_ERR_IO, "cannot write " + p);
        fprintf(fp, "P6\n%d %d\n255\n", f->w, f->h);
        fwrite(img.data(), 1, img.size(), fp);
    }
    fclose(fp);
    return PTX_OK;
}

// Checkpoint / resume of a progressive render (no counterpart in the reference, whose RNG state and image live in device memory
// only: SURVEY.md 5.4).  File: magic, version, width, height, seed base, sample counters, then the running sums (float4 per pixel)
// and the XORWOW streams (uint4 + uint2 per pixel) exactly as the kernels hold them.
namespace {
struct FilmFileHeader {
    char magic[8];                 // "PTXFILM\0"
    uint32_t version, width, height, sample_base;
    uint64_t seed_base;
    float total_spp, frame_spp;
};
}  // namespace

int ptx_film_save(ptx_film* f, const char* path) {
    if (!f || !path) return fail(nullptr, PTX_ERR_INVALID, "ptx_film_save: NULL argument");
    ptx_ctx* c = f->ctx;
    PTX_CUDA(c, cudaSetDevice(c->device));
    { int frc = finalize_pending(c); if (frc != PTX_OK) return frc; }
    const size_t npix = (size_t)f->w * f->h;
    std::vector<float4> acc(npix); std::vector<uint4> ra(npix); std::vector<uint2> rb(npix);
    PTX_CUDA(c, cudaMemcpyAsync(acc.data(), f->d_accum, npix * sizeof(float4), cudaMemcpyDeviceToHost, c->stream));
    PTX_CUDA(c, cudaMemcpyAsync(ra.data(), f->d_rngA, npix * sizeof(uint4), cudaMemcpyDeviceToHost, c->stream));
    PTX_CUDA(c, cudaMemcpyAsync(rb.data(), f->d_rngB, npix * sizeof(uint2), cudaMemcpyDeviceToHost, c->stream));
    PTX_CUDA(c, cudaStreamSynchronize(c->stream));
    FilmFileHeader hd{};
    memcpy(hd.magic, "PTXFILM", 8);
    hd.version = 1; hd.width = (uint32_t)f->w; hd.height = (uint32_t)f->h; hd.sample_base = f->sample_base;
    hd.seed_base = f->seed_base; hd.total_spp = f->total_spp; hd.frame_spp = f->frame_spp;
    FILE* fp = fopen(path, "wb");
    if (!fp) return fail(c, PTX_ERR_IO, std::string("cannot write ") + path);
    bool ok = fwrite(&hd, sizeof hd, 1, fp) == 1 && fwrite(acc.data(), sizeof(float4), npix, fp) == npix &&
              fwrite(ra.data(), sizeof(uint4), npix, fp) == npix && fwrite(rb.data(), sizeof(uint2), npix, fp) == npix;
    ok = (fclose(fp) == 0) && ok;
    if (!ok) return fail(c, PTX_ERR_IO, std::string("short write to ") + path);
    return PTX_OK;
}

int ptx_film_load(ptx_film* f, const char* path) {
    if (!f || !path) return fail(nullptr, PTX_ERR_INVALID, "ptx_film_load: NULL argument");
    ptx_ctx* c = f->ctx;
    PTX_CUDA(c, cudaSetDevice(c->device));
    { int frc = finalize_pending(c); if (frc != PTX_OK) return frc; }
    FILE* fp = fopen(path, "rb");
    if (!fp) return fail(c, PTX_ERR_IO, std::string("cannot open ") + path);
    FilmFileHeader hd{};
    if (fread(&hd, sizeof hd, 1, fp) != 1 || memcmp(hd.magic, "PTXFILM", 8) != 0 || hd.version != 1) {
        fclose(fp);
        return fail(c, PTX_ERR_IO, std::string(path) + ": not a film checkpoint (magic / version)");
    }
    if (hd.width != (uint32_t)f->w || hd.height != (uint32_t)f->h) {
        fclose(fp);
        return fail(c, PTX_ERR_INVALID, std::string(path) + ": the checkpoint is " + std::to_string(hd.width) + " x " + std::to_string(hd.height) +
                                         ", the film " + std::to_string(f->w) + " x " + std::to_string(f->h));
    }
    const size_t npix = (size_t)f->w * f->h;
    std::vector<float4> acc(npix); std::vector<uint4> ra(npix); std::vector<uint2> rb(npix);
    const bool ok = fread(acc.data(), sizeof(float4), npix, fp) == npix && fread(ra.data(), sizeof(uint4), npix, fp) == npix &&
                    fread(rb.data(), sizeof(uint2), npix, fp) == npix;
    fclose(fp);
    if (!ok) return fail(c, PTX_ERR_IO, std::string(path) + ": truncated film checkpoint");
    PTX_CUDA(c, cudaMemcpyAsync(f->d_accum, acc.data(), npix * sizeof(float4), cudaMemcpyHostToDevice, c->stream));
    PTX_CUDA(c, cudaMemcpyAsync(f->d_rngA, ra.data(), npix * sizeof(uint4), cudaMemcpyHostToDevice, c->stream));
    PTX_CUDA(c, cudaMemcpyAsync(f->d_rngB, rb.data(), npix * sizeof(uint2), cudaMemcpyHostToDevice, c->stream));
    PTX_CUDA(c, cudaStreamSynchronize(c->stream));
    f->seed_base = hd.seed_base; f->sample_base = hd.sample_base; f->total_spp = hd.total_spp; f->frame_spp = hd.frame_spp;
    return PTX_OK;
}

int ptx_write_png_rgb8(const char* path, int width, int height, const uint8_t* rgb_host, int flip_y) {
    if (!path || !rgb_host || width <= 0 || height <= 0) return fail(nullptr, PTX_ERR_INVALID, "ptx_write_png_rgb8: bad argument");
    if (!write_png_rgb8(path, width, height, rgb_host, flip_y != 0)) return fail(nullptr, PTX_ERR_IO, std::string("cannot write ") + path);
    return PTX_OK;
}

int ptx_film_device_buffer(ptx_film* f, void** dev_ptr, size_t* bytes) {
    if (!f || !dev_ptr) return fail(nullptr, PTX_ERR_INVALID, "ptx_film_device_buffer: NULL argument");
    *dev_ptr = f->d_accum;
    if (bytes) *bytes = (size_t)f->w * f->h * sizeof(float4);
    return PTX_OK;
}

int ptx_film_set_sample_count(ptx_film* f, float spp_total) {
    if (!f) return fail(nullptr, PTX_ERR_INVALID, "ptx_film_set_sample_count: film is NULL");
    f->total_spp = spp_total;
    return PTX_OK;
}

int ptx_film_rng_state(ptx_film* f, int64_t first, int64_t n, uint32_t* out_host) {
    if (!f || !out_host || first < 0 || n < 0 || first + n > (int64_t)f->w * f->h) return fail(nullptr, PTX_ERR_INVALID, "ptx_film_rng_state: bad range");
    ptx_ctx* c = f->ctx;
    PTX_CUDA(c, cudaSetDevice(c->device));
    std::vector<uint4> a((size_t)n); std::vector<uint2> b((size_t)n);
    PTX_CUDA(c, cudaStreamSynchronize(c->stream));
    if (n) {
        PTX_CUDA(c, cudaMemcpy(a.data(), f->d_rngA + first, (size_t)n * sizeof(uint4), cudaMemcpyDeviceToHost));
        PTX_CUDA(c, cudaMemcpy(b.data(), f->d_rngB + first, (size_t)n * sizeof(uint2), cudaMemcpyDeviceToHost));
    }
    for (int64_t i = 0; i < n; i++) {
        uint32_t* o = out_host + 6 * i;
        o[0] = a[i].x; o[1] = a[i].y; o[2] = a[i].z; o[3] = a[i].w; o[4] = b[i].x; o[5] = b[i].y;
    }
    return PTX_OK;
}

void ptx_film_destroy(ptx_film* f) {
    if (!f) return;
    cudaSetDevice(f->ctx->device);
    cudaStreamSynchronize(f->ctx->stream);
    cudaFree(f->d_accum); cudaFree(f->d_rngA); cudaFree(f->d_rngB); cudaFree(f->d_rgb8); cudaFree(f->d_linear);
    delete f;
}


// ---------------------------------------------------------------------------------------------
// render: the wavefront loop
// ---------------------------------------------------------------------------------------------
namespace {

int ensure_wave_capacity(ptx_ctx* c, uint32_t slots, uint32_t owned) {
    uint32_t need = std::max(slots, owned);
    if (need > c->ws.capacity) {
        cudaFree(c->ws.A); cudaFree(c->ws.B); cudaFree(c->ws.C); cudaFree(c->ws.D); cudaFree(c->ws.E);
        cudaFree(c->ws.queue[0]); cudaFree(c->ws.queue[1]);
        c->ws.A = c->ws.B = c->ws.C = c->ws.E = nullptr; c->ws.D = nullptr; c->ws.queue[0] = c->ws.queue[1] = nullptr;
        c->ws.capacity = 0;
        size_t n = need;
        PTX_CUDA(c, cudaMalloc(&c->ws.A, n * sizeof(float4)));
        PTX_CUDA(c, cudaMalloc(&c->ws.B, n * sizeof(float4)));
        PTX_CUDA(c, cudaMalloc(&c->ws.C, n * sizeof(float4)));
        PTX_CUDA(c, cudaMalloc(&c->ws.E, n * sizeof(float4)));
        PTX_CUDA(c, cudaMalloc(&c->ws.D, n * sizeof(uint2)));
        PTX_CUDA(c, cudaMalloc(&c->ws.queue[0], n * sizeof(uint32_t)));
        PTX_CUDA(c, cudaMalloc(&c->ws.queue[1], n * sizeof(uint32_t)));
        c->ws.capacity = need;
    }
    size_t words = (size_t)need / 1024 + 8;
    if (words > c->scratch_words) {
        cudaFree(c->d_scratch); c->d_scratch = nullptr; c->scratch_words = 0;
        PTX_CUDA(c, cudaMalloc(&c->d_scratch, words * sizeof(uint32_t)));
        c->scratch_words = words;
    }
    return PTX_OK;
}

// slot -> pixel map and per-pixel chunk progress: all the persistent kernel needs besides the film
int ensure_pixmap(ptx_ctx* c, uint32_t owned, uint32_t film_pixels) {
    if (owned > c->pixmap_capacity) {
        PTX_CUDA(c, cudaStreamSynchronize(c->stream));
        cudaFree(c->d_slot_base); cudaFree(c->d_slot_sorted); cudaFree(c->d_slot_active); cudaFree(c->d_cost_keys[0]); cudaFree(c->d_cost_keys[1]); cudaFree(c->d_cost_vals); cudaFree(c->d_cost_tmp);
        c->d_slot_base = c->d_slot_sorted = c->d_slot_active = c->d_cost_keys[0] = c->d_cost_keys[1] = c->d_cost_vals = nullptr; c->d_cost_tmp = nullptr;
        c->ws.slot_pixel = nullptr; c->pixmap_capacity = 0; c->slot_key[0] = -1;
        const size_t bytes = (size_t)owned * sizeof(uint32_t);
        PTX_CUDA(c, cudaMalloc(&c->d_slot_base, bytes)); PTX_CUDA(c, cudaMalloc(&c->d_slot_sorted, bytes)); PTX_CUDA(c, cudaMalloc(&c->d_slot_active, bytes));
        PTX_CUDA(c, cudaMalloc(&c->d_cost_keys[0], bytes)); PTX_CUDA(c, cudaMalloc(&c->d_cost_keys[1], bytes)); PTX_CUDA(c, cudaMalloc(&c->d_cost_vals, bytes));
        size_t need = 0;
        PTX_CUDA(c, cub::DeviceRadixSort::SortPairs(nullptr, need, c->d_cost_keys[0], c->d_cost_keys[1], c->d_cost_vals, c->d_slot_sorted, (int)owned, 0, 8, c->stream));
        PTX_CUDA(c, cudaMalloc(&c->d_cost_tmp, need));
        c->cost_tmp_bytes = need;
        c->pixmap_capacity = owned;
    }
    if (film_pixels > c->progress_capacity) {
        cudaFree(c->ws.progress);
        c->ws.progress = nullptr; c->progress_capacity = 0;
        PTX_CUDA(c, cudaMalloc(&c->ws.progress, (size_t)film_pixels * sizeof(uint32_t)));
        c->progress_capacity = film_pixels;
    }
    return PTX_OK;
}

constexpr uint32_t kPhiloxPool = 1u << 22;   // slots in flight in counter-RNG mode

}  // namespace

int ptx_render(ptx_ctx* c, const ptx_scene* scene, const ptx_camera* cam, ptx_film* film, const ptx_render_opts* opts) {
    if (!c || !scene || !cam || !film || !opts) return fail(c, PTX_ERR_INVALID, "ptx_render: NULL argument");
    if (scene->ctx != c || film->ctx != c) return fail(c, PTX_ERR_INVALID, "ptx_render: scene/film belong to another context");
    if (opts->sampling_mode < PTX_SAMPLING_BSDF || opts->sampling_mode > PTX_SAMPLING_TOPK) return fail(c, PTX_ERR_INVALID, "ptx_render: bad sampling_mode");
    // grid / MIS modes: like the reference they fall back to cosine sampling when the scene carries no radiosity grids
    const bool guided = opts->sampling_mode != PTX_SAMPLING_BSDF && scene->view.guide != nullptr;
    if (guided && opts->rng_mode != PTX_RNG_XORWOW_COMPAT)
        return fail(c, PTX_ERR_UNSUPPORTED, "grid / MIS sampling runs with the reference's per-pixel XORWOW streams only (rng_mode = PTX_RNG_XORWOW_COMPAT)");
    if (guided && opts->pipeline != 0 && opts->pipeline != 3)
        return fail(c, PTX_ERR_UNSUPPORTED, "grid / MIS sampling is implemented in the persistent kernel only (pipeline 0 / 3)");
    if (opts->spp < 0) return fail(c, PTX_ERR_INVALID, "ptx_render: spp < 0");
    int rng_mode = opts->rng_mode;
    if (rng_mode != PTX_RNG_XORWOW_COMPAT && rng_mode != PTX_RNG_PHILOX) return fail(c, PTX_ERR_INVALID, "ptx_render: bad rng_mode");
    int shard_mode = opts->shard_mode, rank = opts->shard_rank, world = opts->shard_world > 0 ? opts->shard_world : 1;
    if (shard_mode == PTX_SHARD_NONE) { rank = 0; world = 1; }
    if (rank < 0 || rank >= world) return fail(c, PTX_ERR_INVALID, "ptx_render: bad shard rank/world");
    if (shard_mode == PTX_SHARD_SAMPLES && rng_mode != PTX_RNG_PHILOX)
        return fail(c, PTX_ERR_INVALID, "sample-range sharding needs the counter-based RNG (the XORWOW stream is sequential per pixel)");
    PTX_CUDA(c, cudaSetDevice(c->device));
    { int frc = finalize_pending(c); if (frc != PTX_OK) return frc; }

    const int w = film->w, h = film->h;
    RenderParams rp{};
    rp.shade.max_depth = opts->max_depth > 0 ? opts->max_depth : 5;
    rp.shade.rr_start_depth = opts->rr_start_depth > 0 ? opts->rr_start_depth : 3;
    rp.shade.rr_cap = opts->rr_cap > 0.f ? opts->rr_cap : 0.95f;
    rp.shade.sampling_mode = guided ? opts->sampling_mode : 0;
    rp.seed = film->seed_base;
    int spp_local = opts->spp;
    if (!opts->accumulate) film->sample_base = 0;
    rp.sample_begin = film->sample_base;                   // progressive frames continue the sample index
    if (shard_mode == PTX_SHARD_SAMPLES) {
        int32_t lo = 0, cnt = 0;
        ptx_shard_sample_range(opts->spp, shard_mode, rank, world, &lo, &cnt);
        spp_local = cnt;
        rp.sample_begin += (uint32_t)lo;
    }
    rp.spp = spp_local;
    rp.n_owned = owned_pixel_count(w, h, shard_mode, rank, world);
    rp.total_paths = (unsigned long long)rp.n_owned * (unsigned long long)spp_local;
    rp.n_slots = rng_mode == PTX_RNG_XORWOW_COMPAT ? rp.n_owned : (uint32_t)std::min<unsigned long long>(rp.total_paths, kPhiloxPool);

    if (!scene->view.nodes) return fail(c, PTX_ERR_INVALID, "ptx_render: the scene has no acceleration structure (a rebuild failed)");
    // 0 = default: the flat-list kernel when the scene is tiny enough to have one (flat.h), else the persistent BVH8 kernel
    const bool has_flat = !scene->flat.units.empty() && scene->d_flat_recs != nullptr;
    if (opts->pipeline == 6 && (!has_flat || guided))
        return fail(c, PTX_ERR_UNSUPPORTED, "pipeline 6 (flat unit list) needs a scene of at most 20 planar units and BSDF sampling");
    const int pipeline = opts->pipeline == 0 ? ((has_flat && !guided) ? 6 : 3) : opts->pipeline;
#if !defined(PTX_EXPERIMENTAL)
    if (pipeline == 5) return fail(c, PTX_ERR_UNSUPPORTED, "pipeline 5 (CTA wavefront) is an experiment: build the library with -DPTX_EXPERIMENTAL");
#endif
    if (scene->view.inst && (pipeline != 3 || guided))
        return fail(c, PTX_ERR_UNSUPPORTED, "two-level (instanced) scenes render with the persistent kernel and BSDF sampling only (pipeline 0 / 3)");
    const bool cta_wave = pipeline == 5;                               // CTA wavefront: same items and hand-off as the persistent kernel
    const bool flat = pipeline == 6;
    const bool persist = pipeline == 3 || cta_wave || flat;
    int rc = ensure_pixmap(c, rp.n_owned, persist ? (uint32_t)w * (uint32_t)h : 0u);
    if (rc == PTX_OK && !persist) rc = ensure_wave_capacity(c, rp.n_slots, rp.n_owned);
    if (rc != PTX_OK) return rc;
    // compat RNG: a pixel's samples form one sequential chain; the persistent kernel splits it into n_chunks work items
    // (the pixel's RNG state and running sum handed on through the film between them), handed out chunk-major.  The
    // boundaries: U equal parts, the last of them halved again and again down to spp / 32, so the tail of the frame is
    // bounded by a 1/32 pixel.  U = 2 (1/2, 1/4, ... : 6 items per pixel) costs the fewest hand-offs and is right when every
    // lane gets many pixels; with few pixels per lane (a tile-sharded 1080p frame on 8 GPUs: 2.3) a pixel whose first half
    // starts late cannot finish its remaining half inside the frame, and more, smaller parts keep all chains moving in
    // step.  opts->chunks overrides: n > 0 = n halving chunks, n < 0 = U = -n.
    {
        const uint32_t spp = (uint32_t)std::max(rp.spp, 0);
        const uint32_t floor_sz = std::max(1u, spp / 32u);
        int U = 2;
        if (opts->chunks < 0) U = -opts->chunks;
        else if (opts->chunks == 0) {
            const double per_lane = (double)rp.n_owned / ((double)c->prop.multiProcessorCount * 768.0);
            // C2 on one rank of 1 / 2 / 4 / 8: 18 / 9 / 4.6 / 2.3 pixels per lane (tools/shard_probe.py).  Parts shorter than 64
            // samples cost more in hand-offs than they gain (C1, 64 spp: 13.8 -> 12.2 G with eight parts)
            U = (per_lane >= 8.0 || spp < 512u) ? 2 : 8;
        }
        U = (int)std::min<uint32_t>((uint32_t)std::max(U, 1), std::max(1u, std::min(spp, 32u)));
        int n = 0;
        rp.chunk_start[0] = 0;
        if (opts->chunks > 0) {
            int log2spp = 0;
            while ((spp >> (log2spp + 1)) > 0) log2spp++;
            const int want = std::min(opts->chunks, log2spp + 1);
            for (n = 1; n < want; n++) rp.chunk_start[n] = spp - (spp >> n);
        } else if (spp > 0) {
            for (n = 1; n < U; n++) rp.chunk_start[n] = (uint32_t)((unsigned long long)spp * (unsigned)n / (unsigned)U);
            uint32_t at = rp.chunk_start[U - 1], left = spp - at;
            while (left > floor_sz && left / 2u >= 1u && n < kMaxChunks - 1) { at += left - left / 2u; left = spp - at; rp.chunk_start[n++] = at; }
        } else n = 1;
        rp.n_chunks = std::max(n, 1);
        for (int i = rp.n_chunks; i <= kMaxChunks; i++) rp.chunk_start[i] = spp;
    }
    rp.shade_batch = opts->shade_batch > 0 ? opts->shade_batch : 12;      // lanes of a warp that must be waiting before it shades
    int key[5] = {w, h, shard_mode, rank, world};
    if (memcmp(key, c->slot_key, sizeof key) != 0) {
        launch_slot_pixels(c->stream, c->d_slot_base, w, h, shard_mode, rank, world, rp.n_owned);
        memcpy(c->slot_key, key, sizeof key);
    }
    c->ws.slot_pixel = c->d_slot_base;
    if (!opts->accumulate) { rc = ptx_film_clear(film); if (rc != PTX_OK) return rc; }

    CameraView cv = camera_view(*cam, w, h);
    FlatScene fs{};
    if (flat) {
        for (size_t k = 0; k < scene->flat.units.size(); k++) flat_set_unit(fs, (int)k, scene->flat.units[k]);
        fs.recs = scene->d_flat_recs;
        fs.n_units = (uint32_t)scene->flat.units.size();
        float scale = scene->flat.scale;                               // rays start at the camera or on the scene
        for (int k = 0; k < 3; k++) scale = std::max(scale, std::fabs(cam->origin[k]));
        flat_margins(scale, fs.m0, fs.m1);
        fs.key_mask = kFlatKeyMask;
    }
    // Whole-pixel culling (persistent pipelines).  A pixel all of whose camera rays miss the scene's bounding box contributes nothing:
    // each of its samples draws the two jitter numbers, misses and ends (integrator.h:384-385, 199-201).  Such pixels are taken off
    // the work list; with the sequential RNG their streams are advanced by 2 * spp draws (k_cull_pixels), so the film AND the RNG state
    // after the frame are what tracing them leaves, bit for bit.  Their paths / segments are counted in the statistics (and
    // separately in culled_paths).  The test is conservative: the screen rectangle of the projected bounds, 2 pixels of margin; any
    // corner behind the camera turns it off (cull.h).
    PTX_CUDA(c, cudaEventRecord(c->ev0, c->stream));       // render_ms covers the cull, the cost probe + sort and the path kernel
    c->culled_paths = 0;
    int pre_launches = 0;                                  // own kernels launched before the path kernel (cull, cost probe)
    CullRect cull{0, 0, w, h};
    if (persist && rp.spp > 0 && rp.n_owned > 0 && !(opts->flags & PTX_RENDER_NO_BOUNDS_CULL) && cull_rect(*cam, scene->stats.bounds_min, scene->stats.bounds_max, w, h, cull)) {
        const int rw = cull.x1 - cull.x0;
        c->h_cull_rows.assign((size_t)h, 0u);
        uint32_t rows = 0;
        for (int y = 0; y < h; y++) {
            c->h_cull_rows[(size_t)y] = rows;
            const bool owned = shard_mode != PTX_SHARD_TILES || world <= 1 || (y / kBandRows) % world == rank;
            if (owned && y >= cull.y0 && y < cull.y1) rows++;
        }
        const uint32_t n_active = rows * (uint32_t)rw;
        if ((unsigned long long)n_active * 16ull < (unsigned long long)rp.n_owned * 15ull) {      // worth it: more than 1/16 of the pixels go
            if ((uint32_t)h > c->cull_rows_capacity) {
                cudaFree(c->d_cull_rows); c->d_cull_rows = nullptr; c->cull_rows_capacity = 0;
                PTX_CUDA(c, cudaMalloc(&c->d_cull_rows, (size_t)h * sizeof(uint32_t)));
                c->cull_rows_capacity = (uint32_t)h;
            }
            PTX_CUDA(c, cudaMemcpyAsync(c->d_cull_rows, c->h_cull_rows.data(), (size_t)h * sizeof(uint32_t), cudaMemcpyHostToDevice, c->stream));
            const uint32_t* skip = nullptr; uint32_t weyl = 0;
            if (rng_mode == PTX_RNG_XORWOW_COMPAT) {
                const uint64_t n = 2ull * (uint64_t)rp.spp;
                if (!c->d_skip) PTX_CUDA(c, cudaMalloc(&c->d_skip, sizeof(uint32_t) * kXorwowJumpWords));
                if (c->skip_n != n) {
                    std::vector<uint32_t> m((size_t)kXorwowJumpWords);
                    xorwow_build_skip_matrix(n, m.data());
                    PTX_CUDA(c, cudaStreamSynchronize(c->stream));       // an earlier frame may still read the old matrix
                    PTX_CUDA(c, cudaMemcpy(c->d_skip, m.data(), sizeof(uint32_t) * kXorwowJumpWords, cudaMemcpyHostToDevice));
                    c->skip_n = n;
                }
                skip = c->d_skip; weyl = (uint32_t)(362437ull * n);
            }
            launch_cull_pixels(c->stream, film->view(), c->d_slot_base, rp.n_owned, cull, c->d_cull_rows, skip, weyl, c->d_slot_active);
            PTX_CUDA(c, cudaGetLastError());
            pre_launches++;
            c->culled_paths = (unsigned long long)(rp.n_owned - n_active) * (unsigned long long)rp.spp;
            c->ws.slot_pixel = c->d_slot_active;
            rp.n_owned = n_active;
            rp.total_paths = (unsigned long long)n_active * (unsigned long long)rp.spp;
        }
    }
    // Work order, sequential RNG: the longest-expected pixels first (kernels.cu k_pixel_cost), so that no long sample chain starts late.
    // Costs 2 probe paths per pixel (0.2 % of a 1 024-spp frame) + an 8-bit radix sort; only the order of the work changes.
    // On a tree (not the flat list) consecutive pixels share nodes in L1 / L2, which the cost order breaks up: C3 at 1080p is 0.9 % faster in
    // film order (profiles/r2s_cost_order_c3.log), so there the order is only used when a lane gets few pixels (sharded / small frames).
    const bool order_pays = flat || (unsigned long long)rp.n_owned < 8ull * (unsigned long long)c->prop.multiProcessorCount * 3ull * 256ull;
    // (from 256 spp: the two probe paths per pixel are 3 % of a 64-spp frame -- C1 2.04 ms with, 1.87 ms without the ordering: profiles/r2ag_c1_ab.log)
    if (persist && !cta_wave && rng_mode == PTX_RNG_XORWOW_COMPAT && rp.spp >= 256 && rp.n_owned > 0 && !scene->view.inst && order_pays && !(opts->flags & PTX_RENDER_NO_COST_ORDER)) {
        launch_pixel_cost(c->stream, scene->view, flat ? &fs : nullptr, cv, rp.shade, c->ws.slot_pixel, rp.n_owned, c->d_cost_keys[0], c->d_cost_vals);
        PTX_CUDA(c, cudaGetLastError());
        size_t tmp = c->cost_tmp_bytes;
        PTX_CUDA(c, cub::DeviceRadixSort::SortPairs(c->d_cost_tmp, tmp, c->d_cost_keys[0], c->d_cost_keys[1], c->d_cost_vals, c->d_slot_sorted, (int)rp.n_owned, 0, 8, c->stream));
        c->ws.slot_pixel = c->d_slot_sorted;
        pre_launches++;                                    // k_pixel_cost (the radix sort's kernels are cub's, not counted)
    }
    FilmView fv = film->view();
    const bool fused = pipeline == 1;
    const bool path_loop = pipeline == 2;
    // bounces per path-loop launch: a compat-mode slot lives for >= spp segments, so long launches waste
    // little on lanes that die early; opts->loop_iters overrides (tuning / tests)
    int loop_iters = opts->loop_iters > 0 ? opts->loop_iters : 2 * std::min(std::max(rp.spp, 16), 256);
    const bool count_work = (opts->flags & 2) != 0;
    const bool stage_timing = (opts->flags & 1) != 0;

    unsigned long long live0 = (rp.spp > 0) ? std::min<unsigned long long>(rp.n_slots, rng_mode == PTX_RNG_PHILOX ? rp.total_paths : rp.n_slots) : 0ull;
    unsigned long long* hc = c->h_counters + CTR_COUNT;   // upload half of the pinned block (the first half is the read-back buffer)
    for (int i = 0; i < CTR_COUNT; i++) hc[i] = 0;
    hc[CTR_LIVE] = live0; hc[CTR_QLEN] = rp.n_slots; hc[CTR_NEXT_PATH] = persist ? 0ull : rp.n_slots;
    PTX_CUDA(c, cudaMemcpyAsync(c->ws.counters, hc, sizeof(unsigned long long) * CTR_COUNT, cudaMemcpyHostToDevice, c->stream));
    hc = c->h_counters;                                   // read-back half from here on
    const bool async = persist && (opts->flags & PTX_RENDER_ASYNC) != 0;

    ptx_render_stats st{};
    st.kernel_launches = (uint64_t)pre_launches;
    if (persist) {      // one launch for the whole frame: items are pulled from CTR_NEXT_PATH, nothing else to drive
        if (rng_mode == PTX_RNG_XORWOW_COMPAT && rp.n_owned) PTX_CUDA(c, cudaMemsetAsync(c->ws.progress, 0, (size_t)w * h * sizeof(uint32_t), c->stream));
        if (flat) {
            int per_sm = opts->ctas_per_sm & 15;
            if (per_sm == 0) {
                // Resident CTAs per SM (256 threads each).  Bulk throughput is best with 4 (64 registers; C2 28.6 against 27.0 G
                // with 3).  With the sequential RNG a frame cannot end before its longest pixel chain, and a chain advances
                // faster the fewer warps share its SM: when a rank owns few pixels per resident lane AND the chains are long
                // (spp >= 512: a tile-sharded 1080p frame on 8 GPUs) 2 CTAs finish sooner; short chains (C1: 64 spp, 2.3
                // pixels per lane) are throughput-bound again and want 3 (23.5 against 19.7 G with 2, 20.5 with 4).
                // With whole-pixel culling n_owned counts only the pixels that can see the scene -- all of them long chains on the Cornell
                // frames -- and the thresholds were re-measured on one rank of 1 / 2 / 4 / 8 (7.7 / 3.8 / 1.9 / 0.96 such pixels per lane;
                // profiles/r2ad_shard_probe_cull.log): 4 CTAs 99.0 / 53.0 / 32.2 / 19.4 ms, 3 CTAs 101.1 / 50.8 / 26.3 / 19.1, 2 CTAs - / 56.9 / 29.3 / 15.3.
                const double px_per_lane = (double)rp.n_owned / ((double)c->prop.multiProcessorCount * 768.0);
                if (rng_mode == 0 && px_per_lane < 1.5 && rp.spp >= 512) per_sm = 2;
                else per_sm = px_per_lane >= 6.0 ? 4 : 3;
            }
            launch_path_flat(c->stream, rng_mode, c->ws, scene->view, fs, fv, cv, rp, c->prop.multiProcessorCount, per_sm);
        }
#if defined(PTX_EXPERIMENTAL)
        else if (cta_wave) launch_path_cta(c->stream, rng_mode, c->ws, scene->view, fv, cv, rp, c->prop.multiProcessorCount, opts->ctas_per_sm);
#endif
        else launch_path_persist(c->stream, rng_mode, c->ws, scene->view, fv, cv, rp, c->prop.multiProcessorCount, opts->ctas_per_sm);
        PTX_CUDA(c, cudaGetLastError());
        st.kernel_launches++;
        live0 = 0;
    } else {
        launch_raygen(c->stream, rng_mode, c->ws, fv, cv, rp);
        st.kernel_launches++;
    }

    const int kBatch = 8;
    if (stage_timing && c->stage_events.size() < (size_t)kBatch * 3) {
        while (c->stage_events.size() < (size_t)kBatch * 3) { cudaEvent_t e; PTX_CUDA(c, cudaEventCreate(&e)); c->stage_events.push_back(e); }
    }
    uint32_t qbound = rp.n_slots;
    unsigned long long live = live0;
    int q = 0;
    // in compat mode no slot can die before it has traced spp segments
    long long first_check = rng_mode == PTX_RNG_XORWOW_COMPAT ? std::max(1, rp.spp) : 1;
    while (live > 0) {
        int iters = (int)std::min<long long>(std::max<long long>(first_check, 1), 4096);
        if (st.iterations > 0) iters = kBatch;
        if (stage_timing) iters = std::min(iters, kBatch);
        if (path_loop) {
            iters = loop_iters;
            if (stage_timing) cudaEventRecord(c->stage_events[1], c->stream);
            launch_path_loop(c->stream, rng_mode, c->ws, scene->view, fv, cv, rp, q, qbound, iters, opts->ctas_per_sm);
            if (stage_timing) cudaEventRecord(c->stage_events[2], c->stream);
            st.kernel_launches++; st.iterations += (uint64_t)iters;
        }
        for (int i = 0; i < iters && !path_loop; i++) {
            if (fused) {
                if (stage_timing) cudaEventRecord(c->stage_events[3 * i + 1], c->stream);
                launch_bounce(c->stream, rng_mode, c->ws, scene->view, fv, cv, rp, q, qbound);
                if (stage_timing) cudaEventRecord(c->stage_events[3 * i + 2], c->stream);
                st.kernel_launches++;
            } else {
                // north_star subsystem 3: reorder the rays before extend / sort the hits by material before shade (opt-in, pipeline 4)
                auto sort_queue = [&](bool by_material) -> int {
                    if (qbound > c->sort_capacity) {
                        PTX_CUDA(c, cudaStreamSynchronize(c->stream));
                        cudaFree(c->d_sort_keys[0]); cudaFree(c->d_sort_keys[1]); cudaFree(c->d_sort_tmp);
                        c->d_sort_keys[0] = c->d_sort_keys[1] = nullptr; c->d_sort_tmp = nullptr; c->sort_capacity = 0;
                        PTX_CUDA(c, cudaMalloc(&c->d_sort_keys[0], (size_t)qbound * 4)); PTX_CUDA(c, cudaMalloc(&c->d_sort_keys[1], (size_t)qbound * 4));
                        size_t need = 0;
                        PTX_CUDA(c, cub::DeviceRadixSort::SortPairs(nullptr, need, c->d_sort_keys[0], c->d_sort_keys[1], c->ws.queue[0], c->ws.queue[1], (int)qbound, 0, 31, c->stream));
                        PTX_CUDA(c, cudaMalloc(&c->d_sort_tmp, need));
                        c->sort_tmp_bytes = need; c->sort_capacity = qbound;
                    }
                    launch_queue_keys(c->stream, c->ws, scene->view, q, qbound, by_material, scene->stats.bounds_min, scene->stats.bounds_max, c->d_sort_keys[0]);
                    size_t tmp = c->sort_tmp_bytes;
                    PTX_CUDA(c, cub::DeviceRadixSort::SortPairs(c->d_sort_tmp, tmp, c->d_sort_keys[0], c->d_sort_keys[1], c->ws.queue[q], c->ws.queue[q ^ 1], (int)qbound, 0, 31, c->stream));
                    q ^= 1; st.kernel_launches += 4;
                    return PTX_OK;
                };
                if (opts->flags & PTX_RENDER_SORT_RAYS) { int src = sort_queue(false); if (src != PTX_OK) return src; }
                if (stage_timing) cudaEventRecord(c->stage_events[3 * i], c->stream);
                launch_extend(c->stream, c->ws, scene->view, q, qbound, count_work);
                if (stage_timing) cudaEventRecord(c->stage_events[3 * i + 1], c->stream);
                if (opts->flags & PTX_RENDER_SORT_MATERIAL) { int src = sort_queue(true); if (src != PTX_OK) return src; }
                launch_shade(c->stream, rng_mode, c->ws, scene->view, fv, cv, rp, q, qbound);
                if (stage_timing) cudaEventRecord(c->stage_events[3 * i + 2], c->stream);
                st.kernel_launches += 2;
            }
            st.iterations++;
        }
        PTX_CUDA(c, cudaGetLastError());
        PTX_CUDA(c, cudaMemcpyAsync(hc, c->ws.counters, sizeof(unsigned long long) * CTR_COUNT, cudaMemcpyDeviceToHost, c->stream));
        PTX_CUDA(c, cudaStreamSynchronize(c->stream));
        if (stage_timing)
            for (int i = 0; i < (path_loop ? 1 : iters); i++) {
                float a = 0.f, b = 0.f;
                if (!fused && !path_loop) cudaEventElapsedTime(&a, c->stage_events[3 * i], c->stage_events[3 * i + 1]);
                cudaEventElapsedTime(&b, c->stage_events[3 * i + 1], c->stage_events[3 * i + 2]);
                st.extend_ms += a; st.shade_ms += b;
            }
        live = hc[CTR_LIVE];
        uint32_t qlen = (uint32_t)hc[CTR_QLEN];
        if (live > 0 && qlen > 8192u && live * 4ull < (unsigned long long)qlen * 3ull) {      // < 75 % of the queue is live: compact
            launch_compact(c->stream, c->ws, q, qbound, c->d_scratch);
            st.kernel_launches += 4; st.compactions++;
            q ^= 1;
            qbound = (uint32_t)live;        // stream order: nothing dies between the read-back and the compaction
        }
    }
    PTX_CUDA(c, cudaEventRecord(c->ev1, c->stream));
    PTX_CUDA(c, cudaMemcpyAsync(hc, c->ws.counters, sizeof(unsigned long long) * CTR_COUNT, cudaMemcpyDeviceToHost, c->stream));
    if (shard_mode == PTX_SHARD_SAMPLES) film->total_spp += (float)spp_local;   // partial: ptx_film_combine (or the caller) sets the total
    else film->total_spp += (float)opts->spp;
    film->frame_spp += (float)opts->spp;
    film->sample_base += (uint32_t)opts->spp;
    if (async) {                                          // completed by finalize_pending()
        PTX_CUDA(c, cudaGetLastError());
        c->stats = st;
        c->stats_pending = true;
        return PTX_OK;
    }
    PTX_CUDA(c, cudaStreamSynchronize(c->stream));
    PTX_CUDA(c, cudaGetLastError());
    cudaEventElapsedTime(&st.render_ms, c->ev0, c->ev1);
    if (persist) st.shade_ms = st.render_ms;               // the one kernel is the whole render
    st.segments = hc[CTR_SEGMENTS] + c->culled_paths; st.paths = hc[CTR_PATHS] + c->culled_paths;
    st.culled_paths = c->culled_paths;
    st.other_ms = st.render_ms - st.extend_ms - st.shade_ms;
    st.node_visits = hc[CTR_NODE_VISITS]; st.tri_tests = hc[CTR_TRI_TESTS]; st.slot_launches = hc[CTR_SLOT_LAUNCHES];
    c->stats = st;
    return PTX_OK;
}

int64_t ptx_shard_pixel_count(int width, int height, int shard_mode, int rank, int world) {
    if (width <= 0 || height <= 0 || world <= 0 || rank < 0 || rank >= world) return fail(nullptr, PTX_ERR_INVALID, "ptx_shard_pixel_count: bad argument");
    return (int64_t)owned_pixel_count(width, height, shard_mode, rank, world);
}

int ptx_shard_owned_pixels(int width, int height, int shard_mode, int rank, int world, uint32_t* pixels_out, int64_t capacity) {
    int64_t n = ptx_shard_pixel_count(width, height, shard_mode, rank, world);
    if (n < 0) return (int)n;
    if (!pixels_out || capacity < n) return fail(nullptr, PTX_ERR_INVALID, "ptx_shard_owned_pixels: buffer too small");
    // same map as k_slot_pixels (kernels.cu)
    const uint32_t band_px = (uint32_t)kBandRows * (uint32_t)width;
    for (int64_t i = 0; i < n; i++) {
        if (shard_mode != PTX_SHARD_TILES || world <= 1) { pixels_out[i] = (uint32_t)i; continue; }
        uint32_t local_band = (uint32_t)i / band_px, within = (uint32_t)i % band_px;
        pixels_out[i] = (local_band * (uint32_t)world + (uint32_t)rank) * band_px + within;
    }
    return PTX_OK;
}

int ptx_shard_sample_range(int spp, int shard_mode, int rank, int world, int32_t* first, int32_t* count) {
    if (spp < 0 || world <= 0 || rank < 0 || rank >= world || !first || !count) return fail(nullptr, PTX_ERR_INVALID, "ptx_shard_sample_range: bad argument");
    if (shard_mode != PTX_SHARD_SAMPLES) { *first = 0; *count = spp; return PTX_OK; }
    int lo = (int)((long long)spp * rank / world), hi = (int)((long long)spp * (rank + 1) / world);
    *first = lo; *count = hi - lo;
    return PTX_OK;
}

int ptx_render_stats_get(const ptx_ctx* c, ptx_render_stats* out) {
    if (!c || !out) return fail(nullptr, PTX_ERR_INVALID, "ptx_render_stats_get: NULL argument");
    { int frc = finalize_pending(const_cast<ptx_ctx*>(c)); if (frc != PTX_OK) return frc; }
    *out = c->stats;
    return PTX_OK;
}

// ---------------------------------------------------------------------------------------------
// queries
// ---------------------------------------------------------------------------------------------
int ptx_trace_primary_ids(ptx_ctx* c, const ptx_scene* scene, const ptx_camera* cam, int width, int height, int32_t* prim_id_host,
                          float* t_host) {
    if (!c || !scene || !cam || width <= 0 || height <= 0 || !prim_id_host || !t_host) return fail(c, PTX_ERR_INVALID, "ptx_trace_primary_ids: bad argument");
    if (!scene->view.nodes) return fail(c, PTX_ERR_INVALID, "ptx_trace_primary_ids: the scene has no acceleration structure (a rebuild failed)");
    PTX_CUDA(c, cudaSetDevice(c->device));
    size_t n = (size_t)width * height;
    int32_t* d_ids = nullptr; float* d_t = nullptr;
    PTX_CUDA(c, cudaMalloc(&d_ids, n * 4));
    cudaError_t e = cudaMalloc(&d_t, n * 4);
    if (e != cudaSuccess) { cudaFree(d_ids); return cuda_fail(c, e, "cudaMalloc"); }
    launch_primary_ids(c->stream, scene->view, camera_view(*cam, width, height), width, height, d_ids, d_t);
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(prim_id_host, d_ids, n * 4, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(t_host, d_t, n * 4, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(d_ids); cudaFree(d_t);
    if (e != cudaSuccess) return cuda_fail(c, e, "ptx_trace_primary_ids");
    return PTX_OK;
}

int ptx_trace_rays(ptx_ctx* c, const ptx_scene* scene, const float* rays_host, size_t n, float t_min, int32_t* prim_id_host, float* t_host) {
    if (!c || !scene || (n && (!rays_host || !prim_id_host || !t_host))) return fail(c, PTX_ERR_INVALID, "ptx_trace_rays: bad argument");
    if (!scene->view.nodes) return fail(c, PTX_ERR_INVALID, "ptx_trace_rays: the scene has no acceleration structure (a rebuild failed)");
    if (!n) return PTX_OK;
    PTX_CUDA(c, cudaSetDevice(c->device));
    float* d_rays = nullptr; int32_t* d_ids = nullptr; float* d_t = nullptr;
    cudaError_t e = cudaMalloc(&d_rays, n * 24);
    if (e == cudaSuccess) e = cudaMalloc(&d_ids, n * 4);
    if (e == cudaSuccess) e = cudaMalloc(&d_t, n * 4);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_rays, rays_host, n * 24, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) { launch_trace_rays(c->stream, scene->view, d_rays, n, t_min, d_ids, d_t); e = cudaGetLastError(); }
    if (e == cudaSuccess) e = cudaMemcpyAsync(prim_id_host, d_ids, n * 4, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(t_host, d_t, n * 4, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(d_rays); cudaFree(d_ids); cudaFree(d_t);
    if (e != cudaSuccess) return cuda_fail(c, e, "ptx_trace_rays");
    return PTX_OK;
}

// No counterpart in the reference.  Same result as ptx_trace_rays, but no
// cudaMalloc / cudaFree (nor any other allocating call) happens here: the device staging is a file-scope array of kernels.cu.
int ptx_trace_rays_noalloc(ptx_ctx* c, const ptx_scene* scene, const float* rays_host, size_t n, float t_min, int32_t* prim_id_host, float* t_host) {
    if (!c || !scene || (n && (!rays_host || !prim_id_host || !t_host))) return fail(c, PTX_ERR_INVALID, "ptx_trace_rays_noalloc: bad argument");
    if (scene->ctx != c) return fail(c, PTX_ERR_INVALID, "ptx_trace_rays_noalloc: the scene belongs to another context");
    if (n > (size_t)PTX_NOALLOC_MAX_RAYS)
        return fail(c, PTX_ERR_INVALID, "ptx_trace_rays_noalloc: at most PTX_NOALLOC_MAX_RAYS (65536) rays per call; split the batch or use ptx_trace_rays");
    if (!scene->view.nodes) return fail(c, PTX_ERR_INVALID, "ptx_trace_rays_noalloc: the scene has no acceleration structure");
    if (!n) return PTX_OK;
    PTX_CUDA(c, cudaSetDevice(c->device));
    cudaError_t e = trace_rays_noalloc(c->stream, c->device, scene->view, rays_host, n, t_min, prim_id_host, t_host);
    if (e != cudaSuccess) return cuda_fail(c, e, "ptx_trace_rays_noalloc");
    return PTX_OK;
}

int ptx_ctx_flush_l2(ptx_ctx* c) {
    if (!c) return fail(nullptr, PTX_ERR_INVALID, "ptx_ctx_flush_l2: ctx is NULL");
    PTX_CUDA(c, cudaSetDevice(c->device));
    const size_t bytes = (size_t)256 << 20;                    // 2 x the 126 MB L2
    if (!c->d_flush) PTX_CUDA(c, cudaMalloc(&c->d_flush, bytes));
    c->flush_value ^= 0xFF;
    PTX_CUDA(c, cudaMemsetAsync(c->d_flush, c->flush_value, bytes, c->stream));
    return PTX_OK;
}

int ptx_trace_occluded(ptx_ctx* c, const ptx_scene* scene, const float* segs_host, size_t n, uint8_t* occluded_host) {
    if (!c || !scene || (n && (!segs_host || !occluded_host))) return fail(c, PTX_ERR_INVALID, "ptx_trace_occluded: bad argument");
    if (!scene->view.nodes) return fail(c, PTX_ERR_INVALID, "ptx_trace_occluded: the scene has no acceleration structure (a rebuild failed)");
    if (!n) return PTX_OK;
    PTX_CUDA(c, cudaSetDevice(c->device));
    float* d_segs = nullptr; uint8_t* d_out = nullptr;
    cudaError_t e = cudaMalloc(&d_segs, n * 24);
    if (e == cudaSuccess) e = cudaMalloc(&d_out, n);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_segs, segs_host, n * 24, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) { launch_occluded(c->stream, scene->view, d_segs, n, d_out); e = cudaGetLastError(); }
    if (e == cudaSuccess) e = cudaMemcpyAsync(occluded_host, d_out, n, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(d_segs); cudaFree(d_out);
    if (e != cudaSuccess) return cuda_fail(c, e, "ptx_trace_occluded");
    return PTX_OK;
}

int ptx_render_visibility(ptx_ctx* c, const ptx_scene* scene, const ptx_camera* cam, ptx_film* film, int spp, const float* radiosity_host,
                          uint8_t* rgb8_host) {
    if (!c || !scene || !cam || !film || !rgb8_host || spp <= 0) return fail(c, PTX_ERR_INVALID, "ptx_render_visibility: bad argument");
    if (scene->ctx != c || film->ctx != c) return fail(c, PTX_ERR_INVALID, "ptx_render_visibility: scene/film belong to another context");
    if (!scene->view.nodes) return fail(c, PTX_ERR_INVALID, "ptx_render_visibility: the scene has no acceleration structure (a rebuild failed)");
    PTX_CUDA(c, cudaSetDevice(c->device));
    const size_t npix = (size_t)film->w * film->h, nprim = scene->view.n_prims;
    float* d_rad = nullptr; uint8_t* d_img = nullptr;
    cudaError_t e = cudaMalloc(&d_img, npix * 3);
    if (e == cudaSuccess && radiosity_host && nprim) {
        e = cudaMalloc(&d_rad, nprim * 12);
        if (e == cudaSuccess) e = cudaMemcpyAsync(d_rad, radiosity_host, nprim * 12, cudaMemcpyHostToDevice, c->stream);
    }
    if (e == cudaSuccess) { launch_visibility(c->stream, scene->view, film->view(), camera_view(*cam, film->w, film->h), spp, d_rad, d_img); e = cudaGetLastError(); }
    if (e == cudaSuccess) e = cudaMemcpyAsync(rgb8_host, d_img, npix * 3, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(d_rad); cudaFree(d_img);
    if (e != cudaSuccess) return cuda_fail(c, e, "ptx_render_visibility");
    return PTX_OK;
}


// ---------------------------------------------------------------------------------------------
// multi-GPU film combine: one process per GPU, the scene replicated, every rank renders its share
// (ptx_render_opts.shard_*) and the films are combined over NCCL / NVLink INSIDE the library, on the context's stream.
// The reference has no multi-GPU path: nothing to cite.
// ---------------------------------------------------------------------------------------------
}  // extern "C"

namespace {
struct NcclApi {
    void* handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    std::string error;
};
NcclApi* nccl_api() {
    static NcclApi api;
    static bool tried = false;
    if (tried) return &api;
    tried = true;
    // a process that already carries NCCL (e.g. PyTorch's bundled copy) gets that very copy: same soname
    const char* names[] = {getenv("PTX_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
        if (!n || !*n) continue;
        api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (api.handle) break;
    }
    if (!api.handle) { api.error = std::string("NCCL not found (set PTX_NCCL_LIB): ") + (dlerror() ? dlerror() : ""); return &api; }
    bool ok = true;
    auto sym = [&](const char* name) { void* p = dlsym(api.handle, name); if (!p) { ok = false; api.error = std::string("NCCL symbol missing: ") + name; } return p; };
    api.GetUniqueId = (decltype(api.GetUniqueId))sym("ncclGetUniqueId");
    api.CommInitRank = (decltype(api.CommInitRank))sym("ncclCommInitRank");
    api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
    api.AllGather = (decltype(api.AllGather))sym("ncclAllGather");
    api.AllReduce = (decltype(api.AllReduce))sym("ncclAllReduce");
    api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
    if (!ok) { dlclose(api.handle); api.handle = nullptr; }
    return &api;
}
}  // namespace

struct ptx_comm {
    ptx_ctx* ctx = nullptr;
    ncclComm_t comm = nullptr;
    int rank = 0, world = 1;
    float4* d_gather = nullptr; size_t gather_capacity = 0;      // tiles: world x (pixels owned by rank 0) float4
};

extern "C" {

int ptx_comm_unique_id(uint8_t* id_out, size_t bytes) {
    if (!id_out || bytes < PTX_COMM_ID_BYTES) return fail(nullptr, PTX_ERR_INVALID, "ptx_comm_unique_id: the id needs PTX_COMM_ID_BYTES (128) bytes");
    static_assert(sizeof(ncclUniqueId) == PTX_COMM_ID_BYTES, "ncclUniqueId is 128 bytes");
    NcclApi* n = nccl_api();
    if (!n->handle) return fail(nullptr, PTX_ERR_UNSUPPORTED, n->error);
    ncclUniqueId id;
    ncclResult_t r = n->GetUniqueId(&id);
    if (r != ncclSuccess) return fail(nullptr, PTX_ERR_CUDA, std::string("ncclGetUniqueId: ") + n->GetErrorString(r));
    memcpy(id_out, &id, sizeof id);
    return PTX_OK;
}

int ptx_comm_create(ptx_ctx* c, const uint8_t* id, int rank, int world, ptx_comm** out) {
    if (!c || !id || !out || world <= 0 || rank < 0 || rank >= world) return fail(c, PTX_ERR_INVALID, "ptx_comm_create: bad argument");
    *out = nullptr;
    NcclApi* n = nccl_api();
    if (!n->handle) return fail(c, PTX_ERR_UNSUPPORTED, n->error);
    PTX_CUDA(c, cudaSetDevice(c->device));
    ncclUniqueId uid;
    memcpy(&uid, id, sizeof uid);
    ptx_comm* m = new ptx_comm();
    m->ctx = c; m->rank = rank; m->world = world;
    ncclResult_t r = n->CommInitRank(&m->comm, world, uid, rank);
    if (r != ncclSuccess) { delete m; return fail(c, PTX_ERR_CUDA, std::string("ncclCommInitRank: ") + n->GetErrorString(r)); }
    *out = m;
    return PTX_OK;
}

void ptx_comm_destroy(ptx_comm* m) {
    if (!m) return;
    cudaSetDevice(m->ctx->device);
    cudaStreamSynchronize(m->ctx->stream);
    if (m->comm) nccl_api()->CommDestroy(m->comm);
    cudaFree(m->d_gather);
    delete m;
}

int ptx_film_combine(ptx_film* f, ptx_comm* m, int shard_mode) {
    if (!f || !m) return fail(nullptr, PTX_ERR_INVALID, "ptx_film_combine: NULL argument");
    ptx_ctx* c = f->ctx;
    if (m->ctx != c) return fail(c, PTX_ERR_INVALID, "ptx_film_combine: film and communicator belong to different contexts");
    if (shard_mode != PTX_SHARD_TILES && shard_mode != PTX_SHARD_SAMPLES) return fail(c, PTX_ERR_INVALID, "ptx_film_combine: shard_mode must be PTX_SHARD_TILES or PTX_SHARD_SAMPLES");
    NcclApi* n = nccl_api();
    PTX_CUDA(c, cudaSetDevice(c->device));
    const int w = f->w, h = f->h;
    FilmView fv = f->view();
    ncclResult_t r = ncclSuccess;
    if (shard_mode == PTX_SHARD_TILES) {
        // every rank sends only what it owns (1 / G of the film, no zeros); rank 0 owns the most pixels, its count pads the slices
        const uint32_t per_rank = owned_pixel_count(w, h, 1, 0, m->world);
        const uint32_t mine = owned_pixel_count(w, h, 1, m->rank, m->world);
        const size_t need = (size_t)per_rank * m->world;
        if (need > m->gather_capacity) {
            PTX_CUDA(c, cudaStreamSynchronize(c->stream));
            cudaFree(m->d_gather); m->d_gather = nullptr; m->gather_capacity = 0;
            PTX_CUDA(c, cudaMalloc(&m->d_gather, need * sizeof(float4)));
            m->gather_capacity = need;
        }
        int rc = ensure_pixmap(c, mine, 0u);
        if (rc != PTX_OK) return rc;
        int key[5] = {w, h, PTX_SHARD_TILES, m->rank, m->world};
        if (memcmp(key, c->slot_key, sizeof key) != 0) {
            launch_slot_pixels(c->stream, c->d_slot_base, w, h, PTX_SHARD_TILES, m->rank, m->world, mine);
            memcpy(c->slot_key, key, sizeof key);
        }
        float4* my_slice = m->d_gather + (size_t)m->rank * per_rank;
        launch_film_pack_owned(c->stream, fv, c->d_slot_base, mine, my_slice);       // canonical slot order (the render may have used the cost order)
        PTX_CUDA(c, cudaGetLastError());
        r = n->AllGather(my_slice, m->d_gather, (size_t)per_rank * 4, ncclFloat32, m->comm, c->stream);     // in place
        if (r != ncclSuccess) return fail(c, PTX_ERR_CUDA, std::string("ncclAllGather: ") + n->GetErrorString(r));
        launch_film_unpack_all(c->stream, fv, m->d_gather, per_rank, m->world);
        PTX_CUDA(c, cudaGetLastError());
    } else {
        // sample ranges: every rank holds partial sums of every pixel
        r = n->AllReduce(f->d_accum, f->d_accum, (size_t)w * h * 4, ncclFloat32, ncclSum, m->comm, c->stream);
        if (r != ncclSuccess) return fail(c, PTX_ERR_CUDA, std::string("ncclAllReduce: ") + n->GetErrorString(r));
        f->total_spp = f->frame_spp;          // what the ranks were asked for together (ptx_render keeps count)
    }
    return PTX_OK;
}

}  // extern "C"

// oracle/ref_harness.cu -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// Headless driver around the UNMODIFIED reference headers under /root/reference/include
// (they are #included where they lie; nothing is copied).  It is built twice by oracle/Makefile:
//
//   oracle/_ref/ref_host   -DREF_HOST : every __device__ function of the reference is also compiled
//                                       for the host (macro technique of SURVEY.md Appendix A.3), so
//                                       Scene::intersect / integrator() / curand_* run on CPU cores.
//                                       This is the "host-compiled scalar build of the reference
//                                       integrator" (BASELINE.json) and the golden-vector generator.
//   oracle/_ref/ref_cuda              : the reference's own kernels on the GPU (render_init + render,
//                                       include/rendering/integrator.h:274-280, 371-408) plus a twin of
//                                       `render` whose only differences are max_depth as a parameter
//                                       (the original hard-codes 5, integrator.h:389) and an extra
//                                       pre-tonemap float output.  This is the R-GPU baseline.
//
// What it re-states (because the originals need OpenGL/GLFW, unavailable on this box):
//   RenderState::allocateBuffers  include/application_state.h:91-123   (camera aspect + render_init)
//   renderFrame                   include/application.h:157-216        (updateCameraOrbit + render)
//   SceneState::loadScene         include/application_state.h:367-464  (loadOBJ + BVHBuilder + Scene)
//
// Sub-commands (all outputs raw little-endian binaries, layout documented per command):
//   render  <scene.obj> <w> <h> <spp> <depth> <frames> <out_prefix> [convert_quads] [threads] [orig]
//   ids     <scene.obj> <w> <h> <out_prefix> [convert_quads]
//   rays    <scene.obj> <rays.bin> <n> <out_prefix> [convert_quads]
//   rng     <seed> <subsequence> <count>
//   dump    <scene.obj> <out_prefix> [convert_quads]

#ifdef REF_HOST
#define QUALIFIERS static __forceinline__ __host__ __device__
#endif
#include <cuda_runtime.h>
#include <curand_kernel.h>
#include <cfloat>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iomanip>
#include <sstream>
#include <string>
#include <vector>
#ifdef REF_HOST
#undef __device__
#define __device__ __location__(host) __location__(device)
#define clock64() 0LL
#define atomicAdd(p, v) (*(p) += (v))
#ifdef _OPENMP
#include <omp.h>
#endif
#endif

#include "core/vector.h"
#include "rendering/sensor.h"
#include "rendering/primitive.h"
#include "rendering/scene.h"
#include "rendering/bvh.h"
#include "rendering/render_config.h"
#include "utils/file_manager.h"
#include "rendering/integrator.h"
#ifdef REF_HOST
#include "rendering/form_factors.h"      // subdivide_primitives (form_factors.h:475-574), host only
#endif

// ---------------------------------------------------------------------------------------------
// helpers
// ---------------------------------------------------------------------------------------------
static void write_file(const std::string& path, const void* data, size_t bytes) {
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) { fprintf(stderr, "cannot write %s\n", path.c_str()); exit(2); }
    fwrite(data, 1, bytes, f);
    fclose(f);
}

struct HostScene {
    std::vector<Primitive> prims;
    std::vector<BVHNode> nodes;
    std::vector<int> indices;
};

// quads -> 2 triangles exactly like convertQuadsToTriangles (application_state.h:323-365):
// (v00,v10,v11) and (v00,v11,v01) with GEOMETRIC normals recomputed by the 4-arg Triangle ctor.
static void split_quads(std::vector<Primitive>& prims) {
    std::vector<Primitive> out;
    for (auto& p : prims) {
        if (p.type == PRIM_QUAD) {
            const Quad& q = p.quad;
            Primitive a; a.type = PRIM_TRIANGLE; new (&a.tri) Triangle(q.v00, q.v10, q.v11, q.bsdf); a.tri.Le = q.Le;
            Primitive b; b.type = PRIM_TRIANGLE; new (&b.tri) Triangle(q.v00, q.v11, q.v01, q.bsdf); b.tri.Le = q.Le;
            out.push_back(a); out.push_back(b);
        } else out.push_back(p);
    }
    prims.swap(out);
}

static int g_subdivision = 0;      // `dump ... [convert_quads] [subdivision]`: loadScene order, application_state.h:411-434
static bool load_scene(const char* path, bool convert_quads, HostScene& hs) {
    Primitive* list = nullptr; int n = 0;
    std::streambuf* old = std::cout.rdbuf(nullptr);   // silence the reference's banners
    bool ok;
#if defined(USE_PBRT_LOADER) && !defined(__CUDA_ARCH__)   // the reference hides loadPBRT from the device pass (file_manager.h:280)
    std::string sp(path);
    if (sp.size() > 5 && sp.substr(sp.size() - 5) == ".pbrt") ok = loadPBRT(sp, &list, n);      // utils/pbrt_loader.h:178
    else
#endif
        ok = loadOBJ(path, &list, n);
    if (ok) {
        hs.prims.assign(list, list + n);
        delete[] list;
        if (convert_quads) split_quads(hs.prims);
#ifdef REF_HOST
        if (g_subdivision > 0) hs.prims = subdivide_primitives(hs.prims, g_subdivision);
#endif
        BVHBuilder b(hs.prims.data(), (int)hs.prims.size());
        hs.nodes = b.nodes; hs.indices = b.primitive_indices;
    }
    std::cout.rdbuf(old);
    return ok;
}

static Sensor make_camera(int w, int h) {
    // RenderState ctor (application_state.h:88) + allocateBuffers (:106-109) + renderFrame's
    // updateCameraOrbit (application.h:162).
    Sensor cam(Vector3f(0.5, 3, 8.5), Vector3f(0, 2.5, 0), Vector3f(0, 1, 0), 40.0f, 1.0f);
    cam.image_width = w; cam.image_height = h;
    cam.aspect = (float)w / (float)h;
    cam.updateCamera();
    cam.updateCameraOrbit();
    return cam;
}

// One pixel of `render` (integrator.h:371-408) with max_depth lifted to a parameter and the
// pre-tonemap mean also returned.  Under REF_HOST this is host code; otherwise device code.
__device__ void pixel_body(int x, int y, Sensor* cam, Scene* scene, curandState* rs, int spp,
                           int max_depth, float* linear, unsigned char* image) {
    int pixel_index = y * cam->image_width + x;
    curandState* local_rng = &rs[pixel_index];
    Vector3f color(0.0f, 0.0f, 0.0f);
    for (int s = 0; s < spp; s++) {
        float u = (x + curand_uniform(local_rng)) / float(cam->image_width);
        float v = (y + curand_uniform(local_rng)) / float(cam->image_height);
        Ray ray = cam->get_ray(u, v);
        Vector3f sample_color(0.0f, 0.0f, 0.0f);
        integrator(scene, ray, sample_color, max_depth, local_rng, SamplingMode::SAMPLING_BSDF);
        color += sample_color;
    }
    color /= float(spp);
    if (linear) { linear[pixel_index * 3 + 0] = color.r(); linear[pixel_index * 3 + 1] = color.g(); linear[pixel_index * 3 + 2] = color.b(); }
    color = color / (color + Vector3f(1.0f, 1.0f, 1.0f));
    const float gamma = 1.0f / 2.2f;
    color.e[0] = powf(color.e[0], gamma);
    color.e[1] = powf(color.e[1], gamma);
    color.e[2] = powf(color.e[2], gamma);
    int idx = pixel_index * 3;
    image[idx + 0] = (unsigned char)(255.99f * fminf(color.r(), 1.0f));
    image[idx + 1] = (unsigned char)(255.99f * fminf(color.g(), 1.0f));
    image[idx + 2] = (unsigned char)(255.99f * fminf(color.b(), 1.0f));
}

__device__ void primary_id_body(int x, int y, int w, int h, Sensor* cam, Scene* scene, int* ids, float* ts) {
    float u = (x + 0.5f) / float(w);
    float v = (y + 0.5f) / float(h);
    Ray ray = cam->get_ray(u, v);
    SurfaceInteractionRecord si;
    bool hit = scene->intersect(ray, 1e-4f, FLT_MAX, si);
    ids[y * w + x] = hit ? (int)(si.prim_ptr - scene->primitives) : -1;
    ts[y * w + x] = hit ? si.t : -1.0f;
}

#ifndef REF_HOST
__global__ void render_twin(float* linear, unsigned char* image, Sensor* cam, Scene* scene,
                            curandState* rs, int spp, int max_depth) {
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= cam->image_width || y >= cam->image_height) return;
    pixel_body(x, y, cam, scene, rs, spp, max_depth, linear, image);
}
__global__ void primary_ids_kernel(int w, int h, Sensor* cam, Scene* scene, int* ids, float* ts) {
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= w || y >= h) return;
    primary_id_body(x, y, w, h, cam, scene, ids, ts);
}
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "CUDA %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(3); } } while (0)

struct DeviceScene {
    Primitive* d_prims = nullptr; BVHNode* d_nodes = nullptr; int* d_idx = nullptr; Scene* d_scene = nullptr;
    void upload(HostScene& hs) {
        CK(cudaMalloc(&d_prims, hs.prims.size() * sizeof(Primitive)));
        CK(cudaMemcpy(d_prims, hs.prims.data(), hs.prims.size() * sizeof(Primitive), cudaMemcpyHostToDevice));
        CK(cudaMalloc(&d_nodes, hs.nodes.size() * sizeof(BVHNode)));
        CK(cudaMemcpy(d_nodes, hs.nodes.data(), hs.nodes.size() * sizeof(BVHNode), cudaMemcpyHostToDevice));
        CK(cudaMalloc(&d_idx, hs.indices.size() * sizeof(int)));
        CK(cudaMemcpy(d_idx, hs.indices.data(), hs.indices.size() * sizeof(int), cudaMemcpyHostToDevice));
        Scene h(d_prims, (int)hs.prims.size(), d_nodes, d_idx);
        CK(cudaMalloc(&d_scene, sizeof(Scene)));
        CK(cudaMemcpy(d_scene, &h, sizeof(Scene), cudaMemcpyHostToDevice));
    }
};
#endif

// ---------------------------------------------------------------------------------------------
// sub-commands
// ---------------------------------------------------------------------------------------------
static int cmd_render(int argc, char** argv) {
    if (argc < 9) { fprintf(stderr, "render <scene> <w> <h> <spp> <depth> <frames> <out_prefix> [convert_quads] [threads] [orig]\n"); return 1; }
    const char* path = argv[2];
    int w = atoi(argv[3]), h = atoi(argv[4]), spp = atoi(argv[5]), depth = atoi(argv[6]), frames = atoi(argv[7]);
    std::string prefix = argv[8];
    bool cq = argc > 9 && atoi(argv[9]) != 0;
    int threads = argc > 10 ? atoi(argv[10]) : 0;
    bool orig = argc > 11 && atoi(argv[11]) != 0;   // device only: launch the reference's own `render` (depth 5)
    HostScene hs;
    if (!load_scene(path, cq, hs)) return 2;
    Sensor cam = make_camera(w, h);
    size_t npix = (size_t)w * h;
    std::vector<float> linear(npix * 3);
    std::vector<unsigned char> image(npix * 3);
    double ms_total = 0.0, ms_init = 0.0;
#ifdef REF_HOST
    (void)orig;
    Scene scene(hs.prims.data(), (int)hs.prims.size(), hs.nodes.data(), hs.indices.data());
    std::vector<curandState> rs(npix);
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
    threads = omp_get_max_threads();
#else
    threads = 1;
#endif
    auto t0 = std::chrono::high_resolution_clock::now();
    #pragma omp parallel for schedule(dynamic, 1)
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) { int p = y * w + x; curand_init(2023 + p, p, 0, &rs[p]); }   // render_init, integrator.h:274-280
    auto t1 = std::chrono::high_resolution_clock::now();
    ms_init = std::chrono::duration<double, std::milli>(t1 - t0).count();
    for (int f = 0; f < frames; f++) {
        auto a = std::chrono::high_resolution_clock::now();
        #pragma omp parallel for schedule(dynamic, 1)
        for (int y = 0; y < h; y++)
            for (int x = 0; x < w; x++)
                pixel_body(x, y, &cam, &scene, rs.data(), spp, depth, linear.data(), image.data());
        auto b = std::chrono::high_resolution_clock::now();
        ms_total += std::chrono::duration<double, std::milli>(b - a).count();
    }
#else
    DeviceScene ds; ds.upload(hs);
    Sensor* d_cam; CK(cudaMalloc(&d_cam, sizeof(Sensor))); CK(cudaMemcpy(d_cam, &cam, sizeof(Sensor), cudaMemcpyHostToDevice));
    curandState* d_rs; CK(cudaMalloc(&d_rs, npix * sizeof(curandState)));
    float* d_lin; CK(cudaMalloc(&d_lin, npix * 3 * sizeof(float)));
    unsigned char* d_img; CK(cudaMalloc(&d_img, npix * 3));
    dim3 block(BLOCK_X, BLOCK_Y), grid((w + block.x - 1) / block.x, (h + block.y - 1) / block.y);
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    CK(cudaEventRecord(e0));
    render_init<<<grid, block>>>(w, h, d_rs);
    CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); ms_init = ms;
    for (int f = 0; f < frames; f++) {
        CK(cudaEventRecord(e0));
        if (orig) render<<<grid, block>>>(d_img, d_cam, ds.d_scene, d_rs, spp, SamplingMode::SAMPLING_BSDF);
        else render_twin<<<grid, block>>>(d_lin, d_img, d_cam, ds.d_scene, d_rs, spp, depth);
        CK(cudaEventRecord(e1)); CK(cudaDeviceSynchronize());
        CK(cudaGetLastError());
        CK(cudaEventElapsedTime(&ms, e0, e1)); ms_total += ms;
    }
    CK(cudaMemcpy(linear.data(), d_lin, npix * 3 * sizeof(float), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(image.data(), d_img, npix * 3, cudaMemcpyDeviceToHost));
    threads = 0;
#endif
    if (prefix != "-") {
        write_file(prefix + ".linear.f32", linear.data(), linear.size() * sizeof(float));   // [h][w][3], row 0 = bottom
        write_file(prefix + ".rgb8.u8", image.data(), image.size());                         // [h][w][3], row 0 = bottom
    }
    double mean[3] = {0, 0, 0};
    for (size_t i = 0; i < npix; i++) for (int c = 0; c < 3; c++) mean[c] += linear[i * 3 + c];
    printf("{\"cmd\":\"render\",\"impl\":\"%s\",\"kernel\":\"%s\",\"w\":%d,\"h\":%d,\"spp\":%d,\"depth\":%d,\"frames\":%d,"
           "\"prims\":%zu,\"nodes\":%zu,\"threads\":%d,\"ms_init\":%.4f,\"ms_per_frame\":%.4f,\"mean\":[%.9g,%.9g,%.9g]}\n",
#ifdef REF_HOST
           "ref_host", "integrator",
#else
           "ref_cuda", orig ? "render" : "render_twin",
#endif
           w, h, spp, orig ? 5 : depth, frames, hs.prims.size(), hs.nodes.size(), threads, ms_init,
           ms_total / (frames > 0 ? frames : 1), mean[0] / npix, mean[1] / npix, mean[2] / npix);
    return 0;
}

static int cmd_ids(int argc, char** argv) {
    if (argc < 6) { fprintf(stderr, "ids <scene> <w> <h> <out_prefix> [convert_quads]\n"); return 1; }
    int w = atoi(argv[3]), h = atoi(argv[4]);
    std::string prefix = argv[5];
    bool cq = argc > 6 && atoi(argv[6]) != 0;
    HostScene hs;
    if (!load_scene(argv[2], cq, hs)) return 2;
    Sensor cam = make_camera(w, h);
    size_t npix = (size_t)w * h;
    std::vector<int> ids(npix); std::vector<float> ts(npix);
#ifdef REF_HOST
    Scene scene(hs.prims.data(), (int)hs.prims.size(), hs.nodes.data(), hs.indices.data());
    for (int y = 0; y < h; y++) for (int x = 0; x < w; x++) primary_id_body(x, y, w, h, &cam, &scene, ids.data(), ts.data());
#else
    DeviceScene ds; ds.upload(hs);
    Sensor* d_cam; CK(cudaMalloc(&d_cam, sizeof(Sensor))); CK(cudaMemcpy(d_cam, &cam, sizeof(Sensor), cudaMemcpyHostToDevice));
    int* d_ids; float* d_ts; CK(cudaMalloc(&d_ids, npix * 4)); CK(cudaMalloc(&d_ts, npix * 4));
    dim3 block(16, 16), grid((w + 15) / 16, (h + 15) / 16);
    primary_ids_kernel<<<grid, block>>>(w, h, d_cam, ds.d_scene, d_ids, d_ts);
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(ids.data(), d_ids, npix * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(ts.data(), d_ts, npix * 4, cudaMemcpyDeviceToHost));
#endif
    write_file(prefix + ".ids.i32", ids.data(), npix * 4);   // [h][w], -1 = miss, row 0 = bottom
    write_file(prefix + ".t.f32", ts.data(), npix * 4);
    long long sum = 0, hits = 0;
    for (size_t i = 0; i < npix; i++) if (ids[i] >= 0) { sum += ids[i]; hits++; }
    printf("{\"cmd\":\"ids\",\"w\":%d,\"h\":%d,\"hits\":%lld,\"id_sum\":%lld}\n", w, h, hits, sum);
    return 0;
}

#ifdef REF_HOST
// rays.bin: n records of 6 floats (o.xyz, d.xyz); d is normalised by the Ray ctor as in the reference.
// Outputs: <prefix>.bvh.i32 / .bvh_t.f32 (Scene::intersect, BVH path) and <prefix>.lin.i32 / .lin_t.f32
// (Scene::intersect_linear, scene.h:113-129), t_min = 1e-4, t_max = FLT_MAX as in integrator.h:199.
static int cmd_rays(int argc, char** argv) {
    if (argc < 6) { fprintf(stderr, "rays <scene> <rays.bin> <n> <out_prefix> [convert_quads]\n"); return 1; }
    size_t n = (size_t)atoll(argv[4]);
    std::string prefix = argv[5];
    bool cq = argc > 6 && atoi(argv[6]) != 0;
    HostScene hs;
    if (!load_scene(argv[2], cq, hs)) return 2;
    std::vector<float> rays(n * 6);
    FILE* f = fopen(argv[3], "rb");
    if (!f || fread(rays.data(), sizeof(float), n * 6, f) != n * 6) { fprintf(stderr, "bad rays file\n"); return 2; }
    fclose(f);
    Scene scene(hs.prims.data(), (int)hs.prims.size(), hs.nodes.data(), hs.indices.data());
    std::vector<int> a(n), b(n); std::vector<float> ta(n), tb(n);
    for (size_t i = 0; i < n; i++) {
        Ray r(Vector3f(rays[i * 6], rays[i * 6 + 1], rays[i * 6 + 2]), Vector3f(rays[i * 6 + 3], rays[i * 6 + 4], rays[i * 6 + 5]));
        SurfaceInteractionRecord s1, s2;
        bool h1 = scene.intersect(r, 1e-4f, FLT_MAX, s1);
        bool h2 = scene.intersect_linear(r, 1e-4f, FLT_MAX, s2);
        a[i] = h1 ? (int)(s1.prim_ptr - scene.primitives) : -1; ta[i] = h1 ? s1.t : -1.0f;
        b[i] = h2 ? (int)(s2.prim_ptr - scene.primitives) : -1; tb[i] = h2 ? s2.t : -1.0f;
    }
    write_file(prefix + ".bvh.i32", a.data(), n * 4); write_file(prefix + ".bvh_t.f32", ta.data(), n * 4);
    write_file(prefix + ".lin.i32", b.data(), n * 4); write_file(prefix + ".lin_t.f32", tb.data(), n * 4);
    size_t diff = 0; for (size_t i = 0; i < n; i++) diff += a[i] != b[i];
    printf("{\"cmd\":\"rays\",\"n\":%zu,\"bvh_vs_linear_mismatch\":%zu}\n", n, diff);
    return 0;
}

static int cmd_rng(int argc, char** argv) {
    if (argc < 5) { fprintf(stderr, "rng <seed> <subsequence> <count>\n"); return 1; }
    unsigned long long seed = strtoull(argv[2], 0, 10), sub = strtoull(argv[3], 0, 10);
    int count = atoi(argv[4]);
    curandState st; curand_init(seed, sub, 0, &st);
    printf("{\"seed\":%llu,\"subsequence\":%llu,\"v\":[%u,%u,%u,%u,%u],\"d\":%u,\"uniform\":[", seed, sub, st.v[0], st.v[1], st.v[2], st.v[3], st.v[4], st.d);
    for (int i = 0; i < count; i++) printf("%s%.9g", i ? "," : "", curand_uniform(&st));
    printf("]}\n");
    return 0;
}

// dump: per primitive 1 int type (0 tri, 1 quad) + 12 floats vertices (tri: v0 v1 v2 + 3 zeros)
// + 3 normal + 3 bsdf + 3 Le  => 22 x 4 bytes; plus the BVH (nodes: 6 floats bbox + 3 ints; indices).
static int cmd_dump(int argc, char** argv) {
    if (argc < 4) { fprintf(stderr, "dump <scene> <out_prefix> [convert_quads]\n"); return 1; }
    std::string prefix = argv[3];
    bool cq = argc > 4 && atoi(argv[4]) != 0;
    g_subdivision = argc > 5 ? atoi(argv[5]) : 0;
    HostScene hs;
    if (!load_scene(argv[2], cq, hs)) return 2;
    std::vector<float> rec;
    for (auto& p : hs.prims) {
        float r[22] = {0};
        int ty = p.type == PRIM_TRIANGLE ? 0 : 1; memcpy(&r[0], &ty, 4);
        if (ty == 0) {
            const Triangle& t = p.tri;
            for (int k = 0; k < 3; k++) { r[1 + k] = t.v0[k]; r[4 + k] = t.v1[k]; r[7 + k] = t.v2[k]; r[13 + k] = t.normal[k]; r[16 + k] = t.bsdf[k]; r[19 + k] = t.Le[k]; }
        } else {
            const Quad& q = p.quad;
            for (int k = 0; k < 3; k++) { r[1 + k] = q.v00[k]; r[4 + k] = q.v10[k]; r[7 + k] = q.v11[k]; r[10 + k] = q.v01[k]; r[13 + k] = q.normal[k]; r[16 + k] = q.bsdf[k]; r[19 + k] = q.Le[k]; }
        }
        rec.insert(rec.end(), r, r + 22);
    }
    write_file(prefix + ".prims.bin", rec.data(), rec.size() * 4);
    std::vector<float> nb;
    for (auto& nd : hs.nodes) {
        float r[9]; for (int k = 0; k < 3; k++) { r[k] = nd.bbox.min[k]; r[3 + k] = nd.bbox.max[k]; }
        memcpy(&r[6], &nd.left_child, 4); memcpy(&r[7], &nd.right_child, 4); memcpy(&r[8], &nd.prim_count, 4);
        nb.insert(nb.end(), r, r + 9);
    }
    write_file(prefix + ".nodes.bin", nb.data(), nb.size() * 4);
    write_file(prefix + ".indices.i32", hs.indices.data(), hs.indices.size() * 4);
    printf("{\"cmd\":\"dump\",\"prims\":%zu,\"nodes\":%zu}\n", hs.prims.size(), hs.nodes.size());
    return 0;
}

// segs.bin: n records of 6 floats (from.xyz, to.xyz).  Output <prefix>.occ.u8: Scene::occluded(from, to) (scene.h:132-149,
// BVH any-hit scene.h:152-190) and <prefix>.occ_lin.u8: the same query by linear scan over all primitives.
static int cmd_occluded(int argc, char** argv) {
    if (argc < 6) { fprintf(stderr, "occluded <scene> <segs.bin> <n> <out_prefix> [convert_quads]\n"); return 1; }
    size_t n = (size_t)atoll(argv[4]);
    std::string prefix = argv[5];
    bool cq = argc > 6 && atoi(argv[6]) != 0;
    HostScene hs;
    if (!load_scene(argv[2], cq, hs)) return 2;
    std::vector<float> segs(n * 6);
    FILE* f = fopen(argv[3], "rb");
    if (!f || fread(segs.data(), sizeof(float), n * 6, f) != n * 6) { fprintf(stderr, "bad segs file\n"); return 2; }
    fclose(f);
    Scene scene(hs.prims.data(), (int)hs.prims.size(), hs.nodes.data(), hs.indices.data());
    Scene flat(hs.prims.data(), (int)hs.prims.size());              // no BVH: occluded() takes its linear branch
    std::vector<unsigned char> a(n), b(n);
    size_t diff = 0, occ = 0;
    for (size_t i = 0; i < n; i++) {
        Vector3f from(segs[i * 6], segs[i * 6 + 1], segs[i * 6 + 2]), to(segs[i * 6 + 3], segs[i * 6 + 4], segs[i * 6 + 5]);
        a[i] = scene.occluded(from, to) ? 1 : 0;
        b[i] = flat.occluded(from, to) ? 1 : 0;
        diff += a[i] != b[i]; occ += a[i];
    }
    write_file(prefix + ".occ.u8", a.data(), n); write_file(prefix + ".occ_lin.u8", b.data(), n);
    printf("{\"cmd\":\"occluded\",\"n\":%zu,\"occluded\":%zu,\"bvh_vs_linear_mismatch\":%zu}\n", n, occ, diff);
    return 0;
}

// One pixel of render_radiosity (integrator.h:460-504) on the host: primary visibility, Le + per-primitive radiosity,
// sqrt "gamma".  Radiosity comes from the (out-of-scope) solver; here primitive i gets rad[3*i..] from a file.
static void radiosity_pixel(int x, int y, Sensor* cam, Scene* scene, curandState* rs, int spp, unsigned char* image) {
    int pixel_index = y * cam->image_width + x;
    curandState local_rng = rs[pixel_index];
    Vector3f color(0.0f, 0.0f, 0.0f);
    for (int s = 0; s < spp; s++) {
        float u = (x + curand_uniform(&local_rng)) / float(cam->image_width);
        float v = (y + curand_uniform(&local_rng)) / float(cam->image_height);
        Ray ray = cam->get_ray(u, v);
        SurfaceInteractionRecord si;
        if (scene->intersect(ray, 1e-4f, FLT_MAX, si)) {
            color += si.Le;
            if (si.prim_ptr != nullptr) color += si.prim_ptr->getRadiosity();
            else color += si.radiosity;
        }
    }
    color /= float(spp);
    color = Vector3f(sqrtf(fminf(color.x(), 1.0f)), sqrtf(fminf(color.y(), 1.0f)), sqrtf(fminf(color.z(), 1.0f)));
    int idx = 3 * pixel_index;
    image[idx + 0] = (unsigned char)(255.99f * color.x());
    image[idx + 1] = (unsigned char)(255.99f * color.y());
    image[idx + 2] = (unsigned char)(255.99f * color.z());
    rs[pixel_index] = local_rng;
}

static int cmd_visibility(int argc, char** argv) {
    if (argc < 8) { fprintf(stderr, "visibility <scene> <w> <h> <spp> <rad.bin|-> <out_prefix> [convert_quads]\n"); return 1; }
    int w = atoi(argv[3]), h = atoi(argv[4]), spp = atoi(argv[5]);
    std::string radf = argv[6], prefix = argv[7];
    bool cq = argc > 8 && atoi(argv[8]) != 0;
    HostScene hs;
    if (!load_scene(argv[2], cq, hs)) return 2;
    if (radf != "-") {
        std::vector<float> rad(hs.prims.size() * 3);
        FILE* f = fopen(radf.c_str(), "rb");
        if (!f || fread(rad.data(), sizeof(float), rad.size(), f) != rad.size()) { fprintf(stderr, "bad radiosity file\n"); return 2; }
        fclose(f);
        for (size_t i = 0; i < hs.prims.size(); i++) hs.prims[i].setRadiosity(Vector3f(rad[3 * i], rad[3 * i + 1], rad[3 * i + 2]));
    }
    Sensor cam = make_camera(w, h);
    size_t npix = (size_t)w * h;
    Scene scene(hs.prims.data(), (int)hs.prims.size(), hs.nodes.data(), hs.indices.data());
    std::vector<curandState> rs(npix);
    for (size_t p = 0; p < npix; p++) curand_init(2023 + p, p, 0, &rs[p]);
    std::vector<unsigned char> image(npix * 3);
    for (int y = 0; y < h; y++) for (int x = 0; x < w; x++) radiosity_pixel(x, y, &cam, &scene, rs.data(), spp, image.data());
    write_file(prefix + ".vis.u8", image.data(), image.size());
    std::vector<unsigned> st(npix * 6);
    for (size_t p = 0; p < npix; p++) { for (int k = 0; k < 5; k++) st[p * 6 + k] = rs[p].v[k]; st[p * 6 + 5] = rs[p].d; }
    write_file(prefix + ".rng.u32", st.data(), st.size() * 4);
    printf("{\"cmd\":\"visibility\",\"w\":%d,\"h\":%d,\"spp\":%d}\n", w, h, spp);
    return 0;
}

// Grid-guided / MIS sampling (integrator.h:226-262, grid.h, sampleMIS integrator.h:112-166) on the host.  The per-primitive
// 16x16 radiosity grids come from the (out-of-scope) radiosity solver; here they are read from a file
// ([n_prims][256][3] floats).  scene.precomputed_cdfs stays NULL, so the reference builds each CDF with its OWN code
// (initGridFromPrimitive -> Grid::initFromRadiosity -> buildCDFs, grid.h:62-137), the same arithmetic SceneState::
// precomputeCDFs (application_state.h:492-585, needs GL headers) performs once per solve.
static int cmd_guided(int argc, char** argv) {
    if (argc < 11) { fprintf(stderr, "guided <scene> <w> <h> <spp> <depth> <mode 0..4> <bsdf_fraction> <grids.f32> <out_prefix> [convert_quads]\n"); return 1; }
    int w = atoi(argv[3]), h = atoi(argv[4]), spp = atoi(argv[5]), depth = atoi(argv[6]), mode = atoi(argv[7]);
    float frac = (float)atof(argv[8]);
    std::string gridf = argv[9], prefix = argv[10];
    bool cq = argc > 11 && atoi(argv[11]) != 0;
    HostScene hs;
    if (!load_scene(argv[2], cq, hs)) return 2;
    {
        std::vector<float> g(hs.prims.size() * GRID_SIZE * 3);
        FILE* f = fopen(gridf.c_str(), "rb");
        if (!f || fread(g.data(), sizeof(float), g.size(), f) != g.size()) { fprintf(stderr, "bad grid file\n"); return 2; }
        fclose(f);
        for (size_t i = 0; i < hs.prims.size(); i++) {
            Vector3f* dst = hs.prims[i].type == PRIM_TRIANGLE ? hs.prims[i].tri.radiosity_grid : hs.prims[i].quad.radiosity_grid;
            for (int c = 0; c < GRID_SIZE; c++) dst[c] = Vector3f(g[(i * GRID_SIZE + c) * 3], g[(i * GRID_SIZE + c) * 3 + 1], g[(i * GRID_SIZE + c) * 3 + 2]);
        }
    }
    Sensor cam = make_camera(w, h);
    size_t npix = (size_t)w * h;
    Scene scene(hs.prims.data(), (int)hs.prims.size(), hs.nodes.data(), hs.indices.data());
    scene.mis_bsdf_fraction = frac;
    std::vector<curandState> rs(npix);
    for (size_t p = 0; p < npix; p++) curand_init(2023 + p, p, 0, &rs[p]);
    std::vector<float> linear(npix * 3);
    #pragma omp parallel for schedule(dynamic, 1)
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {
            int pixel_index = y * w + x;
            curandState* local_rng = &rs[pixel_index];
            Vector3f color(0.0f, 0.0f, 0.0f);
            for (int s = 0; s < spp; s++) {                                   // the sample loop of render, integrator.h:382-392
                float u = (x + curand_uniform(local_rng)) / float(w);
                float v = (y + curand_uniform(local_rng)) / float(h);
                Ray ray = cam.get_ray(u, v);
                Vector3f sample_color(0.0f, 0.0f, 0.0f);
                integrator(&scene, ray, sample_color, depth, local_rng, (SamplingMode)mode);
                color += sample_color;
            }
            color /= float(spp);
            linear[pixel_index * 3 + 0] = color.r(); linear[pixel_index * 3 + 1] = color.g(); linear[pixel_index * 3 + 2] = color.b();
        }
    write_file(prefix + ".linear.f32", linear.data(), linear.size() * sizeof(float));
    std::vector<unsigned> st(npix * 6);
    for (size_t p = 0; p < npix; p++) { for (int k = 0; k < 5; k++) st[p * 6 + k] = rs[p].v[k]; st[p * 6 + 5] = rs[p].d; }
    write_file(prefix + ".rng.u32", st.data(), st.size() * 4);
    double mean[3] = {0, 0, 0};
    for (size_t i = 0; i < npix; i++) for (int c = 0; c < 3; c++) mean[c] += linear[i * 3 + c];
    printf("{\"cmd\":\"guided\",\"w\":%d,\"h\":%d,\"spp\":%d,\"depth\":%d,\"mode\":%d,\"mean\":[%.9g,%.9g,%.9g]}\n", w, h, spp, depth, mode,
           mean[0] / npix, mean[1] / npix, mean[2] / npix);
    return 0;
}
#endif

int main(int argc, char** argv) {
    if (argc < 2) { fprintf(stderr, "usage: %s render|ids|rays|rng|dump ...\n", argv[0]); return 1; }
    std::string c = argv[1];
    if (c == "render") return cmd_render(argc, argv);
    if (c == "ids") return cmd_ids(argc, argv);
#ifdef REF_HOST
    if (c == "rays") return cmd_rays(argc, argv);
    if (c == "rng") return cmd_rng(argc, argv);
    if (c == "dump") return cmd_dump(argc, argv);
    if (c == "occluded") return cmd_occluded(argc, argv);
    if (c == "visibility") return cmd_visibility(argc, argv);
    if (c == "guided") return cmd_guided(argc, argv);
#endif
    fprintf(stderr, "unknown command %s\n", c.c_str());
    return 1;
}

// include/ptx/compat.hpp -- header-only C++ shim over the C ABI (ptx.h) in the reference's own vocabulary.
//
// The reference (USharma002/CUDA-PathTracer) has no FFI layer: its render loop is written against a handful of C++
// structs and free functions.  This header offers the same names, argument meaning and error behaviour on top of
// libptx.so, so that a render loop written against the reference compiles with its `#include`s swapped and its
// kernel launches replaced by the member calls below (INTEGRATION.md shows the edits).  Nothing here computes:
// every call forwards to an `extern "C"` entry point of include/ptx.h.
//
//   reference (file:line under /root/reference)                      here (namespace ptx_compat)
//   ---------------------------------------------------------------  --------------------------------------------
//   class Sensor               include/rendering/sensor.h:14-101     Sensor: same fields (origin, lower_left_corner,
//                                                                     horizontal, vertical, image_width, image_height,
//                                                                     radius, yaw, pitch, vup, lookat, vfov, aspect),
//                                                                     updateCamera / updateCameraOrbit / setPosition /
//                                                                     setLookAt / setFov
//   loadOBJ / loadPBRT / loadScene(filename, Primitive**, int&)      loadOBJ / loadPBRT / loadScene: `bool`, caller owns
//                              include/utils/file_manager.h:93,291    the array and frees it with freePrimitives()
//                              include/utils/pbrt_loader.h:178        (the reference: delete[])
//   RenderState                include/application_state.h:76-149    RenderState: width, height, h_image, h_camera,
//                                                                     allocateBuffers(), updateResolution(w, h), cleanup()
//   SceneState::loadScene(filename, subdivision_count, convert_quads) SceneState: loadScene(...) with the same three
//                              include/application_state.h:367-464    arguments, num_primitives, cleanup()
//   renderFrame()              include/application.h:157-216         renderFrame(render, scene, config): camera orbit
//                                                                     update, render, synchronise, copy RGB8 to h_image
//   AppConfig (spp, sampling_mode, subdivision_step,                 AppConfig: the fields the hot path reads
//              convert_quads_to_triangles) application_state.h:262-283
//   cudaMallocSafe throws std::runtime_error  utils/cuda_utils.h:53-60   allocateBuffers() throws std::runtime_error
#ifndef PTX_COMPAT_HPP
#define PTX_COMPAT_HPP

#include <cstddef>
#include <cstdint>
#include <cstring>
#include <iostream>
#include <stdexcept>
#include <string>

#include "../ptx.h"

namespace ptx_compat {

// the three floats the reference passes around as Vector3f (include/core/vector.h:20-23): e[0..2], x() y() z()
struct Vector3f {
    float e[3];
    Vector3f() : e{0.f, 0.f, 0.f} {}
    Vector3f(float x, float y, float z) : e{x, y, z} {}
    float x() const { return e[0]; }
    float y() const { return e[1]; }
    float z() const { return e[2]; }
};

// rendering/render_config.h:41-47
enum SamplingMode { SAMPLING_BSDF = 0, SAMPLING_FORMFACTOR = 1, SAMPLING_RADIOSITY = 2, SAMPLING_MIS = 3, SAMPLING_TOPK = 4 };

using Primitive = ptx_primitive;       // one reference Primitive minus its radiosity payload (primitive.h:21-29)

// ---- Sensor (rendering/sensor.h:14-101) --------------------------------------------------------------------
class Sensor {
  public:
    Sensor(Vector3f lookfrom, Vector3f lookat_, Vector3f vup_, float vfov_, float aspect_) {
        ptx_camera c{};
        ptx_camera_set(&c, lookfrom.e, lookat_.e, vup_.e, vfov_, aspect_);      // sets radius, yaw = 90, pitch = 0, updateCamera()
        c.image_width = 0; c.image_height = 0;
        from_c(c);
    }
    void updateCamera() { ptx_camera c = to_c(); ptx_camera_update(&c); from_c(c); }             // sensor.h:38-51
    void updateCameraOrbit() { ptx_camera c = to_c(); ptx_camera_orbit(&c); from_c(c); }         // sensor.h:56-67
    void setPosition(const Vector3f& p) { origin = p; updateCamera(); }
    void setLookAt(const Vector3f& t) { lookat = t; updateCamera(); }
    void setFov(float f) { vfov = f; updateCamera(); }

    float radius, yaw, pitch;
    Vector3f vup, lookat;
    Vector3f origin, lower_left_corner, horizontal, vertical;
    float vfov, aspect;
    int image_width, image_height;

    ptx_camera to_c() const {
        ptx_camera c{};
        c.radius = radius; c.yaw = yaw; c.pitch = pitch;
        std::memcpy(c.vup, vup.e, 12); std::memcpy(c.lookat, lookat.e, 12);
        std::memcpy(c.origin, origin.e, 12); std::memcpy(c.lower_left_corner, lower_left_corner.e, 12);
        std::memcpy(c.horizontal, horizontal.e, 12); std::memcpy(c.vertical, vertical.e, 12);
        c.vfov = vfov; c.aspect = aspect; c.image_width = image_width; c.image_height = image_height;
        return c;
    }
    void from_c(const ptx_camera& c) {
        radius = c.radius; yaw = c.yaw; pitch = c.pitch;
        std::memcpy(vup.e, c.vup, 12); std::memcpy(lookat.e, c.lookat, 12);
        std::memcpy(origin.e, c.origin, 12); std::memcpy(lower_left_corner.e, c.lower_left_corner, 12);
        std::memcpy(horizontal.e, c.horizontal, 12); std::memcpy(vertical.e, c.vertical, 12);
        vfov = c.vfov; aspect = c.aspect; image_width = c.image_width; image_height = c.image_height;
    }
};

// ---- loaders (utils/file_manager.h:93, 291-315; utils/pbrt_loader.h:178) -------------------------------------
// Same contract: false + a line on stderr when the file cannot be read or holds no primitives; on success the
// caller owns *h_list_out (free with freePrimitives; the reference uses delete[]).
inline bool loadScene(const std::string& filename, Primitive** h_list_out, int& num_primitives_out) {
    int64_t n = 0;
    int rc = ptx_parse_scene_file(filename.c_str(), nullptr, h_list_out, &n);
    if (rc != PTX_OK) {
        std::cerr << "Error: " << ptx_last_error(nullptr) << std::endl;
        *h_list_out = nullptr; num_primitives_out = 0;
        return false;
    }
    num_primitives_out = (int)n;
    return true;
}
inline bool loadOBJ(const std::string& obj_filename, Primitive** h_list_out, int& n) { return loadScene(obj_filename, h_list_out, n); }
inline bool loadPBRT(const std::string& pbrt_filename, Primitive** h_list_out, int& n) { return loadScene(pbrt_filename, h_list_out, n); }
inline void freePrimitives(Primitive* list) { ptx_free_primitives(list); }

// ---- AppConfig: the fields the hot path reads (application_state.h:262-283) ------------------------------------
struct AppConfig {
    int spp = 1;
    int max_depth = 5;                       // the literal at integrator.h:389
    SamplingMode sampling_mode = SAMPLING_BSDF;
    int subdivision_step = 0;
    bool convert_quads_to_triangles = false;
};

// one device context shared by RenderState and SceneState (the reference: the implicit device + default stream)
class Device {
  public:
    explicit Device(int index = 0) {
        if (ptx_ctx_create(index, &ctx_) != PTX_OK) throw std::runtime_error(std::string("ptx_ctx_create: ") + ptx_last_error(nullptr));
    }
    ~Device() { if (ctx_) ptx_ctx_destroy(ctx_); }
    Device(const Device&) = delete;
    Device& operator=(const Device&) = delete;
    ptx_ctx* ctx() const { return ctx_; }

  private:
    ptx_ctx* ctx_ = nullptr;
};

// ---- RenderState (application_state.h:76-149) -------------------------------------------------------------------
struct RenderState {
    unsigned char* h_image = nullptr;        // host RGB8, row 0 = bottom: what renderOpenGL uploads (application.h:221-246)
    int width = 800, height = 600, img_size = 0;
    Sensor h_camera;
    ptx_film* film = nullptr;                // replaces d_image, d_rand_state (and the never-read d_prev_image)
    ptx_ctx* ctx = nullptr;

    explicit RenderState(Device& dev)
        : h_camera(Vector3f(0.5f, 3.f, 8.5f), Vector3f(0.f, 2.5f, 0.f), Vector3f(0.f, 1.f, 0.f), 40.0f, 1.0f), ctx(dev.ctx()) {}
    ~RenderState() { cleanup(); }
    RenderState(const RenderState&) = delete;
    RenderState& operator=(const RenderState&) = delete;

    void allocateBuffers() {                 // application_state.h:91-123; throws like cudaMallocSafe
        if (film) { ptx_film_destroy(film); film = nullptr; }
        delete[] h_image;
        img_size = width * height * 3;
        h_image = new unsigned char[(size_t)img_size];
        h_camera.image_width = width; h_camera.image_height = height;
        h_camera.aspect = (float)width / (float)height;
        h_camera.updateCamera();
        if (ptx_film_create(ctx, width, height, &film) != PTX_OK)      // cudaMalloc d_image / d_rand_state + render_init<<<>>>
            throw std::runtime_error(std::string("CUDA malloc failed for film: ") + ptx_last_error(ctx));
    }
    void updateResolution(int w, int h) {    // application_state.h:125-129
        if (w == width && h == height && film) return;
        width = w; height = h;
        allocateBuffers();
    }
    void cleanup() {                         // application_state.h:131-143
        if (film) { ptx_film_destroy(film); film = nullptr; }
        delete[] h_image; h_image = nullptr;
    }
};

// ---- SceneState (application_state.h:153-165, 367-490) -----------------------------------------------------------
struct SceneState {
    ptx_scene* scene = nullptr;              // replaces h_primitives, d_primitives, d_bvh_nodes, d_bvh_indices, d_scene, h_scene
    int num_primitives = 0;
    int num_bvh_nodes = 0;
    ptx_ctx* ctx = nullptr;

    explicit SceneState(Device& dev) : ctx(dev.ctx()) {}
    ~SceneState() { cleanup(); }
    SceneState(const SceneState&) = delete;
    SceneState& operator=(const SceneState&) = delete;

    // loadOBJ/loadPBRT + convertQuadsToTriangles + subdivide_primitives + upload + buildAccelStructure + Scene(...)
    bool loadScene(const std::string& filename, int subdivision_count, bool convert_quads) {
        cleanup();
        ptx_load_opts o{};
        o.convert_quads = convert_quads ? 1 : 0;
        o.subdivision = subdivision_count;
        if (ptx_scene_load(ctx, filename.c_str(), &o, &scene) != PTX_OK) {
            std::cerr << "Failed to load scene: " << filename << " (" << ptx_last_error(ctx) << ")" << std::endl;   // :394-397
            scene = nullptr;
            return false;
        }
        ptx_scene_stats st;
        ptx_scene_stats_get(scene, &st);
        num_primitives = (int)st.n_primitives;
        num_bvh_nodes = (int)st.n_nodes;
        return true;
    }
    // SceneState::precomputeCDFs (application_state.h:492-585): the radiosity solver's per-primitive 16 x 16 RGB grids
    // (Triangle/Quad::radiosity_grid, num_primitives * 256 * 3 floats) become the CDFs the grid / MIS sampling modes read;
    // mis_bsdf_fraction = Scene::mis_bsdf_fraction (scene.h:215).  nullptr removes them.
    bool precomputeCDFs(const float* radiosity_grids_rgb, float mis_bsdf_fraction = 0.5f) {
        return scene && ptx_scene_set_radiosity_grids(scene, radiosity_grids_rgb, mis_bsdf_fraction) == PTX_OK;
    }
    // the host primitive list in the reference's numbering (what the radiosity solver reads as h_primitives)
    int getPrimitives(Primitive* out, int capacity) const { return (int)ptx_scene_get_primitives(scene, out, capacity); }
    void cleanup() {                         // application_state.h:466-490
        if (scene) { ptx_scene_destroy(scene); scene = nullptr; }
        num_primitives = num_bvh_nodes = 0;
    }
};

// ---- renderFrame (application.h:157-216) ---------------------------------------------------------------------------
// camera orbit update + by-value upload, render<<<>>>(d_image, d_camera, d_scene, d_rand_state, spp, mode),
// cudaDeviceSynchronize, cudaMemcpy d_image -> h_image.  Returns false (and prints, like the reference's CUDA checks)
// on error; the per-pixel XORWOW streams continue from frame to frame exactly as d_rand_state does.
inline bool renderFrame(RenderState& render, const SceneState& scene, const AppConfig& config, ptx_render_stats* stats = nullptr) {
    if (!scene.scene || !render.film) return false;                       // :158
    render.h_camera.updateCameraOrbit();                                  // :161-162
    ptx_camera cam = render.h_camera.to_c();                              // :163-164 (passed by value instead of cudaMemcpy)
    ptx_render_opts o{};
    o.spp = config.spp;
    o.max_depth = config.max_depth;
    o.rng_mode = PTX_RNG_XORWOW_COMPAT;
    o.sampling_mode = (int)config.sampling_mode;
    if (ptx_render(render.ctx, scene.scene, &cam, render.film, &o) != PTX_OK) {
        std::cerr << "CUDA error: " << ptx_last_error(render.ctx) << std::endl;
        return false;
    }
    if (stats) ptx_render_stats_get(render.ctx, stats);
    return ptx_film_resolve_rgb8(render.film, render.h_image, /*dst_is_device=*/0, /*flip_y=*/0) == PTX_OK;   // :199, :211-212
}

}  // namespace ptx_compat
#endif

// examples/render_loop.cpp -- the reference's main loop, headless, written against include/ptx/compat.hpp.
//
// Mirrors main() -> loadScene -> allocateBuffers -> { renderFrame } of the reference (src/main.cu:63-119,
// include/application.h:157-216) without the GLFW/OpenGL/ImGui shell (unavailable on this box): the frame that
// renderOpenGL would upload as a texture is written to a PNG instead (the "Save PNG" path, ui_windows.h:195-210).
//
//   render_loop <scene.obj|.pbrt> <width> <height> <spp> <frames> <out.png|-> [max_depth] [subdivision] [convert_quads]
//   render_loop --camera <width> <height>       print the default camera after allocateBuffers + updateCameraOrbit
//
// Build: g++ -std=c++17 -Iinclude examples/render_loop.cpp -Lcuda-pathtracer_b200 -lptx -Wl,-rpath,$PWD/cuda-pathtracer_b200
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include <ptx/compat.hpp>

using namespace ptx_compat;

int main(int argc, char** argv) {
    if (argc >= 4 && std::strcmp(argv[1], "--camera") == 0) {      // host-only: no device needed
        Sensor cam(Vector3f(0.5f, 3.f, 8.5f), Vector3f(0.f, 2.5f, 0.f), Vector3f(0.f, 1.f, 0.f), 40.0f, 1.0f);
        cam.image_width = std::atoi(argv[2]); cam.image_height = std::atoi(argv[3]);
        cam.aspect = (float)cam.image_width / (float)cam.image_height;
        cam.updateCamera();
        cam.updateCameraOrbit();
        std::printf("%.9g %.9g %.9g  %.9g %.9g %.9g  %.9g %.9g %.9g  %.9g %.9g %.9g  %.9g\n", cam.origin.x(), cam.origin.y(), cam.origin.z(),
                    cam.lower_left_corner.x(), cam.lower_left_corner.y(), cam.lower_left_corner.z(), cam.horizontal.x(), cam.horizontal.y(),
                    cam.horizontal.z(), cam.vertical.x(), cam.vertical.y(), cam.vertical.z(), cam.radius);
        return 0;
    }
    if (argc < 7) {
        std::fprintf(stderr, "usage: %s <scene> <width> <height> <spp> <frames> <out.png|-> [max_depth] [subdivision] [convert_quads]\n", argv[0]);
        return 2;
    }
    const char* scene_path = argv[1];
    const int width = std::atoi(argv[2]), height = std::atoi(argv[3]), frames = std::atoi(argv[5]);
    AppConfig config;
    config.spp = std::atoi(argv[4]);
    if (argc > 7) config.max_depth = std::atoi(argv[7]);
    if (argc > 8) config.subdivision_step = std::atoi(argv[8]);
    if (argc > 9) config.convert_quads_to_triangles = std::atoi(argv[9]) != 0;
    try {
        Device dev(0);                                   // no CUDA device => throws: the product has no CPU path
        RenderState render(dev);
        SceneState scene(dev);
        if (!scene.loadScene(scene_path, config.subdivision_step, config.convert_quads_to_triangles)) return 1;
        std::printf("Scene loaded: %d primitives, %d BVH nodes\n", scene.num_primitives, scene.num_bvh_nodes);
        render.updateResolution(width, height);          // allocateBuffers(): film + render_init
        unsigned long long segments = 0;
        double ms = 0.0;
        for (int f = 0; f < frames; f++) {               // the while (!glfwWindowShouldClose) loop, minus the window
            ptx_render_stats st{};
            if (!renderFrame(render, scene, config, &st)) return 1;
            segments += st.segments; ms += st.render_ms;
        }
        std::printf("%d frame(s), %llu segments, %.3f ms on the device, %.1f Mpaths*bounces/s\n", frames, segments, ms,
                    ms > 0.0 ? (double)segments / ms / 1e3 : 0.0);
        if (std::strcmp(argv[6], "-") != 0) {            // stbi_flip_vertically_on_write(1) + stbi_write_png (ui_windows.h:205-208)
            if (ptx_write_png_rgb8(argv[6], width, height, render.h_image, /*flip_y=*/1) != PTX_OK) {
                std::fprintf(stderr, "Failed to save image: %s\n", ptx_last_error(nullptr));
                return 1;
            }
            std::printf("Saved image to %s\n", argv[6]);
        }
    } catch (const std::exception& e) {
        std::fprintf(stderr, "fatal: %s\n", e.what());
        return 3;
    }
    return 0;
}

// examples/multi_gpu.cpp -- a C++ host rendering ONE frame on several GPUs through the C ABI only (include/ptx.h):
// one host thread per GPU, each with its own ptx_ctx, the scene replicated, tile sharding, and the film combined INSIDE the
// library (ptx_film_combine: pack owned tiles + ncclAllGather + unpack over NVLink).  The reference has no multi-GPU path.
//
//   multi_gpu <scene.obj|.pbrt> <width> <height> <spp> <depth> <n_gpus> [out.png]
//
// Prints one line per rank and, on rank 0, a checksum of the combined film next to the checksum of the same frame rendered
// unsharded on GPU 0: with the reference's RNG streams and tile sharding they are equal bit for bit.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "ptx.h"

#define CHECK(call)                                                                                   \
    do {                                                                                              \
        int rc_ = (call);                                                                             \
        if (rc_ != PTX_OK) { fprintf(stderr, "%s failed (%d): %s\n", #call, rc_, ptx_last_error(nullptr)); exit(1); } \
    } while (0)

static uint64_t checksum(const std::vector<float>& v) {
    uint64_t h = 1469598103934665603ull;
    for (float f : v) { uint32_t u; memcpy(&u, &f, 4); h = (h ^ u) * 1099511628211ull; }
    return h;
}

int main(int argc, char** argv) {
    if (argc < 7) { fprintf(stderr, "usage: %s scene width height spp depth n_gpus [out.png]\n", argv[0]); return 2; }
    const char* scene_path = argv[1];
    const int w = atoi(argv[2]), h = atoi(argv[3]), spp = atoi(argv[4]), depth = atoi(argv[5]), world = atoi(argv[6]);
    const char* out_png = argc > 7 ? argv[7] : nullptr;
    uint8_t id[PTX_COMM_ID_BYTES];
    if (world > 1) CHECK(ptx_comm_unique_id(id, sizeof id));
    std::vector<uint64_t> sums((size_t)world, 0);
    auto rank_main = [&](int rank) {
        ptx_ctx* ctx = nullptr; ptx_scene* scene = nullptr; ptx_film* film = nullptr; ptx_comm* comm = nullptr;
        CHECK(ptx_ctx_create(rank, &ctx));
        if (world > 1) CHECK(ptx_comm_create(ctx, id, rank, world, &comm));          // collective: every rank's thread calls it
        CHECK(ptx_scene_load(ctx, scene_path, nullptr, &scene));
        CHECK(ptx_film_create(ctx, w, h, &film));
        ptx_camera cam;
        CHECK(ptx_camera_default(&cam, w, h));
        ptx_render_opts o; memset(&o, 0, sizeof o);
        o.spp = spp; o.max_depth = depth; o.rng_mode = PTX_RNG_XORWOW_COMPAT;
        o.shard_mode = world > 1 ? PTX_SHARD_TILES : PTX_SHARD_NONE; o.shard_rank = rank; o.shard_world = world;
        o.flags = PTX_RENDER_ASYNC;                                                   // no host sync between render and combine
        CHECK(ptx_render(ctx, scene, &cam, film, &o));
        if (world > 1) CHECK(ptx_film_combine(film, comm, PTX_SHARD_TILES));
        std::vector<float> lin((size_t)w * h * 3);
        CHECK(ptx_film_read_linear(film, lin.data()));                                // every rank holds the whole frame now
        sums[(size_t)rank] = checksum(lin);
        ptx_render_stats st;
        CHECK(ptx_render_stats_get(ctx, &st));
        printf("rank %d of %d: %llu segments in %.2f ms, film checksum %016llx\n", rank, world, (unsigned long long)st.segments, st.render_ms,
               (unsigned long long)sums[(size_t)rank]);
        if (rank == 0 && out_png) CHECK(ptx_film_write(film, out_png));
        if (rank == 0) {                                                              // the same frame unsharded, for comparison
            CHECK(ptx_film_reset(film, 2023));
            o.shard_mode = PTX_SHARD_NONE; o.shard_rank = 0; o.shard_world = 1; o.flags = 0;
            CHECK(ptx_render(ctx, scene, &cam, film, &o));
            CHECK(ptx_film_read_linear(film, lin.data()));
            printf("unsharded on GPU 0: film checksum %016llx -> %s\n", (unsigned long long)checksum(lin), checksum(lin) == sums[0] ? "IDENTICAL" : "DIFFERENT");
        }
        ptx_comm_destroy(comm); ptx_film_destroy(film); ptx_scene_destroy(scene); ptx_ctx_destroy(ctx);
    };
    std::vector<std::thread> threads;
    for (int r = 0; r < world; r++) threads.emplace_back(rank_main, r);
    for (auto& t : threads) t.join();
    for (int r = 1; r < world; r++) if (sums[(size_t)r] != sums[0]) { fprintf(stderr, "rank %d holds a different film\n", r); return 1; }
    return 0;
}

// multi_rank_driver.cpp — TEST INFRASTRUCTURE (built and launched by tests/test_gpu_native_driver.py).
//
// What a host without Python and without torch.distributed does to render one frame on N GPUs through the C ABI alone
// (include/rrt.h) — the role src/main.rs:59-76 plays in the reference, once per rank:
//     rank 0: rrt_host_frame_create        other ranks: rrt_host_frame_open      (one canvas in shared host memory)
//     every rank, per frame:  rrt_render_device (its interleaved tile rows)  ->  rrt_host_frame_barrier ("previous
//     frame has been read")  ->  rrt_copy_shard_to_host_frame  ->  rrt_synchronize  ->  rrt_host_frame_barrier
//     ("frame complete"); rank 0 then owns the canvas until it enters the next frame's first barrier.
// It links librrt_b200.so and nothing else (no CUDA runtime calls of its own).  Rank 0 checks every assembled frame
// against the whole frame rendered by rrt_render on its own GPU, word for word, and prints one line per frame.
//   usage: multi_rank_driver <rank> <world> <shm name> <width> <height> <device> <frames>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../../include/rrt.h"

#define CHECK(call)                                                                              \
    do {                                                                                         \
        const int rc_ = (call);                                                                  \
        if (rc_ != RRT_OK) {                                                                     \
            std::fprintf(stderr, "rank %d: %s -> %d: %s\n", rank, #call, rc_, rrt_last_error()); \
            return 2;                                                                            \
        }                                                                                        \
    } while (0)

// deterministic scene, identical on every rank: a floor, a mirror wall and a field of small triangles
static void build_scene(std::vector<double>& v, std::vector<double>& vt, std::vector<double>& vn, std::vector<int32_t>& mat) {
    auto tri = [&](const double* a, const double* b, const double* c, int m) {
        const double e1[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]}, e2[3] = {c[0] - a[0], c[1] - a[1], c[2] - a[2]};
        double n[3] = {e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0]};
        const double l = std::sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
        for (double& x : n) x = l > 0 ? x / l : 0.0;
        for (const double* p : {a, b, c}) {
            v.insert(v.end(), p, p + 3);
            vn.insert(vn.end(), n, n + 3);
        }
        const double t[6] = {0.0, 0.0, 2.5, 0.0, 0.0, 2.5};
        vt.insert(vt.end(), t, t + 6);
        mat.push_back(m);
    };
    const double f0[3] = {-9, 0, -12}, f1[3] = {9, 0, -12}, f2[3] = {9, 0, 14}, f3[3] = {-9, 0, 14};
    tri(f0, f1, f2, 0); tri(f0, f2, f3, 0);
    const double w0[3] = {-6, 0, 12}, w1[3] = {6, 0, 12}, w2[3] = {6, 7, 12}, w3[3] = {-6, 7, 12};
    tri(w0, w1, w2, 1); tri(w0, w2, w3, 1);
    uint64_t s = 0x9E3779B97F4A7C15ull;
    auto rnd = [&s]() { s = s * 6364136223846793005ull + 1442695040888963407ull; return (double)(s >> 11) / 9007199254740992.0; };
    for (int i = 0; i < 6000; ++i) {
        const double c[3] = {-7.0 + 14.0 * rnd(), 0.3 + 5.0 * rnd(), -2.0 + 12.0 * rnd()};
        double p[3][3];
        for (auto& q : p) for (int k = 0; k < 3; ++k) q[k] = c[k] + 0.5 * (rnd() - 0.5);
        tri(p[0], p[1], p[2], 0);
    }
}

int main(int argc, char** argv) {
    if (argc < 8) { std::fprintf(stderr, "usage: %s rank world shm_name width height device frames\n", argv[0]); return 64; }
    const int rank = std::atoi(argv[1]), world = std::atoi(argv[2]);
    const char* name = argv[3];
    const int W = std::atoi(argv[4]), H = std::atoi(argv[5]), device = std::atoi(argv[6]), frames = std::atoi(argv[7]);
    if (rrt_device_count() <= device) { std::fprintf(stderr, "rank %d: no CUDA device %d (no CPU fallback)\n", rank, device); return 3; }

    std::vector<double> v, vt, vn;
    std::vector<int32_t> mat;
    build_scene(v, vt, vn, mat);
    std::vector<uint8_t> texels(64 * 64 * 3);
    for (int y = 0; y < 64; ++y)
        for (int x = 0; x < 64; ++x) {
            const bool c = ((x / 8 + y / 8) & 1) != 0;
            uint8_t* t = &texels[3 * (64 * y + x)];
            t[0] = c ? 230 : 70; t[1] = (uint8_t)(4 * x); t[2] = (uint8_t)(4 * y);
        }
    rrt_texture tex{64, 64, texels.data()};
    rrt_material mats[2] = {{{1, 1, 1}, {1, 1, 1}, {1, 1, 1}, 240.0, 0.0, 0, -1}, {{1, 1, 1}, {1, 1, 1}, {1, 1, 1}, 240.0, 0.8, 0, -1}};
    rrt_scene_desc d{};
    d.n_tris = mat.size(); d.v = v.data(); d.vt = vt.data(); d.vn = vn.data(); d.material_id = mat.data();
    d.n_materials = 2; d.materials = mats; d.n_textures = 1; d.textures = &tex;
    for (int k = 0; k < 3; ++k) { d.root_min[k] = -20.0; d.root_max[k] = 20.0; }   // utils.rs:145
    rrt_scene* scene = nullptr;
    CHECK(rrt_scene_create(&d, device, &scene));

    rrt_host_frame* hf = nullptr;
    if (rank == 0) CHECK(rrt_host_frame_create(name, W, H, world, &hf));
    else CHECK(rrt_host_frame_open(name, 60.0, &hf));
    CHECK(rrt_host_frame_pin(hf, device));
    uint32_t* canvas = rrt_host_frame_pixels(hf);

    // this rank's compact rows live in device memory the library hands out (rrt_frame_alloc: no CUDA runtime here)
    const rrt_shard sh{8, world, rank};
    const int rows = rrt_shard_rows(H, &sh);
    uint32_t* d_rows = nullptr;
    CHECK(rrt_frame_alloc(device, W, rows > 0 ? rows : 1, &d_rows));

    std::vector<uint32_t> whole((size_t)W * H);
    int bad_frames = 0;
    for (int f = 0; f < frames; ++f) {
        rrt_render_params p;
        rrt_render_params_default(&p);
        p.width = W; p.height = H;
        p.origin[0] = -2.0 + 1.3 * f;                 // a moving camera: every frame differs from the one before
        p.origin[1] = 2.0 + 0.2 * f;
        CHECK(rrt_render_device(scene, &p, &sh, d_rows, nullptr, nullptr));
        if (f > 0) CHECK(rrt_host_frame_barrier(hf, 120.0));          // rank 0 has finished with frame f - 1
        CHECK(rrt_copy_shard_to_host_frame(device, d_rows, W, H, &sh, canvas, nullptr));
        CHECK(rrt_synchronize(scene, nullptr));
        CHECK(rrt_host_frame_barrier(hf, 120.0));                     // frame f is complete in the shared canvas
        if (rank == 0) {
            CHECK(rrt_render(scene, &p, whole.data(), nullptr, nullptr));
            size_t diff = 0, lit = 0;
            for (size_t i = 0; i < whole.size(); ++i) { diff += canvas[i] != whole[i]; lit += whole[i] != 0x00FFFFFFu && whole[i] != 0u; }
            std::printf("frame %d: %zu of %zu words differ from the single-GPU frame (%zu pixels show geometry)\n", f, diff, whole.size(), lit);
            bad_frames += diff != 0 || lit < whole.size() / 10;
        }
    }
    CHECK(rrt_host_frame_barrier(hf, 120.0));
    CHECK(rrt_frame_free(device, d_rows));
    rrt_host_frame_close(hf);
    rrt_scene_destroy(scene);
    if (rank == 0) std::printf("%s\n", bad_frames ? "MISMATCH" : "ALL FRAMES EQUAL");
    return bad_frames ? 1 : 0;
}

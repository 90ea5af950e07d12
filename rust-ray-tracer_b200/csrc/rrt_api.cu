// rrt_api.cu — the C ABI of include/rrt.h: scene upload, render calls, per-ray seams.
// No CPU rendering path exists in this library: every render entry point needs a CUDA device.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "../../include/rrt.h"
#include "rrt_octree.h"
#include "rrt_render.h"
#include "rrt_shade.cuh"

using namespace rrt;

static_assert(RRT_NOALLOC_MAX_PIXELS == kNoAllocMaxPixels, "include/rrt.h and rrt_render.h state one limit");

namespace {

thread_local std::string g_err = "";

int fail(int code, const std::string& msg) {
    g_err = msg;
    return code;
}
int cuda_fail(cudaError_t e, const char* what) {
    return fail(RRT_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}
#define CU(x)                                                    \
    do {                                                         \
        cudaError_t e_ = (x);                                    \
        if (e_ != cudaSuccess) return cuda_fail(e_, #x);         \
    } while (0)

// lists longer than this get a culling hierarchy; RRT_CULL_THRESHOLD=0 scans every list linearly
uint32_t cull_threshold() {
    const char* s = std::getenv("RRT_CULL_THRESHOLD");
    if (s && *s) return (uint32_t)std::strtoul(s, nullptr, 10);
    return 1;
}
// RRT_OCTREE_BUILD=seq forces the sequential insertion loop (the literal octree.rs algorithm); default: the
// order-exact parallel construction for anything big enough to matter.
bool build_tree(HostOctree& tree, const double* v, uint64_t n, const double* lo, const double* hi, std::string* err) {
    const char* s = std::getenv("RRT_OCTREE_BUILD");
    const bool seq = s ? std::string(s) == "seq" : n < 20000;
    return seq ? tree.build(v, n, lo, hi, err) : tree.build_parallel(v, n, lo, hi, err);
}
uint32_t cull_leaf() {
    const char* s = std::getenv("RRT_CULL_LEAF");
    if (s && *s) return (uint32_t)std::strtoul(s, nullptr, 10);
    return 1;
}

}  // namespace

struct rrt_scene {
    int device = 0;
    DevScene dev{};
    rrt_scene_stats stats{};
    std::vector<void*> allocs;
    // scratch for the host-buffer entry points, grown on demand
    uint32_t* d_fb = nullptr;
    int32_t* d_hit = nullptr;
    uint32_t* d_scratch = nullptr;   // per-pixel accumulators + sample-queue head of the persistent kernel
    size_t fb_cap = 0, hit_cap = 0, scratch_cap = 0;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    RenderSchedule sched;            // dispatch order of the next frame from the measured cost of this one (rrt_render.h)
    uint32_t* d_fstat = nullptr;     // what the searches round the walk returned last frame + their switches (DevParams::fstat)
    int search_mode = 1;             // RRT_SEARCH_MODE at scene creation: 0 walk only, 1 switched per frame (default), 2 always
    // rrt_render_frames: two device frames, a render and a copy stream, created on first use
    uint32_t* d_seq_fb[2] = {nullptr, nullptr};
    size_t seq_cap[2] = {0, 0};
    cudaStream_t seq_render = nullptr, seq_copy = nullptr;
    cudaEvent_t seq_rendered[2] = {nullptr, nullptr}, seq_copied[2] = {nullptr, nullptr}, seq_t0 = nullptr, seq_t1 = nullptr;

    template <class T>
    int upload(const T* host, size_t n, const T** out) {
        void* p = nullptr;
        const size_t bytes = std::max<size_t>(n, 1) * sizeof(T);
        CU(cudaMalloc(&p, bytes));
        allocs.push_back(p);
        if (n) CU(cudaMemcpy(p, host, n * sizeof(T), cudaMemcpyHostToDevice));
        stats.device_bytes += bytes;
        *out = static_cast<const T*>(p);
        return RRT_OK;
    }
    ~rrt_scene() {
        cudaSetDevice(device);
        for (void* p : allocs) cudaFree(p);
        if (d_fb) cudaFree(d_fb);
        if (d_hit) cudaFree(d_hit);
        if (d_scratch) cudaFree(d_scratch);
        if (d_fstat) cudaFree(d_fstat);
        sched.release();
        if (seq_render) { cudaStreamSynchronize(seq_render); cudaStreamDestroy(seq_render); }
        if (seq_copy) { cudaStreamSynchronize(seq_copy); cudaStreamDestroy(seq_copy); }
        for (int b = 0; b < 2; ++b) {
            if (d_seq_fb[b]) cudaFree(d_seq_fb[b]);
            if (seq_rendered[b]) cudaEventDestroy(seq_rendered[b]);
            if (seq_copied[b]) cudaEventDestroy(seq_copied[b]);
        }
        if (seq_t0) cudaEventDestroy(seq_t0);
        if (seq_t1) cudaEventDestroy(seq_t1);
        if (ev0) cudaEventDestroy(ev0);
        if (ev1) cudaEventDestroy(ev1);
    }
};

namespace {

int make_params(const rrt_scene* scene, const rrt_render_params* p, const rrt_shard* sh, DevParams* out) {
    if (!p) return fail(RRT_ERR_INVALID, "null params");
    if (p->width <= 0 || p->height <= 0) return fail(RRT_ERR_INVALID, "width/height must be positive");
    if ((int64_t)p->width * p->height > (int64_t)1 << 30) return fail(RRT_ERR_INVALID, "frame too large");
    if (p->n_lights < 0 || p->n_lights > RRT_MAX_LIGHTS) return fail(RRT_ERR_INVALID, "n_lights out of range");
    if (p->max_depth < 0) return fail(RRT_ERR_INVALID, "max_depth < 0");
    // the reference recurses to whatever depth it is given (raytracer.rs:76); the kernel keeps one mirror level per
    // slot of a fixed per-lane array, so a deeper request is refused rather than rendered differently
    if (p->max_depth > kMaxReflectDepth)
        return fail(RRT_ERR_INVALID, "max_depth > " + std::to_string(kMaxReflectDepth) + " (mirror levels kept per lane)");
    DevParams d{};
    d.W = p->width;
    d.H = p->height;
    d.ox = p->origin[0]; d.oy = p->origin[1]; d.oz = p->origin[2];
    d.xs = p->viewport[0] / (double)p->width;    // engine.rs:189
    d.ys = p->viewport[1] / (double)p->height;   // engine.rs:190
    d.zv = p->viewport[2];                       // engine.rs:191
    d.n_lights = p->n_lights;
    d.max_depth = p->max_depth;
    for (int i = 0; i < p->n_lights; ++i) {
        if (p->lights[i].kind < 0 || p->lights[i].kind > 2) return fail(RRT_ERR_INVALID, "bad light kind");
        d.lights[i] = p->lights[i];
    }
    d.offset = p->surface_offset;
    d.use_rot = p->use_camera_rot;
    for (int i = 0; i < 9; ++i) d.rot[i] = p->camera_rot[i];
    rrt_shard one{1, 1, 0};
    if (!sh) sh = &one;
    if (sh->tile_rows < 1 || sh->n_shards < 1 || sh->shard_index < 0 || sh->shard_index >= sh->n_shards)
        return fail(RRT_ERR_INVALID, "bad shard");
    d.tile_rows = sh->tile_rows;
    d.n_shards = sh->n_shards;
    d.shard_index = sh->shard_index;
    d.n_rows_local = rrt_shard_rows(p->height, sh);
    d.search_mode = scene ? scene->search_mode : 0;
    d.fstat = scene ? scene->d_fstat : nullptr;
    if (!d.fstat && d.search_mode == 1) d.search_mode = 0;
    *out = d;
    return RRT_OK;
}

}  // namespace

extern "C" {

int rrt_abi_version(void) { return RRT_ABI_VERSION; }
void rrt_set_last_error_(const char* msg) { g_err = msg ? msg : ""; }  // used by rrt_loader.cpp
const char* rrt_last_error(void) { return g_err.c_str(); }

int rrt_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

void rrt_render_params_default(rrt_render_params* p) {
    if (!p) return;
    std::memset(p, 0, sizeof *p);
    p->width = 800;   // main.rs:14
    p->height = 800;  // main.rs:15
    p->origin[0] = 0.0; p->origin[1] = 2.0; p->origin[2] = -10.0;  // main.rs:62-66
    p->viewport[0] = p->viewport[1] = p->viewport[2] = 1.0;        // engine.rs:113-119
    p->n_lights = 4;                                               // main.rs:32-58
    p->lights[0] = rrt_light{RRT_LIGHT_AMBIENT, 0, 0.5, {0, 0, 0}};
    p->lights[1] = rrt_light{RRT_LIGHT_POINT, 0, 0.4, {-7.0, 1.0, -15.0}};
    p->lights[2] = rrt_light{RRT_LIGHT_POINT, 0, 0.5, {0.0, 1.0, -41.0}};
    p->lights[3] = rrt_light{RRT_LIGHT_DIRECTIONAL, 0, 0.4, {-5.0, 0.0, 20.0}};
    p->max_depth = 5;           // raytracer.rs:20
    p->surface_offset = 0.0001; // raytracer.rs:17
    p->use_camera_rot = 0;
    p->camera_rot[0] = p->camera_rot[4] = p->camera_rot[8] = 1.0;
}

int rrt_shard_rows(int32_t height, const rrt_shard* sh) {
    if (!sh || sh->n_shards <= 1) return height;
    const int T = sh->tile_rows, G = sh->n_shards, s = sh->shard_index;
    if (T < 1 || s < 0 || s >= G) return 0;
    int rows = 0;
    for (int j = s; (int64_t)j * T < height; j += G) rows += std::min(T, height - j * T);
    return rows;
}

int rrt_scene_create(const rrt_scene_desc* d, int device, rrt_scene** out) {
    if (!d || !out) return fail(RRT_ERR_INVALID, "null argument");
    *out = nullptr;
    if (d->n_tris >= (1ull << 28)) return fail(RRT_ERR_INVALID, "too many triangles (record index is 28 bits)");
    if (d->n_tris && (!d->v || !d->vt || !d->vn || !d->material_id)) return fail(RRT_ERR_INVALID, "null triangle arrays");
    // validation the reference performs by panicking (utils.rs:61 unwrap; Vec index out of range)
    for (uint32_t i = 0; i < d->n_textures; ++i)
        if (d->textures[i].width <= 0 || d->textures[i].height <= 0 || !d->textures[i].rgb)
            return fail(RRT_ERR_INVALID, "empty texture");
    for (uint32_t i = 0; i < d->n_materials; ++i) {
        const rrt_material& m = d->materials[i];
        if (m.tex_id < 0 || (uint32_t)m.tex_id >= d->n_textures) return fail(RRT_ERR_INVALID, "material without texture (utils.rs:61)");
        if (m.bump_id >= (int32_t)d->n_textures) return fail(RRT_ERR_INVALID, "bump_id out of range");
        if (m.bump_id >= 0) {
            const rrt_texture &a = d->textures[m.tex_id], &b = d->textures[m.bump_id];
            // the bump map is indexed with the COLOUR map's (tx,ty) and its own stride (raytracer.rs:128)
            if ((int64_t)b.width * (a.height - 1) + (a.width - 1) >= (int64_t)b.width * b.height)
                return fail(RRT_ERR_INVALID, "bump map smaller than colour map (reference would panic)");
        }
    }
    for (uint64_t i = 0; i < d->n_tris; ++i)
        if (d->material_id[i] < 0 || (uint32_t)d->material_id[i] >= d->n_materials)
            return fail(RRT_ERR_INVALID, "material_id out of range");
    int ndev = rrt_device_count();
    if (ndev <= 0) return fail(RRT_ERR_CUDA, "no CUDA device (this library has no CPU fallback)");
    if (device < 0 || device >= ndev) return fail(RRT_ERR_INVALID, "device index out of range");
    CU(cudaSetDevice(device));

    // host: exact reference octree, then flatten
    HostOctree tree;
    std::string err;
    if (!build_tree(tree, d->v, d->n_tris, d->root_min, d->root_max, &err)) return fail(RRT_ERR_DEPTH, err);
    FlatScene flat;
    // RRT_SEARCH_MODE: 0 = octree walk only (no scene-wide hierarchy is built), 1 = the searches that go round the walk
    // are switched per frame by what they answered (default), 2 = always tried
    int search_mode = 1;
    if (const char* e = std::getenv("RRT_SEARCH_MODE")) search_mode = std::max(0, std::min(2, std::atoi(e)));
    flatten(tree, d->v, cull_threshold(), cull_leaf(), &flat, search_mode != 0);
    if (flat.global_root == 0u) search_mode = 0;

    auto sc = std::make_unique<rrt_scene>();
    sc->search_mode = search_mode;
    sc->device = device;
    if (const char* e = getenv("RRT_ORDER_HINT")) sc->sched.enabled = atoi(e) != 0;
    if (const char* e = getenv("RRT_ORDER_TAIL")) sc->sched.tail_safety = (float)atof(e);   // tuning knob (rrt_render.h)
    rrt_scene_stats& st = sc->stats;
    st.n_tris = d->n_tris;
    st.n_tris_stored = flat.n_stored;
    st.n_tris_dropped = tree.n_dropped;
    st.n_nodes = tree.nodes.size();
    st.n_nodes_nonempty = (tree.nodes[0].triangle_count > 0) ? flat.nodes.size() : 0;
    st.max_depth = tree.max_depth;
    st.root_list = flat.root_list;
    st.max_list = flat.max_list;
    st.n_cull_nodes = flat.cull.size();
    { std::vector<HostNode>().swap(tree.nodes); std::vector<uint32_t>().swap(tree.next); }

    int rc;
    if ((rc = sc->upload(flat.nodes.data(), flat.nodes.size(), &sc->dev.nodes))) return rc;
    if ((rc = sc->upload(flat.tris.data(), flat.tris.size(), &sc->dev.tris))) return rc;
    if ((rc = sc->upload(flat.cull.data(), flat.cull.size(), &sc->dev.cull))) return rc;
    if ((rc = sc->upload(flat.sub_boxes.data(), flat.sub_boxes.size(), &sc->dev.sub_boxes))) return rc;
    sc->dev.n_nodes = (uint32_t)flat.nodes.size();
    sc->dev.root_tri_count = st.n_tris_stored ? 1u : 0u;
    sc->dev.planes_safe = (flat.planes_safe && !std::getenv("RRT_SLOW_DIV")) ? 1u : 0u;
    sc->dev.global_root = flat.global_root;
    sc->dev.sparse_ok = flat.sparse_ok ? 1u : 0u;
    {
        const uint32_t init[12] = {0u, 0u, 1u, 0u, 0u, 0u, 1u, 0u, 0u, 0u, 0u, 0u};   // both searches start switched on
        CU(cudaMalloc((void**)&sc->d_fstat, sizeof init));
        CU(cudaMemcpy(sc->d_fstat, init, sizeof init, cudaMemcpyHostToDevice));
    }
#if RRT_CULL_F32
    sc->dev.cull_o_max = kCullOriginRel * flat.scale;
#else
    sc->dev.cull_o_max = 0.0;
#endif

    // shading records by original triangle id (entities.rs:78-84)
    {
        std::vector<DevShade> sh(d->n_tris);
#pragma omp parallel for schedule(static)
        for (long long i = 0; i < (long long)d->n_tris; ++i) {
            const double *n = d->vn + 9 * i, *t = d->vt + 6 * i;
            DevShade& s = sh[i];
            std::memcpy(s.n1, n, 24); std::memcpy(s.n2, n + 3, 24); std::memcpy(s.n3, n + 6, 24);
            std::memcpy(s.t1, t, 16); std::memcpy(s.t2, t + 2, 16); std::memcpy(s.t3, t + 4, 16);
            s.material = d->material_id[i];
            s._pad = 0;
        }
        if ((rc = sc->upload(sh.data(), sh.size(), &sc->dev.shade))) return rc;
    }
    {
        std::vector<DevMaterial> ms(d->n_materials);
        for (uint32_t i = 0; i < d->n_materials; ++i) {
            const rrt_material& m = d->materials[i];
            std::memcpy(ms[i].ka, m.ka, 24); std::memcpy(ms[i].kd, m.kd, 24); std::memcpy(ms[i].ks, m.ks, 24);
            ms[i].ns = m.ns; ms[i].kr = m.kr; ms[i].tex = m.tex_id; ms[i].bump = m.bump_id;
        }
        if ((rc = sc->upload(ms.data(), ms.size(), &sc->dev.materials))) return rc;
    }
    {
        std::vector<DevTexture> ts(d->n_textures);
        std::vector<uchar4> tmp;
        for (uint32_t i = 0; i < d->n_textures; ++i) {
            const rrt_texture& t = d->textures[i];
            const size_t n = (size_t)t.width * t.height;
            tmp.resize(n);
            for (size_t k = 0; k < n; ++k) tmp[k] = make_uchar4(t.rgb[3 * k], t.rgb[3 * k + 1], t.rgb[3 * k + 2], 0);
            const uchar4* dp = nullptr;
            if ((rc = sc->upload(tmp.data(), n, &dp))) return rc;
            ts[i] = DevTexture{dp, (uint32_t)t.width, (uint32_t)t.height};
        }
        if ((rc = sc->upload(ts.data(), ts.size(), &sc->dev.tex))) return rc;
    }
    CU(cudaEventCreate(&sc->ev0));
    CU(cudaEventCreate(&sc->ev1));
    *out = sc.release();
    return RRT_OK;
}

void rrt_scene_destroy(rrt_scene* s) { delete s; }

int rrt_scene_order_hint(rrt_scene* s, int enable) {
    if (!s) return fail(RRT_ERR_INVALID, "null scene");
    s->sched.enabled = enable != 0;
    s->sched.valid = false;
    s->sched.n_blocks = 0;   // next launch re-arms the cost array
    return RRT_OK;
}

int rrt_scene_search_mode(rrt_scene* s, int mode, int filter_on, int sparse_on) {
    if (!s) return fail(RRT_ERR_INVALID, "null scene");
    if (mode < 0 || mode > 3) return fail(RRT_ERR_INVALID, "search mode must be 0..3");
    if (mode != 0 && s->dev.global_root == 0u)
        return fail(RRT_ERR_INVALID, "scene was created without the scene-wide hierarchy (RRT_SEARCH_MODE=0 or a scene it does not apply to)");
    CU(cudaSetDevice(s->device));
    const uint32_t init[12] = {0u, 0u, mode == 3 ? (filter_on ? 1u : 0u) : 1u, 0u, 0u, 0u, mode == 3 ? (sparse_on ? 1u : 0u) : 1u, 0u, 0u, 0u, 0u, 0u};
    CU(cudaMemcpy(s->d_fstat, init, sizeof init, cudaMemcpyHostToDevice));
    s->search_mode = mode;
    return RRT_OK;
}

int rrt_scene_search_state(rrt_scene* s, uint32_t out[12]) {
    if (!s || !out) return fail(RRT_ERR_INVALID, "null argument");
    CU(cudaSetDevice(s->device));
    CU(cudaDeviceSynchronize());
    CU(cudaMemcpy(out, s->d_fstat, 12 * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    return RRT_OK;
}

int rrt_scene_get_stats(const rrt_scene* s, rrt_scene_stats* out) {
    if (!s || !out) return fail(RRT_ERR_INVALID, "null argument");
    *out = s->stats;
    return RRT_OK;
}

}  // extern "C"

// ---------------------------------------------------------------- render entry points
namespace {

template <class T>
int ensure(T** p, size_t* cap, size_t n) {
    if (*cap >= n) return RRT_OK;
    if (*p) cudaFree(*p);
    *p = nullptr;
    *cap = 0;
    CU(cudaMalloc((void**)p, n * sizeof(T)));
    *cap = n;
    return RRT_OK;
}

// one frame (or shard) of Scene::draw_scene on `stream`
cudaError_t render_frame(rrt_scene* s, const DevParams& P, uint32_t* d_fb, int32_t* d_hit, DevCounters* d_counters,
                         cudaStream_t stream) {
    return launch_render(s->dev, P, d_fb, d_hit, s->d_scratch, d_counters, stream, d_counters ? nullptr : &s->sched);
}

struct DevBuf {  // RAII scratch for the per-ray seams
    void* p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    int alloc(size_t bytes) {
        CU(cudaMalloc(&p, bytes ? bytes : 1));
        return RRT_OK;
    }
};

}  // namespace

extern "C" {

int rrt_render_device(rrt_scene* s, const rrt_render_params* p, const rrt_shard* shard, uint32_t* d_fb,
                      int32_t* d_hit_ids, void* cuda_stream) {
    if (!s || !d_fb) return fail(RRT_ERR_INVALID, "null argument");
    DevParams P;
    int rc = make_params(s, p, shard, &P);
    if (rc) return rc;
    CU(cudaSetDevice(s->device));
    if ((rc = ensure(&s->d_scratch, &s->scratch_cap, render_scratch_words(P.n_rows_local, P.W)))) return rc;
    CU(render_frame(s, P, d_fb, d_hit_ids, nullptr, (cudaStream_t)cuda_stream));
    return RRT_OK;
}

// Allocation-free rrt_render_device (no counterpart in the reference):
// calls no cudaMalloc / cudaFree and no allocating library routine — the kernel's scratch is a static device array
// of rrt_render.cu, so the shard may hold at most RRT_NOALLOC_MAX_PIXELS pixels.
int rrt_render_device_noalloc(rrt_scene* s, const rrt_render_params* p, const rrt_shard* shard, uint32_t* d_fb,
                              int32_t* d_hit_ids, void* cuda_stream) {
    if (!s || !d_fb) return fail(RRT_ERR_INVALID, "null argument");
    DevParams P;
    int rc = make_params(s, p, shard, &P);
    if (rc) return rc;
    const unsigned long long npix = (unsigned long long)P.n_rows_local * (unsigned long long)P.W;
    if (npix > (unsigned long long)RRT_NOALLOC_MAX_PIXELS)
        return fail(RRT_ERR_INVALID, "rrt_render_device_noalloc: " + std::to_string(npix) + " pixels exceed the static scratch of " +
                                         std::to_string((unsigned long long)RRT_NOALLOC_MAX_PIXELS) +
                                         " pixels (RRT_NOALLOC_MAX_PIXELS); render it in shards or use rrt_render_device");
    CU(cudaSetDevice(s->device));
    CU(launch_render_noalloc(s->dev, P, d_fb, d_hit_ids, (cudaStream_t)cuda_stream));
    return RRT_OK;
}

int rrt_scene_reserve(rrt_scene* s, const rrt_render_params* p, const rrt_shard* shard, int with_device_frame) {
    if (!s) return fail(RRT_ERR_INVALID, "null scene");
    DevParams P;
    int rc = make_params(s, p, shard, &P);
    if (rc) return rc;
    CU(cudaSetDevice(s->device));
    if ((rc = ensure(&s->d_scratch, &s->scratch_cap, render_scratch_words(P.n_rows_local, P.W)))) return rc;
    if (with_device_frame && (rc = ensure(&s->d_fb, &s->fb_cap, (size_t)P.W * std::max(P.n_rows_local, 1)))) return rc;
    CU(reserve_render(P, &s->sched, 0));
    CU(cudaStreamSynchronize(0));
    return RRT_OK;
}

int rrt_render_device_into_frame(rrt_scene* s, const rrt_render_params* p, const rrt_shard* shard, uint32_t* d_frame,
                                 void* cuda_stream) {
    if (!s || !d_frame) return fail(RRT_ERR_INVALID, "null argument");
    DevParams P;
    int rc = make_params(s, p, shard, &P);
    if (rc) return rc;
    P.full_frame = 1;
    CU(cudaSetDevice(s->device));
    if ((rc = ensure(&s->d_scratch, &s->scratch_cap, render_scratch_words(P.n_rows_local, P.W)))) return rc;
    CU(render_frame(s, P, d_frame, nullptr, nullptr, (cudaStream_t)cuda_stream));
    return RRT_OK;
}

int rrt_frame_alloc(int device, int32_t width, int32_t height, uint32_t** d_frame) {
    if (!d_frame || width <= 0 || height <= 0) return fail(RRT_ERR_INVALID, "bad argument");
    CU(cudaSetDevice(device));
    void* p = nullptr;
    CU(cudaMalloc(&p, (size_t)width * height * sizeof(uint32_t)));   // its own allocation: IPC handles export whole allocations
    CU(cudaMemset(p, 0, (size_t)width * height * sizeof(uint32_t)));
    *d_frame = (uint32_t*)p;
    return RRT_OK;
}
int rrt_frame_free(int device, uint32_t* d_frame) {
    CU(cudaSetDevice(device));
    if (d_frame) CU(cudaFree(d_frame));
    return RRT_OK;
}
int rrt_ipc_get_handle(int device, const void* d_ptr, unsigned char handle[64]) {
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t size");
    if (!d_ptr || !handle) return fail(RRT_ERR_INVALID, "null argument");
    CU(cudaSetDevice(device));
    cudaIpcMemHandle_t h;
    CU(cudaIpcGetMemHandle(&h, const_cast<void*>(d_ptr)));
    std::memcpy(handle, &h, 64);
    return RRT_OK;
}
int rrt_ipc_open_handle(int device, const unsigned char handle[64], void** d_ptr) {
    if (!d_ptr || !handle) return fail(RRT_ERR_INVALID, "null argument");
    CU(cudaSetDevice(device));
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle, 64);
    CU(cudaIpcOpenMemHandle(d_ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return RRT_OK;
}
int rrt_ipc_close_handle(int device, void* d_ptr) {
    CU(cudaSetDevice(device));
    if (d_ptr) CU(cudaIpcCloseMemHandle(d_ptr));
    return RRT_OK;
}
int rrt_copy_frame_to_host(int device, const uint32_t* d_frame, int32_t width, int32_t height, uint32_t* fb,
                           void* cuda_stream) {
    if (!d_frame || !fb || width <= 0 || height <= 0) return fail(RRT_ERR_INVALID, "bad argument");
    CU(cudaSetDevice(device));
    CU(cudaMemcpyAsync(fb, d_frame, (size_t)width * height * sizeof(uint32_t), cudaMemcpyDeviceToHost,
                       (cudaStream_t)cuda_stream));
    return RRT_OK;
}

// A shard's compact rows leave for their places in a FULL frame in host memory: tile row j of the frame (tile_rows
// scan-lines) is the shard's j / n_shards-th group, so one pitched copy moves every whole group and a second one the
// ragged last group (only the frame's last tile row can be short).
// Device alias of a caller's HOST frame when the GPU can store into it directly: page-locked memory (cudaHostAlloc,
// cudaHostRegister, the shared canvas of rrt_host_frame_pin) is mapped into the device's address space under unified
// addressing.  Then the render kernel's finished pixels go straight to the host frame as posted PCIe writes while the
// frame is still being rendered, and no device->host copy follows the kernel (a 4K frame: 33 MB, 0.6 ms on one link).
// nullptr = pageable memory (or switched off with RRT_HOST_STORES=0): render into device memory and copy.
static uint32_t* host_frame_alias(const uint32_t* host, size_t words) {
    static const bool enabled = [] { const char* e = std::getenv("RRT_HOST_STORES"); return !(e && e[0] == '0'); }();
    if (!enabled || !host || words == 0) return nullptr;
    cudaPointerAttributes a0{}, a1{};
    if (cudaPointerGetAttributes(&a0, host) != cudaSuccess || cudaPointerGetAttributes(&a1, host + (words - 1)) != cudaSuccess) {
        (void)cudaGetLastError();
        return nullptr;
    }
    if (a0.type != cudaMemoryTypeHost || a1.type != cudaMemoryTypeHost || !a0.devicePointer || !a1.devicePointer) return nullptr;
    uint32_t* d0 = static_cast<uint32_t*>(a0.devicePointer);
    if (static_cast<uint32_t*>(a1.devicePointer) != d0 + (words - 1)) return nullptr;   // not one mapped range
    return d0;
}

int rrt_copy_shard_to_host_frame(int device, const uint32_t* d_rows, int32_t width, int32_t height, const rrt_shard* shard,
                                 uint32_t* host_frame, void* cuda_stream) {
    if (!d_rows || !host_frame || width <= 0 || height <= 0) return fail(RRT_ERR_INVALID, "bad argument");
    rrt_shard one{1, 1, 0};
    const rrt_shard* sh = shard ? shard : &one;
    if (sh->tile_rows < 1 || sh->n_shards < 1 || sh->shard_index < 0 || sh->shard_index >= sh->n_shards)
        return fail(RRT_ERR_INVALID, "bad shard");
    CU(cudaSetDevice(device));
    cudaStream_t st = (cudaStream_t)cuda_stream;
    const int rows = rrt_shard_rows(height, sh);
    const size_t row_bytes = (size_t)width * sizeof(uint32_t), group = row_bytes * sh->tile_rows;
    const int n_full = rows / sh->tile_rows, rest = rows % sh->tile_rows;
    uint32_t* dst0 = host_frame + (size_t)sh->shard_index * sh->tile_rows * width;
    if (n_full > 0)
        CU(cudaMemcpy2DAsync(dst0, group * sh->n_shards, d_rows, group, group, (size_t)n_full, cudaMemcpyDeviceToHost, st));
    if (rest > 0)
        CU(cudaMemcpyAsync(dst0 + (size_t)n_full * sh->n_shards * sh->tile_rows * width, d_rows + (size_t)n_full * sh->tile_rows * width,
                           row_bytes * rest, cudaMemcpyDeviceToHost, st));
    return RRT_OK;
}

int rrt_render_to_host_frame(rrt_scene* s, const rrt_render_params* p, const rrt_shard* shard, uint32_t* host_frame,
                             void* cuda_stream) {
    if (!s || !host_frame) return fail(RRT_ERR_INVALID, "null argument");
    DevParams P;
    int rc = make_params(s, p, shard, &P);
    if (rc) return rc;
    CU(cudaSetDevice(s->device));
    cudaStream_t st = (cudaStream_t)cuda_stream;
    if ((rc = ensure(&s->d_scratch, &s->scratch_cap, render_scratch_words(P.n_rows_local, P.W)))) return rc;
    if (uint32_t* alias = host_frame_alias(host_frame, (size_t)P.W * P.H)) {
        P.full_frame = 1;   // owned row k -> its row of the full host frame, stored by the kernel itself
        CU(render_frame(s, P, alias, nullptr, nullptr, st));
        return RRT_OK;
    }
    const size_t n = (size_t)P.W * std::max(P.n_rows_local, 1);
    if ((rc = ensure(&s->d_fb, &s->fb_cap, n))) return rc;
    CU(render_frame(s, P, s->d_fb, nullptr, nullptr, st));
    return rrt_copy_shard_to_host_frame(s->device, s->d_fb, P.W, P.H, shard, host_frame, cuda_stream);
}

int rrt_synchronize(rrt_scene* s, void* cuda_stream) {
    if (!s) return fail(RRT_ERR_INVALID, "null scene");
    CU(cudaSetDevice(s->device));
    CU(cudaStreamSynchronize((cudaStream_t)cuda_stream));
    return RRT_OK;
}

int rrt_render(rrt_scene* s, const rrt_render_params* p, uint32_t* fb, int32_t* hit_ids, float* device_ms) {
    if (!s || !fb) return fail(RRT_ERR_INVALID, "null argument");
    DevParams P;
    int rc = make_params(s, p, nullptr, &P);
    if (rc) return rc;
    CU(cudaSetDevice(s->device));
    const size_t n = (size_t)P.W * P.H;
    uint32_t* alias = host_frame_alias(fb, n);   // page-locked caller buffer: the kernel stores into it, no copy
    if (!alias && (rc = ensure(&s->d_fb, &s->fb_cap, n))) return rc;
    if (hit_ids && (rc = ensure(&s->d_hit, &s->hit_cap, n * 4))) return rc;
    if ((rc = ensure(&s->d_scratch, &s->scratch_cap, render_scratch_words(P.n_rows_local, P.W)))) return rc;
    CU(cudaEventRecord(s->ev0, 0));
    CU(render_frame(s, P, alias ? alias : s->d_fb, hit_ids ? s->d_hit : nullptr, nullptr, 0));
    CU(cudaEventRecord(s->ev1, 0));
    if (alias) CU(cudaStreamSynchronize(0));
    else CU(cudaMemcpy(fb, s->d_fb, n * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    if (hit_ids) CU(cudaMemcpy(hit_ids, s->d_hit, n * 4 * sizeof(int32_t), cudaMemcpyDeviceToHost));
    if (device_ms) CU(cudaEventElapsedTime(device_ms, s->ev0, s->ev1));
    return RRT_OK;
}

int rrt_render_frames(rrt_scene* s, const rrt_render_params* params, int32_t n_frames, uint32_t* const* fbs,
                      float* total_ms) {
    if (!s || !params || !fbs || n_frames < 0) return fail(RRT_ERR_INVALID, "bad argument");
    if (n_frames == 0) { if (total_ms) *total_ms = 0.0f; return RRT_OK; }
    for (int32_t i = 0; i < n_frames; ++i) {
        if (!fbs[i]) return fail(RRT_ERR_INVALID, "null frame buffer");
        if (params[i].width != params[0].width || params[i].height != params[0].height)
            return fail(RRT_ERR_INVALID, "rrt_render_frames: all frames of a sequence have one resolution");
    }
    CU(cudaSetDevice(s->device));
    if (!s->seq_render) {
        CU(cudaStreamCreateWithFlags(&s->seq_render, cudaStreamNonBlocking));
        CU(cudaStreamCreateWithFlags(&s->seq_copy, cudaStreamNonBlocking));
        for (int b = 0; b < 2; ++b) {
            CU(cudaEventCreateWithFlags(&s->seq_rendered[b], cudaEventDisableTiming));
            CU(cudaEventCreateWithFlags(&s->seq_copied[b], cudaEventDisableTiming));
        }
        CU(cudaEventCreate(&s->seq_t0));
        CU(cudaEventCreate(&s->seq_t1));
    }
    // enqueue everything; whatever happens, nothing may still be writing to the caller's buffers on return
    auto enqueue = [&]() -> int {
        int rc;
        CU(cudaEventRecord(s->seq_t0, s->seq_render));
        for (int32_t i = 0; i < n_frames; ++i) {
            DevParams P;
            if ((rc = make_params(s, params + i, nullptr, &P))) return rc;
            const size_t n = (size_t)P.W * P.H;
            const int b = i & 1;
            if ((rc = ensure(&s->d_scratch, &s->scratch_cap, render_scratch_words(P.n_rows_local, P.W)))) return rc;
            if (uint32_t* alias = host_frame_alias(fbs[i], n)) {
                // page-locked destination: the kernel stores frame i into it itself; nothing to copy, nothing to wait for
                CU(render_frame(s, P, alias, nullptr, nullptr, s->seq_render));
                CU(cudaEventRecord(s->seq_rendered[b], s->seq_render));
                CU(cudaStreamWaitEvent(s->seq_copy, s->seq_rendered[b], 0));   // keeps seq_t1 (on the copy stream) behind it
                CU(cudaEventRecord(s->seq_copied[b], s->seq_copy));
                continue;
            }
            if ((rc = ensure(&s->d_seq_fb[b], &s->seq_cap[b], n))) return rc;   // no-op after the first two frames
            if (i >= 2) CU(cudaStreamWaitEvent(s->seq_render, s->seq_copied[b], 0));   // frame i-2 has left this buffer
            CU(render_frame(s, P, s->d_seq_fb[b], nullptr, nullptr, s->seq_render));
            CU(cudaEventRecord(s->seq_rendered[b], s->seq_render));
            CU(cudaStreamWaitEvent(s->seq_copy, s->seq_rendered[b], 0));
            CU(cudaMemcpyAsync(fbs[i], s->d_seq_fb[b], n * sizeof(uint32_t), cudaMemcpyDeviceToHost, s->seq_copy));
            CU(cudaEventRecord(s->seq_copied[b], s->seq_copy));
        }
        CU(cudaEventRecord(s->seq_t1, s->seq_copy));
        return RRT_OK;
    };
    const int rc = enqueue();
    const cudaError_t e_render = cudaStreamSynchronize(s->seq_render), e_copy = cudaStreamSynchronize(s->seq_copy);
    if (rc) return rc;
    CU(e_render);
    CU(e_copy);
    if (total_ms) CU(cudaEventElapsedTime(total_ms, s->seq_t0, s->seq_t1));
    return RRT_OK;
}

int rrt_render_counters(rrt_scene* s, const rrt_render_params* p, uint32_t* fb, rrt_counters* out) {
    if (!s || !out) return fail(RRT_ERR_INVALID, "null argument");
    DevParams P;
    int rc = make_params(s, p, nullptr, &P);
    if (rc) return rc;
    CU(cudaSetDevice(s->device));
    const size_t n = (size_t)P.W * P.H;
    if ((rc = ensure(&s->d_fb, &s->fb_cap, n))) return rc;
    DevBuf c;
    if ((rc = c.alloc(sizeof(DevCounters)))) return rc;
    CU(cudaMemset(c.p, 0, sizeof(DevCounters)));
    if ((rc = ensure(&s->d_scratch, &s->scratch_cap, render_scratch_words(P.n_rows_local, P.W)))) return rc;
    CU(render_frame(s, P, s->d_fb, nullptr, (DevCounters*)c.p, 0));
    DevCounters h;
    CU(cudaMemcpy(&h, c.p, sizeof h, cudaMemcpyDeviceToHost));
    if (fb) CU(cudaMemcpy(fb, s->d_fb, n * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    out->rays_primary = h.rays_primary; out->rays_shadow = h.rays_shadow; out->rays_reflect = h.rays_reflect;
    out->mt_tests = h.mt_tests; out->box_tests = h.box_tests; out->cull_tests = h.cull_tests;
    out->nodes_entered = h.nodes_entered;
    return RRT_OK;
}

int rrt_deinterleave_device(int device, const uint32_t* d_gathered, int32_t rows_per_block, int32_t width,
                            int32_t height, int32_t tile_rows, int32_t n_shards, uint32_t* d_fb, void* cuda_stream) {
    if (!d_gathered || !d_fb) return fail(RRT_ERR_INVALID, "null argument");
    if (width <= 0 || height <= 0 || tile_rows < 1 || n_shards < 1 || rows_per_block < 1)
        return fail(RRT_ERR_INVALID, "bad geometry");
    CU(cudaSetDevice(device));
    CU(launch_deinterleave(d_gathered, rows_per_block, width, height, tile_rows, n_shards, d_fb, (cudaStream_t)cuda_stream));
    return RRT_OK;
}

int rrt_trace_rays(rrt_scene* s, uint64_t n, const double* o, const double* d, const double* max_t, int32_t* out_tri,
                   double* out_tuv) {
    if (!s || (n && (!o || !d || !out_tri))) return fail(RRT_ERR_INVALID, "null argument");
    if (n == 0) return RRT_OK;
    CU(cudaSetDevice(s->device));
    DevBuf bo, bd, bm, bt, buv;
    int rc;
    if ((rc = bo.alloc(n * 24)) || (rc = bd.alloc(n * 24)) || (rc = bt.alloc(n * 4)) || (rc = buv.alloc(n * 24))) return rc;
    if (max_t && (rc = bm.alloc(n * 8))) return rc;
    CU(cudaMemcpy(bo.p, o, n * 24, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(bd.p, d, n * 24, cudaMemcpyHostToDevice));
    if (max_t) CU(cudaMemcpy(bm.p, max_t, n * 8, cudaMemcpyHostToDevice));
    CU(cudaMemset(buv.p, 0, n * 24));
    CU(launch_trace(s->dev, n, (const double*)bo.p, (const double*)bd.p, max_t ? (const double*)bm.p : nullptr,
                    (int32_t*)bt.p, (double*)buv.p, s->search_mode != 0, 0));
    CU(cudaMemcpy(out_tri, bt.p, n * 4, cudaMemcpyDeviceToHost));
    if (out_tuv) {
        // "untouched on miss": copy through a staging vector and merge
        std::vector<double> tmp(3 * n);
        CU(cudaMemcpy(tmp.data(), buv.p, n * 24, cudaMemcpyDeviceToHost));
        for (uint64_t i = 0; i < n; ++i)
            if (out_tri[i] >= 0) std::memcpy(out_tuv + 3 * i, tmp.data() + 3 * i, 24);
    }
    return RRT_OK;
}

int rrt_ray_colours(rrt_scene* s, const rrt_render_params* p, uint64_t n, const double* o, const double* d,
                    uint8_t* out_rgb) {
    if (!s || (n && (!o || !d || !out_rgb))) return fail(RRT_ERR_INVALID, "null argument");
    DevParams P;
    int rc = make_params(s, p, nullptr, &P);
    if (rc) return rc;
    if (n == 0) return RRT_OK;
    CU(cudaSetDevice(s->device));
    DevBuf bo, bd, bc;
    if ((rc = bo.alloc(n * 24)) || (rc = bd.alloc(n * 24)) || (rc = bc.alloc(n * 3))) return rc;
    CU(cudaMemcpy(bo.p, o, n * 24, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(bd.p, d, n * 24, cudaMemcpyHostToDevice));
    CU(launch_colours(s->dev, P, n, (const double*)bo.p, (const double*)bd.p, (uint8_t*)bc.p, 0));
    CU(cudaMemcpy(out_rgb, bc.p, n * 3, cudaMemcpyDeviceToHost));
    return RRT_OK;
}

// ---------------------------------------------------------------- host octree inspection (no GPU)
struct rrt_octree {
    HostOctree t;
};

int rrt_octree_build(const double* v, uint64_t n_tris, const double root_min[3], const double root_max[3],
                     rrt_octree** out) {
    if (!out || (n_tris && !v) || !root_min || !root_max) return fail(RRT_ERR_INVALID, "null argument");
    if (n_tris > 0x7FFFFFF0ull) return fail(RRT_ERR_INVALID, "too many triangles");
    auto t = std::make_unique<rrt_octree>();
    std::string err;
    if (!build_tree(t->t, v, n_tris, root_min, root_max, &err)) return fail(RRT_ERR_DEPTH, err);
    *out = t.release();
    return RRT_OK;
}
void rrt_octree_destroy(rrt_octree* t) { delete t; }
uint64_t rrt_octree_node_count(const rrt_octree* t) { return t ? t->t.nodes.size() : 0; }
uint64_t rrt_octree_digest(const rrt_octree* t) { return t ? t->t.digest() : 0; }
int rrt_octree_node(const rrt_octree* t, uint64_t i, double box[6], uint64_t* triangle_count, uint64_t* first_child,
                    uint64_t* list_len) {
    if (!t || i >= t->t.nodes.size()) return fail(RRT_ERR_INVALID, "node index out of range");
    const HostNode& n = t->t.nodes[i];
    if (box) std::memcpy(box, n.box, sizeof n.box);
    if (triangle_count) *triangle_count = n.triangle_count;
    if (first_child) *first_child = n.first_child;
    if (list_len) *list_len = n.list_len;
    return RRT_OK;
}
int rrt_octree_flatten_stats(const rrt_octree* t, const double* v, rrt_scene_stats* out, double* seconds) {
    if (!t || !out || (t->t.n_tris && !v)) return fail(RRT_ERR_INVALID, "null argument");
    const auto t0 = std::chrono::steady_clock::now();
    FlatScene flat;
    flatten(t->t, v, cull_threshold(), cull_leaf(), &flat, !(std::getenv("RRT_SEARCH_MODE") && std::atoi(std::getenv("RRT_SEARCH_MODE")) == 0));
    if (seconds) *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    std::memset(out, 0, sizeof *out);
    out->n_tris = t->t.n_tris;
    out->n_tris_stored = flat.n_stored;
    out->n_tris_dropped = t->t.n_dropped;
    out->n_nodes = t->t.nodes.size();
    out->n_nodes_nonempty = (t->t.nodes[0].triangle_count > 0) ? flat.nodes.size() : 0;
    out->max_depth = t->t.max_depth;
    out->root_list = flat.root_list;
    out->max_list = flat.max_list;
    out->n_cull_nodes = flat.cull.size();
    out->device_bytes = flat.nodes.size() * sizeof(DevNode) + flat.tris.size() * sizeof(DevTri) + flat.cull.size() * sizeof(DevCull);
    // order-independent digest of the flattened arrays, so tests can compare builds (e.g. 1 thread vs many)
    uint64_t h = 1469598103934665603ull;
    auto mix = [&h](const void* p, size_t n) {
        const unsigned char* b = static_cast<const unsigned char*>(p);
        for (size_t i = 0; i < n; ++i) { h ^= b[i]; h *= 1099511628211ull; }
    };
    mix(flat.nodes.data(), flat.nodes.size() * sizeof(DevNode));
    mix(flat.tris.data(), flat.tris.size() * sizeof(DevTri));
    mix(flat.cull.data(), flat.cull.size() * sizeof(DevCull));
    mix(flat.sub_boxes.data(), flat.sub_boxes.size() * sizeof(DevSubBox));
    out->flat_digest = h;
    return RRT_OK;
}
int64_t rrt_octree_node_list(const rrt_octree* t, uint64_t i, uint32_t* out, uint64_t cap) {
    if (!t || i >= t->t.nodes.size()) return -1;
    uint64_t k = 0;
    for (uint32_t tr = t->t.nodes[i].head; tr != UINT32_MAX; tr = t->t.next[tr], ++k)
        if (out && k < cap) out[k] = tr;
    return (int64_t)k;
}

}  // extern "C"

// emu_trace.cpp — TEST INFRASTRUCTURE ONLY (built and loaded by tests/test_host_emulation.py, nothing else).
//
// Compiles the PRODUCT's traversal headers (csrc/rrt_device.cuh, csrc/rrt_trace.cuh: short-stack walk, FMA
// quotients, f32 culling hierarchies, list / subtree boxes, division-free Moller-Trumbore reject) for the HOST,
// against the product's own host octree + flattener (csrc/rrt_octree.cpp), so that the claim "this walk returns
// what the reference recursion returns"  can be fuzzed against the oracle on millions of rays and
// thousands of random scenes without a GPU.  It is not a CPU fallback: the library (librrt_b200.so) neither
// contains nor loads it, and rendering without a CUDA device fails (tests/test_abi.py).
// Arithmetic: the same IEEE double / float operations — g++ -O2 -ffp-contract=off -mfma, fma()/fmaf() are the
// hardware's single-rounding instructions exactly like DFMA / FFMA, fmin/fmax/fminf/fmaxf follow the same
// "NaN loses" rule as CUDA's.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include <cuda_runtime.h>   // vector types (double2, float4, uint4, ...) and the empty host-side __device__ macros

// ---- the handful of device intrinsics the headers use, restated for the host
template <class T> static inline T __ldg(const T* p) { return *p; }
static inline int __popc(unsigned x) { return __builtin_popcount(x); }
static inline int __ffs(int x) { return __builtin_ffs(x); }
static inline int __clz(int x) { return x ? __builtin_clz((unsigned)x) : 32; }
static inline double __longlong_as_double(long long v) { double d; std::memcpy(&d, &v, 8); return d; }
static inline long long __double_as_longlong(double d) { long long v; std::memcpy(&v, &d, 8); return v; }
static inline float __int_as_float(int v) { float f; std::memcpy(&f, &v, 4); return f; }
static inline int __float_as_int(float f) { int v; std::memcpy(&v, &f, 4); return v; }
static inline float __double2float_rn(double d) { return (float)d; }   // default rounding mode = nearest even
static inline float __double2float_ru(double d) {                      // round toward +inf
    float f = (float)d;
    if ((double)f < d) f = std::nextafterf(f, INFINITY);
    return f;                                                          // NaN stays NaN, +-inf stay
}
static inline int __double2hiint(double d) { return (int)(__double_as_longlong(d) >> 32); }
static inline int __double2loint(double d) { return (int)(__double_as_longlong(d) & 0xFFFFFFFFll); }
#ifndef __noinline__
#define __noinline__ __attribute__((noinline))
#endif
using std::fabs; using std::fma; using std::fmax; using std::fmin; using std::sqrt; using std::pow;

#include "../../rust-ray-tracer_b200/csrc/rrt_raygen.cuh"   // pulls in rrt_shade.cuh, rrt_trace.cuh and rrt_device.cuh

namespace {
struct EmuScene {
    rrt::FlatScene flat;
    rrt::DevScene dev{};
    uint64_t n_nodes_host = 0;
    // shading side (emu_scene_set_shading): packed exactly as rrt_scene_create packs it (csrc/rrt_api.cu)
    std::vector<rrt::DevShade> shade;
    std::vector<rrt::DevMaterial> materials;
    std::vector<rrt::DevTexture> tex;
    std::vector<std::vector<uchar4>> texels;
};
}  // namespace

// The fields of DevParams derived as make_params derives them (csrc/rrt_api.cu: origin, viewport / size, lights,
// max_depth clamp, surface offset, rotation, shard and its row count as rrt_shard_rows counts them).
static rrt::DevParams emu_params(const rrt_render_params* rp, int tile_rows, int n_shards, int shard_index) {
    rrt::DevParams P{};
    P.W = rp->width; P.H = rp->height;
    P.ox = rp->origin[0]; P.oy = rp->origin[1]; P.oz = rp->origin[2];
    P.xs = rp->viewport[0] / (double)rp->width;
    P.ys = rp->viewport[1] / (double)rp->height;
    P.zv = rp->viewport[2];
    P.n_lights = rp->n_lights;
    P.max_depth = rp->max_depth > rrt::kMaxReflectDepth ? rrt::kMaxReflectDepth : rp->max_depth;
    for (int i = 0; i < rp->n_lights; ++i) P.lights[i] = rp->lights[i];
    P.offset = rp->surface_offset;
    P.use_rot = rp->use_camera_rot;
    for (int i = 0; i < 9; ++i) P.rot[i] = rp->camera_rot[i];
    P.tile_rows = tile_rows; P.n_shards = n_shards; P.shard_index = shard_index;
    int rows = 0;
    if (n_shards <= 1) rows = rp->height;
    else for (int j = shard_index; (int64_t)j * tile_rows < rp->height; j += n_shards) rows += std::min(tile_rows, rp->height - j * tile_rows);
    P.n_rows_local = rows;
    return P;
}

// what traverse() may try before the walk (rrt_trace.cuh: kTryFilter | kTrySparse); tests switch it to 0 = walk only
static uint32_t g_allow = rrt::kTryFilter | rrt::kTrySparse;

extern "C" {

void emu_set_allow(uint32_t allow) { g_allow = allow; }

// v: n x 9 doubles; slow_div != 0 forces the true-division slab path (RRT_SLOW_DIV of the library)
int emu_scene_create(const double* v, uint64_t n_tris, const double* root_min, const double* root_max,
                     uint32_t cull_threshold, uint32_t cull_leaf, int parallel_build, int slow_div, void** out) {
    auto s = new EmuScene;
    rrt::HostOctree tree;
    std::string err;
    const bool ok = parallel_build ? tree.build_parallel(v, n_tris, root_min, root_max, &err)
                                   : tree.build(v, n_tris, root_min, root_max, &err);
    if (!ok) { delete s; return 1; }
    rrt::flatten(tree, v, cull_threshold, cull_leaf, &s->flat);
    s->n_nodes_host = tree.nodes.size();
    s->dev.nodes = s->flat.nodes.data();
    s->dev.tris = s->flat.tris.data();
    s->dev.cull = s->flat.cull.data();
    s->dev.sub_boxes = s->flat.sub_boxes.data();
    s->dev.n_nodes = (uint32_t)s->flat.nodes.size();
    s->dev.root_tri_count = s->flat.n_stored ? 1u : 0u;                       // as rrt_scene_create (rrt_api.cu)
    s->dev.planes_safe = (s->flat.planes_safe && !slow_div) ? 1u : 0u;
    s->dev.global_root = s->flat.global_root;
    s->dev.sparse_ok = s->flat.sparse_ok ? 1u : 0u;
#if RRT_CULL_F32
    s->dev.cull_o_max = rrt::kCullOriginRel * s->flat.scale;
#else
    s->dev.cull_o_max = 0.0;
#endif
    *out = s;
    return 0;
}

void emu_scene_destroy(void* p) { delete static_cast<EmuScene*>(p); }

uint64_t emu_scene_stored(void* p) { return static_cast<EmuScene*>(p)->flat.n_stored; }
int emu_scene_planes_safe(void* p) { return (int)static_cast<EmuScene*>(p)->dev.planes_safe; }

// shading records, materials and RGB8 textures, packed as rrt_scene_create does (csrc/rrt_api.cu: shading records by
// original triangle id, DevMaterial field by field, texels widened to RGBX)
int emu_scene_set_shading(void* p, uint64_t n_tris, const double* vt, const double* vn, const int32_t* material_id,
                          uint32_t n_materials, const rrt_material* materials, uint32_t n_textures,
                          const rrt_texture* textures) {
    EmuScene* s = static_cast<EmuScene*>(p);
    s->shade.resize(n_tris);
    for (uint64_t i = 0; i < n_tris; ++i) {
        const double *n = vn + 9 * i, *t = vt + 6 * i;
        rrt::DevShade& sh = s->shade[i];
        std::memcpy(sh.n1, n, 24); std::memcpy(sh.n2, n + 3, 24); std::memcpy(sh.n3, n + 6, 24);
        std::memcpy(sh.t1, t, 16); std::memcpy(sh.t2, t + 2, 16); std::memcpy(sh.t3, t + 4, 16);
        sh.material = material_id[i];
        sh._pad = 0;
    }
    s->materials.resize(n_materials);
    for (uint32_t i = 0; i < n_materials; ++i) {
        const rrt_material& m = materials[i];
        rrt::DevMaterial& o = s->materials[i];
        std::memcpy(o.ka, m.ka, 24); std::memcpy(o.kd, m.kd, 24); std::memcpy(o.ks, m.ks, 24);
        o.ns = m.ns; o.kr = m.kr; o.tex = m.tex_id; o.bump = m.bump_id;
    }
    s->tex.resize(n_textures);
    s->texels.resize(n_textures);
    for (uint32_t i = 0; i < n_textures; ++i) {
        const rrt_texture& t = textures[i];
        const size_t n = (size_t)t.width * t.height;
        s->texels[i].resize(n);
        for (size_t k = 0; k < n; ++k) s->texels[i][k] = make_uchar4(t.rgb[3 * k], t.rgb[3 * k + 1], t.rgb[3 * k + 2], 0);
        s->tex[i] = rrt::DevTexture{s->texels[i].data(), (uint32_t)t.width, (uint32_t)t.height};
    }
    s->dev.shade = s->shade.data();
    s->dev.materials = s->materials.data();
    s->dev.tex = s->tex.data();
    return 0;
}

// RayTracer::get_ray_colour for n rays through the product's RayStateT state machine (what colour_kernel and
// render_kernel run per lane).  The fields of DevParams a single ray reads are derived as make_params derives them
// (csrc/rrt_api.cu: origin, viewport / size, lights, max_depth clamp, surface offset).
int emu_ray_colours(void* p, const rrt_render_params* rp, uint64_t n, const double* o, const double* d, uint8_t* out_rgb) {
    const EmuScene* s = static_cast<EmuScene*>(p);
    if (!s->dev.shade) return 1;
    const rrt::DevParams P = emu_params(rp, 1, 1, 0);
#pragma omp parallel for schedule(dynamic, 256)
    for (long long i = 0; i < (long long)n; ++i) {
        rrt::RayState st;
        rrt::TraceCount tc;
        rrt::Hit h;
        st.begin(rrt::V3{o[3 * i], o[3 * i + 1], o[3 * i + 2]}, rrt::V3{d[3 * i], d[3 * i + 1], d[3 * i + 2]});
        do {
            rrt::traverse<false>(s->dev, st.so, st.sd, st.smax, st.shadow, h, tc, g_allow);
        } while (!st.advance<false>(s->dev, P, h));
        const uint32_t c = st.res;
        out_rgb[3 * i] = (uint8_t)(c & 255u); out_rgb[3 * i + 1] = (uint8_t)((c >> 8) & 255u); out_rgb[3 * i + 2] = (uint8_t)((c >> 16) & 255u);
    }
    return 0;
}

// One shard of one frame the way render_kernel produces it, minus the warps: every position of the sample queue is
// decoded with the kernel's own sample_pixel / buffer_row / primary_ray, traced and shaded through RayStateT, and merged
// through the kernel's packed per-pixel accumulator (sample_add / mix_pack).  full_frame != 0: fb is the whole frame and
// owned rows go to their buffer rows (rrt_render_device_into_frame); else fb holds the shard's rows compactly.
// hit_ids (may be NULL): [rows][W][4] in the same row numbering as fb... of the COMPACT rows, as the kernel writes them.
int emu_render_shard(void* p, const rrt_render_params* rp, int tile_rows, int n_shards, int shard_index, int full_frame,
                     uint32_t* fb, int32_t* hit_ids) {
    const EmuScene* s = static_cast<EmuScene*>(p);
    if (!s->dev.shade) return 1;
    rrt::DevParams P = emu_params(rp, tile_rows, n_shards, shard_index);
    P.full_frame = full_frame;
    const int tiles_x = (P.W + rrt::kTileW - 1) / rrt::kTileW, tiles_y = (P.n_rows_local + rrt::kTileH - 1) / rrt::kTileH;
    const long long n_tiles = (long long)tiles_x * tiles_y * 8;
    std::vector<uint32_t> acc((size_t)P.n_rows_local * P.W, 0u);
#pragma omp parallel for schedule(dynamic, 16)
    for (long long tile = 0; tile < n_tiles; ++tile) {
        for (unsigned l = 0; l < 32; ++l) {
            int col, krow;
            unsigned k;
            rrt::sample_pixel((unsigned long long)(tile >> 3), (unsigned)tile & 7u, l, tiles_x, col, krow, k);
            if (!(col < P.W && krow < P.n_rows_local)) continue;
            const size_t pix = (size_t)krow * P.W + col;
            const int r = rrt::buffer_row(P, krow);
            const size_t out = P.full_frame ? (size_t)r * P.W + col : pix;
            rrt::V3 d;
            uint32_t add;
            if (rrt::primary_ray(P, col, r, k, d)) {
                rrt::RayState st;
                rrt::TraceCount tc;
                rrt::Hit h;
                        st.begin(rrt::V3{P.ox, P.oy, P.oz}, d);
                do {
                    rrt::traverse<false>(s->dev, st.so, st.sd, st.smax, st.shadow, h, tc, g_allow);
                } while (!st.advance<false>(s->dev, P, h));
                if (hit_ids) hit_ids[pix * 4 + k] = st.hit0;
                add = rrt::sample_add(st.res);
            } else {
                if (hit_ids) hit_ids[pix * 4 + k] = -1;
                add = 1u << 30;
            }
            // the four samples of a pixel sit in one tile, i.e. in this thread: no atomics needed here
            const uint32_t old = acc[pix];
            acc[pix] = old + add;
            if ((old >> 30) == 3u) fb[out] = (add == (1u << 30)) ? 0u : rrt::mix_pack(old + add);
        }
    }
    return 0;
}

// work counters of the walk, summed over n rays: out[4] = exact M-T tests, child-slab tests, culling-box tests, nodes
int emu_trace_counts(void* p, uint64_t n, const double* o, const double* d, int any_hit, uint64_t* out) {
    const EmuScene* s = static_cast<EmuScene*>(p);
    uint64_t mt = 0, box = 0, cull = 0, nodes = 0;
#pragma omp parallel for schedule(dynamic, 256) reduction(+ : mt, box, cull, nodes)
    for (long long i = 0; i < (long long)n; ++i) {
        rrt::Hit h;
        rrt::TraceCount tc;
        rrt::traverse<true>(s->dev, rrt::V3{o[3 * i], o[3 * i + 1], o[3 * i + 2]}, rrt::V3{d[3 * i], d[3 * i + 1], d[3 * i + 2]},
                            __longlong_as_double(0x7FF0000000000000ll), any_hit != 0, h, tc, g_allow);
        mt += tc.mt; box += tc.box; cull += tc.cull; nodes += tc.nodes;
    }
    out[0] = mt; out[1] = box; out[2] = cull; out[3] = nodes;
    return 0;
}

// Ray::intersect_with_octant_with_max_t for n rays through the product's traverse(); out_tuv as rrt_trace_rays
int emu_trace_rays(void* p, uint64_t n, const double* o, const double* d, const double* max_t, int any_hit,
                   int32_t* out_tri, double* out_tuv) {
    const EmuScene* s = static_cast<EmuScene*>(p);
#pragma omp parallel for schedule(dynamic, 256)
    for (long long i = 0; i < (long long)n; ++i) {
        rrt::Hit h;
        rrt::TraceCount tc;
        const rrt::V3 oo{o[3 * i], o[3 * i + 1], o[3 * i + 2]}, dd{d[3 * i], d[3 * i + 1], d[3 * i + 2]};
        const double mt = max_t ? max_t[i] : __longlong_as_double(0x7FF0000000000000ll);
        rrt::traverse<false>(s->dev, oo, dd, mt, any_hit != 0, h, tc, g_allow);
        out_tri[i] = (h.id == rrt::kNone) ? -1 : (int32_t)h.id;
        if (out_tuv) {
            out_tuv[3 * i] = out_tuv[3 * i + 1] = out_tuv[3 * i + 2] = 0.0;
            if (h.id != rrt::kNone) {
                double u, v;
                rrt::hit_uv(oo, dd, s->dev.tris + h.rec, u, v);
                out_tuv[3 * i] = h.t; out_tuv[3 * i + 1] = u; out_tuv[3 * i + 2] = v;
            }
        }
    }
    return 0;
}

// what the searches did with n rays: out[0] = answered, out[1] = handed to the walk, out[2] = not tried
int emu_search_outcomes(void* p, uint64_t n, const double* o, const double* d, const double* max_t, int any_hit, uint64_t* out) {
    const EmuScene* s = static_cast<EmuScene*>(p);
    uint64_t a = 0, h2 = 0, z = 0;
#pragma omp parallel for schedule(dynamic, 256) reduction(+ : a, h2, z)
    for (long long i = 0; i < (long long)n; ++i) {
        rrt::Hit h;
        rrt::TraceCount tc;
        uint32_t oc = 0;
        const double mt = max_t ? max_t[i] : __longlong_as_double(0x7FF0000000000000ll);
        rrt::traverse<false>(s->dev, rrt::V3{o[3 * i], o[3 * i + 1], o[3 * i + 2]}, rrt::V3{d[3 * i], d[3 * i + 1], d[3 * i + 2]},
                             mt, any_hit != 0, h, tc, g_allow, &oc);
        a += oc == 1u; h2 += oc == 2u; z += oc == 0u;
    }
    out[0] = a; out[1] = h2; out[2] = z;
    return 0;
}

// Closest accepted record of the WHOLE scene through the scene-wide culling hierarchy (flatten's global hierarchy),
// ordered walk with limit pruning, ties to the smaller triangle id — i.e. a plain closest-hit tracer over the product's
// records.  Development aid for the round-2 design note (profiles/r01_experiments.md): how often is that NOT what the
// reference's octree walk returns?  out_tri[i] = id or -1; returns 1 when the scene has no scene-wide hierarchy.
int emu_closest_global(void* p, uint64_t n, const double* o, const double* d, int32_t* out_tri) {
    const EmuScene* s = static_cast<EmuScene*>(p);
    const uint32_t groot = s->dev.global_root;
    if (groot == 0u) return 1;
#pragma omp parallel for schedule(dynamic, 256)
    for (long long i = 0; i < (long long)n; ++i) {
        const rrt::V3 oo{o[3 * i], o[3 * i + 1], o[3 * i + 2]}, dd{d[3 * i], d[3 * i + 1], d[3 * i + 2]};
        const rrt::AxisRay ax = rrt::make_axis(oo.x, dd.x, false), ay = rrt::make_axis(oo.y, dd.y, false),
                           az = rrt::make_axis(oo.z, dd.z, false);
        const rrt::CullRay cr = rrt::make_cull_ray(s->dev.cull_o_max, oo, dd, ax, ay, az);
        rrt::Hit nb;
        nb.id = rrt::kNone; nb.rec = 0; nb.t = __longlong_as_double(0x7FF0000000000000ll);
        float limit = __int_as_float(0x7f800000);
        uint32_t stk[rrt::kMaxCullDepth + 2];
        int sp = 0;
        uint32_t cur = groot - 1u;
        for (;;) {
            if (cur & rrt::kLeafBit) {
                if (rrt::test_record(oo, dd, s->dev.tris, cur & 0x0FFFFFFFu, nb)) limit = __double2float_ru(nb.t);
                if (sp == 0) break;
                cur = stk[--sp];
            } else {
                const rrt::DevCull& c = s->dev.cull[cur];
                float tn0, tn1;
                const bool go0 = rrt::cull_box(cr, c.box[0][0], c.box[0][1], c.box[0][2], c.box[0][3], c.box[0][4], c.box[0][5], limit, tn0);
                const bool go1 = rrt::cull_box(cr, c.box[1][0], c.box[1][1], c.box[1][2], c.box[1][3], c.box[1][4], c.box[1][5], limit, tn1);
                const bool swap = tn1 < tn0;
                if (go0 && go1) { stk[sp++] = swap ? c.child[0] : c.child[1]; cur = swap ? c.child[1] : c.child[0]; }
                else if (go0 || go1) cur = go0 ? c.child[0] : c.child[1];
                else if (sp != 0) cur = stk[--sp];
                else break;
            }
        }
        out_tri[i] = nb.id == rrt::kNone ? -1 : (int32_t)nb.id;
    }
    return 0;
}

}  // extern "C"

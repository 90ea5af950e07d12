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

#include "../../rust-ray-tracer_b200/csrc/rrt_trace.cuh"

namespace {
struct EmuScene {
    rrt::FlatScene flat;
    rrt::DevScene dev{};
    uint64_t n_nodes_host = 0;
};
}  // namespace

extern "C" {

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
        rrt::traverse<false>(s->dev, oo, dd, mt, any_hit != 0, h, tc);
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

}  // extern "C"

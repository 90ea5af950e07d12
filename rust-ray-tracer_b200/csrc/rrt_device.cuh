// rrt_device.cuh — device-side restatement of the reference's per-ray arithmetic for sm_100a.
// Every expression keeps the reference's association order; the translation unit is compiled with
// -fmad=false -prec-div=true -prec-sqrt=true (Rust never fuses a*b+c).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/rrt.h"
#include "rrt_octree.h"

namespace rrt {

constexpr int kMaxReflectDepth = 16;   // params.max_depth is clamped to this (reference: 5, raytracer.rs:20)
constexpr double kEps = 2.220446049250313e-16;  // f64::EPSILON (ray.rs:66,89)

struct DevMaterial {     // material.rs:11-22
    double ka[3], kd[3], ks[3];
    double ns, kr;
    int32_t tex, bump;   // indices into DevScene::tex; bump = -1 for None
};
struct DevTexture {      // entities.rs:87-91, texels widened to RGBX for aligned 4-byte loads
    const uchar4* texels;
    uint32_t width, height;
};

// Everything the kernels read; lives in HBM, passed by value (pointers only).
struct DevScene {
    const DevNode* nodes;
    const DevTri* tris;
    const DevCull* cull;
    const DevSubBox* sub_boxes;
    const DevShade* shade;
    const DevMaterial* materials;
    const DevTexture* tex;
    uint32_t n_nodes;
    uint32_t root_tri_count;   // nodes[0].triangle_count analogue: 0 => every ray misses (ray.rs:112-114)
    uint32_t planes_safe;      // octree planes satisfy the range/order conditions of the FMA quotient path
    double cull_o_max;         // f32 culling: rays with an |origin component| above this are not culled
    uint32_t global_root;      // 1 + root of the scene-wide culling hierarchy (FlatScene::global_root), 0 = none
    uint32_t sparse_ok;        // FlatScene::sparse_ok
};

struct DevCounters {
    unsigned long long rays_primary, rays_shadow, rays_reflect, mt_tests, box_tests, cull_tests, nodes_entered;
};

// ---------------------------------------------------------------- Vector3d (engine.rs:9-104)
struct V3 {
    double x, y, z;
};
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ V3 operator*(V3 a, double s) { return {a.x * s, a.y * s, a.z * s}; }
// IEEE division / square root as CALLS (shading only: cold code, ~40 / ~25 inlined instructions per site
// otherwise, and render_kernel is instruction-cache sensitive — profiles/r01_experiments.md).  Same
// correctly rounded operations, same bits.
#ifndef RRT_SHADE_CALLS
#define RRT_SHADE_CALLS 1
#endif
#if RRT_SHADE_CALLS
static __device__ __noinline__ double div_rn(double a, double b) { return a / b; }
static __device__ __noinline__ double sqrt_rn(double a) { return sqrt(a); }
static __device__ __noinline__ double pow_call(double a, double b) { return pow(a, b); }
#else
__device__ __forceinline__ double div_rn(double a, double b) { return a / b; }
__device__ __forceinline__ double sqrt_rn(double a) { return sqrt(a); }
__device__ __forceinline__ double pow_call(double a, double b) { return pow(a, b); }
#endif

__device__ __forceinline__ V3 operator/(V3 a, double s) { return {div_rn(a.x, s), div_rn(a.y, s), div_rn(a.z, s)}; }
__device__ __forceinline__ V3 operator-(V3 a) { return {-a.x, -a.y, -a.z}; }
__device__ __forceinline__ double dot(V3 a, V3 b) { return (a.x * b.x) + (a.y * b.y) + (a.z * b.z); }  // engine.rs:85-87
__device__ __forceinline__ double length(V3 a) { return sqrt_rn(a.x * a.x + a.y * a.y + a.z * a.z); }  // engine.rs:89-91
__device__ __forceinline__ V3 cross(V3 a, V3 b) {                                                       // engine.rs:93-99
    return {a.y * b.z - a.z * b.y, -(a.x * b.z - a.z * b.x), a.x * b.y - a.y * b.x};
}
__device__ __forceinline__ V3 normalised(V3 a) { return a / length(a); }                               // engine.rs:101-103

// Rust `as usize` (saturating, NaN -> 0) followed by `% n`   (raytracer.rs:52-53)
__device__ __forceinline__ uint32_t rust_usize_mod(double x, uint32_t n) {
    if (!(x > 0.0)) return 0u;
    if (x < 4294967296.0) return (uint32_t)x % n;
    if (x >= 18446744073709551616.0) return (uint32_t)(0xFFFFFFFFFFFFFFFFull % n);
    return (uint32_t)((unsigned long long)x % n);
}
// f64::clamp(0.0, 255.0) as u8 (raytracer.rs:98-107): NaN -> 0, truncation toward zero
__device__ __forceinline__ uint32_t rust_clamp_u8(double x) {
    if (x != x) return 0u;
    x = x < 0.0 ? 0.0 : x;
    x = x > 255.0 ? 255.0 : x;
    return (uint32_t)x;
}

// What a traversal carries: t, the original triangle id (tie order, hit-id plane) and the record index.
// u and v are NOT carried (4 registers per hit on a register-bound kernel): the closest-hit callers
// re-evaluate them for the one winning record with hit_uv() — the same instruction sequence on the same
// operands as the accepting test, hence the same bits.
struct Hit {
    double t;
    uint32_t id;      // original triangle index; UINT32_MAX = None
    uint32_t rec;     // index of the winning DevTri record
};

// ---------------------------------------------------------------- Moller-Trumbore (ray.rs:56-94)
// Exact, with a division-free early reject whose decision provably equals the reference's:
// the reference rejects when u = (1/a)*(s.h) is < 0 or > 1.  With x = sign(a)*(s.h), aa = |a|:
//   x < -1e-200 and aa < 1e100  =>  (1/a)*(s.h) is a negative number of magnitude > 1e-300 (no
//                                    underflow to -0)                          => u < 0  => reject
//   x > aa*(1+1e-12)            =>  the two roundings (1/a, then the product) lose < 3e-16 relative,
//                                    so the computed u is still > 1 (or +inf)   => reject
// Everything else (including every NaN) takes the reference's own instruction sequence.
// Returns true and fills t,u,v when the reference would return Some.
__device__ __forceinline__ bool moller_trumbore(const V3& o, const V3& d, const DevTri* __restrict__ rec,
                                                double& t, double& u, double& v, uint32_t& id) {
    const double2* p = reinterpret_cast<const double2*>(rec);
    const double2 r0 = __ldg(p + 0), r1 = __ldg(p + 1), r2 = __ldg(p + 2), r3 = __ldg(p + 3), r4 = __ldg(p + 4);
    const V3 v1{r0.x, r0.y, r1.x}, e1{r1.y, r2.x, r2.y}, e2{r3.x, r3.y, r4.x};
    const V3 h = cross(d, e2);                 // ray.rs:62
    const double a = dot(e1, h);               // ray.rs:64
    if (a > -kEps && a < kEps) return false;   // ray.rs:66
    const V3 s = o - v1;                       // ray.rs:72
    const double sh = dot(s, h);
    const double aa = fabs(a);
    const double x = a < 0.0 ? -sh : sh;
    if (x > aa * 1.000000000001) return false;        // certainly u > 1
    if (x < -1e-200 && aa < 1e100) return false;      // certainly u < 0
    const double f = 1.0 / a;                  // ray.rs:71 (inline: a call here forces spills in the hot loop, c4 +15 %)
    u = f * sh;                                // ray.rs:73
    if (u < 0.0 || u > 1.0) return false;      // ray.rs:75
    const V3 q = cross(s, e1);                 // ray.rs:79
    v = f * dot(d, q);                         // ray.rs:80
    if (v < 0.0 || u + v > 1.0) return false;  // ray.rs:82
    t = f * dot(e2, q);                        // ray.rs:87
    if (!(t > kEps)) return false;             // ray.rs:89
    id = __double_as_longlong(r4.y) & 0xFFFFFFFFll;
    return true;
}

// u, v of an ACCEPTED record (ray.rs:60-80), bit-identical to what moller_trumbore computed for it.
__device__ __forceinline__ void hit_uv(const V3& o, const V3& d, const DevTri* __restrict__ rec, double& u, double& v) {
    const double2* p = reinterpret_cast<const double2*>(rec);
    const double2 r0 = __ldg(p + 0), r1 = __ldg(p + 1), r2 = __ldg(p + 2), r3 = __ldg(p + 3), r4 = __ldg(p + 4);
    const V3 v1{r0.x, r0.y, r1.x}, e1{r1.y, r2.x, r2.y}, e2{r3.x, r3.y, r4.x};
    const V3 h = cross(d, e2);
    const double a = dot(e1, h);
    const double f = 1.0 / a;
    const V3 s = o - v1;
    u = f * dot(s, h);
    const V3 q = cross(s, e1);
    v = f * dot(d, q);
}

}  // namespace rrt

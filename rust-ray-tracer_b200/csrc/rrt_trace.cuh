// rrt_trace.cuh — octree traversal (ray.rs:96-168) as an iterative short-stack walk.
//
// Why the iterative form equals the reference recursion (re-derived in DESIGN.md):
// a node that accepts ANY own triangle returns Some, so every ancestor's child loop `break`s once that
// subtree is finished (ray.rs:155-160) — all pending siblings at every level are dead: the stack is
// cleared.  Its own children are still visited (in sorted order, stopping at the first that returns
// Some).  Ancestors are processed before descendants, so a strict `<` on the running best reproduces
// "child wins only if strictly closer" (ray.rs:163).  max_t bounds the ROOT's own scan only
// (ray.rs:117 vs :153 which recurses with +inf); the final answer must still beat max_t (ray.rs:163
// at the root), which `best.t` starting at max_t provides.
#pragma once
#include "rrt_device.cuh"

namespace rrt {

constexpr uint32_t kNone = 0xFFFFFFFFu;
constexpr uint32_t kLeafBit = 0x80000000u;  // cull entry: bit31 leaf | (count-1) << 28 | first record

struct TraceCount {
    uint32_t mt = 0, box = 0, cull = 0, nodes = 0;
};

// Development aid: tools/walk_stats.py compiles this header for the HOST with -DRRT_WALK_STATS and collects
// histograms of the walk (how many children a visited node keeps / the exact slabs hit / survive their subtree
// boxes, how long the culling walks are).  Never defined in a product build: the hooks expand to nothing.
#ifdef RRT_WALK_STATS
void rrt_walk_stat(int kind, unsigned value);
#define RRT_STAT(kind, value) rrt_walk_stat(kind, (unsigned)(value))
#else
#define RRT_STAT(kind, value) ((void)0)
#endif

// min/max for operands known not to be NaN (one compare + selects instead of the IEEE fmin/fmax
// sequence).  Which of +0/-0 comes back is irrelevant: results only feed comparisons.
__device__ __forceinline__ double min_nn(double a, double b) { return b < a ? b : a; }
__device__ __forceinline__ double max_nn(double a, double b) { return b > a ? b : a; }

// Per-ray, per-axis state for the child slab numerators (ray.rs:22-27).
//   fast: the correctly rounded quotient a/d is obtained from y = RN(1/d) (one true division per ray and
//         axis) by two FMA correction steps (div_markstein below) instead of a full division per plane.
//         The range conditions are established ONCE: |d| and |o| in [2^-400, 2^401) (o may also be 0),
//         checked here, and every octree plane in {0} U [2^-400, 2^401), checked at flatten time
//         (DevScene::planes_safe) — so a = plane - o is 0 or has magnitude in [2^-452, 2^402], every
//         quotient is normal and every residual a - d*q is representable (a multiple of 2^(e_a-104)).
//         The three planes of an axis are ordered p0 <= p1 <= p2 (also part of planes_safe), subtraction
//         and correctly rounded division are monotone, so q0 <= q1 <= q2 for d > 0 and q0 >= q1 >= q2
//         for d < 0: the min/max of ray.rs:29-36 over a pair of plane quotients is a SELECT on the sign
//         of d (equal quotients make either choice the same value).
//   otherwise (zero / tiny / huge / non-finite component): true divisions and NaN-aware min/max.
struct AxisRay {
    double o, d, y;
    bool fast, neg;
};

__device__ __forceinline__ bool safe_mag(double x, bool zero_ok) {
    const uint32_t e = ((uint32_t)__double2hiint(x) >> 20) & 0x7FFu;
    return (e - 623u) <= 800u || (zero_ok && x == 0.0);
}

__device__ __forceinline__ AxisRay make_axis(double o, double d, bool planes_safe) {
    AxisRay a;
    a.o = o;
    a.d = d;
    a.fast = planes_safe && safe_mag(d, false) && safe_mag(o, true);
    a.neg = d < 0.0;
    a.y = 1.0 / d;
    return a;
}

// RN(a/d) from y = RN(1/d), no over/underflow (see AxisRay).  q0 = RN(a*y) is within 1.5 ulp of a/d;
// one correction (r0 may need 54 bits, its rounding is harmless) lands within 1/2 ulp + 2^-51 ulp, i.e.
// q1 is a FAITHFUL rounding; for a faithful q1 the residual r1 = a - d*q1 is exact in one FMA and
// Markstein's theorem (Handbook of Floating-Point Arithmetic, FMA-based division: y correctly rounded,
// q faithful => RN(q + r*y) = RN(a/d)) makes q2 the IEEE quotient.  a == 0 gives +-0 (the sign of a zero
// never matters here: quotients only feed comparisons).  Verified against `/` by rrt_selftest_division.
__device__ __forceinline__ double div_markstein(double a, double d, double y) {
    const double q0 = a * y;
    const double q1 = fma(fma(-d, q0, a), y, q0);
    return fma(fma(-d, q1, a), y, q1);
}

// Slow path for one axis: the reference's own divisions and Rust's NaN-ignoring f64::min/max
// (ray.rs:22-36).  A NaN can only survive a pair when BOTH plane quotients are NaN (zero-extent box);
// it is then replaced by the identity of the enclosing max/min chain, which is what NaN-ignoring
// max(max(.,.),.) / min(min(.,.),.) amount to.
static __device__ __noinline__ void slab_axis_slow(double p0, double p1, double p2, double o, double d, double* out) {
    const double t0 = (p0 - o) / d, t1 = (p1 - o) / d, t2 = (p2 - o) / d;
    double mn0 = fmin(t0, t1), mn1 = fmin(t1, t2), mx0 = fmax(t0, t1), mx1 = fmax(t1, t2);
    const double ninf = __longlong_as_double(0xFFF0000000000000ll), pinf = __longlong_as_double(0x7FF0000000000000ll);
    out[0] = mn0 != mn0 ? ninf : mn0;
    out[1] = mn1 != mn1 ? ninf : mn1;
    out[2] = mx0 != mx0 ? pinf : mx0;
    out[3] = mx1 != mx1 ? pinf : mx1;
}

#ifndef RRT_SLAB_SIGN
#define RRT_SLAB_SIGN 1
#endif

// mn[h] / mx[h] = min / max of the two plane quotients bounding half h (0 = low, 1 = high) of one axis.
__device__ __forceinline__ void slab_axis(const AxisRay& a, double p0, double p1, double p2, double& mn0, double& mn1,
                                          double& mx0, double& mx1) {
    if (a.fast) {
        const double a0 = p0 - a.o, a1 = p1 - a.o, a2 = p2 - a.o;
        const double q0 = div_markstein(a0, a.d, a.y), q1 = div_markstein(a1, a.d, a.y), q2 = div_markstein(a2, a.d, a.y);
#if RRT_SLAB_SIGN
        mn0 = a.neg ? q1 : q0; mx0 = a.neg ? q0 : q1;
        mn1 = a.neg ? q2 : q1; mx1 = a.neg ? q1 : q2;
#else
        mn0 = min_nn(q0, q1); mx0 = max_nn(q0, q1);
        mn1 = min_nn(q1, q2); mx1 = max_nn(q1, q2);
#endif
    } else {
        double out[4];
        slab_axis_slow(p0, p1, p2, a.o, a.d, out);
        mn0 = out[0]; mn1 = out[1]; mx0 = out[2]; mx1 = out[3];
    }
}

// ---------------------------------------------------------------- conservative culling of long lists
// The hierarchy's boxes only decide which records get the exact test, so they are tested in FP32 (FFMA +
// FMNMX at full rate, 64-byte nodes) with everything rounded or padded to the safe side.
// Error budget, in units of S = FlatScene::scale (>= 1, >= every |coordinate| of a stored triangle):
//   * boxes are padded by 1e-5 S, then rounded outward to f32;
//   * a ray is ELIGIBLE when every |origin component| <= 16 S and max|direction component| is in
//     [1e-10, 1e10] (anything else — NaN included — is not culled at all: every record is tested);
//   * per plane t~ = fma(b, inv, -(o32*inv)), o32 = RN32(o), inv = RN32(1/d) clamped to +-1e25:
//     |o - o32| <= 2^-24 * 16 S, the product o32*inv and the fma each round once (2^-24 relative to
//     terms of magnitude <= 17 S |inv|), inv carries 2^-24 relative: in space units the plane moves by
//     less than 4.2e-6 S, under half of the pad.  A hit accepted by the exact test lies within 1e-12 S of
//     its triangle, hence (pad) strictly inside the computed slab interval on every axis:
//     tnear~ <= t* <= tfar~, which is all the three comparisons below use (t* > 0; `tnear~ > nb.t` proves
//     t* > nb.t, so an exact tie is never pruned).
//   * a direction component that is zero or below 1e-25 uses inv = +-1e25: the ray cannot leave a slab of
//     thickness >= 2 pad within t <= 17 S / 1e-10, and the computed interval is either astronomically
//     wide around 0 (origin inside) or astronomically far on one side (origin outside by > pad/2).
#if RRT_CULL_F32
struct CullRay {
    float ix, iy, iz, cx, cy, cz;   // inv and -(o32 * inv)
    bool ok;
};
__device__ __forceinline__ CullRay make_cull_ray(const DevScene& sc, const V3& o, const V3& d, const AxisRay& ax,
                                                 const AxisRay& ay, const AxisRay& az) {
    CullRay c;
    const double m = fmax(fmax(fabs(d.x), fabs(d.y)), fabs(d.z));
    c.ok = fabs(o.x) <= sc.cull_o_max && fabs(o.y) <= sc.cull_o_max && fabs(o.z) <= sc.cull_o_max && m >= 1e-10 &&
           m <= 1e10 && d.x == d.x && d.y == d.y && d.z == d.z;
    c.ix = fminf(fmaxf(__double2float_rn(ax.y), -1e25f), 1e25f);
    c.iy = fminf(fmaxf(__double2float_rn(ay.y), -1e25f), 1e25f);
    c.iz = fminf(fmaxf(__double2float_rn(az.y), -1e25f), 1e25f);
    c.cx = -(__double2float_rn(o.x) * c.ix);
    c.cy = -(__double2float_rn(o.y) * c.iy);
    c.cz = -(__double2float_rn(o.z) * c.iz);
    return c;
}
// one child box of a DevCull: returns "may contain an acceptable hit with t <= limit", tnear out
__device__ __forceinline__ bool cull_box(const CullRay& c, float bx0, float by0, float bz0, float bx1, float by1, float bz1,
                                         float limit, float& tnear) {
    const float x0 = fmaf(bx0, c.ix, c.cx), x1 = fmaf(bx1, c.ix, c.cx);
    const float y0 = fmaf(by0, c.iy, c.cy), y1 = fmaf(by1, c.iy, c.cy);
    const float z0 = fmaf(bz0, c.iz, c.cz), z1 = fmaf(bz1, c.iz, c.cz);
    tnear = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fminf(z0, z1));
    const float tfar = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fmaxf(z0, z1));
    // tnear <= tfar && tfar >= 0 && tnear <= limit, as ONE comparison.  The folded form additionally demands
    // limit >= 0, which only ever fails for a caller-supplied max_t < 0 — and then nothing can be accepted anyway
    // (every accepted t is > EPSILON); a NaN limit is ignored by fminf, i.e. it prunes nothing.
    return !c.ok || (fmaxf(tnear, 0.0f) <= fminf(tfar, limit));
}
#endif

// ray.rs:39-46: hit = !(tmax < 0.0) && !(tmin > tmax), as `bit` or 0.  Written as two chained predicate
// compares: left to the compiler the pair is rewritten into an IEEE fmax(tmin, 0) (NaN quieting and all,
// 8 instructions per child box) followed by a compare.
__device__ __forceinline__ uint32_t slab_hit_bit(double tmin, double tmax, uint32_t bit) {
#if defined(__CUDA_ARCH__)
    uint32_t r;
    asm("{\n\t.reg .pred p;\n\tsetp.geu.f64 p, %1, 0d0000000000000000;\n\tsetp.leu.and.f64 p, %2, %1, p;\n\t"
        "selp.u32 %0, %3, 0, p;\n\t}"
        : "=r"(r)
        : "d"(tmax), "d"(tmin), "r"(bit));
    return r;
#else   // host build of this header (tests/host_emulation): the same predicate in C
    return (!(tmax < 0.0) && !(tmin > tmax)) ? bit : 0u;
#endif
}

// One exact triangle test against the node-local best: strict '<', and on an exact t tie the earlier
// list entry wins (ray.rs:123-127). Lists are in ascending triangle id, so "earlier" == smaller id;
// records may therefore be visited in any order.
__device__ __forceinline__ bool test_record(const V3& o, const V3& d, const DevTri* __restrict__ tris, uint32_t r, Hit& nb) {
    double t, u, v;
    uint32_t id;
    if (moller_trumbore(o, d, tris + r, t, u, v, id)) {
        if (t < nb.t || (t == nb.t && nb.id != kNone && id < nb.id)) {
            nb.t = t; nb.id = id; nb.rec = r;
            return true;
        }
    }
    return false;
}

#ifndef RRT_SORT_NOUNROLL
#define RRT_SORT_NOUNROLL 1
#endif
#ifndef RRT_LEAF_NOUNROLL
#define RRT_LEAF_NOUNROLL 1
#endif
#ifndef RRT_SUB_BOX
#define RRT_SUB_BOX 2     // 1: test a node's subtree box on entry; 2: at the parent, for the children its exact slabs hit
#endif
#ifndef RRT_LIST_BOX
#define RRT_LIST_BOX 1
#endif
#ifndef RRT_CULL_LOOP
#define RRT_CULL_LOOP 0   // 0: if leaf / else inner per iteration; 1: inner step then leaf step; 2: descend-until-leaf
                          // 3: EXPERIMENT (round 2 measures it) — leaves are queued while the hierarchy is walked and
                          //    tested afterwards, so that the lanes of a warp meet in ONE exact-test block per list
                          //    instead of one per loop iteration (profiles: Moller-Trumbore runs at 10-16 of 32 lanes).
                          //    Records of a list may be tested in any order (test_record), and the f32 limit only ever
                          //    prunes: testing later can only test more, never decide differently.
#endif
#ifndef RRT_LATE_PRUNE
#define RRT_LATE_PRUNE 0  // EXPERIMENT (round 2 measures it): once nothing is pending at any level (no unvisited sibling of
                          // the current node or of an ancestor: the state after a node accepted an own triangle, and along
                          // every single-child chain below it), the rest of the walk can only matter by IMPROVING `best` —
                          // there is no sibling left for a `Some` to cut off (ray.rs:155-160).  Then (a) the node's own list
                          // is searched with limit min(nb.t, best.t) from the start, and (b) the trailing children of the
                          // sorted order whose subtree boxes begin beyond best.t are not visited (a SUFFIX only: a dropped
                          // child could otherwise have returned Some and cut off a later sibling).
#endif
#ifndef RRT_PEND
#define RRT_PEND 8        // RRT_CULL_LOOP == 3: queued leaves before the walk pauses to test them
#endif

#ifndef RRT_SLAB_PRED
#define RRT_SLAB_PRED 0   // EXPERIMENT (round 2 measures it): which children the exact slabs hit, from pairwise comparisons
                          // of the plane quotients instead of per-child tmin / tmax; tmin / tmax / sort keys only where
                          // two or more children are hit (tools/walk_stats.py: 27 % of the node visits of c4, 33 % of c2)
#endif

#if RRT_SLAB_PRED
// ray.rs:29-46 for one child box, reformulated.  With e_a / x_a the entry / exit quotient of the child's half on axis
// a (mn / mx of slab_axis, never NaN: the slow path replaces a NaN pair by the identities of the chains):
//     tmin = max(e_x, e_y, e_z),  tmax = min(x_x, x_y, x_z)
//     !(tmax < 0)     <=>  x_x >= 0 and x_y >= 0 and x_z >= 0
//     !(tmin > tmax)  <=>  e_a <= x_b for all nine (a, b)
// and e_a <= x_a always holds (fast path: the quotients of an axis are monotone in the plane; slow path: fmin <= fmax
// of the same pair, or -inf <= +inf for a replaced pair).  What is left are the six cross-axis comparisons and three
// signs: per PAIR of axes and pair of halves one predicate chain `e_a <= x_b and e_b <= x_a and x_c >= 0` decides two
// children at once (twelve chains, 36 DSETP, for what per-child min / max chains take ~100 instructions).
__device__ __forceinline__ uint32_t slab_pair_bits(double e1, double x1, double e2, double x2, double s, uint32_t bits) {
#if defined(__CUDA_ARCH__)
    uint32_t r;
    asm("{\n\t.reg .pred p;\n\tsetp.le.f64 p, %1, %2;\n\tsetp.le.and.f64 p, %3, %4, p;\n\t"
        "setp.ge.and.f64 p, %5, 0d0000000000000000, p;\n\tselp.u32 %0, %6, 0, p;\n\t}"
        : "=r"(r)
        : "d"(e1), "d"(x1), "d"(e2), "d"(x2), "d"(s), "r"(bits));
    return r;
#else   // host build of this header (tests/host_emulation): the same predicate in C
    return (e1 <= x1 && e2 <= x2 && s >= 0.0) ? bits : 0u;
#endif
}

// child k: x half = k in {2,3,6,7}, y half = k >= 4, z half = k in {1,2,5,6} (octree.rs:216-225 order, as below)
__device__ __forceinline__ uint32_t children_hit_pred(const double (&mnx)[2], const double (&mxx)[2], const double (&mny)[2],
                                                      const double (&mxy)[2], const double (&mnz)[2], const double (&mxz)[2]) {
    // (x half, y half): children {0,1} {2,3} {4,5} {6,7}; carries the sign of the x exit
    const uint32_t xy = slab_pair_bits(mnx[0], mxy[0], mny[0], mxx[0], mxx[0], 0x03u) |
                        slab_pair_bits(mnx[1], mxy[0], mny[0], mxx[1], mxx[1], 0x0Cu) |
                        slab_pair_bits(mnx[0], mxy[1], mny[1], mxx[0], mxx[0], 0x30u) |
                        slab_pair_bits(mnx[1], mxy[1], mny[1], mxx[1], mxx[1], 0xC0u);
    // (x half, z half): children {0,4} {1,5} {3,7} {2,6}; carries the sign of the z exit
    const uint32_t xz = slab_pair_bits(mnx[0], mxz[0], mnz[0], mxx[0], mxz[0], 0x11u) |
                        slab_pair_bits(mnx[0], mxz[1], mnz[1], mxx[0], mxz[1], 0x22u) |
                        slab_pair_bits(mnx[1], mxz[0], mnz[0], mxx[1], mxz[0], 0x88u) |
                        slab_pair_bits(mnx[1], mxz[1], mnz[1], mxx[1], mxz[1], 0x44u);
    // (y half, z half): children {0,3} {1,2} {4,7} {5,6}; carries the sign of the y exit
    const uint32_t yz = slab_pair_bits(mny[0], mxz[0], mnz[0], mxy[0], mxy[0], 0x09u) |
                        slab_pair_bits(mny[0], mxz[1], mnz[1], mxy[0], mxy[0], 0x06u) |
                        slab_pair_bits(mny[1], mxz[0], mnz[0], mxy[1], mxy[1], 0x90u) |
                        slab_pair_bits(mny[1], mxz[1], mnz[1], mxy[1], mxy[1], 0x60u);
    return xy & xz & yz;
}

#endif

#ifndef RRT_SHADOW_FILTER
#define RRT_SHADOW_FILTER 0   // EXPERIMENT (CPU work counters only so far): shadow queries first ask "is there ANY accepted
                              // triangle with t < max_t in the conservative f32 hull of the walk?" — no f64 slabs, no sort,
                              // any order.  None => the reference returns None (it can only return an accepted triangle
                              // with t < max_t, ray.rs:163 at the root); otherwise the exact walk decides as before.
#endif
#ifndef RRT_PRIMARY_FILTER
#define RRT_PRIMARY_FILTER 0  // with RRT_SHADOW_FILTER: camera rays whose lane just finished a sample that MISSED everything
                              // (the next sample of a lane is the same position of the neighbouring tile) try the filter first
#endif
#if RRT_SHADOW_FILTER && RRT_CULL_F32
// One culling hierarchy (or leaf word) walked in any order: false as soon as a record is accepted with t < max_t.
template <bool COUNT>
__device__ __forceinline__ bool list_clear(const DevScene& sc, const V3& o, const V3& d, const double max_t, const CullRay& cr,
                                           const float limit, uint32_t cur, uint32_t* __restrict__ cstk, TraceCount& tc) {
    int csp = 0;
    for (;;) {
        if (cur & kLeafBit) {
            const uint32_t first = cur & 0x0FFFFFFFu, cnt = ((cur >> 28) & 7u) + 1u;
#pragma unroll 1
            for (uint32_t i = 0; i < cnt; ++i) {
                double t, u, v;
                uint32_t id;
                if (COUNT) tc.mt++;
                RRT_STAT(7, 2);
                if (moller_trumbore(o, d, sc.tris + first + i, t, u, v, id) && t < max_t) { RRT_STAT(7, 4); return false; }
            }
            if (csp == 0) return true;
            cur = cstk[--csp];
        } else {
            const DevCull* __restrict__ c = sc.cull + cur;
            const float4* bp = reinterpret_cast<const float4*>(c);
            const float4 f0 = __ldg(bp), f1 = __ldg(bp + 1), f2 = __ldg(bp + 2);
            const uint2 ch = __ldg(reinterpret_cast<const uint2*>(&c->child[0]));
            float tn0, tn1;
            if (COUNT) tc.cull += 2;
            RRT_STAT(7, 5);
            const bool go0 = cull_box(cr, f0.x, f0.y, f0.z, f0.w, f1.x, f1.y, limit, tn0);
            const bool go1 = cull_box(cr, f1.z, f1.w, f2.x, f2.y, f2.z, f2.w, limit, tn1);
            if (go0 && go1) { cstk[csp++] = ch.y; cur = ch.x; }
            else if (go0 || go1) cur = go0 ? ch.x : ch.y;
            else if (csp != 0) cur = cstk[--csp];
            else return true;
        }
    }
}

// true: provably no triangle of the scene is accepted by the exact test with t < max_t.  Visits a SUPERSET of the nodes
// the reference can reach (subtree boxes instead of exact child slabs), so a `true` is exact; `false` decides nothing.
template <bool COUNT>
__device__ __forceinline__ bool shadow_clear(const DevScene& sc, const V3& o, const V3& d, const double max_t, const CullRay& cr,
                                             uint2* __restrict__ fstk, uint32_t* __restrict__ cstk, TraceCount& tc) {
    // fstk / cstk: the walk's own stacks (it starts after this returns); fstk holds per open level the first child
    // and the mask of the kept children still to visit
    if (!cr.ok) return false;
    const float limit = __double2float_ru(max_t);
#if RRT_SHADOW_FILTER == 2
    // a scene-wide hierarchy over all records (flatten, RRT_GLOBAL_CULL=1) answers the same question without the octree
    const uint32_t groot = __ldg(&sc.nodes[0]._pad2[0]);
    if (groot != 0u) {
        const bool clear = list_clear<COUNT>(sc, o, d, max_t, cr, limit, groot - 1u, cstk, tc);
        if (clear) RRT_STAT(7, 3);
        return clear;
    }
#endif
    int sp = 0;
    uint32_t node = 0;
    for (;;) {
        const DevNode* __restrict__ n = sc.nodes + node;
        const uint4 m = __ldg(reinterpret_cast<const uint4*>(&n->first_child));
        if (COUNT) tc.nodes++;
        RRT_STAT(7, 0);
        bool scan = m.w != 0u;
        if (scan) {
            const float4 lb = __ldg(reinterpret_cast<const float4*>(&n->list_box[0]));
            const float2 lc = __ldg(reinterpret_cast<const float2*>(&n->list_box[4]));
            float tn;
            if (COUNT) tc.cull++;
            RRT_STAT(7, 1);
            scan = cull_box(cr, lb.x, lb.y, lb.z, lb.w, lc.x, lc.y, limit, tn);
        }
        if (scan && !list_clear<COUNT>(sc, o, d, max_t, cr, limit, __ldg(&n->cull_root), cstk, tc)) return false;
        // children: every kept child whose subtree box the segment [0, max_t] may touch
        uint32_t todo = 0;
        for (uint32_t left = m.y, rank = 0; left; left &= left - 1u, ++rank) {
            const float4* sb = reinterpret_cast<const float4*>(sc.sub_boxes + (m.x + rank));
            const float4 s0 = __ldg(sb);
            const float2 s1 = __ldg(reinterpret_cast<const float2*>(sb + 1));
            float tn;
            if (COUNT) tc.cull++;
            RRT_STAT(7, 6);
            if (cull_box(cr, s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, limit, tn)) todo |= 1u << rank;
        }
        if (todo) fstk[sp++] = make_uint2(m.x, todo);
        if (sp == 0) { RRT_STAT(7, 3); return true; }
        uint2& top = fstk[sp - 1];
        const uint32_t r = (uint32_t)__ffs(top.y) - 1u;
        node = top.x + r;
        top.y &= top.y - 1u;
        if (top.y == 0u) --sp;
    }
}
#endif

#ifndef RRT_SPARSE_PRIMARY
#define RRT_SPARSE_PRIMARY 0  // EXPERIMENT (with RRT_SHADOW_FILTER == 2 and a scene built with RRT_GLOBAL_CULL=1): closest-hit
                              // rays with max_t = +inf first look for the closest accepted record of the whole scene in
                              // the scene-wide hierarchy and return it when the octree walk provably returns the same:
                              // see sparse_closest() and the design note in profiles/r01_experiments.md
#endif
#if RRT_SPARSE_PRIMARY && RRT_SHADOW_FILTER == 2 && RRT_CULL_F32
// Is the accepted t of a record inside the slab interval of its triangle's bounding box (widened by slack)?  False for
// the grazing records whose t is rounding noise, and for anything not finite.
__device__ __forceinline__ bool record_t_plausible(const V3& o, const V3& d, const DevTri* __restrict__ rec, double t, double slack) {
    const double* p = reinterpret_cast<const double*>(rec);
    double tn = -1e300, tf = 1e300;
    bool ok = true;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const double a = __ldg(p + k), b = a + __ldg(p + 3 + k), c = a + __ldg(p + 6 + k);
        const double lo = fmin(a, fmin(b, c)), hi = fmax(a, fmax(b, c));
        const double ok_ = k == 0 ? o.x : (k == 1 ? o.y : o.z), dk = k == 0 ? d.x : (k == 1 ? d.y : d.z);
        if (dk != 0.0) {
            const double t0 = div_rn(lo - ok_, dk), t1 = div_rn(hi - ok_, dk);   // as calls: cold code, six divisions
            tn = fmax(tn, fmin(t0, t1));
            tf = fmin(tf, fmax(t0, t1));
        } else {
            ok = ok && ok_ >= lo - 1e-9 * fabs(lo) && ok_ <= hi + 1e-9 * fabs(hi);
        }
    }
    return ok && t >= tn - slack && t <= tf + slack;
}

// The walk's result is the minimum-t own hit over the nodes it VISITS.  C = the closest accepted record of the scene is
// that result when (1) no other accepted record has a t within a margin of C's (ties and octant-boundary cases), (2) every
// box on the way from the root to C's node is hit by the exact slabs with tmin > 0 and tmin < tmax — so the node is
// reachable, and no sibling can be sorted before the path through the origin-inside rule (key = tmax) or a point contact —
// and (3) every stored triangle lies inside its node's box (flatten's flag), which puts the hits of any sibling sorted
// earlier before C's along the ray.  Anything else: false, and the exact walk runs as if this had not happened.
template <bool COUNT>
__device__ __forceinline__ bool sparse_closest(const DevScene& sc, const V3& o, const V3& d, const CullRay& cr, const AxisRay& ax,
                                               const AxisRay& ay, const AxisRay& az, uint32_t* __restrict__ cstk, Hit& best,
                                               TraceCount& tc) {
    const uint32_t groot = __ldg(&sc.nodes[0]._pad2[0]);
    if (!cr.ok || groot == 0u || __ldg(&sc.nodes[0]._pad) == 0u) return false;
    const double kInf = __longlong_as_double(0x7FF0000000000000ll);
    const double kMargin = 1e-6;
    // Every accepted record the search meets must lie where its triangle is: its t inside the slab interval of the
    // triangle's own bounding box, give or take kSlack (1e-9 S along the flattest non-zero direction component) — a
    // grazing record whose computed t is noise hands the ray over.  With that CHECKED (not assumed), and flatten's
    // "inside" flag, a record of an octant sorted BEFORE the path to C's node (left through a plane the path is entered
    // through) is met by the search (its box begins before that plane, hence before the limit) and is accepted, if at
    // all, no later than 2 kSlack after C: inside the window below, which hands over.
    const double ymax = fmax(fmax(d.x != 0.0 ? fabs(ax.y) : 0.0, d.y != 0.0 ? fabs(ay.y) : 0.0), d.z != 0.0 ? fabs(az.y) : 0.0);
    const double kSlack = 1e-9 * (sc.cull_o_max * (1.0 / kCullOriginRel)) * ymax;
    const double kAbs = 4.0 * kSlack;
    bool noisy = false;
    double t1 = kInf, t2 = kInf;          // smallest and second smallest accepted t (t2: of ANOTHER record)
    uint32_t id1 = kNone, rec1 = 0;
    float limit = __int_as_float(0x7f800000);
    uint32_t cur = groot - 1u;
    int csp = 0;
    for (;;) {
        if (cur & kLeafBit) {
            const uint32_t r = cur & 0x0FFFFFFFu;
            double t, u, v;
            uint32_t id;
            if (COUNT) tc.mt++;
            RRT_STAT(7, 8);
            if (moller_trumbore(o, d, sc.tris + r, t, u, v, id)) {
                noisy = noisy || !record_t_plausible(o, d, sc.tris + r, t, kSlack);
                if (t < t1 || (t == t1 && id < id1)) { t2 = t1; t1 = t; id1 = id; rec1 = r; }
                else if (t < t2) t2 = t;
                limit = __double2float_ru(t1 + (kMargin * t1 + kAbs));     // keep looking a little beyond the best
            }
            if (csp == 0) break;
            cur = cstk[--csp];
        } else {
            const DevCull* __restrict__ c = sc.cull + cur;
            const float4* bp = reinterpret_cast<const float4*>(c);
            const float4 f0 = __ldg(bp), f1 = __ldg(bp + 1), f2 = __ldg(bp + 2);
            const uint2 ch = __ldg(reinterpret_cast<const uint2*>(&c->child[0]));
            float tn0, tn1;
            if (COUNT) tc.cull += 2;
            RRT_STAT(7, 9);
            const bool go0 = cull_box(cr, f0.x, f0.y, f0.z, f0.w, f1.x, f1.y, limit, tn0);
            const bool go1 = cull_box(cr, f1.z, f1.w, f2.x, f2.y, f2.z, f2.w, limit, tn1);
            const bool swap = tn1 < tn0;
            if (go0 && go1) { cstk[csp++] = swap ? ch.x : ch.y; cur = swap ? ch.y : ch.x; }
            else if (go0 || go1) cur = go0 ? ch.x : ch.y;
            else if (csp != 0) cur = cstk[--csp];
            else break;
        }
    }
    if (id1 == kNone) { RRT_STAT(7, 10); return true; }        // no accepted record anywhere: None (best is None already)
    if (noisy || !(t2 > t1 + (kMargin * t1 + kAbs))) { RRT_STAT(7, 11); return false; }
    // up from C's node: every box on the path, with the walk's own slab arithmetic
    uint32_t node = (uint32_t)(__double_as_longlong(__ldg(&reinterpret_cast<const double*>(sc.tris + rec1)[9])) >> 32);
    while (node != 0u) {
        const DevNode* __restrict__ nd = sc.nodes + node;
        const uint32_t parent = __ldg(&nd->_pad), k = __ldg(&nd->_pad2[1]) & 7u;
        const DevNode* __restrict__ pn = sc.nodes + parent;
        double mnx[2], mxx[2], mny[2], mxy[2], mnz[2], mxz[2];
        slab_axis(ax, __ldg(&pn->px[0]), __ldg(&pn->px[1]), __ldg(&pn->px[2]), mnx[0], mnx[1], mxx[0], mxx[1]);
        slab_axis(ay, __ldg(&pn->py[0]), __ldg(&pn->py[1]), __ldg(&pn->py[2]), mny[0], mny[1], mxy[0], mxy[1]);
        slab_axis(az, __ldg(&pn->pz[0]), __ldg(&pn->pz[1]), __ldg(&pn->pz[2]), mnz[0], mnz[1], mxz[0], mxz[1]);
        const bool xh = (k & 2u) != 0u, yh = k >= 4u, zh = ((k ^ (k >> 1)) & 1u) != 0u;   // as in the child loop below
        const double tmin = max_nn(max_nn(xh ? mnx[1] : mnx[0], yh ? mny[1] : mny[0]), zh ? mnz[1] : mnz[0]);
        const double tmax = min_nn(min_nn(xh ? mxx[1] : mxx[0], yh ? mxy[1] : mxy[0]), zh ? mxz[1] : mxz[0]);
        if (COUNT) tc.box++;
        RRT_STAT(7, 12);
        if (!(tmin > 0.0 && tmin < tmax)) { RRT_STAT(7, 13); return false; }
        node = parent;
    }
    best.t = t1; best.id = id1; best.rec = rec1;
    RRT_STAT(7, 14);
    return true;
}
#endif

// Ray::intersect_with_octant_with_max_t(octree, 0, max_t).  any_hit: the caller only needs Some/None
// (shadow rays, raytracer.rs:181-187) — `best` only ever improves, so the first update decides.
template <bool COUNT>
__device__ __forceinline__ void traverse(const DevScene& sc, const V3& o, const V3& d, const double max_t,
                                         const bool any_hit, Hit& best, TraceCount& tc, const bool use_filter = true,
                                         uint32_t* __restrict__ sparse_outcome = nullptr) {
    best.id = kNone;
    best.rec = 0;
    best.t = max_t;
    if (sc.root_tri_count == 0) return;  // ray.rs:112-114 at the root

    // Open levels of the walk: the deepest one lives in registers (lv_first = first kept child of the node,
    // lv_ord = remaining visiting order, 3 bits per child, | remaining count << 24); shallower levels that
    // still have children left are on the stack.  Exhausted levels are never pushed.
    uint2 stk[kMaxOctreeDepth + 1];
    uint32_t lv_first = 0, lv_ord = 0;
    uint32_t cstk[kMaxCullDepth + 2];  // culling-hierarchy stack (depth bounded by the builder)
    int sp = 0;
    uint32_t node = 0;
    bool is_root = true;
    const AxisRay ax = make_axis(o.x, d.x, sc.planes_safe != 0u), ay = make_axis(o.y, d.y, sc.planes_safe != 0u),
                  az = make_axis(o.z, d.z, sc.planes_safe != 0u);
    const double kInf = __longlong_as_double(0x7FF0000000000000ll);
#if RRT_CULL_F32
    const CullRay cr = make_cull_ray(sc, o, d, ax, ay, az);
#if RRT_SHADOW_FILTER
    bool try_filter = (any_hit || RRT_PRIMARY_FILTER) && use_filter;
#if RRT_SPARSE_PRIMARY && RRT_SHADOW_FILTER == 2
    if (!any_hit && use_filter && max_t == kInf) {
        const bool answered = sparse_closest<COUNT>(sc, o, d, cr, ax, ay, az, cstk, best, tc);
        if (sparse_outcome) *sparse_outcome = answered ? 1u : 2u;
        if (answered) return;
        try_filter = false;   // handed over: an accepted record exists (or the ray cannot be culled), the filter would say so too
    }
#endif
    // closest-hit rays too when the caller expects a miss (RRT_PRIMARY_FILTER): no accepted triangle below max_t anywhere
    // is None for them just the same
    if (try_filter && shadow_clear<COUNT>(sc, o, d, max_t, cr, stk, cstk, tc)) return;
#endif
#else
    // reciprocal direction for the CONSERVATIVE culling boxes only (never for a reference decision)
    const double ivx = min_nn(max_nn(ax.y, -1e300), 1e300), ivy = min_nn(max_nn(ay.y, -1e300), 1e300),
                 ivz = min_nn(max_nn(az.y, -1e300), 1e300);
#endif

    for (;;) {
#if RRT_CULL_F32 && RRT_SUB_BOX == 1
        // everything below this node (own list + all descendants) lies inside sub_boxes[node]: a ray that
        // misses it gets None from the whole subtree, exactly what walking it would return.
        bool enter = true;
        if (!is_root) {
            const float4* sb = reinterpret_cast<const float4*>(sc.sub_boxes + node);
            const float4 s0 = __ldg(sb);
            const float2 s1 = __ldg(reinterpret_cast<const float2*>(sb + 1));
            float tn;
            if (COUNT) tc.cull++;
            enter = cull_box(cr, s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, __int_as_float(0x7f800000), tn);
        }
        if (enter) {
#endif
        const DevNode* __restrict__ n = sc.nodes + node;
        const uint4 m = __ldg(reinterpret_cast<const uint4*>(&n->first_child));  // first_child, child_mask, tri_first, tri_n
        if (COUNT) tc.nodes++;

        // ---- own triangles (ray.rs:116-129): hierarchy nodes and their leaves; a short list is one leaf word
        Hit nb;
        nb.id = kNone;
        nb.rec = 0;
        nb.t = is_root ? max_t : kInf;
#if RRT_CULL_F32
#if RRT_LATE_PRUNE
        // nothing pending anywhere: own records at t >= best.t can neither improve `best` nor cut anything off
        const bool alone = sp == 0 && (lv_ord >> 24) == 0u;
        float limit = __double2float_ru(alone ? min_nn(nb.t, best.t) : nb.t);
#else
        float limit = __double2float_ru(nb.t);
#endif
        bool scan = m.w != 0u;
#if RRT_LIST_BOX
        if (scan) {  // the node's own list as a whole
            const float4 lb = __ldg(reinterpret_cast<const float4*>(&n->list_box[0]));
            const float2 lc = __ldg(reinterpret_cast<const float2*>(&n->list_box[4]));
            float tn;
            if (COUNT) tc.cull++;
            scan = cull_box(cr, lb.x, lb.y, lb.z, lb.w, lc.x, lc.y, limit, tn);
        }
#endif
        if (scan) {
#else
        if (m.w != 0u) {
#endif
            uint32_t cur = __ldg(&n->cull_root);
            int csp = 0;
            bool live = true;
#ifdef RRT_WALK_STATS
            unsigned ws_inner = 0, ws_leaf = 0, ws_first = 0;
#endif
#if RRT_CULL_LOOP == 3
            uint32_t pend[RRT_PEND];
            int np = 0;
            for (;;) {
              while (live && np < RRT_PEND) {
                if (cur & kLeafBit) {
                    pend[np++] = cur;
                    if (csp != 0) cur = cstk[--csp];
                    else live = false;
                } else {
#else
            while (live) {
#endif
#if RRT_CULL_LOOP == 3
#elif RRT_CULL_LOOP == 2
                while (live && !(cur & kLeafBit)) {
#elif RRT_CULL_LOOP == 1
                if (!(cur & kLeafBit)) {
#else
                const bool was_leaf = (cur & kLeafBit) != 0u;
                if (!was_leaf) {
#endif
                    const DevCull* __restrict__ c = sc.cull + cur;
                    bool go0, go1, swap;
#ifdef RRT_WALK_STATS
                    ws_inner++;
#endif
                    if (COUNT) tc.cull += 2;
#if RRT_CULL_F32
                    const float4* bp = reinterpret_cast<const float4*>(c);
                    const float4 f0 = __ldg(bp), f1 = __ldg(bp + 1), f2 = __ldg(bp + 2);
                    const uint2 ch = __ldg(reinterpret_cast<const uint2*>(&c->child[0]));
                    float tn0, tn1;
                    go0 = cull_box(cr, f0.x, f0.y, f0.z, f0.w, f1.x, f1.y, limit, tn0);
                    go1 = cull_box(cr, f1.z, f1.w, f2.x, f2.y, f2.z, f2.w, limit, tn1);
                    swap = tn1 < tn0;
#else
                    const double2* bp = reinterpret_cast<const double2*>(c);
                    const uint2 ch = __ldg(reinterpret_cast<const uint2*>(&c->child[0]));
                    bool go[2];
                    double tn[2];
#pragma unroll
                    for (int s = 0; s < 2; ++s) {
                        const double2 b0 = __ldg(bp + 3 * s), b1 = __ldg(bp + 3 * s + 1), b2 = __ldg(bp + 3 * s + 2);
                        const double x0 = (b0.x - o.x) * ivx, x1 = (b1.y - o.x) * ivx;
                        const double y0 = (b0.y - o.y) * ivy, y1 = (b2.x - o.y) * ivy;
                        const double z0 = (b1.x - o.z) * ivz, z1 = (b2.y - o.z) * ivz;
                        const double tnear = max_nn(max_nn(min_nn(x0, x1), min_nn(y0, y1)), min_nn(z0, z1));
                        const double tfar = min_nn(min_nn(max_nn(x0, x1), max_nn(y0, y1)), max_nn(z0, z1));
                        // conservative by construction: boxes are padded outward by 1e-7 relative (flatten),
                        // orders of magnitude above the rounding of these products; a NaN (non-finite ray)
                        // fails every comparison and such a ray can never be accepted by the exact test.
                        go[s] = (tnear <= tfar) && (tfar >= 0.0) && (tnear <= nb.t);
                        tn[s] = tnear;
                    }
                    go0 = go[0]; go1 = go[1];
                    swap = tn[1] < tn[0];
#endif
                    if (go0 && go1) {
                        cstk[csp++] = swap ? ch.x : ch.y;
                        cur = swap ? ch.y : ch.x;
                    } else if (go0 || go1) {
                        cur = go0 ? ch.x : ch.y;
                    } else if (csp != 0) {
                        cur = cstk[--csp];
                    } else {
                        live = false;
                    }
                }
#if RRT_CULL_LOOP == 3
              }
              bool upd = false;
#pragma unroll 1
              for (int j = 0; j < np; ++j) {
                    const uint32_t first = pend[j] & 0x0FFFFFFFu;
                    const uint32_t cnt = ((pend[j] >> 28) & 7u) + 1u;
#elif RRT_CULL_LOOP == 0
                if (was_leaf) {
#else
                if (live && (cur & kLeafBit)) {
#endif
#if RRT_CULL_LOOP != 3
                    const uint32_t first = cur & 0x0FFFFFFFu;
                    const uint32_t cnt = ((cur >> 28) & 7u) + 1u;
                    bool upd = false;
#endif
#if RRT_LEAF_NOUNROLL
#pragma unroll 1
#endif
                    for (uint32_t i = 0; i < cnt; ++i) {
                        if (COUNT) tc.mt++;
                        upd |= test_record(o, d, sc.tris, first + i, nb);
                    }
#ifdef RRT_WALK_STATS
                    ws_leaf++;
                    if (upd && !ws_first) ws_first = ws_inner + 1;   // inner steps before the first accepted record (+1)
#endif
#if RRT_CULL_LOOP == 3
              }
              np = 0;
              if (!live || (any_hit && nb.id != kNone && nb.t < best.t)) break;  // list done / best will improve => Some
#if RRT_CULL_F32
              if (upd) limit = fminf(limit, __double2float_ru(nb.t));
#endif
            }
#else
                    if (any_hit && nb.id != kNone && nb.t < best.t) break;  // best will improve => Some
#if RRT_CULL_F32
#if RRT_LATE_PRUNE
                    if (upd) limit = fminf(limit, __double2float_ru(nb.t));
#else
                    if (upd) limit = __double2float_ru(nb.t);
#endif
#endif
                    if (csp != 0) cur = cstk[--csp];
                    else live = false;
                }
            }
#endif
            RRT_STAT(3, ws_inner);
            RRT_STAT(4, ws_leaf);
            RRT_STAT(5, ws_first);
        }
        RRT_STAT(6, (m.w != 0u ? 1u : 0u) | (m.y != 0u ? 2u : 0u) | (nb.id != kNone ? 4u : 0u));
        if (nb.id != kNone) {
            sp = 0;  // this node returns Some: every pending sibling, at every level, is dead
            lv_ord = 0;
            if (nb.t < best.t) {
                best = nb;
                if (any_hit) return;
            }
        }

        // ---- children (ray.rs:131-147): slab-test the kept ones, stable sort by key
        if (m.y != 0u) {
            const double2* pp = reinterpret_cast<const double2*>(&n->px[0]);
            const double2 q0 = __ldg(pp), q1 = __ldg(pp + 1), q2 = __ldg(pp + 2), q3 = __ldg(pp + 3);
            const double q4 = __ldg(&n->pz[2]);
            // (plane - origin) / direction : the SAME operands as ray.rs:22-27 for every child box that
            // uses the plane, so 9 quotients stand for the reference's 48.
            double mnx[2], mxx[2], mny[2], mxy[2], mnz[2], mxz[2];
            slab_axis(ax, q0.x, q0.y, q1.x, mnx[0], mnx[1], mxx[0], mxx[1]);
            slab_axis(ay, q1.y, q2.x, q2.y, mny[0], mny[1], mxy[0], mxy[1]);
            slab_axis(az, q3.x, q3.y, q4, mnz[0], mnz[1], mxz[0], mxz[1]);
#if RRT_SLAB_PRED
            uint32_t hitmask = children_hit_pred(mnx, mxx, mny, mxy, mnz, mxz) & m.y;
            double key[8];
            if (hitmask & (hitmask - 1u)) {   // two or more children hit: the sort below may need their keys
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const int xh = (k == 2 || k == 3 || k == 6 || k == 7) ? 1 : 0, yh = k >= 4 ? 1 : 0,
                              zh = (k == 1 || k == 2 || k == 5 || k == 6) ? 1 : 0;
                    const double tmin = max_nn(max_nn(mnx[xh], mny[yh]), mnz[zh]);  // ray.rs:29-32
                    const double tmax = min_nn(min_nn(mxx[xh], mxy[yh]), mxz[zh]);  // ray.rs:33-36
                    key[k] = (tmin < 0.0) ? tmax : tmin;                             // ray.rs:49-53
                }
            }
#else
            // max(max(x, y), z) / min(min(x, y), z) of ray.rs:29-36: the (x, y) part is shared by the two
            // children that differ only in z
            double nxy[2][2], fxy[2][2];
#pragma unroll
            for (int xh = 0; xh < 2; ++xh)
#pragma unroll
                for (int yh = 0; yh < 2; ++yh) {
                    nxy[xh][yh] = max_nn(mnx[xh], mny[yh]);
                    fxy[xh][yh] = min_nn(mxx[xh], mxy[yh]);
                }
            double key[8];
            uint32_t hitmask = 0;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const int xh = (k == 2 || k == 3 || k == 6 || k == 7) ? 1 : 0, yh = k >= 4 ? 1 : 0,
                          zh = (k == 1 || k == 2 || k == 5 || k == 6) ? 1 : 0;
                const double tmin = max_nn(nxy[xh][yh], mnz[zh]);  // ray.rs:29-32
                const double tmax = min_nn(fxy[xh][yh], mxz[zh]);  // ray.rs:33-36
                key[k] = (tmin < 0.0) ? tmax : tmin;                 // ray.rs:49-53
                hitmask |= slab_hit_bit(tmin, tmax, 1u << k);        // ray.rs:39-46
            }
#endif
            hitmask &= m.y;
            if (COUNT) tc.box += __popc(m.y);
            RRT_STAT(0, __popc(m.y));
            RRT_STAT(1, __popc(hitmask));
#if RRT_CULL_F32 && RRT_SUB_BOX == 2
            // A child whose SUBTREE box (own list + all descendants, padded f32, DevSubBox) the ray misses
            // returns None (ray.rs:104-168 can only return a triangle the exact test accepts); dropping it from
            // the sorted list is invisible (stable sort; None children are skipped by the loop at ray.rs:152-161).
            // Tested here, for the children the exact slabs hit, rather than on entry: no sort slot, no push/pop,
            // and the box loads of one node are independent of each other.
#if RRT_LATE_PRUNE
            // after this node's own list: is anything pending besides the children about to be pushed?
            const bool alone_c = sp == 0 && (lv_ord >> 24) == 0u;
            const float far_limit = alone_c ? __double2float_ru(best.t) : __int_as_float(0x7f800000);
            uint32_t far_r = 0;   // by rank among the kept children: subtree box begins beyond best.t
#endif
            for (uint32_t left = hitmask; left;) {
                const uint32_t kk = (uint32_t)__ffs(left) - 1u;
                left &= left - 1u;
                const uint32_t rank = (uint32_t)__popc(m.y & ((1u << kk) - 1u));
                const float4* sb = reinterpret_cast<const float4*>(sc.sub_boxes + (m.x + rank));
                const float4 s0 = __ldg(sb);
                const float2 s1 = __ldg(reinterpret_cast<const float2*>(sb + 1));
                float tn;
                if (COUNT) tc.cull++;
                if (!cull_box(cr, s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, __int_as_float(0x7f800000), tn)) hitmask &= ~(1u << kk);
#if RRT_LATE_PRUNE
                // same proof as cull_box's limit: every acceptable t* in the box has t* >= tn, so tn > ru(best.t) means
                // t* > best.t — this subtree cannot improve `best` (strict '<', ray.rs:163)
                else if (cr.ok && tn > far_limit) far_r |= 1u << rank;
#endif
            }
#endif
            RRT_STAT(2, __popc(hitmask));
            if (hitmask) {
                // stable sort by key (ray.rs:146-147) as repeated selection of the smallest remaining key,
                // lowest child index first among equals; the last remaining child needs no comparison.
                uint32_t order = 0, rem = hitmask, shift = 0;
                const uint32_t nh = __popc(hitmask);
#if RRT_SORT_NOUNROLL
#pragma unroll 1
#endif
                for (uint32_t r = 1; r < nh; ++r) {
                    double bk = 0.0;
                    uint32_t bi = 8;
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const bool take = ((rem >> k) & 1u) && (bi == 8u || key[k] < bk);
                        bk = take ? key[k] : bk;
                        bi = take ? (uint32_t)k : bi;
                    }
                    rem &= ~(1u << bi);
                    order |= (uint32_t)__popc(m.y & ((1u << bi) - 1u)) << shift;
                    shift += 3;
                }
                const uint32_t last = (uint32_t)__ffs(rem) - 1u;
                order |= (uint32_t)__popc(m.y & ((1u << last) - 1u)) << shift;
#if RRT_LATE_PRUNE
                uint32_t keep = nh;   // drop the trailing children that cannot improve `best` (alone_c only: far_r == 0 otherwise)
                while (keep != 0u && ((far_r >> ((order >> (3u * (keep - 1u))) & 7u)) & 1u)) --keep;
                if (keep != 0u) {
                    if (lv_ord >> 24) stk[sp++] = make_uint2(lv_first, lv_ord);
                    lv_first = m.x;
                    lv_ord = (order & ((1u << (3u * keep)) - 1u)) | (keep << 24);
                }
#else
                if (lv_ord >> 24) stk[sp++] = make_uint2(lv_first, lv_ord);
                lv_first = m.x;
                lv_ord = order | (nh << 24);
#endif
            }
        }

#if RRT_CULL_F32 && RRT_SUB_BOX == 1
        }
#endif
        // ---- next node: first remaining child of the deepest open level (ray.rs:152-161)
        if ((lv_ord >> 24) == 0u) {
            if (sp == 0) return;
            const uint2 top = stk[--sp];
            lv_first = top.x;
            lv_ord = top.y;
        }
        node = lv_first + (lv_ord & 7u);
        lv_ord = ((lv_ord & 0x00FFFFFFu) >> 3) | (((lv_ord >> 24) - 1u) << 24);
        is_root = false;
    }
}

}  // namespace rrt

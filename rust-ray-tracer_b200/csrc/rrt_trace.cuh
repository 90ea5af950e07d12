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
//   otherwise (zero / tiny / huge / non-finite component): true divisions and NaN-aware min/max.
struct AxisRay {
    double o, d, y;
    bool fast;
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

// mn[h] / mx[h] = min / max of the two plane quotients bounding half h (0 = low, 1 = high) of one axis.
__device__ __forceinline__ void slab_axis(const AxisRay& a, double p0, double p1, double p2, double& mn0, double& mn1,
                                          double& mx0, double& mx1) {
    if (a.fast) {
        const double a0 = p0 - a.o, a1 = p1 - a.o, a2 = p2 - a.o;
        const double q0 = div_markstein(a0, a.d, a.y), q1 = div_markstein(a1, a.d, a.y), q2 = div_markstein(a2, a.d, a.y);
        mn0 = min_nn(q0, q1); mx0 = max_nn(q0, q1);
        mn1 = min_nn(q1, q2); mx1 = max_nn(q1, q2);
    } else {
        double out[4];
        slab_axis_slow(p0, p1, p2, a.o, a.d, out);
        mn0 = out[0]; mn1 = out[1]; mx0 = out[2]; mx1 = out[3];
    }
}

// One exact triangle test against the node-local best: strict '<', and on an exact t tie the earlier
// list entry wins (ray.rs:123-127). Lists are in ascending triangle id, so "earlier" == smaller id;
// records may therefore be visited in any order.
__device__ __forceinline__ void test_record(const V3& o, const V3& d, const DevTri* rec, Hit& nb) {
    double t, u, v;
    uint32_t id;
    if (moller_trumbore(o, d, rec, t, u, v, id)) {
        if (t < nb.t || (t == nb.t && nb.id != kNone && id < nb.id)) {
            nb.t = t; nb.u = u; nb.v = v; nb.id = id;
        }
    }
}

// Ray::intersect_with_octant_with_max_t(octree, 0, max_t).  any_hit: the caller only needs Some/None
// (shadow rays, raytracer.rs:181-187) — `best` only ever improves, so the first update decides.
template <bool COUNT>
__device__ __forceinline__ void traverse(const DevScene& sc, const V3& o, const V3& d, const double max_t,
                                         const bool any_hit, Hit& best, TraceCount& tc) {
    best.id = kNone;
    best.t = max_t;
    best.u = 0.0;
    best.v = 0.0;
    if (sc.root_tri_count == 0) return;  // ray.rs:112-114 at the root

    uint2 stk[kMaxOctreeDepth + 1];  // per level: .x = first kept child, .y = remaining order (3 bits each) | count << 24
    uint32_t cstk[kMaxCullDepth + 2];  // culling-hierarchy stack (depth bounded by the builder)
    int sp = 0;
    uint32_t node = 0;
    bool is_root = true;
    const AxisRay ax = make_axis(o.x, d.x, sc.planes_safe != 0u), ay = make_axis(o.y, d.y, sc.planes_safe != 0u),
                  az = make_axis(o.z, d.z, sc.planes_safe != 0u);
    // reciprocal direction for the CONSERVATIVE culling boxes only (never for a reference decision)
    const double ivx = min_nn(max_nn(ax.y, -1e300), 1e300), ivy = min_nn(max_nn(ay.y, -1e300), 1e300),
                 ivz = min_nn(max_nn(az.y, -1e300), 1e300);
    const double kInf = __longlong_as_double(0x7FF0000000000000ll);

    for (;;) {
        const DevNode* __restrict__ n = sc.nodes + node;
        const uint4 m = __ldg(reinterpret_cast<const uint4*>(&n->first_child));  // first_child, child_mask, tri_first, tri_n
        if (COUNT) tc.nodes++;

        // ---- own triangles (ray.rs:116-129): one loop for linear lists, hierarchy nodes and their leaves
        Hit nb;
        nb.id = kNone;
        nb.t = is_root ? max_t : kInf;
        nb.u = nb.v = 0.0;
        if (m.w != 0u) {
            uint32_t cur = __ldg(&n->cull_root);
            int csp = 0;
            for (;;) {
                if (cur & kLeafBit) {
                    const DevTri* rec = sc.tris + (cur & 0x0FFFFFFFu);
                    const uint32_t cnt = ((cur >> 28) & 7u) + 1u;
                    for (uint32_t i = 0; i < cnt; ++i) {
                        if (COUNT) tc.mt++;
                        test_record(o, d, rec + i, nb);
                    }
                    if (any_hit && nb.id != kNone && nb.t < best.t) break;  // best will improve => Some
                    if (csp == 0) break;
                    cur = cstk[--csp];
                } else {
                    const DevCull* __restrict__ c = sc.cull + cur;
                    const double2* bp = reinterpret_cast<const double2*>(c);
                    const uint2 ch = __ldg(reinterpret_cast<const uint2*>(&c->child[0]));
                    bool go[2];
                    double tn[2];
#pragma unroll
                    for (int s = 0; s < 2; ++s) {
                        const double2 b0 = __ldg(bp + 3 * s), b1 = __ldg(bp + 3 * s + 1), b2 = __ldg(bp + 3 * s + 2);
                        if (COUNT) tc.cull++;
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
                    if (go[0] && go[1]) {
                        const bool swap = tn[1] < tn[0];
                        cstk[csp++] = swap ? ch.x : ch.y;
                        cur = swap ? ch.y : ch.x;
                    } else if (go[0] || go[1]) {
                        cur = go[0] ? ch.x : ch.y;
                    } else {
                        if (csp == 0) break;
                        cur = cstk[--csp];
                    }
                }
            }
        }
        if (nb.id != kNone) {
            sp = 0;  // this node returns Some: every pending sibling, at every level, is dead
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
            double key[8];
            uint32_t hitmask = 0;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                key[k] = 0.0;
                if ((m.y >> k) & 1u) {
                    const int xh = (k == 2 || k == 3 || k == 6 || k == 7) ? 1 : 0, yh = k >= 4 ? 1 : 0,
                              zh = (k == 1 || k == 2 || k == 5 || k == 6) ? 1 : 0;
                    const double tmin = max_nn(max_nn(mnx[xh], mny[yh]), mnz[zh]);  // ray.rs:29-32
                    const double tmax = min_nn(min_nn(mxx[xh], mxy[yh]), mxz[zh]);  // ray.rs:33-36
                    const bool hit = !(tmax < 0.0) && !(tmin > tmax);               // ray.rs:39-46
                    key[k] = (tmin < 0.0) ? tmax : tmin;                             // ray.rs:49-53
                    hitmask |= hit ? (1u << k) : 0u;
                }
            }
            if (COUNT) tc.box += __popc(m.y);
            if (hitmask) {
                // stable sort by key (ray.rs:146-147) as repeated selection of the smallest remaining key,
                // lowest child index first among equals; the last remaining child needs no comparison.
                uint32_t order = 0, rem = hitmask, shift = 0;
                const uint32_t nh = __popc(hitmask);
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
                stk[sp++] = make_uint2(m.x, order | (nh << 24));
            }
        }

        // ---- next node: first remaining child of the deepest open level (ray.rs:152-161)
        while (sp > 0 && (stk[sp - 1].y >> 24) == 0u) --sp;
        if (sp == 0) return;
        const uint2 top = stk[sp - 1];
        node = top.x + (top.y & 7u);
        const uint32_t cnt = (top.y >> 24) - 1u;
        stk[sp - 1].y = ((top.y & 0x00FFFFFFu) >> 3) | (cnt << 24);
        is_root = false;
    }
}

}  // namespace rrt

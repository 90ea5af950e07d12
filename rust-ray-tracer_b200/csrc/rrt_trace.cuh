// rrt_trace.cuh — octree traversal (ray.rs:96-168): the exact iterative short-stack walk, and the search of ONE
// scene-wide culling hierarchy that answers most rays without the walk where that is provably the walk's answer.
//
// Why the iterative walk equals the reference recursion:
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
constexpr uint32_t kDone = 0xFFFFFFFFu;     // bottom of a hierarchy stack: "no entry left" (never a real leaf word, see flatten)

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

// mn[h] / mx[h] = min / max of the two plane quotients bounding half h (0 = low, 1 = high) of one axis.
__device__ __forceinline__ void slab_axis(const AxisRay& a, double p0, double p1, double p2, double& mn0, double& mn1,
                                          double& mx0, double& mx1) {
    if (a.fast) {
        const double a0 = p0 - a.o, a1 = p1 - a.o, a2 = p2 - a.o;
        const double q0 = div_markstein(a0, a.d, a.y), q1 = div_markstein(a1, a.d, a.y), q2 = div_markstein(a2, a.d, a.y);
        mn0 = a.neg ? q1 : q0; mx0 = a.neg ? q0 : q1;
        mn1 = a.neg ? q2 : q1; mx1 = a.neg ? q1 : q2;
    } else {
        double out[4];
        slab_axis_slow(p0, p1, p2, a.o, a.d, out);
        mn0 = out[0]; mn1 = out[1]; mx0 = out[2]; mx1 = out[3];
    }
}

// ---------------------------------------------------------------- conservative culling in f32
// The hierarchies' boxes only decide which records get the exact test, so they are tested in FP32 (FFMA +
// FMNMX at full rate, 64-byte nodes) with everything rounded or padded to the safe side.
// Error budget, in units of S = FlatScene::scale (>= 1, >= every |coordinate| of a stored triangle):
//   * boxes are padded by 1e-5 S, then rounded outward to f32;
//   * a ray is ELIGIBLE when every |origin component| <= 16 S and max|direction component| is in
//     [1e-10, 1e10] (anything else — NaN included — is not culled at all: every record is tested);
//   * per plane t~ = fma(b, inv, -(o32*inv)), o32 = RN32(o), inv = RN32(1/d) clamped to +-1e25:
//     |o - o32| <= 2^-24 * 16 S, the product o32*inv and the fma each round once (2^-24 relative to
//     terms of magnitude <= 17 S |inv|), inv carries 2^-24 relative: in space units the plane moves by
//     less than 4.2e-6 S, under half of the pad.  A hit accepted by the exact test lies within 1e-12 S of
//     its triangle (ASSUMPTION (A) of DESIGN.md §5.4: false only for rays within ~1e-9 rad AND ~1e-9 S of a
//     triangle's plane, where the reference's own t is rounding noise), hence (pad) strictly inside the computed
//     slab interval on every axis: tnear~ <= t* <= tfar~, which is all the three comparisons below use
//     (t* > 0; `tnear~ > limit` proves t* > limit, so an exact tie is never pruned).
//   * a direction component that is zero or below 1e-25 uses inv = +-1e25: the ray cannot leave a slab of
//     thickness >= 2 pad within t <= 17 S / 1e-10, and the computed interval is either astronomically
//     wide around 0 (origin inside) or astronomically far on one side (origin outside by > pad/2).
struct CullRay {
    float ix, iy, iz, cx, cy, cz;   // inv and -(o32 * inv)
    bool ok;
};
__device__ __forceinline__ CullRay make_cull_ray(const double cull_o_max, const V3& o, const V3& d, const AxisRay& ax,
                                                 const AxisRay& ay, const AxisRay& az) {
    CullRay c;
    const double m = fmax(fmax(fabs(d.x), fabs(d.y)), fabs(d.z));
    c.ok = fabs(o.x) <= cull_o_max && fabs(o.y) <= cull_o_max && fabs(o.z) <= cull_o_max && m >= 1e-10 &&
           m <= 1e10 && d.x == d.x && d.y == d.y && d.z == d.z;
    c.ix = fminf(fmaxf(__double2float_rn(ax.y), -1e25f), 1e25f);
    c.iy = fminf(fmaxf(__double2float_rn(ay.y), -1e25f), 1e25f);
    c.iz = fminf(fmaxf(__double2float_rn(az.y), -1e25f), 1e25f);
    c.cx = -(__double2float_rn(o.x) * c.ix);
    c.cy = -(__double2float_rn(o.y) * c.iy);
    c.cz = -(__double2float_rn(o.z) * c.iz);
    return c;
}
// one box: returns "may contain an acceptable hit with t <= limit", tnear out
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

// ---------------------------------------------------------------- the exact walk
// Ray::intersect_with_octant_with_max_t(octree, 0, max_t).  any_hit: the caller only needs Some/None
// (shadow rays, raytracer.rs:181-187) — `best` only ever improves, so the first update decides.
template <bool COUNT>
__device__ __forceinline__ void walk(const DevScene& sc, const V3& o, const V3& d, const double max_t, const bool any_hit,
                                     Hit& best, TraceCount& tc) {
    best.id = kNone;
    best.rec = 0;
    best.t = max_t;
    if (sc.root_tri_count == 0) return;  // ray.rs:112-114 at the root

    // Open levels of the walk: the deepest one lives in registers (lv_first = first kept child of the node,
    // lv_ord = remaining visiting order, 3 bits per child, | remaining count << 24); shallower levels that
    // still have children left are on the stack.  Exhausted levels are never pushed.
    uint2 stk[kMaxOctreeDepth + 1];
    uint32_t lv_first = 0, lv_ord = 0;
    uint32_t cstk[kMaxCullDepth + 3];  // culling-hierarchy stack (depth bounded by the builder), bottom entry kDone
    cstk[0] = kDone;
    int sp = 0;
    uint32_t node = 0;
    bool is_root = true;
    const AxisRay ax = make_axis(o.x, d.x, sc.planes_safe != 0u), ay = make_axis(o.y, d.y, sc.planes_safe != 0u),
                  az = make_axis(o.z, d.z, sc.planes_safe != 0u);
    const double kInf = __longlong_as_double(0x7FF0000000000000ll);
    const CullRay cr = make_cull_ray(sc.cull_o_max, o, d, ax, ay, az);

    for (;;) {
        const DevNode* __restrict__ n = sc.nodes + node;
        const uint4 m = __ldg(reinterpret_cast<const uint4*>(&n->first_child));  // first_child, child_mask, tri_first, tri_n
        if (COUNT) tc.nodes++;

        // ---- own triangles (ray.rs:116-129): hierarchy nodes and their leaves; a short list is one leaf word
        Hit nb;
        nb.id = kNone;
        nb.rec = 0;
        nb.t = is_root ? max_t : kInf;
        float limit = __double2float_ru(nb.t);
        bool scan = m.w != 0u;
        if (scan) {  // the node's own list as a whole
            const float4 lb = __ldg(reinterpret_cast<const float4*>(&n->list_box[0]));
            const float2 lc = __ldg(reinterpret_cast<const float2*>(&n->list_box[4]));
            float tn;
            if (COUNT) tc.cull++;
            scan = cull_box(cr, lb.x, lb.y, lb.z, lb.w, lc.x, lc.y, limit, tn);
        }
        if (scan) {
            uint32_t cur = __ldg(&n->cull_root);
            int csp = 1;   // cstk[0] = kDone, see scene_search
            // rounds of "while inner node / while leaf": a lane that reaches a leaf waits for the others of its warp, so
            // the exact tests below run with more lanes than one test per hierarchy step would (profiles/r02_experiments.md)
            while (cur != kDone) {
                while (!(cur & kLeafBit)) {
                    const DevCull* __restrict__ c = sc.cull + cur;
                    if (COUNT) tc.cull += 2;
                    const float4* bp = reinterpret_cast<const float4*>(c);
                    const float4 f0 = __ldg(bp), f1 = __ldg(bp + 1), f2 = __ldg(bp + 2);
                    const uint2 ch = __ldg(reinterpret_cast<const uint2*>(&c->child[0]));
                    float tn0, tn1;
                    const bool go0 = cull_box(cr, f0.x, f0.y, f0.z, f0.w, f1.x, f1.y, limit, tn0);
                    const bool go1 = cull_box(cr, f1.z, f1.w, f2.x, f2.y, f2.z, f2.w, limit, tn1);
                    const bool swap = tn1 < tn0;
                    const bool both = go0 && go1;
                    uint32_t nxt = go0 ? ch.x : ch.y;
                    if (both) {
                        nxt = swap ? ch.y : ch.x;
                        cstk[csp++] = swap ? ch.x : ch.y;
                    }
                    if (!(go0 || go1)) nxt = cstk[--csp];
                    cur = nxt;
                }
                while ((cur & kLeafBit) && cur != kDone) {
                    const uint32_t first = cur & 0x0FFFFFFFu;
                    const uint32_t cnt = ((cur >> 28) & 7u) + 1u;
                    bool upd = false;
#pragma unroll 1
                    for (uint32_t i = 0; i < cnt; ++i) {
                        if (COUNT) tc.mt++;
                        upd |= test_record(o, d, sc.tris, first + i, nb);
                    }
                    if (upd) limit = __double2float_ru(nb.t);
                    cur = (any_hit && nb.id != kNone && nb.t < best.t) ? kDone : cstk[--csp];  // best will improve => Some
                }
            }
        }
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
            hitmask &= m.y;
            if (COUNT) tc.box += __popc(m.y);
            // A child whose SUBTREE box (own list + all descendants, padded f32, DevSubBox) the ray misses
            // returns None (ray.rs:104-168 can only return a triangle the exact test accepts); dropping it from
            // the sorted list is invisible (stable sort; None children are skipped by the loop at ray.rs:152-161).
            // Tested here, for the children the exact slabs hit, rather than on entry: no sort slot, no push/pop,
            // and the box loads of one node are independent of each other.
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
            }
            if (hitmask) {
                // stable sort by key (ray.rs:146-147) as repeated selection of the smallest remaining key,
                // lowest child index first among equals; the last remaining child needs no comparison.
                uint32_t order = 0, rem = hitmask, shift = 0;
                const uint32_t nh = __popc(hitmask);
#pragma unroll 1
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
                if (lv_ord >> 24) stk[sp++] = make_uint2(lv_first, lv_ord);
                lv_first = m.x;
                lv_ord = order | (nh << 24);
            }
        }

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

// ---------------------------------------------------------------- search of the scene-wide hierarchy
// flatten builds ONE more culling hierarchy of the usual kind over ALL records of the scene (DevScene::global_root,
// every leaf a single record): which octree node owns a record does not matter to the two questions below, and a
// binary SAH hierarchy answers them in ~25 f32 box-pair steps where the walk visits ~10 octree nodes at ~300
// instructions each.  ONE loop answers both: a single copy in the instruction cache, and both kinds of ray keep a warp
// converged.  Inlined at traverse()'s one call site: as a called function every trip of the loop re-creates the
// global-memory descriptor (2 LDC + 4 R2UR), ~9 % of a step (measurements: profiles/r02_experiments.md).

// (1) `any` — shadow queries (triangle_exists_between_points, raytracer.rs:164-188; DESIGN.md §5.5 "scene_clear").  The
// walk's result is the minimum-t own hit over the nodes it visits, every one of them a record the exact test ACCEPTS
// with t < max_t at the root (ray.rs:163) — so if no record of the scene is accepted with t < max_t, the reference
// returns None: status 1, id kNone (under the conservative culling above).  An accepted record below max_t decides
// nothing: status 0, and the walk runs.
//
// (2) `!any` — closest-hit rays with max_t = +inf whose origin lies outside the geometry (camera rays; DESIGN.md §5.6
// "sparse_closest").  Let C be the accepted record with the smallest t.  The walk returns C when (proof: DESIGN.md §5.6)
//   (a) every stored triangle lies inside the box of its node, and the root box is not absurdly larger than the
//       geometry (flatten: DevScene::sparse_ok);
//   (b) all three direction components take the fast slab path (finite, non-zero, sane magnitude), the ray is
//       eligible for f32 culling;
//   (c) no OTHER accepted record has t <= t_C + W, W = kSparseWindow * S * max_k |1/d_k| — twice the distance (in t)
//       by which an accepted hit can lie outside its triangle's box under assumption (A), plus rounding;
//   (d) every box on the path from the root to C's node is hit by the walk's own slab arithmetic with
//       0 <= tmin < tmax (reachable, and the origin is not inside it).
// Then a sibling sorted before a path box is left no later than the path box is entered (they share the quotient of
// the plane that separates them, bit for bit), so any record accepted inside it has t <= t_C + W and is excluded by
// (c): the path box is the first child that returns Some at every level, C's node is reached, and nothing the walk
// visits can beat or tie C.  Anything else: status 0, and the walk runs as if this had not happened.
constexpr double kSparseWindow = 4.1e-5;   // 2 x (pad 1e-5 + f32 slop 0.6e-5, rounded up to 2e-5) + 2.5 % for roundings

struct SearchResult {
    double t;
    uint32_t id, rec;
    uint32_t status;   // 0: handed over; 1: answered (id == kNone: no record is accepted (below max_t))
};

// tmin / tmax of reference child k of `pn` with the walk's own arithmetic (slab_axis on the fast path: same operands,
// same bits); only the two planes per axis that bound the child are needed
__device__ __forceinline__ void child_interval(const AxisRay& ax, const AxisRay& ay, const AxisRay& az,
                                               const DevNode* __restrict__ pn, uint32_t k, double& tmin, double& tmax) {
    const uint32_t xh = (k >> 1) & 1u, yh = k >> 2, zh = (k ^ (k >> 1)) & 1u;   // child order octree.rs:216-225
    const double x0 = div_markstein(__ldg(&pn->px[xh]) - ax.o, ax.d, ax.y), x1 = div_markstein(__ldg(&pn->px[xh + 1]) - ax.o, ax.d, ax.y);
    const double y0 = div_markstein(__ldg(&pn->py[yh]) - ay.o, ay.d, ay.y), y1 = div_markstein(__ldg(&pn->py[yh + 1]) - ay.o, ay.d, ay.y);
    const double z0 = div_markstein(__ldg(&pn->pz[zh]) - az.o, az.d, az.y), z1 = div_markstein(__ldg(&pn->pz[zh + 1]) - az.o, az.d, az.y);
    tmin = max_nn(max_nn(ax.neg ? x1 : x0, ay.neg ? y1 : y0), az.neg ? z1 : z0);
    tmax = min_nn(min_nn(ax.neg ? x0 : x1, ay.neg ? y0 : y1), az.neg ? z0 : z1);
}

template <bool COUNT>
__device__ __forceinline__ SearchResult scene_search(const DevScene& sc, const V3& o, const V3& d, const double max_t,
                                                    const bool any, TraceCount& tc) {
    const double kInf = __longlong_as_double(0x7FF0000000000000ll);
    SearchResult res;
    res.t = kInf; res.id = kNone; res.rec = 0; res.status = 0u;
    if (sc.global_root == 0u || (!any && sc.sparse_ok == 0u)) return res;
    const bool ps = sc.planes_safe != 0u;
    const AxisRay ax = make_axis(o.x, d.x, ps), ay = make_axis(o.y, d.y, ps), az = make_axis(o.z, d.z, ps);
    const CullRay cr = make_cull_ray(sc.cull_o_max, o, d, ax, ay, az);
    if (!cr.ok || (!any && !(ax.fast && ay.fast && az.fast))) return res;
    const double ymax = fmax(fmax(fabs(ax.y), fabs(ay.y)), fabs(az.y));
    const double W = kSparseWindow * (sc.cull_o_max * (1.0 / kCullOriginRel)) * ymax;
    double t1 = kInf, t2 = kInf;          // smallest accepted t, and the smallest accepted t of any OTHER record
    uint32_t id1 = kNone, rec1 = 0;
    float limit = any ? __double2float_ru(max_t) : __int_as_float(0x7f800000);
    // The stack's bottom entry is kDone (a leaf word no hierarchy holds: flatten keeps record indices below 2^28 - 1), so a
    // pop needs no emptiness test and "finished" is one more value of `cur` — no flag to carry through the loops.
    uint32_t cstk[kMaxCullDepth + 3];
    cstk[0] = kDone;
    int csp = 1;
    uint32_t cur = sc.global_root - 1u;
    // rounds of "while inner node / while leaf", nearer child first for both kinds of ray (measured: the order pays for
    // shadow queries too, profiles/r02_experiments.md)
    while (cur != kDone) {
        while (!(cur & kLeafBit)) {
            const DevCull* __restrict__ c = sc.cull + cur;
            const float4* bp = reinterpret_cast<const float4*>(c);
            const float4 f0 = __ldg(bp), f1 = __ldg(bp + 1), f2 = __ldg(bp + 2);
            const uint2 ch = __ldg(reinterpret_cast<const uint2*>(&c->child[0]));
            float tn0, tn1;
            if (COUNT) tc.cull += 2;
            const bool go0 = cull_box(cr, f0.x, f0.y, f0.z, f0.w, f1.x, f1.y, limit, tn0);
            const bool go1 = cull_box(cr, f1.z, f1.w, f2.x, f2.y, f2.z, f2.w, limit, tn1);
            const bool swap = tn1 < tn0;
            const bool both = go0 && go1;
            uint32_t nxt = go0 ? ch.x : ch.y;
            if (both) {
                nxt = swap ? ch.y : ch.x;
                cstk[csp++] = swap ? ch.x : ch.y;
            }
            if (!(go0 || go1)) nxt = cstk[--csp];
            cur = nxt;
        }
        while ((cur & kLeafBit) && cur != kDone) {
            const uint32_t r = cur & 0x0FFFFFFFu;
            double t, u, v;
            uint32_t id;
            if (COUNT) tc.mt++;
            if (moller_trumbore(o, d, sc.tris + r, t, u, v, id)) {
                if (any) {
                    if (t < max_t) return res;             // something may be in the way: the walk decides
                } else {
                    if (t < t1) { t2 = t1; t1 = t; id1 = id; rec1 = r; }
                    else t2 = min_nn(t2, t);
                    limit = __double2float_ru(t1 + W);     // keep looking a window beyond the best
                }
            }
            cur = cstk[--csp];
        }
    }
    if (any || id1 == kNone) { res.status = 1u; return res; }   // no accepted record (below max_t) anywhere: None
    if (!(t2 > t1 + W)) return res;                              // (c) fails (a NaN anywhere fails it too)
    uint32_t node = __ldg(&sc.tris[rec1].node);
    while (node != 0u) {                                         // (d): up from C's node
        const DevNode* __restrict__ nd = sc.nodes + node;
        const uint32_t parent = __ldg(&nd->parent), k = __ldg(&nd->child_slot);
        double tmin, tmax;
        child_interval(ax, ay, az, sc.nodes + parent, k, tmin, tmax);
        if (COUNT) tc.box++;
        if (!(tmin >= 0.0 && tmin < tmax)) return res;
        node = parent;
    }
    res.t = t1; res.id = id1; res.rec = rec1; res.status = 1u;
    return res;
}

// What a caller allows traverse() to try before the walk
constexpr uint32_t kTryFilter = 1u;   // any-hit rays: scene_search(any)
constexpr uint32_t kTrySparse = 2u;   // closest-hit rays with max_t = +inf: scene_search(!any)

// Ray::intersect_with_octant_with_max_t(octree, 0, max_t).  outcome (may be null): 1 = answered by a search,
// 2 = a search ran and handed the ray to the walk, 0 = walk only.
template <bool COUNT>
__device__ __forceinline__ void traverse(const DevScene& sc, const V3& o, const V3& d, const double max_t,
                                         const bool any_hit, Hit& best, TraceCount& tc,
                                         const uint32_t allow = kTryFilter | kTrySparse, uint32_t* outcome = nullptr) {
    uint32_t oc = 0u;
    bool need_walk = true;
    const bool try_any = any_hit && (allow & kTryFilter);
    const bool try_closest = !any_hit && (allow & kTrySparse) && max_t == __longlong_as_double(0x7FF0000000000000ll);
    if (sc.root_tri_count == 0) {
        need_walk = false;   // ray.rs:112-114 at the root
        best.id = kNone; best.rec = 0; best.t = max_t;
    } else if (try_any || try_closest) {
        const SearchResult r = scene_search<COUNT>(sc, o, d, max_t, any_hit, tc);
        oc = r.status ? 1u : 2u;
        if (r.status) { need_walk = false; best.t = any_hit ? max_t : r.t; best.id = r.id; best.rec = r.rec; }
    }
    if (outcome) *outcome = oc;
    if (need_walk) walk<COUNT>(sc, o, d, max_t, any_hit, best, tc);
}

}  // namespace rrt

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

struct TraceCount {
    uint32_t mt = 0, box = 0, cull = 0, nodes = 0;
};

// One exact triangle test against the node-local best (strict '<', earlier list entry wins ties,
// ray.rs:123-127; lists are in ascending triangle id, so "earlier" == smaller id).
template <bool ORDERED>
__device__ __forceinline__ void test_record(const V3& o, const V3& d, const DevTri* rec, Hit& nb) {
    double t, u, v;
    uint32_t id;
    if (moller_trumbore(o, d, rec, t, u, v, id)) {
        const bool better = ORDERED ? (t < nb.t) : (t < nb.t || (t == nb.t && nb.id != kNone && id < nb.id));
        if (better) {
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
    int sp = 0;
    uint32_t node = 0;
    bool is_root = true;
    // reciprocal direction for the CONSERVATIVE culling boxes only (never for a reference decision)
    double ivx = 0.0, ivy = 0.0, ivz = 0.0;
    bool have_inv = false;

    for (;;) {
        const DevNode* __restrict__ n = sc.nodes + node;
        const uint4 m = __ldg(reinterpret_cast<const uint4*>(&n->first_child));  // first_child, child_mask, tri_first, tri_n
        const uint32_t cull_root = __ldg(&n->cull_root);
        if (COUNT) tc.nodes++;

        // ---- own triangles (ray.rs:116-129)
        Hit nb;
        nb.id = kNone;
        nb.t = is_root ? max_t : __longlong_as_double(0x7FF0000000000000ll);
        nb.u = nb.v = 0.0;
        if (cull_root == kNone) {
            const DevTri* rec = sc.tris + m.z;
            for (uint32_t i = 0; i < m.w; ++i) {
                if (COUNT) tc.mt++;
                test_record<true>(o, d, rec + i, nb);
            }
        } else {
            if (!have_inv) {
                have_inv = true;
                ivx = fmin(fmax(1.0 / d.x, -1e300), 1e300);
                ivy = fmin(fmax(1.0 / d.y, -1e300), 1e300);
                ivz = fmin(fmax(1.0 / d.z, -1e300), 1e300);
            }
            uint32_t cstk[64];
            int csp = 0;
            uint32_t cur = cull_root;
            for (;;) {
                const DevCull* __restrict__ c = sc.cull + cur;
                const double2* bp = reinterpret_cast<const double2*>(c);
                const uint4 ch = __ldg(reinterpret_cast<const uint4*>(&c->child[0]));  // child0, child1, count0, count1
                uint32_t go[2];
                double tn[2];
#pragma unroll
                for (int s = 0; s < 2; ++s) {
                    const double2 b0 = __ldg(bp + 3 * s), b1 = __ldg(bp + 3 * s + 1), b2 = __ldg(bp + 3 * s + 2);
                    if (COUNT) tc.cull++;
                    const double x0 = (b0.x - o.x) * ivx, x1 = (b1.y - o.x) * ivx;
                    const double y0 = (b0.y - o.y) * ivy, y1 = (b2.x - o.y) * ivy;
                    const double z0 = (b1.x - o.z) * ivz, z1 = (b2.y - o.z) * ivz;
                    const double tnear = fmax(fmax(fmin(x0, x1), fmin(y0, y1)), fmin(z0, z1));
                    const double tfar = fmin(fmin(fmax(x0, x1), fmax(y0, y1)), fmax(z0, z1));
                    // padded boxes + these slacks keep the test conservative w.r.t. rounding
                    const bool hit = (tnear <= tfar + fabs(tfar) * 1e-9 + 1e-300) && (tfar >= 0.0) &&
                                     (tnear - fabs(tnear) * 1e-9 <= nb.t);
                    go[s] = hit ? 1u : 0u;
                    tn[s] = tnear;
                }
                const uint32_t cidx[2] = {ch.x, ch.y}, ccnt[2] = {ch.z, ch.w};
                uint32_t inner[2];
                int n_inner = 0;
#pragma unroll
                for (int s = 0; s < 2; ++s) {
                    if (!go[s]) continue;
                    if (cidx[s] & 0x80000000u) {
                        const DevTri* rec = sc.tris + (cidx[s] & 0x7FFFFFFFu);
                        for (uint32_t i = 0; i < ccnt[s]; ++i) {
                            if (COUNT) tc.mt++;
                            test_record<false>(o, d, rec + i, nb);
                        }
                    } else {
                        inner[n_inner++] = (uint32_t)s;
                    }
                }
                if (n_inner == 2) {
                    const int nearer = tn[1] < tn[0] ? 1 : 0;
                    cstk[csp++] = cidx[1 - nearer];
                    cur = cidx[nearer];
                } else if (n_inner == 1) {
                    cur = cidx[inner[0]];
                } else {
                    if (csp == 0) break;
                    cur = cstk[--csp];
                }
                if (any_hit && nb.id != kNone && nb.t < best.t) break;  // best will improve => Some
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
            // uses the plane, so 9 true divisions stand for the reference's 48.
            const double ax0 = (q0.x - o.x) / d.x, ax1 = (q0.y - o.x) / d.x, ax2 = (q1.x - o.x) / d.x;
            const double ay0 = (q1.y - o.y) / d.y, ay1 = (q2.x - o.y) / d.y, ay2 = (q2.y - o.y) / d.y;
            const double az0 = (q3.x - o.z) / d.z, az1 = (q3.y - o.z) / d.z, az2 = (q4 - o.z) / d.z;
            const double mnx[2] = {fmin(ax0, ax1), fmin(ax1, ax2)}, mxx[2] = {fmax(ax0, ax1), fmax(ax1, ax2)};
            const double mny[2] = {fmin(ay0, ay1), fmin(ay1, ay2)}, mxy[2] = {fmax(ay0, ay1), fmax(ay1, ay2)};
            const double mnz[2] = {fmin(az0, az1), fmin(az1, az2)}, mxz[2] = {fmax(az0, az1), fmax(az1, az2)};
            double key[8];
            uint32_t hitmask = 0;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const int xh = (k == 2 || k == 3 || k == 6 || k == 7) ? 1 : 0, yh = k >= 4 ? 1 : 0,
                          zh = (k == 1 || k == 2 || k == 5 || k == 6) ? 1 : 0;
                const double tmin = fmax(fmax(mnx[xh], mny[yh]), mnz[zh]);  // ray.rs:29-32
                const double tmax = fmin(fmin(mxx[xh], mxy[yh]), mxz[zh]);  // ray.rs:33-36
                const bool kept = (m.y >> k) & 1u;
                const bool hit = kept && !(tmax < 0.0) && !(tmin > tmax);   // ray.rs:39-46
                key[k] = (tmin < 0.0) ? tmax : tmin;                         // ray.rs:49-53
                hitmask |= hit ? (1u << k) : 0u;
            }
            if (COUNT) tc.box += __popc(m.y);
            if (hitmask) {
                // stable sort by key (ray.rs:146-147) as a rank computation: rank_k = #{hit j that must
                // come before k} = #{j<k : !(key_k < key_j)} + #{j>k : key_j < key_k}
                uint32_t order = 0;
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    uint32_t rank = 0;
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        if (j == k) continue;
                        const bool before = (j < k) ? !(key[k] < key[j]) : (key[j] < key[k]);
                        rank += (before && ((hitmask >> j) & 1u)) ? 1u : 0u;
                    }
                    const uint32_t off = __popc(m.y & ((1u << k) - 1u));
                    order |= ((hitmask >> k) & 1u) ? (off << (3u * rank)) : 0u;
                }
                stk[sp++] = make_uint2(m.x, order | ((uint32_t)__popc(hitmask) << 24));
            }
        }

        // ---- next node: first remaining child of the deepest open level (ray.rs:152-161)
        while (sp > 0 && (stk[sp - 1].y >> 24) == 0u) --sp;
        if (sp == 0) return;
        uint2 top = stk[sp - 1];
        node = top.x + (top.y & 7u);
        const uint32_t cnt = (top.y >> 24) - 1u;
        stk[sp - 1].y = ((top.y & 0x00FFFFFFu) >> 3) | (cnt << 24);
        is_root = false;
    }
}

}  // namespace rrt

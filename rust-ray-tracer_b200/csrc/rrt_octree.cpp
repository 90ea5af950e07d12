// rrt_octree.cpp — exact host replay of the reference's incremental octree construction
// (src/collision/octree.rs:41-241, src/collision/aabb.rs:25-60) on flat arrays, plus flattening into
// the device layout (rrt_octree.h).  The tree SHAPE is part of the results contract: it decides which
// triangle wins a t-tie and which triangles a ray can reach at all.
#include "rrt_octree.h"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <cstring>
#include <numeric>

#include <omp.h>

namespace rrt {

namespace {

// aabb.rs:49-60 — closed-interval overlap, a = [amin,amax], b likewise (6 doubles min.xyz max.xyz)
inline bool boxes_overlap(const double* a, const double* b) {
    if (a[3] < b[0] || a[0] > b[3]) return false;
    if (a[4] < b[1] || a[1] > b[4]) return false;
    if (a[5] < b[2] || a[2] > b[5]) return false;
    return true;
}

// aabb.rs:25-47 — Rust f64::min/max ignore a NaN operand, exactly like fmin/fmax.
inline void tri_box(const double* v, double* b) {
    for (int k = 0; k < 3; ++k) {
        b[k] = std::fmin(v[k], std::fmin(v[3 + k], v[6 + k]));
        b[3 + k] = std::fmax(v[k], std::fmax(v[3 + k], v[6 + k]));
    }
}

}  // namespace

bool HostOctree::build(const double* v, uint64_t n, const double root_min[3], const double root_max[3],
                       std::string* err) {
    nodes.clear();
    next.assign(n, UINT32_MAX);
    n_tris = n;
    n_dropped = 0;
    max_depth = 0;
    HostNode root;
    for (int k = 0; k < 3; ++k) {
        root.box[k] = root_min[k];
        root.box[3 + k] = root_max[k];
    }
    nodes.push_back(root);  // octree.rs:23-39
    nodes.reserve(n / 2 + 16);

    auto append = [&](uint32_t node, uint32_t tri) {  // self.nodes[i].triangles.push(tri)
        HostNode& nd = nodes[node];
        if (nd.tail == UINT32_MAX)
            nd.head = tri;
        else
            next[nd.tail] = tri;
        nd.tail = tri;
        nd.list_len++;
    };

    for (uint64_t i = 0; i < n; ++i) {
        double tb[6];
        tri_box(v + 9 * i, tb);
        uint32_t node = 0;
        // push_at_octant (octree.rs:54-108) with the tail recursion unrolled into a loop.
        for (;;) {
            if (!boxes_overlap(tb, nodes[node].box)) {  // octree.rs:65,71-73 (only ever false at the root)
                if (node == 0) n_dropped++;
                break;
            }
            nodes[node].triangle_count += 1;  // octree.rs:75
            const bool leaf = nodes[node].first_child == 0;
            if (leaf && nodes[node].list_len == 0) {  // octree.rs:77-78: first triangle in an empty leaf stays
                append(node, (uint32_t)i);
                break;
            }
            if (leaf) {  // octree.rs:79-80 -> subdivide (octree.rs:121-241)
                if (nodes[node].depth + 1 > (uint32_t)kMaxOctreeDepth) {
                    if (err) *err = "octree deeper than kMaxOctreeDepth (coincident triangles?)";
                    return false;
                }
                const HostNode p = nodes[node];
                const double x0 = p.box[0], y0 = p.box[1], z0 = p.box[2];
                const double x1 = p.box[3], y1 = p.box[4], z1 = p.box[5];
                const double xm = x0 + (x1 - x0) / 2.0;  // octree.rs:136,142
                const double ym = y0 + (y1 - y0) / 2.0;  // octree.rs:137,144
                const double zm = z0 + (z1 - z0) / 2.0;  // octree.rs:138,146
                const uint32_t first = (uint32_t)nodes.size();
                // octree.rs:216-225: bbl, bfl, bfr, bbr, tbl, tfl, tfr, tbr (x-high for k in {2,3,6,7},
                // y-high for k >= 4, z-high for k in {1,2,5,6})
                for (int k = 0; k < 8; ++k) {
                    const bool xh = (k == 2 || k == 3 || k == 6 || k == 7), yh = k >= 4,
                               zh = (k == 1 || k == 2 || k == 5 || k == 6);
                    HostNode c;
                    c.box[0] = xh ? xm : x0;
                    c.box[3] = xh ? x1 : xm;
                    c.box[1] = yh ? ym : y0;
                    c.box[4] = yh ? y1 : ym;
                    c.box[2] = zh ? zm : z0;
                    c.box[5] = zh ? z1 : zm;
                    c.depth = p.depth + 1;
                    nodes.push_back(c);
                }
                nodes[node].first_child = first;
                max_depth = std::max(max_depth, p.depth + 1);
            }
            // octree.rs:82-104: descend iff exactly one child box overlaps the triangle box
            const uint32_t fc = nodes[node].first_child;
            int hits = 0;
            uint32_t only = 0;
            for (int k = 0; k < 8; ++k)
                if (boxes_overlap(nodes[fc + k].box, tb)) {
                    hits++;
                    only = fc + k;
                }
            if (hits == 1) {
                node = only;
                continue;
            }
            append(node, (uint32_t)i);
            break;
        }
    }
    return true;
}

// ---------------------------------------------------------------- parallel, order-exact construction
namespace {

struct TNode {
    double box[6];
    uint32_t depth = 0, count = 0, trigger = 0, first_child = 0;
    std::vector<uint32_t> list;   // own list, insertion order
    TNode* kids = nullptr;        // 8 children (octree.rs:216-225 order) once subdivided
};

struct ParBuild {
    const double* v;
    std::vector<std::vector<TNode*>> events;   // subdivided nodes, one bucket per OpenMP thread
    bool too_deep = false;

    // seq = the triangles that reach this node, in insertion order (all overlap its box)
    void run(TNode* nd, std::vector<uint32_t>& seq) {
        nd->count = (uint32_t)seq.size();  // octree.rs:75, once per arriving triangle
        if (seq.empty()) return;
        nd->list.push_back(seq[0]);  // octree.rs:77-78: the first triangle of an empty leaf stays for ever
        if (seq.size() == 1) return;
        if (nd->depth + 1 > (uint32_t)kMaxOctreeDepth) {
#pragma omp atomic write
            too_deep = true;
            return;
        }
        nd->trigger = seq[1];  // octree.rs:79-80: the second arrival subdivides
        nd->kids = new TNode[8];
        events[(size_t)omp_get_thread_num()].push_back(nd);
        const double x0 = nd->box[0], y0 = nd->box[1], z0 = nd->box[2], x1 = nd->box[3], y1 = nd->box[4], z1 = nd->box[5];
        const double xm = x0 + (x1 - x0) / 2.0, ym = y0 + (y1 - y0) / 2.0, zm = z0 + (z1 - z0) / 2.0;  // octree.rs:136-146
        for (int k = 0; k < 8; ++k) {
            const bool xh = (k == 2 || k == 3 || k == 6 || k == 7), yh = k >= 4, zh = (k == 1 || k == 2 || k == 5 || k == 6);
            TNode& c = nd->kids[k];
            c.box[0] = xh ? xm : x0; c.box[3] = xh ? x1 : xm;
            c.box[1] = yh ? ym : y0; c.box[4] = yh ? y1 : ym;
            c.box[2] = zh ? zm : z0; c.box[5] = zh ? z1 : zm;
            c.depth = nd->depth + 1;
        }
        std::vector<uint32_t> sub[8];
        // octree.rs:82-104 for every later arrival: the routing of a triangle depends on its own box only, so a
        // long sequence is classified in parallel (code 0..7 = the one child it overlaps, 8 = it stays here) and
        // then distributed in order.
        const size_t m = seq.size();
        auto classify = [&](size_t i) -> uint8_t {
            double tb[6];
            tri_box(v + 9 * (size_t)seq[i], tb);
            int hits = 0, only = 0;
            for (int k = 0; k < 8; ++k)
                if (boxes_overlap(nd->kids[k].box, tb)) { hits++; only = k; }
            return (uint8_t)(hits == 1 ? only : 8);
        };
        if (m > 200000) {
            std::vector<uint8_t> code(m);
            uint8_t* cp = code.data();
#pragma omp taskloop grainsize(32768) shared(cp)
            for (long long i = 1; i < (long long)m; ++i) cp[i] = classify((size_t)i);
            size_t cnt[9] = {0};
            for (size_t i = 1; i < m; ++i) cnt[code[i]]++;
            for (int k = 0; k < 8; ++k) sub[k].reserve(cnt[k]);
            nd->list.reserve(1 + cnt[8]);
            for (size_t i = 1; i < m; ++i) {
                if (code[i] < 8) sub[code[i]].push_back(seq[i]);
                else nd->list.push_back(seq[i]);
            }
        } else {
            for (size_t i = 1; i < m; ++i) {
                const uint8_t c = classify(i);
                if (c < 8) sub[c].push_back(seq[i]);
                else nd->list.push_back(seq[i]);
            }
        }
        std::vector<uint32_t>().swap(seq);
        for (int k = 0; k < 8; ++k) {
            if (sub[k].size() > 2048) {
                TNode* c = &nd->kids[k];
                std::vector<uint32_t>* sq = new std::vector<uint32_t>(std::move(sub[k]));
#pragma omp task firstprivate(c, sq)
                {
                    run(c, *sq);
                    delete sq;
                }
            } else {
                run(&nd->kids[k], sub[k]);
            }
        }
    }
};

void free_tree(TNode* kids) {
    if (!kids) return;
    for (int k = 0; k < 8; ++k) free_tree(kids[k].kids);
    delete[] kids;
}

}  // namespace

bool HostOctree::build_parallel(const double* v, uint64_t n, const double root_min[3], const double root_max[3],
                                std::string* err) {
    nodes.clear();
    next.assign(n, UINT32_MAX);
    n_tris = n;
    n_dropped = 0;
    max_depth = 0;
    TNode root;
    for (int k = 0; k < 3; ++k) { root.box[k] = root_min[k]; root.box[3 + k] = root_max[k]; }
    std::vector<uint32_t> seq;
    {   // octree.rs:65,71-73: a triangle whose box misses the root box is dropped
        std::vector<uint8_t> keep(n);
#pragma omp parallel for schedule(static)
        for (long long i = 0; i < (long long)n; ++i) {
            double tb[6];
            tri_box(v + 9 * (size_t)i, tb);
            keep[(size_t)i] = boxes_overlap(tb, root.box) ? 1 : 0;
        }
        seq.reserve(n);
        for (uint64_t i = 0; i < n; ++i) {
            if (keep[i]) seq.push_back((uint32_t)i);
            else n_dropped++;
        }
    }
    const bool verbose = std::getenv("RRT_VERBOSE") != nullptr;
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto secs = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) {
        return std::chrono::duration<double>(b - a).count();
    };
    const auto t0 = now();
    ParBuild pb{v};
    pb.events.resize((size_t)omp_get_max_threads());
#pragma omp parallel
#pragma omp single
    pb.run(&root, seq);
    const auto t1 = now();
    if (pb.too_deep) {
        free_tree(root.kids);
        if (err) *err = "octree deeper than kMaxOctreeDepth (coincident triangles?)";
        return false;
    }
    // subdivision events in the order the sequential build performs them = by triggering triangle
    std::vector<TNode*> ev;
    for (auto& b : pb.events) ev.insert(ev.end(), b.begin(), b.end());
    std::sort(ev.begin(), ev.end(), [](const TNode* a, const TNode* b) { return a->trigger < b->trigger; });
    for (size_t r = 0; r < ev.size(); ++r) ev[r]->first_child = (uint32_t)(1 + 8 * r);
    nodes.resize(1 + 8 * ev.size());
    auto emit = [&](const TNode& t, uint32_t idx) {
        HostNode h;
        std::memcpy(h.box, t.box, sizeof h.box);
        h.first_child = t.first_child;
        h.triangle_count = t.count;
        h.depth = t.depth;
        h.list_len = (uint32_t)t.list.size();
        if (!t.list.empty()) {
            h.head = t.list.front();
            h.tail = t.list.back();
            for (size_t j = 0; j + 1 < t.list.size(); ++j) next[t.list[j]] = t.list[j + 1];
        }
        nodes[idx] = h;
    };
    emit(root, 0);
    uint32_t md = 0;
#pragma omp parallel for schedule(static) reduction(max : md)
    for (long long r = 0; r < (long long)ev.size(); ++r) {
        for (int k = 0; k < 8; ++k) emit(ev[r]->kids[k], ev[r]->first_child + k);
        md = std::max(md, ev[r]->depth + 1);
    }
    max_depth = md;
    const auto t2 = now();
    free_tree(root.kids);
    if (verbose)
        std::fprintf(stderr, "[rrt] build_parallel: partition %.2fs, number+emit %.2fs, free %.2fs\n", secs(t0, t1), secs(t1, t2),
                     secs(t2, now()));
    return true;
}

uint64_t HostOctree::digest() const {
    uint64_t h = 1469598103934665603ull;
    auto mix = [&h](const void* p, size_t n) {
        const unsigned char* b = static_cast<const unsigned char*>(p);
        for (size_t i = 0; i < n; ++i) { h ^= b[i]; h *= 1099511628211ull; }
    };
    for (const HostNode& n : nodes) {
        mix(n.box, sizeof n.box);
        const uint32_t w[6] = {n.first_child, n.triangle_count, n.head, n.tail, n.list_len, n.depth};
        mix(w, sizeof w);
    }
    mix(next.data(), next.size() * sizeof(uint32_t));
    const uint64_t t[3] = {n_tris, n_dropped, max_depth};
    mix(t, sizeof t);
    return h;
}

void HostOctree::list_of(uint32_t node, std::vector<uint32_t>* out) const {
    out->clear();
    for (uint32_t t = nodes[node].head; t != UINT32_MAX; t = next[t]) out->push_back(t);
}

// ---------------------------------------------------------------- culling hierarchy (per long list)
namespace {

struct CullBuilder {
    const double* v;
    std::vector<uint32_t>& ids;       // triangle ids of this list, permuted in place into leaf order
    std::vector<double> boxes;        // per entry of ids (parallel array), 6 doubles
    std::vector<DevCull>& out;
    uint32_t rec_base;                // record index of ids[0]
    bool tight;                       // false: infinite boxes (culling off, every record is tested)
    uint32_t leaf_max;                // records per leaf (<= 8: 3-bit count field)

    double scale;                     // FlatScene::scale
    bool indirect = false;            // ids[] are RECORD indices and every leaf is one record (the scene-wide hierarchy)
    uint32_t* next = nullptr;         // indirect: non-null selects build_tasks (`out` is pre-sized, slots are positional)

#if RRT_CULL_F32
    void pad_box(double* b) const {
        const double p = kCullPadRel * scale;
        for (int k = 0; k < 3; ++k) {
            b[k] -= p;
            b[3 + k] += p;
        }
    }
    static float down(double x) {
        float f = (float)x;
        if ((double)f > x) f = std::nextafterf(f, -INFINITY);
        return f;
    }
    static float up(double x) {
        float f = (float)x;
        if ((double)f < x) f = std::nextafterf(f, INFINITY);
        return f;
    }
    static void store_box(DevCull& c, int s, const double* b) {
        for (int k = 0; k < 3; ++k) {
            c.box[s][k] = down(b[k]);
            c.box[s][3 + k] = up(b[3 + k]);
        }
    }
#else
    void pad_box(double* b) const {
        for (int k = 0; k < 3; ++k) {
            const double m = std::fmax(std::fabs(b[k]), std::fabs(b[3 + k]));
            const double p = 1e-7 * std::fmax(1.0, m);
            b[k] -= p;
            b[3 + k] += p;
        }
    }
    static void store_box(DevCull& c, int s, const double* b) { std::memcpy(c.box[s], b, 6 * sizeof(double)); }
#endif
    void bounds(uint32_t lo, uint32_t hi, double* b) const {
        if (!tight || !(scale < INFINITY)) {
            for (int k = 0; k < 3; ++k) { b[k] = -INFINITY; b[3 + k] = INFINITY; }
            return;
        }
        for (int k = 0; k < 3; ++k) { b[k] = INFINITY; b[3 + k] = -INFINITY; }
        for (uint32_t i = lo; i < hi; ++i)
            for (int k = 0; k < 3; ++k) {
                b[k] = std::fmin(b[k], boxes[6 * (size_t)i + k]);
                b[3 + k] = std::fmax(b[3 + k], boxes[6 * (size_t)i + 3 + k]);
            }
    }
    static double area(const double* b) {
        const double dx = b[3] - b[0], dy = b[4] - b[1], dz = b[5] - b[2];
        return dx * dy + dy * dz + dz * dx;
    }
    // Partition [lo,hi) by a binned surface-area heuristic on centroids; returns mid.
    uint32_t split(uint32_t lo, uint32_t hi) {
        double cb[6];
        for (int k = 0; k < 3; ++k) { cb[k] = INFINITY; cb[3 + k] = -INFINITY; }
        auto cen = [&](uint32_t i, int k) { return 0.5 * (boxes[6 * (size_t)i + k] + boxes[6 * (size_t)i + 3 + k]); };
        for (uint32_t i = lo; i < hi; ++i)
            for (int k = 0; k < 3; ++k) {
                cb[k] = std::fmin(cb[k], cen(i, k));
                cb[3 + k] = std::fmax(cb[3 + k], cen(i, k));
            }
        constexpr int NB = 16;
        double best_cost = INFINITY;
        int best_axis = -1, best_bin = 0;
        for (int ax = 0; ax < 3; ++ax) {
            const double ext = cb[3 + ax] - cb[ax];
            if (!(ext > 0.0)) continue;
            double bb[NB][6];
            uint32_t cnt[NB] = {0};
            for (int b = 0; b < NB; ++b)
                for (int k = 0; k < 3; ++k) { bb[b][k] = INFINITY; bb[b][3 + k] = -INFINITY; }
            for (uint32_t i = lo; i < hi; ++i) {
                int b = (int)((cen(i, ax) - cb[ax]) / ext * NB);
                b = std::min(NB - 1, std::max(0, b));
                cnt[b]++;
                for (int k = 0; k < 3; ++k) {
                    bb[b][k] = std::fmin(bb[b][k], boxes[6 * (size_t)i + k]);
                    bb[b][3 + k] = std::fmax(bb[b][3 + k], boxes[6 * (size_t)i + 3 + k]);
                }
            }
            double la[NB], ra[NB];
            uint32_t lc[NB], rc[NB];
            double acc[6];
            for (int k = 0; k < 3; ++k) { acc[k] = INFINITY; acc[3 + k] = -INFINITY; }
            uint32_t c = 0;
            for (int b = 0; b < NB; ++b) {
                for (int k = 0; k < 3; ++k) { acc[k] = std::fmin(acc[k], bb[b][k]); acc[3 + k] = std::fmax(acc[3 + k], bb[b][3 + k]); }
                c += cnt[b];
                la[b] = c ? area(acc) : 0.0;
                lc[b] = c;
            }
            for (int k = 0; k < 3; ++k) { acc[k] = INFINITY; acc[3 + k] = -INFINITY; }
            c = 0;
            for (int b = NB - 1; b >= 0; --b) {
                for (int k = 0; k < 3; ++k) { acc[k] = std::fmin(acc[k], bb[b][k]); acc[3 + k] = std::fmax(acc[3 + k], bb[b][3 + k]); }
                c += cnt[b];
                ra[b] = c ? area(acc) : 0.0;
                rc[b] = c;
            }
            for (int b = 0; b + 1 < NB; ++b) {
                if (lc[b] == 0 || rc[b + 1] == 0) continue;
                const double cost = la[b] * lc[b] + ra[b + 1] * rc[b + 1];
                if (cost < best_cost) { best_cost = cost; best_axis = ax; best_bin = b; }
            }
        }
        if (best_axis < 0) return lo + (hi - lo) / 2;  // all centroids coincide: split by count
        const double ext = cb[3 + best_axis] - cb[best_axis];
        // in-place partition of ids+boxes
        uint32_t i = lo, j = hi;
        while (i < j) {
            int b = (int)((cen(i, best_axis) - cb[best_axis]) / ext * NB);
            b = std::min(NB - 1, std::max(0, b));
            if (b <= best_bin) {
                ++i;
            } else {
                --j;
                std::swap(ids[i], ids[j]);
                for (int k = 0; k < 6; ++k) std::swap(boxes[6 * (size_t)i + k], boxes[6 * (size_t)j + k]);
            }
        }
        if (i == lo || i == hi) return lo + (hi - lo) / 2;
        return i;
    }
    static uint32_t leaf_word(uint32_t first, uint32_t n) { return 0x80000000u | ((n - 1u) << 28) | first; }
    // Builds the subtree over [lo,hi) (hi-lo > leaf_max) and returns its DevCull index. SAH splits can be
    // arbitrarily unbalanced, so below depth 30 the split is by count: total depth <= 30 + 28 < kMaxCullDepth.
    uint32_t build(uint32_t lo, uint32_t hi, int depth = 0) {
        if (next) return build_tasks(lo, hi, depth, 0u);
        const uint32_t me = (uint32_t)out.size();
        out.emplace_back();
        const uint32_t mid = (tight && depth < 30) ? split(lo, hi) : lo + (hi - lo) / 2;
        const uint32_t r[3] = {lo, mid, hi};
        for (int s = 0; s < 2; ++s) {
            double b[6];
            bounds(r[s], r[s + 1], b);
            pad_box(b);
            const uint32_t n = r[s + 1] - r[s];
            const uint32_t child = (n <= leaf_max) ? (indirect ? leaf_word(ids[r[s]], 1u) : leaf_word(rec_base + r[s], n))
                                                   : build(r[s], r[s + 1], depth + 1);
            DevCull& c = out[me];  // re-fetch: vector may have grown
            store_box(c, s, b);
            c.child[s] = child;
        }
        return me;
    }
    // Same tree, `out` pre-sized: a binary tree with one record per leaf over [lo, hi) has hi - lo - 1 inner nodes, so
    // the subtree over [lo, hi) owns the slots [slot, slot + hi - lo - 1) — its root at `slot`, the left half right
    // behind it, the right half behind that.  The numbering is therefore independent of scheduling, and the two halves
    // of a large range (split() permutes [lo, hi) only) are built as independent OpenMP tasks.
    uint32_t build_tasks(uint32_t lo, uint32_t hi, int depth, uint32_t slot) {
        const uint32_t me = slot;
        const uint32_t mid = (tight && depth < 30) ? split(lo, hi) : lo + (hi - lo) / 2;
        const uint32_t r[3] = {lo, mid, hi};
        const uint32_t child_slot[2] = {slot + 1u, slot + (mid - lo)};
        uint32_t child[2];
        for (int s = 0; s < 2; ++s) {
            double b[6];
            bounds(r[s], r[s + 1], b);
            pad_box(b);
            store_box(out[me], s, b);
            const uint32_t n = r[s + 1] - r[s];
            if (n <= leaf_max) {
                child[s] = leaf_word(ids[r[s]], 1u);
            } else if (n > 50000u) {
                uint32_t* dst = &child[s];
                const uint32_t a = r[s], e = r[s + 1], cs = child_slot[s];
#pragma omp task firstprivate(dst, a, e, depth, cs)
                *dst = build_tasks(a, e, depth + 1, cs);
            } else {
                child[s] = build_tasks(r[s], r[s + 1], depth + 1, child_slot[s]);
            }
        }
#pragma omp taskwait
        out[me].child[0] = child[0];
        out[me].child[1] = child[1];
        return me;
    }
};

}  // namespace

void flatten(const HostOctree& t, const double* v, uint32_t cull_threshold, uint32_t cull_leaf, FlatScene* out,
             bool global_hierarchy) {
    out->nodes.clear();
    out->tris.clear();
    out->cull.clear();
    out->n_stored = out->root_list = out->max_list = 0;
    out->global_root = 0;
    out->sparse_ok = false;
    out->planes_safe = true;
    out->max_depth = t.max_depth;
    if (t.nodes.empty() || t.nodes[0].triangle_count == 0) {
        // ray.rs:112-114: an empty root returns None for every ray; keep one inert node.
        DevNode n{};
        n.cull_root = UINT32_MAX;
        out->nodes.push_back(n);
        return;
    }
    // breadth-first renumbering of kept (triangle_count > 0) nodes; siblings stay consecutive
    std::vector<uint32_t> order;  // device index -> host index
    order.push_back(0);
    out->nodes.emplace_back();
    for (size_t d = 0; d < order.size(); ++d) {
        const HostNode& h = t.nodes[order[d]];
        DevNode dn{};
        dn.cull_root = UINT32_MAX;
        if (h.first_child != 0) {
            const HostNode* c = &t.nodes[h.first_child];
            // planes exactly as subdivide() produced them: child 0 = (min .. min+half), child 6 = (min+half .. max)
            dn.px[0] = c[0].box[0]; dn.px[1] = c[0].box[3]; dn.px[2] = c[6].box[3];
            dn.py[0] = c[0].box[1]; dn.py[1] = c[0].box[4]; dn.py[2] = c[6].box[4];
            dn.pz[0] = c[0].box[2]; dn.pz[1] = c[0].box[5]; dn.pz[2] = c[6].box[5];
            dn.first_child = (uint32_t)order.size();
            for (int k = 0; k < 8; ++k)
                if (c[k].triangle_count > 0) {
                    dn.child_mask |= 1u << k;
                    order.push_back(h.first_child + k);
                    out->nodes.emplace_back();
                }
        }
        out->nodes[d] = dn;
    }
    // range condition of the FMA quotient path (rrt_trace.cuh): every plane is 0 or in [2^-400, 2^401)
    out->planes_safe = true;
    for (const DevNode& dn : out->nodes)
        if (dn.child_mask)
            for (int k = 0; k < 9; ++k) {
                const double p = k < 3 ? dn.px[k] : (k < 6 ? dn.py[k - 3] : dn.pz[k - 6]);
                const double a = std::fabs(p);
                if (!(p == 0.0 || (a >= std::ldexp(1.0, -400) && a < std::ldexp(1.0, 401)))) out->planes_safe = false;
                // the sign-selected near/far planes of the fast path need min <= mid <= max on every axis
                if (k % 3 != 0) {
                    const double prev = k < 3 ? dn.px[k - 1] : (k < 6 ? dn.py[k - 4] : dn.pz[k - 7]);
                    if (!(prev <= p)) out->planes_safe = false;
                }
            }
    // coordinate scale of everything a culling hierarchy can contain
    out->scale = 1.0;
    for (size_t d = 0; d < order.size(); ++d)
        for (uint32_t tr = t.nodes[order[d]].head; tr != UINT32_MAX; tr = t.next[tr])
            for (int k = 0; k < 9; ++k) {
                const double a = std::fabs(v[9 * (size_t)tr + k]);
                if (a > out->scale) out->scale = a;   // NaN / inf coordinates: see below
            }
    if (!(out->scale < 1e12)) out->scale = INFINITY;   // culling then pads to infinite boxes: every record is tested
    // triangle records and culling hierarchies, node by node.  Lists are independent: every list builds
    // its hierarchy into its own vector (indices local to the list) on an OpenMP thread, then the pieces are
    // concatenated in node order, which makes the result independent of the thread count.
    const size_t n_dev = order.size();
    std::vector<uint64_t> tri_off(n_dev + 1, 0);
    for (size_t d = 0; d < n_dev; ++d) tri_off[d + 1] = tri_off[d] + t.nodes[order[d]].list_len;
    out->n_stored = tri_off[n_dev];
    out->tris.resize(tri_off[n_dev]);
    std::vector<std::vector<DevCull>> pieces(n_dev);
    std::vector<double> own_box(6 * n_dev);   // unpadded f64 box of every own list (empty: +inf/-inf)
    const uint32_t leaf_short = cull_threshold ? std::min(cull_threshold, 8u) : 8u;
    const bool tight = cull_threshold != 0;
    const uint32_t leaf_max = tight ? std::max(1u, std::min(cull_leaf, 8u)) : 8u;
#pragma omp parallel
    {
        std::vector<uint32_t> ids;
#pragma omp for schedule(dynamic, 64)
        for (long long dd = 0; dd < (long long)n_dev; ++dd) {
            const size_t d = (size_t)dd;
            t.list_of(order[d], &ids);
            DevNode& dn = out->nodes[d];
            dn.tri_first = (uint32_t)tri_off[d];
            dn.tri_n = (uint32_t)ids.size();
            for (int k = 0; k < 3; ++k) { dn.list_box[k] = -INFINITY; dn.list_box[3 + k] = INFINITY; }
            {
                double* ob = &own_box[6 * d];
                double tb[6];
                for (int k = 0; k < 3; ++k) { ob[k] = INFINITY; ob[3 + k] = -INFINITY; }
                for (uint32_t id : ids) {
                    tri_box(v + 9 * (size_t)id, tb);
                    for (int k = 0; k < 3; ++k) { ob[k] = std::fmin(ob[k], tb[k]); ob[3 + k] = std::fmax(ob[3 + k], tb[3 + k]); }
                }
            }
#if RRT_CULL_F32
            if (!ids.empty() && tight && out->scale < INFINITY) {
                double b[6];
                std::memcpy(b, &own_box[6 * d], sizeof b);
                const double pad = kCullPadRel * out->scale;
                for (int k = 0; k < 3; ++k) {
                    dn.list_box[k] = CullBuilder::down(b[k] - pad);
                    dn.list_box[3 + k] = CullBuilder::up(b[3 + k] + pad);
                }
            }
#endif
            if (ids.empty()) {
                dn.cull_root = UINT32_MAX;
            } else if (ids.size() <= leaf_short) {
                dn.cull_root = CullBuilder::leaf_word(dn.tri_first, (uint32_t)ids.size());  // short list: one leaf
            } else {
                std::vector<double> boxes(6 * ids.size());
                for (size_t i = 0; i < ids.size(); ++i) tri_box(v + 9 * (size_t)ids[i], &boxes[6 * i]);
                CullBuilder cb{v, ids, std::move(boxes), pieces[d], dn.tri_first, tight, leaf_max, out->scale};
                dn.cull_root = cb.build(0, (uint32_t)ids.size());   // index local to pieces[d]; rebased below
            }
            DevTri* rec = out->tris.data() + tri_off[d];
            for (uint32_t id : ids) {
                const double* p = v + 9 * (size_t)id;
                DevTri r{};
                for (int k = 0; k < 3; ++k) {
                    r.v1[k] = p[k];
                    r.e1[k] = p[3 + k] - p[k];  // ray.rs:60
                    r.e2[k] = p[6 + k] - p[k];  // ray.rs:61
                }
                r.id = id;
                *rec++ = r;
            }
        }
    }
    for (size_t d = 0; d < n_dev; ++d) {
        const uint64_t len = t.nodes[order[d]].list_len;
        if (order[d] == 0) out->root_list = len;
        else out->max_list = std::max<uint64_t>(out->max_list, len);
    }
    // subtree boxes, bottom-up: kept children have larger device indices than their parent (breadth-first order)
    out->sub_boxes.assign(n_dev, DevSubBox{});
    for (size_t d = n_dev; d-- > 0;) {
        double* b = &own_box[6 * d];
        const DevNode& dn = out->nodes[d];
        const uint32_t nk = (uint32_t)__builtin_popcount(dn.child_mask);
        for (uint32_t c = 0; c < nk; ++c) {
            const double* cb = &own_box[6 * (size_t)(dn.first_child + c)];   // already the child's SUBTREE box
            for (int k = 0; k < 3; ++k) { b[k] = std::fmin(b[k], cb[k]); b[3 + k] = std::fmax(b[3 + k], cb[3 + k]); }
        }
        DevSubBox& sb = out->sub_boxes[d];
        for (int k = 0; k < 3; ++k) { sb.box[k] = -INFINITY; sb.box[3 + k] = INFINITY; }
#if RRT_CULL_F32
        if (tight && out->scale < INFINITY) {
            const double pad = kCullPadRel * out->scale;
            for (int k = 0; k < 3; ++k) {
                sb.box[k] = CullBuilder::down(b[k] - pad);
                sb.box[3 + k] = CullBuilder::up(b[3 + k] + pad);
            }
        }
#endif
    }
    // concatenate the hierarchies; inner-node references (bit31 clear) move by the piece's base
    size_t n_cull = 0;
    for (const auto& pc : pieces) n_cull += pc.size();
    out->cull.reserve(n_cull);
    for (size_t d = 0; d < n_dev; ++d) {
        std::vector<DevCull>& pc = pieces[d];
        if (pc.empty()) continue;
        const uint32_t base = (uint32_t)out->cull.size();
        out->nodes[d].cull_root += base;   // a list with a hierarchy has an inner node as its root
        for (DevCull& c : pc) {
            for (int k = 0; k < 2; ++k)
                if (!(c.child[k] & 0x80000000u)) c.child[k] += base;
            out->cull.push_back(c);
        }
        std::vector<DevCull>().swap(pc);
    }
    // The way UP from a record: its node, a node's parent and which reference child of it the node is; and whether
    // every stored triangle lies inside the box of the node that stores it (false when a triangle pokes out of the
    // root box, utils.rs:145 / octree.rs:71-73) with a root box of sane size — precondition (a) of scene_search for camera rays.
    {
        bool inside = true;
        for (size_t d = 0; d < out->nodes.size(); ++d) {
            DevNode& dn = out->nodes[d];
            uint32_t rank = 0;
            for (int k = 0; k < 8; ++k)
                if (dn.child_mask & (1u << k)) {
                    DevNode& ch = out->nodes[dn.first_child + rank++];
                    ch.parent = (uint32_t)d;
                    ch.child_slot = (uint32_t)k;
                }
            const double* nb = t.nodes[order[d]].box;
            for (uint32_t r = dn.tri_first; r < dn.tri_first + dn.tri_n; ++r) {
                out->tris[r].node = (uint32_t)d;
                double tb[6];
                tri_box(v + 9 * (size_t)out->tris[r].id, tb);
                for (int k = 0; k < 3; ++k)
                    if (!(tb[k] >= nb[k] && tb[3 + k] <= nb[3 + k])) inside = false;
            }
        }
        double rmax = 0.0;
        for (int k = 0; k < 6; ++k) rmax = std::fmax(rmax, std::fabs(t.nodes[0].box[k]));
        out->sparse_ok = inside && rmax <= 1e4 * out->scale;
    }
    // ONE more culling hierarchy over ALL records of the scene, every leaf a single record, for the searches that go
    // round the octree walk (rrt_trace.cuh: scene_search) — which record belongs to which octree node
    // does not matter to their questions.  n - 1 inner nodes, built as OpenMP tasks on disjoint ranges.
    if (global_hierarchy && tight && out->scale < INFINITY && out->tris.size() > 1 && out->tris.size() < (1u << 28)) {
        const size_t nr = out->tris.size();
        std::vector<uint32_t> recs(nr);
        std::vector<double> boxes(6 * nr);
#pragma omp parallel for schedule(static)
        for (long long r = 0; r < (long long)nr; ++r) {
            recs[r] = (uint32_t)r;
            tri_box(v + 9 * (size_t)out->tris[r].id, &boxes[6 * (size_t)r]);
        }
        std::vector<DevCull> piece(nr - 1);
        uint32_t next = 0;
        CullBuilder cb{v, recs, std::move(boxes), piece, 0u, true, 1u, out->scale, true, &next};
        uint32_t root = 0;
#pragma omp parallel
#pragma omp single
        root = cb.build(0, (uint32_t)nr);
        const uint32_t base = (uint32_t)out->cull.size();
        out->cull.reserve(out->cull.size() + piece.size());
        for (DevCull& c : piece) {
            for (int k = 0; k < 2; ++k)
                if (!(c.child[k] & 0x80000000u)) c.child[k] += base;
            out->cull.push_back(c);
        }
        out->global_root = root + base + 1u;
    }
}

}  // namespace rrt

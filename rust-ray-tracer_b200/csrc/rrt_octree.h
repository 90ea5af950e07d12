// rrt_octree.h — host-side octree of the reference (src/collision/octree.rs) as flat arrays, and the
// pointer-free device layout it is flattened into.  Product code (not the oracle).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace rrt {

constexpr int kMaxOctreeDepth = 64;
constexpr int kMaxCullDepth = 62;     // depth bound of a per-list culling hierarchy (builder enforces it)  // reference recursion is unbounded; deeper trees are refused (RRT_ERR_DEPTH)

// One reference OctantNode (octree.rs:7-12) without heap vectors: children are the 8 consecutive nodes
// starting at first_child (octree.rs:226-237); the triangle list is an intrusive singly linked list
// threaded through HostOctree::next (each triangle lives in at most one list).
struct HostNode {
    double box[6];             // min.xyz, max.xyz  (aabb.rs:4-7)
    uint32_t first_child = 0;  // 0 = leaf
    uint32_t triangle_count = 0;  // octree.rs:11 (everything that passed through, not the list length)
    uint32_t head = UINT32_MAX, tail = UINT32_MAX;  // own list in insertion order
    uint32_t list_len = 0;
    uint32_t depth = 0;
};

struct HostOctree {
    std::vector<HostNode> nodes;
    std::vector<uint32_t> next;  // per triangle: next triangle in the same node list
    uint64_t n_tris = 0, n_dropped = 0;
    uint32_t max_depth = 0;

    // Sequential insertion, bit-identical to Octree::push_triangle for i = 0..n-1 (octree.rs:41-108).
    // v = n x 9 doubles. Returns false (err set) if the tree exceeds kMaxOctreeDepth.
    bool build(const double* v, uint64_t n, const double root_min[3], const double root_max[3], std::string* err);
    // The same tree — same boxes, lists, counts AND node numbering — built top-down on OpenMP tasks.
    // Why it is the same tree: once a node is subdivided, a triangle is routed by its own box alone (into
    // the one child it overlaps, or it stays), so the subsequence of triangles entering a child evolves
    // independently of its siblings; a leaf keeps its first triangle and is subdivided by its second
    // (octree.rs:77-80), which then lands in a fresh empty child or stays — so every subdivision has ONE
    // triggering triangle, and the sequential build appends the 8 children in order of that triangle's index.
    bool build_parallel(const double* v, uint64_t n, const double root_min[3], const double root_max[3], std::string* err);
    uint64_t digest() const;   // FNV-1a over nodes and list links (tests: sequential == parallel)
    void list_of(uint32_t node, std::vector<uint32_t>* out) const;
};

// ---------------------------------------------------------------- device layout

// Node kept on the device (only nodes with triangle_count > 0; empty ones return None at ray.rs:112
// and the child sort is stable, so dropping them cannot change a result).  128 bytes = one cache line.
struct alignas(128) DevNode {
    uint32_t first_child;        // device index of the first kept child (kept children are consecutive)
    uint32_t child_mask;         // bit k = reference child k (order octree.rs:216-225) is kept
    uint32_t tri_first, tri_n;   // own triangle records [tri_first, tri_first + tri_n)
    double px[3], py[3], pz[3];  // the 3 planes per axis shared by the 8 child boxes: min, min+half, max
                                 // (octree.rs:136-210) — 9 distinct slab numerators instead of 48
    uint32_t cull_root;          // entry of this list's culling hierarchy: a DevCull index, or a leaf word
                                 // (bit31 | (count-1) << 28 | first record) when the list is short
    uint32_t parent;             // device index of the parent node (root: 0) ...
    float list_box[6];           // conservative (padded, outward-rounded) f32 box of the node's OWN list:
                                 // a ray that misses it cannot be accepted by any record of the list
    uint32_t child_slot;         // ... and which reference child (0..7, octree.rs:216-225) of it this node is
    uint32_t _pad;
};
static_assert(sizeof(DevNode) == 128, "DevNode layout");

// Intersection record, 80 bytes = 5 x 16-byte vector loads. e1/e2 are the reference's own
// `v2 - v1`, `v3 - v1` (ray.rs:60-61) evaluated once on the host in f64: same operands, same bits.
struct alignas(16) DevTri {
    double v1[3], e1[3], e2[3];
    uint32_t id;   // index in `f`-line order (what hit-id planes report; also the in-list tie order)
    uint32_t node; // device index of the octree node whose list holds this record
};
static_assert(sizeof(DevTri) == 80, "DevTri layout");

// Shading record by ORIGINAL triangle id, 128 bytes, touched only on hits (entities.rs:78-84).
struct alignas(16) DevShade {
    double n1[3], n2[3], n3[3];
    double t1[2], t2[2], t3[2];
    int32_t material;
    int32_t _pad;
};
static_assert(sizeof(DevShade) == 128, "DevShade layout");

// Culling hierarchy over one long node list (binary BVH, both child boxes in the parent). Every candidate it
// yields is still decided by the exact f64 Moller-Trumbore test; the boxes only have to be CONSERVATIVE.
#ifndef RRT_CULL_F32
#define RRT_CULL_F32 1
#endif
#if RRT_CULL_F32
// f32 boxes, rounded outward after padding by kCullPadRel * S (S = scene coordinate scale, FlatScene::scale):
// 64 bytes = 4 x 16-byte loads, tested with FP32 FFMA/FMNMX (full rate) instead of the FP64 pipe.
// Error budget and the ray eligibility rule are in rrt_trace.cuh (CullRay).
using cull_real = float;
struct alignas(16) DevCull {
    float box[2][6];      // child 0 / child 1 : min.xyz max.xyz (padded, rounded outward)
    uint32_t child[2];    // bit31 set: leaf = bit31 | (count-1) << 28 | first record; else index of DevCull
    uint32_t _pad[2];
};
static_assert(sizeof(DevCull) == 64, "DevCull layout");
constexpr double kCullPadRel = 1e-5;
constexpr double kCullOriginRel = 16.0;   // rays with |origin component| > kCullOriginRel * S are not culled
#else
using cull_real = double;
struct alignas(16) DevCull {
    double box[2][6];     // child 0 / child 1 : min.xyz max.xyz (padded by 1e-7 relative)
    uint32_t child[2];
    uint32_t _pad[2];
};
static_assert(sizeof(DevCull) == 112, "DevCull layout");
#endif

// Conservative (padded, outward-rounded) f32 box of EVERYTHING below a node: its own list and the lists of all
// its descendants.  A ray that misses it gets None from the whole subtree (ray.rs:104-168 can only return a
// triangle the exact test accepts), so the node is skipped on entry.  32 bytes, parallel to nodes[].
struct alignas(32) DevSubBox {
    float box[6];
    uint32_t _pad[2];
};
static_assert(sizeof(DevSubBox) == 32, "DevSubBox layout");

struct FlatScene {
    std::vector<DevNode> nodes;
    std::vector<DevSubBox> sub_boxes;
    std::vector<DevTri> tris;
    std::vector<DevCull> cull;
    uint64_t n_stored = 0, root_list = 0, max_list = 0;
    uint32_t max_depth = 0;
    bool planes_safe = true;   // every octree plane is 0 or has magnitude in [2^-400, 2^401), and the three planes
                               // of every axis are ordered min <= mid <= max (rrt_trace.cuh)
    double scale = 1.0;        // S = max(1, largest |coordinate| of a stored triangle): culling pad / eligibility unit
    uint32_t global_root = 0;  // 1 + DevCull index of the root of the scene-wide hierarchy over ALL records (0 = none)
    bool sparse_ok = false;    // every stored triangle lies inside its node's box and the root box is at most 1e4 S
                               // across: precondition (a) of scene_search for camera rays (rrt_trace.cuh)
};

// Flatten: renumber kept nodes breadth-first (siblings consecutive), concatenate node lists.
// cull_threshold (1..8): lists longer than this get a culling hierarchy with tight (padded) boxes whose
// leaves hold at most cull_leaf (1..8) records; shorter lists are one leaf (every record is tested).
// 0 = culling off: long lists still go through the same hierarchy code, but with infinite boxes, so
// every record of the list is tested (the exact linear scan of the reference, in another order).
// global_hierarchy: also build the scene-wide hierarchy (one more DevCull per record) for the searches that go round
// the octree walk (rrt_trace.cuh: scene_search).
void flatten(const HostOctree& t, const double* v, uint32_t cull_threshold, uint32_t cull_leaf, FlatScene* out,
             bool global_hierarchy = true);

}  // namespace rrt

// rrt_oracle.cpp — CPU restatement of the reference's per-pixel path.  TEST INFRASTRUCTURE ONLY.
//
// This file is the checker, not the product: only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may load liboracle.  The product (rust-ray-tracer_b200/)
// never links, imports or calls it.
//
// PARITY PIN STATUS: the reference ships no live tests, golden vectors or fixtures for this path
//  and cannot be compiled here (no Rust toolchain) — so this oracle is pinned only
// by (i) the two known answers recoverable from the dead test module src/collision/octree.rs:244-652
// (first triangle lands at the root; child-box order bbl,bfl,bfr,bbr,tbl,tfl,tfr,tbr), (ii) an
// independent pure-Python transcription (oracle/py_oracle.py) and (iii) a brute-force all-triangle
// closest hit.  Against the real Rust binary: "parity unpinned".
//
// Every function cites the reference file:line it restates (paths under /root/reference).
// Build: g++ -O2 -std=c++17 -ffp-contract=off -fno-fast-math -fopenmp -shared -fPIC (oracle/Makefile).
// Rust never contracts a*b+c into an FMA, hence -ffp-contract=off.

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <memory>
#include <string>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "../include/rrt.h"

namespace {

// ---------------------------------------------------------------- Vector3d  (src/scene/engine.rs:9-104)
struct V3 {
    double x, y, z;
};
inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }      // engine.rs:16-26
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }      // engine.rs:36-46
inline V3 operator*(V3 a, double s) { return {a.x * s, a.y * s, a.z * s}; }        // engine.rs:48-58
inline V3 operator/(V3 a, double s) { return {a.x / s, a.y / s, a.z / s}; }        // engine.rs:72-82
inline V3 operator-(V3 a) { return {-a.x, -a.y, -a.z}; }                           // engine.rs:60-70
inline double dot(V3 a, V3 b) { return (a.x * b.x) + (a.y * b.y) + (a.z * b.z); }  // engine.rs:85-87
inline double length(V3 a) { return std::sqrt(a.x * a.x + a.y * a.y + a.z * a.z); }  // engine.rs:89-91
inline V3 cross(V3 a, V3 b) {                                                       // engine.rs:93-99
    return {a.y * b.z - a.z * b.y, -(a.x * b.z - a.z * b.x), a.x * b.y - a.y * b.x};
}
inline V3 normalised(V3 a) { return a / length(a); }                               // engine.rs:101-103

// Rust f64::min / f64::max: the non-NaN operand wins (std::fmin/fmax have the same contract).
inline double rmin(double a, double b) { return std::fmin(a, b); }
inline double rmax(double a, double b) { return std::fmax(a, b); }

constexpr double kEps = 2.220446049250313e-16;  // f64::EPSILON (ray.rs:66,89)

// ---------------------------------------------------------------- scene data
struct Color {
    uint8_t r, g, b;
};  // entities.rs:11-16
struct Texture {
    std::vector<Color> colours;
    size_t width = 0, height = 0;
};  // entities.rs:87-91
struct Material {                  // material.rs:11-22
    V3 ka, kd, ks;
    double ns, kr;
    int tex, bump;  // indices into textures, bump = -1 for None
};
struct Triangle {                  // entities.rs:73-84
    V3 v1, v2, v3;
    V3 t1, t2, t3;  // tex coords (z unused)
    V3 n1, n2, n3;
    int material;
};
struct Aabb {
    V3 mn, mx;
};  // aabb.rs:4-7

// aabb.rs:25-47
Aabb aabb_from_triangle(const Triangle& t) {
    Aabb b;
    b.mn.x = rmin(t.v1.x, rmin(t.v2.x, t.v3.x));
    b.mx.x = rmax(t.v1.x, rmax(t.v2.x, t.v3.x));
    b.mn.y = rmin(t.v1.y, rmin(t.v2.y, t.v3.y));
    b.mx.y = rmax(t.v1.y, rmax(t.v2.y, t.v3.y));
    b.mn.z = rmin(t.v1.z, rmin(t.v2.z, t.v3.z));
    b.mx.z = rmax(t.v1.z, rmax(t.v2.z, t.v3.z));
    return b;
}
// aabb.rs:49-60 (closed intervals)
bool aabb_intersects(const Aabb& a, const Aabb& o) {
    if (a.mx.x < o.mn.x || a.mn.x > o.mx.x) return false;
    if (a.mx.y < o.mn.y || a.mn.y > o.mx.y) return false;
    if (a.mx.z < o.mn.z || a.mn.z > o.mx.z) return false;
    return true;
}

// ---------------------------------------------------------------- octree  (src/collision/octree.rs:7-241)
struct OctantNode {
    size_t aabb_index;
    std::vector<size_t> triangles;
    std::vector<size_t> children;
    size_t triangle_count;
};
struct Octree {
    std::vector<OctantNode> nodes;
    std::vector<Aabb> aabbs;  // node boxes AND triangle boxes interleaved (octree.rs:44-49,227-229)
    std::vector<Triangle> triangles;
    std::vector<size_t> triangle_aabbs;

    void init(const double mn[3], const double mx[3]) {  // octree.rs:23-39
        nodes.clear();
        aabbs.clear();
        triangles.clear();
        triangle_aabbs.clear();
        aabbs.push_back(Aabb{{mn[0], mn[1], mn[2]}, {mx[0], mx[1], mx[2]}});
        nodes.push_back(OctantNode{0, {}, {}, 0});
    }
    void push_triangle(const Triangle& t) {  // octree.rs:41-52
        Aabb tb = aabb_from_triangle(t);
        size_t aabb_index = aabbs.size();
        size_t triangle_index = triangles.size();
        triangles.push_back(t);
        aabbs.push_back(tb);
        triangle_aabbs.push_back(aabb_index);
        push_at_octant(triangle_index, aabb_index, 0);
    }
    bool octant_hits_tri_box(size_t octant, size_t tri_aabb) const {  // octree.rs:110-119
        return aabb_intersects(aabbs[nodes[octant].aabb_index], aabbs[tri_aabb]);
    }
    void push_at_octant(size_t tri, size_t tri_aabb, size_t octant) {  // octree.rs:54-108
        bool intersects = aabb_intersects(aabbs[tri_aabb], aabbs[nodes[octant].aabb_index]);
        bool has_triangle = !nodes[octant].triangles.empty();
        std::vector<size_t> children = nodes[octant].children;
        bool is_leaf = children.empty();
        if (!intersects) return;
        nodes[octant].triangle_count += 1;
        if (is_leaf && !has_triangle) {
            nodes[octant].triangles.push_back(tri);
            return;
        }
        if (is_leaf) children = subdivide(octant);
        std::vector<size_t> hit;
        for (size_t c : children)
            if (octant_hits_tri_box(c, tri_aabb)) hit.push_back(c);
        if (hit.size() == 1)
            push_at_octant(tri, tri_aabb, hit[0]);
        else
            nodes[octant].triangles.push_back(tri);
    }
    std::vector<size_t> subdivide(size_t octant) {  // octree.rs:121-241
        Aabb b = aabbs[nodes[octant].aabb_index];
        double x0 = b.mn.x, y0 = b.mn.y, z0 = b.mn.z, x1 = b.mx.x, y1 = b.mx.y, z1 = b.mx.z;
        double hx = (x1 - x0) / 2.0, hy = (y1 - y0) / 2.0, hz = (z1 - z0) / 2.0;
        double xm = x0 + hx, ym = y0 + hy, zm = z0 + hz;
        // order octree.rs:216-225: bbl, bfl, bfr, bbr, tbl, tfl, tfr, tbr  (front = +z, top = +y, right = +x)
        Aabb kids[8] = {
            {{x0, y0, z0}, {xm, ym, zm}},  // bottom_back_left   octree.rs:140-147
            {{x0, y0, zm}, {xm, ym, z1}},  // bottom_front_left  octree.rs:149-156
            {{xm, y0, zm}, {x1, ym, z1}},  // bottom_front_right octree.rs:158-165
            {{xm, y0, z0}, {x1, ym, zm}},  // bottom_back_right  octree.rs:167-174
            {{x0, ym, z0}, {xm, y1, zm}},  // top_back_left      octree.rs:176-183
            {{x0, ym, zm}, {xm, y1, z1}},  // top_front_left     octree.rs:185-192
            {{xm, ym, zm}, {x1, y1, z1}},  // top_front_right    octree.rs:194-201
            {{xm, ym, z0}, {x1, y1, zm}},  // top_back_right     octree.rs:203-210
        };
        std::vector<size_t> out;
        nodes[octant].children.clear();
        for (const Aabb& k : kids) {  // octree.rs:216-238
            size_t ni = nodes.size(), ai = aabbs.size();
            aabbs.push_back(k);
            nodes.push_back(OctantNode{ai, {}, {}, 0});
            nodes[octant].children.push_back(ni);
            out.push_back(ni);
        }
        return out;
    }
};

// ---------------------------------------------------------------- work counters
struct Counters {
    uint64_t n_a = 0, n_u = 0, n_v = 0, n_t = 0, n_slab = 0, n_node = 0;
    uint64_t rays_primary = 0, rays_shadow = 0, rays_reflect = 0;
    void add(const Counters& o) {
        n_a += o.n_a; n_u += o.n_u; n_v += o.n_v; n_t += o.n_t; n_slab += o.n_slab; n_node += o.n_node;
        rays_primary += o.rays_primary; rays_shadow += o.rays_shadow; rays_reflect += o.rays_reflect;
    }
};

// ---------------------------------------------------------------- ray  (src/collision/ray.rs)
struct Hit {
    bool some = false;
    double t = 0, u = 0, v = 0;
    size_t tri = 0;
};
struct Ray {
    V3 origin, direction;

    // ray.rs:21-54 — returns the sort key (tmin, or tmax when the origin is inside / tmin<0)
    bool intersect_aabb(const Aabb& b, double* key) const {
        double t1 = (b.mn.x - origin.x) / direction.x;
        double t2 = (b.mx.x - origin.x) / direction.x;
        double t3 = (b.mn.y - origin.y) / direction.y;
        double t4 = (b.mx.y - origin.y) / direction.y;
        double t5 = (b.mn.z - origin.z) / direction.z;
        double t6 = (b.mx.z - origin.z) / direction.z;
        double tmin = rmax(rmax(rmin(t1, t2), rmin(t3, t4)), rmin(t5, t6));
        double tmax = rmin(rmin(rmax(t1, t2), rmax(t3, t4)), rmax(t5, t6));
        if (tmax < 0.0) return false;
        if (tmin > tmax) return false;
        *key = (tmin < 0.0) ? tmax : tmin;
        return true;
    }

    // ray.rs:56-94 — Moller-Trumbore, two-sided
    bool intersect_with_triangle(const Triangle& tr, double* t, double* u, double* v, Counters* c) const {
        if (c) c->n_a++;
        V3 e1 = tr.v2 - tr.v1;
        V3 e2 = tr.v3 - tr.v1;
        V3 h = cross(direction, e2);
        double a = dot(e1, h);
        if (a > -kEps && a < kEps) return false;
        if (c) c->n_u++;
        double f = 1.0 / a;
        V3 s = origin - tr.v1;
        double uu = f * dot(s, h);
        if (uu < 0.0 || uu > 1.0) return false;
        if (c) c->n_v++;
        V3 q = cross(s, e1);
        double vv = f * dot(direction, q);
        if (vv < 0.0 || uu + vv > 1.0) return false;
        if (c) c->n_t++;
        double tt = f * dot(e2, q);
        if (tt > kEps) {
            *t = tt; *u = uu; *v = vv;
            return true;
        }
        return false;
    }

    // ray.rs:104-168 — kept recursive on purpose: this recursion IS the specification.
    Hit intersect_with_octant_with_max_t(const Octree& oc, size_t octant, double max_t, Counters* c) const {
        const OctantNode& node = oc.nodes[octant];
        if (node.triangle_count == 0) return Hit{};
        if (c) c->n_node++;
        Hit own;
        double closest = max_t;
        for (size_t ti : node.triangles) {  // ray.rs:119-129
            double t, u, v;
            if (intersect_with_triangle(oc.triangles[ti], &t, &u, &v, c)) {
                if (t < closest) {
                    closest = t;
                    own.some = true; own.t = t; own.u = u; own.v = v; own.tri = ti;
                }
            }
        }
        std::pair<double, size_t> kids[8];  // ray.rs:132-144
        size_t nk = 0;
        for (size_t coi : node.children) {
            double key;
            if (c) c->n_slab++;
            if (intersect_aabb(oc.aabbs[oc.nodes[coi].aabb_index], &key)) kids[nk++] = {key, coi};
        }
        // ray.rs:146-147: slice::sort_by is a STABLE sort; a NaN key would panic in Rust — here it
        // simply compares false (keys are never NaN for non-degenerate boxes).
        std::stable_sort(kids, kids + nk, [](const std::pair<double, size_t>& a, const std::pair<double, size_t>& b) {
            return a.first < b.first;
        });
        Hit child;
        double child_t = std::numeric_limits<double>::infinity();
        for (size_t i = 0; i < nk; ++i) {  // ray.rs:152-161: first child with ANY hit wins
            Hit r = intersect_with_octant_with_max_t(oc, kids[i].second, std::numeric_limits<double>::infinity(), c);
            if (r.some) {
                child_t = r.t;
                child = r;
                break;
            }
        }
        if (child_t < closest) return child;  // ray.rs:163-167
        return own;
    }
};

struct OScene {
    Octree octree;
    std::vector<Material> materials;
    std::vector<Texture> textures;
};

}  // namespace

namespace {

// ---------------------------------------------------------------- shading  (src/scene/raytracer.rs)
// Rust `as usize` / `as u8`: saturating, NaN -> 0 (raytracer.rs:52-53,98-107).
inline uint64_t rust_as_usize(double x) {
    if (!(x > 0.0)) return 0;  // NaN, negatives, -0
    if (x >= 18446744073709551616.0) return UINT64_MAX;
    return (uint64_t)x;
}
inline uint8_t rust_clamp_u8(double x) {  // f64::clamp(0,255) as u8 ; clamp(NaN) = NaN -> 0
    if (x != x) return 0;
    if (x < 0.0) x = 0.0;
    if (x > 255.0) x = 255.0;
    return (uint8_t)x;
}

struct Tracer {
    const OScene* sc;
    const rrt_render_params* p;

    // raytracer.rs:114-162
    V3 normal_at(const Hit& h, size_t tx, size_t ty) const {
        const Triangle& tr = sc->octree.triangles[h.tri];
        double w = 1.0 - h.u - h.v;
        V3 n = tr.n2 * h.u + tr.n3 * h.v + tr.n1 * w;
        const Material& m = sc->materials[tr.material];
        if (m.bump >= 0) {
            const Texture& bm = sc->textures[m.bump];
            // Vec indexing panics out of range in Rust; the flat C-ABI rejects such scenes up front
            // (bump smaller than colour map), so the index is always valid here.
            Color c = bm.colours[bm.width * ty + tx];
            V3 bv{(double)c.r, (double)c.g, (double)c.b};
            bv = normalised(bv);
            bv = (bv * 2.0) - V3{1.0, 1.0, 1.0};
            V3 t = cross(n, V3{0.0, 1.0, 0.0});
            if (length(t) == 0.0) t = cross(n, V3{0.0, 0.0, 1.0});
            t = normalised(t);
            V3 b = normalised(cross(n, t));
            n = V3{dot(bv, t), dot(bv, b), dot(bv, n)};
        }
        return normalised(n);
    }

    // raytracer.rs:164-188 — true when NOT occluded (the reference's name is inverted)
    bool light_reaches(V3 origin, V3 normal, V3 target, Counters* c) const {
        V3 direction = target - origin;
        V3 new_origin = origin + normal * p->surface_offset;
        Ray ray{new_origin, direction};
        double max_t = length(direction);
        if (c) c->rays_shadow++;
        Hit h = ray.intersect_with_octant_with_max_t(sc->octree, 0, max_t, c);
        return !h.some;
    }

    // raytracer.rs:260-277
    V3 diffuse(double intensity, double n_dot_l, V3 normal, V3 l, const Material& m) const {
        if (n_dot_l <= 0.0) return V3{0.0, 0.0, 0.0};
        return m.kd * intensity * n_dot_l / (length(normal) * length(l));
    }
    // raytracer.rs:279-304
    V3 specular(double s, double intensity, V3 normal, V3 v, V3 l, const Material& m) const {
        if (s != -1.0) {
            V3 r = (normal * 2.0) * dot(normal, l) - l;
            double r_dot_v = dot(r, v);
            if (r_dot_v > 0.0) return m.ks * intensity * std::pow(r_dot_v / (length(r) * length(v)), s);
        }
        return V3{0.0, 0.0, 0.0};
    }
    // raytracer.rs:192-258
    V3 lighting(V3 point, V3 normal, V3 v, const Material& m, Counters* c) const {
        V3 i{0.0, 0.0, 0.0};
        for (int k = 0; k < p->n_lights; ++k) {
            const rrt_light& L = p->lights[k];
            if (L.kind == RRT_LIGHT_AMBIENT) {
                i = i + m.ka * L.intensity;
            } else if (L.kind == RRT_LIGHT_DIRECTIONAL) {
                V3 d{L.vec[0], L.vec[1], L.vec[2]};
                double n_dot_l = dot(normal, d);
                i = i + diffuse(L.intensity, n_dot_l, normal, d, m);
                i = i + specular(m.ns, L.intensity, normal, v, d, m);
            } else {
                V3 pos{L.vec[0], L.vec[1], L.vec[2]};
                if (!light_reaches(point, normal, pos, c)) break;  // raytracer.rs:235-237: ends the WHOLE loop
                V3 l = pos - point;
                double n_dot_l = dot(normal, l);
                i = i + diffuse(L.intensity, n_dot_l, normal, l, m);
                i = i + specular(m.ns, L.intensity, normal, v, l, m);
            }
        }
        return i;
    }

    // raytracer.rs:33-112
    Color colour(V3 origin, V3 direction, uint32_t depth, int32_t* hit0, Counters* c) const {
        Ray ray{origin, direction};
        Hit h = ray.intersect_with_octant_with_max_t(sc->octree, 0, std::numeric_limits<double>::infinity(), c);
        if (depth == 0 && hit0) *hit0 = h.some ? (int32_t)h.tri : -1;
        if (!h.some) return Color{255, 255, 255};  // WHITE raytracer.rs:10-14,110
        const Triangle& tr = sc->octree.triangles[h.tri];
        const Material& m = sc->materials[tr.material];
        V3 pt = origin + direction * h.t;
        const Texture& tex = sc->textures[m.tex];
        double w = 1.0 - h.u - h.v;
        double tex_x = tr.t2.x * h.u + tr.t3.x * h.v + tr.t1.x * w;
        double tex_y = tr.t2.y * h.u + tr.t3.y * h.v + tr.t1.y * w;
        size_t tx = rust_as_usize(tex_x * (double)tex.width) % tex.width;
        size_t ty = rust_as_usize(tex_y * (double)tex.height) % tex.height;
        Color col = tex.colours[tex.width * ty + tx];
        V3 n = normal_at(h, tx, ty);
        V3 I = lighting(pt, n, -direction, m, c);
        V3 local{(double)col.r * I.x, (double)col.g * I.y, (double)col.b * I.z};
        double refl = m.kr;
        if (refl > 0.0 && depth < (uint32_t)p->max_depth) {  // raytracer.rs:76
            double d_dot_n = dot(direction, n);
            V3 rdir = normalised(direction - n * 2.0 * d_dot_n);
            V3 rorg = pt + n * p->surface_offset;
            if (c) c->rays_reflect++;
            Color rc = colour(rorg, rdir, depth + 1, nullptr, c);
            V3 rv{(double)rc.r, (double)rc.g, (double)rc.b};
            V3 fin = local * (1.0 - refl) + rv * refl;
            return Color{rust_clamp_u8(fin.x), rust_clamp_u8(fin.y), rust_clamp_u8(fin.z)};
        }
        return Color{rust_clamp_u8(local.x), rust_clamp_u8(local.y), rust_clamp_u8(local.z)};
    }
};

// entities.rs:49-69
inline Color mix4(const Color c[4]) {
    uint64_t r = 0, g = 0, b = 0;
    for (int i = 0; i < 4; ++i) { r += c[i].r; g += c[i].g; b += c[i].b; }
    return Color{(uint8_t)(r / 4), (uint8_t)(g / 4), (uint8_t)(b / 4)};
}
// entities.rs:32-36
inline uint32_t pack(Color c) { return (uint32_t)c.b + ((uint32_t)c.g << 8) + ((uint32_t)c.r << 16); }

thread_local std::string g_err;
int fail(const std::string& s) { g_err = s; return RRT_ERR_INVALID; }

}  // namespace

extern "C" {

const char* oracle_last_error(void) { return g_err.c_str(); }

struct oracle_counters {  // SURVEY §8d staged counts
    uint64_t n_a, n_u, n_v, n_t, n_slab, n_node, rays_primary, rays_shadow, rays_reflect;
};

int oracle_scene_create(const rrt_scene_desc* d, void** out) {
    if (!d || !out) return fail("null argument");
    auto sc = std::make_unique<OScene>();
    for (uint32_t i = 0; i < d->n_textures; ++i) {
        Texture t;
        t.width = (size_t)d->textures[i].width;
        t.height = (size_t)d->textures[i].height;
        if (d->textures[i].width <= 0 || d->textures[i].height <= 0) return fail("empty texture");
        t.colours.resize(t.width * t.height);
        std::memcpy(t.colours.data(), d->textures[i].rgb, t.width * t.height * 3);
        sc->textures.push_back(std::move(t));
    }
    for (uint32_t i = 0; i < d->n_materials; ++i) {
        const rrt_material& m = d->materials[i];
        if (m.tex_id < 0 || (uint32_t)m.tex_id >= d->n_textures) return fail("material without texture");
        if (m.bump_id >= (int32_t)d->n_textures) return fail("bad bump id");
        if (m.bump_id >= 0) {
            const Texture& a = sc->textures[m.tex_id];
            const Texture& b = sc->textures[m.bump_id];
            if (b.width * (a.height - 1) + (a.width - 1) >= b.colours.size()) return fail("bump map smaller than colour map");
        }
        sc->materials.push_back(Material{{m.ka[0], m.ka[1], m.ka[2]}, {m.kd[0], m.kd[1], m.kd[2]},
                                         {m.ks[0], m.ks[1], m.ks[2]}, m.ns, m.kr, m.tex_id, m.bump_id});
    }
    sc->octree.init(d->root_min, d->root_max);
    for (uint64_t i = 0; i < d->n_tris; ++i) {
        const double* v = d->v + 9 * i;
        const double* t = d->vt + 6 * i;
        const double* n = d->vn + 9 * i;
        Triangle tr;
        tr.v1 = {v[0], v[1], v[2]}; tr.v2 = {v[3], v[4], v[5]}; tr.v3 = {v[6], v[7], v[8]};
        tr.t1 = {t[0], t[1], 0.0}; tr.t2 = {t[2], t[3], 0.0}; tr.t3 = {t[4], t[5], 0.0};
        tr.n1 = {n[0], n[1], n[2]}; tr.n2 = {n[3], n[4], n[5]}; tr.n3 = {n[6], n[7], n[8]};
        tr.material = d->material_id[i];
        if (tr.material < 0 || (uint32_t)tr.material >= d->n_materials) return fail("bad material id");
        sc->octree.push_triangle(tr);  // utils.rs:192-198: inserted in `f`-line order
    }
    *out = sc.release();
    return RRT_OK;
}

void oracle_scene_destroy(void* s) { delete (OScene*)s; }

// Scene::draw_scene  engine.rs:186-255 + Canvas::put_pixel engine.rs:146-158.
// Renders buffer rows r in [row_begin,row_end) with (r - row_begin) % row_step == 0 and columns with
// c % col_step == 0 (steps of 1 = the whole frame).  OpenMP dynamic over rows ~ rayon's task per row
// (engine.rs:201-203).  The y=-H/2 row is skipped: put_pixel would discard it (engine.rs:152-155).
// Returns seconds of wall time spent in the loop through *seconds.
int oracle_render(void* scene, const rrt_render_params* p, uint32_t* fb, int32_t* hit_ids, int n_threads,
                  int row_begin, int row_end, int row_step, int col_step, oracle_counters* counters,
                  double* seconds) {
    if (!scene || !p || !fb) return fail("null argument");
    const OScene* sc = (const OScene*)scene;
    const int W = p->width, H = p->height;
    if (W <= 0 || H <= 0) return fail("bad size");
    if (row_step < 1) row_step = 1;
    if (col_step < 1) col_step = 1;
    if (row_begin < 1) row_begin = 1;  // buffer row 0 is never written
    if (row_end > H) row_end = H;
    const double xs = p->viewport[0] / (double)W;  // engine.rs:189
    const double ys = p->viewport[1] / (double)H;  // engine.rs:190
    const double zv = p->viewport[2];              // engine.rs:191
    const V3 org{p->origin[0], p->origin[1], p->origin[2]};
    Tracer tr{sc, p};
    Counters total;
#ifdef _OPENMP
    if (n_threads <= 0) n_threads = omp_get_max_threads();
#else
    n_threads = 1;
#endif
    auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel num_threads(n_threads)
    {
        Counters local;
        Counters* cp = counters ? &local : nullptr;
#pragma omp for schedule(dynamic, 1)
        for (int r = row_begin; r < row_end; r += row_step) {
            // put_pixel: new_y = H - (y + H/2)  =>  y = H - r - H/2   (engine.rs:149-150)
            const int yy = H - r - H / 2;
            if (yy < -(H / 2) || yy >= H / 2) continue;  // outside the traced range engine.rs:198
            for (int cidx = 0; cidx < 2 * (W / 2); cidx += col_step) {
                const int x = cidx - W / 2;  // new_x = x + W/2 (engine.rs:147)
                V3 d[4] = {
                    {(double)x * xs, (double)yy * ys, zv},                  // engine.rs:207-211
                    {((double)x + 0.5) * xs, (double)yy * ys, zv},          // engine.rs:217-221
                    {(double)x * xs, ((double)yy + 0.5) * ys, zv},          // engine.rs:224-228
                    {((double)x + 0.5) * xs, ((double)yy + 0.5) * ys, zv},  // engine.rs:231-235
                };
                Color c[4];
                for (int k = 0; k < 4; ++k) {
                    V3 dir = d[k];
                    if (p->use_camera_rot) {  // extension, identical in the CUDA kernel
                        const double* R = p->camera_rot;
                        dir = V3{(R[0] * d[k].x + R[1] * d[k].y) + R[2] * d[k].z,
                                 (R[3] * d[k].x + R[4] * d[k].y) + R[5] * d[k].z,
                                 (R[6] * d[k].x + R[7] * d[k].y) + R[8] * d[k].z};
                    }
                    int32_t hid = -1;
                    if (cp) cp->rays_primary++;
                    c[k] = tr.colour(org, dir, 0, &hid, cp);
                    if (hit_ids) hit_ids[((size_t)r * W + cidx) * 4 + k] = hid;
                }
                fb[(size_t)r * W + cidx] = pack(mix4(c));
            }
        }
#pragma omp critical
        total.add(local);
    }
    auto t1 = std::chrono::steady_clock::now();
    if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
    if (counters) {
        counters->n_a = total.n_a; counters->n_u = total.n_u; counters->n_v = total.n_v; counters->n_t = total.n_t;
        counters->n_slab = total.n_slab; counters->n_node = total.n_node;
        counters->rays_primary = total.rays_primary; counters->rays_shadow = total.rays_shadow;
        counters->rays_reflect = total.rays_reflect;
    }
    return RRT_OK;
}

// Ray::intersect_with_octant_with_max_t on arbitrary rays (ray.rs:104-168). max_t NULL = +inf.
int oracle_trace_rays(void* scene, uint64_t n, const double* o, const double* d, const double* max_t,
                      int32_t* out_tri, double* out_tuv) {
    if (!scene) return fail("null scene");
    const OScene* sc = (const OScene*)scene;
#pragma omp parallel for schedule(dynamic, 256)
    for (int64_t i = 0; i < (int64_t)n; ++i) {
        Ray r{{o[3 * i], o[3 * i + 1], o[3 * i + 2]}, {d[3 * i], d[3 * i + 1], d[3 * i + 2]}};
        double mt = max_t ? max_t[i] : std::numeric_limits<double>::infinity();
        Hit h = r.intersect_with_octant_with_max_t(sc->octree, 0, mt, nullptr);
        out_tri[i] = h.some ? (int32_t)h.tri : -1;
        if (h.some && out_tuv) { out_tuv[3 * i] = h.t; out_tuv[3 * i + 1] = h.u; out_tuv[3 * i + 2] = h.v; }
    }
    return RRT_OK;
}

// Cross-check that ignores the octree: closest accepted M-T hit over ALL triangles in index order,
// strict '<' (SURVEY §8c(i)).  Differs from the octree path only on ties / reference quirks.
int oracle_trace_bruteforce(void* scene, uint64_t n, const double* o, const double* d, const double* max_t,
                            int32_t* out_tri, double* out_tuv) {
    if (!scene) return fail("null scene");
    const OScene* sc = (const OScene*)scene;
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t i = 0; i < (int64_t)n; ++i) {
        Ray r{{o[3 * i], o[3 * i + 1], o[3 * i + 2]}, {d[3 * i], d[3 * i + 1], d[3 * i + 2]}};
        double best = max_t ? max_t[i] : std::numeric_limits<double>::infinity();
        int32_t bi = -1;
        double bu = 0, bv = 0;
        for (size_t k = 0; k < sc->octree.triangles.size(); ++k) {
            double t, u, v;
            if (r.intersect_with_triangle(sc->octree.triangles[k], &t, &u, &v, nullptr) && t < best) {
                best = t; bi = (int32_t)k; bu = u; bv = v;
            }
        }
        out_tri[i] = bi;
        if (bi >= 0 && out_tuv) { out_tuv[3 * i] = best; out_tuv[3 * i + 1] = bu; out_tuv[3 * i + 2] = bv; }
    }
    return RRT_OK;
}

// RayTracer::get_ray_colour on arbitrary rays (raytracer.rs:29-31).
int oracle_ray_colours(void* scene, const rrt_render_params* p, uint64_t n, const double* o, const double* d,
                       uint8_t* out_rgb) {
    if (!scene || !p) return fail("null argument");
    const OScene* sc = (const OScene*)scene;
    Tracer tr{sc, p};
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t i = 0; i < (int64_t)n; ++i) {
        Color c = tr.colour({o[3 * i], o[3 * i + 1], o[3 * i + 2]}, {d[3 * i], d[3 * i + 1], d[3 * i + 2]}, 0, nullptr, nullptr);
        out_rgb[3 * i] = c.r; out_rgb[3 * i + 1] = c.g; out_rgb[3 * i + 2] = c.b;
    }
    return RRT_OK;
}

// ---- octree inspection (same shapes as the product's rrt_octree_* so tests can diff the two trees)
uint64_t oracle_octree_node_count(void* scene) { return ((OScene*)scene)->octree.nodes.size(); }
int oracle_octree_node(void* scene, uint64_t i, double box[6], uint64_t* triangle_count, uint64_t* first_child,
                       uint64_t* list_len) {
    const Octree& oc = ((OScene*)scene)->octree;
    if (i >= oc.nodes.size()) return fail("node index out of range");
    const OctantNode& n = oc.nodes[i];
    const Aabb& b = oc.aabbs[n.aabb_index];
    box[0] = b.mn.x; box[1] = b.mn.y; box[2] = b.mn.z; box[3] = b.mx.x; box[4] = b.mx.y; box[5] = b.mx.z;
    *triangle_count = n.triangle_count;
    *first_child = n.children.empty() ? 0 : n.children[0];
    for (size_t k = 0; k < n.children.size(); ++k)
        if (n.children[k] != n.children[0] + k) return fail("children not contiguous");
    *list_len = n.triangles.size();
    return RRT_OK;
}
int64_t oracle_octree_node_list(void* scene, uint64_t i, uint32_t* out, uint64_t cap) {
    const Octree& oc = ((OScene*)scene)->octree;
    if (i >= oc.nodes.size()) return -1;
    const OctantNode& n = oc.nodes[i];
    for (size_t k = 0; k < n.triangles.size() && k < cap; ++k) out[k] = (uint32_t)n.triangles[k];
    return (int64_t)n.triangles.size();
}
int oracle_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

}  // extern "C"

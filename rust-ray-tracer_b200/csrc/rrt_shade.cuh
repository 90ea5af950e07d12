// rrt_shade.cuh — RayTracer::get_ray_colour (raytracer.rs:29-304) as a per-thread state machine with a
// SINGLE traversal call site: primary / reflection rays (closest hit) and shadow rays (any hit) of
// different lanes meet in the same traversal loop instead of serialising through separate copies.
#pragma once
#include "rrt_trace.cuh"

namespace rrt {

struct DevParams {
    int32_t W, H;
    double ox, oy, oz;      // RayTracer.origin
    double xs, ys, zv;      // viewport.width / W, viewport.height / H, viewport.distance (engine.rs:189-191)
    int32_t n_lights, max_depth;
    rrt_light lights[RRT_MAX_LIGHTS];
    double offset;          // SURFACE_OFFSET (raytracer.rs:17)
    int32_t use_rot, _pad;
    double rot[9];
    int32_t tile_rows, n_shards, shard_index, n_rows_local;  // row sharding (rrt_shard)
    int32_t full_frame;     // 1: fb is a FULL frame (possibly peer memory), owned row k goes to its buffer row
    int32_t search_mode;    // searches that go round the octree walk (rrt_trace.cuh): 0 never, 2 always, 1 switched per
                            // scene handle from one launch to the next by how many rays they answered (fstat),
                            // 3 the switches in fstat as the host set them (rrt_scene_search_mode)
    uint32_t* fstat;        // per scene handle, device: [0] shadow queries answered, [1] handed to the walk, [2] filter on,
                            // [3] launches, [4] camera rays answered, [5] handed over, [6] sparse search on, [7] unused
};

struct RayCount {
    uint32_t primary = 0, shadow = 0, reflect = 0;
    TraceCount tc;
};

// raytracer.rs:260-277 + 279-304 for one light vector l: returns diffuse and adds specular separately,
// in the reference's order `i += diffuse; i += specular` (raytracer.rs:216-226, 242-252).
__device__ __forceinline__ void add_light(V3& I, const DevMaterial* __restrict__ m, double intensity, const V3& n,
                                          const V3& v, const V3& l) {
    const double n_dot_l = dot(n, l);
    if (!(n_dot_l <= 0.0)) {  // raytracer.rs:268 returns zero when n_dot_l <= 0 (NaN falls through, as in Rust)
        const V3 kd{__ldg(&m->kd[0]), __ldg(&m->kd[1]), __ldg(&m->kd[2])};
        I = I + kd * intensity * n_dot_l / (length(n) * length(l));  // raytracer.rs:276
    } else {
        I = I + V3{0.0, 0.0, 0.0};
    }
    const double s = __ldg(&m->ns);
    bool spec = false;
    if (s != -1.0) {  // raytracer.rs:288
        const V3 r = (n * 2.0) * dot(n, l) - l;  // raytracer.rs:289
        const double r_dot_v = dot(r, v);
        if (r_dot_v > 0.0) {
            const V3 ks{__ldg(&m->ks[0]), __ldg(&m->ks[1]), __ldg(&m->ks[2])};
            I = I + ks * intensity * pow_call(div_rn(r_dot_v, length(r) * length(v)), s);  // raytracer.rs:293-295
            spec = true;
        }
    }
    if (!spec) I = I + V3{0.0, 0.0, 0.0};
}

// Per-ray state of RayTracer::get_ray_colour_recursive (raytracer.rs:33-112) between traversals.
// Usage:  st.begin(o, d);  do { traverse(st.so, st.sd, st.smax, st.shadow, h); } while (!st.advance(sc, P, h));
// then st.res is the colour packed r | g<<8 | b<<16 and st.hit0 the depth-0 triangle id (-1 = miss).
//
// Register budget: the traversal is the hot loop and the kernel is register-bound (64 per thread), so the
// only doubles this struct needs in registers across a traversal are the ray handed to traverse().  The
// shading state a shadow query interrupts (hit point, normal, accumulated light, view direction: 12
// doubles) is stored explicitly in `parked` while the shadow ray is traced (shared memory for it was measured
// slower: it takes L1 capacity away from the tree data, profiles/r01_experiments.md).
struct RayState {
    // reflection stack: local colour and reflectivity of every open level (raytracer.rs:89-101)
    double lx[kMaxReflectDepth], ly[kMaxReflectDepth], lz[kMaxReflectDepth], lk[kMaxReflectDepth];
    V3 so, sd;            // the ray to hand to traverse(); for primary / reflection rays it IS the current ray
    double smax;
    double parked[12];    // pt, n, I, view direction
    uint32_t texel, res;
    int32_t hit0, mid;    // depth-0 triangle id; material of the point being shaded
    int depth, li;
    bool shadow;
    RayCount rc;

    __device__ __forceinline__ void begin(const V3& o, const V3& d) {
        so = o; sd = d;
        smax = __longlong_as_double(0x7FF0000000000000ll);
        shadow = false;
        depth = 0; li = 0;
        res = 0x00FFFFFFu;
        hit0 = -1;
        mid = 0;
        texel = 0;
    }

    // unwind the open mirror levels: every level re-quantises to u8 (raytracer.rs:89-101)
    __device__ __forceinline__ void unwind() {
        while (depth > 0) {
            --depth;
            const double k = lk[depth];
            const V3 rv{(double)(res & 255u), (double)((res >> 8) & 255u), (double)((res >> 16) & 255u)};
            const V3 fin = V3{lx[depth], ly[depth], lz[depth]} * (1.0 - k) + rv * k;  // raytracer.rs:95
            res = rust_clamp_u8(fin.x) | (rust_clamp_u8(fin.y) << 8) | (rust_clamp_u8(fin.z) << 16);
        }
    }

    __device__ __forceinline__ void park_store(const V3& pt, const V3& n, const V3& I, const V3& rd) {
        parked[0] = pt.x; parked[1] = pt.y; parked[2] = pt.z;
        parked[3] = n.x;  parked[4] = n.y;  parked[5] = n.z;
        parked[6] = I.x;  parked[7] = I.y;  parked[8] = I.z;
        parked[9] = rd.x; parked[10] = rd.y; parked[11] = rd.z;
    }
    __device__ __forceinline__ void park_load(V3& pt, V3& n, V3& I, V3& rd) const {
        pt = V3{parked[0], parked[1], parked[2]};
        n = V3{parked[3], parked[4], parked[5]};
        I = V3{parked[6], parked[7], parked[8]};
        rd = V3{parked[9], parked[10], parked[11]};
    }

    // Consumes the result of the traversal just made; returns true when the colour is final.
    template <bool COUNT>
    __device__ __forceinline__ bool advance(const DevScene& sc, const DevParams& P, const Hit& h) {
        V3 pt, n, I, rd;
        const DevMaterial* __restrict__ mat;
        if (!shadow) {
            if (depth == 0) hit0 = (h.id == kNone) ? -1 : (int32_t)h.id;
            if (h.id == kNone) {  // WHITE raytracer.rs:109-111
                res = 0x00FFFFFFu;
                unwind();
                return true;
            }
            // ---- hit set-up (raytracer.rs:38-57)
            rd = sd;
            double hu, hv;
            hit_uv(so, sd, sc.tris + h.rec, hu, hv);
            const double2* sp2 = reinterpret_cast<const double2*>(sc.shade + h.id);
            const double2 a0 = __ldg(sp2), a1 = __ldg(sp2 + 1), a2 = __ldg(sp2 + 2), a3 = __ldg(sp2 + 3),
                          a4 = __ldg(sp2 + 4), a5 = __ldg(sp2 + 5), a6 = __ldg(sp2 + 6), a7 = __ldg(sp2 + 7);
            const V3 n1{a0.x, a0.y, a1.x}, n2{a1.y, a2.x, a2.y}, n3{a3.x, a3.y, a4.x};
            const double t1x = a4.y, t1y = a5.x, t2x = a5.y, t2y = a6.x, t3x = a6.y, t3y = a7.x;
            mid = (int32_t)(__double_as_longlong(a7.y) & 0xFFFFFFFFll);
            mat = sc.materials + mid;
            pt = so + sd * h.t;  // raytracer.rs:39
            const double w = 1.0 - hu - hv;  // raytracer.rs:43
            const double tex_x = t2x * hu + t3x * hv + t1x * w;  // raytracer.rs:45-47
            const double tex_y = t2y * hu + t3y * hv + t1y * w;  // raytracer.rs:48-50
            const DevTexture tex = sc.tex[__ldg(&mat->tex)];
            const uint32_t tx = rust_usize_mod(tex_x * (double)tex.width, tex.width);    // raytracer.rs:52
            const uint32_t ty = rust_usize_mod(tex_y * (double)tex.height, tex.height);  // raytracer.rs:53
            const uchar4 c = __ldg(tex.texels + (size_t)tex.width * ty + tx);            // raytracer.rs:55
            texel = (uint32_t)c.x | ((uint32_t)c.y << 8) | ((uint32_t)c.z << 16);
            // ---- normal (raytracer.rs:114-162)
            n = n2 * hu + n3 * hv + n1 * w;  // raytracer.rs:122-124
            const int32_t bump = __ldg(&mat->bump);
            if (bump >= 0) {
                const DevTexture bt = sc.tex[bump];
                const uchar4 b = __ldg(bt.texels + (size_t)bt.width * ty + tx);  // raytracer.rs:128 (same tx,ty)
                V3 bv{(double)b.x, (double)b.y, (double)b.z};
                bv = normalised(bv);                       // raytracer.rs:129
                bv = (bv * 2.0) - V3{1.0, 1.0, 1.0};       // raytracer.rs:130-135
                V3 t = cross(n, V3{0.0, 1.0, 0.0});        // raytracer.rs:137-141
                if (length(t) == 0.0) t = cross(n, V3{0.0, 0.0, 1.0});  // raytracer.rs:143-149
                t = normalised(t);
                const V3 bb = normalised(cross(n, t));     // raytracer.rs:152
                n = V3{dot(bv, t), dot(bv, bb), dot(bv, n)};  // raytracer.rs:154-158
            }
            n = normalised(n);  // raytracer.rs:161
            I = V3{0.0, 0.0, 0.0};
            li = 0;
        } else {
            // result of the shadow query for point light `li` (raytracer.rs:232-252)
            shadow = false;
            park_load(pt, n, I, rd);
            mat = sc.materials + mid;
            if (h.id != kNone) {
                li = P.n_lights;  // `break`: ends the WHOLE light loop (raytracer.rs:235-237)
            } else {
                const rrt_light& L = P.lights[li];
                const V3 l = V3{L.vec[0], L.vec[1], L.vec[2]} - pt;  // raytracer.rs:239
                add_light(I, mat, L.intensity, n, -rd, l);
                li++;
            }
        }

        // ---- light loop (raytracer.rs:205-255) until a point light needs its shadow ray
        while (li < P.n_lights) {
            const rrt_light& L = P.lights[li];
            if (L.kind == RRT_LIGHT_AMBIENT) {
                const V3 ka{__ldg(&mat->ka[0]), __ldg(&mat->ka[1]), __ldg(&mat->ka[2])};
                I = I + ka * L.intensity;  // raytracer.rs:208
                li++;
            } else if (L.kind == RRT_LIGHT_DIRECTIONAL) {
                add_light(I, mat, L.intensity, n, -rd, V3{L.vec[0], L.vec[1], L.vec[2]});  // raytracer.rs:214-226
                li++;
            } else {
                // triangle_exists_between_points (raytracer.rs:164-188)
                park_store(pt, n, I, rd);
                sd = V3{L.vec[0], L.vec[1], L.vec[2]} - pt;  // raytracer.rs:170
                so = pt + n * P.offset;                     // raytracer.rs:172
                smax = length(sd);                          // raytracer.rs:179
                shadow = true;
                if (COUNT) rc.shadow++;
                return false;
            }
        }

        // ---- local colour and optional mirror bounce (raytracer.rs:66-108)
        const V3 local{(double)(texel & 255u) * I.x, (double)((texel >> 8) & 255u) * I.y,
                       (double)((texel >> 16) & 255u) * I.z};
        const double kr = __ldg(&mat->kr);
        if (kr > 0.0 && depth < P.max_depth) {  // raytracer.rs:76
            const double d_dot_n = dot(rd, n);  // raytracer.rs:78
            const V3 ndir = normalised(rd - n * 2.0 * d_dot_n);  // raytracer.rs:79
            so = pt + n * P.offset;  // raytracer.rs:82
            sd = ndir;
            lx[depth] = local.x; ly[depth] = local.y; lz[depth] = local.z; lk[depth] = kr;
            depth++;
            smax = __longlong_as_double(0x7FF0000000000000ll);
            if (COUNT) rc.reflect++;
            return false;
        }
        res = rust_clamp_u8(local.x) | (rust_clamp_u8(local.y) << 8) | (rust_clamp_u8(local.z) << 16);  // raytracer.rs:104-108
        unwind();
        return true;
    }
};

}  // namespace rrt

// rrt_render.cu — kernels for Scene::draw_scene (engine.rs:186-255) on sm_100a and their launchers.
// Compiled with -fmad=false -prec-div=true -prec-sqrt=true (no fast-math anywhere).
#include "rrt_render.h"
#include "rrt_shade.cuh"

namespace rrt {

namespace {

constexpr int kBlock = 256;      // 8 warps
constexpr int kTileW = 16;       // pixels per block tile: 4 warps across x 4 px
constexpr int kTileH = 4;        //                        2 warps down   x 2 px

// compact row k of this shard -> buffer row (rrt_shard in rrt.h)
__device__ __forceinline__ int buffer_row(const DevParams& P, int k) {
    if (P.n_shards <= 1) return k;
    const int jl = k / P.tile_rows, within = k - jl * P.tile_rows;
    return (P.shard_index + jl * P.n_shards) * P.tile_rows + within;
}

// One thread per supersample: a warp covers 4x2 pixels x 4 samples (an 8x4 lattice of half-pixel
// steps), lanes 4p..4p+3 are the samples d1..d4 of pixel p (engine.rs:207-236).
template <bool COUNT>
__global__ void __launch_bounds__(kBlock)
render_kernel(const DevScene sc, const __grid_constant__ DevParams P, uint32_t* __restrict__ fb,
              int32_t* __restrict__ hit_ids, DevCounters* __restrict__ counters, const int tiles_x) {
    const int tile = blockIdx.x;
    const int by = tile / tiles_x, bx = tile - by * tiles_x;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int p = lane >> 2, k = lane & 3;
    const int col = bx * kTileW + (warp & 3) * 4 + (p & 3);
    const int krow = by * kTileH + (warp >> 2) * 2 + (p >> 2);
    const bool in_buf = col < P.W && krow < P.n_rows_local;
    const int r = buffer_row(P, krow);
    // put_pixel: new_x = x + W/2, new_y = H - (y + H/2)  (engine.rs:147-150)
    const int x = col - P.W / 2;
    const int y = P.H - r - P.H / 2;
    const bool traced = in_buf && r >= 1 && r < P.H && col < 2 * (P.W / 2) && y >= -(P.H / 2) && y < P.H / 2;

    uint32_t rgb = 0;
    int32_t hid = -1;
    RayCount rc;
    if (traced) {
        const double fx = (k & 1) ? ((double)x + 0.5) : (double)x;   // engine.rs:208,218
        const double fy = (k & 2) ? ((double)y + 0.5) : (double)y;   // engine.rs:209,226
        V3 d{fx * P.xs, fy * P.ys, P.zv};
        if (P.use_rot) {  // extension (rrt.h): d' = R d
            d = V3{(P.rot[0] * d.x + P.rot[1] * d.y) + P.rot[2] * d.z,
                   (P.rot[3] * d.x + P.rot[4] * d.y) + P.rot[5] * d.z,
                   (P.rot[6] * d.x + P.rot[7] * d.y) + P.rot[8] * d.z};
        }
        rgb = ray_colour<COUNT>(sc, P, V3{P.ox, P.oy, P.oz}, d, &hid, rc);
    }
    // Color::mix (entities.rs:49-69): per-channel integer mean of the 4 samples, truncating
    uint32_t cr = rgb & 255u, cg = (rgb >> 8) & 255u, cb = (rgb >> 16) & 255u;
    cr += __shfl_xor_sync(0xFFFFFFFFu, cr, 1); cg += __shfl_xor_sync(0xFFFFFFFFu, cg, 1); cb += __shfl_xor_sync(0xFFFFFFFFu, cb, 1);
    cr += __shfl_xor_sync(0xFFFFFFFFu, cr, 2); cg += __shfl_xor_sync(0xFFFFFFFFu, cg, 2); cb += __shfl_xor_sync(0xFFFFFFFFu, cb, 2);
    if (in_buf) {
        const size_t pix = (size_t)krow * P.W + col;
        if (k == 0) fb[pix] = traced ? ((cb >> 2) + ((cg >> 2) << 8) + ((cr >> 2) << 16)) : 0u;  // entities.rs:32-36
        if (hit_ids) hit_ids[pix * 4 + k] = hid;
    }
    if (COUNT) {
        unsigned long long v[7] = {rc.primary, rc.shadow, rc.reflect, rc.tc.mt, rc.tc.box, rc.tc.cull, rc.tc.nodes};
#pragma unroll
        for (int i = 0; i < 7; ++i) {
            for (int o = 16; o > 0; o >>= 1) v[i] += __shfl_xor_sync(0xFFFFFFFFu, v[i], o);
            if (lane == 0 && v[i]) atomicAdd(reinterpret_cast<unsigned long long*>(counters) + i, v[i]);
        }
    }
}

// Ray::intersect_with_octant_with_max_t for arbitrary rays (ray.rs:104-168)
__global__ void __launch_bounds__(kBlock)
trace_kernel(const DevScene sc, const uint64_t n, const double* __restrict__ o, const double* __restrict__ d,
             const double* __restrict__ max_t, int32_t* __restrict__ out_tri, double* __restrict__ out_tuv) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Hit h;
    TraceCount tc;
    const double mt = max_t ? max_t[i] : __longlong_as_double(0x7FF0000000000000ll);
    traverse<false>(sc, V3{o[3 * i], o[3 * i + 1], o[3 * i + 2]}, V3{d[3 * i], d[3 * i + 1], d[3 * i + 2]}, mt, false, h, tc);
    out_tri[i] = (h.id == kNone) ? -1 : (int32_t)h.id;
    if (h.id != kNone && out_tuv) {
        out_tuv[3 * i] = h.t; out_tuv[3 * i + 1] = h.u; out_tuv[3 * i + 2] = h.v;
    }
}

// RayTracer::get_ray_colour for arbitrary rays (raytracer.rs:29-31)
__global__ void __launch_bounds__(kBlock)
colour_kernel(const DevScene sc, const __grid_constant__ DevParams P, const uint64_t n, const double* __restrict__ o,
              const double* __restrict__ d, uint8_t* __restrict__ out_rgb) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int32_t hid;
    RayCount rc;
    const uint32_t c = ray_colour<false>(sc, P, V3{o[3 * i], o[3 * i + 1], o[3 * i + 2]},
                                         V3{d[3 * i], d[3 * i + 1], d[3 * i + 2]}, &hid, rc);
    out_rgb[3 * i] = (uint8_t)(c & 255u); out_rgb[3 * i + 1] = (uint8_t)((c >> 8) & 255u); out_rgb[3 * i + 2] = (uint8_t)((c >> 16) & 255u);
}

// scatter gathered compact shards into the full frame (rrt_deinterleave_device)
__global__ void deinterleave_kernel(const uint32_t* __restrict__ g, int rows_per_block, int W, int H, int T, int G,
                                    uint32_t* __restrict__ fb) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)W * H) return;
    const int r = (int)(i / W), c = (int)(i - (size_t)r * W);
    const int j = r / T, s = j % G, jl = j / G;
    const int k = jl * T + (r - j * T);
    fb[i] = (r == 0) ? 0u : g[((size_t)s * rows_per_block + k) * W + c];
}

}  // namespace

cudaError_t launch_render(const DevScene& sc, const DevParams& P, uint32_t* d_fb, int32_t* d_hit,
                          DevCounters* d_counters, cudaStream_t stream) {
    const int tiles_x = (P.W + kTileW - 1) / kTileW, tiles_y = (P.n_rows_local + kTileH - 1) / kTileH;
    const long long tiles = (long long)tiles_x * tiles_y;
    if (tiles <= 0) return cudaSuccess;
    if (d_counters)
        render_kernel<true><<<(unsigned)tiles, kBlock, 0, stream>>>(sc, P, d_fb, d_hit, d_counters, tiles_x);
    else
        render_kernel<false><<<(unsigned)tiles, kBlock, 0, stream>>>(sc, P, d_fb, d_hit, nullptr, tiles_x);
    return cudaGetLastError();
}

cudaError_t launch_trace(const DevScene& sc, uint64_t n, const double* o, const double* d, const double* max_t,
                         int32_t* out_tri, double* out_tuv, cudaStream_t stream) {
    if (n == 0) return cudaSuccess;
    trace_kernel<<<(unsigned)((n + kBlock - 1) / kBlock), kBlock, 0, stream>>>(sc, n, o, d, max_t, out_tri, out_tuv);
    return cudaGetLastError();
}

cudaError_t launch_colours(const DevScene& sc, const DevParams& P, uint64_t n, const double* o, const double* d,
                           uint8_t* out_rgb, cudaStream_t stream) {
    if (n == 0) return cudaSuccess;
    colour_kernel<<<(unsigned)((n + kBlock - 1) / kBlock), kBlock, 0, stream>>>(sc, P, n, o, d, out_rgb);
    return cudaGetLastError();
}

cudaError_t launch_deinterleave(const uint32_t* g, int rows_per_block, int W, int H, int T, int G, uint32_t* fb,
                                cudaStream_t stream) {
    const size_t n = (size_t)W * H;
    if (n == 0) return cudaSuccess;
    deinterleave_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(g, rows_per_block, W, H, T, G, fb);
    return cudaGetLastError();
}

}  // namespace rrt

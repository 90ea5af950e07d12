// rrt_raygen.cuh — the per-sample bookkeeping of Scene::draw_scene (engine.rs:186-255) around a traced ray:
// which pixel and supersample a position of the sample queue is, whether the reference traces it and with which
// direction, and how the four samples of a pixel are merged.  Kept apart from the kernel so that the same source is
// compiled for the host by the emulation tests (tests/host_emulation), which have no warps and no queue.
#pragma once
#include "rrt_shade.cuh"

namespace rrt {

constexpr int kTileW = 16;       // samples are enumerated in 16x4-pixel blocks of eight 4x2-pixel tiles,
constexpr int kTileH = 4;        // 4 supersamples per pixel: 32 consecutive indices = one 8x4 half-pixel lattice

// compact row k of this shard -> buffer row (rrt_shard in rrt.h)
__device__ __forceinline__ int buffer_row(const DevParams& P, int k) {
    if (P.n_shards <= 1) return k;
    const int jl = k / P.tile_rows, within = k - jl * P.tile_rows;
    return (P.shard_index + jl * P.n_shards) * P.tile_rows + within;
}

// block `blk` (16x4 pixels, row-major over the shard's compact rows), warp tile w (0..7) of the block, position l
// (0..31) in the tile  ->  column, compact row and supersample index
__device__ __forceinline__ void sample_pixel(unsigned long long blk, unsigned w, unsigned l, int tiles_x, int& col, int& krow,
                                             unsigned& k) {
    const unsigned p = l >> 2;
    const int by = (int)(blk / (unsigned)tiles_x), bx = (int)(blk - (unsigned long long)by * tiles_x);
    col = bx * kTileW + (int)(w & 3u) * 4 + (int)(p & 3u);
    krow = by * kTileH + (int)(w >> 2) * 2 + (int)(p >> 2);
    k = l & 3u;
}

// Does the reference trace supersample k of the pixel at (col, buffer row r), and along which direction?
// put_pixel: new_x = x + W/2, new_y = H - (y + H/2)  (engine.rs:147-150); row 0, an odd last column and an odd last
// row are never written by the reference.
__device__ __forceinline__ bool primary_ray(const DevParams& P, int col, int r, unsigned k, V3& d) {
    const int x = col - P.W / 2;
    const int y = P.H - r - P.H / 2;
    const bool traced = r >= 1 && r < P.H && col < 2 * (P.W / 2) && y >= -(P.H / 2) && y < P.H / 2;
    if (traced) {
        const double fx = (k & 1u) ? ((double)x + 0.5) : (double)x;   // engine.rs:208,218
        const double fy = (k & 2u) ? ((double)y + 0.5) : (double)y;   // engine.rs:209,226
        d = V3{fx * P.xs, fy * P.ys, P.zv};
        if (P.use_rot) {  // extension (rrt.h): d' = R d
            d = V3{(P.rot[0] * d.x + P.rot[1] * d.y) + P.rot[2] * d.z,
                   (P.rot[3] * d.x + P.rot[4] * d.y) + P.rot[5] * d.z,
                   (P.rot[6] * d.x + P.rot[7] * d.y) + P.rot[8] * d.z};
        }
    }
    return traced;
}

// Packed per-pixel accumulator: 10 bits per channel + 2-bit arrival count.  What one finished sample adds ...
__device__ __forceinline__ uint32_t sample_add(uint32_t c) {
    return (c & 255u) | (((c >> 8) & 255u) << 10) | (((c >> 16) & 255u) << 20) | (1u << 30);
}
// ... and the pixel the 4th arrival stores: Color::mix + pack (entities.rs:32-36,49-69)
__device__ __forceinline__ uint32_t mix_pack(uint32_t sum) {
    const uint32_t r = (sum & 1023u) >> 2, g = ((sum >> 10) & 1023u) >> 2, b = ((sum >> 20) & 1023u) >> 2;
    return b + (g << 8) + (r << 16);
}

}  // namespace rrt

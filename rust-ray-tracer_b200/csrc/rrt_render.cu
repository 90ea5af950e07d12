// rrt_render.cu — kernels for Scene::draw_scene (engine.rs:186-255) on sm_100a and their launchers.
// Compiled with -fmad=false -prec-div=true -prec-sqrt=true (no fast-math anywhere).
#include "rrt_render.h"
#include "rrt_raygen.cuh"

#include <cub/device/device_radix_sort.cuh>

#include <mutex>

// Scratch of the allocation-free render call (rrt_render_device_noalloc, no
// counterpart in the reference): the per-pixel accumulators + the sample-queue head that launch_render otherwise gets
// from a cudaMalloc'ed buffer of the scene handle.  File scope, one copy per GPU that loads this library, held until
// the process exits: kNoAllocMaxPixels * 4 + 16 bytes = 8 MiB + 16 B.  In HOST code this name is only a stand-in for
// the device array (not a device address): the launcher below resolves it with cudaGetSymbolAddress once per device.
// The limit itself is rrt::kNoAllocMaxPixels in rrt_render.h (RRT_NOALLOC_MAX_PIXELS in include/rrt.h).
static __device__ __align__(256) uint32_t g_noalloc_scratch[((size_t)1 << 21) + 4];

namespace rrt {

namespace {

#ifndef RRT_BLOCK
#define RRT_BLOCK 128
#endif
#ifndef RRT_MIN_BLOCKS
#define RRT_MIN_BLOCKS 8
#endif
#ifndef RRT_REFILL_MIN
#define RRT_REFILL_MIN 32    // idle lanes a warp waits for before it fetches new samples; 32 = a warp keeps one 4x2-pixel
                             // tile (x4 supersamples) together through its shadow / mirror rays (sweep: profiles/r01_experiments.md)
#endif
constexpr int kBlock = RRT_BLOCK;
#ifndef RRT_TIMELINE
#define RRT_TIMELINE 0      // development aid (tools/drain_timeline.py): when does the queue run dry, when do warps exit
#endif
#if RRT_TIMELINE
__device__ unsigned long long g_tl[4 + 256];   // [0] min start, [1] min queue-dry, [2] max exit, [3] warps; [4..] exits per 5 us after queue-dry
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#endif

// Frame-to-frame switch of the two searches that go round the octree walk (rrt_trace.cuh).  Each pays where it answers
// most of the rays it is tried on (c4: 98 % of the shadow queries, all camera rays; c2: 77 % / 99.9 %) and costs where
// it does not (c5: every shadow ray is occluded and the camera sits inside the soup — a handed-over ray pays the search
// AND the walk).  So every launch counts what the searches returned (DevParams::fstat, per scene handle) and a
// one-thread kernel behind it sets the switches for the next launch: a search stays on while it answers at least half
// of its rays (measured break-even, profiles/r02_experiments.md), and every kProbeEvery-th frame runs with both on
// again, which is how a switched-off search gets back in when the camera has moved.  Results do not depend on the
// switches: both paths are exact.
constexpr uint32_t kProbeEvery = 32u;
__global__ void search_switch_kernel(uint32_t* __restrict__ fs) {
    const uint32_t a = fs[0], h = fs[1];
    if (a + h >= 64u) fs[2] = (2ull * a >= (unsigned long long)a + h) ? 1u : 0u;
    const uint32_t sa = fs[4], sh = fs[5];
    if (sa + sh >= 64u) fs[6] = (2ull * sa >= (unsigned long long)sa + sh) ? 1u : 0u;
    fs[8] = a; fs[9] = h; fs[10] = sa; fs[11] = sh;   // what the decision was made from (rrt_scene_search_state)
    fs[0] = fs[1] = fs[4] = fs[5] = 0u;
    const uint32_t frames = ++fs[3];
    if (frames % kProbeEvery == 0u) { fs[2] = 1u; fs[6] = 1u; }
}

// Persistent render kernel (Scene::draw_scene, engine.rs:186-255).  One lane = one supersample at a time.
// Whenever lanes of a warp run dry — their ray missed, or finished its shadow / mirror chain early — they
// fetch the next sample indices from a global queue (one atomicAdd per warp per refill) instead of idling
// until the slowest lane of the warp is done.  All lanes meet in the single traverse() call site whatever
// kind of ray (primary, shadow, reflection) they carry.
// The 4 samples of a pixel (engine.rs:207-236) may be finished by different lanes at different times:
// each adds its colour into a packed per-pixel accumulator (10 bits per channel + 2-bit arrival count);
// the 4th arrival divides by 4 (Color::mix, entities.rs:49-69) and stores the pixel (entities.rs:32-36).
template <bool COUNT, bool HINT>
__global__ void __launch_bounds__(kBlock, RRT_MIN_BLOCKS)
render_kernel(const DevScene sc, const __grid_constant__ DevParams P, uint32_t* __restrict__ fb,
              int32_t* __restrict__ hit_ids, uint32_t* __restrict__ acc, unsigned long long* __restrict__ next,
              DevCounters* __restrict__ counters, const int tiles_x, const unsigned long long total,
              const uint32_t* __restrict__ order, uint32_t* __restrict__ cost) {
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;
    RayState st;
    bool active = false, drained = false;
    size_t pix = 0, out = 0;   // compact pixel index (accumulator, hit ids) / index the finished pixel is stored at
    unsigned k = 0;
    TraceCount tc_total;
    uint32_t n_primary = 0, n_shadow = 0, n_reflect = 0;
    // dispatch-order hint (RenderSchedule): the 16x4-pixel block this warp is working on and when it fetched it
    // (in shared memory: touched once per tile, and the hot loop has no register to spare)
    __shared__ uint2 tile_info_s[HINT ? kBlock / 32 : 1];
    uint2& tile_info = tile_info_s[HINT ? threadIdx.x >> 5 : 0];
    if (HINT && lane == 0) tile_info = make_uint2(0xFFFFFFFFu, 0u);
    // what this warp's searches returned: shadow queries answered / handed to the walk, camera rays likewise
    __shared__ unsigned int fstat_s[kBlock / 32][4];
    unsigned int* const fstat_w = fstat_s[threadIdx.x >> 5];
    if (lane < 4) fstat_w[lane] = 0u;
    const bool switched = P.search_mode == 1 || P.search_mode == 3;
    const bool filter_on = P.search_mode == 2 || (switched && P.fstat[2] != 0u);
    const bool sparse_on = (P.search_mode == 2 || (switched && P.fstat[6] != 0u)) && sc.sparse_ok != 0u;
    __syncwarp();
#if RRT_TIMELINE
    if (lane == 0) atomicMin(&g_tl[0], gtime());
#endif

    for (;;) {
        // ---- refill idle lanes from the global sample queue
        const unsigned idle = __ballot_sync(0xFFFFFFFFu, !active);
        if (HINT && idle == 0xFFFFFFFFu && lane == 0 && tile_info.x != 0xFFFFFFFFu) {
            // the warp is through with its tile: a block costs what its slowest tile took (SM cycles, fetch to last ray)
            atomicMax(cost + tile_info.x, (uint32_t)clock() - tile_info.y);
            tile_info.x = 0xFFFFFFFFu;
        }
        if (__popc(idle) >= RRT_REFILL_MIN && !drained) {
            const int leader = __ffs(idle) - 1;
            unsigned long long base = 0;
            if ((int)lane == leader) base = atomicAdd(next, (unsigned long long)__popc(idle));
            base = __shfl_sync(0xFFFFFFFFu, base, leader);
            if (base + __popc(idle) >= total) drained = true;  // warp-uniform: the queue is (about to be) empty
#if RRT_TIMELINE
            if (drained && lane == 0) atomicMin(&g_tl[1], gtime());
#endif
            if (!active) {
                const unsigned long long s = base + __popc(idle & lt_mask);
                if (s < total) {
                    const unsigned long long tile = s >> 5;
                    const unsigned l = (unsigned)s & 31u;
                    unsigned long long blk = tile >> 3;
                    if (HINT && order) blk = __ldg(order + blk);   // the last part of the frame: heaviest blocks first
                    if (HINT && lane == 0) tile_info = make_uint2((uint32_t)blk, (uint32_t)clock());   // one tile per fetch
                    int col, krow;
                    sample_pixel(blk, (unsigned)tile & 7u, l, tiles_x, col, krow, k);
                    if (col < P.W && krow < P.n_rows_local) {
                        pix = (size_t)krow * P.W + col;
                        const int r = buffer_row(P, krow);
                        out = P.full_frame ? (size_t)r * P.W + col : pix;
                        V3 d;
                        if (primary_ray(P, col, r, k, d)) {
                            st.begin(V3{P.ox, P.oy, P.oz}, d);
                            active = true;
                            if (COUNT) n_primary++;
                        } else {
                            // never written by the reference (row 0, odd last column / row): stays 0, hit id -1
                            if (hit_ids) hit_ids[pix * 4 + k] = -1;
                            if ((atomicAdd(acc + pix, 1u << 30) >> 30) == 3u) fb[out] = 0u;
                        }
                    }
                }
            }
        }
        if (!__any_sync(0xFFFFFFFFu, active)) {
            if (drained) break;
            continue;
        }
        // ---- one traversal for every lane that carries a ray, whatever its kind
        if (active) {
            Hit h;
            const bool was_shadow = st.shadow;
            const uint32_t allow = was_shadow ? (filter_on ? kTryFilter : 0u) : ((sparse_on && st.depth == 0) ? kTrySparse : 0u);
            uint32_t outcome;
            traverse<COUNT>(sc, st.so, st.sd, st.smax, was_shadow, h, tc_total, allow, &outcome);
            if (P.search_mode == 1) {   // per warp: how many of its searches answered / handed over (lanes here are a subset)
                const unsigned here = __activemask();
                const unsigned tried = __ballot_sync(here, outcome != 0u), ans = __ballot_sync(here, outcome == 1u),
                               shad = __ballot_sync(here, was_shadow);
                if (tried != 0u && (int)lane == __ffs(here) - 1) {
                    fstat_w[0] += __popc(ans & shad);
                    fstat_w[1] += __popc(tried & ~ans & shad);
                    fstat_w[2] += __popc(ans & ~shad);
                    fstat_w[3] += __popc(tried & ~ans & ~shad);
                }
                __syncwarp(here);
            }
            if (st.advance<COUNT>(sc, P, h)) {
                active = false;
                if (hit_ids) hit_ids[pix * 4 + k] = st.hit0;
                const uint32_t add = sample_add(st.res);
                const uint32_t old = atomicAdd(acc + pix, add);
                if ((old >> 30) == 3u) fb[out] = mix_pack(old + add);  // 4th sample of this pixel: Color::mix + pack
            }
        }
    }
    __syncwarp();
    if (P.search_mode == 1 && lane < 4 && fstat_w[lane] != 0u) atomicAdd(P.fstat + (lane < 2 ? lane : lane + 2), fstat_w[lane]);
#if RRT_TIMELINE
    if (lane == 0) {
        const unsigned long long t = gtime();
        atomicMax(&g_tl[2], t);
        atomicAdd(&g_tl[3], 1ull);
        const unsigned long long q = g_tl[1];   // set before any warp can exit
        unsigned long long b = t > q ? (t - q) / 5000ull : 0ull;
        atomicAdd(&g_tl[4 + (b > 255ull ? 255ull : b)], 1ull);
    }
#endif
    if (COUNT) {
        n_shadow = st.rc.shadow; n_reflect = st.rc.reflect;
        unsigned long long v[7] = {n_primary, n_shadow, n_reflect, tc_total.mt, tc_total.box, tc_total.cull, tc_total.nodes};
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
             const double* __restrict__ max_t, int32_t* __restrict__ out_tri, double* __restrict__ out_tuv, const uint32_t allow) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Hit h;
    TraceCount tc;
    const double mt = max_t ? max_t[i] : __longlong_as_double(0x7FF0000000000000ll);
    traverse<false>(sc, V3{o[3 * i], o[3 * i + 1], o[3 * i + 2]}, V3{d[3 * i], d[3 * i + 1], d[3 * i + 2]}, mt, false, h, tc, allow);
    out_tri[i] = (h.id == kNone) ? -1 : (int32_t)h.id;
    if (h.id != kNone && out_tuv) {
        double u, v;
        hit_uv(V3{o[3 * i], o[3 * i + 1], o[3 * i + 2]}, V3{d[3 * i], d[3 * i + 1], d[3 * i + 2]}, sc.tris + h.rec, u, v);
        out_tuv[3 * i] = h.t; out_tuv[3 * i + 1] = u; out_tuv[3 * i + 2] = v;
    }
}

// RayTracer::get_ray_colour for arbitrary rays (raytracer.rs:29-31)
__global__ void __launch_bounds__(kBlock)
colour_kernel(const DevScene sc, const __grid_constant__ DevParams P, const uint64_t n,
              const double* __restrict__ o, const double* __restrict__ d, uint8_t* __restrict__ out_rgb) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    RayState st;
    // every search on every ray it applies to (also mirror rays: their origins lie INSIDE the geometry, which the
    // render kernel leaves to the walk — here they test the hand-over rules) unless the caller asked for the walk only
    const uint32_t allow = P.search_mode != 0 ? (kTryFilter | kTrySparse) : 0u;
    TraceCount tc;
    st.begin(V3{o[3 * i], o[3 * i + 1], o[3 * i + 2]}, V3{d[3 * i], d[3 * i + 1], d[3 * i + 2]});
    Hit h;
    do {
        traverse<false>(sc, st.so, st.sd, st.smax, st.shadow, h, tc, allow);
    } while (!st.advance<false>(sc, P, h));
    const uint32_t c = st.res;
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

// RenderSchedule: {sum, max} of the block costs of the frame just rendered
__global__ void order_stats_kernel(const uint32_t* __restrict__ cost, uint32_t n, unsigned long long* __restrict__ stats) {
    unsigned long long sum = 0ull;
    uint32_t mx = 0u;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t c = cost[i];
        sum += c;
        mx = max(mx, c);
    }
    for (int o = 16; o > 0; o >>= 1) {
        sum += __shfl_xor_sync(0xFFFFFFFFu, sum, o);
        mx = max(mx, __shfl_xor_sync(0xFFFFFFFFu, mx, o));
    }
    if ((threadIdx.x & 31u) == 0u) {
        atomicAdd(stats, sum);
        atomicMax(stats + 1, (unsigned long long)mx);
    }
}

// RenderSchedule: next frame's sort keys.  Blocks are handed out in raster order (good locality: the resident warps
// work on a thin band of the frame) EXCEPT the last part of the sequence, which is sorted by descending cost class
// (half octaves of the block's slowest tile) so that the kernel ends on its cheapest tiles.  How long that part must be
// follows from this frame's own numbers: a tile fetched with less than (slowest tile) of the frame left can end up as
// the tail; the frame lasts about (8 tiles per block x sum of block costs) / resident warps (an over-estimate: a block
// costs its SLOWEST tile), so the last  safety x max / that  of the sequence is sorted (safety <= 0: all of it).
__global__ void order_keys_kernel(uint32_t* __restrict__ cost, uint32_t* __restrict__ keys, uint32_t n,
                                  const unsigned long long* __restrict__ stats, unsigned long long* __restrict__ stats_next,
                                  float resident_warps, float safety) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) { stats_next[0] = 0ull; stats_next[1] = 0ull; }
    if (i >= n) return;
    uint32_t n_keep = 0u;
    if (safety > 0.0f && stats[0] > 0ull) {
        const float frac = safety * (float)stats[1] * resident_warps / (8.0f * (float)stats[0]);
        n_keep = frac >= 1.0f ? 0u : n - (uint32_t)ceilf(frac * (float)n);
    }
    const uint32_t c = cost[i];
    cost[i] = 0u;
    uint32_t key = 0u;
    if (c > 1u) {
        const uint32_t e = 31u - (uint32_t)__clz(c);
        key = 2u * e + ((c >> (e - 1u)) & 1u);   // <= 62
    }
    keys[i] = i < n_keep ? 63u : key;
}

__global__ void iota_kernel(uint32_t* __restrict__ v, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) v[i] = i;
}

}  // namespace

void RenderSchedule::release() {
    if (side) { cudaStreamSynchronize(side); cudaStreamDestroy(side); side = nullptr; }
    if (ev_rendered) { cudaEventDestroy(ev_rendered); ev_rendered = nullptr; }
    if (ev_sorted) { cudaEventDestroy(ev_sorted); ev_sorted = nullptr; }
    void* ps[] = {order, cost, keys, keys_alt, iota, temp, stats};
    for (void* p : ps) if (p) cudaFree(p);
    order = cost = keys = keys_alt = iota = nullptr;
    temp = nullptr; stats = nullptr;
    cap_blocks = 0; temp_bytes = 0; valid = false; pending = false;
}

// The schedule's arrays, streams and events for frames of up to n_blocks blocks (no-op when they are there already).
static cudaError_t schedule_allocate(RenderSchedule* sch, uint32_t n_blocks, cudaStream_t stream) {
    if (n_blocks <= sch->cap_blocks && sch->order) return cudaSuccess;
    cudaError_t e;
    sch->release();
    if ((e = cudaStreamCreateWithFlags(&sch->side, cudaStreamNonBlocking)) != cudaSuccess) return e;
    if ((e = cudaEventCreateWithFlags(&sch->ev_rendered, cudaEventDisableTiming)) != cudaSuccess) return e;
    if ((e = cudaEventCreateWithFlags(&sch->ev_sorted, cudaEventDisableTiming)) != cudaSuccess) return e;
    uint32_t** bufs[] = {&sch->order, &sch->cost, &sch->keys, &sch->keys_alt, &sch->iota};
    for (uint32_t** b : bufs)
        if ((e = cudaMalloc((void**)b, (size_t)n_blocks * sizeof(uint32_t))) != cudaSuccess) return e;
    size_t bytes = 0;
    if ((e = cub::DeviceRadixSort::SortPairsDescending(nullptr, bytes, sch->keys, sch->keys_alt, sch->iota, sch->order,
                                                       (int)n_blocks, 0, 6, stream)) != cudaSuccess) return e;
    if ((e = cudaMalloc(&sch->temp, bytes ? bytes : 1)) != cudaSuccess) return e;
    if ((e = cudaMalloc((void**)&sch->stats, 4 * sizeof(unsigned long long))) != cudaSuccess) return e;
    sch->temp_bytes = bytes;
    sch->cap_blocks = n_blocks;
    sch->n_blocks = 0;   // no shape yet: the next launch arms the cost array
    iota_kernel<<<(n_blocks + 255u) / 256u, 256, 0, stream>>>(sch->iota, n_blocks);
    return cudaGetLastError();
}

// Make the schedule fit this launch; a launch of another shape (resolution, shard) starts from raster order again.
static cudaError_t schedule_prepare(RenderSchedule* sch, const DevParams& P, uint32_t n_blocks, cudaStream_t stream) {
    const bool same = sch->n_blocks == n_blocks && sch->W == P.W && sch->rows == P.n_rows_local && sch->n_shards == P.n_shards &&
                      sch->shard_index == P.shard_index && sch->tile_rows == P.tile_rows;
    if (same && sch->order) return cudaSuccess;
    cudaError_t e;
    if ((e = schedule_allocate(sch, n_blocks, stream)) != cudaSuccess) return e;
    if ((e = cudaMemsetAsync(sch->cost, 0, (size_t)n_blocks * sizeof(uint32_t), stream)) != cudaSuccess) return e;
    if ((e = cudaMemsetAsync(sch->stats, 0, 4 * sizeof(unsigned long long), stream)) != cudaSuccess) return e;
    sch->n_blocks = n_blocks; sch->W = P.W; sch->rows = P.n_rows_local; sch->n_shards = P.n_shards;
    sch->shard_index = P.shard_index; sch->tile_rows = P.tile_rows;
    sch->valid = false;
    return cudaGetLastError();
}

// Occupancy of the three render_kernel variants, per device (a process may hold scenes on several GPUs, and handles
// may be driven from several host threads): which = 0 plain, 1 with the dispatch-order hint, 2 counting.
static cudaError_t render_occupancy(int which, int* n_sm, int* per_sm) {
    constexpr int kMaxDev = 64;
    static std::mutex mu;
    static int sm[kMaxDev], occ[kMaxDev][3];   // zero-initialised: 0 = not asked yet
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    std::lock_guard<std::mutex> lock(mu);
    const bool cached = dev >= 0 && dev < kMaxDev && occ[dev][which] != 0;
    if (!cached) {
        int n = 0, b = 0;
        if ((e = cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
        e = which == 2   ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, render_kernel<true, false>, kBlock, 0)
            : which == 1 ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, render_kernel<false, true>, kBlock, 0)
                         : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, render_kernel<false, false>, kBlock, 0);
        if (e != cudaSuccess) return e;
        b = b > 0 ? b : 1;
        if (dev < 0 || dev >= kMaxDev) { *n_sm = n; *per_sm = b; return cudaSuccess; }
        sm[dev] = n;
        occ[dev][which] = b;
    }
    *n_sm = sm[dev];
    *per_sm = occ[dev][which];
    return cudaSuccess;
}

// cost[] -> order[] on stream `st`: {sum, max}, sort keys (which also clears cost[] for the next recording), one stable
// 6-bit radix sort of the block ids.
static cudaError_t schedule_update(RenderSchedule* sch, uint32_t n_blocks, float resident_warps, cudaStream_t st) {
    unsigned long long* stats = sch->stats + 2 * (sch->parity & 1);
    order_stats_kernel<<<(n_blocks + 1023u) / 1024u, 256, 0, st>>>(sch->cost, n_blocks, stats);
    order_keys_kernel<<<(n_blocks + 255u) / 256u, 256, 0, st>>>(sch->cost, sch->keys, n_blocks, stats,
                                                                sch->stats + 2 * (~sch->parity & 1), resident_warps, sch->tail_safety);
    sch->parity ^= 1;
    return cub::DeviceRadixSort::SortPairsDescending(sch->temp, sch->temp_bytes, sch->keys, sch->keys_alt, sch->iota, sch->order,
                                                     (int)n_blocks, 0, 6, st);
}

// scratch = (n_rows_local*W + 2) u32 words: per-pixel accumulators followed by the 64-bit queue head.
cudaError_t launch_render(const DevScene& sc, const DevParams& P, uint32_t* d_fb, int32_t* d_hit, uint32_t* d_scratch,
                          DevCounters* d_counters, cudaStream_t stream, RenderSchedule* sch) {
    const int tiles_x = (P.W + kTileW - 1) / kTileW, tiles_y = (P.n_rows_local + kTileH - 1) / kTileH;
    const unsigned long long total = (unsigned long long)tiles_x * tiles_y * (kTileW * kTileH * 4);
    if (total == 0) return cudaSuccess;
    const size_t npix = (size_t)P.n_rows_local * P.W;
    const size_t head = (npix + 1) & ~(size_t)1;  // 8-byte aligned
    cudaError_t e = cudaMemsetAsync(d_scratch, 0, (head + 2) * sizeof(uint32_t), stream);
    if (e != cudaSuccess) return e;
    unsigned long long* next = reinterpret_cast<unsigned long long*>(d_scratch + head);
    const uint32_t n_blocks = (uint32_t)tiles_x * (uint32_t)tiles_y;
    bool hint = sch && sch->enabled && !d_counters && RRT_REFILL_MIN == 32 && n_blocks >= sch->min_blocks;
    if (sch && sch->pending) {   // the previous frame's sort runs on the side stream: this frame starts after it
        if ((e = cudaStreamWaitEvent(stream, sch->ev_sorted, 0)) != cudaSuccess) return e;
        sch->pending = false;
    }
    if (hint && schedule_prepare(sch, P, n_blocks, stream) != cudaSuccess) {
        // no memory for the hint's arrays (5 words per block): render in raster order, now and from now on
        (void)cudaGetLastError();
        sch->release();
        sch->enabled = false;
        hint = false;
    }
    // persistent grid: one wave of the variant about to be launched, on the device that launches it
    int n_sm = 0, per_sm = 0;
    if ((e = render_occupancy(d_counters ? 2 : (hint ? 1 : 0), &n_sm, &per_sm)) != cudaSuccess) return e;
    const unsigned long long want = (total + kBlock - 1) / kBlock;
    const unsigned long long resident = (unsigned long long)n_sm * per_sm;
    const unsigned grid = (unsigned)(want < resident ? want : resident);
    const uint32_t* order = hint && sch->valid ? sch->order : nullptr;
    uint32_t* cost = hint ? sch->cost : nullptr;
    if (d_counters)
        render_kernel<true, false><<<grid, kBlock, 0, stream>>>(sc, P, d_fb, d_hit, d_scratch, next, d_counters, tiles_x, total, nullptr, nullptr);
    else if (hint)
        render_kernel<false, true><<<grid, kBlock, 0, stream>>>(sc, P, d_fb, d_hit, d_scratch, next, nullptr, tiles_x, total, order, cost);
    else
        render_kernel<false, false><<<grid, kBlock, 0, stream>>>(sc, P, d_fb, d_hit, d_scratch, next, nullptr, tiles_x, total, nullptr, nullptr);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    if (!d_counters && P.search_mode == 1) search_switch_kernel<<<1, 1, 0, stream>>>(P.fstat);
    if (hint) {
        // next frame's order: stable sort of the blocks by this frame's cost class, on the side stream, so that it
        // overlaps whatever the caller does with the frame next (copy to the host, gather, L2 flush of a benchmark)
        cudaStream_t side = sch->side;
        if ((e = cudaEventRecord(sch->ev_rendered, stream)) != cudaSuccess) return e;
        if ((e = cudaStreamWaitEvent(side, sch->ev_rendered, 0)) != cudaSuccess) return e;
        if ((e = schedule_update(sch, n_blocks, (float)grid * (kBlock / 32), side)) != cudaSuccess) return e;
        if ((e = cudaEventRecord(sch->ev_sorted, side)) != cudaSuccess) return e;
        sch->pending = true;
        sch->valid = true;
    }
    return cudaGetLastError();
}

// Everything launch_render would otherwise set up inside the FIRST frame of shape P (rrt_scene_reserve): the schedule's
// arrays, side stream and events, and the occupancy queries of the kernel variants.
cudaError_t reserve_render(const DevParams& P, RenderSchedule* sch, cudaStream_t stream) {
    int n_sm = 0, per_sm = 0;
    cudaError_t e;
    for (int which = 0; which < 2; ++which)
        if ((e = render_occupancy(which, &n_sm, &per_sm)) != cudaSuccess) return e;
    const uint32_t tiles_x = (uint32_t)((P.W + kTileW - 1) / kTileW), tiles_y = (uint32_t)((P.n_rows_local + kTileH - 1) / kTileH);
    const uint32_t n_blocks = tiles_x * tiles_y;
    if (!sch || !sch->enabled || n_blocks < sch->min_blocks) return cudaSuccess;
    if (schedule_allocate(sch, n_blocks, stream) != cudaSuccess) {   // as in launch_render: no memory, no hint
        (void)cudaGetLastError();
        sch->release();
        sch->enabled = false;
    }
    return cudaSuccess;
}

#if RRT_TIMELINE
extern "C" int rrt_debug_timeline(unsigned long long* out, int reset) {
    if (out && cudaMemcpyFromSymbol(out, g_tl, sizeof(g_tl)) != cudaSuccess) return 1;
    if (reset) {
        unsigned long long z[4 + 256] = {~0ull, ~0ull, 0ull, 0ull};
        if (cudaMemcpyToSymbol(g_tl, z, sizeof(z)) != cudaSuccess) return 1;
    }
    return 0;
}
#endif


size_t render_scratch_words(int n_rows_local, int W) { return (((size_t)n_rows_local * W + 1) & ~(size_t)1) + 2; }

static_assert(sizeof(g_noalloc_scratch) / sizeof(uint32_t) == (size_t)kNoAllocMaxPixels + 4, "static scratch size");

// Device address of g_noalloc_scratch on the CURRENT device, resolved once per device (set-up of the allocation-free
// call): cudaGetSymbolAddress with that device current, return code checked, result cached under a mutex.
static cudaError_t noalloc_scratch_ptr(uint32_t** out) {
    constexpr int kMaxDev = 64;
    static std::mutex mu;
    static uint32_t* ptr[kMaxDev];   // zero-initialised: nullptr = not resolved yet
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= kMaxDev) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(mu);
    if (!ptr[dev]) {
        void* p = nullptr;
        if ((e = cudaGetSymbolAddress(&p, g_noalloc_scratch)) != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        ptr[dev] = static_cast<uint32_t*>(p);
    }
    *out = ptr[dev];
    return cudaSuccess;
}

// launch_render without any allocation: the scratch is the static device array above, the dispatch-order hint (whose
// arrays and cub temporary are allocated on demand) is left out, so the kernel is render_kernel<false, false>.  The
// array keeps the previous call's contents: the part this call uses is cleared first (cudaMemsetAsync on `stream`).
// Every call on a device shares the array — calls that overlap on the GPU race.
cudaError_t launch_render_noalloc(const DevScene& sc, const DevParams& P, uint32_t* d_fb, int32_t* d_hit, cudaStream_t stream) {
    if ((size_t)P.n_rows_local * (size_t)P.W > (size_t)kNoAllocMaxPixels) return cudaErrorInvalidValue;
    const int tiles_x = (P.W + kTileW - 1) / kTileW, tiles_y = (P.n_rows_local + kTileH - 1) / kTileH;
    const unsigned long long total = (unsigned long long)tiles_x * tiles_y * (kTileW * kTileH * 4);
    if (total == 0) return cudaSuccess;
    uint32_t* scratch = nullptr;
    cudaError_t e = noalloc_scratch_ptr(&scratch);
    if (e != cudaSuccess) return e;
    const size_t npix = (size_t)P.n_rows_local * P.W;
    const size_t head = (npix + 1) & ~(size_t)1;  // 8-byte aligned
    if ((e = cudaMemsetAsync(scratch, 0, (head + 2) * sizeof(uint32_t), stream)) != cudaSuccess) return e;
    int n_sm = 0, per_sm = 0;
    if ((e = render_occupancy(0, &n_sm, &per_sm)) != cudaSuccess) return e;
    const unsigned long long want = (total + kBlock - 1) / kBlock;
    const unsigned long long resident = (unsigned long long)n_sm * per_sm;
    const unsigned grid = (unsigned)(want < resident ? want : resident);
    unsigned long long* next = reinterpret_cast<unsigned long long*>(scratch + head);
    render_kernel<false, false><<<grid, kBlock, 0, stream>>>(sc, P, d_fb, d_hit, scratch, next, nullptr, tiles_x, total, nullptr, nullptr);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    if (P.search_mode == 1) search_switch_kernel<<<1, 1, 0, stream>>>(P.fstat);
    return cudaGetLastError();
}

cudaError_t launch_trace(const DevScene& sc, uint64_t n, const double* o, const double* d, const double* max_t,
                         int32_t* out_tri, double* out_tuv, bool searches, cudaStream_t stream) {
    if (n == 0) return cudaSuccess;
    trace_kernel<<<(unsigned)((n + kBlock - 1) / kBlock), kBlock, 0, stream>>>(sc, n, o, d, max_t, out_tri, out_tuv,
                                                                                searches ? (kTryFilter | kTrySparse) : 0u);
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

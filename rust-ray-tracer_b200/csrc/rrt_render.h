// rrt_render.h — host-visible launchers of the kernels in rrt_render.cu.
#pragma once
#include <cuda_runtime.h>

#include "rrt_device.cuh"

namespace rrt {

struct DevParams;

// Dispatch-order hint carried from one frame of a scene handle to the next.  The persistent kernel ends when its
// slowest late-fetched warp tile ends (a heavy tile's dependent chain lasts 0.2-0.6 ms where the mean is < 0.1 ms),
// so every launch records what each 16x4-pixel block cost and the NEXT launch of the same shape hands out the last
// part of its blocks heaviest first: the tail is then made of the cheapest tiles (order_keys_kernel).  Only the
// ORDER of the work changes — every sample is traced in full every frame and lands in the same per-pixel
// accumulator, so frames are bit-identical with and without the hint (tests/test_gpu_properties.py).  The first
// frame of a shape runs in raster order.  Measurements: DESIGN.md §6, profiles/r01_v8_c4_drain_timeline.txt.
struct RenderSchedule {
    uint32_t *order = nullptr, *cost = nullptr, *keys = nullptr, *keys_alt = nullptr, *iota = nullptr;
    void* temp = nullptr;
    unsigned long long* stats = nullptr;        // 2 x {sum, max} of the block costs, used alternately
    int parity = 0;
    cudaStream_t side = nullptr;                // the sort of frame N runs here, behind ev_rendered; frame N+1 waits for ev_sorted
    cudaEvent_t ev_rendered = nullptr, ev_sorted = nullptr;
    bool pending = false;
    float tail_safety = 4.0f;                   // share of the sequence that is sorted = tail_safety x slowest tile / frame (<= 0: all)
    size_t temp_bytes = 0, cap_blocks = 0;
    uint32_t n_blocks = 0, min_blocks = 1024;   // frames smaller than min_blocks blocks are not worth two more launches
    int W = 0, rows = 0, n_shards = 0, shard_index = 0, tile_rows = 0;
    bool valid = false, enabled = true;
    void release();
};

// d_scratch: render_scratch_words(P.n_rows_local, P.W) u32 words of device memory (zeroed by the launcher)
cudaError_t launch_render(const DevScene& sc, const DevParams& P, uint32_t* d_fb, int32_t* d_hit, uint32_t* d_scratch,
                          DevCounters* d_counters, cudaStream_t stream, RenderSchedule* sch = nullptr);
size_t render_scratch_words(int n_rows_local, int W);
// what launch_render would set up inside the first frame of shape P (schedule arrays, occupancy queries), done ahead
cudaError_t reserve_render(const DevParams& P, RenderSchedule* sch, cudaStream_t stream);
// Allocation-free variant (rrt_render_device_noalloc): scratch in a static device array of rrt_render.cu, at most
// kNoAllocMaxPixels pixels per call (cudaErrorInvalidValue beyond), no dispatch-order hint.
constexpr uint32_t kNoAllocMaxPixels = 1u << 21;   // == RRT_NOALLOC_MAX_PIXELS (include/rrt.h): 2 Mi pixels, 8 MiB of scratch
cudaError_t launch_render_noalloc(const DevScene& sc, const DevParams& P, uint32_t* d_fb, int32_t* d_hit, cudaStream_t stream);
// searches: let rays be answered by the scene-wide hierarchy where that is exact (rrt_trace.cuh); false = walk only
cudaError_t launch_trace(const DevScene& sc, uint64_t n, const double* o, const double* d, const double* max_t,
                         int32_t* out_tri, double* out_tuv, bool searches, cudaStream_t stream);
cudaError_t launch_colours(const DevScene& sc, const DevParams& P, uint64_t n, const double* o, const double* d,
                           uint8_t* out_rgb, cudaStream_t stream);
cudaError_t launch_deinterleave(const uint32_t* g, int rows_per_block, int W, int H, int T, int G, uint32_t* fb,
                                cudaStream_t stream);

}  // namespace rrt

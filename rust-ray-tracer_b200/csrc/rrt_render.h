// rrt_render.h — host-visible launchers of the kernels in rrt_render.cu.
#pragma once
#include <cuda_runtime.h>

#include "rrt_device.cuh"

namespace rrt {

struct DevParams;

// d_scratch: render_scratch_words(P.n_rows_local, P.W) u32 words of device memory (zeroed by the launcher)
cudaError_t launch_render(const DevScene& sc, const DevParams& P, uint32_t* d_fb, int32_t* d_hit, uint32_t* d_scratch,
                          DevCounters* d_counters, cudaStream_t stream);
size_t render_scratch_words(int n_rows_local, int W);
cudaError_t launch_trace(const DevScene& sc, uint64_t n, const double* o, const double* d, const double* max_t,
                         int32_t* out_tri, double* out_tuv, cudaStream_t stream);
cudaError_t launch_colours(const DevScene& sc, const DevParams& P, uint64_t n, const double* o, const double* d,
                           uint8_t* out_rgb, cudaStream_t stream);
cudaError_t launch_deinterleave(const uint32_t* g, int rows_per_block, int W, int H, int T, int G, uint32_t* fb,
                                cudaStream_t stream);

}  // namespace rrt

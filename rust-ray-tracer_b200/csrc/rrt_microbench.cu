// rrt_microbench.cu — measures the roofline denominator this path is bound by: FP64 vector throughput
// WITHOUT fused multiply-add (the reference's Rust f64 never contracts a*b+c, so the kernels are built
// with -fmad=false and every multiply and add is its own DMUL / DADD).  MEASURED_PEAKS.json only has
// HBM and bf16 tensor numbers (SURVEY §8d asks for this measurement).
#include <cuda_runtime.h>

#include "../../include/rrt.h"
#include "rrt_trace.cuh"

extern "C" void rrt_set_last_error_(const char* msg);

namespace {

// 8 independent chains per thread, each iteration = 8 DMUL + 8 DADD (16 flops), fully dependent within a
// chain so the compiler cannot hoist; compiled with -fmad=false like the rest of the library.
__global__ void __launch_bounds__(256) fp64_chain(double* out, int iters, double m, double a) {
    double x0 = threadIdx.x * 1e-3, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
#pragma unroll 4
    for (int i = 0; i < iters; ++i) {
        x0 = x0 * m + a; x1 = x1 * m + a; x2 = x2 * m + a; x3 = x3 * m + a;
        x4 = x4 * m + a; x5 = x5 * m + a; x6 = x6 * m + a; x7 = x7 * m + a;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
}

// Self-test of the FMA quotient path (rrt_trace.cuh: div_markstein) against the hardware's IEEE division
// on pseudo-random operand pairs inside the path's range condition, including the classic hard cases
// (divisor / dividend mantissas of all ones, all zeros, near powers of two).
__device__ __forceinline__ unsigned long long splitmix(unsigned long long& s) {
    unsigned long long z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__device__ __forceinline__ double shape(unsigned long long r, unsigned long long m) {
    // exponent in [2^-400, 2^400], mantissa pattern chosen by the low bits of r
    const int e = 1023 - 400 + (int)((r >> 8) % 801u);
    unsigned long long mant = m & 0xFFFFFFFFFFFFFull;
    switch (r & 7u) {
        case 0: mant = 0xFFFFFFFFFFFFFull; break;                 // all ones
        case 1: mant = 0; break;                                   // power of two
        case 2: mant = 0xFFFFFFFFFFFFFull - (m & 0xFF); break;     // just below all ones
        case 3: mant = m & 0xFF; break;                            // just above a power of two
        default: break;
    }
    return __longlong_as_double((long long)(((r >> 63) << 63) | ((unsigned long long)e << 52) | mant));
}
__global__ void division_selftest(unsigned long long seed, int per_thread, unsigned long long* mismatches) {
    unsigned long long s = seed + 0x1234567ull * (blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x);
    unsigned long long bad = 0;
    for (int i = 0; i < per_thread; ++i) {
        const double d = shape(splitmix(s), splitmix(s));
        const double y = 1.0 / d;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const double a = shape(splitmix(s), splitmix(s));
            const double q = rrt::div_markstein(a, d, y), ref = a / d;
            bad += (__double_as_longlong(q) != __double_as_longlong(ref)) ? 1ull : 0ull;
        }
    }
    if (bad) atomicAdd(mismatches, bad);
}

}  // namespace

extern "C" int rrt_selftest_division(int device, uint64_t n_pairs, uint64_t* mismatches) {
    if (!mismatches) { rrt_set_last_error_("null argument"); return RRT_ERR_INVALID; }
    if (cudaSetDevice(device) != cudaSuccess) { rrt_set_last_error_("cudaSetDevice failed"); return RRT_ERR_CUDA; }
    unsigned long long* d = nullptr;
    if (cudaMalloc(&d, 8) != cudaSuccess) { rrt_set_last_error_("cudaMalloc failed"); return RRT_ERR_CUDA; }
    cudaMemset(d, 0, 8);
    const int threads = 256, blocks = 148 * 8;
    const uint64_t per_thread = n_pairs / ((uint64_t)threads * blocks * 4) + 1;
    division_selftest<<<blocks, threads>>>(0xC0FFEEull, (int)per_thread, d);
    unsigned long long h = 0;
    const cudaError_t e = cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    cudaFree(d);
    if (e != cudaSuccess) { rrt_set_last_error_(cudaGetErrorString(e)); return RRT_ERR_CUDA; }
    *mismatches = h;
    return RRT_OK;
}

extern "C" int rrt_microbench_fp64(int device, double* tflops_no_fma, double* sm_clock_mhz) {
    if (!tflops_no_fma) { rrt_set_last_error_("null argument"); return RRT_ERR_INVALID; }
    if (cudaSetDevice(device) != cudaSuccess) { rrt_set_last_error_("cudaSetDevice failed"); return RRT_ERR_CUDA; }
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, device);
    const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 1 << 14;
    double* d = nullptr;
    if (cudaMalloc(&d, sizeof(double) * blocks * threads) != cudaSuccess) { rrt_set_last_error_("cudaMalloc failed"); return RRT_ERR_CUDA; }
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 6; ++rep) {
        cudaEventRecord(e0);
        fp64_chain<<<blocks, threads>>>(d, iters, 0.9999999, 1e-7);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep >= 2 && ms < best) best = ms;
    }
    const cudaError_t e = cudaGetLastError();
    cudaFree(d);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (e != cudaSuccess) { rrt_set_last_error_(cudaGetErrorString(e)); return RRT_ERR_CUDA; }
    const double flops = 16.0 * iters * (double)blocks * threads;
    *tflops_no_fma = flops / (best * 1e-3) / 1e12;
    if (sm_clock_mhz) *sm_clock_mhz = prop.clockRate / 1e3;
    return RRT_OK;
}

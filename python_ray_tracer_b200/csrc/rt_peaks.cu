// rt_peaks.cu -- FMA-rate micro-kernels: the ALU roofline denominators of the tracer.
//
// The trace kernel is bound by the FP64 pipe (geometry and shading are double, see
// DESIGN.md), which MEASURED_PEAKS.json does not cover (it holds HBM and bf16 tensor
// numbers).  rt_measure_fma_peak() times a register-resident chain of independent
// FMAs on every SM and returns the achieved FLOP/s (2 per FMA), so bench.py can state
// the tracer's arithmetic rate as a fraction of a peak measured on the same GPU in the
// same run.
#include <cuda_runtime.h>

#include "../../include/rt_abi.h"

namespace {

template <typename T>
__global__ void __launch_bounds__(256) fma_chain_kernel(T* sink, int iters, T a, T b) {
    // 16 independent accumulators per thread: enough ILP to cover the pipe latency at
    // 8 warps per scheduler.
    T acc[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = (T)(threadIdx.x + i);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[i] = acc[i] * a + b;
    }
    T sum = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) sum += acc[i];
    if (sum == (T)-1.2345) sink[0] = sum;  // never true; keeps the chain alive
}

template <typename T>
int measure(int device, double* flops_out) {
    int previous = -1;
    cudaGetDevice(&previous);
    if (cudaSetDevice(device) != cudaSuccess) return RT_ERR_NO_DEVICE;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return RT_ERR_CUDA;
    T* sink = nullptr;
    if (cudaMalloc(&sink, sizeof(T)) != cudaSuccess) return RT_ERR_CUDA;
    const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 4096;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    double best = 0.0;
    for (int rep = 0; rep < 6; ++rep) {
        cudaEventRecord(e0, 0);
        fma_chain_kernel<T><<<blocks, threads>>>(sink, iters, (T)0.999999, (T)1e-7);
        cudaEventRecord(e1, 0);
        if (cudaEventSynchronize(e1) != cudaSuccess) break;
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        double flops = 2.0 * 16.0 * (double)iters * (double)blocks * threads / (ms * 1e-3);
        if (rep > 0 && flops > best) best = flops;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(sink);
    cudaError_t err = cudaGetLastError();
    if (previous >= 0) cudaSetDevice(previous);
    if (err != cudaSuccess) return RT_ERR_CUDA;
    *flops_out = best;
    return RT_OK;
}

}  // namespace

extern "C" int rt_measure_fma_peak(int device, int32_t dtype, double* flops_out) {
    if (flops_out == nullptr) return RT_ERR_BAD_ARG;
    if (dtype == RT_DTYPE_F64) return measure<double>(device, flops_out);
    if (dtype == RT_DTYPE_F32) return measure<float>(device, flops_out);
    return RT_ERR_BAD_ARG;
}

// rt_launch_impl.cuh -- body shared by trace_exact.cu and trace_fast.cu.  The including
// file defines RT_TU_MODE (rt::kModeExact or rt::kModeFast) and RT_TU_NAME.
#pragma once

#include "rt_kernels.cuh"
#include "rt_launch.h"

namespace rt {
namespace {

constexpr int kMode = RT_TU_MODE;
constexpr size_t kMaxDynamicSmem = 227 * 1024 - 64;  // per-block limit on sm_100a minus statics

template <typename Kernel>
cudaError_t prepare(Kernel kernel, size_t smem) {
    if (smem > kMaxDynamicSmem) return cudaErrorInvalidValue;
    if (smem > 48 * 1024) return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    return cudaSuccess;
}

cudaError_t render_rows(bool stats, const SceneParams& sp, const CameraParams& cp, const OutputParams& out,
                        cudaStream_t stream) {
    const int rows = cp.row1 - cp.row0;
    if (rows <= 0 || cp.width <= 0) return cudaSuccess;
    const size_t smem = staged_bytes(sp.n_spheres);
    dim3 grid((cp.width + kTileW - 1) / kTileW, (rows + kTileH - 1) / kTileH);
    cudaError_t err;
    if (stats) {
        if ((err = prepare(render_rows_kernel<kMode, true>, smem)) != cudaSuccess) return err;
        render_rows_kernel<kMode, true><<<grid, kBlockThreads, smem, stream>>>(sp, cp, out);
    } else {
        if ((err = prepare(render_rows_kernel<kMode, false>, smem)) != cudaSuccess) return err;
        render_rows_kernel<kMode, false><<<grid, kBlockThreads, smem, stream>>>(sp, cp, out);
    }
    return cudaGetLastError();
}

cudaError_t trace_rays(bool stats, const SceneParams& sp, const RayBatch& rb, const OutputParams& out,
                       cudaStream_t stream) {
    if (rb.n <= 0) return cudaSuccess;
    const size_t smem = staged_bytes(sp.n_spheres);
    const unsigned int grid = (unsigned int)((rb.n + kBlockThreads - 1) / kBlockThreads);
    cudaError_t err;
    if (stats) {
        if ((err = prepare(trace_rays_kernel<kMode, true>, smem)) != cudaSuccess) return err;
        trace_rays_kernel<kMode, true><<<grid, kBlockThreads, smem, stream>>>(sp, rb, out);
    } else {
        if ((err = prepare(trace_rays_kernel<kMode, false>, smem)) != cudaSuccess) return err;
        trace_rays_kernel<kMode, false><<<grid, kBlockThreads, smem, stream>>>(sp, rb, out);
    }
    return cudaGetLastError();
}

cudaError_t sphere_distance(const RayBatch& rb, const double* c, double radius, double* out,
                            cudaStream_t stream) {
    if (rb.n <= 0) return cudaSuccess;
    SphereGeom g = {c[0], c[1], c[2], radius * radius};
    double cc = (c[0] * c[0]) + (c[1] * c[1]) + (c[2] * c[2]);
    const unsigned int grid = (unsigned int)((rb.n + 255) / 256);
    sphere_distance_kernel<kMode><<<grid, 256, 0, stream>>>(rb, g, cc, out);
    return cudaGetLastError();
}

cudaError_t ray_directions(const CameraParams& cp, double* out, long long plane_stride, cudaStream_t stream) {
    const long long n = (long long)(cp.row1 - cp.row0) * cp.width;
    if (n <= 0) return cudaSuccess;
    const unsigned int grid = (unsigned int)((n + 255) / 256);
    ray_directions_kernel<kMode><<<grid, 256, 0, stream>>>(cp, out, plane_stride);
    return cudaGetLastError();
}

const Launchers kLaunchers = {render_rows, trace_rays, sphere_distance, ray_directions};

}  // namespace
}  // namespace rt

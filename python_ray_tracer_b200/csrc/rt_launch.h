// rt_launch.h -- host-callable launchers, one set per arithmetic policy.
#pragma once

#include <cuda_runtime.h>

#include "rt_device.cuh"

namespace rt {

struct Launchers {
    // `small`: host-built table of a scene of <= kSmallMax spheres (fast policy), or nullptr
    cudaError_t (*render_rows)(bool stats, const SceneParams&, const CameraParams&, const OutputParams&,
                               const SmallTable* small, cudaStream_t);
    cudaError_t (*trace_rays)(bool stats, const SceneParams&, const RayBatch&, const OutputParams&,
                              cudaStream_t);
    cudaError_t (*sphere_distance)(const RayBatch&, const double* centre, double radius, double* out,
                                   cudaStream_t);
    cudaError_t (*ray_directions)(const CameraParams&, double* out, long long plane_stride, cudaStream_t);
};

const Launchers& exact_launchers();  // trace_exact.cu, compiled with -fmad=false
const Launchers& fast_launchers();   // trace_fast.cu

// One zeroed ticket counter (device memory of the current device) for the next persistent
// launch; see rt_kernels.cuh:Tickets.  Counters come from a per-device ring so launches
// in flight on different streams never share one.  Defined in rt_abi.cu.
cudaError_t ticket_counter(unsigned int** counter_out);
void release_ticket_counters();

// The word kernels of the current device report errors through (DeviceError, rt_device.cuh): pinned host memory
// mapped into the device, so the host reads it without touching a stream.  Defined in rt_abi.cu.
cudaError_t device_error_word(unsigned int** device_pointer_out);

// Largest scene whose sphere records fit one block's shared memory.
int max_staged_spheres();

}  // namespace rt

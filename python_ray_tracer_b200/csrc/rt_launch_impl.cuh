// rt_launch_impl.cuh -- body shared by trace_exact.cu and trace_fast.cu.  The including
// file defines RT_TU_MODE (rt::kModeExact or rt::kModeFast) and RT_TU_NAME.
#pragma once

#include "rt_kernels.cuh"
#include "rt_launch.h"

namespace rt {
namespace {

constexpr int kMode = RT_TU_MODE;
constexpr size_t kMaxDynamicSmem = 227 * 1024 - 64;  // per-block limit on sm_100a minus statics

// One big block per SM once RT_MIN_BLOCKS private copies of the sphere table (plus the 1 KB
// the driver reserves per block) no longer fit an SM's shared memory.
inline bool big_blocks(int n_spheres) { return (staged_bytes(n_spheres) + 1024 + 16) * RT_MIN_BLOCKS > kSmemPerSm; }

// Persistent launch shape: as many blocks as the device keeps resident for this kernel at
// this shared-memory size, but no more than there are tickets to hand out.
template <typename Kernel>
cudaError_t persistent_grid(Kernel kernel, int block_threads, size_t smem, unsigned int n_tickets, unsigned int* grid_out) {
    if (smem > kMaxDynamicSmem) return cudaErrorInvalidValue;
    // the answer of the last query per host thread (kernels of one signature share this
    // function instance, so the kernel address is part of the key)
    static thread_local struct { const void* kernel; int device; size_t smem; unsigned int slots; } cached = {nullptr, -1, 0, 0};
    cudaError_t err;
    int device = 0;
    if ((err = cudaGetDevice(&device)) != cudaSuccess) return err;
    if (cached.kernel != (const void*)kernel || cached.device != device || cached.smem != smem) {
        if (smem > 48 * 1024 &&
            (err = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess)
            return err;
        int sms = 0, per_sm = 0;
        if ((err = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device)) != cudaSuccess) return err;
        if ((err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, block_threads, smem)) != cudaSuccess) return err;
        if (per_sm < 1) return cudaErrorLaunchOutOfResources;
        cached = {(const void*)kernel, device, smem, (unsigned int)sms * (unsigned int)per_sm};
    }
    const unsigned int slots = cached.slots;
    const unsigned int block_warps = (unsigned int)block_threads / 32u;
    const unsigned int wanted = (n_tickets + block_warps - 1) / block_warps;
    *grid_out = wanted < slots ? wanted : slots;
    return cudaSuccess;
}

inline cudaError_t make_tickets(Tickets& tk, const SceneParams& sp, const OutputParams& out, unsigned int n_items, unsigned int grid,
                                int block_threads, bool camera) {
    cudaError_t err;
    if ((err = ticket_counter(&tk.counter)) != cudaSuccess) return err;
    tk.n_warps = grid * ((unsigned int)block_threads / 32u);
    tk.n_items = n_items;
    const unsigned int rounds = n_items / tk.n_warps;
    const bool all_tickets = rounds <= RT_SHORT_LAUNCH_ROUNDS || sp.n_spheres >= RT_COSTLY_ITEM_SPHERES;
    if (all_tickets) {
        tk.reversed = camera ? (out.wide_patches && out.even_stores ? 2u : 1u) : 0u;
        tk.static_rounds = rounds >= 1u ? 1u : 0u;
    } else {
        tk.reversed = 0u;
        tk.static_rounds = rounds - (unsigned int)((unsigned long long)rounds * RT_DYNAMIC_PCT / 100ull);
    }
    tk.n_tickets = n_items - tk.static_rounds * tk.n_warps;
    // big blocks (items of >= 100 us) and scenes of costly items keep single tickets
    const bool chunks = block_threads == kBlockThreads && sp.n_spheres < RT_COSTLY_ITEM_SPHERES;
    const unsigned int T = tk.n_tickets;
    const unsigned int hi = chunks && RT_CHUNK4_ROUNDS > 0 ? RT_CHUNK4_ROUNDS * tk.n_warps : 0xffffffffu;
    const unsigned int mid = chunks && RT_CHUNK2_ROUNDS > 0 ? RT_CHUNK2_ROUNDS * tk.n_warps : 0xffffffffu;
    tk.n4 = T > hi ? (T - hi + 3u) / 4u : 0u;
    tk.cover4 = 4u * tk.n4 < T ? 4u * tk.n4 : T;
    const unsigned int rest = T - tk.cover4;
    tk.n2 = rest > mid ? (rest - mid + 1u) / 2u : 0u;
    tk.cover2 = 2u * tk.n2 < rest ? 2u * tk.n2 : rest;
    tk.n_draws = tk.n4 + tk.n2 + (rest - tk.cover2);
    return cudaSuccess;
}

template <typename Kernel, typename Work>
cudaError_t launch_persistent(Kernel kernel, int block_threads, const SceneParams& sp, const Work& work,
                              const OutputParams& out, unsigned long long n_items64, cudaStream_t stream,
                              bool camera = false) {
    if (n_items64 > 0x7ffffff0ull) return cudaErrorInvalidValue;  // 31-bit items (reciprocal division in the kernels)
    const unsigned int n_items = (unsigned int)n_items64;
    const size_t smem = sp.accel != nullptr ? sizeof(double) * 3 * (size_t)sp.n_lights_used + 16
                                            : staged_bytes(sp.n_spheres, sp.n_lights_used);
    unsigned int grid = 0;
    cudaError_t err = persistent_grid(kernel, block_threads, smem, n_items, &grid);
    if (err != cudaSuccess) return err;
    Tickets tk;
    if ((err = make_tickets(tk, sp, out, n_items, grid, block_threads, camera)) != cudaSuccess) return err;
    kernel<<<grid, block_threads, smem, stream>>>(sp, work, out, tk);
    return cudaGetLastError();
}

template <typename Kernel>
cudaError_t launch_small(Kernel kernel, const SceneParams& sp, const CameraParams& cp, const OutputParams& out,
                         const SmallTable& table, unsigned long long n_items64, cudaStream_t stream) {
    if (n_items64 > 0x7ffffff0ull) return cudaErrorInvalidValue;
    const unsigned int n_items = (unsigned int)n_items64;
    unsigned int grid = 0;
    cudaError_t err = persistent_grid(kernel, kBlockThreads, 0, n_items, &grid);
    if (err != cudaSuccess) return err;
    Tickets tk;
    if ((err = make_tickets(tk, sp, out, n_items, grid, kBlockThreads, true)) != cudaSuccess) return err;
    kernel<<<grid, kBlockThreads, 0, stream>>>(sp, cp, out, tk, table);
    return cudaGetLastError();
}

template <int PW, int N>
cudaError_t render_rows_small(bool stats, const SceneParams& sp, const CameraParams& cp, const OutputParams& out,
                              const SmallTable& table, unsigned long long n, cudaStream_t stream) {
    if constexpr (kMode == kModeFast) {
        if (sp.n_spheres == N) {
            if (stats) return launch_small(render_rows_small_kernel<true, N, PW>, sp, cp, out, table, n, stream);
            return launch_small(render_rows_small_kernel<false, N, PW>, sp, cp, out, table, n, stream);
        }
        if constexpr (N > 1) return render_rows_small<PW, N - 1>(stats, sp, cp, out, table, n, stream);
    }
    return cudaErrorInvalidValue;
}

template <int PW>
cudaError_t render_rows_patch(bool stats, const SceneParams& sp, const CameraParams& cp, const OutputParams& out,
                              const SmallTable* small, cudaStream_t stream) {
    const int rows = cp.n_rows;
    constexpr int kPatchH = 32 / PW;
    const unsigned long long tiles_x = (unsigned long long)(cp.width + PW - 1) / PW;
    const unsigned long long tiles_y = (unsigned long long)(rows + kPatchH - 1) / kPatchH;
    const unsigned long long n = tiles_x * tiles_y;
    if (sp.accel != nullptr) {
        if (stats) return launch_persistent(render_rows_kernel<kMode, true, kBlockThreads, 1, PW>, kBlockThreads, sp, cp, out, n, stream, true);
        return launch_persistent(render_rows_kernel<kMode, false, kBlockThreads, 1, PW>, kBlockThreads, sp, cp, out, n, stream, true);
    }
    if (kMode == kModeFast && small != nullptr && sp.n_spheres <= kSmallMax && sp.n_lights_used == 1)
        return render_rows_small<PW, kSmallMax>(stats, sp, cp, out, *small, n, stream);
    if (big_blocks(sp.n_spheres)) {
        if (stats) return launch_persistent(render_rows_kernel<kMode, true, kBigBlockThreads, 0, PW>, kBigBlockThreads, sp, cp, out, n, stream, true);
        return launch_persistent(render_rows_kernel<kMode, false, kBigBlockThreads, 0, PW>, kBigBlockThreads, sp, cp, out, n, stream, true);
    }
    if (stats) return launch_persistent(render_rows_kernel<kMode, true, kBlockThreads, 0, PW>, kBlockThreads, sp, cp, out, n, stream, true);
    return launch_persistent(render_rows_kernel<kMode, false, kBlockThreads, 0, PW>, kBlockThreads, sp, cp, out, n, stream, true);
}

cudaError_t render_rows(bool stats, const SceneParams& sp, const CameraParams& cp, const OutputParams& out,
                        const SmallTable* small, cudaStream_t stream) {
    if (cp.n_rows <= 0 || cp.width <= 0) return cudaSuccess;
    if (out.wide_patches) return render_rows_patch<kWideTileW>(stats, sp, cp, out, small, stream);
    return render_rows_patch<kWarpTileW>(stats, sp, cp, out, small, stream);
}

cudaError_t trace_rays(bool stats, const SceneParams& sp, const RayBatch& rb, const OutputParams& out,
                       cudaStream_t stream) {
    if (rb.n <= 0) return cudaSuccess;
    const unsigned long long n_items = ((unsigned long long)rb.n + 31ull) / 32ull;
    if (sp.accel != nullptr) {
        if (stats) return launch_persistent(trace_rays_kernel<kMode, true, kBlockThreads, 1>, kBlockThreads, sp, rb, out, n_items, stream);
        return launch_persistent(trace_rays_kernel<kMode, false, kBlockThreads, 1>, kBlockThreads, sp, rb, out, n_items, stream);
    }
    if (big_blocks(sp.n_spheres)) {
        if (stats) return launch_persistent(trace_rays_kernel<kMode, true, kBigBlockThreads, 0>, kBigBlockThreads, sp, rb, out, n_items, stream);
        return launch_persistent(trace_rays_kernel<kMode, false, kBigBlockThreads, 0>, kBigBlockThreads, sp, rb, out, n_items, stream);
    }
    if (stats) return launch_persistent(trace_rays_kernel<kMode, true, kBlockThreads, 0>, kBlockThreads, sp, rb, out, n_items, stream);
    return launch_persistent(trace_rays_kernel<kMode, false, kBlockThreads, 0>, kBlockThreads, sp, rb, out, n_items, stream);
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
    const long long n = (long long)cp.n_rows * cp.width;
    if (n <= 0) return cudaSuccess;
    const unsigned int grid = (unsigned int)((n + 255) / 256);
    ray_directions_kernel<kMode><<<grid, 256, 0, stream>>>(cp, out, plane_stride);
    return cudaGetLastError();
}

const Launchers kLaunchers = {render_rows, trace_rays, sphere_distance, ray_directions};

}  // namespace
}  // namespace rt

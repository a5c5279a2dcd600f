// rt_kernels.cuh -- __global__ entry points, instantiated once per arithmetic policy by
// trace_exact.cu (-fmad=false) and trace_fast.cu.
#pragma once

#include "rt_chain.cuh"

namespace rt {

constexpr int kTileW = 32;        // pixels per block, x
constexpr int kTileH = 4;         // pixels per block, y
constexpr int kBlockThreads = kTileW * kTileH;
constexpr int kWarpTileW = 8;     // each warp owns an 8x4 patch: bounce coherence, and
                                  // every row segment it stores is one full 32-byte sector

__host__ __device__ inline size_t staged_bytes(int n_spheres) {
    return (sizeof(SphereGeom) + sizeof(SphereAux)) * (size_t)n_spheres;
}

// np.linspace sample i (base.py:132-139): i*step + start, last sample = stop exactly.
__device__ __forceinline__ double linspace_at(double start, double stop, double step, int num, int i) {
    if (num <= 1) return start;
    if (i == num - 1) return stop;
    return (double)i * step + start;
}

// get_ray_directions (base.py:123-141) for one pixel.
template <int MODE>
__device__ __forceinline__ Vec3 primary_direction(const CameraParams& cp, int col, int row) {
    double x = linspace_at(cp.x_start, cp.x_stop, cp.x_step, cp.width, col);
    double y = linspace_at(cp.y_start, cp.y_stop, cp.y_step, cp.height, row);
    Vec3 q = {x - cp.cam[0], y - cp.cam[1], 0.0 - cp.cam[2]};
    return MODE == kModeExact ? norm_exact(q) : norm_fast(q);
}

template <bool STATS>
__device__ __forceinline__ void flush_counters(const Counters& k, unsigned long long* stats) {
    if (!STATS || stats == nullptr) return;
    unsigned long long traced = k.traced, hits = k.hits, deepest = k.deepest, tests = k.tests;
    for (int off = 16; off > 0; off >>= 1) {
        traced += __shfl_down_sync(0xffffffffu, traced, off);
        hits += __shfl_down_sync(0xffffffffu, hits, off);
        tests += __shfl_down_sync(0xffffffffu, tests, off);
        unsigned long long other = __shfl_down_sync(0xffffffffu, deepest, off);
        deepest = other > deepest ? other : deepest;
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(stats + 0, traced);
        atomicAdd(stats + 1, hits);
        atomicMax(stats + 2, deepest);
        atomicAdd(stats + 3, tests);
    }
}

__device__ __forceinline__ unsigned int quantize_channel(double v) {
    // base.py:143-151: uint8(255 * clip(v, 0, 1)), truncating
    return (unsigned int)(255.0 * clip01(v));
}

__device__ __forceinline__ void store_pixel(const OutputParams& out, long long p, Vec3 c) {
    if (out.rgb != nullptr) {
        if (out.rgb_is_f64) {
            double* f = static_cast<double*>(out.rgb);
            f[p] = c.x;
            f[out.plane_stride + p] = c.y;
            f[2 * out.plane_stride + p] = c.z;
        } else {
            float* f = static_cast<float*>(out.rgb);
            f[p] = (float)c.x;
            f[out.plane_stride + p] = (float)c.y;
            f[2 * out.plane_stride + p] = (float)c.z;
        }
    }
    if (out.rgba8 != nullptr) {
        uchar4 q = make_uchar4((unsigned char)quantize_channel(c.x), (unsigned char)quantize_channel(c.y),
                               (unsigned char)quantize_channel(c.z), 255);
        reinterpret_cast<uchar4*>(out.rgba8)[p] = q;
    }
}

// Camera rows [row0,row1): ray generation + trace + stores, one thread per pixel.
template <int MODE, bool STATS>
__global__ void __launch_bounds__(kBlockThreads) render_rows_kernel(SceneParams sp, CameraParams cp,
                                                                   OutputParams out) {
    extern __shared__ __align__(16) unsigned char smem[];
    const Staged st = stage_scene<MODE>(sp, smem);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int col = blockIdx.x * kTileW + warp * kWarpTileW + (lane & (kWarpTileW - 1));
    const int row = cp.row0 + blockIdx.y * kTileH + (lane >> 3);
    const bool inside = col < cp.width && row < cp.row1;

    Counters k = {0, 0, 0, 0};
    if (inside) {
        Vec3 o = {cp.cam[0], cp.cam[1], cp.cam[2]};
        Vec3 d = primary_direction<MODE>(cp, col, row);
        Vec3 c = trace_chain<MODE, STATS>(sp, st, o, d, k);
        store_pixel(out, (long long)(row - cp.row0) * cp.width + col, c);
    }
    flush_counters<STATS>(k, out.stats);
}

// Explicit ray batch, one thread per ray.
template <int MODE, bool STATS>
__global__ void __launch_bounds__(kBlockThreads) trace_rays_kernel(SceneParams sp, RayBatch rb,
                                                                  OutputParams out) {
    extern __shared__ __align__(16) unsigned char smem[];
    const Staged st = stage_scene<MODE>(sp, smem);
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    Counters k = {0, 0, 0, 0};
    if (i < rb.n) {
        Vec3 o = {rb.origin0[0], rb.origin0[1], rb.origin0[2]};
        if (rb.origins != nullptr) o = {rb.origins[i], rb.origins[rb.n + i], rb.origins[2 * rb.n + i]};
        Vec3 d = {rb.directions[i], rb.directions[rb.n + i], rb.directions[2 * rb.n + i]};
        Vec3 c = trace_chain<MODE, STATS>(sp, st, o, d, k);
        store_pixel(out, i, c);
    }
    flush_counters<STATS>(k, out.stats);
}

// NumpySphere.intersect (shape.py:28-51) for a ray batch against one sphere.
template <int MODE>
__global__ void sphere_distance_kernel(RayBatch rb, SphereGeom g, double cc, double* out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rb.n) return;
    Vec3 o = {rb.origin0[0], rb.origin0[1], rb.origin0[2]};
    if (rb.origins != nullptr) o = {rb.origins[i], rb.origins[rb.n + i], rb.origins[2 * rb.n + i]};
    Vec3 d = {rb.directions[i], rb.directions[rb.n + i], rb.directions[2 * rb.n + i]};
    if (MODE == kModeExact) {
        out[i] = sphere_distance_exact(o, d, g, cc);
    } else {
        HalfB q = half_b(o, d, g);
        out[i] = can_hit(q) ? near_root(q) : kMiss;
    }
}

// get_ray_directions (base.py:123-141) materialised as three planes.
template <int MODE>
__global__ void ray_directions_kernel(CameraParams cp, double* out, long long plane_stride) {
    const long long n = (long long)(cp.row1 - cp.row0) * cp.width;
    const long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const int row = cp.row0 + (int)(p / cp.width), col = (int)(p % cp.width);
    Vec3 d = primary_direction<MODE>(cp, col, row);
    out[p] = d.x;
    out[plane_stride + p] = d.y;
    out[2 * plane_stride + p] = d.z;
}

}  // namespace rt

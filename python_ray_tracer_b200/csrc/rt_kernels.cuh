// rt_kernels.cuh -- __global__ entry points, instantiated once per arithmetic policy by
// trace_exact.cu (-fmad=false) and trace_fast.cu.
#pragma once

#include "rt_chain.cuh"

namespace rt {

// Tuning knobs (overridable with -D for the variant sweeps of profiles/).
#ifndef RT_ACCEL_MIN_BLOCKS
#define RT_ACCEL_MIN_BLOCKS 6  // tree kernels: latency-bound, more resident warps beat fewer spills (profiles/r01/accel_sweep.txt)
#endif
#ifndef RT_MIN_BLOCKS
#define RT_MIN_BLOCKS 5  // resident 128-thread blocks per SM the register allocator must allow
#endif
#ifndef RT_DYNAMIC_PCT
#define RT_DYNAMIC_PCT 30     // long launches of cheap items: share of the items left to the ticket counter
#endif
#ifndef RT_SHORT_LAUNCH_ROUNDS
#define RT_SHORT_LAUNCH_ROUNDS 64  // launches with fewer items per warp are ticketed entirely
#endif
#ifndef RT_COSTLY_ITEM_SPHERES
#define RT_COSTLY_ITEM_SPHERES 8   // scenes with at least this many spheres are ticketed entirely
#endif

#ifndef RT_PATCH_W
#define RT_PATCH_W 8
#endif
// Each warp renders a patch of 32 pixels.  8x4 by default: bounce coherence, and every row
// segment it stores is one full 32-byte sector.  32x1 (RT_FLAG_WIDE_STORES) when the frame is
// written into another GPU's memory: a warp's store is then one 128-byte line per plane or
// 8-bit row, which is what NVLink moves efficiently (about +4 % tracing time at 8K).
constexpr int kWarpTileW = RT_PATCH_W;
constexpr int kWideTileW = 32;
// Two block shapes.  Small scenes: 128-thread blocks, RT_MIN_BLOCKS resident per SM, each
// with its own copy of the (small) sphere table.  Scenes whose table would not fit that
// many times in an SM's shared memory: one 640-thread block per SM sharing one copy, the
// same 20 warps per SM.
constexpr int kBlockThreads = 128;
constexpr int kBigBlockThreads = 640;
constexpr size_t kSmemPerSm = 227 * 1024;

__host__ __device__ inline size_t staged_bytes(int n_spheres, int n_lights = 1) {
    return (sizeof(SphereGeom) + sizeof(SpherePrimary) + sizeof(SphereAux) + sizeof(SphereMaterial)) * (size_t)n_spheres +
           sizeof(double) * 3 * (size_t)n_lights;
}

// Work distribution of the persistent kernels.  A work item is one 8x4 patch (or 32 rays).
// The first static_rounds * n_warps items are dealt round-robin (warp w takes items w,
// w + n_warps, ...: no atomics); the rest are drawn one at a time from a device counter
// (lane 0, one atomicAdd per ticket), so the launch ends balanced at the granularity of a
// single patch.  Every warp keeps its next ticket in flight while it works and draws
// exactly one ticket >= n_tickets before it leaves, so the draw that returns
// n_tickets + n_warps - 1 is the last of the launch and puts the counter back to zero for
// the next launch that uses it.
//
// How much is ticketed, and in which order, is chosen per launch from measurements on B200
// (profiles/r01/schedule_sweep.txt).  Patch times of the README frame run from 1 us (sky)
// to 28 us (15-bounce chains where the spheres meet the ground): a static deal of
// everything leaves a 15 % tail at 1080p, while ticketing everything costs ~1 G
// same-address atomics/s on an 8K frame (+9 %).  So: short launches and launches of costly
// items (many spheres) are ticketed entirely, after one static round that hides the burst
// of first draws, and taken from the bottom of the frame upwards -- the long chains sit low
// in the frame and the rows that come last are its top, where the reference's fixed camera
// (image plane z = 0, +y up, base.py:130-141) sees the least, so the tail is made of short
// patches; long launches of cheap items deal the first 70 % statically, top-down.  These
// are heuristics about time only, never about results.
struct Tickets {
    unsigned int* counter;
    unsigned int n_tickets;      // items handed out dynamically
    unsigned int n_warps;        // gridDim.x * (block threads / 32)
    unsigned int static_rounds;  // items [0, static_rounds * n_warps) are dealt statically
    unsigned int n_items;
    unsigned int reversed;       // 1: take items from the last one backwards; 2: alternately from both ends (wide-store kernels)
};
// A draw is split in two: issue_ticket() starts the atomic, take_ticket() broadcasts its result
// after the item.  The atomic is atom.inc with the launch's last draw as its bound (every warp
// ends on one draw past the tickets, n_tickets + n_warps draws in all): the counter re-arms
// itself for the next launch, and -- unlike atom.add under a lane == 0 predicate -- ptxas does
// not wrap it in its warp-aggregation sequence (vote, popc, shuffle of the result).
// What this does NOT achieve is hiding the round trip (~0.8 us): at 96 registers the result is
// spilled on the next instruction, and the spill store waits for it.  Measured consequences
// (profiles/r01/schedule_sweep.txt): a 1080p frame of trivial items takes 68 us (22 draws per
// warp), handing out 8/4/2/1 items per draw lowers that floor to 51 us but costs the README
// frame 6 us in balance (its costly patches cluster), so single tickets stay.
__device__ __forceinline__ unsigned int issue_ticket(const Tickets& tk, int lane) {
    unsigned int t = 0;
    if (lane == 0)
        asm volatile("atom.global.inc.u32 %0, [%1], %2;" : "=r"(t) : "l"(tk.counter), "r"(tk.n_tickets + tk.n_warps - 1u));
    return t;
}
__device__ __forceinline__ unsigned int take_ticket(const Tickets& tk, unsigned int t, int lane) {
    return __shfl_sync(0xffffffffu, t, 0);
}

// The item loop of a warp: the static rounds, then tickets, the next one kept in flight under
// the item.  (Drawing only when the ticket is needed -- no draw live across the item body -- was
// measured for the big blocks, whose items take >= 100 us: 17.99 against 18.07 ms on c3, not kept.)
template <typename Item>
__device__ __forceinline__ void run_items(const Tickets& tk, const int lane, const unsigned int warp_id, Item&& item) {
    unsigned int drawn = issue_ticket(tk, lane);  // in flight during the static rounds
#pragma unroll 1
    for (unsigned int r = 0; r < tk.static_rounds; ++r) item(r * tk.n_warps + warp_id);
    const unsigned int first_dynamic = tk.static_rounds * tk.n_warps;
    unsigned int next = take_ticket(tk, drawn, lane);
#pragma unroll 1
    while (next < tk.n_tickets) {
        drawn = issue_ticket(tk, lane);  // in flight during this item
        item(first_dynamic + next);
        next = take_ticket(tk, drawn, lane);
    }
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
    // base.py:143-151: uint8(255 * clip(v, 0, 1)), truncating.  The clip is two tests of the
    // high word (FP64 fmin / fmax cost ~8 instructions each on sm_100a): same bytes for every
    // v that is not NaN.
    const int hi = __double2hiint(v);
    if (hi < 0) return 0u;               // v < 0 and -0.0
    if (hi >= 0x3ff00000) return 255u;   // v >= 1
    return (unsigned int)(255.0 * v);
}

// 3-byte pixels of one patch row, packed into 32-bit words by the warp (Bytes8::kRgb8Words):
// lane px of a row segment of PW pixels stores word px of its 3*PW/4 words, taking the bytes of
// pixels 4*px/3 and 4*px/3 + 1 from their lanes.  The segment starts at a multiple of four
// pixels and the frame width is one too, so every word is aligned and no word straddles two
// rows or two patches.  Called by all 32 lanes; `valid` = this lane owns a pixel of the frame.
template <int PW>
__device__ __forceinline__ void store_rgb8_words(const OutputParams& out, bool valid, long long p, int px, int cols_left,
                                                 Vec3 c) {
    unsigned int q = 0u;
    if (valid) q = quantize_channel(c.x) | (quantize_channel(c.y) << 8) | (quantize_channel(c.z) << 16);
    const int lane = threadIdx.x & 31;
    const int first = (4 * px) / 3, shift = 8 * ((4 * px) % 3);
    const int seg = lane - px;  // lane of the segment's first pixel
    const unsigned int qa = __shfl_sync(0xffffffffu, q, seg + (first < PW ? first : PW - 1));
    const unsigned int qb = __shfl_sync(0xffffffffu, q, seg + (first + 1 < PW ? first + 1 : PW - 1));
    const int n_valid = cols_left < PW ? cols_left : PW;
    if (valid && 4 * px < 3 * n_valid) {
        const unsigned long long both = (unsigned long long)qa | ((unsigned long long)qb << 24);
        *reinterpret_cast<unsigned int*>(out.rgba8 + (p - px) * 3 + 4 * px) = (unsigned int)(both >> shift);
    }
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
        if (out.bytes8 == kRgba8) {
            uchar4 q = make_uchar4((unsigned char)quantize_channel(c.x), (unsigned char)quantize_channel(c.y),
                                   (unsigned char)quantize_channel(c.z), 255);
            reinterpret_cast<uchar4*>(out.rgba8)[p] = q;
        } else if (out.bytes8 == kRgb8Bytes) {
            out.rgba8[3 * p + 0] = (unsigned char)quantize_channel(c.x);
            out.rgba8[3 * p + 1] = (unsigned char)quantize_channel(c.y);
            out.rgba8[3 * p + 2] = (unsigned char)quantize_channel(c.z);
        }  // kRgb8Words: stored by the warp, store_rgb8_words()
    }
}

// Camera rows of one band or interleaved shard (CameraParams): ray generation + trace + stores, one thread per pixel of a
// patch.  Persistent: the grid is sized to the resident block slots of the GPU.
template <int MODE, bool STATS, int BLOCK, bool ACCEL, int PW>
__global__ void __launch_bounds__(BLOCK, BLOCK == kBlockThreads ? (ACCEL ? RT_ACCEL_MIN_BLOCKS : RT_MIN_BLOCKS) : 1)
    render_rows_kernel(SceneParams sp, CameraParams cp, OutputParams out, Tickets tk) {
    constexpr int kWarpTileW = PW, kWarpTileH = 32 / PW;
    extern __shared__ __align__(16) unsigned char smem[];
    AccelView av = {};
    if (ACCEL) av = accel_view(sp.accel);
    const Staged st = ACCEL ? stage_lights_only<MODE>(sp, smem, av) : stage_scene<MODE>(sp, smem, cp.cam);

    const int lane = threadIdx.x & 31;
    const int px = lane & (kWarpTileW - 1), py = lane / kWarpTileW;
    const unsigned int tiles_x = (unsigned int)(cp.width + kWarpTileW - 1) / kWarpTileW;
    const unsigned int tiles_x_recip = tiles_x > 1u ? (unsigned int)(0x100000000ull / tiles_x) : 0xffffffffu;
    const unsigned int warp_id = blockIdx.x * (BLOCK / 32) + (threadIdx.x >> 5);
    const Vec3 eye = {cp.cam[0], cp.cam[1], cp.cam[2]};
    Counters k = {0, 0, 0, 0};

    auto render_patch = [&](unsigned int item) {
#ifdef RT_TIMING_DEBUG
        unsigned long long t_item0 = 0;
        if (STATS) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_item0));
#endif
        // item / tiles_x by a reciprocal formed once per thread: floor(2^32 / d) underestimates
        // the quotient of a 31-bit item by at most one
        unsigned int ty = __umulhi(item, tiles_x_recip);
        unsigned int tx = item - ty * tiles_x;
        if (tx >= tiles_x) {
            tx -= tiles_x;
            ty += 1u;
        }
        const int local_row = (int)ty * kWarpTileH + py;
        const int col = (int)tx * kWarpTileW + px;
        const bool valid = col < cp.width && local_row < cp.n_rows;
        if constexpr (PW == kWideTileW) {
            // the wide-store kernels can pack 3-byte pixels across the warp: all lanes reach the store
            long long p = 0;
            Vec3 c = {0.0, 0.0, 0.0};
            if (valid) {
                const int row = image_row(cp, local_row);
                Vec3 d = primary_direction<MODE>(cp, col, row);
                c = trace_chain<MODE, STATS, !ACCEL, ACCEL>(sp, st, av, eye, d, k);
                p = (long long)(out.image_order ? row : local_row) * cp.width + col;
                store_pixel(out, p, c);
            }
            if (out.bytes8 == kRgb8Words && out.rgba8 != nullptr) store_rgb8_words<PW>(out, valid, p, px, cp.width - col + px, c);
        } else if (valid) {
            const int row = image_row(cp, local_row);
            Vec3 d = primary_direction<MODE>(cp, col, row);
            Vec3 c = trace_chain<MODE, STATS, !ACCEL, ACCEL>(sp, st, av, eye, d, k);
            store_pixel(out, (long long)(out.image_order ? row : local_row) * cp.width + col, c);
        }
#ifdef RT_TIMING_DEBUG
        if (STATS && out.stats != nullptr) {
            __syncwarp();
            unsigned long long t_item1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_item1));
            if (lane == 0) {
                out.stats[8 + 2 * 148 * 32 + 2 * item] = t_item0;
                out.stats[8 + 2 * 148 * 32 + 2 * item + 1] = t_item1;
            }
        }
#endif
    };

#ifdef RT_TIMING_DEBUG
    // experiment only (tools/warp_timeline.py): per-warp start/end stamps behind the counters
    unsigned long long t_start = 0;
    if (STATS) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_start));
#endif
    auto render_item = [&](unsigned int idx) {
        unsigned int item = tk.reversed ? tk.n_items - 1u - idx : idx;
        if constexpr (PW == kWideTileW) {
            // frames that several GPUs write (RT_FLAG_EVEN_STORES): two patches from the bottom, then one
            // from the top, so that cheap rows (many stores per microsecond) and costly ones mix over the
            // launch; the two fronts meet a third of the way down, where the reference camera still sees
            // sky, so the launch still ends on cheap patches
            if (tk.reversed == 2u) {
                const unsigned int third = tk.n_items / 3u, group = idx / 3u, r = idx - 3u * group;
                if (group >= third) item = tk.n_items - 1u - (idx - third);  // the last n_items % 3
                else item = r == 2u ? group : tk.n_items - 1u - (2u * group + r);
            }
        }
        render_patch(item);
    };
    run_items(tk, lane, warp_id, render_item);
#ifdef RT_TIMING_DEBUG
    if (STATS && out.stats != nullptr && lane == 0) {
        unsigned long long t_end;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_end));
        out.stats[8 + 2 * warp_id] = t_start;
        out.stats[8 + 2 * warp_id + 1] = t_end;
    }
#endif
    flush_counters<STATS>(k, out.stats);
}

// Explicit ray batch, one thread per ray, same work distribution (an item = 32 rays).
template <int MODE, bool STATS, int BLOCK, bool ACCEL>
__global__ void __launch_bounds__(BLOCK, BLOCK == kBlockThreads ? (ACCEL ? RT_ACCEL_MIN_BLOCKS : RT_MIN_BLOCKS) : 1)
    trace_rays_kernel(SceneParams sp, RayBatch rb, OutputParams out, Tickets tk) {
    extern __shared__ __align__(16) unsigned char smem[];
    AccelView av = {};
    if (ACCEL) av = accel_view(sp.accel);
    const Staged st = ACCEL ? stage_lights_only<MODE>(sp, smem, av) : stage_scene<MODE>(sp, smem, sp.eye);
    const int lane = threadIdx.x & 31;
    const unsigned int warp_id = blockIdx.x * (BLOCK / 32) + (threadIdx.x >> 5);
    Counters k = {0, 0, 0, 0};

    auto trace_item = [&](unsigned int item) {
        const long long i = (long long)item * 32 + lane;
        if (i < rb.n) {
            Vec3 o = {rb.origin0[0], rb.origin0[1], rb.origin0[2]};
            if (rb.origins != nullptr) o = {rb.origins[i], rb.origins[rb.n + i], rb.origins[2 * rb.n + i]};
            Vec3 d = {rb.directions[i], rb.directions[rb.n + i], rb.directions[2 * rb.n + i]};
            Vec3 c = trace_chain<MODE, STATS, false, ACCEL>(sp, st, av, o, d, k);
            store_pixel(out, i, c);
        }
    };

    run_items(tk, lane, warp_id, trace_item);
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
    const long long n = (long long)cp.n_rows * cp.width;
    const long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const int row = image_row(cp, (int)(p / cp.width)), col = (int)(p % cp.width);
    Vec3 d = primary_direction<MODE>(cp, col, row);
    out[p] = d.x;
    out[plane_stride + p] = d.y;
    out[2 * plane_stride + p] = d.z;
}

}  // namespace rt

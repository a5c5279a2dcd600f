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
#ifndef RT_SMALL_MIN_BLOCKS
#define RT_SMALL_MIN_BLOCKS 5  // the same for the constant-bank kernels of small scenes
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
// Tickets per draw (profiles/r02/ticket_chunks.txt): fours lose (four neighbouring patches where the
// spheres meet the ground outlast the frame), pairs while more than 6 rounds are left win 6 % at 1080p
// and cost nothing on short launches, which stay on single tickets.
#ifndef RT_CHUNK4_ROUNDS
#define RT_CHUNK4_ROUNDS 0   // a draw takes 4 tickets while more than this many per warp are left (0: never) ...
#endif
#ifndef RT_CHUNK2_ROUNDS
#define RT_CHUNK2_ROUNDS 6   // ... 2 while more than this many are left, then 1 (0 0: always single tickets)
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

// Work distribution of the persistent kernels.  A work item is one 8x4 patch (or 32 rays).
// The first static_rounds * n_warps items are dealt round-robin (warp w takes items w,
// w + n_warps, ...: no atomics); the rest are tickets drawn from a device counter by lane 0, so the
// launch ends balanced at the granularity of a single patch.  Every warp keeps its next draw in
// flight while it works.
//
// How much is ticketed, and in which order, is chosen per launch from measurements on B200
// (profiles/r01/schedule_sweep.txt).  Patch times of the README frame run from 1 us (sky)
// to 28 us (15-bounce chains where the spheres meet the ground): a static deal of
// everything leaves a 15 % tail at 1080p.  So: short launches and launches of costly
// items (many spheres) are ticketed entirely, after one static round that hides the burst
// of first draws, and taken from the bottom of the frame upwards -- the long chains sit low
// in the frame and the rows that come last are its top, where the reference's fixed camera
// (image plane z = 0, +y up, base.py:130-141) sees the least, so the tail is made of short
// patches; long launches of cheap items deal the first 70 % statically, top-down.  These
// are heuristics about time only, never about results.
//
// Draws on ONE address are served at about one per nanosecond by the L2 slice that owns it: a
// 1080p frame of trivial patches took 68 us in round 1 (64 800 single-ticket draws), which is the
// whole budget of a frame now.  So a draw may take several consecutive tickets while the launch is
// young and single ones at the end (RT_CHUNK4_ROUNDS / RT_CHUNK2_ROUNDS); with the bottom-up order
// the single tickets are the cheap sky patches, so the launch still ends level.
struct Tickets {
    unsigned int* counter;       // draws so far; zero between launches (the last draw of a launch wraps it)
    unsigned int n_tickets;      // items handed out dynamically
    unsigned int n_warps;        // gridDim.x * (block threads / 32)
    unsigned int static_rounds;  // items [0, static_rounds * n_warps) are dealt statically
    unsigned int n_items;
    unsigned int reversed;       // 1: take items from the last one backwards; 2: alternately from both ends (wide-store kernels)
    // draw k takes tickets [4k, 4k+4) for k < n4, then pairs for the next n2 draws, then single tickets
    unsigned int n4, n2, cover4, cover2, n_draws;
};
// A draw is split in two: issue_draw() starts the atomic, take_draw() broadcasts its result when
// the warp has run out of items.  The atomic is atom.inc with the launch's last draw as its bound
// (every warp ends on one draw past the last chunk, n_draws + n_warps draws in all): the counter
// re-arms itself for the next launch, and -- unlike atom.add under a lane == 0 predicate -- ptxas
// does not wrap it in its warp-aggregation sequence (vote, prefix shuffles, a second atomic).  Which
// tickets a draw takes is a function of its index, so chunks need no atom.add.
__device__ __forceinline__ unsigned int issue_draw(const Tickets& tk, int lane) {
    unsigned int t = 0;
    if (lane == 0)
        asm volatile("atom.global.inc.u32 %0, [%1], %2;" : "=r"(t) : "l"(tk.counter), "r"(tk.n_draws + tk.n_warps - 1u));
    return t;
}
__device__ __forceinline__ unsigned int take_draw(unsigned int t) { return __shfl_sync(0xffffffffu, t, 0); }

// The item loop of a warp: the static rounds, then chunks of tickets, the next draw kept in flight
// under the items of the current one.  One copy of the item body (it is most of the kernel's code,
// and the instruction cache is cold at every launch of the bench: L2 is flushed between frames).
template <typename Item>
__device__ __forceinline__ void run_items(const Tickets& tk, const int lane, const unsigned int warp_id, Item&& item) {
    unsigned int drawn = issue_draw(tk, lane);  // in flight during the static rounds
    const unsigned int first_dynamic = tk.static_rounds * tk.n_warps;
    unsigned int round = 0, pos = 0, end = 0;
#pragma unroll 1
    for (;;) {
        unsigned int idx;
        if (round < tk.static_rounds) {
            idx = round * tk.n_warps + warp_id;
            ++round;
        } else {
            if (pos == end) {
                const unsigned int k = take_draw(drawn);
                if (k >= tk.n_draws) break;
                if (k < tk.n4) {
                    pos = 4u * k;
                    end = pos + 4u < tk.cover4 ? pos + 4u : tk.cover4;
                } else if (k < tk.n4 + tk.n2) {
                    pos = tk.cover4 + 2u * (k - tk.n4);
                    end = pos + 2u < tk.cover4 + tk.cover2 ? pos + 2u : tk.cover4 + tk.cover2;
                } else {
                    pos = tk.cover4 + tk.cover2 + (k - tk.n4 - tk.n2);
                    end = pos + 1u;
                }
                drawn = issue_draw(tk, lane);
            }
            idx = first_dynamic + pos++;
        }
        item(idx);
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
// The fast policy's pixels are FP32: the product rounded towards zero truncates to the same byte as
// the exact product 255 * v (which FP64 holds exactly: 8 x 24 bits), so the fused 8-bit image equals
// rt_quantize_dev of the stored planes, bit for bit.
__device__ __forceinline__ unsigned int quantize_channel(float v) {
    // saturate is a modifier of the multiply on sm_100a: two instructions, no branch
    return __float2uint_rz(__fmul_rz(255.0f, __saturatef(v)));
}

template <typename T>
struct Rgb {
    T x, y, z;
};

// 3-byte pixels of one patch row, packed into 32-bit words by the warp (Bytes8::kRgb8Words):
// lane px of a row segment of PW pixels stores word px of its 3*PW/4 words, taking the bytes of
// pixels 4*px/3 and 4*px/3 + 1 from their lanes.  The segment starts at a multiple of four
// pixels and the frame width is one too, so every word is aligned and no word straddles two
// rows or two patches.  Called by all 32 lanes; `valid` = this lane owns a pixel of the frame.
template <int PW, typename T>
__device__ __forceinline__ void store_rgb8_words(const OutputParams& out, bool valid, long long p, int px, int cols_left,
                                                 Rgb<T> c) {
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

template <typename T>
__device__ __forceinline__ void store_pixel(const OutputParams& out, long long p, Rgb<T> c) {
    if (out.rgb != nullptr) {
        if (out.rgb_is_f64) {
            double* f = static_cast<double*>(out.rgb);
            f[p] = (double)c.x;
            f[out.plane_stride + p] = (double)c.y;
            f[2 * out.plane_stride + p] = (double)c.z;
        } else {
            float* f = static_cast<float*>(out.rgb);
            f[p] = (float)c.x;
            f[out.plane_stride + p] = (float)c.y;
            f[2 * out.plane_stride + p] = (float)c.z;
        }
    }
    if (out.rgba8 != nullptr) {
        if (out.bytes8 == kRgba8) {
            reinterpret_cast<unsigned int*>(out.rgba8)[p] =
                quantize_channel(c.x) | (quantize_channel(c.y) << 8) | (quantize_channel(c.z) << 16) | 0xff000000u;
        } else if (out.bytes8 == kRgb8Bytes) {
            out.rgba8[3 * p + 0] = (unsigned char)quantize_channel(c.x);
            out.rgba8[3 * p + 1] = (unsigned char)quantize_channel(c.y);
            out.rgba8[3 * p + 2] = (unsigned char)quantize_channel(c.z);
        }  // kRgb8Words: stored by the warp, store_rgb8_words()
    }
}

// ---- the two tracers the kernels are written over ----------------------------------------------------
// camera(col, row) / ray(o, d) -> colour of one pixel / ray, counters in k.
struct ExactTracer {
    using Pixel = Rgb<double>;
    const SceneParams& sp;
    Staged st;
    AccelView av;
    template <bool STATS, bool TREE>
    __device__ __forceinline__ Pixel ray(Vec3 o, Vec3 d, Counters& k) const {
        const Vec3 c = trace_chain_exact<STATS, TREE>(sp, st, av, o, d, k);
        return {c.x, c.y, c.z};
    }
};

// Camera rows of one band or interleaved shard (CameraParams): ray generation + trace + stores, one
// thread per pixel of a patch.  Persistent: the grid is sized to the resident block slots of the GPU.
// `pixel(col, row, k)` is the tracer of the instantiation.
template <bool STATS, int BLOCK, int PW, typename PixelFn>
__device__ __forceinline__ void render_rows_body(const CameraParams& cp, const OutputParams& out, const Tickets& tk, PixelFn&& pixel) {
    constexpr int kTileW = PW, kTileH = 32 / PW;
    const int lane = threadIdx.x & 31;
    const int px = lane & (kTileW - 1), py = lane / kTileW;
    const unsigned int tiles_x = (unsigned int)(cp.width + kTileW - 1) / kTileW;
    const unsigned int tiles_x_recip = tiles_x > 1u ? (unsigned int)(0x100000000ull / tiles_x) : 0xffffffffu;
    const unsigned int warp_id = blockIdx.x * (BLOCK / 32) + (threadIdx.x >> 5);
    Counters k = {0, 0, 0, 0};

#ifdef RT_TIMING_DEBUG
    // experiment only (tools/warp_timeline.py): per-warp and per-item start/end stamps behind the counters
    unsigned long long t_start = 0;
    if (STATS) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_start));
#endif
    auto render_item = [&](unsigned int idx) {
        unsigned int item = tk.reversed ? tk.n_items - 1u - idx : idx;
#ifdef RT_TIMING_DEBUG
        unsigned long long t_item0 = 0;
        if (STATS) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_item0));
#endif
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
        // item / tiles_x by a reciprocal formed once per thread: floor(2^32 / d) underestimates
        // the quotient of a 31-bit item by at most one
        unsigned int ty = __umulhi(item, tiles_x_recip);
        unsigned int tx = item - ty * tiles_x;
        if (tx >= tiles_x) {
            tx -= tiles_x;
            ty += 1u;
        }
        const int local_row = (int)ty * kTileH + py;
        const int col = (int)tx * kTileW + px;
        const bool valid = col < cp.width && local_row < cp.n_rows;
        if constexpr (PW == kWideTileW) {
            // the wide-store kernels can pack 3-byte pixels across the warp: all lanes reach the store
            long long p = 0;
            decltype(pixel(0, 0, k)) c = {0, 0, 0};
            if (valid) {
                const int row = image_row(cp, local_row);
                c = pixel(col, row, k);
                p = (long long)(out.image_order ? row : local_row) * cp.width + col;
                store_pixel(out, p, c);
            }
            if (out.bytes8 == kRgb8Words && out.rgba8 != nullptr) store_rgb8_words<PW>(out, valid, p, px, cp.width - col + px, c);
        } else if (valid) {
            const int row = image_row(cp, local_row);
            const auto c = pixel(col, row, k);
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

// camera point of a pixel on the plane z = 0 minus the eye: the UNNORMALISED primary direction
__device__ __forceinline__ Vec3 primary_offset(const CameraParams& cp, int col, int row) {
    const double x = linspace_at(cp.x_start, cp.x_stop, cp.x_step, cp.width, col);
    const double y = linspace_at(cp.y_start, cp.y_stop, cp.y_step, cp.height, row);
    return {x - cp.cam[0], y - cp.cam[1], 0.0 - cp.cam[2]};
}

template <class Scene, bool STATS>
__device__ __forceinline__ Rgb<float> fast_camera_pixel(const Scene& sc, const SceneParams& sp, const CameraParams& cp, int col, int row,
                                                        Counters& k) {
    const Vec3 q = primary_offset(cp, col, row);
    Rgb<float> c;
    trace_chain_fast<Scene, STATS, true>(sc, sp, {cp.cam[0], cp.cam[1], cp.cam[2]}, q, fdot(q, q), c.x, c.y, c.z, k);
    return c;
}

// KIND: 0 = sphere table staged in shared memory, 1 = BVH blob
template <int MODE, bool STATS, int BLOCK, int KIND, int PW>
__global__ void __launch_bounds__(BLOCK, BLOCK == kBlockThreads ? (KIND == 1 ? RT_ACCEL_MIN_BLOCKS : RT_MIN_BLOCKS) : 1)
    render_rows_kernel(const __grid_constant__ SceneParams sp, const __grid_constant__ CameraParams cp,
                       const __grid_constant__ OutputParams out, const __grid_constant__ Tickets tk) {
    extern __shared__ __align__(16) unsigned char smem[];
    if (KIND == 1 && !accel_blob_ok(sp)) return;
    if constexpr (MODE == kModeExact) {
        const ExactTracer tr = {sp, stage_exact(sp, smem), KIND == 1 ? accel_view(sp) : AccelView{}};
        const Vec3 eye = {cp.cam[0], cp.cam[1], cp.cam[2]};
        render_rows_body<STATS, BLOCK, PW>(cp, out, tk, [&](int col, int row, Counters& k) {
            return tr.template ray<STATS, KIND == 1>(eye, primary_direction<kModeExact>(cp, col, row), k);
        });
    } else if constexpr (KIND == 1) {
        const TreeScene sc = stage_tree(sp, smem);
        render_rows_body<STATS, BLOCK, PW>(cp, out, tk, [&](int col, int row, Counters& k) {
            return fast_camera_pixel<TreeScene, STATS>(sc, sp, cp, col, row, k);
        });
    } else {
        const StagedScene sc = stage_fast(sp, smem, cp.cam);
        render_rows_body<STATS, BLOCK, PW>(cp, out, tk, [&](int col, int row, Counters& k) {
            return fast_camera_pixel<StagedScene, STATS>(sc, sp, cp, col, row, k);
        });
    }
}

// Scenes of N <= kSmallMax spheres, fast policy: the table is a kernel parameter.
template <bool STATS, int N, int PW>
__global__ void __launch_bounds__(kBlockThreads, RT_SMALL_MIN_BLOCKS)
    render_rows_small_kernel(const __grid_constant__ SceneParams sp, const __grid_constant__ CameraParams cp,
                             const __grid_constant__ OutputParams out, const __grid_constant__ Tickets tk,
                             const __grid_constant__ SmallTable table) {
    __shared__ __align__(16) unsigned char smem[N * (sizeof(ShadeRec) + sizeof(SphereGeom))];
    const SmallScene<N> sc = stage_small<N>(sp, table, smem);
    render_rows_body<STATS, kBlockThreads, PW>(cp, out, tk, [&](int col, int row, Counters& k) {
        return fast_camera_pixel<SmallScene<N>, STATS>(sc, sp, cp, col, row, k);
    });
}

// Explicit ray batch, one thread per ray, same work distribution (an item = 32 rays).
template <bool STATS, int BLOCK, typename RayFn>
__device__ __forceinline__ void trace_rays_body(const RayBatch& rb, const OutputParams& out, const Tickets& tk, RayFn&& ray) {
    const int lane = threadIdx.x & 31;
    const unsigned int warp_id = blockIdx.x * (BLOCK / 32) + (threadIdx.x >> 5);
    Counters k = {0, 0, 0, 0};
    auto trace_item = [&](unsigned int item) {
        const long long i = (long long)item * 32 + lane;
        if (i < rb.n) {
            Vec3 o = {rb.origin0[0], rb.origin0[1], rb.origin0[2]};
            if (rb.origins != nullptr) o = {rb.origins[i], rb.origins[rb.n + i], rb.origins[2 * rb.n + i]};
            const Vec3 d = {rb.directions[i], rb.directions[rb.n + i], rb.directions[2 * rb.n + i]};
            store_pixel(out, i, ray(o, d, k));
        }
    };
    run_items(tk, lane, warp_id, trace_item);
    flush_counters<STATS>(k, out.stats);
}

template <class Scene, bool STATS>
__device__ __forceinline__ Rgb<float> fast_ray(const Scene& sc, const SceneParams& sp, Vec3 o, Vec3 d, Counters& k) {
    Rgb<float> c;
    trace_chain_fast<Scene, STATS, false>(sc, sp, o, d, 1.0, c.x, c.y, c.z, k);
    return c;
}

template <int MODE, bool STATS, int BLOCK, int KIND>
__global__ void __launch_bounds__(BLOCK, BLOCK == kBlockThreads ? (KIND == 1 ? RT_ACCEL_MIN_BLOCKS : RT_MIN_BLOCKS) : 1)
    trace_rays_kernel(const __grid_constant__ SceneParams sp, const __grid_constant__ RayBatch rb,
                      const __grid_constant__ OutputParams out, const __grid_constant__ Tickets tk) {
    extern __shared__ __align__(16) unsigned char smem[];
    if (KIND == 1 && !accel_blob_ok(sp)) return;
    if constexpr (MODE == kModeExact) {
        const ExactTracer tr = {sp, stage_exact(sp, smem), KIND == 1 ? accel_view(sp) : AccelView{}};
        trace_rays_body<STATS, BLOCK>(rb, out, tk, [&](Vec3 o, Vec3 d, Counters& k) { return tr.template ray<STATS, KIND == 1>(o, d, k); });
    } else if constexpr (KIND == 1) {
        const TreeScene sc = stage_tree(sp, smem);
        trace_rays_body<STATS, BLOCK>(rb, out, tk, [&](Vec3 o, Vec3 d, Counters& k) { return fast_ray<TreeScene, STATS>(sc, sp, o, d, k); });
    } else {
        const StagedScene sc = stage_fast(sp, smem, sp.eye);
        trace_rays_body<STATS, BLOCK>(rb, out, tk, [&](Vec3 o, Vec3 d, Counters& k) { return fast_ray<StagedScene, STATS>(sc, sp, o, d, k); });
    }
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

// rt_abi_host.cu -- the host-buffer entry points of include/rt_abi.h: upload the scene,
// launch through the device entry points, download the result, synchronise.
//
// A per-thread cache keeps one device scratch area (scene tables, rays, outputs,
// counters) per device and a private stream, grown on demand and reused across calls;
// rt_release_cache() frees it.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../include/rt_abi.h"
#include "rt_launch.h"

namespace {

constexpr int kMaxDevices = 64;

struct Scratch {
    void* ptr = nullptr;
    size_t bytes = 0;
};

constexpr int kMaxChunks = 16;

struct DeviceCache {
    cudaStream_t stream = nullptr;       // uploads and kernels
    cudaStream_t copy_stream = nullptr;  // downloads, overlapped with the next chunk's kernel
    cudaEvent_t rendered[kMaxChunks] = {};
    Scratch scene, rays, rgb, rgb8, stats, accel, png;
    // pinned, device-mapped landing zone of rt_render_png_host (file + its size word) for callers whose own
    // buffer the device cannot write
    void* png_stage = nullptr;
    size_t png_stage_bytes = 0;
};

thread_local DeviceCache g_cache[kMaxDevices];
// defined in rt_abi.cu: records the message rt_last_error() returns
extern "C" void rt_internal_set_error(const char* message);
extern "C" int rt_internal_small_scene(const rt_env* env, int32_t n_spheres);

cudaError_t reserve(Scratch& s, size_t bytes) {
    if (bytes <= s.bytes) return cudaSuccess;
    if (s.ptr != nullptr) cudaFree(s.ptr);
    s.ptr = nullptr;
    s.bytes = 0;
    size_t want = bytes + bytes / 4 + 256;
    cudaError_t err = cudaMalloc(&s.ptr, want);
    if (err == cudaSuccess) s.bytes = want;
    return err;
}

void release(Scratch& s) {
    if (s.ptr != nullptr) cudaFree(s.ptr);
    s.ptr = nullptr;
    s.bytes = 0;
}

int host_fail(int status, const char* message) {
    rt_internal_set_error(message);
    return status;
}

struct Restore {
    int previous = -1;
    ~Restore() {
        if (previous >= 0) cudaSetDevice(previous);
    }
};

#define HOST_CUDA(call)                                                                      \
    do {                                                                                     \
        cudaError_t err__ = (call);                                                          \
        if (err__ != cudaSuccess) {                                                          \
            char msg__[256];                                                                 \
            snprintf(msg__, sizeof(msg__), "%s: %s", #call, cudaGetErrorString(err__));      \
            rt_internal_set_error(msg__);                                                    \
            return RT_ERR_CUDA;                                                              \
        }                                                                                    \
    } while (0)

// Uploads the sphere SoA and the light table into the cache; returns device pointers.
int upload_scene(DeviceCache& c, const double* spheres_host, int32_t n_spheres, const double* lights_host,
                 int32_t n_lights, const double** spheres_dev, const double** lights_dev) {
    const size_t sphere_bytes = sizeof(double) * RT_SPHERE_FIELDS * (size_t)n_spheres;
    const size_t light_bytes = sizeof(double) * 3 * (size_t)n_lights;
    HOST_CUDA(reserve(c.scene, sphere_bytes + light_bytes));
    char* base = static_cast<char*>(c.scene.ptr);
    HOST_CUDA(cudaMemcpyAsync(base, spheres_host, sphere_bytes, cudaMemcpyHostToDevice, c.stream));
    HOST_CUDA(cudaMemcpyAsync(base + sphere_bytes, lights_host, light_bytes, cudaMemcpyHostToDevice, c.stream));
    *spheres_dev = reinterpret_cast<const double*>(base);
    *lights_dev = reinterpret_cast<const double*>(base + sphere_bytes);
    return RT_OK;
}

// Scenes of RT_ACCEL_AUTO_SPHERES shapes or more get a bounding-volume tree unless the caller
// brought one (env->accel_dev) or opted out (RT_FLAG_NO_ACCEL): built here on the host, uploaded
// on the call's stream, named in `*use` (a copy of the caller's env).
int attach_accel(DeviceCache& c, const double* spheres_host, int32_t n_spheres, const rt_env* env, rt_env* use) {
    *use = *env;
    if (env->accel_dev != nullptr || (env->flags & RT_FLAG_NO_ACCEL) || n_spheres < RT_ACCEL_AUTO_SPHERES) return RT_OK;
    thread_local std::vector<unsigned char> blob;
    const int64_t capacity = rt_accel_bytes(n_spheres);
    if ((int64_t)blob.size() < capacity) blob.resize((size_t)capacity);
    int64_t used = 0;
    // built inside the call, used once: the quick median build (the call is bound by the
    // copy of the frame to the host; a better tree would not pay for its own construction)
    rt_env build_env = *env;
    build_env.flags |= RT_FLAG_ACCEL_FAST_BUILD;
    const int rc = rt_accel_build_host(spheres_host, n_spheres, &build_env, blob.data(), capacity, &used);
    if (rc != RT_OK) return host_fail(rc, "host entry point: building the sphere tree failed");
    HOST_CUDA(reserve(c.accel, (size_t)used));
    // pageable source: the copy is staged before the call returns, so `blob` may be reused
    HOST_CUDA(cudaMemcpyAsync(c.accel.ptr, blob.data(), (size_t)used, cudaMemcpyHostToDevice, c.stream));
    use->accel_dev = c.accel.ptr;
    use->accel_bytes = used;
    return RT_OK;
}

int enter(int device, Restore& restore, DeviceCache** cache) {
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) return host_fail(RT_ERR_NO_DEVICE, "host entry point: no CUDA device or device index out of range");
    if (device < 0 || device >= count || device >= kMaxDevices) return host_fail(RT_ERR_NO_DEVICE, "host entry point: no CUDA device or device index out of range");
    cudaGetDevice(&restore.previous);
    HOST_CUDA(cudaSetDevice(device));
    DeviceCache& c = g_cache[device];
    if (c.stream == nullptr) {
        HOST_CUDA(cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking));
        HOST_CUDA(cudaStreamCreateWithFlags(&c.copy_stream, cudaStreamNonBlocking));
        for (int i = 0; i < kMaxChunks; ++i) HOST_CUDA(cudaEventCreateWithFlags(&c.rendered[i], cudaEventDisableTiming));
    }
    *cache = &c;
    return RT_OK;
}

}  // namespace

extern "C" {

int rt_release_cache(void) {
    rt::release_ticket_counters();
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess) return RT_OK;
    Restore restore;
    cudaGetDevice(&restore.previous);
    for (int d = 0; d < count && d < kMaxDevices; ++d) {
        DeviceCache& c = g_cache[d];
        if (c.stream == nullptr && c.scene.ptr == nullptr) continue;
        cudaSetDevice(d);
        release(c.scene);
        release(c.rays);
        release(c.rgb);
        release(c.rgb8);
        release(c.stats);
        release(c.accel);
        release(c.png);
        if (c.png_stage != nullptr) cudaFreeHost(c.png_stage);
        c.png_stage = nullptr;
        c.png_stage_bytes = 0;
        if (c.stream != nullptr) {
            cudaStreamDestroy(c.stream);
            cudaStreamDestroy(c.copy_stream);
            for (int i = 0; i < kMaxChunks; ++i) cudaEventDestroy(c.rendered[i]);
        }
        c.stream = nullptr;
        c.copy_stream = nullptr;
    }
    return RT_OK;
}

int rt_render_frame_host(int device, const double* spheres_host, int32_t n_spheres, const double* point_lights_host,
                         const rt_env* env, const rt_camera* cam, int32_t row0, int32_t row1, void* rgb_host,
                         int32_t rgb_dtype, uint8_t* rgb8_host, uint64_t* stats_host) {
    if (spheres_host == nullptr || point_lights_host == nullptr || env == nullptr || cam == nullptr) return host_fail(RT_ERR_BAD_ARG, "host entry point: NULL pointer, bad size or unknown enum value");
    if (n_spheres <= 0 || env->n_point_lights < 1) return host_fail(RT_ERR_BAD_ARG, "host entry point: NULL pointer, bad size or unknown enum value");
    if (rgb_dtype != RT_DTYPE_F32 && rgb_dtype != RT_DTYPE_F64) return host_fail(RT_ERR_BAD_ARG, "host entry point: NULL pointer, bad size or unknown enum value");
    if (cam->width <= 0 || cam->height <= 0 || row0 < 0 || row1 > cam->height || row0 > row1) return host_fail(RT_ERR_BAD_ARG, "host entry point: NULL pointer, bad size or unknown enum value");
    Restore restore;
    DeviceCache* c = nullptr;
    int rc = enter(device, restore, &c);
    if (rc != RT_OK) return rc;

    rt_env with_tree;
    if ((rc = attach_accel(*c, spheres_host, n_spheres, env, &with_tree)) != RT_OK) return rc;
    env = &with_tree;
    // the tables are host memory here by definition: per-scene constants are formed on the host
    with_tree.spheres_host = spheres_host;
    with_tree.point_lights_host = point_lights_host;
    const double *spheres_dev, *lights_dev;
    if (rt_internal_small_scene(env, n_spheres)) {
        // the camera kernel of a small scene reads its table from the kernel parameters: nothing to upload
        HOST_CUDA(reserve(c->scene, 256));
        spheres_dev = lights_dev = static_cast<const double*>(c->scene.ptr);
    } else if ((rc = upload_scene(*c, spheres_host, n_spheres, point_lights_host, env->n_point_lights, &spheres_dev, &lights_dev)) != RT_OK) {
        return rc;
    }
    // the scratch buffers hold the band, chunk by chunk: a caller's whole-frame addressing flag
    // (meant for rt_render_rows_dev with its own frame buffer) must not reach the kernel
    with_tree.flags &= ~RT_FLAG_IMAGE_ORDER;
    const int64_t n = (int64_t)(row1 - row0) * cam->width;
    const size_t elem = rgb_dtype == RT_DTYPE_F64 ? 8 : 4;
    // the 8-bit image is quantised by the trace kernel itself, from the FP64 pixel (what
    // save_image does, base.py:143-151): no linear planes unless the caller wants them
    if (rgb_host != nullptr) HOST_CUDA(reserve(c->rgb, (size_t)n * 3 * elem + 16));
    if (rgb8_host != nullptr) {
        HOST_CUDA(reserve(c->rgb8, (size_t)n * 3 + 16));
        with_tree.flags |= RT_FLAG_RGB8;
    }
    if (stats_host != nullptr) {
        HOST_CUDA(reserve(c->stats, sizeof(uint64_t) * RT_STATS_WORDS));
        HOST_CUDA(cudaMemsetAsync(c->stats.ptr, 0, sizeof(uint64_t) * RT_STATS_WORDS, c->stream));
    }
    // Frames are rendered in row chunks: while chunk i is copied to the host (copy engine,
    // copy_stream) the kernel of chunk i+1 runs.  The copies are the longer leg (PCIe), so
    // the frame costs about one chunk's kernel plus the copy time.  Small frames: one chunk.
    const int rows = row1 - row0;
    int chunks = 1;
    if (n >= (int64_t)1 << 20) chunks = 8;  // (profiles/r02/host_path.txt: 1080p RGB8 0.164 ms with 8, 0.180 with 4, 0.186 with 1)
    int head_div = 32;
    if (const char* e = getenv("RT_HOST_CHUNKS")) {  // measurement knob (tools/host_chunks_sweep.sh): "chunks[,head divisor]"
        int c = 0, d = 0;
        if (sscanf(e, "%d,%d", &c, &d) >= 1 && c >= 1 && c <= kMaxChunks) chunks = c;
        if (d >= 2) head_div = d;
    }
    if (chunks > rows) chunks = rows > 0 ? rows : 1;
    // Nothing is copied while the first chunk renders, so it is a small one: with more than one
    // chunk the first takes 1/16 of the rows and the others share the rest evenly.
    auto chunk_start = [&](int i) -> int {
        if (i <= 0) return 0;
        if (i >= chunks) return rows;
        const int64_t head = rows / head_div;
        return (int)((head + ((int64_t)rows - head) * (i - 1) / (chunks - 1)) & ~3LL);  // multiples of 4 rows (the patch height)
    };
    char* rgb_base = static_cast<char*>(c->rgb.ptr);
    uint64_t* stats_dev = stats_host != nullptr ? static_cast<uint64_t*>(c->stats.ptr) : nullptr;
    // (a failure inside the loop must not return while earlier chunks are still being copied into the caller's
    // buffers: it leaves the loop, both streams are drained below, then the status is returned)
    cudaError_t failed = cudaSuccess;
    const char* failed_call = "";
#define LOOP_CUDA(call)                          \
    if ((failed = (call)) != cudaSuccess) {      \
        failed_call = #call;                     \
        break;                                   \
    }
    for (int i = 0; i < chunks && n > 0; ++i) {
        const int r0 = row0 + chunk_start(i), r1 = row0 + chunk_start(i + 1);
        if (r1 <= r0) continue;
        const int64_t first = (int64_t)(r0 - row0) * cam->width, count = (int64_t)(r1 - r0) * cam->width;
        rc = rt_render_rows_dev(device, spheres_dev, n_spheres, lights_dev, env, cam, r0, r1,
                                rgb_host != nullptr ? rgb_base + first * elem : nullptr, rgb_dtype, n,
                                rgb8_host != nullptr ? static_cast<uint8_t*>(c->rgb8.ptr) + first * 3 : nullptr, stats_dev,
                                c->stream);
        if (rc != RT_OK) break;
        LOOP_CUDA(cudaEventRecord(c->rendered[i], c->stream));
        LOOP_CUDA(cudaStreamWaitEvent(c->copy_stream, c->rendered[i], 0));
        if (rgb_host != nullptr) {
            // the chunk's rows of the three planes: one strided copy (three when a plane is
            // longer than a 2-D copy's pitch may be)
            const size_t pitch = (size_t)n * elem;
            if (pitch <= ((size_t)1 << 30)) {
                LOOP_CUDA(cudaMemcpy2DAsync(static_cast<char*>(rgb_host) + (size_t)first * elem, pitch,
                                            rgb_base + (size_t)first * elem, pitch, (size_t)count * elem, 3,
                                            cudaMemcpyDeviceToHost, c->copy_stream));
            } else {
                for (int ch = 0; ch < 3; ++ch)
                    LOOP_CUDA(cudaMemcpyAsync(static_cast<char*>(rgb_host) + ((size_t)ch * n + first) * elem,
                                              rgb_base + ((size_t)ch * n + first) * elem, (size_t)count * elem,
                                              cudaMemcpyDeviceToHost, c->copy_stream));
            }
        }
        if (rgb8_host != nullptr)
            LOOP_CUDA(cudaMemcpyAsync(rgb8_host + first * 3, static_cast<uint8_t*>(c->rgb8.ptr) + first * 3,
                                      (size_t)count * 3, cudaMemcpyDeviceToHost, c->copy_stream));
    }
#undef LOOP_CUDA
    if (failed != cudaSuccess || rc != RT_OK) {
        cudaStreamSynchronize(c->copy_stream);
        cudaStreamSynchronize(c->stream);
        if (failed != cudaSuccess) {
            char msg[256];
            snprintf(msg, sizeof(msg), "%.160s: %s", failed_call, cudaGetErrorString(failed));
            rt_internal_set_error(msg);
            return RT_ERR_CUDA;
        }
        return rc;
    }
    HOST_CUDA(cudaStreamSynchronize(c->copy_stream));
    if (stats_host != nullptr)
        HOST_CUDA(cudaMemcpyAsync(stats_host, c->stats.ptr, sizeof(uint64_t) * RT_STATS_WORDS, cudaMemcpyDeviceToHost, c->stream));
    HOST_CUDA(cudaStreamSynchronize(c->stream));
    return RT_OK;
}

// save_image's file without the 8-bit image ever leaving the device: trace (8-bit image only), encode the bands,
// pack the file -- into the caller's buffer directly if it is pinned memory the device can address, into a
// pinned staging buffer otherwise -- and wait once.
int rt_render_png_host(int device, const double* spheres_host, int32_t n_spheres, const double* point_lights_host,
                       const rt_env* env, const rt_camera* cam, uint8_t* png_host, int64_t capacity, int64_t* png_bytes_out) {
    if (spheres_host == nullptr || point_lights_host == nullptr || env == nullptr || cam == nullptr || png_host == nullptr ||
        png_bytes_out == nullptr || n_spheres <= 0 || env->n_point_lights < 1 || cam->width <= 0 || cam->height <= 0)
        return host_fail(RT_ERR_BAD_ARG, "rt_render_png_host: NULL pointer or bad size");
    const int64_t bound = rt_png_file_bound(cam->width, cam->height);
    if (bound < 0) return host_fail(RT_ERR_UNSUPPORTED, "rt_render_png_host: rows wider than 21 666 pixels");
    Restore restore;
    DeviceCache* c = nullptr;
    int rc = enter(device, restore, &c);
    if (rc != RT_OK) return rc;
    rt_env use;
    if ((rc = attach_accel(*c, spheres_host, n_spheres, env, &use)) != RT_OK) return rc;
    use.spheres_host = spheres_host;
    use.point_lights_host = point_lights_host;
    use.flags = (use.flags & ~RT_FLAG_IMAGE_ORDER) | RT_FLAG_RGB8;
    const double *spheres_dev, *lights_dev;
    if (rt_internal_small_scene(&use, n_spheres)) {
        HOST_CUDA(reserve(c->scene, 256));
        spheres_dev = lights_dev = static_cast<const double*>(c->scene.ptr);
    } else if ((rc = upload_scene(*c, spheres_host, n_spheres, point_lights_host, env->n_point_lights, &spheres_dev, &lights_dev)) != RT_OK) {
        return rc;
    }
    const int64_t n = (int64_t)cam->width * cam->height;
    const int64_t scratch_bytes = rt_png_scratch_bytes(cam->width, cam->height);
    HOST_CUDA(reserve(c->rgb8, (size_t)n * 3 + 16));
    HOST_CUDA(reserve(c->png, (size_t)scratch_bytes));
    // where the pack kernel writes: the caller's buffer if the device can address it (pinned host memory)
    uint8_t* out_dev = nullptr;
    unsigned long long* size_word = nullptr;
    cudaPointerAttributes attr;
    const bool direct = capacity >= 128 + 8 && cudaPointerGetAttributes(&attr, png_host) == cudaSuccess && attr.type == cudaMemoryTypeHost &&
                        attr.devicePointer != nullptr;
    cudaGetLastError();  // (an unregistered pointer is reported as an error by older runtimes: not ours)
    const size_t stage_need = (size_t)bound + 64;
    if (c->png_stage_bytes < (direct ? 64 : stage_need)) {
        if (c->png_stage != nullptr) cudaFreeHost(c->png_stage);
        c->png_stage = nullptr;
        c->png_stage_bytes = 0;
        HOST_CUDA(cudaHostAlloc(&c->png_stage, direct ? 64 : stage_need, cudaHostAllocMapped));
        c->png_stage_bytes = direct ? 64 : stage_need;
    }
    void* stage_dev = nullptr;
    HOST_CUDA(cudaHostGetDevicePointer(&stage_dev, c->png_stage, 0));
    size_word = static_cast<unsigned long long*>(stage_dev);  // first 8 bytes of the staging area: the size
    volatile unsigned long long* size_host = static_cast<volatile unsigned long long*>(c->png_stage);
    *size_host = ~0ull;
    int64_t out_capacity;
    if (direct) {
        out_dev = static_cast<uint8_t*>(attr.devicePointer);
        out_capacity = capacity;
    } else {
        out_dev = static_cast<uint8_t*>(stage_dev) + 64;
        out_capacity = (int64_t)c->png_stage_bytes - 64;
    }
    rc = rt_render_rows_dev(device, spheres_dev, n_spheres, lights_dev, &use, cam, 0, cam->height, nullptr, RT_DTYPE_F32, n,
                            static_cast<uint8_t*>(c->rgb8.ptr), nullptr, c->stream);
    if (rc != RT_OK) return rc;
    // One launch for frames of many bands (8K: 549 us against 589 for the file in pinned host memory -- its chunks cross
    // PCIe while later bands are encoded); two launches below that (1080p: 71 us against 76: the single launch pays a
    // memset and a per-block epilogue that the separate, very light pack kernel does not).
    const int32_t band_rows = rt_png_rows_per_band(cam->width, cam->height);
    if ((cam->height + band_rows - 1) / band_rows >= RT_PNG_ONE_LAUNCH_BANDS) {
        rc = rt_png_encode_file_dev(device, static_cast<const uint8_t*>(c->rgb8.ptr), 3, (int64_t)cam->width * 3, cam->width, cam->height,
                                    c->png.ptr, scratch_bytes, out_dev, out_capacity, reinterpret_cast<uint64_t*>(size_word), c->stream);
    } else {
        rc = rt_png_encode_bands_dev(device, static_cast<const uint8_t*>(c->rgb8.ptr), 3, (int64_t)cam->width * 3, cam->height, cam->width,
                                     cam->height, 0, cam->height, 1, c->png.ptr, scratch_bytes, c->stream);
        if (rc == RT_OK) rc = rt_png_pack_dev(device, cam->width, cam->height, c->png.ptr, out_dev, out_capacity, reinterpret_cast<uint64_t*>(size_word), c->stream);
    }
    HOST_CUDA(cudaStreamSynchronize(c->stream));
    if (rc != RT_OK) return rc;
    const unsigned long long bytes = *size_host;
    if (bytes == 0ull || bytes == ~0ull || (int64_t)bytes > capacity)
        return host_fail(RT_ERR_BAD_ARG, "rt_render_png_host: png_host is too small for this frame (rt_png_file_bound always suffices)");
    if (!direct) memcpy(png_host, static_cast<const uint8_t*>(c->png_stage) + 64, (size_t)bytes);
    *png_bytes_out = (int64_t)bytes;
    return RT_OK;
}

int rt_trace_rays_host(int device, const double* spheres_host, int32_t n_spheres, const double* point_lights_host,
                       const rt_env* env, const double* origins_host, const double* origin0,
                       const double* directions_host, int64_t n_rays, void* rgb_host, int32_t rgb_dtype,
                       uint64_t* stats_host) {
    if (spheres_host == nullptr || point_lights_host == nullptr || env == nullptr || rgb_host == nullptr) return host_fail(RT_ERR_BAD_ARG, "host entry point: NULL pointer, bad size or unknown enum value");
    if (n_spheres <= 0 || env->n_point_lights < 1 || n_rays < 0) return host_fail(RT_ERR_BAD_ARG, "host entry point: NULL pointer, bad size or unknown enum value");
    if (n_rays > 0 && directions_host == nullptr) return host_fail(RT_ERR_BAD_ARG, "host entry point: NULL pointer, bad size or unknown enum value");
    if (origins_host == nullptr && origin0 == nullptr) return host_fail(RT_ERR_BAD_ARG, "host entry point: NULL pointer, bad size or unknown enum value");
    if (rgb_dtype != RT_DTYPE_F32 && rgb_dtype != RT_DTYPE_F64) return host_fail(RT_ERR_BAD_ARG, "host entry point: NULL pointer, bad size or unknown enum value");
    Restore restore;
    DeviceCache* c = nullptr;
    int rc = enter(device, restore, &c);
    if (rc != RT_OK) return rc;

    const double *spheres_dev, *lights_dev;
    if ((rc = upload_scene(*c, spheres_host, n_spheres, point_lights_host, env->n_point_lights, &spheres_dev, &lights_dev)) != RT_OK) return rc;
    rt_env with_tree;
    if ((rc = attach_accel(*c, spheres_host, n_spheres, env, &with_tree)) != RT_OK) return rc;
    env = &with_tree;
    // the tables are host memory here by definition: per-scene constants are formed on the host
    with_tree.spheres_host = spheres_host;
    with_tree.point_lights_host = point_lights_host;
    const size_t plane = sizeof(double) * 3 * (size_t)n_rays;
    HOST_CUDA(reserve(c->rays, 2 * plane + 16));
    double* dirs_dev = static_cast<double*>(c->rays.ptr);
    double* orig_dev = origins_host != nullptr ? dirs_dev + 3 * n_rays : nullptr;
    if (n_rays > 0) {
        HOST_CUDA(cudaMemcpyAsync(dirs_dev, directions_host, plane, cudaMemcpyHostToDevice, c->stream));
        if (orig_dev != nullptr) HOST_CUDA(cudaMemcpyAsync(orig_dev, origins_host, plane, cudaMemcpyHostToDevice, c->stream));
    }
    const size_t elem = rgb_dtype == RT_DTYPE_F64 ? 8 : 4;
    HOST_CUDA(reserve(c->rgb, (size_t)n_rays * 3 * elem + 16));
    if (stats_host != nullptr) {
        HOST_CUDA(reserve(c->stats, sizeof(uint64_t) * RT_STATS_WORDS));
        HOST_CUDA(cudaMemsetAsync(c->stats.ptr, 0, sizeof(uint64_t) * RT_STATS_WORDS, c->stream));
    }
    rc = rt_trace_rays_dev(device, spheres_dev, n_spheres, lights_dev, env, orig_dev, origin0, dirs_dev, n_rays, c->rgb.ptr,
                           rgb_dtype, n_rays, stats_host != nullptr ? static_cast<uint64_t*>(c->stats.ptr) : nullptr, c->stream);
    if (rc != RT_OK) return rc;
    if (n_rays > 0) HOST_CUDA(cudaMemcpyAsync(rgb_host, c->rgb.ptr, (size_t)n_rays * 3 * elem, cudaMemcpyDeviceToHost, c->stream));
    if (stats_host != nullptr)
        HOST_CUDA(cudaMemcpyAsync(stats_host, c->stats.ptr, sizeof(uint64_t) * RT_STATS_WORDS, cudaMemcpyDeviceToHost, c->stream));
    HOST_CUDA(cudaStreamSynchronize(c->stream));
    return RT_OK;
}

}  // extern "C"

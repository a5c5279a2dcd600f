// rt_abi.cu -- the C ABI of include/rt_abi.h over the kernels of trace_exact.cu /
// trace_fast.cu.  Plain pointers and sizes in, status codes out; no state besides a
// per-thread staging cache used by the *_host entry points.
#include <cuda_runtime.h>

#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>

#include "../../include/rt_abi.h"
#include "rt_accel.h"
#include "rt_launch.h"

namespace {

thread_local char g_error[512] = "";

int fail(int status, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
    return status;
}

int cuda_fail(cudaError_t err, const char* what) {
    return fail(RT_ERR_CUDA, "%s: %s (%s)", what, cudaGetErrorString(err), cudaGetErrorName(err));
}

#define RT_CUDA(call)                                          \
    do {                                                       \
        cudaError_t err__ = (call);                            \
        if (err__ != cudaSuccess) return cuda_fail(err__, #call); \
    } while (0)

// Makes `device` current for the scope and restores the caller's device afterwards, so
// the library never disturbs the host framework's notion of the current device.
struct DeviceScope {
    int previous = -1;
    int status = RT_OK;
    explicit DeviceScope(int device) {
        int count = 0;
        cudaError_t err = cudaGetDeviceCount(&count);
        if (err != cudaSuccess || count == 0) {
            status = fail(RT_ERR_NO_DEVICE, "no CUDA device: %s", cudaGetErrorString(err));
            return;
        }
        if (device < 0 || device >= count) {
            status = fail(RT_ERR_NO_DEVICE, "device %d out of range (0..%d)", device, count - 1);
            return;
        }
        cudaGetDevice(&previous);
        if (previous != device) {
            err = cudaSetDevice(device);
            if (err != cudaSuccess) status = cuda_fail(err, "cudaSetDevice");
        }
    }
    ~DeviceScope() {
        int now = -1;
        if (previous >= 0 && cudaGetDevice(&now) == cudaSuccess && now != previous) cudaSetDevice(previous);
    }
};

int check_env(const rt_env* env, int32_t n_spheres) {
    if (env == nullptr) return fail(RT_ERR_BAD_ARG, "env is NULL");
    if (n_spheres <= 0) return fail(RT_ERR_BAD_ARG, "n_spheres = %d (scene has no shapes)", n_spheres);
    if (env->n_point_lights < 1) return fail(RT_ERR_BAD_ARG, "n_point_lights = %d; lights[0] must be a point light", env->n_point_lights);
    if (env->n_dome_lights < 0 || env->n_dome_lights > RT_MAX_DOME_LIGHTS)
        return fail(RT_ERR_BAD_ARG, "n_dome_lights = %d outside 0..%d", env->n_dome_lights, RT_MAX_DOME_LIGHTS);
    if (env->mode != RT_MODE_EXACT && env->mode != RT_MODE_FAST) return fail(RT_ERR_BAD_ARG, "unknown mode %d", env->mode);
    if (!(env->cutoff >= 0.0)) return fail(RT_ERR_BAD_ARG, "cutoff must be >= 0");
    if (env->accel_dev != nullptr && env->accel_bytes < (int64_t)sizeof(rt::AccelHeader))
        return fail(RT_ERR_BAD_ARG, "accel_bytes = %lld is smaller than a blob header", (long long)env->accel_bytes);
    if (env->accel_dev == nullptr && n_spheres > rt::max_staged_spheres())
        return fail(RT_ERR_UNSUPPORTED, "%d spheres exceed the %d that fit one block's shared memory; "
                    "build a tree with rt_accel_build_host and pass it in rt_env.accel_dev", n_spheres,
                    rt::max_staged_spheres());
    return RT_OK;
}

int check_camera(const rt_camera* cam, int32_t row0, int32_t row1) {
    if (cam == nullptr) return fail(RT_ERR_BAD_ARG, "camera is NULL");
    if (cam->width <= 0 || cam->height <= 0) return fail(RT_ERR_BAD_ARG, "camera size %dx%d", cam->width, cam->height);
    if (row0 < 0 || row1 > cam->height || row0 > row1) return fail(RT_ERR_BAD_ARG, "rows [%d,%d) outside 0..%d", row0, row1, cam->height);
    return RT_OK;
}

rt::SceneParams scene_params(const double* spheres_dev, int32_t n_spheres, const double* lights_dev, const rt_env* env) {
    rt::SceneParams sp;
    memset(&sp, 0, sizeof(sp));
    sp.spheres = spheres_dev;
    sp.lights = lights_dev;
    sp.accel = static_cast<const unsigned char*>(env->accel_dev);
    sp.accel_bytes = env->accel_bytes;
    sp.error_word = nullptr;
    // (the tree kernels and the exact policy's tie stack report through it; current device = the call's: DeviceScope)
    if (sp.accel != nullptr || env->mode == RT_MODE_EXACT) rt::device_error_word(&sp.error_word);
    sp.n_spheres = n_spheres;
    sp.n_lights_used = (env->flags & RT_FLAG_ALL_POINT_LIGHTS) ? env->n_point_lights : 1;
    sp.n_dome = env->n_dome_lights;
    sp.max_depth = env->max_depth;
    double sum = 0.0;
    for (int i = 0; i < 3; ++i) sp.eye[i] = env->eye[i];
    for (int i = 0; i < env->n_dome_lights; ++i) {
        sp.dome_intensity[i] = env->dome_intensity[i];
        sum += env->dome_intensity[i];
    }
    sp.dome_sum = sum;
    for (int i = 0; i < 3; ++i) sp.dome_color[i] = env->dome_color[i];
    sp.cutoff = env->mode == RT_MODE_FAST ? env->cutoff : 0.0;
    // fast policy: per-launch constants formed here instead of per block or per hit
    for (int i = 0; i < 3; ++i) sp.dome_rgb[i] = (float)(env->dome_color[i] * sum);
    sp.cutoff_f = (float)sp.cutoff;
    sp.chain_bound = -1.0f;
    sp.host_bound = 0;
    if (env->spheres_host != nullptr) {
        const double* f = env->spheres_host;
        const size_t n = (size_t)n_spheres;
        double lmax = 0.0, gmax = 0.0;
        for (size_t s = 0; s < n; ++s) {
            lmax = std::fmax(lmax, rt::material_bound(f[RT_F_DIFFUSE_GAIN * n + s], f[RT_F_SPECULAR_GAIN * n + s],
                                                      f[RT_F_SPECULAR_ROUGHNESS * n + s], f[RT_F_IRIDESCENCE_GAIN * n + s],
                                                      f[RT_F_COLOR_R * n + s], f[RT_F_COLOR_G * n + s], f[RT_F_COLOR_B * n + s],
                                                      f[RT_F_TEXTURE * n + s] != 0.0, sum, env->dome_color));
            gmax = std::fmax(gmax, std::fabs(f[RT_F_SPECULAR_GAIN * n + s]));
        }
        sp.chain_bound = rt::chain_bound_of(sp.cutoff, lmax, gmax);
        sp.host_bound = 1;
    }
    if (env->point_lights_host != nullptr)
        for (int i = 0; i < 3; ++i) sp.light0[i] = env->point_lights_host[i];
    return sp;
}

bool small_scene(const rt_env* env, int32_t n_spheres) {
    if (env->mode != RT_MODE_FAST || env->spheres_host == nullptr || env->point_lights_host == nullptr) return false;
    if (n_spheres > rt::kSmallMax || env->accel_dev != nullptr || (env->flags & RT_FLAG_GENERIC_KERNEL)) return false;
    if ((env->flags & RT_FLAG_ALL_POINT_LIGHTS) && env->n_point_lights > 1) return false;
    return true;
}

// Scenes of at most kSmallMax spheres whose table the caller holds on the host too (rt_env.spheres_host,
// point_lights_host): everything the fast policy's kernel reads of the scene, as one kernel parameter.
bool small_table(const rt_env* env, int32_t n_spheres, const double* origin0, rt::SmallTable* table) {
    if (!small_scene(env, n_spheres)) return false;
    memset(table, 0, sizeof(*table));
    const double* f = env->spheres_host;
    const size_t n = (size_t)n_spheres;
    for (size_t s = 0; s < n; ++s) {
        const double cx = f[RT_F_CX * n + s], cy = f[RT_F_CY * n + s], cz = f[RT_F_CZ * n + s], r = f[RT_F_RADIUS * n + s];
        table->geom[s] = {cx, cy, cz, r * r};
        table->prim[s] = rt::make_sphere_primary(origin0, cx, cy, cz, r);
        table->rec[s] = rt::make_shade_rec(cx, cy, cz, r, f[RT_F_DIFFUSE_GAIN * n + s], f[RT_F_SPECULAR_GAIN * n + s],
                                           f[RT_F_SPECULAR_ROUGHNESS * n + s], f[RT_F_IRIDESCENCE_GAIN * n + s],
                                           f[RT_F_COLOR_R * n + s], f[RT_F_COLOR_G * n + s], f[RT_F_COLOR_B * n + s],
                                           f[RT_F_TEXTURE * n + s] != 0.0);
    }
    return true;
}

// The screen rectangle and linspace steps of base.py:130-139, evaluated in the same
// double arithmetic Python uses.
rt::CameraParams camera_params(const rt_camera* cam, int32_t row0, int32_t row1) {
    rt::CameraParams cp;
    for (int i = 0; i < 3; ++i) cp.cam[i] = cam->position[i];
    const double aspect = (double)cam->width / (double)cam->height;
    cp.x_start = -1.0;
    cp.x_stop = 1.0;
    cp.y_start = 1.0 / aspect + 0.25;
    cp.y_stop = -1.0 / aspect + 0.25;
    cp.x_step = cam->width > 1 ? (cp.x_stop - cp.x_start) / (double)(cam->width - 1) : 0.0;
    cp.y_step = cam->height > 1 ? (cp.y_stop - cp.y_start) / (double)(cam->height - 1) : 0.0;
    cp.width = cam->width;
    cp.height = cam->height;
    cp.row0 = row0;
    cp.n_rows = row1 - row0;
    cp.strip_rows = cp.n_rows > 0 ? cp.n_rows : 1;
    cp.strip_step = 1;
    return cp;
}

// Ring of ticket counters per device (rt_launch.h).  A counter is zero whenever no launch
// is using it: the kernels put it back to zero themselves.
constexpr int kMaxDevices = 64;
constexpr unsigned int kRingSlots = 1024;
constexpr unsigned int kRingStride = 8;  // every counter in a 32-byte sector of its own
struct CounterRing {
    unsigned int* base = nullptr;
    std::atomic<unsigned int> next{0};
};
CounterRing g_rings[kMaxDevices];
std::mutex g_ring_mutex;
// one pinned, device-mapped word per device: kernels report through it (rt::DeviceError)
unsigned int* g_error_words[kMaxDevices] = {};

}  // namespace

namespace rt {
cudaError_t ticket_counter(unsigned int** counter_out) {
    int device = 0;
    cudaError_t err = cudaGetDevice(&device);
    if (err != cudaSuccess) return err;
    if (device < 0 || device >= kMaxDevices) return cudaErrorInvalidDevice;
    CounterRing& ring = g_rings[device];
    if (ring.base == nullptr) {
        std::lock_guard<std::mutex> lock(g_ring_mutex);
        if (ring.base == nullptr) {
            unsigned int* p = nullptr;
            if ((err = cudaMalloc(&p, sizeof(unsigned int) * kRingSlots * kRingStride)) != cudaSuccess) return err;
            if ((err = cudaMemset(p, 0, sizeof(unsigned int) * kRingSlots * kRingStride)) != cudaSuccess) {
                cudaFree(p);
                return err;
            }
            ring.base = p;
        }
    }
    *counter_out = ring.base + (size_t)(ring.next.fetch_add(1u) % kRingSlots) * kRingStride;
    return cudaSuccess;
}

cudaError_t device_error_word(unsigned int** device_pointer_out) {
    int device = 0;
    cudaError_t err = cudaGetDevice(&device);
    if (err != cudaSuccess) return err;
    if (device < 0 || device >= kMaxDevices) return cudaErrorInvalidDevice;
    if (g_error_words[device] == nullptr) {
        std::lock_guard<std::mutex> lock(g_ring_mutex);
        if (g_error_words[device] == nullptr) {
            unsigned int* p = nullptr;
            if ((err = cudaHostAlloc(&p, sizeof(unsigned int), cudaHostAllocMapped | cudaHostAllocPortable)) != cudaSuccess) return err;
            *p = 0u;
            g_error_words[device] = p;
        }
    }
    void* dev = nullptr;
    if ((err = cudaHostGetDevicePointer(&dev, g_error_words[device], 0)) != cudaSuccess) return err;
    *device_pointer_out = static_cast<unsigned int*>(dev);
    return cudaSuccess;
}

void release_ticket_counters() {
    std::lock_guard<std::mutex> lock(g_ring_mutex);
    int previous = -1;
    cudaGetDevice(&previous);
    for (int d = 0; d < kMaxDevices; ++d) {
        if (g_rings[d].base == nullptr) continue;
        if (cudaSetDevice(d) == cudaSuccess) cudaFree(g_rings[d].base);
        g_rings[d].base = nullptr;
    }
    if (previous >= 0) cudaSetDevice(previous);
}
}  // namespace rt

namespace {

// What a kernel of an earlier launch on `device` reported (sticky until read here).
int pending_device_error(int device) {
    if (device < 0 || device >= kMaxDevices || g_error_words[device] == nullptr) return RT_OK;
    const unsigned int code = *reinterpret_cast<volatile unsigned int*>(g_error_words[device]);
    if (code == rt::kDevOk) return RT_OK;
    *reinterpret_cast<volatile unsigned int*>(g_error_words[device]) = rt::kDevOk;
    if (code == rt::kDevBadAccelBlob)
        return fail(RT_ERR_UNSUPPORTED, "an earlier launch on device %d was given an rt_env.accel_dev that is not a blob of "
                    "rt_accel_build_host for its sphere table (magic, shape count or size disagree): it rendered nothing", device);
    if (code == rt::kDevAccelStack)
        return fail(RT_ERR_UNSUPPORTED, "an earlier launch on device %d walked a tree deeper than the traversal stack (%d): "
                    "not a blob of rt_accel_build_host; its pixels are wrong", device, rt::kAccelStack);
    if (code == rt::kDevTieStack)
        return fail(RT_ERR_UNSUPPORTED, "an earlier exact-policy launch on device %d met more shapes at exactly equal distances "
                    "(duplicated shapes bouncing into each other) than its stack of pending rays holds: those pixels lack rays", device);
    return fail(RT_ERR_UNSUPPORTED, "an earlier launch on device %d reported device error %u", device, code);
}

const rt::Launchers& launchers(int mode) { return mode == RT_MODE_FAST ? rt::fast_launchers() : rt::exact_launchers(); }

// uint8(255*clip(v,0,1)) truncating (base.py:143-151), planar -> interleaved.
template <typename T, int CH>
__global__ void quantize_kernel(const T* rgb, long long plane_stride, long long n, uint8_t* out) {
    const long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    unsigned int q[3];
    for (int c = 0; c < 3; ++c) {
        double v = (double)rgb[c * plane_stride + p];
        v = fmin(fmax(v, 0.0), 1.0);
        q[c] = (unsigned int)(255.0 * v);
    }
    if (CH == 4) {
        reinterpret_cast<uchar4*>(out)[p] = make_uchar4((unsigned char)q[0], (unsigned char)q[1], (unsigned char)q[2], 255);
    } else {
        out[3 * p + 0] = (uint8_t)q[0];
        out[3 * p + 1] = (uint8_t)q[1];
        out[3 * p + 2] = (uint8_t)q[2];
    }
}

}  // namespace

extern "C" {

int rt_abi_version(void) { return RT_ABI_VERSION; }

const char* rt_last_error(void) { return g_error; }

// not part of the ABI: lets rt_abi_host.cu record a message (hidden from the dynamic table)
__attribute__((visibility("hidden"))) void rt_internal_set_error(const char* message) {
    snprintf(g_error, sizeof(g_error), "%s", message);
}
// not part of the ABI: will a camera launch with this env take the kernel-parameter path (and so never
// read the device copies of the tables)?
__attribute__((visibility("hidden"))) int rt_internal_small_scene(const rt_env* env, int32_t n_spheres) {
    return small_scene(env, n_spheres) ? 1 : 0;
}

int rt_device_status(int device) { return pending_device_error(device); }

int rt_scene_travels_as_parameter(const rt_env* env, int32_t n_spheres) {
    if (env == nullptr || n_spheres <= 0) return 0;
    return small_scene(env, n_spheres) ? 1 : 0;
}

int rt_device_count(int* count_out) {
    if (count_out == nullptr) return fail(RT_ERR_BAD_ARG, "count_out is NULL");
    int count = 0;
    cudaError_t err = cudaGetDeviceCount(&count);
    if (err != cudaSuccess) {
        *count_out = 0;
        return fail(RT_ERR_NO_DEVICE, "cudaGetDeviceCount: %s", cudaGetErrorString(err));
    }
    *count_out = count;
    return RT_OK;
}

namespace {
int render_camera(int device, const double* spheres_dev, int32_t n_spheres, const double* point_lights_dev,
                  const rt_env* env, const rt::CameraParams& cp, void* rgb_dev, int32_t rgb_dtype,
                  int64_t plane_stride, uint8_t* rgba8_dev, uint64_t* stats_dev, void* stream) {
    if (spheres_dev == nullptr || point_lights_dev == nullptr) return fail(RT_ERR_BAD_ARG, "scene pointers are NULL");
    if (rgb_dtype != RT_DTYPE_F32 && rgb_dtype != RT_DTYPE_F64) return fail(RT_ERR_BAD_ARG, "unknown rgb_dtype %d", rgb_dtype);
    const bool image_order = (env->flags & RT_FLAG_IMAGE_ORDER) != 0;
    const int64_t n = image_order ? (int64_t)cp.height * cp.width : (int64_t)cp.n_rows * cp.width;
    if (rgb_dev != nullptr && plane_stride < n) return fail(RT_ERR_BAD_ARG, "plane_stride %lld < %lld pixels", (long long)plane_stride, (long long)n);
    // camera rays leave the eye: the kernels take "towards the camera" (shader.py:76) from the ray itself
    for (int i = 0; i < 3; ++i)
        if (env->eye[i] != cp.cam[i])
            return fail(RT_ERR_BAD_ARG, "camera rays must start at env->eye (scene.camera.position, shader.py:76); "
                        "for rays from elsewhere use rt_trace_rays_dev");
    DeviceScope scope(device);
    if (scope.status != RT_OK) return scope.status;
    // 3-byte pixels go out as packed words from the wide-store kernels when every patch row starts on a word boundary
    int bytes8 = rt::kRgba8;
    if (env->flags & RT_FLAG_RGB8)
        bytes8 = ((env->flags & RT_FLAG_WIDE_STORES) && cp.width % 4 == 0 && reinterpret_cast<uintptr_t>(rgba8_dev) % 4 == 0)
                     ? rt::kRgb8Words : rt::kRgb8Bytes;
    rt::OutputParams out = {rgb_dev, rgba8_dev, reinterpret_cast<unsigned long long*>(stats_dev), plane_stride,
                            rgb_dtype == RT_DTYPE_F64, image_order ? 1 : 0, bytes8,
                            (env->flags & RT_FLAG_WIDE_STORES) ? 1 : 0, (env->flags & RT_FLAG_EVEN_STORES) ? 1 : 0};
    rt::SmallTable table;
    const bool small = small_table(env, n_spheres, cp.cam, &table);
    cudaError_t err = launchers(env->mode).render_rows(stats_dev != nullptr, scene_params(spheres_dev, n_spheres, point_lights_dev, env),
                                                      cp, out, small ? &table : nullptr, static_cast<cudaStream_t>(stream));
    if (err != cudaSuccess) return cuda_fail(err, "render_rows launch");
    return RT_OK;
}
}  // namespace

int rt_render_rows_dev(int device, const double* spheres_dev, int32_t n_spheres, const double* point_lights_dev,
                       const rt_env* env, const rt_camera* cam, int32_t row0, int32_t row1, void* rgb_dev,
                       int32_t rgb_dtype, int64_t plane_stride, uint8_t* rgba8_dev, uint64_t* stats_dev,
                       void* stream) {
    int rc;
    if ((rc = pending_device_error(device)) != RT_OK) return rc;
    if ((rc = check_env(env, n_spheres)) != RT_OK) return rc;
    if ((rc = check_camera(cam, row0, row1)) != RT_OK) return rc;
    return render_camera(device, spheres_dev, n_spheres, point_lights_dev, env, camera_params(cam, row0, row1), rgb_dev,
                         rgb_dtype, plane_stride, rgba8_dev, stats_dev, stream);
}

int64_t rt_strip_shard_rows(int32_t height, int32_t strip_rows, int32_t strip_first, int32_t strip_step) {
    if (height <= 0 || strip_rows <= 0 || strip_first < 0 || strip_step <= 0) return -1;
    int64_t rows = 0;
    for (int64_t k = strip_first; k * strip_rows < height; k += strip_step) {
        const int64_t left = (int64_t)height - k * strip_rows;
        rows += left < strip_rows ? left : strip_rows;
    }
    return rows;
}

int rt_render_strips_dev(int device, const double* spheres_dev, int32_t n_spheres, const double* point_lights_dev,
                         const rt_env* env, const rt_camera* cam, int32_t strip_rows, int32_t strip_first,
                         int32_t strip_step, void* rgb_dev, int32_t rgb_dtype, int64_t plane_stride,
                         uint8_t* rgba8_dev, uint64_t* stats_dev, void* stream) {
    int rc;
    if ((rc = pending_device_error(device)) != RT_OK) return rc;
    if ((rc = check_env(env, n_spheres)) != RT_OK) return rc;
    if ((rc = check_camera(cam, 0, cam != nullptr ? cam->height : 0)) != RT_OK) return rc;
    if (strip_rows <= 0 || strip_step <= 0 || strip_first < 0 || strip_first >= strip_step)
        return fail(RT_ERR_BAD_ARG, "strips: rows %d, first %d, step %d", strip_rows, strip_first, strip_step);
    rt::CameraParams cp = camera_params(cam, 0, cam->height);
    cp.row0 = strip_first * strip_rows;
    cp.n_rows = (int)rt_strip_shard_rows(cam->height, strip_rows, strip_first, strip_step);
    cp.strip_rows = strip_rows;
    cp.strip_step = strip_step;
    return render_camera(device, spheres_dev, n_spheres, point_lights_dev, env, cp, rgb_dev, rgb_dtype, plane_stride,
                         rgba8_dev, stats_dev, stream);
}

int rt_peer_buffer_create(int device, uint64_t bytes, void** ptr_out, uint8_t* handle_out) {
    if (ptr_out == nullptr || handle_out == nullptr || bytes == 0) return fail(RT_ERR_BAD_ARG, "peer buffer: NULL pointer or zero size");
    static_assert(sizeof(cudaIpcMemHandle_t) == RT_IPC_HANDLE_BYTES, "RT_IPC_HANDLE_BYTES must equal the CUDA IPC handle size");
    DeviceScope scope(device);
    if (scope.status != RT_OK) return scope.status;
    void* p = nullptr;
    RT_CUDA(cudaMalloc(&p, bytes));
    cudaIpcMemHandle_t handle;
    cudaError_t err = cudaMemset(p, 0, bytes);
    if (err == cudaSuccess) err = cudaIpcGetMemHandle(&handle, p);
    if (err != cudaSuccess) {
        cudaFree(p);
        return cuda_fail(err, "peer buffer create");
    }
    memcpy(handle_out, &handle, sizeof(handle));
    *ptr_out = p;
    return RT_OK;
}

int rt_peer_buffer_open(int device, const uint8_t* handle_in, void** ptr_out) {
    if (ptr_out == nullptr || handle_in == nullptr) return fail(RT_ERR_BAD_ARG, "peer buffer: NULL pointer");
    DeviceScope scope(device);
    if (scope.status != RT_OK) return scope.status;
    cudaIpcMemHandle_t handle;
    memcpy(&handle, handle_in, sizeof(handle));
    RT_CUDA(cudaIpcOpenMemHandle(ptr_out, handle, cudaIpcMemLazyEnablePeerAccess));
    return RT_OK;
}

int rt_peer_buffer_close(int device, void* ptr) {
    if (ptr == nullptr) return RT_OK;
    DeviceScope scope(device);
    if (scope.status != RT_OK) return scope.status;
    RT_CUDA(cudaIpcCloseMemHandle(ptr));
    return RT_OK;
}

int rt_peer_buffer_destroy(int device, void* ptr) {
    if (ptr == nullptr) return RT_OK;
    DeviceScope scope(device);
    if (scope.status != RT_OK) return scope.status;
    RT_CUDA(cudaFree(ptr));
    return RT_OK;
}

int rt_trace_rays_dev(int device, const double* spheres_dev, int32_t n_spheres, const double* point_lights_dev,
                      const rt_env* env, const double* origins_dev, const double* origin0,
                      const double* directions_dev, int64_t n_rays, void* rgb_dev, int32_t rgb_dtype,
                      int64_t plane_stride, uint64_t* stats_dev, void* stream) {
    int rc;
    if ((rc = pending_device_error(device)) != RT_OK) return rc;
    if ((rc = check_env(env, n_spheres)) != RT_OK) return rc;
    if (spheres_dev == nullptr || point_lights_dev == nullptr) return fail(RT_ERR_BAD_ARG, "scene pointers are NULL");
    if (n_rays < 0) return fail(RT_ERR_BAD_ARG, "n_rays = %lld", (long long)n_rays);
    if (n_rays > 0 && directions_dev == nullptr) return fail(RT_ERR_BAD_ARG, "directions are NULL");
    if (origins_dev == nullptr && origin0 == nullptr) return fail(RT_ERR_BAD_ARG, "no ray origin given");
    if (rgb_dtype != RT_DTYPE_F32 && rgb_dtype != RT_DTYPE_F64) return fail(RT_ERR_BAD_ARG, "unknown rgb_dtype %d", rgb_dtype);
    if (n_rays > 0 && rgb_dev == nullptr) return fail(RT_ERR_BAD_ARG, "rgb_dev is NULL");
    if (plane_stride < n_rays) return fail(RT_ERR_BAD_ARG, "plane_stride %lld < %lld rays", (long long)plane_stride, (long long)n_rays);
    DeviceScope scope(device);
    if (scope.status != RT_OK) return scope.status;
    rt::RayBatch rb;
    rb.origins = origins_dev;
    rb.directions = directions_dev;
    for (int i = 0; i < 3; ++i) rb.origin0[i] = origin0 != nullptr ? origin0[i] : 0.0;
    rb.n = n_rays;
    rt::OutputParams out = {rgb_dev, nullptr, reinterpret_cast<unsigned long long*>(stats_dev), plane_stride,
                            rgb_dtype == RT_DTYPE_F64, 0};
    cudaError_t err = launchers(env->mode).trace_rays(stats_dev != nullptr, scene_params(spheres_dev, n_spheres, point_lights_dev, env), rb,
                                                     out, static_cast<cudaStream_t>(stream));
    if (err != cudaSuccess) return cuda_fail(err, "trace_rays launch");
    return RT_OK;
}

int rt_ray_directions_dev(int device, const rt_camera* cam, int32_t row0, int32_t row1, double* directions_dev,
                          int64_t plane_stride, void* stream) {
    int rc;
    if ((rc = check_camera(cam, row0, row1)) != RT_OK) return rc;
    if (directions_dev == nullptr) return fail(RT_ERR_BAD_ARG, "directions_dev is NULL");
    if (plane_stride < (int64_t)(row1 - row0) * cam->width) return fail(RT_ERR_BAD_ARG, "plane_stride too small");
    DeviceScope scope(device);
    if (scope.status != RT_OK) return scope.status;
    cudaError_t err = rt::exact_launchers().ray_directions(camera_params(cam, row0, row1), directions_dev, plane_stride,
                                                           static_cast<cudaStream_t>(stream));
    if (err != cudaSuccess) return cuda_fail(err, "ray_directions launch");
    return RT_OK;
}

int rt_sphere_distance_dev(int device, const double* origins_dev, const double* origin0, const double* directions_dev,
                           int64_t n_rays, const double* center, double radius, int32_t mode, double* distance_dev,
                           void* stream) {
    if (n_rays < 0) return fail(RT_ERR_BAD_ARG, "n_rays = %lld", (long long)n_rays);
    if (center == nullptr || distance_dev == nullptr || (n_rays > 0 && directions_dev == nullptr))
        return fail(RT_ERR_BAD_ARG, "NULL pointer");
    if (origins_dev == nullptr && origin0 == nullptr) return fail(RT_ERR_BAD_ARG, "no ray origin given");
    if (mode != RT_MODE_EXACT && mode != RT_MODE_FAST) return fail(RT_ERR_BAD_ARG, "unknown mode %d", mode);
    DeviceScope scope(device);
    if (scope.status != RT_OK) return scope.status;
    rt::RayBatch rb;
    rb.origins = origins_dev;
    rb.directions = directions_dev;
    for (int i = 0; i < 3; ++i) rb.origin0[i] = origin0 != nullptr ? origin0[i] : 0.0;
    rb.n = n_rays;
    cudaError_t err = launchers(mode).sphere_distance(rb, center, radius, distance_dev, static_cast<cudaStream_t>(stream));
    if (err != cudaSuccess) return cuda_fail(err, "sphere_distance launch");
    return RT_OK;
}

int rt_quantize_dev(int device, const void* rgb_dev, int32_t rgb_dtype, int64_t plane_stride, int64_t n_pixels,
                    uint8_t* out_dev, int32_t channels_out, void* stream) {
    if (rgb_dev == nullptr || out_dev == nullptr) return fail(RT_ERR_BAD_ARG, "NULL pointer");
    if (n_pixels < 0 || plane_stride < n_pixels) return fail(RT_ERR_BAD_ARG, "bad sizes");
    if (channels_out != 3 && channels_out != 4) return fail(RT_ERR_BAD_ARG, "channels_out must be 3 or 4");
    if (rgb_dtype != RT_DTYPE_F32 && rgb_dtype != RT_DTYPE_F64) return fail(RT_ERR_BAD_ARG, "unknown rgb_dtype %d", rgb_dtype);
    if (n_pixels == 0) return RT_OK;
    DeviceScope scope(device);
    if (scope.status != RT_OK) return scope.status;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const unsigned int grid = (unsigned int)((n_pixels + 255) / 256);
    if (rgb_dtype == RT_DTYPE_F64) {
        if (channels_out == 4) quantize_kernel<double, 4><<<grid, 256, 0, s>>>((const double*)rgb_dev, plane_stride, n_pixels, out_dev);
        else quantize_kernel<double, 3><<<grid, 256, 0, s>>>((const double*)rgb_dev, plane_stride, n_pixels, out_dev);
    } else {
        if (channels_out == 4) quantize_kernel<float, 4><<<grid, 256, 0, s>>>((const float*)rgb_dev, plane_stride, n_pixels, out_dev);
        else quantize_kernel<float, 3><<<grid, 256, 0, s>>>((const float*)rgb_dev, plane_stride, n_pixels, out_dev);
    }
    cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) return cuda_fail(err, "quantize launch");
    return RT_OK;
}

}  // extern "C"

// rt_device.cuh -- device-side ray/sphere tracer shared by the two kernel builds.
//
// Replaces, per ray, NumpyRenderer.raytrace_scene and everything it calls
// (reference: ray_tracer/infrastructure/numpy/base.py:91-141, shape.py:28-51,
// shader.py:63-320).  Two arithmetic policies, selected at compile time:
//
//   Exact  all FP64, the reference's expressions and operation order; its translation
//          unit (trace_exact.cu) is compiled with -fmad=false so nothing is contracted.
//          Geometry decisions (nearest hit, root choice, shadow test, checker parity)
//          therefore agree with the NumPy reference bit for bit; only libm-grade
//          functions (pow, sin) may differ in the last place.
//   Fast   FP64 geometry in half-b form with FMAs, square roots only for accepted hits,
//          a square-root-free any-hit shadow test, one division for the whole GGX term,
//          rsqrt normalisation, and a proven throughput cutoff of the bounce chain.
//
// The reference's recursion is a linear chain (every hit spawns exactly one mirror ray,
// shader.py:143-161) and its colour enters the parent linearly with weight
// 0.5*specular_gain*lit (shader.py:106), so the chain is run as a loop with a running
// weight held in registers:  pixel = sum_k (prod_{j<k} w_j) * local_k.
#pragma once

#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

#include "../../include/rt_abi.h"

namespace rt {

constexpr int kModeExact = 0;
constexpr int kModeFast = 1;
constexpr double kMiss = RT_FARAWAY;  // base.py:12
constexpr double kNudge = 0.0001;     // shader.py:77
constexpr double kAmbient = 0.004;    // shader.py:183-184
constexpr double kEps = 1e-8;         // shader.py:275
constexpr double kPi = 3.141592653589793;
// ((1.5-1)/(1.5+1))**2 as Python evaluates it (shader.py:290)
constexpr double kF0 = 0.04000000000000001;
// (thin_film_ior-1)/2 with thin_film_ior = 1.4 (shader.py:54, :215)
constexpr double kHueShift = (1.4 - 1.0) / 2.0;
constexpr double kFilmWeight = 0.1;      // shader.py:52
constexpr double kFilmThickness = 0.3;   // shader.py:53

// Everything a launch needs besides the rays; passed by value as a kernel parameter.
struct SceneParams {
    const double* spheres;  // field-major sphere SoA (rt_abi.h), device
    const double* lights;   // n_point_lights x 3, device
    const unsigned char* accel;  // BVH blob (rt_accel.h) or nullptr: test every sphere
    long long accel_bytes;       // size the caller gave for it
    unsigned int* error_word;    // host-mapped word a kernel sets when it cannot render what it was given (DeviceError)
    int n_spheres;
    int n_lights_used;      // 1 unless RT_FLAG_ALL_POINT_LIGHTS
    int n_dome;
    int max_depth;          // <= 0: unlimited
    double eye[3];
    double dome_intensity[RT_MAX_DOME_LIGHTS];
    double dome_sum;        // sum of dome_intensity (Fast)
    double dome_color[3];
    double cutoff;          // Fast: chain cutoff in linear units, 0 = off
    // Fast policy, formed on the host (rt_abi.cu: scene_params):
    float dome_rgb[3];      // dome_color * dome_sum
    float cutoff_f;         // (float)cutoff
    float chain_bound;      // sup |local colour| / (1 - sup w) over the materials when the caller gave the host
                            // copy of the table (rt_env.spheres_host); < 0: no cut-off possible; see host_bound
    int host_bound;         // 1: chain_bound is valid; 0: the staging loop computes it (no host copy of the table)
    double light0[3];       // lights[0] when the host copy of the light table was given (small-scene kernels)
};

// What a kernel reports through SceneParams::error_word; the next ABI call on the device fails loudly with it
// (rt_device_status() reads and clears it).  Pixels of a launch that reported one are undefined.
enum DeviceError : unsigned int {
    kDevOk = 0,
    kDevBadAccelBlob = 1,    // rt_env.accel_dev is not a blob of rt_accel_build_host for THIS sphere table
    kDevAccelStack = 2,      // a tree deeper than the traversal stack (not a blob of rt_accel_build_host)
    kDevTieStack = 3         // exact policy: more rays pending from shapes at exactly equal distance than the stack holds
};

// Camera rows -> rays (base.py:123-141).  The linspace steps are divided on the host
// (IEEE division, same value NumPy gets).
struct CameraParams {
    double cam[3];
    double x_start, x_stop, x_step;  // np.linspace(-1, 1, W)
    double y_start, y_stop, y_step;  // np.linspace(1/a + .25, -1/a + .25, H)
    int width, height;
    // Local row r of the launch (r in [0, n_rows)) is image row
    //   row0 + r + (r / strip_rows) * (strip_step - 1) * strip_rows:
    // a contiguous band has strip_step = 1; an interleaved shard owns every
    // strip_step-th strip of strip_rows rows (a multiple of the patch height).
    int row0, n_rows;
    int strip_rows, strip_step;
};
__host__ __device__ inline int image_row(const CameraParams& cp, int local_row) {
    return cp.strip_step == 1 ? cp.row0 + local_row
                              : cp.row0 + local_row + (local_row / cp.strip_rows) * (cp.strip_step - 1) * cp.strip_rows;
}

// Explicit ray batch (base.py:91-96 called with arrays).
struct RayBatch {
    const double* origins;     // 3 planes of n, or nullptr
    const double* directions;  // 3 planes of n
    double origin0[3];
    long long n;
};

enum Bytes8 : int {
    kRgba8 = 0,       // 4 bytes per pixel, one 32-bit store per pixel
    kRgb8Words = 1,   // 3 bytes per pixel, a patch row packed into 32-bit words by the warp (width % 4 == 0)
    kRgb8Bytes = 2    // 3 bytes per pixel, byte stores (any width)
};
struct OutputParams {
    void* rgb;            // planar, may be nullptr
    uint8_t* rgba8;       // 8-bit image (layout: bytes8), may be nullptr
    unsigned long long* stats;
    long long plane_stride;
    int rgb_is_f64;
    int image_order;      // camera kernel: index outputs by image pixel (row*W + col) instead of
                          // the shard's packed order -- the buffers are whole frames, possibly
                          // peer memory of another GPU (RT_FLAG_IMAGE_ORDER)
    int bytes8;           // Bytes8 (RT_FLAG_RGB8)
    int wide_patches;     // camera kernel: 32x1 patches instead of 8x4 (RT_FLAG_WIDE_STORES); host-side choice
    int even_stores;      // wide patches only: items alternately from both ends of the frame (RT_FLAG_EVEN_STORES)
};

// Per-sphere records staged in shared memory.  Reads are warp-uniform (every lane tests
// the same sphere) so each is one broadcast transaction: two 16-byte loads per record.
struct __align__(16) SphereGeom {
    double cx, cy, cz;
    double rr;  // radius * radius
};
struct __align__(16) SphereAux {
    double inv_r;  // 1.0 / radius (shader.py:74)
    double cc;     // |C|^2 summed left to right (shape.py:35 abs(self.position))
};

// Fast policy: what shading one hit needs of its sphere (64 bytes, read as four 16-byte loads).
// Formed once per scene -- on the host for the constant-bank kernels, by the staging loop otherwise --
// so that nothing per-material is computed per hit: the texture colour arrives multiplied by
// diffuse_gain, roughness as alpha^2 = roughness^4, the film weight as 0.1 * iridescence_gain
// (shader.py:52), the bounce weight as 0.5 * specular_gain (shader.py:106).  Single precision
// where the value only enters the FP32 shading tail.
struct __align__(16) ShadeRec {
    double cx, cy, cz, inv_r;
    float kd_r, kd_g, kd_b, spec_gain;
    float a2, film, half_gain;
    unsigned int flags;
};
constexpr unsigned int kRecChecker = 1u;  // TextureChecker (shader.py:22-32)
constexpr unsigned int kRecInward = 2u;   // negative radius: the normal points into the sphere

// rays that start at the camera: O - C and |O - C|^2 - r^2 are per-sphere constants of the launch
struct __align__(16) SpherePrimary {
    double lx, ly, lz;  // origin0 - centre
    double c;           // |origin0 - centre|^2 - r^2
};

__host__ __device__ inline ShadeRec make_shade_rec(double cx, double cy, double cz, double r, double diffuse_gain,
                                                   double specular_gain, double roughness, double irid_gain, double cr,
                                                   double cg, double cb, bool checker) {
    ShadeRec m;
    m.cx = cx, m.cy = cy, m.cz = cz;
    m.inv_r = 1.0 / r;
    m.kd_r = (float)(checker ? diffuse_gain : diffuse_gain * cr);
    m.kd_g = (float)(checker ? diffuse_gain : diffuse_gain * cg);
    m.kd_b = (float)(checker ? diffuse_gain : diffuse_gain * cb);
    m.spec_gain = (float)specular_gain;
    const double alpha = roughness * roughness;
    m.a2 = (float)(alpha * alpha);
    m.film = (float)(kFilmWeight * irid_gain);
    m.half_gain = (float)(0.5 * specular_gain);
    m.flags = (checker ? kRecChecker : 0u) | (r < 0.0 ? kRecInward : 0u);
    return m;
}
__host__ __device__ inline SpherePrimary make_sphere_primary(const double* origin0, double cx, double cy, double cz, double r) {
    const double lx = origin0[0] - cx, ly = origin0[1] - cy, lz = origin0[2] - cz;
    return {lx, ly, lz, fma(lz, lz, fma(ly, ly, fma(lx, lx, -(r * r))))};
}

// Upper bound of |local colour| of one material (fast policy's chain cut-off).  local = ambient +
// diffuse + dome + gain*(spec_base + gain*glint) + film, with F <= 1, G1(L) <= 1,
// D <= a2/(pi*(min(1,a2^2)+eps)) and G1(V)/(4 NdotV + eps) <= 1/(2*(sqrt(min(a2,1))+eps)).
__host__ __device__ inline double material_bound(double diffuse_gain, double specular_gain, double roughness, double irid_gain,
                                                 double cr, double cg, double cb, bool checker, double dome_sum,
                                                 const double* dome_color) {
    double tex = checker ? 1.0 : fmax(fabs(cr), fmax(fabs(cg), fabs(cb)));
    double dome = fabs(dome_sum) * fmax(fabs(dome_color[0]), fmax(fabs(dome_color[1]), fabs(dome_color[2])));
    double alpha = roughness * roughness;
    double a2 = alpha * alpha;
    double dmax = a2 / (kPi * (fmin(1.0, a2 * a2) + kEps));
    double gv = 1.0 / (2.0 * (sqrt(fmin(a2, 1.0)) + kEps));
    double g = fabs(specular_gain);
    double spec = g * (dmax * gv + g);
    double film = 1.4 * kFilmWeight * fabs(irid_gain);
    return kAmbient + fabs(diffuse_gain) * tex + dome + spec + film;
}

// sup |local| / (1 - sup w) of the chain cut-off (rt_fast.cuh), or -1 when it cannot be used
__host__ __device__ inline float chain_bound_of(double cutoff, double local_max, double gain_max) {
    if (!(cutoff > 0.0)) return -1.0f;
    const double wmax = 0.5 * gain_max;
    // remaining <= W * lmax * sum_j wmax^j ; only usable when the series converges
    if (wmax < 1.0 && local_max < 1e300) return (float)(local_max / (1.0 - wmax)) * 1.000001f;
    return -1.0f;
}

#ifndef RT_SMALL_MAX
#define RT_SMALL_MAX 4
#endif
// Scenes of at most kSmallMax spheres whose table the caller also holds on the host: the whole table
// travels as a kernel parameter (constant bank), the loops over it are unrolled at compile time and
// the sphere constants are instruction operands -- no staging loop, no shared-memory sphere loads.
constexpr int kSmallMax = RT_SMALL_MAX;
struct SmallTable {
    SphereGeom geom[kSmallMax];
    SpherePrimary prim[kSmallMax];
    ShadeRec rec[kSmallMax];
};

struct Vec3 {
    double x, y, z;
};

__device__ __forceinline__ Vec3 operator+(Vec3 a, Vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ Vec3 operator-(Vec3 a, Vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ Vec3 operator*(Vec3 a, double k) { return {a.x * k, a.y * k, a.z * k}; }
// base.py:34-35: products summed left to right.
__device__ __forceinline__ double dot(Vec3 a, Vec3 b) { return (a.x * b.x) + (a.y * b.y) + (a.z * b.z); }
__device__ __forceinline__ double clip01(double v) { return fmin(fmax(v, 0.0), 1.0); }

// base.py:61-64 as written: multiply by the reciprocal of the length, 0 -> 1.
__device__ __forceinline__ Vec3 norm_exact(Vec3 a) {
    double mag = sqrt(dot(a, a));
    double inv = 1.0 / (mag == 0.0 ? 1.0 : mag);
    return a * inv;
}

// ---------------------------------------------------------------------------------------
// Fast-policy special functions.  CUDA's double sqrt / rsqrt / division are IEEE-rounded
// and each carries a range-checked slow path (a branch plus a subroutine call).  The fast
// policy only needs results good to a few ulp, so it seeds from the 64-bit MUFU
// approximations (rel. error <= 2^-22, whole exponent range) and refines with two Newton
// steps: branch-free, 5-8 FP64 instructions each.  Arguments must be positive, normal
// and finite; callers guard the zero case with a select.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ double rsqrt_fast(double x) {
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    const double hx = 0.5 * x;
    double e = fma(-hx * y, y, 0.5);  // (1 - x*y*y)/2
    y = fma(y, e, y);
    e = fma(-hx * y, y, 0.5);
    return fma(y, e, y);
}
__device__ __forceinline__ double rcp_fast(double x) {
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    double e = fma(-x, y, 1.0);
    y = fma(y, e, y);
    e = fma(-x, y, 1.0);
    return fma(y, e, y);
}
// sqrt(x) for x > 0: one Newton step on the reciprocal root, then one on the root itself.
__device__ __forceinline__ double sqrt_pos_fast(double x) {
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    double e = fma(-(0.5 * x) * y, y, 0.5);
    y = fma(y, e, y);            // ~2^-44
    double s = x * y;
    double r = fma(-s, s, x);    // x - s*s
    return fma(r, 0.5 * y, s);
}
__device__ __forceinline__ double sqrt_fast(double x) { return __double2hiint(x) > 0 ? sqrt_pos_fast(x) : 0.0; }
// max(v,0) / clip without the NaN plumbing of fmax()/fmin() (inputs here are never NaN).
// The tests read the sign / exponent from the high word on the integer pipe (a written
// "v < 1.0 ? v : 1.0" is turned back into the DSETP.MIN + fix-up sequence of fmin()).
__device__ __forceinline__ double pos_part(double v) { return __double2hiint(v) < 0 ? 0.0 : v; }
__device__ __forceinline__ double clip01_fast(double v) {
    const int hi = __double2hiint(v);
    return hi < 0 ? 0.0 : (hi >= 0x3ff00000 ? 1.0 : v);
}

// Fast normalisation: one reciprocal square root, no division.  A zero vector stays zero
// like the reference's (0 * 1/1).
__device__ __forceinline__ double fdot(Vec3 a, Vec3 b) { return fma(a.z, b.z, fma(a.y, b.y, a.x * b.x)); }
__device__ __forceinline__ Vec3 norm_fast(Vec3 a) {
    double m2 = fdot(a, a);
    double inv = m2 > 1e-290 ? rsqrt_fast(m2) : 1.0;  // (flushes vectors shorter than 1e-145)
    return a * inv;
}

}  // namespace rt

// rt_trace.cuh -- intersection, shadow test and shading terms (device): the exact policy's
// functions and the sphere-test primitives both policies share.  See rt_device.cuh for the policies.
#pragma once

#include "rt_device.cuh"

namespace rt {

// Shared-memory view of the staged scene (filled by stage_scene()).
// Per-sphere records staged in shared memory are addressed through 32-bit shared-window
// addresses and read with explicit ld.shared.v2.f64: every lane reads the same record
// (one broadcast wavefront), the address is base + index * 32 with no per-use
// generic-to-shared conversion.
struct Staged {
    unsigned int geom;   // SphereGeom[n]
    unsigned int aux;    // SphereAux[n]
    unsigned int mat;    // SphereMaterial[n]
    unsigned int lights; // double[3 * n_lights_used]
};
__device__ __forceinline__ void lds_f64x2(unsigned int addr, double& a, double& b) {
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(a), "=d"(b) : "r"(addr));
}
__device__ __forceinline__ SphereGeom load_geom(const Staged& st, int s) {
    SphereGeom g;
    lds_f64x2(st.geom + (unsigned int)s * 32u, g.cx, g.cy);
    lds_f64x2(st.geom + (unsigned int)s * 32u + 16u, g.cz, g.rr);
    return g;
}
// Material of one sphere as staged (64 bytes): NumpyShader's gains (shader.py:36-54), the
// constant texture colour and the texture kind (1.0 = checker).
struct __align__(16) SphereMaterial {
    double diffuse_gain, specular_gain, roughness, irid_gain;
    double r, g, b, checker;
};
__device__ __forceinline__ SphereAux load_aux(const Staged& st, int s) {
    SphereAux a;
    lds_f64x2(st.aux + (unsigned int)s * 16u, a.inv_r, a.cc);
    return a;
}

// ---------------------------------------------------------------------------------------
// NumpySphere.intersect (shape.py:28-51)
// ---------------------------------------------------------------------------------------

// Exact: the reference's expression, term by term.
__device__ __forceinline__ double sphere_distance_exact(Vec3 o, Vec3 d, const SphereGeom& g, double cc) {
    Vec3 c = {g.cx, g.cy, g.cz};
    double b = 2.0 * dot(d, o - c);
    double cterm = cc + dot(o, o) - 2.0 * dot(c, o) - g.rr;
    double disc = (b * b) - (4.0 * cterm);
    double sq = sqrt(fmax(0.0, disc));
    double h0 = (-b - sq) / 2.0;
    double h1 = (-b + sq) / 2.0;
    double h = (h0 > 0.0 && h0 < h1) ? h0 : h1;
    return (disc > 0.0 && h > 0.0) ? h : kMiss;
}

// Fast: half-b form.  With b' = D.(O-C) and c = |O-C|^2 - r^2 the reference's roots are
// -b' -+ sqrt(b'^2 - c).  The far root is positive iff b' < 0 or c < 0, so spheres that
// cannot be hit are rejected before the square root.
struct HalfB {
    double b, c, disc;
};
__device__ __forceinline__ HalfB half_b(Vec3 o, Vec3 d, const SphereGeom& g) {
    double lx = o.x - g.cx, ly = o.y - g.cy, lz = o.z - g.cz;
    HalfB q;
    q.b = fma(d.z, lz, fma(d.y, ly, d.x * lx));
    q.c = fma(lz, lz, fma(ly, ly, fma(lx, lx, -g.rr)));
    q.disc = fma(q.b, q.b, -q.c);
    return q;
}
// The sign and "is positive" tests read the high word of the double on the integer pipe
// instead of spending FP64-pipe slots on DSETP: hi < 0 <=> sign bit set; hi > 0 <=> the
// value is positive and above 2^-1022 * 2^20 (smaller discriminants/roots count as a miss;
// the reference's own expression carries ~1e-6 absolute noise there).
__device__ __forceinline__ bool is_neg(double v) { return __double2hiint(v) < 0; }
__device__ __forceinline__ bool is_pos(double v) { return __double2hiint(v) > 0; }
__device__ __forceinline__ bool can_hit(const HalfB& q) {
    return ((__double2hiint(q.b) | __double2hiint(q.c)) < 0) && is_pos(q.disc);
}
__device__ __forceinline__ double near_root(const HalfB& q) {
    double sq = sqrt_pos_fast(q.disc);  // callers checked disc > 0
    double h0 = -q.b - sq;
    double h = is_pos(h0) ? h0 : sq - q.b;
    return is_pos(h) ? h : kMiss;
}

// ---------------------------------------------------------------------------------------
// NumpyShader._calculate_shadow (shader.py:114-128): lit iff no other sphere is closer
// along the shadow ray than the point's own sphere (FARAWAY when the ray leaves it).
// ---------------------------------------------------------------------------------------
// min of two positive doubles (no NaN plumbing): positive doubles order like their bits
__device__ __forceinline__ double fmin_pos(double a, double b) {
    return __double_as_longlong(b) < __double_as_longlong(a) ? b : a;
}
__device__ __forceinline__ bool in_light_exact(const Staged& st, int n, int s, Vec3 p, Vec3 l, unsigned long long& tests) {
    tests += (unsigned long long)n;
    double own = sphere_distance_exact(p, l, load_geom(st, s), load_aux(st, s).cc);
    bool lit = true;
    for (int j = 0; j < n; ++j) {
        double u = sphere_distance_exact(p, l, load_geom(st, j), load_aux(st, j).cc);
        lit = lit && !(u < own);
    }
    return lit;
}

// ---------------------------------------------------------------------------------------
// Material fetch (one hit): read-only path, 8 doubles of column s.
// ---------------------------------------------------------------------------------------
struct Material {
    double diffuse_gain, specular_gain, roughness, irid_gain;
    double r, g, b;
    bool checker;
};
// from the global SoA (staging only)
__device__ __forceinline__ Material load_material(const SceneParams& sp, int s) {
    const double* f = sp.spheres;
    const size_t n = (size_t)sp.n_spheres;
    Material m;
    m.diffuse_gain = __ldg(f + RT_F_DIFFUSE_GAIN * n + s);
    m.specular_gain = __ldg(f + RT_F_SPECULAR_GAIN * n + s);
    m.roughness = __ldg(f + RT_F_SPECULAR_ROUGHNESS * n + s);
    m.irid_gain = __ldg(f + RT_F_IRIDESCENCE_GAIN * n + s);
    m.r = __ldg(f + RT_F_COLOR_R * n + s);
    m.g = __ldg(f + RT_F_COLOR_G * n + s);
    m.b = __ldg(f + RT_F_COLOR_B * n + s);
    m.checker = __ldg(f + RT_F_TEXTURE * n + s) != 0.0;
    return m;
}
// from shared memory (one shaded hit): four broadcast 16-byte loads
__device__ __forceinline__ Material load_material(const Staged& st, int s) {
    Material m;
    double checker;
    const unsigned int a = st.mat + (unsigned int)s * 64u;
    lds_f64x2(a, m.diffuse_gain, m.specular_gain);
    lds_f64x2(a + 16u, m.roughness, m.irid_gain);
    lds_f64x2(a + 32u, m.r, m.g);
    lds_f64x2(a + 48u, m.b, checker);
    m.checker = checker != 0.0;
    return m;
}
__device__ __forceinline__ Vec3 load_light(const Staged& st, int li) {
    Vec3 p;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(p.x) : "r"(st.lights + (unsigned int)li * 24u));
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(p.y) : "r"(st.lights + (unsigned int)li * 24u + 8u));
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(p.z) : "r"(st.lights + (unsigned int)li * 24u + 16u));
    return p;
}

// TextureChecker.get_color (shader.py:29-32): astype(int) truncates, % 2 is non-negative.
__device__ __forceinline__ bool checker_on(Vec3 m) {
    long long ix = (long long)(m.x * 2.0);
    long long iz = (long long)(m.z * 2.0);
    return ((ix ^ iz) & 1LL) == 0;
}

// ---------------------------------------------------------------------------------------
// NumpyShader._calculate_physical_specular (shader.py:246-320), grey value.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ double specular_exact(Vec3 n, Vec3 to_light, Vec3 to_eye, double rough,
                                                 double gain) {
    Vec3 l = norm_exact(to_light);
    Vec3 v = norm_exact(to_eye);
    Vec3 h = norm_exact(l + v);
    double ndv = clip01(dot(n, v));
    double ndh = clip01(dot(n, h));
    double vdh = clip01(dot(v, h));
    double ndl = clip01(dot(n, l));
    double fres = kF0 + (1.0 - kF0) * pow(1.0 - vdh, 5.0);
    double alpha = rough * rough;
    double a2 = alpha * alpha;
    double den = (ndh * ndh) * (a2 - 1.0) + 1.0;
    double dist = a2 / (kPi * ((den * den) + kEps));
    double g1l = 2.0 * ndl / (ndl + sqrt(a2 + (1.0 - a2) * (ndl * ndl)) + kEps);
    double g1v = 2.0 * ndv / (ndv + sqrt(a2 + (1.0 - a2) * (ndv * ndv)) + kEps);
    double geo = g1l * g1v;
    double base = (fres * dist * geo) / (4.0 * ndv + kEps);
    double glint = pow(1.0 - ndv, 2.5);
    glint = glint * ndl;
    double total = base + gain * glint;
    return ndv <= 0.0 ? 0.0 : total;
}

__device__ __forceinline__ float sqrt_approx(float x) {
    float y;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// ---------------------------------------------------------------------------------------
// NumpyShader._calculate_physical_iridescence (shader.py:186-232), before the gain.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ Vec3 film_color_exact(Vec3 n, Vec3 to_eye) {
    double va = clip01(dot(n, to_eye));
    double af = fabs(va - 0.5) * 2.0;
    double phase = af * kPi * kFilmThickness * 10.0;
    double ip = sin(phase);
    double r = ip * kHueShift + (1.0 - kHueShift) * (1.0 - ip);
    double g = ip * (1.0 - kHueShift) + kHueShift * (1.0 - ip);
    double b = 0.5 + 0.5 * ip;
    return {r, g, b};
}

}  // namespace rt

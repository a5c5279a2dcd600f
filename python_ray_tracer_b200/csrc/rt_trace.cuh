// rt_trace.cuh -- intersection, shadow test, shading and the bounce chain (device).
// See rt_device.cuh for the two arithmetic policies.
#pragma once

#include "rt_device.cuh"

namespace rt {

// Shared-memory view of the staged scene (filled by stage_scene()).
// Per-sphere records staged in shared memory are addressed through 32-bit shared-window
// addresses and read with explicit ld.shared.v2.f64: every lane reads the same record
// (one broadcast wavefront), the address is base + index * 32 with no per-use
// generic-to-shared conversion.
struct __align__(16) SpherePrimary {
    double lx, ly, lz;  // origin0 - centre
    double c;           // |origin0 - centre|^2 - r^2
};
struct Staged {
    unsigned int geom;   // SphereGeom[n]
    unsigned int prim;   // SpherePrimary[n]: constants of rays that start at the camera
    unsigned int aux;    // SphereAux[n]
    unsigned int mat;    // SphereMaterial[n]
    unsigned int lights; // double[3 * n_lights_used]
    double chain_bound;  // Fast: sup of |local colour| over materials / (1 - sup w), or < 0
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
__device__ __forceinline__ SpherePrimary load_prim(const Staged& st, int s) {
    SpherePrimary p;
    lds_f64x2(st.prim + (unsigned int)s * 32u, p.lx, p.ly);
    lds_f64x2(st.prim + (unsigned int)s * 32u + 16u, p.lz, p.c);
    return p;
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
// Camera rays all start at the eye: O-C and |O-C|^2 - r^2 are per-sphere constants of the
// launch (staged once per block), leaving one dot product and one FMA per test.
__device__ __forceinline__ HalfB half_b_primary(Vec3 d, const SpherePrimary& p) {
    HalfB q;
    q.b = fma(d.z, p.lz, fma(d.y, p.ly, d.x * p.lx));
    q.c = p.c;
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
template <int MODE>
__device__ __forceinline__ bool in_light(const Staged& st, int n, int s, Vec3 p, Vec3 l,
                                         unsigned long long& tests) {
    tests += (unsigned long long)n;
    if (MODE == kModeExact) {
        double own = sphere_distance_exact(p, l, load_geom(st, s), load_aux(st, s).cc);
        bool lit = true;
        for (int j = 0; j < n; ++j) {
            double u = sphere_distance_exact(p, l, load_geom(st, j), load_aux(st, j).cc);
            lit = lit && !(u < own);
        }
        return lit;
    } else {
        // nearest hit among the OTHER spheres first: in most shadow tests there is none and
        // the point's own sphere never has to be intersected.  Two spheres per trip.
        double other = kMiss;
        int j = 0;
        for (; j + 1 < n; j += 2) {
            HalfB q0 = half_b(p, l, load_geom(st, j));
            HalfB q1 = half_b(p, l, load_geom(st, j + 1));
            if (j != s && can_hit(q0)) other = fmin_pos(other, near_root(q0));
            if (j + 1 != s && can_hit(q1)) other = fmin_pos(other, near_root(q1));
        }
        if (j < n && j != s) {
            HalfB q = half_b(p, l, load_geom(st, j));
            if (can_hit(q)) other = fmin_pos(other, near_root(q));
        }
        if (other == kMiss) return true;
        HalfB qo = half_b(p, l, load_geom(st, s));
        const double own = can_hit(qo) ? near_root(qo) : kMiss;
        return !(other < own);
        return true;
    }
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

// Fast: l and v arrive normalised (re-normalising a unit vector changes the last bit
// only).  The cancellation-prone quantities are formed in FP64 -- the cosines and
// den = NdotH^2 (a2 - 1) + 1 written as (|L+V|^2 - (N.(L+V))^2 (1 - a2)) / |L+V|^2, whose
// numerator needs no square root -- and then handed to single precision, where Fresnel,
// the GGX distribution, both Smith terms, their common quotient and the glint are smooth
// positive products: a relative error of ~1e-6 in the result, i.e. <= 1e-6 absolute where
// the term is below 1 and 1e-6 relative where the distribution peaks (the parity bar is
// 1e-4 on both readings).  Roots and reciprocals are single MUFU instructions.
__device__ __forceinline__ float sqrt_approx(float x) {
    float y;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ double specular_fast(Vec3 n, Vec3 l, Vec3 v, double rough, double gain) {
    const double ndv_d = fdot(n, v);
    if (__double2hiint(ndv_d) <= 0) return 0.0;  // NdotV <= 0 (or denormal)
    const Vec3 hv = l + v;
    const double h2_d = fdot(hv, hv);
    const double nh_d = fdot(n, hv), vh_d = fdot(v, hv);
    const double alpha_d = rough * rough;
    const double a2_d = alpha_d * alpha_d;
    // |L+V|^2 * den; NdotH <= 0 clips to den = 1
    const double num_d = __double2hiint(nh_d) < 0 ? h2_d : fma(-(nh_d * nh_d), 1.0 - a2_d, h2_d);

    const float ndv = fminf((float)ndv_d, 1.0f);
    const float ndl = fminf(fmaxf((float)fdot(n, l), 0.0f), 1.0f);
    const float h2 = (float)h2_d, a2 = (float)a2_d, oma2 = 1.0f - a2;
    const bool flat = !(h2 > 1e-30f);  // L = -V: the reference's norm() leaves a zero half vector
    const float ih2 = flat ? 0.0f : __frcp_rn(h2);
    const float den = flat ? 1.0f : (float)num_d * ih2;
    const float vdh = flat ? 0.0f : fminf(fmaxf((float)vh_d * rsqrtf(h2), 0.0f), 1.0f);
    const float eps = (float)kEps;
    const float omv = 1.0f - vdh, omv2 = omv * omv;
    const float fres = fmaf(1.0f - (float)kF0, omv2 * omv2 * omv, (float)kF0);
    const float sl = sqrt_approx(fmaf(oma2, ndl * ndl, a2));
    const float sv = sqrt_approx(fmaf(oma2, ndv * ndv, a2));
    const float num = fres * a2 * (4.0f * ndl * ndv);
    const float dn = ((float)kPi * fmaf(den, den, eps)) * ((ndl + sl + eps) * (ndv + sv + eps)) * fmaf(4.0f, ndv, eps);
    const float base = __fdividef(num, dn);
    const float x = 1.0f - ndv;
    const float glint = (x * x) * sqrt_approx(x) * ndl;
    return (double)fmaf((float)gain, glint, base);
}

// ---------------------------------------------------------------------------------------
// NumpyShader._calculate_physical_iridescence (shader.py:186-232), before the gain.
// ---------------------------------------------------------------------------------------
template <int MODE>
__device__ __forceinline__ Vec3 film_color(Vec3 n, Vec3 to_eye) {
    if (MODE == kModeExact) {
        double va = clip01(dot(n, to_eye));
        double af = fabs(va - 0.5) * 2.0;
        double phase = af * kPi * kFilmThickness * 10.0;
        double ip = sin(phase);
        double r = ip * kHueShift + (1.0 - kHueShift) * (1.0 - ip);
        double g = ip * (1.0 - kHueShift) + kHueShift * (1.0 - ip);
        double b = 0.5 + 0.5 * ip;
        return {r, g, b};
    }
    // Fast: the tint is a smooth function of N.V, bounded by 1 and scaled by
    // 0.1 * iridescence_gain afterwards, so single precision (abs. error < 1e-6 before the
    // scale) is far inside the 1e-4 budget.  phase = af * 3*pi: reduce to one period in
    // turns, then the hardware sine on [-pi, pi].
    float va = (float)pos_part(fdot(n, to_eye));  // <= 1 + a few ulp: the upper clip is moot
    float turns = fabsf(va - 0.5f) * 3.0f;  // phase / (2 pi), in [0, 1.5]
    turns -= rintf(turns);                    // [-0.5, 0.5]
    float ip = __sinf(turns * 6.283185307179586f);
    const float hs = (float)kHueShift;
    float r = fmaf(ip, hs, (1.0f - hs) * (1.0f - ip));
    float g = fmaf(ip, 1.0f - hs, hs * (1.0f - ip));
    float b = fmaf(0.5f, ip, 0.5f);
    return {(double)r, (double)g, (double)b};
}

}  // namespace rt

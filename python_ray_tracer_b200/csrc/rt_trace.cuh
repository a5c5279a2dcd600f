// rt_trace.cuh -- intersection, shadow test, shading and the bounce chain (device).
// See rt_device.cuh for the two arithmetic policies.
#pragma once

#include "rt_device.cuh"

namespace rt {

// Shared-memory view of the staged scene (filled by stage_scene()).
struct Staged {
    const SphereGeom* geom;
    const SphereAux* aux;
    double chain_bound;  // Fast: sup of |local colour| over materials / (1 - sup w), or < 0
};

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
__device__ __forceinline__ bool can_hit(const HalfB& q) { return q.disc > 0.0 && (q.b < 0.0 || q.c < 0.0); }
__device__ __forceinline__ double near_root(const HalfB& q) {
    double sq = sqrt(q.disc);
    double h0 = -q.b - sq;
    double h = h0 > 0.0 ? h0 : sq - q.b;
    return h > 0.0 ? h : kMiss;
}

// ---------------------------------------------------------------------------------------
// NumpyShader._calculate_shadow (shader.py:114-128): lit iff no other sphere is closer
// along the shadow ray than the point's own sphere (FARAWAY when the ray leaves it).
// ---------------------------------------------------------------------------------------
template <int MODE>
__device__ __forceinline__ bool in_light(const Staged& st, int n, int s, Vec3 p, Vec3 l,
                                         unsigned long long& tests) {
    tests += (unsigned long long)n;
    if (MODE == kModeExact) {
        double own = sphere_distance_exact(p, l, st.geom[s], st.aux[s].cc);
        bool lit = true;
        for (int j = 0; j < n; ++j) {
            double u = sphere_distance_exact(p, l, st.geom[j], st.aux[j].cc);
            lit = lit && !(u < own);
        }
        return lit;
    } else {
        HalfB qo = half_b(p, l, st.geom[s]);
        double own = can_hit(qo) ? near_root(qo) : kMiss;
        for (int j = 0; j < n; ++j) {
            if (j == s) continue;
            HalfB q = half_b(p, l, st.geom[j]);
            if (can_hit(q)) {
                if (own == kMiss) {
                    // any positive root occludes; near_root() > 0 is implied by can_hit()
                    // except when rounding puts the root at <= 0, so confirm it.
                    if (near_root(q) != kMiss) return false;
                } else if (near_root(q) < own) {
                    return false;
                }
            }
        }
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
// only); the four quotients F*D*G1*G1/(4 NdotV + eps) share one division; x^5 and x^2.5
// are products and one square root.
__device__ __forceinline__ double specular_fast(Vec3 n, Vec3 l, Vec3 v, double rough, double gain) {
    double ndv = clip01(fdot(n, v));
    if (ndv <= 0.0) return 0.0;
    Vec3 hv = l + v;
    double h2 = fdot(hv, hv);
    double ih = h2 > 0.0 ? rsqrt(h2) : 1.0;
    double ndh = clip01(fdot(n, hv) * ih);
    double vdh = clip01(fdot(v, hv) * ih);
    double ndl = clip01(fdot(n, l));
    double omv = 1.0 - vdh;
    double omv2 = omv * omv;
    double fres = fma(1.0 - kF0, omv2 * omv2 * omv, kF0);
    double alpha = rough * rough;
    double a2 = alpha * alpha;
    double oma2 = 1.0 - a2;
    double den = fma(ndh * ndh, a2 - 1.0, 1.0);
    double sl = sqrt(fma(oma2, ndl * ndl, a2));
    double sv = sqrt(fma(oma2, ndv * ndv, a2));
    double num = fres * a2 * (4.0 * ndl * ndv);
    double dn = kPi * fma(den, den, kEps) * (ndl + sl + kEps) * (ndv + sv + kEps) * fma(4.0, ndv, kEps);
    double base = num / dn;
    double x = 1.0 - ndv;
    double glint = x * x * sqrt(x) * ndl;
    return fma(gain, glint, base);
}

// ---------------------------------------------------------------------------------------
// NumpyShader._calculate_physical_iridescence (shader.py:186-232), before the gain.
// ---------------------------------------------------------------------------------------
template <int MODE>
__device__ __forceinline__ Vec3 film_color(Vec3 n, Vec3 to_eye) {
    double va = clip01(MODE == kModeExact ? dot(n, to_eye) : fdot(n, to_eye));
    double af = fabs(va - 0.5) * 2.0;
    double phase = af * kPi * kFilmThickness * 10.0;
    double ip = sin(phase);
    double r = ip * kHueShift + (1.0 - kHueShift) * (1.0 - ip);
    double g = ip * (1.0 - kHueShift) + kHueShift * (1.0 - ip);
    double b = 0.5 + 0.5 * ip;
    return {r, g, b};
}

}  // namespace rt

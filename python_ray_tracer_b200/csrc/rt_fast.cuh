// rt_fast.cuh -- the fast arithmetic policy: one generation of a bounce chain (device).
//
// Same path as rt_chain.cuh's exact body -- nearest hit (base.py:97-105), NumpyShader.create
// (shader.py:63-112), shadow test (shader.py:114-128), mirror ray (shader.py:143-161) -- with the
// arithmetic arranged for the FP64 pipe of sm_100a, which issues one warp instruction every two
// cycles and returns a result ~8 cycles later:
//
//   * Nothing is normalised on the way to a geometric decision.  A ray/sphere test is invariant under
//     the length of the direction if the discriminant is written b^2 - |d|^2 c, so camera rays are
//     tested as the unnormalised P - eye (the reciprocal length, needed for the hit point and the view
//     vector, is formed beside the tests, not before them), shadow rays as the unnormalised L - M, and
//     the mirror direction D - 2(D.N)N of a unit D is used as it is (|N| = 1 to rounding, so its
//     length drifts by parts in 1e16 per bounce and nothing downstream depends on it).
//   * The shadow test never intersects the point's own sphere unless it has to: the nudged point is
//     outside its sphere, so that sphere is met again only when the light is behind the surface
//     (N.L < 0), and it matters only if another sphere is met too.  "No other sphere is met" -- the
//     common case -- needs no square root at all.
//   * Everything after the cosines is single precision: Lambert, dome, iridescence, the colour sums
//     and the bounce weight (relative error ~1e-7 against a 1e-4 bar).  The GGX term keeps its
//     FP64-formed inputs (the denominator NdotH^2 (a2 - 1) + 1 cancels to ~a2 at the highlight).
//   * Per-material constants (diffuse colour * gain, roughness^4, 0.1 * iridescence gain,
//     0.5 * specular gain) are formed once per scene (ShadeRec), not per hit.
//
// Three scene policies feed the same body, so their pictures agree bit for bit:
//   SmallScene<N>  <= kSmallMax spheres as a kernel parameter: constant-bank operands, loops unrolled
//   StagedScene    any table that fits shared memory: the loops of round 1, two spheres per trip
//   TreeScene      BVH blob in global memory (rt_accel.cuh): the tree picks which tests run
#pragma once

#include "rt_accel.cuh"

// SmallScene: leave nearest() on one branch when no sphere can be hit (measured on B200, README 1080p:
// 0.0540 ms with, 0.0558 without; profiles/r02/sweep3_epilogue.txt)
#ifndef RT_COMBINED_MISS
#define RT_COMBINED_MISS 1
#endif

namespace rt {

struct Counters {
    unsigned long long traced, hits, deepest, tests;
};

__device__ __forceinline__ double rsqrt_approx_d(double x) {  // MUFU.RSQ64H, rel. error <= 2^-22
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    return y;
}
__device__ __forceinline__ float rcp_approx_f(float x) {
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// ---- ray/sphere tests ---------------------------------------------------------------------------
// A direction of any length: a = |d|^2.  near_root() of the result is |d|^2 times the ray parameter.
__device__ __forceinline__ HalfB half_b_scaled(Vec3 o, Vec3 d, double a, const SphereGeom& g) {
    const double lx = o.x - g.cx, ly = o.y - g.cy, lz = o.z - g.cz;
    HalfB q;
    q.b = fma(d.z, lz, fma(d.y, ly, d.x * lx));
    q.c = fma(lz, lz, fma(ly, ly, fma(lx, lx, -g.rr)));
    q.disc = fma(q.b, q.b, -(a * q.c));
    return q;
}
// Camera rays all start at the eye: O - C and |O - C|^2 - r^2 are constants of the launch.
__device__ __forceinline__ HalfB half_b_primary(Vec3 q, double a, const SpherePrimary& p) {
    HalfB h;
    h.b = fma(q.z, p.lz, fma(q.y, p.ly, q.x * p.lx));
    h.c = p.c;
    h.disc = fma(h.b, h.b, -(a * p.c));
    return h;
}
__device__ __forceinline__ SpherePrimary primary_of(Vec3 eye, const SphereGeom& g) {  // = make_sphere_primary()
    const double lx = eye.x - g.cx, ly = eye.y - g.cy, lz = eye.z - g.cz;
    return {lx, ly, lz, fma(lz, lz, fma(ly, ly, fma(lx, lx, -g.rr)))};
}
__device__ __forceinline__ void consider(double t, int s, double& tmin, int& sid) {
    if (t < tmin) {
        tmin = t;
        sid = s;
    }
}

__device__ __forceinline__ ShadeRec lds_rec(unsigned int addr) {
    ShadeRec m;
    lds_f64x2(addr, m.cx, m.cy);
    lds_f64x2(addr + 16u, m.cz, m.inv_r);
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(m.kd_r), "=f"(m.kd_g), "=f"(m.kd_b), "=f"(m.spec_gain) : "r"(addr + 32u));
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=f"(m.a2), "=f"(m.film), "=f"(m.half_gain), "=r"(m.flags) : "r"(addr + 48u));
    return m;
}
__device__ __forceinline__ Vec3 lds_vec3(unsigned int addr) {
    Vec3 p;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(p.x) : "r"(addr));
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(p.y) : "r"(addr + 8u));
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(p.z) : "r"(addr + 16u));
    return p;
}

// ---- scene policies -----------------------------------------------------------------------------
template <int N>
struct SmallScene {
    const SmallTable* t;  // the kernel parameter: compile-time indices become constant-bank operands
    unsigned int recs;    // shared-memory copy of t->rec, indexed by the sphere that was hit
    unsigned int geoms;   // and of t->geom, indexed by "the other spheres" of the shadow test
    Vec3 light0;
    float bound;

    __device__ __forceinline__ int n_lights() const { return 1; }
    __device__ __forceinline__ Vec3 light(int) const { return light0; }
    __device__ __forceinline__ float chain_bound() const { return bound; }
    __device__ __forceinline__ ShadeRec rec(int s) const { return lds_rec(recs + (unsigned int)s * 64u); }

    // all N tests first (independent FP64 chains interleave), then the roots of those that can be hit
    __device__ __forceinline__ void nearest_primary(Vec3, Vec3 q, double a, double& tmin, int& sid, unsigned long long& tests) const {
        tests += N;
        HalfB h[N];
#pragma unroll
        for (int s = 0; s < N; ++s) h[s] = half_b_primary(q, a, t->prim[s]);
        bool may[N], any = false;
#pragma unroll
        for (int s = 0; s < N; ++s) any |= (may[s] = can_hit(h[s]));
        if (RT_COMBINED_MISS && !any) return;  // most rays of a sparse scene: one branch instead of N
#pragma unroll
        for (int s = 0; s < N; ++s)
            if (may[s]) consider(near_root(h[s]), s, tmin, sid);
    }
    __device__ __forceinline__ void nearest(Vec3 o, Vec3 d, double& tmin, int& sid, unsigned long long& tests) const {
        tests += N;
        HalfB h[N];
#pragma unroll
        for (int s = 0; s < N; ++s) h[s] = half_b(o, d, t->geom[s]);
        bool may[N], any = false;
#pragma unroll
        for (int s = 0; s < N; ++s) any |= (may[s] = can_hit(h[s]));
        if (RT_COMBINED_MISS && !any) return;
#pragma unroll
        for (int s = 0; s < N; ++s)
            if (may[s]) consider(near_root(h[s]), s, tmin, sid);
    }
    // is any sphere but `own` met by the ray p + u*lv, u > 0?  (no roots needed)  The N - 1 others are
    // own + 1, own + 2, ... (mod N), read from the shared-memory copy: straight-line, nothing wasted on `own`.
    __device__ __forceinline__ bool others_hit(Vec3 p, Vec3 lv, double ll, int own, unsigned long long& tests) const {
        tests += N;
        bool any = false;
        int s = own;
#pragma unroll
        for (int j = 1; j < N; ++j) {
            s = s + 1 == N ? 0 : s + 1;
            SphereGeom g;
            lds_f64x2(geoms + (unsigned int)s * 32u, g.cx, g.cy);
            lds_f64x2(geoms + (unsigned int)s * 32u + 16u, g.cz, g.rr);
            any |= can_hit(half_b_scaled(p, lv, ll, g));
        }
        return any;
    }
    // the whole of shader.py:126-128: lit iff no other sphere is nearer than the point's own
    __device__ __forceinline__ bool lit_slow(Vec3 p, Vec3 lv, double ll, int own, unsigned long long&) const {
        double mine = kMiss, other = kMiss;
#pragma unroll
        for (int s = 0; s < N; ++s) {
            const HalfB h = half_b_scaled(p, lv, ll, t->geom[s]);
            if (can_hit(h)) {
                const double u = near_root(h);
                if (s == own) mine = u;
                else other = fmin_pos(other, u);
            }
        }
        return !(other < mine);
    }
};

struct StagedScene {
    unsigned int geom, prim, recs, lights;  // shared-window addresses: SphereGeom[n], SpherePrimary[n], ShadeRec[n], double[3 n_lights]
    int n, lights_used;
    float bound;

    __device__ __forceinline__ int n_lights() const { return lights_used; }
    __device__ __forceinline__ Vec3 light(int li) const { return lds_vec3(lights + (unsigned int)li * 24u); }
    __device__ __forceinline__ float chain_bound() const { return bound; }
    __device__ __forceinline__ ShadeRec rec(int s) const { return lds_rec(recs + (unsigned int)s * 64u); }
    __device__ __forceinline__ SphereGeom g(int s) const {
        SphereGeom v;
        lds_f64x2(geom + (unsigned int)s * 32u, v.cx, v.cy);
        lds_f64x2(geom + (unsigned int)s * 32u + 16u, v.cz, v.rr);
        return v;
    }
    __device__ __forceinline__ SpherePrimary p0(int s) const {
        SpherePrimary v;
        lds_f64x2(prim + (unsigned int)s * 32u, v.lx, v.ly);
        lds_f64x2(prim + (unsigned int)s * 32u + 16u, v.lz, v.c);
        return v;
    }

    __device__ __forceinline__ void nearest_primary(Vec3, Vec3 q, double a, double& tmin, int& sid, unsigned long long& tests) const {
        tests += (unsigned long long)n;
        int s = 0;
        for (; s + 1 < n; s += 2) {  // two spheres per trip: their FP64 chains interleave
            const HalfB h0 = half_b_primary(q, a, p0(s)), h1 = half_b_primary(q, a, p0(s + 1));
            if (can_hit(h0)) consider(near_root(h0), s, tmin, sid);
            if (can_hit(h1)) consider(near_root(h1), s + 1, tmin, sid);
        }
        if (s < n) {
            const HalfB h = half_b_primary(q, a, p0(s));
            if (can_hit(h)) consider(near_root(h), s, tmin, sid);
        }
    }
    __device__ __forceinline__ void nearest(Vec3 o, Vec3 d, double& tmin, int& sid, unsigned long long& tests) const {
        tests += (unsigned long long)n;
        int s = 0;
        for (; s + 1 < n; s += 2) {
            const HalfB h0 = half_b(o, d, g(s)), h1 = half_b(o, d, g(s + 1));
            if (can_hit(h0)) consider(near_root(h0), s, tmin, sid);
            if (can_hit(h1)) consider(near_root(h1), s + 1, tmin, sid);
        }
        if (s < n) {
            const HalfB h = half_b(o, d, g(s));
            if (can_hit(h)) consider(near_root(h), s, tmin, sid);
        }
    }
    __device__ __forceinline__ bool others_hit(Vec3 p, Vec3 lv, double ll, int own, unsigned long long& tests) const {
        tests += (unsigned long long)n;
        bool any = false;
        int s = 0;
        for (; s + 1 < n; s += 2) {
            const HalfB h0 = half_b_scaled(p, lv, ll, g(s)), h1 = half_b_scaled(p, lv, ll, g(s + 1));
            any = any || (s != own && can_hit(h0)) || (s + 1 != own && can_hit(h1));
            if (any) return true;
        }
        if (s < n && s != own) any = can_hit(half_b_scaled(p, lv, ll, g(s)));
        return any;
    }
    __device__ __forceinline__ bool lit_slow(Vec3 p, Vec3 lv, double ll, int own, unsigned long long&) const {
        const HalfB ho = half_b_scaled(p, lv, ll, g(own));
        const double mine = can_hit(ho) ? near_root(ho) : kMiss;
        for (int s = 0; s < n; ++s) {
            if (s == own) continue;
            const HalfB h = half_b_scaled(p, lv, ll, g(s));
            if (can_hit(h) && near_root(h) < mine) return false;
        }
        return true;
    }
};

struct TreeScene {
    AccelView av;
    unsigned int lights;
    int lights_used;
    float bound;

    __device__ __forceinline__ int n_lights() const { return lights_used; }
    __device__ __forceinline__ Vec3 light(int li) const { return lds_vec3(lights + (unsigned int)li * 24u); }
    __device__ __forceinline__ float chain_bound() const { return bound; }
    __device__ __forceinline__ ShadeRec rec(int s) const {
        const double2 c0 = __ldg(reinterpret_cast<const double2*>(av.rec + s));
        const double2 c1 = __ldg(reinterpret_cast<const double2*>(av.rec + s) + 1);
        const float4 f0 = __ldg(reinterpret_cast<const float4*>(av.rec + s) + 2);
        const float4 f1 = __ldg(reinterpret_cast<const float4*>(av.rec + s) + 3);
        ShadeRec m;
        m.cx = c0.x, m.cy = c0.y, m.cz = c1.x, m.inv_r = c1.y;
        m.kd_r = f0.x, m.kd_g = f0.y, m.kd_b = f0.z, m.spec_gain = f0.w;
        m.a2 = f1.x, m.film = f1.y, m.half_gain = f1.z, m.flags = __float_as_uint(f1.w);
        return m;
    }
    __device__ __forceinline__ void nearest_primary(Vec3 eye, Vec3 q, double a, double& tmin, int& sid, unsigned long long& tests) const {
        // slab distances are ray parameters, roots are |q|^2 times that: scale the limit (rounded up)
        const float inv_a = __frcp_ru(__double2float_rd(a));
        accel_nearest(av, eye, q, inv_a, tmin, sid, tests, [&](const SphereGeom& g, int) {
            const HalfB h = half_b_primary(q, a, primary_of(eye, g));
            return can_hit(h) ? near_root(h) : kMiss;
        });
    }
    __device__ __forceinline__ void nearest(Vec3 o, Vec3 d, double& tmin, int& sid, unsigned long long& tests) const {
        accel_nearest(av, o, d, 1.0f, tmin, sid, tests, [&](const SphereGeom& g, int) {
            const HalfB h = half_b(o, d, g);
            return can_hit(h) ? near_root(h) : kMiss;
        });
    }
    __device__ __forceinline__ bool others_hit(Vec3 p, Vec3 lv, double ll, int own, unsigned long long& tests) const {
        return accel_occluded(av, own, p, lv, kMiss, 1.0f, tests, [&](const SphereGeom& g, int) {
            return can_hit(half_b_scaled(p, lv, ll, g)) ? 0.0 : kMiss;
        });
    }
    __device__ __forceinline__ bool lit_slow(Vec3 p, Vec3 lv, double ll, int own, unsigned long long& tests) const {
        const HalfB ho = half_b_scaled(p, lv, ll, ldg_geom(av.geom + own));
        const double mine = can_hit(ho) ? near_root(ho) : kMiss;
        tests += 1;
        const float inv_a = __frcp_ru(__double2float_rd(ll));
        return !accel_occluded(av, own, p, lv, mine, inv_a, tests, [&](const SphereGeom& g, int) {
            const HalfB h = half_b_scaled(p, lv, ll, g);
            return can_hit(h) ? near_root(h) : kMiss;
        });
    }
};

// ---- NumpyShader._calculate_physical_specular (shader.py:246-320), grey value before the gain ----
// Inputs are dot products of the UNNORMALISED light and view vectors with N and with each other.
// With l = lv/|lv| and v = ev/|ev| unit vectors, H = l + v has |H|^2 = 2 + 2 l.v, N.H = N.l + N.v,
// V.H = 1 + l.v: no half vector is formed.  The cancellation-prone den = NdotH^2 (a2 - 1) + 1 is
// written (|H|^2 - (N.H)^2 (1 - a2)) / |H|^2 with the numerator in FP64; Fresnel, the GGX distribution,
// both Smith terms, their common quotient and the glint are smooth positive FP32 products with
// single-instruction MUFU roots and reciprocals (the tail round 1 validated against every fixture).
// Only called with N.l > 0 and N.v > 0 (otherwise the term is zero), so H is never the zero vector.
__device__ __forceinline__ float specular_tail(const ShadeRec& m, float ndl_raw, float ndv_raw, float h2, float num, float vh) {
    const float a2 = m.a2, oma2 = 1.0f - a2;
    const float ndl = fminf(ndl_raw, 1.0f), ndv = fminf(ndv_raw, 1.0f);
    const float ih2 = rcp_approx_f(h2);
    const float den = fmaxf(num * ih2, a2);  // NdotH clipped at 1 (N is not re-normalised, shader.py:74)
    const float vdh = fminf(fmaxf(vh * rsqrtf(h2), 0.0f), 1.0f);
    const float eps = (float)kEps;
    const float omv = 1.0f - vdh, omv2 = omv * omv;
    const float fres = fmaf(1.0f - (float)kF0, omv2 * omv2 * omv, (float)kF0);
    const float sl = sqrt_approx(fmaf(oma2, ndl * ndl, a2));
    const float sv = sqrt_approx(fmaf(oma2, ndv * ndv, a2));
    const float top = fres * a2 * (4.0f * ndl * ndv);
    const float dn = ((float)kPi * fmaf(den, den, eps)) * ((ndl + sl + eps) * (ndv + sv + eps)) * fmaf(4.0f, ndv, eps);
    const float base = __fdividef(top, dn);
    const float x = 1.0f - ndv;
    const float glint = (x * x) * sqrt_approx(x) * ndl;
    return fmaf(m.spec_gain, glint, base);
}

#ifndef RT_SPEC_F32_A2
#define RT_SPEC_F32_A2 2.0f  // experiment: materials with roughness^4 >= this form |H|^2 and N.H in FP32 (2.0 = never)
#endif

// iridescence tint before the gain (shader.py:186-232) from va = clip(N.V, 0, 1): smooth, bounded by 1
__device__ __forceinline__ void film_tint(float va, float& r, float& g, float& b) {
    float turns = fabsf(va - 0.5f) * 3.0f;  // phase / (2 pi) = |va - .5| * 2 * pi * .3 * 10 / (2 pi), in [0, 1.5]
    turns -= rintf(turns);                    // [-0.5, 0.5]
    const float ip = __sinf(turns * 6.283185307179586f);
    const float hs = (float)kHueShift;
    r = fmaf(ip, hs, (1.0f - hs) * (1.0f - ip));
    g = fmaf(ip, 1.0f - hs, hs * (1.0f - ip));
    b = fmaf(0.5f, ip, 0.5f);
}

// One generation.  PRIMARY: the ray is o = eye, d = P - eye unnormalised, a = |d|^2; otherwise d is a
// unit vector (to rounding) and a is unused.  On `true` (o, d) is the mirror ray, unit length.
template <class Scene, bool PRIMARY, bool STATS>
__device__ __forceinline__ bool fast_generation(const Scene& sc, const SceneParams& sp, const int generation, Vec3& o, Vec3& d,
                                                const double a, float& cr, float& cg, float& cb, float& weight,
                                                unsigned long long& tests, Counters& k) {
    if (STATS) {
        k.traced += 1;
        k.deepest = generation + 1 > (int)k.deepest ? generation + 1 : k.deepest;
    }
    double tmin = kMiss;
    int sid = -1;
    if (PRIMARY) sc.nearest_primary(o, d, a, tmin, sid, tests);
    else sc.nearest(o, d, tmin, sid, tests);
    if (sid < 0) return false;
    if (STATS) k.hits += 1;

    const ShadeRec m = sc.rec(sid);
    double inv_len = 1.0;
    if (PRIMARY) {
        inv_len = rsqrt_fast(a);
        tmin *= inv_len * inv_len;
    }
    const Vec3 hit = {fma(d.x, tmin, o.x), fma(d.y, tmin, o.y), fma(d.z, tmin, o.z)};
    const Vec3 normal = {(hit.x - m.cx) * m.inv_r, (hit.y - m.cy) * m.inv_r, (hit.z - m.cz) * m.inv_r};
    const Vec3 nudged = {fma(normal.x, kNudge, hit.x), fma(normal.y, kNudge, hit.y), fma(normal.z, kNudge, hit.z)};
    // towards the camera, any positive length (shader.py:76: always the camera, also on bounces)
    Vec3 ev;
    double ee;
    if (PRIMARY) {
        ev = {-d.x, -d.y, -d.z};
        ee = a;
    } else {
        ev = {sp.eye[0] - hit.x, sp.eye[1] - hit.y, sp.eye[2] - hit.z};
        ee = fdot(ev, ev);
    }
    const double ne_d = fdot(normal, ev);
    const float ndv_raw = (float)(ne_d * rsqrt_approx_d(ee));  // N.V, unclipped

    float lr = (float)kAmbient, lg = (float)kAmbient, lb = (float)kAmbient;
    bool lit0 = false;
    float spec_sum = 0.0f;
    const int n_lights = sc.n_lights();
    for (int li = 0; li < n_lights; ++li) {
        const Vec3 lp = sc.light(li);
        const Vec3 lv = {lp.x - hit.x, lp.y - hit.y, lp.z - hit.z};
        const double ll = fdot(lv, lv);
        const double nl_d = fdot(normal, lv);
        bool lit = !sc.others_hit(nudged, lv, ll, sid, tests);
        // something else is in the way: the point is still lit if its own sphere comes first (shader.py:126-128),
        // which takes the light behind the surface (or an inward normal)
        if (!lit && (__double2hiint(nl_d) < 0 || (m.flags & kRecInward))) lit = sc.lit_slow(nudged, lv, ll, sid, tests);
        if (li == 0) lit0 = lit;
        if (lit && __double2hiint(nl_d) > 0) {
            const float ndl_raw = (float)(nl_d * rsqrt_approx_d(ll));
            float lam = ndl_raw;
            if (m.flags & kRecChecker) lam = checker_on(hit) ? ndl_raw : 0.0f;
            lr = fmaf(m.kd_r, lam, lr);
            lg = fmaf(m.kd_g, lam, lg);
            lb = fmaf(m.kd_b, lam, lb);
            if (m.spec_gain != 0.0f && __double2hiint(ne_d) > 0) {
                const double le_d = fdot(lv, ev);
                float h2, num, vh;
                if (m.a2 >= RT_SPEC_F32_A2) {
                    const float lvd = (float)le_d * rsqrtf((float)ll) * rsqrtf((float)ee);
                    const float nh = ndl_raw + ndv_raw;
                    h2 = fmaf(2.0f, lvd, 2.0f);
                    num = fmaf(-(nh * nh), 1.0f - m.a2, h2);
                    vh = 1.0f + lvd;
                } else {
                    const double il = rsqrt_fast(ll);
                    const double ie = PRIMARY ? inv_len : rsqrt_fast(ee);
                    const double lvd = le_d * (il * ie);
                    const double nh = fma(nl_d, il, ne_d * ie);
                    const double h2_d = fma(2.0, lvd, 2.0);
                    h2 = (float)h2_d;
                    num = (float)fma(-(nh * nh), 1.0 - (double)m.a2, h2_d);
                    vh = (float)(1.0 + lvd);
                }
                spec_sum += specular_tail(m, ndl_raw, ndv_raw, h2, num, vh);
            }
        }
    }

    // dome light (shader.py:234-244): colour of the last dome * sum of intensities * max(N.y, 0)
    const float ny = fmaxf((float)normal.y, 0.0f);
    lr = fmaf(sp.dome_rgb[0], ny, lr);
    lg = fmaf(sp.dome_rgb[1], ny, lg);
    lb = fmaf(sp.dome_rgb[2], ny, lb);
    // specular enters as spec * gain * lit (shader.py:106)
    const float spec_term = spec_sum * m.spec_gain;
    lr += spec_term;
    lg += spec_term;
    lb += spec_term;
    // iridescence (shader.py:110)
    if (m.film != 0.0f) {
        float fr, fg, fb;
        film_tint(fmaxf(ndv_raw, 0.0f), fr, fg, fb);
        lr = fmaf(fr, m.film, lr);
        lg = fmaf(fg, m.film, lg);
        lb = fmaf(fb, m.film, lb);
    }
    cr = fmaf(weight, lr, cr);
    cg = fmaf(weight, lg, cg);
    cb = fmaf(weight, lb, cb);

    // the mirror ray enters the parent as refl * 0.5 * gain * lit (shader.py:106): children of weight 0
    // add exactly 0 and are not traced; neither are chains that provably stay below the cut-off
    weight *= lit0 ? m.half_gain : 0.0f;
    if (weight == 0.0f) return false;
    if (sc.chain_bound() > 0.0f && fabsf(weight) * sc.chain_bound() < sp.cutoff_f) return false;
    if (sp.max_depth > 0 && generation + 1 >= sp.max_depth) return false;

    // mirror direction (shader.py:151): D - 2 (D.N) N for the unit D
    const double dn2 = 2.0 * fdot(d, normal);
    Vec3 r = {fma(-dn2, normal.x, d.x), fma(-dn2, normal.y, d.y), fma(-dn2, normal.z, d.z)};
    if (PRIMARY) r = r * inv_len;
    d = r;
    o = nudged;
    return true;
}

template <class Scene, bool STATS, bool CAMERA>
__device__ __forceinline__ void trace_chain_fast(const Scene& sc, const SceneParams& sp, Vec3 o, Vec3 d, double a, float& cr,
                                                 float& cg, float& cb, Counters& k) {
    cr = cg = cb = 0.0f;
    float weight = 1.0f;
    unsigned long long tests = 0;
    int generation = 0;
    bool alive = true;
    if (CAMERA) {
        alive = fast_generation<Scene, true, STATS>(sc, sp, 0, o, d, a, cr, cg, cb, weight, tests, k);
        generation = 1;
    }
    if (alive) {
#pragma unroll 1
        for (;; ++generation)
            if (!fast_generation<Scene, false, STATS>(sc, sp, generation, o, d, 1.0, cr, cg, cb, weight, tests, k)) break;
    }
    if (STATS) k.tests += tests;
}

}  // namespace rt

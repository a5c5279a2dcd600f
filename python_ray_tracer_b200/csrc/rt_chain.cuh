// rt_chain.cuh -- scene staging and the register-resident bounce chain (device).
#pragma once

#include "rt_accel.cuh"

#ifndef RT_PEEL_PRIMARY
#define RT_PEEL_PRIMARY 1
#endif
#ifndef RT_FUSE_NEXT_HIT
#define RT_FUSE_NEXT_HIT 0  // experiment (rt_chain_fused.cuh): shadow test and the mirror ray's nearest hit in one loop
#endif

namespace rt {

struct Counters {
    unsigned long long traced, hits, deepest, tests;
};

// Upper bound of |local colour| of one material (Fast cutoff).  local = ambient +
// diffuse + dome + gain*(spec_base + gain*glint) + film, with F <= 1, G1(L) <= 1,
// D <= a2/(pi*(min(1,a2^2)+eps)) and G1(V)/(4 NdotV + eps) <= 1/(2*(sqrt(min(a2,1))+eps)).
__device__ __forceinline__ double material_bound(const SceneParams& sp, int s) {
    Material m = load_material(sp, s);
    double tex = m.checker ? 1.0 : fmax(fabs(m.r), fmax(fabs(m.g), fabs(m.b)));
    double dome = fabs(sp.dome_sum) * fmax(fabs(sp.dome_color[0]), fmax(fabs(sp.dome_color[1]), fabs(sp.dome_color[2])));
    double alpha = m.roughness * m.roughness;
    double a2 = alpha * alpha;
    double dmax = a2 / (kPi * (fmin(1.0, a2 * a2) + kEps));
    double gv = 1.0 / (2.0 * (sqrt(fmin(a2, 1.0)) + kEps));
    double g = fabs(m.specular_gain);
    double spec = g * (dmax * gv + g);
    double film = 1.4 * kFilmWeight * fabs(m.irid_gain);
    return kAmbient + fabs(m.diffuse_gain) * tex + dome + spec + film;
}

// Fill the shared-memory sphere records from the SoA in global memory.  The kernels are
// persistent (one resident block per slot, looping over tiles), so every block does this
// once per launch; the table is 144 bytes per sphere plus the point lights in use.  Must be called by all threads.
// ACCEL launches keep the spheres in global memory (BVH blob): only the lights are staged, and
// the cut-off bound comes from the blob header (computed once per scene on the host).
template <int MODE>
__device__ __forceinline__ Staged stage_lights_only(const SceneParams& sp, unsigned char* smem, const AccelView& av) {
    double* lights = reinterpret_cast<double*>(smem);
    for (int i = threadIdx.x; i < 3 * sp.n_lights_used; i += blockDim.x) lights[i] = __ldg(sp.lights + i);
    __syncthreads();
    Staged st;
    st.geom = st.prim = st.aux = st.mat = 0u;
    st.lights = (unsigned int)__cvta_generic_to_shared(lights);
    st.chain_bound = -1.0;
    if (MODE == kModeFast && sp.cutoff > 0.0) {
        const double wmax = 0.5 * av.gain_bound;
        if (wmax < 1.0 && isfinite(av.local_bound)) st.chain_bound = av.local_bound / (1.0 - wmax);
    }
    return st;
}

template <int MODE>
__device__ __forceinline__ Staged stage_scene(const SceneParams& sp, unsigned char* smem, const double* origin0) {
    const int n = sp.n_spheres;
    SphereGeom* geom = reinterpret_cast<SphereGeom*>(smem);
    SpherePrimary* prim = reinterpret_cast<SpherePrimary*>(smem + sizeof(SphereGeom) * (size_t)n);
    SphereAux* aux = reinterpret_cast<SphereAux*>(smem + (sizeof(SphereGeom) + sizeof(SpherePrimary)) * (size_t)n);
    SphereMaterial* mat = reinterpret_cast<SphereMaterial*>(
        smem + (sizeof(SphereGeom) + sizeof(SpherePrimary) + sizeof(SphereAux)) * (size_t)n);
    double* lights = reinterpret_cast<double*>(
        smem + (sizeof(SphereGeom) + sizeof(SpherePrimary) + sizeof(SphereAux) + sizeof(SphereMaterial)) * (size_t)n);
    for (int i = threadIdx.x; i < 3 * sp.n_lights_used; i += blockDim.x) lights[i] = __ldg(sp.lights + i);
    __shared__ unsigned long long s_bound_bits, s_gain_bits;
    if (threadIdx.x == 0) {
        s_bound_bits = 0ull;
        s_gain_bits = 0ull;
    }
    __syncthreads();
    double bound = 0.0, gain = 0.0;
    for (int s = threadIdx.x; s < n; s += blockDim.x) {
        double cx = __ldg(sp.spheres + (size_t)RT_F_CX * n + s);
        double cy = __ldg(sp.spheres + (size_t)RT_F_CY * n + s);
        double cz = __ldg(sp.spheres + (size_t)RT_F_CZ * n + s);
        double r = __ldg(sp.spheres + (size_t)RT_F_RADIUS * n + s);
        geom[s] = {cx, cy, cz, r * r};
        aux[s] = {1.0 / r, (cx * cx) + (cy * cy) + (cz * cz)};
        {
            // rays of generation 0 of the camera kernel all start at origin0
            double lx = origin0[0] - cx, ly = origin0[1] - cy, lz = origin0[2] - cz;
            prim[s] = {lx, ly, lz, fma(lz, lz, fma(ly, ly, fma(lx, lx, -(r * r))))};
        }
        {
            Material m = load_material(sp, s);
            mat[s] = {m.diffuse_gain, m.specular_gain, m.roughness, m.irid_gain, m.r, m.g, m.b, m.checker ? 1.0 : 0.0};
        }
        if (MODE == kModeFast) {
            bound = fmax(bound, material_bound(sp, s));
            gain = fmax(gain, fabs(__ldg(sp.spheres + (size_t)RT_F_SPECULAR_GAIN * n + s)));
        }
    }
    if (MODE == kModeFast) {
        // non-negative doubles order like their bit patterns
        atomicMax(&s_bound_bits, (unsigned long long)__double_as_longlong(bound));
        atomicMax(&s_gain_bits, (unsigned long long)__double_as_longlong(gain));
    }
    __syncthreads();
    Staged st;
    st.geom = (unsigned int)__cvta_generic_to_shared(geom);
    st.prim = (unsigned int)__cvta_generic_to_shared(prim);
    st.aux = (unsigned int)__cvta_generic_to_shared(aux);
    st.mat = (unsigned int)__cvta_generic_to_shared(mat);
    st.lights = (unsigned int)__cvta_generic_to_shared(lights);
    st.chain_bound = -1.0;
    if (MODE == kModeFast && sp.cutoff > 0.0) {
        double lmax = __longlong_as_double((long long)s_bound_bits);
        double wmax = 0.5 * __longlong_as_double((long long)s_gain_bits);
        // remaining <= W * lmax * sum_j wmax^j ; only usable when the series converges
        if (wmax < 1.0 && isfinite(lmax)) st.chain_bound = lmax / (1.0 - wmax);
    }
    return st;
}

}  // namespace rt
#if RT_FUSE_NEXT_HIT
#include "rt_chain_fused.cuh"
#endif
namespace rt {

// One camera or user ray through the scene: raytrace_scene (base.py:91-121) with the
// recursion of shader.py:143-161 unrolled into a loop.  Exact ties between two shapes
// (both contribute in the reference) resolve to the lowest shape index here.
// CAMERA: generation 0 starts at sp.eye (the camera kernel), so the fast policy uses the
// staged per-sphere constants of that origin.
// ACCEL: nearest hit and shadow test walk the BVH blob (av) instead of the staged table.
// One generation of a chain: nearest hit, NumpyShader.create, the mirror ray.  Returns false
// when the chain ends here.  PRIMARY: the ray starts at the camera eye (generation 0 of the
// camera kernel, fast policy): staged per-sphere constants of that origin, to_eye = -d.
template <int MODE, bool STATS, bool PRIMARY, bool ACCEL>
__device__ __forceinline__ bool bounce(const SceneParams& sp, const Staged& st, const AccelView& av, const int generation,
                                       Vec3& o, Vec3& d, Vec3& color, double& weight, unsigned long long& tests,
                                       Counters& k) {
    const int n = sp.n_spheres;
    const Vec3 eye = {sp.eye[0], sp.eye[1], sp.eye[2]};
    if (sp.max_depth > 0 && generation >= sp.max_depth) return false;
    if (STATS) {
        k.traced += 1;
        k.deepest = generation + 1 > (int)k.deepest ? generation + 1 : k.deepest;
        if (!ACCEL) tests += (unsigned long long)n;
    }

    // ---- nearest hit (base.py:97-105)
    double tmin = kMiss;
    int sid = -1;
    if (ACCEL) {
        accel_nearest<MODE>(av, o, d, tmin, sid, tests);
    } else if (MODE == kModeExact) {
        for (int s = 0; s < n; ++s) {
            double t = sphere_distance_exact(o, d, load_geom(st, s), load_aux(st, s).cc);
            if (t < tmin) {
                tmin = t;
                sid = s;
            }
        }
    } else {
        // two spheres per trip: their FP64 chains are independent and interleave
        constexpr bool primary = PRIMARY;
        int s = 0;
        for (; s + 1 < n; s += 2) {
            HalfB q0 = primary ? half_b_primary(d, load_prim(st, s)) : half_b(o, d, load_geom(st, s));
            HalfB q1 = primary ? half_b_primary(d, load_prim(st, s + 1)) : half_b(o, d, load_geom(st, s + 1));
            if (can_hit(q0)) {
                double t = near_root(q0);
                if (t < tmin) {
                    tmin = t;
                    sid = s;
                }
            }
            if (can_hit(q1)) {
                double t = near_root(q1);
                if (t < tmin) {
                    tmin = t;
                    sid = s + 1;
                }
            }
        }
        if (s < n) {
            HalfB q = primary ? half_b_primary(d, load_prim(st, s)) : half_b(o, d, load_geom(st, s));
            if (can_hit(q)) {
                double t = near_root(q);
                if (t < tmin) {
                    tmin = t;
                    sid = s;
                }
            }
        }
    }
    if (sid < 0) return false;
    if (STATS) k.hits += 1;

    // ---- NumpyShader.create (shader.py:63-112)
    const SphereGeom g = ACCEL ? ldg_geom(av.geom + sid) : load_geom(st, sid);
    const double inv_r = ACCEL ? __ldg(&av.aux[sid].inv_r) : load_aux(st, sid).inv_r;
    const Material m = ACCEL ? load_material(sp, sid) : load_material(st, sid);
    const Vec3 centre = {g.cx, g.cy, g.cz};
    Vec3 hit, normal, nudged, to_eye;
    if (MODE == kModeExact) {
        hit = o + d * tmin;
        normal = (hit - centre) * inv_r;
        to_eye = norm_exact(eye - hit);
        nudged = hit + normal * kNudge;
    } else {
        hit = {fma(d.x, tmin, o.x), fma(d.y, tmin, o.y), fma(d.z, tmin, o.z)};
        normal = (hit - centre) * inv_r;
        // a primary ray leaves the eye: the direction back to it is -d (to rounding; fast policy only)
        to_eye = PRIMARY ? Vec3{-d.x, -d.y, -d.z} : norm_fast(eye - hit);
        nudged = {fma(normal.x, kNudge, hit.x), fma(normal.y, kNudge, hit.y), fma(normal.z, kNudge, hit.z)};
    }

    Vec3 local = {kAmbient, kAmbient, kAmbient};
    Vec3 tex = {m.r, m.g, m.b};
    if (m.checker) {
        double on = checker_on(hit) ? 1.0 : 0.0;
        tex = {on, on, on};
    }

    bool lit0 = false;
    double spec_sum = 0.0;
    for (int li = 0; li < sp.n_lights_used; ++li) {
        Vec3 lp = load_light(st, li);
        Vec3 to_light = MODE == kModeExact ? norm_exact(lp - hit) : norm_fast(lp - hit);
        bool lit;
        if (ACCEL) {
            const double own = accel_sphere_t<MODE>(nudged, to_light, g, MODE == kModeExact ? __ldg(&av.aux[sid].cc) : 0.0);
            tests += 1;
            lit = !accel_occluded<MODE>(av, sid, nudged, to_light, own, tests);
        } else {
            lit = in_light<MODE>(st, n, sid, nudged, to_light, tests);
        }
        if (li == 0) lit0 = lit;
        if (MODE == kModeExact) {
            double litf = lit ? 1.0 : 0.0;
            double lambert = fmax(dot(normal, to_light), 0.0);
            local.x += tex.x * lambert * litf * m.diffuse_gain;
            local.y += tex.y * lambert * litf * m.diffuse_gain;
            local.z += tex.z * lambert * litf * m.diffuse_gain;
            double spec = specular_exact(normal, to_light, to_eye, m.roughness, m.specular_gain);
            spec_sum += spec * litf;
        } else if (lit) {
            double lambert = pos_part(fdot(normal, to_light)) * m.diffuse_gain;
            local.x = fma(tex.x, lambert, local.x);
            local.y = fma(tex.y, lambert, local.y);
            local.z = fma(tex.z, lambert, local.z);
            if (m.specular_gain != 0.0)
                spec_sum += specular_fast(normal, to_light, to_eye, m.roughness, m.specular_gain);
        }
    }

    // dome light (shader.py:234-244)
    double ny = MODE == kModeExact ? fmax((normal.x * 0.0) + (normal.y * 1.0) + (normal.z * 0.0), 0.0) : pos_part(normal.y);
    double dome;
    if (MODE == kModeExact) {
        dome = 0.0;
        for (int li = 0; li < sp.n_dome; ++li) dome += sp.dome_intensity[li] * ny;
    } else {
        dome = sp.dome_sum * ny;
    }
    local.x += sp.dome_color[0] * dome;
    local.y += sp.dome_color[1] * dome;
    local.z += sp.dome_color[2] * dome;

    // specular enters as spec * gain * lit (shader.py:106)
    double spec_term = spec_sum * m.specular_gain;
    local.x += spec_term;
    local.y += spec_term;
    local.z += spec_term;

    // iridescence (shader.py:110)
    if (MODE == kModeExact || m.irid_gain != 0.0) {
        Vec3 film = film_color<MODE>(normal, to_eye);
        local.x += film.x * kFilmWeight * m.irid_gain;
        local.y += film.y * kFilmWeight * m.irid_gain;
        local.z += film.z * kFilmWeight * m.irid_gain;
    }

    color.x += weight * local.x;
    color.y += weight * local.y;
    color.z += weight * local.z;

    // the mirror ray enters the parent as refl * 0.5 * gain * lit (shader.py:106)
    weight *= (lit0 ? 0.5 : 0.0) * m.specular_gain;
    // Exact keeps tracing a weightless chain, as the reference's unconditional recursion
    // does (shader.py:90): its children add exactly 0 but count in the ray totals.
    if (MODE != kModeExact && weight == 0.0) return false;
    if (MODE == kModeFast && st.chain_bound > 0.0 && fabs(weight) * st.chain_bound < sp.cutoff) return false;

    // mirror direction (shader.py:151): D - (N*2)*(D.N), normalised
    if (MODE == kModeExact) {
        d = norm_exact(d - (normal * 2.0) * dot(d, normal));
    } else {
        double dn2 = 2.0 * fdot(d, normal);
        d = norm_fast({fma(-dn2, normal.x, d.x), fma(-dn2, normal.y, d.y), fma(-dn2, normal.z, d.z)});
    }
    o = nudged;
    return true;
}

template <int MODE, bool STATS, bool CAMERA, bool ACCEL>
__device__ __forceinline__ Vec3 trace_chain(const SceneParams& sp, const Staged& st, const AccelView& av, Vec3 o,
                                            Vec3 d, Counters& k) {
#if RT_FUSE_NEXT_HIT
    if constexpr (MODE == kModeFast && !ACCEL) return trace_chain_fused<STATS, CAMERA>(sp, st, o, d, k);
#endif
    Vec3 color = {0.0, 0.0, 0.0};
    double weight = 1.0;
    unsigned long long tests = 0;
    int generation = 0;
    bool alive = true;
    // generation 0 of the camera kernel is four fifths of the 1080p frame: its own copy of the
    // body, with what is known about a primary ray folded in at compile time
    if (RT_PEEL_PRIMARY && CAMERA && MODE == kModeFast) {
        alive = bounce<MODE, STATS, true, ACCEL>(sp, st, av, 0, o, d, color, weight, tests, k);
        generation = 1;
    }
    if (alive) {
#pragma unroll 1
        for (;; ++generation)
            if (!bounce<MODE, STATS, false, ACCEL>(sp, st, av, generation, o, d, color, weight, tests, k)) break;
    }
    if (STATS) k.tests += tests;
    return color;
}

}  // namespace rt

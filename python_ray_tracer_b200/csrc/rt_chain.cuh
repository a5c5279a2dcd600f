// rt_chain.cuh -- scene staging, and the bounce chain of the EXACT policy (device).
// The fast policy's chain is rt_fast.cuh.
#pragma once

#include "rt_fast.cuh"

namespace rt {

// Bytes of shared memory a block needs for a scene (both policies keep within this).
//   exact: SphereGeom + SphereAux + SphereMaterial = 112 B per sphere
//   fast:  SphereGeom + SpherePrimary + ShadeRec   = 128 B per sphere
__host__ __device__ inline size_t staged_bytes(int n_spheres, int n_lights = 1) {
    return 128 * (size_t)n_spheres + sizeof(double) * 3 * (size_t)n_lights;
}

// ---- exact policy: records in shared memory, all FP64 ------------------------------------------------
// Fill the shared-memory sphere records from the SoA in global memory.  The kernels are persistent
// (one resident block per slot, looping over patches), so every block does this once per launch.
// Must be called by all threads.
__device__ __forceinline__ Staged stage_exact(const SceneParams& sp, unsigned char* smem) {
    const int n = sp.n_spheres;
    SphereGeom* geom = reinterpret_cast<SphereGeom*>(smem);
    SphereAux* aux = reinterpret_cast<SphereAux*>(smem + sizeof(SphereGeom) * (size_t)n);
    SphereMaterial* mat = reinterpret_cast<SphereMaterial*>(smem + (sizeof(SphereGeom) + sizeof(SphereAux)) * (size_t)n);
    const bool tree = sp.accel != nullptr;  // the tree kernels read spheres from the blob: lights only
    double* lights = reinterpret_cast<double*>(smem + (tree ? 0 : 128 * (size_t)n));
    for (int i = threadIdx.x; i < 3 * sp.n_lights_used; i += blockDim.x) lights[i] = __ldg(sp.lights + i);
    for (int s = threadIdx.x; s < n && !tree; s += blockDim.x) {
        double cx = __ldg(sp.spheres + (size_t)RT_F_CX * n + s);
        double cy = __ldg(sp.spheres + (size_t)RT_F_CY * n + s);
        double cz = __ldg(sp.spheres + (size_t)RT_F_CZ * n + s);
        double r = __ldg(sp.spheres + (size_t)RT_F_RADIUS * n + s);
        geom[s] = {cx, cy, cz, r * r};
        aux[s] = {1.0 / r, (cx * cx) + (cy * cy) + (cz * cz)};
        Material m = load_material(sp, s);
        mat[s] = {m.diffuse_gain, m.specular_gain, m.roughness, m.irid_gain, m.r, m.g, m.b, m.checker ? 1.0 : 0.0};
    }
    __syncthreads();
    Staged st;
    st.geom = (unsigned int)__cvta_generic_to_shared(geom);
    st.aux = (unsigned int)__cvta_generic_to_shared(aux);
    st.mat = (unsigned int)__cvta_generic_to_shared(mat);
    st.lights = (unsigned int)__cvta_generic_to_shared(lights);
    return st;
}

// ---- fast policy: the three scene policies of rt_fast.cuh ---------------------------------------------
__device__ __forceinline__ unsigned int stage_lights(const SceneParams& sp, double* lights) {
    for (int i = threadIdx.x; i < 3 * sp.n_lights_used; i += blockDim.x) lights[i] = __ldg(sp.lights + i);
    return (unsigned int)__cvta_generic_to_shared(lights);
}

__device__ __forceinline__ StagedScene stage_fast(const SceneParams& sp, unsigned char* smem, const double* origin0) {
    const int n = sp.n_spheres;
    SphereGeom* geom = reinterpret_cast<SphereGeom*>(smem);
    SpherePrimary* prim = reinterpret_cast<SpherePrimary*>(smem + 32 * (size_t)n);
    ShadeRec* rec = reinterpret_cast<ShadeRec*>(smem + 64 * (size_t)n);
    StagedScene sc;
    sc.lights = stage_lights(sp, reinterpret_cast<double*>(smem + 128 * (size_t)n));
    __shared__ unsigned long long s_bound_bits, s_gain_bits;
    if (!sp.host_bound) {
        if (threadIdx.x == 0) s_bound_bits = 0ull, s_gain_bits = 0ull;
        __syncthreads();
    }
    double bound = 0.0, gain = 0.0;
    for (int s = threadIdx.x; s < n; s += blockDim.x) {
        const double* f = sp.spheres;
        const double cx = __ldg(f + (size_t)RT_F_CX * n + s), cy = __ldg(f + (size_t)RT_F_CY * n + s);
        const double cz = __ldg(f + (size_t)RT_F_CZ * n + s), r = __ldg(f + (size_t)RT_F_RADIUS * n + s);
        const Material m = load_material(sp, s);
        geom[s] = {cx, cy, cz, r * r};
        prim[s] = make_sphere_primary(origin0, cx, cy, cz, r);
        rec[s] = make_shade_rec(cx, cy, cz, r, m.diffuse_gain, m.specular_gain, m.roughness, m.irid_gain, m.r, m.g, m.b, m.checker);
        if (!sp.host_bound) {
            bound = fmax(bound, material_bound(m.diffuse_gain, m.specular_gain, m.roughness, m.irid_gain, m.r, m.g, m.b, m.checker,
                                               sp.dome_sum, sp.dome_color));
            gain = fmax(gain, fabs(m.specular_gain));
        }
    }
    if (!sp.host_bound) {
        // no host copy of the table: the bound is formed here; non-negative doubles order like their bit patterns
        atomicMax(&s_bound_bits, (unsigned long long)__double_as_longlong(bound));
        atomicMax(&s_gain_bits, (unsigned long long)__double_as_longlong(gain));
    }
    __syncthreads();
    sc.geom = (unsigned int)__cvta_generic_to_shared(geom);
    sc.prim = (unsigned int)__cvta_generic_to_shared(prim);
    sc.recs = (unsigned int)__cvta_generic_to_shared(rec);
    sc.n = n;
    sc.lights_used = sp.n_lights_used;
    sc.bound = sp.host_bound ? sp.chain_bound
                             : chain_bound_of(sp.cutoff, __longlong_as_double((long long)s_bound_bits),
                                              __longlong_as_double((long long)s_gain_bits));
    return sc;
}

__device__ __forceinline__ TreeScene stage_tree(const SceneParams& sp, unsigned char* smem) {
    TreeScene sc;
    sc.av = accel_view(sp);
    sc.lights = stage_lights(sp, reinterpret_cast<double*>(smem));
    __syncthreads();
    sc.lights_used = sp.n_lights_used;
    sc.bound = chain_bound_of(sp.cutoff, sc.av.local_bound, sc.av.gain_bound);
    return sc;
}

template <int N>
__device__ __forceinline__ SmallScene<N> stage_small(const SceneParams& sp, const SmallTable& table, unsigned char* smem) {
    // what is indexed by a sphere number only known at run time (the sphere that was hit, the others of
    // its shadow test): a shared-memory copy, N * (16 + 8) words
    unsigned int* dst = reinterpret_cast<unsigned int*>(smem);
    const unsigned int* rec = reinterpret_cast<const unsigned int*>(table.rec);
    const unsigned int* geom = reinterpret_cast<const unsigned int*>(table.geom);
    for (int i = threadIdx.x; i < N * 16; i += blockDim.x) dst[i] = rec[i];
    for (int i = threadIdx.x; i < N * 8; i += blockDim.x) dst[N * 16 + i] = geom[i];
    __syncthreads();
    SmallScene<N> sc;
    sc.t = &table;
    sc.recs = (unsigned int)__cvta_generic_to_shared(smem);
    sc.geoms = sc.recs + N * 64u;
    sc.light0 = {sp.light0[0], sp.light0[1], sp.light0[2]};
    sc.bound = sp.chain_bound;
    return sc;
}

// ---- exact policy: one ray through the scene ----------------------------------------------------------
// raytrace_scene (base.py:91-121) with the recursion of shader.py:143-161 unrolled into a loop: every hit
// spawns exactly one mirror ray whose colour enters the parent with weight 0.5*specular_gain*lit
// (shader.py:106).  All FP64, the reference's expressions in the reference's order (this translation unit
// is compiled with -fmad=false).  Children of weight 0 are still traced, as the reference's unconditional
// recursion does (shader.py:90): they add exactly 0 but count in the ray totals.
// TREE: nearest hit and shadow test walk the BVH blob instead of the staged table.

// Nearest hit (base.py:97-105): smallest distance, the lowest shape index that has it, and how many shapes do.
template <bool TREE>
__device__ __forceinline__ void nearest_exact(const SceneParams& sp, const Staged& st, const AccelView& av, Vec3 o, Vec3 d,
                                              double& tmin, int& sid, int& ties, unsigned long long& tests) {
    tmin = kMiss;
    sid = -1;
    ties = 0;
    if (TREE) {
        accel_nearest(av, o, d, 1.0f, tmin, sid, tests,
                      [&](const SphereGeom& g, int s) { return sphere_distance_exact(o, d, g, __ldg(&av.aux[s].cc)); }, &ties);
    } else {
        for (int s = 0; s < sp.n_spheres; ++s) {
            const double t = sphere_distance_exact(o, d, load_geom(st, s), load_aux(st, s).cc);
            if (t < tmin) {
                tmin = t;
                sid = s;
                ties = 1;
            } else if (t == tmin && t != kMiss) {
                ++ties;
            }
        }
    }
}

// NumpyShader.create (shader.py:63-112) for the hit of ray (o, d) on shape sid at distance t: adds weight * local
// colour to `color`; (o, d, weight) become the mirror ray and its weight.
template <bool STATS, bool TREE>
__device__ __forceinline__ void shade_exact(const SceneParams& sp, const Staged& st, const AccelView& av, const int sid, const double t,
                                            Vec3& o, Vec3& d, double& weight, Vec3& color, unsigned long long& tests, Counters& k) {
    const int n = sp.n_spheres;
    const Vec3 eye = {sp.eye[0], sp.eye[1], sp.eye[2]};
    if (STATS) k.hits += 1;
    const SphereGeom g = TREE ? ldg_geom(av.geom + sid) : load_geom(st, sid);
    const double inv_r = TREE ? __ldg(&av.aux[sid].inv_r) : load_aux(st, sid).inv_r;
    const Material m = TREE ? load_material(sp, sid) : load_material(st, sid);
    const Vec3 centre = {g.cx, g.cy, g.cz};
    const Vec3 hit = o + d * t;
    const Vec3 normal = (hit - centre) * inv_r;
    const Vec3 to_eye = norm_exact(eye - hit);
    const Vec3 nudged = hit + normal * kNudge;

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
        Vec3 to_light = norm_exact(lp - hit);
        bool lit;
        if (TREE) {
            const double own = sphere_distance_exact(nudged, to_light, g, __ldg(&av.aux[sid].cc));
            tests += 1;
            lit = !accel_occluded(av, sid, nudged, to_light, own, 1.0f, tests, [&](const SphereGeom& gj, int j) {
                return sphere_distance_exact(nudged, to_light, gj, __ldg(&av.aux[j].cc));
            });
        } else {
            lit = in_light_exact(st, n, sid, nudged, to_light, tests);
        }
        if (li == 0) lit0 = lit;
        double litf = lit ? 1.0 : 0.0;
        double lambert = fmax(dot(normal, to_light), 0.0);
        local.x += tex.x * lambert * litf * m.diffuse_gain;
        local.y += tex.y * lambert * litf * m.diffuse_gain;
        local.z += tex.z * lambert * litf * m.diffuse_gain;
        double spec = specular_exact(normal, to_light, to_eye, m.roughness, m.specular_gain);
        spec_sum += spec * litf;
    }

    // dome light (shader.py:234-244)
    double ny = fmax((normal.x * 0.0) + (normal.y * 1.0) + (normal.z * 0.0), 0.0);
    double dome = 0.0;
    for (int li = 0; li < sp.n_dome; ++li) dome += sp.dome_intensity[li] * ny;
    local.x += sp.dome_color[0] * dome;
    local.y += sp.dome_color[1] * dome;
    local.z += sp.dome_color[2] * dome;

    // specular enters as spec * gain * lit (shader.py:106)
    double spec_term = spec_sum * m.specular_gain;
    local.x += spec_term;
    local.y += spec_term;
    local.z += spec_term;

    // iridescence (shader.py:110)
    Vec3 film = film_color_exact(normal, to_eye);
    local.x += film.x * kFilmWeight * m.irid_gain;
    local.y += film.y * kFilmWeight * m.irid_gain;
    local.z += film.z * kFilmWeight * m.irid_gain;

    color.x += weight * local.x;
    color.y += weight * local.y;
    color.z += weight * local.z;

    // the mirror ray enters the parent as refl * 0.5 * gain * lit (shader.py:106)
    weight *= (lit0 ? 0.5 : 0.0) * m.specular_gain;
    // mirror direction (shader.py:151): D - (N*2)*(D.N), normalised
    d = norm_exact(d - (normal * 2.0) * dot(d, normal));
    o = nudged;
}

// Rays waiting their turn.  A ray normally has one child, so the chain is a loop; where several shapes lie at
// exactly the nearest distance (a shape listed twice, coincident surfaces) the reference shades every one of
// them and recurses from each (base.py:102-119): the first child continues the loop, the others wait here.
struct PendingRay {
    Vec3 o, d;
    double weight;
    int generation;
};
constexpr int kTieStack = 32;

template <bool STATS, bool TREE>
__device__ __forceinline__ Vec3 trace_chain_exact(const SceneParams& sp, const Staged& st, const AccelView& av, Vec3 o, Vec3 d,
                                                  Counters& k) {
    Vec3 color = {0.0, 0.0, 0.0};
    double weight = 1.0;
    unsigned long long tests = 0;
    PendingRay pending[kTieStack];
    int top = 0, generation = 0;
#pragma unroll 1
    for (;;) {
        bool goes_on = false;
        if (!(sp.max_depth > 0 && generation >= sp.max_depth)) {
            if (STATS) {
                k.traced += 1;
                k.deepest = generation + 1 > (int)k.deepest ? generation + 1 : k.deepest;
                if (!TREE) tests += (unsigned long long)sp.n_spheres;
            }
            double tmin;
            int sid, ties;
            nearest_exact<TREE>(sp, st, av, o, d, tmin, sid, ties, tests);
            if (sid >= 0) {
                goes_on = true;
                if (ties > 1) {
                    // every other shape at that distance, highest index first onto the stack (so they come off in
                    // the reference's order); found by testing all shapes again: ties are rare
                    for (int s = sp.n_spheres - 1; s > sid; --s) {
                        const SphereGeom g = TREE ? ldg_geom(av.geom + s) : load_geom(st, s);
                        const double cc = TREE ? __ldg(&av.aux[s].cc) : load_aux(st, s).cc;
                        if (sphere_distance_exact(o, d, g, cc) != tmin) continue;
                        PendingRay child = {o, d, weight, generation + 1};
                        shade_exact<STATS, TREE>(sp, st, av, s, tmin, child.o, child.d, child.weight, color, tests, k);
                        if (top < kTieStack) pending[top++] = child;
                        else if (sp.error_word != nullptr) *reinterpret_cast<volatile unsigned int*>(sp.error_word) = kDevTieStack;
                    }
                }
                shade_exact<STATS, TREE>(sp, st, av, sid, tmin, o, d, weight, color, tests, k);
                ++generation;
            }
        }
        if (goes_on) continue;
        if (top == 0) break;
        --top;
        o = pending[top].o, d = pending[top].d, weight = pending[top].weight, generation = pending[top].generation;
    }
    if (STATS) k.tests += tests;
    return color;
}

}  // namespace rt

// rt_chain_fused.cuh -- EXPERIMENT, off by default (-DRT_FUSE_NEXT_HIT=1 builds it in).  Run once
// on a B200 (47 parity tests green, counters identical, 1080p 0.0805 ms: DESIGN.md section 6c);
// tools/round2_sweep.sh runs the full suite and the other workloads.  Fast policy, staged sphere
// table only.
//
// The shadow ray of a hit and its mirror ray leave the same point (the nudged hit point,
// shader.py:77, :117, :152).  trace_chain() walks the sphere table once for the shadow test of
// generation g and once more for the nearest hit of generation g + 1; here one loop serves both
// rays: O - C and |O - C|^2 - r^2 are formed once per sphere (7 of the 13 FP64 instructions of a
// test, one pair of shared-memory loads), and the two dot products are independent chains.
// The mirror ray is intersected before it is known whether the chain goes on (it does after
// 94 % of the hits of the README frame); a chain that ends drops the result.
//
// Same expressions on the same operands in the same order as trace_chain()'s fast path, so the
// colours, the nearest-hit decisions and the counters are meant to be bit-identical to it; the
// counters count the rays of the chain (the algorithmic work), not the dropped intersection.
// One body serves every generation (no peeled generation 0: half the code of trace_chain()).
#pragma once

namespace rt {

#ifndef RT_FUSE_PAIRS
#define RT_FUSE_PAIRS 0  // 1: two spheres per trip of the shared loop (four independent chains instead of two)
#endif

// Both rays against sphere j: the quadratic of the shadow ray `l` and of the mirror ray `dm`.
__device__ __forceinline__ void two_rays_one_sphere(const Staged& st, const int j, const Vec3 p, const Vec3 l, const Vec3 dm,
                                                    HalfB& qs, HalfB& qm) {
    const SphereGeom g = load_geom(st, j);
    const double lx = p.x - g.cx, ly = p.y - g.cy, lz = p.z - g.cz;
    const double c = fma(lz, lz, fma(ly, ly, fma(lx, lx, -g.rr)));
    qs.b = fma(l.z, lz, fma(l.y, ly, l.x * lx));
    qm.b = fma(dm.z, lz, fma(dm.y, ly, dm.x * lx));
    qs.c = c;
    qm.c = c;
    qs.disc = fma(qs.b, qs.b, -c);
    qm.disc = fma(qm.b, qm.b, -c);
}
__device__ __forceinline__ void settle(const int j, const int s, const HalfB& qs, const HalfB& qm, double& other, double& t_next,
                                       int& s_next) {
    if (j != s && can_hit(qs)) other = fmin_pos(other, near_root(qs));
    if (can_hit(qm)) {
        const double t = near_root(qm);
        if (t < t_next) {
            t_next = t;
            s_next = j;
        }
    }
}

// in_light() (fast) for light direction `l` and the nearest hit of the ray `dm`, both from `p`.
__device__ __forceinline__ void shadow_and_next_hit(const Staged& st, const int n, const int s, const Vec3 p, const Vec3 l,
                                                    const Vec3 dm, bool& lit, double& t_next, int& s_next) {
    double other = kMiss;
    t_next = kMiss;
    s_next = -1;
    int j = 0;
#if RT_FUSE_PAIRS
#pragma unroll 1
    for (; j + 1 < n; j += 2) {
        HalfB qs0, qm0, qs1, qm1;
        two_rays_one_sphere(st, j, p, l, dm, qs0, qm0);
        two_rays_one_sphere(st, j + 1, p, l, dm, qs1, qm1);
        settle(j, s, qs0, qm0, other, t_next, s_next);  // ascending j: ties go to the lower index
        settle(j + 1, s, qs1, qm1, other, t_next, s_next);
    }
#endif
#pragma unroll 1
    for (; j < n; ++j) {
        HalfB qs, qm;
        two_rays_one_sphere(st, j, p, l, dm, qs, qm);
        settle(j, s, qs, qm, other, t_next, s_next);
    }
    if (other == kMiss) {
        lit = true;
        return;
    }
    const HalfB qo = half_b(p, l, load_geom(st, s));
    const double own = can_hit(qo) ? near_root(qo) : kMiss;
    lit = !(other < own);
}

template <bool STATS, bool CAMERA>
__device__ __forceinline__ Vec3 trace_chain_fused(const SceneParams& sp, const Staged& st, Vec3 o, Vec3 d, Counters& k) {
    const int n = sp.n_spheres;
    const Vec3 eye = {sp.eye[0], sp.eye[1], sp.eye[2]};
    Vec3 color = {0.0, 0.0, 0.0};
    double weight = 1.0;
    unsigned long long tests = 0;

    // ---- generation 0: nearest hit (base.py:97-105), as trace_chain() does it
    if (STATS) {
        k.traced += 1;
        k.deepest = k.deepest < 1 ? 1 : k.deepest;
        tests += (unsigned long long)n;
    }
    double tmin = kMiss;
    int sid = -1;
    {
        int s = 0;
        for (; s + 1 < n; s += 2) {
            HalfB q0 = CAMERA ? half_b_primary(d, load_prim(st, s)) : half_b(o, d, load_geom(st, s));
            HalfB q1 = CAMERA ? half_b_primary(d, load_prim(st, s + 1)) : half_b(o, d, load_geom(st, s + 1));
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
            HalfB q = CAMERA ? half_b_primary(d, load_prim(st, s)) : half_b(o, d, load_geom(st, s));
            if (can_hit(q)) {
                double t = near_root(q);
                if (t < tmin) {
                    tmin = t;
                    sid = s;
                }
            }
        }
    }

#pragma unroll 1
    for (int generation = 0; sid >= 0; ++generation) {
        if (STATS) k.hits += 1;

        // ---- NumpyShader.create (shader.py:63-112)
        const SphereGeom g = load_geom(st, sid);
        const double inv_r = load_aux(st, sid).inv_r;
        const Material m = load_material(st, sid);
        const Vec3 centre = {g.cx, g.cy, g.cz};
        const Vec3 hit = {fma(d.x, tmin, o.x), fma(d.y, tmin, o.y), fma(d.z, tmin, o.z)};
        const Vec3 normal = (hit - centre) * inv_r;
        // a primary ray leaves the eye: the direction back to it is -d (to rounding), as in trace_chain()
        Vec3 to_eye = {-d.x, -d.y, -d.z};
        if (!(CAMERA && generation == 0)) to_eye = norm_fast(eye - hit);
        const Vec3 nudged = {fma(normal.x, kNudge, hit.x), fma(normal.y, kNudge, hit.y), fma(normal.z, kNudge, hit.z)};
        // mirror direction (shader.py:151): D - (N*2)*(D.N), normalised -- needed by the shared loop
        const double dn2 = 2.0 * fdot(d, normal);
        const Vec3 dm = norm_fast({fma(-dn2, normal.x, d.x), fma(-dn2, normal.y, d.y), fma(-dn2, normal.z, d.z)});

        Vec3 local = {kAmbient, kAmbient, kAmbient};
        Vec3 tex = {m.r, m.g, m.b};
        if (m.checker) {
            double on = checker_on(hit) ? 1.0 : 0.0;
            tex = {on, on, on};
        }

        bool lit0 = false;
        double spec_sum = 0.0;
        double t_next = kMiss;
        int s_next = -1;
        for (int li = 0; li < sp.n_lights_used; ++li) {
            const Vec3 lp = load_light(st, li);
            const Vec3 to_light = norm_fast(lp - hit);
            bool lit;
            if (li == 0) {
                shadow_and_next_hit(st, n, sid, nudged, to_light, dm, lit, t_next, s_next);
                tests += (unsigned long long)n;
                lit0 = lit;
            } else {
                lit = in_light<kModeFast>(st, n, sid, nudged, to_light, tests);
            }
            if (lit) {
                double lambert = pos_part(fdot(normal, to_light)) * m.diffuse_gain;
                local.x = fma(tex.x, lambert, local.x);
                local.y = fma(tex.y, lambert, local.y);
                local.z = fma(tex.z, lambert, local.z);
                if (m.specular_gain != 0.0) spec_sum += specular_fast(normal, to_light, to_eye, m.roughness, m.specular_gain);
            }
        }

        // dome light (shader.py:234-244)
        const double dome = sp.dome_sum * pos_part(normal.y);
        local.x += sp.dome_color[0] * dome;
        local.y += sp.dome_color[1] * dome;
        local.z += sp.dome_color[2] * dome;

        // specular enters as spec * gain * lit (shader.py:106)
        const double spec_term = spec_sum * m.specular_gain;
        local.x += spec_term;
        local.y += spec_term;
        local.z += spec_term;

        // iridescence (shader.py:110)
        if (m.irid_gain != 0.0) {
            Vec3 film = film_color<kModeFast>(normal, to_eye);
            local.x += film.x * kFilmWeight * m.irid_gain;
            local.y += film.y * kFilmWeight * m.irid_gain;
            local.z += film.z * kFilmWeight * m.irid_gain;
        }

        color.x += weight * local.x;
        color.y += weight * local.y;
        color.z += weight * local.z;

        // the mirror ray enters the parent as refl * 0.5 * gain * lit (shader.py:106)
        weight *= (lit0 ? 0.5 : 0.0) * m.specular_gain;
        if (weight == 0.0) break;
        if (st.chain_bound > 0.0 && fabs(weight) * st.chain_bound < sp.cutoff) break;
        if (sp.max_depth > 0 && generation + 1 >= sp.max_depth) break;

        // the mirror ray is generation + 1 of the chain; its nearest hit is already known
        if (STATS) {
            k.traced += 1;
            k.deepest = (unsigned long long)(generation + 2) > k.deepest ? (unsigned long long)(generation + 2) : k.deepest;
            tests += (unsigned long long)n;
        }
        o = nudged;
        d = dm;
        tmin = t_next;
        sid = s_next;
    }
    if (STATS) k.tests += tests;
    return color;
}

}  // namespace rt

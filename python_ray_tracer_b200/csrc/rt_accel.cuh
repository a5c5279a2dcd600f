// rt_accel.cuh -- device traversal of the sphere BVH blob (rt_accel.h).
//
// Used for scenes too large to be worth testing exhaustively (and for those whose table does
// not fit shared memory).  Every sphere test here is the same function on the same record as
// in the exhaustive loops of rt_chain.cuh / rt_trace.cuh; the tree only decides which tests
// are skipped, and it skips a test only when the ray misses a box that contains the sphere
// (boxes are padded, so the slab test's rounding never pre-empts the sphere test).
#pragma once

#include "rt_accel.h"
#include "rt_trace.cuh"

namespace rt {

// Measurement builds only (-DRT_ACCEL_PROBE=k, tools/c3_probe.sh): the `evaluated` counter counts something else
// than sphere tests.  1 / 2: inner-node visits per thread / per warp;  3 / 4: leaf visits per thread / per warp;
// 5 / 6: traversals per thread / per warp.
#ifndef RT_ACCEL_PROBE
#define RT_ACCEL_PROBE 0
#endif
__device__ __forceinline__ void accel_probe(unsigned long long& evaluated, int what, unsigned long long sphere_tests) {
#if RT_ACCEL_PROBE == 0
    if (what == 0) evaluated += sphere_tests;
#else
    const bool leader = (threadIdx.x & 31) == (unsigned)(__ffs(__activemask()) - 1);
    if (what == (RT_ACCEL_PROBE + 1) / 2 && ((RT_ACCEL_PROBE & 1) || leader)) evaluated += 1ull;
#endif
}

struct AccelView {
    const AccelNode* nodes;
    const int* order;
    const SphereGeom* leaf_geom;
    const SphereGeom* geom;
    const SphereAux* aux;
    const ShadeRec* rec;
    const int* large;
    unsigned int* error;  // SceneParams::error_word
    int n_nodes, n_large;
    double local_bound, gain_bound;
    float extent;
};

// Is sp.accel a blob built for this sphere table?  A stale or foreign blob would render a mixed scene or
// read out of bounds: the launch reports it and renders nothing.  Same answer in every thread.
__device__ __forceinline__ bool accel_blob_ok(const SceneParams& sp) {
    const AccelHeader* h = reinterpret_cast<const AccelHeader*>(sp.accel);
    const bool ok = sp.accel_bytes >= (long long)sizeof(AccelHeader) && __ldg(&h->magic) == kAccelMagic &&
                    __ldg(&h->n_spheres) == sp.n_spheres && __ldg(&h->bytes) <= sp.accel_bytes;
    if (!ok && sp.error_word != nullptr && threadIdx.x == 0) *reinterpret_cast<volatile unsigned int*>(sp.error_word) = kDevBadAccelBlob;
    return ok;
}

__device__ __forceinline__ AccelView accel_view(const SceneParams& sp) {
    const unsigned char* blob = sp.accel;
    const AccelHeader* h = reinterpret_cast<const AccelHeader*>(blob);
    AccelView v;
    v.error = sp.error_word;
    v.nodes = reinterpret_cast<const AccelNode*>(blob + h->off_nodes);
    v.order = reinterpret_cast<const int*>(blob + h->off_order);
    v.leaf_geom = reinterpret_cast<const SphereGeom*>(blob + h->off_leaf_geom);
    v.geom = reinterpret_cast<const SphereGeom*>(blob + h->off_geom);
    v.aux = reinterpret_cast<const SphereAux*>(blob + h->off_aux);
    v.rec = reinterpret_cast<const ShadeRec*>(blob + h->off_rec);
    v.large = reinterpret_cast<const int*>(blob + h->off_large);
    v.n_nodes = h->n_nodes;
    v.n_large = h->n_large;
    v.local_bound = h->local_bound;
    v.gain_bound = h->gain_bound;
    v.extent = h->extent;
    return v;
}

__device__ __forceinline__ SphereGeom ldg_geom(const SphereGeom* p) {
    const double2 a = __ldg(reinterpret_cast<const double2*>(p));
    const double2 b = __ldg(reinterpret_cast<const double2*>(p) + 1);
    return {a.x, a.y, b.x, b.y};
}

// ---- the slab test: FP32, conservative ------------------------------------------------------
// FP64 min/max cost ~8 instructions each on sm_100a (NaN rules emulated), and a slab test needs a
// dozen: in FP64 it was half of all instructions the tree kernels issued.  In FP32 it is one
// FMNMX each -- but the test may then never reject a box the FP64 sphere test would hit in.
//
// Per axis, with A = plane - origin and t = A / d:   computed t_lo must not exceed the true
// one, computed t_hi must not fall below it (whatever the sign of d, that is "A_lo too small,
// A_hi too large").  t = fma(plane, inv, c) with inv = rcp(float(d)) and c = -(o -+ eps) * inv,
// one FFMA per plane.  Roundings on the way, as displacements of the plane: origin to FP32
// (2^-24 |o|), o +- eps (2^-24 (|o| + eps)), the product c (the same again), the fma
// (2^-24 |A|), inv (float(d) 2^-24 + rcp.approx 2^-23, relative to t, i.e. 3 * 2^-24 |A|), and
// |A| <= |o| + extent + eps: below 7 * 2^-24 (|o| + extent) in all.  The origin is shifted by
// eps = 2^-20 (|o|_max + extent) = 16 * 2^-24 (...) towards each plane.  Boxes are already
// rounded outward by the builder.
//   |inv| is capped at 2^64 (d = +-0 included): the test then sees a ray whose direction
//   component is +-2^-64 instead of something smaller, i.e. displaced by less than 2^-64 t at
//   distance t -- nothing against the margin -- and no product is ever inf * 0 or inf - inf.
//   (Uncapped, a ray exactly parallel to an axis got NaN for one plane and +inf for the other,
//   and dropping the NaN rejected boxes the ray runs through: tests/test_accel_slabs.py.)
//   Products stay finite because the builder refuses coordinates beyond kAccelMaxCoordinate
//   (1e15) and rays that start further out than that are not culled at all.
struct RaySlabs {
    float ix, iy, iz;     // 1 / direction
    float lox, loy, loz;  // -(origin + eps) / direction: added to low plane / direction
    float hix, hiy, hiz;  // -(origin - eps) / direction: added to high plane / direction
};
__device__ __forceinline__ float rcp_approx(float x) {
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ RaySlabs ray_slabs(Vec3 o, Vec3 d, float extent) {
    const float ox = __double2float_rn(o.x), oy = __double2float_rn(o.y), oz = __double2float_rn(o.z);
    const float reach = fmaxf(fmaxf(fabsf(ox), fabsf(oy)), fabsf(oz)) + extent;
    const float eps = fmaf(reach, 9.5367431640625e-07f, 1e-30f);  // 2^-20
    RaySlabs r;
    constexpr float kCap = 18446744073709551616.0f;  // 2^64
    r.ix = fminf(fmaxf(rcp_approx(__double2float_rn(d.x)), -kCap), kCap);
    r.iy = fminf(fmaxf(rcp_approx(__double2float_rn(d.y)), -kCap), kCap);
    r.iz = fminf(fmaxf(rcp_approx(__double2float_rn(d.z)), -kCap), kCap);
    r.lox = -(ox + eps) * r.ix, r.loy = -(oy + eps) * r.iy, r.loz = -(oz + eps) * r.iz;
    r.hix = -(ox - eps) * r.ix, r.hiy = -(oy - eps) * r.iy, r.hiz = -(oz - eps) * r.iz;
    if (!(reach <= (float)kAccelMaxCoordinate)) {
        // an origin FP32 slabs cannot place (or NaN): every box is accepted, nothing is culled
        r.ix = r.iy = r.iz = 0.0f;
        r.lox = r.loy = r.loz = -INFINITY;
        r.hix = r.hiy = r.hiz = INFINITY;
    }
    return r;
}

struct NodeView {
    float lox, loy, loz, hix, hiy, hiz;
    unsigned int ref;
};
__device__ __forceinline__ NodeView ldg_node(const AccelNode* p) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(p));
    const float4 b = __ldg(reinterpret_cast<const float4*>(p) + 1);
    return {a.x, a.y, a.z, a.w, b.x, b.y, __float_as_uint(b.z)};
}

// Distance at which the ray enters the box (0 if it starts inside); `hit` = it may meet the box
// at a distance in (0, limit]  (limit rounded up from the FP64 one).
__device__ __forceinline__ float slab_enter(const RaySlabs& r, const NodeView& n, float limit, bool& hit) {
    const float ax = fmaf(n.lox, r.ix, r.lox), bx = fmaf(n.hix, r.ix, r.hix);
    const float ay = fmaf(n.loy, r.iy, r.loy), by = fmaf(n.hiy, r.iy, r.hiy);
    const float az = fmaf(n.loz, r.iz, r.loz), bz = fmaf(n.hiz, r.iz, r.hiz);
    const float t_near = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), 0.0f));
    const float t_far = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), limit));
    hit = !(t_near > t_far);
    return t_near;
}

// Nearest hit (base.py:97-105): smallest t, lowest shape index among equal t.
// Both children of an inner node are tested together (their boxes are adjacent in memory, so
// the four loads are in flight at once); the one entered first is visited next, the other is
// pushed with its entry distance only if the ray meets it, and dropped on the way back up if a
// closer hit has been found meanwhile.
// `sphere_t(geom, shape)`: the policy's own sphere test, kMiss on a miss.  Its distances may be a
// multiple of the ray parameter the slabs measure (an unnormalised direction): limit_scale converts,
// rounded up.
template <class SphereT>
__device__ __forceinline__ void accel_nearest(const AccelView& av, Vec3 o, Vec3 d, float limit_scale, double& tmin, int& sid,
                                              unsigned long long& evaluated, SphereT&& sphere_t, int* ties = nullptr) {
    auto slab_limit = [&](double t) { return t < kMiss ? __fmul_ru(__double2float_ru(t), limit_scale) : INFINITY; };
    // (`ties`: how many shapes lie at exactly the nearest distance; boxes entered at that distance are not culled --
    // the cut is "entered beyond the limit" -- so every one of them is visited)
    auto consider = [&](int s, double t) {
        if (ties != nullptr && t != kMiss) *ties = t < tmin ? 1 : (t == tmin ? *ties + 1 : *ties);
        if (t < tmin || (t == tmin && t != kMiss && s < sid)) {
            tmin = t;
            sid = s;
        }
    };
    for (int i = 0; i < av.n_large; ++i) {
        const int s = __ldg(av.large + i);
        consider(s, sphere_t(ldg_geom(av.geom + s), s));
    }
    accel_probe(evaluated, 0, (unsigned long long)av.n_large);
    accel_probe(evaluated, 3, 0);
    if (av.n_nodes == 0) return;
    const RaySlabs ray = ray_slabs(o, d, av.extent);
    float stack_t[kAccelStack];  // entry distance of a pushed box
    unsigned int stack_ref[kAccelStack];
    int top = 0;
    float limit = slab_limit(tmin);
    bool hit;
    const NodeView root = ldg_node(av.nodes);
    slab_enter(ray, root, limit, hit);
    if (!hit) return;
    unsigned int ref = root.ref;
    while (true) {
        const int first = (int)(ref >> 3), count = (int)(ref & 7u);
        bool popped = true;
        if (count > 0) {
            for (int k = 0; k < count; ++k) {
                const int slot = first + k;
                const int s = __ldg(av.order + slot);
                consider(s, sphere_t(ldg_geom(av.leaf_geom + slot), s));
            }
            accel_probe(evaluated, 0, (unsigned long long)count);
            accel_probe(evaluated, 2, 0);
            limit = slab_limit(tmin);
        } else {
            accel_probe(evaluated, 1, 0);
            const NodeView a = ldg_node(av.nodes + first), b = ldg_node(av.nodes + first + 1);
            bool hit_a, hit_b;
            const float ta = slab_enter(ray, a, limit, hit_a), tb = slab_enter(ray, b, limit, hit_b);
            const unsigned int ref_a = a.ref, ref_b = b.ref;
            if (hit_a && hit_b) {
                const bool b_first = tb < ta;
                if (top < kAccelStack) {
                    stack_t[top] = b_first ? ta : tb;
                    stack_ref[top] = b_first ? ref_a : ref_b;
                    ++top;
                } else if (av.error != nullptr) {
                    // deeper than any tree rt_accel_build_host makes: the far child is lost -- say so
                    *reinterpret_cast<volatile unsigned int*>(av.error) = kDevAccelStack;
                }
                ref = b_first ? ref_b : ref_a;
                popped = false;
            } else if (hit_a || hit_b) {
                ref = hit_a ? ref_a : ref_b;
                popped = false;
            }
        }
        if (popped) {
            while (true) {
                if (top == 0) return;
                --top;
                if (!(stack_t[top] > limit)) {
                    ref = stack_ref[top];
                    break;
                }
            }
        }
    }
}

// NumpyShader._calculate_shadow (shader.py:114-128): is any OTHER sphere hit closer than
// `own` (the distance to the point's own sphere, kMiss if the shadow ray leaves it)?
template <class SphereT>
__device__ __forceinline__ bool accel_occluded(const AccelView& av, int s_own, Vec3 p, Vec3 l, double own, float limit_scale,
                                               unsigned long long& evaluated, SphereT&& sphere_t) {
    for (int i = 0; i < av.n_large; ++i) {
        const int s = __ldg(av.large + i);
        if (s == s_own) continue;
        const double t = sphere_t(ldg_geom(av.geom + s), s);
        if (t < own) return true;
    }
    accel_probe(evaluated, 0, (unsigned long long)av.n_large);
    accel_probe(evaluated, 3, 0);
    if (av.n_nodes == 0) return false;
    const RaySlabs ray = ray_slabs(p, l, av.extent);
    const float limit = own < kMiss ? __fmul_ru(__double2float_ru(own), limit_scale) : INFINITY;
    unsigned int stack[kAccelStack];
    int top = 0;
    bool hit;
    const NodeView root = ldg_node(av.nodes);
    slab_enter(ray, root, limit, hit);
    if (!hit) return false;
    unsigned int ref = root.ref;
    while (true) {
        const int first = (int)(ref >> 3), count = (int)(ref & 7u);
        bool popped = true;
        if (count > 0) {
            for (int k = 0; k < count; ++k) {
                const int slot = first + k;
                const int s = __ldg(av.order + slot);
                if (s == s_own) continue;
                const double t = sphere_t(ldg_geom(av.leaf_geom + slot), s);
                if (t < own) return true;
            }
            accel_probe(evaluated, 0, (unsigned long long)count);
            accel_probe(evaluated, 2, 0);
        } else {
            accel_probe(evaluated, 1, 0);
            const NodeView a = ldg_node(av.nodes + first), b = ldg_node(av.nodes + first + 1);
            bool hit_a, hit_b;
            slab_enter(ray, a, limit, hit_a);
            slab_enter(ray, b, limit, hit_b);
            if (hit_a || hit_b) {
                if (hit_a && hit_b) {
                    if (top < kAccelStack) stack[top++] = b.ref;
                    else if (av.error != nullptr) *reinterpret_cast<volatile unsigned int*>(av.error) = kDevAccelStack;
                }
                ref = hit_a ? a.ref : b.ref;
                popped = false;
            }
        }
        if (popped) {
            if (top == 0) return false;
            ref = stack[--top];
        }
    }
}

}  // namespace rt

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

struct AccelView {
    const AccelNode* nodes;
    const int* order;
    const SphereGeom* leaf_geom;
    const SphereGeom* geom;
    const SphereAux* aux;
    const int* large;
    int n_nodes, n_large;
    double local_bound, gain_bound;
};

__device__ __forceinline__ AccelView accel_view(const unsigned char* blob) {
    const AccelHeader* h = reinterpret_cast<const AccelHeader*>(blob);
    AccelView v;
    v.nodes = reinterpret_cast<const AccelNode*>(blob + h->off_nodes);
    v.order = reinterpret_cast<const int*>(blob + h->off_order);
    v.leaf_geom = reinterpret_cast<const SphereGeom*>(blob + h->off_leaf_geom);
    v.geom = reinterpret_cast<const SphereGeom*>(blob + h->off_geom);
    v.aux = reinterpret_cast<const SphereAux*>(blob + h->off_aux);
    v.large = reinterpret_cast<const int*>(blob + h->off_large);
    v.n_nodes = h->n_nodes;
    v.n_large = h->n_large;
    v.local_bound = h->local_bound;
    v.gain_bound = h->gain_bound;
    return v;
}

__device__ __forceinline__ SphereGeom ldg_geom(const SphereGeom* p) {
    const double2 a = __ldg(reinterpret_cast<const double2*>(p));
    const double2 b = __ldg(reinterpret_cast<const double2*>(p) + 1);
    return {a.x, a.y, b.x, b.y};
}

// distance along the ray to one sphere, kMiss on a miss: the policy's own sphere test
template <int MODE>
__device__ __forceinline__ double accel_sphere_t(Vec3 o, Vec3 d, const SphereGeom& g, double cc) {
    if (MODE == kModeExact) return sphere_distance_exact(o, d, g, cc);
    HalfB q = half_b(o, d, g);
    return can_hit(q) ? near_root(q) : kMiss;
}

struct RaySlabs {
    Vec3 o, inv;  // origin and 1/direction (IEEE division: a zero component gives +-inf)
};
__device__ __forceinline__ RaySlabs ray_slabs(Vec3 o, Vec3 d) { return {o, {1.0 / d.x, 1.0 / d.y, 1.0 / d.z}}; }

// does the ray enter the box before `limit`?  fmin/fmax drop the NaN of 0 * inf
__device__ __forceinline__ bool slab_hit(const RaySlabs& r, const AccelNode& n, double limit) {
    const double ax = (n.lo[0] - r.o.x) * r.inv.x, bx = (n.hi[0] - r.o.x) * r.inv.x;
    const double ay = (n.lo[1] - r.o.y) * r.inv.y, by = (n.hi[1] - r.o.y) * r.inv.y;
    const double az = (n.lo[2] - r.o.z) * r.inv.z, bz = (n.hi[2] - r.o.z) * r.inv.z;
    const double t_near = fmax(fmax(fmin(ax, bx), fmin(ay, by)), fmin(az, bz));
    const double t_far = fmin(fmin(fmax(ax, bx), fmax(ay, by)), fmax(az, bz));
    return t_far >= fmax(t_near, 0.0) && t_near < limit;
}

__device__ __forceinline__ AccelNode ldg_node(const AccelNode* p) {
    AccelNode n;
    const double2* q = reinterpret_cast<const double2*>(p);
    const double2 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2);
    const int4 t = __ldg(reinterpret_cast<const int4*>(p) + 3);
    n.lo[0] = a.x, n.lo[1] = a.y, n.lo[2] = b.x;
    n.hi[0] = b.y, n.hi[1] = c.x, n.hi[2] = c.y;
    n.first = t.x;
    n.count = t.y;
    return n;
}

// Nearest hit (base.py:97-105): smallest t, lowest shape index among equal t.
template <int MODE>
__device__ __forceinline__ void accel_nearest(const AccelView& av, Vec3 o, Vec3 d, double& tmin, int& sid,
                                              unsigned long long& evaluated) {
    auto consider = [&](int s, double t) {
        if (t < tmin || (t == tmin && t != kMiss && s < sid)) {
            tmin = t;
            sid = s;
        }
    };
    for (int i = 0; i < av.n_large; ++i) {
        const int s = __ldg(av.large + i);
        consider(s, accel_sphere_t<MODE>(o, d, ldg_geom(av.geom + s), MODE == kModeExact ? __ldg(&av.aux[s].cc) : 0.0));
    }
    evaluated += (unsigned long long)av.n_large;
    if (av.n_nodes == 0) return;
    const RaySlabs ray = ray_slabs(o, d);
    int stack[kAccelStack];
    int top = 0, node = 0;
    while (true) {
        const AccelNode n = ldg_node(av.nodes + node);
        bool descend = false;
        if (slab_hit(ray, n, tmin)) {
            if (n.count > 0) {
                for (int k = 0; k < n.count; ++k) {
                    const int slot = n.first + k;
                    const int s = __ldg(av.order + slot);
                    consider(s, accel_sphere_t<MODE>(o, d, ldg_geom(av.leaf_geom + slot),
                                                     MODE == kModeExact ? __ldg(&av.aux[s].cc) : 0.0));
                }
                evaluated += (unsigned long long)n.count;
            } else {
                // near child first: its hits shrink tmin before the far child is looked at
                const int axis = -n.count - 1;
                const bool backwards = is_neg(axis == 0 ? d.x : (axis == 1 ? d.y : d.z));
                const int near_child = n.first + (backwards ? 1 : 0);
                if (top < kAccelStack) stack[top++] = n.first + (backwards ? 0 : 1);
                node = near_child;
                descend = true;
            }
        }
        if (!descend) {
            if (top == 0) break;
            node = stack[--top];
        }
    }
}

// NumpyShader._calculate_shadow (shader.py:114-128): is any OTHER sphere hit closer than
// `own` (the distance to the point's own sphere, kMiss if the shadow ray leaves it)?
template <int MODE>
__device__ __forceinline__ bool accel_occluded(const AccelView& av, int s_own, Vec3 p, Vec3 l, double own,
                                               unsigned long long& evaluated) {
    for (int i = 0; i < av.n_large; ++i) {
        const int s = __ldg(av.large + i);
        if (s == s_own) continue;
        const double t = accel_sphere_t<MODE>(p, l, ldg_geom(av.geom + s), MODE == kModeExact ? __ldg(&av.aux[s].cc) : 0.0);
        if (t < own) return true;
    }
    evaluated += (unsigned long long)av.n_large;
    if (av.n_nodes == 0) return false;
    const RaySlabs ray = ray_slabs(p, l);
    int stack[kAccelStack];
    int top = 0, node = 0;
    while (true) {
        const AccelNode n = ldg_node(av.nodes + node);
        bool descend = false;
        if (slab_hit(ray, n, own)) {
            if (n.count > 0) {
                for (int k = 0; k < n.count; ++k) {
                    const int slot = n.first + k;
                    const int s = __ldg(av.order + slot);
                    if (s == s_own) continue;
                    const double t = accel_sphere_t<MODE>(p, l, ldg_geom(av.leaf_geom + slot),
                                                          MODE == kModeExact ? __ldg(&av.aux[s].cc) : 0.0);
                    if (t < own) return true;
                }
                evaluated += (unsigned long long)n.count;
            } else {
                if (top < kAccelStack) stack[top++] = n.first + 1;
                node = n.first;
                descend = true;
            }
        }
        if (!descend) {
            if (top == 0) break;
            node = stack[--top];
        }
    }
    return false;
}

}  // namespace rt

// rt_accel.h -- layout of the sphere BVH blob (host builder <-> device traversal).
//
// Nothing in the reference corresponds to this: NumpyRenderer.raytrace_scene intersects every
// ray with every shape (base.py:97) and every shadow ray with every shape again
// (shader.py:126).  The BVH only skips tests that provably miss; the tests it does run are the
// same arithmetic on the same records, so the nearest hit, the shadow decision and the ray
// totals are the ones brute force finds (ties still go to the lowest shape index).
//
// One position-independent blob in device memory (offsets from its start):
//   AccelHeader
//   AccelNode  nodes[n_nodes]        binary tree, root = 0, children of an inner node adjacent
//   int32      order[n_in_tree]      shape index of every leaf slot
//   SphereGeom leaf_geom[n_in_tree]  {cx,cy,cz,r*r} in leaf-slot order (what traversal reads)
//   SphereGeom geom[n_spheres]       the same records by shape index (what shading reads)
//   SphereAux  aux[n_spheres]        {1/r, |C|^2} by shape index
//   ShadeRec   rec[n_spheres]        what the fast policy's shading reads (rt_device.cuh), by shape index
//   int32      large[n_large]        shapes kept out of the tree and tested for every ray
//                                    (a ground sphere of radius 1e5 would swallow every box)
#pragma once

#include <stdint.h>

namespace rt {

constexpr uint32_t kAccelMagic = 0x32425452u;  // "RTB2" (second layout: + the fast policy's shading records)
#ifndef RT_ACCEL_LEAF
#define RT_ACCEL_LEAF 2  // measured: 2 beats 4 by 6 % at 1 025 shapes (fewer FP64 sphere tests per leaf visited)
#endif
constexpr int kAccelLeafSize = RT_ACCEL_LEAF;
constexpr int kAccelMaxLarge = 16;
constexpr int kAccelSahDepth = 36;  // surface-area splits above this depth, median splits below:
constexpr int kAccelStack = 64;     // depth <= 36 + log2(2^26 shapes) = 62 pending far children
constexpr double kAccelMaxCoordinate = 1e15;  // beyond this plane * (1/d capped at 2^64) could overflow FP32: builder refuses

struct AccelHeader {
    uint32_t magic;
    int32_t n_spheres;
    int32_t n_nodes;
    int32_t n_in_tree;
    int32_t n_large;
    int32_t reserved;
    int64_t off_nodes, off_order, off_leaf_geom, off_geom, off_aux, off_large, off_rec;
    int64_t bytes;
    double local_bound;  // sup over materials of |local colour| (material_bound(), rt_chain.cuh)
    double gain_bound;   // sup over materials of |specular_gain|
    float extent;        // sup of |box coordinate| over the tree (scales the slab test's margin)
    float reserved2[3];
};

// 32 bytes.  Boxes are the union of the spheres' FP64 boxes rounded OUTWARD to FP32.  The slab
// test runs in FP32 against an origin shifted by a per-ray margin that bounds every rounding
// in it (rt_accel.cuh: ray_slabs), so a ray the FP64 sphere test accepts is never rejected.
struct alignas(16) AccelNode {
    float lo[3];
    float hi[3];
    uint32_t ref;    // first << 3 | count.  Inner node: first = index of the left child (the right
                     // one follows it), count = 0.  Leaf: first = first leaf slot, count = slots used
                     // (1..kAccelLeafSize <= 7)
    int32_t axis;    // inner: split axis (0..2); leaf: -1.  Not read by the traversal.
};
static_assert(kAccelLeafSize >= 1 && kAccelLeafSize <= 7, "leaf count is packed into 3 bits");

}  // namespace rt

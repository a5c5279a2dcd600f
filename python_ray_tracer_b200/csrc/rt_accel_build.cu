// rt_accel_build.cu -- host-side builder of the sphere BVH blob (rt_accel.h) and its C ABI.
// Plain C++ (no device code): a top-down median split over sphere centres, leaves of up to
// kAccelLeafSize shapes.  Building 1 025 spheres takes tens of microseconds; the blob is
// uploaded once per scene.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include "../../include/rt_abi.h"
#include "rt_accel.h"
#include "rt_device.cuh"

namespace {

using rt::AccelHeader;
using rt::AccelNode;

struct Builder {
    const double* f;  // sphere SoA
    int n;
    std::vector<AccelNode> nodes;
    std::vector<int32_t> order;  // leaf slots
    std::vector<int32_t> items;  // shapes in the tree, permuted in place while splitting

    double centre(int s, int axis) const { return f[(size_t)axis * n + s]; }
    double radius(int s) const { return std::fabs(f[(size_t)RT_F_RADIUS * n + s]); }

    // largest float <= x / smallest float >= x, then one more step outward
    static float round_down(double x) {
        float f = (float)x;
        if ((double)f > x) f = std::nextafterf(f, -INFINITY);
        return std::nextafterf(f, -INFINITY);
    }
    static float round_up(double x) {
        float f = (float)x;
        if ((double)f < x) f = std::nextafterf(f, INFINITY);
        return std::nextafterf(f, INFINITY);
    }

    void bounds(int begin, int end, float* lo_out, float* hi_out) const {
        double lo[3], hi[3];
        for (int a = 0; a < 3; ++a) {
            lo[a] = INFINITY;
            hi[a] = -INFINITY;
        }
        for (int i = begin; i < end; ++i) {
            const int s = items[i];
            const double r = radius(s);
            for (int a = 0; a < 3; ++a) {
                lo[a] = std::min(lo[a], centre(s, a) - r);
                hi[a] = std::max(hi[a], centre(s, a) + r);
            }
        }
        // c -+ r rounds in FP64 (relative 1e-16); the outward FP32 rounding covers it a millionfold
        for (int a = 0; a < 3; ++a) {
            lo_out[a] = round_down(lo[a]);
            hi_out[a] = round_up(hi[a]);
        }
    }

    // fills nodes[node] for items[begin, end)
    void build(int node, int begin, int end) {
        AccelNode nd;
        bounds(begin, end, nd.lo, nd.hi);
        const int count = end - begin;
        if (count <= rt::kAccelLeafSize) {
            nd.ref = ((uint32_t)order.size() << 3) | (uint32_t)count;
            nd.axis = -1;
            for (int i = begin; i < end; ++i) order.push_back(items[i]);
            nodes[node] = nd;
            return;
        }
        double clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY};
        for (int i = begin; i < end; ++i)
            for (int a = 0; a < 3; ++a) {
                clo[a] = std::min(clo[a], centre(items[i], a));
                chi[a] = std::max(chi[a], centre(items[i], a));
            }
        int axis = 0;
        for (int a = 1; a < 3; ++a)
            if (chi[a] - clo[a] > chi[axis] - clo[axis]) axis = a;
        const int mid = begin + count / 2;
        std::nth_element(items.begin() + begin, items.begin() + mid, items.begin() + end, [&](int x, int y) {
            const double cx = centre(x, axis), cy = centre(y, axis);
            return cx < cy || (cx == cy && x < y);
        });
        const int left = (int)nodes.size();
        nodes.push_back(AccelNode());
        nodes.push_back(AccelNode());
        nd.ref = (uint32_t)left << 3;
        nd.axis = axis;
        nodes[node] = nd;
        build(left, begin, mid);
        build(left + 1, mid, end);
    }
};

size_t align16(size_t v) { return (v + 15) & ~(size_t)15; }

// rt_chain.cuh:material_bound() on the host (same expression; it only has to be an upper bound)
double material_bound_host(const double* f, int n, int s, const rt_env* env) {
    const double dg = f[(size_t)RT_F_DIFFUSE_GAIN * n + s], sg = f[(size_t)RT_F_SPECULAR_GAIN * n + s];
    const double rough = f[(size_t)RT_F_SPECULAR_ROUGHNESS * n + s], ig = f[(size_t)RT_F_IRIDESCENCE_GAIN * n + s];
    const bool checker = f[(size_t)RT_F_TEXTURE * n + s] != 0.0;
    const double tex = checker ? 1.0
                               : std::max(std::fabs(f[(size_t)RT_F_COLOR_R * n + s]),
                                          std::max(std::fabs(f[(size_t)RT_F_COLOR_G * n + s]), std::fabs(f[(size_t)RT_F_COLOR_B * n + s])));
    double dome_sum = 0.0;
    for (int i = 0; i < env->n_dome_lights; ++i) dome_sum += env->dome_intensity[i];
    const double dome = std::fabs(dome_sum) * std::max(std::fabs(env->dome_color[0]), std::max(std::fabs(env->dome_color[1]), std::fabs(env->dome_color[2])));
    const double alpha = rough * rough, a2 = alpha * alpha;
    const double dmax = a2 / (rt::kPi * (std::min(1.0, a2 * a2) + rt::kEps));
    const double gv = 1.0 / (2.0 * (std::sqrt(std::min(a2, 1.0)) + rt::kEps));
    const double g = std::fabs(sg);
    return rt::kAmbient + std::fabs(dg) * tex + dome + g * (dmax * gv + g) + 1.4 * rt::kFilmWeight * std::fabs(ig);
}

}  // namespace

extern "C" {

int64_t rt_accel_bytes(int32_t n_spheres) {
    if (n_spheres <= 0) return -1;
    const size_t n = (size_t)n_spheres;
    return (int64_t)(align16(sizeof(AccelHeader)) + align16(sizeof(AccelNode) * (2 * n + 1)) + align16(4 * n) +
                     align16(32 * n) + align16(32 * n) + align16(16 * n) + align16(4 * rt::kAccelMaxLarge) + 64);
}

int rt_accel_build_host(const double* spheres_host, int32_t n_spheres, const rt_env* env, void* blob_out,
                        int64_t capacity, int64_t* bytes_out) {
    if (spheres_host == nullptr || env == nullptr || blob_out == nullptr || bytes_out == nullptr || n_spheres <= 0)
        return RT_ERR_BAD_ARG;
    if (capacity < rt_accel_bytes(n_spheres)) return RT_ERR_BAD_ARG;
    const int n = n_spheres;
    if (n > (1 << 26)) return RT_ERR_UNSUPPORTED;  // traversal packs a node index and a leaf count into 32 bits
    for (int s = 0; s < n; ++s) {
        double reach = std::fabs(spheres_host[(size_t)RT_F_RADIUS * n + s]);
        for (int a = 0; a < 3; ++a) reach = std::max(reach, std::fabs(spheres_host[(size_t)a * n + s]));
        if (!(reach <= rt::kAccelMaxCoordinate)) return RT_ERR_UNSUPPORTED;  // (NaN lands here too)
    }
    Builder b;
    b.f = spheres_host;
    b.n = n;
    // shapes far larger than the typical one stay out of the tree
    std::vector<double> radii(n);
    for (int s = 0; s < n; ++s) radii[s] = b.radius(s);
    std::vector<double> sorted = radii;
    std::nth_element(sorted.begin(), sorted.begin() + n / 2, sorted.end());
    const double median = sorted[n / 2];
    std::vector<int32_t> large;
    for (int s = 0; s < n; ++s) {
        if ((int)large.size() < rt::kAccelMaxLarge && radii[s] > 64.0 * median)
            large.push_back(s);
        else
            b.items.push_back(s);
    }
    if (!b.items.empty()) {
        b.nodes.reserve(2 * b.items.size() + 1);
        b.nodes.push_back(AccelNode());
        b.build(0, 0, (int)b.items.size());
    }

    AccelHeader h;
    memset(&h, 0, sizeof(h));
    h.magic = rt::kAccelMagic;
    h.n_spheres = n;
    h.n_nodes = (int32_t)b.nodes.size();
    h.n_in_tree = (int32_t)b.order.size();
    h.n_large = (int32_t)large.size();
    size_t off = align16(sizeof(AccelHeader));
    h.off_nodes = (int64_t)off;
    off += align16(sizeof(AccelNode) * b.nodes.size());
    h.off_order = (int64_t)off;
    off += align16(4 * b.order.size());
    h.off_leaf_geom = (int64_t)off;
    off += align16(32 * b.order.size());
    h.off_geom = (int64_t)off;
    off += align16(32 * (size_t)n);
    h.off_aux = (int64_t)off;
    off += align16(16 * (size_t)n);
    h.off_large = (int64_t)off;
    off += align16(4 * large.size());
    h.bytes = (int64_t)off;
    if ((int64_t)off > capacity) return RT_ERR_BAD_ARG;
    double lmax = 0.0, gmax = 0.0;
    for (int s = 0; s < n; ++s) {
        lmax = std::max(lmax, material_bound_host(spheres_host, n, s, env));
        gmax = std::max(gmax, std::fabs(spheres_host[(size_t)RT_F_SPECULAR_GAIN * n + s]));
    }
    h.local_bound = lmax;
    h.gain_bound = gmax;
    float extent = 0.0f;
    for (const AccelNode& nd : b.nodes)
        for (int a = 0; a < 3; ++a) extent = std::max(extent, std::max(std::fabs(nd.lo[a]), std::fabs(nd.hi[a])));
    h.extent = extent;

    char* base = static_cast<char*>(blob_out);
    memset(base, 0, off);
    memcpy(base, &h, sizeof(h));
    if (!b.nodes.empty()) memcpy(base + h.off_nodes, b.nodes.data(), sizeof(AccelNode) * b.nodes.size());
    if (!b.order.empty()) memcpy(base + h.off_order, b.order.data(), 4 * b.order.size());
    auto geom_of = [&](int s, double* out) {
        const double r = spheres_host[(size_t)RT_F_RADIUS * n + s];
        out[0] = b.centre(s, 0);
        out[1] = b.centre(s, 1);
        out[2] = b.centre(s, 2);
        out[3] = r * r;
    };
    double* leaf_geom = reinterpret_cast<double*>(base + h.off_leaf_geom);
    for (size_t k = 0; k < b.order.size(); ++k) geom_of(b.order[k], leaf_geom + 4 * k);
    double* geom = reinterpret_cast<double*>(base + h.off_geom);
    double* aux = reinterpret_cast<double*>(base + h.off_aux);
    for (int s = 0; s < n; ++s) {
        geom_of(s, geom + 4 * (size_t)s);
        const double r = spheres_host[(size_t)RT_F_RADIUS * n + s];
        const double cx = b.centre(s, 0), cy = b.centre(s, 1), cz = b.centre(s, 2);
        aux[2 * (size_t)s] = 1.0 / r;                                 // shader.py:74
        aux[2 * (size_t)s + 1] = (cx * cx) + (cy * cy) + (cz * cz);  // shape.py:35
    }
    if (!large.empty()) memcpy(base + h.off_large, large.data(), 4 * large.size());
    *bytes_out = h.bytes;
    return RT_OK;
}

}  // extern "C"

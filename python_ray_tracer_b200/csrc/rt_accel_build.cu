// rt_accel_build.cu -- host-side builder of the sphere BVH blob (rt_accel.h) and its C ABI.
// Plain C++ (no device code): top-down binned surface-area splits (median splits as the fallback
// and below kAccelSahDepth), leaves of up to kAccelLeafSize shapes, all on FP32 boxes rounded
// outward from the FP64 spheres.  The blob is uploaded once per scene.
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

// One shape while the tree is built: its FP64 box rounded OUTWARD to FP32 (node boxes are
// unions of these, and min / max do not round) and its index in the caller's table.  The
// records themselves are permuted while splitting, so every node works on a contiguous run.
struct Item {
    float lo[3];
    float hi[3];
    int32_t index;
    int32_t pad;
};

struct Box {
    float lo[3], hi[3];
    void clear() {
        for (int a = 0; a < 3; ++a) lo[a] = INFINITY, hi[a] = -INFINITY;
    }
    void grow(const float* l, const float* h) {
        for (int a = 0; a < 3; ++a) lo[a] = std::min(lo[a], l[a]), hi[a] = std::max(hi[a], h[a]);
    }
    float half_area() const {
        const float x = hi[0] - lo[0], y = hi[1] - lo[1], z = hi[2] - lo[2];
        return x * y + y * z + z * x;
    }
};

struct Builder {
    std::vector<AccelNode> nodes;
    std::vector<int32_t> order;  // leaf slots
    std::vector<Item> items;     // shapes in the tree
    int sah_depth = rt::kAccelSahDepth;  // 0: median splits only (RT_FLAG_ACCEL_FAST_BUILD)
    int deepest = 0;                     // depth of the deepest node (root = 0)

    // largest float <= x / smallest float >= x, then one more step outward (c -+ r itself
    // rounds in FP64, relative 1e-16: covered a millionfold)
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
    void add(const double* f, int n, int s) {
        Item it;
        const double r = std::fabs(f[(size_t)RT_F_RADIUS * n + s]);
        for (int a = 0; a < 3; ++a) {
            it.lo[a] = round_down(f[(size_t)a * n + s] - r);
            it.hi[a] = round_up(f[(size_t)a * n + s] + r);
        }
        it.index = s;
        it.pad = 0;
        items.push_back(it);
    }
    Box bounds(int begin, int end) const {
        Box b;
        b.clear();
        for (int i = begin; i < end; ++i) b.grow(items[i].lo, items[i].hi);
        return b;
    }

    static constexpr int kBins = 16;
    static int bin_of(const Item& it, int axis, float origin2, float scale) {
        // twice the centroid: (lo + hi) against twice the lower bound
        const int k = (int)(((it.lo[axis] + it.hi[axis]) - origin2) * scale);
        return std::min(kBins - 1, std::max(0, k));
    }

    // Binned surface-area heuristic over all three axes in one pass: the split (axis, plane
    // between two of kBins centroid bins) minimising  area(left) n_left + area(right) n_right.
    // Partitions items[begin, end) and returns the first index of the right half, or -1 if no
    // plane separates the centroids.
    int sah_partition(int begin, int end, int* axis_out, Box* left_out, Box* right_out) {
        float clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY};
        for (int i = begin; i < end; ++i)
            for (int a = 0; a < 3; ++a) {
                const float c2 = items[i].lo[a] + items[i].hi[a];
                clo[a] = std::min(clo[a], c2);
                chi[a] = std::max(chi[a], c2);
            }
        float scale[3];
        bool usable[3];
        for (int a = 0; a < 3; ++a) {
            const float width = chi[a] - clo[a];
            usable[a] = width > 0.0f && std::isfinite(width);
            scale[a] = usable[a] ? kBins / width : 0.0f;
        }
        Box bins[3][kBins];
        int counts[3][kBins];
        for (int a = 0; a < 3; ++a)
            for (int k = 0; k < kBins; ++k) bins[a][k].clear(), counts[a][k] = 0;
        for (int i = begin; i < end; ++i) {
            const Item& it = items[i];
            for (int a = 0; a < 3; ++a) {
                const int k = bin_of(it, a, clo[a], scale[a]);
                bins[a][k].grow(it.lo, it.hi);
                counts[a][k] += 1;
            }
        }
        float best = INFINITY;
        int best_axis = -1, best_plane = 0;
        for (int a = 0; a < 3; ++a) {
            if (!usable[a]) continue;
            float right_cost[kBins];
            Box acc;
            acc.clear();
            int n_acc = 0;
            for (int k = kBins - 1; k > 0; --k) {
                if (counts[a][k]) acc.grow(bins[a][k].lo, bins[a][k].hi), n_acc += counts[a][k];
                right_cost[k] = n_acc ? acc.half_area() * (float)n_acc : -1.0f;
            }
            acc.clear();
            n_acc = 0;
            for (int k = 1; k < kBins; ++k) {  // plane between bins k-1 and k
                if (counts[a][k - 1]) acc.grow(bins[a][k - 1].lo, bins[a][k - 1].hi), n_acc += counts[a][k - 1];
                if (n_acc == 0 || right_cost[k] < 0.0f) continue;
                const float cost = acc.half_area() * (float)n_acc + right_cost[k];
                if (cost < best) best = cost, best_axis = a, best_plane = k;
            }
        }
        if (best_axis < 0) return -1;
        left_out->clear();
        right_out->clear();
        for (int k = 0; k < kBins; ++k)
            if (counts[best_axis][k]) (k < best_plane ? left_out : right_out)->grow(bins[best_axis][k].lo, bins[best_axis][k].hi);
        const float origin2 = clo[best_axis], sc = scale[best_axis];
        auto mid = std::partition(items.begin() + begin, items.begin() + end,
                                  [&](const Item& it) { return bin_of(it, best_axis, origin2, sc) < best_plane; });
        *axis_out = best_axis;
        return (int)(mid - items.begin());
    }

    // Few shapes: the exact surface-area sweep (every split of the centroid order on every axis);
    // binning would spend more time clearing bins than looking at shapes.
    static constexpr int kSmall = 8;
    int small_partition(int begin, int end, int* axis_out, Box* left_out, Box* right_out) {
        const int count = end - begin;
        Item sorted[kSmall], best_order[kSmall];
        Box suffix[kSmall];
        float best = INFINITY;
        int best_axis = -1, best_k = 0;
        for (int a = 0; a < 3; ++a) {
            for (int i = 0; i < count; ++i) {  // insertion sort by (centroid, index)
                const Item it = items[begin + i];
                const float c = it.lo[a] + it.hi[a];
                int j = i;
                while (j > 0) {
                    const float cj = sorted[j - 1].lo[a] + sorted[j - 1].hi[a];
                    if (cj < c || (cj == c && sorted[j - 1].index < it.index)) break;
                    sorted[j] = sorted[j - 1];
                    --j;
                }
                sorted[j] = it;
            }
            Box acc;
            acc.clear();
            for (int i = count - 1; i > 0; --i) {
                acc.grow(sorted[i].lo, sorted[i].hi);
                suffix[i] = acc;
            }
            acc.clear();
            bool improved = false;
            for (int k = 1; k < count; ++k) {  // left = sorted[0, k)
                acc.grow(sorted[k - 1].lo, sorted[k - 1].hi);
                const float cost = acc.half_area() * (float)k + suffix[k].half_area() * (float)(count - k);
                if (cost < best) best = cost, best_axis = a, best_k = k, improved = true, *left_out = acc, *right_out = suffix[k];
            }
            if (improved)
                for (int i = 0; i < count; ++i) best_order[i] = sorted[i];
        }
        if (best_axis < 0) return -1;  // (non-finite boxes only)
        for (int i = 0; i < count; ++i) items[begin + i] = best_order[i];
        *axis_out = best_axis;
        return begin + best_k;
    }

    // median of the centroids along the widest axis: the fallback, and the rule below
    // kAccelSahDepth (it bounds the depth of what remains by log2 of the count)
    int median_partition(int begin, int end, int* axis_out, Box* left_out, Box* right_out) {
        float clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY};
        for (int i = begin; i < end; ++i)
            for (int a = 0; a < 3; ++a) {
                const float c2 = items[i].lo[a] + items[i].hi[a];
                clo[a] = std::min(clo[a], c2);
                chi[a] = std::max(chi[a], c2);
            }
        int axis = 0;
        for (int a = 1; a < 3; ++a)
            if (chi[a] - clo[a] > chi[axis] - clo[axis]) axis = a;
        const int mid = begin + (end - begin) / 2;
        std::nth_element(items.begin() + begin, items.begin() + mid, items.begin() + end, [&](const Item& x, const Item& y) {
            const float cx = x.lo[axis] + x.hi[axis], cy = y.lo[axis] + y.hi[axis];
            return cx < cy || (cx == cy && x.index < y.index);
        });
        *left_out = bounds(begin, mid);
        *right_out = bounds(mid, end);
        *axis_out = axis;
        return mid;
    }

    // fills nodes[node] for items[begin, end), whose box is `box`
    void build(int node, int begin, int end, int depth, const Box& box) {
        AccelNode nd;
        for (int a = 0; a < 3; ++a) nd.lo[a] = box.lo[a], nd.hi[a] = box.hi[a];
        deepest = std::max(deepest, depth);
        const int count = end - begin;
        if (count <= rt::kAccelLeafSize) {
            nd.ref = ((uint32_t)order.size() << 3) | (uint32_t)count;
            nd.axis = -1;
            for (int i = begin; i < end; ++i) order.push_back(items[i].index);
            nodes[node] = nd;
            return;
        }
        int axis = 0;
        Box left_box, right_box;
        int mid = -1;
        if (depth < sah_depth)
            mid = count <= kSmall ? small_partition(begin, end, &axis, &left_box, &right_box)
                                  : sah_partition(begin, end, &axis, &left_box, &right_box);
        if (mid <= begin || mid >= end) mid = median_partition(begin, end, &axis, &left_box, &right_box);
        const int left = (int)nodes.size();
        nodes.push_back(AccelNode());
        nodes.push_back(AccelNode());
        nd.ref = (uint32_t)left << 3;
        nd.axis = axis;
        nodes[node] = nd;
        build(left, begin, mid, depth + 1, left_box);
        build(left + 1, mid, end, depth + 1, right_box);
    }
};

size_t align16(size_t v) { return (v + 15) & ~(size_t)15; }

extern "C" void rt_internal_set_error(const char* message);  // rt_abi.cu: what rt_last_error() returns
int build_fail(int status, const char* message) {
    rt_internal_set_error(message);
    return status;
}

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
                     align16(32 * n) + align16(32 * n) + align16(16 * n) + align16(sizeof(rt::ShadeRec) * n) +
                     align16(4 * rt::kAccelMaxLarge) + 64);
}

int rt_accel_build_host(const double* spheres_host, int32_t n_spheres, const rt_env* env, void* blob_out,
                        int64_t capacity, int64_t* bytes_out) {
    if (spheres_host == nullptr || env == nullptr || blob_out == nullptr || bytes_out == nullptr || n_spheres <= 0)
        return build_fail(RT_ERR_BAD_ARG, "rt_accel_build_host: NULL pointer or no spheres");
    if (capacity < rt_accel_bytes(n_spheres))
        return build_fail(RT_ERR_BAD_ARG, "rt_accel_build_host: capacity is smaller than rt_accel_bytes(n_spheres)");
    const int n = n_spheres;
    if (n > (1 << 26))  // traversal packs a node index and a leaf count into 32 bits
        return build_fail(RT_ERR_UNSUPPORTED, "rt_accel_build_host: more than 2^26 spheres");
    for (int s = 0; s < n; ++s) {
        double reach = std::fabs(spheres_host[(size_t)RT_F_RADIUS * n + s]);
        for (int a = 0; a < 3; ++a) reach = std::max(reach, std::fabs(spheres_host[(size_t)a * n + s]));
        if (!(reach <= rt::kAccelMaxCoordinate))  // (NaN lands here too)
            return build_fail(RT_ERR_UNSUPPORTED, "rt_accel_build_host: a coordinate or radius beyond 1e15 (or NaN): the tree's boxes are FP32");
    }
    Builder b;
    if (env->flags & RT_FLAG_ACCEL_FAST_BUILD) b.sah_depth = 0;
    // shapes far larger than the typical one stay out of the tree
    std::vector<double> radii(n);
    for (int s = 0; s < n; ++s) radii[s] = std::fabs(spheres_host[(size_t)RT_F_RADIUS * n + s]);
    std::vector<double> sorted = radii;
    std::nth_element(sorted.begin(), sorted.begin() + n / 2, sorted.end());
    const double median = sorted[n / 2];
    std::vector<int32_t> large;
    for (int s = 0; s < n; ++s) {
        if ((int)large.size() < rt::kAccelMaxLarge && radii[s] > 64.0 * median)
            large.push_back(s);
        else
            b.add(spheres_host, n, s);
    }
    if (!b.items.empty()) {
        b.nodes.reserve(2 * b.items.size() + 1);
        b.nodes.push_back(AccelNode());
        b.build(0, 0, (int)b.items.size(), 0, b.bounds(0, (int)b.items.size()));
        // the traversal keeps at most one pending sibling per level on its stack
        if (b.deepest + 1 > rt::kAccelStack)
            return build_fail(RT_ERR_UNSUPPORTED, "rt_accel_build_host: the tree came out deeper than the traversal stack of 64");
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
    h.off_rec = (int64_t)off;
    off += align16(sizeof(rt::ShadeRec) * (size_t)n);
    h.off_large = (int64_t)off;
    off += align16(4 * large.size());
    h.bytes = (int64_t)off;
    if ((int64_t)off > capacity) return build_fail(RT_ERR_BAD_ARG, "rt_accel_build_host: capacity too small for the blob");
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
        out[0] = spheres_host[(size_t)RT_F_CX * n + s];
        out[1] = spheres_host[(size_t)RT_F_CY * n + s];
        out[2] = spheres_host[(size_t)RT_F_CZ * n + s];
        out[3] = r * r;
    };
    double* leaf_geom = reinterpret_cast<double*>(base + h.off_leaf_geom);
    for (size_t k = 0; k < b.order.size(); ++k) geom_of(b.order[k], leaf_geom + 4 * k);
    double* geom = reinterpret_cast<double*>(base + h.off_geom);
    double* aux = reinterpret_cast<double*>(base + h.off_aux);
    for (int s = 0; s < n; ++s) {
        geom_of(s, geom + 4 * (size_t)s);
        const double r = spheres_host[(size_t)RT_F_RADIUS * n + s];
        const double cx = geom[4 * (size_t)s], cy = geom[4 * (size_t)s + 1], cz = geom[4 * (size_t)s + 2];
        aux[2 * (size_t)s] = 1.0 / r;                                 // shader.py:74
        aux[2 * (size_t)s + 1] = (cx * cx) + (cy * cy) + (cz * cz);  // shape.py:35
    }
    rt::ShadeRec* rec = reinterpret_cast<rt::ShadeRec*>(base + h.off_rec);
    auto field = [&](int f, int s) { return spheres_host[(size_t)f * n + s]; };
    for (int s = 0; s < n; ++s)
        rec[s] = rt::make_shade_rec(field(RT_F_CX, s), field(RT_F_CY, s), field(RT_F_CZ, s), field(RT_F_RADIUS, s),
                                    field(RT_F_DIFFUSE_GAIN, s), field(RT_F_SPECULAR_GAIN, s), field(RT_F_SPECULAR_ROUGHNESS, s),
                                    field(RT_F_IRIDESCENCE_GAIN, s), field(RT_F_COLOR_R, s), field(RT_F_COLOR_G, s),
                                    field(RT_F_COLOR_B, s), field(RT_F_TEXTURE, s) != 0.0);
    if (!large.empty()) memcpy(base + h.off_large, large.data(), 4 * large.size());
    *bytes_out = h.bytes;
    return RT_OK;
}

}  // extern "C"

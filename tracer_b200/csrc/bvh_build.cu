// Host-side binned-SAH BVH builder over per-surface global AABBs.
//
// Replaces KdTree.build_tree (accel_tree.py:42-146): the reference builds a
// kd-tree over object boundaries with an exact-candidate SAH in Python
// (O(nodes x boundaries)).  Parity for the acceleration structure is defined
// on *results* (accel on/off give identical bundles), so the layout
// is free: a binary BVH, children of a node adjacent, leaves of <= leaf_size
// surfaces.  SAH matters for scenes like the NSTTF field, where one surface
// (the tower receiver) sits far from the rest: a median split would inflate
// every ancestor box of that surface across the empty space in between.
//
// Pure host code (no device calls): the CPU test-suite exercises it directly.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include "common.cuh"

namespace {

constexpr int kBins = 16;
constexpr int kMaxSahDepth = 32;   // deeper levels use median splits: depth <= 32 + log2(n) < 64 (the walk's stack)

struct Box {
    double lo[3], hi[3];
    void reset() { for (int k = 0; k < 3; ++k) { lo[k] = INFINITY; hi[k] = -INFINITY; } }
    void grow(const double *l, const double *h) {
        for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], l[k]); hi[k] = std::max(hi[k], h[k]); }
    }
    void grow(const Box &b) { grow(b.lo, b.hi); }
    double half_area() const {
        double dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
        if (!(dx >= 0.0) || !(dy >= 0.0) || !(dz >= 0.0)) return 0.0;
        return dx * dy + dy * dz + dz * dx;
    }
};

struct Work { int node, begin, end, depth; };

}  // namespace

extern "C" int trc_bvh_build(const double *lo3, const double *hi3, int32_t n, int32_t leaf_size,
                             TrcBvhNode *nodes, int32_t node_capacity, int32_t *n_nodes_out,
                             int32_t *leaf_order) {
    if (!lo3 || !hi3 || n < 1 || leaf_size < 1 || leaf_size > TRC_BVH_MAX_LEAF || !nodes || !n_nodes_out ||
        !leaf_order || node_capacity < 2 * (int64_t)n)
        return fail(TRC_ERR_BAD_ARGUMENT, "trc_bvh_build: bad argument%s");
    std::vector<double> cen((size_t)n * 3);
    for (int64_t i = 0; i < (int64_t)n * 3; ++i) cen[i] = 0.5 * (lo3[i] + hi3[i]);
    std::vector<int32_t> order(n), tmp(n);
    for (int i = 0; i < n; ++i) order[i] = i;
    std::vector<Work> stack;
    stack.push_back({0, 0, n, 0});
    int n_nodes = 1;
    while (!stack.empty()) {
        Work w = stack.back();
        stack.pop_back();
        Box nb; nb.reset();
        Box cb; cb.reset();
        for (int i = w.begin; i < w.end; ++i) {
            int p = order[i];
            nb.grow(lo3 + 3 * (size_t)p, hi3 + 3 * (size_t)p);
            cb.grow(&cen[3 * (size_t)p], &cen[3 * (size_t)p]);
        }
        TrcBvhNode &nd = nodes[w.node];
        memset(&nd, 0, sizeof(nd));
        for (int k = 0; k < 3; ++k) { nd.lo[k] = nb.lo[k]; nd.hi[k] = nb.hi[k]; }
        int cnt = w.end - w.begin;
        if (cnt <= leaf_size) {
            nd.first = w.begin;
            nd.count = cnt;
            continue;
        }
        int mid = -1;
        if (w.depth < kMaxSahDepth) {
            double best_cost = INFINITY;
            int best_axis = -1, best_split = -1;
            double best_scale = 0.0;
            for (int ax = 0; ax < 3; ++ax) {
                double ext = cb.hi[ax] - cb.lo[ax];
                if (!(ext > 0.0)) continue;
                double scale = kBins / ext;
                Box bb[kBins];
                int bc[kBins];
                for (int b = 0; b < kBins; ++b) { bb[b].reset(); bc[b] = 0; }
                for (int i = w.begin; i < w.end; ++i) {
                    int p = order[i];
                    int b = (int)((cen[3 * (size_t)p + ax] - cb.lo[ax]) * scale);
                    b = b < 0 ? 0 : (b >= kBins ? kBins - 1 : b);
                    bb[b].grow(lo3 + 3 * (size_t)p, hi3 + 3 * (size_t)p);
                    bc[b]++;
                }
                double right_area[kBins];
                int right_cnt[kBins];
                Box acc; acc.reset();
                int c = 0;
                for (int b = kBins - 1; b > 0; --b) {
                    acc.grow(bb[b]); c += bc[b];
                    right_area[b] = acc.half_area(); right_cnt[b] = c;
                }
                acc.reset(); c = 0;
                for (int b = 0; b < kBins - 1; ++b) {   // split after bin b
                    acc.grow(bb[b]); c += bc[b];
                    if (c == 0 || right_cnt[b + 1] == 0) continue;
                    double cost = acc.half_area() * c + right_area[b + 1] * right_cnt[b + 1];
                    if (cost < best_cost) { best_cost = cost; best_axis = ax; best_split = b; best_scale = scale; }
                }
            }
            if (best_axis >= 0) {
                // stable partition keeps the builder deterministic
                int a = w.begin, z = 0;
                for (int i = w.begin; i < w.end; ++i) {
                    int p = order[i];
                    int b = (int)((cen[3 * (size_t)p + best_axis] - cb.lo[best_axis]) * best_scale);
                    b = b < 0 ? 0 : (b >= kBins ? kBins - 1 : b);
                    if (b <= best_split) order[a++] = p; else tmp[z++] = p;
                }
                memcpy(&order[a], tmp.data(), sizeof(int32_t) * (size_t)z);
                mid = a;
            }
        }
        if (mid <= w.begin || mid >= w.end) {
            // median split on the widest centroid axis (or by index when all centroids coincide)
            int ax = 0;
            for (int k = 1; k < 3; ++k)
                if (cb.hi[k] - cb.lo[k] > cb.hi[ax] - cb.lo[ax]) ax = k;
            mid = w.begin + cnt / 2;
            std::nth_element(order.begin() + w.begin, order.begin() + mid, order.begin() + w.end,
                             [&](int32_t p, int32_t q) {
                                 double cp = cen[3 * (size_t)p + ax], cq = cen[3 * (size_t)q + ax];
                                 return cp < cq || (cp == cq && p < q);
                             });
        }
        if (n_nodes + 2 > node_capacity) return fail(TRC_ERR_CAPACITY, "trc_bvh_build: node capacity exceeded%s");
        int left = n_nodes;
        n_nodes += 2;
        nd.first = left;
        nd.count = 0;
        stack.push_back({left + 1, mid, w.end, w.depth + 1});
        stack.push_back({left, w.begin, mid, w.depth + 1});
    }
    memcpy(leaf_order, order.data(), sizeof(int32_t) * (size_t)n);
    *n_nodes_out = n_nodes;
    return TRC_OK;
}

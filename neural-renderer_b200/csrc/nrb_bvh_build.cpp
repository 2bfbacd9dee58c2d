// nrb_bvh_build.cpp — host SAH BVH builder behind nrb_build_bvh.
//
// Produces what the reference's Mesh::buildBvh yields (src/mesh_bvh.cpp:27-162): a flattened
// binary BvhNode[] whose inner nodes have right == left + 1, leaves hold [first, first+count) of an
// index buffer permuted into leaf order.  The reference delegates the build to the vendored
// madmann91/bvh library; this is an independent top-down binned-SAH builder (16 bins, leaves of at
// most 4 triangles) for callers that do not link that library (synthetic scenes, the bench).  Tree
// shape differs from ext/bvh's, which is fine: primitive ids are defined by whichever BVH is handed
// to nrb_update_mesh, and the parity tests hand the SAME nodes to the oracle and to the device.
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <numeric>
#include <vector>

#include "../../include/nrb200.h"

namespace {

struct Box {
    float mn[3], mx[3];
    void reset() {
        for (int k = 0; k < 3; ++k) { mn[k] = FLT_MAX; mx[k] = -FLT_MAX; }
    }
    void grow(const float* p) {
        for (int k = 0; k < 3; ++k) { mn[k] = std::min(mn[k], p[k]); mx[k] = std::max(mx[k], p[k]); }
    }
    void grow(const Box& b) {
        for (int k = 0; k < 3; ++k) { mn[k] = std::min(mn[k], b.mn[k]); mx[k] = std::max(mx[k], b.mx[k]); }
    }
    float half_area() const {
        const float dx = mx[0] - mn[0], dy = mx[1] - mn[1], dz = mx[2] - mn[2];
        return dx < 0.0f ? 0.0f : dx * dy + dy * dz + dz * dx;
    }
};

constexpr int kBins = 16;
constexpr int kMaxLeaf = 4;

struct Task {
    int node, begin, end;
};

}  // namespace

extern "C" int nrb_build_bvh(const float* vertices, int num_vertices, uint32_t* indices, int num_triangles,
                             nrb_bvh_node* nodes, int* num_nodes, uint32_t* prim_ids) {
    if (!vertices || !indices || !nodes || !num_nodes || num_triangles <= 0 || num_vertices <= 0) return NRB_ERR_INVALID;
    const int n = num_triangles;
    std::vector<Box> boxes(n);
    std::vector<float> centers(3 * static_cast<size_t>(n));
    for (int i = 0; i < n; ++i) {
        Box b;
        b.reset();
        for (int c = 0; c < 3; ++c) {
            const uint32_t vi = indices[3 * i + c];
            if (vi >= static_cast<uint32_t>(num_vertices)) return NRB_ERR_INVALID;
            b.grow(vertices + 3 * static_cast<size_t>(vi));
        }
        boxes[i] = b;
        for (int k = 0; k < 3; ++k) centers[3 * static_cast<size_t>(i) + k] = 0.5f * (b.mn[k] + b.mx[k]);
    }
    std::vector<uint32_t> order(n);
    std::iota(order.begin(), order.end(), 0u);

    int count = 1;
    std::vector<Task> todo;
    todo.push_back({0, 0, n});
    while (!todo.empty()) {
        const Task t = todo.back();
        todo.pop_back();
        Box bounds, cbounds;
        bounds.reset();
        cbounds.reset();
        for (int i = t.begin; i < t.end; ++i) {
            bounds.grow(boxes[order[i]]);
            cbounds.grow(&centers[3 * static_cast<size_t>(order[i])]);
        }
        nrb_bvh_node& out = nodes[t.node];
        std::memcpy(out.bounds_min, bounds.mn, 12);
        std::memcpy(out.bounds_max, bounds.mx, 12);
        const int m = t.end - t.begin;
        auto make_leaf = [&]() {
            out.is_leaf = 1;
            out.first = t.begin;
            out.count = m;
            out.left = out.right = -1;
        };
        if (m <= kMaxLeaf) {
            // a leaf unless splitting is clearly cheaper; small leaves keep traversal short
            if (m <= 2) { make_leaf(); continue; }
        }
        // binned SAH over the three axes
        int bestAxis = -1, bestBin = -1;
        float bestCost = FLT_MAX;
        for (int axis = 0; axis < 3; ++axis) {
            const float lo = cbounds.mn[axis], hi = cbounds.mx[axis];
            if (!(hi > lo)) continue;
            const float scale = kBins / (hi - lo);
            Box bb[kBins];
            int bc[kBins] = {0};
            for (auto& b : bb) b.reset();
            for (int i = t.begin; i < t.end; ++i) {
                const uint32_t p = order[i];
                int b = static_cast<int>((centers[3 * static_cast<size_t>(p) + axis] - lo) * scale);
                b = std::min(kBins - 1, std::max(0, b));
                bb[b].grow(boxes[p]);
                ++bc[b];
            }
            float rightArea[kBins];
            int rightCount[kBins];
            Box acc;
            acc.reset();
            int c = 0;
            for (int b = kBins - 1; b > 0; --b) {
                acc.grow(bb[b]);
                c += bc[b];
                rightArea[b] = acc.half_area();
                rightCount[b] = c;
            }
            acc.reset();
            c = 0;
            for (int b = 0; b < kBins - 1; ++b) {
                acc.grow(bb[b]);
                c += bc[b];
                if (c == 0 || rightCount[b + 1] == 0) continue;
                const float cost = acc.half_area() * c + rightArea[b + 1] * rightCount[b + 1];
                if (cost < bestCost) { bestCost = cost; bestAxis = axis; bestBin = b; }
            }
        }
        int mid;
        if (bestAxis < 0) {
            if (m <= kMaxLeaf) { make_leaf(); continue; }
            mid = t.begin + m / 2;  // coincident centroids: split by count
        } else {
            const float leafCost = bounds.half_area() * m;
            if (m <= kMaxLeaf && bestCost >= leafCost) { make_leaf(); continue; }
            const float lo = cbounds.mn[bestAxis], scale = kBins / (cbounds.mx[bestAxis] - lo);
            auto it = std::partition(order.begin() + t.begin, order.begin() + t.end, [&](uint32_t p) {
                int b = static_cast<int>((centers[3 * static_cast<size_t>(p) + bestAxis] - lo) * scale);
                b = std::min(kBins - 1, std::max(0, b));
                return b <= bestBin;
            });
            mid = static_cast<int>(it - order.begin());
            if (mid == t.begin || mid == t.end) mid = t.begin + m / 2;
        }
        out.is_leaf = 0;
        out.first = out.count = 0;
        out.left = count;
        out.right = count + 1;
        count += 2;
        todo.push_back({out.right, mid, t.end});
        todo.push_back({out.left, t.begin, mid});
    }
    // permute the index buffer into leaf order
    std::vector<uint32_t> tmp(indices, indices + 3 * static_cast<size_t>(n));
    for (int i = 0; i < n; ++i) {
        const uint32_t src = order[i];
        indices[3 * i] = tmp[3 * static_cast<size_t>(src)];
        indices[3 * i + 1] = tmp[3 * static_cast<size_t>(src) + 1];
        indices[3 * i + 2] = tmp[3 * static_cast<size_t>(src) + 2];
        if (prim_ids) prim_ids[i] = src;
    }
    *num_nodes = count;
    return NRB_OK;
}

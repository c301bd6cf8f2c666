// Host-side set-up helpers of the histogram path for callers that do not link the reference's own
// libs/geom + libs/acc (the Python mirror, bench.py, C users): the direction set of the spherical histogram
// and a KD-tree over it.  In the reference both are built by the apps before any kernel runs
// (apps/optimize_trajectory/optimize_trajectory.cu:157-164: generate_sphere_mesh(1.0f, 3u), acc::KDTree<3>).
#include <algorithm>
#include <deque>
#include <unordered_map>

#include "common.cuh"

namespace {

struct P3 { float x, y, z; };

inline P3 midpoint(P3 a, P3 b) { return P3{(a.x + b.x) / 2.0f, (a.y + b.y) / 2.0f, (a.z + b.z) / 2.0f}; }

}  // namespace

extern "C" {

// The vertex list of generate_sphere_mesh(radius, subdivisions), libs/geom/sphere.h:17-115, in its order.
// Every face contributes three midpoint vertices whether or not the neighbour across the edge already made one; the
// connectivity only ever uses the FIRST vertex registered for an edge, so each shared edge leaves an unreferenced
// coincident twin behind: 12 + 3 (20 + 80 + 320) = 1272 vertices after three subdivisions, 642 distinct positions.
// The histogram kernels bin into all of them (the apps size sphere_hist by verts.size()), so the list -- twins and
// order included -- is an input of K4 / K5 / K6 / K7.
int uavmvs_sphere_vertices(float radius, uint32_t subdivisions, float * verts, uint32_t capacity, uint32_t * count)
{
    UAV_REQUIRE(count, "null argument");
    UAV_REQUIRE(subdivisions <= 7, "more than 7 subdivisions");
    const float s = 0.5257311f, t = 0.8506508f;
    std::vector<P3> v = {{0, -s, t}, {0, s, t}, {0, -s, -t}, {0, s, -t}, {t, 0, s}, {t, 0, -s},
                         {-t, 0, s}, {-t, 0, -s}, {s, t, 0}, {s, -t, 0}, {-s, t, 0}, {-s, -t, 0}};
    std::vector<uint32_t> f = {0, 4, 1, 0, 9, 4, 9, 5, 4, 4, 5, 8, 4, 8, 1, 8, 10, 1, 8, 3, 10, 5, 3, 8, 5, 2, 3, 2, 7, 3,
                               7, 10, 3, 7, 6, 10, 7, 11, 6, 11, 0, 6, 0, 1, 6, 6, 1, 10, 9, 0, 11, 9, 11, 2, 9, 2, 5, 7, 2, 11};
    for (uint32_t level = 0; level < subdivisions; ++level) {
        std::unordered_map<uint64_t, uint32_t> first_on_edge;
        auto key = [](uint32_t a, uint32_t b) { return ((uint64_t)std::min(a, b) << 32) | std::max(a, b); };
        const size_t nf = f.size() / 3;
        for (size_t j = 0; j < nf; ++j)
            for (int e = 0; e < 3; ++e) {
                const uint32_t a = f[3 * j + e], b = f[3 * j + (e + 1) % 3];
                first_on_edge.emplace(key(a, b), (uint32_t)v.size());     // keeps the earlier entry, if any
                v.push_back(midpoint(v[a], v[b]));                        // pushed regardless: the twin
            }
        for (size_t j = 0; j < nf; ++j) {
            const uint32_t a = f[3 * j], b = f[3 * j + 1], c = f[3 * j + 2];
            const uint32_t ab = first_on_edge[key(a, b)], bc = first_on_edge[key(b, c)], ca = first_on_edge[key(c, a)];
            const uint32_t add[9] = {ab, b, bc, bc, c, ca, ca, a, ab};
            f.insert(f.end(), add, add + 9);
            f[3 * j] = ab; f[3 * j + 1] = bc; f[3 * j + 2] = ca;
        }
    }
    *count = (uint32_t)v.size();
    if (!verts) return 0;
    UAV_REQUIRE(capacity >= v.size(), "vertex buffer too small");
    for (size_t i = 0; i < v.size(); ++i) {
        // normalized() = v / norm(v), norm = sqrt(((0 + x x) + y y) + z z); then * radius  (sphere.h:100-103)
        float acc = 0.0f;
        acc = acc + v[i].x * v[i].x; acc = acc + v[i].y * v[i].y; acc = acc + v[i].z * v[i].z;
        const float len = std::sqrt(acc);
        verts[3 * i] = (v[i].x / len) * radius; verts[3 * i + 1] = (v[i].y / len) * radius; verts[3 * i + 2] = (v[i].z / len) * radius;
    }
    return 0;
}

// Median-split KD-tree over n points in the node layout of uavmvs_kdtree_create ({vid, dim, left, right}); the rule of
// acc::KDTree (libs/acc/kd_tree.h:92-158, single-threaded form): split dimension cycles x, y, z with depth, the node
// keeps the median of its range (std::nth_element on the coordinate), children are numbered in creation order, breadth
// first.  Any tree of this shape answers find_nn identically except between coincident points, where the first one met
// in depth-first order wins (nnsearch.h:97).
int uavmvs_kdtree_build(const float * verts, uint32_t n, uavmvs_kdtree ** out)
{
    UAV_REQUIRE(verts && out && n > 0, "null / empty argument");
    struct Range { uint32_t id, first, last, dim; };
    std::vector<uint32_t> order(n), nodes(4 * (size_t)n);
    for (uint32_t i = 0; i < n; ++i) order[i] = i;
    std::deque<Range> todo;
    uint32_t next_id = 0;
    todo.push_back(Range{next_id++, 0u, n, 0u});
    while (!todo.empty()) {
        const Range r = todo.front(); todo.pop_front();
        const uint32_t mid = (r.first + r.last) / 2, d = r.dim;
        std::nth_element(order.begin() + r.first, order.begin() + mid, order.begin() + r.last,
            [verts, d](uint32_t a, uint32_t b) { return verts[3 * (size_t)a + d] < verts[3 * (size_t)b + d]; });
        uint32_t left = UAV_NAI, right = UAV_NAI;
        if (mid > r.first) { left = next_id++; todo.push_back(Range{left, r.first, mid, (d + 1) % 3}); }
        if (r.last > mid + 1) { right = next_id++; todo.push_back(Range{right, mid + 1, r.last, (d + 1) % 3}); }
        nodes[4 * (size_t)r.id] = order[mid]; nodes[4 * (size_t)r.id + 1] = d;
        nodes[4 * (size_t)r.id + 2] = left; nodes[4 * (size_t)r.id + 3] = right;
    }
    return uavmvs_kdtree_create(nodes.data(), verts, n, out);
}

}  // extern "C"

/*
 * oracle/_ref/libref_acc.so — the REFERENCE's own CPU code for the tree steps
 * of the hot path, compiled from where it lies under /root/reference:
 *   libs/acc/bvh_tree.h   (SAH build :335-399, intersect :419-467)
 *   libs/acc/primitives.h (slab :119-132, triangle :214-248)
 *   libs/acc/kd_tree.h    (build :92-158, find_nn(s) :160-256)
 * No reference source is copied; this file only #includes those headers and
 * re-exports them through a C ABI so tests can pin oracle/uavmvs_oracle.cpp
 * against them.  TEST INFRASTRUCTURE ONLY (never linked by the product).
 *
 * The two `convert<>` specialisations below are the extension point the
 * reference itself uses (libs/cacc/bvh_tree.cu:17-48, kd_tree.cu:15-40) to
 * reach the private arrays.
 */
#include <memory>
#include <tuple>
#include <cstdint>
#include <cstring>
#include <vector>

#include <math/vector.h>
#include "acc/bvh_tree.h"
#include "acc/kd_tree.h"

typedef acc::BVHTree<unsigned, math::Vec3f> RefBVH;
typedef acc::KDTree<3, unsigned> RefKD;

struct BVHDump {
    std::vector<uint32_t> nodes;   /* first,last,left,right per node */
    std::vector<float> aabbs;      /* min xyz, max xyz per node */
    std::vector<float> tris;       /* a,b,c per triangle (leaf order) */
    std::vector<uint32_t> indices; /* original face id per leaf-order triangle */
};

struct KDDump {
    std::vector<uint32_t> nodes;   /* vid,dim,left,right per node */
    std::vector<float> verts;      /* xyz per vertex */
};

namespace acc {
template <> template <>
BVHDump RefBVH::convert<BVHDump>(RefBVH const & t) {
    BVHDump d;
    std::size_t n = t.num_nodes;
    d.nodes.resize(4 * n); d.aabbs.resize(6 * n);
    for (std::size_t i = 0; i < n; ++i) {
        d.nodes[4 * i + 0] = t.nodes[i].first;
        d.nodes[4 * i + 1] = t.nodes[i].last;
        d.nodes[4 * i + 2] = t.nodes[i].left;
        d.nodes[4 * i + 3] = t.nodes[i].right;
        for (int k = 0; k < 3; ++k) {
            d.aabbs[6 * i + k] = t.nodes[i].aabb.min[k];
            d.aabbs[6 * i + 3 + k] = t.nodes[i].aabb.max[k];
        }
    }
    std::size_t m = t.tris.size();
    d.tris.resize(9 * m); d.indices.resize(m);
    for (std::size_t i = 0; i < m; ++i) {
        d.indices[i] = t.indices[i];
        for (int k = 0; k < 3; ++k) {
            d.tris[9 * i + k] = t.tris[i].a[k];
            d.tris[9 * i + 3 + k] = t.tris[i].b[k];
            d.tris[9 * i + 6 + k] = t.tris[i].c[k];
        }
    }
    return d;
}

template <> template <>
KDDump RefKD::convert<KDDump>(RefKD const & t) {
    KDDump d;
    std::size_t n = t.num_nodes;
    d.nodes.resize(4 * n); d.verts.resize(3 * t.vertices.size());
    for (std::size_t i = 0; i < n; ++i) {
        d.nodes[4 * i + 0] = t.nodes[i].vertex_id;
        d.nodes[4 * i + 1] = t.nodes[i].d;
        d.nodes[4 * i + 2] = t.nodes[i].left;
        d.nodes[4 * i + 3] = t.nodes[i].right;
    }
    for (std::size_t i = 0; i < t.vertices.size(); ++i)
        for (int k = 0; k < 3; ++k) d.verts[3 * i + k] = t.vertices[i][k];
    return d;
}
}  // namespace acc

struct RefBVHHandle { RefBVH::Ptr tree; BVHDump dump; };
struct RefKDHandle { RefKD::Ptr tree; KDDump dump; };

extern "C" {

void * ref_bvh_build(float const * verts, uint32_t num_verts,
    uint32_t const * faces, uint32_t num_faces, int threads)
{
    std::vector<math::Vec3f> v(num_verts);
    for (uint32_t i = 0; i < num_verts; ++i) v[i] = math::Vec3f(verts + 3 * i);
    std::vector<unsigned> f(faces, faces + 3 * (std::size_t)num_faces);
    RefBVHHandle * h = new RefBVHHandle;
    h->tree = RefBVH::create(f, v, threads);
    h->dump = RefBVH::convert<BVHDump>(*h->tree);
    return h;
}

uint32_t ref_bvh_num_nodes(void * hp) { return ((RefBVHHandle *)hp)->dump.nodes.size() / 4; }
uint32_t ref_bvh_num_tris(void * hp) { return ((RefBVHHandle *)hp)->dump.indices.size(); }

void ref_bvh_export(void * hp, uint32_t * nodes, float * aabbs, float * tris, uint32_t * indices)
{
    BVHDump const & d = ((RefBVHHandle *)hp)->dump;
    std::memcpy(nodes, d.nodes.data(), d.nodes.size() * 4);
    std::memcpy(aabbs, d.aabbs.data(), d.aabbs.size() * 4);
    std::memcpy(tris, d.tris.data(), d.tris.size() * 4);
    std::memcpy(indices, d.indices.data(), d.indices.size() * 4);
}

/* rays: origin xyz, dir xyz, tmin, tmax (8 floats each).
 * hit[i] = 1/0; face[i] = original face id or 0xffffffff; t[i] = hit distance. */
void ref_bvh_intersect(void * hp, float const * rays, uint64_t num_rays,
    uint8_t * hit, uint32_t * face, float * t)
{
    RefBVH const & tree = *((RefBVHHandle *)hp)->tree;
    #pragma omp parallel for schedule(dynamic, 1024)
    for (int64_t i = 0; i < (int64_t)num_rays; ++i) {
        RefBVH::Ray ray;
        ray.origin = math::Vec3f(rays + 8 * i);
        ray.dir = math::Vec3f(rays + 8 * i + 3);
        ray.tmin = rays[8 * i + 6];
        ray.tmax = rays[8 * i + 7];
        RefBVH::Hit h;
        h.idx = 0xffffffffu; h.t = 0.0f;
        bool r = tree.intersect(ray, &h);
        if (hit) hit[i] = r ? 1 : 0;
        if (face) face[i] = r ? h.idx : 0xffffffffu;
        if (t) t[i] = r ? h.t : 0.0f;
    }
}

/* acc::BVHTree::closest_point(vertex, &cp, max_dist), libs/acc/bvh_tree.h:492-548 */
void ref_bvh_closest_point(void * hp, float const * points, uint64_t n, float max_dist,
    uint8_t * found, uint32_t * idx, float * cp)
{
    RefBVH const & tree = *((RefBVHHandle *)hp)->tree;
    #pragma omp parallel for schedule(dynamic, 256)
    for (int64_t i = 0; i < (int64_t)n; ++i) {
        std::pair<unsigned, math::Vec3f> r;
        r.first = 0xffffffffu; r.second = math::Vec3f(0.0f);
        bool f = tree.closest_point(math::Vec3f(points + 3 * i), &r, max_dist);
        found[i] = f ? 1 : 0; idx[i] = f ? r.first : 0xffffffffu;
        for (int k = 0; k < 3; ++k) cp[3 * i + k] = f ? r.second[k] : 0.0f;
    }
}

void ref_bvh_destroy(void * hp) { delete (RefBVHHandle *)hp; }

void * ref_kd_build(float const * verts, uint32_t num_verts, int threads)
{
    std::vector<math::Vec3f> v(num_verts);
    for (uint32_t i = 0; i < num_verts; ++i) v[i] = math::Vec3f(verts + 3 * i);
    RefKDHandle * h = new RefKDHandle;
    h->tree = RefKD::create(v, threads);
    h->dump = RefKD::convert<KDDump>(*h->tree);
    return h;
}

uint32_t ref_kd_num_nodes(void * hp) { return ((RefKDHandle *)hp)->dump.nodes.size() / 4; }

void ref_kd_export(void * hp, uint32_t * nodes, float * verts)
{
    KDDump const & d = ((RefKDHandle *)hp)->dump;
    std::memcpy(nodes, d.nodes.data(), d.nodes.size() * 4);
    std::memcpy(verts, d.verts.data(), d.verts.size() * 4);
}

void ref_kd_find_nn(void * hp, float const * queries, uint64_t n, uint32_t * idx, float * dist)
{
    RefKD const & tree = *((RefKDHandle *)hp)->tree;
    for (uint64_t i = 0; i < n; ++i) {
        std::pair<unsigned, float> nn(0xffffffffu, 0.0f);
        tree.find_nn(math::Vec3f(queries + 3 * i), &nn);
        idx[i] = nn.first;
        if (dist) dist[i] = nn.second;
    }
}

void ref_kd_destroy(void * hp) { delete (RefKDHandle *)hp; }

}  // extern "C"

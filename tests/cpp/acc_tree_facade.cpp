// An application that builds its trees with the reference's own libs/acc (acc::BVHTree, acc::KDTree) and hands them to
// the device through the ORIGINAL factory calls of libs/cacc -- cacc::BVHTree<DEVICE>::create<uint, math::Vec3f>(tree),
// cacc::KDTree<3, DEVICE>::create<uint>(kd) -- now provided by include/uavmvs_b200/cacc.hpp.  Self-checking: the device
// answers (closest-hit trace through the C ABI, nearest neighbour) must equal acc's own CPU answers bit for bit.
// Built by oracle/Makefile (target accdemo) into oracle/_ref/acc_tree_facade, since it needs the reference's headers.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <tuple>
#include <vector>

#include "math/vector.h"
#include "acc/bvh_tree.h"
#include "acc/kd_tree.h"
#include "uavmvs_b200/cacc.hpp"

typedef unsigned int uint;

static float frand(unsigned & s) { s = s * 1664525u + 1013904223u; return (s >> 8) * (1.0f / 16777216.0f); }

int main() {
    unsigned seed = 12345u;
    // a bumpy 40 x 40 height field: 3200 triangles
    const int G = 40;
    std::vector<math::Vec3f> verts;
    std::vector<uint> faces;
    for (int y = 0; y <= G; ++y) for (int x = 0; x <= G; ++x) verts.push_back(math::Vec3f(x * 0.5f, y * 0.5f, 2.0f * frand(seed)));
    for (int y = 0; y < G; ++y) for (int x = 0; x < G; ++x) {
        uint a = y * (G + 1) + x, b = a + 1, c = a + G + 1, d = c + 1;
        faces.push_back(a); faces.push_back(b); faces.push_back(d);
        faces.push_back(a); faces.push_back(d); faces.push_back(c);
    }
    typedef acc::BVHTree<uint, math::Vec3f> Tree;
    Tree::Ptr tree = Tree::create(faces, verts, 2);
    acc::KDTree<3, uint>::Ptr kd = acc::KDTree<3, uint>::create(verts, 1);

    cacc::select_cuda_device(3, 5);
    cacc::BVHTree<cacc::DEVICE>::Ptr dtree = cacc::BVHTree<cacc::DEVICE>::create<uint, math::Vec3f>(tree);
    cacc::KDTree<3u, cacc::DEVICE>::Ptr dkd = cacc::KDTree<3u, cacc::DEVICE>::create<uint>(kd);

    const int R = 20000;
    std::vector<float> rays(8 * R), t(R), q(3 * R);
    std::vector<unsigned char> hit(R);
    std::vector<uint32_t> face(R), nn(R);
    int mism = 0, hits = 0;
    for (int i = 0; i < R; ++i) {
        math::Vec3f o(20.0f * frand(seed), 20.0f * frand(seed), 3.0f + 5.0f * frand(seed));
        math::Vec3f d(frand(seed) - 0.5f, frand(seed) - 0.5f, -0.2f - frand(seed));
        d = d / d.norm();
        for (int k = 0; k < 3; ++k) { rays[8 * i + k] = o[k]; rays[8 * i + 3 + k] = d[k]; q[3 * i + k] = o[k]; }
        rays[8 * i + 6] = 0.0f; rays[8 * i + 7] = 100.0f;
    }
    UAVMVS_CHECK(uavmvs_bvh_trace(nullptr, dtree->accessor().h, rays.data(), R, 0, hit.data(), face.data(), t.data()));
    UAVMVS_CHECK(uavmvs_kdtree_find_nn(nullptr, const_cast<uavmvs_kdtree *>(dkd->accessor().h), q.data(), R, 0, nn.data()));
    for (int i = 0; i < R; ++i) {
        Tree::Ray ray;
        ray.origin = math::Vec3f(rays[8 * i], rays[8 * i + 1], rays[8 * i + 2]);
        ray.dir = math::Vec3f(rays[8 * i + 3], rays[8 * i + 4], rays[8 * i + 5]);
        ray.tmin = 0.0f; ray.tmax = 100.0f;
        Tree::Hit h;
        const bool ok = tree->intersect(ray, &h);
        hits += ok;
        if (ok != (hit[i] != 0)) { ++mism; continue; }
        if (ok && (h.idx != face[i] || std::memcmp(&h.t, &t[i], 4) != 0)) ++mism;
        std::pair<uint, float> cpu_nn;
        kd->find_nn(math::Vec3f(q[3 * i], q[3 * i + 1], q[3 * i + 2]), &cpu_nn);
        if (cpu_nn.first != nn[i]) {
            // acc breaks distance ties with <=, the device search with <: only coincident vertices may differ
            math::Vec3f a = verts[cpu_nn.first] - verts[nn[i]];
            if (a.norm() != 0.0f) ++mism;
        }
    }
    std::printf("%s rays %d hits %d mismatches %d\n", mism == 0 ? "OK" : "FAIL", R, hits, mism);
    return mism == 0 ? 0 : 1;
}

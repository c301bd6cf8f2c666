// The launch sequence of apps/evaluate_trajectory/evaluate_trajectory.cu:125-202 written against
// include/uavmvs_b200/cacc.hpp: per-view update_observation_rays on a stream, then
// process / evaluate / calculate_func_recons / reduction.  Inputs come from a flat binary file
// written by tests/test_cpp_facade.py; results go to stdout for the test to compare with the oracle.
#include <cstdint>
#include <cstdio>
#include <vector>

#include "uavmvs_b200/cacc.hpp"

template <typename T> static std::vector<T> rd(FILE * f, size_t n) {
    std::vector<T> v(n);
    if (n && fread(v.data(), sizeof(T), n, f) != n) { std::fprintf(stderr, "short read\n"); std::exit(2); }
    return v;
}

int main(int argc, char ** argv) {
    if (argc < 2) return 2;
    FILE * f = fopen(argv[1], "rb");
    if (!f) return 2;
    std::vector<uint32_t> hdr = rd<uint32_t>(f, 6);   // nodes, tris, points, views, width, height
    uint32_t nn = hdr[0], nt = hdr[1], np = hdr[2], nv = hdr[3];
    std::vector<float> par = rd<float>(f, 2);          // max_distance, target_recon
    std::vector<uint32_t> nodes = rd<uint32_t>(f, 4 * (size_t)nn);
    std::vector<float> aabbs = rd<float>(f, 6 * (size_t)nn), tris = rd<float>(f, 9 * (size_t)nt);
    std::vector<uint32_t> indices = rd<uint32_t>(f, nt);
    std::vector<float> pts = rd<float>(f, 3 * (size_t)np), nrm = rd<float>(f, 3 * (size_t)np);
    std::vector<float> views = rd<float>(f, 28 * (size_t)nv);
    fclose(f);

    cacc::select_cuda_device(3, 5);
    cacc::BVHTree<cacc::DEVICE>::Ptr dbvh_tree = cacc::BVHTree<cacc::DEVICE>::create(nodes.data(), aabbs.data(),
        tris.data(), indices.data(), nn, nt);
    cacc::PointCloud<cacc::HOST>::Ptr cloud = cacc::PointCloud<cacc::HOST>::create(np);
    cacc::PointCloud<cacc::HOST>::Data data = cloud->cdata();
    for (size_t i = 0; i < 3 * (size_t)np; ++i) { data.vertices_ptr[i] = pts[i]; data.normals_ptr[i] = nrm[i]; }
    cacc::PointCloud<cacc::DEVICE>::Ptr dcloud = cacc::PointCloud<cacc::DEVICE>::create<cacc::HOST>(cloud);
    cacc::VectorArray<cacc::Vec3f, cacc::DEVICE>::Ptr dobs_rays = cacc::VectorArray<cacc::Vec3f, cacc::DEVICE>::create(np, 32);
    cacc::Array<float, cacc::DEVICE>::Ptr drecons = cacc::Array<float, cacc::DEVICE>::create(np);
    cacc::Array<float, cacc::DEVICE>::Ptr dwrecons = cacc::Array<float, cacc::DEVICE>::create(np);
    cacc::Stream stream = nullptr;

    for (uint32_t k = 0; k < nv; ++k) {
        const float * v = views.data() + 28 * (size_t)k;
        cacc::Vec3f pos; cacc::Mat4f w2c; cacc::Mat3f calib;
        for (int i = 0; i < 3; ++i) pos[i] = v[i];
        for (int i = 0; i < 16; ++i) w2c.m[i] = v[3 + i];
        for (int i = 0; i < 9; ++i) calib.m[i] = v[19 + i];
        update_observation_rays(stream, true, pos, par[0], w2c, calib, (int)hdr[4], (int)hdr[5],
            dbvh_tree->accessor(), dcloud->cdata(), dobs_rays->cdata());
    }
    process_observation_rays(stream, dobs_rays->cdata());
    evaluate_observation_rays(stream, dobs_rays->cdata(), drecons->cdata());
    calculate_func_recons(stream, drecons->cdata(), par[1], dwrecons->cdata());
    cacc::sync(stream);

    std::vector<uint32_t> rows(np);
    std::vector<float> recons(np);
    dobs_rays->download(nullptr, rows.data());
    drecons->download(recons.data());
    std::printf("%.9g %.9g %.9g\n", cacc::reduction::sum(dwrecons), cacc::reduction::min(dwrecons), cacc::reduction::max(dwrecons));
    for (uint32_t i = 0; i < np; ++i) std::printf("%u %.9g\n", rows[i], recons[i]);
    return 0;
}

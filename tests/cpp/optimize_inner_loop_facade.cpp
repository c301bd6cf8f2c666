// One iteration of the optimiser's inner loop (apps/optimize_trajectory/optimize_trajectory.cu:258-428) written
// against include/uavmvs_b200/cacc.hpp: populate all views, remove the selected ones (populate = false), K2, K3,
// then score the Nelder-Mead candidate positions of every removed view - here in ONE batched K5 + K6 + arg-min
// call instead of the reference's launch / copy / host-loop per candidate.  Input: flat binary file written by
// tests/test_cpp_facade.py; output (stdout): per candidate "min argmin", then the per-point counts and recons.
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
    // nodes, tris, points, views, width, height, kd nodes, removed views, candidates
    std::vector<uint32_t> hdr = rd<uint32_t>(f, 9);
    uint32_t nn = hdr[0], nt = hdr[1], np = hdr[2], nv = hdr[3], nk = hdr[6], nr = hdr[7], nc = hdr[8];
    std::vector<float> par = rd<float>(f, 2);          // max_distance, target_recon
    std::vector<uint32_t> nodes = rd<uint32_t>(f, 4 * (size_t)nn);
    std::vector<float> aabbs = rd<float>(f, 6 * (size_t)nn), tris = rd<float>(f, 9 * (size_t)nt);
    std::vector<uint32_t> indices = rd<uint32_t>(f, nt);
    std::vector<float> pts = rd<float>(f, 3 * (size_t)np), nrm = rd<float>(f, 3 * (size_t)np);
    std::vector<float> views = rd<float>(f, 28 * (size_t)nv);
    std::vector<uint32_t> kd_nodes = rd<uint32_t>(f, 4 * (size_t)nk);
    std::vector<float> kd_verts = rd<float>(f, 3 * (size_t)nk);
    std::vector<uint32_t> removed = rd<uint32_t>(f, nr);
    std::vector<float> cand = rd<float>(f, 3 * (size_t)nc);
    fclose(f);

    cacc::select_cuda_device(3, 5);
    cacc::BVHTree<cacc::DEVICE>::Ptr dbvh_tree = cacc::BVHTree<cacc::DEVICE>::create(nodes.data(), aabbs.data(),
        tris.data(), indices.data(), nn, nt);
    cacc::KDTree<3u, cacc::DEVICE>::Ptr dkd_tree = cacc::KDTree<3u, cacc::DEVICE>::create(kd_nodes.data(), kd_verts.data(), nk);
    cacc::PointCloud<cacc::HOST>::Ptr cloud = cacc::PointCloud<cacc::HOST>::create(np);
    cacc::PointCloud<cacc::HOST>::Data data = cloud->cdata();
    for (size_t i = 0; i < 3 * (size_t)np; ++i) { data.vertices_ptr[i] = pts[i]; data.normals_ptr[i] = nrm[i]; }
    cacc::PointCloud<cacc::DEVICE>::Ptr dcloud = cacc::PointCloud<cacc::DEVICE>::create<cacc::HOST>(cloud);
    cacc::VectorArray<cacc::Vec3f, cacc::DEVICE>::Ptr dobs_rays = cacc::VectorArray<cacc::Vec3f, cacc::DEVICE>::create(np, 32);
    cacc::Array<float, cacc::DEVICE>::Ptr drecons = cacc::Array<float, cacc::DEVICE>::create(np);
    cacc::Array<float, cacc::DEVICE>::Ptr dcon_hists = cacc::Array<float, cacc::DEVICE>::create(nc * nk);
    cacc::Stream stream = nullptr;
    int width = (int)hdr[4], height = (int)hdr[5];

    struct View { cacc::Vec3f pos; cacc::Mat4f w2c; cacc::Mat3f calib; };
    std::vector<View> cams(nv);
    for (uint32_t k = 0; k < nv; ++k) {
        const float * v = views.data() + 28 * (size_t)k;
        for (int i = 0; i < 3; ++i) cams[k].pos[i] = v[i];
        for (int i = 0; i < 16; ++i) cams[k].w2c.m[i] = v[3 + i];
        for (int i = 0; i < 9; ++i) cams[k].calib.m[i] = v[19 + i];
    }
    for (uint32_t k = 0; k < nv; ++k)                                  /* :258-278 */
        update_observation_rays(stream, true, cams[k].pos, par[0], cams[k].w2c, cams[k].calib, width, height,
            dbvh_tree->accessor(), dcloud->cdata(), dobs_rays->cdata());
    for (uint32_t r = 0; r < nr; ++r) {                                /* :313-333 */
        const View & c = cams[removed[r]];
        update_observation_rays(stream, false, c.pos, par[0], c.w2c, c.calib, width, height,
            dbvh_tree->accessor(), dcloud->cdata(), dobs_rays->cdata());
    }
    process_observation_rays(stream, dobs_rays->cdata());              /* :337-354 */
    evaluate_observation_rays(stream, dobs_rays->cdata(), drecons->cdata());

    dcon_hists->null(stream);                                          /* :389-407, all candidates at once */
    populate_spherical_histogram_batch(stream, cand.data(), nc, par[0], par[1], dbvh_tree->accessor(), dcloud->cdata(),
        dkd_tree->accessor(), dobs_rays->cdata(), drecons->cdata(), dcon_hists->cdata());
    std::vector<float> vmin(nc); std::vector<uint32_t> imin(nc);
    cacc::Array<float, cacc::DEVICE>::Data no_images = {nullptr};
    evaluate_spherical_histogram_batch(stream, cams[0].calib, width, height, dkd_tree->accessor(), dcon_hists->cdata(), nc,
        128, 45, no_images, vmin.data(), imin.data(), nullptr, nullptr);           /* :409-428 */
    cacc::sync(stream);

    for (uint32_t c = 0; c < nc; ++c) std::printf("%.9g %u\n", vmin[c], imin[c]);
    std::vector<uint32_t> rows(np);
    std::vector<float> recons(np);
    dobs_rays->download(nullptr, rows.data());
    drecons->download(recons.data());
    for (uint32_t i = 0; i < np; ++i) std::printf("%u %.9g\n", rows[i], recons[i]);
    return 0;
}

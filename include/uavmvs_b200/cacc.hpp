/*
 * cacc.hpp -- header-compatible facade of the reference's device containers and of the
 * libs/eval kernel entry points over the C ABI of libuavmvs_b200 (include/uavmvs_b200.h).
 *
 * An app of the reference (apps/evaluate_trajectory/evaluate_trajectory.cu and siblings)
 * keeps its source: it includes this header instead of "cacc/..." and "eval/kernels.h",
 * and a kernel launch
 *
 *     update_observation_rays<<<grid, block, 0, stream>>>(true, pos, max_distance, w2c, calib,
 *         width, height, dbvh_tree->accessor(), dcloud->cdata(), dobs_rays->cdata());
 *
 * becomes the plain call
 *
 *     update_observation_rays(stream, true, pos, max_distance, w2c, calib,
 *         width, height, dbvh_tree->accessor(), dcloud->cdata(), dobs_rays->cdata());
 *
 * (same argument list, the launch configuration replaced by the stream).  Ownership follows
 * the reference: static create() returns a std::shared_ptr (libs/cacc/bvh_tree.h:190-194),
 * accessor()/cdata() return non-owning PODs passed by value (bvh_tree.h:44-49,
 * point_cloud.h:26-32, vector_array.h:23-29).  Errors keep the reference's convention:
 * print "file:line (call): message" and exit (libs/cacc/defines.h:20-29).
 *
 * Host-side inputs are plain arrays (the acc tree's nodes/aabbs/tris/indices, exactly what
 * cacc::BVHTree<HOST>::convert produces, libs/cacc/bvh_tree.cu:17-48) or, when "acc/bvh_tree.h" / "acc/kd_tree.h" were
 * included first, the acc trees themselves through the reference's own create<...>(tree) calls.
 * Plain C++11, no CUDA headers needed by the caller.
 */
#ifndef UAVMVS_B200_CACC_HPP
#define UAVMVS_B200_CACC_HPP

#include <cstdio>
#include <cstdlib>
#include <memory>
#include <vector>

#include "../uavmvs_b200.h"

#define UAVMVS_CHECK(call)                                                                  \
    do {                                                                                    \
        if ((call) != 0) {                                                                  \
            std::fprintf(stderr, "%s:%d (%s): %s\n", __FILE__, __LINE__, #call, uavmvs_last_error()); \
            std::exit(EXIT_FAILURE);                                                        \
        }                                                                                   \
    } while (0)

/* ---- construction from the reference's own host trees ------------------------------------------------------------
 * When the application has included the reference's libs/acc headers BEFORE this one ("acc/bvh_tree.h",
 * "acc/kd_tree.h"), the original factory calls keep working unchanged:
 *     cacc::BVHTree<cacc::DEVICE>::create<uint, math::Vec3f>(bvh_tree)      (libs/cacc/bvh_tree.h:190-194)
 *     cacc::KDTree<3, cacc::DEVICE>::create<uint>(kd_tree)                  (libs/cacc/kd_tree.h:150-154)
 * acc declares `template <class C> static C convert(tree)` and leaves the definition to the consumer; the reference
 * specialises it in libs/cacc/bvh_tree.cu:17-48 and libs/cacc/kd_tree.cu:15-36 for <unsigned int, math::Vec3f>.  The two
 * specialisations below do the same job for a plain-array target, so no .cu file of the reference has to be linked. */
#if defined(ACC_BVHTREE_HEADER) || defined(ACC_KDTREE_HEADER)
namespace uavmvs_b200 {
struct HostTree { std::vector<uint32_t> nodes, indices; std::vector<float> aabbs, tris; };
struct HostKd { std::vector<uint32_t> nodes; std::vector<float> verts; };
}
#endif
#ifdef ACC_BVHTREE_HEADER
ACC_NAMESPACE_BEGIN
template <> template <> inline uavmvs_b200::HostTree
BVHTree<unsigned int, math::Vec3f>::convert<uavmvs_b200::HostTree>(BVHTree<unsigned int, math::Vec3f> const & tree) {
    uavmvs_b200::HostTree h;
    const std::size_t n = tree.num_nodes, t = tree.tris.size();
    h.nodes.resize(4 * n); h.aabbs.resize(6 * n); h.tris.resize(9 * t); h.indices.resize(t);
    for (std::size_t i = 0; i < n; ++i) {
        Node const & nd = tree.nodes[i];
        h.nodes[4 * i] = nd.first; h.nodes[4 * i + 1] = nd.last; h.nodes[4 * i + 2] = nd.left; h.nodes[4 * i + 3] = nd.right;
        for (int k = 0; k < 3; ++k) { h.aabbs[6 * i + k] = nd.aabb.min[k]; h.aabbs[6 * i + 3 + k] = nd.aabb.max[k]; }
    }
    for (std::size_t i = 0; i < t; ++i) {
        h.indices[i] = tree.indices[i];
        for (int k = 0; k < 3; ++k) { h.tris[9 * i + k] = tree.tris[i].a[k]; h.tris[9 * i + 3 + k] = tree.tris[i].b[k]; h.tris[9 * i + 6 + k] = tree.tris[i].c[k]; }
    }
    return h;
}
ACC_NAMESPACE_END
#endif
#ifdef ACC_KDTREE_HEADER
ACC_NAMESPACE_BEGIN
template <> template <> inline uavmvs_b200::HostKd
KDTree<3u, unsigned int>::convert<uavmvs_b200::HostKd>(KDTree<3u, unsigned int> const & tree) {
    uavmvs_b200::HostKd h;
    const std::size_t n = tree.num_nodes;
    h.nodes.resize(4 * n); h.verts.resize(3 * n);
    for (std::size_t i = 0; i < n; ++i) {
        h.nodes[4 * i] = tree.nodes[i].vertex_id; h.nodes[4 * i + 1] = tree.nodes[i].d;
        h.nodes[4 * i + 2] = tree.nodes[i].left; h.nodes[4 * i + 3] = tree.nodes[i].right;
        for (int k = 0; k < 3; ++k) h.verts[3 * i + k] = tree.vertices[i][k];
    }
    return h;
}
ACC_NAMESPACE_END
#endif

namespace cacc {

enum Location { HOST, DEVICE };
typedef void * Stream;   /* cudaStream_t */

struct Vec3f { float v[3]; float & operator[](int i) { return v[i]; } float operator[](int i) const { return v[i]; } };
struct Mat3f { float m[9]; };    /* row-major, libs/cacc/matrix.h:36-44 */
struct Mat4f { float m[16]; };

inline void select_cuda_device(int, int) { UAVMVS_CHECK(uavmvs_set_device(0)); }   /* libs/cacc/util.h:82-87 */
inline void set_cuda_device(int device) { UAVMVS_CHECK(uavmvs_set_device(device)); }
inline void sync(Stream stream) { UAVMVS_CHECK(uavmvs_synchronize(stream)); }

template <Location L> class BVHTree;
template <> class BVHTree<DEVICE> {
public:
    typedef std::shared_ptr<BVHTree> Ptr;
    struct Accessor { const uavmvs_bvh * h; };
    /* cacc::BVHTree<DEVICE>::create(acc_tree): arrays as convert<HOST> lays them out */
    static Ptr create(const uint32_t * nodes, const float * aabbs, const float * tris,
        const uint32_t * indices, uint32_t num_nodes, uint32_t num_tris) {
        Ptr p(new BVHTree());
        UAVMVS_CHECK(uavmvs_bvh_create_from_arrays(nodes, aabbs, tris, indices, num_nodes, num_tris, &p->h));
        return p;
    }
    /* B200-native alternative to the host SAH build (acc::BVHTree constructor, libs/acc/bvh_tree.h:335-399):
     * the same sweep-SAH cost model evaluated on the GPU, from an indexed mesh */
    static Ptr build(const float * verts, uint32_t num_verts, const uint32_t * faces, uint32_t num_faces) {
        Ptr p(new BVHTree());
        UAVMVS_CHECK(uavmvs_bvh_build_sah(verts, num_verts, faces, num_faces, &p->h));
        return p;
    }
#ifdef ACC_BVHTREE_HEADER
    /* cacc::BVHTree<DEVICE>::create<IdxType, Vec3fType>(acc_tree), libs/cacc/bvh_tree.h:190-194 + bvh_tree.cu:54-57 */
    template <typename IdxType, typename Vec3fType>
    static Ptr create(typename acc::BVHTree<IdxType, Vec3fType>::ConstPtr tree) {
        uavmvs_b200::HostTree t = acc::BVHTree<IdxType, Vec3fType>::template convert<uavmvs_b200::HostTree>(*tree);
        return create(t.nodes.data(), t.aabbs.data(), t.tris.data(), t.indices.data(), (uint32_t)(t.nodes.size() / 4),
            (uint32_t)t.indices.size());
    }
#endif
    Accessor accessor() const { Accessor a = {h}; return a; }
    ~BVHTree() { uavmvs_bvh_destroy(h); }
private:
    BVHTree() : h(nullptr) {}
    uavmvs_bvh * h;
};

template <unsigned K, Location L> class KDTree;
template <> class KDTree<3u, DEVICE> {
public:
    typedef std::shared_ptr<KDTree> Ptr;
    struct Accessor { const uavmvs_kdtree * h; };
    static Ptr create(const uint32_t * nodes, const float * verts, uint32_t num_nodes) {
        Ptr p(new KDTree());
        UAVMVS_CHECK(uavmvs_kdtree_create(nodes, verts, num_nodes, &p->h));
        return p;
    }
#ifdef ACC_KDTREE_HEADER
    /* cacc::KDTree<3, DEVICE>::create<IdxType>(acc_kd_tree), libs/cacc/kd_tree.h:150-154 + kd_tree.cu:42-45 */
    template <typename IdxType>
    static Ptr create(typename acc::KDTree<3u, IdxType>::ConstPtr tree) {
        uavmvs_b200::HostKd t = acc::KDTree<3u, IdxType>::template convert<uavmvs_b200::HostKd>(*tree);
        return create(t.nodes.data(), t.verts.data(), (uint32_t)(t.nodes.size() / 4));
    }
#endif
    Accessor accessor() const { Accessor a = {h}; return a; }
    ~KDTree() { uavmvs_kdtree_destroy(h); }
private:
    KDTree() : h(nullptr) {}
    uavmvs_kdtree * h;
};

template <Location L> class PointCloud;
template <> class PointCloud<HOST> {
public:
    typedef std::shared_ptr<PointCloud> Ptr;
    struct Data { float * vertices_ptr, * normals_ptr, * values_ptr, * qualities_ptr; uint32_t num_vertices; };
    static Ptr create(uint32_t n) { return Ptr(new PointCloud(n)); }
    Data cdata() { Data d = {v.data(), nrm.data(), val.data(), q.data(), n}; return d; }
private:
    explicit PointCloud(uint32_t n) : n(n), v(3 * (size_t)n), nrm(3 * (size_t)n), val(n, 0.0f), q(n, 1.0f) {}
    uint32_t n; std::vector<float> v, nrm, val, q;   /* xyz packed; defaults of libs/util/cio.h:45-46 */
};
template <> class PointCloud<DEVICE> {
public:
    typedef std::shared_ptr<PointCloud> Ptr;
    struct Data { uavmvs_cloud * h; };
    template <Location O> static Ptr create(typename PointCloud<O>::Ptr host) {   /* point_cloud.h:95-114 */
        typename PointCloud<O>::Data d = host->cdata();
        Ptr p(new PointCloud());
        UAVMVS_CHECK(uavmvs_cloud_create(d.vertices_ptr, d.normals_ptr, d.values_ptr, d.qualities_ptr, d.num_vertices, &p->h));
        return p;
    }
    void download(PointCloud<HOST>::Ptr host) const {   /* values/qualities written by estimate_capture_difficulty */
        PointCloud<HOST>::Data d = host->cdata();
        UAVMVS_CHECK(uavmvs_cloud_download_attributes(h, d.values_ptr, d.qualities_ptr));
    }
    Data cdata() const { Data d = {h}; return d; }
    ~PointCloud() { uavmvs_cloud_destroy(h); }
private:
    PointCloud() : h(nullptr) {}
    uavmvs_cloud * h;
};

template <typename T, Location L> class VectorArray;
template <> class VectorArray<Vec3f, DEVICE> {
public:
    typedef std::shared_ptr<VectorArray> Ptr;
    struct Data { uavmvs_obs * h; };
    static Ptr create(uint32_t num_cols, uint32_t max_rows, Stream = nullptr) {   /* vector_array.h:87-90 */
        Ptr p(new VectorArray());
        UAVMVS_CHECK(uavmvs_obs_create(num_cols, max_rows, &p->h));
        return p;
    }
    void null(Stream stream = nullptr) { UAVMVS_CHECK(uavmvs_obs_clear(stream, h)); }
    void download(float * data, uint32_t * num_rows) const { UAVMVS_CHECK(uavmvs_obs_download(h, data, num_rows)); }
    Data cdata() const { Data d = {h}; return d; }
    ~VectorArray() { uavmvs_obs_destroy(h); }
private:
    VectorArray() : h(nullptr) {}
    uavmvs_obs * h;
};

template <typename T, Location L> class Array;
template <> class Array<float, DEVICE> {
public:
    typedef std::shared_ptr<Array> Ptr;
    struct Data { uavmvs_array * h; };
    static Ptr create(uint32_t n, Stream = nullptr) {   /* array.h:73-76 */
        Ptr p(new Array()); p->n = n;
        UAVMVS_CHECK(uavmvs_array_create(n, &p->h));
        return p;
    }
    void null(Stream stream = nullptr) { UAVMVS_CHECK(uavmvs_array_null(stream, h)); }   /* array.h:123-128 */
    void download(float * host) const { UAVMVS_CHECK(uavmvs_array_download(h, host)); }   /* operator= HOST <- DEVICE */
    void upload(const float * host) { UAVMVS_CHECK(uavmvs_array_upload(h, host)); }
    uint32_t size() const { return n; }
    Data cdata() const { Data d = {h}; return d; }
    ~Array() { uavmvs_array_destroy(h); }
private:
    Array() : h(nullptr), n(0) {}
    uavmvs_array * h; uint32_t n;
};

template <typename T, Location L> class Image;
template <> class Image<float, DEVICE> {
public:
    typedef std::shared_ptr<Image> Ptr;
    struct Data { uavmvs_image * h; };
    static Ptr create(uint32_t width, uint32_t height, Stream = nullptr) {   /* image.h:84-86 */
        Ptr p(new Image());
        UAVMVS_CHECK(uavmvs_image_create(width, height, &p->h));
        return p;
    }
    void download(float * host) const { UAVMVS_CHECK(uavmvs_image_download(h, host)); }
    Data cdata() const { Data d = {h}; return d; }
    ~Image() { uavmvs_image_destroy(h); }
private:
    Image() : h(nullptr) {}
    uavmvs_image * h;
};

namespace reduction {   /* libs/cacc/reduction.h:132-162 */
inline void all(Array<float, DEVICE>::Ptr a, float out3[3], Stream stream = nullptr) {
    UAVMVS_CHECK(uavmvs_reduce(stream, a->cdata().h, out3));
}
inline float sum(Array<float, DEVICE>::Ptr a, Stream s = nullptr) { float o[3]; all(a, o, s); return o[0]; }
inline float min(Array<float, DEVICE>::Ptr a, Stream s = nullptr) { float o[3]; all(a, o, s); return o[1]; }
inline float max(Array<float, DEVICE>::Ptr a, Stream s = nullptr) { float o[3]; all(a, o, s); return o[2]; }
}  // namespace reduction

}  // namespace cacc

/* ---- libs/eval/kernels.h:26-93, same argument lists, launch configuration -> stream ---- */
inline void configure_heuristic(float m_k, float m_x0, float t_k, float t_x0) {
    UAVMVS_CHECK(uavmvs_configure_heuristic(m_k, m_x0, t_k, t_x0));
}
inline void update_observation_rays(cacc::Stream stream, bool populate, cacc::Vec3f view_pos, float max_distance,
    cacc::Mat4f w2c, cacc::Mat3f calib, int width, int height, cacc::BVHTree<cacc::DEVICE>::Accessor bvh_tree,
    cacc::PointCloud<cacc::DEVICE>::Data cloud, cacc::VectorArray<cacc::Vec3f, cacc::DEVICE>::Data obs_rays) {
    UAVMVS_CHECK(uavmvs_update_observation_rays(stream, populate ? 1 : 0, view_pos.v, max_distance, w2c.m, calib.m,
        width, height, bvh_tree.h, cloud.h, obs_rays.h));
}
inline void process_observation_rays(cacc::Stream stream, cacc::VectorArray<cacc::Vec3f, cacc::DEVICE>::Data obs_rays) {
    UAVMVS_CHECK(uavmvs_process_observation_rays(stream, obs_rays.h));
}
inline void evaluate_observation_rays(cacc::Stream stream, cacc::VectorArray<cacc::Vec3f, cacc::DEVICE>::Data obs_rays,
    cacc::Array<float, cacc::DEVICE>::Data recons) {
    UAVMVS_CHECK(uavmvs_evaluate_observation_rays(stream, obs_rays.h, recons.h));
}
/* process + evaluate + calculate_func_recons on the columns update_observation_rays changed since the last call
 * (the optimiser's per-iteration re-evaluation, optimize_trajectory.cu:313-345, :499-529); same arrays every call. */
inline void reevaluate_observation_rays(cacc::Stream stream, cacc::VectorArray<cacc::Vec3f, cacc::DEVICE>::Data obs_rays,
    cacc::Array<float, cacc::DEVICE>::Data recons, float target_recon, cacc::Array<float, cacc::DEVICE>::Data wrecons) {
    UAVMVS_CHECK(uavmvs_reevaluate_observation_rays(stream, obs_rays.h, recons.h, target_recon, wrecons.h));
}
inline void populate_spherical_histogram(cacc::Stream stream, cacc::Vec3f view_pos, float max_distance,
    cacc::BVHTree<cacc::DEVICE>::Accessor bvh_tree, cacc::PointCloud<cacc::DEVICE>::Data cloud,
    cacc::KDTree<3u, cacc::DEVICE>::Accessor kd_tree, cacc::Array<float, cacc::DEVICE>::Data sphere_hist) {
    UAVMVS_CHECK(uavmvs_populate_spherical_histogram(stream, view_pos.v, max_distance, bvh_tree.h, cloud.h, kd_tree.h,
        sphere_hist.h));
}
inline void populate_spherical_histogram(cacc::Stream stream, cacc::Vec3f view_pos, float max_distance,
    float target_recon, cacc::BVHTree<cacc::DEVICE>::Accessor bvh_tree, cacc::PointCloud<cacc::DEVICE>::Data cloud,
    cacc::KDTree<3u, cacc::DEVICE>::Accessor kd_tree, cacc::VectorArray<cacc::Vec3f, cacc::DEVICE>::Data obs_rays,
    cacc::Array<float, cacc::DEVICE>::Data recons, cacc::Array<float, cacc::DEVICE>::Data sphere_hist) {
    UAVMVS_CHECK(uavmvs_populate_spherical_histogram_recon(stream, view_pos.v, max_distance, target_recon, bvh_tree.h,
        cloud.h, kd_tree.h, obs_rays.h, recons.h, sphere_hist.h));
}
inline void evaluate_spherical_histogram(cacc::Stream stream, cacc::Mat3f calib, int width, int height,
    cacc::KDTree<3u, cacc::DEVICE>::Accessor kd_tree, cacc::Array<float, cacc::DEVICE>::Data sphere_hist,
    cacc::Image<float, cacc::DEVICE>::Data hist) {
    UAVMVS_CHECK(uavmvs_evaluate_spherical_histogram(stream, calib.m, width, height, kd_tree.h, sphere_hist.h, hist.h));
}
/* Batched forms (no reference counterpart: they replace the per-candidate launch loops of
 * apps/generate_guidance_volume/generate_guidance_volume.cu:236-253 and
 * apps/optimize_trajectory/optimize_trajectory.cu:388-428).  positions: host [n][3]; sphere_hists: [n][bins]. */
inline void populate_spherical_histogram_batch(cacc::Stream stream, const float * positions, uint32_t n,
    float max_distance, cacc::BVHTree<cacc::DEVICE>::Accessor bvh_tree, cacc::PointCloud<cacc::DEVICE>::Data cloud,
    cacc::KDTree<3u, cacc::DEVICE>::Accessor kd_tree, cacc::Array<float, cacc::DEVICE>::Data sphere_hists) {
    UAVMVS_CHECK(uavmvs_populate_spherical_histogram_batch(stream, positions, n, max_distance, 0.0f, bvh_tree.h, cloud.h,
        kd_tree.h, nullptr, nullptr, sphere_hists.h));
}
inline void populate_spherical_histogram_batch(cacc::Stream stream, const float * positions, uint32_t n,
    float max_distance, float target_recon, cacc::BVHTree<cacc::DEVICE>::Accessor bvh_tree,
    cacc::PointCloud<cacc::DEVICE>::Data cloud, cacc::KDTree<3u, cacc::DEVICE>::Accessor kd_tree,
    cacc::VectorArray<cacc::Vec3f, cacc::DEVICE>::Data obs_rays, cacc::Array<float, cacc::DEVICE>::Data recons,
    cacc::Array<float, cacc::DEVICE>::Data sphere_hists) {
    UAVMVS_CHECK(uavmvs_populate_spherical_histogram_batch(stream, positions, n, max_distance, target_recon, bvh_tree.h,
        cloud.h, kd_tree.h, obs_rays.h, recons.h, sphere_hists.h));
}
/* images may be a null Data ({nullptr}); h_* are host arrays of n entries or nullptr */
inline void evaluate_spherical_histogram_batch(cacc::Stream stream, cacc::Mat3f calib, int width, int height,
    cacc::KDTree<3u, cacc::DEVICE>::Accessor kd_tree, cacc::Array<float, cacc::DEVICE>::Data sphere_hists, uint32_t n,
    uint32_t image_width, uint32_t image_height, cacc::Array<float, cacc::DEVICE>::Data images,
    float * h_min, uint32_t * h_argmin, float * h_max, uint32_t * h_argmax) {
    UAVMVS_CHECK(uavmvs_evaluate_spherical_histogram_batch(stream, calib.m, width, height,
        const_cast<uavmvs_kdtree *>(kd_tree.h), sphere_hists.h, n, image_width, image_height, images.h,
        h_min, h_argmin, h_max, h_argmax));
}
inline void estimate_capture_difficulty(cacc::Stream stream, float max_distance,
    cacc::BVHTree<cacc::DEVICE>::Accessor bvh_tree, uint32_t mesh_size, cacc::KDTree<3u, cacc::DEVICE>::Accessor kd_tree,
    cacc::PointCloud<cacc::DEVICE>::Data cloud) {
    UAVMVS_CHECK(uavmvs_estimate_capture_difficulty(stream, max_distance, bvh_tree.h, mesh_size, kd_tree.h, cloud.h));
}
inline void calculate_func_recons(cacc::Stream stream, cacc::Array<float, cacc::DEVICE>::Data recons, float target_recon,
    cacc::Array<float, cacc::DEVICE>::Data wrecons) {
    UAVMVS_CHECK(uavmvs_calculate_func_recons(stream, recons.h, target_recon, wrecons.h));
}

#endif /* UAVMVS_B200_CACC_HPP */

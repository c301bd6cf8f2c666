/*
 * oracle/_ref/libref_cuda_{exact,fast}.so — the REFERENCE's own device code for the hot path,
 * compiled for sm_100a from where it lies under /root/reference:
 *   libs/eval/kernels.cu    K1 :138-202, K2 :204-264, K3 :266-291, K4 :293-336, K5 :338-390,
 *                           K6 :392-446, K7 :448-494, K8 :496-510
 *   libs/cacc/bvh_tree.cu, kd_tree.cu, tracing.h, nnsearch.h, primitives.h and the containers
 *   libs/acc/bvh_tree.h, kd_tree.h (host builders the cacc trees are converted from)
 * No reference source is copied: this file only #includes those translation units and re-exports
 * them through a C ABI, so that on the GPU box (a) tests can pin oracle/uavmvs_oracle.cpp and the
 * product kernels against the reference's real kernels and (b) bench/tools can time the reference's
 * per-view launch loop on the same B200 ("second side reference").
 * TEST INFRASTRUCTURE ONLY (never linked by the product).
 *
 * Two builds (oracle/Makefile):
 *   exact: --fmad=false, IEEE div/sqrt, no FTZ  -> the canonical operation order of SURVEY appendix B
 *   fast : the reference's own flags (premake5.lua:19-22: --ftz=true --prec-div=false --prec-sqrt=false)
 *
 * Shims: math/vector.h stand-in (oracle/ref_shim/math), <memory>/<tuple>, and
 * the pre-Volta warp intrinsics mapped onto their *_sync forms with a full mask (every call site in
 * kernels.cu:240-258 and reduction.h:99 runs with all 32 lanes of the warp converged).
 */
#include <memory>
#include <tuple>
#include <cstdint>
#include <cstring>
#include <vector>
#include <cuda_runtime.h>

#define __shfl_xor(v, s) __shfl_xor_sync(0xffffffffu, (v), (s))
#define __shfl_down(v, s) __shfl_down_sync(0xffffffffu, (v), (s))
#define __shfl(v, l) __shfl_sync(0xffffffffu, (v), (l))
#define __ballot(p) __ballot_sync(0xffffffffu, (p))

#include <math/vector.h>
#include "cacc/math.h"          /* include order of the apps (optimize_trajectory.cu:25-28) */
#include "cacc/bvh_tree.cu"
#include "cacc/kd_tree.cu"
#include "cacc/point_cloud.h"
#include "cacc/vector_array.h"
#include "cacc/array.h"
#include "cacc/image.h"
#include "cacc/reduction.h"
#include "eval/kernels.cu"

typedef acc::BVHTree<unsigned, math::Vec3f> AccBVH;
typedef acc::KDTree<3, unsigned> AccKD;

struct RefCuBVH { AccBVH::Ptr host; cacc::BVHTree<cacc::DEVICE>::Ptr dev; };
struct RefCuKD { AccKD::Ptr host; cacc::KDTree<3, cacc::DEVICE>::Ptr dev; };
struct RefCuCloud { cacc::PointCloud<cacc::DEVICE>::Ptr dev; uint32_t n; };
struct RefCuObs { cacc::VectorArray<cacc::Vec3f, cacc::DEVICE>::Ptr dev; uint32_t n, rows; };

static inline cacc::Vec3f v3(float const * p) { return cacc::Vec3f(p[0], p[1], p[2]); }

extern "C" {

int refcu_device_count(void) { int n = 0; return cudaGetDeviceCount(&n) == cudaSuccess ? n : 0; }

void * refcu_bvh_create(float const * verts, uint32_t num_verts, uint32_t const * faces, uint32_t num_faces, int threads)
{
    std::vector<math::Vec3f> v(num_verts);
    for (uint32_t i = 0; i < num_verts; ++i) v[i] = math::Vec3f(verts + 3 * i);
    std::vector<unsigned> f(faces, faces + 3 * (std::size_t)num_faces);
    RefCuBVH * h = new RefCuBVH;
    h->host = AccBVH::create(f, v, threads);
    h->dev = cacc::BVHTree<cacc::DEVICE>::create<unsigned, math::Vec3f>(h->host);
    return h;
}
void refcu_bvh_destroy(void * h) { delete (RefCuBVH *)h; }

void * refcu_kd_create(float const * verts, uint32_t num_verts)
{
    std::vector<math::Vec3f> v(num_verts);
    for (uint32_t i = 0; i < num_verts; ++i) v[i] = math::Vec3f(verts + 3 * i);
    RefCuKD * h = new RefCuKD;
    h->host = AccKD::create(v);
    h->dev = cacc::KDTree<3, cacc::DEVICE>::create<unsigned>(h->host);
    return h;
}
void refcu_kd_destroy(void * h) { delete (RefCuKD *)h; }

void * refcu_cloud_create(uint32_t n, float const * verts, float const * normals, float const * values, float const * qualities)
{
    cacc::PointCloud<cacc::HOST>::Ptr host = cacc::PointCloud<cacc::HOST>::create(n);
    cacc::PointCloud<cacc::HOST>::Data d = host->cdata();
    for (uint32_t i = 0; i < n; ++i) {
        d.vertices_ptr[i] = v3(verts + 3 * i);
        d.normals_ptr[i] = v3(normals + 3 * i);
        d.values_ptr[i] = values ? values[i] : 0.0f;
        d.qualities_ptr[i] = qualities ? qualities[i] : 1.0f;
    }
    RefCuCloud * h = new RefCuCloud;
    h->dev = cacc::PointCloud<cacc::DEVICE>::create<cacc::HOST>(host);
    h->n = n;
    return h;
}
void refcu_cloud_download(void * hp, float * values, float * qualities)
{
    RefCuCloud * h = (RefCuCloud *)hp;
    cacc::PointCloud<cacc::DEVICE>::Data d = h->dev->cdata();
    if (values) CHECK(cudaMemcpy(values, d.values_ptr, h->n * sizeof(float), cudaMemcpyDeviceToHost));
    if (qualities) CHECK(cudaMemcpy(qualities, d.qualities_ptr, h->n * sizeof(float), cudaMemcpyDeviceToHost));
}
void refcu_cloud_destroy(void * h) { delete (RefCuCloud *)h; }

void * refcu_obs_create(uint32_t n, uint32_t max_rows)
{
    RefCuObs * h = new RefCuObs;
    h->dev = cacc::VectorArray<cacc::Vec3f, cacc::DEVICE>::create(n, max_rows);
    h->dev->null(); h->dev->clear(); h->dev->sync();
    h->n = n; h->rows = max_rows;
    return h;
}
void refcu_obs_clear(void * hp) { RefCuObs * h = (RefCuObs *)hp; h->dev->null(); h->dev->clear(); h->dev->sync(); }
/* rows: [max_rows][n][4] floats (row-major like the device array without its pitch) */
void refcu_obs_download(void * hp, uint32_t * num_rows, float * rows)
{
    RefCuObs * h = (RefCuObs *)hp;
    cacc::VectorArray<cacc::Vec3f, cacc::DEVICE>::Data d = h->dev->cdata();
    if (num_rows) CHECK(cudaMemcpy(num_rows, d.num_rows_ptr, h->n * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    if (rows) CHECK(cudaMemcpy2D(rows, h->n * sizeof(cacc::Vec3f), d.data_ptr, d.pitch, h->n * sizeof(cacc::Vec3f),
        h->rows, cudaMemcpyDeviceToHost));
}
void refcu_obs_destroy(void * h) { delete (RefCuObs *)h; }

void refcu_configure_heuristic(float m_k, float m_x0, float t_k, float t_x0) { configure_heuristic(m_k, m_x0, t_k, t_x0); }

/* views: 28 floats each (pos 3, w2c 16 row-major, calib 9 row-major); one launch per view exactly like
 * apps/evaluate_trajectory/evaluate_trajectory.cu:150-166 */
void refcu_update_observation_rays(int populate, float const * views, uint32_t num_views, float max_distance,
    int width, int height, void * bvh, void * cloud, void * obs)
{
    RefCuCloud * c = (RefCuCloud *)cloud;
    dim3 grid(cacc::divup(c->n, KERNEL_BLOCK_SIZE));
    dim3 block(KERNEL_BLOCK_SIZE);
    for (uint32_t k = 0; k < num_views; ++k) {
        float const * v = views + 28 * (std::size_t)k;
        update_observation_rays<<<grid, block>>>(populate != 0, v3(v), max_distance, cacc::Mat4f(v + 3),
            cacc::Mat3f(v + 19), width, height, ((RefCuBVH *)bvh)->dev->accessor(), c->dev->cdata(),
            ((RefCuObs *)obs)->dev->cdata());
    }
    CHECK(cudaDeviceSynchronize());
}

void refcu_process_observation_rays(void * obs)
{
    RefCuObs * o = (RefCuObs *)obs;
    dim3 grid(cacc::divup(o->n, 2));
    dim3 block(32, 2);
    process_observation_rays<<<grid, block>>>(o->dev->cdata());
    CHECK(cudaDeviceSynchronize());
}

static void launch_evaluate(RefCuObs * o, cacc::Array<float, cacc::DEVICE>::Ptr recons)
{
    dim3 grid(cacc::divup(o->n, KERNEL_BLOCK_SIZE));
    dim3 block(KERNEL_BLOCK_SIZE);
    evaluate_observation_rays<<<grid, block>>>(o->dev->cdata(), recons->cdata());
}

void refcu_evaluate_observation_rays(void * obs, float * recons_out)
{
    RefCuObs * o = (RefCuObs *)obs;
    cacc::Array<float, cacc::DEVICE>::Ptr recons = cacc::Array<float, cacc::DEVICE>::create(o->n);
    launch_evaluate(o, recons);
    CHECK(cudaDeviceSynchronize());
    CHECK(cudaMemcpy(recons_out, recons->cdata().data_ptr, o->n * sizeof(float), cudaMemcpyDeviceToHost));
}

/* wrecons = func(recons) (K8) and sum/min/max through cacc::reduction (K9); out3 = {sum, min, max} */
void refcu_calculate_func_recons(float const * recons_in, uint32_t n, float target_recon, float * wrecons_out, float * out3)
{
    cacc::Array<float, cacc::DEVICE>::Ptr recons = cacc::Array<float, cacc::DEVICE>::create(n);
    cacc::Array<float, cacc::DEVICE>::Ptr wrecons = cacc::Array<float, cacc::DEVICE>::create(n);
    CHECK(cudaMemcpy(recons->cdata().data_ptr, recons_in, n * sizeof(float), cudaMemcpyHostToDevice));
    dim3 grid(cacc::divup(n, KERNEL_BLOCK_SIZE));
    dim3 block(KERNEL_BLOCK_SIZE);
    calculate_func_recons<<<grid, block>>>(recons->cdata(), target_recon, wrecons->cdata());
    CHECK(cudaDeviceSynchronize());
    if (wrecons_out) CHECK(cudaMemcpy(wrecons_out, wrecons->cdata().data_ptr, n * sizeof(float), cudaMemcpyDeviceToHost));
    if (out3) {
        out3[0] = cacc::reduction::sum(wrecons);
        out3[1] = cacc::reduction::min(wrecons);
        out3[2] = cacc::reduction::max(wrecons);
    }
}

/* K4 (recons_in == NULL) or K5 for ONE position; hist_out[num sphere verts] */
void refcu_populate_spherical_histogram(float const * view_pos, float max_distance, float target_recon,
    void * bvh, void * cloud, void * kd, void * obs, float const * recons_in, uint32_t num_bins, float * hist_out)
{
    RefCuCloud * c = (RefCuCloud *)cloud;
    cacc::Array<float, cacc::DEVICE>::Ptr hist = cacc::Array<float, cacc::DEVICE>::create(num_bins);
    hist->null();
    dim3 grid(cacc::divup(c->n, KERNEL_BLOCK_SIZE));
    dim3 block(KERNEL_BLOCK_SIZE);
    if (recons_in) {
        cacc::Array<float, cacc::DEVICE>::Ptr recons = cacc::Array<float, cacc::DEVICE>::create(c->n);
        CHECK(cudaMemcpy(recons->cdata().data_ptr, recons_in, c->n * sizeof(float), cudaMemcpyHostToDevice));
        populate_spherical_histogram<<<grid, block>>>(v3(view_pos), max_distance, target_recon,
            ((RefCuBVH *)bvh)->dev->accessor(), c->dev->cdata(), ((RefCuKD *)kd)->dev->accessor(),
            ((RefCuObs *)obs)->dev->cdata(), recons->cdata(), hist->cdata());
        CHECK(cudaDeviceSynchronize());
    } else {
        populate_spherical_histogram<<<grid, block>>>(v3(view_pos), max_distance,
            ((RefCuBVH *)bvh)->dev->accessor(), c->dev->cdata(), ((RefCuKD *)kd)->dev->accessor(), hist->cdata());
        CHECK(cudaDeviceSynchronize());
    }
    CHECK(cudaMemcpy(hist_out, hist->cdata().data_ptr, num_bins * sizeof(float), cudaMemcpyDeviceToHost));
}

/* K6 launched as apps/optimize_trajectory/optimize_trajectory.cu:409-417; image_out[45][128] */
void refcu_evaluate_spherical_histogram(float const * calib, int width, int height, void * kd,
    float const * hist_in, uint32_t num_bins, float * image_out)
{
    cacc::Array<float, cacc::DEVICE>::Ptr hist = cacc::Array<float, cacc::DEVICE>::create(num_bins);
    CHECK(cudaMemcpy(hist->cdata().data_ptr, hist_in, num_bins * sizeof(float), cudaMemcpyHostToDevice));
    cacc::Image<float, cacc::DEVICE>::Ptr img = cacc::Image<float, cacc::DEVICE>::create(128, 45);
    dim3 grid(cacc::divup(128, KERNEL_BLOCK_SIZE), 45);
    dim3 block(KERNEL_BLOCK_SIZE);
    evaluate_spherical_histogram<<<grid, block>>>(cacc::Mat3f(calib), width, height,
        ((RefCuKD *)kd)->dev->accessor(), hist->cdata(), img->cdata());
    CHECK(cudaDeviceSynchronize());
    cacc::Image<float, cacc::DEVICE>::Data d = img->cdata();
    CHECK(cudaMemcpy2D(image_out, 128 * sizeof(float), d.data_ptr, d.pitch, 128 * sizeof(float), 45, cudaMemcpyDeviceToHost));
}

/* K7 as apps/estimate_capture_difficulty/estimate_capture_difficulty.cu:136-146 */
void refcu_estimate_capture_difficulty(float max_distance, void * bvh, uint32_t mesh_size, void * kd, void * cloud)
{
    RefCuCloud * c = (RefCuCloud *)cloud;
    dim3 grid(cacc::divup(c->n, KERNEL_BLOCK_SIZE));
    dim3 block(KERNEL_BLOCK_SIZE);
    estimate_capture_difficulty<<<grid, block>>>(max_distance, ((RefCuBVH *)bvh)->dev->accessor(), mesh_size,
        ((RefCuKD *)kd)->dev->accessor(), c->dev->cdata());
    CHECK(cudaDeviceSynchronize());
}

/* The reference's evaluation loop timed on the device: clear, V launches of K1, K2, K3 (+K8).
 * Returns milliseconds (CUDA events on the legacy stream). */
float refcu_time_evaluate(float const * views, uint32_t num_views, float max_distance, int width, int height,
    float target_recon, void * bvh, void * cloud, void * obs)
{
    RefCuCloud * c = (RefCuCloud *)cloud;
    RefCuObs * o = (RefCuObs *)obs;
    cacc::Array<float, cacc::DEVICE>::Ptr recons = cacc::Array<float, cacc::DEVICE>::create(o->n);
    cacc::Array<float, cacc::DEVICE>::Ptr wrecons = cacc::Array<float, cacc::DEVICE>::create(o->n);
    o->dev->clear();
    CHECK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1;
    CHECK(cudaEventCreate(&e0)); CHECK(cudaEventCreate(&e1));
    CHECK(cudaEventRecord(e0));
    dim3 grid(cacc::divup(c->n, KERNEL_BLOCK_SIZE));
    dim3 block(KERNEL_BLOCK_SIZE);
    for (uint32_t k = 0; k < num_views; ++k) {
        float const * v = views + 28 * (std::size_t)k;
        update_observation_rays<<<grid, block>>>(true, v3(v), max_distance, cacc::Mat4f(v + 3), cacc::Mat3f(v + 19),
            width, height, ((RefCuBVH *)bvh)->dev->accessor(), c->dev->cdata(), o->dev->cdata());
    }
    process_observation_rays<<<dim3(cacc::divup(o->n, 2)), dim3(32, 2)>>>(o->dev->cdata());
    launch_evaluate(o, recons);
    calculate_func_recons<<<grid, block>>>(recons->cdata(), target_recon, wrecons->cdata());
    CHECK(cudaEventRecord(e1));
    CHECK(cudaEventSynchronize(e1));
    float ms = 0.0f;
    CHECK(cudaEventElapsedTime(&ms, e0, e1));
    CHECK(cudaEventDestroy(e0)); CHECK(cudaEventDestroy(e1));
    return ms;
}

}  /* extern "C" */

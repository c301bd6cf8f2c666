/*
 * uavmvs_b200.h -- C ABI of the B200-native visibility-and-reconstructability
 * evaluation (drop-in for the libs/eval + libs/cacc path of nmoehrle/uavmvs).
 *
 * Every entry point cites the reference interface it replaces (paths relative
 * to the reference checkout).  Conventions:
 *   - plain pointers and sizes only; HOST pointers unless the name says device;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream);
 *   - every function returns 0 on success, non-zero on failure;
 *     uavmvs_last_error() gives the message (thread local).  The reference's
 *     convention is print-and-exit (CHECK(), libs/cacc/defines.h:20-29); the
 *     C++ facade (include/uavmvs_b200/cacc.hpp) maps non-zero back to that.
 *   - matrices are row-major float (libs/cacc/matrix.h:36-44);
 *   - there is no CPU fallback: every call needs a CUDA device.
 */
#ifndef UAVMVS_B200_H
#define UAVMVS_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define UAVMVS_NAI 0xffffffffu
#define UAVMVS_VIEW_FLOATS 28 /* pos[3], w2c[16], calib[9] */

typedef struct uavmvs_bvh uavmvs_bvh;         /* cacc::BVHTree<DEVICE>            */
typedef struct uavmvs_kdtree uavmvs_kdtree;   /* cacc::KDTree<3, DEVICE>          */
typedef struct uavmvs_cloud uavmvs_cloud;     /* cacc::PointCloud<DEVICE>         */
typedef struct uavmvs_obs uavmvs_obs;         /* cacc::VectorArray<Vec3f, DEVICE> */
typedef struct uavmvs_array uavmvs_array;     /* cacc::Array<float, DEVICE>       */
typedef struct uavmvs_image uavmvs_image;     /* cacc::Image<float, DEVICE>       */

const char * uavmvs_last_error(void);
/* cacc::select_cuda_device / set_cuda_device, libs/cacc/util.h:36-100 */
int uavmvs_set_device(int device);
int uavmvs_device_count(int * count);
int uavmvs_synchronize(void * stream);

/* ---- BVH: cacc::BVHTree<DEVICE>::create(acc_tree), libs/cacc/bvh_tree.h:190-194,
 * libs/cacc/bvh_tree.cu:17-57.  Arrays are exactly what convert<HOST> hands over:
 * nodes[4n] = {first,last,left,right}, aabbs[6n] = {min,max}, tris[9t] (leaf order),
 * indices[t] = original face ids. */
int uavmvs_bvh_create_from_arrays(const uint32_t * nodes, const float * aabbs, const float * tris,
    const uint32_t * indices, uint32_t num_nodes, uint32_t num_tris, uavmvs_bvh ** out);
/* GPU LBVH build (Morton sort + bottom-up refit) from an indexed mesh; replaces the host
 * acc::BVHTree constructor (libs/acc/bvh_tree.h:335-399) for large scenes. */
int uavmvs_bvh_build_lbvh(const float * verts, uint32_t num_verts, const uint32_t * faces,
    uint32_t num_faces, uavmvs_bvh ** out);
/* Same inputs; the topology comes from parallel locally-ordered clustering over the Morton order
 * (nearest neighbour by union surface area, mutual pairs merge): close to SAH quality, built on the GPU. */
int uavmvs_bvh_build_ploc(const float * verts, uint32_t num_verts, const uint32_t * faces,
    uint32_t num_faces, uavmvs_bvh ** out);
/* Same inputs; full-sweep surface-area-heuristic build on the GPU: the reference builder's own cost model and
 * termination rule (libs/acc/bvh_tree.h:277-333, cost = 1.2 + (A_l/A n_l + A_r/A n_r), leaf when not below n),
 * evaluated for every split position of every axis of every node of a level at once (three pre-sorted orders,
 * segmented box scans, stable segmented partition).  Best traversal quality of the three GPU builders. */
int uavmvs_bvh_build_sah(const float * verts, uint32_t num_verts, const uint32_t * faces,
    uint32_t num_faces, uavmvs_bvh ** out);
int uavmvs_bvh_sizes(const uavmvs_bvh * bvh, uint32_t * num_nodes, uint32_t * num_tris);
/* Export in the reference layout (for the oracle / for acc-side consumers). */
int uavmvs_bvh_export(const uavmvs_bvh * bvh, uint32_t * nodes, float * aabbs, float * tris, uint32_t * indices);
int uavmvs_bvh_destroy(uavmvs_bvh * bvh);
/* cacc::tracing::trace, libs/cacc/tracing.h:37-116 -- batch form for tests and tools.
 * rays[8n] = origin, dir, tmin, tmax.  any_hit != 0: hit flag only (face = first accepted).
 * Otherwise closest hit; face[i] = original face id or UAVMVS_NAI. */
int uavmvs_bvh_trace(void * stream, const uavmvs_bvh * bvh, const float * rays, uint64_t num_rays,
    int any_hit, uint8_t * hit, uint32_t * face, float * t);

/* Closest point on the mesh for num_points HOST queries [n][3]: acc::BVHTree::closest_point(vertex, &cp,
 * max_dist), libs/acc/bvh_tree.h:492-548 (the optimiser's airspace test, apps/optimize_trajectory/
 * optimize_trajectory.cu:370-386, :456).  found[i] = 1 iff a triangle lies closer than max_dist (pass INFINITY
 * for none); tri_index[i] = its index in LEAF ORDER (acc's cp.first, the argument of get_triangle) or
 * 0xffffffff; closest[i] = the point.  Bit-identical to the reference on the same tree.  Synchronises. */
int uavmvs_bvh_closest_point(void * stream, const uavmvs_bvh * bvh, const float * points, uint64_t num_points,
    float max_dist, uint8_t * found, uint32_t * tri_index, float * closest);

/* ---- KD-tree: cacc::KDTree<3,DEVICE>::create(acc_kd), libs/cacc/kd_tree.h:150-154,
 * libs/cacc/kd_tree.cu:15-45.  nodes[4n] = {vid,dim,left,right}, verts[3n]. */
int uavmvs_kdtree_create(const uint32_t * nodes, const float * verts, uint32_t num_nodes, uavmvs_kdtree ** out);
int uavmvs_kdtree_destroy(uavmvs_kdtree * kd);
/* Host-side median-split build over n points (the rule of acc::KDTree<3>, libs/acc/kd_tree.h:92-158, single-threaded
 * numbering) followed by uavmvs_kdtree_create: for callers that do not link libs/acc. */
int uavmvs_kdtree_build(const float * verts, uint32_t n, uavmvs_kdtree ** out);
/* The vertex list of generate_sphere_mesh(radius, subdivisions) (libs/geom/sphere.h:17-115) in its order, coincident
 * twins included: 1272 vertices / 642 distinct directions at 3 subdivisions -- the bins of sphere_hist
 * (apps/optimize_trajectory/optimize_trajectory.cu:157-164).  verts == NULL: only *count is written. */
int uavmvs_sphere_vertices(float radius, uint32_t subdivisions, float * verts, uint32_t capacity, uint32_t * count);
/* cacc::nnsearch::find_nn<3> (libs/cacc/nnsearch.h:67-155) for n host queries [n][3] -> vertex indices (host).
 * certified_lookup != 0 answers through the cube-map / neighbour-list shortcut the batched histogram path uses
 * (kd.cuh); it must and does return the same indices, duplicates included - the parity tests compare the two. */
int uavmvs_kdtree_find_nn(void * stream, uavmvs_kdtree * kd, const float * queries, uint64_t n, int certified_lookup,
    uint32_t * out_idx);

/* Proxy-cloud sampler on the device: apps/generate_proxy_cloud/generate_proxy_cloud.cpp:84-184 (area * density samples
 * per face from std::mt19937 seeded with 3 * face, face normals, sorted by 20-bit Z-order index with a lexicographic
 * tie-break).  Same samples, same order as the reference tool; feed them to uavmvs_cloud_create. */
typedef struct uavmvs_samples uavmvs_samples;
int uavmvs_proxy_cloud_sample(const float * verts, uint32_t num_verts, const uint32_t * faces, uint32_t num_faces,
    float samples_per_unit_area, uavmvs_samples ** out, uint32_t * num_samples);
int uavmvs_proxy_cloud_download(const uavmvs_samples * samples, float * verts, float * normals);
int uavmvs_proxy_cloud_destroy(uavmvs_samples * samples);

/* ---- Point cloud: cacc::PointCloud<HOST>::create(n) + fill + <DEVICE>::create<HOST>,
 * libs/cacc/point_cloud.h:26-32, :91-114.  values/qualities may be NULL
 * (defaults 0 / 1, libs/util/cio.h:45-46). */
int uavmvs_cloud_create(const float * vertices, const float * normals, const float * values,
    const float * qualities, uint32_t n, uavmvs_cloud ** out);
int uavmvs_cloud_download_attributes(const uavmvs_cloud * cloud, float * values, float * qualities);
int uavmvs_cloud_upload_attributes(uavmvs_cloud * cloud, const float * values, const float * qualities);
int uavmvs_cloud_size(const uavmvs_cloud * cloud, uint32_t * n);
int uavmvs_cloud_destroy(uavmvs_cloud * cloud);

/* ---- Observation rays: cacc::VectorArray<Vec3f,DEVICE>::create(num_cols, max_rows),
 * libs/cacc/vector_array.h:23-50, :87-90.  Device layout kept: row-major float4 rows of
 * `pitch` bytes, element (row r, col i) at data + r*pitch + 16*i; num_rows[num_cols]. */
int uavmvs_obs_create(uint32_t num_cols, uint32_t max_rows, uavmvs_obs ** out);
int uavmvs_obs_clear(void * stream, uavmvs_obs * obs);              /* VectorArray::null() */
/* data: [max_rows][num_cols][4] float (dense), num_rows: [num_cols]; either may be NULL */
int uavmvs_obs_download(const uavmvs_obs * obs, float * data, uint32_t * num_rows);
int uavmvs_obs_upload(uavmvs_obs * obs, const float * data, const uint32_t * num_rows);
int uavmvs_obs_destroy(uavmvs_obs * obs);

/* ---- cacc::Array<float,DEVICE>, libs/cacc/array.h:73-128 */
int uavmvs_array_create(uint32_t n, uavmvs_array ** out);
int uavmvs_array_null(void * stream, uavmvs_array * a);
int uavmvs_array_download(const uavmvs_array * a, float * host);
int uavmvs_array_upload(uavmvs_array * a, const float * host);
int uavmvs_array_destroy(uavmvs_array * a);

/* ---- cacc::Image<float,DEVICE>, libs/cacc/image.h:84-86 */
int uavmvs_image_create(uint32_t width, uint32_t height, uavmvs_image ** out);
int uavmvs_image_download(const uavmvs_image * img, float * host); /* dense [h][w] */
int uavmvs_image_destroy(uavmvs_image * img);

/* ---- configure_heuristic, libs/eval/kernels.h:93, kernels.cu:93-103.
 * Process-global like the reference's __constant__ symbols; defaults 8, 4, 32, 16. */
int uavmvs_configure_heuristic(float m_k, float m_x0, float t_k, float t_x0);
/* The same constants bound to ONE observation array: every kernel that evaluates the heuristic on it (K3, K5, the fused
 * evaluations) then uses these instead of the process-wide ones, so host threads working on different arrays -- or one
 * thread alternating between them -- do not share state (SURVEY.md 8b "per-call or per-context, not global"). */
int uavmvs_obs_configure_heuristic(uavmvs_obs * obs, float m_k, float m_x0, float t_k, float t_x0);

/* ---- K1 update_observation_rays, libs/eval/kernels.h:26-35, kernels.cu:138-202.
 * One view, all points; same arguments as the kernel plus the stream. */
int uavmvs_update_observation_rays(void * stream, int populate, const float view_pos[3],
    float max_distance, const float w2c[16], const float calib[9], int width, int height,
    const uavmvs_bvh * bvh, const uavmvs_cloud * cloud, uavmvs_obs * obs);
/* Batched form of the launch loop apps/evaluate_trajectory/evaluate_trajectory.cu:154-164:
 * `num_views` views in stream order, views[28*k] = {pos, w2c, calib} (HOST memory).
 * Result identical to num_views single calls in order. */
int uavmvs_update_observation_rays_batch(void * stream, int populate, const float * views,
    uint32_t num_views, float max_distance, int width, int height,
    const uavmvs_bvh * bvh, const uavmvs_cloud * cloud, uavmvs_obs * obs);
/* Same with the view array already on the device (device pointer). */
int uavmvs_update_observation_rays_batch_device(void * stream, int populate, const float * d_views,
    uint32_t num_views, float max_distance, int width, int height,
    const uavmvs_bvh * bvh, const uavmvs_cloud * cloud, uavmvs_obs * obs);

/* ---- K2 process_observation_rays, kernels.h:37-39, kernels.cu:204-264 */
int uavmvs_process_observation_rays(void * stream, uavmvs_obs * obs);
/* ---- K3 evaluate_observation_rays, kernels.h:41-44, kernels.cu:266-291 */
int uavmvs_evaluate_observation_rays(void * stream, const uavmvs_obs * obs, uavmvs_array * recons);
/* ---- K2 + K3 + K8 for the columns that changed: the optimiser's re-evaluation after it removed or re-added a few
 * views (apps/optimize_trajectory/optimize_trajectory.cu:313-345 and :499-529 run process_observation_rays,
 * evaluate_observation_rays and calculate_func_recons over ALL points every iteration).  Every observation array keeps
 * one flag per 32 columns, set by update_observation_rays (either direction) when a column of the block received or
 * lost a row, and by clear / upload / mark_dirty for all blocks; this call runs the three kernels on the flagged blocks
 * only and clears the flags.  Result identical to the three full launches -- except that rows holding EQUAL sort keys may
 * come out in another order within their column: the reference's network is not idempotent on ties (kernels.cu:241), and
 * the full launches re-sort unchanged columns too -- PROVIDED `recons` (and `wrecons`, may be NULL)
 * are the arrays, and target_recon the value, of the previous call for this observation array (the first call after
 * create / clear / upload evaluates everything). */
int uavmvs_reevaluate_observation_rays(void * stream, uavmvs_obs * obs, uavmvs_array * recons, float target_recon,
    uavmvs_array * wrecons);
/* Flags every column as changed (after writing through uavmvs_obs_device_ptr pointers). */
int uavmvs_obs_mark_dirty(void * stream, uavmvs_obs * obs);
/* ---- K8 calculate_func_recons, kernels.h:86-91, kernels.cu:496-510 */
int uavmvs_calculate_func_recons(void * stream, const uavmvs_array * recons, float target_recon,
    uavmvs_array * wrecons);
/* ---- K9 cacc::reduction::{sum,min,max}, libs/cacc/reduction.h:132-162.
 * out3 = {sum, min, max} (HOST); synchronises the stream. */
int uavmvs_reduce(void * stream, const uavmvs_array * a, float out3[3]);
/* ---- K9 without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same {sum, min, max} as uavmvs_reduce, written to out3->data[0..2] -- a DEVICE array of >= 3 floats the caller created
 * beforehand (uavmvs_array_create) -- asynchronously on `stream` (no synchronisation, no host copy).  The call performs
 * no cudaMalloc / cudaFree / cudaMallocHost and calls no allocating routine: its scratch (one partial per 2048 elements)
 * is a static __device__ array of the library, 128 KB on every GPU that loads it, held until the process exits.
 * Limits: 1 <= a->n <= UAVMVS_REDUCE_NOALLOC_MAX_N, larger inputs are rejected (return 2, message in uavmvs_last_error).
 * The additions have a fixed order: repeated calls on the same data give the same bits (sum within 1e-6 relative of
 * uavmvs_reduce, whose atomics fix no order; min / max identical).  ALL calls on one device share the scratch array:
 * calls that overlap on the GPU (different streams or host threads) race; calls on one stream are ordered by it. */
#define UAVMVS_REDUCE_NOALLOC_MAX_N (1u << 24)
int uavmvs_reduce_noalloc(void * stream, const uavmvs_array * a, uavmvs_array * out3);

/* ---- Whole evaluation as apps/evaluate_trajectory/evaluate_trajectory.cu:125-202 runs it:
 * clear obs, K1 over all views, K2, K3, K8, K9.  views are HOST memory (copied inside).
 * stats (HOST, may be NULL): [0] = occlusion rays traced, [1] = visible observations,
 * [2] = (point,view) pairs evaluated, [3] = pairs surviving the block-level cull.
 * out3 = {sum(wrecon), min, max}; score = out3[0] / num_points. */
int uavmvs_evaluate_trajectory(void * stream, const float * views, uint32_t num_views,
    float max_distance, float target_recon, int width, int height,
    const uavmvs_bvh * bvh, const uavmvs_cloud * cloud, uavmvs_obs * obs,
    uavmvs_array * recons, uavmvs_array * wrecons, float out3[3], uint64_t stats[4]);
int uavmvs_evaluate_trajectory_device(void * stream, const float * d_views, uint32_t num_views,
    float max_distance, float target_recon, int width, int height,
    const uavmvs_bvh * bvh, const uavmvs_cloud * cloud, uavmvs_obs * obs,
    uavmvs_array * recons, uavmvs_array * wrecons, float out3[3], uint64_t stats[4]);

/* The same evaluation followed by the download the app does next (the HOST copy of the per-point arrays,
 * apps/evaluate_trajectory/evaluate_trajectory.cu:188-194 and :215-240): h_recons[num_points] and h_num_rows[num_points]
 * (either may be NULL; pinned memory lets the copies overlap) are filled while the evaluation is still running - the
 * counters beside K2 + K3, the scores in four chunks each copied while the next is computed - and are complete when the
 * call returns.  Same results as uavmvs_evaluate_trajectory + two downloads. */
int uavmvs_evaluate_trajectory_to_host(void * stream, const float * views, uint32_t num_views,
    float max_distance, float target_recon, int width, int height,
    const uavmvs_bvh * bvh, const uavmvs_cloud * cloud, uavmvs_obs * obs,
    uavmvs_array * recons, uavmvs_array * wrecons, float * h_recons, uint32_t * h_num_rows,
    float out3[3], uint64_t stats[4]);

/* ---- Multi-GPU: the proxy cloud is sharded by points, one process per GPU, mesh / KD-tree / trajectory replicated
 * (every output of the path is per point).  The only exchange of the evaluation is the trajectory score, which the
 * reference computes on one GPU with cacc::reduction::sum / min / max (libs/cacc/reduction.h:132-162) and
 * avg_wrecon = sum / N (apps/optimize_trajectory/optimize_trajectory.cu:529).
 * A communicator wraps an ncclComm_t.  libnccl.so.2 is resolved with dlopen at the first use (the copy already
 * loaded into the process, e.g. PyTorch's, else the system's): single-GPU users never need NCCL. */
typedef struct uavmvs_comm uavmvs_comm;
#define UAVMVS_COMM_ID_BYTES 128
/* ncclGetUniqueId on one rank; the caller ships the 128 bytes to the other ranks (MPI, torch.distributed, a file). */
int uavmvs_comm_unique_id(uint8_t id[UAVMVS_COMM_ID_BYTES]);
/* ncclCommInitRank on the calling thread's current device (collective over all `world` ranks). */
int uavmvs_comm_create(const uint8_t id[UAVMVS_COMM_ID_BYTES], int world, int rank, uavmvs_comm ** out);
/* Wraps an ncclComm_t the application already owns (not destroyed by uavmvs_comm_destroy). */
int uavmvs_comm_adopt(void * nccl_comm, int world, int rank, uavmvs_comm ** out);
int uavmvs_comm_info(const uavmvs_comm * comm, int * world, int * rank);
int uavmvs_comm_destroy(uavmvs_comm * comm);

#define UAVMVS_SCORE_DOUBLES 8
/* The evaluation of uavmvs_evaluate_trajectory* on this rank's shard of the cloud, followed on the same stream --
 * no host synchronisation in between -- by ONE collective: an all-gather of a packed 64-byte record per rank
 * {sum(wrecon) (double), points, rays, observations, pairs after the block cull, min, max, pairs}, folded on the device
 * in rank order, so every rank returns the same bits:
 *   score[0] = sum over all shards of wrecon   [1] = number of points   -> avg_wrecon = score[0] / score[1]
 *   score[2] = min, [3] = max of wrecon        [4] = occlusion rays, [5] = visible observations,
 *   score[6] = (point, view) pairs surviving the block-level cull, [7] = (point, view) pairs evaluated.
 * views: HOST [num_views][28] when views_on_device == 0, else a device pointer.  h_recons / h_num_rows: optional HOST
 * buffers of this rank's shard, filled as by uavmvs_evaluate_trajectory_to_host.  comm == NULL: single GPU (no NCCL).
 * local_out3 (may be NULL): this shard's {sum, min, max}. */
int uavmvs_evaluate_trajectory_sharded(void * stream, uavmvs_comm * comm, const float * views, int views_on_device,
    uint32_t num_views, float max_distance, float target_recon, int width, int height,
    const uavmvs_bvh * bvh, const uavmvs_cloud * cloud, uavmvs_obs * obs,
    uavmvs_array * recons, uavmvs_array * wrecons, float * h_recons, uint32_t * h_num_rows,
    double score[UAVMVS_SCORE_DOUBLES], float local_out3[3]);
/* SUM all-reduce (ncclAllReduce, in place, on `stream`) of the first `count` floats of a device array: the partial
 * [B][bins] sphere histograms of the batched K4 / K5 when the cloud is sharded. */
int uavmvs_allreduce_array(void * stream, uavmvs_comm * comm, uavmvs_array * a, uint64_t count);
/* All-gather of per-point vectors in shard order: every rank contributes `count` floats (device array `local`) and
 * receives world * count floats in `all` (device array); shards shorter than `count` are padded by the caller. */
int uavmvs_allgather_array(void * stream, uavmvs_comm * comm, const uavmvs_array * local, uint64_t count,
    uavmvs_array * all);

/* A trajectory resident on the device for repeated evaluation ([num_views][28] floats); the device pointer goes to the
 * *_device / views_on_device entry points.  (The apps keep the cameras on the host and pass one view per launch,
 * apps/evaluate_trajectory/evaluate_trajectory.cu:154-164.) */
int uavmvs_views_upload(void * stream, const float * views, uint32_t num_views, float ** d_views);
/* Overwrites the first num_views views of a device array (stream-ordered; `views` must stay valid until the copy ran). */
int uavmvs_views_update(void * stream, float * d_views, const float * views, uint32_t num_views);
int uavmvs_views_download(void * stream, const float * d_views, uint32_t num_views, float * views);
int uavmvs_views_free(float * d_views);

/* Kernel launches issued by the evaluation path (K1 pipeline, K2/K3/K8, K9, histogram and closest-point kernels) in this
 * process since the last reset; diagnostics for bench.py's gpu_launches (no reference counterpart). */
int uavmvs_launch_count(uint64_t * out, int reset);

/* Device time (ms, CUDA events on the call's stream) of the stages of the calling thread's last
 * uavmvs_evaluate_trajectory*: ms[0] = K1 (all views), [1] = K2, [2] = K3+K8, [3] = K9.  The
 * reference only wall-clocks the K1 loop (apps/evaluate_trajectory/evaluate_trajectory.cu:144-171). */
int uavmvs_profile_last(float ms[4]);
/* Per-kernel device time of the K1 pipeline (no reference counterpart).  After uavmvs_profile_k1_enable(1) every K1
 * launch records CUDA events at its kernel boundaries; uavmvs_profile_k1 returns, for the calling thread's last K1
 * launch, ms[0] = k_view_masks, [1] = k_block_caps + k_sum_caps + scan, [2] = k_cull_emit, [3] = k_trace,
 * [4] = k_retire (summed over the view batches). */
int uavmvs_profile_k1_enable(int on);
int uavmvs_profile_k1(float ms[5]);
/* Device time (ms) between the end of K9 and the end of the score collective (pack, all-gather, fold) of the calling
 * thread's last uavmvs_evaluate_trajectory_sharded. */
int uavmvs_profile_last_collective(float * ms);

/* Diagnostics of the calling thread's K1 launches since the last reset (no reference counterpart):
 * out[0] = (block, view) packets tested by the visibility-shaft pass, [1] = packets certified
 * visible as a whole, [2] = occlusion rays answered by a certificate, [3] = occlusion rays traced. */
int uavmvs_k1_counters(uint64_t out[4], int reset);

/* ---- K4 populate_spherical_histogram (6-arg), kernels.h:46-50, kernels.cu:293-336 */
int uavmvs_populate_spherical_histogram(void * stream, const float view_pos[3], float max_distance,
    const uavmvs_bvh * bvh, const uavmvs_cloud * cloud, const uavmvs_kdtree * kd,
    uavmvs_array * sphere_hist);
/* ---- K5 populate_spherical_histogram (9-arg), kernels.h:52-67, kernels.cu:338-390 */
int uavmvs_populate_spherical_histogram_recon(void * stream, const float view_pos[3],
    float max_distance, float target_recon, const uavmvs_bvh * bvh, const uavmvs_cloud * cloud,
    const uavmvs_kdtree * kd, const uavmvs_obs * obs, const uavmvs_array * recons,
    uavmvs_array * sphere_hist);
/* ---- K6 evaluate_spherical_histogram, kernels.h:69-79, kernels.cu:392-446 */
int uavmvs_evaluate_spherical_histogram(void * stream, const float calib[9], int width, int height,
    const uavmvs_kdtree * kd, const uavmvs_array * sphere_hist, uavmvs_image * hist);
/* ---- batched K4 / K5: num_positions candidate camera positions in one call.  Replaces the per-candidate
 * launch loops of apps/generate_guidance_volume/generate_guidance_volume.cu:236-253 (K4: obs = recons = NULL)
 * and apps/optimize_trajectory/optimize_trajectory.cu:388-398 (K5).  positions: HOST [num_positions][3].
 * sphere_hists: device array of >= num_positions * bins floats, row b = the sphere_hist of position b;
 * contributions are ADDED (the caller nulls it first, like dcon_hist->null()).  Candidates go through the
 * same ray-list pipeline as K1 (block-level range / back-face cull, compacted rays, persistent trace). */
int uavmvs_populate_spherical_histogram_batch(void * stream, const float * positions, uint32_t num_positions,
    float max_distance, float target_recon, const uavmvs_bvh * bvh, const uavmvs_cloud * cloud,
    const uavmvs_kdtree * kd, const uavmvs_obs * obs, const uavmvs_array * recons, uavmvs_array * sphere_hists);

/* ---- batched K6 (+ the host arg-min loop of optimize_trajectory.cu:416-428 as a device reduction).
 * For every position b: image[y][x] of kernels.cu:392-446 from row b of sphere_hists.  images: NULL or a
 * device array of >= num_positions * image_height * image_width floats (dense [b][y][x]).  h_min / h_argmin /
 * h_max / h_argmax: HOST [num_positions] or NULL; arg = y * image_width + x of the FIRST extremum in
 * row-major order.  When any of them is given the call synchronises the stream.  The (pixel, bin) frustum
 * matrix is cached in the KD-tree handle per (calib, width, height, image size). */
int uavmvs_evaluate_spherical_histogram_batch(void * stream, const float calib[9], int width, int height,
    uavmvs_kdtree * kd, const uavmvs_array * sphere_hists, uint32_t num_positions, uint32_t image_width,
    uint32_t image_height, uavmvs_array * images, float * h_min, uint32_t * h_argmin, float * h_max, uint32_t * h_argmax);

/* ---- One Nelder-Mead step of the optimiser for a batch of views, on the device: libs/opti/nelder_mead.h:29-121 driven by
 * the objective of apps/optimize_trajectory/optimize_trajectory.cu:366-438 (altitude test, airspace test through
 * acc::BVHTree::closest_point within min_distance, K5, K6, first arg-min of w * image).  The reference runs this per view
 * on a host thread with a D2H copy and a host arg-min per evaluation; here the views advance in lock-step -- three batched
 * evaluations (4 vertices, reflected point, expansion / contraction candidate) with the decisions taken by device kernels
 * in between -- and nothing is copied to the host inside a step.
 * A handle keeps the simplices ([num_views][4][3] floats, Simplex<3>::verts) and the results of the last step:
 * ret_vid / ret_value = nelder_mead()'s return pair, vmin / vtheta / vphi = the closure's best orientation (:430-434).
 * `active`: HOST indices of the views that take the step (mutually independent ones, :286-310).  airspace = the tree
 * closest_point runs on (the apps use the proxy mesh itself).  num_bins = vertices of the KD-tree (1272). */
typedef struct uavmvs_nm uavmvs_nm;
int uavmvs_nm_create(uint32_t num_views, uint32_t num_bins, uavmvs_nm ** out);
int uavmvs_nm_upload(uavmvs_nm * nm, const float * simplices);
int uavmvs_nm_download(void * stream, const uavmvs_nm * nm, float * simplices, uint32_t * ret_vid, float * ret_value,
    float * vmin, float * vtheta, float * vphi);
int uavmvs_nm_step(void * stream, uavmvs_nm * nm, const uint32_t * active, uint32_t num_active, float min_distance,
    float max_distance, float target_recon, const float calib[9], int width, int height, const uavmvs_bvh * bvh,
    const uavmvs_bvh * airspace, const uavmvs_cloud * cloud, uavmvs_kdtree * kd, const uavmvs_obs * obs, const uavmvs_array * recons);
/* The poses the views of the last uavmvs_nm_step take (optimize_trajectory.cu:441-467): position = the returned simplex
 * vertex, orientation = utp::rotation_from_spherical(vtheta, vphi) (libs/utp/trajectory.cpp:14-39), w2c = [rot | -rot pos];
 * views whose step returned a value >= 0 keep their pose (their re-initialisation, :446-461, stays with the application).
 * d_views: DEVICE [num_views][28] in the order of that step's `active`, updated in place -- ready for the re-adding
 * uavmvs_update_observation_rays_batch_device, so the optimiser's iteration needs no copy to the host in its middle. */
int uavmvs_nm_apply_views(void * stream, const uavmvs_nm * nm, float * d_views, uint32_t num_views);
int uavmvs_nm_destroy(uavmvs_nm * nm);

/* ---- K7 estimate_capture_difficulty, kernels.h:81-84, kernels.cu:448-494 */
int uavmvs_estimate_capture_difficulty(void * stream, float max_distance, const uavmvs_bvh * bvh,
    uint32_t mesh_size, const uavmvs_kdtree * kd, uavmvs_cloud * cloud);

/* Device pointer escape hatch for callers that launch their own kernels on the data
 * (the reference's cdata()/accessor() PODs): which = 0 obs data, 1 obs num_rows,
 * 2 cloud vertices (float4), 3 cloud normals (float4). */
int uavmvs_obs_device_ptr(const uavmvs_obs * obs, int which, void ** ptr, uint64_t * pitch_bytes);

#ifdef __cplusplus
}
#endif
#endif /* UAVMVS_B200_H */

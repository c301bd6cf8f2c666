// Shared types of libuavmvs_b200: opaque handle layouts, error plumbing and the
// fp32 helpers that fix the operation order of the reference.
// Compiled with --fmad=false, IEEE div/sqrt, no FTZ: every expression below is
// evaluated exactly as written.
#pragma once
#include <vector>
#include <mutex>
#include <cuda_runtime.h>
#include <stdint.h>
#include <cfloat>
#include <cstdio>
#include <cstring>
#include <string>
#include <atomic>

#include "../../include/uavmvs_b200.h"

#define UAV_NAI 0xffffffffu
#define UAV_INF __int_as_float(0x7f800000)

namespace uav {

void set_error(const char * file, int line, const char * what, const char * detail);
int fail(const char * file, int line, const char * what);

#define UAV_CUDA(call)                                                              \
    do {                                                                            \
        cudaError_t e_ = (call);                                                    \
        if (e_ != cudaSuccess) {                                                    \
            uav::set_error(__FILE__, __LINE__, #call, cudaGetErrorString(e_));      \
            return 1;                                                               \
        }                                                                           \
    } while (0)
#define UAV_REQUIRE(cond, msg)                                                      \
    do { if (!(cond)) { uav::set_error(__FILE__, __LINE__, #cond, msg); return 2; } } while (0)

#ifndef UAV_STACK
#define UAV_STACK 96              // binary walks (trace.cuh, closest.cu): one entry per level
#endif
constexpr int K1_STACK = 160;     // 4-wide walk of k_trace: <= 3 pushes per wide level (= two binary levels)

static inline unsigned divup(uint64_t a, uint64_t b) { return (unsigned)((a + b - 1) / b); }

// Launch counter of the evaluation path (uavmvs_launch_count): every `<<<..., counted(s)>>>` adds one.
extern std::atomic<unsigned long long> g_launches;
static inline cudaStream_t counted(cudaStream_t s, unsigned n = 1) { g_launches.fetch_add(n, std::memory_order_relaxed); return s; }

// ---------------------------------------------------------------- device layouts
// Two-child node, 64 B, one aligned 4 x LDG.128: both child boxes (exact floats, the
// reference's boxes bit for bit) and both child references.
//   q0 = lmin.xyz, lmax.x   q1 = lmax.yz, rmin.xy   q2 = rmin.z, rmax.xyz
//   q3 = {left_ref, right_ref, left_count, right_count}
// count == 0: ref = index of an inner node;  count > 0: ref = first triangle of a leaf.
struct __align__(64) Node64 { float4 q0, q1, q2; uint4 q3; };
// Four-child node, 128 B (one cache line): the up-to-four GRANDCHILDREN of a binary node (a leaf
// child stays one child), boxes as the reference's exact floats.
//   q[0..2] = boxes of children 0,1 (same packing as Node64 q0..q2), q[3..5] = children 2,3
//   cnt[c] bit 31 set: inner child, cnt[c] & 0x7fffffff = ref[c] = its node index
//   cnt[c] bit 31 clear: leaf child, cnt[c] = first_triangle << 6 | min(size, 31) (bit 5 is left free
//             for the traversal's "box not verified" flag; < 2^25 triangles), ref[c] = first triangle
//   empty slot: point box at 1e18 (no segment reaches it), ref = 0, cnt = 0x80000000
// The traversal kernels read q[0..5] and cnt only (7 of the 8 quads).
// Every inner binary node owns one (indexed like Node64); a traversal that starts at the root only
// ever touches those of even depth.
struct __align__(128) Node128 { float4 q[6]; uint4 ref; uint4 cnt; };
// Four-child node for k_trace, 80 B = 5 quads (the traversal is bound by the L1 data pipe: every quad
// read per node step costs ~0.24 ms on config 2): child boxes as fp16 OFFSETS from the node's centre c,
// rounded outwards (minima down, maxima up), so the decoded box contains the exact one.  That is all the
// traversal needs: inner boxes only have to be conservative, and a leaf reached through a conservative
// box is re-checked against its exact box once one of its triangles is accepted (see k_trace).
//   c = {cx, cy, cz, unused};  h[0..2] = 24 halves: child k -> mn.x mn.y mn.z mx.x mx.y mx.z at 6k..6k+5
//   cnt as in Node128;  empty slot: all offsets +inf
//   K1_PRMT layout: the 24 halves are 12 words (min, max), word 3k + a = child k, axis a, so that one PRMT orders a
//   word (near, far) for the ray's direction sign (k1.cu:k_trace)
struct __align__(16) Node80 { float4 c; uint4 h[3]; uint4 cnt; };
// K1_BF16 (with K1_PRMT): the offsets are stored as bf16 instead of fp16.  A bf16 is the upper half of an fp32, so ONE PRMT
// both picks the near (or far) half of a (min, max) word for the ray's direction sign and places it in the upper half of
// a zero word = the decoded fp32 plane offset: 2 instructions per word instead of 3 (PRMT + two HADD2.F32).  8 mantissa
// bits instead of 11: the outward-rounded boxes are up to 2^-8 of the offset larger, still conservative.
// Measured on config 2 (profiles/r02_x_bf16_nodes.txt): node step 156 -> 145 instructions, parity suite bit-exact, but
// k_trace 2.448 -> 2.577 ms: the looser boxes cost more node visits than the decode saves.  Off by default.
#ifndef K1_BF16
#define K1_BF16 0
#endif
#ifndef K1_PRMT
#define K1_PRMT 1       // measured on config 2 (profiles/r02_b_sweep.txt): k_trace 2.62 -> 2.54 ms at 7 CTAs per SM
#endif
// Pre-reduced triangle, 64 B: a, ab, ac, unit normal (N/|N|), d00, d01, d11, denom.
// All derived values are computed with the reference's operation order, so using them
// gives bit-identical operands (libs/acc/primitives.h:214-248, :134-153).
struct __align__(64) Tri64 { float4 q0, q1, q2, q3; };

}  // namespace uav

struct uavmvs_bvh {
    uint32_t num_nodes, num_tris, num_inner, num_wide;
    // reference layout (kept for export and for the oracle)
    uint4 * nodes;      // {first,last,left,right}
    float * aabbs;      // 6 per node
    float * tris;       // 9 per triangle, leaf order
    uint32_t * indices; // face ids
    // traversal layout
    uav::Node64 * inner;
    uav::Node128 * wide;
    uav::Node80 * wide80;
    float bound_abs[3];   // max |coordinate| of the root box per axis (error band of the fp16 node test)
    uav::Tri64 * ptris;
    float4 * leafbox;     // nested trees: the stored (exact) box of the leaf whose first triangle is t at [2t], [2t + 1]:
                          // {min.xyz, max.x}, {max.y, max.z, 0, 0}; the re-check of k_trace's deferred leaves
    uint32_t root_is_leaf, root_first, root_count;
    uint32_t depth;     // upper bound of the binary depth (2 x levels of the wide collapse), checked against the stacks at upload
    uint32_t nested;    // child boxes inside parent boxes, leaf boxes tight, leaves <= 31 triangles (checked at upload)
    uint32_t max_leaf;
    float max_leaf_diag = 0.0f;   // nested trees: largest leaf-box diagonal (k7.cu: how far behind a hit its leaf's box can begin)
};

struct uavmvs_kdtree {
    uint32_t num_nodes;
    uint4 * nodes;   // {vid, dim, left, right}
    float4 * verts;  // xyz
    // cache of the batched K6: bit (pixel, bin) = "sphere vertex bin projects inside the image of
    // orientation pixel" (kernels.cu:409-441), valid for one (calib, width, height, image size)
    // certified nearest-vertex lookup (kd.cuh: kd_find_nn_fast), built lazily by kd_fast_ensure (containers.cu)
    std::vector<uint32_t> h_nodes; std::vector<float> h_verts;     // host copies for that build
    int fast_state = 0;                 // 0 not built yet, 1 ready, -1 unsuitable point set (plain search only)
    uint32_t fast_m = 0, fast_res = 0;  // distinct positions, cube-map resolution per face
    uint16_t * fast_table = nullptr;    // [6][res][res] -> distinct position nearest to the cell centre
    float4 * fast_pos = nullptr;        // [m] xyz, w = half the certified radius
    uint4 * fast_rule = nullptr;        // [m] {vertex id A, vertex id B, dim (3 = always A), coordinate bits}
    uint16_t * fast_nbr = nullptr;      // [m][8] neighbouring distinct positions, 0xffff = none
    // (host threads sharing the handle: looked up / added under fr_mutex; one entry per (calib, width, height, image
    // size), every entry stays allocated until destroy since kernels of other threads may be reading it)
    struct FrustumBits { float calib[9]; int w, h; uint32_t iw, ih; uint32_t * bits; };   // bits: [words][pixels]
    std::mutex fr_mutex; std::vector<FrustumBits> fr_cache;
};

struct uavmvs_cloud {
    uint32_t n;
    float bound_abs[3];   // max |coordinate| of the points per axis (error bound of the fast image-reject proof, k1.cu)
    float4 * vertices, * normals;
    float * values, * qualities;
    // per 32-point block: bounding sphere (xyz, r) for the conservative view cull
    float4 * block_sphere;
    // per block: axis-aligned bounds of its points (exact min / max of the coordinates), the base
    // of the visibility shaft towards a camera (k1.cu:k_shaft)
    float4 * block_lo, * block_hi;
    // per block: oriented bounds of its points, 5 float4: rows 0..2 = orthonormal frame vector f_i (xyz)
    // and reference point c (w of rows 0..2 = c.x, c.y, c.z); row 3 = {lo0, hi0, lo1, hi1}, row 4 =
    // {lo2, hi2, 0, 0}: f_i . (p - c) in [lo_i, hi_i] for every point p of the block (padded).  f_0 is the
    // axis of the block's normal cone, so points sampled on one plane get a (near) zero-thickness slab.
    float4 * block_obb;
    // per block: cone of its normals (unit axis xyz, half-angle w in radians; w >= pi: unusable)
    // and the largest normal length, for the conservative back-face cull
    float4 * block_cone;
    float * block_nmax;
    uint32_t num_blocks;
    // per super-block (32 blocks = 1024 points): sphere enclosing its block spheres
    float4 * super_sphere;
    uint32_t num_super;
};

struct uavmvs_obs {
    uint32_t num_cols, max_rows;
    uint64_t pitch;          // bytes
    float4 * data;
    uint32_t * num_rows;
    uint32_t * dirty;        // one word per 32 columns: a column of the block changed since the last incremental evaluation
    // heuristic constants of THIS observation array (uavmvs_obs_configure_heuristic); has_heur == 0: the process-wide
    // ones of uavmvs_configure_heuristic at the time of the call, the reference's convention
    float heur[4]; int has_heur;
};

struct uavmvs_array { uint32_t n; float * data; };
// ncclComm_t + a device staging buffer for the score all-gather (eval.cu); comm.cu
struct uavmvs_comm { void * nccl = nullptr; int world = 1, rank = 0, owned = 0; double * gather = nullptr; };
struct uavmvs_image { uint32_t width, height; uint64_t pitch; float * data; };

namespace uav {

// ---------------------------------------------------------------- fp32 helpers
struct V3 { float x, y, z; };

__device__ __forceinline__ V3 mk(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator*(V3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ V3 operator/(V3 a, float s) { return mk(a.x / s, a.y / s, a.z / s); }
__device__ __forceinline__ V3 operator-(V3 a) { return mk(-a.x, -a.y, -a.z); }
// libs/cacc/vector.h:189-198: accumulate from zero, left to right
__device__ __forceinline__ float dot(V3 a, V3 b) {
    float r = 0.0f;
    r = r + a.x * b.x;
    r = r + a.y * b.y;
    r = r + a.z * b.z;
    return r;
}
// same value up to the sign of a zero result; used where only comparisons consume it
__device__ __forceinline__ float dot_nz(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
__device__ __forceinline__ float norm(V3 a) { return sqrtf(dot(a, a)); }
__device__ __forceinline__ V3 cross(V3 a, V3 b) {
    return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
__device__ __forceinline__ V3 normalized(V3 a) { return a / norm(a); }
// std::min/std::max NaN semantics (libs/acc/primitives.h:127-128)
__device__ __forceinline__ float smin(float a, float b) { return (b < a) ? b : a; }
__device__ __forceinline__ float smax(float a, float b) { return (a < b) ? b : a; }
__device__ __forceinline__ V3 ld3(float4 v) { return mk(v.x, v.y, v.z); }

struct Ray { V3 o, d; float tmin, tmax; };

// heuristic constants (configure_heuristic, libs/eval/kernels.cu:93-103); host-side global,
// passed to kernels by value
struct Heur { float m_k, m_x0, t_k, t_x0; };
Heur current_heur();
Heur heur_of(const uavmvs_obs * obs);     // the array's own constants, else the process-wide ones

// Batched spherical-histogram job run through the K1 ray-list pipeline (k1.cu): the "views" are bare
// candidate positions (no frustum), visible rays are binned into hist[view][bin] instead of becoming
// observation rows.  with_recon = 0: K4 (kernels.cu:293-336); 1: K5 (kernels.cu:338-390).
struct KdFastView {               // null table = not available
    const uint16_t * __restrict__ table; const float4 * __restrict__ pos; const uint4 * __restrict__ rule;
    const uint16_t * __restrict__ nbr; uint32_t res;
};
int kd_fast_ensure(uavmvs_kdtree * kd);       // containers.cu; 0 on success (also when the set is unsuitable)
KdFastView kd_fast_view(const uavmvs_kdtree * kd);

struct HistJob {
    const uint4 * kd_nodes; const float4 * kd_verts; uint32_t kd_n;
    KdFastView kd_fast;
    const float * values; const float * qualities;
    int with_recon; float target_recon;
    const float4 * obs; const uint32_t * num_rows; uint32_t max_rows, stride; const float * recons;
    float * hist;
    Heur heur;
    // order-independent accumulation (k1.cu:k_retire_hist): fixed-point accumulators [views][kd_n] and fx = {scale, bound}
    unsigned long long * acc; const float * fx;
    // 1: size the ray buffers for the worst case (every block a candidate of every position) instead of reading the
    // bound back from the device -- no host synchronisation at all; used by the device-resident optimiser step (nm.cu),
    // whose batches are a few dozen positions
    int no_readback;
};

// K1 pipeline (k1.cu).  d_stats (device, may be null): [0] += occlusion rays, [1] += visible
// observations, [3] += (point, view) pairs that survive the block-level cull.
int launch_k1(cudaStream_t s, int populate, const float * d_views, uint32_t num_views,
    float max_distance, int width, int height, const uavmvs_bvh * bvh, const uavmvs_cloud * cloud,
    uavmvs_obs * obs, unsigned long long * d_stats, const HistJob * hist_job = nullptr, const struct K1AfterRetire * after = nullptr);
// Optional continuation of a populate launch (eval.cu: K2 + K3 + K8 of a block range while the next range retires).  On the
// LAST view batch k_retire runs in `chunks` block ranges on the caller's stream; after each range `fn(ctx, side, block0,
// nblocks)` enqueues its work on `side`, which waits for that range; the caller's stream waits for `side` before
// launch_k1 returns.  ev: chunks + 1 events (no timing).
struct K1AfterRetire {
    uint32_t chunks; cudaStream_t side; cudaEvent_t * ev;
    int (*fn)(void * ctx, cudaStream_t side, uint32_t block0, uint32_t nblocks); void * ctx;
};
// k7.cu: K7 on the 4-wide tree (nested trees); 1 / 0 and the launch
int k7_wide_available(const uavmvs_bvh * bvh);
int k7_wide(cudaStream_t s, float max_distance, const uavmvs_bvh * bvh, uint32_t mesh_size, const uavmvs_kdtree * kd, uavmvs_cloud * cloud);
int k1_counters(unsigned long long out[4], int reset);
int k1_profile_enable(int on);
int k1_profile(float ms[5]);

// closest.cu: acc::BVHTree::closest_point for n DEVICE queries [n][3] -> device found / leaf-order triangle index / point
int launch_closest_point(cudaStream_t s, const uavmvs_bvh * bvh, const float * d_points, uint32_t n, float max_dist,
    uint8_t * d_found, uint32_t * d_tri_index, float * d_closest);
// hist.cu: batched K4 / K5 for DEVICE positions [n][3]; batched K6 leaving the packed extrema on the device
int populate_hist_batch_device(cudaStream_t s, const float * d_positions, uint32_t num_positions, float max_distance,
    float target_recon, const uavmvs_bvh * bvh, const uavmvs_cloud * cloud, const uavmvs_kdtree * kd, const uavmvs_obs * obs,
    const uavmvs_array * recons, float * d_sphere_hists, int no_readback);
// d_best: [2 * num_positions] packed keys (min keys, then max keys; see k_evaluate_spherical_histogram_batch), initialised
// here; d_weights: NULL or one factor per position applied to every pixel before the comparison
int evaluate_hist_batch_device(cudaStream_t s, const float calib[9], int width, int height, uavmvs_kdtree * kd,
    const float * d_sphere_hists, uint32_t num_positions, uint32_t image_width, uint32_t image_height, float * d_images,
    const float * d_weights, unsigned long long * d_best);

// comm.cu: NCCL collectives on a stream (f64 != 0: doubles, else floats; counts in elements per rank)
int comm_allgather(cudaStream_t s, uavmvs_comm * c, const void * send, void * recv, size_t count, int f64);
int comm_allreduce_sum_f32(cudaStream_t s, uavmvs_comm * c, void * buf, size_t count);

}  // namespace uav

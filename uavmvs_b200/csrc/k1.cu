// K1: update_observation_rays over a whole trajectory (libs/eval/kernels.cu:138-202 launched
// once per view by apps/evaluate_trajectory/evaluate_trajectory.cu:154-164), as a pipeline of
// small kernels over an explicit ray list:
//
//   k_view_masks   conservative (super-block, view) then (block, view) range/frustum cull
//                  -> masks[sb * V + v], bit b = "view v may see a point of block 32 sb + b"
//   k_block_caps   per block: 32 x candidate views = upper bound of its rays -> exclusive scan
//   k_cull_emit    warp per 32-point block: exact cull of kernels.cu:156-167 for the candidate
//                  views in view order; every surviving (point, view) pair becomes one ray
//                  record (unit direction, distance, point, view) in the block's slice
//   k_trace        persistent warps stream over the ray list; idle lanes are refilled while
//                  the others keep traversing the 4-wide BVH; leaf triangles are deferred
//                  until several lanes hold one; one visibility byte per ray
//   k_retire       warp per block, rays in (view, point) order: visible rays become
//                  observation rows.  In-order retirement reproduces the row order and the
//                  un-clamped counters of the reference's serial launch loop (:175-181).
//
// The ray list is why lanes never wait for each other's rays (rays binned by camera: the
// rays of one block towards one view are adjacent) and why the traversal kernel is small.
#include <cub/cub.cuh>
#include <algorithm>

#include "common.cuh"
#include "trace.cuh"
#include "evalmath.cuh"
#include "kd.cuh"

namespace uav {

constexpr int K1_WARPS = 4;
#ifndef K1_SB_PER_WARP_N
#define K1_SB_PER_WARP_N 1      // measured: 8 -> 0.26 ms, 4 -> 0.20, 2 -> 0.16, 1 -> 0.15 (the per-warp candidate loop is the critical path)
#endif
constexpr int K1_SB_PER_WARP = K1_SB_PER_WARP_N;
#ifndef K1_REFILL_N
#define K1_REFILL_N 20
#endif
#ifndef K1_TRI_N
#define K1_TRI_N 12
#endif
#ifndef K1_STEPS_N
#define K1_STEPS_N 6
#endif
#ifndef K1_PREFETCH
#define K1_PREFETCH 0             // 0 none (measured best), 1 L1, 2 L2
#endif
constexpr int K1_PEND = 6;        // deferred leaves per lane
constexpr int K1_REFILL = K1_REFILL_N;   // refill when at least this many lanes are idle
constexpr int K1_TRI = K1_TRI_N;  // run the triangle phase when this many lanes hold a leaf
constexpr int K1_STEPS = K1_STEPS_N;     // node steps between two rounds of warp-level bookkeeping

// Conservative (group of points, view) test: false only if NO point inside the sphere
// (centre c, radius r) can pass the exact cull of kernels.cu:164-167 (range, image bounds).
// Margins are 100x above fp32 rounding of the exact test, so a rejected group is rejected
// point by point by the exact arithmetic as well.
// Upper bound of sigma_max(R)^2, R = the 3 x 3 part of w2c: largest Gershgorin row sum of R^T R.  A point at distance <= r
// from the sphere's centre moves by at most sigma_max r in camera space; for a rotation R^T R = I and the bound is 1 (the
// Frobenius norm used before is sqrt 3: every block counted as 1.73x its size, and 41 % of the pairs queued for the exact
// cull then failed its image test).  Valid for any matrix; the caller's 1.001 covers the rounding.
__device__ __forceinline__ float view_sigma2(const float * __restrict__ view)
{
    const float * W = view + 3;
    const float m00 = W[0] * W[0] + W[4] * W[4] + W[8] * W[8], m11 = W[1] * W[1] + W[5] * W[5] + W[9] * W[9];
    const float m22 = W[2] * W[2] + W[6] * W[6] + W[10] * W[10];
    const float m01 = fabsf(W[0] * W[1] + W[4] * W[5] + W[8] * W[9]), m02 = fabsf(W[0] * W[2] + W[4] * W[6] + W[8] * W[10]);
    const float m12 = fabsf(W[1] * W[2] + W[5] * W[6] + W[9] * W[10]);
    const float g = fmaxf(fmaxf(m00 + m01 + m02, m01 + m11 + m12), m02 + m12 + m22);
    return (g == g) ? g : UAV_INF;                        // NaN: no bound, nothing is ever rejected
}

// sig2 = view_sigma2(view) (a per-view constant: k_view_masks computes it once per view and hands it round by shuffle)
__device__ __forceinline__ bool block_view_maybe(const float * __restrict__ view, float4 sph,
    float max_distance, int width, int height, float sig2, bool * inside = nullptr)
{
    if (inside) *inside = false;
    float dx = view[0] - sph.x, dy = view[1] - sph.y, dz = view[2] - sph.z;
    float D = sqrtf(dx * dx + dy * dy + dz * dz);
    if (D * 0.99999f - sph.w >= max_distance * 1.00001f) return false;
    if (width <= 0) return true;                          // bare candidate position (histogram jobs): no image test
    const float * W = view + 3; const float * K = view + 19;
    float q[3], e2 = 0.0f;
    #pragma unroll
    for (int i = 0; i < 3; ++i) {
        float a = W[4 * i] * sph.x, b = W[4 * i + 1] * sph.y, c = W[4 * i + 2] * sph.z, d = W[4 * i + 3];
        q[i] = a + b + c + d;
        float e = 1e-6f * (fabsf(a) + fabsf(b) + fabsf(c) + fabsf(d));
        e2 += e * e;
    }
    float qn = sqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2]);
    float rho = sph.w * sqrtf(sig2) * 1.001f + sqrtf(e2) + 1e-4f * (qn + 1.0f);
    float k2n = sqrtf(K[6] * K[6] + K[7] * K[7] + K[8] * K[8]);
    float p2 = K[6] * q[0] + K[7] * q[1] + K[8] * q[2];
    if (!(p2 - k2n * rho > 0.0f)) return true;          // ball not entirely in front: undecided
    float lim[4] = {0.5f, (float)width + 0.5f, 0.5f, (float)height + 0.5f};
    bool all_in = true;                                   // every point of the sphere surely projects inside the image
    #pragma unroll
    for (int s = 0; s < 4; ++s) {
        int r = s >> 1;                                   // image row of K: 0 -> x, 1 -> y
        float a0 = K[3 * r] - lim[s] * K[6], a1 = K[3 * r + 1] - lim[s] * K[7], a2 = K[3 * r + 2] - lim[s] * K[8];
        float an = sqrtf(a0 * a0 + a1 * a1 + a2 * a2);
        float v = a0 * q[0] + a1 * q[1] + a2 * q[2];
        if ((s & 1) == 0) { if (v + an * rho < 0.0f) return false; all_in = all_in && (v - an * rho > 0.0f); }   // coordinate < 0 everywhere
        else { if (v - an * rho > 0.0f) return false; all_in = all_in && (v + an * rho < 0.0f); }                // coordinate >= size everywhere
    }
    if (inside) *inside = all_in;
    return true;
}

// Conservative back-face test for a block: true only if EVERY point p in the sphere and every
// normal n in the cone (|n| <= nmax) gives dot((c - p) / |c - p|, n) < 0.087 (kernels.cu:160-162).
// The direction to the camera deviates from the centre's by at most asin(r / D); angles add up.
__device__ __forceinline__ bool block_backfacing(const float * __restrict__ view, float4 sph, float4 cone, float nmax)
{
    if (!(cone.w < 3.2f)) return false;
    float dx = view[0] - sph.x, dy = view[1] - sph.y, dz = view[2] - sph.z;
    float D = sqrtf(dx * dx + dy * dy + dz * dz);
    if (!(D > sph.w * 1.001f + 1e-6f)) return false;
    float c = (dx * cone.x + dy * cone.y + dz * cone.z) / D;
    float theta = acosf(fminf(fmaxf(c, -1.0f), 1.0f));
    float delta = asinf(fminf(sph.w * 1.001f / D, 1.0f));
    float amin = theta - cone.w - delta - 0.002f;        // smallest possible angle between direction and normal
    if (amin >= 1.5708f + 0.001f) return true;           // cos <= 0 for every pair
    if (amin <= 0.0f) return false;
    return nmax * cosf(amin) < 0.0865f;
}

// One warp = one chunk of 32 views (a view per lane, kept in registers) x K1_SB_PER_WARP
// super-blocks.  Surviving views are re-tested against the 32 blocks (a block per lane).
__global__ void __launch_bounds__(K1_WARPS * 32)
k_view_masks(const float * __restrict__ views, uint32_t num_views, const float4 * __restrict__ super_spheres,
    uint32_t num_super, const float4 * __restrict__ block_spheres, const float4 * __restrict__ block_cones,
    const float * __restrict__ block_nmax, uint32_t num_blocks,
    float max_distance, int width, int height, uint32_t * __restrict__ masks, uint32_t * __restrict__ inside_masks)
{
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t v = blockIdx.x * 32 + lane;
    const uint32_t sb0 = (blockIdx.y * K1_WARPS + w) * K1_SB_PER_WARP;
    if (sb0 >= num_super) return;
    float view[UAVMVS_VIEW_FLOATS];
    const float * src = views + (size_t)(v < num_views ? v : num_views - 1) * UAVMVS_VIEW_FLOATS;
    #pragma unroll
    for (int i = 0; i < UAVMVS_VIEW_FLOATS / 4; ++i) {
        float4 t = __ldg(reinterpret_cast<const float4 *>(src) + i);
        view[4 * i] = t.x; view[4 * i + 1] = t.y; view[4 * i + 2] = t.z; view[4 * i + 3] = t.w;
    }
    const float sig2 = width > 0 ? view_sigma2(view) : 0.0f;          // bare positions (histogram jobs) have no camera
    for (uint32_t sb = sb0; sb < min(sb0 + K1_SB_PER_WARP, num_super); ++sb) {
        bool cand = v < num_views && block_view_maybe(view, __ldg(super_spheres + sb), max_distance, width, height, sig2);
        uint32_t m = __ballot_sync(0xffffffffu, cand);
        uint32_t mine = 0, mine_in = 0;
        const uint32_t b = sb * 32 + lane;
        const float4 bs = __ldg(block_spheres + (b < num_blocks ? b : num_blocks - 1));
        const float4 bc = __ldg(block_cones + (b < num_blocks ? b : num_blocks - 1));
        const float bn = __ldg(block_nmax + (b < num_blocks ? b : num_blocks - 1));
        while (m) {
            const int j = __ffs(m) - 1; m &= m - 1;
            const float * vj = views + (size_t)(blockIdx.x * 32 + j) * UAVMVS_VIEW_FLOATS;   // uniform address
            bool in = false;
            const float sj = __shfl_sync(0xffffffffu, sig2, j);
            bool c = b < num_blocks && block_view_maybe(vj, bs, max_distance, width, height, sj, &in)
                && !block_backfacing(vj, bs, bc, bn);
            uint32_t word = __ballot_sync(0xffffffffu, c);
            uint32_t word_in = __ballot_sync(0xffffffffu, c && in);
            if ((int)lane == j) { mine = word; mine_in = word_in; }
        }
        if (v < num_views) { masks[(size_t)sb * num_views + v] = mine; inside_masks[(size_t)sb * num_views + v] = mine_in; }
    }
}

// caps[b] = 32 * number of candidate views of block b within [v0, v1).  Warp per super-block (32 blocks = the 32 bits
// of a mask word): every lane adds its view words into bit-sliced counters (7 planes, carry-save: ~14 logic ops per
// word count all 32 bits at once), eight loads in flight; the planes are then read out per bit and summed over the
// lanes with one redux each.  (A thread per block walking the views one dependent load at a time took 0.058 ms on
// config 2, all of it load latency.)
// Blocks with >= early_at candidate views are also appended to the early list (EarlyList below): the cull kernel starts
// them first and cuts the largest into pieces.
#define K1_EARLY_MAX 512u      // listed blocks per batch (the rest keeps its place in the grid)
#define K1_EARLY_PIECES 8u     // CTAs per listed block
// uint32 words of the list: [0] k_trace's work counter, [1] listed blocks, [2, 2 + MAX) pieces done per block -- zeroed per
// batch -- then block ids [MAX], slice base [MAX x PIECES] and ray count [MAX x PIECES] per piece
#define K1_EARLY_ZEROED (2u + K1_EARLY_MAX)
#define K1_EARLY_WORDS (2u + 2u * K1_EARLY_MAX + 2u * K1_EARLY_MAX * K1_EARLY_PIECES)
__global__ void k_block_caps(const uint32_t * __restrict__ masks, uint32_t num_views, uint32_t v0, uint32_t v1,
    uint32_t num_blocks, uint32_t * __restrict__ caps, uint32_t early_at, uint32_t * __restrict__ early)
{
    const uint32_t sb = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (sb >= (num_blocks + 31) / 32) return;                         // warp-uniform
    const uint32_t * row = masks + (size_t)sb * num_views;
    uint32_t c = 0;
    for (uint32_t r0 = v0; r0 < v1; r0 += 32u * 120u) {               // <= 120 words per lane and round: 7 planes hold 127
        const uint32_t r1 = min(v1, r0 + 32u * 120u);
        uint32_t p[7] = {0u, 0u, 0u, 0u, 0u, 0u, 0u};
        for (uint32_t base = r0; base < r1; base += 32u * 8u) {
            uint32_t w[8];
            #pragma unroll
            for (int j = 0; j < 8; ++j) { const uint32_t v = base + 32u * j + lane; w[j] = v < r1 ? __ldg(row + v) : 0u; }
            #pragma unroll
            for (int j = 0; j < 8; ++j) {
                uint32_t carry = w[j];
                #pragma unroll
                for (int k = 0; k < 7; ++k) { const uint32_t t = p[k] & carry; p[k] ^= carry; carry = t; }
            }
        }
        #pragma unroll
        for (uint32_t bit = 0; bit < 32; ++bit) {
            uint32_t cnt = 0;
            #pragma unroll
            for (int k = 0; k < 7; ++k) cnt |= ((p[k] >> bit) & 1u) << k;
            const uint32_t sum = __reduce_add_sync(0xffffffffu, cnt);
            if (lane == bit) c += sum;
        }
    }
    const uint32_t b = sb * 32 + lane;
    if (b < num_blocks) {
        caps[b] = 32u * c;
        if (early && c >= early_at) {
            const uint32_t slot = atomicAdd(early + 1, 1u);
            if (slot < K1_EARLY_MAX) early[K1_EARLY_ZEROED + slot] = b;
        }
    }
}

__global__ void k_sum_caps(const uint32_t * __restrict__ caps, uint32_t n, unsigned long long * total)
{
    unsigned long long s = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) s += caps[i];
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0 && s) atomicAdd(total, s);
}

__global__ void k_add_u64(unsigned long long * dst, const unsigned long long * src) { *dst += *src; }
__global__ void k_add_val(unsigned long long * dst, unsigned long long val) { *dst += val; }

// vcaps[v] = 32 x number of candidate blocks of view v = the ray-record bound of that view alone; the bound of a batch
// [v0, v1) is the sum of its views' (what k_block_caps + k_sum_caps give per batch).  One pass over the mask matrix lets
// the host cut a long trajectory into batches that fit the ray budget with ONE synchronisation instead of one per
// (tentative) batch.  grid = (views / 128, super-blocks / 64); coalesced along the view index.
__global__ void k_view_caps(const uint32_t * __restrict__ masks, uint32_t num_views, uint32_t num_super, uint32_t * __restrict__ vcaps)
{
    const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= num_views) return;
    const uint32_t sb0 = blockIdx.y * 64u, sb1 = min(num_super, sb0 + 64u);
    uint32_t c = 0;
    for (uint32_t sb = sb0; sb < sb1; ++sb) c += __popc(__ldg(masks + (size_t)sb * num_views + v));
    if (c) atomicAdd(vcaps + v, 32u * c);
}

// ---- visibility shafts ----------------------------------------------------------------------
// A (block, view) packet is CERTIFIED when no ray from any point of the block to the camera can be
// occluded, proven without tracing the rays one by one:
//   every ray of the packet lies in the shaft  S = { (1 - s) p + s a : p in [lo, hi], s in [s0, 1] }
//   (a = camera position, [lo, hi] = bounds of the block's points, s = t / l, and the reference
//   only accepts t in [0.001 l, l], kernels.cu:54-58).  At parameter s the cross-section of S is
//   the box [lo + s (a - lo), hi + s (a - hi)], so S meets a box B iff six inequalities LINEAR in s
//   share a solution in [s0, 1].  On a nested tree the reference reports "occluded" iff some leaf's
//   own box passes its slab test and holds an accepted triangle (see k_trace); if S misses the
//   eps-grown box of a node it misses every leaf box below it, and the exact fp32 slab test of any
//   ray inside S rejects those leaves (eps = 2e-6 max|coordinate| + 1.4e-5 is > 10x the rounding of
//   fl(fl(b - o) / d) and of the shaft arithmetic itself).  Hence: a traversal of the 4-wide tree
//   with the shaft that never touches a leaf box proves all rays of the packet visible.
// Packets that touch a leaf are left to k_trace (no claim is made about them).
struct Shaft {
    float lx, ly, lz, hx, hy, hz;         // lo - eps, hi + eps
    float rux, ruy, ruz, rwx, rwy, rwz;   // 1 / (a - lo), 1 / (a - hi)   (|.| clamped to >= 1e-6)
    bool upx, upy, upz, wpx, wpy, wpz;    // a - lo > 0, a - hi > 0
};

__device__ __forceinline__ float clamp_away(float u) { return fabsf(u) < 1e-6f ? 1e-6f : u; }

__device__ __forceinline__ bool shaft_hits(const Shaft & f, float mnx, float mny, float mnz, float mxx, float mxy, float mxz)
{
    // lo + s u <= bmax + eps   and   hi + s w >= bmin - eps, per axis
    const float ax = (mxx - f.lx) * f.rux, bx = (mnx - f.hx) * f.rwx;
    const float ay = (mxy - f.ly) * f.ruy, by = (mny - f.hy) * f.rwy;
    const float az = (mxz - f.lz) * f.ruz, bz = (mnz - f.hz) * f.rwz;
    float s_lo = 0.0009f, s_hi = 1.0001f;
    s_hi = fminf(s_hi, f.upx ? ax : UAV_INF); s_lo = fmaxf(s_lo, f.upx ? -UAV_INF : ax);
    s_lo = fmaxf(s_lo, f.wpx ? bx : -UAV_INF); s_hi = fminf(s_hi, f.wpx ? UAV_INF : bx);
    s_hi = fminf(s_hi, f.upy ? ay : UAV_INF); s_lo = fmaxf(s_lo, f.upy ? -UAV_INF : ay);
    s_lo = fmaxf(s_lo, f.wpy ? by : -UAV_INF); s_hi = fminf(s_hi, f.wpy ? UAV_INF : by);
    s_hi = fminf(s_hi, f.upz ? az : UAV_INF); s_lo = fmaxf(s_lo, f.upz ? -UAV_INF : az);
    s_lo = fmaxf(s_lo, f.wpz ? bz : -UAV_INF); s_hi = fminf(s_hi, f.wpz ? UAV_INF : bz);
    return !(s_lo > s_hi + 2e-6f);        // NaN (never produced: all operands finite) would count as a hit
}

// A triangle inside the shaft does no harm if its PLANE is not crossed by any ray of the packet
// within the accepted parameter range: with h(p) = n . (p - a_tri) the signed height of an origin
// and H that of the camera, the height along a ray is (1 - s) h + s H, and the reference rejects
// the triangle when the crossing has t < 0.001 l or t > l (libs/acc/primitives.h:228-231).  The
// block's oriented bounds give [h_lo, h_hi] over all its points; the triangle is cleared iff the
// height keeps the sign of H, by more than mu, at s = 0.00099 and s = 1.0001 for both ends of the
// interval (mu > 10x the rounding of the reference's own fl-dot(n, o - a) and t = -h / cos).
// This is what lets packets leave the surface they were sampled on: coplanar and concave
// neighbours have h >= 0 = sign of H.
__device__ __forceinline__ bool tri_cleared(const Tri64 * __restrict__ tp, const float4 * ob, float px, float py, float pz, float eps)
{
    const float4 q0 = __ldg(&tp->q0), q2 = __ldg(&tp->q2);
    const float ax = q0.x, ay = q0.y, az = q0.z, nx = q2.y, ny = q2.z, nz = q2.w;
    const float4 o0 = ob[0], o1 = ob[1], o2 = ob[2], o3 = ob[3], o4 = ob[4];
    const float g0 = nx * o0.x + ny * o0.y + nz * o0.z;
    const float g1 = nx * o1.x + ny * o1.y + nz * o1.z;
    const float g2 = nx * o2.x + ny * o2.y + nz * o2.z;
    const float rx = o0.w - ax, ry = o1.w - ay, rz = o2.w - az;          // c - a_tri
    const float base = nx * rx + ny * ry + nz * rz;
    const float a0 = g0 * o3.x, b0 = g0 * o3.y, a1 = g1 * o3.z, b1 = g1 * o3.w, a2 = g2 * o4.x, b2 = g2 * o4.y;
    const float h_lo = base + (fminf(a0, b0) + fminf(a1, b1) + fminf(a2, b2));
    const float h_hi = base + (fmaxf(a0, b0) + fmaxf(a1, b1) + fmaxf(a2, b2));
    const float ex = px - ax, ey = py - ay, ez = pz - az;
    const float H = nx * ex + ny * ey + nz * ez;
    const float d1 = (fabsf(rx) + fabsf(ry) + fabsf(rz)) + (fabsf(ex) + fabsf(ey) + fabsf(ez))
        + ((o3.y - o3.x) + (o3.w - o3.z) + (o4.y - o4.x));
    const float mu = eps + 1e-6f * d1;
    const float n0 = 0.99901f, n1 = 0.00099f, f0 = 1.0001f, f1 = 0.0001f;
    const bool pos = H > 0.0f && (n0 * h_lo + n1 * H > mu) && (f0 * H - f1 * h_hi > mu);
    const bool neg = H < 0.0f && (n0 * h_hi + n1 * H < -mu) && (f0 * H - f1 * h_lo < -mu);
    return pos || neg;                                                    // NaN anywhere: not cleared
}

constexpr int SH_STACK = 160;

struct ShaftArgs {
    const Node128 * __restrict__ wide; const Tri64 * __restrict__ tris;
    const float4 * __restrict__ lo; const float4 * __restrict__ hi; const float4 * __restrict__ obb;
    uint32_t num_blocks;
    const float * __restrict__ views; uint32_t num_views, v0, v1;
    const uint32_t * __restrict__ masks;
    uint32_t * cert; uint32_t nchunks;       // cert[blk * nchunks + (v - v0) / 32], bit (v - v0) % 32
    unsigned long long * counters;           // [0] packets tested, [1] packets certified
    uint32_t max_steps;
};

// One warp per block; candidate views (bits of k_view_masks) are gathered 32 at a time and each
// lane walks the tree with the shaft of its own view.
__global__ void __launch_bounds__(K1_WARPS * 32)
k_shaft(ShaftArgs a)
{
    __shared__ uint32_t s_q[K1_WARPS][64];
    __shared__ float4 s_obb[K1_WARPS][5];
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t blk = blockIdx.x * K1_WARPS + w;
    if (blk >= a.num_blocks) return;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const float4 lo = __ldg(a.lo + blk), hi = __ldg(a.hi + blk);
    if (lane < 5) s_obb[w][lane] = __ldg(a.obb + 5 * (size_t)blk + lane);
    __syncwarp();
    const uint32_t * mrow = a.masks + (size_t)(blk >> 5) * a.num_views;
    const uint32_t mbit = blk & 31;
    uint32_t stack[SH_STACK];
    uint32_t qn = 0, tested = 0, certified = 0;
    unsigned long long dbg_sc = 0, dbg_sf = 0, dbg_tt = 0;
    for (uint32_t c0 = a.v0; c0 < a.v1 || qn > 0; c0 += 32) {
        if (c0 < a.v1) {
            const uint32_t vi = c0 + lane;
            const bool cand = vi < a.v1 && ((__ldg(mrow + vi) >> mbit) & 1u);
            const uint32_t m = __ballot_sync(0xffffffffu, cand);
            if (cand) s_q[w][qn + __popc(m & lt_mask)] = vi;
            qn += __popc(m);
            __syncwarp();
            if (qn < 32 && c0 + 32 < a.v1) continue;
        }
        // ---- up to 32 queued views: one shaft per lane ----
        const uint32_t take = min(qn, 32u);
        if (take == 0) continue;
        const bool active = lane < take;
        const uint32_t v = s_q[w][active ? lane : 0];
        const float * vp = a.views + (size_t)v * UAVMVS_VIEW_FLOATS;
        const float px = __ldg(vp), py = __ldg(vp + 1), pz = __ldg(vp + 2);
        Shaft f; float eps;
        {
            float M = fmaxf(fmaxf(fabsf(px), fabsf(py)), fabsf(pz));
            M = fmaxf(M, fmaxf(fmaxf(fabsf(lo.x), fabsf(lo.y)), fabsf(lo.z)));
            M = fmaxf(M, fmaxf(fmaxf(fabsf(hi.x), fabsf(hi.y)), fabsf(hi.z)));
            eps = 2e-6f * M + 1.4e-5f;
            f.lx = lo.x - eps; f.ly = lo.y - eps; f.lz = lo.z - eps;
            f.hx = hi.x + eps; f.hy = hi.y + eps; f.hz = hi.z + eps;
            const float ux = clamp_away(px - lo.x), uy = clamp_away(py - lo.y), uz = clamp_away(pz - lo.z);
            const float wx = clamp_away(px - hi.x), wy = clamp_away(py - hi.y), wz = clamp_away(pz - hi.z);
            f.rux = 1.0f / ux; f.ruy = 1.0f / uy; f.ruz = 1.0f / uz;
            f.rwx = 1.0f / wx; f.rwy = 1.0f / wy; f.rwz = 1.0f / wz;
            f.upx = ux > 0.0f; f.upy = uy > 0.0f; f.upz = uz > 0.0f;
            f.wpx = wx > 0.0f; f.wpy = wy > 0.0f; f.wpz = wz > 0.0f;
        }
        bool ok = active && (fabsf(px + py + pz) < 1e30f);
        bool done = !ok;
        uint32_t cur = 0; int sp = 0; uint32_t steps = 0, tt = 0;
        while (__any_sync(0xffffffffu, !done)) {
            if (!done) {
                ++steps;
                const Node128 * np = a.wide + cur;
                const float4 q0 = __ldg(&np->q[0]), q1 = __ldg(&np->q[1]), q2 = __ldg(&np->q[2]);
                const float4 q3 = __ldg(&np->q[3]), q4 = __ldg(&np->q[4]), q5 = __ldg(&np->q[5]);
                const uint4 cnt = __ldg(&np->cnt);
                const bool h0 = shaft_hits(f, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y);
                const bool h1 = shaft_hits(f, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w);
                const bool h2 = shaft_hits(f, q3.x, q3.y, q3.z, q3.w, q4.x, q4.y);
                const bool h3 = shaft_hits(f, q4.z, q4.w, q5.x, q5.y, q5.z, q5.w);
                // leaf boxes inside the shaft: every triangle of theirs has to be cleared by its plane
                bool clear = sp <= SH_STACK - 4 && steps <= a.max_steps;
                #pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const bool hc = c == 0 ? h0 : (c == 1 ? h1 : (c == 2 ? h2 : h3));
                    const uint32_t cc = c == 0 ? cnt.x : (c == 1 ? cnt.y : (c == 2 ? cnt.z : cnt.w));
                    if (hc && (int)cc >= 0 && clear) {
                        const uint32_t first = cc >> 6, count = cc & 31u;
                        for (uint32_t i = 0; i < count && clear; ++i, ++tt)
                            clear = tri_cleared(a.tris + first + i, s_obb[w], px, py, pz, eps);
                    }
                }
                if (!clear) {
                    ok = false; done = true;              // no certificate
                } else {
                    if (h0 && (int)cnt.x < 0) stack[sp++] = cnt.x & 0x7fffffffu;
                    if (h1 && (int)cnt.y < 0) stack[sp++] = cnt.y & 0x7fffffffu;
                    if (h2 && (int)cnt.z < 0) stack[sp++] = cnt.z & 0x7fffffffu;
                    if (h3 && (int)cnt.w < 0) stack[sp++] = cnt.w & 0x7fffffffu;
                    if (sp > 0) cur = stack[--sp]; else done = true;
                }
            }
        }
        if (ok) atomicOr(a.cert + (size_t)blk * a.nchunks + ((v - a.v0) >> 5), 1u << ((v - a.v0) & 31u));
        tested += take; certified += __popc(__ballot_sync(0xffffffffu, ok));
        if (active) { if (ok) dbg_sc += steps; else dbg_sf += steps; dbg_tt += tt; }
        // ---- keep the views that did not fit ----
        __syncwarp();
        const uint32_t rest = qn - take;
        const uint32_t carry = lane < rest ? s_q[w][32 + lane] : 0u;
        __syncwarp();
        if (lane < rest) s_q[w][lane] = carry;
        qn = rest;
        __syncwarp();
    }
    if (a.counters && lane == 0 && tested) {
        atomicAdd(a.counters + 0, (unsigned long long)tested);
        atomicAdd(a.counters + 1, (unsigned long long)certified);
    }
    if (a.counters) {
        for (int o = 16; o > 0; o >>= 1) {
            dbg_sc += __shfl_xor_sync(0xffffffffu, dbg_sc, o); dbg_sf += __shfl_xor_sync(0xffffffffu, dbg_sf, o);
            dbg_tt += __shfl_xor_sync(0xffffffffu, dbg_tt, o);
        }
        if (lane == 0) { atomicAdd(a.counters + 4, dbg_sc); atomicAdd(a.counters + 5, dbg_sf); atomicAdd(a.counters + 6, dbg_tt); }
    }
}

// ---- packet entry points ----------------------------------------------------------------------
// The rays of one (block, view) packet all lie in the packet's shaft (see above), and they all walk
// the same upper levels of the tree.  k_entry walks those levels ONCE per packet with the shaft and
// records the frontier: a short list of wide nodes whose subtrees contain every leaf box the shaft
// touches.  A ray of the packet then starts with that list on its stack instead of the root.  The
// answer is unchanged: on a nested tree the reference's result is  OR over leaves (leaf box passes
// the exact slab test AND a triangle is accepted); a leaf whose box passes for some ray of the packet
// is touched by the shaft, so are all its ancestors, hence the walk below either emits one of them
// or reaches the leaf's parent, which is emitted itself.  Subtrees outside the shaft hold no such leaf.
//   rule: a touched inner child whose box is at most `tau` long is emitted, a longer one is expanded;
//   a node with a touched LEAF child is emitted as a whole; more than ENTRY_MAX entries or ENTRY_STEPS
//   expansions: the packet falls back to {root}.
constexpr int ENTRY_MAX = 15;          // entries per packet; record = 16 words: count, entries
constexpr int ENTRY_STEPS = 48;
constexpr int ENTRY_STACK = 48;

struct EntryArgs {
    const Node128 * __restrict__ wide;
    const float4 * __restrict__ lo; const float4 * __restrict__ hi;
    uint32_t num_blocks;
    const float * __restrict__ views; uint32_t num_views, v0, v1;
    const uint32_t * __restrict__ masks;
    const uint32_t * __restrict__ off;       // exclusive scan of 32 x candidate views: off[b] / 32 = first packet of block b
    uint32_t * fr;                           // [packets][16]
    float tau;
    unsigned long long * counters;           // [4] += frontier entries, [5] += packets that fell back to the root
};

__global__ void __launch_bounds__(K1_WARPS * 32)
k_entry(EntryArgs a)
{
    __shared__ uint32_t s_q[K1_WARPS][64];
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t blk = blockIdx.x * K1_WARPS + w;
    if (blk >= a.num_blocks) return;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const float4 lo = __ldg(a.lo + blk), hi = __ldg(a.hi + blk);
    const uint32_t * mrow = a.masks + (size_t)(blk >> 5) * a.num_views;
    const uint32_t mbit = blk & 31;
    const uint32_t pk0 = __ldg(a.off + blk) >> 5;
    uint32_t stack[ENTRY_STACK];
    uint32_t qn = 0, pdone = 0;
    unsigned long long n_entries = 0, n_fallback = 0;
    for (uint32_t c0 = a.v0; c0 < a.v1 || qn > 0; c0 += 32) {
        if (c0 < a.v1) {
            const uint32_t vi = c0 + lane;
            const bool cand = vi < a.v1 && ((__ldg(mrow + vi) >> mbit) & 1u);
            const uint32_t m = __ballot_sync(0xffffffffu, cand);
            if (cand) s_q[w][qn + __popc(m & lt_mask)] = vi;
            qn += __popc(m);
            __syncwarp();
            if (qn < 32 && c0 + 32 < a.v1) continue;
        }
        const uint32_t take = min(qn, 32u);
        if (take == 0) continue;
        const bool active = lane < take;
        const uint32_t v = s_q[w][active ? lane : 0];
        const float * vp = a.views + (size_t)v * UAVMVS_VIEW_FLOATS;
        const float px = __ldg(vp), py = __ldg(vp + 1), pz = __ldg(vp + 2);
        Shaft f;
        {
            float M = fmaxf(fmaxf(fabsf(px), fabsf(py)), fabsf(pz));
            M = fmaxf(M, fmaxf(fmaxf(fabsf(lo.x), fabsf(lo.y)), fabsf(lo.z)));
            M = fmaxf(M, fmaxf(fmaxf(fabsf(hi.x), fabsf(hi.y)), fabsf(hi.z)));
            const float eps = 2e-6f * M + 1.4e-5f;
            f.lx = lo.x - eps; f.ly = lo.y - eps; f.lz = lo.z - eps;
            f.hx = hi.x + eps; f.hy = hi.y + eps; f.hz = hi.z + eps;
            const float ux = clamp_away(px - lo.x), uy = clamp_away(py - lo.y), uz = clamp_away(pz - lo.z);
            const float wx = clamp_away(px - hi.x), wy = clamp_away(py - hi.y), wz = clamp_away(pz - hi.z);
            f.rux = 1.0f / ux; f.ruy = 1.0f / uy; f.ruz = 1.0f / uz;
            f.rwx = 1.0f / wx; f.rwy = 1.0f / wy; f.rwz = 1.0f / wz;
            f.upx = ux > 0.0f; f.upy = uy > 0.0f; f.upz = uz > 0.0f;
            f.wpx = wx > 0.0f; f.wpy = wy > 0.0f; f.wpz = wz > 0.0f;
        }
        uint32_t * rec = a.fr + ((size_t)(pk0 + pdone + lane) << 4);
        bool fallback = active && !(fabsf(px + py + pz) < 1e30f);
        bool done = !active || fallback;
        uint32_t cur = 0, nout = 0, steps = 0; int sp = 0;
        while (__any_sync(0xffffffffu, !done)) {
            if (!done) {
                const Node128 * np = a.wide + cur;
                const float4 q0 = __ldg(&np->q[0]), q1 = __ldg(&np->q[1]), q2 = __ldg(&np->q[2]);
                const float4 q3 = __ldg(&np->q[3]), q4 = __ldg(&np->q[4]), q5 = __ldg(&np->q[5]);
                const uint4 cnt = __ldg(&np->cnt);
                const bool h0 = shaft_hits(f, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y);
                const bool h1 = shaft_hits(f, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w);
                const bool h2 = shaft_hits(f, q3.x, q3.y, q3.z, q3.w, q4.x, q4.y);
                const bool h3 = shaft_hits(f, q4.z, q4.w, q5.x, q5.y, q5.z, q5.w);
                const bool leaf_hit = (h0 && (int)cnt.x >= 0) || (h1 && (int)cnt.y >= 0) || (h2 && (int)cnt.z >= 0) || (h3 && (int)cnt.w >= 0);
                if (leaf_hit) {
                    if (nout < ENTRY_MAX) rec[1 + nout] = cur;
                    ++nout;
                } else {
                    const float e0 = fmaxf(fmaxf(q0.w - q0.x, q1.x - q0.y), q1.y - q0.z);
                    const float e1 = fmaxf(fmaxf(q2.y - q1.z, q2.z - q1.w), q2.w - q2.x);
                    const float e2 = fmaxf(fmaxf(q3.w - q3.x, q4.x - q3.y), q4.y - q3.z);
                    const float e3 = fmaxf(fmaxf(q5.y - q4.z, q5.z - q4.w), q5.w - q5.x);
                    #pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        const bool hc = c == 0 ? h0 : (c == 1 ? h1 : (c == 2 ? h2 : h3));
                        const uint32_t cc = (c == 0 ? cnt.x : (c == 1 ? cnt.y : (c == 2 ? cnt.z : cnt.w))) & 0x7fffffffu;
                        const float ec = c == 0 ? e0 : (c == 1 ? e1 : (c == 2 ? e2 : e3));
                        if (hc) {
                            if (ec <= a.tau) { if (nout < ENTRY_MAX) rec[1 + nout] = cc; ++nout; }
                            else if (sp < ENTRY_STACK) stack[sp++] = cc;
                            else nout = ENTRY_MAX + 1;
                        }
                    }
                }
                ++steps;
                if (nout > ENTRY_MAX || steps > ENTRY_STEPS) { fallback = true; done = true; }
                else if (sp > 0) cur = stack[--sp];
                else done = true;
            }
        }
        if (active) {
            if (fallback) { rec[0] = 1u; rec[1] = 0u; ++n_fallback; n_entries += 1; }
            else { rec[0] = nout; n_entries += nout; }
        }
        __syncwarp();
        const uint32_t rest = qn - take;
        const uint32_t carry = lane < rest ? s_q[w][32 + lane] : 0u;
        __syncwarp();
        if (lane < rest) s_q[w][lane] = carry;
        qn = rest; pdone += take;
        __syncwarp();
    }
    if (a.counters) {
        for (int o = 16; o > 0; o >>= 1) {
            n_entries += __shfl_xor_sync(0xffffffffu, n_entries, o); n_fallback += __shfl_xor_sync(0xffffffffu, n_fallback, o);
        }
        if (lane == 0 && n_entries) { atomicAdd(a.counters + 4, n_entries); atomicAdd(a.counters + 5, n_fallback); }
    }
}

struct RayList {
    float4 * dirl;        // every ray, (view, point) order per block: unit point->camera direction, w = distance l
    uint8_t * vis;        // every ray: (point's index within its block) << 2 | status: 1 occluded, 2 visible
    float4 * tdir;        // rays that have to be traced (not certified by k_shaft): direction, w = l
    float4 * torg;        // ... origin (the point), w = bits of (ray index << 5 | point's index within its block)
    float4 * trd;         // ... traversal constants set up by the cull kernel on full warps: fl(1 / dir), w = error band em
                          //     (ray_fma_setup80), or w < 0: the ray takes the reference walk (axis-parallel direction,
                          //     single-leaf tree).  NULL: k_trace sets them up itself (packet modes)
    float bx, by, bz; uint32_t root_is_leaf;      // scene bounds of the error band, single-leaf tree flag
    const float * vproj;  // k_view_proj rows + margins per view, or NULL (histogram jobs: no image test)
    uint32_t reverse;     // dev: k_cull_emit_q walks the blocks from the last to the first
    const uint32_t * caps;   // [num_blocks] 32 x candidate views of the block in this batch (k_block_caps)
    uint32_t transpose_at;   // k_cull_emit_q: candidate views from which a block takes the transposed phase 1
    uint32_t * early;        // K1_EARLY_WORDS words (k_block_caps), or NULL: blocks the cull kernel starts first
    uint32_t early_at, split_at;   // candidate views from which a block is listed / views per piece of a listed block
    uint32_t * dbg_blocks;   // UAVMVS_K1_DEBUG: per block {cycles, rays, candidate views} of k_cull_emit_q
    uint32_t debug;       // UAVMVS_K1_DEBUG: k_cull_emit_q counts candidates / queued / rejected pairs into counters[4..7]
    const uint32_t * off; // [num_blocks] first entry of each block's slice (both lists)
    uint32_t * cnt;       // [num_blocks] rays in the slice
    uint32_t * tcnt;      // [num_blocks] rays of the slice that have to be traced
    uint32_t * rview;     // histogram jobs only: view (candidate position) index of every ray
    uint32_t * tpk;       // packet (block, candidate view) of every ray that has to be traced; NULL: entry points off
    const uint32_t * fr;  // [packets][16] entry points of k_entry: count, wide-node indices
    // speculative launch (launch_k1): the kernels of a batch are enqueued before the host knows the batch's ray-record
    // bound; they do nothing when the bound, already on the device, exceeds what the buffers hold
    const unsigned long long * g_total; unsigned long long g_cap;
};
#define K1_SPECULATION_GUARD(rl) do { if ((rl).g_total && *(rl).g_total > (rl).g_cap) return; } while (0)

// Exact cull (kernels.cu:153-167) + ray set-up values (:54-58).  One warp per block, views in
// order; lanes = points.  A slice is filled in (view, point) order.
template <bool HIST>
__global__ void __launch_bounds__(K1_WARPS * 32)
k_cull_emit(const float4 * __restrict__ verts, const float4 * __restrict__ normals, uint32_t n, uint32_t num_blocks,
    const float * __restrict__ views, uint32_t num_views, uint32_t v0, uint32_t v1,
    const uint32_t * __restrict__ masks, const uint32_t * __restrict__ inside_masks,
    const uint32_t * __restrict__ cert, uint32_t nchunks,
    float max_distance, int width, int height, RayList rl, unsigned long long * counters)
{
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t blk = blockIdx.x * K1_WARPS + (threadIdx.x >> 5);
    if (blk >= num_blocks) return;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const uint32_t id = blk * 32 + lane;
    const bool valid = id < n;
    const V3 v = ld3(verts[valid ? id : n - 1]), nrm = ld3(normals[valid ? id : n - 1]);
    const uint32_t * mrow = masks + (size_t)(blk >> 5) * num_views;
    const uint32_t * irow = inside_masks + (size_t)(blk >> 5) * num_views;
    const uint32_t mbit = blk & 31;
    const uint32_t base = rl.off[blk];
    uint32_t packet = base >> 5;                           // caps are 32 x candidate views: same numbering as k_entry
    uint32_t count = 0, tcount = 0;
    for (uint32_t g0 = v0; g0 < v1; g0 += 128) {
      // four mask words per lane in flight: the scan of the block's row is latency bound otherwise
      uint32_t mw4[4], iw4[4];
      #pragma unroll
      for (int u = 0; u < 4; ++u) {
          const uint32_t vi = g0 + 32 * u + lane;
          mw4[u] = vi < v1 ? __ldg(mrow + vi) : 0u;
          iw4[u] = (!HIST && vi < v1) ? __ldg(irow + vi) : 0u;
      }
      #pragma unroll
      for (int u = 0; u < 4; ++u) {
        const uint32_t c0 = g0 + 32 * u;
        if (c0 >= v1) break;
        uint32_t word = __ballot_sync(0xffffffffu, (mw4[u] >> mbit) & 1u);
        const uint32_t iword = __ballot_sync(0xffffffffu, (iw4[u] >> mbit) & 1u);   // whole block surely inside the image
        const uint32_t cword = cert ? __ldg(cert + (size_t)blk * nchunks + ((c0 - v0) >> 5)) : 0u;
        while (word) {
            const uint32_t kb = __ffs(word) - 1;
            const uint32_t k = c0 + kb; word &= word - 1;
            const bool certified = (cword >> kb) & 1u;             // every ray of this packet is visible (k_shaft)
            const float * view = views + UAVMVS_VIEW_FLOATS * (size_t)k;
            float vw[UAVMVS_VIEW_FLOATS];
            #pragma unroll
            for (int i = 0; i < UAVMVS_VIEW_FLOATS / 4; ++i) {          // uniform address: broadcast
                float4 t = __ldg(reinterpret_cast<const float4 *>(view) + i);
                vw[4 * i] = t.x; vw[4 * i + 1] = t.y; vw[4 * i + 2] = t.z; vw[4 * i + 3] = t.w;
            }
            V3 v2c = mk(vw[0], vw[1], vw[2]) - v;
            float l = norm(v2c);
            V3 v2cn = v2c / l;
            float ctheta = dot(v2cn, nrm);
            bool pass;
            if (HIST) {                                   // kernels.cu:310-316 / :358-364: scale first, no image test
                const float scale = 1.0f - (l / max_distance);
                pass = valid && !(scale <= 0.0f) && !(ctheta < 0.087f);
            } else {
                pass = valid && !(ctheta < 0.087f) && !(l >= max_distance);
                if (!((iword >> kb) & 1u) && pass) pass = in_image(vw, v, width, height);   // warp-uniform skip
            }
            uint32_t m = __ballot_sync(0xffffffffu, pass);
            if (pass) {
                const uint32_t r = __popc(m & lt_mask);
                const uint32_t idx = base + count + r;
                const float4 d4 = make_float4(v2cn.x, v2cn.y, v2cn.z, l);
                __stcs(rl.dirl + idx, d4);
                if (HIST) rl.rview[idx] = k;
                if (certified) rl.vis[idx] = (uint8_t)(lane << 2 | 2u);
                else {
                    const uint32_t j = base + tcount + r;
                    __stcs(rl.tdir + j, d4);
                    __stcs(rl.torg + j, make_float4(v.x, v.y, v.z, __uint_as_float(idx << 5 | lane)));
                    if (rl.tpk) rl.tpk[j] = packet;
                }
            }
            ++packet;
            count += __popc(m);
            if (!certified) tcount += __popc(m);
        }
      }
    }
    if (lane == 0) {
        rl.cnt[blk] = count; rl.tcnt[blk] = tcount;
        if (counters && count) { atomicAdd(counters + 2, (unsigned long long)(count - tcount)); atomicAdd(counters + 3, (unsigned long long)tcount); }
    }
}

// Per view: M = calib * w2c[0..2] (3 x 4, so that p = M [v, 1] is the reference's calib * (w2c * v), kernels.cu:21-27,
// :165-167, in ONE step) and a margin per row: both the reference's two-stage unfused evaluation and the fused
// evaluation with the rounded M differ from the real number p_i by less than 7.8e-7 A_i (13 roundings of 2^-24) and
// 3e-7 A_i (5 roundings) respectively, A_i = sum_jk |K_ij| |W_jk| |v_k| + sum_j |K_ij| |W_j3|; with |v_k| <= B_k (the cloud's bounds) A_i is a per-view constant and
// d_i = 2e-6 A_i bounds |p_i(reference) - P_i(fast)|.  vproj[v][16] = M (12), d0, d1, d2, 0.
__global__ void k_view_proj(const float * __restrict__ views, uint32_t num_views, float bx, float by, float bz, float * __restrict__ vproj)
{
    const uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= num_views) return;
    const float * W = views + (size_t)v * UAVMVS_VIEW_FLOATS + 3, * K = W + 16;
    const double B[4] = {(double)bx, (double)by, (double)bz, 1.0};
    float * o = vproj + 16 * (size_t)v;
    for (int i = 0; i < 3; ++i) {
        double A = 0.0;
        for (int j = 0; j < 4; ++j) {
            double m = 0.0, ma = 0.0;
            for (int k = 0; k < 3; ++k) { m += (double)K[3 * i + k] * (double)W[4 * k + j]; ma += fabs((double)K[3 * i + k] * (double)W[4 * k + j]); }
            o[4 * i + j] = (float)m;
            A += ma * B[j];
        }
        const double d = 2e-6 * A;
        o[12 + i] = (d == d && d < 1e30) ? (float)(d * 1.0000002) : __int_as_float(0x7f800000);   // NaN / inf: no proof
    }
    o[15] = 0.0f;
}

// Same result as k_cull_emit, fewer instructions: the exact cull costs ~400 instructions per 32 pairs
// (IEEE sqrt, 8 IEEE divisions, unfused multiply-adds) and only ~40 % of the candidate pairs survive it,
// most of the others failing the back-face test.  Phase 1 therefore proves rejections with a dozen
// fast-math instructions from the view position alone (back-face and range tests with a margin three
// orders of magnitude above the rounding error of either evaluation; NaNs never compare as rejected);
// every pair it cannot reject is queued, and phase 2 runs the exact arithmetic on FULL warps of queued
// pairs (lane = pair, not point).  The queue is filled in (view, point) order and drained in order, so
// the slice comes out in exactly the order k_cull_emit writes.  No packets: the opt-in shaft / entry-point
// passes keep using k_cull_emit.
template <bool HIST, bool DBG, int CW>
__global__ void __launch_bounds__(CW * 32, 32 / CW)
k_cull_emit_q(const float4 * __restrict__ verts, const float4 * __restrict__ normals, uint32_t n, uint32_t num_blocks,
    const float * __restrict__ views, uint32_t num_views, uint32_t v0, uint32_t v1,
    const uint32_t * __restrict__ masks, const uint32_t * __restrict__ inside_masks,
    float max_distance, float cull_distance, int width, int height, RayList rl, unsigned long long * counters)
{
    K1_SPECULATION_GUARD(rl);
    __shared__ float4 s_v[CW][32], s_n[CW][32];
    __shared__ uint32_t s_q[CW][64], s_vq[CW][64];
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t lt_mask = (1u << lane) - 1u;
    // Early list (CW == 1): the first K1_EARLY_MAX x K1_EARLY_PIECES CTAs of the grid belong to the listed blocks, so the
    // blocks with the most candidate views start at once instead of wherever their index puts them, and a block with
    // thousands of candidate views (a jump of the Z-order curve: on a config-5 shard ONE such block ran for 0.8 ms of a
    // 0.88 ms kernel) is cut into pieces by view range.  A piece writes at the slice position its candidate views give
    // it; the last piece to finish moves the pieces together, in view order, so the slice is what one warp would have
    // written.
    uint32_t blk = rl.reverse ? num_blocks - 1u - (blockIdx.x * CW + w) : blockIdx.x * CW + w;
    uint32_t piece = 0, npieces = 1, eslot = 0;
    if (CW == 1 && rl.early) {
        const uint32_t EG = K1_EARLY_MAX * K1_EARLY_PIECES;
        const uint32_t en = min(rl.early[1], K1_EARLY_MAX);
        if (blockIdx.x < EG) {
            eslot = blockIdx.x / K1_EARLY_PIECES; piece = blockIdx.x % K1_EARLY_PIECES;
            if (eslot >= en) return;
            blk = rl.early[K1_EARLY_ZEROED + eslot];
            npieces = min(K1_EARLY_PIECES, 1u + (rl.caps[blk] >> 5) / rl.split_at);
            if (piece >= npieces) return;
        } else {
            blk = rl.reverse ? num_blocks - 1u - (blockIdx.x - EG) : blockIdx.x - EG;
            if (blk >= num_blocks) return;
            if ((rl.caps[blk] >> 5) >= rl.early_at) {              // listed?  then its CTAs are at the front of the grid
                bool found = false;
                for (uint32_t i = lane; i < en; i += 32) found = found || rl.early[K1_EARLY_ZEROED + i] == blk;
                if (__any_sync(0xffffffffu, found)) return;
            }
        }
    }
    if (blk >= num_blocks) return;
    // this CTA's views: all of the batch, or the piece's share (multiples of 128 views: the loop below reads 4 x 32)
    uint32_t lv0 = v0, lv1 = v1;
    if (npieces > 1) {
        const uint32_t per = ((v1 - v0 + npieces - 1) / npieces + 127u) / 128u * 128u;
        lv0 = min(v1, v0 + piece * per); lv1 = min(v1, lv0 + per);
    }
    const uint32_t id = blk * 32 + lane;
    const bool valid = id < n;
    const float4 v4 = verts[valid ? id : n - 1];
    float4 n4 = normals[valid ? id : n - 1];
    n4.w = 0.087f - 1e-5f * (1.0f + fabsf(n4.x) + fabsf(n4.y) + fabsf(n4.z));   // phase 1's back-face limit of this point
    s_v[w][lane] = v4; s_n[w][lane] = n4;
    __syncwarp();
    const uint32_t * mrow = masks + (size_t)(blk >> 5) * num_views;
    const uint32_t * irow = inside_masks + (size_t)(blk >> 5) * num_views;
    const uint32_t mbit = blk & 31;
    uint32_t base = rl.off[blk];
    if (piece) {                                          // skip the records of the candidate views before this piece
        uint32_t c = 0;
        for (uint32_t vi = v0 + lane; vi < lv0; vi += 32) c += (__ldg(mrow + vi) >> mbit) & 1u;
        base += 32u * __reduce_add_sync(0xffffffffu, c);
    }
    const float md2 = cull_distance * cull_distance * 1.00001f;
    const float wph = ((float)width + 0.5f) * 1.000001f, hph = ((float)height + 0.5f) * 1.000001f;
    uint32_t count = 0, qn = 0;
    const bool dbg = DBG && counters;                     // UAVMVS_K1_DEBUG: where the candidate pairs go
    const long long dbg_t0 = DBG ? clock64() : 0ll;
    uint32_t dbg_views = 0;
    uint32_t n_cand = 0, n_queued = 0, n_rej_exact = 0, n_rej_image = 0;

    // exact cull + emit of the first min(qn, 32) queued pairs
    auto flush = [&]() {
        const uint32_t take = qn < 32u ? qn : 32u;
        const uint32_t e = s_q[w][lane < take ? lane : 0];
        const uint32_t pt = e & 31u, k = e >> 6;
        const V3 v = ld3(s_v[w][pt]), nrm = ld3(s_n[w][pt]);
        const float * view = views + UAVMVS_VIEW_FLOATS * (size_t)k;
        const float4 p4 = __ldg(reinterpret_cast<const float4 *>(view));
        V3 v2c = mk(p4.x, p4.y, p4.z) - v;
        float l = norm(v2c);
        V3 v2cn = v2c / l;
        float ctheta = dot(v2cn, nrm);
        bool pass;
        if (HIST) {                                   // kernels.cu:310-316 / :358-364: scale first, no image test
            const float scale = 1.0f - (l / max_distance);
            pass = lane < take && !(scale <= 0.0f) && !(ctheta < 0.087f);
        } else {
            pass = lane < take && !(ctheta < 0.087f) && !(l >= max_distance);
            if (dbg && lane < take) { ++n_queued; if (!pass) ++n_rej_exact; }
            const bool need = pass && !((e >> 5) & 1u);               // block not surely inside this view's image
            if (__any_sync(0xffffffffu, need)) {
                if (need) {
                    float vw[UAVMVS_VIEW_FLOATS];
                    vw[0] = p4.x; vw[1] = p4.y; vw[2] = p4.z; vw[3] = p4.w;
                    #pragma unroll
                    for (int i = 1; i < UAVMVS_VIEW_FLOATS / 4; ++i) {
                        float4 t = __ldg(reinterpret_cast<const float4 *>(view) + i);
                        vw[4 * i] = t.x; vw[4 * i + 1] = t.y; vw[4 * i + 2] = t.z; vw[4 * i + 3] = t.w;
                    }
                    pass = in_image(vw, v, width, height);
                    if (dbg && !pass) ++n_rej_image;
                }
            }
        }
        const uint32_t m = __ballot_sync(0xffffffffu, pass);
        if (pass) {
            const uint32_t idx = base + count + __popc(m & lt_mask);
            __stcs(rl.dirl + idx, make_float4(v2cn.x, v2cn.y, v2cn.z, l));      // rl.tdir aliases rl.dirl (nothing certified)
            if (HIST) rl.rview[idx] = k;
            __stcs(rl.torg + idx, make_float4(v.x, v.y, v.z, __uint_as_float(idx << 5 | pt)));
            if (rl.trd) {
                // k_trace's per-ray set-up (three IEEE reciprocals, the error band), done here where every lane holds a ray
                RayT rt; rt.ox = v.x; rt.oy = v.y; rt.oz = v.z; rt.tmax = l;
                rt.rdx = 1.0f / v2cn.x; rt.rdy = 1.0f / v2cn.y; rt.rdz = 1.0f / v2cn.z;
                ray_fma_setup80(rt, rl.bx, rl.by, rl.bz);
                const float mn = fminf(fminf(fabsf(v2cn.x), fabsf(v2cn.y)), fabsf(v2cn.z));
                const float mx = fmaxf(fmaxf(fabsf(v2cn.x), fabsf(v2cn.y)), fabsf(v2cn.z));
                const bool fast = mn > 1e-4f && mx < 2.0f && !rl.root_is_leaf && rt.em >= 0.0f;   // below: em would be uselessly wide
                __stcs(rl.trd + idx, make_float4(rt.rdx, rt.rdy, rt.rdz, fast ? rt.em : -1.0f));
            }
        }
        count += __popc(m);
        __syncwarp();
        const uint32_t rest = qn - take;
        const uint32_t carry = lane < rest ? s_q[w][32 + lane] : 0u;
        __syncwarp();
        if (lane < rest) s_q[w][lane] = carry;
        qn = rest;
        __syncwarp();
    };

    // Phase 1, transposed: candidate views of the block are collected 32 at a time (lane = view, its position -- and, with
    // the image proof, its projection rows -- in registers), and the 32 points are looped over from shared memory: 32
    // independent iterations of a dozen instructions instead of one latency-bound iteration per candidate view (a block
    // whose bounding sphere is huge -- a jump of the Z-order curve -- is a candidate for thousands of views; walked one view
    // at a time such a block alone set the duration of the kernel: 0.5 us per view, 5 ms for the 9978 views of one block of
    // a config-5 shard).  Lane p keeps the word "views of this group that may see point p"; the pairs are then queued in
    // (view, point) order exactly as before.
    const uint32_t nvalid = min(32u, n - blk * 32u);                       // points of this block
    uint32_t vqn = 0;
    auto process_views = [&](uint32_t take) {
        const uint32_t ve = s_vq[w][lane < take ? lane : 0];               // view index | block surely inside its image << 31
        const uint32_t k = ve & 0x7fffffffu;
        const float * view = views + UAVMVS_VIEW_FLOATS * (size_t)k;
        const float px = __ldg(view), py = __ldg(view + 1), pz = __ldg(view + 2);
        const bool proof = !HIST && rl.vproj != nullptr && !(ve >> 31);
        float4 m0 = make_float4(0.f, 0.f, 0.f, 0.f), m1 = m0, m2 = m0, dm = m0;
        if (!HIST && rl.vproj) {
            const float4 * vp = reinterpret_cast<const float4 *>(rl.vproj + 16 * (size_t)k);
            m0 = __ldg(vp); m1 = __ldg(vp + 1); m2 = __ldg(vp + 2); dm = __ldg(vp + 3);
        }
        uint32_t mybits = 0;
        #pragma unroll 4
        for (uint32_t p = 0; p < nvalid; ++p) {
            const float4 V = s_v[w][p], N = s_n[w][p];                     // broadcast; N.w = back-face limit of the point
            // fast-math proof of rejection (any doubt, any NaN -> queued for the exact test)
            const float dx = px - V.x, dy = py - V.y, dz = pz - V.z;
            const float l2 = __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, dx * dx));
            const float dn = __fmaf_rn(dz, N.z, __fmaf_rn(dy, N.y, dx * N.x));
            const float c = dn * rsqrtf(l2);
            bool reject = (c < N.w) || (l2 > md2);
            if (!HIST && rl.vproj) {
                // fast proof that the reference's image test (kernels.cu:165-167) rejects the pair: the point lies in front
                // of the camera (p2 surely positive) and projects outside the image by more than the margins allow
                const float P0 = __fmaf_rn(m0.z, V.z, __fmaf_rn(m0.y, V.y, __fmaf_rn(m0.x, V.x, m0.w)));
                const float P1 = __fmaf_rn(m1.z, V.z, __fmaf_rn(m1.y, V.y, __fmaf_rn(m1.x, V.x, m1.w)));
                const float P2 = __fmaf_rn(m2.z, V.z, __fmaf_rn(m2.y, V.y, __fmaf_rn(m2.x, V.x, m2.w)));
                const float lo2 = P2 - dm.z, hi2 = P2 + dm.z;              // p2 of the reference lies in [lo2, hi2]
                // x = fl(p0 / p2) - 0.5 < 0  <=  p0 < 0.4999995 p2 ;  x >= width  <=  p0 > (width + 0.5)(1 + 1e-6) p2   (p2 > 0)
                const bool out_x = (P0 + dm.x < 0.4999995f * lo2) || (P0 - dm.x > wph * hi2);
                const bool out_y = (P1 + dm.y < 0.4999995f * lo2) || (P1 - dm.y > hph * hi2);
                reject = reject || (proof && lo2 > 0.0f && (out_x || out_y));   // NaN anywhere: no proof
            }
            const uint32_t m = __ballot_sync(0xffffffffu, lane < take && !reject);
            if (lane == p) mybits = m;
        }
        if (dbg && lane == 0) n_cand += take * nvalid;
        if (DBG) dbg_views += take;
        // queue the surviving pairs in (view, point) order
        uint32_t anyv = __reduce_or_sync(0xffffffffu, mybits);
        while (anyv) {
            const uint32_t j = __ffs(anyv) - 1; anyv &= anyv - 1;
            const uint32_t vj = s_vq[w][j];
            const bool has = (mybits >> j) & 1u;
            const uint32_t m = __ballot_sync(0xffffffffu, has);
            if (has) s_q[w][qn + __popc(m & lt_mask)] = ((vj & 0x7fffffffu) << 6) | ((vj >> 31) << 5) | lane;
            qn += __popc(m);
            __syncwarp();
            if (qn >= 32u) flush();
        }
        // keep the views that did not fit
        __syncwarp();
        const uint32_t rest = vqn - take;
        const uint32_t carry = lane < rest ? s_vq[w][32 + lane] : 0u;
        __syncwarp();
        if (lane < rest) s_vq[w][lane] = carry;
        vqn = rest;
        __syncwarp();
    };
    // Blocks with few candidate views keep the direct form (lane = point, one candidate view per iteration: fewer
    // instructions when most views of a group queue something anyway); rl.transpose_at = candidate views from which a
    // block takes the transposed form (k_block_caps knows the count before this kernel runs).
    const bool transposed = (rl.caps[blk] >> 5) >= rl.transpose_at || (!HIST && rl.vproj != nullptr);
    const V3 pv = ld3(v4); const float bf_lim = n4.w;
    for (uint32_t g0 = lv0; g0 < lv1; g0 += 128) {
      uint32_t mw4[4], iw4[4];
      #pragma unroll
      for (int u = 0; u < 4; ++u) {
          const uint32_t vi = g0 + 32 * u + lane;
          mw4[u] = vi < lv1 ? __ldg(mrow + vi) : 0u;
          iw4[u] = (!HIST && vi < lv1) ? __ldg(irow + vi) : 0u;
      }
      #pragma unroll
      for (int u = 0; u < 4; ++u) {
        const uint32_t c0 = g0 + 32 * u;
        if (c0 >= lv1) break;
        const bool cand = (mw4[u] >> mbit) & 1u;
        uint32_t word = __ballot_sync(0xffffffffu, cand);
        if (!word) continue;
        if (transposed) {
            if (cand) s_vq[w][vqn + __popc(word & lt_mask)] = (c0 + lane) | (((iw4[u] >> mbit) & 1u) << 31);
            vqn += __popc(word);
            __syncwarp();
            if (vqn >= 32u) process_views(32u);
            continue;
        }
        const uint32_t iword = __ballot_sync(0xffffffffu, (iw4[u] >> mbit) & 1u);   // whole block surely inside the image
        while (word) {
            const uint32_t kb = __ffs(word) - 1;
            const uint32_t k = c0 + kb; word &= word - 1;
            const float * view = views + UAVMVS_VIEW_FLOATS * (size_t)k;
            const float px = __ldg(view), py = __ldg(view + 1), pz = __ldg(view + 2);
            // fast-math proof of rejection (any doubt, any NaN -> queued for the exact test)
            const float dx = px - pv.x, dy = py - pv.y, dz = pz - pv.z;
            const float l2 = __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, dx * dx));
            const float dn = __fmaf_rn(dz, n4.z, __fmaf_rn(dy, n4.y, dx * n4.x));
            const float c = dn * rsqrtf(l2);
            const bool maybe = valid && !((c < bf_lim) || (l2 > md2));
            if (DBG) { if (dbg && valid) ++n_cand; ++dbg_views; }
            const uint32_t m = __ballot_sync(0xffffffffu, maybe);
            if (maybe) s_q[w][qn + __popc(m & lt_mask)] = (k << 6) | (((iword >> kb) & 1u) << 5) | lane;
            qn += __popc(m);
            __syncwarp();
            if (qn >= 32u) flush();
        }
      }
    }
    if (vqn) process_views(vqn);
    while (qn) flush();
    if (npieces > 1) {
        uint32_t * pb = rl.early + K1_EARLY_ZEROED + K1_EARLY_MAX + eslot * K1_EARLY_PIECES, * pc = pb + K1_EARLY_MAX * K1_EARLY_PIECES;
        __threadfence();                                   // this piece's records before its count
        __syncwarp();
        uint32_t old = 0;
        if (lane == 0) { pb[piece] = base; pc[piece] = count; __threadfence(); old = atomicAdd(rl.early + 2 + eslot, 1u); }
        old = __shfl_sync(0xffffffffu, old, 0);
        if (old != npieces - 1) return;
        __threadfence();
        uint32_t total = 0;
        const uint32_t b0 = rl.off[blk];
        for (uint32_t p = 0; p < npieces; ++p) {
            const uint32_t src = *(volatile uint32_t *)(pb + p), cn = *(volatile uint32_t *)(pc + p), dst = b0 + total;
            if (src != dst) for (uint32_t i = 0; i < cn; i += 32) {       // dst < src: a step never overwrites what a later one reads
                const bool in = i + lane < cn;
                float4 d = make_float4(0.f, 0.f, 0.f, 0.f), o = d, t = d;
                if (in) { d = __ldcg(rl.dirl + src + i + lane); o = __ldcg(rl.torg + src + i + lane); if (rl.trd) t = __ldcg(rl.trd + src + i + lane); }
                __syncwarp();
                if (in) {
                    const uint32_t idx = dst + i + lane;
                    o.w = __uint_as_float(idx << 5 | (__float_as_uint(o.w) & 31u));
                    __stcs(rl.dirl + idx, d); __stcs(rl.torg + idx, o); if (rl.trd) __stcs(rl.trd + idx, t);
                }
                __syncwarp();
            }
            total += cn;
        }
        count = total;
    }
    if (lane == 0) {
        rl.cnt[blk] = count; rl.tcnt[blk] = count;
        if (counters && count) atomicAdd(counters + 3, (unsigned long long)count);
    }
    if (DBG && rl.dbg_blocks && lane == 0) {
        rl.dbg_blocks[3 * (size_t)blk] = (uint32_t)(clock64() - dbg_t0); rl.dbg_blocks[3 * (size_t)blk + 1] = count; rl.dbg_blocks[3 * (size_t)blk + 2] = dbg_views;
    }
    if (dbg) {
        for (int o = 16; o > 0; o >>= 1) {
            n_cand += __shfl_xor_sync(0xffffffffu, n_cand, o); n_queued += __shfl_xor_sync(0xffffffffu, n_queued, o);
            n_rej_exact += __shfl_xor_sync(0xffffffffu, n_rej_exact, o); n_rej_image += __shfl_xor_sync(0xffffffffu, n_rej_image, o);
        }
        if (lane == 0) {
            atomicAdd(counters + 4, (unsigned long long)n_cand); atomicAdd(counters + 5, (unsigned long long)n_queued);
            atomicAdd(counters + 6, (unsigned long long)n_rej_exact); atomicAdd(counters + 7, (unsigned long long)n_rej_image);
        }
    }
}

struct TraceArgs {
    BvhView bvh;
    const float4 * __restrict__ verts;
    RayList rl;
    uint32_t num_blocks;
    uint32_t * work;              // next block slice to hand out
    int refill, tri, steps;       // K1_REFILL / K1_TRI / K1_STEPS (env overrides for sweeps)
};

__device__ __forceinline__ void prefetch_l2(const void * p)
{
#if K1_PREFETCH == 2
    asm volatile("prefetch.global.L2 [%0];" :: "l"(p));
#elif K1_PREFETCH == 1
    asm volatile("prefetch.global.L1 [%0];" :: "l"(p));
#endif
}

#ifndef K1_RAY_PREFETCH
#define K1_RAY_PREFETCH 0      // measured: no change (3.80-3.85 ms either way)
#endif
#ifndef K1_MIN_BLOCKS
#define K1_MIN_BLOCKS 7        // measured (r02_b): the PRMT node test needs 70 registers; 8 CTAs (64) spill and gain nothing
#endif

// The leaf's own box (the reference's stored box; for nested trees verified at upload to be the tight bound of the leaf's
// triangles) tested with the reference's decision: slab_bool() proves accept / reject from t~ = (b - o) fl(1 / d) when the
// margin allows (nearly always) and runs the reference's own divisions otherwise.
static __device__ __noinline__ bool leaf_box_exact(const float4 * __restrict__ leafbox, uint32_t first,
    float ox, float oy, float oz, float dx, float dy, float dz, float rdx, float rdy, float rdz, float tmin, float tmax)
{
    const float4 b0 = __ldg(leafbox + 2 * (size_t)first), b1 = __ldg(leafbox + 2 * (size_t)first + 1);
    RayQ q; q.r.o = mk(ox, oy, oz); q.r.d = mk(dx, dy, dz); q.r.tmin = tmin; q.r.tmax = tmax;
    q.rd = mk(rdx, rdy, rdz); q.fast = true;             // callers only hold rays with 1e-4 < |d_k| < 2
    float t;
    return slab_bool(q, b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, t);
}

// Occlusion rays against the 4-wide BVH of a tree whose boxes are nested (every child box
// inside its parent's box, every leaf box the tight bound of its triangles: exact-union trees
// such as acc's SAH tree and our LBVH; verified at upload).  The slab test is monotone in the
// box bounds (each of fl(b - o), fl(x / d), min, max is), so a ray that passes a leaf's box
// passes every ancestor's box: the reference reaches a leaf iff the leaf's OWN box passes, and
// its answer is  OR over leaves (leaf box passes AND some triangle of the leaf is accepted).
// Inner boxes therefore only need a conservative test (never reject a box the exact test
// accepts).  Leaf boxes the filtered test cannot decide are deferred with a flag; the exact
// box test then runs only if a triangle of that leaf is accepted (both orders of the AND agree).
__global__ void __launch_bounds__(K1_WARPS * 32, K1_MIN_BLOCKS)
k_trace(TraceArgs a)
{
    K1_SPECULATION_GUARD(a.rl);
    __shared__ uint32_t s_pend[K1_PEND][K1_WARPS * 32];     // deferred leaves: first << 6 | unverified << 5 | count
    const uint32_t tid = threadIdx.x, lane = tid & 31;
    const uint32_t lt_mask = (1u << lane) - 1u;
    uint32_t stack[K1_STACK];

    uint32_t pos = 0, end = 0; bool drained = false;         // warp-uniform: current slice of the ray list
    bool has_ray = false, traversing = false, hit = false;
    uint32_t idx = 0, cur = 0, npend = 0, ptl = 0; int sp = 0;
    RayT rt; rt.ox = rt.oy = rt.oz = rt.rdx = rt.rdy = rt.rdz = rt.tmin = rt.tmax = 0.0f;
    rt.nox = rt.noy = rt.noz = rt.em = 0.0f;
    V3 rd = mk(0.f, 0.f, 0.f);
#if K1_PRMT
#if K1_BF16
    // PRMT selectors of the NEAR plane per axis: bytes {0, 0, lo, hi} of the min half (0x1044) or of the max half (0x3244) of a
    // (min, max) word = that bf16 offset as an fp32; the far plane's selector is the other one (^ 0x2200)
    uint32_t selx = 0x1044u, sely = 0x1044u, selz = 0x1044u;
#else
    uint32_t selx = 0x3210u, sely = 0x3210u, selz = 0x3210u;     // PRMT selectors: (min, max) -> (near, far) per axis
#endif
#endif
    uint32_t serviced_idle = 0;

    for (;;) {
        uint32_t idle = __ballot_sync(0xffffffffu, !has_ray);
        if (idle != serviced_idle && __popc(idle) >= a.refill) {
            // ---------------- refill idle lanes from the ray list ----------------
            const uint32_t need = __popc(idle), rank = __popc(idle & lt_mask);
            uint32_t given = 0; bool mine = false;
            while (given < need && !drained) {
                if (pos == end) {
                    uint32_t b = 0;
                    if (lane == 0) b = atomicAdd(a.work, 1u);
                    b = __shfl_sync(0xffffffffu, b, 0);
                    if (b >= a.num_blocks) { drained = true; break; }
                    pos = __ldg(a.rl.off + b); end = pos + a.rl.tcnt[b];
                    continue;
                }
                const uint32_t take = min(need - given, end - pos);
                if (!has_ray && rank >= given && rank < given + take) { idx = pos + (rank - given); mine = true; }
                pos += take; given += take;
            }
#if K1_RAY_PREFETCH
            // the records the next refills will hand out: in L2 by then (DRAM latency otherwise stalls the whole warp)
            if (pos + lane < end) {
                asm volatile("prefetch.global.L2 [%0];" :: "l"(a.rl.tdir + pos + lane));
                asm volatile("prefetch.global.L2 [%0];" :: "l"(a.rl.torg + pos + lane));
            }
#endif
            if (mine) {
                const uint32_t slot = idx;
                const float4 d4 = __ldcs(a.rl.tdir + idx);             // streamed once: keep L2 for the tree
                const float4 o4 = __ldcs(a.rl.torg + idx);
                const V3 o = ld3(o4);
                idx = __float_as_uint(o4.w) >> 5;                      // position in the ordered list
                ptl = (__float_as_uint(o4.w) & 31u) << 2;
                rd = ld3(d4);
                rt.ox = o.x; rt.oy = o.y; rt.oz = o.z;
                rt.tmin = d4.w * 0.001f; rt.tmax = d4.w;                 // kernels.cu:54-58
                bool fast;
                if (a.rl.trd) {                                           // constants set up by k_cull_emit_q
                    const float4 t4 = __ldcs(a.rl.trd + slot);
                    rt.rdx = t4.x; rt.rdy = t4.y; rt.rdz = t4.z; rt.em = t4.w;
                    rt.nox = -(rt.ox * rt.rdx); rt.noy = -(rt.oy * rt.rdy); rt.noz = -(rt.oz * rt.rdz);
                    fast = t4.w >= 0.0f;
                } else {
                    const float mn = fminf(fminf(fabsf(rd.x), fabsf(rd.y)), fabsf(rd.z));
                    const float mx = fmaxf(fmaxf(fabsf(rd.x), fabsf(rd.y)), fabsf(rd.z));
                    fast = mn > 1e-4f && mx < 2.0f && !a.bvh.root_is_leaf;     // below: em would be uselessly wide
                    if (fast) {
                        rt.rdx = 1.0f / rd.x; rt.rdy = 1.0f / rd.y; rt.rdz = 1.0f / rd.z;
                        ray_fma_setup80(rt, a.bvh.bx, a.bvh.by, a.bvh.bz);
                    }
                }
                if (fast) {
#if K1_PRMT
#if K1_BF16
                    selx = rt.rdx < 0.0f ? 0x3244u : 0x1044u; sely = rt.rdy < 0.0f ? 0x3244u : 0x1044u; selz = rt.rdz < 0.0f ? 0x3244u : 0x1044u;
#else
                    selx = rt.rdx < 0.0f ? 0x1032u : 0x3210u; sely = rt.rdy < 0.0f ? 0x1032u : 0x3210u; selz = rt.rdz < 0.0f ? 0x1032u : 0x3210u;
#endif
#endif
                    has_ray = true; hit = false; sp = 0; npend = 0; traversing = true; cur = 0;
                    if (a.rl.tpk) {                                      // start from the packet's entry points
                        const uint32_t * rec = a.rl.fr + ((size_t)__ldg(a.rl.tpk + slot) << 4);
                        const uint32_t ne = __ldg(rec);
                        for (uint32_t e = 1; e <= ne; ++e) stack[sp++] = __ldg(rec + e);
                        if (sp > 0) cur = stack[--sp]; else traversing = false;
                    }
                } else {
                    // axis-parallel (or degenerate) ray, single-leaf tree: the reference walk, right away
                    bool h = trace_any_exact(a.bvh, o.x, o.y, o.z, rd.x, rd.y, rd.z, rt.tmin, rt.tmax);
                    a.rl.vis[idx] = (uint8_t)(ptl | (h ? 1u : 2u));
                }
            }
            serviced_idle = __ballot_sync(0xffffffffu, !has_ray);
            if (serviced_idle == 0xffffffffu) {
                if (drained) break;
                serviced_idle = 0;        // only such rays were handed out: refill again
                continue;
            }
        }

        // ---------------- K1_STEPS traversal steps (4-wide nodes) ----------------
        #pragma unroll 1
        for (int step = 0; step < a.steps; ++step) {
            if (traversing && npend <= K1_PEND - 4) {
                const Node80 * np = a.bvh.wide80 + cur;
                const float4 c4 = __ldg(&np->c);
                const uint4 ha = __ldg(&np->h[0]), hb = __ldg(&np->h[1]), hc = __ldg(&np->h[2]);
                const uint4 cnt = __ldg(&np->cnt);
                const float crx = __fmaf_rn(c4.x, rt.rdx, rt.nox), cry = __fmaf_rn(c4.y, rt.rdy, rt.noy), crz = __fmaf_rn(c4.z, rt.rdz, rt.noz);
#if K1_PRMT
                // Node80 holds one word (min, max) per child and axis; one PRMT per word orders it (near, far) for this ray's
                // direction signs, so the near planes only feed the max chain and the far planes the min chain (no min/max
                // pairs).  The error band is folded into the far side: far constants and the ray's tmax are raised by em,
                // and "tmax~ + em < tmin~" (= g < -em) proves the rejection.
                const float cfx = crx + rt.em, cfy = cry + rt.em, cfz = crz + rt.em, tmaxe = rt.tmax + rt.em;
#if K1_BF16
                const uint32_t fsx = selx ^ 0x2200u, fsy = sely ^ 0x2200u, fsz = selz ^ 0x2200u;
                #define UAV_SLAB(wx, wy, wz, hit) { \
                    const float nx = __uint_as_float(__byte_perm((wx), 0u, selx)), fx = __uint_as_float(__byte_perm((wx), 0u, fsx)); \
                    const float ny = __uint_as_float(__byte_perm((wy), 0u, sely)), fy = __uint_as_float(__byte_perm((wy), 0u, fsy)); \
                    const float nz = __uint_as_float(__byte_perm((wz), 0u, selz)), fz = __uint_as_float(__byte_perm((wz), 0u, fsz)); \
                    const float tn = fmaxf(fmaxf(__fmaf_rn(nx, rt.rdx, crx), __fmaf_rn(ny, rt.rdy, cry)), fmaxf(__fmaf_rn(nz, rt.rdz, crz), rt.tmin)); \
                    const float tf = fminf(fminf(__fmaf_rn(fx, rt.rdx, cfx), __fmaf_rn(fy, rt.rdy, cfy)), fminf(__fmaf_rn(fz, rt.rdz, cfz), tmaxe)); \
                    hit = !(tf < tn); }
#else
                #define UAV_SLAB(wx, wy, wz, hit) { \
                    const uint32_t ux = __byte_perm((wx), 0u, selx), uy = __byte_perm((wy), 0u, sely), uz = __byte_perm((wz), 0u, selz); \
                    const float2 fx = __half22float2(*reinterpret_cast<const __half2 *>(&ux)); \
                    const float2 fy = __half22float2(*reinterpret_cast<const __half2 *>(&uy)); \
                    const float2 fz = __half22float2(*reinterpret_cast<const __half2 *>(&uz)); \
                    const float tn = fmaxf(fmaxf(__fmaf_rn(fx.x, rt.rdx, crx), __fmaf_rn(fy.x, rt.rdy, cry)), fmaxf(__fmaf_rn(fz.x, rt.rdz, crz), rt.tmin)); \
                    const float tf = fminf(fminf(__fmaf_rn(fx.y, rt.rdx, cfx), __fmaf_rn(fy.y, rt.rdy, cfy)), fminf(__fmaf_rn(fz.y, rt.rdz, cfz), tmaxe)); \
                    hit = !(tf < tn); }
#endif
                bool h0, h1, h2, h3;
                UAV_SLAB(ha.x, ha.y, ha.z, h0)
                UAV_SLAB(ha.w, hb.x, hb.y, h1)
                UAV_SLAB(hb.z, hb.w, hc.x, h2)
                UAV_SLAB(hc.y, hc.z, hc.w, h3)
                #undef UAV_SLAB
#else
                #define UAV_H2(w) __half22float2(*reinterpret_cast<const __half2 *>(&(w)))
                const float2 p0 = UAV_H2(ha.x), p1 = UAV_H2(ha.y), p2 = UAV_H2(ha.z), p3 = UAV_H2(ha.w);
                const float2 p4 = UAV_H2(hb.x), p5 = UAV_H2(hb.y), p6 = UAV_H2(hb.z), p7 = UAV_H2(hb.w);
                const float2 p8 = UAV_H2(hc.x), p9 = UAV_H2(hc.y), p10 = UAV_H2(hc.z), p11 = UAV_H2(hc.w);
                #undef UAV_H2
                const float g0 = slab_fma80(rt, crx, cry, crz, p0.x, p0.y, p1.x, p1.y, p2.x, p2.y);
                const float g1 = slab_fma80(rt, crx, cry, crz, p3.x, p3.y, p4.x, p4.y, p5.x, p5.y);
                const float g2 = slab_fma80(rt, crx, cry, crz, p6.x, p6.y, p7.x, p7.y, p8.x, p8.y);
                const float g3 = slab_fma80(rt, crx, cry, crz, p9.x, p9.y, p10.x, p10.y, p11.x, p11.y);
                const float m0 = rt.em, m1 = rt.em, m2 = rt.em, m3 = rt.em;
                // visit unless surely missed; leaf children are deferred (any-hit is order independent)
                // with the "box not verified" flag when the filtered test cannot prove the hit
                const bool h0 = !(g0 < -m0), h1 = !(g1 < -m1), h2 = !(g2 < -m2), h3 = !(g3 < -m3);
#endif
                const bool f0 = h0 && (int)cnt.x >= 0, f1 = h1 && (int)cnt.y >= 0, f2 = h2 && (int)cnt.z >= 0, f3 = h3 && (int)cnt.w >= 0;
                const bool i0 = h0 && (int)cnt.x < 0, i1 = h1 && (int)cnt.y < 0, i2 = h2 && (int)cnt.z < 0, i3 = h3 && (int)cnt.w < 0;
                const uint4 ref = make_uint4(cnt.x & 0x7fffffffu, cnt.y & 0x7fffffffu, cnt.z & 0x7fffffffu, cnt.w & 0x7fffffffu);
                if (f0) { s_pend[npend][tid] = cnt.x | 32u; ++npend; }
                if (f1) { s_pend[npend][tid] = cnt.y | 32u; ++npend; }
                if (f2) { s_pend[npend][tid] = cnt.z | 32u; ++npend; }
                if (f3) { s_pend[npend][tid] = cnt.w | 32u; ++npend; }
                // inner children: one is visited next, the others go on the stack (and are prefetched:
                // on a visible ray every pushed node is popped later)
                uint32_t nxt = i0 ? ref.x : UAV_NAI;
                if (i1 && nxt != UAV_NAI) { stack[sp++] = nxt; prefetch_l2(a.bvh.wide80 + nxt); }
                if (i1) nxt = ref.y;
                if (i2 && nxt != UAV_NAI) { stack[sp++] = nxt; prefetch_l2(a.bvh.wide80 + nxt); }
                if (i2) nxt = ref.z;
                if (i3 && nxt != UAV_NAI) { stack[sp++] = nxt; prefetch_l2(a.bvh.wide80 + nxt); }
                if (i3) nxt = ref.w;
                if (nxt == UAV_NAI) {
                    if (sp > 0) nxt = stack[--sp]; else traversing = false;
                }
                cur = nxt;
            }
        }
        // ---------------- triangle phase ----------------
        {
            const uint32_t pmask = __ballot_sync(0xffffffffu, npend > 0);
            if (pmask) {
                const uint32_t tmask = __ballot_sync(0xffffffffu, traversing);
                const uint32_t full = __ballot_sync(0xffffffffu, npend > K1_PEND - 4);
                if (__popc(pmask) >= a.tri || full || !tmask) {
                    if (npend > 0) {
                        Ray ray; ray.o = mk(rt.ox, rt.oy, rt.oz); ray.d = rd; ray.tmin = rt.tmin; ray.tmax = rt.tmax;
                        for (uint32_t p = 0; p < npend && !hit; ++p) {
                            const uint32_t r = s_pend[p][tid];
                            const uint32_t first = r >> 6, count = r & 31u;
                            float t; bool h = false;
                            for (uint32_t i = 0; i < count; ++i) h |= hit_tri(ray, a.bvh.tris + first + i, t);
                            if (h && (r & 32u))      // rare: the leaf's box still has to pass the exact test
                                h = leaf_box_exact(a.bvh.leafbox, first, rt.ox, rt.oy, rt.oz, rd.x, rd.y, rd.z, rt.rdx, rt.rdy, rt.rdz, rt.tmin, rt.tmax);
                            hit = h;
                        }
                        npend = 0;
                    }
                }
            }
        }
        // ---------------- finished lanes post their result ----------------
        if (has_ray && (hit || (!traversing && npend == 0))) {
            __stcs(a.rl.vis + idx, (uint8_t)(ptl | (hit ? 1u : 2u)));
            has_ray = false; traversing = false; npend = 0;
        }
    }
}

// Trees that are not nested (user-supplied boxes) or hold leaves of more than 31 triangles:
// every ray takes the reference walk with the exact test at every node.
__global__ void k_trace_exact(TraceArgs a, uint32_t total)
{
    const uint32_t b = blockIdx.x * K1_WARPS + (threadIdx.x >> 5);
    if (b >= a.num_blocks) return;
    const uint32_t base = a.rl.off[b], count = a.rl.tcnt[b];
    for (uint32_t i = threadIdx.x & 31; i < count; i += 32) {
        const float4 d4 = a.rl.tdir[base + i];
        const float4 o4 = a.rl.torg[base + i];
        const V3 o = ld3(o4);
        bool h = trace_any_exact(a.bvh, o.x, o.y, o.z, d4.x, d4.y, d4.z, d4.w * 0.001f, d4.w);
        a.rl.vis[__float_as_uint(o4.w) >> 5] = (uint8_t)(((__float_as_uint(o4.w) & 31u) << 2) | (h ? 1u : 2u));
    }
}

struct RetireArgs {
    const float4 * __restrict__ verts; const float4 * __restrict__ normals;
    uint32_t n, num_blocks;
    uint32_t block0 = 0;             // k_retire: first block of this launch (num_blocks = end of its range)
    RayList rl;
    float max_distance;
    float4 * obs; uint32_t * num_rows; uint32_t max_rows; uint32_t stride;   // stride in float4
    uint32_t * dirty;                // per 32-point block: some column received or lost a row (incremental K2/K3)
    unsigned long long * stats;
};

// Warp per block; the block's rays are walked in (view, point) order, 32 at a time.
// POPULATE: kernels.cu:171-181 (row = running counter of the point, stored iff row < max_rows).  The reference takes
//   every row with atomicAdd(num_rows, 1), which is what lets its optimiser run K1 for different views from several
//   host threads on the SAME observation array (optimize_trajectory.cu:470-495).  Here a first pass over the block's
//   visibility bytes counts the visible rays per point and ONE atomicAdd per point reserves that many rows; the second
//   pass fills them in view order.  Concurrent launches on one array therefore never hand out a row twice and never
//   lose an increment -- their rows interleave per (launch, block) instead of per ray, one of the orders the
//   reference's own concurrent launches can produce.
// !POPULATE: kernels.cu:183-200 (mark the first stored ray equal within 1e-5).
template <bool POPULATE>
__global__ void __launch_bounds__(K1_WARPS * 32)
k_retire(RetireArgs a)
{
    K1_SPECULATION_GUARD(a.rl);
    __shared__ float4 s_nrm[K1_WARPS][32], s_rx[K1_WARPS][32], s_ry[K1_WARPS][32];
    __shared__ uint32_t s_cnt[K1_WARPS][32];
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t blk = a.block0 + blockIdx.x * K1_WARPS + w;
    if (blk >= a.num_blocks) return;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const uint32_t id = blk * 32 + lane;
    const bool valid = id < a.n;
    const uint32_t count = a.rl.cnt[blk];
    if (count == 0) return;
    const uint32_t base = a.rl.off[blk];
    {
        V3 nrm = ld3(a.normals[valid ? id : a.n - 1]);
        V3 rx = normalized(orthogonal(nrm));              // relative_direction(), kernels.cu:63-76
        V3 ry = normalized(cross(nrm, rx));
        s_nrm[w][lane] = make_float4(nrm.x, nrm.y, nrm.z, 0.0f);
        s_rx[w][lane] = make_float4(rx.x, rx.y, rx.z, 0.0f);
        s_ry[w][lane] = make_float4(ry.x, ry.y, ry.z, 0.0f);
        s_cnt[w][lane] = POPULATE ? 0u : (valid ? a.num_rows[id] : 0u);
    }
    __syncwarp();
    if (POPULATE) {
        // pass 1: visible rays per point of this block (the status bytes stay in L1 / L2 for pass 2)
        for (uint32_t c0 = 0; c0 < count; c0 += 128) {
            uint32_t vb4[4];
            #pragma unroll
            for (int u = 0; u < 4; ++u) { const uint32_t i = c0 + 32 * u + lane; vb4[u] = i < count ? (uint32_t)__ldg(a.rl.vis + base + i) : 0u; }
            #pragma unroll
            for (int u = 0; u < 4; ++u) if ((vb4[u] & 3u) == 2u) atomicAdd(&s_cnt[w][vb4[u] >> 2], 1u);
        }
        __syncwarp();
        const uint32_t mine = s_cnt[w][lane];
        uint32_t first = 0;
        if (mine && valid) first = atomicAdd(a.num_rows + id, mine);     // rows [first, first + mine) are this launch's
        __syncwarp();
        s_cnt[w][lane] = first;
        __syncwarp();
    }
    unsigned long long st_vis = 0;
    bool touched = false;
    // the next 32 records are fetched while the current ones are processed: an iteration is a chain of two
    // DRAM round trips (status byte, then direction) otherwise, and the kernel is latency bound
    uint32_t vb_next = lane < count ? (uint32_t)__ldcs(a.rl.vis + base + lane) : 0u;
    float4 d4_next = lane < count ? __ldcs(a.rl.dirl + base + lane) : make_float4(0.f, 0.f, 0.f, 0.f);
    for (uint32_t c0 = 0; c0 < count; c0 += 32) {
        const uint32_t vb = vb_next;
        const float4 d4 = d4_next;
        {
            const uint32_t i_next = c0 + 32 + lane;
            vb_next = 0u;
            if (i_next < count) { vb_next = (uint32_t)__ldcs(a.rl.vis + base + i_next); d4_next = __ldcs(a.rl.dirl + base + i_next); }
        }
        const bool vis = (vb & 3u) == 2u;
        const uint32_t vis_mask = __ballot_sync(0xffffffffu, vis);
        if (!vis_mask) continue;
        touched = true;
        uint32_t pt = lane;
        V3 rel = mk(0.f, 0.f, 0.f); float scale = 0.f;
        if (vis) {
            pt = vb >> 2;
            const V3 v2cn = ld3(d4);
            rel = normalized(mk(dot(ld3(s_rx[w][pt]), v2cn), dot(ld3(s_ry[w][pt]), v2cn), dot(ld3(s_nrm[w][pt]), v2cn)));
            scale = 1.0f - (d4.w / a.max_distance);
            ++st_vis;
        }
        const uint32_t peers = __match_any_sync(0xffffffffu, pt) & vis_mask;
        const uint32_t row_pt = blk * 32 + pt;
        if (POPULATE) {
            if (vis) {
                uint32_t row = s_cnt[w][pt] + __popc(peers & lt_mask);
                if (row < a.max_rows) a.obs[(size_t)row * a.stride + row_pt] = make_float4(rel.x, rel.y, rel.z, scale);
            }
            __syncwarp();
            if (vis && (peers >> lane) == 1u) s_cnt[w][pt] += __popc(peers);   // highest peer updates
            __syncwarp();
        } else {
            // entries of the same point are applied in view order
            const uint32_t rank = __popc(peers & lt_mask);
            uint32_t max_rank = 0;
            for (uint32_t m = vis_mask; m; m &= m - 1)
                max_rank = max(max_rank, __shfl_sync(0xffffffffu, rank, __ffs(m) - 1));
            for (uint32_t r = 0; r <= max_rank; ++r) {
                if (vis && rank == r) {
                    const uint32_t rows = min(s_cnt[w][pt], a.max_rows);
                    volatile float4 * col = a.obs + row_pt;
                    for (uint32_t k = 0; k < rows; ++k) {
                        volatile float4 * p = col + (size_t)k * a.stride;
                        float ox = p->x, oy = p->y, oz = p->z, ow = p->w;
                        bool equal = fabsf(rel.x - ox) < 1e-5f && fabsf(rel.y - oy) < 1e-5f
                            && fabsf(rel.z - oz) < 1e-5f && fabsf(scale - ow) < 1e-5f;
                        if (equal) { p->w = -1.0f; break; }
                    }
                }
                __syncwarp();
            }
        }
    }
    if (touched && lane == 0 && a.dirty) a.dirty[blk] = 1u;
    if (a.stats) {
        for (int o = 16; o > 0; o >>= 1) st_vis += __shfl_xor_sync(0xffffffffu, st_vis, o);
        if (lane == 0) { atomicAdd(a.stats + 0, (unsigned long long)count); atomicAdd(a.stats + 1, st_vis); }
    }
}

// Histogram jobs: every visible ray adds its score to hist[view][nearest sphere bin].
// WITH_RECON = false: K4, kernels.cu:318-335.  true: K5, kernels.cu:366-389.
//
// The reference adds every contribution with a float atomicAdd (kernels.cu:335, :389), so its histogram depends on the
// arrival order in the last bits.  Here the sum is ORDER INDEPENDENT, hence bit-reproducible from run to run: a
// contribution c with |c| <= bound (a bound known before the launch: the largest quality for K4, 4 (|target| + 1)^2 for
// K5; fx = {scale, bound}, scale = 2^(36 - e) with bound <= 2^e) is rounded once to the fixed-point integer
// q = rn(c * scale) -- an absolute error <= bound 2^-37, five orders of magnitude below the float rounding of the sums
// it replaces -- and integers add associatively.  Rays of one block towards one candidate position mostly fall into the
// same bin (32 neighbouring points seen from tens of metres away), so the warp first adds the q of equal (position, bin)
// keys with match_any + redux (q split into two 18-bit halves: redux is a 32-bit instruction) and issues ONE 64-bit
// atomic per distinct key.  k_hist_finish turns the accumulators back into floats and adds them to the caller's array.
// Contributions outside the bound (NaN / inf inputs, recons that are not recons) keep the reference's float atomic.
template <bool WITH_RECON>
__global__ void __launch_bounds__(K1_WARPS * 32)
k_retire_hist(RetireArgs a, HistJob h)
{
    K1_SPECULATION_GUARD(a.rl);
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t blk = blockIdx.x * K1_WARPS + w;
    if (blk >= a.num_blocks) return;
    const uint32_t count = a.rl.cnt[blk];
    if (count == 0) return;
    const uint32_t base = a.rl.off[blk];
    KdView kd; kd.nodes = h.kd_nodes; kd.verts = h.kd_verts; kd.n = h.kd_n;
    const float fx_scale = __ldg(h.fx), fx_bound = __ldg(h.fx + 1);
    unsigned long long st_vis = 0;
    for (uint32_t c0 = 0; c0 < count; c0 += 32) {
        const uint32_t i = c0 + lane;
        const uint32_t vb = i < count ? (uint32_t)__ldcs(a.rl.vis + base + i) : 0u;
        bool act = (vb & 3u) == 2u;                       // visible (not occluded, not past the end)
        if (!__any_sync(0xffffffffu, act)) continue;
        float contrib = 0.0f;
        uint32_t key = 0xffffffffu - lane;                // lanes without a contribution: a key nobody shares
        if (act) {
            ++st_vis;
            const float4 d4 = __ldcs(a.rl.dirl + base + i);
            const uint32_t view = a.rl.rview[base + i];
            const uint32_t id = blk * 32 + (vb >> 2);
            const V3 v2cn = ld3(d4), n = ld3(a.normals[id]);
            const float l = d4.w;
            const float scale = 1.0f - (l / a.max_distance);
            const float ctheta = dot(v2cn, n);
            if (!WITH_RECON) {
                const float capture_difficulty = fmaxf(h.qualities[id], 0.0f);
                const float min_theta = fminf(h.values[id], 1.484f);
                const float scaling = (kPi / 2.0f) / ((kPi / 2.0f) - min_theta);
                const float theta = acosf(__saturatef(ctheta));
                const float rel_theta = fmaxf(theta - min_theta, 0.0f) * scaling;
                contrib = capture_difficulty * cosf(rel_theta) * scale;
            } else {
                const uint32_t rows = h.num_rows[id];
                if (rows >= h.max_rows) act = false;      // kernels.cu:373
                else {
                    const float recon = h.recons[id];
                    float new_recon = 0.0f;
                    if (rows >= 1) {
                        const V3 rel = relative_direction(v2cn, n);
                        const float m_k = h.heur.m_k, m_x0 = kPi / h.heur.m_x0, t_k = h.heur.t_k, t_x0 = kPi / h.heur.t_x0;
                        float sum = 0.0f;
                        for (uint32_t r = 0; r < rows; ++r) {     // heuristic(), kernels.cu:111-136
                            const float4 o = h.obs[(size_t)r * h.stride + id];
                            const float calpha = dot(rel, mk(o.x, o.y, o.z));
                            const float alpha = acosf(fmaxf(-1.0f, fminf(calpha, 1.0f)));
                            const float sc = fminf(o.w, scale);
                            const float ct = fminf(o.z, rel.z);
                            const float matchability = (1.0f - logistic(alpha, m_k, m_x0)) * ct;
                            const float triangulation = logistic(alpha, t_k, t_x0) * sc;
                            sum += matchability * triangulation;
                        }
                        new_recon = recon + sum;
                    }
                    contrib = func(new_recon, h.target_recon) - func(recon, h.target_recon);
                }
            }
            if (act) {
                const uint32_t bin = h.kd_fast.table ? kd_find_nn_fast(h.kd_fast, kd, -v2cn) : kd_find_nn(kd, -v2cn);
                key = view * h.kd_n + bin;
            }
        }
        const bool fin = act && fabsf(contrib) <= fx_bound;                   // NaN: false
        if (act && !fin) { atomicAdd(h.hist + key, contrib); key = 0xffffffffu - lane; }
        const long long q = fin ? __float2ll_rn(contrib * fx_scale) : 0ll;    // |q| <= 2^36; c * 2^k is exact
        const uint32_t peers = __match_any_sync(0xffffffffu, key);
        const int s_hi = __reduce_add_sync(peers, (int)(q >> 18));
        const uint32_t s_lo = __reduce_add_sync(peers, (uint32_t)(q & 0x3ffffll));
        if (fin && lane == (uint32_t)__ffs(peers) - 1u)
            atomicAdd(h.acc + key, (unsigned long long)((long long)s_hi * 262144ll + (long long)s_lo));
    }
    if (a.stats) {
        for (int o = 16; o > 0; o >>= 1) st_vis += __shfl_xor_sync(0xffffffffu, st_vis, o);
        if (lane == 0) { atomicAdd(a.stats + 0, (unsigned long long)count); atomicAdd(a.stats + 1, st_vis); }
    }
}

// ---- host side ---------------------------------------------------------------------------
struct K1Scratch {
    uint32_t * masks = nullptr; size_t masks_cap = 0;
    uint32_t * caps = nullptr, * off = nullptr, * cnt = nullptr; size_t blocks_cap = 0;
    void * cub_tmp = nullptr; size_t cub_cap = 0;
    float4 * dirl = nullptr, * tdir = nullptr, * torg = nullptr, * trd = nullptr; uint8_t * vis = nullptr; size_t rays_cap = 0, trd_cap = 0;
    uint32_t * tcnt = nullptr;
    uint32_t * rview = nullptr; size_t rview_cap = 0;
    uint32_t * tpk = nullptr; size_t tpk_cap = 0;
    uint32_t * fr = nullptr; size_t fr_cap = 0;
    uint32_t * cert = nullptr; size_t cert_cap = 0;
    uint32_t * work = nullptr; unsigned long long * total = nullptr, * h_total = nullptr;
    unsigned long long * counters = nullptr;      // diagnostics: shaft packets tested / certified, rays certified / traced
    float * vproj = nullptr; size_t vproj_cap = 0;
    uint32_t * dbg_blocks = nullptr; size_t dbg_blocks_cap = 0;
    uint32_t * vcaps = nullptr, * h_vcaps = nullptr; size_t vcaps_cap = 0, h_vcaps_cap = 0;   // per-view bounds (batch planning)
    uint32_t plan_views = 0, plan_blocks = 0;      // the last call with this shape needed a plan: plan at once
    int grid_trace = 0;
    int tune_refill = K1_REFILL, tune_tri = K1_TRI, tune_steps = K1_STEPS, tune_cull_warps = 1, tune_transpose_at = 512, tune_early_at = 256, tune_split_at = 1024;
    cudaEvent_t ev_total = nullptr;
    // per-kernel timing of the last launch_k1 (uavmvs_profile_k1): events at the kernel boundaries of every view batch
    std::vector<cudaEvent_t> ev_pool; size_t ev_used = 0;
    std::vector<int> ev_kind;                      // what starts at event i: 0 masks, 1 caps + scan, 2 cull, 3 trace, 4 retire, 5 end
};
static thread_local K1Scratch g_k1;
static std::atomic<int> g_k1_profile{0};

static int k1_mark(K1Scratch & k, cudaStream_t s, int kind)
{
    if (!g_k1_profile.load(std::memory_order_relaxed)) return 0;
    if (k.ev_used == k.ev_pool.size()) {
        cudaEvent_t e;
        UAV_CUDA(cudaEventCreate(&e));
        k.ev_pool.push_back(e); k.ev_kind.push_back(0);
    }
    k.ev_kind[k.ev_used] = kind;
    UAV_CUDA(cudaEventRecord(k.ev_pool[k.ev_used++], s));
    return 0;
}

template <typename T>
static int grow(T ** p, size_t * cap, size_t need, cudaStream_t s)
{
    if (need <= *cap) return 0;
    if (*p) UAV_CUDA(cudaFreeAsync(*p, s));        // stream-ordered: earlier kernels finish first
    UAV_CUDA(cudaMallocAsync(p, need * sizeof(T), s));
    *cap = need;
    return 0;
}

// Opt-in (UAVMVS_SHAFTS=1, read at every call).  Measured on config 2 (profiles/r01_c_shafts.txt): 23 % of the
// packets / 31 % of the rays are certified, but the shaft walk costs more than the tracing it saves
// (K1 8.1 ms with shafts vs 5.3 ms without), so the default pipeline does not run it.
static bool shafts_enabled()
{
    const char * e = getenv("UAVMVS_SHAFTS");
    return e && e[0] == '1';
}

// Packet entry points (k_entry): opt-in (UAVMVS_ENTRY=1).  Measured on config 2 (profiles/r01_c_shafts.txt):
// k_trace 3.4 -> 2.9 ms, but k_entry itself costs 0.8 ms (tau 32) to 1.1 ms (tau 8): a net loss.
static bool entry_enabled() { const char * e = getenv("UAVMVS_ENTRY"); return e && e[0] == '1'; }
static float entry_tau() { const char * e = getenv("UAVMVS_ENTRY_TAU"); return e ? (float)atof(e) : 32.0f; }

static size_t ray_budget()
{
    // upper bound of ray records (24 B + 1 B each) kept at a time; views are processed in
    // batches when a whole trajectory would need more
    const char * e = getenv("UAVMVS_RAY_BUDGET");             // read at every call (tests shrink it)
    size_t v = e ? (size_t)strtoull(e, nullptr, 10) : ((size_t)1 << 27);
    if (v < 1024) v = 1024;
    if (v > ((size_t)1 << 27)) v = (size_t)1 << 27;          // ray index << 5 | lane must fit 32 bits
    return v;
}

int launch_k1(cudaStream_t s, int populate, const float * d_views, uint32_t num_views,
    float max_distance, int width, int height, const uavmvs_bvh * bvh, const uavmvs_cloud * cloud,
    uavmvs_obs * obs, unsigned long long * d_stats, const HistJob * hj, const K1AfterRetire * after)
{
    UAV_REQUIRE(bvh && cloud && (obs || hj), "null handle");
    UAV_REQUIRE(hj || obs->num_cols == cloud->n, "obs_rays columns != cloud vertices");
    if (hj) { width = 0; height = 0; }                   // bare positions: no image test anywhere
    UAV_REQUIRE(bvh->num_tris < (1u << 25), "more than 2^25 triangles");
    if (num_views == 0) return (populate && after) ? after->fn(after->ctx, s, 0, cloud->num_blocks) : 0;
    K1Scratch & k = g_k1;
    const uint32_t nb = cloud->num_blocks;
    if (!k.work) {
        UAV_CUDA(cudaMalloc(&k.work, sizeof(uint32_t) * K1_EARLY_WORDS));   // word 0: k_trace's counter, then the early list
        UAV_CUDA(cudaMalloc(&k.total, sizeof(unsigned long long)));
        UAV_CUDA(cudaMalloc(&k.counters, 8 * sizeof(unsigned long long)));
        UAV_CUDA(cudaMemset(k.counters, 0, 8 * sizeof(unsigned long long)));
        UAV_CUDA(cudaMallocHost(&k.h_total, sizeof(unsigned long long)));
        int dev = 0, sms = 0, occ_a = 0, occ_b = 0;
        UAV_CUDA(cudaGetDevice(&dev));
        UAV_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        UAV_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_a, k_trace, K1_WARPS * 32, 0));
        occ_b = occ_a;
        int occ = occ_a < occ_b ? occ_a : occ_b;
        { const char * e = getenv("UAVMVS_TRACE_CTAS"); if (e && atoi(e) > 0 && atoi(e) < occ) occ = atoi(e); }   // dev: leave room for co-running kernels
        k.grid_trace = sms * (occ > 0 ? occ : 1);
        // dev sweeps: thresholds of the traversal kernel's bookkeeping
        if (const char * e = getenv("UAVMVS_K1_REFILL")) if (atoi(e) >= 1 && atoi(e) <= 32) k.tune_refill = atoi(e);
        if (const char * e = getenv("UAVMVS_K1_TRI")) if (atoi(e) >= 1 && atoi(e) <= 32) k.tune_tri = atoi(e);
        if (const char * e = getenv("UAVMVS_CULL_WARPS")) if (atoi(e) == 1 || atoi(e) == 2 || atoi(e) == 4) k.tune_cull_warps = atoi(e);
        if (const char * e = getenv("UAVMVS_CULL_TRANSPOSE_AT")) k.tune_transpose_at = atoi(e) >= 0 ? atoi(e) : 256;
        if (const char * e = getenv("UAVMVS_K1_STEPS")) if (atoi(e) >= 1 && atoi(e) <= 64) k.tune_steps = atoi(e);
    }
    // read per call (tests lower them to push small scenes through the piece path)
    { const char * e = getenv("UAVMVS_CULL_EARLY_AT"); k.tune_early_at = e ? atoi(e) : 256; }                     // 0: no early list
    { const char * e = getenv("UAVMVS_CULL_SPLIT_AT"); k.tune_split_at = (e && atoi(e) > 0) ? atoi(e) : 1024; }
    if (grow(&k.masks, &k.masks_cap, 2 * (size_t)cloud->num_super * num_views, s)) return 1;   // candidate bits, then "surely inside the image" bits
    uint32_t * inside_masks = k.masks + (size_t)cloud->num_super * num_views;
    if (nb + 1 > k.blocks_cap) {
        size_t c0 = k.blocks_cap, c1 = k.blocks_cap, c2 = k.blocks_cap, c3 = k.blocks_cap;
        if (grow(&k.caps, &c0, (size_t)nb + 1, s) || grow(&k.off, &c1, (size_t)nb + 1, s) || grow(&k.cnt, &c2, (size_t)nb + 1, s)
            || grow(&k.tcnt, &c3, (size_t)nb + 1, s)) return 1;
        k.blocks_cap = nb + 1;
    }
    size_t tmp_need = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, tmp_need, k.caps, k.off, (int)nb, s);
    if (tmp_need > k.cub_cap) {
        if (k.cub_tmp) UAV_CUDA(cudaFreeAsync(k.cub_tmp, s));
        UAV_CUDA(cudaMallocAsync(&k.cub_tmp, tmp_need, s));
        k.cub_cap = tmp_need;
    }

    k.ev_used = 0;
    if (k1_mark(k, s, 0)) return 1;
    // Opt-in (UAVMVS_IMAGE_PROOF=1).  Measured on config 2 (profiles/r02_d_image_proof.txt): 16.1 M of the 39.0 M pairs queued
    // for the exact cull fail only the image test, and the proof removes them -- but it lengthens the latency-bound
    // per-view loop of phase 1 (four uniform loads + a 9-FMA chain per candidate): k_cull_emit_q 0.79 -> 1.08 ms.
    const bool img_proof = !hj && getenv("UAVMVS_IMAGE_PROOF");
    if (img_proof) {
        if (grow(&k.vproj, &k.vproj_cap, 16 * (size_t)num_views, s)) return 1;
        k_view_proj<<<divup(num_views, 128), 128, 0, uav::counted(s)>>>(d_views, num_views, cloud->bound_abs[0], cloud->bound_abs[1],
            cloud->bound_abs[2], k.vproj);
    }
    dim3 mgrid(divup(num_views, 32), divup(cloud->num_super, K1_WARPS * K1_SB_PER_WARP));
    // Histogram jobs test `1 - l / max_distance <= 0` (kernels.cu:310-311): with max_distance <= 0 (the planner's
    // swapped arguments can pass a negative average there, plan_trajectory.cu:270-274) nothing is out of range, so the
    // conservative range culls are switched off and the exact test decides alone.
    const float cull_distance = (hj && !(max_distance > 0.0f)) ? INFINITY : max_distance;
    k_view_masks<<<mgrid, K1_WARPS * 32, 0, uav::counted(s)>>>(d_views, num_views, cloud->super_sphere, cloud->num_super,
        cloud->block_sphere, cloud->block_cone, cloud->block_nmax, nb, cull_distance, width, height, k.masks, inside_masks);
    UAV_CUDA(cudaGetLastError());

    // view batches: [v0, v1) such that the upper bound of ray records fits the budget
    // Speculation: the bound of a batch's ray records is needed on the host only to size the buffers.  When buffers
    // from an earlier call exist, the whole batch is enqueued at once against their capacity - every kernel that
    // touches them first checks the bound on the device - and the host waits only for the 8-byte copy of the bound,
    // which arrives while the cull and the traversal are already running (the stream no longer drains between
    // k_sum_caps and k_cull_emit: ~0.05 ms per evaluation of config 2).  A bound above the capacity means the
    // guarded kernels did nothing; the batch is then redone the synchronous way.
    bool speculate = !getenv("UAVMVS_NO_SPECULATION") && k.rays_cap > 0 && bvh->nested && !bvh->root_is_leaf
        && !entry_enabled() && !shafts_enabled() && !getenv("UAVMVS_CULL_PACKETS");
    if (speculate && !k.ev_total) UAV_CUDA(cudaEventCreateWithFlags(&k.ev_total, cudaEventDisableTiming));
    // A trajectory whose bound exceeds the budget is cut into batches of consecutive views from the per-view bounds
    // (k_view_caps): one synchronisation for the whole call, every batch then enqueued without waiting for the device.
    bool planned = false;
    auto make_plan = [&]() -> int {
        if (grow(&k.vcaps, &k.vcaps_cap, (size_t)num_views, s)) return 1;
        if (num_views > k.h_vcaps_cap) {
            if (k.h_vcaps) cudaFreeHost(k.h_vcaps);
            UAV_CUDA(cudaMallocHost(&k.h_vcaps, sizeof(uint32_t) * (size_t)num_views));
            k.h_vcaps_cap = num_views;
        }
        UAV_CUDA(cudaMemsetAsync(k.vcaps, 0, sizeof(uint32_t) * (size_t)num_views, s));
        k_view_caps<<<dim3(divup(num_views, 128), divup(cloud->num_super, 64)), 128, 0, uav::counted(s)>>>(k.masks, num_views, cloud->num_super, k.vcaps);
        UAV_CUDA(cudaMemcpyAsync(k.h_vcaps, k.vcaps, sizeof(uint32_t) * (size_t)num_views, cudaMemcpyDeviceToHost, s));
        UAV_CUDA(cudaStreamSynchronize(s));
        planned = true; k.plan_views = num_views; k.plan_blocks = nb;
        return 0;
    };
    // histogram jobs of the device-resident optimiser: buffers for the worst case, nothing read back
    const unsigned long long worst = 32ull * nb * num_views;
    const bool no_readback = hj && hj->no_readback && worst <= min((unsigned long long)ray_budget(), 1ull << 27);
    if (no_readback) speculate = false;
    else if (num_views > 1 && k.plan_views == num_views && k.plan_blocks == nb) { speculate = false; if (make_plan()) return 1; }
    bool after_done = false;
    uint32_t v0 = 0;
    while (v0 < num_views) {
        uint32_t v1 = num_views;
        unsigned long long total = 0;
        if (no_readback) total = worst;
        if (planned) {
            v1 = v0;
            while (v1 < num_views && (v1 == v0 || total + k.h_vcaps[v1] <= ray_budget())) total += k.h_vcaps[v1++];
            UAV_REQUIRE(total <= (1ull << 27), "a single view yields more than 2^27 candidate pairs");
        }
        if (k1_mark(k, s, 1)) return 1;
        // early list of the cull kernel (CW == 1, evaluation jobs): zeroed together with k_trace's work counter
        const bool early = !hj && k.tune_cull_warps == 1 && k.tune_early_at > 0;
        UAV_CUDA(cudaMemsetAsync(k.work, 0, sizeof(uint32_t) * K1_EARLY_ZEROED, s));
        k_block_caps<<<divup((size_t)cloud->num_super * 32, 128), 128, 0, uav::counted(s)>>>(k.masks, num_views, v0, v1, nb, k.caps,
            (uint32_t)k.tune_early_at, early ? k.work : nullptr);
        const bool spec = speculate && !planned;
        speculate = false;                                   // one attempt per call
        const unsigned long long spec_cap = min((unsigned long long)k.rays_cap, min((unsigned long long)ray_budget(), 1ull << 27));
        if (no_readback) {
        } else if (planned) {
            if (d_stats && total) k_add_val<<<1, 1, 0, uav::counted(s)>>>(d_stats + 3, total);   // pairs surviving the block-level cull
        } else {
            UAV_CUDA(cudaMemsetAsync(k.total, 0, sizeof(unsigned long long), s));
            k_sum_caps<<<min(divup(nb, 256), 1024u), 256, 0, uav::counted(s)>>>(k.caps, nb, k.total);
            UAV_CUDA(cudaMemcpyAsync(k.h_total, k.total, sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
            if (spec) {
                UAV_CUDA(cudaEventRecord(k.ev_total, s));
                total = spec_cap;                            // sizes nothing: the buffers already hold this many records
            } else {
                UAV_CUDA(cudaStreamSynchronize(s));
                total = *k.h_total;
                if (total > ray_budget() && v1 - v0 > 1) {   // too many candidates for one batch: plan the batches
                    if (make_plan()) return 1;
                    continue;
                }
                UAV_REQUIRE(total <= (1ull << 27), "a single view yields more than 2^27 candidate pairs");
                if (d_stats) k_add_u64<<<1, 1, 0, uav::counted(s)>>>(d_stats + 3, k.total);   // pairs surviving the block-level cull
            }
        }
        if (total > 0) {
            if (total > k.rays_cap) {
                size_t c0 = k.rays_cap, c1 = k.rays_cap, c2 = k.rays_cap, c3 = k.rays_cap;
                if (grow(&k.dirl, &c0, (size_t)total, s) || grow(&k.tdir, &c1, (size_t)total, s) || grow(&k.torg, &c3, (size_t)total, s)
                    || grow(&k.vis, &c2, (size_t)total, s)) return 1;
                k.rays_cap = (size_t)total;
            }
            cub::DeviceScan::ExclusiveSum(k.cub_tmp, tmp_need, k.caps, k.off, (int)nb, uav::counted(s, 2));   // cub: DeviceScanInitKernel + DeviceScanKernel
            if (hj && grow(&k.rview, &k.rview_cap, k.rays_cap, s)) return 1;
            const bool entry = entry_enabled() && bvh->nested && !bvh->root_is_leaf;
            if (entry) {
                if (grow(&k.tpk, &k.tpk_cap, k.rays_cap, s) || grow(&k.fr, &k.fr_cap, (size_t)(total / 32 + 1) * 16, s)) return 1;
                EntryArgs ea; ea.wide = bvh->wide; ea.lo = cloud->block_lo; ea.hi = cloud->block_hi; ea.num_blocks = nb;
                ea.views = d_views; ea.num_views = num_views; ea.v0 = v0; ea.v1 = v1; ea.masks = k.masks; ea.off = k.off;
                ea.fr = k.fr; ea.tau = entry_tau(); ea.counters = k.counters;
                k_entry<<<divup(nb, K1_WARPS), K1_WARPS * 32, 0, uav::counted(s)>>>(ea);
            }
            RayList rl; rl.g_total = spec ? k.total : nullptr; rl.g_cap = spec_cap; rl.tpk = entry ? k.tpk : nullptr; rl.fr = k.fr; rl.rview = k.rview; rl.dirl = k.dirl; rl.tdir = k.tdir; rl.torg = k.torg; rl.vis = k.vis; rl.off = k.off; rl.cnt = k.cnt; rl.tcnt = k.tcnt;
            // shafts: whole (block, view) packets proven visible without tracing (nested trees only)
            const uint32_t nchunks = divup(v1 - v0, 32);
            const bool shafts = shafts_enabled() && bvh->nested && !bvh->root_is_leaf;
            if (shafts) {
                if (grow(&k.cert, &k.cert_cap, (size_t)nb * nchunks, s)) return 1;
                UAV_CUDA(cudaMemsetAsync(k.cert, 0, sizeof(uint32_t) * (size_t)nb * nchunks, s));
                ShaftArgs sa; sa.wide = bvh->wide; sa.tris = bvh->ptris; sa.lo = cloud->block_lo; sa.hi = cloud->block_hi;
                sa.obb = cloud->block_obb; sa.num_blocks = nb;
                sa.views = d_views; sa.num_views = num_views; sa.v0 = v0; sa.v1 = v1; sa.masks = k.masks;
                sa.cert = k.cert; sa.nchunks = nchunks; sa.counters = k.counters;
                { const char * e = getenv("UAVMVS_SHAFT_STEPS"); sa.max_steps = e ? (uint32_t)atoi(e) : 1000000u; }
                k_shaft<<<divup(nb, K1_WARPS), K1_WARPS * 32, 0, uav::counted(s)>>>(sa);
            }
            const bool queued = !shafts && !entry && !getenv("UAVMVS_CULL_PACKETS");
            if (k1_mark(k, s, 2)) return 1;
            if (queued) rl.tdir = rl.dirl;            // one record per ray: every ray is traced, in list order
            // measured on config 2 (profiles/r02_b_sweep.txt): the set-up in the cull kernel costs it 0.05 ms and gains k_trace
            // nothing (the refill is not where its issue slots go), so it stays opt-in
            const bool cull_setup = queued && getenv("UAVMVS_CULL_SETUP");
            if (cull_setup && grow(&k.trd, &k.trd_cap, k.rays_cap, s)) return 1;
            rl.trd = cull_setup ? k.trd : nullptr;
            rl.debug = getenv("UAVMVS_K1_DEBUG") ? 1u : 0u;
            rl.reverse = getenv("UAVMVS_CULL_REVERSE") ? 1u : 0u;
            // histogram jobs (thousands of candidate positions per block) are faster in the direct form throughout
            // (config 4, 32 k positions: 8.7 us per position against 9.4 with the transposed form from 512 views)
            rl.caps = k.caps; rl.transpose_at = (hj && !getenv("UAVMVS_CULL_TRANSPOSE_AT")) ? 0x7fffffffu : (uint32_t)k.tune_transpose_at;
            rl.dbg_blocks = nullptr;
            if (rl.debug) { if (grow(&k.dbg_blocks, &k.dbg_blocks_cap, 3 * (size_t)nb, s)) return 1; rl.dbg_blocks = k.dbg_blocks; }
            rl.vproj = img_proof ? k.vproj : nullptr;
            rl.bx = bvh->bound_abs[0]; rl.by = bvh->bound_abs[1]; rl.bz = bvh->bound_abs[2]; rl.root_is_leaf = bvh->root_is_leaf;
            if (queued) {
                // CTA size of the cull kernel: a CTA holds its SM slot until its slowest warp (= 32-point block) is done, and
                // the work per block varies by orders of magnitude, so small CTAs pack better (UAVMVS_CULL_WARPS: 1, 2 or 4).
                // Measured on config 2 (profiles/r02_k_sweep.txt): 4 warps 0.71 ms, 1 warp 0.68; transposed phase 1 from 512
                // candidate views 0.65, from 128 1.00, never 0.83 (the slowest block then sets the duration of the kernel);
                // the two forms as separate kernels 0.72 - 0.86 (r02_l: the heavy blocks' kernel is one long tail)
                // r02_q: phase 2 as two queued stages (exact image test first, then back-face / range + emit on the 59 % that
                // survive it): 0.616 -> 0.653 ms, the second queue costs more than the arithmetic it saves; early list from
                // 128 candidate views instead of 256: 0.685
                const int cw = k.tune_cull_warps;
                rl.early = (early && !(rl.debug && cw != 1)) ? k.work : nullptr;
                rl.early_at = (uint32_t)k.tune_early_at; rl.split_at = (uint32_t)k.tune_split_at;
                const uint32_t eg = rl.early ? K1_EARLY_MAX * K1_EARLY_PIECES : 0u;
                #define UAV_CULL(H, D, W) k_cull_emit_q<H, D, W><<<divup(nb, W) + (W == 1 ? eg : 0u), W * 32, 0, uav::counted(s)>>>(cloud->vertices, cloud->normals, \
                    cloud->n, nb, d_views, num_views, v0, v1, k.masks, inside_masks, max_distance, cull_distance, width, height, rl, k.counters)
                if (rl.debug && !hj) { if (cw == 1) UAV_CULL(false, true, 1); else UAV_CULL(false, true, 4); }
                else if (hj) { if (cw == 1) UAV_CULL(true, false, 1); else if (cw == 2) UAV_CULL(true, false, 2); else UAV_CULL(true, false, 4); }
                else { if (cw == 1) UAV_CULL(false, false, 1); else if (cw == 2) UAV_CULL(false, false, 2); else UAV_CULL(false, false, 4); }
                #undef UAV_CULL
            }
            else if (hj) k_cull_emit<true><<<divup(nb, K1_WARPS), K1_WARPS * 32, 0, uav::counted(s)>>>(cloud->vertices, cloud->normals, cloud->n, nb,
                d_views, num_views, v0, v1, k.masks, inside_masks, shafts ? k.cert : nullptr, nchunks, max_distance, width, height, rl, k.counters);
            else k_cull_emit<false><<<divup(nb, K1_WARPS), K1_WARPS * 32, 0, uav::counted(s)>>>(cloud->vertices, cloud->normals, cloud->n, nb,
                d_views, num_views, v0, v1, k.masks, inside_masks, shafts ? k.cert : nullptr, nchunks, max_distance, width, height, rl, k.counters);
            if (k1_mark(k, s, 3)) return 1;
            TraceArgs ta; ta.bvh = view_of(bvh); ta.verts = cloud->vertices; ta.rl = rl; ta.num_blocks = nb; ta.work = k.work;
            ta.refill = k.tune_refill; ta.tri = k.tune_tri; ta.steps = k.tune_steps;
            unsigned grid = (unsigned)k.grid_trace;
            if (bvh->nested) k_trace<<<grid, K1_WARPS * 32, 0, uav::counted(s)>>>(ta);
            else k_trace_exact<<<divup(nb, K1_WARPS), K1_WARPS * 32, 0, uav::counted(s)>>>(ta, (uint32_t)total);
            if (k1_mark(k, s, 4)) return 1;
            RetireArgs ra; ra.verts = cloud->vertices; ra.normals = cloud->normals; ra.n = cloud->n; ra.num_blocks = nb;
            ra.rl = rl; ra.max_distance = max_distance;
            ra.stats = d_stats;
            if (hj) {
                ra.obs = nullptr; ra.num_rows = nullptr; ra.max_rows = 0; ra.stride = 0; ra.dirty = nullptr;
                if (hj->with_recon) k_retire_hist<true><<<divup(nb, K1_WARPS), K1_WARPS * 32, 0, uav::counted(s)>>>(ra, *hj);
                else k_retire_hist<false><<<divup(nb, K1_WARPS), K1_WARPS * 32, 0, uav::counted(s)>>>(ra, *hj);
                UAV_CUDA(cudaGetLastError());
            } else {
                ra.obs = obs->data; ra.num_rows = obs->num_rows; ra.max_rows = obs->max_rows;
                ra.stride = (uint32_t)(obs->pitch / sizeof(float4));
                ra.dirty = obs->dirty;
                if (populate && after && after->chunks > 1 && v1 == num_views && nb >= 64 * after->chunks) {
                    // last batch: every column is complete once its block has retired -> hand the ranges on one by one
                    for (uint32_t c = 0; c < after->chunks; ++c) {
                        const uint32_t b0 = (uint32_t)((uint64_t)nb * c / after->chunks), b1 = (uint32_t)((uint64_t)nb * (c + 1) / after->chunks);
                        ra.block0 = b0; ra.num_blocks = b1;
                        k_retire<true><<<divup(b1 - b0, K1_WARPS), K1_WARPS * 32, 0, uav::counted(s)>>>(ra);
                        UAV_CUDA(cudaGetLastError());
                        UAV_CUDA(cudaEventRecord(after->ev[c], s));
                        UAV_CUDA(cudaStreamWaitEvent(after->side, after->ev[c], 0));
                        if (int rc = after->fn(after->ctx, after->side, b0, b1 - b0)) return rc;
                    }
                    after_done = true;
                    UAV_CUDA(cudaEventRecord(after->ev[after->chunks], after->side));
                    UAV_CUDA(cudaStreamWaitEvent(s, after->ev[after->chunks], 0));
                } else {
                    if (populate && after && v1 == num_views) {          // too small to split: the continuation follows on the same stream
                        k_retire<true><<<divup(nb, K1_WARPS), K1_WARPS * 32, 0, uav::counted(s)>>>(ra);
                        UAV_CUDA(cudaGetLastError());
                        if (int rc = after->fn(after->ctx, s, 0, nb)) return rc;
                        after_done = true;
                    }
                    else if (populate) k_retire<true><<<divup(nb, K1_WARPS), K1_WARPS * 32, 0, uav::counted(s)>>>(ra);
                    else k_retire<false><<<divup(nb, K1_WARPS), K1_WARPS * 32, 0, uav::counted(s)>>>(ra);
                    UAV_CUDA(cudaGetLastError());
                }
            }
        }
        if (k1_mark(k, s, 5)) return 1;
        if (spec) {
            UAV_CUDA(cudaEventSynchronize(k.ev_total));      // the bound only; the batch keeps running
            if (*k.h_total > spec_cap) {                     // guarded kernels were no-ops: same views again, sized properly
                if (*k.h_total > ray_budget() && num_views > 1 && make_plan()) return 1;
                continue;
            }
            if (d_stats) k_add_u64<<<1, 1, 0, uav::counted(s)>>>(d_stats + 3, k.total);
        }
        v0 = v1;
    }
    // the continuation runs exactly when the columns are final, also when the last batch held no candidate pair
    if (populate && after && !after_done) { if (int rc = after->fn(after->ctx, s, 0, nb)) return rc; }
    return 0;
}

int k1_profile_enable(int on) { g_k1_profile.store(on ? 1 : 0); return 0; }

// ms[0] = k_view_masks, [1] = k_block_caps + k_sum_caps + scan, [2] = k_cull_emit*, [3] = k_trace*, [4] = k_retire*,
// summed over the view batches of the calling thread's last launch_k1.
int k1_profile(float ms[5])
{
    K1Scratch & k = g_k1;
    for (int i = 0; i < 5; ++i) ms[i] = 0.0f;
    UAV_REQUIRE(k.ev_used >= 2, "no profiled K1 launch on this thread (uavmvs_profile_k1_enable)");
    UAV_CUDA(cudaEventSynchronize(k.ev_pool[k.ev_used - 1]));
    for (size_t i = 0; i + 1 < k.ev_used; ++i) {
        const int kind = k.ev_kind[i];
        if (kind >= 5) continue;                             // "end" of a batch: the gap to the next batch is host time
        float t = 0.0f;
        UAV_CUDA(cudaEventElapsedTime(&t, k.ev_pool[i], k.ev_pool[i + 1]));
        ms[kind] += t;
    }
    return 0;
}

int k1_counters(unsigned long long out[4], int reset)
{
    K1Scratch & k = g_k1;
    for (int i = 0; i < 4; ++i) out[i] = 0;
    if (!k.counters) return 0;
    UAV_CUDA(cudaMemcpy(out, k.counters, 4 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    if (getenv("UAVMVS_K1_DEBUG")) {
        unsigned long long d[8];
        UAV_CUDA(cudaMemcpy(d, k.counters, 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
        if (k.dbg_blocks && k.dbg_blocks_cap) {
            const size_t nbk = k.dbg_blocks_cap / 3;
            std::vector<uint32_t> h(3 * nbk);
            UAV_CUDA(cudaMemcpy(h.data(), k.dbg_blocks, sizeof(uint32_t) * h.size(), cudaMemcpyDeviceToHost));
            std::vector<uint32_t> order(nbk);
            for (size_t i = 0; i < nbk; ++i) order[i] = (uint32_t)i;
            std::sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return h[3 * a] > h[3 * b]; });
            double sum = 0.0; for (size_t i = 0; i < nbk; ++i) sum += h[3 * i];
            fprintf(stderr, "[k1] cull per block: %zu blocks, mean %.0f cycles, median %u, p99 %u; slowest:", nbk, sum / nbk, h[3 * order[nbk / 2]], h[3 * order[nbk / 100]]);
            for (int i = 0; i < 6 && i < (int)nbk; ++i) fprintf(stderr, " [blk %u: %u cycles, %u rays, %u views]", order[i], h[3 * order[i]], h[3 * order[i] + 1], h[3 * order[i] + 2]);
            fprintf(stderr, "\n");
        }
        fprintf(stderr, "[k1] cull: candidate pairs %llu, queued for the exact cull %llu, rejected there by back-face / range %llu, by the image test %llu, rays %llu\n",
            d[4], d[5], d[6], d[7], d[3]);
    }
    if (getenv("UAVMVS_SHAFT_DEBUG")) {
        unsigned long long d[8];
        UAV_CUDA(cudaMemcpy(d, k.counters, 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
        fprintf(stderr, "[k1] counters 4..6: %llu %llu %llu (entry: frontier entries, root fallbacks; shaft: steps ok / failed, triangle tests)\n", d[4], d[5], d[6]);
    }
    if (reset) UAV_CUDA(cudaMemset(k.counters, 0, 8 * sizeof(unsigned long long)));
    return 0;
}

}  // namespace uav

// Observation-ray path: K1 update_observation_rays (single view and batched over a
// trajectory), K2 process, K3 evaluate, K8 func, K9 reduction, and the fused
// evaluate_trajectory entry point.  Reference: libs/eval/kernels.cu:21-291, :496-510.
#include "common.cuh"
#include "trace.cuh"

namespace uav {

static Heur g_heur = {8.0f, 4.0f, 32.0f, 16.0f};   // m_k, m_x0, t_k, t_x0 (kernels.cu:93-96)
Heur current_heur() { return g_heur; }
constexpr float kPi = 3.14159265f;                            // kernels.cu:78

// ---- helpers restating kernels.cu:21-40, :63-91, :111-136 --------------------------
__device__ __forceinline__ V3 orthogonal(V3 v) {
    V3 n0 = mk(1.0f, 0.0f, 0.0f), n1 = mk(0.0f, 1.0f, 0.0f);
    if (fabsf(dot(n0, v)) < fabsf(dot(n1, v))) return cross(n0, v);
    return cross(n1, v);
}
__device__ __forceinline__ V3 relative_direction(V3 v2cn, V3 n) {
    V3 rx = normalized(orthogonal(n));
    V3 ry = normalized(cross(n, rx));
    V3 rel = mk(dot(rx, v2cn), dot(ry, v2cn), dot(n, v2cn));
    return normalized(rel);
}
// view: pos[3] w2c[16] calib[9]; returns true if the point projects inside the image
__device__ __forceinline__ bool in_image(const float * __restrict__ view, V3 v, int width, int height) {
    const float * W = view + 3; const float * K = view + 19;
    float q[3];
    #pragma unroll
    for (int i = 0; i < 3; ++i) {           // mult(w2c, v, 1.0f), libs/cacc/matrix.h:112-126
        float s = 0.0f;
        s += W[4 * i + 0] * v.x; s += W[4 * i + 1] * v.y; s += W[4 * i + 2] * v.z;
        q[i] = s + W[4 * i + 3] * 1.0f;
    }
    float p[3];
    #pragma unroll
    for (int i = 0; i < 3; ++i) {           // calib * q, libs/cacc/matrix.h:69-82
        float s = 0.0f;
        s += K[3 * i + 0] * q[0]; s += K[3 * i + 1] * q[1]; s += K[3 * i + 2] * q[2];
        p[i] = s;
    }
    float x = p[0] / p[2] - 0.5f, y = p[1] / p[2] - 0.5f;   // project(), kernels.cu:21-27
    return !(x < 0.0f || (float)width <= x || y < 0.0f || (float)height <= y);
}
__device__ __forceinline__ float logistic(float x, float k, float x0) {
    return 1.0f / (1.0f + expf(-k * (x - x0)));
}
__device__ __forceinline__ float func(float recon, float target) {
    float a = fabsf(fminf(recon - target, 0.0f));
    return a * a;
}
__device__ __forceinline__ float pair_term(float4 a, float4 b, float mk_, float mx0, float tk, float tx0) {
    float calpha = dot(mk(b.x, b.y, b.z), mk(a.x, a.y, a.z));
    float alpha = acosf(fmaxf(-1.0f, fminf(calpha, 1.0f)));
    float scale = fminf(a.w, b.w);
    float ctheta = fminf(a.z, b.z);
    float matchability = (1.0f - logistic(alpha, mk_, mx0)) * ctheta;
    float triangulation = logistic(alpha, tk, tx0) * scale;
    return matchability * triangulation;
}

// ---- K1 ----------------------------------------------------------------------------
// Two launches per call:
//   k_view_masks   conservative (super-block, view) and (block, view) frustum/range cull ->
//                  masks[sb * V + v] = bit b set iff view v may see a point of block 32 sb + b.
//   k_update_observation_rays
//                  persistent warps; each warp pulls 32-point blocks from a global counter,
//                  runs the exact cull of kernels.cu:156-167 for the candidate views, and feeds
//                  the surviving (point, view) pairs through a per-warp reorder ring: idle lanes
//                  are refilled from the ring while other lanes still traverse, leaf triangles
//                  are deferred until several lanes have some, and finished pairs retire IN
//                  ORDER, which reproduces the row order of the reference's serial one-launch-
//                  per-view loop (kernels.cu:175-181) and its un-clamped counters.
struct K1Args {
    BvhView bvh;
    const float4 * __restrict__ verts; const float4 * __restrict__ normals;
    uint32_t n, num_blocks;
    const float * __restrict__ views; uint32_t num_views;
    const uint32_t * __restrict__ masks;
    float max_distance; int width, height;
    float4 * obs; uint32_t * num_rows; uint32_t max_rows; uint32_t stride;   // stride in float4
    uint32_t * work;              // next block to hand out
    unsigned long long * stats;   // [rays, visible, pairs, pairs after block cull] or null
};

constexpr int K1_WARPS = 4;
constexpr int K1_RING = 128;      // reorder ring entries per warp
constexpr int K1_PEND = 6;        // deferred leaves per lane
constexpr int K1_REFILL = 8;      // refill when at least this many lanes are idle
constexpr int K1_TRI = 8;         // run the triangle phase when this many lanes hold a leaf
constexpr int K1_SB_PER_WARP = 8;
constexpr int K1_STACK = 160;     // <= 3 pushes per two binary levels; LBVH depth <= 96
constexpr int K1_STEPS = 4;       // node steps between two rounds of warp-level bookkeeping

// Conservative (group of points, view) test: false only if NO point inside the sphere
// (centre c, radius r) can pass the exact cull of kernels.cu:164-167 (range, image bounds).
// Margins are 100x above fp32 rounding of the exact test, so a rejected group is rejected
// point by point by the exact arithmetic as well.
__device__ __forceinline__ bool block_view_maybe(const float * __restrict__ view, float4 sph,
    float max_distance, int width, int height)
{
    float dx = view[0] - sph.x, dy = view[1] - sph.y, dz = view[2] - sph.z;
    float D = sqrtf(dx * dx + dy * dy + dz * dz);
    if (D * 0.99999f - sph.w >= max_distance * 1.00001f) return false;
    const float * W = view + 3; const float * K = view + 19;
    float q[3], e2 = 0.0f, fro2 = 0.0f;
    #pragma unroll
    for (int i = 0; i < 3; ++i) {
        float a = W[4 * i] * sph.x, b = W[4 * i + 1] * sph.y, c = W[4 * i + 2] * sph.z, d = W[4 * i + 3];
        q[i] = a + b + c + d;
        float e = 1e-6f * (fabsf(a) + fabsf(b) + fabsf(c) + fabsf(d));
        e2 += e * e;
        fro2 += W[4 * i] * W[4 * i] + W[4 * i + 1] * W[4 * i + 1] + W[4 * i + 2] * W[4 * i + 2];
    }
    float qn = sqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2]);
    float rho = sph.w * sqrtf(fro2) * 1.001f + sqrtf(e2) + 1e-4f * (qn + 1.0f);
    float k2n = sqrtf(K[6] * K[6] + K[7] * K[7] + K[8] * K[8]);
    float p2 = K[6] * q[0] + K[7] * q[1] + K[8] * q[2];
    if (!(p2 - k2n * rho > 0.0f)) return true;          // ball not entirely in front: undecided
    float lim[4] = {0.5f, (float)width + 0.5f, 0.5f, (float)height + 0.5f};
    #pragma unroll
    for (int s = 0; s < 4; ++s) {
        int r = s >> 1;                                   // image row of K: 0 -> x, 1 -> y
        float a0 = K[3 * r] - lim[s] * K[6], a1 = K[3 * r + 1] - lim[s] * K[7], a2 = K[3 * r + 2] - lim[s] * K[8];
        float an = sqrtf(a0 * a0 + a1 * a1 + a2 * a2);
        float v = a0 * q[0] + a1 * q[1] + a2 * q[2];
        if ((s & 1) == 0) { if (v + an * rho < 0.0f) return false; }   // coordinate < 0 everywhere
        else { if (v - an * rho > 0.0f) return false; }                // coordinate >= size everywhere
    }
    return true;
}

// One warp = one chunk of 32 views (a view per lane, kept in registers) x K1_SB_PER_WARP
// super-blocks.  Surviving views are re-tested against the 32 blocks (a block per lane).
__global__ void __launch_bounds__(K1_WARPS * 32)
k_view_masks(const float * __restrict__ views, uint32_t num_views, const float4 * __restrict__ super_spheres,
    uint32_t num_super, const float4 * __restrict__ block_spheres, uint32_t num_blocks,
    float max_distance, int width, int height, uint32_t * __restrict__ masks, unsigned long long * stats)
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
    unsigned long long cand_pairs = 0;
    for (uint32_t sb = sb0; sb < min(sb0 + K1_SB_PER_WARP, num_super); ++sb) {
        bool cand = v < num_views && block_view_maybe(view, __ldg(super_spheres + sb), max_distance, width, height);
        uint32_t m = __ballot_sync(0xffffffffu, cand);
        uint32_t mine = 0;
        const uint32_t b = sb * 32 + lane;
        const float4 bs = __ldg(block_spheres + (b < num_blocks ? b : num_blocks - 1));
        while (m) {
            const int j = __ffs(m) - 1; m &= m - 1;
            const float * vj = views + (size_t)(blockIdx.x * 32 + j) * UAVMVS_VIEW_FLOATS;   // uniform address
            bool c = b < num_blocks && block_view_maybe(vj, bs, max_distance, width, height);
            uint32_t word = __ballot_sync(0xffffffffu, c);
            if ((int)lane == j) mine = word;
        }
        if (v < num_views) masks[(size_t)sb * num_views + v] = mine;
        cand_pairs += __popc(mine);
    }
    if (stats) {
        for (int o = 16; o > 0; o >>= 1) cand_pairs += __shfl_xor_sync(0xffffffffu, cand_pairs, o);
        if (lane == 0) atomicAdd(stats + 3, cand_pairs * 32ull);
    }
}

struct K1Warp {
    float4 pos[32], nrm[32], rx[32], ry[32];
    float4 dir[K1_RING];                  // unit point->camera direction, w = distance l
    uint32_t cnt[32];
    uint32_t entry[K1_RING];              // lane | view << 5
    uint32_t pend[K1_PEND][32];           // deferred leaves: first << 5 | count   (count <= 31)
    uint8_t status[K1_RING];              // 0 in flight, 1 occluded, 2 visible
};

// EXACT_INNER = false relies on the tree's boxes being nested (every child box inside its
// parent's box, which holds for exact-union trees such as acc's SAH tree and our LBVH and is
// verified at upload).  The slab test is monotone in the box bounds (each of fl(b - o),
// fl(x / d), min, max is), so a ray that passes a leaf's box passes every ancestor's box: the
// reference reaches a leaf iff the leaf's OWN box passes.  Inner boxes then only have to be
// tested conservatively (never reject a box the exact test accepts); the exact decision, with
// its IEEE-division fallback, is needed for leaf boxes alone.  Rays for which the argument
// does not hold (a zero / tiny / non-finite direction component: NaN in the slab chain) have
// eabs = inf and take the exact test at every node.
#ifndef K1_MIN_BLOCKS
#define K1_MIN_BLOCKS 5
#endif
template <bool POPULATE, bool EXACT_INNER>
__global__ void __launch_bounds__(K1_WARPS * 32, K1_MIN_BLOCKS)
k_update_observation_rays(K1Args a)
{
    __shared__ K1Warp s_warp[K1_WARPS];
    K1Warp & S = s_warp[threadIdx.x >> 5];
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const uint32_t V = a.num_views;
    uint32_t stack[K1_STACK];
    unsigned long long st_rays = 0, st_vis = 0;

    for (;;) {
        uint32_t blk = 0;
        if (lane == 0) blk = atomicAdd(a.work, 1u);
        blk = __shfl_sync(0xffffffffu, blk, 0);
        if (blk >= a.num_blocks) break;
        const uint32_t id = blk * 32 + lane;
        const bool valid = id < a.n;
        __syncwarp();
        {
            float4 v4 = a.verts[valid ? id : a.n - 1], n4 = a.normals[valid ? id : a.n - 1];
            S.pos[lane] = v4; S.nrm[lane] = n4;
            V3 nrm = ld3(n4);
            V3 rx = normalized(orthogonal(nrm));          // relative_direction(), kernels.cu:63-76
            V3 ry = normalized(cross(nrm, rx));
            S.rx[lane] = make_float4(rx.x, rx.y, rx.z, 0.0f);
            S.ry[lane] = make_float4(ry.x, ry.y, ry.z, 0.0f);
            S.cnt[lane] = valid ? a.num_rows[id] : 0u;
        }
        const V3 v = ld3(S.pos[lane]), nrm = ld3(S.nrm[lane]);
        const uint32_t * mrow = a.masks + (size_t)(blk >> 5) * V;
        const uint32_t mbit = blk & 31;
        __syncwarp();

        uint32_t head = 0, issue = 0, tail = 0;          // ring pointers (monotonic)
        uint32_t chunk = 0, word = 0; bool exhausted = false;
        // lane state
        bool has_ray = false, traversing = false, hit = false;
        uint32_t slot = 0, cur = 0, npend = 0; int sp = 0;
        RayF rf; rf.rdx = rf.rdy = rf.rdz = rf.nox = rf.noy = rf.noz = rf.tmin = rf.tmax = rf.eabs = 0.0f;
        V3 rd = mk(0.f, 0.f, 0.f), ro = mk(0.f, 0.f, 0.f);
        uint32_t serviced_idle = 0;                      // idle mask after the last service
        // a leaf reference packs first << 5 | count; the (rare) part of a leaf beyond 31 triangles is
        // tested at once
        auto defer_leaf = [&](uint32_t first, uint32_t count) {
            if (count > 31u) {
                Ray ray; ray.o = ro; ray.d = rd; ray.tmin = rf.tmin; ray.tmax = rf.tmax;
                float t;
                for (uint32_t i = 31u; i < count; ++i) hit |= hit_tri(ray, a.bvh.tris + first + i, t);
                count = 31u;
            }
            S.pend[npend][lane] = (first << 5) | count; ++npend;
        };

        for (;;) {
            uint32_t idle = __ballot_sync(0xffffffffu, !has_ray);
            if (idle != serviced_idle && __popc(idle) >= K1_REFILL) {
                __syncwarp();                             // status bytes written by finishing lanes
                // ---------------- retire finished pairs in order ----------------
                for (;;) {
                    uint32_t e_idx = head + lane;
                    bool in = e_idx < issue;
                    uint32_t stt = in ? S.status[e_idx & (K1_RING - 1)] : 0u;
                    uint32_t done_mask = __ballot_sync(0xffffffffu, stt != 0u);
                    uint32_t nret = (done_mask == 0xffffffffu) ? 32u : (uint32_t)(__ffs(~done_mask) - 1);
                    if (nret == 0) break;
                    bool vis = lane < nret && stt == 2u;
                    uint32_t e = S.entry[e_idx & (K1_RING - 1)];
                    uint32_t pt = vis ? (e & 31u) : lane;
                    uint32_t vis_mask = __ballot_sync(0xffffffffu, vis);
                    uint32_t peers = __match_any_sync(0xffffffffu, pt) & vis_mask;
                    if (vis) ++st_vis;
                    V3 rel = mk(0.f, 0.f, 0.f); float scale = 0.f;
                    const uint32_t row_pt = blk * 32 + pt;
                    if (vis) {
                        float4 d4 = S.dir[e_idx & (K1_RING - 1)];
                        V3 v2cn = ld3(d4);
                        rel = normalized(mk(dot(ld3(S.rx[pt]), v2cn), dot(ld3(S.ry[pt]), v2cn), dot(ld3(S.nrm[pt]), v2cn)));
                        scale = 1.0f - (d4.w / a.max_distance);
                    }
                    if (POPULATE) {
                        if (vis) {
                            uint32_t row = S.cnt[pt] + __popc(peers & lt_mask);
                            if (row < a.max_rows)
                                a.obs[(size_t)row * a.stride + row_pt] = make_float4(rel.x, rel.y, rel.z, scale);
                        }
                        __syncwarp();
                        if (vis && (peers >> lane) == 1u) S.cnt[pt] += __popc(peers);   // highest peer updates
                    } else {
                        // kernels.cu:183-200: mark the first stored ray equal within 1e-5; entries of the
                        // same point are applied in view order
                        uint32_t rank = __popc(peers & lt_mask);
                        uint32_t max_rank = 0;
                        for (uint32_t m = vis_mask; m; m &= m - 1)
                            max_rank = max(max_rank, __shfl_sync(0xffffffffu, rank, __ffs(m) - 1));
                        for (uint32_t r = 0; r <= max_rank; ++r) {
                            if (vis && rank == r) {
                                uint32_t rows = min(S.cnt[pt], a.max_rows);
                                volatile float4 * col = a.obs + row_pt;
                                for (uint32_t i = 0; i < rows; ++i) {
                                    volatile float4 * p = col + (size_t)i * a.stride;
                                    float ox = p->x, oy = p->y, oz = p->z, ow = p->w;
                                    bool equal = fabsf(rel.x - ox) < 1e-5f && fabsf(rel.y - oy) < 1e-5f
                                        && fabsf(rel.z - oz) < 1e-5f && fabsf(scale - ow) < 1e-5f;
                                    if (equal) { p->w = -1.0f; break; }
                                }
                            }
                            __syncwarp();
                        }
                    }
                    __syncwarp();
                    head += nret;
                    if (nret < 32) break;
                }
                // ---------------- produce: exact cull of the next candidate views ----------------
                while (!exhausted && tail - issue < 32 && tail - head <= K1_RING - 32) {
                    if (!word) {
                        if (chunk * 32 >= V) { exhausted = true; break; }
                        uint32_t vi = chunk * 32 + lane;
                        uint32_t mw = vi < V ? __ldg(mrow + vi) : 0u;
                        word = __ballot_sync(0xffffffffu, (mw >> mbit) & 1u);
                        ++chunk;
                        continue;
                    }
                    const uint32_t k = (chunk - 1) * 32 + (__ffs(word) - 1); word &= word - 1;
                    const float * view = a.views + UAVMVS_VIEW_FLOATS * (size_t)k;
                    float vw[UAVMVS_VIEW_FLOATS];
                    #pragma unroll
                    for (int i = 0; i < UAVMVS_VIEW_FLOATS / 4; ++i) {          // uniform address: broadcast
                        float4 t = __ldg(reinterpret_cast<const float4 *>(view) + i);
                        vw[4 * i] = t.x; vw[4 * i + 1] = t.y; vw[4 * i + 2] = t.z; vw[4 * i + 3] = t.w;
                    }
                    // exact cull, kernels.cu:156-167
                    V3 v2c = mk(vw[0], vw[1], vw[2]) - v;
                    float l = norm(v2c);
                    V3 v2cn = v2c / l;
                    float ctheta = dot(v2cn, nrm);
                    bool pass = valid && !(ctheta < 0.087f) && !(l >= a.max_distance);
                    if (pass) pass = in_image(vw, v, a.width, a.height);
                    uint32_t m = __ballot_sync(0xffffffffu, pass);
                    if (pass) {
                        uint32_t t = (tail + __popc(m & lt_mask)) & (K1_RING - 1);
                        S.entry[t] = lane | (k << 5);
                        S.dir[t] = make_float4(v2cn.x, v2cn.y, v2cn.z, l);
                        S.status[t] = 0;
                    }
                    tail += __popc(m);
                }
                __syncwarp();
                // ---------------- issue: idle lanes take the next entries ----------------
                {
                    uint32_t navail = tail - issue;
                    uint32_t r = __popc(idle & lt_mask);
                    if (!has_ray && r < navail) {
                        slot = (issue + r) & (K1_RING - 1);
                        uint32_t e = S.entry[slot];
                        float4 d4 = S.dir[slot];
                        Ray ray; ray.o = ld3(S.pos[e & 31u]); ray.d = ld3(d4);
                        ray.tmin = d4.w * 0.001f; ray.tmax = d4.w;                  // kernels.cu:54-58
                        rf = make_rayf(ray); ro = ray.o; rd = ray.d;
                        has_ray = true; hit = false; sp = 0; npend = 0; ++st_rays;
                        if (a.bvh.root_is_leaf) { traversing = false; defer_leaf(a.bvh.root_first, a.bvh.root_count); }
                        else { traversing = true; cur = 0; }
                    }
                    issue += min((uint32_t)__popc(idle), navail);
                }
                serviced_idle = __ballot_sync(0xffffffffu, !has_ray);
                if (serviced_idle == 0xffffffffu) {
                    if (exhausted && head == tail) break;     // block finished
                    serviced_idle = 0;                        // (cannot happen; forces another service)
                    continue;
                }
            }

            // ---------------- K1_STEPS traversal steps (4-wide nodes) ----------------
            #pragma unroll 1
            for (int step = 0; step < K1_STEPS; ++step) {
                if (traversing && npend <= K1_PEND - 4) {
                    const Node128 * np = a.bvh.wide + cur;
                    const float4 q0 = __ldg(&np->q[0]), q1 = __ldg(&np->q[1]), q2 = __ldg(&np->q[2]);
                    const float4 q3 = __ldg(&np->q[3]), q4 = __ldg(&np->q[4]), q5 = __ldg(&np->q[5]);
                    const uint4 ref = __ldg(&np->ref), cnt = __ldg(&np->cnt);
                    bool s0, s1, s2, s3, d0, d1, d2, d3; float t0, t1, t2, t3;   // sure hit, decided, entry
                    slab_fast(rf, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, s0, d0, t0);
                    slab_fast(rf, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, s1, d1, t1);
                    slab_fast(rf, q3.x, q3.y, q3.z, q3.w, q4.x, q4.y, s2, d2, t2);
                    slab_fast(rf, q4.z, q4.w, q5.x, q5.y, q5.z, q5.w, s3, d3, t3);
                    const bool always_exact = EXACT_INNER || rf.eabs == UAV_INF;
                    const bool l0 = cnt.x != 0u, l1 = cnt.y != 0u, l2 = cnt.z != 0u, l3 = cnt.w != 0u;   // leaf (or empty)
                    // undecided boxes: exact test where the decision matters, conservative visit elsewhere
                    bool h0 = (s0 || !d0) && cnt.x != UAV_NAI, h1 = (s1 || !d1) && cnt.y != UAV_NAI;
                    bool h2 = (s2 || !d2) && cnt.z != UAV_NAI, h3 = (s3 || !d3) && cnt.w != UAV_NAI;
                    const bool x0 = h0 && !d0 && (always_exact || l0), x1 = h1 && !d1 && (always_exact || l1);
                    const bool x2 = h2 && !d2 && (always_exact || l2), x3 = h3 && !d3 && (always_exact || l3);
                    if (x0 || x1 || x2 || x3) {               // rare
                        if (x0) h0 = slab_exact_call(ro.x, ro.y, ro.z, rd.x, rd.y, rd.z, rf.tmin, rf.tmax,
                                                     q0.x, q0.y, q0.z, q0.w, q1.x, q1.y);
                        if (x1) h1 = slab_exact_call(ro.x, ro.y, ro.z, rd.x, rd.y, rd.z, rf.tmin, rf.tmax,
                                                     q1.z, q1.w, q2.x, q2.y, q2.z, q2.w);
                        if (x2) h2 = slab_exact_call(ro.x, ro.y, ro.z, rd.x, rd.y, rd.z, rf.tmin, rf.tmax,
                                                     q3.x, q3.y, q3.z, q3.w, q4.x, q4.y);
                        if (x3) h3 = slab_exact_call(ro.x, ro.y, ro.z, rd.x, rd.y, rd.z, rf.tmin, rf.tmax,
                                                     q4.z, q4.w, q5.x, q5.y, q5.z, q5.w);
                    }
                    // leaf children are deferred (any-hit is order independent)
                    if (h0 && l0) defer_leaf(ref.x, cnt.x);
                    if (h1 && l1) defer_leaf(ref.y, cnt.y);
                    if (h2 && l2) defer_leaf(ref.z, cnt.z);
                    if (h3 && l3) defer_leaf(ref.w, cnt.w);
                    // inner children: nearest first, the others go on the stack
                    const float k0 = (h0 && !l0) ? t0 : UAV_INF, k1 = (h1 && !l1) ? t1 : UAV_INF;
                    const float k2 = (h2 && !l2) ? t2 : UAV_INF, k3 = (h3 && !l3) ? t3 : UAV_INF;
                    // (an entry distance is never +inf for a box that can be hit: tmin <= tmax <= ray.tmax)
                    float kb = k0; uint32_t rb = ref.x;
                    if (k1 < kb) { kb = k1; rb = ref.y; }
                    if (k2 < kb) { kb = k2; rb = ref.z; }
                    if (k3 < kb) { kb = k3; rb = ref.w; }
                    if (k0 < UAV_INF && ref.x != rb) stack[sp++] = ref.x;
                    if (k1 < UAV_INF && ref.y != rb) stack[sp++] = ref.y;
                    if (k2 < UAV_INF && ref.z != rb) stack[sp++] = ref.z;
                    if (k3 < UAV_INF && ref.w != rb) stack[sp++] = ref.w;
                    if (kb < UAV_INF) cur = rb;
                    else if (sp > 0) cur = stack[--sp];
                    else traversing = false;
                }
            }
            // ---------------- triangle phase ----------------
            {
                uint32_t pmask = __ballot_sync(0xffffffffu, npend > 0);
                if (pmask) {
                    uint32_t tmask = __ballot_sync(0xffffffffu, traversing);
                    uint32_t full = __ballot_sync(0xffffffffu, npend > K1_PEND - 4);
                    if (__popc(pmask) >= K1_TRI || full || !tmask) {
                        if (npend > 0) {
                            Ray ray; ray.o = ro; ray.d = rd; ray.tmin = rf.tmin; ray.tmax = rf.tmax;
                            for (uint32_t p = 0; p < npend && !hit; ++p) {
                                uint32_t ref = S.pend[p][lane];
                                uint32_t first = ref >> 5, count = ref & 31u;
                                float t;
                                for (uint32_t i = 0; i < count; ++i) hit |= hit_tri(ray, a.bvh.tris + first + i, t);
                            }
                            npend = 0;
                        }
                        __syncwarp();
                    }
                }
            }
            // ---------------- finished lanes post their result ----------------
            if (has_ray && (hit || (!traversing && npend == 0))) {
                S.status[slot] = hit ? 1 : 2;
                has_ray = false; traversing = false; npend = 0;
            }
        }
        __syncwarp();
        if (POPULATE && valid) a.num_rows[id] = S.cnt[lane];
    }
    if (a.stats) {
        for (int o = 16; o > 0; o >>= 1) {
            st_rays += __shfl_xor_sync(0xffffffffu, st_rays, o);
            st_vis += __shfl_xor_sync(0xffffffffu, st_vis, o);
        }
        if (lane == 0) { atomicAdd(a.stats + 0, st_rays); atomicAdd(a.stats + 1, st_vis); }
    }
}

// ---- K2: process_observation_rays, kernels.cu:204-264 (one warp per point) ----------
__global__ void __launch_bounds__(128)
k_process_observation_rays(float4 * obs, uint32_t * num_rows_ptr, uint32_t max_rows, uint32_t stride, uint32_t n)
{
    const uint32_t tx = threadIdx.x & 31;
    const uint32_t id = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (id >= n) return;
    uint32_t num_rows = min(num_rows_ptr[id], max_rows);
    float4 * col = obs + id;
    float4 ray = make_float4(0.f, 0.f, 0.f, 0.f);
    int key = (int)tx;
    float theta = -1.0f;
    if (tx < num_rows) {
        ray = col[(size_t)tx * stride];
        theta = (ray.w >= 0.0f) ? ray.z : -1.0f;
    }
    for (int i = 0; (1u << i) < num_rows; ++i) {
        bool asc = (tx >> (i + 1)) % 2;
        for (int s = 1 << i; s > 0; s >>= 1) {
            bool dir = (tx % (s << 1)) < s;
            float otheta = __shfl_xor_sync(0xffffffffu, theta, s);
            int okey = __shfl_xor_sync(0xffffffffu, key, s);
            if (otheta != theta && ((otheta > theta) == (asc ^ dir))) { theta = otheta; key = okey; }
        }
    }
    uint32_t num_valid = __popc(__ballot_sync(0xffffffffu, theta >= 0.0f));
    if (num_valid < num_rows && tx == 0) num_rows_ptr[id] = num_valid;
    ray.x = __shfl_sync(0xffffffffu, ray.x, key);
    ray.y = __shfl_sync(0xffffffffu, ray.y, key);
    ray.z = __shfl_sync(0xffffffffu, ray.z, key);
    ray.w = __shfl_sync(0xffffffffu, ray.w, key);
    if (tx < num_valid) col[(size_t)tx * stride] = ray;
}

// ---- K3: evaluate_observation_rays, kernels.cu:266-291 (one thread per point) -------
// The point's column is staged in shared memory once; the pair sum keeps the reference's
// order (inner j < i ascending, outer i ascending).
constexpr int K3_THREADS = 128;
__global__ void __launch_bounds__(K3_THREADS)
k_evaluate_observation_rays(const float4 * __restrict__ obs, const uint32_t * __restrict__ num_rows_ptr,
    uint32_t max_rows, uint32_t stride, uint32_t n, float * __restrict__ recons,
    float target, float * __restrict__ wrecons, Heur h)
{
    extern __shared__ float4 s_col[];   // [max_rows][K3_THREADS]
    const uint32_t id = blockIdx.x * K3_THREADS + threadIdx.x;
    if (id >= n) return;
    const uint32_t rows = min(num_rows_ptr[id], max_rows);
    for (uint32_t i = 0; i < rows; ++i) s_col[i * K3_THREADS + threadIdx.x] = obs[(size_t)i * stride + id];
    const float m_k = h.m_k, m_x0 = kPi / h.m_x0, t_k = h.t_k, t_x0 = kPi / h.t_x0;
    float recon = rows >= 1 ? 0.0f : -1.0f;
    for (uint32_t i = 1; i < rows; ++i) {
        float4 ri = s_col[i * K3_THREADS + threadIdx.x];
        float sum = 0.0f;
        for (uint32_t j = 0; j < i; ++j)
            sum += pair_term(s_col[j * K3_THREADS + threadIdx.x], ri, m_k, m_x0, t_k, t_x0);
        recon += sum;
    }
    recons[id] = recon;
    if (wrecons) wrecons[id] = func(recon, target);
}

// ---- K8: calculate_func_recons, kernels.cu:496-510 -----------------------------------
__global__ void k_calculate_func_recons(const float * __restrict__ recons, uint32_t n, float target,
    float * __restrict__ wrecons)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) wrecons[i] = func(recons[i], target);
}

// ---- K9: sum/min/max, libs/cacc/reduction.h:104-162 ----------------------------------
// acc[0] = sum (double), acc[1] = min, acc[2] = max as order-preserving ints.
__device__ __forceinline__ int f2ord(float f) { int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; }
__host__ __device__ __forceinline__ float ord2f(int i) {
    int b = i >= 0 ? i : i ^ 0x7fffffff;
#ifdef __CUDA_ARCH__
    return __int_as_float(b);
#else
    float f; memcpy(&f, &b, 4); return f;
#endif
}
struct ReduceAcc { double sum; int mn, mx; int pad; };

__global__ void k_reduce_init(ReduceAcc * acc) { acc->sum = 0.0; acc->mn = 0x7fffffff; acc->mx = (int)0x80000000; }

__global__ void __launch_bounds__(256)
k_reduce(const float * __restrict__ x, uint32_t n, ReduceAcc * acc)
{
    float s = 0.0f, mn = UAV_INF, mx = -UAV_INF;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        float v = x[i]; s += v; mn = fminf(mn, v); mx = fmaxf(mx, v);
    }
    for (int o = 16; o > 0; o >>= 1) {
        s += __shfl_xor_sync(0xffffffffu, s, o);
        mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    __shared__ float ss[8], smn[8], smx[8];
    uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) { ss[w] = s; smn[w] = mn; smx[w] = mx; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int i = 0; i < 8; ++i) { t += (double)ss[i]; mn = fminf(mn, smn[i]); mx = fmaxf(mx, smx[i]); }
        atomicAdd(&acc->sum, t);
        atomicMin(&acc->mn, f2ord(mn));
        atomicMax(&acc->mx, f2ord(mx));
    }
}

// ---- launch helpers --------------------------------------------------------------------
struct Scratch {   // per-process scratch on the device (views, reduction accumulator, stats)
    float * views = nullptr; size_t views_cap = 0;
    ReduceAcc * acc = nullptr;
    unsigned long long * stats = nullptr;
    uint32_t * work = nullptr;                       // K1 block counter
    uint32_t * masks = nullptr; size_t masks_cap = 0;   // K1 (super-block, view) block masks
    ReduceAcc * h_acc = nullptr; unsigned long long * h_stats = nullptr;   // pinned
    cudaEvent_t ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};     // K1 | K2 | K3+K8 | K9 boundaries
    bool ev_valid = false;
};
static thread_local Scratch g_scratch;

static int ensure_scratch(size_t view_floats)
{
    Scratch & s = g_scratch;
    if (!s.acc) {
        UAV_CUDA(cudaMalloc(&s.acc, sizeof(ReduceAcc)));
        UAV_CUDA(cudaMalloc(&s.stats, 4 * sizeof(unsigned long long)));
        UAV_CUDA(cudaMalloc(&s.work, sizeof(uint32_t)));
        UAV_CUDA(cudaMallocHost(&s.h_acc, sizeof(ReduceAcc)));
        UAV_CUDA(cudaMallocHost(&s.h_stats, 4 * sizeof(unsigned long long)));
        for (int i = 0; i < 5; ++i) UAV_CUDA(cudaEventCreate(&s.ev[i]));
    }
    if (view_floats > s.views_cap) {
        if (s.views) cudaFree(s.views);
        UAV_CUDA(cudaMalloc(&s.views, view_floats * sizeof(float)));
        s.views_cap = view_floats;
    }
    return 0;
}

static int launch_k1(cudaStream_t s, int populate, const float * d_views, uint32_t num_views,
    float max_distance, int width, int height, const uavmvs_bvh * bvh, const uavmvs_cloud * cloud,
    uavmvs_obs * obs, unsigned long long * d_stats)
{
    UAV_REQUIRE(bvh && cloud && obs, "null handle");
    UAV_REQUIRE(obs->num_cols == cloud->n, "obs_rays columns != cloud vertices");
    UAV_REQUIRE(bvh->num_tris < (1u << 27), "more than 2^27 triangles");
    UAV_REQUIRE(num_views < (1u << 27), "more than 2^27 views");
    if (num_views == 0) return 0;
    if (ensure_scratch(0)) return 1;
    Scratch & sc = g_scratch;
    size_t need = (size_t)cloud->num_super * num_views;
    if (need > sc.masks_cap) {
        // stream-ordered: kernels still using the old buffer finish first
        if (sc.masks) UAV_CUDA(cudaFreeAsync(sc.masks, s));
        UAV_CUDA(cudaMallocAsync(&sc.masks, need * sizeof(uint32_t), s));
        sc.masks_cap = need;
    }
    static int grid_persistent = 0;
    if (!grid_persistent) {
        int dev = 0, sms = 0, occ_t = 0, occ_f = 0;
        UAV_CUDA(cudaGetDevice(&dev));
        UAV_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        UAV_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_t, k_update_observation_rays<true, true>, K1_WARPS * 32, 0));
        UAV_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_f, k_update_observation_rays<false, true>, K1_WARPS * 32, 0));
        int occ = occ_t < occ_f ? occ_t : occ_f;
        grid_persistent = sms * (occ > 0 ? occ : 1);
    }
    dim3 mgrid(divup(num_views, 32), divup(cloud->num_super, K1_WARPS * K1_SB_PER_WARP));
    k_view_masks<<<mgrid, K1_WARPS * 32, 0, s>>>(d_views, num_views, cloud->super_sphere, cloud->num_super,
        cloud->block_sphere, cloud->num_blocks, max_distance, width, height, sc.masks, d_stats);
    UAV_CUDA(cudaGetLastError());
    UAV_CUDA(cudaMemsetAsync(sc.work, 0, sizeof(uint32_t), s));
    K1Args a;
    a.bvh = view_of(bvh);
    a.verts = cloud->vertices; a.normals = cloud->normals; a.n = cloud->n; a.num_blocks = cloud->num_blocks;
    a.views = d_views; a.num_views = num_views; a.masks = sc.masks;
    a.max_distance = max_distance; a.width = width; a.height = height;
    a.obs = obs->data; a.num_rows = obs->num_rows; a.max_rows = obs->max_rows;
    a.stride = (uint32_t)(obs->pitch / sizeof(float4));
    a.work = sc.work; a.stats = d_stats;
    unsigned grid = (unsigned)grid_persistent;
    unsigned want = divup(cloud->num_blocks, K1_WARPS);
    if (grid > want) grid = want;
    const bool exact_inner = !bvh->nested;
    if (populate) {
        if (exact_inner) k_update_observation_rays<true, true><<<grid, K1_WARPS * 32, 0, s>>>(a);
        else k_update_observation_rays<true, false><<<grid, K1_WARPS * 32, 0, s>>>(a);
    } else {
        if (exact_inner) k_update_observation_rays<false, true><<<grid, K1_WARPS * 32, 0, s>>>(a);
        else k_update_observation_rays<false, false><<<grid, K1_WARPS * 32, 0, s>>>(a);
    }
    UAV_CUDA(cudaGetLastError());
    return 0;
}

static int launch_k3(cudaStream_t s, const uavmvs_obs * obs, uavmvs_array * recons, float target, uavmvs_array * wrecons)
{
    UAV_REQUIRE(obs && recons, "null handle");
    UAV_REQUIRE(recons->n == obs->num_cols, "recons size != obs columns");
    size_t smem = sizeof(float4) * obs->max_rows * K3_THREADS;
    static bool attr_set = false;
    if (!attr_set) {
        UAV_CUDA(cudaFuncSetAttribute(k_evaluate_observation_rays, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
        attr_set = true;
    }
    k_evaluate_observation_rays<<<divup(obs->num_cols, K3_THREADS), K3_THREADS, smem, s>>>(
        obs->data, obs->num_rows, obs->max_rows, (uint32_t)(obs->pitch / sizeof(float4)), obs->num_cols,
        recons->data, target, wrecons ? wrecons->data : nullptr, g_heur);
    UAV_CUDA(cudaGetLastError());
    return 0;
}

static int launch_reduce(cudaStream_t s, const uavmvs_array * a)
{
    k_reduce_init<<<1, 1, 0, s>>>(g_scratch.acc);
    unsigned grid = min(divup(a->n, 256 * 8), 148u * 8u);
    k_reduce<<<grid ? grid : 1, 256, 0, s>>>(a->data, a->n, g_scratch.acc);
    UAV_CUDA(cudaGetLastError());
    UAV_CUDA(cudaMemcpyAsync(g_scratch.h_acc, g_scratch.acc, sizeof(ReduceAcc), cudaMemcpyDeviceToHost, s));
    return 0;
}

}  // namespace uav

using namespace uav;

extern "C" {

int uavmvs_configure_heuristic(float m_k, float m_x0, float t_k, float t_x0)
{
    g_heur.m_k = m_k; g_heur.m_x0 = m_x0; g_heur.t_k = t_k; g_heur.t_x0 = t_x0;
    return 0;
}

int uavmvs_update_observation_rays_batch_device(void * stream, int populate, const float * d_views,
    uint32_t num_views, float max_distance, int width, int height,
    const uavmvs_bvh * bvh, const uavmvs_cloud * cloud, uavmvs_obs * obs)
{
    return launch_k1((cudaStream_t)stream, populate, d_views, num_views, max_distance, width, height,
        bvh, cloud, obs, nullptr);
}

int uavmvs_update_observation_rays_batch(void * stream, int populate, const float * views,
    uint32_t num_views, float max_distance, int width, int height,
    const uavmvs_bvh * bvh, const uavmvs_cloud * cloud, uavmvs_obs * obs)
{
    UAV_REQUIRE(views || num_views == 0, "null views");
    if (num_views == 0) return 0;
    cudaStream_t s = (cudaStream_t)stream;
    if (ensure_scratch((size_t)num_views * UAVMVS_VIEW_FLOATS)) return 1;
    UAV_CUDA(cudaMemcpyAsync(g_scratch.views, views, sizeof(float) * UAVMVS_VIEW_FLOATS * (size_t)num_views,
        cudaMemcpyHostToDevice, s));
    int rc = launch_k1(s, populate, g_scratch.views, num_views, max_distance, width, height, bvh, cloud, obs, nullptr);
    if (rc) return rc;
    // the scratch view buffer may be reused by the next call: keep stream order, no host sync needed
    return 0;
}

int uavmvs_update_observation_rays(void * stream, int populate, const float view_pos[3],
    float max_distance, const float w2c[16], const float calib[9], int width, int height,
    const uavmvs_bvh * bvh, const uavmvs_cloud * cloud, uavmvs_obs * obs)
{
    UAV_REQUIRE(view_pos && w2c && calib, "null argument");
    float view[UAVMVS_VIEW_FLOATS];
    for (int i = 0; i < 3; ++i) view[i] = view_pos[i];
    for (int i = 0; i < 16; ++i) view[3 + i] = w2c[i];
    for (int i = 0; i < 9; ++i) view[19 + i] = calib[i];
    // one private device copy per call so concurrent streams never share the staging buffer
    cudaStream_t s = (cudaStream_t)stream;
    float * d_view;
    UAV_CUDA(cudaMallocAsync(&d_view, sizeof(view), s));
    UAV_CUDA(cudaMemcpyAsync(d_view, view, sizeof(view), cudaMemcpyHostToDevice, s));
    int rc = launch_k1(s, populate, d_view, 1, max_distance, width, height, bvh, cloud, obs, nullptr);
    UAV_CUDA(cudaFreeAsync(d_view, s));
    return rc;
}

int uavmvs_process_observation_rays(void * stream, uavmvs_obs * obs)
{
    UAV_REQUIRE(obs, "null obs");
    k_process_observation_rays<<<divup((uint64_t)obs->num_cols * 32, 128), 128, 0, (cudaStream_t)stream>>>(
        obs->data, obs->num_rows, obs->max_rows, (uint32_t)(obs->pitch / sizeof(float4)), obs->num_cols);
    UAV_CUDA(cudaGetLastError());
    return 0;
}

int uavmvs_evaluate_observation_rays(void * stream, const uavmvs_obs * obs, uavmvs_array * recons)
{
    return launch_k3((cudaStream_t)stream, obs, recons, 0.0f, nullptr);
}

int uavmvs_calculate_func_recons(void * stream, const uavmvs_array * recons, float target_recon, uavmvs_array * wrecons)
{
    UAV_REQUIRE(recons && wrecons && recons->n == wrecons->n, "bad arrays");
    k_calculate_func_recons<<<divup(recons->n, 256), 256, 0, (cudaStream_t)stream>>>(recons->data, recons->n,
        target_recon, wrecons->data);
    UAV_CUDA(cudaGetLastError());
    return 0;
}

int uavmvs_reduce(void * stream, const uavmvs_array * a, float out3[3])
{
    UAV_REQUIRE(a && out3, "null argument");
    if (ensure_scratch(0)) return 1;
    if (launch_reduce((cudaStream_t)stream, a)) return 1;
    UAV_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    out3[0] = (float)g_scratch.h_acc->sum; out3[1] = ord2f(g_scratch.h_acc->mn); out3[2] = ord2f(g_scratch.h_acc->mx);
    return 0;
}

int uavmvs_evaluate_trajectory_device(void * stream, const float * d_views, uint32_t num_views,
    float max_distance, float target_recon, int width, int height,
    const uavmvs_bvh * bvh, const uavmvs_cloud * cloud, uavmvs_obs * obs,
    uavmvs_array * recons, uavmvs_array * wrecons, float out3[3], uint64_t stats[4])
{
    UAV_REQUIRE(bvh && cloud && obs && recons && wrecons, "null handle");
    cudaStream_t s = (cudaStream_t)stream;
    if (ensure_scratch(0)) return 1;
    UAV_CUDA(cudaMemsetAsync(obs->num_rows, 0, 4 * (size_t)obs->num_cols, s));
    if (stats) UAV_CUDA(cudaMemsetAsync(g_scratch.stats, 0, 4 * sizeof(unsigned long long), s));
    g_scratch.ev_valid = false;
    UAV_CUDA(cudaEventRecord(g_scratch.ev[0], s));
    int rc = launch_k1(s, 1, d_views, num_views, max_distance, width, height, bvh, cloud, obs,
        stats ? g_scratch.stats : nullptr);
    if (rc) return rc;
    UAV_CUDA(cudaEventRecord(g_scratch.ev[1], s));
    rc = uavmvs_process_observation_rays(stream, obs);
    if (rc) return rc;
    UAV_CUDA(cudaEventRecord(g_scratch.ev[2], s));
    rc = launch_k3(s, obs, recons, target_recon, wrecons);
    if (rc) return rc;
    UAV_CUDA(cudaEventRecord(g_scratch.ev[3], s));
    if (launch_reduce(s, wrecons)) return 1;
    UAV_CUDA(cudaEventRecord(g_scratch.ev[4], s));
    g_scratch.ev_valid = true;
    if (stats) UAV_CUDA(cudaMemcpyAsync(g_scratch.h_stats, g_scratch.stats, 4 * sizeof(unsigned long long),
        cudaMemcpyDeviceToHost, s));
    if (out3 || stats) {
        UAV_CUDA(cudaStreamSynchronize(s));
        if (out3) { out3[0] = (float)g_scratch.h_acc->sum; out3[1] = ord2f(g_scratch.h_acc->mn); out3[2] = ord2f(g_scratch.h_acc->mx); }
        if (stats) {
            for (int i = 0; i < 4; ++i) stats[i] = g_scratch.h_stats[i];
            stats[2] = (uint64_t)cloud->n * num_views;
        }
    }
    return 0;
}

int uavmvs_profile_last(float ms[4])
{
    UAV_REQUIRE(ms, "null argument");
    UAV_REQUIRE(g_scratch.ev_valid, "no evaluate_trajectory call recorded on this thread");
    UAV_CUDA(cudaEventSynchronize(g_scratch.ev[4]));
    for (int i = 0; i < 4; ++i) UAV_CUDA(cudaEventElapsedTime(&ms[i], g_scratch.ev[i], g_scratch.ev[i + 1]));
    return 0;
}

int uavmvs_evaluate_trajectory(void * stream, const float * views, uint32_t num_views,
    float max_distance, float target_recon, int width, int height,
    const uavmvs_bvh * bvh, const uavmvs_cloud * cloud, uavmvs_obs * obs,
    uavmvs_array * recons, uavmvs_array * wrecons, float out3[3], uint64_t stats[4])
{
    UAV_REQUIRE(views || num_views == 0, "null views");
    cudaStream_t s = (cudaStream_t)stream;
    if (ensure_scratch((size_t)num_views * UAVMVS_VIEW_FLOATS)) return 1;
    if (num_views)
        UAV_CUDA(cudaMemcpyAsync(g_scratch.views, views, sizeof(float) * UAVMVS_VIEW_FLOATS * (size_t)num_views,
            cudaMemcpyHostToDevice, s));
    return uavmvs_evaluate_trajectory_device(stream, g_scratch.views, num_views, max_distance, target_recon,
        width, height, bvh, cloud, obs, recons, wrecons, out3, stats);
}

}  // extern "C"

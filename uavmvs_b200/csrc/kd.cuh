// KD-tree nearest neighbour on the device (libs/cacc/nnsearch.h:67-155), shared by the
// spherical-histogram kernels (hist.cu) and the batched histogram retire pass (k1.cu).
#pragma once
#include "common.cuh"

namespace uav {

constexpr int KD_STACK = 40;     // checked against the tree's depth at upload (containers.cu: kd_validate)

struct KdView { const uint4 * __restrict__ nodes; const float4 * __restrict__ verts; uint32_t n; };

// libs/cacc/nnsearch.h:67-155: depth-first nearest neighbour, strict '<' (first found wins)
__device__ __forceinline__ uint32_t kd_find_nn(const KdView & kd, V3 q)
{
    uint32_t idx = UAV_NAI; float max_dist = UAV_INF;
    uint32_t node_idx = 0; bool down = true;
    uint32_t stack[KD_STACK]; int sp = 0;
    while (true) {
        uint4 node = __ldg(kd.nodes + node_idx);          // vid, dim, left, right
        float4 v4 = __ldg(kd.verts + node.x);
        V3 vert = ld3(v4);
        float qd = node.y == 0 ? q.x : (node.y == 1 ? q.y : q.z);
        float vd = node.y == 0 ? vert.x : (node.y == 1 ? vert.y : vert.z);
        float diff = qd - vd;
        if (down) {
            float dist = norm(q - vert);
            if (dist < max_dist) { idx = node.x; max_dist = dist; }
            if (node.z != UAV_NAI || node.w != UAV_NAI) {
                if (node.z != UAV_NAI && node.w != UAV_NAI) stack[sp++] = node_idx;
                uint32_t next = (diff < 0.0f) ? node.z : node.w;
                uint32_t other = (diff < 0.0f) ? node.w : node.z;
                node_idx = (next != UAV_NAI) ? next : other;
            } else { down = false; node_idx = UAV_NAI; }
        } else {
            if (fabsf(diff) < max_dist) { down = true; node_idx = (diff < 0.0f) ? node.w : node.z; }
            else { down = false; node_idx = UAV_NAI; }
        }
        if (node_idx == UAV_NAI) { if (sp == 0) break; node_idx = stack[--sp]; }
    }
    return idx;
}

// Certified shortcut for kd_find_nn on queries near the unit sphere (the histogram kernels ask for the icosphere
// vertex nearest to a unit direction, ~25 node visits each).  A cube-map cell proposes a distinct position c; c is
// the reference's answer when (exact-arithmetic argument, margins three orders of magnitude above the fp32 rounding
// of the distances, which are a few 1e-8 here):
//   * d(q,c) + 1e-5 < d(q,u) for every u of c's neighbour list N(c) = all distinct positions closer than D(c) to c,
//   * d(q,c) < D(c)/2 - 1e-5: any position outside N(c) is then farther than D(c) - d(q,c) > d(q,c) from q.
// The reference search returns the FIRST vertex of minimal computed distance in its depth-first order.  With the margin
// above only the (bitwise coincident) duplicates of c attain the minimum; the icosphere stores most positions twice
// (SURVEY.md Appendix quirk 5).  Two duplicates are either ancestor and descendant in the KD-tree (the ancestor is met
// first on the way down) or lie on either side of a node A whose split plane passes through them; the walk enters
// the side `q[dim] - A[dim] < 0 ? left : right` first (nnsearch.h:118-121) and cannot prune the subtree holding the
// duplicate before reaching it, since every plane between q and c is closer to q than c is and the running bound is
// still above d(q,c) + 1e-5.  The rule {A, B, dim, coordinate} is precomputed per position (containers.cu).
// A closer neighbour becomes the new proposal (at most 3 hops); everything else takes the plain search.
__device__ __forceinline__ uint32_t kd_find_nn_fast(const KdFastView & f, const KdView & kd, V3 q)
{
    const float ax = fabsf(q.x), ay = fabsf(q.y), az = fabsf(q.z);
    uint32_t face; float ma, u, v;
    if (ax >= ay && ax >= az) { face = q.x < 0.0f ? 1u : 0u; ma = ax; u = q.y; v = q.z; }
    else if (ay >= az) { face = q.y < 0.0f ? 3u : 2u; ma = ay; u = q.x; v = q.z; }
    else { face = q.z < 0.0f ? 5u : 4u; ma = az; u = q.x; v = q.y; }
    if (!(ma > 1e-20f) || !(ma < 1e20f)) return kd_find_nn(kd, q);          // zero, huge or NaN query
    const float inv = 0.5f * (float)f.res / ma, half = 0.5f * (float)f.res;
    const int iu = min((int)f.res - 1, max(0, (int)(u * inv + half)));
    const int iv = min((int)f.res - 1, max(0, (int)(v * inv + half)));
    uint32_t c = f.table[(face * f.res + (uint32_t)iv) * f.res + (uint32_t)iu];
    for (int hop = 0; hop < 3; ++hop) {
        const float4 p = __ldg(f.pos + c);
        const float d = norm(q - ld3(p));
        float dn = UAV_INF; uint32_t cn = c;
        const uint4 nb = __ldg(reinterpret_cast<const uint4 *>(f.nbr) + c);     // 8 x uint16
        const uint32_t w[4] = { nb.x, nb.y, nb.z, nb.w };
        #pragma unroll
        for (int k = 0; k < 8; ++k) {
            const uint32_t n = (w[k >> 1] >> ((k & 1) * 16)) & 0xffffu;
            if (n != 0xffffu) {
                const float dk = norm(q - ld3(__ldg(f.pos + n)));
                if (dk < dn) { dn = dk; cn = n; }
            }
        }
        if (d + 1e-5f < dn) {
            if (!(d < p.w)) break;
            const uint4 r = __ldg(f.rule + c);
            if (r.z == 3u) return r.x;
            const float qd = r.z == 0 ? q.x : (r.z == 1 ? q.y : q.z);
            return (qd - __uint_as_float(r.w) < 0.0f) ? r.x : r.y;
        }
        if (!(dn < d)) break;                      // inside the margin of a bisector (or NaN)
        c = cn;
    }
    return kd_find_nn(kd, q);
}

static inline KdView kd_view(const uavmvs_kdtree * k) { KdView v; v.nodes = k->nodes; v.verts = k->verts; v.n = k->num_nodes; return v; }

}  // namespace uav

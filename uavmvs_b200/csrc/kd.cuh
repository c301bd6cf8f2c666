// KD-tree nearest neighbour on the device (libs/cacc/nnsearch.h:67-155), shared by the
// spherical-histogram kernels (hist.cu) and the batched histogram retire pass (k1.cu).
#pragma once
#include "common.cuh"

namespace uav {

struct KdView { const uint4 * __restrict__ nodes; const float4 * __restrict__ verts; uint32_t n; };

// libs/cacc/nnsearch.h:67-155: depth-first nearest neighbour, strict '<' (first found wins)
__device__ __forceinline__ uint32_t kd_find_nn(const KdView & kd, V3 q)
{
    uint32_t idx = UAV_NAI; float max_dist = UAV_INF;
    uint32_t node_idx = 0; bool down = true;
    uint32_t stack[40]; int sp = 0;
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

static inline KdView kd_view(const uavmvs_kdtree * k) { KdView v; v.nodes = k->nodes; v.verts = k->verts; v.n = k->num_nodes; return v; }

}  // namespace uav

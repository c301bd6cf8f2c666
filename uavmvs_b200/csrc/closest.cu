// Closest point on the mesh: acc::BVHTree::closest_point (libs/acc/bvh_tree.h:469-548) with the box and
// triangle primitives of libs/acc/primitives.h:155-163, :174-212.  The reference runs it on the CPU inside
// the optimiser's objective (apps/optimize_trajectory/optimize_trajectory.cu:370-386, :456: altitude /
// airspace test of every Nelder-Mead candidate) and in evaluate_reconstruction / trajectory_tools; here it
// is a batched device query so that the candidate test stays next to the batched histogram kernels.
// One thread per query, the reference's visiting order (nearer child first, ties to the right child,
// strict '<' everywhere) and operation order: found flag, triangle index and point are bit-identical.
#include "common.cuh"
#include "trace.cuh"

namespace uav {

__device__ __forceinline__ float sqdist(V3 a, V3 b) { V3 d = a - b; return dot(d, d); }

__device__ __forceinline__ V3 closest_on_box(V3 v, float mnx, float mny, float mnz, float mxx, float mxy, float mxz)
{
    return mk(smax(mnx, smin(v.x, mxx)), smax(mny, smin(v.y, mxy)), smax(mnz, smin(v.z, mxz)));
}

// libs/acc/primitives.h:174-212 on the raw triangle (9 floats)
__device__ __forceinline__ V3 closest_on_tri(V3 vertex, const float * __restrict__ t9)
{
    V3 a = mk(__ldg(t9), __ldg(t9 + 1), __ldg(t9 + 2)), b = mk(__ldg(t9 + 3), __ldg(t9 + 4), __ldg(t9 + 5));
    V3 c = mk(__ldg(t9 + 6), __ldg(t9 + 7), __ldg(t9 + 8));
    V3 ab = b - a, ac = c - a;
    V3 normal = cross(ab, ac);
    float n = norm(normal);
    normal = normal / n;
    V3 p = vertex - normal * dot(normal, vertex - a);
    V3 ap = p - a;
    float d00 = dot(ab, ab), d01 = dot(ab, ac), d11 = dot(ac, ac);
    float d20 = dot(ap, ab), d21 = dot(ap, ac);
    float denom = d00 * d11 - d01 * d01;         // reference: double of this float; float / double rounds the same
    float b1 = (d11 * d20 - d01 * d21) / denom;
    float b2 = (d00 * d21 - d01 * d20) / denom;
    float b0 = 1.0f - b1 - b2;
    if (b0 < 0.0f) {
        V3 bc = c - b;
        float len = norm(bc);
        float t = smax(0.0f, smin(dot(bc, p - b) / len, len));
        return b + bc * (t / len);
    }
    if (b1 < 0.0f) {
        V3 ca = a - c;
        float len = norm(ca);
        float t = smax(0.0f, smin(dot(ca, p - c) / len, len));
        return c + ca * (t / len);
    }
    if (b2 < 0.0f) {
        float len = norm(ab);
        float t = smax(0.0f, smin(dot(ab, p - a) / len, len));
        return a + ab * (t / len);
    }
    return (a * b0 + b * b1) + c * b2;
}

__global__ void __launch_bounds__(128)
k_closest_point(BvhView bvh, const float * __restrict__ points, uint32_t n, float max_dist,
    uint8_t * __restrict__ found, uint32_t * __restrict__ idx_out, float * __restrict__ cp_out)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const V3 vertex = mk(points[3 * (size_t)i], points[3 * (size_t)i + 1], points[3 * (size_t)i + 2]);
    float dist = max_dist * max_dist;
    uint32_t idx = UAV_NAI; V3 cp = mk(0.0f, 0.0f, 0.0f);
    uint2 stack[UAV_STACK]; int sp = 0;
    // (ref, count): count == 0 -> inner node `ref`; count > 0 -> leaf holding triangles [ref, ref + count)
    uint2 cur = bvh.root_is_leaf ? make_uint2(bvh.root_first, bvh.root_count) : make_uint2(0u, 0u);
    for (;;) {
        bool pop = false;
        if (cur.y == 0) {
            const Node64 * np = bvh.inner + cur.x;
            const float4 q0 = __ldg(&np->q0), q1 = __ldg(&np->q1), q2 = __ldg(&np->q2);
            const uint4 q3 = __ldg(&np->q3);
            const float dl = sqdist(closest_on_box(vertex, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y), vertex);
            const float dr = sqdist(closest_on_box(vertex, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w), vertex);
            const bool left = dl < dist, right = dr < dist;
            const uint2 L = make_uint2(q3.x, q3.z), R = make_uint2(q3.y, q3.w);
            if (left && right) {
                if (dl < dr) { stack[sp++] = R; cur = L; } else { stack[sp++] = L; cur = R; }
            } else if (left) cur = L;
            else if (right) cur = R;
            else pop = true;
        } else {
            uint32_t idx_leaf = UAV_NAI; V3 cp_leaf = mk(0.0f, 0.0f, 0.0f); float d_leaf = UAV_INF;
            for (uint32_t t = cur.x; t < cur.x + cur.y; ++t) {
                const V3 c = closest_on_tri(vertex, bvh.raw_tris + 9 * (size_t)t);
                const float d = sqdist(c, vertex);
                if (d < d_leaf) { cp_leaf = c; d_leaf = d; idx_leaf = t; }
            }
            if (idx_leaf != UAV_NAI && d_leaf < dist) { dist = d_leaf; idx = idx_leaf; cp = cp_leaf; }
            pop = true;
        }
        if (pop) { if (sp == 0) break; cur = stack[--sp]; }
    }
    const bool f = idx != UAV_NAI;
    found[i] = f ? 1 : 0; idx_out[i] = idx;
    cp_out[3 * (size_t)i] = f ? cp.x : 0.0f; cp_out[3 * (size_t)i + 1] = f ? cp.y : 0.0f; cp_out[3 * (size_t)i + 2] = f ? cp.z : 0.0f;
}

}  // namespace uav

using namespace uav;

int uav::launch_closest_point(cudaStream_t s, const uavmvs_bvh * bvh, const float * d_points, uint32_t n, float max_dist,
    uint8_t * d_found, uint32_t * d_tri_index, float * d_closest)
{
    UAV_REQUIRE(bvh && d_points && d_found && d_tri_index && d_closest, "null argument");
    if (n == 0) return 0;
    k_closest_point<<<divup(n, 128), 128, 0, uav::counted(s)>>>(view_of(bvh), d_points, n, max_dist, d_found, d_tri_index, d_closest);
    UAV_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int uavmvs_bvh_closest_point(void * stream, const uavmvs_bvh * bvh, const float * points, uint64_t num_points,
    float max_dist, uint8_t * found, uint32_t * tri_index, float * closest)
{
    UAV_REQUIRE(bvh && points && found && tri_index && closest, "null argument");
    UAV_REQUIRE(num_points < (1ull << 32), "too many queries");
    UAV_REQUIRE(bvh->max_leaf < (1u << 31), "leaf too large");
    if (num_points == 0) return 0;
    cudaStream_t s = (cudaStream_t)stream;
    const uint32_t n = (uint32_t)num_points;
    float * d_pts, * d_cp; uint8_t * d_found; uint32_t * d_idx;
    UAV_CUDA(cudaMallocAsync(&d_pts, sizeof(float) * 3 * (size_t)n, s));
    UAV_CUDA(cudaMallocAsync(&d_cp, sizeof(float) * 3 * (size_t)n, s));
    UAV_CUDA(cudaMallocAsync(&d_found, (size_t)n, s));
    UAV_CUDA(cudaMallocAsync(&d_idx, sizeof(uint32_t) * (size_t)n, s));
    UAV_CUDA(cudaMemcpyAsync(d_pts, points, sizeof(float) * 3 * (size_t)n, cudaMemcpyHostToDevice, s));
    k_closest_point<<<divup(n, 128), 128, 0, uav::counted(s)>>>(view_of(bvh), d_pts, n, max_dist, d_found, d_idx, d_cp);
    UAV_CUDA(cudaGetLastError());
    UAV_CUDA(cudaMemcpyAsync(found, d_found, (size_t)n, cudaMemcpyDeviceToHost, s));
    UAV_CUDA(cudaMemcpyAsync(tri_index, d_idx, sizeof(uint32_t) * (size_t)n, cudaMemcpyDeviceToHost, s));
    UAV_CUDA(cudaMemcpyAsync(closest, d_cp, sizeof(float) * 3 * (size_t)n, cudaMemcpyDeviceToHost, s));
    UAV_CUDA(cudaStreamSynchronize(s));
    cudaFreeAsync(d_pts, s); cudaFreeAsync(d_cp, s); cudaFreeAsync(d_found, s); cudaFreeAsync(d_idx, s);
    return 0;
}

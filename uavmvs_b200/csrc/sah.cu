// GPU full-sweep SAH build.  Same objective and termination rule as the reference's host builder
// (libs/acc/bvh_tree.h:277-333: cost = 1.2 + (A_l/A·n_l + A_r/A·n_r) over every split position of every
// axis, leaf when the best cost is not below n), but evaluated for ALL nodes of a tree level at once with
// data-parallel primitives instead of a recursive per-node sort:
//
//   * the triangles are sorted ONCE by box centre along x, y and z (three radix sorts);
//   * all three orders are kept partitioned by node: a node owns the same position range [first, last) in
//     each, sorted along that order's axis.  One segmented prefix scan and one segmented suffix scan of
//     the boxes (segments = nodes) over the 3N entries give, for every split position of every axis of
//     every node, the exact boxes of both sides; one 64-bit atomicMin per warp and node picks the best;
//   * splitting is a stable segmented partition of the three orders (one scan of the "goes left" flags),
//     which keeps each of them sorted, so the next level needs no sort at all.
//
// Per level: ~12 launches streaming O(N) data; a 10^6-triangle scene is built in a few tens of ms.
// The reference sweeps only nodes of <= 64 triangles and bins larger ones (64 bins); sweeping everywhere
// is the same cost model evaluated more finely.  The result is a valid reference-layout tree (nodes,
// aabbs, tris, indices) and goes through the same finalize_layout as every other tree.
#include <cub/cub.cuh>
#include <thrust/iterator/reverse_iterator.h>

#include <cfloat>
#include <cstdlib>
#include "common.cuh"

namespace uav {

int finalize_layout(uavmvs_bvh * b);
int alloc_reference_arrays(uavmvs_bvh * b);

struct SegBox { float mn[3], mx[3]; uint32_t seg; };   // seg = node id | axis << 28, NAI = inactive

struct SegUnion {
    __device__ __forceinline__ SegBox operator()(const SegBox & a, const SegBox & b) const {
        if (a.seg != b.seg) return b;
        SegBox r;
        for (int k = 0; k < 3; ++k) { r.mn[k] = fminf(a.mn[k], b.mn[k]); r.mx[k] = fmaxf(a.mx[k], b.mx[k]); }
        r.seg = b.seg;
        return r;
    }
};

__device__ __forceinline__ float seg_area(const SegBox & b) {
    float e0 = b.mx[0] - b.mn[0], e1 = b.mx[1] - b.mn[1], e2 = b.mx[2] - b.mn[2];
    return 2.0f * (e0 * e1 + e1 * e2 + e2 * e0);
}

constexpr uint32_t SAH_AXIS_SHIFT = 28;            // node ids < 2^28 (triangles < 2^25 anyway)
constexpr uint32_t SAH_LEAF = 0xfffffffeu;        // split[] value: node became a leaf this level

__global__ void k_sah_boxes(const float * __restrict__ verts, const uint32_t * __restrict__ faces, uint32_t n,
    float * __restrict__ boxes, float * __restrict__ kx, float * __restrict__ ky, float * __restrict__ kz,
    uint32_t * __restrict__ ids)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float mn[3] = {UAV_INF, UAV_INF, UAV_INF}, mx[3] = {-UAV_INF, -UAV_INF, -UAV_INF};
    for (int v = 0; v < 3; ++v) {
        const float * p = verts + 3 * (size_t)faces[3 * (size_t)i + v];
        for (int k = 0; k < 3; ++k) { mn[k] = fminf(mn[k], p[k]); mx[k] = fmaxf(mx[k], p[k]); }
    }
    for (int k = 0; k < 3; ++k) { boxes[6 * (size_t)i + k] = mn[k]; boxes[6 * (size_t)i + 3 + k] = mx[k]; }
    kx[i] = (mn[0] + mx[0]) / 2.0f; ky[i] = (mn[1] + mx[1]) / 2.0f; kz[i] = (mn[2] + mx[2]) / 2.0f;   // primitives.h:108-112
    ids[i] = i;
}

// in[a][p] = box of the triangle at position p of order a, tagged with its node
__global__ void k_sah_gather(const uint32_t * __restrict__ ord, const uint32_t * __restrict__ seg,
    const float * __restrict__ boxes, uint32_t n, SegBox * __restrict__ in)
{
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x, a = blockIdx.y;
    if (p >= n) return;
    uint32_t s = seg[p];
    SegBox o;
    const float * b = boxes + 6 * (size_t)ord[(size_t)a * n + p];
    for (int k = 0; k < 3; ++k) { o.mn[k] = b[k]; o.mx[k] = b[3 + k]; }
    o.seg = (s == UAV_NAI) ? UAV_NAI : (s | (a << SAH_AXIS_SHIFT));
    in[(size_t)a * n + p] = o;
}

// best[node] = min over (axis, split position) of (cost, ~(axis, position)): the reference's `cost <= min_cost`
// keeps the LAST best candidate in (axis, position) order (bvh_tree.h:318), hence the complement.
__global__ void k_sah_cost(const SegBox * __restrict__ L, const SegBox * __restrict__ R,
    const uint32_t * __restrict__ seg, const uint4 * __restrict__ nodes, uint32_t n, float ct,
    unsigned long long * __restrict__ best)
{
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x, a = blockIdx.y;
    uint32_t s = (p < n) ? seg[p] : UAV_NAI;
    unsigned long long key = ~0ull;
    if (s != UAV_NAI) {
        uint4 nd = nodes[s];
        if (p > nd.x) {
            const SegBox * La = L + (size_t)a * n, * Ra = R + (size_t)a * n;
            float area = seg_area(La[nd.y - 1]);
            float nl = (float)(p - nd.x), nr = (float)(nd.y - p);
            float cost = ct + (seg_area(La[p - 1]) / area * nl + seg_area(Ra[p]) / area * nr);
            if (!(cost >= 0.0f) || cost > FLT_MAX) cost = FLT_MAX;           // point-like node boxes: 0/0
            key = ((unsigned long long)__float_as_uint(cost) << 32) | (uint32_t)~((a << 30) | p);
        }
    }
    // one atomic per warp when the whole warp sits in one node (the common case high in the tree)
    uint32_t s0 = __shfl_sync(0xffffffffu, s, 0);
    if (__all_sync(0xffffffffu, s == s0)) {
        for (int o = 16; o > 0; o >>= 1) {
            unsigned long long other = __shfl_xor_sync(0xffffffffu, key, o);
            key = other < key ? other : key;
        }
        if ((threadIdx.x & 31) == 0 && key != ~0ull) atomicMin(best + s0, key);
    } else if (key != ~0ull) atomicMin(best + s, key);
}

// one thread per active node: leaf or split?  Also stores the node's own box.
__global__ void k_sah_decide(const uint32_t * __restrict__ active, uint32_t count, const SegBox * __restrict__ L,
    uint32_t n, const unsigned long long * __restrict__ best, const uint4 * __restrict__ nodes,
    float * __restrict__ aabbs, uint32_t max_leaf, uint32_t * __restrict__ split, uint32_t * __restrict__ flag)
{
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= count) return;
    uint32_t id = active[k];
    uint4 nd = nodes[id];
    uint32_t cnt = nd.y - nd.x;
    const SegBox & box = L[nd.y - 1];                        // order 0: prefix over the whole node
    for (int j = 0; j < 3; ++j) { aabbs[6 * (size_t)id + j] = box.mn[j]; aabbs[6 * (size_t)id + 3 + j] = box.mx[j]; }
    unsigned long long b = best[id];
    uint32_t sp = SAH_LEAF;
    if (cnt > 1) {
        float cost = __uint_as_float((uint32_t)(b >> 32));
        uint32_t ap = ~(uint32_t)b;                          // axis << 30 | position
        if (cost < (float)cnt) sp = ap;                      // bvh_tree.h:321 `if (min_cost >= n) return false`
        else if (cnt > max_leaf) sp = (cost < FLT_MAX) ? ap : (nd.x + cnt / 2);   // never keep an oversized leaf
    }
    split[id] = sp;
    flag[k] = (sp != SAH_LEAF) ? 1u : 0u;
}

// children of the splitting nodes, ids handed out in active-list order (deterministic)
__global__ void k_sah_children(const uint32_t * __restrict__ active, uint32_t count, const uint32_t * __restrict__ flag,
    const uint32_t * __restrict__ rank, const uint32_t * __restrict__ split, uint32_t next_id,
    uint4 * __restrict__ nodes, unsigned long long * __restrict__ best, uint32_t * __restrict__ next_active)
{
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= count || !flag[k]) return;
    uint32_t id = active[k];
    uint4 nd = nodes[id];
    uint32_t pos = split[id] & 0x3fffffffu;
    uint32_t lc = next_id + 2 * rank[k], rc = lc + 1;
    nodes[lc] = make_uint4(nd.x, pos, UAV_NAI, UAV_NAI);
    nodes[rc] = make_uint4(pos, nd.y, UAV_NAI, UAV_NAI);
    nd.z = lc; nd.w = rc;
    nodes[id] = nd;
    best[lc] = ~0ull; best[rc] = ~0ull;
    next_active[2 * rank[k]] = lc; next_active[2 * rank[k] + 1] = rc;
}

// side[t] = 1 when triangle t goes to the right child (its position in the chosen axis' order >= split)
__global__ void k_sah_side(const uint32_t * __restrict__ ord, const uint32_t * __restrict__ seg,
    const uint32_t * __restrict__ split, uint32_t n, uint8_t * __restrict__ side)
{
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    uint32_t s = seg[p];
    if (s == UAV_NAI) return;
    uint32_t sp = split[s];
    if (sp == SAH_LEAF) return;
    uint32_t a = sp >> 30, pos = sp & 0x3fffffffu;
    side[ord[(size_t)a * n + p]] = (p >= pos) ? 1 : 0;
}

__global__ void k_sah_left_flags(const uint32_t * __restrict__ ord, const uint32_t * __restrict__ seg,
    const uint32_t * __restrict__ split, const uint8_t * __restrict__ side, uint32_t n, uint32_t * __restrict__ left)
{
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x, a = blockIdx.y;
    if (p >= n) return;
    uint32_t s = seg[p], f = 0;
    if (s != UAV_NAI && split[s] != SAH_LEAF) f = side[ord[(size_t)a * n + p]] ? 0u : 1u;
    left[(size_t)a * n + p] = f;
}

// stable partition of every splitting node's range in all three orders
__global__ void k_sah_scatter(const uint32_t * __restrict__ ord, const uint32_t * __restrict__ seg,
    const uint32_t * __restrict__ split, const uint4 * __restrict__ nodes, const uint32_t * __restrict__ left,
    const uint32_t * __restrict__ lscan, uint32_t n, uint32_t * __restrict__ ord_out)
{
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x, a = blockIdx.y;
    if (p >= n) return;
    size_t base = (size_t)a * n;
    uint32_t s = seg[p], t = ord[base + p], q = p;
    if (s != UAV_NAI) {
        uint32_t sp = split[s];
        if (sp != SAH_LEAF) {
            uint32_t first = nodes[s].x, pos = sp & 0x3fffffffu;
            uint32_t lbefore = lscan[base + p] - lscan[base + first];
            q = left[base + p] ? first + lbefore : pos + ((p - first) - lbefore);
        }
    }
    ord_out[base + q] = t;
}

__global__ void k_sah_reseg(const uint32_t * __restrict__ split, const uint4 * __restrict__ nodes, uint32_t n,
    uint32_t * __restrict__ seg)
{
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    uint32_t s = seg[p];
    if (s == UAV_NAI) return;
    uint32_t sp = split[s];
    if (sp == SAH_LEAF) { seg[p] = UAV_NAI; return; }
    uint4 nd = nodes[s];                                     // children already linked
    seg[p] = (p < (sp & 0x3fffffffu)) ? nd.z : nd.w;
}

__global__ void k_sah_emit_tris(const float * __restrict__ verts, const uint32_t * __restrict__ faces,
    const uint32_t * __restrict__ ord, uint32_t n, float * __restrict__ tris, uint32_t * __restrict__ indices)
{
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    uint32_t t = ord[p];
    indices[p] = t;
    for (int v = 0; v < 3; ++v) {
        const float * q = verts + 3 * (size_t)faces[3 * (size_t)t + v];
        for (int k = 0; k < 3; ++k) tris[9 * (size_t)p + 3 * v + k] = q[k];
    }
}

}  // namespace uav

using namespace uav;

extern "C" int uavmvs_bvh_build_sah(const float * verts, uint32_t num_verts, const uint32_t * faces,
    uint32_t num_faces, uavmvs_bvh ** out)
{
    UAV_REQUIRE(verts && faces && out, "null argument");
    UAV_REQUIRE(num_faces >= 1 && num_verts >= 1, "empty mesh");
    UAV_REQUIRE(num_faces < (1u << 25), "too many triangles");
    const uint32_t n = num_faces;
    float ct = 1.2f;                                          // traversal cost of the reference's builder
    uint32_t max_leaf = 8;
    if (const char * e = getenv("UAVMVS_SAH_CT")) ct = (float)atof(e);
    if (const char * e = getenv("UAVMVS_SAH_MAX_LEAF")) max_leaf = (uint32_t)atoi(e);
    if (max_leaf < 1) max_leaf = 1;
    if (max_leaf > 31) max_leaf = 31;
    uavmvs_bvh * b = new uavmvs_bvh();
    b->num_nodes = 2 * n - 1; b->num_tris = n;
    if (alloc_reference_arrays(b)) return 1;
    float * d_verts, * boxes, * keys, * keys2; uint32_t * d_faces, * ids, * ord, * ord2, * seg, * split, * left, * lscan;
    uint32_t * active, * next_active, * flag, * rank; uint8_t * side; SegBox * in, * L, * R; unsigned long long * best;
    UAV_CUDA(cudaMalloc(&d_verts, sizeof(float) * 3 * (size_t)num_verts));
    UAV_CUDA(cudaMalloc(&d_faces, sizeof(uint32_t) * 3 * (size_t)n));
    UAV_CUDA(cudaMalloc(&boxes, sizeof(float) * 6 * (size_t)n));
    UAV_CUDA(cudaMalloc(&keys, sizeof(float) * 3 * (size_t)n)); UAV_CUDA(cudaMalloc(&keys2, sizeof(float) * (size_t)n));
    UAV_CUDA(cudaMalloc(&ids, 4 * (size_t)n));
    UAV_CUDA(cudaMalloc(&ord, 12 * (size_t)n)); UAV_CUDA(cudaMalloc(&ord2, 12 * (size_t)n));
    UAV_CUDA(cudaMalloc(&left, 12 * (size_t)n)); UAV_CUDA(cudaMalloc(&lscan, 12 * (size_t)n));
    UAV_CUDA(cudaMalloc(&seg, 4 * (size_t)n)); UAV_CUDA(cudaMalloc(&side, (size_t)n));
    UAV_CUDA(cudaMalloc(&split, 4 * (size_t)b->num_nodes)); UAV_CUDA(cudaMalloc(&best, 8 * (size_t)b->num_nodes));
    UAV_CUDA(cudaMalloc(&active, 4 * (size_t)n)); UAV_CUDA(cudaMalloc(&next_active, 4 * (size_t)n));
    UAV_CUDA(cudaMalloc(&flag, 4 * (size_t)n)); UAV_CUDA(cudaMalloc(&rank, 4 * (size_t)n));
    UAV_CUDA(cudaMalloc(&in, sizeof(SegBox) * 3 * (size_t)n)); UAV_CUDA(cudaMalloc(&L, sizeof(SegBox) * 3 * (size_t)n));
    UAV_CUDA(cudaMalloc(&R, sizeof(SegBox) * 3 * (size_t)n));
    UAV_CUDA(cudaMemcpy(d_verts, verts, sizeof(float) * 3 * (size_t)num_verts, cudaMemcpyHostToDevice));
    UAV_CUDA(cudaMemcpy(d_faces, faces, sizeof(uint32_t) * 3 * (size_t)n, cudaMemcpyHostToDevice));
    k_sah_boxes<<<divup(n, 256), 256>>>(d_verts, d_faces, n, boxes, keys, keys + n, keys + 2 * (size_t)n, ids);

    const int n3 = (int)(3 * n);
    thrust::reverse_iterator<SegBox *> rin(in + n3), rout(R + n3);
    size_t sort_bytes = 0, scan_bytes = 0, rscan_bytes = 0, sum_bytes = 0; void * tmp = nullptr;
    cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, keys, keys2, ids, ord, (int)n);
    cub::DeviceScan::InclusiveScan(nullptr, scan_bytes, in, L, SegUnion(), n3);
    cub::DeviceScan::InclusiveScan(nullptr, rscan_bytes, rin, rout, SegUnion(), n3);
    cub::DeviceScan::ExclusiveSum(nullptr, sum_bytes, left, lscan, n3);
    size_t tmp_bytes = sort_bytes;
    if (scan_bytes > tmp_bytes) tmp_bytes = scan_bytes;
    if (rscan_bytes > tmp_bytes) tmp_bytes = rscan_bytes;
    if (sum_bytes > tmp_bytes) tmp_bytes = sum_bytes;
    UAV_CUDA(cudaMalloc(&tmp, tmp_bytes));
    for (int a = 0; a < 3; ++a)     // stable: equal centres keep ascending triangle ids, like the reference's comparator
        cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, keys + (size_t)a * n, keys2, ids, ord + (size_t)a * n, (int)n);

    // root
    uint4 root = make_uint4(0u, n, UAV_NAI, UAV_NAI);
    unsigned long long none = ~0ull; uint32_t zero = 0;
    UAV_CUDA(cudaMemcpy(b->nodes, &root, sizeof(root), cudaMemcpyHostToDevice));
    UAV_CUDA(cudaMemcpy(best, &none, 8, cudaMemcpyHostToDevice));
    UAV_CUDA(cudaMemcpy(active, &zero, 4, cudaMemcpyHostToDevice));
    UAV_CUDA(cudaMemset(seg, 0, 4 * (size_t)n));
    uint32_t count = 1, next_id = 1, levels = 0;
    const dim3 g3(divup(n, 256), 3);
    while (count > 0) {
        k_sah_gather<<<g3, 256>>>(ord, seg, boxes, n, in);
        cub::DeviceScan::InclusiveScan(tmp, tmp_bytes, in, L, SegUnion(), n3);
        cub::DeviceScan::InclusiveScan(tmp, tmp_bytes, rin, rout, SegUnion(), n3);
        k_sah_cost<<<g3, 256>>>(L, R, seg, b->nodes, n, ct, best);
        k_sah_decide<<<divup(count, 256), 256>>>(active, count, L, n, best, b->nodes, b->aabbs, max_leaf, split, flag);
        cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, flag, rank, (int)count);
        uint32_t last[2];
        UAV_CUDA(cudaMemcpy(&last[0], flag + (count - 1), 4, cudaMemcpyDeviceToHost));
        UAV_CUDA(cudaMemcpy(&last[1], rank + (count - 1), 4, cudaMemcpyDeviceToHost));
        const uint32_t m = last[0] + last[1];
        if (m) {
            UAV_REQUIRE(next_id + 2 * m <= b->num_nodes, "SAH build: node count overflow");
            k_sah_children<<<divup(count, 256), 256>>>(active, count, flag, rank, split, next_id, b->nodes, best, next_active);
            k_sah_side<<<divup(n, 256), 256>>>(ord, seg, split, n, side);
            k_sah_left_flags<<<g3, 256>>>(ord, seg, split, side, n, left);
            cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, left, lscan, n3);
            k_sah_scatter<<<g3, 256>>>(ord, seg, split, b->nodes, left, lscan, n, ord2);
            uint32_t * t = ord; ord = ord2; ord2 = t;
        }
        k_sah_reseg<<<divup(n, 256), 256>>>(split, b->nodes, n, seg);
        UAV_CUDA(cudaGetLastError());
        next_id += 2 * m;
        count = 2 * m;
        uint32_t * t = active; active = next_active; next_active = t;
        UAV_REQUIRE(++levels < 4096, "SAH build: runaway depth");
    }
    k_sah_emit_tris<<<divup(n, 256), 256>>>(d_verts, d_faces, ord, n, b->tris, b->indices);
    UAV_CUDA(cudaGetLastError());
    UAV_CUDA(cudaDeviceSynchronize());
    b->num_nodes = next_id;
    cudaFree(tmp); cudaFree(d_verts); cudaFree(d_faces); cudaFree(boxes); cudaFree(keys); cudaFree(keys2); cudaFree(ids);
    cudaFree(ord); cudaFree(ord2); cudaFree(left); cudaFree(lscan); cudaFree(seg); cudaFree(side); cudaFree(split);
    cudaFree(best); cudaFree(active); cudaFree(next_active); cudaFree(flag); cudaFree(rank); cudaFree(in); cudaFree(L); cudaFree(R);
    if (finalize_layout(b)) return 1;
    *out = b;
    return 0;
}

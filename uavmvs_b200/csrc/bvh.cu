// BVH handles: upload of an acc-built tree (reference topology and boxes, bit for bit),
// GPU LBVH build, conversion into the 64-byte traversal layouts, export, batch trace.
#include <cub/cub.cuh>

#include <cuda_fp16.h>
#include "common.cuh"
#include "trace.cuh"

namespace uav {

static thread_local std::string g_error;

void set_error(const char * file, int line, const char * what, const char * detail) {
    char buf[1024];
    snprintf(buf, sizeof(buf), "%s:%d (%s): %s", file, line, what, detail ? detail : "");
    g_error = buf;
}

// ------------------------------------------------------------------ conversion
__global__ void k_flag_inner(const uint4 * __restrict__ nodes, uint32_t n, uint32_t * __restrict__ flag)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint4 nd = nodes[i];
    flag[i] = (nd.z != UAV_NAI && nd.w != UAV_NAI) ? 1u : 0u;
}

__global__ void k_build_inner(const uint4 * __restrict__ nodes, const float * __restrict__ aabbs,
    const uint32_t * __restrict__ flag, const uint32_t * __restrict__ slot, uint32_t n,
    Node64 * __restrict__ inner)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || !flag[i]) return;
    uint4 nd = nodes[i];
    uint32_t l = nd.z, r = nd.w;
    const float * bl = aabbs + 6 * (size_t)l;
    const float * br = aabbs + 6 * (size_t)r;
    Node64 o;
    o.q0 = make_float4(bl[0], bl[1], bl[2], bl[3]);
    o.q1 = make_float4(bl[4], bl[5], br[0], br[1]);
    o.q2 = make_float4(br[2], br[3], br[4], br[5]);
    uint4 ln = nodes[l], rn = nodes[r];
    uint4 q3;
    if (flag[l]) { q3.x = slot[l]; q3.z = 0u; } else { q3.x = ln.x; q3.z = ln.y - ln.x; }
    if (flag[r]) { q3.y = slot[r]; q3.w = 0u; } else { q3.y = rn.x; q3.w = rn.y - rn.x; }
    o.q3 = q3;
    inner[slot[i]] = o;
}

// 4-wide collapse, top down and breadth first; only reachable wide nodes are stored, in BFS order.
// mode 0 (default): a wide node holds the grandchildren of its binary node (a leaf child stays as it is).
// mode 1 (UAVMVS_COLLAPSE=1): start from the two children and keep replacing the inner child of largest
// surface area by its two children until there are four; measured slightly slower on config 2.
// (Any-hit results do not depend on the collapse: child boxes are the binary nodes' exact boxes.)
__device__ __forceinline__ float box_area(const float * b) {
    float e0 = b[3] - b[0], e1 = b[4] - b[1], e2 = b[5] - b[2];
    return e0 * e1 + e1 * e2 + e2 * e0;
}

__global__ void k_wide_expand(const uint32_t * __restrict__ list, uint32_t count, const uint4 * __restrict__ nodes,
    const float * __restrict__ aabbs, const uint32_t * __restrict__ flag, int mode, uint4 * __restrict__ kids_out,
    uint32_t * __restrict__ ninner)
{
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= count) return;
    uint4 nd = nodes[list[k]];
    uint32_t kids[4] = {nd.z, nd.w, UAV_NAI, UAV_NAI};
    int nk = 2;
    if (mode == 0) {            // two binary levels per wide node
        uint32_t l = nd.z, r = nd.w; nk = 0;
        if (flag[l]) { uint4 cn = nodes[l]; kids[nk++] = cn.z; kids[nk++] = cn.w; } else kids[nk++] = l;
        if (flag[r]) { uint4 cn = nodes[r]; kids[nk++] = cn.z; kids[nk++] = cn.w; } else kids[nk++] = r;
        for (int j = nk; j < 4; ++j) kids[j] = UAV_NAI;
    }
    while (mode != 0 && nk < 4) {
        int best = -1; float best_area = -1.0f;
        for (int j = 0; j < nk; ++j)
            if (flag[kids[j]]) {
                float a = box_area(aabbs + 6 * (size_t)kids[j]);
                if (!(a <= best_area)) { best_area = a; best = j; }    // NaN areas still get opened
            }
        if (best < 0) break;
        uint4 cn = nodes[kids[best]];
        kids[best] = cn.z; kids[nk++] = cn.w;
    }
    uint32_t ni = 0;
    for (int j = 0; j < nk; ++j) ni += flag[kids[j]];
    kids_out[k] = make_uint4(kids[0], kids[1], kids[2], kids[3]);
    ninner[k] = ni;
}

__global__ void k_wide_emit(uint32_t count, const uint4 * __restrict__ kids_in, const uint32_t * __restrict__ off,
    uint32_t level_base, uint32_t next_base, const uint4 * __restrict__ nodes, const float * __restrict__ aabbs,
    const uint32_t * __restrict__ flag, Node128 * __restrict__ wide, uint32_t * __restrict__ next_list)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    uint4 kk = kids_in[i];
    uint32_t kids[4] = {kk.x, kk.y, kk.z, kk.w};
    float b[24]; uint32_t ref[4], cnt[4];
    uint32_t o_in = off[i];
    for (int k = 0; k < 4; ++k) {
        uint32_t ch = kids[k];
        if (ch != UAV_NAI) {
            for (int j = 0; j < 6; ++j) b[6 * k + j] = aabbs[6 * (size_t)ch + j];
            if (flag[ch]) {
                uint32_t slot = next_base + o_in;
                next_list[o_in++] = ch;
                ref[k] = slot; cnt[k] = 0x80000000u | slot;
            } else {   // leaf: cnt = ready-made deferred-leaf reference, ref = first triangle
                uint4 cn = nodes[ch]; uint32_t c = cn.y - cn.x;
                ref[k] = cn.x; cnt[k] = (cn.x << 6) | (c < 31u ? c : 31u);
            }
        } else {
            // empty slot: a far-away point box that every ray surely misses (|t| ~ 1e18 / |d|),
            // flagged as "inner" so that no leaf logic ever looks at it
            for (int j = 0; j < 6; ++j) b[6 * k + j] = 1e18f;
            ref[k] = 0u; cnt[k] = 0x80000000u;
        }
    }
    Node128 o;
    for (int j = 0; j < 6; ++j) o.q[j] = make_float4(b[4 * j], b[4 * j + 1], b[4 * j + 2], b[4 * j + 3]);
    o.ref = make_uint4(ref[0], ref[1], ref[2], ref[3]);
    o.cnt = make_uint4(cnt[0], cnt[1], cnt[2], cnt[3]);
    wide[level_base + i] = o;
}

// Directed fp32 -> bf16 (bits): the largest bf16 <= x / the smallest bf16 >= x.  Truncating the low 16 bits rounds towards
// zero; one unit is added to the magnitude when bits were dropped and the direction points away from zero (a carry into
// the exponent, up to infinity, is still on the right side).  NaN stays NaN.
__device__ __forceinline__ unsigned short f2bf_rd(float x)
{
    const uint32_t u = __float_as_uint(x);
    uint32_t t = u >> 16;
    if ((u & 0xffffu) && (u >> 31)) t += 1u;
    return (unsigned short)t;
}
__device__ __forceinline__ unsigned short f2bf_ru(float x)
{
    const uint32_t u = __float_as_uint(x);
    uint32_t t = u >> 16;
    if ((u & 0xffffu) && !(u >> 31)) t += 1u;
    return (unsigned short)t;
}

// Node80 from Node128: centre of the union of the occupied children, outward-rounded fp16 offsets.
__global__ void k_build_wide80(const Node128 * __restrict__ wide, uint32_t n, Node80 * __restrict__ out)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const Node128 nd = wide[i];
    const float * q = reinterpret_cast<const float *>(nd.q);          // 24 floats, 6 per child
    const uint32_t cnt[4] = {nd.cnt.x, nd.cnt.y, nd.cnt.z, nd.cnt.w};
    float mn[3] = {UAV_INF, UAV_INF, UAV_INF}, mx[3] = {-UAV_INF, -UAV_INF, -UAV_INF};
    bool occ[4];
    for (int k = 0; k < 4; ++k) {
        occ[k] = !(cnt[k] == 0x80000000u && q[6 * k] == 1e18f);       // empty slots: inner index 0 at 1e18
        if (!occ[k]) continue;
        for (int a = 0; a < 3; ++a) { mn[a] = fminf(mn[a], q[6 * k + a]); mx[a] = fmaxf(mx[a], q[6 * k + 3 + a]); }
    }
    float c[3];
    for (int a = 0; a < 3; ++a) { c[a] = 0.5f * mn[a] + 0.5f * mx[a]; if (!(fabsf(c[a]) < 1e30f)) c[a] = 0.0f; }
    __half h[24];
    for (int k = 0; k < 4; ++k)
        for (int a = 0; a < 3; ++a) {
#if K1_PRMT
            const int lo = 2 * (3 * k + a), hi = lo + 1;                          // word (min, max) per child and axis
#else
            const int lo = 6 * k + a, hi = 6 * k + 3 + a;
#endif
#if K1_PRMT && K1_BF16
            if (occ[k]) {
                h[lo] = __ushort_as_half(f2bf_rd(__fsub_rd(q[6 * k + a], c[a])));      // bf16 bits in the 16-bit slots
                h[hi] = __ushort_as_half(f2bf_ru(__fsub_ru(q[6 * k + 3 + a], c[a])));
            } else {
                h[lo] = __ushort_as_half((unsigned short)0x7f80);      // +inf: no segment reaches it
                h[hi] = __ushort_as_half((unsigned short)0x7f80);
            }
#else
            if (occ[k]) {
                h[lo] = __float2half_rd(__fsub_rd(q[6 * k + a], c[a]));
                h[hi] = __float2half_ru(__fsub_ru(q[6 * k + 3 + a], c[a]));
            } else {
                h[lo] = __ushort_as_half((unsigned short)0x7c00);      // +inf: no segment reaches it
                h[hi] = __ushort_as_half((unsigned short)0x7c00);
            }
#endif
        }
    Node80 o;
    o.c = make_float4(c[0], c[1], c[2], 0.0f);
    const uint32_t * hw = reinterpret_cast<const uint32_t *>(h);
    for (int j = 0; j < 3; ++j) o.h[j] = make_uint4(hw[4 * j], hw[4 * j + 1], hw[4 * j + 2], hw[4 * j + 3]);
    o.cnt = nd.cnt;
    out[i] = o;
}

// out[0] += violations of "every child box lies inside its parent's box and every leaf box is the
// exact bound of its triangles" (NaNs count as violations); out[1] = largest leaf
__global__ void k_check_tree(const uint4 * __restrict__ nodes, const float * __restrict__ aabbs,
    const float * __restrict__ tris, uint32_t n, uint32_t * __restrict__ out)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint4 nd = nodes[i];
    const float * p = aabbs + 6 * (size_t)i;
    bool ok = true;
    if (nd.z == UAV_NAI || nd.w == UAV_NAI) {
        float mn[3] = {UAV_INF, UAV_INF, UAV_INF}, mx[3] = {-UAV_INF, -UAV_INF, -UAV_INF};
        for (uint32_t t = nd.x; t < nd.y; ++t)
            for (int v = 0; v < 3; ++v)
                for (int k = 0; k < 3; ++k) {
                    float x = tris[9 * (size_t)t + 3 * v + k];
                    mn[k] = fminf(mn[k], x); mx[k] = fmaxf(mx[k], x);
                    ok = ok && (x == x);
                }
        for (int k = 0; k < 3; ++k) ok = ok && (p[k] == mn[k]) && (p[3 + k] == mx[k]);
        ok = ok && nd.y > nd.x;
        atomicMax(out + 1, nd.y - nd.x);
    } else {
        for (int c = 0; c < 2; ++c) {
            const float * q = aabbs + 6 * (size_t)(c ? nd.w : nd.z);
            for (int k = 0; k < 3; ++k) ok = ok && (q[k] >= p[k]) && (q[3 + k] <= p[3 + k]) && (q[k] <= q[3 + k]);
        }
    }
    if (!ok) atomicAdd(out, 1u);
}

// Exact leaf boxes addressed by the leaf's first triangle (k_trace re-checks a deferred leaf against its own box once one
// of its triangles is accepted; the pending entry only carries first / count).
// max_diag: largest box diagonal over the leaves, as float bits (positive floats order like their bits); k7.cu's margin
__global__ void k_leaf_boxes(const uint4 * __restrict__ nodes, const float * __restrict__ aabbs, uint32_t n, float4 * __restrict__ out,
    uint32_t * __restrict__ max_diag)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint4 nd = nodes[i];
    if (nd.z != UAV_NAI && nd.w != UAV_NAI) return;
    const float * p = aabbs + 6 * (size_t)i;
    out[2 * (size_t)nd.x] = make_float4(p[0], p[1], p[2], p[3]);
    out[2 * (size_t)nd.x + 1] = make_float4(p[4], p[5], 0.0f, 0.0f);
    const float dx = p[3] - p[0], dy = p[4] - p[1], dz = p[5] - p[2];
    const float dg = sqrtf(dx * dx + dy * dy + dz * dz);
    if (dg == dg) atomicMax(max_diag, __float_as_uint(fmaxf(dg, 0.0f)));
}

// Per-triangle constants with the reference's operation order (libs/acc/primitives.h:
// 217-225 normal, :139-145 Gram terms).  Degenerate triangles (|N| < eps, rejected at
// :222) get a zero normal, which the |cosine| < eps test rejects.
__global__ void k_build_tris(const float * __restrict__ tris, uint32_t n, Tri64 * __restrict__ out)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float * t = tris + 9 * (size_t)i;
    V3 a = mk(t[0], t[1], t[2]), b = mk(t[3], t[4], t[5]), c = mk(t[6], t[7], t[8]);
    V3 ab = b - a, ac = c - a;
    V3 nrm = cross(ab, ac);
    float len = norm(nrm);
    V3 un = (len < FLT_EPSILON) ? mk(0.0f, 0.0f, 0.0f) : nrm / len;
    float d00 = dot(ab, ab), d01 = dot(ab, ac), d11 = dot(ac, ac);
    float denom = d00 * d11 - d01 * d01;
    Tri64 o;
    o.q0 = make_float4(a.x, a.y, a.z, ab.x);
    o.q1 = make_float4(ab.y, ab.z, ac.x, ac.y);
    o.q2 = make_float4(ac.z, un.x, un.y, un.z);
    o.q3 = make_float4(d00, d01, d11, denom);
    out[i] = o;
}

int finalize_layout(uavmvs_bvh * b)
{
    uint32_t n = b->num_nodes, t = b->num_tris;
    uint32_t * flag = nullptr, * slot = nullptr;
    UAV_CUDA(cudaMalloc(&flag, sizeof(uint32_t) * (size_t)n));
    UAV_CUDA(cudaMalloc(&slot, sizeof(uint32_t) * (size_t)n));
    k_flag_inner<<<divup(n, 256), 256>>>(b->nodes, n, flag);
    size_t tmp_bytes = 0; void * tmp = nullptr;
    cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, flag, slot, (int)n);
    UAV_CUDA(cudaMalloc(&tmp, tmp_bytes));
    cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, flag, slot, (int)n);
    uint32_t last_flag = 0, last_slot = 0;
    UAV_CUDA(cudaMemcpy(&last_flag, flag + (n - 1), 4, cudaMemcpyDeviceToHost));
    UAV_CUDA(cudaMemcpy(&last_slot, slot + (n - 1), 4, cudaMemcpyDeviceToHost));
    b->num_inner = last_flag + last_slot;
    uint4 root;
    UAV_CUDA(cudaMemcpy(&root, b->nodes, sizeof(uint4), cudaMemcpyDeviceToHost));
    b->root_is_leaf = !(root.z != UAV_NAI && root.w != UAV_NAI);
    b->root_first = root.x; b->root_count = root.y - root.x;
    UAV_CUDA(cudaMalloc(&b->inner, sizeof(Node64) * (size_t)(b->num_inner ? b->num_inner : 1)));
    UAV_CUDA(cudaMalloc(&b->wide, sizeof(Node128) * (size_t)(b->num_inner ? b->num_inner : 1)));
    UAV_CUDA(cudaMalloc(&b->ptris, sizeof(Tri64) * (size_t)t));
    {
        uint32_t * chk = nullptr, h_chk[2] = {1, 0};
        UAV_CUDA(cudaMalloc(&chk, 8));
        UAV_CUDA(cudaMemset(chk, 0, 8));
        k_check_tree<<<divup(n, 256), 256>>>(b->nodes, b->aabbs, b->tris, n, chk);
        UAV_CUDA(cudaMemcpy(h_chk, chk, 8, cudaMemcpyDeviceToHost));
        cudaFree(chk);
        b->max_leaf = h_chk[1];
        // the fast traversal needs nested inner boxes, tight leaf boxes and leaves of <= 31 triangles
        b->nested = (h_chk[0] == 0 && h_chk[1] <= 31u) ? 1u : 0u;
    }
    k_build_inner<<<divup(n, 256), 256>>>(b->nodes, b->aabbs, flag, slot, n, b->inner);
    b->num_wide = 0;
    if (b->num_inner) {
        uint32_t * list_a, * list_b, * ninner, * off; uint4 * kids;
        const size_t cap = b->num_inner;
        UAV_CUDA(cudaMalloc(&list_a, 4 * cap)); UAV_CUDA(cudaMalloc(&list_b, 4 * cap));
        UAV_CUDA(cudaMalloc(&ninner, 4 * cap)); UAV_CUDA(cudaMalloc(&off, 4 * cap));
        UAV_CUDA(cudaMalloc(&kids, sizeof(uint4) * cap));
        UAV_CUDA(cudaMemset(list_a, 0, 4));                  // the root's wide node is wide[0]
        int mode = 0;     // measured on config 2: two binary levels per wide node 4.95 ms / step, greedy by area 5.01 ms
        if (const char * e = getenv("UAVMVS_COLLAPSE")) mode = atoi(e);
        uint32_t count = 1, level_base = 0, wide_levels = 0;
        while (count) {
            k_wide_expand<<<divup(count, 128), 128>>>(list_a, count, b->nodes, b->aabbs, flag, mode, kids, ninner);
            cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, ninner, off, (int)count);    // count <= n: tmp is large enough
            uint32_t last[2];
            UAV_CUDA(cudaMemcpy(&last[0], ninner + (count - 1), 4, cudaMemcpyDeviceToHost));
            UAV_CUDA(cudaMemcpy(&last[1], off + (count - 1), 4, cudaMemcpyDeviceToHost));
            const uint32_t next_base = level_base + count;
            UAV_REQUIRE((size_t)next_base + last[0] + last[1] <= cap, "wide collapse overflow");
            k_wide_emit<<<divup(count, 128), 128>>>(count, kids, off, level_base, next_base, b->nodes, b->aabbs, flag,
                b->wide, list_b);
            UAV_CUDA(cudaGetLastError());
            level_base = next_base; count = last[0] + last[1];
            uint32_t * t = list_a; list_a = list_b; list_b = t;
            ++wide_levels;
        }
        b->num_wide = level_base;
        cudaFree(list_a); cudaFree(list_b); cudaFree(ninner); cudaFree(off); cudaFree(kids);
        // Every traversal keeps a fixed local stack (the reference too: 8 + 64 slots, unchecked, libs/cacc/tracing.h:28-33).
        // A wide level spans at most two binary levels, so the binary depth is <= 2 wide_levels: the binary walks push one
        // entry per level (UAV_STACK), the 4-wide walk of K1 up to three per wide level (K1_STACK).  Deeper trees are
        // refused here instead of overflowing there.
        b->depth = 2 * wide_levels;
        UAV_REQUIRE(2 * wide_levels <= (uint32_t)UAV_STACK && 3 * wide_levels + 4 <= (uint32_t)K1_STACK,
            "BVH deeper than the traversal stacks (96 binary levels): rebuild it with uavmvs_bvh_build_sah / _ploc");
    }
    {
        const uint32_t ni = b->num_wide ? b->num_wide : 1;
        UAV_CUDA(cudaMalloc(&b->wide80, sizeof(Node80) * (size_t)ni));
        UAV_CUDA(cudaMemset(b->wide80, 0, sizeof(Node80) * (size_t)ni));
        if (b->num_wide) k_build_wide80<<<divup(b->num_wide, 256), 256>>>(b->wide, b->num_wide, b->wide80);
        float root[6];
        UAV_CUDA(cudaMemcpy(root, b->aabbs, sizeof(root), cudaMemcpyDeviceToHost));
        for (int a = 0; a < 3; ++a) {
            float m = fmaxf(fabsf(root[a]), fabsf(root[3 + a]));
            b->bound_abs[a] = (m == m && m < 1e30f) ? m : 1e30f;
        }
    }
    if (b->nested) {
        UAV_CUDA(cudaMalloc(&b->leafbox, sizeof(float4) * 2 * (size_t)t));
        UAV_CUDA(cudaMemset(b->leafbox, 0, sizeof(float4) * 2 * (size_t)t));
        uint32_t * d_diag = nullptr, h_diag = 0;
        UAV_CUDA(cudaMalloc(&d_diag, sizeof(uint32_t)));
        UAV_CUDA(cudaMemset(d_diag, 0, sizeof(uint32_t)));
        k_leaf_boxes<<<divup(n, 256), 256>>>(b->nodes, b->aabbs, n, b->leafbox, d_diag);
        UAV_CUDA(cudaMemcpy(&h_diag, d_diag, sizeof(uint32_t), cudaMemcpyDeviceToHost));
        cudaFree(d_diag);
        memcpy(&b->max_leaf_diag, &h_diag, sizeof(float));
    }
    k_build_tris<<<divup(t, 256), 256>>>(b->tris, t, b->ptris);
    UAV_CUDA(cudaGetLastError());
    UAV_CUDA(cudaDeviceSynchronize());
    cudaFree(tmp); cudaFree(flag); cudaFree(slot);
    return 0;
}

int alloc_reference_arrays(uavmvs_bvh * b)
{
    UAV_CUDA(cudaMalloc(&b->nodes, sizeof(uint4) * (size_t)b->num_nodes));
    UAV_CUDA(cudaMalloc(&b->aabbs, sizeof(float) * 6 * (size_t)b->num_nodes));
    UAV_CUDA(cudaMalloc(&b->tris, sizeof(float) * 9 * (size_t)b->num_tris));
    UAV_CUDA(cudaMalloc(&b->indices, sizeof(uint32_t) * (size_t)b->num_tris));
    return 0;
}

// ------------------------------------------------------------------ LBVH build
__device__ __forceinline__ uint64_t spread21(uint64_t x)
{
    x &= 0x1fffffull;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

__global__ void k_scene_bounds(const float * __restrict__ verts, uint32_t nv, float * __restrict__ bounds)
{
    // bounds: min xyz as ordered ints [0..2], max xyz [3..5]; tiny, atomics are fine
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nv) return;
    for (int k = 0; k < 3; ++k) {
        float v = verts[3 * (size_t)i + k];
        int iv = __float_as_int(v);
        int ord = iv >= 0 ? iv : iv ^ 0x7fffffff;
        atomicMin((int *)bounds + k, ord);
        atomicMax((int *)bounds + 3 + k, ord);
    }
}

__global__ void k_morton(const float * __restrict__ verts, const uint32_t * __restrict__ faces,
    uint32_t nf, const float * __restrict__ bounds, uint64_t * __restrict__ keys, uint32_t * __restrict__ ids)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nf) return;
    float lo[3], ext = 0.0f;
    for (int k = 0; k < 3; ++k) {
        int a = ((const int *)bounds)[k], b = ((const int *)bounds)[3 + k];
        float mn = __int_as_float(a >= 0 ? a : a ^ 0x7fffffff);
        float mx = __int_as_float(b >= 0 ? b : b ^ 0x7fffffff);
        lo[k] = mn; ext = fmaxf(ext, mx - mn);
    }
    ext = fmaxf(ext, 1e-20f);
    uint64_t code = 0;
    for (int k = 0; k < 3; ++k) {
        float c = 0.0f;
        for (int j = 0; j < 3; ++j) c += verts[3 * (size_t)faces[3 * (size_t)i + j] + k];
        c = (c / 3.0f - lo[k]) / ext;
        c = fminf(fmaxf(c, 0.0f), 1.0f);
        uint64_t q = (uint64_t)(c * 2097151.0f);
        code |= spread21(q) << k;
    }
    keys[i] = code; ids[i] = i;
}

__device__ __forceinline__ int lcp(const uint64_t * __restrict__ keys, int n, int i, int j)
{
    if (j < 0 || j >= n) return -1;
    uint64_t a = keys[i], b = keys[j];
    if (a != b) return __clzll(a ^ b);
    return 64 + __clz((uint32_t)i ^ (uint32_t)j);   // duplicate codes: fall back to the position
}

// Karras 2012: one thread per internal node.  Node ids: internal i -> i (root 0),
// leaf j -> (n-1)+j.
__global__ void k_radix_tree(const uint64_t * __restrict__ keys, int n, uint4 * __restrict__ nodes,
    uint32_t * __restrict__ parent)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    int d = (lcp(keys, n, i, i + 1) - lcp(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    int dmin = lcp(keys, n, i, i - d);
    int lmax = 2;
    while (lcp(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
        if (lcp(keys, n, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = lcp(keys, n, i, j);
    int s = 0;
    for (int t = (l + 1) / 2; ; t = (t + 1) / 2) {
        if (lcp(keys, n, i, i + (s + t) * d) > dnode) s += t;
        if (t <= 1) break;
    }
    int gamma = i + s * d + min(d, 0);
    int lo = min(i, j), hi = max(i, j);
    uint32_t left = (lo == gamma) ? (uint32_t)(n - 1 + gamma) : (uint32_t)gamma;
    uint32_t right = (hi == gamma + 1) ? (uint32_t)(n - 1 + gamma + 1) : (uint32_t)(gamma + 1);
    nodes[i] = make_uint4((uint32_t)lo, (uint32_t)hi + 1u, left, right);
    parent[left] = (uint32_t)i;
    parent[right] = (uint32_t)i;
    if (i == 0) parent[0] = UAV_NAI;
}

__global__ void k_leaves_refit(const float * __restrict__ verts, const uint32_t * __restrict__ faces,
    const uint32_t * __restrict__ ids, int n, uint4 * __restrict__ nodes, float * __restrict__ aabbs,
    float * __restrict__ tris, uint32_t * __restrict__ indices, const uint32_t * __restrict__ parent,
    uint32_t * __restrict__ visits)
{
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    uint32_t f = ids[j];
    float t[9];
    for (int k = 0; k < 3; ++k)
        for (int c = 0; c < 3; ++c) t[3 * k + c] = verts[3 * (size_t)faces[3 * (size_t)f + k] + c];
    for (int c = 0; c < 9; ++c) tris[9 * (size_t)j + c] = t[c];
    indices[j] = f;
    uint32_t me = (uint32_t)(n - 1 + j);
    nodes[me] = make_uint4((uint32_t)j, (uint32_t)j + 1u, UAV_NAI, UAV_NAI);
    float * bb = aabbs + 6 * (size_t)me;
    for (int c = 0; c < 3; ++c) {   // libs/acc/primitives.h:74-81
        bb[c] = fminf(t[c], fminf(t[3 + c], t[6 + c]));
        bb[3 + c] = fmaxf(t[c], fmaxf(t[3 + c], t[6 + c]));
    }
    __threadfence();
    uint32_t p = (n > 1) ? parent[me] : UAV_NAI;
    while (p != UAV_NAI) {
        if (atomicAdd(visits + p, 1u) == 0u) return;   // first child to arrive stops
        __threadfence();
        uint4 nd = nodes[p];
        volatile const float * l = aabbs + 6 * (size_t)nd.z;
        volatile const float * r = aabbs + 6 * (size_t)nd.w;
        float * o = aabbs + 6 * (size_t)p;
        for (int c = 0; c < 3; ++c) { o[c] = fminf(l[c], r[c]); o[3 + c] = fmaxf(l[3 + c], r[3 + c]); }
        __threadfence();
        p = parent[p];
    }
}

// ------------------------------------------------------------------ PLOC build
// Parallel locally-ordered clustering (Meister & Bittner 2017) over the Morton order: every round
// each cluster looks at its PLOC_R neighbours on either side, picks the one whose union box has the
// smallest surface area, and mutual pairs merge.  Bottom-up like the LBVH, but the topology follows
// surface area instead of Morton-code bits, which brings the tree close to SAH quality.
// Node ids as in the LBVH layout: inner nodes [0, n-1) with the root at 0 (ids are handed out from
// n-2 downwards, the root is created last), leaf j = n-1+j (one triangle, Morton order).
constexpr int PLOC_R = 24;

__device__ __forceinline__ float union_area(const float * a, const float * b)
{
    float dx = fmaxf(a[3], b[3]) - fminf(a[0], b[0]);
    float dy = fmaxf(a[4], b[4]) - fminf(a[1], b[1]);
    float dz = fmaxf(a[5], b[5]) - fminf(a[2], b[2]);
    return dx * dy + dy * dz + dz * dx;
}

__global__ void k_ploc_leaves(const float * __restrict__ verts, const uint32_t * __restrict__ faces,
    const uint32_t * __restrict__ ids, int n, uint4 * __restrict__ nodes, float * __restrict__ aabbs,
    float * __restrict__ tris, uint32_t * __restrict__ indices, uint32_t * __restrict__ cluster)
{
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    uint32_t f = ids[j];
    float t[9];
    for (int k = 0; k < 3; ++k)
        for (int c = 0; c < 3; ++c) t[3 * k + c] = verts[3 * (size_t)faces[3 * (size_t)f + k] + c];
    for (int c = 0; c < 9; ++c) tris[9 * (size_t)j + c] = t[c];
    indices[j] = f;
    uint32_t me = (uint32_t)(n - 1 + j);
    nodes[me] = make_uint4((uint32_t)j, (uint32_t)j + 1u, UAV_NAI, UAV_NAI);
    float * bb = aabbs + 6 * (size_t)me;
    for (int c = 0; c < 3; ++c) {   // libs/acc/primitives.h:74-81
        bb[c] = fminf(t[c], fminf(t[3 + c], t[6 + c]));
        bb[3 + c] = fmaxf(t[c], fmaxf(t[3 + c], t[6 + c]));
    }
    cluster[j] = me;
}

__global__ void k_ploc_nearest(const uint32_t * __restrict__ cluster, int n, const float * __restrict__ aabbs,
    uint32_t * __restrict__ nn)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float bi[6];
    const float * p = aabbs + 6 * (size_t)cluster[i];
    for (int k = 0; k < 6; ++k) bi[k] = p[k];
    float best = UAV_INF; int bj = -1;
    int lo = max(0, i - PLOC_R), hi = min(n - 1, i + PLOC_R);
    for (int j = lo; j <= hi; ++j) {
        if (j == i) continue;
        float a = union_area(bi, aabbs + 6 * (size_t)cluster[j]);
        if (a < best) { best = a; bj = j; }          // ties: the lower position wins
    }
    nn[i] = (uint32_t)bj;
}

// keep[i] = 0 for the higher member of a mutual pair (absorbed), 1 otherwise; made[i] = 1 for the lower
// member (creates a node)
__global__ void k_ploc_flags(const uint32_t * __restrict__ nn, int n, uint32_t * __restrict__ keep,
    uint32_t * __restrict__ made)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int j = (int)nn[i];
    bool mutual = j >= 0 && (int)nn[j] == i;
    keep[i] = (mutual && i > j) ? 0u : 1u;
    made[i] = (mutual && i < j) ? 1u : 0u;
}

__global__ void k_ploc_merge(const uint32_t * __restrict__ cluster, const uint32_t * __restrict__ nn, int n,
    const uint32_t * __restrict__ keep, const uint32_t * __restrict__ keep_pos,
    const uint32_t * __restrict__ made, const uint32_t * __restrict__ made_pos, uint32_t id_base,
    uint4 * __restrict__ nodes, float * __restrict__ aabbs, uint32_t * __restrict__ out_cluster)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || !keep[i]) return;
    uint32_t c = cluster[i];
    if (made[i]) {
        uint32_t l = c, r = cluster[nn[i]];
        uint32_t id = id_base + made_pos[i];
        const float * a = aabbs + 6 * (size_t)l, * b = aabbs + 6 * (size_t)r;
        float * o = aabbs + 6 * (size_t)id;
        for (int k = 0; k < 3; ++k) { o[k] = fminf(a[k], b[k]); o[3 + k] = fmaxf(a[3 + k], b[3 + k]); }
        uint4 ln = nodes[l], rn = nodes[r];
        nodes[id] = make_uint4(min(ln.x, rn.x), max(ln.y, rn.y), l, r);   // covering range (not a partition)
        c = id;
    }
    out_cluster[keep_pos[i]] = c;
}

// ------------------------------------------------------------------ batch trace
__global__ void k_trace(BvhView bvh, const float * __restrict__ rays, uint64_t n, int any_hit,
    uint8_t * __restrict__ hit, uint32_t * __restrict__ face, float * __restrict__ t)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool active = i < n;
    const float * r = rays + 8 * (active ? i : 0);
    Ray ray;
    ray.o = mk(r[0], r[1], r[2]); ray.d = mk(r[3], r[4], r[5]); ray.tmin = r[6]; ray.tmax = r[7];
    if (any_hit) {
        bool h = trace_any_warp(bvh, ray, active);
        if (!active) return;
        hit[i] = h ? 1 : 0;
        if (face) face[i] = UAV_NAI;
        if (t) t[i] = 0.0f;
    } else {
        uint32_t f; float tt;
        bool h = trace_closest_warp(bvh, ray, active, f, tt);
        if (!active) return;
        hit[i] = h ? 1 : 0;
        if (face) face[i] = h ? f : UAV_NAI;
        if (t) t[i] = h ? tt : 0.0f;
    }
}

}  // namespace uav

using namespace uav;

extern "C" {

const char * uavmvs_last_error(void) { return g_error.c_str(); }

int uavmvs_set_device(int device) { UAV_CUDA(cudaSetDevice(device)); return 0; }
int uavmvs_device_count(int * count) { UAV_CUDA(cudaGetDeviceCount(count)); return 0; }
int uavmvs_synchronize(void * stream) { UAV_CUDA(cudaStreamSynchronize((cudaStream_t)stream)); return 0; }

// User-supplied topology (the reference's converter trusts acc, libs/cacc/bvh_tree.cu:17-48): every index in range, every
// node reachable at most once from the root (no cycles, no shared subtrees), triangle ranges inside the array.  Depth is
// checked by finalize_layout for every tree, GPU-built ones included.
static int validate_tree_arrays(const uint32_t * nd, uint32_t n, uint32_t t)
{
    std::vector<uint32_t> order; order.reserve(n);
    std::vector<char> seen(n, 0);
    order.push_back(0); seen[0] = 1;
    for (size_t h = 0; h < order.size(); ++h) {
        const uint32_t * q = nd + 4 * (size_t)order[h];
        const bool inner = q[2] != UAV_NAI && q[3] != UAV_NAI;
        if (!inner) {
            UAV_REQUIRE(q[0] < q[1] && q[1] <= t, "leaf triangle range outside the triangle array");
            continue;
        }
        for (int c = 2; c < 4; ++c) {
            UAV_REQUIRE(q[c] < n, "child index outside the node array");
            UAV_REQUIRE(!seen[q[c]], "nodes do not form a tree (a node is reachable twice)");
            seen[q[c]] = 1; order.push_back(q[c]);
        }
    }
    return 0;
}

static int destroy_on_error(uavmvs_bvh * b, int rc) { if (rc) uavmvs_bvh_destroy(b); return rc; }

static int upload_tree_arrays(uavmvs_bvh * b, const uint32_t * nodes, const float * aabbs, const float * tris, const uint32_t * indices)
{
    if (alloc_reference_arrays(b)) return 1;
    UAV_CUDA(cudaMemcpy(b->nodes, nodes, sizeof(uint4) * (size_t)b->num_nodes, cudaMemcpyHostToDevice));
    UAV_CUDA(cudaMemcpy(b->aabbs, aabbs, sizeof(float) * 6 * (size_t)b->num_nodes, cudaMemcpyHostToDevice));
    UAV_CUDA(cudaMemcpy(b->tris, tris, sizeof(float) * 9 * (size_t)b->num_tris, cudaMemcpyHostToDevice));
    UAV_CUDA(cudaMemcpy(b->indices, indices, sizeof(uint32_t) * (size_t)b->num_tris, cudaMemcpyHostToDevice));
    return finalize_layout(b);
}

int uavmvs_bvh_create_from_arrays(const uint32_t * nodes, const float * aabbs, const float * tris,
    const uint32_t * indices, uint32_t num_nodes, uint32_t num_tris, uavmvs_bvh ** out)
{
    UAV_REQUIRE(nodes && aabbs && tris && indices && out, "null argument");
    UAV_REQUIRE(num_nodes >= 1 && num_tris >= 1, "empty tree");
    if (int rc = validate_tree_arrays(nodes, num_nodes, num_tris)) return rc;
    uavmvs_bvh * b = new uavmvs_bvh();
    b->num_nodes = num_nodes; b->num_tris = num_tris;
    if (int rc = destroy_on_error(b, upload_tree_arrays(b, nodes, aabbs, tris, indices))) return rc;
    *out = b;
    return 0;
}

int uavmvs_bvh_build_lbvh(const float * verts, uint32_t num_verts, const uint32_t * faces,
    uint32_t num_faces, uavmvs_bvh ** out)
{
    UAV_REQUIRE(verts && faces && out, "null argument");
    UAV_REQUIRE(num_faces >= 1 && num_verts >= 1, "empty mesh");
    int n = (int)num_faces;
    uavmvs_bvh * b = new uavmvs_bvh();
    b->num_nodes = 2 * num_faces - 1; b->num_tris = num_faces;
    if (alloc_reference_arrays(b)) return 1;
    float * d_verts; uint32_t * d_faces, * ids, * ids2, * parent, * visits; uint64_t * keys, * keys2; float * bounds;
    UAV_CUDA(cudaMalloc(&d_verts, sizeof(float) * 3 * (size_t)num_verts));
    UAV_CUDA(cudaMalloc(&d_faces, sizeof(uint32_t) * 3 * (size_t)num_faces));
    UAV_CUDA(cudaMalloc(&ids, 4 * (size_t)n)); UAV_CUDA(cudaMalloc(&ids2, 4 * (size_t)n));
    UAV_CUDA(cudaMalloc(&keys, 8 * (size_t)n)); UAV_CUDA(cudaMalloc(&keys2, 8 * (size_t)n));
    UAV_CUDA(cudaMalloc(&parent, 4 * (size_t)b->num_nodes));
    UAV_CUDA(cudaMalloc(&visits, 4 * (size_t)n));
    UAV_CUDA(cudaMalloc(&bounds, 6 * sizeof(float)));
    UAV_CUDA(cudaMemcpy(d_verts, verts, sizeof(float) * 3 * (size_t)num_verts, cudaMemcpyHostToDevice));
    UAV_CUDA(cudaMemcpy(d_faces, faces, sizeof(uint32_t) * 3 * (size_t)num_faces, cudaMemcpyHostToDevice));
    int init[6] = {0x7fffffff, 0x7fffffff, 0x7fffffff, (int)0x80000000, (int)0x80000000, (int)0x80000000};
    UAV_CUDA(cudaMemcpy(bounds, init, sizeof(init), cudaMemcpyHostToDevice));
    UAV_CUDA(cudaMemset(visits, 0, 4 * (size_t)n));
    k_scene_bounds<<<divup(num_verts, 256), 256>>>(d_verts, num_verts, bounds);
    k_morton<<<divup(n, 256), 256>>>(d_verts, d_faces, num_faces, bounds, keys, ids);
    size_t tmp_bytes = 0; void * tmp = nullptr;
    cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys, keys2, ids, ids2, n, 0, 63);
    UAV_CUDA(cudaMalloc(&tmp, tmp_bytes));
    cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, keys, keys2, ids, ids2, n, 0, 63);
    if (n > 1) k_radix_tree<<<divup(n - 1, 256), 256>>>(keys2, n, b->nodes, parent);
    k_leaves_refit<<<divup(n, 256), 256>>>(d_verts, d_faces, ids2, n, b->nodes, b->aabbs, b->tris,
        b->indices, parent, visits);
    UAV_CUDA(cudaGetLastError());
    UAV_CUDA(cudaDeviceSynchronize());
    cudaFree(tmp); cudaFree(d_verts); cudaFree(d_faces); cudaFree(ids); cudaFree(ids2);
    cudaFree(keys); cudaFree(keys2); cudaFree(parent); cudaFree(visits); cudaFree(bounds);
    if (finalize_layout(b)) return 1;
    *out = b;
    return 0;
}

int uavmvs_bvh_build_ploc(const float * verts, uint32_t num_verts, const uint32_t * faces,
    uint32_t num_faces, uavmvs_bvh ** out)
{
    UAV_REQUIRE(verts && faces && out, "null argument");
    UAV_REQUIRE(num_faces >= 1 && num_verts >= 1, "empty mesh");
    int n = (int)num_faces;
    uavmvs_bvh * b = new uavmvs_bvh();
    b->num_nodes = 2 * num_faces - 1; b->num_tris = num_faces;
    if (alloc_reference_arrays(b)) return 1;
    float * d_verts; uint32_t * d_faces, * ids, * ids2; uint64_t * keys, * keys2; float * bounds;
    uint32_t * cl_a, * cl_b, * nn, * keep, * keep_pos, * made, * made_pos;
    UAV_CUDA(cudaMalloc(&d_verts, sizeof(float) * 3 * (size_t)num_verts));
    UAV_CUDA(cudaMalloc(&d_faces, sizeof(uint32_t) * 3 * (size_t)num_faces));
    UAV_CUDA(cudaMalloc(&ids, 4 * (size_t)n)); UAV_CUDA(cudaMalloc(&ids2, 4 * (size_t)n));
    UAV_CUDA(cudaMalloc(&keys, 8 * (size_t)n)); UAV_CUDA(cudaMalloc(&keys2, 8 * (size_t)n));
    UAV_CUDA(cudaMalloc(&bounds, 6 * sizeof(float)));
    UAV_CUDA(cudaMalloc(&cl_a, 4 * (size_t)n)); UAV_CUDA(cudaMalloc(&cl_b, 4 * (size_t)n));
    UAV_CUDA(cudaMalloc(&nn, 4 * (size_t)n));
    UAV_CUDA(cudaMalloc(&keep, 4 * (size_t)n)); UAV_CUDA(cudaMalloc(&keep_pos, 4 * (size_t)n));
    UAV_CUDA(cudaMalloc(&made, 4 * (size_t)n)); UAV_CUDA(cudaMalloc(&made_pos, 4 * (size_t)n));
    UAV_CUDA(cudaMemcpy(d_verts, verts, sizeof(float) * 3 * (size_t)num_verts, cudaMemcpyHostToDevice));
    UAV_CUDA(cudaMemcpy(d_faces, faces, sizeof(uint32_t) * 3 * (size_t)num_faces, cudaMemcpyHostToDevice));
    int init[6] = {0x7fffffff, 0x7fffffff, 0x7fffffff, (int)0x80000000, (int)0x80000000, (int)0x80000000};
    UAV_CUDA(cudaMemcpy(bounds, init, sizeof(init), cudaMemcpyHostToDevice));
    k_scene_bounds<<<divup(num_verts, 256), 256>>>(d_verts, num_verts, bounds);
    k_morton<<<divup(n, 256), 256>>>(d_verts, d_faces, num_faces, bounds, keys, ids);
    size_t sort_bytes = 0, scan_bytes = 0; void * tmp = nullptr;
    cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, keys, keys2, ids, ids2, n, 0, 63);
    cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, keep, keep_pos, n);
    size_t tmp_bytes = sort_bytes > scan_bytes ? sort_bytes : scan_bytes;
    UAV_CUDA(cudaMalloc(&tmp, tmp_bytes));
    cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, keys, keys2, ids, ids2, n, 0, 63);
    k_ploc_leaves<<<divup(n, 256), 256>>>(d_verts, d_faces, ids2, n, b->nodes, b->aabbs, b->tris, b->indices, cl_a);
    UAV_CUDA(cudaGetLastError());
    int count = n;
    uint32_t next_id = (uint32_t)(n - 1);          // inner ids are handed out downwards; the root gets 0
    while (count > 1) {
        k_ploc_nearest<<<divup(count, 128), 128>>>(cl_a, count, b->aabbs, nn);
        k_ploc_flags<<<divup(count, 256), 256>>>(nn, count, keep, made);
        cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, keep, keep_pos, count);
        cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, made, made_pos, count);
        uint32_t last[2];
        UAV_CUDA(cudaMemcpy(&last[0], made + (count - 1), 4, cudaMemcpyDeviceToHost));
        UAV_CUDA(cudaMemcpy(&last[1], made_pos + (count - 1), 4, cudaMemcpyDeviceToHost));
        uint32_t m = last[0] + last[1];
        UAV_REQUIRE(m >= 1 && m <= next_id, "PLOC made no progress");   // the global nearest pair is always mutual
        next_id -= m;
        k_ploc_merge<<<divup(count, 256), 256>>>(cl_a, nn, count, keep, keep_pos, made, made_pos, next_id,
            b->nodes, b->aabbs, cl_b);
        UAV_CUDA(cudaGetLastError());
        count -= (int)m;
        uint32_t * t = cl_a; cl_a = cl_b; cl_b = t;
    }
    UAV_CUDA(cudaDeviceSynchronize());
    UAV_REQUIRE(next_id == 0 || n == 1, "PLOC node count mismatch");
    cudaFree(tmp); cudaFree(d_verts); cudaFree(d_faces); cudaFree(ids); cudaFree(ids2);
    cudaFree(keys); cudaFree(keys2); cudaFree(bounds); cudaFree(cl_a); cudaFree(cl_b); cudaFree(nn);
    cudaFree(keep); cudaFree(keep_pos); cudaFree(made); cudaFree(made_pos);
    if (finalize_layout(b)) return 1;
    *out = b;
    return 0;
}

int uavmvs_bvh_sizes(const uavmvs_bvh * b, uint32_t * num_nodes, uint32_t * num_tris)
{
    UAV_REQUIRE(b, "null bvh");
    if (num_nodes) *num_nodes = b->num_nodes;
    if (num_tris) *num_tris = b->num_tris;
    return 0;
}

int uavmvs_bvh_export(const uavmvs_bvh * b, uint32_t * nodes, float * aabbs, float * tris, uint32_t * indices)
{
    UAV_REQUIRE(b, "null bvh");
    if (nodes) UAV_CUDA(cudaMemcpy(nodes, b->nodes, sizeof(uint4) * (size_t)b->num_nodes, cudaMemcpyDeviceToHost));
    if (aabbs) UAV_CUDA(cudaMemcpy(aabbs, b->aabbs, sizeof(float) * 6 * (size_t)b->num_nodes, cudaMemcpyDeviceToHost));
    if (tris) UAV_CUDA(cudaMemcpy(tris, b->tris, sizeof(float) * 9 * (size_t)b->num_tris, cudaMemcpyDeviceToHost));
    if (indices) UAV_CUDA(cudaMemcpy(indices, b->indices, sizeof(uint32_t) * (size_t)b->num_tris, cudaMemcpyDeviceToHost));
    return 0;
}

int uavmvs_bvh_destroy(uavmvs_bvh * b)
{
    if (!b) return 0;
    cudaFree(b->nodes); cudaFree(b->aabbs); cudaFree(b->tris); cudaFree(b->indices);
    cudaFree(b->inner); cudaFree(b->wide); cudaFree(b->wide80); cudaFree(b->ptris); cudaFree(b->leafbox);
    delete b;
    return 0;
}

int uavmvs_bvh_trace(void * stream, const uavmvs_bvh * b, const float * rays, uint64_t n,
    int any_hit, uint8_t * hit, uint32_t * face, float * t)
{
    UAV_REQUIRE(b && rays && hit, "null argument");
    if (n == 0) return 0;
    cudaStream_t s = (cudaStream_t)stream;
    float * d_rays, * d_t; uint8_t * d_hit; uint32_t * d_face;
    UAV_CUDA(cudaMalloc(&d_rays, 32 * n)); UAV_CUDA(cudaMalloc(&d_t, 4 * n));
    UAV_CUDA(cudaMalloc(&d_hit, n)); UAV_CUDA(cudaMalloc(&d_face, 4 * n));
    UAV_CUDA(cudaMemcpyAsync(d_rays, rays, 32 * n, cudaMemcpyHostToDevice, s));
    k_trace<<<divup(n, 128), 128, 0, s>>>(view_of(b), d_rays, n, any_hit, d_hit, d_face, d_t);
    UAV_CUDA(cudaGetLastError());
    UAV_CUDA(cudaMemcpyAsync(hit, d_hit, n, cudaMemcpyDeviceToHost, s));
    if (face) UAV_CUDA(cudaMemcpyAsync(face, d_face, 4 * n, cudaMemcpyDeviceToHost, s));
    if (t) UAV_CUDA(cudaMemcpyAsync(t, d_t, 4 * n, cudaMemcpyDeviceToHost, s));
    UAV_CUDA(cudaStreamSynchronize(s));
    cudaFree(d_rays); cudaFree(d_t); cudaFree(d_hit); cudaFree(d_face);
    return 0;
}

}  // extern "C"

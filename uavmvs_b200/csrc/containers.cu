// Device containers behind the C ABI: point cloud, observation-ray array, float array,
// image, KD-tree.  They mirror the memory layouts of the reference's cacc containers
// (libs/cacc/point_cloud.h:26-32, vector_array.h:23-50, array.h, image.h, kd_tree.h).
#include "common.cuh"
#include "kd.cuh"
#include <algorithm>
#include <cmath>
#include <cstdlib>

namespace uav {

// Bounding sphere of every 32-point block (centre = box centre, radius padded so that the
// block-level range cull in eval.cu is conservative w.r.t. fp32 rounding).
__global__ void k_block_spheres(const float4 * __restrict__ verts, uint32_t n, float4 * __restrict__ out,
    float4 * __restrict__ lo, float4 * __restrict__ hi)
{
    uint32_t lane = threadIdx.x & 31;
    uint32_t blk = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    uint32_t id = blk * 32 + lane;
    if (blk * 32 >= n) return;
    bool valid = id < n;
    float4 v = verts[valid ? id : n - 1];
    float mn[3] = {v.x, v.y, v.z}, mx[3] = {v.x, v.y, v.z};
    for (int k = 0; k < 3; ++k)
        for (int o = 16; o > 0; o >>= 1) {
            mn[k] = fminf(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], o));
            mx[k] = fmaxf(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o));
        }
    float cx = 0.5f * (mn[0] + mx[0]), cy = 0.5f * (mn[1] + mx[1]), cz = 0.5f * (mn[2] + mx[2]);
    float dx = v.x - cx, dy = v.y - cy, dz = v.z - cz;
    float r = sqrtf(dx * dx + dy * dy + dz * dz);
    for (int o = 16; o > 0; o >>= 1) r = fmaxf(r, __shfl_xor_sync(0xffffffffu, r, o));
    if (lane == 0) {
        out[blk] = make_float4(cx, cy, cz, r * 1.0001f + 1e-6f);
        lo[blk] = make_float4(mn[0], mn[1], mn[2], 0.0f);
        hi[blk] = make_float4(mx[0], mx[1], mx[2], 0.0f);
    }
}

// Cone of the block's normals: axis = normalised sum, half-angle = largest angle to the axis
// (padded); blocks with a degenerate axis get half-angle 4 (> pi: the cull never uses them).
__global__ void k_block_cones(const float4 * __restrict__ normals, uint32_t n, float4 * __restrict__ cone,
    float * __restrict__ nmax)
{
    uint32_t lane = threadIdx.x & 31;
    uint32_t blk = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    uint32_t id = blk * 32 + lane;
    if (blk * 32 >= n) return;
    float4 v = normals[id < n ? id : n - 1];
    float len = sqrtf(v.x * v.x + v.y * v.y + v.z * v.z);
    float ux = v.x / len, uy = v.y / len, uz = v.z / len;
    bool bad = !(len > 1e-20f) || !(len < 1e20f);
    float sx = bad ? 0.0f : ux, sy = bad ? 0.0f : uy, sz = bad ? 0.0f : uz, mx = len;
    for (int o = 16; o > 0; o >>= 1) {
        sx += __shfl_xor_sync(0xffffffffu, sx, o); sy += __shfl_xor_sync(0xffffffffu, sy, o);
        sz += __shfl_xor_sync(0xffffffffu, sz, o); mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    float sl = sqrtf(sx * sx + sy * sy + sz * sz);
    float ax = sx / sl, ay = sy / sl, az = sz / sl;
    float c = ax * ux + ay * uy + az * uz;
    float ang = bad ? 4.0f : acosf(fminf(fmaxf(c, -1.0f), 1.0f));
    for (int o = 16; o > 0; o >>= 1) ang = fmaxf(ang, __shfl_xor_sync(0xffffffffu, ang, o));
    uint32_t anybad = __ballot_sync(0xffffffffu, bad);
    if (!(sl > 1e-3f) || anybad) ang = 4.0f;
    if (lane == 0) { cone[blk] = make_float4(ax, ay, az, ang * 1.0001f + 1e-4f); nmax[blk] = mx * 1.0001f; }
}

// Oriented bounds of every block in the frame of its normal cone (see common.cuh:block_obb).
__global__ void k_block_obb(const float4 * __restrict__ verts, uint32_t n, const float4 * __restrict__ sphere,
    const float4 * __restrict__ cone, float4 * __restrict__ obb)
{
    uint32_t lane = threadIdx.x & 31;
    uint32_t blk = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    uint32_t id = blk * 32 + lane;
    if (blk * 32 >= n) return;
    float4 v = verts[id < n ? id : n - 1];
    float4 sp = sphere[blk], cn = cone[blk];
    float ax = cn.x, ay = cn.y, az = cn.z;
    float al = sqrtf(ax * ax + ay * ay + az * az);
    if (!(fabsf(al - 1.0f) < 1e-3f)) { ax = 0.0f; ay = 0.0f; az = 1.0f; }     // degenerate cone: world frame
    // f1 = unit vector orthogonal to the axis (cross with the world axis of the smallest component)
    float bx, by, bz;
    if (fabsf(ax) <= fabsf(ay) && fabsf(ax) <= fabsf(az)) { bx = 0.0f; by = -az; bz = ay; }
    else if (fabsf(ay) <= fabsf(az)) { bx = az; by = 0.0f; bz = -ax; }
    else { bx = -ay; by = ax; bz = 0.0f; }
    float bl = sqrtf(bx * bx + by * by + bz * bz);
    bx /= bl; by /= bl; bz /= bl;
    float cx = ay * bz - az * by, cy = az * bx - ax * bz, cz = ax * by - ay * bx;
    float dx = v.x - sp.x, dy = v.y - sp.y, dz = v.z - sp.z;
    float d[3] = {ax * dx + ay * dy + az * dz, bx * dx + by * dy + bz * dz, cx * dx + cy * dy + cz * dz};
    float mn[3], mx[3];
    for (int k = 0; k < 3; ++k) {
        mn[k] = d[k]; mx[k] = d[k];
        for (int o = 16; o > 0; o >>= 1) {
            mn[k] = fminf(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], o));
            mx[k] = fmaxf(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o));
        }
    }
    if (lane == 0) {
        const float pad = 2e-6f * (1.0f + sp.w);      // rounding of p - c and of the dot products (|p - c| <= r)
        float4 * o = obb + 5 * (size_t)blk;
        o[0] = make_float4(ax, ay, az, sp.x);
        o[1] = make_float4(bx, by, bz, sp.y);
        o[2] = make_float4(cx, cy, cz, sp.z);
        o[3] = make_float4(mn[0] - pad, mx[0] + pad, mn[1] - pad, mx[1] + pad);
        o[4] = make_float4(mn[2] - pad, mx[2] + pad, 0.0f, 0.0f);
    }
}

// Sphere around the 32 block spheres of a super-block (centre = centre of the box of the block
// centres; radius = max over blocks of |c - c_b| + r_b, padded against rounding).
__global__ void k_super_spheres(const float4 * __restrict__ blocks, uint32_t num_blocks, float4 * __restrict__ out)
{
    uint32_t lane = threadIdx.x & 31;
    uint32_t sb = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (sb * 32 >= num_blocks) return;
    uint32_t b = sb * 32 + lane;
    float4 s = blocks[b < num_blocks ? b : num_blocks - 1];
    float mn[3] = {s.x, s.y, s.z}, mx[3] = {s.x, s.y, s.z};
    for (int k = 0; k < 3; ++k)
        for (int o = 16; o > 0; o >>= 1) {
            mn[k] = fminf(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], o));
            mx[k] = fmaxf(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o));
        }
    float cx = 0.5f * (mn[0] + mx[0]), cy = 0.5f * (mn[1] + mx[1]), cz = 0.5f * (mn[2] + mx[2]);
    float dx = s.x - cx, dy = s.y - cy, dz = s.z - cz;
    float r = sqrtf(dx * dx + dy * dy + dz * dz) * 1.0001f + s.w;
    for (int o = 16; o > 0; o >>= 1) r = fmaxf(r, __shfl_xor_sync(0xffffffffu, r, o));
    if (lane == 0) out[sb] = make_float4(cx, cy, cz, r * 1.0001f + 1e-6f);
}

__global__ void k_pad3(const float * __restrict__ in, uint32_t n, float4 * __restrict__ out)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = make_float4(in[3 * (size_t)i], in[3 * (size_t)i + 1], in[3 * (size_t)i + 2], 0.0f);
}

__global__ void k_fill(float * __restrict__ p, uint32_t n, float v)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

static int upload3(const float * host, uint32_t n, float4 * dst)
{
    float * tmp;
    UAV_CUDA(cudaMalloc(&tmp, sizeof(float) * 3 * (size_t)n));
    UAV_CUDA(cudaMemcpy(tmp, host, sizeof(float) * 3 * (size_t)n, cudaMemcpyHostToDevice));
    k_pad3<<<divup(n, 256), 256>>>(tmp, n, dst);
    UAV_CUDA(cudaGetLastError());
    UAV_CUDA(cudaDeviceSynchronize());
    cudaFree(tmp);
    return 0;
}


// ---- certified nearest-vertex lookup tables (device side and the argument: kd.cuh, kd_find_nn_fast) ----------
KdFastView kd_fast_view(const uavmvs_kdtree * kd)
{
    KdFastView v; v.table = nullptr; v.pos = nullptr; v.rule = nullptr; v.nbr = nullptr; v.res = 0;
    if (kd && kd->fast_state == 1) { v.table = kd->fast_table; v.pos = kd->fast_pos; v.rule = kd->fast_rule; v.nbr = kd->fast_nbr; v.res = kd->fast_res; }
    return v;
}

int kd_fast_ensure(uavmvs_kdtree * kd)
{
    if (!kd || kd->fast_state != 0) return 0;
    kd->fast_state = -1;                                   // until proven suitable
    if (getenv("UAVMVS_NO_KD_FAST")) return 0;
    const uint32_t n = kd->num_nodes;
    const std::vector<uint32_t> & nd = kd->h_nodes; const std::vector<float> & vt = kd->h_verts;
    // the neighbour lists below are all-pairs on one host thread: direction sets only (the icosphere has 1272 vertices);
    // larger trees keep the plain search
    if (n < 2 || n > 8192 || nd.size() != 4 * (size_t)n) return 0;
    // topology: every node reachable once from the root, vertex ids a permutation
    std::vector<uint32_t> parent(n, UAV_NAI), depth(n, 0), node_of(n, UAV_NAI), order; order.reserve(n);
    std::vector<char> seen(n, 0);
    order.push_back(0); seen[0] = 1;
    for (size_t h = 0; h < order.size(); ++h) {
        const uint32_t i = order[h];
        if (nd[4 * i + 1] > 2 || nd[4 * i] >= n || node_of[nd[4 * i]] != UAV_NAI) return 0;
        node_of[nd[4 * i]] = i;
        for (int c = 2; c < 4; ++c) {
            const uint32_t ch = nd[4 * i + c];
            if (ch == UAV_NAI) continue;
            if (ch >= n || seen[ch]) return 0;
            seen[ch] = 1; parent[ch] = i; depth[ch] = depth[i] + 1; order.push_back(ch);
        }
    }
    if (order.size() != n) return 0;
    // distinct positions (bitwise after -0 -> +0); at most two vertices per position
    auto key = [&](uint32_t v, int k) { float f = vt[3 * (size_t)v + k] + 0.0f; uint32_t u; memcpy(&u, &f, 4); return u; };
    for (size_t i = 0; i < 3 * (size_t)n; ++i) if (!(fabsf(vt[i]) < 1e18f)) return 0;
    std::vector<uint32_t> ids(n);
    for (uint32_t i = 0; i < n; ++i) ids[i] = i;
    std::sort(ids.begin(), ids.end(), [&](uint32_t a, uint32_t b) {
        for (int k = 0; k < 3; ++k) { if (key(a, k) != key(b, k)) return key(a, k) < key(b, k); }
        return a < b; });
    std::vector<float4> pos; std::vector<uint4> rule;
    for (uint32_t i = 0; i < n;) {
        uint32_t j = i + 1;
        while (j < n && key(ids[j], 0) == key(ids[i], 0) && key(ids[j], 1) == key(ids[i], 1) && key(ids[j], 2) == key(ids[i], 2)) ++j;
        if (j - i > 2) return 0;
        const uint32_t a = ids[i];
        uint4 r = make_uint4(a, a, 3u, 0u);
        if (j - i == 2) {
            const uint32_t b = ids[i + 1];
            uint32_t na = node_of[a], nb = node_of[b], ca = UAV_NAI, cb = UAV_NAI;     // ca / cb: child of the meeting node
            while (depth[na] > depth[nb]) { ca = na; na = parent[na]; }
            while (depth[nb] > depth[na]) { cb = nb; nb = parent[nb]; }
            while (na != nb) { ca = na; na = parent[na]; cb = nb; nb = parent[nb]; }
            if (na == node_of[a]) r = make_uint4(a, a, 3u, 0u);              // a is b's ancestor: met first on the way down
            else if (na == node_of[b]) r = make_uint4(b, b, 3u, 0u);
            else {
                const uint32_t dim = nd[4 * na + 1];
                float coord = vt[3 * (size_t)nd[4 * na] + dim]; uint32_t cbits; memcpy(&cbits, &coord, 4);
                const bool a_left = nd[4 * na + 2] == ca;
                if (!a_left && nd[4 * na + 2] != cb) return 0;
                r = a_left ? make_uint4(a, b, dim, cbits) : make_uint4(b, a, dim, cbits);
            }
        }
        pos.push_back(make_float4(vt[3 * (size_t)a], vt[3 * (size_t)a + 1], vt[3 * (size_t)a + 2], 0.0f));
        rule.push_back(r);
        i = j;
    }
    const uint32_t m = (uint32_t)pos.size();
    if (m < 2 || m > 65000) return 0;
    // neighbour lists: everything closer than D = distance of the first position left out; at most 8 entries
    std::vector<uint16_t> nbr(8 * (size_t)m, 0xffffu);
    std::vector<std::pair<double, uint32_t>> ds(m);
    for (uint32_t i = 0; i < m; ++i) {
        for (uint32_t j = 0; j < m; ++j) {
            const double dx = (double)pos[i].x - pos[j].x, dy = (double)pos[i].y - pos[j].y, dz = (double)pos[i].z - pos[j].z;
            ds[j] = std::make_pair(j == i ? 1e300 : sqrt(dx * dx + dy * dy + dz * dz), j);
        }
        const uint32_t top = std::min<uint32_t>(m - 1, 9);
        std::partial_sort(ds.begin(), ds.begin() + top, ds.end());
        uint32_t k = 0;
        while (k < top && k < 8 && ds[k].first <= 1.6 * ds[0].first) ++k;
        const double D = k < top ? ds[k].first : (k < m - 1 ? ds[k].first : 1e30);
        for (uint32_t t = 0; t < k; ++t) nbr[8 * (size_t)i + t] = (uint16_t)ds[t].second;
        float halfD = (float)std::min(0.5 * D - 1e-5, 1e30);
        halfD = nextafterf(halfD, -1.0f);
        pos[i].w = halfD > 0.0f ? halfD : 0.0f;
    }
    // cube-map proposals: nearest distinct position of every cell centre
    const uint32_t res = 128;
    std::vector<uint16_t> table(6 * (size_t)res * res);
    for (uint32_t face = 0; face < 6; ++face)
        for (uint32_t iv = 0; iv < res; ++iv) {
            uint32_t last = 0;
            for (uint32_t iu = 0; iu < res; ++iu) {
                const double u = (iu + 0.5) / res * 2.0 - 1.0, v = (iv + 0.5) / res * 2.0 - 1.0, sgn = (face & 1) ? -1.0 : 1.0;
                double d[3];
                if (face < 2) { d[0] = sgn; d[1] = u; d[2] = v; } else if (face < 4) { d[0] = u; d[1] = sgn; d[2] = v; } else { d[0] = u; d[1] = v; d[2] = sgn; }
                const double l = sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
                for (int k = 0; k < 3; ++k) d[k] /= l;
                // greedy walk over the neighbour lists from the previous cell's answer (a proposal only: a poor one costs a
                // hop or the plain search on the device, never a wrong answer)
                auto dist2 = [&](uint32_t j) { const double a = d[0] - pos[j].x, b = d[1] - pos[j].y, c = d[2] - pos[j].z; return a * a + b * b + c * c; };
                uint32_t best = last; double bd = dist2(best);
                for (bool moved = true; moved;) {
                    moved = false;
                    for (int t = 0; t < 8; ++t) {
                        const uint16_t j = nbr[8 * (size_t)best + t];
                        if (j != 0xffffu && dist2(j) < bd) { bd = dist2(j); best = j; moved = true; break; }
                    }
                }
                table[((size_t)face * res + iv) * res + iu] = (uint16_t)best;
                last = best;
            }
        }
    UAV_CUDA(cudaMalloc(&kd->fast_table, sizeof(uint16_t) * table.size()));
    UAV_CUDA(cudaMalloc(&kd->fast_pos, sizeof(float4) * (size_t)m));
    UAV_CUDA(cudaMalloc(&kd->fast_rule, sizeof(uint4) * (size_t)m));
    UAV_CUDA(cudaMalloc(&kd->fast_nbr, sizeof(uint16_t) * nbr.size()));
    UAV_CUDA(cudaMemcpy(kd->fast_table, table.data(), sizeof(uint16_t) * table.size(), cudaMemcpyHostToDevice));
    UAV_CUDA(cudaMemcpy(kd->fast_pos, pos.data(), sizeof(float4) * (size_t)m, cudaMemcpyHostToDevice));
    UAV_CUDA(cudaMemcpy(kd->fast_rule, rule.data(), sizeof(uint4) * (size_t)m, cudaMemcpyHostToDevice));
    UAV_CUDA(cudaMemcpy(kd->fast_nbr, nbr.data(), sizeof(uint16_t) * nbr.size(), cudaMemcpyHostToDevice));
    kd->fast_m = m; kd->fast_res = res; kd->fast_state = 1;
    return 0;
}

}  // namespace uav

using namespace uav;

extern "C" {

// The device search keeps one stack slot per level with two children (kd.cuh: stack[KD_STACK]; the reference has 6 + 64,
// libs/cacc/nnsearch.h:28-33, unchecked).  User-supplied topology is validated here instead of overflowing there:
// indices in range, every node reached at most once from the root, depth within the stack.
static int kd_validate(const uint32_t * nd, uint32_t n)
{
    std::vector<uint32_t> depth(n, 0), order; order.reserve(n);
    std::vector<char> seen(n, 0);
    order.push_back(0); seen[0] = 1;
    for (size_t h = 0; h < order.size(); ++h) {
        const uint32_t i = order[h];
        UAV_REQUIRE(nd[4 * (size_t)i] < n, "KD node refers to a vertex outside the array");
        UAV_REQUIRE(nd[4 * (size_t)i + 1] < 3, "KD node split dimension outside 0..2");
        for (int c = 2; c < 4; ++c) {
            const uint32_t ch = nd[4 * (size_t)i + c];
            if (ch == UAV_NAI) continue;
            UAV_REQUIRE(ch < n, "KD child index outside the array");
            UAV_REQUIRE(!seen[ch], "KD nodes do not form a tree (a node is reachable twice)");
            seen[ch] = 1; depth[ch] = depth[i] + 1; order.push_back(ch);
            UAV_REQUIRE(depth[ch] < (uint32_t)KD_STACK, "KD-tree deeper than the search stack (40 levels)");
        }
    }
    return 0;
}

static int kdtree_fill(uavmvs_kdtree * k, const uint32_t * nodes, const float * verts, uint32_t n)
{
    k->num_nodes = n;
    UAV_CUDA(cudaMalloc(&k->nodes, sizeof(uint4) * (size_t)n));
    UAV_CUDA(cudaMalloc(&k->verts, sizeof(float4) * (size_t)n));
    UAV_CUDA(cudaMemcpy(k->nodes, nodes, sizeof(uint4) * (size_t)n, cudaMemcpyHostToDevice));
    if (upload3(verts, n, k->verts)) return 1;
    k->h_nodes.assign(nodes, nodes + 4 * (size_t)n);
    k->h_verts.assign(verts, verts + 3 * (size_t)n);
    return kd_fast_ensure(k);                 // built here, read-only afterwards: host threads share the handle
}

int uavmvs_kdtree_create(const uint32_t * nodes, const float * verts, uint32_t n, uavmvs_kdtree ** out)
{
    UAV_REQUIRE(nodes && verts && out && n >= 1, "bad argument");
    if (int rc = kd_validate(nodes, n)) return rc;
    uavmvs_kdtree * k = new uavmvs_kdtree();
    if (int rc = kdtree_fill(k, nodes, verts, n)) { uavmvs_kdtree_destroy(k); return rc; }
    *out = k;
    return 0;
}

int uavmvs_kdtree_destroy(uavmvs_kdtree * k)
{
    if (!k) return 0;
    cudaFree(k->nodes); cudaFree(k->verts);
    for (const uavmvs_kdtree::FrustumBits & e : k->fr_cache) cudaFree(e.bits);
    cudaFree(k->fast_table); cudaFree(k->fast_pos); cudaFree(k->fast_rule); cudaFree(k->fast_nbr);
    delete k;
    return 0;
}

int uavmvs_cloud_create(const float * vertices, const float * normals, const float * values,
    const float * qualities, uint32_t n, uavmvs_cloud ** out)
{
    UAV_REQUIRE(vertices && normals && out, "null argument");
    UAV_REQUIRE(n >= 1, "empty cloud");
    uavmvs_cloud * c = new uavmvs_cloud();
    c->n = n; c->num_blocks = divup(n, 32);
    UAV_CUDA(cudaMalloc(&c->vertices, sizeof(float4) * (size_t)n));
    UAV_CUDA(cudaMalloc(&c->normals, sizeof(float4) * (size_t)n));
    UAV_CUDA(cudaMalloc(&c->values, sizeof(float) * (size_t)n));
    UAV_CUDA(cudaMalloc(&c->qualities, sizeof(float) * (size_t)n));
    UAV_CUDA(cudaMalloc(&c->block_sphere, sizeof(float4) * (size_t)c->num_blocks));
    UAV_CUDA(cudaMalloc(&c->block_lo, sizeof(float4) * (size_t)c->num_blocks));
    UAV_CUDA(cudaMalloc(&c->block_hi, sizeof(float4) * (size_t)c->num_blocks));
    UAV_CUDA(cudaMalloc(&c->block_cone, sizeof(float4) * (size_t)c->num_blocks));
    UAV_CUDA(cudaMalloc(&c->block_obb, sizeof(float4) * 5 * (size_t)c->num_blocks));
    UAV_CUDA(cudaMalloc(&c->block_nmax, sizeof(float) * (size_t)c->num_blocks));
    c->num_super = divup(c->num_blocks, 32);
    UAV_CUDA(cudaMalloc(&c->super_sphere, sizeof(float4) * (size_t)c->num_super));
    if (upload3(vertices, n, c->vertices)) return 1;
    if (upload3(normals, n, c->normals)) return 1;
    for (int a = 0; a < 3; ++a) c->bound_abs[a] = 0.0f;
    for (size_t i = 0; i < 3 * (size_t)n; ++i) {
        const float m = fabsf(vertices[i]);
        if (!(m <= c->bound_abs[i % 3])) c->bound_abs[i % 3] = m;     // NaN sticks: the proof then never fires
    }
    if (values) UAV_CUDA(cudaMemcpy(c->values, values, 4 * (size_t)n, cudaMemcpyHostToDevice));
    else k_fill<<<divup(n, 256), 256>>>(c->values, n, 0.0f);        /* libs/util/cio.h:45 */
    if (qualities) UAV_CUDA(cudaMemcpy(c->qualities, qualities, 4 * (size_t)n, cudaMemcpyHostToDevice));
    else k_fill<<<divup(n, 256), 256>>>(c->qualities, n, 1.0f);     /* libs/util/cio.h:46 */
    k_block_spheres<<<divup((uint64_t)c->num_blocks * 32, 128), 128>>>(c->vertices, n, c->block_sphere,
        c->block_lo, c->block_hi);
    k_block_cones<<<divup((uint64_t)c->num_blocks * 32, 128), 128>>>(c->normals, n, c->block_cone, c->block_nmax);
    k_block_obb<<<divup((uint64_t)c->num_blocks * 32, 128), 128>>>(c->vertices, n, c->block_sphere, c->block_cone, c->block_obb);
    k_super_spheres<<<divup((uint64_t)c->num_super * 32, 128), 128>>>(c->block_sphere, c->num_blocks, c->super_sphere);
    UAV_CUDA(cudaGetLastError());
    UAV_CUDA(cudaDeviceSynchronize());
    *out = c;
    return 0;
}

int uavmvs_cloud_download_attributes(const uavmvs_cloud * c, float * values, float * qualities)
{
    UAV_REQUIRE(c, "null cloud");
    if (values) UAV_CUDA(cudaMemcpy(values, c->values, 4 * (size_t)c->n, cudaMemcpyDeviceToHost));
    if (qualities) UAV_CUDA(cudaMemcpy(qualities, c->qualities, 4 * (size_t)c->n, cudaMemcpyDeviceToHost));
    return 0;
}

int uavmvs_cloud_upload_attributes(uavmvs_cloud * c, const float * values, const float * qualities)
{
    UAV_REQUIRE(c, "null cloud");
    if (values) UAV_CUDA(cudaMemcpy(c->values, values, 4 * (size_t)c->n, cudaMemcpyHostToDevice));
    if (qualities) UAV_CUDA(cudaMemcpy(c->qualities, qualities, 4 * (size_t)c->n, cudaMemcpyHostToDevice));
    return 0;
}

int uavmvs_cloud_size(const uavmvs_cloud * c, uint32_t * n) { UAV_REQUIRE(c && n, "null"); *n = c->n; return 0; }

int uavmvs_cloud_destroy(uavmvs_cloud * c)
{
    if (!c) return 0;
    cudaFree(c->vertices); cudaFree(c->normals); cudaFree(c->values); cudaFree(c->qualities);
    cudaFree(c->block_sphere); cudaFree(c->block_lo); cudaFree(c->block_hi); cudaFree(c->block_obb); cudaFree(c->super_sphere); cudaFree(c->block_cone); cudaFree(c->block_nmax);
    delete c;
    return 0;
}

int uavmvs_obs_create(uint32_t num_cols, uint32_t max_rows, uavmvs_obs ** out)
{
    UAV_REQUIRE(out && num_cols >= 1 && max_rows >= 1 && max_rows <= 32, "bad argument (max_rows must be 1..32)");
    uavmvs_obs * o = new uavmvs_obs();
    o->num_cols = num_cols; o->max_rows = max_rows;
    o->pitch = (((uint64_t)num_cols * 16 + 511) / 512) * 512;   /* cudaMallocPitch-like row alignment */
    UAV_CUDA(cudaMalloc(&o->data, o->pitch * max_rows));
    UAV_CUDA(cudaMalloc(&o->num_rows, 4 * (size_t)num_cols));
    UAV_CUDA(cudaMemset(o->data, 0, o->pitch * max_rows));
    UAV_CUDA(cudaMemset(o->num_rows, 0, 4 * (size_t)num_cols));
    UAV_CUDA(cudaMalloc(&o->dirty, 4 * (size_t)divup(num_cols, 32)));
    UAV_CUDA(cudaMemset(o->dirty, 0xff, 4 * (size_t)divup(num_cols, 32)));
    *out = o;
    return 0;
}

int uavmvs_obs_clear(void * stream, uavmvs_obs * o)
{
    UAV_REQUIRE(o, "null obs");
    UAV_CUDA(cudaMemsetAsync(o->data, 0, o->pitch * o->max_rows, (cudaStream_t)stream));
    UAV_CUDA(cudaMemsetAsync(o->num_rows, 0, 4 * (size_t)o->num_cols, (cudaStream_t)stream));
    UAV_CUDA(cudaMemsetAsync(o->dirty, 0xff, 4 * (size_t)divup(o->num_cols, 32), (cudaStream_t)stream));
    return 0;
}

int uavmvs_obs_mark_dirty(void * stream, uavmvs_obs * o)
{
    UAV_REQUIRE(o, "null obs");
    UAV_CUDA(cudaMemsetAsync(o->dirty, 0xff, 4 * (size_t)divup(o->num_cols, 32), (cudaStream_t)stream));
    return 0;
}

int uavmvs_obs_download(const uavmvs_obs * o, float * data, uint32_t * num_rows)
{
    UAV_REQUIRE(o, "null obs");
    if (data) UAV_CUDA(cudaMemcpy2D(data, (size_t)o->num_cols * 16, o->data, o->pitch,
        (size_t)o->num_cols * 16, o->max_rows, cudaMemcpyDeviceToHost));
    if (num_rows) UAV_CUDA(cudaMemcpy(num_rows, o->num_rows, 4 * (size_t)o->num_cols, cudaMemcpyDeviceToHost));
    return 0;
}

int uavmvs_obs_upload(uavmvs_obs * o, const float * data, const uint32_t * num_rows)
{
    UAV_REQUIRE(o, "null obs");
    if (data) UAV_CUDA(cudaMemcpy2D(o->data, o->pitch, data, (size_t)o->num_cols * 16,
        (size_t)o->num_cols * 16, o->max_rows, cudaMemcpyHostToDevice));
    if (num_rows) UAV_CUDA(cudaMemcpy(o->num_rows, num_rows, 4 * (size_t)o->num_cols, cudaMemcpyHostToDevice));
    UAV_CUDA(cudaMemset(o->dirty, 0xff, 4 * (size_t)divup(o->num_cols, 32)));
    return 0;
}

int uavmvs_obs_device_ptr(const uavmvs_obs * o, int which, void ** ptr, uint64_t * pitch_bytes)
{
    UAV_REQUIRE(o && ptr, "null argument");
    if (which == 0) { *ptr = o->data; if (pitch_bytes) *pitch_bytes = o->pitch; }
    else if (which == 1) { *ptr = o->num_rows; if (pitch_bytes) *pitch_bytes = 0; }
    else UAV_REQUIRE(false, "which must be 0 or 1");
    return 0;
}

int uavmvs_obs_destroy(uavmvs_obs * o)
{
    if (!o) return 0;
    cudaFree(o->data); cudaFree(o->num_rows); cudaFree(o->dirty);
    delete o;
    return 0;
}

int uavmvs_array_create(uint32_t n, uavmvs_array ** out)
{
    UAV_REQUIRE(out && n >= 1, "bad argument");
    uavmvs_array * a = new uavmvs_array();
    a->n = n;
    UAV_CUDA(cudaMalloc(&a->data, 4 * (size_t)n));
    UAV_CUDA(cudaMemset(a->data, 0, 4 * (size_t)n));
    *out = a;
    return 0;
}

int uavmvs_array_null(void * stream, uavmvs_array * a)
{
    UAV_REQUIRE(a, "null array");
    UAV_CUDA(cudaMemsetAsync(a->data, 0, 4 * (size_t)a->n, (cudaStream_t)stream));
    return 0;
}

int uavmvs_array_download(const uavmvs_array * a, float * host)
{
    UAV_REQUIRE(a && host, "null argument");
    UAV_CUDA(cudaMemcpy(host, a->data, 4 * (size_t)a->n, cudaMemcpyDeviceToHost));
    return 0;
}

int uavmvs_array_upload(uavmvs_array * a, const float * host)
{
    UAV_REQUIRE(a && host, "null argument");
    UAV_CUDA(cudaMemcpy(a->data, host, 4 * (size_t)a->n, cudaMemcpyHostToDevice));
    return 0;
}

int uavmvs_array_destroy(uavmvs_array * a) { if (a) { cudaFree(a->data); delete a; } return 0; }

int uavmvs_image_create(uint32_t width, uint32_t height, uavmvs_image ** out)
{
    UAV_REQUIRE(out && width >= 1 && height >= 1, "bad argument");
    uavmvs_image * im = new uavmvs_image();
    im->width = width; im->height = height;
    im->pitch = (((uint64_t)width * 4 + 511) / 512) * 512;
    UAV_CUDA(cudaMalloc(&im->data, im->pitch * height));
    UAV_CUDA(cudaMemset(im->data, 0, im->pitch * height));
    *out = im;
    return 0;
}

int uavmvs_image_download(const uavmvs_image * im, float * host)
{
    UAV_REQUIRE(im && host, "null argument");
    UAV_CUDA(cudaMemcpy2D(host, (size_t)im->width * 4, im->data, im->pitch, (size_t)im->width * 4,
        im->height, cudaMemcpyDeviceToHost));
    return 0;
}

int uavmvs_image_destroy(uavmvs_image * im) { if (im) { cudaFree(im->data); delete im; } return 0; }

}  // extern "C"

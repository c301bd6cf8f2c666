// K7 estimate_capture_difficulty (kernels.cu:448-494) on the 4-wide fp16 tree of K1.
//
// The reference casts, per point, one CLOSEST-HIT ray per icosphere direction of its hemisphere and asks one thing of the
// result: is the closest hit a face of the proxy mesh (face id < mesh_size: `sum += ctheta`) or not (no hit, or a face of
// the airspace mesh appended behind it: `max_ctheta`).  The reference's walk (bvh_tree.h:419-467) shrinks tmax as it goes,
// so which triangles it tests depends on its visiting order.  What it RETURNS does not, except in two rare situations:
//
//   Let C be the triangles the reference accepts at the ORIGINAL interval [tmin, tmax] whose own leaf box passes the
//   exact slab test at that interval.  On a nested tree (k_check_tree: child boxes inside parent boxes, leaf boxes tight)
//   the slab test is monotone in the box bounds and in tmax, so every triangle the reference ever tests lies in C, and
//   its result has t >= T* = min over C.  If a triangle with t = T* has a leaf box that passes at [tmin, T*] ("tight"),
//   the reference reaches that leaf whatever it has found before (its current bound is >= T*, the leaf and hence all its
//   ancestors pass) and ends with t = T*; ties replace (`!(t > best)`), so its face is SOME triangle with t = T*.
//   Hence: class of the reference's hit = class of the triangles at T*, provided (a) they are all of one class, i.e.
//   min t over the mesh faces != min t over the other faces, and (b) one of them is tight.
//
// k_k7_trace enumerates the part of C that matters with the conservative walk of k_trace and a bound that shrinks WITH A
// MARGIN: nodes are tested against [tmin, min(tmax, t_found + M)], t_found the smallest accepted t so far.  A tight triangle
// at T* <= t_found is never cut off (its leaf passes at [tmin, T*]).  A triangle of C that is not tight was accepted
// outside its own leaf box, which the +-1e-3 barycentric slack of primitives.h:240-244 allows: its hit point lies within
// 1e-3 (|ab| + |ac|) of the triangle, hence of the box, so the ray enters that box no later than t + diag(box) + slack;
// with M = 1.02 max leaf-box diagonal + 1e-3 every such triangle below t_found is still reached.  So the smallest t per
// class and its tightness come out as if all of C had been enumerated.  The kernel keeps the smallest t and its tightness per class, and writes one status byte per (point,
// direction): 1 mesh, 2 free, 3 "order matters here" (a tie between the classes, or a closest hit outside its own leaf
// box: the +-1e-3 barycentric slack of primitives.h:240-244 allows that).  k_k7_finish then walks the directions of a
// point in the reference's order with the reference's additions, and runs the reference walk itself for status 3.
//
// Directions with a component |d_k| <= 1e-4 (15 % of an icosphere: its vertices on the coordinate planes, d_k = 0) cannot use
// fl(1 / d_k).  Along such a ray coordinate k moves by at most 1e-4 tmax, so the reference can only accept a box whose slab
// [mn_k, mx_k], widened by dlt = 1.001e-4 tmax, holds o_k (for d_k = 0 its (b - o) / 0 = +-inf says the same of the closed
// slab).  The node test keeps its form with rd_k = +-1e30: the two plane values of that axis are then either hugely
// negative / positive (origin inside: no constraint) or both hugely positive / negative (outside: rejected); the slab is
// widened by dlt + 2e-4 + 1e-6 (|o_k| + B_k), which also covers the cancellation error of the 1e30-scaled planes (3e-5 m),
// through the ray constant nox (near side) and a per-axis band em_k (far side).  Dropping the axis instead would be
// conservative too, but a ray then visits every node of the scene along it (measured: 100x slower).
#include "common.cuh"
#include "trace.cuh"
#include "kd.cuh"

namespace uav {

constexpr int K7_WARPS = 4;
constexpr int K7_PEND = 6;            // deferred leaves per lane
constexpr int K7_REFILL = 20;         // refill when this many lanes are idle (k_trace's measured optimum)
constexpr int K7_TRI = 12;            // triangle phase when this many lanes hold a leaf
constexpr int K7_STEPS = 12;          // node steps between two rounds of bookkeeping (tools/sweep_k7.sh: 6 -> 40.8 ms per 2^18 points, 12 -> 39.0, 24 -> 39.1;
                                      // refill 4 / 8 / 12 / 20 / 24 -> 43.4 / 42.2 / 41.6 / 40.8 / 39.0 with 16 steps)
constexpr uint32_t K7_UNITS_PER_GRAB = 8;   // (block, direction) units per atomic = blocks per group
// Unit order: groups of 8 consecutive 32-point blocks; within a group direction-major, so that one grab is ONE direction over
// 8 neighbouring blocks: the rays a warp holds are parallel and start next to each other (Z-ordered cloud), like the rays of
// one block towards one camera in K1.  (Block-major order, a warp walking one block through consecutive icosphere indices,
// mixes unrelated directions in a warp.)
__device__ __forceinline__ void k7_unit(uint32_t u, uint32_t S, uint32_t & b, uint32_t & i)
{
    const uint32_t per = K7_UNITS_PER_GRAB * S, g = u / per, r = u - g * per;
    i = r / K7_UNITS_PER_GRAB;
    b = g * K7_UNITS_PER_GRAB + (r - i * K7_UNITS_PER_GRAB);
}

struct K7Args {
    BvhView bvh;
    KdView kd;
    const float4 * __restrict__ verts; const float4 * __restrict__ normals;
    uint32_t n;                       // points of the cloud
    uint32_t p0, np;                  // this launch: points [p0, p0 + np), p0 a multiple of 32
    uint32_t mesh_size;
    float max_distance;
    float margin;                     // M above: 1.02 x the largest leaf-box diagonal of the tree + 1e-3
    uint8_t * status;                 // [kd.n][np] : 0 not cast, 1 mesh, 2 free, 3 reference walk needed
    uint32_t * work;                  // unit counter
    int refill, steps, tri;           // K7_REFILL / K7_STEPS / K7_TRI (UAVMVS_K7_REFILL / _STEPS / _TRI for sweeps)
};

// the leaf's stored box against the reference's slab test at [tmin, tmax]; `fast` = the filtered form may be used
static __device__ __noinline__ bool k7_leaf_box(const float4 * __restrict__ leafbox, uint32_t first, const Ray & ray, V3 rd,
    bool fast, float tmax)
{
    const float4 b0 = __ldg(leafbox + 2 * (size_t)first), b1 = __ldg(leafbox + 2 * (size_t)first + 1);
    RayQ q; q.r = ray; q.r.tmax = tmax; q.rd = rd; q.fast = fast;
    float t;
    return slab_bool(q, b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, t);
}


__global__ void __launch_bounds__(K7_WARPS * 32, 6)
k_k7_trace(K7Args a)
{
    __shared__ uint32_t s_pend[K7_PEND][K7_WARPS * 32];     // deferred leaves: first << 6 | count
    const uint32_t tid = threadIdx.x, lane = tid & 31;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const uint32_t S = a.kd.n;
    const uint32_t total_units = (((a.np + 31) / 32 + K7_UNITS_PER_GRAB - 1) / K7_UNITS_PER_GRAB) * K7_UNITS_PER_GRAB * S;   // blocks beyond the last hold no points
    const float l = a.max_distance;
    uint32_t stack[K1_STACK];
    RayT rt; V3 rd = mk(0.f, 0.f, 1.f), rq = rd;
    float emx = 0.f, emy = 0.f, emz = 0.f;                  // per-axis far-side band: rt.em, + twice the widening on a parallel axis
    rt.ox = rt.oy = rt.oz = 0.f; rt.rdx = rt.rdy = rt.rdz = 1.f; rt.nox = rt.noy = rt.noz = 0.f; rt.tmin = 0.f; rt.tmax = 0.f; rt.em = 0.f;
    bool has_ray = false, traversing = false, fastq = false, drained = false;
    uint32_t sp = 0, npend = 0, cur = 0, out_idx = 0, flags = 0;
    float best_m = UAV_INF, best_o = UAV_INF;               // smallest accepted t: mesh faces / the others
    float tcut = 0.f;                                       // upper end of the node tests' interval: min(tmax, smallest accepted t + M)
    uint32_t selx = 0x3210u, sely = 0x3210u, selz = 0x3210u;
    uint32_t unit = 0, unit_end = 0, cu = 0, um = 0;         // this warp's grab of units, current unit, its points still to hand out
    uint32_t serviced_idle = 0;

    for (;;) {
        const uint32_t idle = __ballot_sync(0xffffffffu, !has_ray);
        if (idle != serviced_idle && __popc(idle) >= a.refill) {
            // ---------------- refill idle lanes: the next cast (point, direction) pairs in unit order ----------------
            const uint32_t need = __popc(idle), rank = __popc(idle & lt_mask);
            uint32_t given = 0, my_unit = 0, my_p = 0; bool mine = false;
            while (given < need && !drained) {
                if (um == 0) {
                    if (unit == unit_end) {
                        uint32_t g = 0;
                        if (lane == 0) g = atomicAdd(a.work, K7_UNITS_PER_GRAB);
                        g = __shfl_sync(0xffffffffu, g, 0);
                        if (g >= total_units) { drained = true; break; }
                        unit = g; unit_end = min(g + K7_UNITS_PER_GRAB, total_units);
                    }
                    cu = unit++;
                    uint32_t b, i; k7_unit(cu, S, b, i);
                    const uint32_t pl = b * 32 + lane;                   // lane = point of the block
                    bool cast = false;
                    if (pl < a.np) {                                     // kernels.cu:468-470
                        const V3 dir = normalized(ld3(__ldg(a.kd.verts + i)));
                        cast = !(dot(dir, ld3(a.normals[a.p0 + pl])) < 0.087f);
                    }
                    um = __ballot_sync(0xffffffffu, cast);
                    continue;
                }
                const uint32_t have = __popc(um), take = min(need - given, have);
                if (!has_ray && rank >= given && rank < given + take) { my_p = __fns(um, 0, (int)(rank - given) + 1); my_unit = cu; mine = true; }
                if (take == have) um = 0;
                else um &= ~((2u << __fns(um, 0, (int)take)) - 1u);      // drop its first `take` points
                given += take;
            }
            if (mine) {
                uint32_t b, i; k7_unit(my_unit, S, b, i);
                const uint32_t pl = b * 32 + my_p;
                out_idx = i * a.np + pl;
                const V3 o = ld3(a.verts[a.p0 + pl]);
                rd = normalized(ld3(__ldg(a.kd.verts + i)));
                rt.ox = o.x; rt.oy = o.y; rt.oz = o.z;
                rt.tmin = l * 0.001f; rt.tmax = l;                       // kernels.cu:475-476
                const bool kx = fabsf(rd.x) > 1e-4f, ky = fabsf(rd.y) > 1e-4f, kz = fabsf(rd.z) > 1e-4f;
                rq = mk(1.0f / rd.x, 1.0f / rd.y, 1.0f / rd.z);          // for the filtered leaf-box test (fastq only)
                const float R = 1e30f;                                   // parallel axis: plane values only carry a sign
                rt.rdx = kx ? rq.x : (rd.x < 0.0f ? -R : R); rt.rdy = ky ? rq.y : (rd.y < 0.0f ? -R : R); rt.rdz = kz ? rq.z : (rd.z < 0.0f ? -R : R);
                rt.nox = -(rt.ox * rt.rdx); rt.noy = -(rt.oy * rt.rdy); rt.noz = -(rt.oz * rt.rdz);
                // the band of ray_fma_setup80 over the axes that carry real plane parameters
                rt.em = 4.8e-7f * ((kx ? (fabsf(rt.ox) + a.bvh.bx) * fabsf(rt.rdx) : 0.0f) + (ky ? (fabsf(rt.oy) + a.bvh.by) * fabsf(rt.rdy) : 0.0f)
                    + (kz ? (fabsf(rt.oz) + a.bvh.bz) * fabsf(rt.rdz) : 0.0f)) + 4e-6f * rt.tmax;
                const float dl = 1.001e-4f * l + 2e-4f;
                const float wx = kx ? 0.0f : (dl + 1e-6f * (fabsf(rt.ox) + a.bvh.bx)) * R;   // widening of a parallel axis, in plane units
                const float wy = ky ? 0.0f : (dl + 1e-6f * (fabsf(rt.oy) + a.bvh.by)) * R;
                const float wz = kz ? 0.0f : (dl + 1e-6f * (fabsf(rt.oz) + a.bvh.bz)) * R;
                rt.nox -= wx; rt.noy -= wy; rt.noz -= wz;                // near planes lowered by w ...
                emx = rt.em + 2.0f * wx; emy = rt.em + 2.0f * wy; emz = rt.em + 2.0f * wz;   // ... far planes raised by w
                fastq = kx && ky && kz;
                selx = rt.rdx < 0.0f ? 0x1032u : 0x3210u; sely = rt.rdy < 0.0f ? 0x1032u : 0x3210u; selz = rt.rdz < 0.0f ? 0x1032u : 0x3210u;
                has_ray = true; traversing = true; sp = 0; npend = 0; cur = 0; flags = 0;
                best_m = UAV_INF; best_o = UAV_INF; tcut = rt.tmax;
            }
            serviced_idle = __ballot_sync(0xffffffffu, !has_ray);
            if (serviced_idle == 0xffffffffu) {
                if (drained) break;
                serviced_idle = 0;
                continue;
            }
        }

        // ---------------- node steps: k_trace's conservative test on the Node80 tree, original interval throughout ----------------
        #pragma unroll 1
        for (int step = 0; step < a.steps; ++step) {
            if (traversing && npend <= K7_PEND - 4) {
                const Node80 * np = a.bvh.wide80 + cur;
                const float4 c4 = __ldg(&np->c);
                const uint4 ha = __ldg(&np->h[0]), hb = __ldg(&np->h[1]), hc = __ldg(&np->h[2]);
                const uint4 cnt = __ldg(&np->cnt);
                const float crx = __fmaf_rn(c4.x, rt.rdx, rt.nox), cry = __fmaf_rn(c4.y, rt.rdy, rt.noy), crz = __fmaf_rn(c4.z, rt.rdz, rt.noz);
                const float cfx = crx + emx, cfy = cry + emy, cfz = crz + emz, tmaxe = tcut + rt.em;
                #define UAV_SLAB(wx, wy, wz, hit) { \
                    const uint32_t ux = __byte_perm((wx), 0u, selx), uy = __byte_perm((wy), 0u, sely), uz = __byte_perm((wz), 0u, selz); \
                    const float2 fx = __half22float2(*reinterpret_cast<const __half2 *>(&ux)); \
                    const float2 fy = __half22float2(*reinterpret_cast<const __half2 *>(&uy)); \
                    const float2 fz = __half22float2(*reinterpret_cast<const __half2 *>(&uz)); \
                    const float tn = fmaxf(fmaxf(__fmaf_rn(fx.x, rt.rdx, crx), __fmaf_rn(fy.x, rt.rdy, cry)), fmaxf(__fmaf_rn(fz.x, rt.rdz, crz), rt.tmin)); \
                    const float tf = fminf(fminf(__fmaf_rn(fx.y, rt.rdx, cfx), __fmaf_rn(fy.y, rt.rdy, cfy)), fminf(__fmaf_rn(fz.y, rt.rdz, cfz), tmaxe)); \
                    hit = !(tf < tn); }
                bool h0, h1, h2, h3;
                UAV_SLAB(ha.x, ha.y, ha.z, h0)
                UAV_SLAB(ha.w, hb.x, hb.y, h1)
                UAV_SLAB(hb.z, hb.w, hc.x, h2)
                UAV_SLAB(hc.y, hc.z, hc.w, h3)
                #undef UAV_SLAB
                const bool f0 = h0 && (int)cnt.x >= 0, f1 = h1 && (int)cnt.y >= 0, f2 = h2 && (int)cnt.z >= 0, f3 = h3 && (int)cnt.w >= 0;
                const bool i0 = h0 && (int)cnt.x < 0, i1 = h1 && (int)cnt.y < 0, i2 = h2 && (int)cnt.z < 0, i3 = h3 && (int)cnt.w < 0;
                const uint4 ref = make_uint4(cnt.x & 0x7fffffffu, cnt.y & 0x7fffffffu, cnt.z & 0x7fffffffu, cnt.w & 0x7fffffffu);
                if (f0) { s_pend[npend][tid] = cnt.x; ++npend; }
                if (f1) { s_pend[npend][tid] = cnt.y; ++npend; }
                if (f2) { s_pend[npend][tid] = cnt.z; ++npend; }
                if (f3) { s_pend[npend][tid] = cnt.w; ++npend; }
                uint32_t nxt = i0 ? ref.x : UAV_NAI;
                if (i1 && nxt != UAV_NAI) stack[sp++] = nxt;
                if (i1) nxt = ref.y;
                if (i2 && nxt != UAV_NAI) stack[sp++] = nxt;
                if (i2) nxt = ref.z;
                if (i3 && nxt != UAV_NAI) stack[sp++] = nxt;
                if (i3) nxt = ref.w;
                if (nxt == UAV_NAI) {
                    if (sp > 0) nxt = stack[--sp]; else traversing = false;
                }
                cur = nxt;
            }
        }
        // ---------------- triangle phase: every accepted triangle of every deferred leaf ----------------
        {
            const uint32_t pmask = __ballot_sync(0xffffffffu, npend > 0);
            if (pmask) {
                const uint32_t tmask = __ballot_sync(0xffffffffu, traversing);
                const uint32_t full = __ballot_sync(0xffffffffu, npend > K7_PEND - 4);
                if (__popc(pmask) >= a.tri || full || !tmask) {
                    if (npend > 0) {
                        Ray ray; ray.o = mk(rt.ox, rt.oy, rt.oz); ray.d = rd; ray.tmin = rt.tmin; ray.tmax = rt.tmax;
                        for (uint32_t p = 0; p < npend; ++p) {
                            const uint32_t r = s_pend[p][tid];
                            const uint32_t first = r >> 6, count = r & 31u;
                            for (uint32_t i = 0; i < count; ++i) {
                                float t;
                                if (!hit_tri(ray, a.bvh.tris + first + i, t)) continue;
                                // in C only if its leaf's own box passes at the original interval; tight if it passes at [tmin, t]
                                const bool tight = k7_leaf_box(a.bvh.leafbox, first, ray, rq, fastq, t);
                                if (!tight && !k7_leaf_box(a.bvh.leafbox, first, ray, rq, fastq, ray.tmax)) continue;
                                tcut = fminf(tcut, t + a.margin);
                                const bool other = __ldg(a.bvh.indices + first + i) >= a.mesh_size;
                                const uint32_t bit = other ? 2u : 1u;
                                const float bt = other ? best_o : best_m;
                                if (t < bt) {
                                    if (other) best_o = t; else best_m = t;
                                    flags = (flags & ~bit) | (tight ? bit : 0u);
                                } else if (t == bt && tight) flags |= bit;
                            }
                        }
                        npend = 0;
                    }
                }
            }
        }
        // ---------------- finished lanes post their status ----------------
        if (has_ray && !traversing && npend == 0) {
            uint32_t st;
            if (best_m == UAV_INF && best_o == UAV_INF) st = 2u;                     // no hit
            else if (best_m == best_o) st = 3u;                                     // tie between the classes: order decides
            else if (best_m < best_o) st = (flags & 1u) ? 1u : 3u;
            else st = (flags & 2u) ? 2u : 3u;
            a.status[out_idx] = (uint8_t)st;
            has_ray = false;
        }
    }
}

// Per point, directions in the reference's order with the reference's additions (kernels.cu:463-493); the status byte
// replaces the trace, except for status 3 (and 0, which k_k7_trace never leaves on a cast pair): those run the
// reference walk here.  counters[0] += pairs that needed it.
__global__ void __launch_bounds__(64)
k_k7_finish(K7Args a, float * __restrict__ values, float * __restrict__ qualities, unsigned long long * counters)
{
    const uint32_t pl = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = pl < a.np;                              // whole warps stay in the collective trace
    const uint32_t id = a.p0 + (valid ? pl : a.np - 1);
    const V3 v = ld3(a.verts[id]), nrm = ld3(a.normals[id]);
    const float l = a.max_distance;
    float sum = 0.0f, max_ctheta = 0.0f;
    uint32_t walked = 0;
    for (uint32_t i = 0; i < a.kd.n; ++i) {
        const V3 dir = normalized(ld3(__ldg(a.kd.verts + i)));
        const float ctheta = dot(dir, nrm);
        const bool cast = valid && !(ctheta < 0.087f);
        if (!__any_sync(0xffffffffu, cast)) continue;
        const uint32_t st = cast ? a.status[(size_t)i * a.np + pl] : 2u;
        const bool walk = cast && (st == 3u || st == 0u);
        bool mesh = st == 1u;
        if (__any_sync(0xffffffffu, walk)) {
            Ray ray; ray.o = v; ray.d = dir; ray.tmin = l * 0.001f; ray.tmax = l;
            uint32_t face; float t;
            const bool hit = trace_closest_warp(a.bvh, ray, walk, face, t);
            if (walk) { mesh = hit && face < a.mesh_size; ++walked; }
        }
        if (cast) {
            if (!mesh) { if (ctheta > max_ctheta) max_ctheta = ctheta; }
            else sum += ctheta;
        }
    }
    if (valid) {
        const float mx = (a.kd.n / 2.0f) * 0.5f;
        values[id] = acosf(__saturatef(max_ctheta));
        qualities[id] = sum / mx;
    }
    if (counters) {
        walked = __reduce_add_sync(0xffffffffu, walked);
        if ((threadIdx.x & 31) == 0 && walked) atomicAdd(counters, (unsigned long long)walked);
    }
}

// ---- host side -----------------------------------------------------------------------------------------------------
struct K7Scratch { uint8_t * status = nullptr; size_t status_cap = 0; uint32_t * work = nullptr; unsigned long long * counters = nullptr; int grid = 0; };
static thread_local K7Scratch g_k7;

// 1 if the tree can take the wide walk (nested, not a single leaf, Node80 + leaf boxes present, fp16 PRMT layout)
int k7_wide_available(const uavmvs_bvh * bvh)
{
#if K1_PRMT && !K1_BF16
    return bvh->nested && !bvh->root_is_leaf && bvh->wide80 && bvh->leafbox;
#else
    (void)bvh; return 0;
#endif
}

int k7_wide(cudaStream_t s, float max_distance, const uavmvs_bvh * bvh, uint32_t mesh_size, const uavmvs_kdtree * kd, uavmvs_cloud * cloud)
{
    K7Scratch & k = g_k7;
    const uint32_t S = kd->num_nodes, n = cloud->n;
    if (!n || !S) return 0;
    if (!k.work) {
        UAV_CUDA(cudaMalloc(&k.work, sizeof(uint32_t)));
        UAV_CUDA(cudaMalloc(&k.counters, sizeof(unsigned long long)));
        UAV_CUDA(cudaMemset(k.counters, 0, sizeof(unsigned long long)));
        int dev = 0, sms = 0, occ = 0;
        UAV_CUDA(cudaGetDevice(&dev));
        UAV_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        UAV_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_k7_trace, K7_WARPS * 32, 0));
        k.grid = sms * (occ > 0 ? occ : 1);
    }
    // points per launch: one status byte per (direction, point); 2^18 points x 1272 directions = 333 MB
    uint32_t batch = 1u << 18;
    { const char * e = getenv("UAVMVS_K7_BATCH"); if (e && atoi(e) >= 32) batch = ((uint32_t)atoi(e) + 31u) / 32u * 32u; }
    while ((unsigned long long)(batch + 32ull * K7_UNITS_PER_GRAB) * S >= (1ull << 32)) batch /= 2;      // out_idx and the unit counter are 32-bit
    const size_t need = (size_t)std::min(batch, (n + 31u) / 32u * 32u) * S;
    if (need > k.status_cap) {
        if (k.status) UAV_CUDA(cudaFree(k.status));
        k.status = nullptr; k.status_cap = 0;
        UAV_CUDA(cudaMalloc(&k.status, need));
        k.status_cap = need;
    }
    for (uint32_t p0 = 0; p0 < n; p0 += batch) {
        K7Args a; a.bvh = view_of(bvh); a.kd = kd_view(kd); a.verts = cloud->vertices; a.normals = cloud->normals; a.n = n;
        a.p0 = p0; a.np = std::min(batch, n - p0); a.mesh_size = mesh_size; a.max_distance = max_distance;
        a.margin = 1.02f * bvh->max_leaf_diag + 1e-3f;
        a.status = k.status; a.work = k.work;
        a.refill = K7_REFILL; a.steps = K7_STEPS; a.tri = K7_TRI;
        if (const char * e = getenv("UAVMVS_K7_REFILL")) if (atoi(e) >= 1 && atoi(e) <= 32) a.refill = atoi(e);
        if (const char * e = getenv("UAVMVS_K7_STEPS")) if (atoi(e) >= 1 && atoi(e) <= 64) a.steps = atoi(e);
        if (const char * e = getenv("UAVMVS_K7_TRI")) if (atoi(e) >= 1 && atoi(e) <= 32) a.tri = atoi(e);
        UAV_CUDA(cudaMemsetAsync(k.status, 0, (size_t)a.np * S, s));
        UAV_CUDA(cudaMemsetAsync(k.work, 0, sizeof(uint32_t), s));
        k_k7_trace<<<k.grid, K7_WARPS * 32, 0, uav::counted(s)>>>(a);
        k_k7_finish<<<divup(a.np, 64), 64, 0, uav::counted(s)>>>(a, cloud->values, cloud->qualities, k.counters);
        UAV_CUDA(cudaGetLastError());
    }
    if (getenv("UAVMVS_K7_DEBUG")) {
        unsigned long long c = 0;
        UAV_CUDA(cudaStreamSynchronize(s));
        UAV_CUDA(cudaMemcpy(&c, k.counters, sizeof(c), cudaMemcpyDeviceToHost));
        UAV_CUDA(cudaMemset(k.counters, 0, sizeof(c)));
        fprintf(stderr, "[k7] pairs that took the reference walk: %llu\n", c);
    }
    return 0;
}

}  // namespace uav

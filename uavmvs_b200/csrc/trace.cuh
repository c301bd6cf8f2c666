// BVH traversal device functions.  Canon = libs/acc/bvh_tree.h:419-467 with the ray
// primitives of libs/acc/primitives.h:119-132 (slab) and :214-248 (triangle), which is
// also what libs/cacc/tracing.h:37-116 walks on the device.
//
// Any-hit result is independent of the visiting order: a leaf is reached iff every box
// on its root path (root excluded) passes the slab test against the ORIGINAL
// [tmin,tmax], and the query is true iff some reached leaf holds an accepted triangle.
// trace_any() exploits that (leaf children are tested at once, only inner nodes are
// stacked); trace_closest() keeps the reference's near-first order and its shrinking
// tmax because those decide which slack-zone hit wins.
#pragma once
#include <cuda_fp16.h>
#include "common.cuh"

namespace uav {

struct BvhView {
    const Node64 * __restrict__ inner;
    const Node128 * __restrict__ wide;
    const Node80 * __restrict__ wide80;
    float bx, by, bz;                         // max |coordinate| of the scene per axis
    const Tri64 * __restrict__ tris;
    const float * __restrict__ raw_tris;     // 9 floats per triangle, leaf order (reference layout)
    const uint32_t * __restrict__ indices;
    const float4 * __restrict__ leafbox;     // see uavmvs_bvh::leafbox (null for trees that are not nested)
    uint32_t root_is_leaf, root_first, root_count;
};

__host__ inline BvhView view_of(const uavmvs_bvh * b) {
    BvhView v;
    v.inner = b->inner; v.wide = b->wide; v.wide80 = b->wide80;
    v.bx = b->bound_abs[0]; v.by = b->bound_abs[1]; v.bz = b->bound_abs[2]; v.tris = b->ptris; v.raw_tris = b->tris; v.indices = b->indices; v.leafbox = b->leafbox;
    v.root_is_leaf = b->root_is_leaf; v.root_first = b->root_first; v.root_count = b->root_count;
    return v;
}

// libs/acc/primitives.h:119-132, true divisions, std::min/max NaN behaviour
__device__ __forceinline__ bool slab(const Ray & r, float mnx, float mny, float mnz,
    float mxx, float mxy, float mxz, float & tmin_out)
{
    float tmin = r.tmin, tmax = r.tmax;
    float t1, t2;
    t1 = (mnx - r.o.x) / r.d.x; t2 = (mxx - r.o.x) / r.d.x;
    tmin = smax(tmin, smin(smin(t1, t2), UAV_INF)); tmax = smin(tmax, smax(smax(t1, t2), -UAV_INF));
    t1 = (mny - r.o.y) / r.d.y; t2 = (mxy - r.o.y) / r.d.y;
    tmin = smax(tmin, smin(smin(t1, t2), UAV_INF)); tmax = smin(tmax, smax(smax(t1, t2), -UAV_INF));
    t1 = (mnz - r.o.z) / r.d.z; t2 = (mxz - r.o.z) / r.d.z;
    tmin = smax(tmin, smin(smin(t1, t2), UAV_INF)); tmax = smin(tmax, smax(smax(t1, t2), -UAV_INF));
    tmin_out = tmin;
    return tmax >= smax(tmin, 0.0f);
}

// libs/acc/primitives.h:214-248 on the pre-reduced triangle
__device__ __forceinline__ bool hit_tri(const Ray & r, const Tri64 * __restrict__ tp, float & t_out)
{
    float4 q0 = __ldg(&tp->q0), q1 = __ldg(&tp->q1), q2 = __ldg(&tp->q2), q3 = __ldg(&tp->q3);
    V3 a = mk(q0.x, q0.y, q0.z), ab = mk(q0.w, q1.x, q1.y), ac = mk(q1.z, q1.w, q2.x);
    V3 n = mk(q2.y, q2.z, q2.w);
    float d00 = q3.x, d01 = q3.y, d11 = q3.z, denom = q3.w;
    float cosine = dot_nz(n, r.d);
    if (fabsf(cosine) < FLT_EPSILON) return false;      // also rejects degenerate (n == 0)
    float t = -dot_nz(n, r.o - a) / cosine;
    if (t < r.tmin || r.tmax < t) return false;
    V3 p = r.o + r.d * t;
    V3 ap = p - a;
    float d20 = dot_nz(ap, ab), d21 = dot_nz(ap, ac);
    float b1 = (d11 * d20 - d01 * d21) / denom;
    float b2 = (d00 * d21 - d01 * d20) / denom;
    float b0 = 1.0f - b1 - b2;
    const float eps = 1e-3f;
    if (-eps > b0 || b0 > 1.0f + eps) return false;
    if (-eps > b1 || b1 > 1.0f + eps) return false;
    if (-eps > b2 || b2 > 1.0f + eps) return false;
    t_out = t;
    return true;
}

#ifndef UAV_STACK
#define UAV_STACK 96
#endif

// ---- filtered slab predicate -----------------------------------------------------------
// The boolean of slab() is all the any-hit query needs (the entry distance only orders the
// visit).  t~ = (b - o) * fl(1/d) differs from the reference's fl((b - o) / d) by < 2 ulp,
// and x -> x(1 +- eps) is monotone, so the bound survives the min/max chain:
//   |Tmin~ - Tmin| <= eps |Tmin~|,  |Tmax~ - Tmax| <= eps |Tmax~|      (eps = 1e-6 >> 2 ulp)
// With g = Tmax~ - max(Tmin~, 0) and m = eps (|Tmin~| + |Tmax~|):  g > m proves the exact
// test accepts, g < -m proves it rejects; anything else (incl. inf/NaN) re-runs the exact
// IEEE-division test.  Rays with a direction component below 1e-20 in magnitude (or not
// finite) never take the fast path, so no inf/NaN can be produced by it.
struct RayQ { Ray r; V3 rd; bool fast; };

__device__ __forceinline__ RayQ make_rayq(const Ray & r)
{
    RayQ q; q.r = r;
    q.rd = mk(1.0f / r.d.x, 1.0f / r.d.y, 1.0f / r.d.z);
    float m = fminf(fminf(fabsf(r.d.x), fabsf(r.d.y)), fabsf(r.d.z));
    float M = fmaxf(fmaxf(fabsf(r.d.x), fabsf(r.d.y)), fabsf(r.d.z));
    q.fast = (m > 1e-20f) && (M < 1e20f);
    return q;
}

// Returns the reference's accept/reject decision exactly; tnear is only an ordering hint.
__device__ __forceinline__ bool slab_bool(const RayQ & q, float mnx, float mny, float mnz,
    float mxx, float mxy, float mxz, float & tnear)
{
    if (q.fast) {
        float ax = (mnx - q.r.o.x) * q.rd.x, bx = (mxx - q.r.o.x) * q.rd.x;
        float ay = (mny - q.r.o.y) * q.rd.y, by = (mxy - q.r.o.y) * q.rd.y;
        float az = (mnz - q.r.o.z) * q.rd.z, bz = (mxz - q.r.o.z) * q.rd.z;
        float tmin = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), q.r.tmin));
        float tmax = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), q.r.tmax));
        float g = tmax - fmaxf(tmin, 0.0f);
        float m = 1e-6f * (fabsf(tmin) + fabsf(tmax)) + 1e-30f;
        tnear = tmin;
        if (g > m) return true;
        if (g < -m) return false;
    }
    return slab(q.r, mnx, mny, mnz, mxx, mxy, mxz, tnear);
}

// Both queries are WARP-COLLECTIVE: all 32 lanes of the warp must call them together
// (lanes without a ray pass active = false).  The loop condition is a warp vote, which
// gives the hardware one reconvergence point per traversal step; without it the
// data-dependent exits let the lanes of a warp drift apart and run almost serially
// (measured: 4.6 active threads per instruction).

// Occlusion query.  Returns true iff the segment is blocked.
__device__ __forceinline__ bool trace_any_warp(const BvhView & bvh, const Ray & ray, bool active)
{
    bool hit = false;
    if (bvh.root_is_leaf) {
        float t;
        if (active)
            for (uint32_t i = 0; i < bvh.root_count; ++i)
                if (hit_tri(ray, bvh.tris + bvh.root_first + i, t)) hit = true;
        return hit;
    }
    const RayQ q = make_rayq(ray);
    uint32_t stack[UAV_STACK];
    int sp = 0;
    uint32_t cur = 0;
    bool done = !active;
    while (__any_sync(0xffffffffu, !done)) {
        if (!done) {
            const Node64 * np = bvh.inner + cur;
            float4 q0 = __ldg(&np->q0), q1 = __ldg(&np->q1), q2 = __ldg(&np->q2);
            uint4 q3 = __ldg(&np->q3);
            float tl, tr;
            bool hl = slab_bool(q, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, tl);
            bool hr = slab_bool(q, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, tr);
            // leaf children are tested at once (any-hit is order independent)
            uint32_t first = 0, count = 0, first2 = 0, count2 = 0;
            if (hl && q3.z) { first = q3.x; count = q3.z; hl = false; }
            if (hr && q3.w) {
                if (count) { first2 = q3.y; count2 = q3.w; } else { first = q3.y; count = q3.w; }
                hr = false;
            }
            float t;
            for (uint32_t i = 0; i < count; ++i) hit |= hit_tri(ray, bvh.tris + first + i, t);
            for (uint32_t i = 0; i < count2; ++i) hit |= hit_tri(ray, bvh.tris + first2 + i, t);
            if (hit) done = true;
            else if (hl && hr) {
                bool lf = tl < tr;
                stack[sp++] = lf ? q3.y : q3.x;
                cur = lf ? q3.x : q3.y;
            } else if (hl) cur = q3.x;
            else if (hr) cur = q3.y;
            else if (sp == 0) done = true;
            else cur = stack[--sp];
        }
    }
    return hit;
}

// Closest hit in the reference's order; face id through indices[] (bvh_tree.h:409).
__device__ __forceinline__ bool trace_closest_warp(const BvhView & bvh, Ray ray, bool active,
    uint32_t & face, float & t_hit)
{
    float best = UAV_INF; uint32_t best_face = UAV_NAI;
    uint2 stack[UAV_STACK];
    int sp = 0;
    uint2 cur = bvh.root_is_leaf ? make_uint2(bvh.root_first, bvh.root_count) : make_uint2(0u, 0u);
    bool done = !active;
    while (__any_sync(0xffffffffu, !done)) {
        if (!done) {
            bool pop = false;
            if (cur.y == 0) {
                const Node64 * np = bvh.inner + cur.x;
                float4 q0 = __ldg(&np->q0), q1 = __ldg(&np->q1), q2 = __ldg(&np->q2);
                uint4 q3 = __ldg(&np->q3);
                float tl, tr;
                bool hl = slab(ray, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, tl);
                bool hr = slab(ray, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, tr);
                uint2 L = make_uint2(q3.x, q3.z), R = make_uint2(q3.y, q3.w);
                if (hl && hr) {
                    bool lf = tl < tr;
                    stack[sp++] = lf ? R : L;
                    cur = lf ? L : R;
                } else if (hl) cur = L;
                else if (hr) cur = R;
                else pop = true;
            } else {
                bool any = false;
                for (uint32_t i = 0; i < cur.y; ++i) {
                    float t;
                    if (hit_tri(ray, bvh.tris + cur.x + i, t) && !(t > best)) {
                        best = t; best_face = __ldg(bvh.indices + cur.x + i); any = true;
                    }
                }
                if (any) ray.tmax = best;
                pop = true;
            }
            if (pop) { if (sp == 0) done = true; else cur = stack[--sp]; }
        }
    }
    face = best_face; t_hit = best;
    return best < UAV_INF;
}

}  // namespace uav

// ---- lean filtered node test (K1 traversal) ---------------------------------------------
// Same idea as slab_bool() with one FFMA per plane: t~ = fma(b, rd, -(o * rd)).  Against the
// reference's fl(fl(b - o) / d) this carries a relative error < 4 * 2^-24 plus an ABSOLUTE
// error <= 2^-24 |o / d| (the rounding of o * rd is not cancelled when b is close to o), so
// the undecided band is m = 1e-6 (|Tmin~| + |Tmax~|) + eabs with
// eabs = 3e-7 (|o.x rd.x| + |o.y rd.y| + |o.z rd.z|) >= 2 * 2^-23 max_k |o_k rd_k|.
// Rays that may not use the fast path get eabs = +inf: every test of theirs is "undecided".
namespace uav {

struct RayF {            // per-lane traversal constants
    float rdx, rdy, rdz;       // fl(1 / d)
    float nox, noy, noz;       // -(o * rd)
    float tmin, tmax, eabs;
};

__device__ __forceinline__ RayF make_rayf(const Ray & r)
{
    RayF f;
    f.rdx = 1.0f / r.d.x; f.rdy = 1.0f / r.d.y; f.rdz = 1.0f / r.d.z;
    float ox = r.o.x * f.rdx, oy = r.o.y * f.rdy, oz = r.o.z * f.rdz;
    f.nox = -ox; f.noy = -oy; f.noz = -oz;
    f.tmin = r.tmin; f.tmax = r.tmax;
    float m = fminf(fminf(fabsf(r.d.x), fabsf(r.d.y)), fabsf(r.d.z));
    float M = fmaxf(fmaxf(fabsf(r.d.x), fabsf(r.d.y)), fabsf(r.d.z));
    bool fast = (m > 1e-12f) && (M < 1e12f) && (fabsf(ox) + fabsf(oy) + fabsf(oz) < 1e30f);
    f.eabs = fast ? 3e-7f * (fabsf(ox) + fabsf(oy) + fabsf(oz)) : UAV_INF;
    return f;
}

// decided: hit is the reference's decision.  !decided: caller must run slab() (exact).
__device__ __forceinline__ void slab_fast(const RayF & f, float mnx, float mny, float mnz,
    float mxx, float mxy, float mxz, bool & hit, bool & decided, float & tnear)
{
    float ax = __fmaf_rn(mnx, f.rdx, f.nox), bx = __fmaf_rn(mxx, f.rdx, f.nox);
    float ay = __fmaf_rn(mny, f.rdy, f.noy), by = __fmaf_rn(mxy, f.rdy, f.noy);
    float az = __fmaf_rn(mnz, f.rdz, f.noz), bz = __fmaf_rn(mxz, f.rdz, f.noz);
    float tmin = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), f.tmin));
    float tmax = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), f.tmax));
    float lo = fmaxf(tmin, 0.0f);
    float g = tmax - lo;
    float m = __fmaf_rn(1e-6f, fabsf(tmin) + fabsf(tmax), f.eabs);
    hit = g > m;
    decided = hit || (g < -m);     // NaN / inf anywhere: neither holds -> undecided
    tnear = tmin;
}

static __device__ __noinline__ bool slab_exact_call(float ox, float oy, float oz, float dx, float dy, float dz,
    float tmin, float tmax, float mnx, float mny, float mnz, float mxx, float mxy, float mxz)
{
    Ray r; r.o = mk(ox, oy, oz); r.d = mk(dx, dy, dz); r.tmin = tmin; r.tmax = tmax;
    float t;
    return slab(r, mnx, mny, mnz, mxx, mxy, mxz, t);
}

}  // namespace uav

// ---- origin-relative filtered slab test (k1.cu traversal) ---------------------------------
// t~ = fl(fl(b - o) * fl(1 / d)).  The first operation is the reference's own fl(b - o); only
// the division is replaced, so |t~ - t_ref| <= 3 * 2^-24 |t~| and the undecided band is purely
// relative: m = 1e-6 (|Tmin~| + |Tmax~|).  g > m proves the reference accepts the box, g < -m
// proves it rejects; a box that only has to be tested conservatively is visited unless g < -m.
// Only rays with |d_k| > 1e-12 on all axes may use it (no inf / NaN can arise).
namespace uav {

struct RayT { float ox, oy, oz, rdx, rdy, rdz, tmin, tmax; float nox, noy, noz, em; };

// One FFMA per plane: t~ = fma(b, rd, -(o rd)).  Against the reference's fl(fl(b - o) / d) this has a
// relative error < 4 * 2^-24 and an ABSOLUTE error <= 2^-24 |o_k rd_k| per axis (the rounding of
// o rd does not cancel).  The min/max chain is monotone and 1-Lipschitz and clamps the results into
// [ray.tmin, ...] and [..., ray.tmax] with ray.tmin > 0, so with
//     em = 2.4e-7 (|o.x rd.x| + |o.y rd.y| + |o.z rd.z|) + 4e-6 ray.tmax          (one constant per ray)
// g = Tmax~ - Tmin~ > em proves that the reference accepts the box and g < -em proves that it rejects
// (if Tmin~ > 2 ray.tmax then |g| > Tmin~ / 2, which dominates the relative error of Tmin~ itself).
__device__ __forceinline__ void ray_fma_setup(RayT & f)
{
    const float ax = f.ox * f.rdx, ay = f.oy * f.rdy, az = f.oz * f.rdz;
    f.nox = -ax; f.noy = -ay; f.noz = -az;
    f.em = 2.4e-7f * (fabsf(ax) + fabsf(ay) + fabsf(az)) + 4e-6f * f.tmax;
}

// Node80 variant: the plane is c + off (off = outward-rounded fp16 offset), t~ = fma(off, rd, cr) with
// cr = fma(c, rd, -(o rd)) once per node and axis.  cr carries the rounding of o rd and its own, both
// <= 2^-24 (|o_k| + |c_k|) |rd_k| with |c_k| <= B_k = max |coordinate| of the scene on axis k, so the
// per-ray band is  em80 = 4.8e-7 sum_k (|o_k| + B_k) |rd_k| + 4e-6 ray.tmax  (same argument as above).
// The decoded box CONTAINS the exact one, so only "g < -em80 => the reference rejects" is ever used.
__device__ __forceinline__ void ray_fma_setup80(RayT & f, float bx, float by, float bz)
{
    const float ax = f.ox * f.rdx, ay = f.oy * f.rdy, az = f.oz * f.rdz;
    f.nox = -ax; f.noy = -ay; f.noz = -az;
    f.em = 4.8e-7f * ((fabsf(f.ox) + bx) * fabsf(f.rdx) + (fabsf(f.oy) + by) * fabsf(f.rdy) + (fabsf(f.oz) + bz) * fabsf(f.rdz))
        + 4e-6f * f.tmax;
}

__device__ __forceinline__ float slab_fma80(const RayT & f, float crx, float cry, float crz,
    float mnx, float mny, float mnz, float mxx, float mxy, float mxz)
{
    const float ax = __fmaf_rn(mnx, f.rdx, crx), bx = __fmaf_rn(mxx, f.rdx, crx);
    const float ay = __fmaf_rn(mny, f.rdy, cry), by = __fmaf_rn(mxy, f.rdy, cry);
    const float az = __fmaf_rn(mnz, f.rdz, crz), bz = __fmaf_rn(mxz, f.rdz, crz);
    const float tmin = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), f.tmin));
    const float tmax = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), f.tmax));
    return tmax - tmin;
}

__device__ __forceinline__ float slab_fma(const RayT & f, float mnx, float mny, float mnz,
    float mxx, float mxy, float mxz)
{
    const float ax = __fmaf_rn(mnx, f.rdx, f.nox), bx = __fmaf_rn(mxx, f.rdx, f.nox);
    const float ay = __fmaf_rn(mny, f.rdy, f.noy), by = __fmaf_rn(mxy, f.rdy, f.noy);
    const float az = __fmaf_rn(mnz, f.rdz, f.noz), bz = __fmaf_rn(mxz, f.rdz, f.noz);
    const float tmin = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), f.tmin));
    const float tmax = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), f.tmax));
    return tmax - tmin;
}

__device__ __forceinline__ void slab_rel(const RayT & f, float mnx, float mny, float mnz,
    float mxx, float mxy, float mxz, float & g, float & m)
{
    float ax = (mnx - f.ox) * f.rdx, bx = (mxx - f.ox) * f.rdx;
    float ay = (mny - f.oy) * f.rdy, by = (mxy - f.oy) * f.rdy;
    float az = (mnz - f.oz) * f.rdz, bz = (mxz - f.oz) * f.rdz;
    float tmin = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), f.tmin));
    float tmax = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), f.tmax));
    g = tmax - fmaxf(tmin, 0.0f);
    m = 1e-6f * (fabsf(tmin) + fabsf(tmax));
}

// The reference's any-hit walk with the exact test at every node (libs/acc/bvh_tree.h:419-467);
// used for the rare rays the filtered test may not handle.
static __device__ __noinline__ bool trace_any_exact(const BvhView bvh, float ox, float oy, float oz,
    float dx, float dy, float dz, float tmin, float tmax)
{
    Ray ray; ray.o = mk(ox, oy, oz); ray.d = mk(dx, dy, dz); ray.tmin = tmin; ray.tmax = tmax;
    float t;
    if (bvh.root_is_leaf) {
        for (uint32_t i = 0; i < bvh.root_count; ++i)
            if (hit_tri(ray, bvh.tris + bvh.root_first + i, t)) return true;
        return false;
    }
    uint32_t stack[UAV_STACK];
    int sp = 0;
    uint32_t cur = 0;
    for (;;) {
        const Node64 * np = bvh.inner + cur;
        float4 q0 = __ldg(&np->q0), q1 = __ldg(&np->q1), q2 = __ldg(&np->q2);
        uint4 q3 = __ldg(&np->q3);
        float tl, tr;
        bool hl = slab(ray, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, tl);
        bool hr = slab(ray, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, tr);
        if (hl && q3.z) { for (uint32_t i = 0; i < q3.z; ++i) if (hit_tri(ray, bvh.tris + q3.x + i, t)) return true; hl = false; }
        if (hr && q3.w) { for (uint32_t i = 0; i < q3.w; ++i) if (hit_tri(ray, bvh.tris + q3.y + i, t)) return true; hr = false; }
        if (hl && hr) { stack[sp++] = q3.y; cur = q3.x; }
        else if (hl) cur = q3.x;
        else if (hr) cur = q3.y;
        else if (sp == 0) return false;
        else cur = stack[--sp];
    }
}

}  // namespace uav

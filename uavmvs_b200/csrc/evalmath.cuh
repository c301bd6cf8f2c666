// Restated helper math of libs/eval/kernels.cu:21-40, :63-91, :111-136, shared by the
// observation-ray kernels (k1.cu, eval.cu).  Operation order is the reference's (SURVEY.md
// Appendix B); compiled with --fmad=false.
#pragma once
#include "common.cuh"

namespace uav {

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

// Same expression with the device intrinsics the reference's own logistic() uses (__expf,
// libs/eval/kernels.cu:80-84) and an approximate reciprocal; relative error ~1e-6 per term, inside
// the 1e-5 gate on scores.  acosf stays the full-precision one (the logistics amplify angle errors).
__device__ __forceinline__ float logistic_fast(float x, float k, float x0) {
    return __fdividef(1.0f, 1.0f + __expf(-k * (x - x0)));
}
// acos(x) = sqrt(1 - |x|) P7(|x|), reflected for x < 0 (Abramowitz & Stegun 4.4.46, |error| <= 2e-8 on [0, 1]);
// the square root is MUFU.RSQ plus one Newton step.  Absolute error < 1e-7 rad: the same size as the 2 ulp of
// CUDA's acosf at these magnitudes, at half the instructions.  x must already be clamped to [-1, 1].
__device__ __forceinline__ float acos_fast(float x) {
    const float a = fabsf(x);
    const float y = fmaxf(1.0f - a, 1e-30f);
    const float r = rsqrtf(y);
    float s = y * r;
    s = __fmaf_rn(0.5f * r, __fmaf_rn(-s, s, y), s);
    float p = -0.0012624911f;
    p = __fmaf_rn(p, a, 0.0066700901f);
    p = __fmaf_rn(p, a, -0.0170881256f);
    p = __fmaf_rn(p, a, 0.0308918810f);
    p = __fmaf_rn(p, a, -0.0501743046f);
    p = __fmaf_rn(p, a, 0.0889789874f);
    p = __fmaf_rn(p, a, -0.2145988016f);
    p = __fmaf_rn(p, a, 1.5707963050f);
    const float v = s * p;
    return x < 0.0f ? 3.14159265358979f - v : v;
}
// logistic(x; k, x0) = 1 / (1 + 2^(c1 x + c0)) with c1 = -k log2(e), c0 = k x0 log2(e) folded once per launch:
// one FFMA, ex2.approx, one add, rcp.approx (the reference's own logistic is __expf + an approximate division,
// libs/eval/kernels.cu:80-84 under --prec-div=false).  The rounding of the folded exponent is the same
// half-ulp-of-the-argument that __expf's x * log2(e) carries.
struct LogisticC { float c1, c0; };
__host__ __device__ __forceinline__ LogisticC logistic_fold(float k, float x0) {
    LogisticC c; c.c1 = -k * 1.4426950408889634f; c.c0 = k * x0 * 1.4426950408889634f; return c;
}
__device__ __forceinline__ float logistic_folded(float x, LogisticC c) {
    float e, r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(__fmaf_rn(c.c1, x, c.c0)));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.0f + e));
    return r;
}
__device__ __forceinline__ float pair_term_fast(float4 a, float4 b, LogisticC m, LogisticC t) {
    float calpha = b.x * a.x + b.y * a.y + b.z * a.z;
    float alpha = acos_fast(fmaxf(-1.0f, fminf(calpha, 1.0f)));
    float scale = fminf(a.w, b.w);
    float ctheta = fminf(a.z, b.z);
    float matchability = (1.0f - logistic_folded(alpha, m)) * ctheta;
    float triangulation = logistic_folded(alpha, t) * scale;
    return matchability * triangulation;
}

}  // namespace uav

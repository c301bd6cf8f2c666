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
__device__ __forceinline__ float pair_term_fast(float4 a, float4 b, float mk_, float mx0, float tk, float tx0) {
    float calpha = b.x * a.x + b.y * a.y + b.z * a.z;
    float alpha = acosf(fmaxf(-1.0f, fminf(calpha, 1.0f)));
    float scale = fminf(a.w, b.w);
    float ctheta = fminf(a.z, b.z);
    float matchability = (1.0f - logistic_fast(alpha, mk_, mx0)) * ctheta;
    float triangulation = logistic_fast(alpha, tk, tx0) * scale;
    return matchability * triangulation;
}

}  // namespace uav

// Observation-ray path: K1 update_observation_rays (single view and batched over a
// trajectory), K2 process, K3 evaluate, K8 func, K9 reduction, and the fused
// evaluate_trajectory entry point.  Reference: libs/eval/kernels.cu:21-291, :496-510.
#include "common.cuh"
#include "trace.cuh"
#include "evalmath.cuh"

namespace uav {

static Heur g_heur = {8.0f, 4.0f, 32.0f, 16.0f};   // m_k, m_x0, t_k, t_x0 (kernels.cu:93-96)
Heur current_heur() { return g_heur; }

// ---- K2: process_observation_rays, kernels.cu:204-264 (one warp per point) ----------
__global__ void __launch_bounds__(128)
k_process_observation_rays(float4 * obs, uint32_t * num_rows_ptr, uint32_t max_rows, uint32_t stride, uint32_t n)
{
    const uint32_t tx = threadIdx.x & 31;
    const uint32_t id = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (id >= n) return;
    uint32_t num_rows = min(num_rows_ptr[id], max_rows);
    float4 * col = obs + id;
    float4 ray = make_float4(0.f, 0.f, 0.f, 0.f);
    int key = (int)tx;
    float theta = -1.0f;
    if (tx < num_rows) {
        ray = col[(size_t)tx * stride];
        theta = (ray.w >= 0.0f) ? ray.z : -1.0f;
    }
    for (int i = 0; (1u << i) < num_rows; ++i) {
        bool asc = (tx >> (i + 1)) % 2;
        for (int s = 1 << i; s > 0; s >>= 1) {
            bool dir = (tx % (s << 1)) < s;
            float otheta = __shfl_xor_sync(0xffffffffu, theta, s);
            int okey = __shfl_xor_sync(0xffffffffu, key, s);
            if (otheta != theta && ((otheta > theta) == (asc ^ dir))) { theta = otheta; key = okey; }
        }
    }
    uint32_t num_valid = __popc(__ballot_sync(0xffffffffu, theta >= 0.0f));
    if (num_valid < num_rows && tx == 0) num_rows_ptr[id] = num_valid;
    ray.x = __shfl_sync(0xffffffffu, ray.x, key);
    ray.y = __shfl_sync(0xffffffffu, ray.y, key);
    ray.z = __shfl_sync(0xffffffffu, ray.z, key);
    ray.w = __shfl_sync(0xffffffffu, ray.w, key);
    if (tx < num_valid) col[(size_t)tx * stride] = ray;
}

// ---- K3: evaluate_observation_rays, kernels.cu:266-291 (one thread per point) -------
// The point's column is staged in shared memory once; the pair sum keeps the reference's
// order (inner j < i ascending, outer i ascending).
constexpr int K3_THREADS = 128;
__global__ void __launch_bounds__(K3_THREADS)
k_evaluate_observation_rays(const float4 * __restrict__ obs, const uint32_t * __restrict__ num_rows_ptr,
    uint32_t max_rows, uint32_t stride, uint32_t n, float * __restrict__ recons,
    float target, float * __restrict__ wrecons, Heur h)
{
    extern __shared__ float4 s_col[];   // [max_rows][K3_THREADS]
    const uint32_t id = blockIdx.x * K3_THREADS + threadIdx.x;
    if (id >= n) return;
    const uint32_t rows = min(num_rows_ptr[id], max_rows);
    for (uint32_t i = 0; i < rows; ++i) s_col[i * K3_THREADS + threadIdx.x] = obs[(size_t)i * stride + id];
    const float m_k = h.m_k, m_x0 = kPi / h.m_x0, t_k = h.t_k, t_x0 = kPi / h.t_x0;
    float recon = rows >= 1 ? 0.0f : -1.0f;
    for (uint32_t i = 1; i < rows; ++i) {
        float4 ri = s_col[i * K3_THREADS + threadIdx.x];
        float sum = 0.0f;
        for (uint32_t j = 0; j < i; ++j)
            sum += pair_term(s_col[j * K3_THREADS + threadIdx.x], ri, m_k, m_x0, t_k, t_x0);
        recon += sum;
    }
    recons[id] = recon;
    if (wrecons) wrecons[id] = func(recon, target);
}

// ---- K8: calculate_func_recons, kernels.cu:496-510 -----------------------------------
__global__ void k_calculate_func_recons(const float * __restrict__ recons, uint32_t n, float target,
    float * __restrict__ wrecons)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) wrecons[i] = func(recons[i], target);
}

// ---- K9: sum/min/max, libs/cacc/reduction.h:104-162 ----------------------------------
// acc[0] = sum (double), acc[1] = min, acc[2] = max as order-preserving ints.
__device__ __forceinline__ int f2ord(float f) { int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; }
__host__ __device__ __forceinline__ float ord2f(int i) {
    int b = i >= 0 ? i : i ^ 0x7fffffff;
#ifdef __CUDA_ARCH__
    return __int_as_float(b);
#else
    float f; memcpy(&f, &b, 4); return f;
#endif
}
struct ReduceAcc { double sum; int mn, mx; int pad; };

__global__ void k_reduce_init(ReduceAcc * acc) { acc->sum = 0.0; acc->mn = 0x7fffffff; acc->mx = (int)0x80000000; }

__global__ void __launch_bounds__(256)
k_reduce(const float * __restrict__ x, uint32_t n, ReduceAcc * acc)
{
    float s = 0.0f, mn = UAV_INF, mx = -UAV_INF;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        float v = x[i]; s += v; mn = fminf(mn, v); mx = fmaxf(mx, v);
    }
    for (int o = 16; o > 0; o >>= 1) {
        s += __shfl_xor_sync(0xffffffffu, s, o);
        mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    __shared__ float ss[8], smn[8], smx[8];
    uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) { ss[w] = s; smn[w] = mn; smx[w] = mx; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int i = 0; i < 8; ++i) { t += (double)ss[i]; mn = fminf(mn, smn[i]); mx = fmaxf(mx, smx[i]); }
        atomicAdd(&acc->sum, t);
        atomicMin(&acc->mn, f2ord(mn));
        atomicMax(&acc->mx, f2ord(mx));
    }
}

// ---- launch helpers --------------------------------------------------------------------
struct Scratch {   // per-process scratch on the device (views, reduction accumulator, stats)
    float * views = nullptr; size_t views_cap = 0;
    ReduceAcc * acc = nullptr;
    unsigned long long * stats = nullptr;
    ReduceAcc * h_acc = nullptr; unsigned long long * h_stats = nullptr;   // pinned
    cudaEvent_t ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};     // K1 | K2 | K3+K8 | K9 boundaries
    bool ev_valid = false;
};
static thread_local Scratch g_scratch;

static int ensure_scratch(size_t view_floats)
{
    Scratch & s = g_scratch;
    if (!s.acc) {
        UAV_CUDA(cudaMalloc(&s.acc, sizeof(ReduceAcc)));
        UAV_CUDA(cudaMalloc(&s.stats, 4 * sizeof(unsigned long long)));
        UAV_CUDA(cudaMallocHost(&s.h_acc, sizeof(ReduceAcc)));
        UAV_CUDA(cudaMallocHost(&s.h_stats, 4 * sizeof(unsigned long long)));
        for (int i = 0; i < 5; ++i) UAV_CUDA(cudaEventCreate(&s.ev[i]));
    }
    if (view_floats > s.views_cap) {
        if (s.views) cudaFree(s.views);
        UAV_CUDA(cudaMalloc(&s.views, view_floats * sizeof(float)));
        s.views_cap = view_floats;
    }
    return 0;
}

static int launch_k3(cudaStream_t s, const uavmvs_obs * obs, uavmvs_array * recons, float target, uavmvs_array * wrecons)
{
    UAV_REQUIRE(obs && recons, "null handle");
    UAV_REQUIRE(recons->n == obs->num_cols, "recons size != obs columns");
    size_t smem = sizeof(float4) * obs->max_rows * K3_THREADS;
    static bool attr_set = false;
    if (!attr_set) {
        UAV_CUDA(cudaFuncSetAttribute(k_evaluate_observation_rays, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
        attr_set = true;
    }
    k_evaluate_observation_rays<<<divup(obs->num_cols, K3_THREADS), K3_THREADS, smem, s>>>(
        obs->data, obs->num_rows, obs->max_rows, (uint32_t)(obs->pitch / sizeof(float4)), obs->num_cols,
        recons->data, target, wrecons ? wrecons->data : nullptr, g_heur);
    UAV_CUDA(cudaGetLastError());
    return 0;
}

static int launch_reduce(cudaStream_t s, const uavmvs_array * a)
{
    k_reduce_init<<<1, 1, 0, s>>>(g_scratch.acc);
    unsigned grid = min(divup(a->n, 256 * 8), 148u * 8u);
    k_reduce<<<grid ? grid : 1, 256, 0, s>>>(a->data, a->n, g_scratch.acc);
    UAV_CUDA(cudaGetLastError());
    UAV_CUDA(cudaMemcpyAsync(g_scratch.h_acc, g_scratch.acc, sizeof(ReduceAcc), cudaMemcpyDeviceToHost, s));
    return 0;
}

}  // namespace uav

using namespace uav;

extern "C" {

int uavmvs_configure_heuristic(float m_k, float m_x0, float t_k, float t_x0)
{
    g_heur.m_k = m_k; g_heur.m_x0 = m_x0; g_heur.t_k = t_k; g_heur.t_x0 = t_x0;
    return 0;
}

int uavmvs_update_observation_rays_batch_device(void * stream, int populate, const float * d_views,
    uint32_t num_views, float max_distance, int width, int height,
    const uavmvs_bvh * bvh, const uavmvs_cloud * cloud, uavmvs_obs * obs)
{
    return launch_k1((cudaStream_t)stream, populate, d_views, num_views, max_distance, width, height,
        bvh, cloud, obs, nullptr);
}

int uavmvs_update_observation_rays_batch(void * stream, int populate, const float * views,
    uint32_t num_views, float max_distance, int width, int height,
    const uavmvs_bvh * bvh, const uavmvs_cloud * cloud, uavmvs_obs * obs)
{
    UAV_REQUIRE(views || num_views == 0, "null views");
    if (num_views == 0) return 0;
    cudaStream_t s = (cudaStream_t)stream;
    if (ensure_scratch((size_t)num_views * UAVMVS_VIEW_FLOATS)) return 1;
    UAV_CUDA(cudaMemcpyAsync(g_scratch.views, views, sizeof(float) * UAVMVS_VIEW_FLOATS * (size_t)num_views,
        cudaMemcpyHostToDevice, s));
    int rc = launch_k1(s, populate, g_scratch.views, num_views, max_distance, width, height, bvh, cloud, obs, nullptr);
    if (rc) return rc;
    // the scratch view buffer may be reused by the next call: keep stream order, no host sync needed
    return 0;
}

int uavmvs_update_observation_rays(void * stream, int populate, const float view_pos[3],
    float max_distance, const float w2c[16], const float calib[9], int width, int height,
    const uavmvs_bvh * bvh, const uavmvs_cloud * cloud, uavmvs_obs * obs)
{
    UAV_REQUIRE(view_pos && w2c && calib, "null argument");
    float view[UAVMVS_VIEW_FLOATS];
    for (int i = 0; i < 3; ++i) view[i] = view_pos[i];
    for (int i = 0; i < 16; ++i) view[3 + i] = w2c[i];
    for (int i = 0; i < 9; ++i) view[19 + i] = calib[i];
    // one private device copy per call so concurrent streams never share the staging buffer
    cudaStream_t s = (cudaStream_t)stream;
    float * d_view;
    UAV_CUDA(cudaMallocAsync(&d_view, sizeof(view), s));
    UAV_CUDA(cudaMemcpyAsync(d_view, view, sizeof(view), cudaMemcpyHostToDevice, s));
    int rc = launch_k1(s, populate, d_view, 1, max_distance, width, height, bvh, cloud, obs, nullptr);
    UAV_CUDA(cudaFreeAsync(d_view, s));
    return rc;
}

int uavmvs_process_observation_rays(void * stream, uavmvs_obs * obs)
{
    UAV_REQUIRE(obs, "null obs");
    k_process_observation_rays<<<divup((uint64_t)obs->num_cols * 32, 128), 128, 0, (cudaStream_t)stream>>>(
        obs->data, obs->num_rows, obs->max_rows, (uint32_t)(obs->pitch / sizeof(float4)), obs->num_cols);
    UAV_CUDA(cudaGetLastError());
    return 0;
}

int uavmvs_evaluate_observation_rays(void * stream, const uavmvs_obs * obs, uavmvs_array * recons)
{
    return launch_k3((cudaStream_t)stream, obs, recons, 0.0f, nullptr);
}

int uavmvs_calculate_func_recons(void * stream, const uavmvs_array * recons, float target_recon, uavmvs_array * wrecons)
{
    UAV_REQUIRE(recons && wrecons && recons->n == wrecons->n, "bad arrays");
    k_calculate_func_recons<<<divup(recons->n, 256), 256, 0, (cudaStream_t)stream>>>(recons->data, recons->n,
        target_recon, wrecons->data);
    UAV_CUDA(cudaGetLastError());
    return 0;
}

int uavmvs_reduce(void * stream, const uavmvs_array * a, float out3[3])
{
    UAV_REQUIRE(a && out3, "null argument");
    if (ensure_scratch(0)) return 1;
    if (launch_reduce((cudaStream_t)stream, a)) return 1;
    UAV_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    out3[0] = (float)g_scratch.h_acc->sum; out3[1] = ord2f(g_scratch.h_acc->mn); out3[2] = ord2f(g_scratch.h_acc->mx);
    return 0;
}

int uavmvs_evaluate_trajectory_device(void * stream, const float * d_views, uint32_t num_views,
    float max_distance, float target_recon, int width, int height,
    const uavmvs_bvh * bvh, const uavmvs_cloud * cloud, uavmvs_obs * obs,
    uavmvs_array * recons, uavmvs_array * wrecons, float out3[3], uint64_t stats[4])
{
    UAV_REQUIRE(bvh && cloud && obs && recons && wrecons, "null handle");
    cudaStream_t s = (cudaStream_t)stream;
    if (ensure_scratch(0)) return 1;
    UAV_CUDA(cudaMemsetAsync(obs->num_rows, 0, 4 * (size_t)obs->num_cols, s));
    if (stats) UAV_CUDA(cudaMemsetAsync(g_scratch.stats, 0, 4 * sizeof(unsigned long long), s));
    g_scratch.ev_valid = false;
    UAV_CUDA(cudaEventRecord(g_scratch.ev[0], s));
    int rc = launch_k1(s, 1, d_views, num_views, max_distance, width, height, bvh, cloud, obs,
        stats ? g_scratch.stats : nullptr);
    if (rc) return rc;
    UAV_CUDA(cudaEventRecord(g_scratch.ev[1], s));
    rc = uavmvs_process_observation_rays(stream, obs);
    if (rc) return rc;
    UAV_CUDA(cudaEventRecord(g_scratch.ev[2], s));
    rc = launch_k3(s, obs, recons, target_recon, wrecons);
    if (rc) return rc;
    UAV_CUDA(cudaEventRecord(g_scratch.ev[3], s));
    if (launch_reduce(s, wrecons)) return 1;
    UAV_CUDA(cudaEventRecord(g_scratch.ev[4], s));
    g_scratch.ev_valid = true;
    if (stats) UAV_CUDA(cudaMemcpyAsync(g_scratch.h_stats, g_scratch.stats, 4 * sizeof(unsigned long long),
        cudaMemcpyDeviceToHost, s));
    if (out3 || stats) {
        UAV_CUDA(cudaStreamSynchronize(s));
        if (out3) { out3[0] = (float)g_scratch.h_acc->sum; out3[1] = ord2f(g_scratch.h_acc->mn); out3[2] = ord2f(g_scratch.h_acc->mx); }
        if (stats) {
            for (int i = 0; i < 4; ++i) stats[i] = g_scratch.h_stats[i];
            stats[2] = (uint64_t)cloud->n * num_views;
        }
    }
    return 0;
}

int uavmvs_profile_last(float ms[4])
{
    UAV_REQUIRE(ms, "null argument");
    UAV_REQUIRE(g_scratch.ev_valid, "no evaluate_trajectory call recorded on this thread");
    UAV_CUDA(cudaEventSynchronize(g_scratch.ev[4]));
    for (int i = 0; i < 4; ++i) UAV_CUDA(cudaEventElapsedTime(&ms[i], g_scratch.ev[i], g_scratch.ev[i + 1]));
    return 0;
}

int uavmvs_evaluate_trajectory(void * stream, const float * views, uint32_t num_views,
    float max_distance, float target_recon, int width, int height,
    const uavmvs_bvh * bvh, const uavmvs_cloud * cloud, uavmvs_obs * obs,
    uavmvs_array * recons, uavmvs_array * wrecons, float out3[3], uint64_t stats[4])
{
    UAV_REQUIRE(views || num_views == 0, "null views");
    cudaStream_t s = (cudaStream_t)stream;
    if (ensure_scratch((size_t)num_views * UAVMVS_VIEW_FLOATS)) return 1;
    if (num_views)
        UAV_CUDA(cudaMemcpyAsync(g_scratch.views, views, sizeof(float) * UAVMVS_VIEW_FLOATS * (size_t)num_views,
            cudaMemcpyHostToDevice, s));
    return uavmvs_evaluate_trajectory_device(stream, g_scratch.views, num_views, max_distance, target_recon,
        width, height, bvh, cloud, obs, recons, wrecons, out3, stats);
}

}  // extern "C"

// Observation-ray path: K1 update_observation_rays (single view and batched over a
// trajectory), K2 process, K3 evaluate, K8 func, K9 reduction, and the fused
// evaluate_trajectory entry point.  Reference: libs/eval/kernels.cu:21-291, :496-510.
#include "common.cuh"
#include "trace.cuh"
#include "evalmath.cuh"

namespace uav {

static Heur g_heur = {8.0f, 4.0f, 32.0f, 16.0f};   // m_k, m_x0, t_k, t_x0 (kernels.cu:93-96)
Heur current_heur() { return g_heur; }
Heur heur_of(const uavmvs_obs * obs)
{
    if (obs && obs->has_heur) { Heur h = {obs->heur[0], obs->heur[1], obs->heur[2], obs->heur[3]}; return h; }
    return g_heur;
}

// ---- K2 + K3 (+K8): process_observation_rays, kernels.cu:204-264, evaluate_observation_rays,
// :266-291 (heuristic :111-136), calculate_func_recons :496-510 -------------------------------
// One CTA = 32 consecutive points.  The [rows x 32 points] tile of the observation array is
// moved through shared memory with coalesced 512-B row segments (the reference reads and writes
// one 16-B element per lane at stride `pitch`); each warp then handles 8 points, one at a time:
//   DO_SORT  lane = row: the reference's warp bitonic network (same compare/exchange sequence, so
//            ties and the final order are those of kernels.cu:233-248), compaction of invalidated
//            rows, counter rewritten only if something was invalid (:251-254)
//   DO_EVAL  the n(n-1)/2 pairs are spread over the lanes (pair p -> lane p % 32), partial sums
//            combined by a shuffle tree.  The summation order differs from the reference's
//            serial double loop; scores are gated at 1e-5 relative (north_star), not bitwise.
constexpr int SE_THREADS = 128;
constexpr int SE_TILE_STRIDE = 33;      // float4 per tile row: padding against bank conflicts

template <bool DO_SORT, bool DO_EVAL>
__global__ void __launch_bounds__(SE_THREADS)
k_sort_eval(float4 * obs, uint32_t * num_rows_ptr, uint32_t max_rows, uint32_t stride, uint32_t n,
    float * __restrict__ recons, float target, float * __restrict__ wrecons, Heur h, uint32_t * dirty, uint32_t block0)
{
    const uint32_t bx = blockIdx.x + block0;                   // chunked launches (overlapped result copies)
    // incremental form: only blocks in which K1 added or invalidated a row since the last incremental pass
    if (dirty && dirty[bx] == 0u) return;
    __shared__ float4 s_tile[32 * SE_TILE_STRIDE];
    __shared__ float4 s_col[SE_THREADS / 32][32];
    __shared__ __align__(16) float s_key[SE_THREADS / 32][32];
    __shared__ uint32_t s_rows[32], s_valid[32];
    __shared__ uint16_t s_pairs[496];                      // pair q -> (i << 8 | j), i > j, q = i (i - 1) / 2 + j
    const uint32_t tx = threadIdx.x & 31, w = threadIdx.x >> 5;
    const uint32_t pt0 = bx * 32;
    if (DO_EVAL) {
        for (uint32_t q = threadIdx.x; q < 496; q += SE_THREADS) {
            uint32_t i = (uint32_t)((1.0f + sqrtf(1.0f + 8.0f * (float)q)) * 0.5f);
            while (i * (i - 1) / 2 > q) --i;
            while ((i + 1) * i / 2 <= q) ++i;
            s_pairs[q] = (uint16_t)(i << 8 | (q - i * (i - 1) / 2));
        }
    }
    if (threadIdx.x < 32) {
        uint32_t id = pt0 + threadIdx.x;
        uint32_t r = id < n ? min(num_rows_ptr[id], max_rows) : 0u;
        s_rows[threadIdx.x] = r; s_valid[threadIdx.x] = r;
    }
    __syncthreads();
    uint32_t rmax = s_rows[tx];
    for (int o = 16; o > 0; o >>= 1) rmax = max(rmax, __shfl_xor_sync(0xffffffffu, rmax, o));
    // coalesced load of rows [0, rmax): warp w takes rows w, w + 4, ...
    for (uint32_t r = w; r < rmax; r += SE_THREADS / 32)
        if (pt0 + tx < n) s_tile[r * SE_TILE_STRIDE + tx] = obs[(size_t)r * stride + pt0 + tx];
    __syncthreads();

    const LogisticC lm = logistic_fold(h.m_k, kPi / h.m_x0), lt = logistic_fold(h.t_k, kPi / h.t_x0);
    for (uint32_t p = w * 8; p < w * 8 + 8; ++p) {
        const uint32_t id = pt0 + p;
        if (id >= n) break;                                   // warp-uniform
        const uint32_t num_rows = s_rows[p];
        float4 ray = make_float4(0.f, 0.f, 0.f, 0.f);
        if (tx < num_rows) ray = s_tile[tx * SE_TILE_STRIDE + p];
        uint32_t nv = num_rows;
        if (DO_SORT) {
            float theta = -1.0f;
            if (tx < num_rows) theta = (ray.w >= 0.0f) ? ray.z : -1.0f;
            // Descending order of the valid rows by counting: rank = number of larger keys.  For distinct keys
            // every correct sort gives the reference's order; only the arrangement of EQUAL keys depends on the
            // reference's bitonic network (it never exchanges equal keys, kernels.cu:241), so a column with a
            // tie among its valid rows runs that network instead.  Invalid rows and padding (-1) sort behind
            // every valid row (cos(theta) >= 0.087) and are never written back.
            // The keys go through shared memory, four per load; two valid rows hold equal keys exactly when they get
            // equal ranks (a larger key counts towards every smaller one), so one match_any finds the ties.
            __syncwarp();
            s_key[w][tx] = theta;
            __syncwarp();
            uint32_t rank = 0;
            const float4 * kp = reinterpret_cast<const float4 *>(s_key[w]);
            for (uint32_t k4 = 0; 4u * k4 < num_rows; ++k4) {
                const float4 o = kp[k4];
                rank += (uint32_t)(o.x > theta) + (uint32_t)(o.y > theta) + (uint32_t)(o.z > theta) + (uint32_t)(o.w > theta);
            }
            const bool valid_row = theta >= 0.0f;
            nv = __popc(__ballot_sync(0xffffffffu, valid_row));
            const uint32_t peers = __match_any_sync(0xffffffffu, valid_row ? rank : 64u + tx);
            const bool tie = (valid_row && peers != (1u << tx)) || !(theta == theta);      // a NaN key: leave it to the network
            if (!__any_sync(0xffffffffu, tie)) {
                __syncwarp();
                if (valid_row) { s_tile[rank * SE_TILE_STRIDE + p] = ray; if (DO_EVAL) s_col[w][rank] = ray; }
                __syncwarp();
                if (DO_EVAL) ray = s_col[w][tx];                      // rows >= nv: stale, never used
            } else {
                int key = (int)tx;
                for (int i = 0; (1u << i) < num_rows; ++i) {
                    bool asc = (tx >> (i + 1)) % 2;
                    for (int s = 1 << i; s > 0; s >>= 1) {
                        bool dir = (tx % (s << 1)) < s;
                        float otheta = __shfl_xor_sync(0xffffffffu, theta, s);
                        int okey = __shfl_xor_sync(0xffffffffu, key, s);
                        if (otheta != theta && ((otheta > theta) == (asc ^ dir))) { theta = otheta; key = okey; }
                    }
                }
                ray.x = __shfl_sync(0xffffffffu, ray.x, key);
                ray.y = __shfl_sync(0xffffffffu, ray.y, key);
                ray.z = __shfl_sync(0xffffffffu, ray.z, key);
                ray.w = __shfl_sync(0xffffffffu, ray.w, key);
                if (tx < nv) s_tile[tx * SE_TILE_STRIDE + p] = ray;      // rows >= nv keep their old content
            }
            if (tx == 0) s_valid[p] = nv;
        }
        if (DO_EVAL) {
            __syncwarp();
            s_col[w][tx] = ray;
            __syncwarp();
            const uint32_t rows = nv;
            const uint32_t npairs = rows * (rows - (rows ? 1u : 0u)) / 2u;
            float sum = 0.0f;
            // pair q -> (i, j), i > j, enumerated row by row (table lookup: the carries of an incremental
            // enumeration cost 11 % of the kernel)
            for (uint32_t q = tx; q < npairs; q += 32) {
                const uint32_t ij = s_pairs[q];
                sum += pair_term_fast(s_col[w][ij & 255u], s_col[w][ij >> 8], lm, lt);
            }
            for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
            if (tx == 0) {
                float recon = rows >= 1 ? sum : -1.0f;            // kernels.cu:283-288: none -> -1, one -> 0
                recons[id] = recon;
                if (wrecons) wrecons[id] = func(recon, target);
            }
        }
    }
    if (dirty && threadIdx.x == 0) dirty[bx] = 0u;     // every thread read the flag before the first barrier
    if (DO_SORT) {
        __syncthreads();
        for (uint32_t r = w; r < rmax; r += SE_THREADS / 32)
            if (pt0 + tx < n && r < s_valid[tx]) obs[(size_t)r * stride + pt0 + tx] = s_tile[r * SE_TILE_STRIDE + tx];
        if (threadIdx.x < 32 && pt0 + threadIdx.x < n && s_valid[threadIdx.x] < s_rows[threadIdx.x])
            num_rows_ptr[pt0 + threadIdx.x] = s_valid[threadIdx.x];
    }
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
    cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};     // K1 | K2 | K3+K8 | K9 | score collective boundaries
    double * score_dev = nullptr, * h_score = nullptr;                     // fold result + one record (single GPU); pinned copy
    bool ev_valid = false;
    cudaStream_t copy_stream = nullptr;                                    // result copies of evaluate_trajectory_to_host
    cudaStream_t side_stream = nullptr; cudaEvent_t ev_side[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};   // K2 + K3 + K8 beside k_retire (UAVMVS_OVERLAP_K3)
    cudaEvent_t ev_copy[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
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
        for (int i = 0; i < 6; ++i) UAV_CUDA(cudaEventCreate(&s.ev[i]));
        UAV_CUDA(cudaMalloc(&s.score_dev, 16 * sizeof(double)));
        UAV_CUDA(cudaMallocHost(&s.h_score, 8 * sizeof(double)));
    }
    if (view_floats > s.views_cap) {
        if (s.views) cudaFree(s.views);
        UAV_CUDA(cudaMalloc(&s.views, view_floats * sizeof(float)));
        s.views_cap = view_floats;
    }
    return 0;
}

static int launch_sort_eval(cudaStream_t s, bool do_sort, bool do_eval, uavmvs_obs * obs, uavmvs_array * recons,
    float target, uavmvs_array * wrecons, bool incremental = false, uint32_t block0 = 0, uint32_t nblocks = 0)
{
    uint32_t * dirty = (incremental && obs) ? obs->dirty : nullptr;
    UAV_REQUIRE(obs, "null handle");
    UAV_REQUIRE(!do_eval || (recons && recons->n == obs->num_cols), "recons size != obs columns");
    UAV_REQUIRE(!wrecons || wrecons->n == obs->num_cols, "wrecons size != obs columns");
    const unsigned grid = nblocks ? nblocks : divup(obs->num_cols, 32);
    const uint32_t stride = (uint32_t)(obs->pitch / sizeof(float4));
    float * r = recons ? recons->data : nullptr, * wr = wrecons ? wrecons->data : nullptr;
    if (do_sort && do_eval)
        k_sort_eval<true, true><<<grid, SE_THREADS, 0, uav::counted(s)>>>(obs->data, obs->num_rows, obs->max_rows, stride, obs->num_cols, r, target, wr, heur_of(obs), dirty, block0);
    else if (do_sort)
        k_sort_eval<true, false><<<grid, SE_THREADS, 0, uav::counted(s)>>>(obs->data, obs->num_rows, obs->max_rows, stride, obs->num_cols, r, target, wr, heur_of(obs), dirty, block0);
    else
        k_sort_eval<false, true><<<grid, SE_THREADS, 0, uav::counted(s)>>>(obs->data, obs->num_rows, obs->max_rows, stride, obs->num_cols, r, target, wr, heur_of(obs), dirty, block0);
    UAV_CUDA(cudaGetLastError());
    return 0;
}

static int launch_reduce(cudaStream_t s, const uavmvs_array * a)
{
    k_reduce_init<<<1, 1, 0, uav::counted(s)>>>(g_scratch.acc);
    unsigned grid = min(divup(a->n, 256 * 8), 148u * 8u);
    k_reduce<<<grid ? grid : 1, 256, 0, uav::counted(s)>>>(a->data, a->n, g_scratch.acc);
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

int uavmvs_obs_configure_heuristic(uavmvs_obs * obs, float m_k, float m_x0, float t_k, float t_x0)
{
    UAV_REQUIRE(obs, "null obs");
    obs->heur[0] = m_k; obs->heur[1] = m_x0; obs->heur[2] = t_k; obs->heur[3] = t_x0; obs->has_heur = 1;
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
    return launch_sort_eval((cudaStream_t)stream, true, false, obs, nullptr, 0.0f, nullptr);
}

int uavmvs_evaluate_observation_rays(void * stream, const uavmvs_obs * obs, uavmvs_array * recons)
{
    return launch_sort_eval((cudaStream_t)stream, false, true, const_cast<uavmvs_obs *>(obs), recons, 0.0f, nullptr);
}

int uavmvs_reevaluate_observation_rays(void * stream, uavmvs_obs * obs, uavmvs_array * recons, float target_recon,
    uavmvs_array * wrecons)
{
    UAV_REQUIRE(obs && recons, "null argument");
    return launch_sort_eval((cudaStream_t)stream, true, true, obs, recons, target_recon, wrecons, true);
}

int uavmvs_calculate_func_recons(void * stream, const uavmvs_array * recons, float target_recon, uavmvs_array * wrecons)
{
    UAV_REQUIRE(recons && wrecons && recons->n == wrecons->n, "bad arrays");
    k_calculate_func_recons<<<divup(recons->n, 256), 256, 0, uav::counted((cudaStream_t)stream)>>>(recons->data, recons->n,
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

// ---- the whole evaluation (apps/evaluate_trajectory/evaluate_trajectory.cu:125-202) ------------------------
// One body for every entry point: clear counters, K1 over all views, K2 + K3 + K8 (optionally in four block ranges whose
// results are copied to the host while the next range is computed), K9, and -- for the sharded form -- the score record
// of this shard packed on the device, all-gathered over the ranks and folded in rank order, still on the same stream.
struct EvalCall {
    cudaStream_t s; const float * d_views; uint32_t num_views; float max_distance, target_recon; int width, height;
    const uavmvs_bvh * bvh; const uavmvs_cloud * cloud; uavmvs_obs * obs; uavmvs_array * recons, * wrecons;
    float * h_recons = nullptr; uint32_t * h_num_rows = nullptr;
    bool want_stats = false;
    bool sharded = false; uavmvs_comm * comm = nullptr;
};

// record = {sum wrecon, points, rays, observations, pairs after the block cull, min, max, pairs}
__global__ void k_pack_score(const ReduceAcc * acc, const unsigned long long * stats, double points, double pairs, double * rec)
{
    rec[0] = acc->sum; rec[1] = points; rec[2] = (double)stats[0]; rec[3] = (double)stats[1]; rec[4] = (double)stats[3];
    rec[5] = (double)ord2f(acc->mn); rec[6] = (double)ord2f(acc->mx); rec[7] = pairs;
}

// Fold in rank order (the same additions on every rank: bit-identical results everywhere).
// out = {sum, points, min, max, rays, observations, pairs after the block cull, pairs}
__global__ void k_fold_score(const double * recs, int world, double * out)
{
    double sum = 0.0, pts = 0.0, rays = 0.0, obs = 0.0, cull = 0.0, pairs = 0.0, mn = INFINITY, mx = -INFINITY;
    for (int r = 0; r < world; ++r) {
        const double * q = recs + 8 * r;
        sum += q[0]; pts += q[1]; rays += q[2]; obs += q[3]; cull += q[4]; pairs += q[7];
        mn = fmin(mn, q[5]); mx = fmax(mx, q[6]);
    }
    out[0] = sum; out[1] = pts; out[2] = mn; out[3] = mx; out[4] = rays; out[5] = obs; out[6] = cull; out[7] = pairs;
}

// K2 + K3 + K8 of a block range as soon as that range has retired (K1AfterRetire), and the range's recon copy behind it
struct AfterCtx { const EvalCall * c; uint32_t k; };
static int run_evaluation_after_retire(void * ctx, cudaStream_t side, uint32_t b0, uint32_t nbk)
{
    AfterCtx * a = static_cast<AfterCtx *>(ctx);
    const EvalCall & c = *a->c;
    Scratch & g = g_scratch;
    if (int rc = launch_sort_eval(side, true, true, c.obs, c.recons, c.target_recon, c.wrecons, false, b0, nbk)) return rc;
    if (c.h_recons && a->k < 4) {
        const size_t p0 = (size_t)b0 * 32, p1 = min((size_t)(b0 + nbk) * 32, (size_t)c.obs->num_cols);
        UAV_CUDA(cudaEventRecord(g.ev_copy[1 + a->k], side));
        UAV_CUDA(cudaStreamWaitEvent(g.copy_stream, g.ev_copy[1 + a->k], 0));
        UAV_CUDA(cudaMemcpyAsync(c.h_recons + p0, c.recons->data + p0, 4 * (p1 - p0), cudaMemcpyDeviceToHost, g.copy_stream));
    }
    ++a->k;
    return 0;
}

static int run_evaluation(const EvalCall & c, float out3[3], uint64_t stats[4], double score[UAVMVS_SCORE_DOUBLES])
{
    UAV_REQUIRE(c.bvh && c.cloud && c.obs && c.recons && c.wrecons, "null handle");
    cudaStream_t s = c.s;
    Scratch & g = g_scratch;
    const bool need_stats = c.want_stats || c.sharded;
    const bool to_host = c.h_recons || c.h_num_rows;
    if (to_host && !g.copy_stream) {
        UAV_CUDA(cudaStreamCreateWithFlags(&g.copy_stream, cudaStreamNonBlocking));
        for (auto & e : g.ev_copy) UAV_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    }
    UAV_CUDA(cudaMemsetAsync(c.obs->num_rows, 0, 4 * (size_t)c.obs->num_cols, s));
    if (need_stats) UAV_CUDA(cudaMemsetAsync(g.stats, 0, 4 * sizeof(unsigned long long), s));
    g.ev_valid = false;
    UAV_CUDA(cudaEventRecord(g.ev[0], s));
    // Opt-in, measured slower (profiles/r02_aa_overlap_k3.txt): k_retire in four block ranges with K2 + K3 + K8 of each finished
    // range on a side stream.  retire + K3 0.887 ms back to back, 0.957 ms overlapped (step 4.226 -> 4.305 ms): k_sort_eval
    // already fills the issue slots, and four launches of each cost more than the overlap returns.
    const bool overlap = getenv("UAVMVS_OVERLAP_K3") != nullptr;
    AfterCtx actx{&c, 0};
    K1AfterRetire after{4, nullptr, g.ev_side, run_evaluation_after_retire, &actx};
    if (overlap) {
        if (!g.side_stream) {
            UAV_CUDA(cudaStreamCreateWithFlags(&g.side_stream, cudaStreamNonBlocking));
            for (auto & e : g.ev_side) UAV_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        }
        after.side = g.side_stream;
    }
    int rc = launch_k1(s, 1, c.d_views, c.num_views, c.max_distance, c.width, c.height, c.bvh, c.cloud, c.obs,
        need_stats ? g.stats : nullptr, nullptr, overlap ? &after : nullptr);
    if (rc) return rc;
    UAV_CUDA(cudaEventRecord(g.ev[1], s));
    UAV_CUDA(cudaEventRecord(g.ev[2], s));      // K2 is fused into the next launch
    // counters are final after K1 here (nothing was invalidated, so K2 rewrites none: kernels.cu:251-254): their copy
    // runs beside K2 + K3; recon goes out in four chunks, each copied while the next is computed
    if (c.h_num_rows) {
        UAV_CUDA(cudaEventRecord(g.ev_copy[0], s));
        UAV_CUDA(cudaStreamWaitEvent(g.copy_stream, g.ev_copy[0], 0));
        UAV_CUDA(cudaMemcpyAsync(c.h_num_rows, c.obs->num_rows, 4 * (size_t)c.obs->num_cols, cudaMemcpyDeviceToHost, g.copy_stream));
    }
    const uint32_t nblocks = divup(c.obs->num_cols, 32), chunks = overlap ? 0 : (c.h_recons && nblocks >= 64 ? 4 : 1);
    // four ranges of decreasing size: what stays exposed at the end is the copy of the LAST range only (10 % of the points)
    static const uint32_t cut4[5] = {0u, 36u, 66u, 90u, 100u};
    for (uint32_t k = 0; k < chunks; ++k) {
        const uint32_t b0 = chunks == 4 ? (uint32_t)((uint64_t)nblocks * cut4[k] / 100u) : (uint32_t)((uint64_t)nblocks * k / chunks);
        const uint32_t b1 = chunks == 4 ? (uint32_t)((uint64_t)nblocks * cut4[k + 1] / 100u) : (uint32_t)((uint64_t)nblocks * (k + 1) / chunks);
        rc = launch_sort_eval(s, true, true, c.obs, c.recons, c.target_recon, c.wrecons, false, b0, b1 - b0);
        if (rc) return rc;
        if (c.h_recons) {
            const size_t p0 = (size_t)b0 * 32, p1 = min((size_t)b1 * 32, (size_t)c.obs->num_cols);
            UAV_CUDA(cudaEventRecord(g.ev_copy[1 + k], s));
            UAV_CUDA(cudaStreamWaitEvent(g.copy_stream, g.ev_copy[1 + k], 0));
            UAV_CUDA(cudaMemcpyAsync(c.h_recons + p0, c.recons->data + p0, 4 * (p1 - p0), cudaMemcpyDeviceToHost, g.copy_stream));
        }
    }
    UAV_CUDA(cudaEventRecord(g.ev[3], s));
    if (launch_reduce(s, c.wrecons)) return 1;
    UAV_CUDA(cudaEventRecord(g.ev[4], s));
    if (c.sharded) {
        // the collective follows K9 on the stream: no host synchronisation in between
        const int world = c.comm ? c.comm->world : 1, rank = c.comm ? c.comm->rank : 0;
        double * recs = g.score_dev;                        // [8] fold result, then [8 * world] records
        if (c.comm) {
            if (!c.comm->gather) UAV_CUDA(cudaMalloc(&c.comm->gather, sizeof(double) * 8 * (size_t)(world + 1)));
            recs = c.comm->gather;
        }
        k_pack_score<<<1, 1, 0, uav::counted(s)>>>(g.acc, g.stats, (double)c.cloud->n, (double)c.cloud->n * (double)c.num_views,
            recs + 8 + 8 * (size_t)rank);
        UAV_CUDA(cudaGetLastError());
        if (c.comm && comm_allgather(s, c.comm, recs + 8 + 8 * (size_t)rank, recs + 8, 8, 1)) return 3;
        k_fold_score<<<1, 1, 0, uav::counted(s)>>>(recs + 8, world, recs);
        UAV_CUDA(cudaGetLastError());
        UAV_CUDA(cudaMemcpyAsync(g.h_score, recs, 8 * sizeof(double), cudaMemcpyDeviceToHost, s));
    }
    UAV_CUDA(cudaEventRecord(g.ev[5], s));
    g.ev_valid = true;
    if (need_stats) UAV_CUDA(cudaMemcpyAsync(g.h_stats, g.stats, 4 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
    if (to_host) {                                           // the caller's stream owns the host buffers when this returns
        UAV_CUDA(cudaEventRecord(g.ev_copy[5], g.copy_stream));
        UAV_CUDA(cudaStreamWaitEvent(s, g.ev_copy[5], 0));
    }
    if (out3 || stats || score || to_host) {
        UAV_CUDA(cudaStreamSynchronize(s));
        if (out3) { out3[0] = (float)g.h_acc->sum; out3[1] = ord2f(g.h_acc->mn); out3[2] = ord2f(g.h_acc->mx); }
        if (stats) {
            for (int i = 0; i < 4; ++i) stats[i] = g.h_stats[i];
            stats[2] = (uint64_t)c.cloud->n * c.num_views;
        }
        if (score) for (int i = 0; i < UAVMVS_SCORE_DOUBLES; ++i) score[i] = g.h_score[i];
    }
    return 0;
}

static EvalCall make_call(void * stream, const float * d_views, uint32_t num_views, float max_distance, float target_recon,
    int width, int height, const uavmvs_bvh * bvh, const uavmvs_cloud * cloud, uavmvs_obs * obs, uavmvs_array * recons,
    uavmvs_array * wrecons)
{
    EvalCall c; c.s = (cudaStream_t)stream; c.d_views = d_views; c.num_views = num_views; c.max_distance = max_distance;
    c.target_recon = target_recon; c.width = width; c.height = height; c.bvh = bvh; c.cloud = cloud; c.obs = obs;
    c.recons = recons; c.wrecons = wrecons;
    return c;
}

static int stage_views(cudaStream_t s, const float * views, uint32_t num_views)
{
    UAV_REQUIRE(views || num_views == 0, "null views");
    if (ensure_scratch((size_t)num_views * UAVMVS_VIEW_FLOATS)) return 1;
    if (num_views)
        UAV_CUDA(cudaMemcpyAsync(g_scratch.views, views, sizeof(float) * UAVMVS_VIEW_FLOATS * (size_t)num_views,
            cudaMemcpyHostToDevice, s));
    return 0;
}

int uavmvs_evaluate_trajectory_device(void * stream, const float * d_views, uint32_t num_views,
    float max_distance, float target_recon, int width, int height,
    const uavmvs_bvh * bvh, const uavmvs_cloud * cloud, uavmvs_obs * obs,
    uavmvs_array * recons, uavmvs_array * wrecons, float out3[3], uint64_t stats[4])
{
    if (ensure_scratch(0)) return 1;
    EvalCall c = make_call(stream, d_views, num_views, max_distance, target_recon, width, height, bvh, cloud, obs, recons, wrecons);
    c.want_stats = stats != nullptr;
    return run_evaluation(c, out3, stats, nullptr);
}

int uavmvs_evaluate_trajectory_to_host(void * stream, const float * views, uint32_t num_views,
    float max_distance, float target_recon, int width, int height,
    const uavmvs_bvh * bvh, const uavmvs_cloud * cloud, uavmvs_obs * obs,
    uavmvs_array * recons, uavmvs_array * wrecons, float * h_recons, uint32_t * h_num_rows,
    float out3[3], uint64_t stats[4])
{
    if (int rc = stage_views((cudaStream_t)stream, views, num_views)) return rc;
    EvalCall c = make_call(stream, g_scratch.views, num_views, max_distance, target_recon, width, height, bvh, cloud, obs, recons, wrecons);
    c.want_stats = stats != nullptr; c.h_recons = h_recons; c.h_num_rows = h_num_rows;
    float dummy[3];
    return run_evaluation(c, out3 ? out3 : dummy, stats, nullptr);      // always synchronises: the host buffers are the result
}

int uavmvs_evaluate_trajectory_sharded(void * stream, uavmvs_comm * comm, const float * views, int views_on_device,
    uint32_t num_views, float max_distance, float target_recon, int width, int height,
    const uavmvs_bvh * bvh, const uavmvs_cloud * cloud, uavmvs_obs * obs,
    uavmvs_array * recons, uavmvs_array * wrecons, float * h_recons, uint32_t * h_num_rows,
    double score[UAVMVS_SCORE_DOUBLES], float local_out3[3])
{
    UAV_REQUIRE(score, "null score");
    const float * d_views = views;
    if (!views_on_device) {
        if (int rc = stage_views((cudaStream_t)stream, views, num_views)) return rc;
        d_views = g_scratch.views;
    } else if (ensure_scratch(0)) return 1;
    EvalCall c = make_call(stream, d_views, num_views, max_distance, target_recon, width, height, bvh, cloud, obs, recons, wrecons);
    c.sharded = true; c.comm = comm; c.h_recons = h_recons; c.h_num_rows = h_num_rows;
    return run_evaluation(c, local_out3, nullptr, score);
}

int uavmvs_views_upload(void * stream, const float * views, uint32_t num_views, float ** d_views)
{
    UAV_REQUIRE(views && d_views && num_views > 0, "null argument");
    float * d = nullptr;
    const size_t bytes = sizeof(float) * UAVMVS_VIEW_FLOATS * (size_t)num_views;
    UAV_CUDA(cudaMalloc(&d, bytes));
    cudaError_t e = cudaMemcpyAsync(d, views, bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize((cudaStream_t)stream);      // `views` may be pageable / reused
    if (e != cudaSuccess) { cudaFree(d); UAV_CUDA(e); }
    *d_views = d;
    return 0;
}

int uavmvs_views_update(void * stream, float * d_views, const float * views, uint32_t num_views)
{
    UAV_REQUIRE(views && d_views, "null argument");
    UAV_CUDA(cudaMemcpyAsync(d_views, views, sizeof(float) * UAVMVS_VIEW_FLOATS * (size_t)num_views, cudaMemcpyHostToDevice, (cudaStream_t)stream));
    return 0;
}

int uavmvs_views_download(void * stream, const float * d_views, uint32_t num_views, float * views)
{
    UAV_REQUIRE(views && d_views, "null argument");
    UAV_CUDA(cudaMemcpyAsync(views, d_views, sizeof(float) * UAVMVS_VIEW_FLOATS * (size_t)num_views, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    UAV_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return 0;
}

int uavmvs_views_free(float * d_views)
{
    if (d_views) UAV_CUDA(cudaFree(d_views));
    return 0;
}

int uavmvs_k1_counters(uint64_t out[4], int reset)
{
    UAV_REQUIRE(out, "null argument");
    unsigned long long tmp[4];
    if (uav::k1_counters(tmp, reset)) return 1;
    for (int i = 0; i < 4; ++i) out[i] = tmp[i];
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

int uavmvs_profile_k1_enable(int on) { return uav::k1_profile_enable(on); }

int uavmvs_profile_k1(float ms[5])
{
    UAV_REQUIRE(ms, "null argument");
    return uav::k1_profile(ms);
}

int uavmvs_profile_last_collective(float * ms)
{
    UAV_REQUIRE(ms, "null argument");
    UAV_REQUIRE(g_scratch.ev_valid, "no evaluate_trajectory call recorded on this thread");
    UAV_CUDA(cudaEventSynchronize(g_scratch.ev[5]));
    UAV_CUDA(cudaEventElapsedTime(ms, g_scratch.ev[4], g_scratch.ev[5]));
    return 0;
}

int uavmvs_evaluate_trajectory(void * stream, const float * views, uint32_t num_views,
    float max_distance, float target_recon, int width, int height,
    const uavmvs_bvh * bvh, const uavmvs_cloud * cloud, uavmvs_obs * obs,
    uavmvs_array * recons, uavmvs_array * wrecons, float out3[3], uint64_t stats[4])
{
    if (int rc = stage_views((cudaStream_t)stream, views, num_views)) return rc;
    return uavmvs_evaluate_trajectory_device(stream, g_scratch.views, num_views, max_distance, target_recon,
        width, height, bvh, cloud, obs, recons, wrecons, out3, stats);
}

}  // extern "C"

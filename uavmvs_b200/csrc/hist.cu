// Spherical-histogram path: K4/K5 populate_spherical_histogram, K6
// evaluate_spherical_histogram, K7 estimate_capture_difficulty.
// Reference: libs/eval/kernels.cu:293-494, libs/cacc/nnsearch.h:67-155.
#include "common.cuh"
#include "trace.cuh"
#include "kd.cuh"

#include <vector>

namespace uav {

constexpr float kPi = 3.14159265f;

__device__ __forceinline__ float logistic(float x, float k, float x0) { return 1.0f / (1.0f + expf(-k * (x - x0))); }
__device__ __forceinline__ float func(float recon, float target) { float a = fabsf(fminf(recon - target, 0.0f)); return a * a; }
__device__ __forceinline__ V3 orthogonal(V3 v) {
    V3 n0 = mk(1.0f, 0.0f, 0.0f), n1 = mk(0.0f, 1.0f, 0.0f);
    if (fabsf(dot(n0, v)) < fabsf(dot(n1, v))) return cross(n0, v);
    return cross(n1, v);
}
__device__ __forceinline__ V3 relative_direction(V3 v2cn, V3 n) {
    V3 rx = normalized(orthogonal(n));
    V3 ry = normalized(cross(n, rx));
    return normalized(mk(dot(rx, v2cn), dot(ry, v2cn), dot(n, v2cn)));
}

struct HistArgs {
    BvhView bvh; KdView kd; KdFastView kd_fast;
    const float4 * __restrict__ verts; const float4 * __restrict__ normals;
    const float * __restrict__ values; const float * __restrict__ qualities;
    uint32_t n;
    float px, py, pz, max_distance, target_recon;
    const float4 * __restrict__ obs; const uint32_t * __restrict__ num_rows; uint32_t max_rows, stride;
    const float * __restrict__ recons;
    float * hist;
    Heur heur;
};

constexpr int HIST_THREADS = 128;

// K4 (WITH_RECON = false, kernels.cu:293-336) and K5 (true, kernels.cu:338-390): one
// thread per point; contributions are first accumulated in a per-CTA shared histogram
// and flushed with one global atomic per touched bin.
template <bool WITH_RECON>
__global__ void __launch_bounds__(HIST_THREADS)
k_populate_spherical_histogram(HistArgs a)
{
    extern __shared__ float s_hist[];   // kd.n bins
    for (uint32_t i = threadIdx.x; i < a.kd.n; i += HIST_THREADS) s_hist[i] = 0.0f;
    __syncthreads();
    uint32_t id = blockIdx.x * HIST_THREADS + threadIdx.x;
    const bool valid = id < a.n;
    V3 v = ld3(a.verts[valid ? id : 0]), n = ld3(a.normals[valid ? id : 0]);
    V3 v2c = mk(a.px, a.py, a.pz) - v;
    float l = norm(v2c);
    float scale = 1.0f - (l / a.max_distance);
    V3 v2cn = v2c / l;
    float ctheta = dot(v2cn, n);
    bool cast = valid && scale > 0.0f && !(ctheta < 0.087f);
    Ray ray; ray.o = v; ray.d = v2cn; ray.tmin = l * 0.001f; ray.tmax = l;
    bool blocked = trace_any_warp(a.bvh, ray, cast);   // collective: every lane calls
    bool visible = cast && !blocked;
    if (visible) {
        float contrib = 0.0f; bool add = true;
        if (!WITH_RECON) {
            float capture_difficulty = fmaxf(a.qualities[id], 0.0f);
            float min_theta = fminf(a.values[id], 1.484f);
            float scaling = (kPi / 2.0f) / ((kPi / 2.0f) - min_theta);
            float theta = acosf(__saturatef(ctheta));
            float rel_theta = fmaxf(theta - min_theta, 0.0f) * scaling;
            contrib = capture_difficulty * cosf(rel_theta) * scale;
        } else {
            uint32_t rows = a.num_rows[id];
            if (rows >= a.max_rows) add = false;
            else {
                float recon = a.recons[id];
                float new_recon = 0.0f;
                if (rows >= 1) {
                    V3 rel = relative_direction(v2cn, n);
                    const float m_k = a.heur.m_k, m_x0 = kPi / a.heur.m_x0, t_k = a.heur.t_k, t_x0 = kPi / a.heur.t_x0;
                    float sum = 0.0f;
                    for (uint32_t i = 0; i < rows; ++i) {       // heuristic(), kernels.cu:111-136
                        float4 r = a.obs[(size_t)i * a.stride + id];
                        float calpha = dot(rel, mk(r.x, r.y, r.z));
                        float alpha = acosf(fmaxf(-1.0f, fminf(calpha, 1.0f)));
                        float sc = fminf(r.w, scale);
                        float ct = fminf(r.z, rel.z);
                        float matchability = (1.0f - logistic(alpha, m_k, m_x0)) * ct;
                        float triangulation = logistic(alpha, t_k, t_x0) * sc;
                        sum += matchability * triangulation;
                    }
                    new_recon = recon + sum;
                }
                contrib = func(new_recon, a.target_recon) - func(recon, a.target_recon);
            }
        }
        if (add) {
            uint32_t bin = a.kd_fast.table ? kd_find_nn_fast(a.kd_fast, a.kd, -v2cn) : kd_find_nn(a.kd, -v2cn);
            atomicAdd(&s_hist[bin], contrib);
        }
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < a.kd.n; i += HIST_THREADS) {
        float h = s_hist[i];
        if (h != 0.0f) atomicAdd(a.hist + i, h);
    }
}

// K6: evaluate_spherical_histogram, kernels.cu:392-446.  One thread per output pixel.
__global__ void k_evaluate_spherical_histogram(const float * __restrict__ K9, int width, int height,
    KdView kd, const float * __restrict__ sphere_hist, float * __restrict__ hist, uint32_t hw, uint32_t hh,
    uint32_t hstride)
{
    extern __shared__ float4 s_dir[];   // normalised sphere vertices + histogram value
    for (uint32_t i = threadIdx.x; i < kd.n; i += blockDim.x) {
        V3 d = normalized(ld3(kd.verts[i]));
        s_dir[i] = make_float4(d.x, d.y, d.z, sphere_hist[i]);
    }
    __syncthreads();
    uint32_t x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= hw || y >= hh) return;
    float K[9];
    #pragma unroll
    for (int i = 0; i < 9; ++i) K[i] = K9[i];
    float phi = (x / (float)hw) * 2.0f * kPi;
    float theta = (0.5f + (y / (float)hh) / 2.0f) * kPi;
    float stheta = sinf(theta);
    V3 view_dir = mk(stheta * cosf(phi), stheta * sinf(phi), cosf(theta));
    V3 rz = -view_dir;
    V3 up = mk(0.0f, 0.0f, 1.0f);
    bool stable = fabsf(dot(up, rz)) < 0.99f;
    if (!stable) up = mk(cosf(phi), sinf(phi), 0.0f);
    V3 rx = normalized(cross(up, rz));
    V3 ry = normalized(cross(rz, rx));
    float sum = 0.0f;
    for (uint32_t i = 0; i < kd.n; ++i) {
        float4 d4 = s_dir[i];
        V3 dir = ld3(d4);
        if (dot(view_dir, dir) < 0.0f) continue;
        V3 v = mk(dot(rx, dir), dot(ry, dir), dot(rz, dir));
        float p[3];
        #pragma unroll
        for (int r = 0; r < 3; ++r) {
            float s = 0.0f;
            s += K[3 * r + 0] * v.x; s += K[3 * r + 1] * v.y; s += K[3 * r + 2] * v.z;
            p[r] = s;
        }
        float px = p[0] / p[2] - 0.5f, py = p[1] / p[2] - 0.5f;
        if (px < 0.0f || (float)width <= px || py < 0.0f || (float)height <= py) continue;
        sum += d4.w;
    }
    hist[(size_t)y * hstride + x] = sum;
}

// K7: estimate_capture_difficulty, kernels.cu:448-494.  One thread per point like the reference, so a warp
// holds 32 neighbouring points (the cloud is Z-order sorted) that all cast the SAME direction at a time:
// parallel rays from nearby origins walk the same nodes (a warp per point with the directions over the
// lanes, the earlier mapping, gave 32 unrelated rays per warp: 0.66 vs 1.43 Grays/s at 1 M points).  Small CTAs
// keep small clouds spread over all SMs.  Closest-hit traversal
// in the reference order; sums accumulate in the reference's order (ascending direction index).
__global__ void __launch_bounds__(64)
k_estimate_capture_difficulty(BvhView bvh, KdView kd, uint32_t mesh_size, float max_distance,
    const float4 * __restrict__ verts, const float4 * __restrict__ normals, uint32_t n,
    float * __restrict__ values, float * __restrict__ qualities)
{
    const uint32_t id = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = id < n;                                  // whole warps stay in the collective trace
    const V3 v = ld3(verts[valid ? id : n - 1]), nrm = ld3(normals[valid ? id : n - 1]);
    const float l = max_distance;
    float sum = 0.0f, max_ctheta = 0.0f;
    for (uint32_t i = 0; i < kd.n; ++i) {
        const V3 dir = normalized(ld3(__ldg(kd.verts + i)));
        const float ctheta = dot(dir, nrm);
        const bool cast = valid && !(ctheta < 0.087f);
        if (!__any_sync(0xffffffffu, cast)) continue;
        Ray ray; ray.o = v; ray.d = dir; ray.tmin = l * 0.001f; ray.tmax = l;
        uint32_t face; float t;
        const bool hit = trace_closest_warp(bvh, ray, cast, face, t);
        if (cast) {
            if (!hit || face >= mesh_size) { if (ctheta > max_ctheta) max_ctheta = ctheta; }
            else sum += ctheta;
        }
    }
    if (valid) {
        const float mx = (kd.n / 2.0f) * 0.5f;
        values[id] = acosf(__saturatef(max_ctheta));
        qualities[id] = sum / mx;
    }
}

// ---- batched K6 ------------------------------------------------------------------------------
// Whether sphere vertex i lies inside the image of orientation pixel (x, y) does not depend on the
// histogram values (kernels.cu:409-441), so it is computed once per (calib, width, height) as a bit
// matrix with exactly the arithmetic of k_evaluate_spherical_histogram; the per-position work is then
// "sum the selected bins in ascending order" - the same additions in the same order as the reference.
__global__ void k_frustum_bits(const float * __restrict__ K9, int width, int height, KdView kd,
    uint32_t hw, uint32_t hh, uint32_t * __restrict__ bits)
{
    const uint32_t pix = blockIdx.x * blockDim.x + threadIdx.x, npix = hw * hh;
    if (pix >= npix) return;
    const uint32_t x = pix % hw, y = pix / hw;
    float K[9];
    #pragma unroll
    for (int i = 0; i < 9; ++i) K[i] = K9[i];
    float phi = (x / (float)hw) * 2.0f * kPi;
    float theta = (0.5f + (y / (float)hh) / 2.0f) * kPi;
    float stheta = sinf(theta);
    V3 view_dir = mk(stheta * cosf(phi), stheta * sinf(phi), cosf(theta));
    V3 rz = -view_dir;
    V3 up = mk(0.0f, 0.0f, 1.0f);
    bool stable = fabsf(dot(up, rz)) < 0.99f;
    if (!stable) up = mk(cosf(phi), sinf(phi), 0.0f);
    V3 rx = normalized(cross(up, rz));
    V3 ry = normalized(cross(rz, rx));
    const uint32_t words = (kd.n + 31) / 32;
    for (uint32_t wd = 0; wd < words; ++wd) {
        uint32_t m = 0;
        for (uint32_t b = 0; b < 32; ++b) {
            const uint32_t i = wd * 32 + b;
            if (i >= kd.n) break;
            V3 dir = normalized(ld3(__ldg(kd.verts + i)));
            if (dot(view_dir, dir) < 0.0f) continue;
            V3 v = mk(dot(rx, dir), dot(ry, dir), dot(rz, dir));
            float p[3];
            #pragma unroll
            for (int r = 0; r < 3; ++r) {
                float s = 0.0f;
                s += K[3 * r + 0] * v.x; s += K[3 * r + 1] * v.y; s += K[3 * r + 2] * v.z;
                p[r] = s;
            }
            float px = p[0] / p[2] - 0.5f, py = p[1] / p[2] - 0.5f;
            if (px < 0.0f || (float)width <= px || py < 0.0f || (float)height <= py) continue;
            m |= 1u << b;
        }
        bits[(size_t)wd * npix + pix] = m;
    }
}

// gridDim = (positions, parts): a CTA stages its position's histogram row in shared memory and takes one
// contiguous part of the orientation pixels, a thread per pixel (strided); optional image output; first minimum /
// first maximum in row-major order (the host loop of apps/optimize_trajectory/optimize_trajectory.cu:416-428 made a
// device reduction).  Parts are combined by one 64-bit atomic per CTA on packed keys: min key = ordered(value) << 32 |
// pixel (atomicMin: smallest value, then smallest pixel), max key = ordered(value) << 32 | ~pixel (atomicMax: largest
// value, then smallest pixel).  Small batches (the optimiser's ~100 candidates) thus fill the GPU instead of
// running 100 CTAs of 23 pixels per thread.
constexpr int EVB_THREADS = 256;
__device__ __forceinline__ uint32_t f2ord(float f) { uint32_t u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
__global__ void __launch_bounds__(EVB_THREADS)
k_evaluate_spherical_histogram_batch(const uint32_t * __restrict__ bits, uint32_t words, uint32_t npix,
    const float * __restrict__ sphere_hists, uint32_t S, float * __restrict__ images,
    unsigned long long * __restrict__ best_min, unsigned long long * __restrict__ best_max, const float * __restrict__ weights)
{
    extern __shared__ float s_h[];
    __shared__ float s_v[2][EVB_THREADS / 32]; __shared__ uint32_t s_i[2][EVB_THREADS / 32];
    const uint32_t b = blockIdx.x;
    const uint32_t chunk = (npix + gridDim.y - 1) / gridDim.y;
    const uint32_t p0 = blockIdx.y * chunk, p1 = min(npix, p0 + chunk);
    for (uint32_t i = threadIdx.x; i < words * 32; i += EVB_THREADS) s_h[i] = i < S ? sphere_hists[(size_t)b * S + i] : 0.0f;
    __syncthreads();
    float vmin = UAV_INF, vmax = -UAV_INF; uint32_t imin = UAV_NAI, imax = UAV_NAI;
    const float wgt = weights ? __ldg(weights + b) : 1.0f;
    for (uint32_t pix = p0 + threadIdx.x; pix < p1; pix += EVB_THREADS) {
        float sum = 0.0f;
        for (uint32_t wd = 0; wd < words; ++wd) {
            uint32_t m = __ldg(bits + (size_t)wd * npix + pix);
            while (m) { const uint32_t k = __ffs(m) - 1; m &= m - 1; sum += s_h[wd * 32 + k]; }
        }
        if (images) images[(size_t)b * npix + pix] = sum;
        if (weights) sum = wgt * sum;                        // the optimiser's `w * data` (optimize_trajectory.cu:417)
        if (sum < vmin) { vmin = sum; imin = pix; }          // pixels ascend per thread: first occurrence
        if (sum > vmax) { vmax = sum; imax = pix; }
    }
    if (!best_min) return;                                   // images only (single-position entry point): uniform over the CTA
    // lexicographic (value, index) reduction keeps the first occurrence
    for (int o = 16; o > 0; o >>= 1) {
        float ov = __shfl_xor_sync(0xffffffffu, vmin, o); uint32_t oi = __shfl_xor_sync(0xffffffffu, imin, o);
        if (ov < vmin || (ov == vmin && oi < imin)) { vmin = ov; imin = oi; }
        ov = __shfl_xor_sync(0xffffffffu, vmax, o); oi = __shfl_xor_sync(0xffffffffu, imax, o);
        if (ov > vmax || (ov == vmax && oi < imax)) { vmax = ov; imax = oi; }
    }
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) { s_v[0][w] = vmin; s_i[0][w] = imin; s_v[1][w] = vmax; s_i[1][w] = imax; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < EVB_THREADS / 32; ++k) {
            if (s_v[0][k] < vmin || (s_v[0][k] == vmin && s_i[0][k] < imin)) { vmin = s_v[0][k]; imin = s_i[0][k]; }
            if (s_v[1][k] > vmax || (s_v[1][k] == vmax && s_i[1][k] < imax)) { vmax = s_v[1][k]; imax = s_i[1][k]; }
        }
        // a part without a comparable value (empty, NaN, or nothing below +inf / above -inf) leaves the initial key
        if (imin != UAV_NAI) atomicMin(best_min + b, ((unsigned long long)f2ord(vmin + 0.0f) << 32) | imin);
        if (imax != UAV_NAI) atomicMax(best_max + b, ((unsigned long long)f2ord(vmax + 0.0f) << 32) | (uint32_t)~imax);
    }
}

// cacc::nnsearch::find_nn<3> for a batch of queries (libs/cacc/nnsearch.h:67-155); fast.table != null: through the
// certified lookup of kd.cuh (must return the same indices)
__global__ void k_kd_find_nn(KdView kd, KdFastView fast, const float * __restrict__ q, uint64_t n, uint32_t * __restrict__ out)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const V3 v = mk(q[3 * i], q[3 * i + 1], q[3 * i + 2]);
    out[i] = fast.table ? kd_find_nn_fast(fast, kd, v) : kd_find_nn(kd, v);
}

__global__ void k_fill_positions(const float * __restrict__ pos, uint32_t n, float * __restrict__ views)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float * v = views + (size_t)i * UAVMVS_VIEW_FLOATS;
    for (int k = 0; k < UAVMVS_VIEW_FLOATS; ++k) v[k] = k < 3 ? pos[3 * (size_t)i + k] : 0.0f;
}

// ---- fixed-point scale of the batched K4 / K5 (k1.cu:k_retire_hist) -------------------------------------------
// fx[1] = bound of |contribution|, fx[0] = scale = 2^(36 - e) with bound <= 2^e.  K4: bound = the largest quality of the
// cloud (score = max(quality, 0) cos(..) scale with both other factors in [0, 1], kernels.cu:318-333); K5: given.
__global__ void k_fx_init(float * fx, float bound) { fx[0] = 1.0f; fx[1] = bound; }
__global__ void k_fx_max_quality(const float * __restrict__ q, uint32_t n, float * fx)
{
    float m = 0.0f;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float v = q[i];
        if (v > m && v < UAV_INF) m = v;                  // NaN / inf qualities take the float path of k_retire_hist
    }
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && m > 0.0f) atomicMax(reinterpret_cast<int *>(fx + 1), __float_as_int(m));   // non-negative floats order like ints
}
__global__ void k_fx_scale(float * fx)
{
    const float bound = fx[1];
    int e = 0;
    if (bound > 0.0f && bound < UAV_INF) { frexpf(bound, &e); }          // bound = m 2^e, m in [0.5, 1)
    e = max(-60, min(e, 90));
    fx[0] = ldexpf(1.0f, 36 - e);
}
__global__ void k_hist_finish(const unsigned long long * __restrict__ acc, float * __restrict__ hist, size_t n, const float * __restrict__ fx)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const long long a = (long long)acc[i];
    if (a != 0) hist[i] += (float)((double)a / (double)fx[0]);
}

static int launch_populate(cudaStream_t s, bool with_recon, const float view_pos[3], float max_distance,
    float target_recon, const uavmvs_bvh * bvh, const uavmvs_cloud * cloud, const uavmvs_kdtree * kd,
    const uavmvs_obs * obs, const uavmvs_array * recons, uavmvs_array * hist)
{
    UAV_REQUIRE(view_pos && bvh && cloud && kd && hist, "null argument");
    UAV_REQUIRE(hist->n >= kd->num_nodes, "sphere_hist smaller than the KD-tree");
    UAV_REQUIRE(kd->num_nodes * sizeof(float) <= 48 * 1024, "sphere histogram does not fit shared memory");
    HistArgs a;
    a.bvh = view_of(bvh); a.kd = kd_view(kd);
    if (kd_fast_ensure(const_cast<uavmvs_kdtree *>(kd))) return 1;
    a.kd_fast = kd_fast_view(kd);
    a.verts = cloud->vertices; a.normals = cloud->normals; a.values = cloud->values; a.qualities = cloud->qualities;
    a.n = cloud->n;
    a.px = view_pos[0]; a.py = view_pos[1]; a.pz = view_pos[2];
    a.max_distance = max_distance; a.target_recon = target_recon;
    a.obs = nullptr; a.num_rows = nullptr; a.max_rows = 0; a.stride = 0; a.recons = nullptr;
    if (with_recon) {
        UAV_REQUIRE(obs && recons, "null obs/recons");
        UAV_REQUIRE(obs->num_cols == cloud->n && recons->n == cloud->n, "size mismatch");
        a.obs = obs->data; a.num_rows = obs->num_rows; a.max_rows = obs->max_rows;
        a.stride = (uint32_t)(obs->pitch / sizeof(float4)); a.recons = recons->data;
    }
    a.hist = hist->data; a.heur = heur_of(obs);
    size_t smem = kd->num_nodes * sizeof(float);
    unsigned grid = divup(cloud->n, HIST_THREADS);
    if (with_recon) k_populate_spherical_histogram<true><<<grid, HIST_THREADS, smem, uav::counted(s)>>>(a);
    else k_populate_spherical_histogram<false><<<grid, HIST_THREADS, smem, uav::counted(s)>>>(a);
    UAV_CUDA(cudaGetLastError());
    return 0;
}

// Bit matrix of the batched K6 for one (calibration, image size), cached in the KD-tree handle.
static int frustum_bits(uavmvs_kdtree * kd, const float calib[9], int width, int height, uint32_t image_width,
    uint32_t image_height, cudaStream_t s, const uint32_t ** out)
{
    std::lock_guard<std::mutex> lock(kd->fr_mutex);
    const uint32_t npix = image_width * image_height, words = (kd->num_nodes + 31) / 32;
    for (const uavmvs_kdtree::FrustumBits & e : kd->fr_cache)
        if (e.w == width && e.h == height && e.iw == image_width && e.ih == image_height && memcmp(e.calib, calib, 9 * sizeof(float)) == 0) {
            *out = e.bits;
            return 0;
        }
    UAV_REQUIRE(kd->fr_cache.size() < 256, "more than 256 distinct (calibration, image size) combinations on one KD-tree handle");
    uavmvs_kdtree::FrustumBits e;
    memcpy(e.calib, calib, 9 * sizeof(float)); e.w = width; e.h = height; e.iw = image_width; e.ih = image_height; e.bits = nullptr;
    float * d_K = nullptr;
    UAV_CUDA(cudaMalloc(&e.bits, sizeof(uint32_t) * (size_t)words * npix));
    UAV_CUDA(cudaMalloc(&d_K, 9 * sizeof(float)));
    UAV_CUDA(cudaMemcpy(d_K, calib, 9 * sizeof(float), cudaMemcpyHostToDevice));
    k_frustum_bits<<<divup(npix, 128), 128, 0, uav::counted(s)>>>(d_K, width, height, kd_view(kd), image_width, image_height, e.bits);
    UAV_CUDA(cudaGetLastError());
    UAV_CUDA(cudaStreamSynchronize(s));                    // other streams may use the matrix as soon as the lock drops
    UAV_CUDA(cudaFree(d_K));
    kd->fr_cache.push_back(e);
    *out = e.bits;
    return 0;
}


int populate_hist_batch_device(cudaStream_t s, const float * d_pos, uint32_t num_positions, float max_distance,
    float target_recon, const uavmvs_bvh * bvh, const uavmvs_cloud * cloud, const uavmvs_kdtree * kd, const uavmvs_obs * obs,
    const uavmvs_array * recons, float * d_sphere_hists, int no_readback)
{
    UAV_REQUIRE(d_pos && bvh && cloud && kd && d_sphere_hists, "null argument");
    UAV_REQUIRE((obs == nullptr) == (recons == nullptr), "obs and recons go together (both NULL: K4, both set: K5)");
    if (num_positions == 0) return 0;
    HistJob hj;
    hj.kd_nodes = kd->nodes; hj.kd_verts = kd->verts; hj.kd_n = kd->num_nodes;
    if (kd_fast_ensure(const_cast<uavmvs_kdtree *>(kd))) return 1;      // built at create; a no-op here
    hj.kd_fast = kd_fast_view(kd);
    hj.values = cloud->values; hj.qualities = cloud->qualities;
    hj.with_recon = obs ? 1 : 0; hj.target_recon = target_recon;
    hj.obs = nullptr; hj.num_rows = nullptr; hj.max_rows = 0; hj.stride = 0; hj.recons = nullptr;
    if (obs) {
        UAV_REQUIRE(obs->num_cols == cloud->n && recons->n == cloud->n, "size mismatch");
        hj.obs = obs->data; hj.num_rows = obs->num_rows; hj.max_rows = obs->max_rows;
        hj.stride = (uint32_t)(obs->pitch / sizeof(float4)); hj.recons = recons->data;
    }
    hj.hist = d_sphere_hists; hj.heur = heur_of(obs); hj.no_readback = no_readback;
    const size_t cells = (size_t)num_positions * kd->num_nodes;
    UAV_REQUIRE(cells < 0xffffff00ull, "positions x bins does not fit the 32-bit accumulator key");
    UAV_REQUIRE(cloud->n < (1u << 26), "more than 2^26 points per call (fixed-point head room)");
    unsigned long long * d_acc = nullptr; float * d_fx = nullptr, * d_views = nullptr;
    UAV_CUDA(cudaMallocAsync(&d_acc, sizeof(unsigned long long) * cells, s));
    UAV_CUDA(cudaMallocAsync(&d_fx, 2 * sizeof(float), s));
    UAV_CUDA(cudaMallocAsync(&d_views, sizeof(float) * UAVMVS_VIEW_FLOATS * (size_t)num_positions, s));
    UAV_CUDA(cudaMemsetAsync(d_acc, 0, sizeof(unsigned long long) * cells, s));
    if (obs) {
        const float t1 = fabsf(target_recon) + 1.0f;
        k_fx_init<<<1, 1, 0, uav::counted(s)>>>(d_fx, 4.0f * t1 * t1);
    } else {
        k_fx_init<<<1, 1, 0, uav::counted(s)>>>(d_fx, 0.0f);
        k_fx_max_quality<<<min(divup(cloud->n, 256 * 8), 1024u), 256, 0, uav::counted(s)>>>(cloud->qualities, cloud->n, d_fx);
    }
    k_fx_scale<<<1, 1, 0, uav::counted(s)>>>(d_fx);
    hj.acc = d_acc; hj.fx = d_fx;
    k_fill_positions<<<divup(num_positions, 256), 256, 0, uav::counted(s)>>>(d_pos, num_positions, d_views);
    UAV_CUDA(cudaGetLastError());
    int rc = launch_k1(s, 1, d_views, num_positions, max_distance, 0, 0, bvh, cloud, nullptr, nullptr, &hj);
    if (!rc) {
        k_hist_finish<<<divup(cells, 256), 256, 0, uav::counted(s)>>>(d_acc, d_sphere_hists, cells, d_fx);
        if (cudaGetLastError() != cudaSuccess) rc = 1;
    }
    cudaFreeAsync(d_views, s); cudaFreeAsync(d_acc, s); cudaFreeAsync(d_fx, s);
    return rc;
}

int evaluate_hist_batch_device(cudaStream_t s, const float calib[9], int width, int height, uavmvs_kdtree * kd,
    const float * d_sphere_hists, uint32_t num_positions, uint32_t image_width, uint32_t image_height, float * d_images,
    const float * d_weights, unsigned long long * d_best)
{
    UAV_REQUIRE(calib && kd && d_sphere_hists && d_best, "null argument");
    UAV_REQUIRE(image_width >= 1 && image_height >= 1, "empty image");
    const uint32_t npix = image_width * image_height, S = kd->num_nodes, words = (S + 31) / 32;
    UAV_REQUIRE(words * 32 * sizeof(float) <= 48 * 1024, "sphere histogram does not fit shared memory");
    if (num_positions == 0) return 0;
    const uint32_t * bits = nullptr;
    if (frustum_bits(kd, calib, width, height, image_width, image_height, s, &bits)) return 1;
    // pixel parts: enough CTAs for ~8 per SM when the batch is small; every thread keeps at least one pixel
    uint32_t parts = divup(148u * 8u, num_positions);
    parts = max(1u, min(parts, max(1u, npix / EVB_THREADS)));
    UAV_CUDA(cudaMemsetAsync(d_best, 0xff, sizeof(unsigned long long) * (size_t)num_positions, s));
    UAV_CUDA(cudaMemsetAsync(d_best + num_positions, 0, sizeof(unsigned long long) * (size_t)num_positions, s));
    k_evaluate_spherical_histogram_batch<<<dim3(num_positions, parts), EVB_THREADS, words * 32 * sizeof(float), uav::counted(s)>>>(bits, words, npix,
        d_sphere_hists, S, d_images, d_best, d_best + num_positions, d_weights);
    UAV_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace uav

using namespace uav;

extern "C" {

int uavmvs_populate_spherical_histogram(void * stream, const float view_pos[3], float max_distance,
    const uavmvs_bvh * bvh, const uavmvs_cloud * cloud, const uavmvs_kdtree * kd, uavmvs_array * sphere_hist)
{
    return launch_populate((cudaStream_t)stream, false, view_pos, max_distance, 0.0f, bvh, cloud, kd,
        nullptr, nullptr, sphere_hist);
}

int uavmvs_populate_spherical_histogram_recon(void * stream, const float view_pos[3],
    float max_distance, float target_recon, const uavmvs_bvh * bvh, const uavmvs_cloud * cloud,
    const uavmvs_kdtree * kd, const uavmvs_obs * obs, const uavmvs_array * recons, uavmvs_array * sphere_hist)
{
    return launch_populate((cudaStream_t)stream, true, view_pos, max_distance, target_recon, bvh, cloud, kd,
        obs, recons, sphere_hist);
}

int uavmvs_evaluate_spherical_histogram(void * stream, const float calib[9], int width, int height,
    const uavmvs_kdtree * kd, const uavmvs_array * sphere_hist, uavmvs_image * hist)
{
    UAV_REQUIRE(calib && kd && sphere_hist && hist, "null argument");
    UAV_REQUIRE(sphere_hist->n >= kd->num_nodes, "sphere_hist smaller than the KD-tree");
    UAV_REQUIRE(kd->num_nodes * sizeof(float4) <= 48 * 1024, "sphere does not fit shared memory");
    cudaStream_t s = (cudaStream_t)stream;
    const uint32_t S = kd->num_nodes, words = (S + 31) / 32, npix = hist->width * hist->height;
    if (hist->pitch == (uint64_t)hist->width * 4 && words * 32 * sizeof(float) <= 48 * 1024 && !getenv("UAVMVS_K6_REFERENCE_LOOP")) {
        // contiguous rows (any width that is a multiple of 128, e.g. the apps' 128 x 45): the batched kernel with one
        // position - same additions in the same order per pixel, the 5760 x 1272 frustum tests replaced by the cached
        // bit matrix
        const uint32_t * bits = nullptr;
        if (frustum_bits(const_cast<uavmvs_kdtree *>(kd), calib, width, height, hist->width, hist->height, s, &bits)) return 1;
        const uint32_t parts = max(1u, npix / EVB_THREADS);
        k_evaluate_spherical_histogram_batch<<<dim3(1, parts), EVB_THREADS, words * 32 * sizeof(float), uav::counted(s)>>>(bits, words, npix,
            sphere_hist->data, S, hist->data, nullptr, nullptr, nullptr);   // the image only: no extrema, nothing to allocate
        UAV_CUDA(cudaGetLastError());
        return 0;
    }
    float * d_K;
    UAV_CUDA(cudaMallocAsync(&d_K, 9 * sizeof(float), s));
    UAV_CUDA(cudaMemcpyAsync(d_K, calib, 9 * sizeof(float), cudaMemcpyHostToDevice, s));
    dim3 grid(divup(hist->width, 128), hist->height);
    k_evaluate_spherical_histogram<<<grid, 128, kd->num_nodes * sizeof(float4), uav::counted(s)>>>(d_K, width, height,
        kd_view(kd), sphere_hist->data, hist->data, hist->width, hist->height, (uint32_t)(hist->pitch / 4));
    UAV_CUDA(cudaGetLastError());
    UAV_CUDA(cudaFreeAsync(d_K, s));
    return 0;
}

int uavmvs_estimate_capture_difficulty(void * stream, float max_distance, const uavmvs_bvh * bvh,
    uint32_t mesh_size, const uavmvs_kdtree * kd, uavmvs_cloud * cloud)
{
    UAV_REQUIRE(bvh && kd && cloud, "null argument");
    // nested trees: the wide walk of k7.cu (same results; UAVMVS_K7_EXACT=1 keeps the reference walk for every ray)
    if (uav::k7_wide_available(bvh) && !getenv("UAVMVS_K7_EXACT"))
        return uav::k7_wide((cudaStream_t)stream, max_distance, bvh, mesh_size, kd, cloud);
    k_estimate_capture_difficulty<<<divup(cloud->n, 64), 64, 0, uav::counted((cudaStream_t)stream)>>>(
        view_of(bvh), kd_view(kd), mesh_size, max_distance, cloud->vertices, cloud->normals, cloud->n,
        cloud->values, cloud->qualities);
    UAV_CUDA(cudaGetLastError());
    return 0;
}

int uavmvs_populate_spherical_histogram_batch(void * stream, const float * positions, uint32_t num_positions,
    float max_distance, float target_recon, const uavmvs_bvh * bvh, const uavmvs_cloud * cloud,
    const uavmvs_kdtree * kd, const uavmvs_obs * obs, const uavmvs_array * recons, uavmvs_array * sphere_hists)
{
    UAV_REQUIRE(positions && kd && sphere_hists, "null argument");
    UAV_REQUIRE((uint64_t)sphere_hists->n >= (uint64_t)num_positions * kd->num_nodes, "sphere_hists smaller than [positions][bins]");
    if (num_positions == 0) return 0;
    cudaStream_t s = (cudaStream_t)stream;
    float * d_pos = nullptr;
    UAV_CUDA(cudaMallocAsync(&d_pos, sizeof(float) * 3 * (size_t)num_positions, s));
    UAV_CUDA(cudaMemcpyAsync(d_pos, positions, sizeof(float) * 3 * (size_t)num_positions, cudaMemcpyHostToDevice, s));
    const int rc = populate_hist_batch_device(s, d_pos, num_positions, max_distance, target_recon, bvh, cloud, kd, obs, recons,
        sphere_hists->data, 0);
    cudaFreeAsync(d_pos, s);
    return rc;
}

int uavmvs_kdtree_find_nn(void * stream, uavmvs_kdtree * kd, const float * queries, uint64_t n, int certified_lookup,
    uint32_t * out_idx)
{
    UAV_REQUIRE(kd && (queries || n == 0) && (out_idx || n == 0), "null argument");
    if (n == 0) return 0;
    cudaStream_t s = (cudaStream_t)stream;
    KdFastView fv = kd_fast_view(nullptr);
    if (certified_lookup) {
        if (kd_fast_ensure(kd)) return 1;
        fv = kd_fast_view(kd);
        UAV_REQUIRE(fv.table, "this vertex set has no certified lookup (more than two coincident vertices, or not a tree)");
    }
    float * d_q; uint32_t * d_o;
    UAV_CUDA(cudaMallocAsync(&d_q, sizeof(float) * 3 * n, s));
    UAV_CUDA(cudaMallocAsync(&d_o, sizeof(uint32_t) * n, s));
    UAV_CUDA(cudaMemcpyAsync(d_q, queries, sizeof(float) * 3 * n, cudaMemcpyHostToDevice, s));
    k_kd_find_nn<<<divup(n, 256), 256, 0, uav::counted(s)>>>(kd_view(kd), fv, d_q, n, d_o);
    UAV_CUDA(cudaGetLastError());
    UAV_CUDA(cudaMemcpyAsync(out_idx, d_o, sizeof(uint32_t) * n, cudaMemcpyDeviceToHost, s));
    UAV_CUDA(cudaStreamSynchronize(s));
    UAV_CUDA(cudaFreeAsync(d_q, s)); UAV_CUDA(cudaFreeAsync(d_o, s));
    return 0;
}

int uavmvs_evaluate_spherical_histogram_batch(void * stream, const float calib[9], int width, int height,
    uavmvs_kdtree * kd, const uavmvs_array * sphere_hists, uint32_t num_positions, uint32_t image_width,
    uint32_t image_height, uavmvs_array * images, float * h_min, uint32_t * h_argmin, float * h_max, uint32_t * h_argmax)
{
    UAV_REQUIRE(calib && kd && sphere_hists, "null argument");
    UAV_REQUIRE((uint64_t)sphere_hists->n >= (uint64_t)num_positions * kd->num_nodes, "sphere_hists smaller than [positions][bins]");
    const uint32_t npix = image_width * image_height;
    UAV_REQUIRE(!images || (uint64_t)images->n >= (uint64_t)num_positions * npix, "images smaller than [positions][h][w]");
    if (num_positions == 0) return 0;
    cudaStream_t s = (cudaStream_t)stream;
    unsigned long long * d_best;                      // [min keys | max keys]
    UAV_CUDA(cudaMallocAsync(&d_best, 2 * sizeof(unsigned long long) * (size_t)num_positions, s));
    if (int rc = evaluate_hist_batch_device(s, calib, width, height, kd, sphere_hists->data, num_positions, image_width, image_height,
            images ? images->data : nullptr, nullptr, d_best)) { cudaFreeAsync(d_best, s); return rc; }
    if (h_min || h_argmin || h_max || h_argmax) {
        std::vector<unsigned long long> hb(2 * (size_t)num_positions);
        UAV_CUDA(cudaMemcpyAsync(hb.data(), d_best, sizeof(unsigned long long) * hb.size(), cudaMemcpyDeviceToHost, s));
        UAV_CUDA(cudaStreamSynchronize(s));
        auto ord2float = [](uint32_t o) { uint32_t u = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o; float f; memcpy(&f, &u, 4); return f; };
        for (uint32_t i = 0; i < num_positions; ++i) {
            const unsigned long long kmin = hb[i], kmax = hb[num_positions + i];
            const bool has_min = kmin != ~0ull, has_max = kmax != 0ull;
            if (h_min) h_min[i] = has_min ? ord2float((uint32_t)(kmin >> 32)) : INFINITY;
            if (h_argmin) h_argmin[i] = has_min ? (uint32_t)kmin : 0xFFFFFFFFu;
            if (h_max) h_max[i] = has_max ? ord2float((uint32_t)(kmax >> 32)) : -INFINITY;
            if (h_argmax) h_argmax[i] = has_max ? ~(uint32_t)kmax : 0xFFFFFFFFu;
        }
    }
    UAV_CUDA(cudaFreeAsync(d_best, s));
    return 0;
}

}  // extern "C"

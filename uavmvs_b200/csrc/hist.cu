// Spherical-histogram path: K4/K5 populate_spherical_histogram, K6
// evaluate_spherical_histogram, K7 estimate_capture_difficulty.
// Reference: libs/eval/kernels.cu:293-494, libs/cacc/nnsearch.h:67-155.
#include "common.cuh"
#include "trace.cuh"

namespace uav {

constexpr float kPi = 3.14159265f;

struct KdView { const uint4 * __restrict__ nodes; const float4 * __restrict__ verts; uint32_t n; };

// libs/cacc/nnsearch.h:67-155: depth-first nearest neighbour, strict '<' (first found wins)
__device__ __forceinline__ uint32_t kd_find_nn(const KdView & kd, V3 q)
{
    uint32_t idx = UAV_NAI; float max_dist = UAV_INF;
    uint32_t node_idx = 0; bool down = true;
    uint32_t stack[40]; int sp = 0;
    while (true) {
        uint4 node = __ldg(kd.nodes + node_idx);          // vid, dim, left, right
        float4 v4 = __ldg(kd.verts + node.x);
        V3 vert = ld3(v4);
        float qd = node.y == 0 ? q.x : (node.y == 1 ? q.y : q.z);
        float vd = node.y == 0 ? vert.x : (node.y == 1 ? vert.y : vert.z);
        float diff = qd - vd;
        if (down) {
            float dist = norm(q - vert);
            if (dist < max_dist) { idx = node.x; max_dist = dist; }
            if (node.z != UAV_NAI || node.w != UAV_NAI) {
                if (node.z != UAV_NAI && node.w != UAV_NAI) stack[sp++] = node_idx;
                uint32_t next = (diff < 0.0f) ? node.z : node.w;
                uint32_t other = (diff < 0.0f) ? node.w : node.z;
                node_idx = (next != UAV_NAI) ? next : other;
            } else { down = false; node_idx = UAV_NAI; }
        } else {
            if (fabsf(diff) < max_dist) { down = true; node_idx = (diff < 0.0f) ? node.w : node.z; }
            else { down = false; node_idx = UAV_NAI; }
        }
        if (node_idx == UAV_NAI) { if (sp == 0) break; node_idx = stack[--sp]; }
    }
    return idx;
}

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
    BvhView bvh; KdView kd;
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
            uint32_t bin = kd_find_nn(a.kd, -v2cn);
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

// K7: estimate_capture_difficulty, kernels.cu:448-494.  One warp per point, lanes stride
// over the sphere directions; closest-hit traversal in the reference order.
__global__ void __launch_bounds__(128)
k_estimate_capture_difficulty(BvhView bvh, KdView kd, uint32_t mesh_size, float max_distance,
    const float4 * __restrict__ verts, const float4 * __restrict__ normals, uint32_t n,
    float * __restrict__ values, float * __restrict__ qualities)
{
    uint32_t lane = threadIdx.x & 31;
    uint32_t id = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (id >= n) return;
    V3 v = ld3(verts[id]), nrm = ld3(normals[id]);
    float l = max_distance;
    float sum = 0.0f, max_ctheta = 0.0f;
    for (uint32_t base = 0; base < kd.n; base += 32) {
        uint32_t i = base + lane;
        bool in = i < kd.n;
        V3 dir = normalized(ld3(__ldg(kd.verts + (in ? i : 0))));
        float ctheta = dot(dir, nrm);
        bool cast = in && !(ctheta < 0.087f);
        Ray ray; ray.o = v; ray.d = dir; ray.tmin = l * 0.001f; ray.tmax = l;
        uint32_t face; float t;
        bool hit = trace_closest_warp(bvh, ray, cast, face, t);
        if (cast) {
            if (!hit || face >= mesh_size) { if (ctheta > max_ctheta) max_ctheta = ctheta; }
            else sum += ctheta;
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        sum += __shfl_xor_sync(0xffffffffu, sum, o);
        max_ctheta = fmaxf(max_ctheta, __shfl_xor_sync(0xffffffffu, max_ctheta, o));
    }
    if (lane == 0) {
        float mx = (kd.n / 2.0f) * 0.5f;
        values[id] = acosf(__saturatef(max_ctheta));
        qualities[id] = sum / mx;
    }
}

static KdView kd_view(const uavmvs_kdtree * k) { KdView v; v.nodes = k->nodes; v.verts = k->verts; v.n = k->num_nodes; return v; }

static int launch_populate(cudaStream_t s, bool with_recon, const float view_pos[3], float max_distance,
    float target_recon, const uavmvs_bvh * bvh, const uavmvs_cloud * cloud, const uavmvs_kdtree * kd,
    const uavmvs_obs * obs, const uavmvs_array * recons, uavmvs_array * hist)
{
    UAV_REQUIRE(view_pos && bvh && cloud && kd && hist, "null argument");
    UAV_REQUIRE(hist->n >= kd->num_nodes, "sphere_hist smaller than the KD-tree");
    UAV_REQUIRE(kd->num_nodes * sizeof(float) <= 48 * 1024, "sphere histogram does not fit shared memory");
    HistArgs a;
    a.bvh = view_of(bvh); a.kd = kd_view(kd);
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
    a.hist = hist->data; a.heur = current_heur();
    size_t smem = kd->num_nodes * sizeof(float);
    unsigned grid = divup(cloud->n, HIST_THREADS);
    if (with_recon) k_populate_spherical_histogram<true><<<grid, HIST_THREADS, smem, s>>>(a);
    else k_populate_spherical_histogram<false><<<grid, HIST_THREADS, smem, s>>>(a);
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
    float * d_K;
    UAV_CUDA(cudaMallocAsync(&d_K, 9 * sizeof(float), s));
    UAV_CUDA(cudaMemcpyAsync(d_K, calib, 9 * sizeof(float), cudaMemcpyHostToDevice, s));
    dim3 grid(divup(hist->width, 128), hist->height);
    k_evaluate_spherical_histogram<<<grid, 128, kd->num_nodes * sizeof(float4), s>>>(d_K, width, height,
        kd_view(kd), sphere_hist->data, hist->data, hist->width, hist->height, (uint32_t)(hist->pitch / 4));
    UAV_CUDA(cudaGetLastError());
    UAV_CUDA(cudaFreeAsync(d_K, s));
    return 0;
}

int uavmvs_estimate_capture_difficulty(void * stream, float max_distance, const uavmvs_bvh * bvh,
    uint32_t mesh_size, const uavmvs_kdtree * kd, uavmvs_cloud * cloud)
{
    UAV_REQUIRE(bvh && kd && cloud, "null argument");
    k_estimate_capture_difficulty<<<divup((uint64_t)cloud->n * 32, 128), 128, 0, (cudaStream_t)stream>>>(
        view_of(bvh), kd_view(kd), mesh_size, max_distance, cloud->vertices, cloud->normals, cloud->n,
        cloud->values, cloud->qualities);
    UAV_CUDA(cudaGetLastError());
    return 0;
}

}  // extern "C"

// One simplex-downhill step for a batch of views, entirely on the device: the optimiser's inner loop
// (apps/optimize_trajectory/optimize_trajectory.cu:357-470) with libs/opti/nelder_mead.h:29-121 as the update rule.
//
// The reference runs one host thread per selected view; every objective evaluation is K5 + K6 + a 23 KB D2H + a host
// arg-min, and Nelder-Mead's decisions are taken on the host between them (5-6 evaluations per step).  Here the selected
// views advance in lock-step and the decisions are three tiny kernels between three batched evaluations:
//
//   batch 1   the 4 vertices of every simplex            -> values        (nelder_mead.h:35-37)
//   k_nm_reflect  ranks (insertion order of std::sort on 4 items), centroid of the 3 best, reflected point (:39-64)
//   batch 2   the reflected points                       -> refl_value    (:66)
//   k_nm_decide   expansion / outside contraction / inside contraction candidate, or none (:68-119)
//   batch 3   the candidates (views without one get a position beyond reach: no rays)
//   k_nm_finish   vertex replacement or shrink, returned (vertex id, value), best orientation seen (:396-436 of the app)
//
// An evaluation (`func`, optimize_trajectory.cu:366-438) = altitude test, closest_point airspace test (closest.cu), batched
// K5 with a weight per position, batched K6 + first arg-min on the device.  Nothing is copied to the host inside a step;
// the ray buffers of the three K5 launches are sized for the worst case instead of reading the bound back (k1.cu).
#include "common.cuh"

namespace uav {

constexpr float kPiNm = 3.14159265f;
constexpr float kFar = 1.0e18f;            // a position no point is in range of: the K5 launch does nothing for it

struct NmState {               // per view, device
    float verts[4][3];         // the simplex (Simplex<3>::verts)
    float values[4];           // func at the vertices, this step
    int ranks[4];              // indices by ascending value
    float hpc[3], refl[3], cand[3];
    float refl_value, cand_value;
    int action;                // 0 none (reflection kept / already decided), 1 expansion, 2 outside contraction, 3 inside contraction
    // results of the step
    int ret_vid; float ret_value;
    float vmin, vtheta, vphi;  // best orientation over every evaluation of this step (vmin starts at 0: all values are negative)
};

// ---- objective evaluation ------------------------------------------------------------------------------------
// pre: altitude and airspace tests; writes the position the histogram is computed from (or kFar), the weight w, and for
// positions that return early the value itself.  pos: [n][3]; mode[i] = 0: histogram, value = min(w * image); 1: early value.
__global__ void k_nm_pre(const float * __restrict__ pos, uint32_t n, float min_distance, const uint8_t * __restrict__ found,
    const uint32_t * __restrict__ tri_index, const float * __restrict__ closest, const float * __restrict__ raw_tris,
    float * __restrict__ hist_pos, float * __restrict__ weight, float * __restrict__ early, int * __restrict__ mode)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const V3 p = mk(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]);
    float w = 1.0f, e = 0.0f; int m = 0;
    if (!(fabsf(p.x) < 1e17f && fabsf(p.y) < 1e17f && fabsf(p.z) < 1e17f)) { m = 1; e = 0.0f; }   // placeholder (no candidate)
    else if (p.z < min_distance) { m = 1; e = min_distance - p.z; }                           // optimize_trajectory.cu:373
    else if (found[i]) {                                                                        // :376-386
        const float * t = raw_tris + 9 * (size_t)tri_index[i];
        const V3 a = mk(t[0], t[1], t[2]), b = mk(t[3], t[4], t[5]), c = mk(t[6], t[7], t[8]);
        const V3 normal = cross(b - a, c - a);                                                  // acc calculate_normal, primitives.h:165-171
        const V3 cpp = p - mk(closest[3 * i], closest[3 * i + 1], closest[3 * i + 2]);
        const float dist = norm(cpp);
        if (dot(normal, cpp) < 0.0f) { m = 1; e = dist; }
        else w = 1.0f - (min_distance - dist) / min_distance;
    }
    mode[i] = m; early[i] = e; weight[i] = w;
    hist_pos[3 * i] = m ? kFar : p.x; hist_pos[3 * i + 1] = m ? kFar : p.y; hist_pos[3 * i + 2] = m ? kFar : p.z;
}

__device__ __forceinline__ float nm_ord2float(uint32_t o) { const uint32_t u = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o; return __uint_as_float(u); }

// post: value and orientation of every evaluated position from the packed first-minimum keys of the batched K6
// (optimize_trajectory.cu:411-437: min starts at 0, only strictly smaller products count, first occurrence wins)
__global__ void k_nm_post(const unsigned long long * __restrict__ best_min, uint32_t n, const int * __restrict__ mode,
    const float * __restrict__ early, uint32_t image_width, uint32_t image_height,
    float * __restrict__ value, float * __restrict__ theta, float * __restrict__ phi, int * __restrict__ from_hist)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float v = 0.0f, th = 0.0f, ph = 0.0f; int fh = 0;
    if (mode[i]) v = early[i];
    else {
        fh = 1;
        const unsigned long long k = best_min[i];
        if (k != ~0ull) {
            const float mn = nm_ord2float((uint32_t)(k >> 32));
            if (mn < 0.0f) {
                const uint32_t pix = (uint32_t)k, y = pix / image_width, x = pix % image_width;
                v = mn;
                th = (0.5f + (y / (float)image_height) / 2.0f) * kPiNm;
                ph = (x / (float)image_width) * 2.0f * kPiNm;
            }
        }
    }
    value[i] = v; theta[i] = th; phi[i] = ph; from_hist[i] = fh;
}

// the app's closure keeps the best orientation over all histogram evaluations of the step (`if (min < vmin)`, :430-434)
__device__ __forceinline__ void nm_track(NmState & st, float v, float th, float ph, int from_hist)
{
    if (from_hist && v < st.vmin) { st.vmin = v; st.vtheta = th; st.vphi = ph; }
}

__global__ void k_nm_gather_verts(const NmState * __restrict__ states, const uint32_t * __restrict__ active, uint32_t n, float * __restrict__ pos)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 4 * n) return;
    const NmState & st = states[active[i >> 2]];
    for (int k = 0; k < 3; ++k) pos[3 * i + k] = st.verts[i & 3][k];
}

__global__ void k_nm_reflect(NmState * __restrict__ states, const uint32_t * __restrict__ active, uint32_t n,
    const float * __restrict__ value, const float * __restrict__ theta, const float * __restrict__ phi, const int * __restrict__ from_hist,
    float * __restrict__ pos_out)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    NmState & st = states[active[i]];
    st.vmin = 0.0f; st.vtheta = 0.0f; st.vphi = 0.0f;
    for (int k = 0; k < 4; ++k) { st.values[k] = value[4 * i + k]; nm_track(st, value[4 * i + k], theta[4 * i + k], phi[4 * i + k], from_hist[4 * i + k]); }
    // std::sort of 4 indices by value: libstdc++ sorts fewer than 16 elements by insertion, which leaves equal values in
    // index order
    int r[4] = {0, 1, 2, 3};
    for (int a = 1; a < 4; ++a) {
        const int x = r[a]; int b = a;
        while (b > 0 && st.values[x] < st.values[r[b - 1]]) { r[b] = r[b - 1]; --b; }
        r[b] = x;
    }
    for (int k = 0; k < 4; ++k) st.ranks[k] = r[k];
    V3 hpc = mk(0.0f, 0.0f, 0.0f);
    for (int k = 0; k < 3; ++k) hpc = hpc + mk(st.verts[r[k]][0], st.verts[r[k]][1], st.verts[r[k]][2]);
    hpc = hpc / 3.0f;                                                                            // hpc /= N
    const V3 worst = mk(st.verts[r[3]][0], st.verts[r[3]][1], st.verts[r[3]][2]);
    const V3 refl = hpc + (hpc - worst);
    st.hpc[0] = hpc.x; st.hpc[1] = hpc.y; st.hpc[2] = hpc.z;
    st.refl[0] = refl.x; st.refl[1] = refl.y; st.refl[2] = refl.z;
    pos_out[3 * i] = refl.x; pos_out[3 * i + 1] = refl.y; pos_out[3 * i + 2] = refl.z;
}

__global__ void k_nm_decide(NmState * __restrict__ states, const uint32_t * __restrict__ active, uint32_t n,
    const float * __restrict__ value, const float * __restrict__ theta, const float * __restrict__ phi, const int * __restrict__ from_hist,
    float * __restrict__ pos_out)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    NmState & st = states[active[i]];
    st.refl_value = value[i];
    nm_track(st, value[i], theta[i], phi[i], from_hist[i]);
    const float best = st.values[st.ranks[0]], lousy = st.values[st.ranks[2]], worst_v = st.values[st.ranks[3]];
    const V3 hpc = mk(st.hpc[0], st.hpc[1], st.hpc[2]), refl = mk(st.refl[0], st.refl[1], st.refl[2]);
    const V3 worst = mk(st.verts[st.ranks[3]][0], st.verts[st.ranks[3]][1], st.verts[st.ranks[3]][2]);
    V3 cand = mk(kFar, kFar, kFar); int action = 0;
    if (st.refl_value < best) { cand = refl + (refl - hpc); action = 1; }
    else if (st.refl_value < worst_v) {
        if (!(st.refl_value < lousy)) { cand = hpc + (hpc - worst) * 0.5f; action = 2; }
    } else { cand = hpc - (hpc - worst) * 0.5f; action = 3; }
    st.action = action;
    st.cand[0] = cand.x; st.cand[1] = cand.y; st.cand[2] = cand.z;
    pos_out[3 * i] = cand.x; pos_out[3 * i + 1] = cand.y; pos_out[3 * i + 2] = cand.z;
}

__global__ void k_nm_finish(NmState * __restrict__ states, const uint32_t * __restrict__ active, uint32_t n,
    const float * __restrict__ value, const float * __restrict__ theta, const float * __restrict__ phi, const int * __restrict__ from_hist)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    NmState & st = states[active[i]];
    const int bi = st.ranks[0], wi = st.ranks[3];
    const float best = st.values[bi], worst_v = st.values[wi];
    int vid = bi; float val = best;
    auto set_vert = [&](int k, const float * p) { st.verts[k][0] = p[0]; st.verts[k][1] = p[1]; st.verts[k][2] = p[2]; };
    auto shrink = [&]() {                       // nelder_mead.h:20-27: towards the best vertex, every vertex
        const V3 t = mk(st.verts[bi][0], st.verts[bi][1], st.verts[bi][2]);
        for (int k = 0; k < 4; ++k) {
            const V3 v = mk(st.verts[k][0], st.verts[k][1], st.verts[k][2]);
            const V3 nv = t + (v - t) * 0.5f;
            st.verts[k][0] = nv.x; st.verts[k][1] = nv.y; st.verts[k][2] = nv.z;
        }
    };
    if (st.action == 0) {                       // reflection between best and lousy
        set_vert(wi, st.refl);
    } else {
        st.cand_value = value[i];
        nm_track(st, value[i], theta[i], phi[i], from_hist[i]);
        if (st.action == 1) {
            if (st.cand_value < best) { set_vert(wi, st.cand); vid = wi; val = st.cand_value; }      // expansion
            else { set_vert(wi, st.refl); vid = wi; val = st.refl_value; }                           // reflection
        } else {                                // outside / inside contraction, else shrink
            if (st.cand_value < worst_v) {
                set_vert(wi, st.cand);
                if (st.cand_value < best) { vid = wi; val = st.cand_value; }
            } else shrink();
        }
    }
    st.ret_vid = vid; st.ret_value = val;
}

// The pose an optimised view takes after the step (optimize_trajectory.cu:441-467): position = the returned simplex vertex,
// orientation = utp::rotation_from_spherical(vtheta, vphi) (libs/utp/trajectory.cpp:14-39), w2c = [rot | -rot pos].
// Views whose step returned a non-negative value ("does not contribute", :446) are left as they are: their
// re-initialisation (:447-461) is the application's business.  views: [n][28] in the order of `active`, updated in place.
__global__ void k_nm_apply_views(const NmState * __restrict__ states, const uint32_t * __restrict__ active, uint32_t n, float * __restrict__ views)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const NmState & st = states[active[i]];
    if (!(st.ret_value < 0.0f)) return;
    const V3 pos = mk(st.verts[st.ret_vid][0], st.verts[st.ret_vid][1], st.verts[st.ret_vid][2]);
    const float ct = cosf(st.vtheta), sn = sinf(st.vtheta), cp = cosf(st.vphi), sp = sinf(st.vphi);
    const V3 rz = normalized(mk(sn * cp, sn * sp, ct));
    V3 up = mk(0.0f, 0.0f, -1.0f);
    if (!(fabsf(dot(up, rz)) < 0.99f)) up = mk(cp, sp, 0.0f);
    const V3 rx = normalized(cross(up, rz));
    const V3 ry = normalized(cross(rz, rx));
    float * v = views + (size_t)i * UAVMVS_VIEW_FLOATS;
    v[0] = pos.x; v[1] = pos.y; v[2] = pos.z;
    const V3 rows[3] = {rx, ry, rz};
    for (int r = 0; r < 3; ++r) {
        v[3 + 4 * r] = rows[r].x; v[3 + 4 * r + 1] = rows[r].y; v[3 + 4 * r + 2] = rows[r].z;
        v[3 + 4 * r + 3] = -dot(rows[r], pos);                                      // trans = -rot * pos
    }
    v[15] = 0.0f; v[16] = 0.0f; v[17] = 0.0f; v[18] = 1.0f;
}

}  // namespace uav

// ---- host side --------------------------------------------------------------------------------------------------
struct uavmvs_nm {
    uint32_t n = 0, bins = 0, last_active = 0;
    uav::NmState * states = nullptr;
    uint32_t * active = nullptr;
    // scratch for up to 4 n positions
    float * pos = nullptr, * hist_pos = nullptr, * weight = nullptr, * early = nullptr, * closest = nullptr, * hists = nullptr;
    float * value = nullptr, * theta = nullptr, * phi = nullptr;
    int * mode = nullptr, * from_hist = nullptr;
    uint8_t * found = nullptr; uint32_t * tri = nullptr;
    unsigned long long * best = nullptr;
};

namespace uav {

struct NmCall {
    cudaStream_t s; uavmvs_nm * h; float min_distance, max_distance, target_recon; const float * calib; int width, height;
    const uavmvs_bvh * bvh; const uavmvs_bvh * airspace; const uavmvs_cloud * cloud; uavmvs_kdtree * kd;
    const uavmvs_obs * obs; const uavmvs_array * recons;
};

// func() of optimize_trajectory.cu:366-438 for the m positions in h->pos
static int nm_evaluate(const NmCall & c, uint32_t m)
{
    uavmvs_nm * h = c.h; cudaStream_t s = c.s;
    if (launch_closest_point(s, c.airspace, h->pos, m, c.min_distance, h->found, h->tri, h->closest)) return 1;
    k_nm_pre<<<divup(m, 128), 128, 0, counted(s)>>>(h->pos, m, c.min_distance, h->found, h->tri, h->closest, c.airspace->tris,
        h->hist_pos, h->weight, h->early, h->mode);
    UAV_CUDA(cudaGetLastError());
    UAV_CUDA(cudaMemsetAsync(h->hists, 0, sizeof(float) * (size_t)m * h->bins, s));                      // dcon_hist->null()
    if (populate_hist_batch_device(s, h->hist_pos, m, c.max_distance, c.target_recon, c.bvh, c.cloud, c.kd, c.obs, c.recons, h->hists, 1)) return 1;
    if (evaluate_hist_batch_device(s, c.calib, c.width, c.height, c.kd, h->hists, m, 128, 45, nullptr, h->weight, h->best)) return 1;
    k_nm_post<<<divup(m, 128), 128, 0, counted(s)>>>(h->best, m, h->mode, h->early, 128, 45, h->value, h->theta, h->phi, h->from_hist);
    UAV_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace uav

using namespace uav;

extern "C" {

int uavmvs_nm_create(uint32_t num_views, uint32_t num_bins, uavmvs_nm ** out)
{
    UAV_REQUIRE(out && num_views >= 1 && num_bins >= 1, "bad argument");
    uavmvs_nm * h = new uavmvs_nm();
    h->n = num_views; h->bins = num_bins;
    const size_t m = 4 * (size_t)num_views;
    auto fail = [&](cudaError_t e) { if (e != cudaSuccess) { uav::set_error(__FILE__, __LINE__, "cudaMalloc", cudaGetErrorString(e)); return true; } return false; };
    bool bad = false;
    bad |= fail(cudaMalloc(&h->states, sizeof(NmState) * (size_t)num_views));
    bad |= fail(cudaMalloc(&h->active, sizeof(uint32_t) * (size_t)num_views));
    bad |= fail(cudaMalloc(&h->pos, 12 * m)); bad |= fail(cudaMalloc(&h->hist_pos, 12 * m)); bad |= fail(cudaMalloc(&h->closest, 12 * m));
    bad |= fail(cudaMalloc(&h->weight, 4 * m)); bad |= fail(cudaMalloc(&h->early, 4 * m)); bad |= fail(cudaMalloc(&h->value, 4 * m));
    bad |= fail(cudaMalloc(&h->theta, 4 * m)); bad |= fail(cudaMalloc(&h->phi, 4 * m)); bad |= fail(cudaMalloc(&h->mode, 4 * m));
    bad |= fail(cudaMalloc(&h->from_hist, 4 * m)); bad |= fail(cudaMalloc(&h->found, m)); bad |= fail(cudaMalloc(&h->tri, 4 * m));
    bad |= fail(cudaMalloc(&h->best, 16 * m)); bad |= fail(cudaMalloc(&h->hists, sizeof(float) * m * num_bins));
    if (!bad) bad |= fail(cudaMemset(h->states, 0, sizeof(NmState) * (size_t)num_views));
    if (bad) { uavmvs_nm_destroy(h); return 1; }
    *out = h;
    return 0;
}

int uavmvs_nm_destroy(uavmvs_nm * h)
{
    if (!h) return 0;
    cudaFree(h->states); cudaFree(h->active); cudaFree(h->pos); cudaFree(h->hist_pos); cudaFree(h->closest); cudaFree(h->weight);
    cudaFree(h->early); cudaFree(h->value); cudaFree(h->theta); cudaFree(h->phi); cudaFree(h->mode); cudaFree(h->from_hist);
    cudaFree(h->found); cudaFree(h->tri); cudaFree(h->best); cudaFree(h->hists);
    delete h;
    return 0;
}

int uavmvs_nm_upload(uavmvs_nm * h, const float * simplices)
{
    UAV_REQUIRE(h && simplices, "null argument");
    std::vector<NmState> st(h->n);
    memset(st.data(), 0, sizeof(NmState) * st.size());
    for (uint32_t i = 0; i < h->n; ++i) memcpy(st[i].verts, simplices + 12 * (size_t)i, 12 * sizeof(float));
    UAV_CUDA(cudaMemcpy(h->states, st.data(), sizeof(NmState) * st.size(), cudaMemcpyHostToDevice));
    return 0;
}

int uavmvs_nm_download(void * stream, const uavmvs_nm * h, float * simplices, uint32_t * ret_vid, float * ret_value,
    float * vmin, float * vtheta, float * vphi)
{
    UAV_REQUIRE(h, "null argument");
    std::vector<NmState> st(h->n);
    UAV_CUDA(cudaMemcpyAsync(st.data(), h->states, sizeof(NmState) * st.size(), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    UAV_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    for (uint32_t i = 0; i < h->n; ++i) {
        if (simplices) memcpy(simplices + 12 * (size_t)i, st[i].verts, 12 * sizeof(float));
        if (ret_vid) ret_vid[i] = (uint32_t)st[i].ret_vid;
        if (ret_value) ret_value[i] = st[i].ret_value;
        if (vmin) vmin[i] = st[i].vmin;
        if (vtheta) vtheta[i] = st[i].vtheta;
        if (vphi) vphi[i] = st[i].vphi;
    }
    return 0;
}

int uavmvs_nm_step(void * stream, uavmvs_nm * h, const uint32_t * active, uint32_t num_active, float min_distance,
    float max_distance, float target_recon, const float calib[9], int width, int height, const uavmvs_bvh * bvh,
    const uavmvs_bvh * airspace, const uavmvs_cloud * cloud, uavmvs_kdtree * kd, const uavmvs_obs * obs, const uavmvs_array * recons)
{
    UAV_REQUIRE(h && active && calib && bvh && airspace && cloud && kd && obs && recons, "null argument");
    UAV_REQUIRE(num_active <= h->n, "more active views than the handle holds");
    UAV_REQUIRE(kd->num_nodes == h->bins, "the handle was created for another number of histogram bins");
    UAV_REQUIRE(min_distance > 0.0f, "min_distance must be positive");
    for (uint32_t i = 0; i < num_active; ++i) UAV_REQUIRE(active[i] < h->n, "active view index outside the handle");
    if (num_active == 0) return 0;
    cudaStream_t s = (cudaStream_t)stream;
    NmCall c = {s, h, min_distance, max_distance, target_recon, calib, width, height, bvh, airspace, cloud, kd, obs, recons};
    const uint32_t n = num_active;
    UAV_CUDA(cudaMemcpyAsync(h->active, active, sizeof(uint32_t) * (size_t)n, cudaMemcpyHostToDevice, s));
    k_nm_gather_verts<<<divup(4 * n, 128), 128, 0, counted(s)>>>(h->states, h->active, n, h->pos);
    UAV_CUDA(cudaGetLastError());
    if (int rc = nm_evaluate(c, 4 * n)) return rc;                                              // the four vertices
    k_nm_reflect<<<divup(n, 128), 128, 0, counted(s)>>>(h->states, h->active, n, h->value, h->theta, h->phi, h->from_hist, h->pos);
    UAV_CUDA(cudaGetLastError());
    if (int rc = nm_evaluate(c, n)) return rc;                                                  // the reflected points
    k_nm_decide<<<divup(n, 128), 128, 0, counted(s)>>>(h->states, h->active, n, h->value, h->theta, h->phi, h->from_hist, h->pos);
    UAV_CUDA(cudaGetLastError());
    if (int rc = nm_evaluate(c, n)) return rc;                                                  // expansion / contraction candidates
    k_nm_finish<<<divup(n, 128), 128, 0, counted(s)>>>(h->states, h->active, n, h->value, h->theta, h->phi, h->from_hist);
    UAV_CUDA(cudaGetLastError());
    h->last_active = n;
    return 0;
}

int uavmvs_nm_apply_views(void * stream, const uavmvs_nm * h, float * d_views, uint32_t num_views)
{
    UAV_REQUIRE(h && d_views, "null argument");
    UAV_REQUIRE(num_views == h->last_active, "the view array must hold the views of the last uavmvs_nm_step, in its order");
    if (num_views == 0) return 0;
    k_nm_apply_views<<<divup(num_views, 128), 128, 0, counted((cudaStream_t)stream)>>>(h->states, h->active, num_views, d_views);
    UAV_CUDA(cudaGetLastError());
    return 0;
}

}  // extern "C"

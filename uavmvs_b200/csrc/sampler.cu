// Proxy-cloud sampler on the GPU: apps/generate_proxy_cloud/generate_proxy_cloud.cpp:84-184.
// The reference walks the faces on host threads: std::mt19937 re-seeded with the face's first index slot
// (3 * face, `gen.seed(i)` :107), area * density samples per face (+1 with the probability of the fractional
// part, :118-123), barycentric samples from two uniform draws (:125-137), then a sort by 20-bit Z-order index
// with a lexicographic tie-break (:144-175).  Every face is independent and the final sort makes the output
// independent of the thread schedule, so the GPU version reproduces it bit for bit:
//   k_face_counts   one thread per face: MT19937 state in local memory, first draw decides the count
//   cub scan        offsets
//   k_face_samples  the same generator again, samples + face normals written at the face's offset
//   k_zindex        doubles like the reference (:148-162); four stable radix sorts (z, y, x, Z-order key)
// std::uniform_real_distribution<float>(0, 1) over a 32-bit engine is libstdc++'s generate_canonical<float, 24>:
// float(u32) / 2^32, and a result that rounds to 1 is replaced by the largest float below 1.
// math::geom::triangle_area lives in MVE (not under /root/reference): |(b - a) x (c - a)| / 2.
#include <cub/cub.cuh>

#include "common.cuh"

struct uavmvs_samples {
    uint32_t n;
    float * verts;     // 3 per sample, Z-order sorted
    float * normals;   // 3 per sample
};

namespace uav {

struct MT19937 {
    uint32_t mt[624];
    int idx;
    __device__ void seed(uint32_t s) {
        mt[0] = s;
        for (int i = 1; i < 624; ++i) mt[i] = 1812433253u * (mt[i - 1] ^ (mt[i - 1] >> 30)) + (uint32_t)i;
        idx = 624;
    }
    __device__ void twist() {
        for (int k = 0; k < 624; ++k) {
            uint32_t y = (mt[k] & 0x80000000u) | (mt[(k + 1) % 624] & 0x7fffffffu);
            mt[k] = mt[(k + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
        }
        idx = 0;
    }
    __device__ uint32_t next() {
        if (idx >= 624) twist();
        uint32_t y = mt[idx++];
        y ^= y >> 11; y ^= (y << 7) & 0x9d2c5680u; y ^= (y << 15) & 0xefc60000u; y ^= y >> 18;
        return y;
    }
    __device__ float uniform() {                       // std::uniform_real_distribution<float>(0.0f, 1.0f)
        float r = __uint2float_rn(next()) / 4294967296.0f;
        return r >= 1.0f ? __uint_as_float(0x3f7fffffu) : r;
    }
};

struct FaceGeom { V3 v0, v1, v2, normal; float area; };

__device__ __forceinline__ FaceGeom face_geom(const float * __restrict__ verts, const uint32_t * __restrict__ faces, uint32_t f)
{
    FaceGeom g;
    const float * a = verts + 3 * (size_t)faces[3 * (size_t)f], * b = verts + 3 * (size_t)faces[3 * (size_t)f + 1],
        * c = verts + 3 * (size_t)faces[3 * (size_t)f + 2];
    g.v0 = mk(a[0], a[1], a[2]); g.v1 = mk(b[0], b[1], b[2]); g.v2 = mk(c[0], c[1], c[2]);
    V3 n = cross(g.v1 - g.v0, g.v2 - g.v0);
    float len = norm(n);
    g.normal = n / len;                                  // :111
    g.area = len / 2.0f;
    return g;
}

__device__ __forceinline__ uint32_t face_count(MT19937 & gen, float area, float density)
{
    float face_samples = area * density;                 // :115-123
    uint32_t n = (uint32_t)face_samples;
    if (gen.uniform() < (face_samples - (float)n)) n += 1;
    return n;
}

__global__ void k_face_counts(const float * __restrict__ verts, const uint32_t * __restrict__ faces, uint32_t nf,
    float density, uint32_t * __restrict__ counts)
{
    uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= nf) return;
    MT19937 gen;
    gen.seed(3u * f);
    FaceGeom g = face_geom(verts, faces, f);
    counts[f] = face_count(gen, g.area, density);
}

__global__ void k_face_samples(const float * __restrict__ verts, const uint32_t * __restrict__ faces, uint32_t nf,
    float density, const uint32_t * __restrict__ offsets, float * __restrict__ out_v, float * __restrict__ out_n)
{
    uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= nf) return;
    MT19937 gen;
    gen.seed(3u * f);
    FaceGeom g = face_geom(verts, faces, f);
    const uint32_t n = face_count(gen, g.area, density);
    size_t o = offsets[f];
    for (uint32_t j = 0; j < n; ++j, ++o) {               // :125-137
        float r1 = gen.uniform();
        float r2 = gen.uniform();
        float tmp = sqrtf(r1);
        float u = 1.0f - tmp;
        float v = r2 * tmp;
        float w = 1.0f - v - u;
        V3 p = (g.v0 * u + g.v1 * v) + g.v2 * w;
        out_v[3 * o] = p.x; out_v[3 * o + 1] = p.y; out_v[3 * o + 2] = p.z;
        out_n[3 * o] = g.normal.x; out_n[3 * o + 1] = g.normal.y; out_n[3 * o + 2] = g.normal.z;
    }
}

__device__ __forceinline__ unsigned long long spread3(unsigned long long x, unsigned long long mask)
{
    x &= mask;                                           // libs/acc/math.h:19-45
    x = (x * 0x100000001ull) & 0x003F00000000FFFFull;
    x = (x * 0x10001ull) & 0x003F0000FF0000FFull;
    x = (x * 0x101ull) & 0x300F00F00F00F00Full;
    x = (x * 0x11ull) & 0x30C30C30C30C30C3ull;
    return (x * 0x5ull) & 0x9249249249249249ull;
}

struct ZScale { double scale[3], bias[3]; };

__global__ void k_zindex(const float * __restrict__ v, uint32_t n, ZScale z, unsigned long long * __restrict__ keys,
    float * __restrict__ kx, float * __restrict__ ky, float * __restrict__ kz, uint32_t * __restrict__ ids)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t p[3];
    for (int j = 0; j < 3; ++j) p[j] = (uint32_t)(z.scale[j] * (double)v[3 * (size_t)i + j] + z.bias[j]);   // :156-160
    keys[i] = spread3(p[0], 0x3FFFFFull) | (spread3(p[1], 0x1FFFFFull) << 1) | (spread3(p[2], 0x1FFFFFull) << 2);
    kx[i] = v[3 * (size_t)i]; ky[i] = v[3 * (size_t)i + 1]; kz[i] = v[3 * (size_t)i + 2];
    ids[i] = i;
}

template <typename T>
__global__ void k_gather_keys(const T * __restrict__ src, const uint32_t * __restrict__ ids, uint32_t n, T * __restrict__ dst)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[ids[i]];
}

__global__ void k_gather3(const float * __restrict__ src, const uint32_t * __restrict__ ids, uint32_t n, float * __restrict__ dst)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float * s = src + 3 * (size_t)ids[i];
    dst[3 * (size_t)i] = s[0]; dst[3 * (size_t)i + 1] = s[1]; dst[3 * (size_t)i + 2] = s[2];
}

struct MinMax3 { float mn[3], mx[3]; };
struct MinMax3Op {
    __device__ __forceinline__ MinMax3 operator()(const MinMax3 & a, const MinMax3 & b) const {
        MinMax3 r;
        for (int k = 0; k < 3; ++k) { r.mn[k] = fminf(a.mn[k], b.mn[k]); r.mx[k] = fmaxf(a.mx[k], b.mx[k]); }
        return r;
    }
};
__global__ void k_as_minmax(const float * __restrict__ v, uint32_t n, MinMax3 * __restrict__ out)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    MinMax3 m;
    for (int k = 0; k < 3; ++k) m.mn[k] = m.mx[k] = v[3 * (size_t)i + k];
    out[i] = m;
}

}  // namespace uav

using namespace uav;

extern "C" {

int uavmvs_proxy_cloud_sample(const float * verts, uint32_t num_verts, const uint32_t * faces, uint32_t num_faces,
    float samples_per_unit_area, uavmvs_samples ** out, uint32_t * num_samples)
{
    UAV_REQUIRE(verts && faces && out && num_samples, "null argument");
    UAV_REQUIRE(num_faces >= 1 && num_verts >= 1, "empty mesh");
    UAV_REQUIRE(num_faces < (1u << 30), "too many faces");          // the seed 3 * face must not wrap
    float * d_verts; uint32_t * d_faces, * counts, * offsets;
    UAV_CUDA(cudaMalloc(&d_verts, sizeof(float) * 3 * (size_t)num_verts));
    UAV_CUDA(cudaMalloc(&d_faces, sizeof(uint32_t) * 3 * (size_t)num_faces));
    UAV_CUDA(cudaMalloc(&counts, 4 * (size_t)num_faces)); UAV_CUDA(cudaMalloc(&offsets, 4 * (size_t)num_faces));
    UAV_CUDA(cudaMemcpy(d_verts, verts, sizeof(float) * 3 * (size_t)num_verts, cudaMemcpyHostToDevice));
    UAV_CUDA(cudaMemcpy(d_faces, faces, sizeof(uint32_t) * 3 * (size_t)num_faces, cudaMemcpyHostToDevice));
    k_face_counts<<<divup(num_faces, 64), 64>>>(d_verts, d_faces, num_faces, samples_per_unit_area, counts);
    size_t scan_bytes = 0; void * tmp = nullptr;
    cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, counts, offsets, (int)num_faces);
    UAV_CUDA(cudaMalloc(&tmp, scan_bytes));
    cub::DeviceScan::ExclusiveSum(tmp, scan_bytes, counts, offsets, (int)num_faces);
    uint32_t last[2];
    UAV_CUDA(cudaMemcpy(&last[0], counts + (num_faces - 1), 4, cudaMemcpyDeviceToHost));
    UAV_CUDA(cudaMemcpy(&last[1], offsets + (num_faces - 1), 4, cudaMemcpyDeviceToHost));
    cudaFree(tmp); tmp = nullptr;
    const uint64_t total = (uint64_t)last[0] + last[1];
    UAV_REQUIRE(total < (1ull << 31), "too many samples");
    const uint32_t n = (uint32_t)total;
    uavmvs_samples * s = new uavmvs_samples();
    s->n = n; s->verts = nullptr; s->normals = nullptr;
    if (n) {
        float * raw_v, * raw_n, * kx, * ky, * kz, * kf; uint32_t * ids, * ids2; unsigned long long * keys, * keys2, * kg;
        MinMax3 * mm, * mm_out;
        UAV_CUDA(cudaMalloc(&raw_v, 12 * (size_t)n)); UAV_CUDA(cudaMalloc(&raw_n, 12 * (size_t)n));
        UAV_CUDA(cudaMalloc(&s->verts, 12 * (size_t)n)); UAV_CUDA(cudaMalloc(&s->normals, 12 * (size_t)n));
        k_face_samples<<<divup(num_faces, 64), 64>>>(d_verts, d_faces, num_faces, samples_per_unit_area, offsets, raw_v, raw_n);
        // bounding box of the samples (acc::calculate_aabb, :147)
        UAV_CUDA(cudaMalloc(&mm, sizeof(MinMax3) * (size_t)n)); UAV_CUDA(cudaMalloc(&mm_out, sizeof(MinMax3)));
        k_as_minmax<<<divup(n, 256), 256>>>(raw_v, n, mm);
        MinMax3 init; for (int k = 0; k < 3; ++k) { init.mn[k] = 3.402823466e+38f; init.mx[k] = -3.402823466e+38f; }
        size_t red_bytes = 0;
        cub::DeviceReduce::Reduce(nullptr, red_bytes, mm, mm_out, (int)n, MinMax3Op(), init);
        UAV_CUDA(cudaMalloc(&tmp, red_bytes));
        cub::DeviceReduce::Reduce(tmp, red_bytes, mm, mm_out, (int)n, MinMax3Op(), init);
        MinMax3 box;
        UAV_CUDA(cudaMemcpy(&box, mm_out, sizeof(box), cudaMemcpyDeviceToHost));
        cudaFree(tmp); tmp = nullptr; cudaFree(mm); cudaFree(mm_out);
        ZScale z;
        for (int i = 0; i < 3; ++i) {                                                   // :149-153
            double div = (box.mx[i] - box.mn[i]);            // float subtraction, then widened (as the reference)
            z.scale[i] = 1.0 / div * (double(1 << 20) - 1.0);
            z.bias[i] = - box.mn[i] / div * (double(1 << 20) - 1.0);
        }
        UAV_CUDA(cudaMalloc(&kx, 4 * (size_t)n)); UAV_CUDA(cudaMalloc(&ky, 4 * (size_t)n)); UAV_CUDA(cudaMalloc(&kz, 4 * (size_t)n));
        UAV_CUDA(cudaMalloc(&kf, 4 * (size_t)n)); UAV_CUDA(cudaMalloc(&kg, 8 * (size_t)n));
        UAV_CUDA(cudaMalloc(&ids, 4 * (size_t)n)); UAV_CUDA(cudaMalloc(&ids2, 4 * (size_t)n));
        UAV_CUDA(cudaMalloc(&keys, 8 * (size_t)n)); UAV_CUDA(cudaMalloc(&keys2, 8 * (size_t)n));
        k_zindex<<<divup(n, 256), 256>>>(raw_v, n, z, keys, kx, ky, kz, ids);
        size_t sort_f = 0, sort_k = 0;
        cub::DeviceRadixSort::SortPairs(nullptr, sort_f, kf, kx, ids, ids2, (int)n);
        cub::DeviceRadixSort::SortPairs(nullptr, sort_k, kg, keys2, ids, ids2, (int)n);
        UAV_CUDA(cudaMalloc(&tmp, sort_f > sort_k ? sort_f : sort_k));
        size_t tb = sort_f > sort_k ? sort_f : sort_k;
        // LSD: least significant criterion first (z, then y, then x of the position, then the Z-order key), all stable
        const float * crit[3] = {kz, ky, kx};
        float * scratch = nullptr;
        UAV_CUDA(cudaMalloc(&scratch, 4 * (size_t)n));
        for (int c = 0; c < 3; ++c) {
            k_gather_keys<float><<<divup(n, 256), 256>>>(crit[c], ids, n, kf);
            cub::DeviceRadixSort::SortPairs(tmp, tb, kf, scratch, ids, ids2, (int)n);
            uint32_t * t = ids; ids = ids2; ids2 = t;
        }
        k_gather_keys<unsigned long long><<<divup(n, 256), 256>>>(keys, ids, n, kg);
        cub::DeviceRadixSort::SortPairs(tmp, tb, kg, keys2, ids, ids2, (int)n);
        k_gather3<<<divup(n, 256), 256>>>(raw_v, ids2, n, s->verts);
        k_gather3<<<divup(n, 256), 256>>>(raw_n, ids2, n, s->normals);
        UAV_CUDA(cudaGetLastError());
        UAV_CUDA(cudaDeviceSynchronize());
        cudaFree(tmp); cudaFree(scratch); cudaFree(raw_v); cudaFree(raw_n); cudaFree(kx); cudaFree(ky); cudaFree(kz); cudaFree(kf);
        cudaFree(kg); cudaFree(ids); cudaFree(ids2); cudaFree(keys); cudaFree(keys2);
    }
    UAV_CUDA(cudaDeviceSynchronize());
    cudaFree(d_verts); cudaFree(d_faces); cudaFree(counts); cudaFree(offsets);
    *out = s; *num_samples = n;
    return 0;
}

int uavmvs_proxy_cloud_download(const uavmvs_samples * s, float * verts, float * normals)
{
    UAV_REQUIRE(s, "null samples");
    if (s->n && verts) UAV_CUDA(cudaMemcpy(verts, s->verts, 12 * (size_t)s->n, cudaMemcpyDeviceToHost));
    if (s->n && normals) UAV_CUDA(cudaMemcpy(normals, s->normals, 12 * (size_t)s->n, cudaMemcpyDeviceToHost));
    return 0;
}

int uavmvs_proxy_cloud_destroy(uavmvs_samples * s)
{
    if (!s) return 0;
    cudaFree(s->verts); cudaFree(s->normals);
    delete s;
    return 0;
}

}  // extern "C"

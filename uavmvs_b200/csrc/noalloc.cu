// K9 (sum / min / max of an array) for callers that cannot allocate: code under CUDA graph capture, tight serving loops,
// harnesses that forbid cudaMalloc.  No counterpart in the reference.
//
// uavmvs_reduce (eval.cu) accumulates into a cudaMalloc'd record and reads it back through a cudaMallocHost'd one.  This
// variant needs neither: its scratch -- one partial per 2048-element chunk -- is the file-scope device array below, and the
// result goes to a device array the CALLER created.  The entry point calls no cudaMalloc / cudaFree / cudaMallocHost, no
// allocating library routine, and does not synchronise.
//
//   k_reduce_chunks   CTA c reduces elements [2048 c, 2048 (c + 1)) and writes partial[c]
//   k_reduce_fold     one CTA adds the partials in a fixed order (thread t: c = t, t + 256, ...; then a fixed tree)
//
// Every partial that the fold reads is written by the same call, so nothing has to start at zero.  The additions have a
// fixed order (no atomics): two calls on the same array give the same bits.
//
// LIMITS.  n <= UAVMVS_REDUCE_NOALLOC_MAX_N (2^24 elements = 8192 partials x 16 B = 128 KB of device memory, held by
// every GPU that loads the library until the process exits).  All calls on a device share the array: calls that overlap
// on the GPU (different streams, different host threads) RACE; calls on one stream are ordered by the stream.
#include "common.cuh"

#include <mutex>

namespace uav {

constexpr uint32_t NA_CHUNK = 2048;                                   // elements per CTA = 256 threads x 8
constexpr uint32_t NA_MAX_N = 1u << 24;                               // = UAVMVS_REDUCE_NOALLOC_MAX_N of the header
constexpr uint32_t NA_MAX_PARTIALS = NA_MAX_N / NA_CHUNK;             // 8192
struct NaPartial { double sum; float mn, mx; };                       // 16 B

// In HOST code this name is a host-side stand-in, not a device address: kernels get the pointer that
// cudaGetSymbolAddress returns (na_scratch() below), never `g_na_scratch` itself.
static __device__ __align__(256) NaPartial g_na_scratch[NA_MAX_PARTIALS];

__global__ void __launch_bounds__(256)
k_reduce_chunks(const float * __restrict__ x, uint32_t n, NaPartial * __restrict__ partial)
{
    const uint32_t base = blockIdx.x * NA_CHUNK;
    float s = 0.0f, mn = UAV_INF, mx = -UAV_INF;
    #pragma unroll
    for (uint32_t j = 0; j < NA_CHUNK / 256; ++j) {
        const uint32_t i = base + j * 256 + threadIdx.x;
        if (i < n) { const float v = x[i]; s += v; mn = fminf(mn, v); mx = fmaxf(mx, v); }
    }
    for (int o = 16; o > 0; o >>= 1) {
        s += __shfl_xor_sync(0xffffffffu, s, o);
        mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    __shared__ float ss[8], smn[8], smx[8];
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) { ss[w] = s; smn[w] = mn; smx[w] = mx; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int i = 0; i < 8; ++i) { t += (double)ss[i]; mn = fminf(mn, smn[i]); mx = fmaxf(mx, smx[i]); }
        NaPartial p; p.sum = t; p.mn = mn; p.mx = mx;
        partial[blockIdx.x] = p;
    }
}

__global__ void __launch_bounds__(256)
k_reduce_fold(const NaPartial * __restrict__ partial, uint32_t num_partials, float * __restrict__ out3)
{
    double s = 0.0; float mn = UAV_INF, mx = -UAV_INF;
    for (uint32_t c = threadIdx.x; c < num_partials; c += 256) {
        const NaPartial p = partial[c];
        s += p.sum; mn = fminf(mn, p.mn); mx = fmaxf(mx, p.mx);
    }
    __shared__ double ss[256]; __shared__ float smn[256], smx[256];
    ss[threadIdx.x] = s; smn[threadIdx.x] = mn; smx[threadIdx.x] = mx;
    __syncthreads();
    for (uint32_t h = 128; h > 0; h >>= 1) {
        if (threadIdx.x < h) {
            ss[threadIdx.x] += ss[threadIdx.x + h];
            smn[threadIdx.x] = fminf(smn[threadIdx.x], smn[threadIdx.x + h]);
            smx[threadIdx.x] = fmaxf(smx[threadIdx.x], smx[threadIdx.x + h]);
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) { out3[0] = (float)ss[0]; out3[1] = smn[0]; out3[2] = smx[0]; }
}

// Device address of g_na_scratch on the CURRENT device, looked up once per device (set-up, not an allocation: the array
// is part of the module image).  A table indexed by device ordinal; the mutex makes the first calls of two threads safe.
constexpr int NA_MAX_DEVICES = 64;
static NaPartial * g_na_ptr[NA_MAX_DEVICES] = {};
static std::mutex g_na_mutex;

static int na_scratch(NaPartial ** out)
{
    int dev = -1;
    UAV_CUDA(cudaGetDevice(&dev));
    UAV_REQUIRE(dev >= 0 && dev < NA_MAX_DEVICES, "device ordinal beyond the no-allocation entry point's table (64 devices)");
    std::lock_guard<std::mutex> lock(g_na_mutex);
    if (!g_na_ptr[dev]) {
        void * p = nullptr;
        UAV_CUDA(cudaGetSymbolAddress(&p, g_na_scratch));              // with `dev` current: this device's copy of the array
        UAV_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned a null address for the scratch array");
        g_na_ptr[dev] = static_cast<NaPartial *>(p);
    }
    *out = g_na_ptr[dev];
    return 0;
}

}  // namespace uav

using namespace uav;

extern "C" int uavmvs_reduce_noalloc(void * stream, const uavmvs_array * a, uavmvs_array * out3)
{
    UAV_REQUIRE(a && out3, "null argument");
    UAV_REQUIRE(a->data && out3->data, "array without device storage");
    UAV_REQUIRE(out3->n >= 3, "out3 needs room for {sum, min, max} (3 floats)");
    UAV_REQUIRE(a->n >= 1, "empty input array");
    UAV_REQUIRE(a->n <= NA_MAX_N,
        "input larger than UAVMVS_REDUCE_NOALLOC_MAX_N (2^24 elements): the static scratch array holds 8192 partials; "
        "use uavmvs_reduce, or reduce in pieces");
    NaPartial * scratch = nullptr;
    if (int rc = na_scratch(&scratch)) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    const uint32_t chunks = divup(a->n, NA_CHUNK);
    k_reduce_chunks<<<chunks, 256, 0, uav::counted(s)>>>(a->data, a->n, scratch);
    k_reduce_fold<<<1, 256, 0, uav::counted(s)>>>(scratch, chunks, out3->data);
    UAV_CUDA(cudaGetLastError());
    return 0;
}

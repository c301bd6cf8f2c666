// Multi-GPU plumbing of the point-sharded evaluation: a thin communicator over NCCL.
//
// The reference is single-GPU; what it computes over all points on one device -- cacc::reduction::sum/min/max
// (libs/cacc/reduction.h:132-162) feeding avg_wrecon = sum / N (apps/optimize_trajectory/optimize_trajectory.cu:529)
// -- becomes one small collective over the shards here (eval.cu: uavmvs_evaluate_trajectory_sharded).
//
// libnccl.so.2 is resolved at run time (dlopen): the library keeps loading on hosts without NCCL, and inside a
// PyTorch process the copy PyTorch already loaded is reused (RTLD_NOLOAD first), so one process never holds two NCCLs.
// The few NCCL declarations needed are restated below from its public header (stable C ABI since NCCL 2.0).
#include <dlfcn.h>

#include "common.cuh"

namespace uav {

std::atomic<unsigned long long> g_launches{0};

namespace nccl {
typedef int result_t;                                // ncclResult_t, 0 = ncclSuccess
struct unique_id { char internal[128]; };            // ncclUniqueId (NCCL_UNIQUE_ID_BYTES)
enum { kSum = 0 };                                   // ncclRedOp_t
enum { kUint64 = 5, kFloat32 = 7, kFloat64 = 8 };    // ncclDataType_t

struct Api {
    result_t (*GetUniqueId)(unique_id *) = nullptr;
    result_t (*CommInitRank)(void **, int, unique_id, int) = nullptr;
    result_t (*CommDestroy)(void *) = nullptr;
    const char * (*GetErrorString)(result_t) = nullptr;
    result_t (*AllReduce)(const void *, void *, size_t, int, int, void *, cudaStream_t) = nullptr;
    result_t (*AllGather)(const void *, void *, size_t, int, void *, cudaStream_t) = nullptr;
    void * handle = nullptr;
    bool ok = false;
    std::string why;
};

static Api & api()
{
    static Api a;
    static std::once_flag once;
    std::call_once(once, []() {
        const char * override_path = getenv("UAVMVS_NCCL_LIB");
        const char * names[] = {override_path, "libnccl.so.2", "libnccl.so"};
        for (int pass = 0; pass < 2 && !a.handle; ++pass)           // pass 0: a copy already in the process
            for (const char * n : names) {
                if (!n) continue;
                a.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL | (pass == 0 ? RTLD_NOLOAD : 0));
                if (a.handle) break;
            }
        if (!a.handle) { a.why = "libnccl.so.2 not found (set UAVMVS_NCCL_LIB or LD_LIBRARY_PATH)"; return; }
        bool all = true;
        auto sym = [&](const char * name) { void * p = dlsym(a.handle, name); if (!p) { all = false; a.why = std::string("missing symbol ") + name; } return p; };
        a.GetUniqueId = (decltype(a.GetUniqueId))sym("ncclGetUniqueId");
        a.CommInitRank = (decltype(a.CommInitRank))sym("ncclCommInitRank");
        a.CommDestroy = (decltype(a.CommDestroy))sym("ncclCommDestroy");
        a.GetErrorString = (decltype(a.GetErrorString))sym("ncclGetErrorString");
        a.AllReduce = (decltype(a.AllReduce))sym("ncclAllReduce");
        a.AllGather = (decltype(a.AllGather))sym("ncclAllGather");
        a.ok = all;
    });
    return a;
}
}  // namespace nccl

#define UAV_NCCL(call)                                                                   \
    do {                                                                                 \
        uav::nccl::result_t r_ = (call);                                                 \
        if (r_ != 0) { uav::set_error(__FILE__, __LINE__, #call, uav::nccl::api().GetErrorString(r_)); return 3; } \
    } while (0)

static int need_nccl()
{
    nccl::Api & a = nccl::api();
    if (!a.ok) { set_error(__FILE__, __LINE__, "dlopen(libnccl.so.2)", a.why.c_str()); return 3; }
    return 0;
}

int comm_allgather(cudaStream_t s, uavmvs_comm * c, const void * send, void * recv, size_t count, int f64)
{
    if (need_nccl()) return 3;
    UAV_NCCL(nccl::api().AllGather(send, recv, count, f64 ? nccl::kFloat64 : nccl::kFloat32, c->nccl, s));
    return 0;
}

int comm_allreduce_sum_f32(cudaStream_t s, uavmvs_comm * c, void * buf, size_t count)
{
    if (need_nccl()) return 3;
    UAV_NCCL(nccl::api().AllReduce(buf, buf, count, nccl::kFloat32, nccl::kSum, c->nccl, s));
    return 0;
}

}  // namespace uav

using namespace uav;

extern "C" {

int uavmvs_comm_unique_id(uint8_t id[UAVMVS_COMM_ID_BYTES])
{
    UAV_REQUIRE(id, "null argument");
    if (need_nccl()) return 3;
    nccl::unique_id u;
    UAV_NCCL(nccl::api().GetUniqueId(&u));
    memcpy(id, u.internal, UAVMVS_COMM_ID_BYTES);
    return 0;
}

int uavmvs_comm_create(const uint8_t id[UAVMVS_COMM_ID_BYTES], int world, int rank, uavmvs_comm ** out)
{
    UAV_REQUIRE(id && out, "null argument");
    UAV_REQUIRE(world >= 1 && rank >= 0 && rank < world, "rank outside [0, world)");
    if (need_nccl()) return 3;
    nccl::unique_id u;
    memcpy(u.internal, id, UAVMVS_COMM_ID_BYTES);
    void * nc = nullptr;
    UAV_NCCL(nccl::api().CommInitRank(&nc, world, u, rank));
    uavmvs_comm * c = new uavmvs_comm();
    c->nccl = nc; c->world = world; c->rank = rank; c->owned = 1;
    *out = c;
    return 0;
}

int uavmvs_comm_adopt(void * nccl_comm, int world, int rank, uavmvs_comm ** out)
{
    UAV_REQUIRE(nccl_comm && out, "null argument");
    UAV_REQUIRE(world >= 1 && rank >= 0 && rank < world, "rank outside [0, world)");
    if (need_nccl()) return 3;
    uavmvs_comm * c = new uavmvs_comm();
    c->nccl = nccl_comm; c->world = world; c->rank = rank; c->owned = 0;
    *out = c;
    return 0;
}

int uavmvs_comm_info(const uavmvs_comm * comm, int * world, int * rank)
{
    UAV_REQUIRE(comm, "null comm");
    if (world) *world = comm->world;
    if (rank) *rank = comm->rank;
    return 0;
}

int uavmvs_comm_destroy(uavmvs_comm * comm)
{
    if (!comm) return 0;
    if (comm->gather) cudaFree(comm->gather);
    if (comm->owned && comm->nccl && nccl::api().ok) nccl::api().CommDestroy(comm->nccl);
    delete comm;
    return 0;
}

int uavmvs_allreduce_array(void * stream, uavmvs_comm * comm, uavmvs_array * a, uint64_t count)
{
    UAV_REQUIRE(a && count <= a->n, "count exceeds the array");
    if (!comm || count == 0) return 0;                       // single GPU: the partial histogram is the histogram
    return comm_allreduce_sum_f32((cudaStream_t)stream, comm, a->data, (size_t)count);
}

int uavmvs_allgather_array(void * stream, uavmvs_comm * comm, const uavmvs_array * local, uint64_t count, uavmvs_array * all)
{
    UAV_REQUIRE(local && all && count <= local->n, "count exceeds the local array");
    const int world = comm ? comm->world : 1;
    UAV_REQUIRE((uint64_t)world * count <= all->n, "gathered array too small");
    if (!comm) {
        UAV_CUDA(cudaMemcpyAsync(all->data, local->data, 4 * (size_t)count, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
        return 0;
    }
    return comm_allgather((cudaStream_t)stream, comm, local->data, all->data, (size_t)count, 0);
}

int uavmvs_launch_count(uint64_t * out, int reset)
{
    UAV_REQUIRE(out, "null argument");
    *out = reset ? g_launches.exchange(0) : g_launches.load();
    return 0;
}

}  // extern "C"

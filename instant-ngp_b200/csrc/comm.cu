#include "comm.h"
#include <dlfcn.h>
#include <cstring>
#include <mutex>

namespace ngp {

namespace {
struct NcclId { char internal[NCCL_ID_BYTES]; };
struct Api {
    int (*GetUniqueId)(NcclId*) = nullptr;
    int (*CommInitRank)(void**, int, NcclId, int) = nullptr;
    int (*CommDestroy)(void*) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*ReduceScatter)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, void*, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};
Api& api(const char* lib_path) {
    static Api a; static std::once_flag once; static std::string err;
    std::call_once(once, [&] {
        void* h = nullptr;
        const char* env = getenv("NGP_NCCL_LIB");
        if (lib_path && *lib_path) h = dlopen(lib_path, RTLD_NOW | RTLD_GLOBAL);
        if (!h && env && *env) h = dlopen(env, RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);      // the copy already in the process, else the system one
        if (!h) { err = std::string("cannot load libnccl.so.2: ") + dlerror(); return; }
        a.GetUniqueId = reinterpret_cast<decltype(a.GetUniqueId)>(dlsym(h, "ncclGetUniqueId"));
        a.CommInitRank = reinterpret_cast<decltype(a.CommInitRank)>(dlsym(h, "ncclCommInitRank"));
        a.CommDestroy = reinterpret_cast<decltype(a.CommDestroy)>(dlsym(h, "ncclCommDestroy"));
        a.AllReduce = reinterpret_cast<decltype(a.AllReduce)>(dlsym(h, "ncclAllReduce"));
        a.GetErrorString = reinterpret_cast<decltype(a.GetErrorString)>(dlsym(h, "ncclGetErrorString"));
        a.ReduceScatter = reinterpret_cast<decltype(a.ReduceScatter)>(dlsym(h, "ncclReduceScatter"));
        a.AllGather = reinterpret_cast<decltype(a.AllGather)>(dlsym(h, "ncclAllGather"));
        if (!a.ReduceScatter || !a.AllGather) err = "libnccl.so.2 lacks ncclReduceScatter / ncclAllGather";
        if (!a.GetUniqueId || !a.CommInitRank || !a.CommDestroy || !a.AllReduce || !a.GetErrorString) err = "libnccl.so.2 lacks a required symbol";
    });
    if (!err.empty()) throw Error(err);
    return a;
}
void check(Api& a, int rc, const char* what) {
    if (rc != 0) throw Error(std::string(what) + ": " + a.GetErrorString(rc));
}
}  // namespace

void Comm::unique_id(void* out128, const char* lib_path) {
    Api& a = api(lib_path);
    NcclId id; check(a, a.GetUniqueId(&id), "ncclGetUniqueId");
    memcpy(out128, id.internal, NCCL_ID_BYTES);
}

Comm::Comm(const void* id128, int rank, int world, const char* lib_path) : m_rank(rank), m_world(world) {
    Api& a = api(lib_path);
    NcclId id; memcpy(id.internal, id128, NCCL_ID_BYTES);
    check(a, a.CommInitRank(&m_comm, world, id, rank), "ncclCommInitRank");
}

Comm::~Comm() {
    if (m_comm) { try { api(nullptr).CommDestroy(m_comm); } catch (...) {} }
}

void Comm::all_reduce_sum(void* buf, size_t count, CommDtype dtype, cudaStream_t s) {
    Api& a = api(nullptr);
    check(a, a.AllReduce(buf, buf, count, (int)dtype, /*ncclSum*/ 0, m_comm, s), "ncclAllReduce");
}

void Comm::reduce_scatter_sum(void* buf, size_t count_per_rank, CommDtype dtype, cudaStream_t s) {
    Api& a = api(nullptr);
    const size_t esz = dtype == COMM_F16 ? 2 : 4;
    check(a, a.ReduceScatter(buf, static_cast<char*>(buf) + (size_t)m_rank * count_per_rank * esz, count_per_rank, (int)dtype, /*ncclSum*/ 0, m_comm, s), "ncclReduceScatter");
}

void Comm::all_gather(void* buf, size_t count_per_rank, CommDtype dtype, cudaStream_t s) {
    Api& a = api(nullptr);
    const size_t esz = dtype == COMM_F16 ? 2 : 4;
    check(a, a.AllGather(static_cast<const char*>(buf) + (size_t)m_rank * count_per_rank * esz, buf, count_per_rank, (int)dtype, m_comm, s), "ncclAllGather");
}

}  // namespace ngp

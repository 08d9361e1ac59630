// NCCL over NVLink for the data-parallel training step: the gradient all-reduce is issued from inside the step, per
// hash-table level chunk, on a communication stream that runs beside the remaining backward kernels.
// NCCL is bound at run time (dlopen of the libnccl.so.2 the host process already carries — torch's — or NGP_NCCL_LIB), so the
// library neither links nor includes NCCL; the handful of prototypes used are restated here from nccl.h (2.27/2.28 ABI).
#pragma once
#include "ngp_common.cuh"
#include <string>

namespace ngp {

constexpr int NCCL_ID_BYTES = 128;
enum CommDtype : int { COMM_U32 = 3, COMM_F16 = 6 };   // ncclUint32, ncclFloat16

class Comm {
public:
    static void unique_id(void* out128, const char* lib_path);          // ncclGetUniqueId (rank 0; broadcast by the caller)
    Comm(const void* id128, int rank, int world, const char* lib_path); // ncclCommInitRank: collective over all ranks
    ~Comm();
    void all_reduce_sum(void* buf, size_t count, CommDtype dtype, cudaStream_t s);   // in place
    // in place over a buffer of world * count_per_rank elements: rank r ends up with the sum of everybody's slice r in ITS slice r
    void reduce_scatter_sum(void* buf, size_t count_per_rank, CommDtype dtype, cudaStream_t s);
    // in place: every rank contributes its slice r of the buffer, all ranks end up with all slices
    void all_gather(void* buf, size_t count_per_rank, CommDtype dtype, cudaStream_t s);
    int rank() const { return m_rank; }
    int world() const { return m_world; }
private:
    void* m_comm = nullptr; int m_rank = 0, m_world = 1;
};

}  // namespace ngp

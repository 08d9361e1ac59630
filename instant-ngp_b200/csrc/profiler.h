// Opt-in per-kernel timing with CUDA events on the launching stream (bench.py's roofline leg).  Disabled by default:
// NGP_PROFILE_SCOPE costs one relaxed atomic load when off.
#pragma once
#include <cuda_runtime.h>
#include <string>
namespace ngp {
struct ProfileScope {
    ProfileScope(cudaStream_t s, const char* name);
    ~ProfileScope();
    cudaStream_t s; int slot;
};
void profile_enable(bool on);
std::string profile_report_json(bool reset);   // {"name": {"count": n, "ms": total}, ...}; synchronises the device
}
#define NGP_PROFILE_SCOPE(stream, name) ngp::ProfileScope _ngp_prof_scope((stream), (name))

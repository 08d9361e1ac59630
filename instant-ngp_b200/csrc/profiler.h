// Opt-in per-kernel timing with CUDA events on the launching stream (bench.py's roofline leg).  Disabled by default:
// NGP_PROFILE_SCOPE costs one relaxed atomic load when off.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <string>
namespace ngp {
struct ProfileScope {
    ProfileScope(cudaStream_t s, const char* name, uint64_t units = 0);   // units: work items of this launch (points, samples, params)
    ~ProfileScope();
    cudaStream_t s; int slot;
};
void profile_enable(bool on);
std::string profile_report_json(bool reset);
std::string profile_timeline_json();            // [["name", start_us, duration_us], ...] of the records since the last reset   // {"name": {"count": n, "ms": total, "units": total}, ...}; synchronises the device
}
#define NGP_PROFILE_SCOPE(stream, ...) ngp::ProfileScope _ngp_prof_scope((stream), __VA_ARGS__)

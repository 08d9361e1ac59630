#include "profiler.h"
#include <atomic>
#include <map>
#include <mutex>
#include <vector>
namespace ngp {
namespace {
std::atomic<bool> g_on{false};
std::mutex g_mu;
struct Rec { int name_id; cudaEvent_t a, b; uint64_t units; };
std::vector<Rec> g_recs;
std::vector<std::string> g_names;
std::vector<cudaEvent_t> g_pool;
int name_id(const char* n) { for (size_t i = 0; i < g_names.size(); ++i) if (g_names[i] == n) return (int)i; g_names.push_back(n); return (int)g_names.size() - 1; }
cudaEvent_t get_event() { if (!g_pool.empty()) { cudaEvent_t e = g_pool.back(); g_pool.pop_back(); return e; } cudaEvent_t e; cudaEventCreate(&e); return e; }
}
void profile_enable(bool on) { g_on.store(on); }
ProfileScope::ProfileScope(cudaStream_t s_, const char* name, uint64_t units) : s(s_), slot(-1) {
    if (!g_on.load(std::memory_order_relaxed)) return;
    std::lock_guard<std::mutex> l(g_mu);
    Rec r{name_id(name), get_event(), get_event(), units};
    cudaEventRecord(r.a, s);
    g_recs.push_back(r); slot = (int)g_recs.size() - 1;
}
ProfileScope::~ProfileScope() {
    if (slot < 0) return;
    std::lock_guard<std::mutex> l(g_mu);
    if (slot < (int)g_recs.size()) cudaEventRecord(g_recs[slot].b, s);
}
std::string profile_report_json(bool reset) {
    cudaDeviceSynchronize();
    std::lock_guard<std::mutex> l(g_mu);
    struct Acc { int count = 0; double ms = 0; uint64_t units = 0; };
    std::map<int, Acc> acc;
    for (auto& r : g_recs) { float ms = 0.f; if (cudaEventElapsedTime(&ms, r.a, r.b) == cudaSuccess) { Acc& a = acc[r.name_id]; a.count++; a.ms += ms; a.units += r.units; } }
    std::string out = "{";
    bool first = true;
    for (auto& kv : acc) {
        if (!first) out += ", "; first = false;
        out += "\"" + g_names[kv.first] + "\": {\"count\": " + std::to_string(kv.second.count) + ", \"ms\": " + std::to_string(kv.second.ms) + ", \"units\": " + std::to_string(kv.second.units) + "}";
    }
    out += "}";
    if (reset) { for (auto& r : g_recs) { g_pool.push_back(r.a); g_pool.push_back(r.b); } g_recs.clear(); }
    return out;
}
// absolute timeline of the recorded scopes (debug / tools): [["name", start_us, duration_us], ...] relative to the first record;
// events of different streams are comparable, so the overlap of the ray prefetch with the training stream shows
std::string profile_timeline_json() {
    cudaDeviceSynchronize();
    std::lock_guard<std::mutex> l(g_mu);
    std::string out = "[";
    bool first = true;
    for (auto& r : g_recs) {
        float t0 = 0.f, d = 0.f;
        if (cudaEventElapsedTime(&t0, g_recs[0].a, r.a) != cudaSuccess || cudaEventElapsedTime(&d, r.a, r.b) != cudaSuccess) { cudaGetLastError(); continue; }
        if (!first) out += ", "; first = false;
        out += "[\"" + g_names[r.name_id] + "\", " + std::to_string(t0 * 1e3) + ", " + std::to_string(d * 1e3) + "]";
    }
    return out + "]";
}
}  // namespace ngp

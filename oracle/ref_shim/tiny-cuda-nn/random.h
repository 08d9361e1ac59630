// ORACLE BUILD STUB (test infrastructure).  tiny-cuda-nn is an empty submodule in the reference mount; the reference's
// random_val.cuh only needs `tcnn::default_rng_t` with next_uint / next_float / advance: pcg32 (M. O'Neill, pcg-random.org;
// the generator tiny-cuda-nn vendors), restated here.
#pragma once
#include <cstdint>
#include <cstring>
namespace tcnn {
struct pcg32 {
    uint64_t state, inc;
    pcg32(uint64_t initstate = 0x853c49e6748fea9bULL, uint64_t initseq = 1u) { seed(initstate, initseq); }
    void seed(uint64_t initstate, uint64_t initseq = 1) { state = 0u; inc = (initseq << 1u) | 1u; next_uint(); state += initstate; next_uint(); }
    uint32_t next_uint() { uint64_t old = state; state = old * 0x5851f42d4c957f2dULL + inc; uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u), rot = (uint32_t)(old >> 59u); return (xs >> rot) | (xs << ((~rot + 1u) & 31)); }
    float next_float() { uint32_t u = (next_uint() >> 9) | 0x3f800000u; float f; memcpy(&f, &u, 4); return f - 1.0f; }
    void advance(int64_t delta_ = (1ll << 32)) { uint64_t cm = 0x5851f42d4c957f2dULL, cp = inc, am = 1u, ap = 0u, d = (uint64_t)delta_; while (d > 0) { if (d & 1) { am *= cm; ap = ap * cm + cp; } cp = (cm + 1) * cp; cm *= cm; d /= 2; } state = am * state + ap; }
};
using default_rng_t = pcg32;
}

// Shared device/host helpers for the sm_100a kernels: error handling, pcg32, morton codes, fp16 packing,
// small vector math.  All float math that feeds an integer decision (cell indices, occupancy traversal)
// is written without implicit FMA contraction: the library is compiled with --fmad=false and uses
// explicit fmaf() where a fused multiply-add is intended, so the CPU oracle can match bit for bit.
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdexcept>
#include <string>
#include <atomic>
#include "profiler.h"

#define NGP_HD __host__ __device__ __forceinline__
#define NGP_D __device__ __forceinline__

namespace ngp {

struct Error : std::runtime_error { using std::runtime_error::runtime_error; };

#define NGP_CUDA_CHECK(x)                                                                                   \
    do {                                                                                                     \
        cudaError_t _e = (x);                                                                                \
        if (_e != cudaSuccess)                                                                               \
            throw ngp::Error(std::string(__FILE__) + ":" + std::to_string(__LINE__) + " " #x " failed: " +  \
                             cudaGetErrorString(_e));                                                        \
    } while (0)

#define NGP_CHECK(cond, msg)                                                                  \
    do {                                                                                       \
        if (!(cond)) throw ngp::Error(std::string(__FILE__) + ":" + std::to_string(__LINE__) + " " + (msg)); \
    } while (0)

// every kernel launch of this library goes through NGP_POST_LAUNCH(): error check + launch counter (ngp_kernel_launch_count)
extern std::atomic<uint64_t> g_launch_count;
#define NGP_POST_LAUNCH()                        \
    do {                                         \
        NGP_CUDA_CHECK(cudaGetLastError());      \
        ngp::g_launch_count.fetch_add(1, std::memory_order_relaxed); \
    } while (0)

constexpr uint32_t BATCH_GRANULARITY = 256;   // tcnn::batch_size_granularity as used by the reference callers
constexpr uint32_t TILE_ROWS = 128;           // samples per tensor-core tile (UMMA M)

inline uint32_t div_round_up(uint32_t a, uint32_t b) { return (a + b - 1) / b; }
inline uint64_t next_multiple(uint64_t a, uint64_t b) { return (a + b - 1) / b * b; }

// ------------------------------------------------------------------------------------------------
// pcg32 (O'Neill) — the reference's tcnn::default_rng_t; kernels take it by value and advance(i*k)
// (reference usage: testbed_nerf.cu:649,1421,1758).
// ------------------------------------------------------------------------------------------------
struct Pcg32 {
    uint64_t state, inc;
    static constexpr uint64_t MULT = 0x5851f42d4c957f2dULL;
    NGP_HD Pcg32() : state(0x853c49e6748fea9bULL), inc(0xda3e39cb94b95bdbULL) {}
    NGP_HD explicit Pcg32(uint64_t initstate, uint64_t initseq = 1u) { seed(initstate, initseq); }
    NGP_HD void seed(uint64_t initstate, uint64_t initseq = 1u) {
        state = 0u; inc = (initseq << 1u) | 1u; next_uint(); state += initstate; next_uint();
    }
    NGP_HD uint32_t next_uint() {
        uint64_t old = state;
        state = old * MULT + inc;
        uint32_t xorshifted = (uint32_t)(((old >> 18u) ^ old) >> 27u);
        uint32_t rot = (uint32_t)(old >> 59u);
        return (xorshifted >> rot) | (xorshifted << ((~rot + 1u) & 31));
    }
    NGP_HD float next_float() {
        uint32_t u = (next_uint() >> 9) | 0x3f800000u;
#ifdef __CUDA_ARCH__
        return __uint_as_float(u) - 1.0f;
#else
        float f; memcpy(&f, &u, 4); return f - 1.0f;
#endif
    }
    NGP_HD void advance(int64_t delta_ = (1ll << 32)) {
        uint64_t cur_mult = MULT, cur_plus = inc, acc_mult = 1u, acc_plus = 0u, delta = (uint64_t)delta_;
        while (delta > 0) {
            if (delta & 1) { acc_mult *= cur_mult; acc_plus = acc_plus * cur_mult + cur_plus; }
            cur_plus = (cur_mult + 1) * cur_plus;
            cur_mult *= cur_mult;
            delta /= 2;
        }
        state = acc_mult * state + acc_plus;
    }
};

// ------------------------------------------------------------------------------------------------
// morton codes (128^3 occupancy grid is stored in morton order: testbed_nerf.cu:433-447,666-668)
// ------------------------------------------------------------------------------------------------
NGP_HD uint32_t expand_bits(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu; v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u; v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
NGP_HD uint32_t morton3D(uint32_t x, uint32_t y, uint32_t z) { return expand_bits(x) | (expand_bits(y) << 1) | (expand_bits(z) << 2); }
NGP_HD uint32_t morton3D_invert(uint32_t x) {
    x = x & 0x49249249u; x = (x | (x >> 2)) & 0xc30c30c3u; x = (x | (x >> 4)) & 0x0f00f00fu;
    x = (x | (x >> 8)) & 0xff0000ffu; x = (x | (x >> 16)) & 0x0000ffffu;
    return x;
}

// ------------------------------------------------------------------------------------------------
// tiny vec3
// ------------------------------------------------------------------------------------------------
struct V3 { float x, y, z; };
NGP_HD V3 v3(float x, float y, float z) { return {x, y, z}; }
NGP_HD V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
NGP_HD V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
NGP_HD V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
NGP_HD V3 operator*(float s, V3 a) { return {a.x * s, a.y * s, a.z * s}; }
NGP_HD V3 operator*(V3 a, V3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
NGP_HD V3 operator/(V3 a, V3 b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
NGP_HD float dot3(V3 a, V3 b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)); }
NGP_HD V3 normalize3(V3 a) { float inv = 1.0f / sqrtf(dot3(a, a)); return a * inv; }
NGP_HD float length3(V3 a) { return sqrtf(dot3(a, a)); }
NGP_HD V3 ray_at(V3 o, V3 d, float t) { return {fmaf(t, d.x, o.x), fmaf(t, d.y, o.y), fmaf(t, d.z, o.z)}; }
NGP_HD V3 ld3(const float* p) { return {p[0], p[1], p[2]}; }
// mat3(cam) * v, cam = column-major 4x3 (x-axis, y-axis, z-axis, position)
NGP_HD V3 mat3_mul(const float* m, V3 v) {
    return {fmaf(m[6], v.z, fmaf(m[3], v.y, m[0] * v.x)), fmaf(m[7], v.z, fmaf(m[4], v.y, m[1] * v.x)),
            fmaf(m[8], v.z, fmaf(m[5], v.y, m[2] * v.x))};
}
NGP_HD float clampf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }
NGP_HD int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

struct Aabb { float min[3]; float max[3]; };

// ------------------------------------------------------------------------------------------------
// "blk8" activation layout: matrix X[rows][C] fp16, C % 8 == 0.  Element (r, c) lives at
//     ((r/8) * (C/8) + c/8) * 64 + (r%8) * 8 + (c%8)
// i.e. 8x8 core matrices of 128 contiguous bytes.  A 128-row tile is one contiguous block that a single
// bulk-async copy (TMA, UBLKCP) brings to shared memory already in the tcgen05 no-swizzle canonical layout
// (K-major: SBO = C/8*128 B, LBO = 128 B; MN-major: SBO = 128 B, LBO = C/8*128 B).
// ------------------------------------------------------------------------------------------------
NGP_HD size_t blk8_chunk_offset(uint32_t r, uint32_t chunk, uint32_t C) {   // offset in halves of the 8-half chunk
    return ((size_t)(r >> 3) * (C >> 3) + chunk) * 64 + (size_t)(r & 7) * 8;
}
NGP_HD size_t blk8_offset(uint32_t r, uint32_t c, uint32_t C) { return blk8_chunk_offset(r, c >> 3, C) + (c & 7); }

}  // namespace ngp

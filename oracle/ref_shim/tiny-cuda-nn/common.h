// ORACLE BUILD STUB (test infrastructure): the few tiny-cuda-nn symbols the reference's helper headers mention.
#pragma once
#include <cstdint>
#include <cstddef>
#include <cmath>
#include <string>
#include <stdexcept>
#define TCNN_HOST_DEVICE
#define TCNN_NAMESPACE_BEGIN namespace tcnn {
#define TCNN_NAMESPACE_END }
#define TCNN_MIN_GPU_ARCH 100
namespace tcnn {
using network_precision_t = __half;
template <typename T> T clamp(T val, T lower, T upper) { return val < lower ? lower : (upper < val ? upper : val); }
template <typename T> T div_round_up(T val, T divisor) { return (val + divisor - 1) / divisor; }
template <typename T> T next_multiple(T val, T divisor) { return div_round_up(val, divisor) * divisor; }
template <typename T> void host_device_swap(T& a, T& b) { T c(a); a = b; b = c; }
template <typename T, uint32_t N> struct vector_t { T data[N]; T& operator[](uint32_t i) { return data[i]; } };
template <typename T> struct PitchedPtr { T* ptr; size_t stride_in_bytes; };
// morton3D (restated: 10-bit interleave, the public "expand bits" construction)
inline uint32_t expand_bits(uint32_t v) { v = (v * 0x00010001u) & 0xFF0000FFu; v = (v * 0x00000101u) & 0x0F00F00Fu; v = (v * 0x00000011u) & 0xC30C30C3u; v = (v * 0x00000005u) & 0x49249249u; return v; }
inline uint32_t morton3D(uint32_t x, uint32_t y, uint32_t z) { return expand_bits(x) | (expand_bits(y) << 1) | (expand_bits(z) << 2); }
static constexpr float PI = 3.14159265358979323846f;
static constexpr float SQRT2 = 1.41421356237309504880f;
inline float logistic(const float x) { return 1.0f / (1.0f + expf(-x)); }
inline float logit(const float x) { return -logf(1.0f / (fminf(fmaxf(x, 1e-9f), 1.0f - 1e-9f)) - 1.0f); }
}

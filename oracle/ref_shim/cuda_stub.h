// ORACLE BUILD STUB (test infrastructure): lets g++ compile the reference's __host__ __device__ helper headers.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <algorithm>
#define __host__
#define __device__
#define __global__
#define __forceinline__ inline
#define __restrict__
#define __launch_bounds__(...)
struct __half { _Float16 v; __half() = default; __half(float f) : v((_Float16)f) {} operator float() const { return (float)v; } };
struct __half2 { __half x, y; };
inline __half2 __floats2half2_rn(float a, float b) { return {__half(a), __half(b)}; }
inline float __half2float(__half h) { return (float)h; }
struct dim3_stub { unsigned x = 0, y = 0, z = 0; };
static dim3_stub threadIdx, blockIdx, blockDim, gridDim;
struct float2 { float x, y; }; struct float3 { float x, y, z; }; struct float4 { float x, y, z, w; };
struct uint2 { unsigned x, y; }; struct uint3 { unsigned x, y, z; }; struct uint4 { unsigned x, y, z, w; };
struct int2 { int x, y; }; struct int3 { int x, y, z; };
using cudaStream_t = void*;
inline float __saturatef(float x) { return std::fmin(std::fmax(x, 0.0f), 1.0f); }
inline float copysignf_(float a, float b) { return std::copysign(a, b); }

// Roofline denominators for the hash-grid kernels, measured on the device the library runs on (SURVEY.md 8(d), config 1:
// "denominator = a measured random-4-B-gather microbenchmark over a 24 MB persisting-L2 buffer on the same GPU").
//   l2_gather : every thread keeps LOADS_IN_FLIGHT independent 4-byte loads in flight at addresses hashed in registers
//               (no index array to read, no dependent chain) over a buffer of table size -> peak random-gather rate
//               (gathers/s; x32 B = L2 sector bandwidth when every gather is its own sector)
//   l2_red    : every thread issues scattered red.global.add.noftz.{f16x2 | v2.f16x2 | v4.f16x2} at hashed addresses
//               -> peak scattered-reduction rate (lanes/s), the bound of the encoder's backward scatter
// Both are what the grid kernels would reach with no arithmetic at all; bench.py reports grid_forward / grid_backward as
// fractions of them next to the HBM-copy fraction.
#include "ngp_internal.h"
#include <algorithm>
#include <vector>

namespace ngp {

NGP_D uint32_t mix32(uint32_t x) {   // lowbias32 (Wellons): full-avalanche integer hash, 6 instructions
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}

constexpr uint32_t LOADS_IN_FLIGHT = 32;

__global__ void __launch_bounds__(256) l2_gather_kernel(const uint32_t* __restrict__ buf, uint32_t n_words_mask, uint32_t iters, uint32_t* __restrict__ sink) {
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t acc = 0, key = mix32(tid * 2654435761u + 1u);
    for (uint32_t it = 0; it < iters; ++it) {
        uint32_t v[LOADS_IN_FLIGHT];
#pragma unroll
        for (uint32_t k = 0; k < LOADS_IN_FLIGHT; ++k) { key = mix32(key + k + 1u); v[k] = __ldg(buf + (key & n_words_mask)); }
#pragma unroll
        for (uint32_t k = 0; k < LOADS_IN_FLIGHT; ++k) acc += v[k];
    }
    if (acc == 0x12345678u) sink[0] = acc;   // keeps the loads alive
}

template <int VEC>   // 1: f16x2 (4 B), 2: v2.f16x2 (8 B), 4: v4.f16x2 (16 B) per lane
__global__ void __launch_bounds__(256) l2_red_kernel(__half* __restrict__ buf, uint32_t n_slots_mask, uint32_t iters) {
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t key = mix32(tid * 2654435761u + 7u);
    const uint32_t one = 0x00010001u;   // two fp16 denormals: the sum never overflows
    for (uint32_t it = 0; it < iters; ++it) {
#pragma unroll
        for (uint32_t k = 0; k < 8; ++k) {
            key = mix32(key + k + 1u);
            __half* p = buf + (size_t)(key & n_slots_mask) * (2 * VEC);
            if (VEC == 1) asm volatile("red.global.add.noftz.f16x2 [%0], %1;" ::"l"(p), "r"(one) : "memory");
            else if (VEC == 2) asm volatile("red.global.add.noftz.v2.f16x2 [%0], {%1, %2};" ::"l"(p), "r"(one), "r"(one) : "memory");
            else asm volatile("red.global.add.noftz.v4.f16x2 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(one), "r"(one), "r"(one), "r"(one) : "memory");
        }
    }
}

static uint32_t floor_pow2(uint64_t v) { uint32_t p = 1; while ((uint64_t)p * 2 <= v) p *= 2; return p; }

// returns gathers per second (median of `reps` timed launches after one warm-up)
double microbench_l2_gather(cudaStream_t s, uint64_t buffer_bytes, uint32_t reps) {
    NGP_CHECK(buffer_bytes >= 4096 && reps >= 1 && reps <= 64, "microbench_l2_gather: bad arguments");
    const uint32_t n_words = floor_pow2(buffer_bytes / 4);
    uint32_t* buf = nullptr; uint32_t* sink = nullptr;
    NGP_CUDA_CHECK(cudaMalloc(&buf, (size_t)n_words * 4)); NGP_CUDA_CHECK(cudaMalloc(&sink, 4));
    NGP_CUDA_CHECK(cudaMemsetAsync(buf, 0, (size_t)n_words * 4, s));
    int dev = 0, sms = 148; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const uint32_t blocks = (uint32_t)sms * 8u, iters = 16;   // 8 x 256 threads per SM resident, 512 gathers per thread
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    std::vector<float> ms(reps);
    for (uint32_t r = 0; r <= reps; ++r) {
        cudaEventRecord(e0, s);
        l2_gather_kernel<<<blocks, 256, 0, s>>>(buf, n_words - 1, iters, sink);
        NGP_POST_LAUNCH();
        cudaEventRecord(e1, s);
        NGP_CUDA_CHECK(cudaEventSynchronize(e1));
        if (r > 0) cudaEventElapsedTime(&ms[r - 1], e0, e1);
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(buf); cudaFree(sink);
    std::sort(ms.begin(), ms.end());
    const double n = (double)blocks * 256.0 * iters * LOADS_IN_FLIGHT;
    return n / (ms[reps / 2] * 1e-3);
}

// returns scattered reduction lanes per second; vec = entries of 4 bytes per lane (1, 2 or 4)
double microbench_l2_red(cudaStream_t s, uint64_t buffer_bytes, int vec, uint32_t reps) {
    NGP_CHECK(buffer_bytes >= 4096 && reps >= 1 && reps <= 64 && (vec == 1 || vec == 2 || vec == 4), "microbench_l2_red: bad arguments");
    const uint32_t n_slots = floor_pow2(buffer_bytes / (4 * (uint64_t)vec));
    __half* buf = nullptr;
    NGP_CUDA_CHECK(cudaMalloc(&buf, (size_t)n_slots * 4 * vec));
    NGP_CUDA_CHECK(cudaMemsetAsync(buf, 0, (size_t)n_slots * 4 * vec, s));
    int dev = 0, sms = 148; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const uint32_t blocks = (uint32_t)sms * 8u, iters = 8;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    std::vector<float> ms(reps);
    for (uint32_t r = 0; r <= reps; ++r) {
        cudaEventRecord(e0, s);
        if (vec == 1) l2_red_kernel<1><<<blocks, 256, 0, s>>>(buf, n_slots - 1, iters);
        else if (vec == 2) l2_red_kernel<2><<<blocks, 256, 0, s>>>(buf, n_slots - 1, iters);
        else l2_red_kernel<4><<<blocks, 256, 0, s>>>(buf, n_slots - 1, iters);
        NGP_POST_LAUNCH();
        cudaEventRecord(e1, s);
        NGP_CUDA_CHECK(cudaEventSynchronize(e1));
        if (r > 0) cudaEventElapsedTime(&ms[r - 1], e0, e1);
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(buf);
    std::sort(ms.begin(), ms.end());
    const double n = (double)blocks * 256.0 * iters * 8.0;
    return n / (ms[reps / 2] * 1e-3);
}

}  // namespace ngp

// Tensor-core / TMEM micro-benchmarks for the fused-MLP design (tools, not product): per-SM rates of
//   tcgen05.ld / tcgen05.st (TMEM <-> registers), tcgen05.mma kind::f16 with A from shared memory (SS) and from TMEM (TS)
// at the tile shapes the 64-neuron MLP uses (M128 x N{16,64} x K16; M64 wgrad with MN-major operands), plus a numeric
// check that an A operand written with tcgen05.st.32x32b (two fp16 per 32-bit column, lane = row) is what TS-mode reads.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I../instant-ngp_b200/csrc -o ubench_tc ubench_tc.cu
#include "tc05.cuh"
#include <cstdio>
#include <cstdlib>
#include <vector>

using namespace ngp;
using namespace ngp::tc;

__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t* r) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
                 ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
                   "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]) : "memory");
}
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// D[tmem] (+)= A[tmem] * B[smem]
__device__ __forceinline__ void umma_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}

struct Result { long long cyc[16]; float check[128 * 16]; };

constexpr int ITERS = 256;

// mode of the MMA loop
enum { SS_N64 = 0, SS_N16, SS_N128, SS_N256, SS_M64_MN_N64, TS_N64, TS_N16, SS_N32, N_MODES };

__global__ void __launch_bounds__(128) ubench_kernel(Result* out, int active_warps) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~(uintptr_t)127);
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_slot;
    const uint32_t tid = threadIdx.x, warp = tid >> 5;
    for (uint32_t i = tid; i < 48 * 1024 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0u;
    if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    if (warp == 0) tmem_alloc<512>(&tmem_slot);
    fence_proxy_async(); fence_before_sync(); __syncthreads(); fence_after_sync();
    const uint32_t tbase = tmem_slot, trow = tbase + ((warp * 32u) << 16);
    Result* res = out + blockIdx.x;
    uint32_t parity = 0;

    // ---- 1. tcgen05.ld throughput: ITERS x (4 x x16) per warp ----
    float sink = 0.f;
    __syncthreads();
    long long t0 = clock64();
    if ((int)warp < active_warps) {
        for (int it = 0; it < ITERS; ++it) {
            float v[64];
#pragma unroll
            for (int j = 0; j < 64; j += 16) tmem_ld16(trow + j + (it & 3) * 64, v + j);
            tmem_wait_ld();
            sink += v[0] + v[17] + v[34] + v[51];
        }
    }
    __syncthreads();
    long long t1 = clock64();
    if (tid == 0) res->cyc[0] = t1 - t0;

    // ---- 2. tcgen05.st throughput ----
    uint32_t regs[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) regs[j] = tid + j;
    __syncthreads();
    t0 = clock64();
    if ((int)warp < active_warps) {
        for (int it = 0; it < ITERS; ++it) {
#pragma unroll
            for (int j = 0; j < 64; j += 16) tmem_st16(trow + 256 + j + (it & 3) * 64, regs);
            tmem_wait_st();
        }
    }
    __syncthreads();
    t1 = clock64();
    if (tid == 0) res->cyc[1] = t1 - t0;

    // ---- 3. MMA issue loops ----
    const uint32_t sbase = smem_u32(smem);
    const Operand A = {sbase, 64}, B = {sbase + 16384, 64}, B256 = {sbase + 16384, 64};
    for (int mode = 0; mode < N_MODES; ++mode) {
        fence_proxy_async(); fence_before_sync(); __syncthreads();
        t0 = clock64();
        if (tid == 0) {
            fence_after_sync();
            for (int it = 0; it < ITERS; ++it) {
                const uint32_t k = (it & 3) * 16;
                switch (mode) {
                    case SS_N64: umma(tbase, desc_kmajor(A, k), desc_kmajor(B, k), make_idesc(128, 64, 0, 0), it > 0); break;
                    case SS_N16: umma(tbase, desc_kmajor(A, k), desc_kmajor(B, k), make_idesc(128, 16, 0, 0), it > 0); break;
                    case SS_N32: umma(tbase, desc_kmajor(A, k), desc_kmajor(B, k), make_idesc(128, 32, 0, 0), it > 0); break;
                    case SS_N128: umma(tbase, desc_kmajor(A, k), desc_kmajor(B256, k), make_idesc(128, 128, 0, 0), it > 0); break;
                    case SS_N256: umma(tbase, desc_kmajor(A, k), desc_kmajor(B256, k), make_idesc(128, 256, 0, 0), it > 0); break;
                    case SS_M64_MN_N64: umma(tbase, desc_mnmajor(A, 0, k), desc_mnmajor(B, 0, k), make_idesc(64, 64, 1, 1), it > 0); break;
                    case TS_N64: umma_ts(tbase, tbase + 256 + (it & 3) * 8, desc_kmajor(B, k), make_idesc(128, 64, 0, 0), it > 0); break;
                    case TS_N16: umma_ts(tbase, tbase + 256 + (it & 3) * 8, desc_kmajor(B, k), make_idesc(128, 16, 0, 0), it > 0); break;
                }
            }
            umma_commit(&bar);
        }
        mbar_wait(&bar, parity); parity ^= 1u;
        fence_after_sync();
        t1 = clock64();
        if (tid == 0) res->cyc[2 + mode] = t1 - t0;
    }

    // ---- 4. numeric check of the TS-mode A layout: A[r][k] = (r % 5) + k/4 written by tcgen05.st (packed pairs), B[n][k] = (n == k) ----
    {
        __half* b = reinterpret_cast<__half*>(smem + 16384);
        for (uint32_t i = tid; i < 16 * 16; i += 128) { const uint32_t n = i / 16, k = i % 16; b[blk8_offset(n, k, 16)] = __float2half(n == k ? 1.f : 0.f); }
        uint32_t a[16];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const __half2 h = __floats2half2_rn((float)(tid % 5) + (float)(2 * j) * 0.25f, (float)(tid % 5) + (float)(2 * j + 1) * 0.25f);
            a[j] = *reinterpret_cast<const uint32_t*>(&h);
        }
#pragma unroll
        for (int j = 8; j < 16; ++j) a[j] = 0;
        tmem_st16(trow + 384, a);
        tmem_wait_st();
        fence_proxy_async(); fence_before_sync(); __syncthreads();
        if (tid == 0) {
            fence_after_sync();
            umma_ts(tbase + 128, tbase + 384, desc_kmajor(Operand{sbase + 16384, 16}, 0), make_idesc(128, 16, 0, 0), 0);
            umma_commit(&bar);
        }
        mbar_wait(&bar, parity); parity ^= 1u;
        fence_after_sync();
        float v[16];
        tmem_ld16(trow + 128, v); tmem_wait_ld();
        if (blockIdx.x == 0) for (int j = 0; j < 16; ++j) res->check[tid * 16 + j] = v[j];
    }
    if (sink == 123.456f) res->cyc[15] = 1;
    fence_before_sync(); __syncthreads();
    if (warp == 0) tmem_dealloc<512>(tbase);
}

int main(int argc, char** argv) {
    const int grid = argc > 1 ? atoi(argv[1]) : 148;
    Result* d; cudaMalloc(&d, sizeof(Result) * grid);
    cudaFuncSetAttribute(ubench_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    std::vector<Result> h(grid);
    const char* names[N_MODES] = {"SS M128 N64 K16", "SS M128 N16 K16", "SS M128 N128 K16", "SS M128 N256 K16", "SS M64 N64 K16 (MN-major both)", "TS M128 N64 K16", "TS M128 N16 K16", "SS M128 N32 K16"};
    for (int aw : {4, 1}) {
        cudaMemset(d, 0, sizeof(Result) * grid);
        ubench_kernel<<<grid, 128, 64 * 1024>>>(d, aw);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return 1; }
        cudaMemcpy(h.data(), d, sizeof(Result) * grid, cudaMemcpyDeviceToHost);
        const Result& r = h[0];
        const double ld_bytes = (double)aw * ITERS * 64 * 32 * 4;
        printf("grid %d, %d active warps per CTA\n", grid, aw);
        printf("  tcgen05.ld 32x32b.x16 : %lld cyc for %.0f B  -> %.1f B/cyc/SM (%.1f per warp)\n", r.cyc[0], ld_bytes, ld_bytes / r.cyc[0], ld_bytes / r.cyc[0] / aw);
        printf("  tcgen05.st 32x32b.x16 : %lld cyc for %.0f B  -> %.1f B/cyc/SM (%.1f per warp)\n", r.cyc[1], ld_bytes, ld_bytes / r.cyc[1], ld_bytes / r.cyc[1] / aw);
        if (aw == 4) for (int m = 0; m < N_MODES; ++m) printf("  mma %-32s: %lld cyc / %d = %.1f cyc per instruction\n", names[m], r.cyc[2 + m], ITERS, (double)r.cyc[2 + m] / ITERS);
    }
    int bad = 0;
    for (int t = 0; t < 128; ++t) for (int j = 0; j < 16; ++j) { const float want = (float)(t % 5) + j * 0.25f; if (h[0].check[t * 16 + j] != want) { if (bad < 8) printf("  TS layout mismatch row %d col %d: got %g want %g\n", t, j, h[0].check[t * 16 + j], want); ++bad; } }
    printf("TS-mode A layout check (tcgen05.st 32x32b packed fp16 pairs, lane = row): %s (%d mismatches)\n", bad ? "FAIL" : "OK", bad);
    return 0;
}

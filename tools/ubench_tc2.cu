// Tensor-core issue / latency micro-benchmarks, second set (tools, not product).  Questions answered for the fused-MLP design:
//   Q1  what does ONE tcgen05.mma cost when issued (a) by a divergent `tid == 0` thread with descriptors computed per call (the
//       round-1 kernels) and (b) from a warp-uniform path (elect.sync) with descriptors precomputed — issue-side or tensor-side?
//   Q2  do dependent accumulations into the same TMEM columns serialise (chain vs 4 independent accumulators)?
//   Q3  latency of mma -> commit -> mbarrier wait for 1, 2, 4 MMAs (one MLP layer), and of the epilogue pieces.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -I../instant-ngp_b200/csrc -o ubench_tc2 ubench_tc2.cu
#include "tc05.cuh"
#include <cstdio>
#include <cstdlib>
#include <vector>

using namespace ngp;
using namespace ngp::tc;

__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}

constexpr int NRES = 32;
struct Result { long long cyc[NRES]; };
constexpr int REP = 16;     // outer repetitions of every measured block

#define SYNC_ALL() do { fence_proxy_async(); fence_before_sync(); __syncthreads(); fence_after_sync(); } while (0)

template <uint32_t N>
__global__ void __launch_bounds__(128) ubench2_kernel(Result* out) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~(uintptr_t)127);
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_slot;
    const uint32_t tid = threadIdx.x;
    const uint32_t warp = __shfl_sync(0xffffffffu, tid >> 5, 0);      // warp-uniform for the compiler
    for (uint32_t i = tid; i < 48 * 1024 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0u;
    if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    if (warp == 0) tmem_alloc<512>(&tmem_slot);
    SYNC_ALL();
    const uint32_t tbase = tmem_slot, trow = tbase + ((warp * 32u) << 16);
    Result* res = out + blockIdx.x;
    uint32_t parity = 0;
    const uint32_t sbase = smem_u32(smem);
    const Operand A = {sbase, 64}, B = {sbase + 16384, 64};
    constexpr uint32_t idesc = make_idesc(128, N, 0, 0);
    long long t0, t1;
    int slot = 0;

    // ---- V1: divergent single-thread issue, descriptors built per call (round-1 style), 16 dependent MMAs, one commit ----
    for (int rep = 0; rep < 2; ++rep) {
        SYNC_ALL(); t0 = clock64();
        if (tid == 0) {
            for (int r = 0; r < REP; ++r)
                for (uint32_t k = 0; k < 64; k += 16) umma(tbase, desc_kmajor(A, k), desc_kmajor(B, k), idesc, (k | r) > 0);
            umma_commit(&bar);
        }
        mbar_wait(&bar, parity); parity ^= 1u; fence_after_sync(); t1 = clock64();
        if (tid == 0 && rep == 1) res->cyc[slot] = t1 - t0;
    }
    ++slot;   // 0: V1 total for REP*4 MMAs

    // ---- V2: warp-uniform issue (elect.sync), descriptors precomputed, fully unrolled, dependent chain ----
    uint64_t ad[4], bd[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) { ad[k] = desc_kmajor(A, k * 16); bd[k] = desc_kmajor(B, k * 16); }
    for (int rep = 0; rep < 2; ++rep) {
        SYNC_ALL(); t0 = clock64();
        if (warp == 0) {
            if (elect_one()) {
#pragma unroll
                for (int r = 0; r < REP; ++r) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) umma(tbase, ad[k], bd[k], idesc, (k | r) > 0);
                }
                umma_commit(&bar);
            }
            __syncwarp();
        }
        mbar_wait(&bar, parity); parity ^= 1u; fence_after_sync(); t1 = clock64();
        if (tid == 0 && rep == 1) res->cyc[slot] = t1 - t0;
    }
    ++slot;   // 1: V2 total for REP*4 MMAs

    // ---- V3: as V2 with four independent accumulators (columns 0, N, 2N, 3N... capped at 4 x 64) ----
    for (int rep = 0; rep < 2; ++rep) {
        SYNC_ALL(); t0 = clock64();
        if (warp == 0) {
            if (elect_one()) {
#pragma unroll
                for (int r = 0; r < REP; ++r) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) umma(tbase + (N > 64 ? (k & 1) * 256 : k * 64), ad[k], bd[k], idesc, (N > 64 ? (k >> 1) | r : r) > 0);
                }
                umma_commit(&bar);
            }
            __syncwarp();
        }
        mbar_wait(&bar, parity); parity ^= 1u; fence_after_sync(); t1 = clock64();
        if (tid == 0 && rep == 1) res->cyc[slot] = t1 - t0;
    }
    ++slot;   // 2: V3 total

    // ---- V4: issue -> commit -> wait round trip for n = 1, 2, 4 MMAs (uniform issue), averaged over REP ----
    for (int n = 1; n <= 4; n *= 2) {
        SYNC_ALL(); t0 = clock64();
        for (int r = 0; r < REP; ++r) {
            if (warp == 0) {
                if (elect_one()) {
                    for (int k = 0; k < n; ++k) umma(tbase, ad[k], bd[k], idesc, k > 0);
                    umma_commit(&bar);
                }
                __syncwarp();
            }
            mbar_wait(&bar, parity); parity ^= 1u; fence_after_sync();
        }
        t1 = clock64();
        if (tid == 0) res->cyc[slot] = t1 - t0;
        ++slot;   // 3,4,5
    }

    // ---- V5: the round-1 StageSync round trip: fence + __syncthreads + tid==0 issue of 4 MMAs + commit + all wait ----
    SYNC_ALL(); t0 = clock64();
    for (int r = 0; r < REP; ++r) {
        fence_proxy_async(); fence_before_sync(); __syncthreads();
        if (tid == 0) { fence_after_sync(); for (uint32_t k = 0; k < 64; k += 16) umma(tbase, desc_kmajor(A, k), desc_kmajor(B, k), idesc, k > 0); umma_commit(&bar); }
        mbar_wait(&bar, parity); parity ^= 1u; fence_after_sync();
    }
    t1 = clock64();
    if (tid == 0) res->cyc[slot] = t1 - t0;
    ++slot;   // 6

    // ---- V6: epilogue pieces per 128 x 64 tile, all 4 warps: (a) tcgen05.ld 64 cols + wait, (b) + relu/pack + 8 x st.shared.v4, (c) + tcgen05.st 32 cols + wait ----
    __half* hs = reinterpret_cast<__half*>(smem + 32768);
    float keep = 0.f;
    for (int variant = 0; variant < 3; ++variant) {
        SYNC_ALL(); t0 = clock64();
        for (int r = 0; r < REP; ++r) {
            float v[64];
#pragma unroll
            for (int j = 0; j < 64; j += 16) tmem_ld16(trow + j, v + j);
            tmem_wait_ld();
            if (variant == 0) { keep += v[0] + v[21] + v[42] + v[63]; continue; }
            uint32_t pk[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) asm("cvt.rn.relu.f16x2.f32 %0, %1, %2;" : "=r"(pk[j]) : "f"(v[2 * j + 1]), "f"(v[2 * j]));
            if (variant == 1) {
#pragma unroll
                for (int c = 0; c < 8; ++c) *reinterpret_cast<uint4*>(hs + blk8_chunk_offset(tid, c, 64)) = make_uint4(pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
            } else {
                asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
                             ::"r"(trow + 256), "r"(pk[0]), "r"(pk[1]), "r"(pk[2]), "r"(pk[3]), "r"(pk[4]), "r"(pk[5]), "r"(pk[6]), "r"(pk[7]), "r"(pk[8]), "r"(pk[9]), "r"(pk[10]),
                               "r"(pk[11]), "r"(pk[12]), "r"(pk[13]), "r"(pk[14]), "r"(pk[15]), "r"(pk[16]), "r"(pk[17]), "r"(pk[18]), "r"(pk[19]), "r"(pk[20]), "r"(pk[21]),
                               "r"(pk[22]), "r"(pk[23]), "r"(pk[24]), "r"(pk[25]), "r"(pk[26]), "r"(pk[27]), "r"(pk[28]), "r"(pk[29]), "r"(pk[30]), "r"(pk[31]) : "memory");
                asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
            }
        }
        __syncthreads(); t1 = clock64();
        if (tid == 0) res->cyc[slot] = t1 - t0;
        ++slot;   // 7,8,9
    }

    // ---- V7: TS-mode chain: A from TMEM (cols 256..), B smem, uniform issue, dependent ----
    for (int rep = 0; rep < 2; ++rep) {
        SYNC_ALL(); t0 = clock64();
        if (warp == 0) {
            if (elect_one()) {
#pragma unroll
                for (int r = 0; r < REP; ++r) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) umma_ts(tbase, tbase + 256 + k * 8, bd[k], idesc, (k | r) > 0);
                }
                umma_commit(&bar);
            }
            __syncwarp();
        }
        mbar_wait(&bar, parity); parity ^= 1u; fence_after_sync(); t1 = clock64();
        if (tid == 0 && rep == 1) res->cyc[slot] = t1 - t0;
    }
    ++slot;   // 10

    // ---- V8: weight-gradient form (gemm_pT_q): M = 64, A and B both MN-major over a [128 x 64] tile, K = 128 in 8 steps, dependent ----
    // ---- V9: dgrad form (gemm_act_w): M = 128, A K-major, B MN-major, 4 steps ----
    {
        constexpr uint32_t NW = N > 64 ? 64 : N;
        constexpr uint32_t idesc_w = make_idesc(64, NW, 1, 1), idesc_g = make_idesc(128, NW, 0, 1);
        uint64_t pa[8], qb[8], gb[4];
#pragma unroll
        for (int r = 0; r < 8; ++r) { pa[r] = desc_mnmajor(A, 0, r * 16); qb[r] = desc_mnmajor(B, 0, r * 16); }
#pragma unroll
        for (int k = 0; k < 4; ++k) gb[k] = desc_mnmajor(B, 0, k * 16);
        for (int rep = 0; rep < 2; ++rep) {
            SYNC_ALL(); t0 = clock64();
            if (warp == 0) {
                if (elect_one()) {
#pragma unroll
                    for (int r = 0; r < REP; ++r) {
#pragma unroll
                        for (int k = 0; k < 8; ++k) umma(tbase + 64, pa[k], qb[k], idesc_w, (k | r) > 0);
                    }
                    umma_commit(&bar);
                }
                __syncwarp();
            }
            mbar_wait(&bar, parity); parity ^= 1u; fence_after_sync(); t1 = clock64();
            if (tid == 0 && rep == 1) res->cyc[slot] = t1 - t0;
        }
        ++slot;   // 11
        for (int rep = 0; rep < 2; ++rep) {
            SYNC_ALL(); t0 = clock64();
            if (warp == 0) {
                if (elect_one()) {
#pragma unroll
                    for (int r = 0; r < REP; ++r) {
#pragma unroll
                        for (int k = 0; k < 4; ++k) umma(tbase, ad[k], gb[k], idesc_g, (k | r) > 0);
                    }
                    umma_commit(&bar);
                }
                __syncwarp();
            }
            mbar_wait(&bar, parity); parity ^= 1u; fence_after_sync(); t1 = clock64();
            if (tid == 0 && rep == 1) res->cyc[slot] = t1 - t0;
        }
        ++slot;   // 12
    }

    if (keep == 123.456f) res->cyc[NRES - 1] = 1;
    fence_before_sync(); __syncthreads();
    if (warp == 0) tmem_dealloc<512>(tbase);
}

template <uint32_t N>
static void run(int grid) {
    Result* d; cudaMalloc(&d, sizeof(Result) * grid); cudaMemset(d, 0, sizeof(Result) * grid);
    cudaFuncSetAttribute(ubench2_kernel<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    ubench2_kernel<N><<<grid, 128, 64 * 1024>>>(d);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); exit(1); }
    std::vector<Result> h(grid); cudaMemcpy(h.data(), d, sizeof(Result) * grid, cudaMemcpyDeviceToHost);
    const long long* c = h[0].cyc; const double n = REP * 4.0;
    printf("M128 N%u K16, grid %d (cycles, CTA 0)\n", N, grid);
    printf("  V1 divergent issue, descriptors per call, %d dependent MMAs + commit + wait : %lld  (%.1f / MMA)\n", (int)n, c[0], c[0] / n);
    printf("  V2 uniform issue (elect), precomputed descriptors, dependent chain           : %lld  (%.1f / MMA)\n", c[1], c[1] / n);
    printf("  V3 same, 4 independent accumulators                                          : %lld  (%.1f / MMA)\n", c[2], c[2] / n);
    printf("  V4 issue -> commit -> wait round trip: 1 MMA %.0f, 2 MMAs %.0f, 4 MMAs %.0f cycles\n", c[3] / (double)REP, c[4] / (double)REP, c[5] / (double)REP);
    printf("  V5 round-1 StageSync round trip (fence, __syncthreads, tid==0 issues 4 MMAs, commit, all wait): %.0f cycles\n", c[6] / (double)REP);
    printf("  V6 epilogue per 128x64 tile (4 warps): ld+wait %.0f, +relu/pack+st.shared %.0f, +relu/pack+tcgen05.st %.0f cycles\n", c[7] / (double)REP, c[8] / (double)REP, c[9] / (double)REP);
    printf("  V7 TS mode (A in TMEM) uniform issue dependent chain                        : %lld  (%.1f / MMA)\n", c[10], c[10] / n);
    printf("  V8 wgrad form M64 N%u, A and B MN-major, 8 K-steps, dependent chain             : %lld  (%.1f / MMA)\n", N > 64 ? 64 : N, c[11], c[11] / (REP * 8.0));
    printf("  V9 dgrad form M128 N%u, A K-major, B MN-major, dependent chain                   : %lld  (%.1f / MMA)\n", N > 64 ? 64 : N, c[12], c[12] / n);
    cudaFree(d);
}

int main(int argc, char** argv) {
    const int grid = argc > 1 ? atoi(argv[1]) : 148;
    run<64>(grid); run<16>(grid); run<32>(grid); run<256>(grid);
    return 0;
}

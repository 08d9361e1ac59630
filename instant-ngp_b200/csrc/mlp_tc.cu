// Fully fused 64-neuron MLPs on the 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM).
// Replaces tcnn::FullyFusedMLP<__half,64> and the NerfNetwork glue kernels extract_density / extract_rgb /
// add_density_gradient (nerf_network.h:31-74) — the whole density-MLP -> [density16 | SH16] -> rgb-MLP chain runs
// per 128-sample tile with every activation resident in shared memory, input tiles fed by bulk-async (TMA) copies.
//
// Structure (one warpgroup = 128 threads per CTA, several CTAs per SM, persistent over tiles):
//   thread 0 issues tcgen05.mma for a layer and commits to an mbarrier; all 128 threads wait, then thread t reads
//   TMEM lane t (= sample row t of the tile) with tcgen05.ld, applies ReLU, rounds to fp16 (the same storage points
//   as the reference's fp16 activations) and writes its row as 16-byte chunks into the next layer's operand tile.
#include "ngp_internal.h"
#include "tc05.cuh"
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>

namespace ngp {
using namespace tc;

// ---- shared-memory carve-up helpers ---------------------------------------------------------------------------
struct SmemCursor {
    uint32_t off = 0;
    __host__ __device__ uint32_t take(uint32_t bytes) { uint32_t o = off; off += (bytes + 127u) & ~127u; return o; }
};

// copy a row-major fp16 weight matrix [rows x cols] from global into a blk8 tile in shared memory
NGP_D void load_weights_blk8(__half* dst, const __half* __restrict__ src, uint32_t rows, uint32_t cols, uint32_t tid, uint32_t nthreads) {
    const uint32_t cpr = cols >> 3;
    for (uint32_t i = tid; i < rows * cpr; i += nthreads) {
        const uint32_t r = i / cpr, c = i % cpr;
        *reinterpret_cast<uint4*>(dst + blk8_chunk_offset(r, c, cols)) = __ldg(reinterpret_cast<const uint4*>(src + (size_t)r * cols + c * 8));
    }
}

// the same with compile-time shapes: every 16-byte load of the thread is issued before the first store (the rolled loop above
// exposes one global-load latency per iteration: ~10 x 800 cycles of prologue per CTA for the 20 KB of the NeRF networks)
template <uint32_t ROWS, uint32_t COLS, uint32_t NT>
NGP_D void load_weights_blk8_t(__half* dst, const __half* __restrict__ src, uint32_t tid) {
    constexpr uint32_t CPR = COLS / 8, N = ROWS * CPR, IT = (N + NT - 1) / NT;
    uint4 v[IT];
#pragma unroll
    for (uint32_t k = 0; k < IT; ++k) { const uint32_t i = tid + k * NT; if (i < N) v[k] = __ldg(reinterpret_cast<const uint4*>(src + (size_t)(i / CPR) * COLS + (i % CPR) * 8)); }
#pragma unroll
    for (uint32_t k = 0; k < IT; ++k) { const uint32_t i = tid + k * NT; if (i < N) *reinterpret_cast<uint4*>(dst + blk8_chunk_offset(i / CPR, i % CPR, COLS)) = v[k]; }
}
// all layers of the density / plain chain and (MODE_NERF) the rgb chain into their shared-memory slots
template <uint32_t PW, uint32_t DH, uint32_t RH, bool NERF, uint32_t NT>
NGP_D void load_all_weights(uint8_t* smem, const uint32_t* wd, const uint32_t* wr, const __half* __restrict__ w_density, const __half* __restrict__ w_rgb, uint32_t tid) {
    const __half* w = w_density;
    load_weights_blk8_t<MLP_WIDTH, PW, NT>(reinterpret_cast<__half*>(smem + wd[0]), w, tid); w += MLP_WIDTH * PW;
#pragma unroll
    for (uint32_t l = 1; l < DH; ++l) { load_weights_blk8_t<MLP_WIDTH, MLP_WIDTH, NT>(reinterpret_cast<__half*>(smem + wd[l]), w, tid); w += MLP_WIDTH * MLP_WIDTH; }
    load_weights_blk8_t<MLP_OUT_PAD, MLP_WIDTH, NT>(reinterpret_cast<__half*>(smem + wd[DH]), w, tid);
    if (NERF) {
        w = w_rgb;
        load_weights_blk8_t<MLP_WIDTH, 32, NT>(reinterpret_cast<__half*>(smem + wr[0]), w, tid); w += MLP_WIDTH * 32;
#pragma unroll
        for (uint32_t l = 1; l < RH; ++l) { load_weights_blk8_t<MLP_WIDTH, MLP_WIDTH, NT>(reinterpret_cast<__half*>(smem + wr[l]), w, tid); w += MLP_WIDTH * MLP_WIDTH; }
        load_weights_blk8_t<MLP_OUT_PAD, MLP_WIDTH, NT>(reinterpret_cast<__half*>(smem + wr[RH]), w, tid);
    }
}

NGP_D uint4 pack8(const float* v, bool relu) {
    uint32_t h[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {   // (lo, hi) = (v[2i], v[2i+1]); ReLU rides on the conversion (cvt.rn.relu.f16x2.f32)
        if (relu) asm("cvt.rn.relu.f16x2.f32 %0, %1, %2;" : "=r"(h[i]) : "f"(v[2 * i + 1]), "f"(v[2 * i]));
        else asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(h[i]) : "f"(v[2 * i + 1]), "f"(v[2 * i]));
    }
    return make_uint4(h[0], h[1], h[2], h[3]);
}

// TMEM accumulator [128 x N] (row = this thread) -> fp16 blk8 tile `dst` (C_dst columns) at column c0
template <uint32_t N, bool RELU, uint32_t BATCH = N>
NGP_D void epilogue_to_smem(uint32_t tmem_row_addr, __half* dst, uint32_t C_dst, uint32_t c0, uint32_t row) {
    // BATCH columns per tcgen05.wait::ld: the whole row at once in the backward kernel (2 CTAs per SM, latency-bound);
    // 16 in the forward kernel (4 CTAs per SM, where the smaller register footprint and finer interleave measured faster)
#pragma unroll
    for (uint32_t b = 0; b < N; b += BATCH) {
        float v[BATCH];
#pragma unroll
        for (uint32_t j = 0; j < BATCH; j += 16) tmem_ld16(tmem_row_addr + b + j, v + j);
        tmem_wait_ld();
#pragma unroll
        for (uint32_t j = 0; j < BATCH; j += 8)
            *reinterpret_cast<uint4*>(dst + blk8_chunk_offset(row, (c0 + b + j) >> 3, C_dst)) = pack8(v + j, RELU);
    }
}

// SH degree 4 of 2*d-1 (replaces tcnn SphericalHarmonicsEncoding under Composite; configs/nerf/base.json:37-49)
NGP_D void sh4(float dx, float dy, float dz, float* o) {
    const float x = dx * 2.f - 1.f, y = dy * 2.f - 1.f, z = dz * 2.f - 1.f;
    const float xy = x * y, xz = x * z, yz = y * z, x2 = x * x, y2 = y * y, z2 = z * z;
    o[0] = 0.28209479177387814f;
    o[1] = -0.48860251190291987f * y;
    o[2] = 0.48860251190291987f * z;
    o[3] = -0.48860251190291987f * x;
    o[4] = 1.0925484305920792f * xy;
    o[5] = -1.0925484305920792f * yz;
    o[6] = 0.94617469575755997f * z2 - 0.31539156525251999f;
    o[7] = -1.0925484305920792f * xz;
    o[8] = 0.54627421529603959f * x2 - 0.54627421529603959f * y2;
    o[9] = 0.59004358992664352f * y * (-3.0f * x2 + y2);
    o[10] = 2.8906114426405538f * xy * z;
    o[11] = 0.45704579946446572f * y * (1.0f - 5.0f * z2);
    o[12] = 0.3731763325901154f * z * (5.0f * z2 - 3.0f);
    o[13] = 0.45704579946446572f * x * (1.0f - 5.0f * z2);
    o[14] = 1.4453057213202769f * z * (x2 - y2);
    o[15] = 0.59004358992664352f * x * (-x2 + 3.0f * y2);
}

// dL/d(direction input) from dL/d(SH4 outputs): the analytic derivatives of the 16 polynomials above, times 2 for x = 2 d - 1
// (tcnn SphericalHarmonicsEncoding backward with input gradients; NerfNetwork::backward_impl, nerf_network.h:282-296)
NGP_D void sh4_input_gradient(float dx, float dy, float dz, const float* g, float* out) {
    const float x = dx * 2.f - 1.f, y = dy * 2.f - 1.f, z = dz * 2.f - 1.f;
    const float xy = x * y, xz = x * z, yz = y * z, x2 = x * x, y2 = y * y, z2 = z * z;
    const float c1 = 0.48860251190291987f, c2 = 1.0925484305920792f, c3 = 0.94617469575755997f, c4 = 0.54627421529603959f, c5 = 0.59004358992664352f,
                c6 = 2.8906114426405538f, c7 = 0.45704579946446572f, c8 = 0.3731763325901154f, c9 = 1.4453057213202769f;
    const float gx = -c1 * g[3] + c2 * y * g[4] - c2 * z * g[7] + 2.f * c4 * x * g[8] - 6.f * c5 * xy * g[9] + c6 * yz * g[10] + c7 * (1.f - 5.f * z2) * g[13] +
                     2.f * c9 * xz * g[14] + c5 * (-3.f * x2 + 3.f * y2) * g[15];
    const float gy = -c1 * g[1] + c2 * x * g[4] - c2 * z * g[5] - 2.f * c4 * y * g[8] + c5 * (-3.f * x2 + 3.f * y2) * g[9] + c6 * xz * g[10] + c7 * (1.f - 5.f * z2) * g[11] -
                     2.f * c9 * yz * g[14] + 6.f * c5 * xy * g[15];
    const float gz = c1 * g[2] - c2 * y * g[5] + 2.f * c3 * z * g[6] - c2 * x * g[7] + c6 * xy * g[10] - 10.f * c7 * yz * g[11] + c8 * (15.f * z2 - 3.f) * g[12] -
                     10.f * c7 * xz * g[13] + c9 * (x2 - y2) * g[14];
    out[0] = 2.f * gx; out[1] = 2.f * gy; out[2] = 2.f * gz;
}

// One MMA stage: make this thread's smem writes visible to the async proxy, sync the CTA, let thread 0 issue.
// `issue` is a lambda run by thread 0 only.  All threads return once the MMAs have completed.
struct StageSync {
    uint64_t* bar; uint32_t parity;
    template <class F>
    NGP_D void run(F&& issue) {
        fence_proxy_async();
        fence_before_sync();
        __syncthreads();
        if (threadIdx.x == 0) {
            fence_after_sync();
            issue();
            umma_commit(bar);
        }
        mbar_wait(bar, parity);
        parity ^= 1u;
        fence_after_sync();
    }
};

enum MlpMode : int { MODE_NERF = 0, MODE_DENSITY = 1, MODE_PLAIN = 2 };

struct FwdParams {
    const __half* w_density; const __half* w_rgb;
    const __half* enc;        // [n x PW] blk8
    const float* coords;      // [n][7] (MODE_NERF)
    __half* out; uint32_t out_stride;
    uint32_t n_tiles;
    const uint32_t* n_dyn;    // optional device-side row count: tiles beyond ceil(*n_dyn / 128) are skipped (n_tiles sizes the grid)
    long long* probe;         // debug (NGP_MLP_PROBE=1): per-phase clock64 deltas of CTA 0 / thread 0, summed over its tiles; null in normal runs
};

template <uint32_t PW, uint32_t DH, uint32_t RH, int MODE>
struct FwdSmem {
    // byte offsets from the 128 B aligned dynamic smem base
    uint32_t wd[DH + 1], wr[RH + 1], x0, ha, r0, bars, total;
    __host__ __device__ FwdSmem() {
        SmemCursor c;
        wd[0] = c.take(MLP_WIDTH * PW * 2);
        for (uint32_t l = 1; l < DH; ++l) wd[l] = c.take(MLP_WIDTH * MLP_WIDTH * 2);
        wd[DH] = c.take(MLP_OUT_PAD * MLP_WIDTH * 2);
        if (MODE == MODE_NERF) {
            wr[0] = c.take(MLP_WIDTH * 32 * 2);
            for (uint32_t l = 1; l < RH; ++l) wr[l] = c.take(MLP_WIDTH * MLP_WIDTH * 2);
            wr[RH] = c.take(MLP_OUT_PAD * MLP_WIDTH * 2);
            r0 = c.take(TILE_ROWS * 32 * 2);
        } else { for (uint32_t l = 0; l <= RH; ++l) wr[l] = 0; r0 = 0; }
        x0 = c.take(TILE_ROWS * PW * 2);
        ha = c.take(TILE_ROWS * MLP_WIDTH * 2);   // one hidden-activation tile, rewritten in place layer after layer
        bars = c.take(64);
        total = c.off + 128;
    }
};

template <uint32_t PW, uint32_t DH, uint32_t RH, int MODE>
__global__ void __launch_bounds__(128) mlp_forward_kernel(const FwdParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~(uintptr_t)127);
    const FwdSmem<PW, DH, RH, MODE> L;
    const uint32_t tid = threadIdx.x, warp = tid >> 5;
    uint64_t* bar_full = reinterpret_cast<uint64_t*>(smem + L.bars);       // [2]
    uint64_t* bar_mma = bar_full + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_full + 3);

    // ---- one-time setup: barriers, TMEM, weights ----
    if (tid == 0) { mbar_init(&bar_full[0], 1); mbar_init(&bar_full[1], 1); mbar_init(bar_mma, 1); mbar_fence_init(); }
    if (warp == 0) tmem_alloc<64>(tmem_slot);
    load_all_weights<PW, DH, RH, MODE == MODE_NERF, 128>(smem, L.wd, L.wr, p.w_density, p.w_rgb, tid);
    fence_proxy_async();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t tmem_row = tmem_base + ((warp * 32u) << 16);   // this warp's lane quadrant, column 0

    const uint32_t sbase = smem_u32(smem);
    const Operand X0 = {sbase + L.x0, PW};
    const Operand HA = {sbase + L.ha, MLP_WIDTH}, R0 = {sbase + L.r0, 32};
    __half* ha = reinterpret_cast<__half*>(smem + L.ha);
    __half* r0 = reinterpret_cast<__half*>(smem + L.r0);
    constexpr uint32_t X_BYTES = TILE_ROWS * PW * 2;

    // n_dyn: the row count lives in device memory (tracer: rays alive x steps, known only after the compaction kernel has run)
    const uint32_t n_tiles = p.n_dyn ? min(p.n_tiles, (__ldg(p.n_dyn) + TILE_ROWS - 1) / TILE_ROWS) : p.n_tiles;
    StageSync stage{bar_mma, 0};
    uint32_t full_parity = 0;
    if (tid == 0 && blockIdx.x < n_tiles) {
        mbar_expect_tx(&bar_full[0], X_BYTES);
        bulk_g2s(smem + L.x0, p.enc + (size_t)blockIdx.x * TILE_ROWS * PW, X_BYTES, &bar_full[0]);
    }
    for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const uint32_t next = tile + gridDim.x;
        const uint32_t row_g = tile * TILE_ROWS + tid;
        float dir[3] = {0.f, 0.f, 0.f};
        if (MODE == MODE_NERF) { const float* c = p.coords + (size_t)row_g * 7 + 4; dir[0] = __ldg(c); dir[1] = __ldg(c + 1); dir[2] = __ldg(c + 2); }
        mbar_wait(&bar_full[0], full_parity); full_parity ^= 1u;

        // ---- density / plain network ----
        stage.run([&] { gemm_act_wT<64>(tmem_base, X0, 0, Operand{sbase + L.wd[0], PW}, PW); });
        // the first layer's MMAs have completed: the input tile is free -> fetch the next one under the remaining layers
        if (tid == 0 && next < n_tiles) {
            mbar_expect_tx(&bar_full[0], X_BYTES);
            bulk_g2s(smem + L.x0, p.enc + (size_t)next * TILE_ROWS * PW, X_BYTES, &bar_full[0]);
        }
        epilogue_to_smem<64, true, 16>(tmem_row, ha, MLP_WIDTH, 0, tid);
        for (uint32_t l = 1; l < DH; ++l) {   // hidden activations are rewritten in place: the MMAs that read them have completed
            stage.run([&] { gemm_act_wT<64>(tmem_base, HA, 0, Operand{sbase + L.wd[l], MLP_WIDTH}, MLP_WIDTH); });
            epilogue_to_smem<64, true, 16>(tmem_row, ha, MLP_WIDTH, 0, tid);
        }
        stage.run([&] { gemm_act_wT<16>(tmem_base, HA, 0, Operand{sbase + L.wd[DH], MLP_WIDTH}, MLP_WIDTH); });
        float dout[16];
        tmem_ld16(tmem_row, dout); tmem_wait_ld();
        if (MODE == MODE_DENSITY) { p.out[row_g] = __float2half_rn(dout[0]); continue; }
        if (MODE == MODE_PLAIN) {
            uint4* o = reinterpret_cast<uint4*>(p.out + (size_t)row_g * p.out_stride);
            o[0] = pack8(dout, false); o[1] = pack8(dout + 8, false);
            continue;
        }
        // ---- rgb network: input = [density out (16) | SH4(dir) (16)] ----
        const __half sigma = __float2half_rn(dout[0]);
        *reinterpret_cast<uint4*>(r0 + blk8_chunk_offset(tid, 0, 32)) = pack8(dout, false);
        *reinterpret_cast<uint4*>(r0 + blk8_chunk_offset(tid, 1, 32)) = pack8(dout + 8, false);
        float sh[16]; sh4(dir[0], dir[1], dir[2], sh);
        *reinterpret_cast<uint4*>(r0 + blk8_chunk_offset(tid, 2, 32)) = pack8(sh, false);
        *reinterpret_cast<uint4*>(r0 + blk8_chunk_offset(tid, 3, 32)) = pack8(sh + 8, false);
        stage.run([&] { gemm_act_wT<64>(tmem_base, R0, 0, Operand{sbase + L.wr[0], 32}, 32); });
        epilogue_to_smem<64, true, 16>(tmem_row, ha, MLP_WIDTH, 0, tid);
        for (uint32_t l = 1; l < RH; ++l) {
            stage.run([&] { gemm_act_wT<64>(tmem_base, HA, 0, Operand{sbase + L.wr[l], MLP_WIDTH}, MLP_WIDTH); });
            epilogue_to_smem<64, true, 16>(tmem_row, ha, MLP_WIDTH, 0, tid);
        }
        stage.run([&] { gemm_act_wT<16>(tmem_base, HA, 0, Operand{sbase + L.wr[RH], MLP_WIDTH}, MLP_WIDTH); });
        float rgb[16];
        tmem_ld16(tmem_row, rgb); tmem_wait_ld();
        __half2 o2[2] = {__floats2half2_rn(rgb[0], rgb[1]), __halves2half2(__float2half_rn(rgb[2]), sigma)};
        *reinterpret_cast<uint2*>(p.out + (size_t)row_g * 4) = *reinterpret_cast<uint2*>(o2);
    }
    // ---- teardown ----
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc<64>(tmem_base);
}


// -----------------------------------------------------------------------------------------------------------------
// Forward, "TS" form (round 2): hidden activations never touch shared memory.  The epilogue of a layer reads its TMEM
// accumulator row (tcgen05.ld), applies ReLU, rounds to fp16 and writes the row straight back to TMEM (tcgen05.st) as
// the A operand of the next layer's MMA, which is issued in the A-from-TMEM form.  Measured on B200
// (tools/ubench_tc*.cu, profiles/r02_ubench_tc.txt): a K=16 M128xN64 instruction costs 50 cycles with A in shared memory
// (its 4 KB A read at 128 B/cycle is the bound) and 33.6 with A in TMEM; N=16: 40.5 vs 10.6; the 128x64 epilogue is 148
// cycles into TMEM against 322 into shared memory.  Per 128-sample tile of the NeRF network the tensor pipe is busy
// 720 cycles in the SS form and 354 in this one.  Only the first layer (input tile delivered by TMA) reads A from smem.
// TMEM per CTA: 64 accumulator columns + 32 operand columns (128 allocated) -> 4 CTAs per SM.
// -----------------------------------------------------------------------------------------------------------------
template <uint32_t PW, uint32_t DH, uint32_t RH, int MODE>
struct FwdTsSmem {
    uint32_t wd[DH + 1], wr[RH + 1], x0, bars, total;
    __host__ __device__ FwdTsSmem() {
        SmemCursor c;
        wd[0] = c.take(MLP_WIDTH * PW * 2);
        for (uint32_t l = 1; l < DH; ++l) wd[l] = c.take(MLP_WIDTH * MLP_WIDTH * 2);
        wd[DH] = c.take(MLP_OUT_PAD * MLP_WIDTH * 2);
        if (MODE == MODE_NERF) {
            wr[0] = c.take(MLP_WIDTH * 32 * 2);
            for (uint32_t l = 1; l < RH; ++l) wr[l] = c.take(MLP_WIDTH * MLP_WIDTH * 2);
            wr[RH] = c.take(MLP_OUT_PAD * MLP_WIDTH * 2);
        } else { for (uint32_t l = 0; l <= RH; ++l) wr[l] = 0; }
        x0 = c.take(TILE_ROWS * PW * 2);
        bars = c.take(64);
        total = c.off + 128;
    }
};
constexpr uint32_t TS_ACC = 0, TS_OPA = 64, TS_COLS = 128;   // TMEM columns: fp32 accumulator, packed-fp16 A operand, allocation

// accumulator row [64 fp32] -> ReLU -> fp16 pairs -> the A-operand columns of the same lane
NGP_D void epilogue_to_tmem64(uint32_t trow) {
    float v[64];
#pragma unroll
    for (uint32_t j = 0; j < 64; j += 16) tmem_ld16(trow + TS_ACC + j, v + j);
    tmem_wait_ld();
    uint32_t pk[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) asm("cvt.rn.relu.f16x2.f32 %0, %1, %2;" : "=r"(pk[i]) : "f"(v[2 * i + 1]), "f"(v[2 * i]));
    tmem_st32(trow + TS_OPA, pk);
    tmem_wait_st();
}

// One MMA stage of the TS kernels: every thread's TMEM stores / loads are complete (wait::st / wait::ld done by the caller),
// sync the CTA, thread 0 issues and commits, everyone waits for the MMAs.
struct StageTs {
    uint64_t* bar; uint32_t parity;
    template <class F>
    NGP_D void run(F&& issue) {
        fence_before_sync();
        __syncthreads();
        if (threadIdx.x == 0) {
            fence_after_sync();
            issue();
            umma_commit(bar);
        }
        mbar_wait(bar, parity);
        parity ^= 1u;
        fence_after_sync();
    }
};

template <uint32_t PW, uint32_t DH, uint32_t RH, int MODE>
__global__ void __launch_bounds__(128) mlp_forward_ts_kernel(const FwdParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~(uintptr_t)127);
    const FwdTsSmem<PW, DH, RH, MODE> L;
    const uint32_t tid = threadIdx.x, warp = tid >> 5;
    uint64_t* bar_full = reinterpret_cast<uint64_t*>(smem + L.bars);
    uint64_t* bar_mma = bar_full + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_full + 2);

    if (tid == 0) { mbar_init(bar_full, 1); mbar_init(bar_mma, 1); mbar_fence_init(); }
    if (warp == 0) tmem_alloc<TS_COLS>(tmem_slot);
    load_all_weights<PW, DH, RH, MODE == MODE_NERF, 128>(smem, L.wd, L.wr, p.w_density, p.w_rgb, tid);
    fence_proxy_async();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t trow = tmem_base + ((warp * 32u) << 16);   // this warp's lane quadrant, column 0
    const uint32_t acc = tmem_base + TS_ACC, opa = tmem_base + TS_OPA;
    const uint32_t sbase = smem_u32(smem);
    const Operand X0 = {sbase + L.x0, PW};
    constexpr uint32_t X_BYTES = TILE_ROWS * PW * 2;

    // n_dyn: the row count lives in device memory (tracer: rays alive x steps, known only after the compaction kernel has run)
    const uint32_t n_tiles = p.n_dyn ? min(p.n_tiles, (__ldg(p.n_dyn) + TILE_ROWS - 1) / TILE_ROWS) : p.n_tiles;
    StageTs stage{bar_mma, 0};
    uint32_t full_parity = 0;
    if (tid == 0 && blockIdx.x < n_tiles) {
        mbar_expect_tx(bar_full, X_BYTES);
        bulk_g2s(smem + L.x0, p.enc + (size_t)blockIdx.x * TILE_ROWS * PW, X_BYTES, bar_full);
    }
    const bool probing = p.probe != nullptr && blockIdx.x == 0 && tid == 0;
    long long tp = 0; uint32_t ph = 0;
#define PROBE() do { if (probing) { const long long now = clock64(); p.probe[ph++] += now - tp; tp = now; } } while (0)
    for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        if (probing) { tp = clock64(); ph = 0; p.probe[15] += 1; }
        const uint32_t next = tile + gridDim.x;
        const uint32_t row_g = tile * TILE_ROWS + tid;
        uint4 sh_lo = make_uint4(0, 0, 0, 0), sh_hi = sh_lo;   // SH4 of the direction, evaluated while the input tile is still in flight
        if (MODE == MODE_NERF) {
            const float* c = p.coords + (size_t)row_g * 7 + 4;
            float sh[16]; sh4(__ldg(c), __ldg(c + 1), __ldg(c + 2), sh);
            sh_lo = pack8(sh, false); sh_hi = pack8(sh + 8, false);
        }
        mbar_wait(bar_full, full_parity); full_parity ^= 1u;
        PROBE();   // 0: wait for the input tile

        // ---- density / plain network: first layer reads the TMA-delivered input tile from shared memory ----
        stage.run([&] { gemm_act_wT<64>(acc, X0, 0, Operand{sbase + L.wd[0], PW}, PW); });
        if (tid == 0 && next < n_tiles) {   // the input tile is free: fetch the next one under the remaining layers
            mbar_expect_tx(bar_full, X_BYTES);
            bulk_g2s(smem + L.x0, p.enc + (size_t)next * TILE_ROWS * PW, X_BYTES, bar_full);
        }
        PROBE();   // 1: layer-0 stage
        epilogue_to_tmem64(trow);
        PROBE();   // 2: epilogue
        for (uint32_t l = 1; l < DH; ++l) {
            stage.run([&] { gemm_tmem_wT<64>(acc, opa, Operand{sbase + L.wd[l], MLP_WIDTH}, MLP_WIDTH); });
            epilogue_to_tmem64(trow);
        }
        stage.run([&] { gemm_tmem_wT<16>(acc, opa, Operand{sbase + L.wd[DH], MLP_WIDTH}, MLP_WIDTH); });
        float dout[16];
        tmem_ld16(trow + TS_ACC, dout); tmem_wait_ld();
        PROBE();   // 3: density output stage + ld
        if (MODE == MODE_DENSITY) { p.out[row_g] = __float2half_rn(dout[0]); continue; }
        if (MODE == MODE_PLAIN) {
            uint4* o = reinterpret_cast<uint4*>(p.out + (size_t)row_g * p.out_stride);
            o[0] = pack8(dout, false); o[1] = pack8(dout + 8, false);
            continue;
        }
        // ---- rgb network: input row = [density out (16) | SH4(dir) (16)] as 16 packed columns ----
        const __half sigma = __float2half_rn(dout[0]);
        {
            uint32_t pk[16];
            const uint4 a0 = pack8(dout, false), a1 = pack8(dout + 8, false), a2 = sh_lo, a3 = sh_hi;
            pk[0] = a0.x; pk[1] = a0.y; pk[2] = a0.z; pk[3] = a0.w; pk[4] = a1.x; pk[5] = a1.y; pk[6] = a1.z; pk[7] = a1.w;
            pk[8] = a2.x; pk[9] = a2.y; pk[10] = a2.z; pk[11] = a2.w; pk[12] = a3.x; pk[13] = a3.y; pk[14] = a3.z; pk[15] = a3.w;
            tmem_st16(trow + TS_OPA, pk);
            tmem_wait_st();
        }
        PROBE();   // 4: SH + rgb input row
        stage.run([&] { gemm_tmem_wT<64>(acc, opa, Operand{sbase + L.wr[0], 32}, 32); });
        PROBE();   // 5: rgb layer-0 stage
        epilogue_to_tmem64(trow);
        PROBE();   // 6: epilogue
        for (uint32_t l = 1; l < RH; ++l) {
            stage.run([&] { gemm_tmem_wT<64>(acc, opa, Operand{sbase + L.wr[l], MLP_WIDTH}, MLP_WIDTH); });
            epilogue_to_tmem64(trow);
        }
        PROBE();   // 7: rgb hidden stages + epilogues
        stage.run([&] { gemm_tmem_wT<16>(acc, opa, Operand{sbase + L.wr[RH], MLP_WIDTH}, MLP_WIDTH); });
        float rgb[16];
        tmem_ld16(trow + TS_ACC, rgb); tmem_wait_ld();
        PROBE();   // 8: rgb output stage + ld
        __half2 o2[2] = {__floats2half2_rn(rgb[0], rgb[1]), __halves2half2(__float2half_rn(rgb[2]), sigma)};
        *reinterpret_cast<uint2*>(p.out + (size_t)row_g * 4) = *reinterpret_cast<uint2*>(o2);
        PROBE();   // 9: output store
    }
#undef PROBE
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc<TS_COLS>(tmem_base);
}


// =================================================================================================================
// Backward: forward recompute (activations stay in shared memory) + dgrad chain + wgrad, all on tcgen05.
//   dgrad  D[128 x in]  = G[128 x out] * W[out x in]      (A = G K-major, B = W MN-major)
//   wgrad  dW[out x in] += G^T[out x 128] * In[128 x in]  (A = G MN-major, M = 64; B = In MN-major); output layers are
//          accumulated transposed (dW^T[64 x 16] = In^T * G) because M = 16 is not a legal UMMA shape.
// wgrad accumulators stay in TMEM for the CTA's lifetime and are flushed once with fp32 atomics.
// =================================================================================================================
struct BwdParams {
    const __half* w_density; const __half* w_rgb;
    const __half* enc; const float* coords;
    const __half* dL_dout; uint32_t dout_stride;
    __half* dL_denc;      // nullable
    float* wgrad;         // [n_density_params + n_rgb_params] fp32, atomically accumulated
    uint32_t n_tiles;
    float* dL_dcoords;    // nullable, NeRF mode: [n][7] fp32 gradient of the loss w.r.t. the network input rows; this kernel writes the direction part (cols 4..6)
    unsigned long long* probe;   // debug (NGP_MLP_PROBE=1): per CTA [4] = %globaltimer at start / after the prologue / after the tile loop / at exit; null otherwise
};

template <uint32_t PW, uint32_t DH, uint32_t RH, int MODE, bool TS = false>
struct BwdCfg {
    static constexpr uint32_t ACC_D0 = 64;                                  // density layer-0 wgrad [64 x PW]
    static constexpr uint32_t ACC_DH = ACC_D0 + PW;                         // density hidden wgrads, 64 cols each
    static constexpr uint32_t ACC_DO = ACC_DH + 64 * (DH - 1);              // density output wgrad^T [64 x 16]
    static constexpr uint32_t ACC_R0 = ACC_DO + 16;                         // rgb layer-0 wgrad [64 x 32]
    static constexpr uint32_t ACC_RH = ACC_R0 + (MODE == MODE_NERF ? 32 : 0);
    static constexpr uint32_t ACC_RO = ACC_RH + (MODE == MODE_NERF ? 64 * (RH - 1) : 0);
    static constexpr uint32_t ACC_END = ACC_RO + (MODE == MODE_NERF ? 16 : 0);
    static constexpr uint32_t OPA = ACC_END;                                // TS form: 32 columns of packed fp16 pairs, the A operand of the next MMA
    static constexpr uint32_t COLS_USED = ACC_END + (TS ? 32 : 0);
    static constexpr uint32_t TMEM_COLS = COLS_USED <= 128 ? 128 : (COLS_USED <= 256 ? 256 : 512);
    static_assert(COLS_USED <= 512, "network too large for the TMEM-resident wgrad accumulators");
    uint32_t wd[DH + 1], wr[RH + 1], x0, hd[DH], hr[RH], r0, g, go, bars, total;
    __host__ __device__ BwdCfg() {
        SmemCursor c;
        wd[0] = c.take(MLP_WIDTH * PW * 2);
        for (uint32_t l = 1; l < DH; ++l) wd[l] = c.take(MLP_WIDTH * MLP_WIDTH * 2);
        wd[DH] = c.take(MLP_OUT_PAD * MLP_WIDTH * 2);
        for (uint32_t l = 0; l <= RH; ++l) wr[l] = 0;
        for (uint32_t l = 0; l < RH; ++l) hr[l] = 0;
        r0 = 0;
        if (MODE == MODE_NERF) {
            wr[0] = c.take(MLP_WIDTH * 32 * 2);
            for (uint32_t l = 1; l < RH; ++l) wr[l] = c.take(MLP_WIDTH * MLP_WIDTH * 2);
            wr[RH] = c.take(MLP_OUT_PAD * MLP_WIDTH * 2);
            r0 = c.take(TILE_ROWS * 32 * 2);
            for (uint32_t l = 0; l < RH; ++l) hr[l] = c.take(TILE_ROWS * MLP_WIDTH * 2);
        }
        x0 = c.take(TILE_ROWS * PW * 2);
        for (uint32_t l = 0; l < DH; ++l) hd[l] = c.take(TILE_ROWS * MLP_WIDTH * 2);
        g = c.take(TILE_ROWS * MLP_WIDTH * 2);
        go = c.take(TILE_ROWS * 16 * 2);
        bars = c.take(64);
        total = c.off + 128;
    }
};

// TMEM [128 x 64] dgrad result -> ReLU-masked by the saved activation tile -> fp16 G tile (in place over the old G)
NGP_D void epilogue_masked(uint32_t tmem_row_addr, __half* g, const __half* act, uint32_t row) {
    float v[64];   // the whole row: every tcgen05.ld is issued before the single wait
#pragma unroll
    for (uint32_t j = 0; j < 64; j += 16) tmem_ld16(tmem_row_addr + j, v + j);
    uint4 a[8];    // the saved activations of the row load meanwhile
#pragma unroll
    for (uint32_t c = 0; c < 8; ++c) a[c] = *reinterpret_cast<const uint4*>(act + blk8_chunk_offset(row, c, MLP_WIDTH));
    tmem_wait_ld();
#pragma unroll
    for (uint32_t c = 0; c < 8; ++c) {
        uint4 packed = pack8(v + c * 8, false);
        // ReLU': keep the gradient where the activation is > 0 (fp16 compare on the packed pairs, mask applied with AND)
        const __half2* ah = reinterpret_cast<const __half2*>(&a[c]);
        uint32_t* pk = reinterpret_cast<uint32_t*>(&packed);
        const __half2 zero = __float2half2_rn(0.0f);
#pragma unroll
        for (int i = 0; i < 4; ++i) pk[i] &= __hgt2_mask(ah[i], zero);
        *reinterpret_cast<uint4*>(g + blk8_chunk_offset(row, c, MLP_WIDTH)) = packed;
    }
}

// ---- "TS" form of the backward kernel (round 2) -------------------------------------------------------------------
// Every [128 x 64] fp16 tile that is the A operand of the NEXT MMA (forward recompute: the activations; dgrad: the masked
// gradient G) is written to tensor memory as well as to shared memory, and that MMA reads it from there (tcgen05.mma with
// A in TMEM): its shared-memory traffic is the B operand alone.  The shared-memory copy stays because the weight-gradient
// MMAs read both G and the activations MN-major, which only shared memory can serve.  With KEEP the packed activations of
// the forward recompute also stay in registers until the layer's ReLU' mask has been applied, so the masked epilogue does
// not read them back from shared memory.  Per 128-sample tile of the NeRF network: 414 -> 286 KB of shared-memory traffic.
template <bool ST, bool KEEP>
NGP_D void epilogue_fwd_ts(uint32_t tmem_row_addr, uint32_t opa_row, __half* dst, uint32_t row, uint32_t* keep) {
    float v[64];
#pragma unroll
    for (uint32_t j = 0; j < 64; j += 16) tmem_ld16(tmem_row_addr + j, v + j);
    tmem_wait_ld();
    uint32_t pk[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) asm("cvt.rn.relu.f16x2.f32 %0, %1, %2;" : "=r"(pk[i]) : "f"(v[2 * i + 1]), "f"(v[2 * i]));
    if (ST) tmem_st32(opa_row, pk);
#pragma unroll
    for (uint32_t c = 0; c < 8; ++c)
        *reinterpret_cast<uint4*>(dst + blk8_chunk_offset(row, c, MLP_WIDTH)) = make_uint4(pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
    if (KEEP) {
#pragma unroll
        for (int i = 0; i < 32; ++i) keep[i] = pk[i];
    }
    if (ST) tmem_wait_st();
}
template <bool KEEP>
NGP_D void epilogue_masked_ts(uint32_t tmem_row_addr, uint32_t opa_row, __half* g, const __half* act, const uint32_t* keep, uint32_t row) {
    float v[64];
#pragma unroll
    for (uint32_t j = 0; j < 64; j += 16) tmem_ld16(tmem_row_addr + j, v + j);
    uint32_t a[32];
    if (KEEP) {
#pragma unroll
        for (int i = 0; i < 32; ++i) a[i] = keep[i];
    } else {
#pragma unroll
        for (uint32_t c = 0; c < 8; ++c) {
            const uint4 t = *reinterpret_cast<const uint4*>(act + blk8_chunk_offset(row, c, MLP_WIDTH));
            a[4 * c] = t.x; a[4 * c + 1] = t.y; a[4 * c + 2] = t.z; a[4 * c + 3] = t.w;
        }
    }
    tmem_wait_ld();
    uint32_t pk[32];
    const __half2 zero = __float2half2_rn(0.0f);
#pragma unroll
    for (int i = 0; i < 32; ++i) {
        asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(pk[i]) : "f"(v[2 * i + 1]), "f"(v[2 * i]));
        pk[i] &= __hgt2_mask(*reinterpret_cast<const __half2*>(&a[i]), zero);
    }
    tmem_st32(opa_row, pk);
#pragma unroll
    for (uint32_t c = 0; c < 8; ++c)
        *reinterpret_cast<uint4*>(g + blk8_chunk_offset(row, c, MLP_WIDTH)) = make_uint4(pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
    tmem_wait_st();
}

// Backward through one chain (H hidden layers, input tile IN of CIN columns).  On return, if NEED_DIN, the TMEM work
// accumulator (cols [0, CIN)) holds dL/dIN for this tile.
template <uint32_t H, uint32_t CIN, bool NEED_DIN, bool TS = false, bool KEEP = false>
NGP_D void bwd_chain(StageSync& stage, uint32_t tmem_base, uint32_t tmem_row, uint32_t sbase, uint8_t* smem, uint32_t acc0, uint32_t acch,
                     uint32_t acco, uint32_t in_off, const uint32_t* act_off, uint32_t g_off, uint32_t go_off, const uint32_t* w_off,
                     uint32_t accumulate, uint32_t tid, uint32_t opa_col = 0, const uint32_t* keep = nullptr) {
    const Operand G = {sbase + g_off, MLP_WIDTH}, GO = {sbase + go_off, 16}, IN = {sbase + in_off, CIN};
    __half* g = reinterpret_cast<__half*>(smem + g_off);
    const uint32_t opa = tmem_base + opa_col, opa_row = tmem_row + opa_col;
    {   // output layer: wgrad^T [64 x 16] += act[H-1]^T * GO ; dgrad [128 x 64] = GO * W_out
        const Operand A = {sbase + act_off[H - 1], MLP_WIDTH}, W = {sbase + w_off[H], MLP_WIDTH};
        stage.run([&] { gemm_pT_q<16>(tmem_base + acco, A, 0, GO, 0, accumulate); gemm_act_w<64>(tmem_base, GO, W, 16); });
        if (TS) epilogue_masked_ts<KEEP>(tmem_row, opa_row, g, reinterpret_cast<const __half*>(smem + act_off[H - 1]), keep + (H - 1) * 32, tid);
        else epilogue_masked(tmem_row, g, reinterpret_cast<const __half*>(smem + act_off[H - 1]), tid);
    }
#pragma unroll
    for (uint32_t l = H - 1; l >= 1; --l) {   // hidden layer l: wgrad [64 x 64] += G^T * act[l-1] ; dgrad = G * W_l
        const Operand A = {sbase + act_off[l - 1], MLP_WIDTH}, W = {sbase + w_off[l], MLP_WIDTH};
        stage.run([&] {
            gemm_pT_q<64>(tmem_base + acch + 64 * (l - 1), G, 0, A, 0, accumulate);
            if (TS) gemm_tmem_w<64>(tmem_base, opa, W, 64); else gemm_act_w<64>(tmem_base, G, W, 64);
        });
        if (TS) epilogue_masked_ts<KEEP>(tmem_row, opa_row, g, reinterpret_cast<const __half*>(smem + act_off[l - 1]), keep + (l - 1) * 32, tid);
        else epilogue_masked(tmem_row, g, reinterpret_cast<const __half*>(smem + act_off[l - 1]), tid);
    }
    {   // first layer: wgrad [64 x CIN] += G^T * IN ; dgrad [128 x CIN] = G * W_0
        const Operand W = {sbase + w_off[0], CIN};
        stage.run([&] {
            gemm_pT_q<CIN>(tmem_base + acc0, G, 0, IN, 0, accumulate);
            if (NEED_DIN) { if (TS) gemm_tmem_w<CIN>(tmem_base, opa, W, 64); else gemm_act_w<CIN>(tmem_base, G, W, 64); }
        });
    }
}

// flush one TMEM wgrad accumulator (M = 64 layout: row m lives in lane 32*(m/16) + m%16) to global with fp32 atomics
NGP_D void red_add_f32x4(float* p, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
template <uint32_t N, bool TRANSPOSED>
NGP_D void flush_wgrad(uint32_t tmem_row, uint32_t col, float* dst, uint32_t ld, uint32_t tid) {
    const uint32_t lane = tid & 31, m = (tid >> 5) * 16 + lane;
    float v[N];
#pragma unroll
    for (uint32_t j = 0; j < N; j += 16) tmem_ld16(tmem_row + col + j, v + j);
    tmem_wait_ld();
    if (lane < 16) {
        if (TRANSPOSED) {
#pragma unroll
            for (uint32_t i = 0; i < N; ++i) atomicAdd(dst + (size_t)i * ld + m, v[i]);
        } else {      // row m of the gradient is contiguous: 16-byte vector reductions, a quarter of the atomic traffic
#pragma unroll
            for (uint32_t i = 0; i < N; i += 4) red_add_f32x4(dst + (size_t)m * ld + i, v[i], v[i + 1], v[i + 2], v[i + 3]);
        }
    }
}

template <uint32_t PW, uint32_t DH, uint32_t RH, int MODE, bool NEED_DENC, bool TS = false, bool KEEP = false>
__global__ void __launch_bounds__(128) mlp_backward_kernel(const BwdParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~(uintptr_t)127);
    using Cfg = BwdCfg<PW, DH, RH, MODE, TS>;
    const Cfg L;
    const uint32_t tid = threadIdx.x, warp = tid >> 5;
    uint64_t* bar_full = reinterpret_cast<uint64_t*>(smem + L.bars);
    uint64_t* bar_mma = bar_full + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_full + 2);
    auto stamp = [&](uint32_t i) { if (p.probe && tid == 0) { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); p.probe[(size_t)blockIdx.x * 4 + i] = t; } };
    stamp(0);

    if (tid == 0) { mbar_init(bar_full, 1); mbar_init(bar_mma, 1); mbar_fence_init(); }
    if (warp == 0) tmem_alloc<Cfg::TMEM_COLS>(tmem_slot);
    load_all_weights<PW, DH, RH, MODE == MODE_NERF, 128>(smem, L.wd, L.wr, p.w_density, p.w_rgb, tid);
    fence_proxy_async();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t tmem_row = tmem_base + ((warp * 32u) << 16);
    const uint32_t opa = tmem_base + Cfg::OPA, opa_row = tmem_row + Cfg::OPA;   // TS form: the A-operand columns
    const uint32_t sbase = smem_u32(smem);
    constexpr uint32_t X_BYTES = TILE_ROWS * PW * 2;
    __half* r0 = reinterpret_cast<__half*>(smem + L.r0);
    __half* go = reinterpret_cast<__half*>(smem + L.go);
    uint32_t keep_d[KEEP ? DH * 32 : 1], keep_r[KEEP ? RH * 32 : 1];            // KEEP: the packed activations of the forward recompute

    stamp(1);
    StageSync stage{bar_mma, 0};
    uint32_t full_parity = 0, it = 0;
    if (tid == 0 && blockIdx.x < p.n_tiles) {
        mbar_expect_tx(bar_full, X_BYTES);
        bulk_g2s(smem + L.x0, p.enc + (size_t)blockIdx.x * TILE_ROWS * PW, X_BYTES, bar_full);
    }
    for (uint32_t tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++it) {
        const uint32_t row_g = tile * TILE_ROWS + tid;
        const uint32_t accumulate = it > 0;
        float dir[3] = {0.f, 0.f, 0.f};
        if (MODE == MODE_NERF) { const float* c = p.coords + (size_t)row_g * 7 + 4; dir[0] = __ldg(c); dir[1] = __ldg(c + 1); dir[2] = __ldg(c + 2); }
        mbar_wait(bar_full, full_parity); full_parity ^= 1u;

        // ---- forward recompute: density / plain chain ----
        stage.run([&] { gemm_act_wT<64>(tmem_base, Operand{sbase + L.x0, PW}, 0, Operand{sbase + L.wd[0], PW}, PW); });
        // TS: a layer's activations also go to the TMEM operand columns when a forward layer follows (its A operand)
        if (TS) { if (DH > 1 || MODE == MODE_NERF) epilogue_fwd_ts<true, KEEP>(tmem_row, opa_row, reinterpret_cast<__half*>(smem + L.hd[0]), tid, keep_d);
                  else epilogue_fwd_ts<false, KEEP>(tmem_row, opa_row, reinterpret_cast<__half*>(smem + L.hd[0]), tid, keep_d); }
        else epilogue_to_smem<64, true>(tmem_row, reinterpret_cast<__half*>(smem + L.hd[0]), MLP_WIDTH, 0, tid);
#pragma unroll
        for (uint32_t l = 1; l < DH; ++l) {
            stage.run([&] {
                if (TS) gemm_tmem_wT<64>(tmem_base, opa, Operand{sbase + L.wd[l], MLP_WIDTH}, MLP_WIDTH);
                else gemm_act_wT<64>(tmem_base, Operand{sbase + L.hd[l - 1], MLP_WIDTH}, 0, Operand{sbase + L.wd[l], MLP_WIDTH}, MLP_WIDTH);
            });
            if (TS) { if (l + 1 < DH || MODE == MODE_NERF) epilogue_fwd_ts<true, KEEP>(tmem_row, opa_row, reinterpret_cast<__half*>(smem + L.hd[l]), tid, keep_d + (KEEP ? l * 32 : 0));
                      else epilogue_fwd_ts<false, KEEP>(tmem_row, opa_row, reinterpret_cast<__half*>(smem + L.hd[l]), tid, keep_d + (KEEP ? l * 32 : 0)); }
            else epilogue_to_smem<64, true>(tmem_row, reinterpret_cast<__half*>(smem + L.hd[l]), MLP_WIDTH, 0, tid);
        }
        if (MODE == MODE_NERF) {
            stage.run([&] {
                if (TS) gemm_tmem_wT<16>(tmem_base, opa, Operand{sbase + L.wd[DH], MLP_WIDTH}, MLP_WIDTH);
                else gemm_act_wT<16>(tmem_base, Operand{sbase + L.hd[DH - 1], MLP_WIDTH}, 0, Operand{sbase + L.wd[DH], MLP_WIDTH}, MLP_WIDTH);
            });
            float dout[16];
            tmem_ld16(tmem_row, dout); tmem_wait_ld();
            float sh[16]; sh4(dir[0], dir[1], dir[2], sh);
            const uint4 q0 = pack8(dout, false), q1 = pack8(dout + 8, false), q2 = pack8(sh, false), q3 = pack8(sh + 8, false);
            if (TS) {
                const uint32_t pk[16] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, q3.x, q3.y, q3.z, q3.w};
                tmem_st16(opa_row, pk);
            }
            *reinterpret_cast<uint4*>(r0 + blk8_chunk_offset(tid, 0, 32)) = q0;
            *reinterpret_cast<uint4*>(r0 + blk8_chunk_offset(tid, 1, 32)) = q1;
            *reinterpret_cast<uint4*>(r0 + blk8_chunk_offset(tid, 2, 32)) = q2;
            *reinterpret_cast<uint4*>(r0 + blk8_chunk_offset(tid, 3, 32)) = q3;
            if (TS) tmem_wait_st();
            // ---- forward recompute: rgb hidden layers (the rgb output itself is not needed) ----
            stage.run([&] {
                if (TS) gemm_tmem_wT<64>(tmem_base, opa, Operand{sbase + L.wr[0], 32}, 32);
                else gemm_act_wT<64>(tmem_base, Operand{sbase + L.r0, 32}, 0, Operand{sbase + L.wr[0], 32}, 32);
            });
            if (TS) { if (RH > 1) epilogue_fwd_ts<true, KEEP>(tmem_row, opa_row, reinterpret_cast<__half*>(smem + L.hr[0]), tid, keep_r);
                      else epilogue_fwd_ts<false, KEEP>(tmem_row, opa_row, reinterpret_cast<__half*>(smem + L.hr[0]), tid, keep_r); }
            else epilogue_to_smem<64, true>(tmem_row, reinterpret_cast<__half*>(smem + L.hr[0]), MLP_WIDTH, 0, tid);
#pragma unroll
            for (uint32_t l = 1; l < RH; ++l) {
                stage.run([&] {
                    if (TS) gemm_tmem_wT<64>(tmem_base, opa, Operand{sbase + L.wr[l], MLP_WIDTH}, MLP_WIDTH);
                    else gemm_act_wT<64>(tmem_base, Operand{sbase + L.hr[l - 1], MLP_WIDTH}, 0, Operand{sbase + L.wr[l], MLP_WIDTH}, MLP_WIDTH);
                });
                if (TS) { if (l + 1 < RH) epilogue_fwd_ts<true, KEEP>(tmem_row, opa_row, reinterpret_cast<__half*>(smem + L.hr[l]), tid, keep_r + (KEEP ? l * 32 : 0));
                          else epilogue_fwd_ts<false, KEEP>(tmem_row, opa_row, reinterpret_cast<__half*>(smem + L.hr[l]), tid, keep_r + (KEEP ? l * 32 : 0)); }
                else epilogue_to_smem<64, true>(tmem_row, reinterpret_cast<__half*>(smem + L.hr[l]), MLP_WIDTH, 0, tid);
            }
            // ---- dL/d(rgb out): extract_rgb (nerf_network.h:45-60) — only channels 0..2 feed the rgb net ----
            const uint2 graw = __ldg(reinterpret_cast<const uint2*>(p.dL_dout + (size_t)row_g * 4));
            const __half* gh = reinterpret_cast<const __half*>(&graw);
            {
                __half z = __float2half_rn(0.0f);
                __half2 c0[4] = {__halves2half2(gh[0], gh[1]), __halves2half2(gh[2], z), __halves2half2(z, z), __halves2half2(z, z)};
                *reinterpret_cast<uint4*>(go + blk8_chunk_offset(tid, 0, 16)) = *reinterpret_cast<uint4*>(c0);
                *reinterpret_cast<uint4*>(go + blk8_chunk_offset(tid, 1, 16)) = make_uint4(0, 0, 0, 0);
            }
            bwd_chain<RH, 32, true, TS, KEEP>(stage, tmem_base, tmem_row, sbase, smem, Cfg::ACC_R0, Cfg::ACC_RH, Cfg::ACC_RO, L.r0, L.hr, L.g, L.go, L.wr, accumulate, tid, Cfg::OPA, keep_r);
            if (p.dL_dcoords) {   // d(rgb input)[16:32] = dL/d(SH4(dir)), stored in network precision by the reference, -> dL/d(dir)
                float gsh[16];
                tmem_ld16(tmem_row + 16, gsh); tmem_wait_ld();
#pragma unroll
                for (int k = 0; k < 16; ++k) gsh[k] = __half2float(__float2half_rn(gsh[k]));
                float gd[3]; sh4_input_gradient(dir[0], dir[1], dir[2], gsh, gd);
                float* dst = p.dL_dcoords + (size_t)row_g * 7 + 4;
                dst[0] = gd[0]; dst[1] = gd[1]; dst[2] = gd[2];
            }
            // ---- d(density out) = d(rgb input)[0:16], + dL/dsigma on channel 0: add_density_gradient (nerf_network.h:62-74) ----
            float dd[16];
            tmem_ld16(tmem_row, dd); tmem_wait_ld();
            dd[0] = __half2float(__float2half_rn(dd[0])) + __half2float(gh[3]);
            *reinterpret_cast<uint4*>(go + blk8_chunk_offset(tid, 0, 16)) = pack8(dd, false);
            *reinterpret_cast<uint4*>(go + blk8_chunk_offset(tid, 1, 16)) = pack8(dd + 8, false);
        } else {
            const uint4* gsrc = reinterpret_cast<const uint4*>(p.dL_dout + (size_t)row_g * p.dout_stride);
            *reinterpret_cast<uint4*>(go + blk8_chunk_offset(tid, 0, 16)) = __ldg(gsrc);
            *reinterpret_cast<uint4*>(go + blk8_chunk_offset(tid, 1, 16)) = __ldg(gsrc + 1);
        }
        bwd_chain<DH, PW, NEED_DENC, TS, KEEP>(stage, tmem_base, tmem_row, sbase, smem, Cfg::ACC_D0, Cfg::ACC_DH, Cfg::ACC_DO, L.x0, L.hd, L.g, L.go, L.wd, accumulate, tid, Cfg::OPA, keep_d);
        // X0 is free now: prefetch the next tile's input while the last epilogue runs
        const uint32_t next = tile + gridDim.x;
        if (tid == 0 && next < p.n_tiles) {
            mbar_expect_tx(bar_full, X_BYTES);
            bulk_g2s(smem + L.x0, p.enc + (size_t)next * TILE_ROWS * PW, X_BYTES, bar_full);
        }
        if (NEED_DENC) {
#pragma unroll
            for (uint32_t j = 0; j < PW; j += 16) {
                float v[16];
                tmem_ld16(tmem_row + j, v); tmem_wait_ld();
                *reinterpret_cast<uint4*>(p.dL_denc + blk8_chunk_offset(row_g, j >> 3, PW)) = pack8(v, false);
                *reinterpret_cast<uint4*>(p.dL_denc + blk8_chunk_offset(row_g, (j >> 3) + 1, PW)) = pack8(v + 8, false);
            }
        }
    }
    // ---- flush the TMEM-resident weight gradients ----
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    stamp(2);
    if (it > 0) {
        float* wg = p.wgrad;
        flush_wgrad<PW, false>(tmem_row, Cfg::ACC_D0, wg, PW, tid); wg += MLP_WIDTH * PW;
        for (uint32_t l = 1; l < DH; ++l) { flush_wgrad<64, false>(tmem_row, Cfg::ACC_DH + 64 * (l - 1), wg, 64, tid); wg += 64 * 64; }
        flush_wgrad<16, true>(tmem_row, Cfg::ACC_DO, wg, 64, tid); wg += 16 * 64;
        if (MODE == MODE_NERF) {
            flush_wgrad<32, false>(tmem_row, Cfg::ACC_R0, wg, 32, tid); wg += 64 * 32;
            for (uint32_t l = 1; l < RH; ++l) { flush_wgrad<64, false>(tmem_row, Cfg::ACC_RH + 64 * (l - 1), wg, 64, tid); wg += 64 * 64; }
            flush_wgrad<16, true>(tmem_row, Cfg::ACC_RO, wg, 64, tid);
        }
    }
    fence_before_sync();
    __syncthreads();
    stamp(3);
    if (warp == 0) tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
}


// =================================================================================================================
// Host launchers
// =================================================================================================================
static int sm_count() {
    static int n = 0;
    if (!n) { int dev; NGP_CUDA_CHECK(cudaGetDevice(&dev)); NGP_CUDA_CHECK(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev)); }
    return n;
}

// CTAs per SM of a persistent fused-MLP kernel.  cudaOccupancyMaxActiveBlocksPerMultiprocessor reports 1 for any kernel
// that allocates tensor memory (measured on B200, driver 580), although the SM co-schedules such CTAs as long as their
// TMEM columns (512 per SM), shared memory (227 KB + 1 KB reserved per CTA) and threads fit — so the limit is computed here.
template <class K>
static uint32_t persistent_grid(K kernel, uint32_t smem_bytes, uint32_t tmem_cols, uint32_t n_tiles) {
    NGP_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
    NGP_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared));
    NGP_CHECK(smem_bytes + 1024 <= 227 * 1024, "fused MLP kernel does not fit on an SM");
    uint32_t per_sm = std::min<uint32_t>((227u * 1024u) / (smem_bytes + 1024u), 512u / tmem_cols);
    per_sm = std::max(1u, std::min(per_sm, 8u));
    if (const char* e = getenv("NGP_MLP_CTAS_PER_SM")) per_sm = (uint32_t)std::max(1, atoi(e));
    if (getenv("NGP_DEBUG")) fprintf(stderr, "[ngp] fused MLP kernel: smem %u B, tmem %u cols, %u CTAs/SM, %u tiles\n", smem_bytes, tmem_cols, per_sm, n_tiles);
    return std::min<uint32_t>(n_tiles, per_sm * (uint32_t)sm_count());
}

static bool mlp_ts_form() {   // NGP_MLP_TS=0: the round-1 kernels (every operand in shared memory), for A/B measurements
    static const bool v = [] { const char* e = getenv("NGP_MLP_TS"); return e ? atoi(e) != 0 : true; }();
    return v;
}

template <uint32_t PW, uint32_t DH, uint32_t RH, int MODE>
static void launch_fwd(cudaStream_t s, const FwdParams& p) {
    if (mlp_ts_form()) {
        const FwdTsSmem<PW, DH, RH, MODE> L;
        auto k = mlp_forward_ts_kernel<PW, DH, RH, MODE>;
        const uint32_t grid = persistent_grid(k, L.total, TS_COLS, p.n_tiles);
        static const bool probe = getenv("NGP_MLP_PROBE") != nullptr;
        if (probe && MODE == MODE_NERF && p.n_tiles > 1024) {   // debug: phase breakdown of one CTA, printed to stderr
            FwdParams q = p; long long* d = nullptr; long long h[16];
            NGP_CUDA_CHECK(cudaMalloc(&d, sizeof(h))); NGP_CUDA_CHECK(cudaMemsetAsync(d, 0, sizeof(h), s));
            q.probe = d;
            k<<<grid, 128, L.total, s>>>(q);
            NGP_POST_LAUNCH();
            NGP_CUDA_CHECK(cudaMemcpyAsync(h, d, sizeof(h), cudaMemcpyDeviceToHost, s)); NGP_CUDA_CHECK(cudaStreamSynchronize(s)); cudaFree(d);
            fprintf(stderr, "[ngp] mlp_forward_ts probe: %lld tiles of CTA 0 (grid %u, %u tiles), cycles per tile by phase:", h[15], grid, p.n_tiles);
            for (int i = 0; i < 10; ++i) fprintf(stderr, " %lld", h[15] ? h[i] / h[15] : 0);
            fprintf(stderr, "\n");
            return;
        }
        k<<<grid, 128, L.total, s>>>(p);
        NGP_POST_LAUNCH();
        return;
    }
    const FwdSmem<PW, DH, RH, MODE> L;
    auto k = mlp_forward_kernel<PW, DH, RH, MODE>;
    const uint32_t grid = persistent_grid(k, L.total, 64, p.n_tiles);
    k<<<grid, 128, L.total, s>>>(p);
    NGP_POST_LAUNCH();
}

// NGP_MLP_BWD = 0 (default): every operand in shared memory; 1: TS form; 2 (builds with -DNGP_MLP_BWD_KEEP only): TS form + activations kept in
// registers.  Measured on B200 (profiles/r02_mlp_bwd_forms.txt): 0.0605 / 0.0598 / 0.0618 ms per 2^18 samples — the kernel is bound by the
// length of its per-tile dependency chain (9 MMA stages x ~1.3 k cycles at 2 CTAs per SM), not by shared-memory traffic, so form 0 stays.
static int mlp_bwd_form() {
    static const int v = [] { const char* e = getenv("NGP_MLP_BWD"); return e ? atoi(e) : 0; }();
    return v;
}

template <uint32_t PW, uint32_t DH, uint32_t RH, int MODE, bool TS, bool KEEP>
static void launch_bwd_form(cudaStream_t s, const BwdParams& p) {
    using Cfg = BwdCfg<PW, DH, RH, MODE, TS>;
    const Cfg L;
    static const bool probe = getenv("NGP_MLP_PROBE") != nullptr;
    if (probe && p.dL_denc && p.n_tiles >= 1024) {   // debug: where the kernel's wall time goes, from per-CTA %globaltimer stamps (ns), to stderr
        auto k = mlp_backward_kernel<PW, DH, RH, MODE, true, TS, KEEP>;
        const uint32_t grid = persistent_grid(k, L.total, Cfg::TMEM_COLS, p.n_tiles);
        BwdParams q = p; unsigned long long* d = nullptr; std::vector<unsigned long long> h((size_t)grid * 4);
        NGP_CUDA_CHECK(cudaMalloc(&d, h.size() * 8)); q.probe = d;
        k<<<grid, 128, L.total, s>>>(q);
        NGP_POST_LAUNCH();
        NGP_CUDA_CHECK(cudaMemcpyAsync(h.data(), d, h.size() * 8, cudaMemcpyDeviceToHost, s)); NGP_CUDA_CHECK(cudaStreamSynchronize(s)); cudaFree(d);
        unsigned long long t0 = ~0ull, t3 = 0; double pro = 0, loop = 0, fl = 0, start_spread = 0;
        for (uint32_t b = 0; b < grid; ++b) { t0 = std::min(t0, h[b * 4]); t3 = std::max(t3, h[b * 4 + 3]); }
        for (uint32_t b = 0; b < grid; ++b) { pro += h[b * 4 + 1] - h[b * 4]; loop += h[b * 4 + 2] - h[b * 4 + 1]; fl += h[b * 4 + 3] - h[b * 4 + 2]; start_spread = std::max(start_spread, (double)(h[b * 4] - t0)); }
        fprintf(stderr, "[ngp] mlp_backward probe (grid %u, %u tiles): first start -> last exit %.1f us; per CTA mean: prologue %.1f us, tile loop %.1f us, wgrad flush %.1f us; last CTA started %.1f us after the first\n",
                grid, p.n_tiles, (t3 - t0) * 1e-3, pro / grid * 1e-3, loop / grid * 1e-3, fl / grid * 1e-3, start_spread * 1e-3);
        return;
    }
    if (p.dL_denc) {
        auto k = mlp_backward_kernel<PW, DH, RH, MODE, true, TS, KEEP>;
        const uint32_t grid = persistent_grid(k, L.total, Cfg::TMEM_COLS, p.n_tiles);
        k<<<grid, 128, L.total, s>>>(p);
    } else {
        auto k = mlp_backward_kernel<PW, DH, RH, MODE, false, TS, KEEP>;
        const uint32_t grid = persistent_grid(k, L.total, Cfg::TMEM_COLS, p.n_tiles);
        k<<<grid, 128, L.total, s>>>(p);
    }
    NGP_POST_LAUNCH();
}

template <uint32_t PW, uint32_t DH, uint32_t RH, int MODE>
static void launch_bwd(cudaStream_t s, const BwdParams& p) {
    switch (mlp_bwd_form()) {
        case 1: launch_bwd_form<PW, DH, RH, MODE, true, false>(s, p); break;
#ifdef NGP_MLP_BWD_KEEP   // not built by default: 224 registers per thread and measured slower (profiles/r02_mlp_bwd_forms.txt)
        case 2: launch_bwd_form<PW, DH, RH, MODE, true, true>(s, p); break;
#endif
        default: launch_bwd_form<PW, DH, RH, MODE, false, false>(s, p); break;
    }
}

#define NERF_DISPATCH(FN, MODE, d, ...)                                                                          \
    do {                                                                                                          \
        const uint32_t key = (d).pos_width * 100 + (d).density_hidden * 10 + (d).rgb_hidden;                      \
        switch (key) {                                                                                            \
            case 1611: FN<16, 1, 1, MODE>(__VA_ARGS__); break;                                                    \
            case 1612: FN<16, 1, 2, MODE>(__VA_ARGS__); break;                                                    \
            case 1622: FN<16, 2, 2, MODE>(__VA_ARGS__); break;                                                    \
            case 3211: FN<32, 1, 1, MODE>(__VA_ARGS__); break;                                                    \
            case 3212: FN<32, 1, 2, MODE>(__VA_ARGS__); break;                                                    \
            case 3222: FN<32, 2, 2, MODE>(__VA_ARGS__); break;                                                    \
            default: throw Error("FullyFusedMLP: unsupported NeRF network shape (encoder width 16/32, density hidden 1-2, rgb hidden 1-2, 64 neurons)"); \
        }                                                                                                         \
    } while (0)

#define PLAIN_DISPATCH(FN, in_w, nh, ...)                                                                        \
    do {                                                                                                          \
        const uint32_t key = (in_w) * 10 + (nh);                                                                  \
        switch (key) {                                                                                            \
            case 161: FN<16, 1, 1, MODE_PLAIN>(__VA_ARGS__); break;                                               \
            case 162: FN<16, 2, 1, MODE_PLAIN>(__VA_ARGS__); break;                                               \
            case 321: FN<32, 1, 1, MODE_PLAIN>(__VA_ARGS__); break;                                               \
            case 322: FN<32, 2, 1, MODE_PLAIN>(__VA_ARGS__); break;                                               \
            case 323: FN<32, 3, 1, MODE_PLAIN>(__VA_ARGS__); break;                                               \
            default: throw Error("FullyFusedMLP: unsupported shape (input width 16/32, 1-3 hidden layers, 64 neurons)"); \
        }                                                                                                         \
    } while (0)

static uint32_t tiles_of(uint32_t n) {
    NGP_CHECK(n % TILE_ROWS == 0, "batch size must be a multiple of 128 (callers round to batch_size_granularity = 256)");
    return n / TILE_ROWS;
}

void nerf_mlp_forward(cudaStream_t s, const NerfNetDims& d, const __half* wd, const __half* wr, const __half* enc, const float* coords,
                      uint32_t n, __half* out, const uint32_t* n_dyn) {
    if (n == 0) return;
    NGP_PROFILE_SCOPE(s, "nerf_mlp_forward", n);
    FwdParams p{wd, wr, enc, coords, out, 4, tiles_of(n), n_dyn, nullptr};
    NERF_DISPATCH(launch_fwd, MODE_NERF, d, s, p);
}

void nerf_mlp_density(cudaStream_t s, const NerfNetDims& d, const __half* wd, const __half* enc, uint32_t n, __half* out) {
    if (n == 0) return;
    NGP_PROFILE_SCOPE(s, "nerf_mlp_density", n);
    FwdParams p{wd, nullptr, enc, nullptr, out, 1, tiles_of(n), nullptr, nullptr};
    NerfNetDims dd = d; dd.rgb_hidden = 1;
    NERF_DISPATCH(launch_fwd, MODE_DENSITY, dd, s, p);
}

void nerf_mlp_backward(cudaStream_t s, const NerfNetDims& d, const __half* wd, const __half* wr, const __half* enc, const float* coords,
                       uint32_t n, const __half* dL_dout, __half* dL_denc, float* wgrad, float* dL_dcoords) {
    if (n == 0) return;
    NGP_PROFILE_SCOPE(s, "nerf_mlp_backward", n);
    BwdParams p{wd, wr, enc, coords, dL_dout, 4, dL_denc, wgrad, tiles_of(n), dL_dcoords, nullptr};
    NERF_DISPATCH(launch_bwd, MODE_NERF, d, s, p);
}

void mlp_forward(cudaStream_t s, uint32_t in_w, uint32_t nh, const __half* w, const __half* in, uint32_t n, __half* out, uint32_t out_stride) {
    if (n == 0) return;
    NGP_CHECK(out_stride % 8 == 0 && out_stride >= 16, "mlp_forward: out_stride must be a multiple of 8 and >= 16");
    NGP_PROFILE_SCOPE(s, "mlp_forward", n);
    FwdParams p{w, nullptr, in, nullptr, out, out_stride, tiles_of(n), nullptr, nullptr};
    PLAIN_DISPATCH(launch_fwd, in_w, nh, s, p);
}

void mlp_backward(cudaStream_t s, uint32_t in_w, uint32_t nh, const __half* w, const __half* in, uint32_t n, const __half* dL_dout,
                  uint32_t dout_stride, __half* dL_din, float* wgrad) {
    if (n == 0) return;
    NGP_CHECK(dout_stride % 8 == 0 && dout_stride >= 16, "mlp_backward: dout_stride must be a multiple of 8 and >= 16");
    NGP_PROFILE_SCOPE(s, "mlp_backward", n);
    BwdParams p{w, nullptr, in, nullptr, dL_dout, dout_stride, dL_din, wgrad, tiles_of(n), nullptr, nullptr};
    PLAIN_DISPATCH(launch_bwd, in_w, nh, s, p);
}

}  // namespace ngp

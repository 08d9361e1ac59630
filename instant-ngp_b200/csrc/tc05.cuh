// Thin inline-PTX layer over the Blackwell (sm_100a) tensor-core path used by the fused MLP kernels:
// tcgen05.mma (single-thread issue, accumulators in TMEM), tcgen05.ld, TMEM alloc, mbarrier, bulk-async (TMA) copies.
// Operand tiles live in shared memory in the no-swizzle canonical layout of 8x8 fp16 core matrices ("blk8",
// ngp_common.cuh), which serves K-major and MN-major operands from the same bytes:
//   K-major  (rows = M/N index, cols = K):  LBO = 128 B (next core matrix along K), SBO = C/8*128 B (next 8 rows)
//   MN-major (rows = K index, cols = M/N):  SBO = 128 B (next 8 cols),             LBO = C/8*128 B (next 8 rows)
#pragma once
#include "ngp_common.cuh"

namespace ngp { namespace tc {

NGP_D uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- mbarrier ----
NGP_D void mbar_init(uint64_t* bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count)); }
NGP_D void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
NGP_D void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
NGP_D void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_LOOP:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra WAIT_DONE;\n\t"
        "bra WAIT_LOOP;\n\t"
        "WAIT_DONE:\n\t}"
        ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}

// ---- bulk async copy global -> shared (TMA, SASS UBLKCP), completion on an mbarrier ----
NGP_D void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// generic-proxy smem writes -> visible to the async proxy (UMMA operand reads)
NGP_D void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---- TMEM ----
template <uint32_t COLS>
NGP_D void tmem_alloc(uint32_t* dst_smem) {   // one full warp
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <uint32_t COLS>
NGP_D void tmem_dealloc(uint32_t taddr) {     // same warp that allocated
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(COLS) : "memory");
}
NGP_D void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
NGP_D void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
NGP_D void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// 32 lanes x 16 consecutive fp32 columns: thread t of warp w reads TMEM lane 32*(w%4)+t
NGP_D void tmem_ld16(uint32_t taddr, float* v) {
    uint32_t r[16];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr) : "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// registers -> TMEM, 32 lanes x N consecutive 32-bit columns (thread t of warp w writes lane 32*(w%4)+t).  Used to hand the
// fp16 activations of a layer to the next layer's MMA as its A operand WITHOUT a round trip through shared memory: for
// kind::f16 an A operand in TMEM is [M = lane][K/2 columns], two consecutive K elements packed per 32-bit column (low half
// first) — verified numerically on B200 by tools/ubench_tc.cu.  Measured (tools/ubench_tc*.cu): tcgen05.st 758 B/cycle/SM,
// tcgen05.ld 335 B/cycle/SM; a 128x64 epilogue (ld, relu+pack, st) takes 148 cycles into TMEM against 322 into shared memory.
NGP_D void tmem_st16(uint32_t taddr, const uint32_t* r) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
                 ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
                   "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]) : "memory");
}
NGP_D void tmem_st32(uint32_t taddr, const uint32_t* r) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
                 ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]),
                   "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]),
                   "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31]) : "memory");
}
NGP_D void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// ---- UMMA descriptors ----
// shared-memory matrix descriptor, SWIZZLE_NONE, sm_100 version field = 1 (bits 46-47)
NGP_D uint64_t make_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    return (uint64_t)((smem_addr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo_bytes >> 4) << 16) | ((uint64_t)(sbo_bytes >> 4) << 32) | (1ull << 46);
}
// instruction descriptor for kind::f16: fp16 A/B, fp32 D
constexpr uint32_t make_idesc(uint32_t M, uint32_t N, uint32_t a_mn_major, uint32_t b_mn_major) {
    return (1u << 4) | (a_mn_major << 15) | (b_mn_major << 16) | ((N >> 3) << 17) | ((M >> 4) << 24);
}
// D[tmem] (+)= A[smem] * B[smem], one K=16 step; issued by ONE thread
NGP_D void umma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem] ("TS" form): the A operand is read from tensor memory, so the instruction's shared-memory
// traffic is B alone.  Measured per K=16 instruction on B200: M128 N64 33.6 cycles (SS form: 50, bound by the 4 KB A read at
// 128 B/cycle), M128 N16 10.6 cycles (SS form: 40.5).
NGP_D void umma_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// arrive on an mbarrier when all previously issued MMAs of this thread have completed
NGP_D void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// Operand views over a blk8 buffer [rows x C] at shared address `base`.
struct Operand { uint32_t base, C; };
// K-major: rows = M (or N) index, K along columns starting at column k0 (multiple of 16)
NGP_D uint64_t desc_kmajor(Operand o, uint32_t k0) { return make_desc(o.base + (k0 >> 3) * 128u, 128u, (o.C >> 3) * 128u); }
// MN-major: columns [c0, ..) = M (or N) index, K along rows starting at row r0 (multiple of 16)
NGP_D uint64_t desc_mnmajor(Operand o, uint32_t c0, uint32_t r0) {
    return make_desc(o.base + (r0 >> 3) * (o.C >> 3) * 128u + (c0 >> 3) * 128u, (o.C >> 3) * 128u, 128u);
}

// D[128 x N] = A[128 x K] * W^T, A K-major (activations, K columns from a_k0), W = [N x K] blk8 weights (K-major B)
template <uint32_t N>
NGP_D void gemm_act_wT(uint32_t tmem_d, Operand a, uint32_t a_k0, Operand w, uint32_t K) {
    constexpr uint32_t idesc = make_idesc(128, N, 0, 0);
    for (uint32_t k = 0; k < K; k += 16) umma(tmem_d, desc_kmajor(a, a_k0 + k), desc_kmajor(w, k), idesc, k > 0);
}
// D[128 x N] = A[128 x K] * W^T with A in TMEM (packed fp16 pairs from column tmem_a on), W = [N x K] blk8 weights (K-major B)
template <uint32_t N>
NGP_D void gemm_tmem_wT(uint32_t tmem_d, uint32_t tmem_a, Operand w, uint32_t K) {
    constexpr uint32_t idesc = make_idesc(128, N, 0, 0);
    for (uint32_t k = 0; k < K; k += 16) umma_ts(tmem_d, tmem_a + (k >> 1), desc_kmajor(w, k), idesc, k > 0);
}
// D[128 x N] = G[128 x K] * W with G in TMEM, W = [K x N] blk8 weights read MN-major (dgrad)
template <uint32_t N>
NGP_D void gemm_tmem_w(uint32_t tmem_d, uint32_t tmem_g, Operand w, uint32_t K) {
    constexpr uint32_t idesc = make_idesc(128, N, 0, 1);
    for (uint32_t k = 0; k < K; k += 16) umma_ts(tmem_d, tmem_g + (k >> 1), desc_mnmajor(w, 0, k), idesc, k > 0);
}
// D[128 x N] = G[128 x K] * W, G K-major (K columns = W's rows), W = [K x N] blk8 weights read MN-major (dgrad)
template <uint32_t N>
NGP_D void gemm_act_w(uint32_t tmem_d, Operand g, Operand w, uint32_t K) {
    constexpr uint32_t idesc = make_idesc(128, N, 0, 1);
    for (uint32_t k = 0; k < K; k += 16) umma(tmem_d, desc_kmajor(g, k), desc_mnmajor(w, 0, k), idesc, k > 0);
}
// D[64 x N] (+)= P^T[64 x 128] * Q[128 x N]: P cols [p_c0, p_c0+64) -> M, Q cols [q_c0, q_c0+N) -> N, K = 128 tile rows (wgrad)
template <uint32_t N>
NGP_D void gemm_pT_q(uint32_t tmem_d, Operand p, uint32_t p_c0, Operand q, uint32_t q_c0, uint32_t accumulate_first) {
    constexpr uint32_t idesc = make_idesc(64, N, 1, 1);
    for (uint32_t r = 0; r < 128; r += 16) umma(tmem_d, desc_mnmajor(p, p_c0, r), desc_mnmajor(q, q_c0, r), idesc, (r > 0) | accumulate_first);
}

}}  // namespace ngp::tc

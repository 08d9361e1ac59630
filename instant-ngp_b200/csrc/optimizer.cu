// Fused Adam (+ EMA of the fp16 weights) over the flat parameter buffer [density MLP | rgb MLP | hash table].
// Replaces tcnn's Ema(ExponentialDecay(Adam)) optimizer stack driven by Trainer::optimizer_step
// (testbed_nerf.cu:3678, testbed_image.cu:283; hyper-parameters configs/nerf/base.json:5-22).  Semantics kept:
//   g = grad / loss_scale; hash ("non-matrix") entries with g == 0 are skipped and keep their own step count;
//   L2 regularisation only on MLP ("matrix") weights; fp32 master weights -> fp16 copy; EMA over every entry.
// One pass, pure HBM streaming: 2 B grad + 4+4+4+4 B (master, m, v, step) r/w + 2 B param (+ 4+2 B EMA) per entry.
#include "ngp_internal.h"
#include <cmath>
#include <mutex>
#include <vector>

namespace ngp {

// 4 consecutive parameters per thread (16-byte accesses on the fp32 arrays, 8-byte on the fp16 ones).  The pass has to stay
// under the HBM time of its ~30 B/parameter, i.e. ~25 instructions per parameter, so nothing transcendental is evaluated
// per entry: the bias correction sqrt(1 - beta2^t) / (1 - beta1^t) depends on the entry's own step count t only (untouched
// hash entries do not step) and is read from a table indexed by t, built once per (beta1, beta2) with
// -expm1f(t * ln(beta)) (no cancellation; the reference's per-entry powf costs ~200 instructions per parameter);
// 1 / (sqrt(v) + eps) uses the hardware approximations (2 ulp), divisions by launch constants are multiplications
// by their reciprocals.  All within 1e-6 relative of the reference arithmetic (tests/test_ops_gpu.py: rtol 1e-5).
constexpr uint32_t BIAS_TABLE_N = 1u << 15;
struct AdamConsts { float ln_beta1, ln_beta2, ema_old_debias, inv_ema_new_debias, inv_loss_scale; const float* bias; };

__device__ __noinline__ float bias_correction(float tf, float ln_beta1, float ln_beta2) {
    return sqrtf(-expm1f(tf * ln_beta2)) / (-expm1f(tf * ln_beta1));   // sqrt(1 - beta2^t) / (1 - beta1^t)
}
__global__ void bias_table_kernel(float ln_beta1, float ln_beta2, float* __restrict__ tab) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < BIAS_TABLE_N) tab[t] = t == 0 ? 0.0f : bias_correction((float)t, ln_beta1, ln_beta2);
}
__device__ __forceinline__ float approx_sqrt(float x) { float y; asm("sqrt.approx.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float approx_rcp(float x) { float y; asm("rcp.approx.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

__device__ __forceinline__ void adam_one(const AdamHyper& h, const AdamConsts& c, bool matrix, float g_raw, float& w, float& m, float& v, uint32_t& t,
                                         bool& stepped) {
    float g = g_raw * c.inv_loss_scale;
    const bool skip = matrix ? !h.optimize_matrix : (!h.optimize_non_matrix || g == 0.0f);
    stepped = !skip;
    if (skip) return;
    if (matrix) g += h.l2_reg * w;
    m = h.beta1 * m + (1.0f - h.beta1) * g;
    v = h.beta2 * v + (1.0f - h.beta2) * (g * g);
    ++t;
    const float corr = t < BIAS_TABLE_N ? __ldg(c.bias + t) : bias_correction((float)t, c.ln_beta1, c.ln_beta2);
    w = w - h.learning_rate * corr * approx_rcp(approx_sqrt(v) + h.epsilon) * m;
}

__global__ void __launch_bounds__(256) adam_kernel(const AdamHyper h, const AdamConsts c, uint64_t n, float* __restrict__ master, __half* __restrict__ params,
                                                   const __half* __restrict__ grads, float* __restrict__ m, float* __restrict__ v,
                                                   uint32_t* __restrict__ steps, float* __restrict__ ema, __half* __restrict__ ema_half) {
    const uint64_t i0 = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i0 >= n) return;
    if (i0 + 4 <= n) {
        // every load that does not depend on the gradient goes out first (gradient, fp16 copy, EMA): the pass is a pure
        // stream and lives on bytes in flight; only the optimizer state of touched entries is fetched behind the test
        const uint2 graw = __ldg(reinterpret_cast<const uint2*>(grads + i0));
        const uint2 praw = *reinterpret_cast<const uint2*>(params + i0);
        // fp32 optimizer state is touched once per step and is ~10x the fp16 table: streaming (evict-first) accesses, so that it does
        // not push the table (params + grads, under the persisting L2 window of Trainer::apply_l2_policy) out of L2
        float4 e4 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (ema) e4 = __ldcs(reinterpret_cast<const float4*>(ema + i0));
        const __half* gh = reinterpret_cast<const __half*>(&graw);
        const __half* ph = reinterpret_cast<const __half*>(&praw);
        const bool matrix = i0 < h.n_matrix_params;       // n_matrix_params is a multiple of 4 (MLP layers are 16 x 64 blocks)
        const bool nonzero = ((graw.x & 0x7fff7fffu) | (graw.y & 0x7fff7fffu)) != 0u;
        const bool any = matrix ? h.optimize_matrix : (h.optimize_non_matrix && nonzero);
        if (h.clear_consumed_grads && !matrix && nonzero) *reinterpret_cast<uint2*>(const_cast<__half*>(grads) + i0) = make_uint2(0u, 0u);
        __half wh[4];
        *reinterpret_cast<uint2*>(wh) = praw;
        if (any) {
            float4 w4 = __ldcs(reinterpret_cast<const float4*>(master + i0)), m4 = __ldcs(reinterpret_cast<const float4*>(m + i0)), v4 = __ldcs(reinterpret_cast<const float4*>(v + i0));
            uint4 t4 = __ldcs(reinterpret_cast<const uint4*>(steps + i0));
            float* w = reinterpret_cast<float*>(&w4); float* mm = reinterpret_cast<float*>(&m4); float* vv = reinterpret_cast<float*>(&v4);
            uint32_t* tt = reinterpret_cast<uint32_t*>(&t4);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                bool stepped;
                adam_one(h, c, matrix, __half2float(gh[k]), w[k], mm[k], vv[k], tt[k], stepped);
                wh[k] = stepped ? __float2half_rn(w[k]) : ph[k];
            }
            __stcs(reinterpret_cast<float4*>(master + i0), w4); __stcs(reinterpret_cast<float4*>(m + i0), m4); __stcs(reinterpret_cast<float4*>(v + i0), v4);
            __stcs(reinterpret_cast<uint4*>(steps + i0), t4);
            *reinterpret_cast<uint2*>(params + i0) = *reinterpret_cast<uint2*>(wh);
        }
        if (ema) {
            float* e = reinterpret_cast<float*>(&e4);
            __half eh[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                e[k] = (e[k] * h.ema_decay * c.ema_old_debias + __half2float(wh[k]) * (1.0f - h.ema_decay)) * c.inv_ema_new_debias;
                eh[k] = __float2half_rn(e[k]);
            }
            __stcs(reinterpret_cast<float4*>(ema + i0), e4);
            __stcs(reinterpret_cast<uint2*>(ema_half + i0), *reinterpret_cast<uint2*>(eh));
        }
        return;
    }
    for (uint64_t i = i0; i < n; ++i) {   // tail (n not a multiple of 4)
        const bool matrix = i < h.n_matrix_params;
        float w = master[i], mm = m[i], vv = v[i]; uint32_t t = steps[i]; bool stepped;
        adam_one(h, c, matrix, __half2float(grads[i]), w, mm, vv, t, stepped);
        if (h.clear_consumed_grads && !matrix) const_cast<__half*>(grads)[i] = __float2half_rn(0.0f);
        __half wh = params[i];
        if (stepped) { master[i] = w; m[i] = mm; v[i] = vv; steps[i] = t; wh = __float2half_rn(w); params[i] = wh; }
        if (ema) {
            const float f = (ema[i] * h.ema_decay * c.ema_old_debias + __half2float(wh) * (1.0f - h.ema_decay)) * c.inv_ema_new_debias;
            ema[i] = f; ema_half[i] = __float2half_rn(f);
        }
    }
}

// one table per (device, beta1, beta2), built on first use and kept for the life of the process
static const float* bias_table(cudaStream_t s, float ln_beta1, float ln_beta2) {
    struct Entry { int device; float l1, l2; float* tab; };
    static std::mutex mu; static std::vector<Entry> cache;
    int dev = 0; NGP_CUDA_CHECK(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lock(mu);
    for (const Entry& e : cache) if (e.device == dev && e.l1 == ln_beta1 && e.l2 == ln_beta2) return e.tab;
    float* tab = nullptr;
    NGP_CUDA_CHECK(cudaMalloc(&tab, BIAS_TABLE_N * sizeof(float)));
    bias_table_kernel<<<BIAS_TABLE_N / 256, 256, 0, s>>>(ln_beta1, ln_beta2, tab);
    NGP_POST_LAUNCH();
    NGP_CUDA_CHECK(cudaStreamSynchronize(s));      // once: later steps may come from other streams
    cache.push_back({dev, ln_beta1, ln_beta2, tab});
    return tab;
}

void adam_step(cudaStream_t s, const AdamHyper& h, uint64_t n, float* master, __half* params, const __half* grads, float* m, float* v,
               uint32_t* steps, float* ema_f32, __half* ema_half) {
    if (n == 0) return;
    NGP_CHECK(h.n_matrix_params % 4 == 0, "adam_step: n_matrix_params must be a multiple of 4");
    AdamConsts c{(float)std::log((double)h.beta1), (float)std::log((double)h.beta2), 0.f, 1.f, 1.0f / h.loss_scale, nullptr};
    c.bias = bias_table(s, c.ln_beta1, c.ln_beta2);
    NGP_PROFILE_SCOPE(s, "adam", n);
    if (ema_f32) { c.ema_old_debias = 1.0f - powf(h.ema_decay, (float)(h.ema_step - 1)); c.inv_ema_new_debias = 1.0f / (1.0f - powf(h.ema_decay, (float)h.ema_step)); }
    const uint64_t threads = (n + 3) / 4;
    adam_kernel<<<(uint32_t)((threads + 255) / 256), 256, 0, s>>>(h, c, n, master, params, grads, m, v, steps, ema_f32, ema_half);
    NGP_POST_LAUNCH();
}

__global__ void cast_kernel(const float* __restrict__ in, __half* __restrict__ out, uint64_t n) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = __float2half_rn(in[i]);
}
void cast_f32_to_half(cudaStream_t s, const float* in, __half* out, uint64_t n) {
    if (n == 0) return;
    cast_kernel<<<(uint32_t)((n + 255) / 256), 256, 0, s>>>(in, out, n);
    NGP_POST_LAUNCH();
}

}  // namespace ngp

// Fused Adam (+ EMA of the fp16 weights) over the flat parameter buffer [density MLP | rgb MLP | hash table].
// Replaces tcnn's Ema(ExponentialDecay(Adam)) optimizer stack driven by Trainer::optimizer_step
// (testbed_nerf.cu:3678, testbed_image.cu:283; hyper-parameters configs/nerf/base.json:5-22).  Semantics kept:
//   g = grad / loss_scale; hash ("non-matrix") entries with g == 0 are skipped and keep their own step count;
//   L2 regularisation only on MLP ("matrix") weights; fp32 master weights -> fp16 copy; EMA over every entry.
// One pass, pure HBM streaming: 2 B grad + 4+4+4+4 B (master, m, v, step) r/w + 2 B param (+ 4+2 B EMA) per entry.
#include "ngp_internal.h"

namespace ngp {

__global__ void __launch_bounds__(256) adam_kernel(const AdamHyper h, uint64_t n, float* __restrict__ master, __half* __restrict__ params,
                                                   const __half* __restrict__ grads, float* __restrict__ m, float* __restrict__ v,
                                                   uint32_t* __restrict__ steps, float* __restrict__ ema, __half* __restrict__ ema_half,
                                                   float ema_old_debias, float ema_new_debias) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float g = __half2float(grads[i]) / h.loss_scale;
    const bool matrix = i < h.n_matrix_params;
    const bool skip = matrix ? !h.optimize_matrix : (!h.optimize_non_matrix || g == 0.0f);
    __half wh;
    if (!skip) {
        const float w = master[i];
        if (matrix) g += h.l2_reg * w;
        const float fm = h.beta1 * m[i] + (1.0f - h.beta1) * g;
        const float sm = h.beta2 * v[i] + (1.0f - h.beta2) * (g * g);
        m[i] = fm; v[i] = sm;
        const uint32_t t = ++steps[i];
        const float lr = h.learning_rate * sqrtf(1.0f - powf(h.beta2, (float)t)) / (1.0f - powf(h.beta1, (float)t));
        const float nw = w - lr / (sqrtf(sm) + h.epsilon) * fm;
        master[i] = nw;
        wh = __float2half_rn(nw);
        params[i] = wh;
    } else if (ema) {
        wh = params[i];
    }
    if (ema) {
        const float f = (ema[i] * h.ema_decay * ema_old_debias + __half2float(wh) * (1.0f - h.ema_decay)) / ema_new_debias;
        ema[i] = f;
        ema_half[i] = __float2half_rn(f);
    }
}

void adam_step(cudaStream_t s, const AdamHyper& h, uint64_t n, float* master, __half* params, const __half* grads, float* m, float* v,
               uint32_t* steps, float* ema_f32, __half* ema_half) {
    if (n == 0) return;
    NGP_PROFILE_SCOPE(s, "adam", n);
    float od = 0.f, nd = 1.f;
    if (ema_f32) { od = 1.0f - powf(h.ema_decay, (float)(h.ema_step - 1)); nd = 1.0f - powf(h.ema_decay, (float)h.ema_step); }
    adam_kernel<<<(uint32_t)((n + 255) / 256), 256, 0, s>>>(h, n, master, params, grads, m, v, steps, ema_f32, ema_half, od, nd);
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

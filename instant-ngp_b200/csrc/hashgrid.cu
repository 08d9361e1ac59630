// Multiresolution hash-grid encoding for sm_100a: forward gather + D-linear interpolation, backward scatter-add,
// input gradients.  Replaces tcnn::GridEncoding<__half> behind the reference's NerfNetwork / NetworkWithInputEncoding
// (nerf_network.h:93-95,143-148,324-333; testbed.cu:3961-4005,4101).
//
// Bottleneck: random 4..16-byte gathers (one 32 B L2 sector per corner), L2-latency/LTS-throughput bound when the
// table fits B200's 126 MB L2 and HBM-sector bound otherwise.  Mapping: one thread per (point, 8-feature chunk) —
// a chunk is 8/F consecutive levels — so every thread keeps (8/F)*2^D independent vector loads in flight and the
// warp's result is written as full 16-byte chunks of the blk8 activation layout (ngp_common.cuh) that the tcgen05
// MLP kernels pull in with one bulk-async copy per 128-sample tile.
#include "ngp_internal.h"
#include <cmath>
#include <algorithm>

namespace ngp {

void grid_make_layout(GridLayout& g, uint32_t n_dims, uint32_t n_levels, uint32_t n_features, uint32_t log2T,
                      uint32_t base_resolution, float per_level_scale) {
    NGP_CHECK(n_dims == 2 || n_dims == 3, "HashGrid: n_pos_dims must be 2 or 3");
    NGP_CHECK(n_levels >= 1 && n_levels <= GRID_MAX_LEVELS, "HashGrid: n_levels must be in [1,32]");
    NGP_CHECK(n_features == 1 || n_features == 2 || n_features == 4 || n_features == 8, "HashGrid: n_features_per_level must be 1,2,4 or 8");
    NGP_CHECK(log2T >= 4 && log2T <= 28, "HashGrid: log2_hashmap_size out of range");
    g = GridLayout{};
    g.n_dims = n_dims; g.n_levels = n_levels; g.n_features = n_features; g.log2_hashmap_size = log2T;
    g.base_resolution = base_resolution; g.per_level_scale = per_level_scale;
    g.padded_width = (uint32_t)next_multiple(n_levels * n_features, 16);
    NGP_CHECK(g.padded_width <= 64, "HashGrid: n_levels*n_features_per_level must be <= 64");
    // Level geometry is computed ONCE on the host so that device code and the CPU oracle see identical floats.
    float log2b = std::log2(per_level_scale);
    uint32_t offset = 0;
    for (uint32_t l = 0; l < n_levels; ++l) {
        float scale = exp2f((float)l * log2b) * (float)base_resolution - 1.0f;
        uint32_t res = (uint32_t)ceilf(scale) + 1u;
        uint32_t max_params = 0xFFFFFFFFu / 2;
        double dense = std::pow((double)res, (double)n_dims);
        uint32_t n = dense > (double)max_params ? max_params : (uint32_t)(dense + 0.5);
        n = (n + GRID_PARAM_ALIGN - 1) / GRID_PARAM_ALIGN * GRID_PARAM_ALIGN;
        n = std::min(n, 1u << log2T);
        g.offsets[l] = offset; g.sizes[l] = n; g.resolutions[l] = res; g.scales[l] = scale;
        // the (uint32-wrapping) stride walk of the index function does not depend on the cell: do it here
        uint32_t stride = 1;
        for (uint32_t d = 0; d < n_dims && stride <= n; ++d) { g.dense_stride[l][d] = stride; stride *= res; }
        if (n < stride) g.hashed_mask |= 1u << l;
        offset += n;
    }
    g.offsets[n_levels] = offset;
    g.n_params = (uint64_t)offset * n_features;
}

// ------------------------------------------------------------------------------------------------
template <uint32_t D>
struct Cell { uint32_t c[D]; float f[D]; };

template <uint32_t D>
NGP_D Cell<D> locate(const float* __restrict__ x, float scale) {
    Cell<D> r;
#pragma unroll
    for (uint32_t d = 0; d < D; ++d) {
        float p = fmaf(scale, x[d], 0.5f);
        float fl = floorf(p);
        r.c[d] = (uint32_t)(int)fl;
        r.f[d] = p - fl;
    }
    return r;
}

template <uint32_t D>
NGP_D uint32_t corner_index(const GridLayout& g, uint32_t level, const Cell<D>& cell, uint32_t corner, float& w) {
    uint32_t cl[D];
    w = 1.0f;
#pragma unroll
    for (uint32_t d = 0; d < D; ++d) {
        if ((corner & (1u << d)) == 0) { w *= 1.0f - cell.f[d]; cl[d] = cell.c[d]; }
        else { w *= cell.f[d]; cl[d] = cell.c[d] + 1; }
    }
    const uint32_t size = g.sizes[level];
    uint32_t index;
    if (g.hashed_mask & (1u << level)) {
        index = cl[0];
        if (D > 1) index ^= cl[1] * GRID_PRIME_1;
        if (D > 2) index ^= cl[2] * GRID_PRIME_2;
        index &= size - 1;   // hashed levels always have size == 2^log2_hashmap_size
    } else {
        index = 0;
#pragma unroll
        for (uint32_t d = 0; d < D; ++d) index += cl[d] * g.dense_stride[level][d];
        if (index >= size) index %= size;
    }
    return index;
}

template <uint32_t F> struct Feat;
template <> struct Feat<1> { using T = unsigned short; };
template <> struct Feat<2> { using T = uint32_t; };
template <> struct Feat<4> { using T = uint2; };
template <> struct Feat<8> { using T = uint4; };

template <uint32_t F>
NGP_D void accumulate(float* acc, float w, typename Feat<F>::T raw) {
    const __half* h = reinterpret_cast<const __half*>(&raw);
#pragma unroll
    for (uint32_t f = 0; f < F; ++f) acc[f] = fmaf(w, __half2float(h[f]), acc[f]);
}

// All 2^D corner indices and weights of one level with the per-axis work factored out: two candidate cells per axis
// (c, c+1) give two hash terms / dense offsets and two weights per axis; a corner combines one of each.  The products are
// formed in the reference order ((wx*wy)*wz), so weights and indices are bit-identical to corner_index().
template <uint32_t D>
struct LevelCorners {
    uint32_t term[D][2]; float wt[D][2]; uint32_t size; bool hashed;
    NGP_D LevelCorners(const GridLayout& g, uint32_t level, const float* __restrict__ x) {
        const float scale = g.scales[level];
        size = g.sizes[level]; hashed = (g.hashed_mask >> level) & 1u;
#pragma unroll
        for (uint32_t d = 0; d < D; ++d) {
            const float p = fmaf(scale, x[d], 0.5f), fl = floorf(p);
            const uint32_t c = (uint32_t)(int)fl; const float f = p - fl;
            wt[d][0] = 1.0f - f; wt[d][1] = f;
            const uint32_t mul = hashed ? (d == 0 ? 1u : (d == 1 ? GRID_PRIME_1 : GRID_PRIME_2)) : g.dense_stride[level][d];
            term[d][0] = c * mul; term[d][1] = (c + 1u) * mul;
        }
    }
    NGP_D uint32_t index(uint32_t corner) const {
        uint32_t i;
        if (hashed) {
            i = term[0][corner & 1u];
            if (D > 1) i ^= term[1][(corner >> 1) & 1u];
            if (D > 2) i ^= term[2][(corner >> 2) & 1u];
            return i & (size - 1u);
        }
        i = term[0][corner & 1u];
        if (D > 1) i += term[1][(corner >> 1) & 1u];
        if (D > 2) i += term[2][(corner >> 2) & 1u];
        return i >= size ? i % size : i;
    }
    NGP_D float weight(uint32_t corner) const {
        float w = wt[0][corner & 1u];
        if (D > 1) w *= wt[1][(corner >> 1) & 1u];
        if (D > 2) w *= wt[2][(corner >> 2) & 1u];
        return w;
    }
};

// Mapping: blockIdx.y = 8-feature chunk (8/F consecutive levels), so the level — and with it every constant-bank read of
// the layout and the hashed/dense branch — is uniform across the block; threads are consecutive points, so a warp's
// 16-byte stores fill four whole 128-byte core matrices of the blk8 output.
template <uint32_t D, uint32_t F>
__global__ void __launch_bounds__(256) grid_forward_kernel(const GridLayout g, const __half* __restrict__ table, const float* __restrict__ x,
                                                           uint32_t x_stride, uint32_t n, __half* __restrict__ y, int y_layout,
                                                           const uint32_t* __restrict__ n_dyn, const float* __restrict__ max_level) {
    constexpr uint32_t LPC = 8 / F;          // levels per 8-feature chunk
    constexpr uint32_t NC = 1u << D;
    const uint32_t row = blockIdx.x * blockDim.x + threadIdx.x, chunk = blockIdx.y;
    if (n_dyn) n = min(n, __ldg(n_dyn));
    if (row >= n) return;
    const float* xp = x + (size_t)row * x_stride;
    float xs[D];
#pragma unroll
    for (uint32_t d = 0; d < D; ++d) xs[d] = __ldg(xp + d);
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.0f;
    // per-row level cut-off (GridEncoding::set_max_level_gpu, used by max_level_rand_training): row i keeps the levels below
    // max_level[i] * n_levels (+1e-3), the others encode to zero
    const float ml = max_level ? __ldg(max_level + row) * (float)g.n_levels + 1e-3f : 3.0e38f;
#pragma unroll
    for (uint32_t li = 0; li < LPC; ++li) {
        const uint32_t level = chunk * LPC + li;
        if (level >= g.n_levels) break;
        if ((float)level >= ml) continue;
        const LevelCorners<D> lc(g, level, xs);
        const typename Feat<F>::T* t = reinterpret_cast<const typename Feat<F>::T*>(table) + g.offsets[level];
        typename Feat<F>::T raw[NC];
        // (merging the two x-neighbour corners into one aligned 2-entry load — they share a slot half of the time, see
        // grid_backward_kernel — was measured SLOWER here: the gather is bound by L2 sector traffic, which merging does not change)
#pragma unroll
        for (uint32_t k = 0; k < NC; ++k) raw[k] = __ldg(t + lc.index(k));
#pragma unroll
        for (uint32_t k = 0; k < NC; ++k) accumulate<F>(acc + li * F, lc.weight(k), raw[k]);
    }
    __half2 o[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) o[i] = __floats2half2_rn(acc[2 * i], acc[2 * i + 1]);
    size_t off = (y_layout == LAYOUT_BLK8) ? blk8_chunk_offset(row, chunk, g.padded_width) : (size_t)row * g.padded_width + chunk * 8;
    *reinterpret_cast<uint4*>(y + off) = *reinterpret_cast<const uint4*>(o);
}

// vector reductions into the fp16 gradient table (RED.E.ADD.F16x2; one instruction per table entry)
// (explicit .global: atomicAdd on a generic pointer compiles to an address-space test + shared-memory CAS loop + ATOM)
NGP_D void red_add_h2(__half* p, float a, float b) {
    __half2 v = __floats2half2_rn(a, b);
    asm volatile("red.global.add.noftz.f16x2 [%0], %1;" ::"l"(p), "r"(*reinterpret_cast<uint32_t*>(&v)) : "memory");
}
NGP_D void red_add_h1(__half* p, float a) {
    __half v = __float2half_rn(a);
    asm volatile("red.global.add.noftz.f16 [%0], %1;" ::"l"(p), "h"(*reinterpret_cast<unsigned short*>(&v)) : "memory");
}
NGP_D void red_add_h4(__half* p, float a, float b, float c, float d) {
    __half2 lo = __floats2half2_rn(a, b), hi = __floats2half2_rn(c, d);
    asm volatile("red.global.add.noftz.v2.f16x2 [%0], {%1, %2};" ::"l"(p), "r"(*reinterpret_cast<uint32_t*>(&lo)), "r"(*reinterpret_cast<uint32_t*>(&hi)) : "memory");
}

NGP_D void red_add_h8(__half* p, float a, float b, float c, float d, float e, float f, float g, float h) {
    __half2 v0 = __floats2half2_rn(a, b), v1 = __floats2half2_rn(c, d), v2 = __floats2half2_rn(e, f), v3 = __floats2half2_rn(g, h);
    asm volatile("red.global.add.noftz.v4.f16x2 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(*reinterpret_cast<uint32_t*>(&v0)), "r"(*reinterpret_cast<uint32_t*>(&v1)),
                 "r"(*reinterpret_cast<uint32_t*>(&v2)), "r"(*reinterpret_cast<uint32_t*>(&v3)) : "memory");
}
template <uint32_t F>
NGP_D void red_add_entry(__half* p, float w, const float* d) {
    if (F == 1) red_add_h1(p, w * d[0]);
    else if (F == 2) red_add_h2(p, w * d[0], w * d[1 % F]);
    else if (F == 4) red_add_h4(p, w * d[0], w * d[1 % F], w * d[2 % F], w * d[3 % F]);
    else { red_add_h4(p, w * d[0], w * d[1 % F], w * d[2 % F], w * d[3 % F]); red_add_h4(p + 4, w * d[4 % F], w * d[5 % F], w * d[6 % F], w * d[7 % F]); }
}

// Samples arrive in ray order, so neighbouring lanes of a warp sit in the same cell of the coarser levels (a level-0 cell holds
// ~100 consecutive samples of a ray, a level-8 cell two or three): their reductions would all go to the same eight table entries
// and serialise in the L2 atomic units.  Per level the warp looks for runs of consecutive lanes with identical cell coordinates;
// when there are few runs (agg_max_runs) the eight corner contributions are summed along each run with a segmented shuffle
// reduction (fp32) and only the first lane of a run issues reductions.  Fine levels (every lane its own cell) take the direct path.

template <uint32_t D, uint32_t F>
NGP_D void scatter_corners(const LevelCorners<D>& lc, __half* __restrict__ gl, const float (*v)[F], bool wide) {
    constexpr uint32_t NC = 1u << D;
    if (F == 2 && wide) {
        // Scattered reductions cost one LSU slot per LANE whatever their width (measured: merging aligned x-neighbour pairs into
        // one 8-byte RED took 0.136 -> 0.106 ms).  The two corners along x are one entry apart (hashed levels: x enters the XOR
        // unmultiplied; dense levels: stride 1), so they sit in the same aligned group of FOUR entries (16 bytes) unless x % 4 == 3:
        // one red.v4.f16x2 carries both in 3 of 4 cases (the aligned pair only in 1 of 2), the two idle slots add +0.
#pragma unroll
        for (uint32_t k = 0; k < NC; k += 2) {
            const uint32_t ia = lc.index(k), ib = lc.index(k + 1);
            if ((ia ^ ib) < 4u) {
                const uint32_t sa = ia & 3u, sb = ib & 3u;
                float o[8];
#pragma unroll
                for (uint32_t sl = 0; sl < 4; ++sl) {
                    o[2 * sl] = (sa == sl ? v[k][0] : 0.0f) + (sb == sl ? v[k + 1][0] : 0.0f);
                    o[2 * sl + 1] = (sa == sl ? v[k][1 % F] : 0.0f) + (sb == sl ? v[k + 1][1 % F] : 0.0f);
                }
                red_add_h8(gl + (size_t)(ia & ~3u) * F, o[0], o[1], o[2], o[3], o[4], o[5], o[6], o[7]);
            } else {
                red_add_entry<F>(gl + (size_t)ia * F, 1.0f, v[k]);
                red_add_entry<F>(gl + (size_t)ib * F, 1.0f, v[k + 1]);
            }
        }
    } else if (F <= 4) {
        // The two corners along x are table neighbours (index ^ 1) whenever the cell's x is even (hashed: the x term enters the
        // XOR unmultiplied) or the dense index is even: they then go out as ONE vector reduction.
#pragma unroll
        for (uint32_t k = 0; k < NC; k += 2) {
            const uint32_t ia = lc.index(k), ib = lc.index(k + 1);
            if ((ia ^ ib) == 1u) {
                const bool a_hi = ia & 1u;
                const float* lo = a_hi ? v[k + 1] : v[k]; const float* hi = a_hi ? v[k] : v[k + 1];
                __half* p = gl + (size_t)(ia & ~1u) * F;
                if (F == 1) red_add_h2(p, lo[0], hi[0]);
                else if (F == 2) red_add_h4(p, lo[0], lo[1 % F], hi[0], hi[1 % F]);
                else red_add_h8(p, lo[0], lo[1 % F], lo[2 % F], lo[3 % F], hi[0], hi[1 % F], hi[2 % F], hi[3 % F]);
            } else {
                red_add_entry<F>(gl + (size_t)ia * F, 1.0f, v[k]);
                red_add_entry<F>(gl + (size_t)ib * F, 1.0f, v[k + 1]);
            }
        }
    } else {
#pragma unroll
        for (uint32_t k = 0; k < NC; ++k) red_add_entry<F>(gl + (size_t)lc.index(k) * F, 1.0f, v[k]);
    }
}

template <uint32_t D, uint32_t F>
__global__ void __launch_bounds__(256) grid_backward_kernel(const GridLayout g, const float* __restrict__ x, uint32_t x_stride, uint32_t n,
                                                            const __half* __restrict__ dy, int dy_layout, __half* __restrict__ grad, uint32_t chunk0, uint32_t agg_max_runs, bool wide,
                                                            const float* __restrict__ max_level) {
    constexpr uint32_t LPC = 8 / F;
    constexpr uint32_t NC = 1u << D;
    const uint32_t row = blockIdx.x * blockDim.x + threadIdx.x, chunk = chunk0 + blockIdx.y, lane = threadIdx.x & 31u;
    const bool in_range = row < n;
    uint4 raw = make_uint4(0u, 0u, 0u, 0u);
    float xs[D];
#pragma unroll
    for (uint32_t d = 0; d < D; ++d) xs[d] = 0.0f;
    float ml = 3.0e38f;                       // per-row level cut-off: levels ABOVE max_level * n_levels (+1e-3) receive no gradient from this row
    if (in_range) {
        if (max_level) ml = __ldg(max_level + row) * (float)g.n_levels + 1e-3f;
        const size_t off = (dy_layout == LAYOUT_BLK8) ? blk8_chunk_offset(row, chunk, g.padded_width) : (size_t)row * g.padded_width + chunk * 8;
        raw = __ldg(reinterpret_cast<const uint4*>(dy + off));
        const float* xp = x + (size_t)row * x_stride;
#pragma unroll
        for (uint32_t d = 0; d < D; ++d) xs[d] = __ldg(xp + d);
    }
    if (__ballot_sync(0xffffffffu, (raw.x | raw.y | raw.z | raw.w) != 0u) == 0u) return;   // padding rows / untouched levels contribute nothing
    const __half* dh = reinterpret_cast<const __half*>(&raw);
#pragma unroll
    for (uint32_t li = 0; li < LPC; ++li) {
        const uint32_t level = chunk * LPC + li;
        if (level >= g.n_levels) break;
        float d[F]; bool any = false;
#pragma unroll
        for (uint32_t f = 0; f < F; ++f) { d[f] = (float)level > ml ? 0.0f : __half2float(dh[li * F + f]); any |= d[f] != 0.0f; }
        const uint32_t any_mask = __ballot_sync(0xffffffffu, any);
        if (any_mask == 0u) continue;
        const LevelCorners<D> lc(g, level, xs);
        __half* gl = grad + (size_t)g.offsets[level] * F;
        float v[NC][F];
#pragma unroll
        for (uint32_t k = 0; k < NC; ++k) {
            const float w = lc.weight(k);
#pragma unroll
            for (uint32_t f = 0; f < F; ++f) v[k][f] = any ? w * d[f] : 0.0f;
        }
        // runs of consecutive lanes in the same cell (the per-axis terms identify the cell within a level)
        bool same = any;
#pragma unroll
        for (uint32_t dd = 0; dd < D; ++dd) same &= __shfl_up_sync(0xffffffffu, lc.term[dd][0], 1) == lc.term[dd][0];
        same &= lane != 0 && ((any_mask >> (lane - 1u)) & 1u);
        const uint32_t heads = __ballot_sync(0xffffffffu, !same);
        if ((uint32_t)__popc(heads) <= agg_max_runs) {
            const uint32_t above = lane == 31u ? 0u : (heads & (0xffffffffu << (lane + 1u)));
            const uint32_t run_end = above ? (uint32_t)__ffs((int)above) - 2u : 31u;      // last lane of this lane's run
#pragma unroll
            for (uint32_t o = 1; o < 32; o <<= 1) {
                const bool take = lane + o <= run_end;
#pragma unroll
                for (uint32_t k = 0; k < NC; ++k)
#pragma unroll
                    for (uint32_t f = 0; f < F; ++f) { const float t = __shfl_down_sync(0xffffffffu, v[k][f], o); if (take) v[k][f] += t; }
            }
            if (!same && any) scatter_corners<D, F>(lc, gl, v, wide);
        } else if (any) {
            scatter_corners<D, F>(lc, gl, v, wide);
        }
    }
}

// dL/dx_d = sum_levels scale_l * sum_{corner pairs along d} w_others * (v(c_d+1) - v(c_d)) . dy_l
template <uint32_t D, uint32_t F>
__global__ void __launch_bounds__(256) grid_backward_input_kernel(const GridLayout g, const __half* __restrict__ table, const float* __restrict__ x,
                                                                  uint32_t x_stride, uint32_t n, const __half* __restrict__ dy, int dy_layout,
                                                                  float* __restrict__ dx, uint32_t dx_stride, const float* __restrict__ max_level) {
    constexpr uint32_t NC = 1u << D;
    const uint32_t row = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= n) return;
    const float* xp = x + (size_t)row * x_stride;
    float out[D];
#pragma unroll
    for (uint32_t d = 0; d < D; ++d) out[d] = 0.0f;
    const float ml = max_level ? __ldg(max_level + row) * (float)g.n_levels + 1e-3f : 3.0e38f;   // levels cut off in the forward pass have dy/dx = 0
    for (uint32_t level = 0; level < g.n_levels; ++level) {
        if ((float)level >= ml) break;
        const Cell<D> cell = locate<D>(xp, g.scales[level]);
        const __half* t = table + (size_t)g.offsets[level] * F;
        float v[NC][F];
#pragma unroll
        for (uint32_t k = 0; k < NC; ++k) {
            float w; const uint32_t idx = corner_index<D>(g, level, cell, k, w);
#pragma unroll
            for (uint32_t f = 0; f < F; ++f) v[k][f] = __half2float(t[(size_t)idx * F + f]);
        }
        float dyl[F];
#pragma unroll
        for (uint32_t f = 0; f < F; ++f) {
            const uint32_t c = level * F + f;
            size_t off = (dy_layout == LAYOUT_BLK8) ? blk8_offset(row, c, g.padded_width) : (size_t)row * g.padded_width + c;
            dyl[f] = __half2float(dy[off]);
        }
#pragma unroll
        for (uint32_t gd = 0; gd < D; ++gd) {
            float acc = 0.0f;
#pragma unroll
            for (uint32_t k = 0; k < NC; ++k) {
                if (k & (1u << gd)) continue;
                float wo = 1.0f;
#pragma unroll
                for (uint32_t d = 0; d < D; ++d) if (d != gd) wo *= (k & (1u << d)) ? cell.f[d] : 1.0f - cell.f[d];
#pragma unroll
                for (uint32_t f = 0; f < F; ++f) acc += wo * (v[k | (1u << gd)][f] - v[k][f]) * dyl[f];
            }
            out[gd] += acc * g.scales[level];
        }
    }
#pragma unroll
    for (uint32_t d = 0; d < D; ++d) dx[(size_t)row * dx_stride + d] = out[d];
}

template <uint32_t D>
__global__ void grid_corners_kernel(const GridLayout g, const float* __restrict__ x, uint32_t x_stride, uint32_t n, uint32_t level,
                                    uint32_t* __restrict__ idx, float* __restrict__ wout) {
    const uint32_t row = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= n) return;
    const Cell<D> cell = locate<D>(x + (size_t)row * x_stride, g.scales[level]);
    for (uint32_t k = 0; k < (1u << D); ++k) {
        float w; idx[(size_t)row * (1u << D) + k] = corner_index<D>(g, level, cell, k, w);
        wout[(size_t)row * (1u << D) + k] = w;
    }
}

__global__ void grid_init_kernel(uint64_t n, Pcg32 rng, float* __restrict__ p) {
    const uint64_t i0 = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i0 >= n) return;
    rng.advance((int64_t)i0);
    for (uint64_t i = i0; i < i0 + 4 && i < n; ++i) p[i] = rng.next_float() * (2.0f * GRID_INIT_RANGE) - GRID_INIT_RANGE;
}

#define GRID_DISPATCH(KERNEL, ...)                                                                      \
    do {                                                                                                 \
        const uint32_t key = g.n_dims * 16 + g.n_features;                                               \
        switch (key) {                                                                                   \
            case 2 * 16 + 1: KERNEL<2, 1><<<blocks, 256, 0, s>>>(__VA_ARGS__); break;                   \
            case 2 * 16 + 2: KERNEL<2, 2><<<blocks, 256, 0, s>>>(__VA_ARGS__); break;                   \
            case 2 * 16 + 4: KERNEL<2, 4><<<blocks, 256, 0, s>>>(__VA_ARGS__); break;                   \
            case 2 * 16 + 8: KERNEL<2, 8><<<blocks, 256, 0, s>>>(__VA_ARGS__); break;                   \
            case 3 * 16 + 1: KERNEL<3, 1><<<blocks, 256, 0, s>>>(__VA_ARGS__); break;                   \
            case 3 * 16 + 2: KERNEL<3, 2><<<blocks, 256, 0, s>>>(__VA_ARGS__); break;                   \
            case 3 * 16 + 4: KERNEL<3, 4><<<blocks, 256, 0, s>>>(__VA_ARGS__); break;                   \
            case 3 * 16 + 8: KERNEL<3, 8><<<blocks, 256, 0, s>>>(__VA_ARGS__); break;                   \
            default: throw Error("HashGrid: unsupported (n_dims, n_features_per_level)");               \
        }                                                                                                \
        NGP_POST_LAUNCH();                                                              \
    } while (0)

void grid_forward(cudaStream_t s, const GridLayout& g, const __half* table, const float* x, uint32_t x_stride, uint32_t n,
                  __half* y, int y_layout, const uint32_t* n_dyn, const float* max_level) {
    if (n == 0) return;
    NGP_PROFILE_SCOPE(s, "grid_forward", n);
    const dim3 blocks(div_round_up(n, 256), g.padded_width / 8, 1);
    GRID_DISPATCH(grid_forward_kernel, g, table, x, x_stride, n, y, y_layout, n_dyn, max_level);
}

// runs per warp up to which a level is aggregated (0 = never); NGP_AGG_MAX_RUNS overrides for experiments.  Measured (tools/sweep_agg.sh,
// grid_backward ms per 2^18 samples, synthetic scene / fox): 0: 0.119 / 0.359, 4: 0.109 / 0.216, 8: 0.113 / 0.184, 12: 0.108 / 0.163, 16: 0.126 / 0.141
static bool red_wide() {   // NGP_RED_V4=1: 16-byte group merging.  Measured on B200 (profiles/r02_ab_l2_red.json): 0.086 -> 0.095 ms per 2^18 samples, i.e.
    // SLOWER than aligned-pair merging — the scattered-RED lane rate is the same for 4 / 8 / 16-byte lanes (179 G lanes/s, ngp_microbench_l2_red)
    // and the slot selects cost more issue slots than the 1/6 fewer lanes return.  Off by default, kept for experiments.
    static const bool v = [] { const char* e = getenv("NGP_RED_V4"); return e ? atoi(e) != 0 : false; }();
    return v;
}
static uint32_t agg_max_runs() {
    static const uint32_t v = [] { const char* e = getenv("NGP_AGG_MAX_RUNS"); return e ? (uint32_t)atoi(e) : 12u; }();
    return v;
}

void grid_backward(cudaStream_t s, const GridLayout& g, const float* x, uint32_t x_stride, uint32_t n, const __half* dy,
                   int dy_layout, __half* grad, const float* max_level) {
    if (n == 0) return;
    NGP_PROFILE_SCOPE(s, "grid_backward", n);
    const dim3 blocks(div_round_up(n, 256), g.padded_width / 8, 1);
    GRID_DISPATCH(grid_backward_kernel, g, x, x_stride, n, dy, dy_layout, grad, 0u, agg_max_runs(), red_wide(), max_level);
}

// the same pass restricted to 8-feature chunks [chunk0, chunk0 + n_chunks): the data-parallel step launches it chunk by chunk
// so that each chunk's slice of the gradient table can be all-reduced while the next chunk is still being scattered
void grid_backward_chunks(cudaStream_t s, const GridLayout& g, const float* x, uint32_t x_stride, uint32_t n, const __half* dy,
                          int dy_layout, __half* grad, uint32_t chunk0, uint32_t n_chunks, const float* max_level) {
    if (n == 0 || n_chunks == 0) return;
    NGP_PROFILE_SCOPE(s, "grid_backward", n);
    const dim3 blocks(div_round_up(n, 256), n_chunks, 1);
    GRID_DISPATCH(grid_backward_kernel, g, x, x_stride, n, dy, dy_layout, grad, chunk0, agg_max_runs(), red_wide(), max_level);
}

void grid_backward_input(cudaStream_t s, const GridLayout& g, const __half* table, const float* x, uint32_t x_stride,
                         uint32_t n, const __half* dy, int dy_layout, float* dx, uint32_t dx_stride, const float* max_level) {
    if (n == 0) return;
    const uint32_t blocks = div_round_up(n, 256);
    GRID_DISPATCH(grid_backward_input_kernel, g, table, x, x_stride, n, dy, dy_layout, dx, dx_stride ? dx_stride : g.n_dims, max_level);
}

void grid_corners(cudaStream_t s, const GridLayout& g, const float* x, uint32_t x_stride, uint32_t n, uint32_t level,
                  uint32_t* idx, float* w) {
    NGP_CHECK(level < g.n_levels, "grid_corners: level out of range");
    if (n == 0) return;
    if (g.n_dims == 2) grid_corners_kernel<2><<<div_round_up(n, 256), 256, 0, s>>>(g, x, x_stride, n, level, idx, w);
    else grid_corners_kernel<3><<<div_round_up(n, 256), 256, 0, s>>>(g, x, x_stride, n, level, idx, w);
    NGP_POST_LAUNCH();
}

void grid_init(cudaStream_t s, const GridLayout& g, Pcg32 rng, float* params) {
    const uint64_t n = g.n_params;
    grid_init_kernel<<<(uint32_t)((n / 4 + 255) / 256 + 1), 256, 0, s>>>(n, rng, params);
    NGP_POST_LAUNCH();
}

}  // namespace ngp

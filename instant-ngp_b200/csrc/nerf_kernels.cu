// NeRF-specific kernels for sm_100a: occupancy grid maintenance, training ray marcher with deterministic
// (prefix-scan) sample compaction, volume-rendering loss + analytic gradients, wavefront inference tracer.
// Each kernel cites the reference kernel it replaces (src/testbed_nerf.cu).  The reference compacts with global
// atomicAdd (nondeterministic sample order); here block-level warp-shuffle scans + a one-block scan of
// the block totals give the same set of samples in ray order, so results are reproducible and comparable per slot.
#include "nerf_kernels.h"
#include <cstdlib>

namespace ngp {

constexpr uint32_t SCAN_BLOCK = 256;

// exclusive scan of one value per thread over a 256-thread block; returns the prefix, `total` = block sum (all threads)
NGP_D uint32_t block_exclusive_scan(uint32_t v, uint32_t& total) {
    __shared__ uint32_t warp_sums[SCAN_BLOCK / 32];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t n = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += n; }
    if (lane == 31) warp_sums[warp] = inc;
    __syncthreads();
    uint32_t ws = (lane < SCAN_BLOCK / 32) ? warp_sums[lane] : 0;
#pragma unroll
    for (int o = 1; o < SCAN_BLOCK / 32; o <<= 1) { uint32_t n = __shfl_up_sync(0xffffffffu, ws, o); if (lane >= o) ws += n; }
    const uint32_t warp_prefix = __shfl_sync(0xffffffffu, ws, warp) - __shfl_sync(0xffffffffu, (lane < SCAN_BLOCK / 32) ? warp_sums[lane] : 0, warp);
    total = __shfl_sync(0xffffffffu, ws, SCAN_BLOCK / 32 - 1);
    __syncthreads();
    return warp_prefix + inc - v;
}

// one block: in-place exclusive scan of up to 2 arrays of block totals; grand totals written to out0/out1 (optional)
NGP_D void scan_block_totals(uint32_t n, uint32_t* a, uint32_t* b, uint32_t* out_a, uint32_t* out_b, bool accumulate_b) {
    __shared__ uint32_t sh[2][32];
    uint32_t carry_a = 0, carry_b = 0;
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (uint32_t base = 0; base < n; base += 1024) {
        const uint32_t i = base + threadIdx.x;
        uint32_t va = (i < n) ? a[i] : 0, vb = (b && i < n) ? b[i] : 0;
        uint32_t ia = va, ib = vb;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t na = __shfl_up_sync(0xffffffffu, ia, o), nb = __shfl_up_sync(0xffffffffu, ib, o);
            if (lane >= o) { ia += na; ib += nb; }
        }
        if (lane == 31) { sh[0][warp] = ia; sh[1][warp] = ib; }
        __syncthreads();
        uint32_t wa = sh[0][lane], wb = sh[1][lane];
        uint32_t sa = wa, sb = wb;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t na = __shfl_up_sync(0xffffffffu, sa, o), nb = __shfl_up_sync(0xffffffffu, sb, o);
            if (lane >= o) { sa += na; sb += nb; }
        }
        const uint32_t pa = __shfl_sync(0xffffffffu, sa - wa, warp), pb = __shfl_sync(0xffffffffu, sb - wb, warp);
        const uint32_t ta = __shfl_sync(0xffffffffu, sa, 31), tb = __shfl_sync(0xffffffffu, sb, 31);
        if (i < n) { a[i] = carry_a + pa + ia - va; if (b) b[i] = carry_b + pb + ib - vb; }
        carry_a += ta; carry_b += tb;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        if (out_a) *out_a = carry_a;
        if (out_b) { if (accumulate_b) *out_b += carry_b; else *out_b = carry_b; }
    }
}

__global__ void __launch_bounds__(1024) scan_block_totals_kernel(uint32_t n, uint32_t* a, uint32_t* b, uint32_t* out_a, uint32_t* out_b, bool accumulate_b) {
    scan_block_totals(n, a, b, out_a, out_b, accumulate_b);
}
// the same over ceil(dyn->n_rays / rays_per_block) block totals (device-side step parameters)
__global__ void __launch_bounds__(1024) scan_block_totals_dyn_kernel(const StepParams* __restrict__ dyn, uint32_t rays_per_block, uint32_t* a, uint32_t* b, uint32_t* out_a) {
    scan_block_totals((dyn->n_rays + rays_per_block - 1) / rays_per_block, a, b, out_a, nullptr, false);
}
NGP_D uint32_t dyn_rays(const SamplerBuffers& b, uint32_t n) { return b.dyn ? b.dyn->n_rays : n; }
NGP_D uint32_t dyn_samples(const SamplerBuffers& b, uint32_t m) { return b.dyn ? b.dyn->max_inference : m; }
NGP_D uint32_t dyn_scratch(const SamplerBuffers& b) { return b.dyn ? b.dyn->scratch_per_ray : b.scratch_per_ray; }

// NerfCounters::update_after_training + the inference size of the next step (testbed_nerf.cu:3583-3609,3900-3930), on the device; the same
// fp32 operations in the same order as Testbed::train_end evaluates on the host (no contraction: --fmad=false), hence the same integers
__global__ void control_law_kernel(const uint32_t* __restrict__ counters, uint32_t rays_per_batch, uint32_t target, uint32_t rays_bound, uint32_t samples_bound,
                                   StepParams* __restrict__ out) {
    const uint32_t before = counters[1], measured = counters[2];
    StepParams p{0u, 0u, 0u, 1u};
    if (before != 0u && measured != 0u) {
        uint32_t r = (uint32_t)((float)rays_per_batch * (float)target / (float)measured);
        r = min((r + BATCH_GRANULARITY - 1u) / BATCH_GRANULARITY * BATCH_GRANULARITY, 1u << 18);
        const uint32_t max_samples = target * 16u;
        const uint32_t mi = (min(before, max_samples) + BATCH_GRANULARITY - 1u) / BATCH_GRANULARITY * BATCH_GRANULARITY;
        if (r != 0u && r <= rays_bound && mi <= samples_bound) { p.n_rays = r; p.max_inference = mi; p.scratch_per_ray = min(NERF_STEPS, max_samples / r); p.overflow = 0u; }
    }
    *out = p;
}
void nerf_control_law(cudaStream_t s, const uint32_t* counters, uint32_t rays_per_batch, uint32_t target, uint32_t rays_bound, uint32_t samples_bound, StepParams* out) {
    control_law_kernel<<<1, 1, 0, s>>>(counters, rays_per_batch, target, rays_bound, samples_bound, out);
    NGP_POST_LAUNCH();
}

// =================================================================================================================
// Occupancy grid
// =================================================================================================================
// mark_untrained_density_grid (testbed_nerf.cu:503-592)
__global__ void mark_untrained_kernel(uint32_t n_elements, float* __restrict__ grid, uint32_t n_images, const ImageMeta* __restrict__ metas, bool clear_visible) {
    const uint32_t i = threadIdx.x + blockIdx.x * blockDim.x;
    if (i >= n_elements) return;
    if (grid[i] == -1.0f) return;
    const uint32_t level = i / NERF_GRID_N_CELLS, pos_idx = i % NERF_GRID_N_CELLS;
    const uint32_t x = morton3D_invert(pos_idx >> 0), y = morton3D_invert(pos_idx >> 1), z = morton3D_invert(pos_idx >> 2);
    const float voxel_size = scalbnf(1.0f / NERF_GRIDSIZE, (int)level);
    const V3 pos = (v3((float)x, (float)y, (float)z) * (1.0f / (float)NERF_GRIDSIZE) - v3(0.5f, 0.5f, 0.5f)) * scalbnf(1.0f, (int)level) + v3(0.5f, 0.5f, 0.5f);
    uint32_t count = 0;
    for (uint32_t j = 0; j < n_images && count < 1; ++j) {
        const ImageMeta& m = metas[j];
        const V3 campos = v3(m.xform[9], m.xform[10], m.xform[11]), fwd = v3(m.xform[6], m.xform[7], m.xform[8]);
        for (uint32_t k = 0; k < 8; ++k) {
            const V3 corner = pos + v3((k & 1) ? voxel_size : 0.f, (k & 2) ? voxel_size : 0.f, (k & 4) ? voxel_size : 0.f);
            const V3 dir = normalize3(corner - campos);
            if (dot3(dir, fwd) < 1e-4f) continue;
            float uv[2]; pos_to_uv(m, corner, uv);
            V3 ro, rd; uv_to_ray(m.xform, m.res, m.focal_length, m.principal_point, m.lens_mode, m.lens_params, uv[0], uv[1], ro, rd);
            const V3 rdn = normalize3(rd);
            if (uv[0] > 0.0f && uv[1] > 0.0f && uv[0] < 1.0f && uv[1] < 1.0f && length3(rdn - dir) < 1e-3f) { ++count; break; }
        }
    }
    if (clear_visible || (grid[i] < 0) != (count < 1)) grid[i] = (count >= 1) ? 1.f : -1.f;
    else grid[i] = 1.0f;
}

// generate_grid_samples_nerf_nonuniform (testbed_nerf.cu:635-676)
__global__ void grid_samples_kernel(uint32_t n_elements, Pcg32 rng, uint32_t step, Aabb aabb, const float* __restrict__ grid_in,
                                    float* __restrict__ pos_out, uint32_t* __restrict__ indices, uint32_t n_cascades, float thresh) {
    const uint32_t i = threadIdx.x + blockIdx.x * blockDim.x;
    if (i >= n_elements) return;
    rng.advance((int64_t)((uint64_t)i * 4));
    const uint32_t level = (uint32_t)(rng.next_float() * n_cascades) % n_cascades;
    uint32_t idx = 0;
    for (int j = 0; j < 10; ++j) {
        idx = ((i + step * n_elements) * 56924617u + (uint32_t)j * 19349663u + 96925573u) % NERF_GRID_N_CELLS;
        idx += level * NERF_GRID_N_CELLS;
        if (grid_in[idx] > thresh) break;
    }
    const uint32_t pos_idx = idx % NERF_GRID_N_CELLS;
    const uint32_t x = morton3D_invert(pos_idx >> 0), y = morton3D_invert(pos_idx >> 1), z = morton3D_invert(pos_idx >> 2);
    const float r0 = rng.next_float(), r1 = rng.next_float(), r2 = rng.next_float();
    const V3 pos = ((v3((float)x, (float)y, (float)z) + v3(r0, r1, r2)) * (1.0f / (float)NERF_GRIDSIZE) - v3(0.5f, 0.5f, 0.5f)) * scalbnf(1.0f, (int)level) + v3(0.5f, 0.5f, 0.5f);
    const V3 w = warp_position(pos, aabb);
    pos_out[(size_t)i * 3] = w.x; pos_out[(size_t)i * 3 + 1] = w.y; pos_out[(size_t)i * 3 + 2] = w.z;
    indices[i] = idx;
}

// splat_grid_samples_nerf_max_nearest_neighbor (testbed_nerf.cu:678-702)
__global__ void splat_kernel(uint32_t n, const uint32_t* __restrict__ indices, const __half* __restrict__ density_raw, float* __restrict__ grid_tmp, int dact) {
    const uint32_t i = threadIdx.x + blockIdx.x * blockDim.x;
    if (i >= n) return;
    const float mlp = network_to_density(__half2float(density_raw[i]), dact);
    const float optical_thickness = mlp * scalbnf(MIN_CONE_STEPSIZE(), 0);
    atomicMax(reinterpret_cast<uint32_t*>(&grid_tmp[indices[i]]), __float_as_uint(optical_thickness));
}

// ema_grid_samples_nerf (testbed_nerf.cu:731-754)
__global__ void ema_kernel(uint32_t n, float decay, float* __restrict__ grid, const float* __restrict__ tmp) {
    const uint32_t i = threadIdx.x + blockIdx.x * blockDim.x;
    if (i >= n) return;
    const float p = grid[i];
    grid[i] = (p < 0.f) ? p : fmaxf(p * decay, tmp[i]);
}

// deterministic sum: 1024 block partials (fixed tree order), then one block adds the partials in fixed order
template <bool MEAN_CLAMP>
__global__ void __launch_bounds__(256) partial_sum_kernel(const float* __restrict__ in, uint32_t n, float scale, float* __restrict__ partials) {
    __shared__ float sh[256];
    float acc = 0.f;
    for (uint32_t i = blockIdx.x * 256 + threadIdx.x; i < n; i += gridDim.x * 256) acc += MEAN_CLAMP ? fmaxf(in[i], 0.f) * scale : in[i];
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (uint32_t s = 128; s > 0; s >>= 1) { if (threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s]; __syncthreads(); }
    if (threadIdx.x == 0) partials[blockIdx.x] = sh[0];
}
__global__ void __launch_bounds__(1024) final_sum_kernel(const float* __restrict__ partials, uint32_t n, float* __restrict__ out) {
    __shared__ float sh[1024];
    sh[threadIdx.x] = threadIdx.x < n ? partials[threadIdx.x] : 0.f;
    __syncthreads();
    for (uint32_t s = 512; s > 0; s >>= 1) { if (threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s]; __syncthreads(); }
    if (threadIdx.x == 0) *out = sh[0];
}

// grid_to_bitfield (testbed_nerf.cu:762-786)
__global__ void grid_to_bitfield_kernel(uint32_t n_elements, uint32_t n_nonzero, const float* __restrict__ grid, uint8_t* __restrict__ bitfield,
                                        const float* __restrict__ mean_density) {
    const uint32_t i = threadIdx.x + blockIdx.x * blockDim.x;
    if (i >= n_elements) return;
    if (i >= n_nonzero) { bitfield[i] = 0; return; }
    const float thresh = fminf(NERF_MIN_OPTICAL_THICKNESS, *mean_density);
    const float4 a = reinterpret_cast<const float4*>(grid)[(size_t)i * 2], b = reinterpret_cast<const float4*>(grid)[(size_t)i * 2 + 1];
    uint8_t bits = 0;
    bits |= a.x > thresh ? 1 : 0; bits |= a.y > thresh ? 2 : 0; bits |= a.z > thresh ? 4 : 0; bits |= a.w > thresh ? 8 : 0;
    bits |= b.x > thresh ? 16 : 0; bits |= b.y > thresh ? 32 : 0; bits |= b.z > thresh ? 64 : 0; bits |= b.w > thresh ? 128 : 0;
    bitfield[i] = bits;
}

// bitfield_max_pool (testbed_nerf.cu:788-809)
__global__ void bitfield_max_pool_kernel(uint32_t n_elements, const uint8_t* __restrict__ prev, uint8_t* __restrict__ next) {
    const uint32_t i = threadIdx.x + blockIdx.x * blockDim.x;
    if (i >= n_elements) return;
    const uint2 p = reinterpret_cast<const uint2*>(prev)[i];
    uint8_t bits = 0;
#pragma unroll
    for (uint32_t j = 0; j < 8; ++j) {
        const uint32_t byte = ((j < 4 ? p.x : p.y) >> (8 * (j & 3))) & 0xFFu;
        bits |= byte > 0 ? (uint8_t)(1u << j) : 0;
    }
    const uint32_t x = morton3D_invert(i >> 0) + NERF_GRIDSIZE / 8, y = morton3D_invert(i >> 1) + NERF_GRIDSIZE / 8, z = morton3D_invert(i >> 2) + NERF_GRIDSIZE / 8;
    next[morton3D(x, y, z)] |= bits;
}

void nerf_mark_untrained_grid(cudaStream_t s, uint32_t n, float* grid, uint32_t n_images, const ImageMeta* metas, bool clear_visible) {
    mark_untrained_kernel<<<div_round_up(n, 128), 128, 0, s>>>(n, grid, n_images, metas, clear_visible);
    NGP_POST_LAUNCH();
}
void nerf_generate_grid_samples(cudaStream_t s, uint32_t n, Pcg32 rng, uint32_t step, Aabb aabb, const float* grid_in, float* pos_out,
                                uint32_t* indices, uint32_t n_cascades, float thresh) {
    if (n == 0) return;
    grid_samples_kernel<<<div_round_up(n, 256), 256, 0, s>>>(n, rng, step, aabb, grid_in, pos_out, indices, n_cascades, thresh);
    NGP_POST_LAUNCH();
}
void nerf_splat_grid_samples(cudaStream_t s, uint32_t n, const uint32_t* indices, const __half* density_raw, float* grid_tmp, int dact) {
    if (n == 0) return;
    splat_kernel<<<div_round_up(n, 256), 256, 0, s>>>(n, indices, density_raw, grid_tmp, dact);
    NGP_POST_LAUNCH();
}
void nerf_ema_grid(cudaStream_t s, uint32_t n, float decay, float* grid, const float* tmp) {
    ema_kernel<<<div_round_up(n, 256), 256, 0, s>>>(n, decay, grid, tmp);
    NGP_POST_LAUNCH();
}
void reduce_sum_f32(cudaStream_t s, const float* in, uint32_t n, float* out_dev, float* partials) {
    partial_sum_kernel<false><<<1024, 256, 0, s>>>(in, n, 1.0f, partials); NGP_POST_LAUNCH();
    final_sum_kernel<<<1, 1024, 0, s>>>(partials, 1024, out_dev); NGP_POST_LAUNCH();
}
void nerf_update_bitfield(cudaStream_t s, const float* grid, uint32_t n_cascades, uint8_t* bitfield, float* mean_dev, float* partials) {
    // update_density_grid_mean_and_bitfield (testbed_nerf.cu:3538-3567): mean of max(v,0) over cascade 0
    partial_sum_kernel<true><<<1024, 256, 0, s>>>(grid, NERF_GRID_N_CELLS, 1.0f / (float)NERF_GRID_N_CELLS, partials); NGP_POST_LAUNCH();
    final_sum_kernel<<<1, 1024, 0, s>>>(partials, 1024, mean_dev); NGP_POST_LAUNCH();
    const uint32_t n_bytes = NERF_GRID_N_CELLS / 8 * NERF_CASCADES, n_nonzero = NERF_GRID_N_CELLS / 8 * n_cascades;
    grid_to_bitfield_kernel<<<div_round_up(n_bytes, 256), 256, 0, s>>>(n_bytes, n_nonzero, grid, bitfield, mean_dev); NGP_POST_LAUNCH();
    for (uint32_t level = 1; level < NERF_CASCADES; ++level) {
        bitfield_max_pool_kernel<<<div_round_up(NERF_GRID_N_CELLS / 64, 256), 256, 0, s>>>(NERF_GRID_N_CELLS / 64, bitfield + (size_t)NERF_GRID_N_CELLS * (level - 1) / 8,
                                                                                          bitfield + (size_t)NERF_GRID_N_CELLS * level / 8);
        NGP_POST_LAUNCH();
    }
}


// ---- dilated coarse occupancy (ray classification) -----------------------------------------------------------------
constexpr uint32_t COARSE_RES = 32, COARSE_MIP = 2;   // 32^3 cells of 4^3 fine cells each = mip 2 of the bitfield pyramid
__global__ void coarse_occupancy_kernel(const uint8_t* __restrict__ bitfield, uint32_t* __restrict__ coarse) {
    const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;            // x fastest
    const int x = c % COARSE_RES, y = (c / COARSE_RES) % COARSE_RES, z = c / (COARSE_RES * COARSE_RES);
    bool occ = false;
    for (int dz = -1; dz <= 1; ++dz) for (int dy = -1; dy <= 1; ++dy) for (int dx = -1; dx <= 1; ++dx) {
        const int xx = x + dx, yy = y + dy, zz = z + dz;
        if (xx < 0 || yy < 0 || zz < 0 || xx >= (int)COARSE_RES || yy >= (int)COARSE_RES || zz >= (int)COARSE_RES) continue;
        const V3 centre = v3(((float)xx + 0.5f) / COARSE_RES, ((float)yy + 0.5f) / COARSE_RES, ((float)zz + 0.5f) / COARSE_RES);
        occ = occ || density_grid_occupied_at(centre, bitfield, COARSE_MIP);
    }
    const uint32_t word = __ballot_sync(0xffffffffu, occ);
    if ((threadIdx.x & 31u) == 0) coarse[c >> 5] = word;
}
void nerf_build_coarse_occupancy(cudaStream_t s, const uint8_t* bitfield, uint32_t* coarse) {
    coarse_occupancy_kernel<<<COARSE_RES * COARSE_RES * COARSE_RES / 256, 256, 0, s>>>(bitfield, coarse);
    NGP_POST_LAUNCH();
}
// Conservative test "can the exact walk of this ray ever stand in an occupied cell?": the ray is sampled every 0.05 inside
// the aabb (any point of it is within 0.025 of a sample), the coarse cells are 1/32 = 0.03125 wide and dilated by one cell,
// so a walk position in an occupied fine cell always has a sample in a set coarse cell.  false => the walk yields 0 samples.
NGP_D bool ray_touches_coarse(const uint32_t* __restrict__ coarse, V3 o, V3 d, float tmin, float tmax, uint32_t part, uint32_t n_parts) {
    if (!(tmax >= tmin)) return true;    // degenerate interval / NaN: let the walk decide
    const float t0 = tmin - 0.01f, t1 = tmax + 0.01f;
    const uint32_t n = (uint32_t)fminf(ceilf((t1 - t0) * 20.0f), 64.0f) + 1u;
    if (n > 64u) return true;
    for (uint32_t k = part; k < n; k += n_parts) {     // the samples are dealt out to the n_parts lanes that share the ray
        const float t = fminf(t0 + (float)k * 0.05f, t1);
        const V3 p = ray_at(o, d, t);
        const uint32_t x = (uint32_t)clampi((int)floorf(p.x * COARSE_RES), 0, COARSE_RES - 1), y = (uint32_t)clampi((int)floorf(p.y * COARSE_RES), 0, COARSE_RES - 1),
                       z = (uint32_t)clampi((int)floorf(p.z * COARSE_RES), 0, COARSE_RES - 1);
        const uint32_t c = (z * COARSE_RES + y) * COARSE_RES + x;
        if ((__ldg(coarse + (c >> 5)) >> (c & 31u)) & 1u) return true;
    }
    return false;
}

// =================================================================================================================
// Training sampler: generate_training_samples_nerf (testbed_nerf.cu:1382-1658)
//   pass 1 (count)  : pick image + pixel, build the ray, clip to the aabb, march the occupancy bitfield counting samples
//   scan            : exclusive prefix of the per-ray counts (replaces atomicAdd(numsteps_counter, numsteps))
//   pass 2 (write)  : march again, write NerfCoordinate{warped pos, warped dt, warped dir} at base + j
// =================================================================================================================
// lower bound in a CDF, clamped to the last bin (common.h:268-291)
NGP_D uint32_t cdf_binary_search(float val, const float* __restrict__ data, uint32_t length) {
    if (length == 0) return 0;
    uint32_t first = 0, count = length;
    while (count > 0) {
        const uint32_t step = count / 2, it = first + step;
        if (data[it] < val) { first = it + 1; count -= step + 1; } else count = step;
    }
    return min(first, length - 1);
}
// image_idx (:1317-1338): views in ray order, or - sample_image_proportional_to_error - drawn from the view CDF with a low-discrepancy number
NGP_D uint32_t image_idx(uint32_t base_idx, uint32_t n_rays, uint32_t n_images, const float* __restrict__ cdf_img = nullptr, float* pdf = nullptr) {
    if (cdf_img) {
        const float sample = ld_random_val(base_idx, 0xdeadbeefu);
        const uint32_t img = cdf_binary_search(sample, cdf_img, n_images);
        if (pdf) { const float prev = img > 0 ? cdf_img[img - 1] : 0.0f; *pdf = (cdf_img[img] - prev) * (float)n_images; }
        return img;
    }
    if (pdf) *pdf = 1.0f;
    return ((base_idx * n_images) / n_rays) % n_images;
}
// sample_cdf_2d (:1232-1263): half of the samples stay uniform, the other half picks a row from cdf_y and a column from that row's CDF
NGP_D void sample_cdf_2d(float* uv, uint32_t img, const ErrorMapCdf& c, float* pdf) {
    const float UNIFORM_SAMPLING_FRACTION = 0.5f;
    if (uv[0] < UNIFORM_SAMPLING_FRACTION) { uv[0] /= UNIFORM_SAMPLING_FRACTION; return; }   // (pdf keeps the caller's 1: the reference leaves it untouched here)
    uv[0] = (uv[0] - UNIFORM_SAMPLING_FRACTION) / (1.0f - UNIFORM_SAMPLING_FRACTION);
    const uint32_t rx = (uint32_t)c.res[0], ry = (uint32_t)c.res[1];
    const float* cy = c.cdf_y + (size_t)img * ry;
    const uint32_t y = cdf_binary_search(uv[1], cy, ry);
    float prev = y > 0 ? cy[y - 1] : 0.0f;
    const float pmf_y = cy[y] - prev;
    uv[1] = (uv[1] - prev) / pmf_y;
    const float* cx = c.cdf_x_cond_y + ((size_t)img * ry + y) * rx;
    const uint32_t x = cdf_binary_search(uv[0], cx, rx);
    prev = x > 0 ? cx[x - 1] : 0.0f;
    const float pmf_x = cx[x] - prev;
    uv[0] = (uv[0] - prev) / pmf_x;
    if (pdf) *pdf = pmf_x * pmf_y * (float)(rx * ry);
    uv[0] = ((float)x + uv[0]) / (float)rx; uv[1] = ((float)y + uv[1]) / (float)ry;
}
NGP_D void random_image_pos_training(Pcg32& rng, const int* res, bool snap, float* uv, const ErrorMapCdf* cdf = nullptr, uint32_t img = 0, float* pdf = nullptr) {   // :1291-1315
    uv[0] = rng.next_float(); uv[1] = rng.next_float();
    if (pdf) *pdf = 1.0f;
    if (cdf && cdf->cdf_x_cond_y) sample_cdf_2d(uv, img, *cdf, pdf);
    if (snap) {
        uv[0] = ((float)clampi((int)(uv[0] * (float)res[0]), 0, res[0] - 1) + 0.5f) / (float)res[0];
        uv[1] = ((float)clampi((int)(uv[1] * (float)res[1]), 0, res[1] - 1) + 0.5f) / (float)res[1];
    }
}

// Per-candidate arithmetic of the training walk.  FLAT = (cone_angle <= 1e-5 and max_cascade == 0), the aabb_scale 1 case:
// the mip is always 0, the voxel resolution is the constant 128 and to/from_stepping_space are t / S and n * S, so the
// generic helpers collapse to the expressions below — the same float operations in the same order (the division by the
// power-of-two resolution is written as the exact multiplication by its inverse), hence the same bits.
// (div_by_stepsize: nerf_device.cuh)
// where the counting walk leaves (t, dt) of the samples it finds: a fixed window per ray plus, for the few rays that
// outgrow it, a whole-ray slot (NERF_STEPS entries) taken from a pool with one atomicAdd
struct SampleLog {
    float2* window; uint32_t cap; float2* pool; uint32_t pool_slots; uint32_t* pool_counter; uint32_t slot;
    NGP_D float2* at(uint32_t k) const { return k < cap ? window + k : (slot < pool_slots ? pool + (size_t)slot * NERF_STEPS + k : nullptr); }
};
constexpr uint32_t NO_SLOT = 0xFFFFFFFFu;
struct RayConst { V3 o, d, idir, hs; };   // hs = 0.5 * sign(d): the +-0.5 of distance_to_next_voxel
NGP_D RayConst make_ray_const(V3 o, V3 d) {
    return {o, d, v3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z), v3(0.5f * copysignf(1.0f, d.x), 0.5f * copysignf(1.0f, d.y), 0.5f * copysignf(1.0f, d.z))};
}
template <bool FLAT>
NGP_D void candidate_dt_mip(float t, V3 pos, const Cone& cone, uint32_t max_mip, float& dt, uint32_t& mip, float& n) {
    if (FLAT) { n = div_by_stepsize(t); dt = (n + 1.0f) * MIN_CONE_STEPSIZE() - t; mip = 0; }
    else { n = 0.f; dt = calc_dt(t, cone); mip = mip_from_dt(dt, pos, max_mip); }
}
template <bool FLAT>
NGP_D float candidate_skip(float t, float n, V3 pos, const RayConst& r, const Cone& cone, uint32_t mip) {
    if (FLAT) {
        const V3 p = (float)NERF_GRIDSIZE * (pos - v3(0.5f, 0.5f, 0.5f));
        const float tx = (floorf(p.x + 0.5f + r.hs.x) - p.x) * r.idir.x;
        const float ty = (floorf(p.y + 0.5f + r.hs.y) - p.y) * r.idir.y;
        const float tz = (floorf(p.z + 0.5f + r.hs.z) - p.z) * r.idir.z;
        const float dist = fmaxf(fminf(fminf(tx, ty), tz) * (1.0f / (float)NERF_GRIDSIZE), 0.0f);
        const float nt = div_by_stepsize(t + dist);
        return (n + ceilf(fmaxf(nt - n, 0.5f))) * MIN_CONE_STEPSIZE();
    }
    return advance_to_next_voxel(t, cone, pos, r.d, r.idir, mip);
}

template <bool WRITE>
NGP_D uint32_t march_training_ray(V3 o, V3 d, float startt, const Cone& cone, const Aabb& aabb, const uint8_t* __restrict__ bitfield, uint32_t max_mip,
                                  uint32_t limit, float* __restrict__ coords) {
    const V3 idir = v3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    const V3 wd = warp_direction(d);
    uint32_t j = 0; float t = startt; V3 pos;
    while (aabb_contains(aabb, pos = ray_at(o, d, t)) && j < limit) {
        const float dt = calc_dt(t, cone);
        const uint32_t mip = mip_from_dt(dt, pos, max_mip);
        if (density_grid_occupied_at(pos, bitfield, mip)) {
            if (WRITE) {
                float* c = coords + (size_t)j * 7;
                const V3 wp = warp_position(pos, aabb);
                c[0] = wp.x; c[1] = wp.y; c[2] = wp.z; c[3] = warp_dt(dt); c[4] = wd.x; c[5] = wd.y; c[6] = wd.z;
            }
            ++j; t += dt;
        } else {
            t = advance_to_next_voxel(t, cone, pos, d, idir, mip);
        }
    }
    return j;
}

// ---- one thread per ray, K look-ups in flight -----------------------------------------------------------------------
// Same run speculation as the group marcher, but inside one thread: the chain is advanced K candidates ahead (registers),
// the K occupancy bytes are fetched as independent loads (one exposed L2 latency per K candidates instead of one per
// candidate), then resolved in order.  No redundant chain work, no cross-lane traffic; bit-identical by construction.
template <bool WRITE, bool FLAT, uint32_t K>
NGP_D uint32_t march_training_ray_ilp(V3 o, V3 d, float startt, const Cone& cone, const Aabb& aabb, const uint8_t* __restrict__ bitfield, uint32_t max_mip,
                                      uint32_t limit, float* __restrict__ coords, SampleLog* log = nullptr) {
    if (limit == 0) return 0;
    const RayConst rc = make_ray_const(o, d);
    const V3 wd = warp_direction(d);
    uint32_t j = 0; float t = startt; bool step_mode = false;
    while (true) {
        float ts[K], dts[K]; uint32_t cell[K]; uint32_t n_valid = K;
#pragma unroll
        for (uint32_t k = 0; k < K; ++k) {
            const V3 pos = ray_at(o, d, t);
            if (n_valid == K && !aabb_contains(aabb, pos)) n_valid = k;
            float dt, n; uint32_t mip;
            candidate_dt_mip<FLAT>(t, pos, cone, max_mip, dt, mip, n);
            const uint32_t idx = cascaded_grid_idx_at(pos, mip);
            ts[k] = t; dts[k] = dt; cell[k] = idx == 0xFFFFFFFFu ? idx : idx + mip * NERF_GRID_N_CELLS;
            t = step_mode ? t + dt : candidate_skip<FLAT>(t, n, pos, rc, cone, mip);
        }
        uint32_t occ_mask = 0;
#pragma unroll
        for (uint32_t k = 0; k < K; ++k) {
            const bool look = k < n_valid && cell[k] != 0xFFFFFFFFu;
            const uint32_t byte = look ? (uint32_t)__ldg(bitfield + (cell[k] >> 3)) : 0u;
            occ_mask |= ((byte >> (cell[k] & 7u)) & 1u) << k;
        }
        if (step_mode) {
            const uint32_t bad = ~occ_mask & ((1u << K) - 1u);
            const uint32_t f = bad ? (uint32_t)__ffs((int)bad) - 1u : K;
            const uint32_t n_emit = min(f, limit - j);
            if (WRITE) {
#pragma unroll
                for (uint32_t k = 0; k < K; ++k) {
                    if (k < n_emit) {
                        float* c = coords + (size_t)(j + k) * 7;
                        const V3 wp = warp_position(ray_at(o, d, ts[k]), aabb);
                        c[0] = wp.x; c[1] = wp.y; c[2] = wp.z; c[3] = warp_dt(dts[k]); c[4] = wd.x; c[5] = wd.y; c[6] = wd.z;
                    }
                }
            }
            if (!WRITE && n_emit) {
                if (j + n_emit > log->cap && log->slot == NO_SLOT) log->slot = atomicAdd(log->pool_counter, 1u);
#pragma unroll
                for (uint32_t k = 0; k < K; ++k) {
                    float2* dst = k < n_emit ? log->at(j + k) : nullptr;
                    if (dst) *dst = make_float2(ts[k], dts[k]);
                }
            }
            j += n_emit;
            if (j >= limit) break;
            if (f == K) continue;                  // t already is candidate K
            if (f >= n_valid) break;               // candidate f left the aabb
            float tf = ts[0];
#pragma unroll
            for (uint32_t k = 1; k < K; ++k) if (k == f) tf = ts[k];
            const V3 pos = ray_at(o, d, tf);
            float dt, n; uint32_t mip;
            candidate_dt_mip<FLAT>(tf, pos, cone, max_mip, dt, mip, n);
            t = candidate_skip<FLAT>(tf, n, pos, rc, cone, mip);
            step_mode = false;
        } else {
            if (occ_mask) {
                const uint32_t f = (uint32_t)__ffs((int)occ_mask) - 1u;
                float tf = ts[0];
#pragma unroll
                for (uint32_t k = 1; k < K; ++k) if (k == f) tf = ts[k];
                t = tf; step_mode = true; continue;
            }
            if (n_valid < K) break;
        }
    }
    return j;
}

constexpr uint32_t ILP_K = 8;

template <bool FLAT>
__global__ void __launch_bounds__(SCAN_BLOCK) sampler_count_kernel(uint32_t n_rays, Aabb aabb, Pcg32 rng, uint32_t n_images, const ImageMeta* __restrict__ metas,
                                                                   const uint8_t* __restrict__ bitfield, uint32_t max_mip, bool snap, float cone_angle_constant,
                                                                   uint32_t ray_offset, uint32_t n_rays_global, SamplerBuffers b) {
    const uint32_t i = threadIdx.x + blockIdx.x * blockDim.x;
    uint32_t numsteps = 0;
    if (i < n_rays) {
        const uint32_t ig = ray_offset + i;
        const uint32_t img = image_idx(ig, n_rays_global, n_images, b.cdf.cdf_img);
        const ImageMeta& m = metas[img];
        rng.advance((int64_t)((uint64_t)ig * N_MAX_RANDOM_SAMPLES_PER_RAY));
        float uv[2]; random_image_pos_training(rng, m.res, snap, uv, &b.cdf, img);
        float texel[4]; read_rgba(m, uv[0], uv[1], texel);
        if (texel[0] >= 0.0f) {   // negative = masked-away pixel (:1430-1435)
            if (b.max_level_rand) rng.advance(1);             // :1438 the ray's level cut-off (nerf_fill_max_level hands it to the samples)
            const float motionblur_time = rng.next_float();   // :1440
            float xf[12]; xform_given_rolling_shutter(m.xform, m.xform_end, m.rolling_shutter, uv[0], uv[1], motionblur_time, xf);   // :1447-1451
            V3 o, du; uv_to_ray(xf, m.res, m.focal_length, m.principal_point, m.lens_mode, m.lens_params, uv[0], uv[1], o, du, &b.distortion);
            if (du.x == 0.f && du.y == 0.f && du.z == 0.f) { o = v3(xf[9], xf[10], xf[11]); du = v3(xf[6], xf[7], xf[8]); }
            const V3 d = normalize3(du);
            float tmin, tmax; aabb_ray_intersect(aabb, o, d, tmin, tmax);
            const Cone cone(cone_angle_constant);   // calc_cone_angle (:114-122)
            tmin = fmaxf(tmin, 0.0f);
            const float startt = advance_n_steps(tmin, cone, rng.next_float());
            SampleLog log{b.sample_scratch + (size_t)i * b.scratch_per_ray, b.scratch_per_ray, b.sample_pool, b.pool_slots, &b.counters[3], NO_SLOT};
            numsteps = march_training_ray_ilp<false, FLAT, ILP_K>(o, d, startt, cone, aabb, bitfield, max_mip, NERF_STEPS, nullptr, &log);
            b.overflow_slot[i] = log.slot;
            float* rt = b.ray_tmp + (size_t)i * 8;
            rt[0] = o.x; rt[1] = o.y; rt[2] = o.z; rt[3] = du.x; rt[4] = du.y; rt[5] = du.z; rt[6] = startt; rt[7] = cone.angle;
        }
        b.numsteps_all[i] = numsteps;
    }
    uint32_t total_steps, total_rays;
    block_exclusive_scan(numsteps, total_steps);
    block_exclusive_scan(numsteps > 0 ? 1u : 0u, total_rays);
    if (threadIdx.x == 0) { b.block_steps[blockIdx.x] = total_steps; b.block_rays[blockIdx.x] = total_rays; }
}

template <bool FLAT>
__global__ void __launch_bounds__(SCAN_BLOCK) sampler_write_kernel(uint32_t n_rays, Aabb aabb, uint32_t max_samples, const uint8_t* __restrict__ bitfield,
                                                                   uint32_t max_mip, SamplerBuffers b) {
    const uint32_t i = threadIdx.x + blockIdx.x * blockDim.x;
    const uint32_t numsteps = i < n_rays ? b.numsteps_all[i] : 0;
    uint32_t tot;
    const uint32_t base = b.block_steps[blockIdx.x] + block_exclusive_scan(numsteps, tot);
    const uint32_t ray_idx = b.block_rays[blockIdx.x] + block_exclusive_scan(numsteps > 0 ? 1u : 0u, tot);
    const bool keep = numsteps > 0 && base + numsteps <= max_samples;   // :1616-1619
    uint32_t kept_total;
    block_exclusive_scan(keep ? 1u : 0u, kept_total);
    if (threadIdx.x == 0 && kept_total) atomicAdd(&b.counters[0], kept_total);
    if (!keep) return;
    const float* rt = b.ray_tmp + (size_t)i * 8;
    const V3 o = ld3(rt), du = ld3(rt + 3);
    b.ray_indices[ray_idx] = i;
    float* r = b.rays + (size_t)ray_idx * 6;
    r[0] = o.x; r[1] = o.y; r[2] = o.z; r[3] = du.x; r[4] = du.y; r[5] = du.z;
    b.numsteps[ray_idx * 2 + 0] = numsteps; b.numsteps[ray_idx * 2 + 1] = base;
    const V3 d = normalize3(du);
    float* coords = b.coords + (size_t)base * 7;
    const SampleLog log{b.sample_scratch + (size_t)i * b.scratch_per_ray, b.scratch_per_ray, b.sample_pool, b.pool_slots, nullptr,
                        numsteps > b.scratch_per_ray ? b.overflow_slot[i] : NO_SLOT};
    if (numsteps <= b.scratch_per_ray || log.slot < b.pool_slots) {
        const V3 wd = warp_direction(d);
        for (uint32_t k = 0; k < numsteps; ++k) {
            const float2 td = *log.at(k);
            const V3 wp = warp_position(ray_at(o, d, td.x), aabb);
            float* c = coords + (size_t)k * 7;
            c[0] = wp.x; c[1] = wp.y; c[2] = wp.z; c[3] = warp_dt(td.y); c[4] = wd.x; c[5] = wd.y; c[6] = wd.z;
        }
        return;
    }
    march_training_ray_ilp<true, FLAT, ILP_K>(o, d, rt[6], Cone(rt[7]), aabb, bitfield, max_mip, numsteps, coords);
}

// ---- warp-cooperative marcher: G lanes per ray ------------------------------------------------------------------
// The reference walks a ray serially (one thread per ray); t_{k+1} depends on t_k through float rounding, so the chain
// itself cannot be re-associated without changing traversal decisions.  What CAN be taken off the serial path are the
// occupancy look-ups: the successor of a candidate is step(t) if its cell is occupied and skip(t) if it is not, and both
// are pure functions of t.  The G lanes of a group therefore run the chain in lock-step under the assumption that the
// current run (occupied / empty) continues, lane k keeping candidate k, then test all G cells with one ballot.  The
// candidates up to the first misprediction are exactly the serial walk's; the group restarts from there in the other
// mode.  Every float operation is the serial one, so positions, sample counts and coordinates are bit-identical.
// ---- G lanes per ray, K candidates per lane ------------------------------------------------------------------------
// A round speculates C = G*K candidates: the lanes of a group advance the (serial, bit-exact) t chain C steps in lock
// step, under the current hypothesis (step mode: "every candidate is an occupied sample", skip mode: "every candidate is
// empty and the walk jumps to the next voxel"); lane g keeps t of candidates g, G+g, ..; then each lane evaluates its own
// K candidates (position, aabb test, dt, occupancy byte: K independent loads) and ballots tell where the hypothesis first
// fails.  The chain loop carries no aabb test and no capture beyond one select: candidates past the exit are computed and
// ignored.  Warp instructions per ray scale with 1/rays-per-warp = G/32 at a fixed C, which is why K > 1 pays.
// FAST: the round starts at t in [1e-30, 1024); a round moves t by at most C * (one voxel diagonal / 128 + one step) < 1,
// so every t of the round is inside the range where div_by_stepsize takes its branch-free form, and the test is made once
// per round instead of once per division.
NGP_D float div_by_stepsize_in_range(float t) {
    constexpr float S = SQRT3 / (float)NERF_STEPS, RS = 1.0f / S;
    const float q = t * RS; return fmaf(fmaf(-q, S, t), RS, q);
}
// General path, skip candidates: the cell a skip leaves is found from mip_from_dt(dt, pos), and dt = expf((logf(t) / L + 1) L) - t puts a whole
// logf -> expf pair in front of the skip's own logf -> expf on the serial chain.  But the mip depends on dt only through the binary exponent of
// 256 dt, and in the logarithmic regime dt = t * cone_angle up to the rounding of that evaluation: |x| <= 6.1 and n <= 1563 bound the error of the
// exponent argument by 1.3e-6, i.e. dt by 3.4e-4 relative.  So with a 1e-3 margin either side the exponent is decided by t * cone_angle alone unless
// 256 dt lies within 0.1 % of a power of two (or t within 1 % of a regime boundary) — then, and only then, the exact dt is computed.  Same mip, hence
// the same skip, bit for bit (test_sampler_with_cascades_and_cone_angle); the A/B build (make ab, NGP_AB_OLD_MATH) keeps the exact evaluation.
NGP_D bool mip_without_dt(float t, V3 pos, const Cone& c, uint32_t max_mip, uint32_t& mip) {
#ifdef NGP_AB_OLD_MATH
    return false;
#endif
    if (!(t > c.at * 1.01f && t < c.bt * 0.99f)) return false;
    const float d = t * c.angle * (float)(2 * NERF_GRIDSIZE);
    const float lo = d * 0.999f, hi = d * 1.001f;
    mip = mip_from_pos(pos, max_mip);
    if (hi < 1.0f) return true;            // dt * 256 < 1: mip_from_pos decides (:626-628)
    if (lo < 1.0f) return false;
    const int e = frexp_exponent(lo);
    if (e != frexp_exponent(hi)) return false;
    int v = (int)mip; const int top = (int)max_mip;
    v = v < e ? e : (top < v ? top : v);   // tcnn::clamp(mip, e, max_cascade) (:632)
    mip = (uint32_t)v;
    return true;
}
template <bool FLAT, bool STEP, bool FAST>
NGP_D float chain_next(float t, const RayConst& rc, const Cone& cone, uint32_t max_mip) {
    if (FLAT && FAST) {
        const float n = div_by_stepsize_in_range(t);
        if (STEP) return t + ((n + 1.0f) * MIN_CONE_STEPSIZE() - t);
        const V3 pos = ray_at(rc.o, rc.d, t);
        const V3 p = (float)NERF_GRIDSIZE * (pos - v3(0.5f, 0.5f, 0.5f));
        const float tx = (floorf(p.x + 0.5f + rc.hs.x) - p.x) * rc.idir.x;
        const float ty = (floorf(p.y + 0.5f + rc.hs.y) - p.y) * rc.idir.y;
        const float tz = (floorf(p.z + 0.5f + rc.hs.z) - p.z) * rc.idir.z;
        const float dist = fmaxf(fminf(fminf(tx, ty), tz) * (1.0f / (float)NERF_GRIDSIZE), 0.0f);
        const float nt = div_by_stepsize_in_range(t + dist);
        return (n + ceilf(fmaxf(nt - n, 0.5f))) * MIN_CONE_STEPSIZE();
    }
    if (FLAT && STEP) { const float n = div_by_stepsize(t); return t + ((n + 1.0f) * MIN_CONE_STEPSIZE() - t); }
    const V3 pos = ray_at(rc.o, rc.d, t);
    float dt, n; uint32_t mip;
    if (!FLAT && !STEP && mip_without_dt(t, pos, cone, max_mip, mip)) return advance_to_next_voxel(t, cone, pos, rc.d, rc.idir, mip);
    candidate_dt_mip<FLAT>(t, pos, cone, max_mip, dt, mip, n);
    return STEP ? t + dt : candidate_skip<FLAT>(t, n, pos, rc, cone, mip);
}
template <uint32_t G, uint32_t K, bool FLAT, bool STEP, bool FAST>
NGP_D float chain_round(float t, float* my_t, uint32_t lane, const RayConst& rc, const Cone& cone, uint32_t max_mip) {
    // general path (cascades / cone angle): one chain step is ~250 instructions (logf, expf, IEEE divisions, frexp); unrolled
    // eight-fold and instantiated per mode it outgrows the instruction cache, and with ~8 warps per SM nothing hides the misses
    // (ncu: 73 % of the warp cycles in "no instruction").  The loop stays rolled there.
    constexpr int UNROLL = FLAT ? (G < 8 ? (int)G : 8) : 1;
#pragma unroll
    for (uint32_t k = 0; k < K; ++k) {
#pragma unroll UNROLL
        for (uint32_t g = 0; g < G; ++g) {
            if (g == lane) my_t[k] = t;
            t = chain_next<FLAT, STEP, FAST>(t, rc, cone, max_mip);
        }
    }
    return t;
}
template <uint32_t G, bool WRITE, bool FLAT, uint32_t K = 1>
NGP_D uint32_t march_training_ray_group(V3 o, V3 d, float startt, const Cone& cone, const Aabb& aabb, const uint8_t* __restrict__ bitfield, uint32_t max_mip,
                                        uint32_t limit, float* __restrict__ coords, uint32_t lane, uint32_t gshift, uint32_t gmask, SampleLog* log = nullptr) {
    constexpr uint32_t C = G * K;
    static_assert(C <= 32, "at most 32 candidates per round");
    constexpr uint32_t FULL_G = G == 32 ? 0xffffffffu : ((1u << G) - 1u);
    constexpr uint32_t FULL_C = C == 32 ? 0xffffffffu : ((1u << C) - 1u);
    if (limit == 0) return 0;
    const RayConst rc = make_ray_const(o, d);
    const V3 wd = warp_direction(d);
    uint32_t j = 0; float t_base = startt; bool step_mode = false;
    while (true) {
        float my_t[K], my_dt[K];
        float t_end;
        if (FLAT && t_base >= 1e-30f && t_base < 1024.0f)
            t_end = step_mode ? chain_round<G, K, FLAT, true, true>(t_base, my_t, lane, rc, cone, max_mip)
                              : chain_round<G, K, FLAT, false, true>(t_base, my_t, lane, rc, cone, max_mip);
        else
            t_end = step_mode ? chain_round<G, K, FLAT, true, false>(t_base, my_t, lane, rc, cone, max_mip)
                              : chain_round<G, K, FLAT, false, false>(t_base, my_t, lane, rc, cone, max_mip);
        uint32_t in_mask = 0, occ_mask = 0;
#pragma unroll
        for (uint32_t k = 0; k < K; ++k) {
            const V3 pos = ray_at(o, d, my_t[k]);
            const bool inside = aabb_contains(aabb, pos);
            float n; uint32_t mip;
            candidate_dt_mip<FLAT>(my_t[k], pos, cone, max_mip, my_dt[k], mip, n);
            const bool occ = inside && density_grid_occupied_at(pos, bitfield, mip);
            in_mask |= ((__ballot_sync(gmask, inside) >> gshift) & FULL_G) << (k * G);
            occ_mask |= ((__ballot_sync(gmask, occ) >> gshift) & FULL_G) << (k * G);
        }
        const uint32_t out = ~in_mask & FULL_C;
        const uint32_t n_valid = out ? (uint32_t)__ffs((int)out) - 1u : C;    // first candidate outside the aabb ends the walk
        occ_mask &= n_valid == 32 ? 0xffffffffu : ((1u << n_valid) - 1u);
        if (step_mode) {
            const uint32_t bad = ~occ_mask & FULL_C;
            const uint32_t f = bad ? (uint32_t)__ffs((int)bad) - 1u : C;       // first candidate that is not an occupied sample
            const uint32_t n_emit = min(f, limit - j);
            if (WRITE) {
#pragma unroll
                for (uint32_t k = 0; k < K; ++k) {
                    if (k * G + lane < n_emit) {
                        float* c = coords + (size_t)(j + k * G + lane) * 7;
                        const V3 wp = warp_position(ray_at(o, d, my_t[k]), aabb);
                        c[0] = wp.x; c[1] = wp.y; c[2] = wp.z; c[3] = warp_dt(my_dt[k]); c[4] = wd.x; c[5] = wd.y; c[6] = wd.z;
                    }
                }
            }
            if (!WRITE && n_emit) {
                if (j + n_emit > log->cap && log->slot == NO_SLOT) {   // first sample beyond the window: take a pool slot (group-uniform)
                    uint32_t sl = 0;
                    if (lane == 0) sl = atomicAdd(log->pool_counter, 1u);
                    log->slot = __shfl_sync(gmask, sl, gshift);
                }
#pragma unroll
                for (uint32_t k = 0; k < K; ++k) {
                    float2* dst = k * G + lane < n_emit ? log->at(j + k * G + lane) : nullptr;
                    if (dst) *dst = make_float2(my_t[k], my_dt[k]);
                }
            }
            j += n_emit;
            if (j >= limit) break;
            if (f == C) { t_base = t_end; continue; }
            if (f >= n_valid) break;                                           // candidate f left the aabb: end of the walk
            float tf = my_t[0];
#pragma unroll
            for (uint32_t k = 1; k < K; ++k) if (k == f / G) tf = my_t[k];
            const float my_skip = chain_next<FLAT, false, false>(tf, rc, cone, max_mip);
            t_base = __shfl_sync(gmask, my_skip, gshift + f % G);
            step_mode = false;
        } else {
            if (occ_mask) {
                const uint32_t f = (uint32_t)__ffs((int)occ_mask) - 1u;
                float tf = my_t[0];
#pragma unroll
                for (uint32_t k = 1; k < K; ++k) if (k == f / G) tf = my_t[k];
                t_base = __shfl_sync(gmask, tf, gshift + f % G); step_mode = true; continue;
            }
            if (n_valid < C) break;
            t_base = t_end;
        }
    }
    return j;
}

// Pass 0 of the group marchers, one thread per ray: pick image + pixel, build the ray, clip it to the aabb, and decide whether
// it has to be walked at all.  About two thirds of the training rays never meet an occupied cell; the walk would still
// follow their exact t chain through ~160 empty voxels.  Rays that fail the conservative coarse test have 0 samples by
// construction and are not walked; the others are appended to the candidate list (order irrelevant: results are per ray).
__global__ void __launch_bounds__(128) sampler_classify_kernel(uint32_t n_rays, Aabb aabb, Pcg32 rng, uint32_t n_images, const ImageMeta* __restrict__ metas,
                                                               bool snap, float cone_angle_constant, uint32_t ray_offset, uint32_t n_rays_global, uint32_t rays_per_block,
                                                               SamplerBuffers b) {
    constexpr uint32_t Q = 4;                                            // lanes per ray: the coarse test is split four ways
    if (b.dyn) { n_rays = b.dyn->n_rays; n_rays_global = n_rays; }       // device-side step parameters (single rank: offset 0)
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, i = tid / Q, part = tid % Q;
    if (tid < (n_rays + rays_per_block - 1) / rays_per_block) { b.block_steps[tid] = 0; b.block_rays[tid] = 0; }   // totals are accumulated by the walk
    bool candidate = false;
    if (i < n_rays) {   // the Q lanes of a ray build it redundantly (uniform), lane 0 stores
        const uint32_t ig = ray_offset + i;
        const uint32_t img = image_idx(ig, n_rays_global, n_images, b.cdf.cdf_img);
        const ImageMeta& m = metas[img];
        rng.advance((int64_t)((uint64_t)ig * N_MAX_RANDOM_SAMPLES_PER_RAY));
        float uv[2]; random_image_pos_training(rng, m.res, snap, uv, &b.cdf, img);
        float texel[4]; read_rgba(m, uv[0], uv[1], texel);
        if (texel[0] >= 0.0f) {   // negative = masked-away pixel (:1430-1435)
            if (b.max_level_rand) rng.advance(1);             // :1438 the ray's level cut-off (nerf_fill_max_level hands it to the samples)
            const float motionblur_time = rng.next_float();   // :1440
            float xf[12]; xform_given_rolling_shutter(m.xform, m.xform_end, m.rolling_shutter, uv[0], uv[1], motionblur_time, xf);   // :1447-1451
            V3 o, du; uv_to_ray(xf, m.res, m.focal_length, m.principal_point, m.lens_mode, m.lens_params, uv[0], uv[1], o, du, &b.distortion);
            if (du.x == 0.f && du.y == 0.f && du.z == 0.f) { o = v3(xf[9], xf[10], xf[11]); du = v3(xf[6], xf[7], xf[8]); }
            const V3 d = normalize3(du);
            float tmin, tmax; aabb_ray_intersect(aabb, o, d, tmin, tmax);
            const Cone cone(cone_angle_constant);   // calc_cone_angle (:114-122)
            tmin = fmaxf(tmin, 0.0f);
            if (part == 0) {
                const float startt = advance_n_steps(tmin, cone, rng.next_float());
                float* rt = b.ray_tmp + (size_t)i * 8;
                rt[0] = o.x; rt[1] = o.y; rt[2] = o.z; rt[3] = du.x; rt[4] = du.y; rt[5] = du.z; rt[6] = startt; rt[7] = cone.angle;
            }
            candidate = b.coarse == nullptr || ray_touches_coarse(b.coarse, o, d, tmin, tmax, part, Q);
        }
        if (part == 0) { b.numsteps_all[i] = 0; b.overflow_slot[i] = NO_SLOT; }
    }
    // a ray is a candidate if any of its lanes found a set cell; warp-aggregated append by the first lane of each ray
    const uint32_t any = __ballot_sync(0xffffffffu, candidate), lane = threadIdx.x & 31u;
    const bool mine = part == 0 && ((any >> (lane & ~(Q - 1u))) & ((1u << Q) - 1u)) != 0u;
    const uint32_t mask = __ballot_sync(0xffffffffu, mine);
    uint32_t base = 0;
    if (lane == 0 && mask) base = atomicAdd(&b.counters[4], (uint32_t)__popc(mask));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (mine) b.cand_list[base + __popc(mask & ((1u << lane) - 1u))] = i;
}

template <uint32_t G, bool FLAT, uint32_t K>
__global__ void __launch_bounds__(SCAN_BLOCK) sampler_count_group_kernel(uint32_t n_rays, Aabb aabb, const uint8_t* __restrict__ bitfield, uint32_t max_mip, SamplerBuffers b) {
    constexpr uint32_t RPB = SCAN_BLOCK / G;           // rays per block of the ray-order scan (sampler_write_group_kernel), not of this launch
    const uint32_t lane = threadIdx.x % G, gshift = (threadIdx.x & 31u) / G * G;
    const uint32_t gmask = G == 32 ? 0xffffffffu : (((1u << G) - 1u) << gshift);
    const uint32_t c = blockIdx.x * (blockDim.x / G) + threadIdx.x / G;
    if (c >= b.counters[4]) return;   // group-uniform
    const uint32_t i = b.cand_list[c];
    const float* rt = b.ray_tmp + (size_t)i * 8;
    const V3 o = ld3(rt), du = ld3(rt + 3);
    const float startt = rt[6]; const Cone cone(rt[7]);
    const uint32_t spr = dyn_scratch(b);
    SampleLog log{b.sample_scratch + (size_t)i * spr, spr, b.sample_pool, b.pool_slots, &b.counters[3], NO_SLOT};
    const uint32_t numsteps = march_training_ray_group<G, false, FLAT, K>(o, normalize3(du), startt, cone, aabb, bitfield, max_mip, NERF_STEPS, nullptr, lane, gshift, gmask, &log);
    if (lane == 0) {
        b.overflow_slot[i] = log.slot; b.numsteps_all[i] = numsteps;
        if (numsteps) { atomicAdd(&b.block_steps[i / RPB], numsteps); atomicAdd(&b.block_rays[i / RPB], 1u); }   // per-block totals of the ray-order scan
    }
}

template <uint32_t G, bool FLAT>
__global__ void __launch_bounds__(SCAN_BLOCK) sampler_write_group_kernel(uint32_t n_rays, Aabb aabb, uint32_t max_samples, const uint8_t* __restrict__ bitfield,
                                                                         uint32_t max_mip, SamplerBuffers b) {
    constexpr uint32_t RPB = SCAN_BLOCK / G;
    n_rays = dyn_rays(b, n_rays); max_samples = dyn_samples(b, max_samples);
    const uint32_t spr = dyn_scratch(b);
    const uint32_t lane = threadIdx.x % G, gshift = (threadIdx.x & 31u) / G * G;
    const uint32_t gmask = G == 32 ? 0xffffffffu : (((1u << G) - 1u) << gshift);
    const uint32_t i = blockIdx.x * RPB + threadIdx.x / G;
    const uint32_t numsteps = i < n_rays ? b.numsteps_all[i] : 0;
    const uint32_t v = lane == 0 ? numsteps : 0u;
    uint32_t tot;
    const uint32_t base = __shfl_sync(gmask, b.block_steps[blockIdx.x] + block_exclusive_scan(v, tot), gshift);
    const uint32_t ray_idx = __shfl_sync(gmask, b.block_rays[blockIdx.x] + block_exclusive_scan(v > 0 ? 1u : 0u, tot), gshift);
    const bool keep = numsteps > 0 && base + numsteps <= max_samples;
    uint32_t kept_total;
    block_exclusive_scan(keep && lane == 0 ? 1u : 0u, kept_total);
    if (threadIdx.x == 0 && kept_total) atomicAdd(&b.counters[0], kept_total);
    V3 o = v3(0, 0, 0), du = v3(0, 0, 1);
    uint32_t slot = NO_SLOT;
    if (keep) {
        const float* rt = b.ray_tmp + (size_t)i * 8;
        o = ld3(rt); du = ld3(rt + 3);
        if (lane == 0) {
            b.ray_indices[ray_idx] = i;
            float* r = b.rays + (size_t)ray_idx * 6;
            r[0] = o.x; r[1] = o.y; r[2] = o.z; r[3] = du.x; r[4] = du.y; r[5] = du.z;
            b.numsteps[ray_idx * 2 + 0] = numsteps; b.numsteps[ray_idx * 2 + 1] = base;
        }
        if (numsteps > spr) slot = b.overflow_slot[i];
    }
    // rays whose samples the counting walk logged are expanded by sampler_expand_kernel; the few that outgrew window and pool walk again
    const bool logged = keep && (numsteps <= spr || slot < b.pool_slots);
    if (!keep || logged) return;
    const float* rt = b.ray_tmp + (size_t)i * 8;
    march_training_ray_group<G, true, FLAT>(o, normalize3(du), rt[6], Cone(rt[7]), aabb, bitfield, max_mip, numsteps, b.coords + (size_t)base * 7, lane, gshift, gmask);
}

// Log -> NerfCoordinates, one thread per SAMPLE: the owning ray is found by bisection over the (ascending) base offsets of the
// kept rays, so the work is flat — no per-ray loop whose longest ray (~10x the mean) would bound the kernel.
__global__ void __launch_bounds__(256) sampler_expand_kernel(uint32_t max_samples, Aabb aabb, SamplerBuffers b) {
    __shared__ __align__(16) float stage[256 * 7];
    const uint32_t sidx = blockIdx.x * blockDim.x + threadIdx.x;
    max_samples = dyn_samples(b, max_samples);
    const uint32_t spr = dyn_scratch(b);
    const uint32_t n_kept = b.counters[0];
    if (n_kept == 0) return;
    const uint32_t last = n_kept - 1;
    bool mine = sidx < min(max_samples, b.numsteps[last * 2 + 1] + b.numsteps[last * 2 + 0]);
    float v[7] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (mine) {
        uint32_t lo = 0, hi = last;                       // largest r with base_r <= sidx
        while (lo < hi) {
            const uint32_t mid = (lo + hi + 1) >> 1;
            if (__ldg(b.numsteps + mid * 2 + 1) <= sidx) lo = mid; else hi = mid - 1;
        }
        const uint32_t numsteps = b.numsteps[lo * 2 + 0], k = sidx - b.numsteps[lo * 2 + 1];
        const uint32_t i = b.ray_indices[lo];
        const uint32_t slot = numsteps > spr ? b.overflow_slot[i] : NO_SLOT;
        // (rays that outgrew window and pool were written by the re-walk in sampler_write_group_kernel)
        mine = k < numsteps && (numsteps <= spr || slot < b.pool_slots);
        if (mine) {
            const SampleLog log{b.sample_scratch + (size_t)i * spr, spr, b.sample_pool, b.pool_slots, nullptr, slot};
            const float* r = b.rays + (size_t)lo * 6;
            const V3 o = ld3(r), d = normalize3(ld3(r + 3)), wd = warp_direction(d);
            const float2 td = *log.at(k);
            const V3 wp = warp_position(ray_at(o, d, td.x), aabb);
            v[0] = wp.x; v[1] = wp.y; v[2] = wp.z; v[3] = warp_dt(td.y); v[4] = wd.x; v[5] = wd.y; v[6] = wd.z;
        }
    }
    // a full block owns 256 * 28 contiguous bytes of the coordinate array: staged through shared memory they leave as
    // 16-byte vectors (448 per block) instead of seven 4-byte stores at a 28-byte stride per thread
    if (__syncthreads_and(mine)) {
#pragma unroll
        for (int e = 0; e < 7; ++e) stage[threadIdx.x * 7 + e] = v[e];
        __syncthreads();
        float4* dst = reinterpret_cast<float4*>(b.coords + (size_t)blockIdx.x * 256 * 7);
        for (uint32_t q = threadIdx.x; q < 448; q += 256) dst[q] = reinterpret_cast<const float4*>(stage)[q];
    } else if (mine) {
        float* c = b.coords + (size_t)sidx * 7;
#pragma unroll
        for (int e = 0; e < 7; ++e) c[e] = v[e];
    }
}

// lanes per ray: few long rays -> a whole warp per ray; the usual tens of thousands of rays -> 8 lanes; beyond that the
// lock-step chain's redundancy costs more issue slots than the look-up latency it hides -> one thread per ray
uint32_t nerf_sampler_lanes_per_ray(uint32_t n_rays) { return n_rays <= 1024 ? 32u : (n_rays <= 65536 ? 8u : 1u); }   // measured: tools/phase_profile.py

void nerf_generate_training_samples(cudaStream_t s, uint32_t n_rays, Aabb aabb, uint32_t max_samples, uint32_t n_rays_total, Pcg32 rng,
                                    uint32_t n_images, const ImageMeta* metas, const uint8_t* bitfield, uint32_t max_mip, bool snap,
                                    float cone_angle_constant, uint32_t ray_offset, uint32_t n_rays_global, const SamplerBuffers& b, int phase) {
    (void)n_rays_total;
    NGP_PROFILE_SCOPE(s, phase == SAMPLER_EMIT ? "sampler_emit" : "sampler", n_rays);
    const bool walk = phase != SAMPLER_EMIT, emit = phase != SAMPLER_WALK;
    const uint32_t G = b.lanes_per_ray ? b.lanes_per_ray : nerf_sampler_lanes_per_ray(n_rays);
    NGP_CHECK(G == 1 || G == 2 || G == 4 || G == 8 || G == 32, "sampler_lanes_per_ray must be 0 (automatic), 1, 2, 4, 8 or 32");
    NGP_CHECK(!b.dyn || G != 1, "device-side step parameters need a group marcher (lanes per ray > 1)");
    NGP_CHECK(phase == SAMPLER_ALL || G != 1, "the two-phase sampler needs a group marcher (lanes per ray > 1)");
    uint32_t K = G == 32 ? 1u : 8u / G;                         // 8 candidates per round (32 with a whole warp per ray)
    if (const char* e = getenv("NGP_SAMPLER_K")) K = (uint32_t)atoi(e);
    const bool flat = cone_angle_constant <= 1e-5f && max_mip == 0;
    const uint32_t blocks = div_round_up(n_rays, SCAN_BLOCK / G);
    uint32_t count_block = 64;
    if (const char* e = getenv("NGP_COUNT_BLOCK")) count_block = (uint32_t)atoi(e);
// the walk has no block-wide step: small blocks (2 warps) spread the candidate rays evenly over the SMs
#define NGP_SAMPLER_COUNT(GG, FF, KK) sampler_count_group_kernel<GG, FF, KK><<<div_round_up(n_rays, count_block / GG), count_block, 0, s>>>(n_rays, aabb, bitfield, max_mip, b)
#define NGP_SAMPLER_COUNT_F(GG, KK) do { if (flat) NGP_SAMPLER_COUNT(GG, true, KK); else NGP_SAMPLER_COUNT(GG, false, KK); } while (0)
#define NGP_SAMPLER_WRITE(GG) do { if (flat) sampler_write_group_kernel<GG, true><<<blocks, SCAN_BLOCK, 0, s>>>(n_rays, aabb, max_samples, bitfield, max_mip, b); \
                                   else sampler_write_group_kernel<GG, false><<<blocks, SCAN_BLOCK, 0, s>>>(n_rays, aabb, max_samples, bitfield, max_mip, b); } while (0)
    if (G != 1 && walk) {
        sampler_classify_kernel<<<div_round_up(n_rays * 4, 128), 128, 0, s>>>(n_rays, aabb, rng, n_images, metas, snap, cone_angle_constant, ray_offset, n_rays_global, SCAN_BLOCK / G, b);
        NGP_POST_LAUNCH();
    }
    if (G == 1) { if (flat) sampler_count_kernel<true><<<blocks, SCAN_BLOCK, 0, s>>>(n_rays, aabb, rng, n_images, metas, bitfield, max_mip, snap, cone_angle_constant, ray_offset, n_rays_global, b);
                  else sampler_count_kernel<false><<<blocks, SCAN_BLOCK, 0, s>>>(n_rays, aabb, rng, n_images, metas, bitfield, max_mip, snap, cone_angle_constant, ray_offset, n_rays_global, b); }
    else if (!walk) {}
    else if (G == 32) NGP_SAMPLER_COUNT_F(32, 1);
    else if (G == 8) { if (K == 2) NGP_SAMPLER_COUNT_F(8, 2); else NGP_SAMPLER_COUNT_F(8, 1); }
    else if (G == 4) { if (K == 1) NGP_SAMPLER_COUNT_F(4, 1); else if (K == 4) NGP_SAMPLER_COUNT_F(4, 4); else NGP_SAMPLER_COUNT_F(4, 2); }
    else { if (K == 8) NGP_SAMPLER_COUNT_F(2, 8); else NGP_SAMPLER_COUNT_F(2, 4); }
    NGP_POST_LAUNCH();
    if (!walk) {}
    else if (b.dyn) scan_block_totals_dyn_kernel<<<1, 1024, 0, s>>>(b.dyn, SCAN_BLOCK / G, b.block_steps, b.block_rays, &b.counters[1]);
    else scan_block_totals_kernel<<<1, 1024, 0, s>>>(blocks, b.block_steps, b.block_rays, &b.counters[1], nullptr, false);
    NGP_POST_LAUNCH();
    if (!emit) return;
    if (G == 1) { if (flat) sampler_write_kernel<true><<<blocks, SCAN_BLOCK, 0, s>>>(n_rays, aabb, max_samples, bitfield, max_mip, b);
                  else sampler_write_kernel<false><<<blocks, SCAN_BLOCK, 0, s>>>(n_rays, aabb, max_samples, bitfield, max_mip, b); }
    else if (G == 32) NGP_SAMPLER_WRITE(32);
    else if (G == 8) NGP_SAMPLER_WRITE(8);
    else if (G == 4) NGP_SAMPLER_WRITE(4);
    else NGP_SAMPLER_WRITE(2);
    if (G != 1) {
        NGP_POST_LAUNCH();
        sampler_expand_kernel<<<div_round_up(max_samples, 256), 256, 0, s>>>(max_samples, aabb, b);
    }
    NGP_POST_LAUNCH();
}


// =================================================================================================================
// Loss / compositor / gradient: compute_loss_kernel_train_nerf (testbed_nerf.cu:1660-2012) with the default-off features
// (envmap, error map, depth supervision, exposure, extra dims, max_level_rand_training) left out.
//   pass 1: front-to-back composite of each ray (T, alpha, rgb), target pixel on the replayed random background
//   scan  : exclusive prefix of the compacted step counts (replaces atomicAdd(numsteps_counter_compacted, ...))
//   pass 2: loss + analytic dL/d(rgb,sigma) per kept sample, coordinates copied to the compacted batch
// =================================================================================================================
NGP_D void copy_enc_row(const __half* __restrict__ src, __half* __restrict__ dst, uint32_t src_row, uint32_t dst_row, uint32_t width) {
    for (uint32_t c = 0; c < width / 8; ++c)
        *reinterpret_cast<uint4*>(dst + blk8_chunk_offset(dst_row, c, width)) = __ldg(reinterpret_cast<const uint4*>(src + blk8_chunk_offset(src_row, c, width)));
}
struct LG { V3 loss, grad; };
NGP_D float pick3(float x, float y, float z, uint32_t c);
// per-image exposure (optimize_exposure; testbed_nerf.cu:1794,1969-1983): the target is the pixel times 2^exposure; dL/d exposure of a ray is
// -dL/d prediction (symmetric loss) taken back through the sRGB encoding of the target, times 2^e ln 2, summed per image with atomics
NGP_D float exposure_scale(float e) { return ngp_expf(0.6931471805599453f * e); }
NGP_D float srgb_to_linear_derivative(float s) { return s <= 0.04045f ? 1.0f / 12.92f : 2.4f / 1.055f * powf((s + 0.055f) / 1.055f, 1.4f); }   // common_device.cuh:87-93
NGP_D void accumulate_exposure_gradient(const LossConfig& cfg, uint32_t img, uint32_t c, float grad_c, float target_c, float loss_scale) {
    float dgt = -grad_c;
    if (!cfg.linear_colors) dgt /= srgb_to_linear_derivative(target_c);
    atomicAdd(cfg.exposure_gradient + img * 3 + c, loss_scale * dgt * exposure_scale(cfg.exposure[img * 3 + c]) * 0.6931471805599453f);
}
NGP_D LG loss_and_gradient(V3 target, V3 pred, int type) {   // :207-283,1340-1355
    const float d[3] = {pred.x - target.x, pred.y - target.y, pred.z - target.z}, p[3] = {pred.x, pred.y, pred.z}, tg[3] = {target.x, target.y, target.z};
    float l[3], g[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        switch (type) {
            case LOSS_RELL2: { float den = p[c] * p[c] + 1e-2f; l[c] = d[c] * d[c] / den; g[c] = 2.0f * d[c] / den; break; }
            case LOSS_L1: l[c] = fabsf(d[c]); g[c] = copysignf(1.0f, d[c]); break;
            case LOSS_MAPE: { float den = fabsf(p[c]) + 1e-2f; l[c] = fabsf(d[c]) / den; g[c] = copysignf(1.0f / den, d[c]); break; }
            case LOSS_SMAPE: { float den = 0.5f * (fabsf(p[c]) + fabsf(tg[c])) + 1e-2f; l[c] = fabsf(d[c]) / den; g[c] = copysignf(1.0f / den, d[c]); break; }
            case LOSS_HUBER: { const float alpha = 0.1f; float ad = fabsf(d[c]); float sq = 0.5f / alpha * d[c] * d[c];
                l[c] = (ad > alpha ? (ad - 0.5f * alpha) : sq) / 5.0f;
                g[c] = (ad > alpha ? (d[c] > 0 ? 1.0f : -1.0f) : (d[c] / alpha)) / 5.0f; break; }
            case LOSS_LOGL1: { float dv = fabsf(d[c]) + 1.0f; l[c] = logf(dv); g[c] = copysignf(1.0f / dv, d[c]); break; }
            default: l[c] = d[c] * d[c]; g[c] = 2.0f * d[c]; break;
        }
    }
    return {v3(l[0], l[1], l[2]), v3(g[0], g[1], g[2])};
}

NGP_D void load_out4(const __half* p, float* o) {
    const uint2 raw = __ldg(reinterpret_cast<const uint2*>(p));
    const __half* h = reinterpret_cast<const __half*>(&raw);
    o[0] = __half2float(h[0]); o[1] = __half2float(h[1]); o[2] = __half2float(h[2]); o[3] = __half2float(h[3]);
}

// target depth of the ray's pixel along the (unnormalised) ray, loss of the composited depth against it, times lambda (:1848-1856); 0 without a map
NGP_D float depth_loss_gradient(const LossConfig& cfg, const ImageMeta& m, const float* uv, V3 d_unnormalised, float depth_ray) {
    const float target_depth = length3(d_unnormalised) * (m.depth ? read_depth(m, uv[0], uv[1]) : -1.0f);
    const LG lg = loss_and_gradient(v3(target_depth, target_depth, target_depth), v3(depth_ray, depth_ray, depth_ray), cfg.depth_loss_type);
    return target_depth > 0.0f ? cfg.depth_supervision_lambda * lg.grad.x : 0.0f;
}

// EXTRAS: the default-off per-ray extras (depth supervision, max_level_rand_training) are compiled into a second instantiation, so that the
// kernels of an ordinary step carry neither their registers nor their branches (measured: 0.057 -> 0.063 ms with them in line)
template <bool EXTRAS>
__global__ void __launch_bounds__(SCAN_BLOCK) loss_forward_kernel(uint32_t n_rays, Aabb aabb, Pcg32 rng, LossConfig cfg, uint32_t n_images, const ImageMeta* __restrict__ metas,
                                                                  uint32_t ray_offset, uint32_t n_rays_global, SamplerBuffers sb, LossBuffers lb) {
    const uint32_t i = threadIdx.x + blockIdx.x * blockDim.x;
    uint32_t cn = 0;
    if (i < sb.counters[0]) {
        const uint32_t numsteps = sb.numsteps[i * 2 + 0], base = sb.numsteps[i * 2 + 1];
        const float* cin = sb.coords + (size_t)base * 7;
        const __half* no = lb.network_output + (size_t)base * 4;
        float T = 1.f; const float epsilon = 1e-4f;
        V3 rgb_ray = v3(0, 0, 0);
        const bool with_depth = EXTRAS && cfg.depth_supervision_lambda > 0.0f;
        float depth_ray = 0.f;
        const V3 ray_o = ld3(sb.rays + (size_t)i * 6);
        for (; cn < numsteps; ++cn) {
            if (T < epsilon) break;
            float o[4]; load_out4(no + (size_t)cn * 4, o);
            const V3 rgb = v3(network_to_rgb(o[0], cfg.rgb_activation), network_to_rgb(o[1], cfg.rgb_activation), network_to_rgb(o[2], cfg.rgb_activation));
            const float dt = unwarp_dt(__ldg(cin + (size_t)cn * 7 + 3));
            const float density = network_to_density(o[3], cfg.density_activation);
            const float alpha = alpha_from_density(density, dt);
            const float weight = alpha * T;
            rgb_ray = rgb_ray + weight * rgb;
            if (with_depth) {   // :1743,1752
                const V3 pos = unwarp_position(v3(__ldg(cin + (size_t)cn * 7), __ldg(cin + (size_t)cn * 7 + 1), __ldg(cin + (size_t)cn * 7 + 2)), aabb);
                depth_ray += weight * length3(pos - ray_o);
            }
            T *= (1.f - alpha);
        }
        // replay the sampler's RNG stream to recover pixel + background colour (:1757-1780)
        const uint32_t ig = ray_offset + sb.ray_indices[i];
        rng.advance((int64_t)((uint64_t)ig * N_MAX_RANDOM_SAMPLES_PER_RAY));
        float img_pdf = 1.0f, uv_pdf = 1.0f;
        const uint32_t img = image_idx(ig, n_rays_global, n_images, EXTRAS ? sb.cdf.cdf_img : nullptr, &img_pdf);
        const ImageMeta& m = metas[img];
        float uv[2]; random_image_pos_training(rng, m.res, cfg.snap_to_pixel_centers, uv, EXTRAS ? &sb.cdf : nullptr, img, &uv_pdf);
        const float max_level = (EXTRAS && cfg.max_level_rand) ? rng.next_float() * 2.0f : 1.0f;   // :1775 "Multiply by 2 to ensure 50% of training is at max level"
        rng.advance(1);   // motionblur_time
        V3 bg = ld3(cfg.background_color);
        if (cfg.random_bg_color) { const float a = rng.next_float(), b = rng.next_float(), c = rng.next_float(); bg = v3(a, b, c); }
        bg = v3(srgb_to_linear(bg.x), srgb_to_linear(bg.y), srgb_to_linear(bg.z));
        float tex[4]; read_rgba(m, uv[0], uv[1], tex);
        V3 trgb = v3(tex[0], tex[1], tex[2]); const float ta = tex[3];
        if (cfg.exposure) trgb = v3(exposure_scale(cfg.exposure[img * 3]) * trgb.x, exposure_scale(cfg.exposure[img * 3 + 1]) * trgb.y, exposure_scale(cfg.exposure[img * 3 + 2]) * trgb.z);
        V3 target;
        if (cfg.linear_colors || cfg.color_space == 0) {   // :1801-1808 (m_color_space is Linear unless the caller changed it)
            target = trgb + (1.0f - ta) * bg;
            if (!cfg.linear_colors) {
                target = v3(linear_to_srgb(target.x), linear_to_srgb(target.y), linear_to_srgb(target.z));
                bg = v3(linear_to_srgb(bg.x), linear_to_srgb(bg.y), linear_to_srgb(bg.z));
            }
        } else {                                            // color_space SRGB (:1809-1818)
            bg = v3(linear_to_srgb(bg.x), linear_to_srgb(bg.y), linear_to_srgb(bg.z));
            if (ta > 0) {
                const V3 q = v3(linear_to_srgb(trgb.x / ta), linear_to_srgb(trgb.y / ta), linear_to_srgb(trgb.z / ta));
                target = q * ta + (1.0f - ta) * bg;
            } else target = bg;
        }
        if (cn == numsteps) rgb_ray = rgb_ray + T * bg;
        float* rs = lb.ray_scratch + (size_t)i * RAY_SCRATCH;
        rs[0] = rgb_ray.x; rs[1] = rgb_ray.y; rs[2] = rgb_ray.z; rs[3] = target.x; rs[4] = target.y; rs[5] = target.z; rs[6] = __uint_as_float(img);
        if (EXTRAS) { rs[7] = max_level; rs[10] = uv[0]; rs[11] = uv[1]; rs[12] = img_pdf; rs[13] = uv_pdf; rs[14] = __int_as_float(m.res[0]); rs[15] = __int_as_float(m.res[1]); }
        if (with_depth) { rs[8] = depth_ray; rs[9] = depth_loss_gradient(cfg, m, uv, ld3(sb.rays + (size_t)i * 6 + 3), depth_ray); }
        lb.compacted_numsteps[i] = cn;
    }
    uint32_t total;
    block_exclusive_scan(cn, total);
    if (threadIdx.x == 0) sb.block_steps[blockIdx.x] = total;
}

// error map (:1869-1900, without the sharpness weighting - sharpened data sets are not loaded): the ray's mean loss splatted bilinearly around its
// pixel on the error_map lattice of its view.  The clamp of the integer position uses the IMAGE resolution, as the reference does.
NGP_D void deposit_error(const LossBuffers& lb, uint32_t img, const int* image_res, float u, float v, float mean_loss) {
    const int ex = lb.error_map_res[0], ey = lb.error_map_res[1];
    const float px = fminf(fmaxf(u * (float)ex - 0.5f, 0.0f), (float)ex - (1.0f + 1e-4f)), py = fminf(fmaxf(v * (float)ey - 0.5f, 0.0f), (float)ey - (1.0f + 1e-4f));
    const int ix = (int)px, iy = (int)py;
    const float wx = px - (float)ix, wy = py - (float)iy;
    const int x = clampi(ix, 0, image_res[0] - 2), y = clampi(iy, 0, image_res[1] - 2);
    float* e = lb.error_map + (size_t)img * ex * ey;
    atomicAdd(e + (size_t)y * ex + x, (1 - wx) * (1 - wy) * mean_loss);
    atomicAdd(e + (size_t)y * ex + x + 1, wx * (1 - wy) * mean_loss);
    atomicAdd(e + (size_t)(y + 1) * ex + x, (1 - wx) * wy * mean_loss);
    atomicAdd(e + (size_t)(y + 1) * ex + x + 1, wx * wy * mean_loss);
}

template <bool EXTRAS>
__global__ void __launch_bounds__(SCAN_BLOCK) loss_backward_kernel(uint32_t n_rays, Aabb aabb, uint32_t max_compacted, LossConfig cfg,
                                                                   const float* __restrict__ mean_density, uint32_t n_rays_global, SamplerBuffers sb, LossBuffers lb) {
    const uint32_t i = threadIdx.x + blockIdx.x * blockDim.x;
    const bool active = i < sb.counters[0];
    uint32_t cn = active ? lb.compacted_numsteps[i] : 0;
    uint32_t total;
    const uint32_t cbase = sb.block_steps[blockIdx.x] + block_exclusive_scan(cn, total);
    if (!active) return;
    const uint32_t base = sb.numsteps[i * 2 + 1];
    cn = min(max_compacted - min(max_compacted, cbase), cn);   // :1831-1833
    sb.numsteps[i * 2 + 0] = cn; sb.numsteps[i * 2 + 1] = cbase;
    if (cn == 0) return;
    const float* cin = sb.coords + (size_t)base * 7;
    const __half* no = lb.network_output + (size_t)base * 4;
    const float* rs = lb.ray_scratch + (size_t)i * RAY_SCRATCH;
    const V3 rgb_ray = ld3(rs), target = ld3(rs + 3);
    const V3 ray_o = ld3(sb.rays + (size_t)i * 6);

    const LG lg = loss_and_gradient(target, rgb_ray, cfg.loss_type);
    float mean_loss = (lg.loss.x + lg.loss.y + lg.loss.z) / 3.0f;
    const float uv_pdf = EXTRAS ? rs[13] : 1.0f;
    if (EXTRAS) mean_loss = ((lg.loss.x / (rs[12] * uv_pdf)) + (lg.loss.y / (rs[12] * uv_pdf)) + (lg.loss.z / (rs[12] * uv_pdf))) / 3.0f;   // lg.loss /= img_pdf * uv_pdf (:1846)
    if (lb.loss) lb.loss[i] = mean_loss / (float)n_rays_global;
    if (EXTRAS && lb.error_map) { const int ires[2] = {__float_as_int(rs[14]), __float_as_int(rs[15])}; deposit_error(lb, __float_as_uint(rs[6]), ires, rs[10], rs[11], mean_loss); }
    const float loss_scale = cfg.loss_scale / (float)n_rays_global;   // :1903
    const float output_l2_reg = cfg.rgb_activation == ACT_EXPONENTIAL ? 1e-4f : 0.0f;
    const float output_l1_reg_density = *mean_density < NERF_MIN_OPTICAL_THICKNESS ? 1e-4f : 0.0f;

    V3 rgb_ray2 = v3(0, 0, 0); float T = 1.f;
    const bool with_depth = EXTRAS && cfg.depth_supervision_lambda > 0.0f;
    const float depth_ray = with_depth ? rs[8] : 0.f, depth_gradient = with_depth ? rs[9] : 0.f; float depth_ray2 = 0.f;
    __half* gout = lb.dloss_doutput + (size_t)cbase * 4;
    float* cout = lb.coords_compacted + (size_t)cbase * 7;
    for (uint32_t j = 0; j < cn; ++j) {
        float c[7];
#pragma unroll
        for (int k = 0; k < 7; ++k) { c[k] = __ldg(cin + (size_t)j * 7 + k); cout[(size_t)j * 7 + k] = c[k]; }
        if (lb.enc_out) copy_enc_row(lb.enc_in, lb.enc_out, base + j, cbase + j, lb.enc_width);
        if (EXTRAS && lb.max_level_compacted) lb.max_level_compacted[cbase + j] = rs[7];   // :1915-1917
        const V3 pos = unwarp_position(v3(c[0], c[1], c[2]), aabb);
        const float depth = length3(pos - ray_o);
        const float dt = unwarp_dt(c[3]);
        float o[4]; load_out4(no + (size_t)j * 4, o);
        const V3 rgb = v3(network_to_rgb(o[0], cfg.rgb_activation), network_to_rgb(o[1], cfg.rgb_activation), network_to_rgb(o[2], cfg.rgb_activation));
        const float density = network_to_density(o[3], cfg.density_activation);
        const float alpha = alpha_from_density(density, dt);
        const float weight = alpha * T;
        rgb_ray2 = rgb_ray2 + weight * rgb;
        T *= (1.f - alpha);
        const V3 suffix = rgb_ray - rgb_ray2;
        const V3 dl_drgb = weight * lg.grad;
        const float g0 = loss_scale * (dl_drgb.x * network_to_rgb_derivative(o[0], cfg.rgb_activation) + fmaxf(0.0f, output_l2_reg * o[0]));
        const float g1 = loss_scale * (dl_drgb.y * network_to_rgb_derivative(o[1], cfg.rgb_activation) + fmaxf(0.0f, output_l2_reg * o[1]));
        const float g2 = loss_scale * (dl_drgb.z * network_to_rgb_derivative(o[2], cfg.rgb_activation) + fmaxf(0.0f, output_l2_reg * o[2]));
        const float dd = network_to_density_derivative(o[3], cfg.density_activation);
        const V3 tr = T * rgb - suffix;
        float chain = lg.grad.x * tr.x + lg.grad.y * tr.y + lg.grad.z * tr.z;
        if (with_depth) { depth_ray2 += weight * depth; chain += depth_gradient * (T * depth - (depth_ray - depth_ray2)); }   // :1926,1934-1939
        const float dl_dmlp = dd * (dt * chain);
        const float g3 = loss_scale * dl_dmlp + (o[3] < 0.0f ? -output_l1_reg_density : 0.0f) + (o[3] > -10.0f && depth < cfg.near_distance ? 1e-4f : 0.0f);
        __half2 h[2] = {__floats2half2_rn(g0, g1), __floats2half2_rn(g2, g3)};
        *reinterpret_cast<uint2*>(gout + (size_t)j * 4) = *reinterpret_cast<uint2*>(h);
    }
    if (cfg.exposure_gradient) {
#pragma unroll
        for (uint32_t c = 0; c < 3; ++c) accumulate_exposure_gradient(cfg, __float_as_uint(rs[6]), c, pick3(lg.grad.x, lg.grad.y, lg.grad.z, c) / uv_pdf, pick3(target.x, target.y, target.z, c), loss_scale);
    }
}

// ---- G lanes per ray -------------------------------------------------------------------------------------------------
// Pass 1 keeps exactly one thing serial: the transmittance chain T *= (1 - alpha_k) with its cut-off test, replayed in
// sample order with one group shuffle per sample — the same float operations in the same order as the reference loop, so
// the number of samples that survive compaction is bit-identical.  Everything else is per-lane: samples are loaded and
// activated in parallel (coalesced), lane k keeps T before its own sample (written to lb.t_before for pass 2), the ray
// colour is a per-lane partial sum reduced once at the end, and the target / background colour work of the epilogue is
// spread over three lanes (one per channel).  Pass 2 has no serial part at all: weight and T come from t_before, the
// colour prefix from a group scan.  Colour sums are re-associated (float, ~1e-7 relative), counts are not.
template <uint32_t G>
NGP_D float group_sum(float v, uint32_t gmask) {
#pragma unroll
    for (uint32_t o = G / 2; o > 0; o >>= 1) v += __shfl_xor_sync(gmask, v, o);
    return v;
}
template <uint32_t G>
NGP_D float group_inclusive_scan(float v, uint32_t lane, uint32_t gmask) {
#pragma unroll
    for (uint32_t o = 1; o < G; o <<= 1) { const float u = __shfl_up_sync(gmask, v, o, G); if (lane >= o) v += u; }
    return v;
}
NGP_D float pick3(float x, float y, float z, uint32_t c) { return c == 0 ? x : (c == 1 ? y : z); }   // (declared above)

template <uint32_t G, bool EXTRAS>
__global__ void __launch_bounds__(SCAN_BLOCK) loss_forward_group_kernel(uint32_t n_rays, Aabb aabb, Pcg32 rng, LossConfig cfg, uint32_t n_images, const ImageMeta* __restrict__ metas,
                                                                        uint32_t ray_offset, uint32_t n_rays_global, SamplerBuffers sb, LossBuffers lb) {
    constexpr uint32_t RPB = SCAN_BLOCK / G;
    const uint32_t lane = threadIdx.x % G, gshift = (threadIdx.x & 31u) / G * G;
    const uint32_t gmask = G == 32 ? 0xffffffffu : (((1u << G) - 1u) << gshift);
    const uint32_t i = blockIdx.x * RPB + threadIdx.x / G;
    uint32_t cn = 0;
    if (i < sb.counters[0]) {
        const uint32_t numsteps = sb.numsteps[i * 2 + 0], base = sb.numsteps[i * 2 + 1];
        const float* cin = sb.coords + (size_t)base * 7;
        const __half* no = lb.network_output + (size_t)base * 4;
        float* tb_out = lb.t_before + base;
        float T = 1.f; const float epsilon = 1e-4f;
        V3 acc = v3(0, 0, 0);
        const bool with_depth = EXTRAS && cfg.depth_supervision_lambda > 0.0f;
        float depth_acc = 0.f;
        const V3 ray_o = ld3(sb.rays + (size_t)i * 6);
        bool stop = false;
        // software pipeline: the raw loads of chunk c+1 are in flight while the chain of chunk c is replayed
        uint2 raw_next = make_uint2(0u, 0u); float dtw_next = 0.f;
        if (lane < numsteps) { raw_next = __ldg(reinterpret_cast<const uint2*>(no + (size_t)lane * 4)); dtw_next = __ldg(cin + (size_t)lane * 7 + 3); }
        for (uint32_t c0 = 0; c0 < numsteps && !stop; c0 += G) {
            const uint32_t idx = c0 + lane;
            const uint2 raw = raw_next; const float dtw = dtw_next;
            if (idx + G < numsteps) { raw_next = __ldg(reinterpret_cast<const uint2*>(no + (size_t)(idx + G) * 4)); dtw_next = __ldg(cin + (size_t)(idx + G) * 7 + 3); }
            float alpha = 0.f; V3 rgb = v3(0, 0, 0);
            if (idx < numsteps) {
                const __half* h = reinterpret_cast<const __half*>(&raw);
                const float o[4] = {__half2float(h[0]), __half2float(h[1]), __half2float(h[2]), __half2float(h[3])};
                rgb = v3(network_to_rgb(o[0], cfg.rgb_activation), network_to_rgb(o[1], cfg.rgb_activation), network_to_rgb(o[2], cfg.rgb_activation));
                const float dt = unwarp_dt(dtw);
                const float density = network_to_density(o[3], cfg.density_activation);
                alpha = alpha_from_density(density, dt);
            }
            const uint32_t n = min(G, numsteps - c0);
            float Tb = 0.f; uint32_t done = 0;
            for (uint32_t k = 0; k < n; ++k) {
                if (T < epsilon) { stop = true; break; }
                const float a = __shfl_sync(gmask, alpha, gshift + k);
                if (lane == k) Tb = T;
                T *= (1.f - a);
                ++done;
            }
            cn += done;
            if (lane < done) {
                const float weight = alpha * Tb;
                acc = acc + weight * rgb;
                tb_out[idx] = Tb;
                if (with_depth) {
                    const V3 pos = unwarp_position(v3(__ldg(cin + (size_t)idx * 7), __ldg(cin + (size_t)idx * 7 + 1), __ldg(cin + (size_t)idx * 7 + 2)), aabb);
                    depth_acc += weight * length3(pos - ray_o);
                }
            }
        }
        V3 rgb_ray = v3(group_sum<G>(acc.x, gmask), group_sum<G>(acc.y, gmask), group_sum<G>(acc.z, gmask));
        const float depth_ray = with_depth ? group_sum<G>(depth_acc, gmask) : 0.f;
        // replay the sampler's RNG stream to recover pixel + background colour (:1757-1780); uniform across the group
        const uint32_t ig = ray_offset + sb.ray_indices[i];
        rng.advance((int64_t)((uint64_t)ig * N_MAX_RANDOM_SAMPLES_PER_RAY));
        float img_pdf = 1.0f, uv_pdf = 1.0f;
        const uint32_t img = image_idx(ig, n_rays_global, n_images, EXTRAS ? sb.cdf.cdf_img : nullptr, &img_pdf);
        const ImageMeta& m = metas[img];
        float uv[2]; random_image_pos_training(rng, m.res, cfg.snap_to_pixel_centers, uv, EXTRAS ? &sb.cdf : nullptr, img, &uv_pdf);
        const float max_level = (EXTRAS && cfg.max_level_rand) ? rng.next_float() * 2.0f : 1.0f;   // :1775
        rng.advance(1);
        V3 bg3 = ld3(cfg.background_color);
        if (cfg.random_bg_color) { const float a = rng.next_float(), b = rng.next_float(), c = rng.next_float(); bg3 = v3(a, b, c); }
        float tex[4]; read_rgba(m, uv[0], uv[1], tex);
        if (lane < 3) {   // one colour channel per lane: the transcendental conversions run once per warp instruction, not three times
            float bg = srgb_to_linear(pick3(bg3.x, bg3.y, bg3.z, lane));
            float trgb = pick3(tex[0], tex[1], tex[2], lane); const float ta = tex[3];
            if (cfg.exposure) trgb = exposure_scale(cfg.exposure[img * 3 + lane]) * trgb;
            float target;
            if (cfg.linear_colors || cfg.color_space == 0) {   // :1801-1808
                target = trgb + (1.0f - ta) * bg;
                if (!cfg.linear_colors) { target = linear_to_srgb(target); bg = linear_to_srgb(bg); }
            } else {                                            // color_space SRGB (:1809-1818)
                bg = linear_to_srgb(bg);
                target = ta > 0 ? linear_to_srgb(trgb / ta) * ta + (1.0f - ta) * bg : bg;
            }
            float ray_c = pick3(rgb_ray.x, rgb_ray.y, rgb_ray.z, lane);
            if (cn == numsteps) ray_c = ray_c + T * bg;
            float* rs = lb.ray_scratch + (size_t)i * RAY_SCRATCH;
            rs[lane] = ray_c; rs[3 + lane] = target;
            if (lane == 0) {
                rs[6] = __uint_as_float(img);
                if (EXTRAS) { rs[7] = max_level; rs[10] = uv[0]; rs[11] = uv[1]; rs[12] = img_pdf; rs[13] = uv_pdf; rs[14] = __int_as_float(m.res[0]); rs[15] = __int_as_float(m.res[1]); }
                if (with_depth) { rs[8] = depth_ray; rs[9] = depth_loss_gradient(cfg, m, uv, ld3(sb.rays + (size_t)i * 6 + 3), depth_ray); }
            }
        }
        if (lane == 0) lb.compacted_numsteps[i] = cn;
    }
    uint32_t total;
    block_exclusive_scan(lane == 0 ? cn : 0u, total);
    if (threadIdx.x == 0) sb.block_steps[blockIdx.x] = total;
}

template <uint32_t G, bool EXTRAS>
__global__ void __launch_bounds__(SCAN_BLOCK) loss_backward_group_kernel(uint32_t n_rays, Aabb aabb, uint32_t max_compacted, LossConfig cfg,
                                                                         const float* __restrict__ mean_density, uint32_t n_rays_global, SamplerBuffers sb, LossBuffers lb) {
    constexpr uint32_t RPB = SCAN_BLOCK / G;
    const uint32_t lane = threadIdx.x % G, gshift = (threadIdx.x & 31u) / G * G;
    const uint32_t gmask = G == 32 ? 0xffffffffu : (((1u << G) - 1u) << gshift);
    const uint32_t i = blockIdx.x * RPB + threadIdx.x / G;
    const bool active = i < sb.counters[0];
    uint32_t cn = active ? lb.compacted_numsteps[i] : 0;
    uint32_t total;
    const uint32_t cbase = __shfl_sync(gmask, sb.block_steps[blockIdx.x] + block_exclusive_scan(lane == 0 ? cn : 0u, total), gshift);
    if (!active) return;
    const uint32_t base = sb.numsteps[i * 2 + 1];
    __syncwarp(gmask);   // every lane has read (numsteps, base) before lane 0 overwrites them
    cn = min(max_compacted - min(max_compacted, cbase), cn);
    if (lane == 0) { sb.numsteps[i * 2 + 0] = cn; sb.numsteps[i * 2 + 1] = cbase; }
    if (cn == 0) return;
    const float* cin = sb.coords + (size_t)base * 7;
    const __half* no = lb.network_output + (size_t)base * 4;
    const float* tb_in = lb.t_before + base;
    const float* rs = lb.ray_scratch + (size_t)i * RAY_SCRATCH;
    const V3 rgb_ray = ld3(rs), target = ld3(rs + 3);
    const V3 ray_o = ld3(sb.rays + (size_t)i * 6);

    const LG lg = loss_and_gradient(target, rgb_ray, cfg.loss_type);
    float mean_loss = (lg.loss.x + lg.loss.y + lg.loss.z) / 3.0f;
    const float uv_pdf = EXTRAS ? rs[13] : 1.0f;
    if (EXTRAS) mean_loss = ((lg.loss.x / (rs[12] * uv_pdf)) + (lg.loss.y / (rs[12] * uv_pdf)) + (lg.loss.z / (rs[12] * uv_pdf))) / 3.0f;   // lg.loss /= img_pdf * uv_pdf (:1846)
    if (lb.loss && lane == 0) lb.loss[i] = mean_loss / (float)n_rays_global;
    if (EXTRAS && lb.error_map && lane == 0) { const int ires[2] = {__float_as_int(rs[14]), __float_as_int(rs[15])}; deposit_error(lb, __float_as_uint(rs[6]), ires, rs[10], rs[11], mean_loss); }
    const float loss_scale = cfg.loss_scale / (float)n_rays_global;
    const float output_l2_reg = cfg.rgb_activation == ACT_EXPONENTIAL ? 1e-4f : 0.0f;
    const float output_l1_reg_density = *mean_density < NERF_MIN_OPTICAL_THICKNESS ? 1e-4f : 0.0f;

    V3 carry = v3(0, 0, 0);   // colour composited by the chunks before this one
    const bool with_depth = EXTRAS && cfg.depth_supervision_lambda > 0.0f;
    const float depth_ray = with_depth ? rs[8] : 0.f, depth_gradient = with_depth ? rs[9] : 0.f; float depth_carry = 0.f;
    __half* gout = lb.dloss_doutput + (size_t)cbase * 4;
    float* cout = lb.coords_compacted + (size_t)cbase * 7;
    for (uint32_t c0 = 0; c0 < cn; c0 += G) {
        const uint32_t j = c0 + lane;
        const bool valid = j < cn;
        float o[4] = {0.f, 0.f, 0.f, 0.f}; V3 rgb = v3(0, 0, 0); float alpha = 0.f, dt = 0.f, depth = 0.f, Tb = 0.f;
        if (valid) {
            float c[7];
#pragma unroll
            for (int k = 0; k < 7; ++k) { c[k] = __ldg(cin + (size_t)j * 7 + k); cout[(size_t)j * 7 + k] = c[k]; }
            const uint2 raw = __ldg(reinterpret_cast<const uint2*>(no + (size_t)j * 4));
            Tb = tb_in[j];
            if (lb.enc_out) copy_enc_row(lb.enc_in, lb.enc_out, base + j, cbase + j, lb.enc_width);
            if (EXTRAS && lb.max_level_compacted) lb.max_level_compacted[cbase + j] = rs[7];
            const V3 pos = unwarp_position(v3(c[0], c[1], c[2]), aabb);
            depth = length3(pos - ray_o);
            dt = unwarp_dt(c[3]);
            { const __half* h = reinterpret_cast<const __half*>(&raw); o[0] = __half2float(h[0]); o[1] = __half2float(h[1]); o[2] = __half2float(h[2]); o[3] = __half2float(h[3]); }
            rgb = v3(network_to_rgb(o[0], cfg.rgb_activation), network_to_rgb(o[1], cfg.rgb_activation), network_to_rgb(o[2], cfg.rgb_activation));
            const float density = network_to_density(o[3], cfg.density_activation);
            alpha = alpha_from_density(density, dt);
        }
        const float weight = alpha * Tb, T_after = Tb * (1.f - alpha);   // the chain's own products, from its stored state
        const V3 wr = weight * rgb;                                        // 0 on lanes past the end
        const V3 r2 = carry + v3(group_inclusive_scan<G>(wr.x, lane, gmask), group_inclusive_scan<G>(wr.y, lane, gmask), group_inclusive_scan<G>(wr.z, lane, gmask));
        carry = v3(__shfl_sync(gmask, r2.x, gshift + G - 1), __shfl_sync(gmask, r2.y, gshift + G - 1), __shfl_sync(gmask, r2.z, gshift + G - 1));
        float depth_ray2 = 0.f;
        if (with_depth) {   // uniform across the group
            depth_ray2 = depth_carry + group_inclusive_scan<G>(weight * depth, lane, gmask);
            depth_carry = __shfl_sync(gmask, depth_ray2, gshift + G - 1);
        }
        if (valid) {
            const V3 suffix = rgb_ray - r2;
            const V3 dl_drgb = weight * lg.grad;
            const float g0 = loss_scale * (dl_drgb.x * network_to_rgb_derivative(o[0], cfg.rgb_activation) + fmaxf(0.0f, output_l2_reg * o[0]));
            const float g1 = loss_scale * (dl_drgb.y * network_to_rgb_derivative(o[1], cfg.rgb_activation) + fmaxf(0.0f, output_l2_reg * o[1]));
            const float g2 = loss_scale * (dl_drgb.z * network_to_rgb_derivative(o[2], cfg.rgb_activation) + fmaxf(0.0f, output_l2_reg * o[2]));
            const float dd = network_to_density_derivative(o[3], cfg.density_activation);
            const V3 tr = T_after * rgb - suffix;
            float chain = lg.grad.x * tr.x + lg.grad.y * tr.y + lg.grad.z * tr.z;
            if (with_depth) chain += depth_gradient * (T_after * depth - (depth_ray - depth_ray2));
            const float dl_dmlp = dd * (dt * chain);
            const float g3 = loss_scale * dl_dmlp + (o[3] < 0.0f ? -output_l1_reg_density : 0.0f) + (o[3] > -10.0f && depth < cfg.near_distance ? 1e-4f : 0.0f);
            __half2 h[2] = {__floats2half2_rn(g0, g1), __floats2half2_rn(g2, g3)};
            *reinterpret_cast<uint2*>(gout + (size_t)j * 4) = *reinterpret_cast<uint2*>(h);
        }
    }
    if (cfg.exposure_gradient && lane < 3)   // one channel per lane
        accumulate_exposure_gradient(cfg, __float_as_uint(rs[6]), lane, pick3(lg.grad.x, lg.grad.y, lg.grad.z, lane) / uv_pdf, pick3(target.x, target.y, target.z, lane), loss_scale);
}

// one block: the image count is tens to a few thousand, the work is 3 * n_images Adam updates and three means
__global__ void __launch_bounds__(256) exposure_step_kernel(uint32_t n_images, const float* __restrict__ gradient, float* __restrict__ exposure, float* __restrict__ m1,
                                                             float* __restrict__ m2, float actual_lr, float l2_reg, float scale) {
    __shared__ float sums[3][8];
    float local[3] = {0.f, 0.f, 0.f};
    for (uint32_t i = threadIdx.x; i < n_images; i += blockDim.x) {
#pragma unroll
        for (uint32_t c = 0; c < 3; ++c) {
            const uint32_t k = i * 3 + c;
            const float g = gradient[k] * scale + exposure[k] * l2_reg;
            const float a = 0.9f * m1[k] + (1.0f - 0.9f) * g, b = 0.99f * m2[k] + (1.0f - 0.99f) * g * g;
            m1[k] = a; m2[k] = b;
            const float v = exposure[k] - actual_lr * a / (sqrtf(b) + 1e-8f);
            exposure[k] = v; local[c] += v;
        }
    }
#pragma unroll
    for (uint32_t c = 0; c < 3; ++c) {
        float v = local[c];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31u) == 0) sums[c][threadIdx.x >> 5] = v;
    }
    __syncthreads();
    float mean[3];
#pragma unroll
    for (uint32_t c = 0; c < 3; ++c) { float t = 0.f; for (int w = 0; w < 8; ++w) t += sums[c][w]; mean[c] = t / (float)n_images; }
    for (uint32_t i = threadIdx.x; i < n_images; i += blockDim.x) {      // renormalise: exposures are relative to their mean
#pragma unroll
        for (uint32_t c = 0; c < 3; ++c) exposure[i * 3 + c] -= mean[c];
    }
}
void nerf_exposure_step(cudaStream_t s, uint32_t n_images, const float* gradient, float* exposure, float* first_moment, float* second_moment, uint32_t iter,
                        float learning_rate, float l2_reg, float per_camera_loss_scale) {
    if (!n_images) return;
    const float actual_lr = learning_rate * std::sqrt(1.0f - std::pow(0.99f, (float)iter)) / (1.0f - std::pow(0.9f, (float)iter));
    exposure_step_kernel<<<1, 256, 0, s>>>(n_images, gradient, exposure, first_moment, second_moment, actual_lr, l2_reg, per_camera_loss_scale);
    NGP_POST_LAUNCH();
}

void nerf_compute_loss(cudaStream_t s, uint32_t n_rays, Aabb aabb, uint32_t n_rays_total, Pcg32 rng, uint32_t max_compacted, const LossConfig& cfg,
                       uint32_t n_images, const ImageMeta* metas, const float* mean_density_dev, uint32_t ray_offset, uint32_t n_rays_global,
                       const SamplerBuffers& sb, const LossBuffers& lb) {
    (void)n_rays_total;
    NGP_PROFILE_SCOPE(s, "loss", n_rays);
    // a whole warp per ray unless rays are plentiful: the slowest (longest) ray bounds the kernel, and 32 lanes cut its chunk count 4x
    uint32_t G = sb.lanes_per_ray == 1 ? 1u : (n_rays <= 65536 ? 32u : 8u);
    if (const char* e = getenv("NGP_LOSS_LANES")) G = (uint32_t)atoi(e);
    const uint32_t blocks = div_round_up(n_rays, SCAN_BLOCK / G);
    const bool extras = cfg.depth_supervision_lambda > 0.0f || cfg.max_level_rand || lb.max_level_compacted || lb.error_map || sb.cdf.cdf_x_cond_y || sb.cdf.cdf_img;
#define NGP_LOSS_FWD(K) K<<<blocks, SCAN_BLOCK, 0, s>>>(n_rays, aabb, rng, cfg, n_images, metas, ray_offset, n_rays_global, sb, lb)
#define NGP_LOSS_BWD(K) K<<<blocks, SCAN_BLOCK, 0, s>>>(n_rays, aabb, max_compacted, cfg, mean_density_dev, n_rays_global, sb, lb)
    if (G == 32) { if (extras) NGP_LOSS_FWD((loss_forward_group_kernel<32, true>)); else NGP_LOSS_FWD((loss_forward_group_kernel<32, false>)); }
    else if (G == 8) { if (extras) NGP_LOSS_FWD((loss_forward_group_kernel<8, true>)); else NGP_LOSS_FWD((loss_forward_group_kernel<8, false>)); }
    else { if (extras) NGP_LOSS_FWD(loss_forward_kernel<true>); else NGP_LOSS_FWD(loss_forward_kernel<false>); }
    NGP_POST_LAUNCH();
    scan_block_totals_kernel<<<1, 1024, 0, s>>>(blocks, sb.block_steps, nullptr, &sb.counters[2], nullptr, false);
    NGP_POST_LAUNCH();
    if (G == 32) { if (extras) NGP_LOSS_BWD((loss_backward_group_kernel<32, true>)); else NGP_LOSS_BWD((loss_backward_group_kernel<32, false>)); }
    else if (G == 8) { if (extras) NGP_LOSS_BWD((loss_backward_group_kernel<8, true>)); else NGP_LOSS_BWD((loss_backward_group_kernel<8, false>)); }
    else { if (extras) NGP_LOSS_BWD(loss_backward_kernel<true>); else NGP_LOSS_BWD(loss_backward_kernel<false>); }
    NGP_POST_LAUNCH();
#undef NGP_LOSS_FWD
#undef NGP_LOSS_BWD
}

__global__ void cam_gradient_kernel(Aabb aabb, Pcg32 rng, bool snap, const ImageMeta* __restrict__ metas, uint32_t n_images, uint32_t ray_offset, uint32_t n_rays_global, SamplerBuffers sb,
                                    const float* __restrict__ coords, const float* __restrict__ coords_gradient, float* __restrict__ cam_pos_gradient,
                                    float* __restrict__ cam_rot_gradient, float* __restrict__ distortion_gradient, float* __restrict__ distortion_gradient_weight) {
    const uint32_t i = threadIdx.x + blockIdx.x * blockDim.x;
    if (i >= sb.counters[0]) return;
    const uint32_t numsteps = sb.numsteps[i * 2 + 0];
    if (numsteps == 0) return;                                  // "The ray doesn't matter. So no gradient onto the camera" (:2045-2048)
    const uint32_t base = sb.numsteps[i * 2 + 1];
    const uint32_t ig = ray_offset + sb.ray_indices[i];
    const uint32_t img = image_idx(ig, n_rays_global, n_images, sb.cdf.cdf_img);
    float uv_pdf = 1.0f; float uv[2] = {0.f, 0.f};
    if (sb.cdf.cdf_x_cond_y || distortion_gradient) {   // the pixel (and its pdf) of the ray, from the replayed rng stream (:2083-2086)
        rng.advance((int64_t)((uint64_t)ig * N_MAX_RANDOM_SAMPLES_PER_RAY));
        random_image_pos_training(rng, metas[img].res, snap, uv, &sb.cdf, img, &uv_pdf);
    }
    const V3 ray_o = ld3(sb.rays + (size_t)i * 6), ray_d = normalize3(ld3(sb.rays + (size_t)i * 6 + 3));
    const V3 wpd = v3(1.0f / (aabb.max[0] - aabb.min[0]), 1.0f / (aabb.max[1] - aabb.min[1]), 1.0f / (aabb.max[2] - aabb.min[2]));   // warp_position_derivative (:397-405)
    V3 go = v3(0, 0, 0), gd = v3(0, 0, 0);
    for (uint32_t j = 0; j < numsteps; ++j) {
        const float* c = coords + (size_t)(base + j) * 7; const float* g = coords_gradient + (size_t)(base + j) * 7;
        const V3 pos_gradient = v3(g[0] * wpd.x, g[1] * wpd.y, g[2] * wpd.z);
        go = go + pos_gradient;
        const V3 pos = unwarp_position(v3(c[0], c[1], c[2]), aabb);
        const float t = length3(pos - ray_o);                   // "further-away objects' position changes more rapidly as the direction changes"
        const V3 dir_gradient = v3(g[4] * 0.5f, g[5] * 0.5f, g[6] * 0.5f);   // warp_direction_derivative (:415-417)
        gd = gd + t * pos_gradient + dir_gradient;
    }
    if (distortion_gradient) {   // :2088-2100
        // "Projection of the raydir gradient onto the plane normal to raydir, because that's the only degree of motion that the raydir has",
        // then rotated into the image plane of the view (inverse of the rotation block, by cofactors as glm does) and splatted at the ray's pixel
        const float dp = gd.x * ray_d.x + gd.y * ray_d.y + gd.z * ray_d.z;
        const V3 og = gd - dp * ray_d;
        const float* m = metas[img].xform;
        const float a = m[0], b = m[3], c = m[6], d_ = m[1], e = m[4], f = m[7], g = m[2], h = m[5], k = m[8];
        const float id = 1.0f / (a * (e * k - f * h) - b * (d_ * k - f * g) + c * (d_ * h - e * g));
        const float ix = ((e * k - f * h) * og.x + (c * h - b * k) * og.y + (b * f - c * e) * og.z) * id;
        const float iy = ((f * g - d_ * k) * og.x + (a * k - c * g) * og.y + (c * d_ - a * f) * og.z) * id;
        const float val[2] = {ix / uv_pdf, iy / uv_pdf};
        const int rx = sb.distortion.res[0], ry = sb.distortion.res[1];
        const float xf = (float)rx * uv[0], yf = (float)ry * uv[1];
        const int tx = (int)xf, ty = (int)yf;
        const float wx = xf - (float)tx, wy = yf - (float)ty;
        const float w4[4] = {(1 - wx) * (1 - wy), wx * (1 - wy), (1 - wx) * wy, wx * wy};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int px = clampi(tx + (q & 1), 0, rx - 1), py = clampi(ty + (q >> 1), 0, ry - 1);
            const size_t o = 2 * ((size_t)px + (size_t)py * rx);
            atomicAdd(distortion_gradient + o, val[0] * w4[q]); atomicAdd(distortion_gradient_weight + o, w4[q]);
            atomicAdd(distortion_gradient + o + 1, val[1] * w4[q]); atomicAdd(distortion_gradient_weight + o + 1, w4[q]);
        }
    }
    if (cam_pos_gradient) { atomicAdd(cam_pos_gradient + img * 3, go.x / uv_pdf); atomicAdd(cam_pos_gradient + img * 3 + 1, go.y / uv_pdf); atomicAdd(cam_pos_gradient + img * 3 + 2, go.z / uv_pdf); }
    if (cam_rot_gradient) {                                     // averaged in log space: angle_axis = cross(ray.d, ray_gradient.d) (:2110-2122)
        const V3 aa = v3(ray_d.y * gd.z - ray_d.z * gd.y, ray_d.z * gd.x - ray_d.x * gd.z, ray_d.x * gd.y - ray_d.y * gd.x);
        atomicAdd(cam_rot_gradient + img * 3, aa.x / uv_pdf); atomicAdd(cam_rot_gradient + img * 3 + 1, aa.y / uv_pdf); atomicAdd(cam_rot_gradient + img * 3 + 2, aa.z / uv_pdf);
    }
}
void nerf_compute_cam_gradient(cudaStream_t s, uint32_t n_rays_bound, Aabb aabb, Pcg32 rng, bool snap_to_pixel_centers, const ImageMeta* metas, uint32_t n_images,
                               uint32_t ray_offset, uint32_t n_rays_global, const SamplerBuffers& sb, const float* coords_compacted, const float* coords_gradient,
                               float* cam_pos_gradient, float* cam_rot_gradient, float* distortion_gradient, float* distortion_gradient_weight) {
    if (n_rays_bound == 0) return;
    NGP_PROFILE_SCOPE(s, "cam_gradient", n_rays_bound);
    cam_gradient_kernel<<<div_round_up(n_rays_bound, 128), 128, 0, s>>>(aabb, rng, snap_to_pixel_centers, metas, n_images, ray_offset, n_rays_global, sb, coords_compacted,
                                                                      coords_gradient, cam_pos_gradient, cam_rot_gradient, distortion_gradient, distortion_gradient_weight);
    NGP_POST_LAUNCH();
}

// safe_divide (testbed_nerf.cu:2411-2417) fused with tiny-cuda-nn's Adam step on fp32 parameters that are all non-matrix (a TrainableBuffer has no
// layers): entries whose gradient is zero are skipped and keep their step count, no L2 term, lr bias-corrected per entry (recalled; unpinned)
__global__ void distortion_step_kernel(uint32_t n, float* __restrict__ params, float* __restrict__ gradient, const float* __restrict__ weight, float* __restrict__ m1,
                                       float* __restrict__ m2, uint32_t* __restrict__ steps, float lr, float beta1, float beta2, float epsilon, float loss_scale) {
    const uint32_t i = threadIdx.x + blockIdx.x * blockDim.x;
    if (i >= n) return;
    const float w = weight[i];
    const float divided = w > 0.0f ? gradient[i] / w : 0.0f;
    gradient[i] = divided;
    const float g = divided / loss_scale;
    if (g == 0.0f) return;
    const uint32_t step = ++steps[i];
    const float learning_rate = lr * sqrtf(1.0f - powf(beta2, (float)step)) / (1.0f - powf(beta1, (float)step));
    const float first = m1[i] = beta1 * m1[i] + (1.0f - beta1) * g;
    const float second = m2[i] = beta2 * m2[i] + (1.0f - beta2) * g * g;
    params[i] = params[i] - learning_rate / (sqrtf(second) + epsilon) * first;
}
void nerf_distortion_step(cudaStream_t s, uint32_t n_params, float* params, float* gradient, const float* gradient_weight, float* first_moment, float* second_moment,
                          uint32_t* steps, float learning_rate, float beta1, float beta2, float epsilon, float loss_scale) {
    if (n_params == 0) return;
    distortion_step_kernel<<<div_round_up(n_params, 256), 256, 0, s>>>(n_params, params, gradient, gradient_weight, first_moment, second_moment, steps, learning_rate, beta1, beta2,
                                                                     epsilon, loss_scale);
    NGP_POST_LAUNCH();
}

// construct_cdf_2d / construct_cdf_1d (testbed_nerf.cu:2354-2409)
__global__ void construct_cdf_2d_kernel(uint32_t n_images, uint32_t height, uint32_t width, const float* __restrict__ data, float* __restrict__ cdf_x_cond_y, float* __restrict__ cdf_y) {
    const uint32_t y = threadIdx.x + blockIdx.x * blockDim.x, img = threadIdx.y + blockIdx.y * blockDim.y;
    if (y >= height || img >= n_images) return;
    const float MIN_PDF = 0.01f;
    const size_t off = ((size_t)img * height + y) * width;
    float cum = 0;
    for (uint32_t x = 0; x < width; ++x) { cum += data[off + x] + 1e-10f; cdf_x_cond_y[off + x] = cum; }
    cdf_y[(size_t)img * height + y] = cum;
    const float norm = 1.0f / cum;                           // __frcp_rn
    for (uint32_t x = 0; x < width; ++x) cdf_x_cond_y[off + x] = (1.0f - MIN_PDF) * cdf_x_cond_y[off + x] * norm + MIN_PDF * (float)(x + 1) / (float)width;
}
__global__ void construct_cdf_1d_kernel(uint32_t n_images, uint32_t height, float* __restrict__ cdf_y, float* __restrict__ cdf_img) {
    const uint32_t img = threadIdx.x + blockIdx.x * blockDim.x;
    if (img >= n_images) return;
    const float MIN_PDF = 0.01f;
    float* c = cdf_y + (size_t)img * height;
    float cum = 0;
    for (uint32_t y = 0; y < height; ++y) { cum += c[y]; c[y] = cum; }
    cdf_img[img] = cum;
    const float norm = 1.0f / cum;
    for (uint32_t y = 0; y < height; ++y) c[y] = (1.0f - MIN_PDF) * c[y] * norm + MIN_PDF * (float)(y + 1) / (float)height;
}
void nerf_construct_error_cdfs(cudaStream_t s, uint32_t n_images, uint32_t height, uint32_t width, const float* error_map, float* cdf_x_cond_y, float* cdf_y, float* cdf_img) {
    const dim3 threads(16, 8, 1), blocks(div_round_up(height, threads.x), div_round_up(n_images, threads.y), 1);
    construct_cdf_2d_kernel<<<blocks, threads, 0, s>>>(n_images, height, width, error_map, cdf_x_cond_y, cdf_y);
    NGP_POST_LAUNCH();
    construct_cdf_1d_kernel<<<div_round_up(n_images, 128), 128, 0, s>>>(n_images, height, cdf_y, cdf_img);
    NGP_POST_LAUNCH();
}

// max_level of the un-compacted samples (testbed_nerf.cu:1438,1652-1657), one thread per kept ray: the rng stream of the ray is replayed up to
// the draw that follows the pixel position
__global__ void fill_max_level_kernel(Pcg32 rng, uint32_t n_images, const ImageMeta* __restrict__ metas, bool snap, uint32_t ray_offset, uint32_t n_rays_global,
                                      SamplerBuffers sb, float* __restrict__ max_level) {
    const uint32_t i = threadIdx.x + blockIdx.x * blockDim.x;
    if (i >= sb.counters[0]) return;
    const uint32_t ig = ray_offset + sb.ray_indices[i];
    rng.advance((int64_t)((uint64_t)ig * N_MAX_RANDOM_SAMPLES_PER_RAY));
    const ImageMeta& m = metas[image_idx(ig, n_rays_global, n_images, sb.cdf.cdf_img)];
    float uv[2]; random_image_pos_training(rng, m.res, snap, uv);   // (the draws are what matters here, not where they land)
    const float v = rng.next_float() * 2.0f;
    const uint32_t numsteps = sb.numsteps[i * 2 + 0], base = sb.numsteps[i * 2 + 1];
    for (uint32_t j = 0; j < numsteps; ++j) max_level[base + j] = v;
}
void nerf_fill_max_level(cudaStream_t s, uint32_t n_rays_bound, Pcg32 rng, uint32_t n_images, const ImageMeta* metas, bool snap_to_pixel_centers,
                         uint32_t ray_offset, uint32_t n_rays_global, const SamplerBuffers& sb, float* max_level) {
    if (n_rays_bound == 0) return;
    fill_max_level_kernel<<<div_round_up(n_rays_bound, 256), 256, 0, s>>>(rng, n_images, metas, snap_to_pixel_centers, ray_offset, n_rays_global, sb, max_level);
    NGP_POST_LAUNCH();
}

// fill_rollover_and_rescale<T> / fill_rollover<T> (tcnn, recalled; testbed_nerf.cu:4061-4069): entries [n_input, n_target) are
// cyclic copies of [0, n_input); the gradient copies are multiplied by n_input/n_target.
__global__ void fill_rollover_kernel(uint32_t n_target, const uint32_t* __restrict__ n_input_ptr, __half* __restrict__ dloss, float* __restrict__ coords,
                                     __half* __restrict__ enc, uint32_t enc_width, float* __restrict__ max_level) {
    const uint32_t i = threadIdx.x + blockIdx.x * blockDim.x;
    const uint32_t n_input = min(*n_input_ptr, n_target);
    if (i < n_input || i >= n_target || n_input == 0) return;
    const uint32_t src = i % n_input;
    // element-wise semantics of the reference: element e of the flattened [n x stride] array copies element e % (n_input*stride)
    const float scale = ((float)n_input * 4.0f) / ((float)n_target * 4.0f);
    const uint2 raw = *reinterpret_cast<const uint2*>(dloss + (size_t)src * 4);
    const __half* h = reinterpret_cast<const __half*>(&raw);
    __half2 o[2] = {__floats2half2_rn(__half2float(h[0]) * scale, __half2float(h[1]) * scale), __floats2half2_rn(__half2float(h[2]) * scale, __half2float(h[3]) * scale)};
    *reinterpret_cast<uint2*>(dloss + (size_t)i * 4) = *reinterpret_cast<uint2*>(o);
#pragma unroll
    for (int k = 0; k < 7; ++k) coords[(size_t)i * 7 + k] = coords[(size_t)src * 7 + k];
    if (enc) copy_enc_row(enc, enc, src, i, enc_width);
    if (max_level) max_level[i] = max_level[src];   // fill_rollover<float>(.., max_level_compacted) (:4066-4069)
}
void nerf_fill_rollover(cudaStream_t s, uint32_t n_target, const uint32_t* n_input_dev, __half* dloss, float* coords, __half* enc, uint32_t enc_width, float* max_level) {
    NGP_PROFILE_SCOPE(s, "fill_rollover", n_target);
    fill_rollover_kernel<<<div_round_up(n_target, 256), 256, 0, s>>>(n_target, n_input_dev, dloss, coords, enc, enc_width, max_level);
    NGP_POST_LAUNCH();
}


// =================================================================================================================
// Inference tracer (wavefront): init_rays_with_payload_kernel_nerf (:2228-2352) + advance_pos_nerf (:850-893),
// compact_kernel_nerf (:2198-2226), generate_next_nerf_network_inputs (:948-1022), composite_kernel_nerf (:1024-1230,
// Shade mode only), shade_kernel_nerf (:2164-2196).
// =================================================================================================================
// coarse bit of the 32^3 cell that holds pos (unit cube only; positions outside clamp to the border cell, whose dilated bit covers them)
NGP_D bool coarse_bit_at(const uint32_t* __restrict__ coarse, V3 p) {
    const uint32_t x = (uint32_t)clampi((int)floorf(p.x * COARSE_RES), 0, COARSE_RES - 1), y = (uint32_t)clampi((int)floorf(p.y * COARSE_RES), 0, COARSE_RES - 1),
                   z = (uint32_t)clampi((int)floorf(p.z * COARSE_RES), 0, COARSE_RES - 1);
    const uint32_t c = (z * COARSE_RES + y) * COARSE_RES + x;
    return (__ldg(coarse + (c >> 5)) >> (c & 31u)) & 1u;
}
// occupancy look-up of the tracer.  With the coarse table (max_mip == 0, walk inside the unit cube) a clear coarse bit answers "empty"
// from the 4 KB table in L1 instead of the 2 MB bitfield in L2: the dilated coarse cell is clear only if every fine cell in and around
// it is, so the answer is the bitfield's.
NGP_D bool tracer_occupied(V3 pos, const uint8_t* __restrict__ bitfield, uint32_t mip, const uint32_t* __restrict__ coarse) {
    if (coarse && !coarse_bit_at(coarse, pos)) return false;
    return density_grid_occupied_at(pos, bitfield, mip);
}
NGP_D float advance_to_occupied(float t, const Cone& cone, V3 o, V3 d, V3 idir, const uint8_t* __restrict__ bitfield, uint32_t min_mip, uint32_t max_mip, const Aabb& aabb,
                                const uint32_t* __restrict__ coarse = nullptr) {
    while (true) {   // if_unoccupied_advance_to_next_occupied_voxel (:811-845)
        const V3 pos = ray_at(o, d, t);
        if (t >= MAX_DEPTH || !aabb_contains(aabb, pos)) return MAX_DEPTH;
        uint32_t mip = (uint32_t)clampi((int)mip_from_pos(pos, NERF_CASCADES - 1), (int)min_mip, (int)max_mip);
        if (!bitfield || tracer_occupied(pos, bitfield, mip, coarse)) return t;
        while (mip < max_mip && !density_grid_occupied_at(pos, bitfield, mip + 1)) ++mip;
        t = advance_to_next_voxel(t, cone, pos, d, idir, mip);
    }
}

__global__ void init_rays_kernel(RenderCamera cam, uint32_t sample_index, bool snap, Aabb render_aabb, const uint8_t* __restrict__ bitfield,
                                 uint32_t max_mip, float cone_angle, RaysSoa rays, float* __restrict__ frame_rgba, float* __restrict__ frame_depth, float near_distance,
                                 const uint32_t* __restrict__ coarse) {
    const Cone cone(cone_angle);
    const uint32_t x = threadIdx.x + blockDim.x * blockIdx.x, yl = threadIdx.y + blockDim.y * blockIdx.y;
    const uint32_t W = cam.res[0], y = cam.y0 + yl * (uint32_t)cam.y_stride;
    if (x >= W || y >= (uint32_t)cam.y1) return;
    const uint32_t local = x + W * yl, idx = x + W * y;
    float off[2]; ld_random_pixel_offset(snap ? 0 : sample_index, off);
    const float u = ((float)x + off[0]) / (float)W, v = ((float)y + off[1]) / (float)cam.res[1];
    V3 o, d;
    uv_to_ray(cam.camera, cam.res, cam.focal_length, cam.screen_center, cam.lens_mode, cam.lens_params, u, v, o, d, &cam.distortion);
    if (near_distance != 0.f) o = ray_at(o, d, near_distance);   // uv_to_ray: origin += dir * near_distance (dir not yet normalised)
    NerfPayload& p = rays.payload[local];
    p.max_weight = 0.0f; p.idx = idx; p.n_steps = 0; p.alive = 0; p.pad = 0;
    p.origin[0] = o.x; p.origin[1] = o.y; p.origin[2] = o.z;
    frame_depth[local] = MAX_DEPTH;
    reinterpret_cast<float4*>(frame_rgba)[local] = make_float4(0, 0, 0, 0);
    reinterpret_cast<float4*>(rays.rgba)[local] = make_float4(0, 0, 0, 0);
    rays.depth[local] = 0.0f;
    d = normalize3(d);
    p.dir[0] = d.x; p.dir[1] = d.y; p.dir[2] = d.z;
    float tmin, tmax; aabb_ray_intersect(render_aabb, o, d, tmin, tmax);
    float t = fmaxf(tmin, 0.0f) + 1e-6f;
    p.t = t;
    if (!aabb_contains(render_aabb, ray_at(o, d, t))) return;
    // Four fifths of the pixels of a typical frame see nothing.  The exact walk would still follow each of those rays through ~160 empty
    // voxels, one dependent look-up per voxel; the conservative coarse test (see ray_touches_coarse) proves "no occupied cell on the ray"
    // from ~35 look-ups in a 4 KB table instead, and the ray ends up exactly as the walk would leave it: not alive, t at the aabb entry.
    if (coarse && !ray_touches_coarse(coarse, o, d, fmaxf(tmin, 0.0f), tmax, 0, 1)) return;
    // advance_pos_nerf: jittered first step, then skip to the first occupied voxel
    const V3 idir = v3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    t = advance_n_steps(t, cone, ld_random_val(sample_index, idx * 786433u));
    t = advance_to_occupied(t, cone, o, d, idir, bitfield, 0, max_mip, render_aabb, coarse);
    if (t >= MAX_DEPTH) return;
    p.t = t; p.alive = 1;
}

__global__ void __launch_bounds__(SCAN_BLOCK) compact_count_kernel(const TraceCtl* __restrict__ prev, const NerfPayload* __restrict__ payloads, uint32_t* __restrict__ block_alive) {
    const uint32_t i = threadIdx.x + blockIdx.x * blockDim.x, n = prev->n_alive;
    if (blockIdx.x * blockDim.x >= n) { if (threadIdx.x == 0) block_alive[blockIdx.x] = 0; return; }   // block-uniform: launches are sized by a host-side bound
    uint32_t total;
    block_exclusive_scan((i < n && payloads[i].alive) ? 1u : 0u, total);
    if (threadIdx.x == 0) block_alive[blockIdx.x] = total;
}
__global__ void __launch_bounds__(SCAN_BLOCK) compact_scatter_kernel(const TraceCtl* __restrict__ prev, TraceCtl* __restrict__ cur, TraceCtl* __restrict__ host_copy,
                                                                     uint32_t capacity, uint32_t target_samples, RaysSoa src, RaysSoa dst, RaysSoa hit, const uint32_t* __restrict__ block_alive,
                                                                     uint32_t* __restrict__ counters) {
    const uint32_t i = threadIdx.x + blockIdx.x * blockDim.x, n = prev->n_alive;
    if (i == 0) {   // loop control of the iteration that follows (counters[0] = total alive, written by the scan kernel before this launch)
        TraceCtl c;
        c.n_alive = counters[0];
        c.n_steps = trace_steps(c.n_alive, capacity, target_samples);
        c.n_elements = (c.n_alive * c.n_steps + (BATCH_GRANULARITY - 1u)) / BATCH_GRANULARITY * BATCH_GRANULARITY;   // n_alive * n_steps <= capacity < 2^32
        c.current_step = prev->current_step + prev->n_steps;
        *cur = c;
        *host_copy = c;     // host-mapped: read by the host after the event recorded behind this kernel
    }
    if (blockIdx.x * blockDim.x >= n) return;
    const bool in = i < n;
    const bool alive = in && src.payload[i].alive;
    uint32_t total;
    const uint32_t pos = block_alive[blockIdx.x] + block_exclusive_scan(alive ? 1u : 0u, total);
    if (!in) return;
    const float4 c = reinterpret_cast<const float4*>(src.rgba)[i];
    if (alive) {
        dst.payload[pos] = src.payload[i]; reinterpret_cast<float4*>(dst.rgba)[pos] = c; dst.depth[pos] = src.depth[i];
    } else if (c.w > 0.001f) {
        const uint32_t h = atomicAdd(&counters[1], 1u);   // order of finished rays is irrelevant (one pixel each)
        hit.payload[h] = src.payload[i]; reinterpret_cast<float4*>(hit.rgba)[h] = c; hit.depth[h] = src.depth[i];
    }
}

// generate_next_nerf_network_inputs (:948-1022): n_steps times { skip to the next occupied voxel; emit a sample; step }.
// Written as ONE loop whose body either emits a sample or skips a voxel.  The nested form (a skip loop inside the step loop) makes a warp
// pay the longest skip of any of its lanes once per step: lanes reach their long skips — above all the final walk out of the scene, up to
// ~160 voxels — at different step indices, and the warp runs those walks one after the other (measured: 300 us per iteration in the tail
// of a 3840x2160 frame, 37 us per step).  In the flat loop every lane advances on every trip whatever it is doing, and a warp takes
// max over lanes of (samples + skips) trips.  The last bitfield byte is kept in a register: consecutive samples are 1/1024 * sqrt(3)
// apart, a voxel is 1/128 wide and a byte covers a 2x2x2 block of voxels, so most look-ups of a ray inside a surface repeat the byte.
// Same float operations per sample and per skip as the reference's two routines, hence the same samples.
__global__ void next_inputs_kernel(const TraceCtl* __restrict__ ctl, Aabb render_aabb, Aabb train_aabb, RaysSoa rays, float* __restrict__ input,
                                   const uint8_t* __restrict__ bitfield, uint32_t max_mip, float cone_angle, const uint32_t* __restrict__ coarse) {
    const Cone cone(cone_angle);
    const uint32_t i = threadIdx.x + blockIdx.x * blockDim.x, n_alive = ctl->n_alive, n_steps = ctl->n_steps;
    if (i >= n_alive) return;
    NerfPayload& p = rays.payload[i];
    if (!p.alive) return;
    const V3 o = ld3(p.origin), d = ld3(p.dir), idir = v3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    const V3 wd = warp_direction(d);
    float t = p.t;
    uint32_t j = 0, cached_byte_idx = 0xFFFFFFFFu, cached_byte = 0;
    while (j < n_steps) {
        const V3 pos = ray_at(o, d, t);
        if (t >= MAX_DEPTH || !aabb_contains(render_aabb, pos)) { p.n_steps = (uint16_t)j; return; }   // advance_to_occupied returned MAX_DEPTH (:1003-1007)
        uint32_t mip = (uint32_t)clampi((int)mip_from_pos(pos, NERF_CASCADES - 1), 0, (int)max_mip);
        bool occupied = true;
        if (bitfield) {
            occupied = !(coarse && !coarse_bit_at(coarse, pos));
            if (occupied) {
                const uint32_t idx = cascaded_grid_idx_at(pos, mip);
                if (idx == 0xFFFFFFFFu) occupied = false;
                else {
                    const uint32_t byte_idx = idx / 8 + NERF_GRID_N_CELLS / 8 * mip;
                    if (byte_idx != cached_byte_idx) { cached_byte_idx = byte_idx; cached_byte = __ldg(bitfield + byte_idx); }
                    occupied = (cached_byte >> (idx % 8)) & 1u;
                }
            }
        }
        if (occupied) {
            const float dt = calc_dt(t, cone);
            float* c = input + (size_t)(i + j * n_alive) * 7;
            const V3 wp = warp_position(pos, train_aabb);
            c[0] = wp.x; c[1] = wp.y; c[2] = wp.z; c[3] = warp_dt(dt); c[4] = wd.x; c[5] = wd.y; c[6] = wd.z;
            t += dt; ++j;
        } else {
            while (mip < max_mip && !density_grid_occupied_at(pos, bitfield, mip + 1)) ++mip;
            t = advance_to_next_voxel(t, cone, pos, d, idir, mip);
        }
    }
    p.t = t; p.n_steps = (uint16_t)n_steps;
}

__global__ void composite_kernel(const TraceCtl* __restrict__ ctl, Aabb aabb, float cam_fwd_x, float cam_fwd_y, float cam_fwd_z, float cam_x, float cam_y, float cam_z,
                                 RaysSoa rays, const float* __restrict__ input, const __half* __restrict__ output, int rgb_act, int density_act,
                                 float min_transmittance, int render_mode, float depth_scale) {
    const uint32_t i = threadIdx.x + blockIdx.x * blockDim.x, n_alive = ctl->n_alive, n_steps = ctl->n_steps, current_step = ctl->current_step;
    if (i >= n_alive) return;
    NerfPayload& p = rays.payload[i];
    if (!p.alive) return;
    float4 c = reinterpret_cast<float4*>(rays.rgba)[i];
    float local_depth = rays.depth[i];
    const V3 cam_fwd = v3(cam_fwd_x, cam_fwd_y, cam_fwd_z), cam_pos = v3(cam_x, cam_y, cam_z);
    const uint32_t actual = p.n_steps;
    uint32_t j = 0;
    for (; j < actual; ++j) {
        float o[4]; load_out4(output + (size_t)(i + j * n_alive) * 4, o);
        const float* in = input + (size_t)(i + j * n_alive) * 7;
        const V3 pos = unwarp_position(v3(in[0], in[1], in[2]), aabb);
        const float T = 1.f - c.w;
        const float dt = unwarp_dt(in[3]);
        const float alpha = alpha_from_density(network_to_density(o[3], density_act), dt);
        const float weight = alpha * T;
        V3 rgb = v3(network_to_rgb(o[0], rgb_act), network_to_rgb(o[1], rgb_act), network_to_rgb(o[2], rgb_act));
        if (render_mode != RM_SHADE) {   // :1182-1209
            if (render_mode == RM_POSITIONS) rgb = v3((pos.x - 0.5f) / 2.0f + 0.5f, (pos.y - 0.5f) / 2.0f + 0.5f, (pos.z - 0.5f) / 2.0f + 0.5f);
            else if (render_mode == RM_ENCODING_VIS) rgb = v3(in[0], in[1], in[2]);
            else if (render_mode == RM_DEPTH) { const float dd = dot3(cam_fwd, pos - cam_pos) * depth_scale; rgb = v3(dd, dd, dd); }
            else if (render_mode == RM_AO) rgb = v3(alpha, alpha, alpha);
        }
        c.x += rgb.x * weight; c.y += rgb.y * weight; c.z += rgb.z * weight;
        c.w += weight;
        if (weight > p.max_weight) { p.max_weight = weight; local_depth = dot3(cam_fwd, pos - cam_pos); }
        if (c.w > (1.0f - min_transmittance)) { const float a = c.w; c.x /= a; c.y /= a; c.z /= a; c.w /= a; break; }
    }
    if (j < n_steps) { p.alive = 0; p.n_steps = (uint16_t)(j + current_step); }
    reinterpret_cast<float4*>(rays.rgba)[i] = c;
    rays.depth[i] = local_depth;
}

__global__ void shade_kernel(const uint32_t* __restrict__ n_hit_dev, RaysSoa hit, bool linear_colors, uint32_t W, uint32_t y0, uint32_t y_stride,
                             float* __restrict__ frame_rgba, float* __restrict__ frame_depth, int render_mode) {
    const uint32_t i = threadIdx.x + blockIdx.x * blockDim.x;
    if (i >= *n_hit_dev) return;
    float4 t = reinterpret_cast<const float4*>(hit.rgba)[i];
    if (render_mode == RM_COST) { const float col = (float)hit.payload[i].n_steps / 128; t = make_float4(col, col, col, 1.0f); }
    if (!linear_colors && (render_mode == RM_SHADE || render_mode == RM_SLICE)) { t.x = srgb_to_linear(t.x); t.y = srgb_to_linear(t.y); t.z = srgb_to_linear(t.z); }   // accumulate in linear colours
    const uint32_t idx = hit.payload[i].idx, local = idx % W + W * ((idx / W - y0) / y_stride);
    const float4 f = reinterpret_cast<const float4*>(frame_rgba)[local];
    reinterpret_cast<float4*>(frame_rgba)[local] = make_float4(t.x + f.x * (1.0f - t.w), t.y + f.y * (1.0f - t.w), t.z + f.z * (1.0f - t.w), t.w + f.w * (1.0f - t.w));
    if (render_mode != RM_SLICE && t.w > 0.2f) frame_depth[local] = hit.depth[i];
}

void nerf_init_rays(cudaStream_t s, const RenderCamera& cam, uint32_t sample_index, bool snap, Aabb render_aabb, const uint8_t* bitfield,
                    uint32_t max_mip, float cone, RaysSoa rays, float* frame_rgba, float* frame_depth, float near_distance, const uint32_t* coarse) {
    NGP_PROFILE_SCOPE(s, "render_init_rays", (uint64_t)cam.res[0] * (uint64_t)((cam.y1 - cam.y0 + cam.y_stride - 1) / cam.y_stride));
    const dim3 threads = {16, 8, 1};
    const dim3 blocks = {div_round_up((uint32_t)cam.res[0], 16), div_round_up(div_round_up((uint32_t)(cam.y1 - cam.y0), (uint32_t)cam.y_stride), 8), 1};
    init_rays_kernel<<<blocks, threads, 0, s>>>(cam, sample_index, snap, render_aabb, bitfield, max_mip, cone, rays, frame_rgba, frame_depth, near_distance, coarse);
    NGP_POST_LAUNCH();
}
void nerf_compact_rays(cudaStream_t s, uint32_t n_bound, RaysSoa src, RaysSoa dst, RaysSoa hit, uint32_t* counters, uint32_t* scan_scratch, const TraceCtl* prev,
                       TraceCtl* cur, TraceCtl* host_copy, uint32_t capacity, uint32_t target_samples) {
    NGP_PROFILE_SCOPE(s, "render_compact", n_bound);
    const uint32_t blocks = std::max(div_round_up(n_bound, SCAN_BLOCK), 1u);
    compact_count_kernel<<<blocks, SCAN_BLOCK, 0, s>>>(prev, src.payload, scan_scratch); NGP_POST_LAUNCH();
    scan_block_totals_kernel<<<1, 1024, 0, s>>>(blocks, scan_scratch, nullptr, &counters[0], nullptr, false); NGP_POST_LAUNCH();
    compact_scatter_kernel<<<blocks, SCAN_BLOCK, 0, s>>>(prev, cur, host_copy, capacity, target_samples, src, dst, hit, scan_scratch, counters); NGP_POST_LAUNCH();
}
void nerf_generate_next_inputs(cudaStream_t s, uint32_t n_bound, const TraceCtl* ctl, Aabb render_aabb, Aabb train_aabb, RaysSoa rays, float* input,
                               const uint8_t* bitfield, uint32_t max_mip, float cone, const uint32_t* coarse) {
    if (n_bound == 0) return;
    NGP_PROFILE_SCOPE(s, "render_next_inputs", n_bound);
    next_inputs_kernel<<<div_round_up(n_bound, 128), 128, 0, s>>>(ctl, render_aabb, train_aabb, rays, input, bitfield, max_mip, cone, coarse);
    NGP_POST_LAUNCH();
}
void nerf_composite(cudaStream_t s, uint32_t n_bound, const TraceCtl* ctl, Aabb train_aabb, const float* camera, RaysSoa rays, const float* input,
                    const __half* output, int rgb_act, int density_act, float min_transmittance, int render_mode, float depth_scale) {
    if (n_bound == 0) return;
    NGP_PROFILE_SCOPE(s, "render_composite", n_bound);
    composite_kernel<<<div_round_up(n_bound, 128), 128, 0, s>>>(ctl, train_aabb, camera[6], camera[7], camera[8], camera[9], camera[10], camera[11], rays, input, output,
                                                                rgb_act, density_act, min_transmittance, render_mode, depth_scale);
    NGP_POST_LAUNCH();
}
void nerf_shade(cudaStream_t s, uint32_t n_bound, const uint32_t* n_hit_dev, RaysSoa hit, bool linear_colors, const RenderCamera& cam, float* frame_rgba, float* frame_depth,
                int render_mode) {
    if (n_bound == 0) return;
    shade_kernel<<<div_round_up(n_bound, 128), 128, 0, s>>>(n_hit_dev, hit, linear_colors, (uint32_t)cam.res[0], (uint32_t)cam.y0, (uint32_t)cam.y_stride, frame_rgba, frame_depth,
                                                            render_mode);
    NGP_POST_LAUNCH();
}

}  // namespace ngp

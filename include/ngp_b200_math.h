/* ngp_b200_math.h — explicitly rounded fp32 logf / expf shared by the sm_100a kernels and the CPU oracle.
 *
 * Why: the exponential stepping of the NeRF marcher (reference src/testbed_nerf.cu:124-184, `to_stepping_space` /
 * `from_stepping_space`) feeds logf/expf results into integer decisions (which occupancy cell a sample falls in), and
 * the compositor's transmittance cut-off (`T < 1e-4`, src/testbed_nerf.cu:1710-1722) decides how many samples a ray
 * keeps.  libm and the CUDA math library differ in the last ulp, so "bit-exact traversal" cannot be asserted across
 * them.  The reference itself is compiled with --use_fast_math (CMakeLists.txt:84), i.e. its own results are those of
 * the ~2 ulp hardware approximations; any fp32 evaluation within a couple of ulp is as faithful as another.
 *
 * This file is the ONE evaluation both sides use: plain IEEE fp32 operations and explicit fmaf() only (classic
 * Cephes-style range reduction + minimax polynomial evaluated in Estrin form, < 2 ulp), no library call, no contraction left to the compiler
 * (the library is built with nvcc --fmad=false, the oracle with g++ -ffp-contract=off).  Same source, same bits.
 */
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define NGP_MATH_HD __host__ __device__ __forceinline__
#else
#define NGP_MATH_HD static inline
#endif

NGP_MATH_HD uint32_t ngp_f2u(float x) {
#if defined(__CUDA_ARCH__)
    return __float_as_uint(x);
#else
    uint32_t u; memcpy(&u, &x, 4); return u;
#endif
}
NGP_MATH_HD float ngp_u2f(uint32_t u) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    float x; memcpy(&x, &u, 4); return x;
#endif
}

/* natural logarithm, x > 0 finite (normal or denormal); other inputs follow IEEE conventions (0 -> -inf, <0 / nan -> nan, inf -> inf) */
NGP_MATH_HD float ngp_logf(float x) {
    uint32_t u = ngp_f2u(x);
    if (u == 0u || u == 0x80000000u) return ngp_u2f(0xff800000u);          /* -inf */
    if (u >> 31) return ngp_u2f(0x7fc00000u);                                /* nan */
    if (u >= 0x7f800000u) return x;                                          /* inf / nan */
    int e = 0;
    if (u < 0x00800000u) { x = x * 8388608.0f; u = ngp_f2u(x); e = -23; }   /* denormal: scale by 2^23 (exact) */
    e += (int)(u >> 23) - 126;                                               /* x = m * 2^e, m in [0.5, 1) */
    float m = ngp_u2f((u & 0x007fffffu) | 0x3f000000u);
    if (m < 0.707106781186547524f) { e -= 1; m = m + m - 1.0f; } else { m = m - 1.0f; }
#ifdef NGP_AB_OLD_MATH   /* A/B timing builds only (csrc/Makefile target `ab`): the Horner form this file had before */
    const float z = m * m;
    float y = 7.0376836292e-2f;
    y = fmaf(y, m, -1.1514610310e-1f); y = fmaf(y, m, 1.1676998740e-1f); y = fmaf(y, m, -1.2420140846e-1f); y = fmaf(y, m, 1.4249322787e-1f);
    y = fmaf(y, m, -1.6668057665e-1f); y = fmaf(y, m, 2.0000714765e-1f); y = fmaf(y, m, -2.4999993993e-1f); y = fmaf(y, m, 3.3333331174e-1f);
    y = y * m * z;
    const float fe = (float)e;
    y = fmaf(-2.12194440e-4f, fe, y);
    y = fmaf(-0.5f, z, y);
    float r = m + y;
    r = fmaf(0.693359375f, fe, r);
    return r;
#else
    /* degree-8 minimax polynomial in Estrin form: the marcher evaluates this on the serial chain of every ray, where the depth of the
       dependency chain (4 fused operations here, 8 in Horner form) is what costs time, not the operation count */
    const float z = m * m, z2 = z * z, mz = m * z, fe = (float)e;
    const float b1 = fmaf(-1.1514610310e-1f, m, 1.1676998740e-1f);
    const float b2 = fmaf(-1.2420140846e-1f, m, 1.4249322787e-1f);
    const float b3 = fmaf(-1.6668057665e-1f, m, 2.0000714765e-1f);
    const float b4 = fmaf(-2.4999993993e-1f, m, 3.3333331174e-1f);
    const float w = fmaf(7.0376836292e-2f, z2, fmaf(b1, z, b2));
    const float poly = fmaf(w, z2, fmaf(b3, z, b4));
    const float tail = fmaf(-0.5f, z, -2.12194440e-4f * fe);
    float r = m + fmaf(poly, mz, tail);
    r = fmaf(0.693359375f, fe, r);
    return r;
#endif
}

/* e^x; overflow -> +inf, underflow -> 0 (through denormals), nan -> nan */
NGP_MATH_HD float ngp_expf(float x) {
    if (!(x == x)) return x;
    if (x > 88.72283905206835f) return ngp_u2f(0x7f800000u);
    if (x < -103.972084045410f) return 0.0f;
    const float fn = floorf(fmaf(1.44269504088896341f, x, 0.5f));
    float r = fmaf(fn, -0.693359375f, x);
    r = fmaf(fn, 2.12194440e-4f, r);
    const float z = r * r;
#ifdef NGP_AB_OLD_MATH
    float p = 1.9875691500e-4f;
    p = fmaf(p, r, 1.3981999507e-3f); p = fmaf(p, r, 8.3334519073e-3f); p = fmaf(p, r, 4.1665795894e-2f); p = fmaf(p, r, 1.6666665459e-1f); p = fmaf(p, r, 5.0000001201e-1f);
#else
    const float a = fmaf(1.9875691500e-4f, r, 1.3981999507e-3f), b = fmaf(8.3334519073e-3f, r, 4.1665795894e-2f), c = fmaf(1.6666665459e-1f, r, 5.0000001201e-1f);
    const float p = fmaf(fmaf(a, z, b), z, c);      /* degree 5, Estrin form (depth 3 instead of 5) */
#endif
    float y = fmaf(p, z, r) + 1.0f;
    /* y * 2^n in two exact power-of-two factors (n in [-150, 128]): the second multiplication rounds once when the
       result is denormal — the same single rounding on every IEEE machine */
    const int n = (int)fn, n1 = n / 2, n2 = n - n1;
    y = y * ngp_u2f((uint32_t)(n1 + 127) << 23);
    y = y * ngp_u2f((uint32_t)(n2 + 127) << 23);
    return y;
}

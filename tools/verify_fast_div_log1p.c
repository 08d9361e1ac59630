#include <stdio.h>
#include <stdint.h>
#include <string.h>
#include <math.h>
#include <omp.h>
#include <stdlib.h>
#include "../include/ngp_b200_math.h"   /* the library's own logf: log1p_c = ngp_logf(1 + cone) */
// Exhaustive check for the general marcher's second division (to_stepping_space: logf(t) / log1p_c, testbed_nerf.cu:124-140):
// for every float x with 1e-30 <= |x| < 8 (logf(t) over the whole logarithmic regime at < t <= bt lies in [-0.84, 4.02] at the default
// cone angle 1/256), is  q2 = fmaf(fmaf(-q, L, x), rL, q), q = x * rL  ==  x / L  (IEEE round-to-nearest)?   L = logf(1 + cone).
// usage: verify_fast_div_log1p [cone_angle]      (gcc -O2 -fopenmp -ffp-contract=off verify_fast_div_log1p.c -lm; last run: cone 1/256, L = 0x3b7f8055, 1 723 562 816 floats, 0 mismatches)
int main(int argc, char** argv) {
    const float cone = argc > 1 ? (float)atof(argv[1]) : 1.0f / 256.0f;
    const float L = ngp_logf(1.0f + cone);
    const float rL = 1.0f / L;
    uint32_t hi, lo; float top = 8.0f, bot = 1e-30f; memcpy(&hi, &top, 4); memcpy(&lo, &bot, 4);
    long long bad = 0, bad1 = 0;
    #pragma omp parallel for reduction(+:bad,bad1) schedule(static)
    for (uint32_t b = lo; b < hi; ++b) {
        for (int sgn = 0; sgn < 2; ++sgn) {
            uint32_t bits = b | ((uint32_t)sgn << 31);
            float x; memcpy(&x, &bits, 4);
            float ref = x / L;
            float q = x * rL;
            float r = fmaf(-q, L, x);
            float q2 = fmaf(r, rL, q);
            if (q2 != ref) ++bad;
            if (q != ref) ++bad1;
        }
    }
    printf("cone=%.9g L=%.9g rL=%.9g  checked %lld floats: refined mismatches %lld, plain x*rL mismatches %lld\n", cone, L, rL, 2LL * (hi - lo), bad, bad1);
    return bad != 0;
}

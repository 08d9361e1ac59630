#include <stdio.h>
#include <stdint.h>
#include <string.h>
#include <math.h>
#include <omp.h>
// exhaustive check: for every float t in [0, 2^11), is  q2 = fmaf(fmaf(-q, S, t), rS, q), q = t*rS  ==  t / S  (IEEE RN) ?
int main() {
    const float S = 1.73205080757f / 1024.0f;
    const float rS = 1.0f / S;
    uint32_t hi, lo; float top = 2048.0f, bot = 1e-30f; memcpy(&hi, &top, 4); memcpy(&lo, &bot, 4);
    long long bad = 0, bad1 = 0;
    #pragma omp parallel for reduction(+:bad,bad1) schedule(static)
    for (uint32_t b = lo; b < hi; ++b) {
        float t; memcpy(&t, &b, 4);
        float ref = t / S;
        float q = t * rS;
        float r = fmaf(-q, S, t);
        float q2 = fmaf(r, rS, q);
        if (q2 != ref) ++bad;
        if (q != ref) ++bad1;
    }
    printf("S=%.9g rS=%.9g  checked %u floats: refined mismatches %lld, plain t*rS mismatches %lld\n", S, rS, hi, bad, bad1);
    return 0;
}

/* Exhaustive check of the CUDA kernels' dequantise step (emb_kernels.cuh dequant()) against the reference's
 * arithmetic, upmem/include/emb_host.h:210: final = (float)s / pow(10,9) -- int32 -> float (RNE), divide in
 * double, round to float; equal to one IEEE fp32 divide by 1e9f (checked here as well).  The kernels compute
 * that divide as q0 = RN(x*r), e = fma(-q0, 1e9f, x), q = fma(e, r, q0) with r = RN(1/1e9f); fmaf() is the same
 * correctly rounded fused multiply-add the GPU's fma.rn.f32 is.  All 2^32 int32 inputs.  Prints the two
 * mismatch counts; exit status 0 iff both are 0. */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

int main(void)
{
    const float b = 1.0e9f;
    const float r = 1.0f / 1.0e9f;
    unsigned long long bad_fast = 0, bad_div = 0;
#pragma omp parallel for reduction(+ : bad_fast, bad_div) schedule(static)
    for (long long k = -2147483648LL; k <= 2147483647LL; k++) {
        const float x = (float)(int32_t)k;
        const float ref = (float)((double)x / pow(10, 9));  /* emb_host.h:210 */
        const float dv = x / b;
        const float q0 = x * r;
        const float e = fmaf(-q0, b, x);
        const float q = fmaf(e, r, q0);
        uint32_t a, c, d;
        memcpy(&a, &ref, 4);
        memcpy(&c, &q, 4);
        memcpy(&d, &dv, 4);
        bad_fast += a != c;
        bad_div += a != d;
    }
    uint32_t rb;
    memcpy(&rb, &r, 4);
    printf("reciprocal 0x%08x fast_mismatches %llu fp32_divide_mismatches %llu\n", rb, bad_fast, bad_div);
    return (bad_fast || bad_div) ? 1 : 0;
}

/* hostbw.c — host memory-bandwidth probe for tools/e2e_scaling.py: OpenMP copy and add loops over
 * 32-bit words, nothing else.  Developer tool; not part of the product and not the oracle. */
#include <stddef.h>
#include <stdint.h>
#include <omp.h>

void hostbw_copy(uint32_t *restrict dst, const uint32_t *restrict src, size_t n, int threads)
{
#pragma omp parallel for schedule(static) num_threads(threads)
    for (size_t i = 0; i < n; ++i) dst[i] = src[i];
}

void hostbw_add(float *restrict dst, const float *restrict a, const float *restrict b, size_t n, int threads)
{
#pragma omp parallel for schedule(static) num_threads(threads)
    for (size_t i = 0; i < n; ++i) dst[i] = a[i] + b[i];
}

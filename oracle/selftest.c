/*
 * selftest.c — drives every oracle entry point over small and ragged sizes and prints one checksum
 * line per case.  TEST INFRASTRUCTURE (see oracle.h).  Built twice by the Makefile: plain, and with
 * -fsanitize=address,undefined; tests/test_oracle.py requires both outputs to be identical and
 * sanitizer-clean, and compares the lines with the library loaded through ctypes.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "oracle.h"

static void line(const char *what, const char *dtype, int mode, size_t n, const void *buf, size_t nbytes)
{
    uint64_t c[2];
    oracle_checksum(buf, nbytes, 3, c, 1);
    printf("%s %s mode=%d n=%zu %016llx:%016llx\n", what, dtype, mode, n, (unsigned long long)c[0], (unsigned long long)c[1]);
}

int main(void)
{
    const size_t sizes[] = {0, 1, 2, 7, 4099, 100003};
    const float s = -0.333251953125f;
    for (int mode = 0; mode < 3; ++mode) {
        for (size_t k = 0; k < sizeof sizes / sizeof sizes[0]; ++k) {
            const size_t n = sizes[k];
            /* exact-size heap blocks so the sanitizer sees any access past the end */
            float *x = malloc(n * 4 + 1), *y = malloc(n * 4 + 1), *d = malloc(n * 4 + 1);
            uint16_t *hx = malloc(n * 2 + 1), *hy = malloc(n * 2 + 1), *hd = malloc(n * 2 + 1);
            if (!x || !y || !d || !hx || !hy || !hd) return 2;
            oracle_fill_f32(x, ((uint64_t)1 << 32) - 5, n, 7, 0, mode);
            oracle_fill_f32(y, ((uint64_t)1 << 32) - 5, n, 7, 1, mode);
            oracle_fill_bf16(hx, 11, n, 7, 0, mode);
            oracle_fill_bf16(hy, 11, n, 7, 1, mode);
            for (int threads = 1; threads <= 3; threads += 2) {
                oracle_add_f32(d, x, y, n, threads);      line("add", "f32", mode, n, d, n * 4);
                oracle_copy_f32(d, x, n, threads);        line("copy", "f32", mode, n, d, n * 4);
                oracle_scale_f32(d, s, x, n, threads);    line("scale", "f32", mode, n, d, n * 4);
                oracle_triad_f32(d, s, x, y, n, threads); line("triad", "f32", mode, n, d, n * 4);
                oracle_add_bf16(hd, hx, hy, n, threads);      line("add", "bf16", mode, n, hd, n * 2);
                oracle_copy_bf16(hd, hx, n, threads);         line("copy", "bf16", mode, n, hd, n * 2);
                oracle_scale_bf16(hd, s, hx, n, threads);     line("scale", "bf16", mode, n, hd, n * 2);
                oracle_triad_bf16(hd, s, hx, hy, n, threads); line("triad", "bf16", mode, n, hd, n * 2);
            }
            free(x); free(y); free(d); free(hx); free(hy); free(hd);
        }
    }
    printf("selftest %d %d %d\n", oracle_vector_add_selftest(50000, 20261106u, -1), oracle_vector_add_selftest(50000, 20261106u, 123),
           oracle_vector_add_selftest(0, 1, -1));
    return 0;
}

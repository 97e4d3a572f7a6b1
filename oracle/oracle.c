/*
 * oracle.c — see oracle.h.  TEST INFRASTRUCTURE ONLY; PARITY UNPINNED at the
 * reference level (the reference's Step 9 payload is `nvidia-smi`,
 * /root/reference/README.md:313-314, and computes nothing).
 *
 * Each loop restates the semantics BASELINE.json's north_star names
 * ("cuda-samples vectorAdd ... fp32 C=A+B", "STREAM-style copy, scale, add
 * and triad", "triad using explicit fmaf on both sides").
 *
 * Build: gcc -O2 -std=c11 -fPIC -shared -fopenmp -ffp-contract=off
 */
#include "oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

/* ------------------------------------------------------------------------ */
/* generator                                                                 */

static inline uint32_t mix32(uint32_t x)
{
    x ^= x >> 16;
    x *= 0x7FEB352Du;
    x ^= x >> 15;
    x *= 0x846CA68Bu;
    x ^= x >> 16;
    return x;
}

uint32_t oracle_stream_key(uint32_t seed, uint32_t stream)
{
    return mix32(seed ^ (0x9E3779B9u * (stream + 1u)));
}

uint32_t oracle_hash32(uint32_t key, uint64_t index)
{
    uint32_t lo = (uint32_t)index;
    uint32_t hi = (uint32_t)(index >> 32);
    return mix32(lo ^ key ^ (hi * 0x9E3779B9u));
}

uint32_t oracle_gen_f32_bits(uint32_t key, uint64_t index, int mode)
{
    uint32_t h = oracle_hash32(key, index);
    if (mode == ORACLE_GEN_UNIT) {
        /* k * 2^-24, k < 2^24: exact in binary32 */
        float f = (float)(h >> 8) * 0x1p-24f;
        uint32_t bits;
        memcpy(&bits, &f, 4);
        return bits;
    }
    if (mode == ORACLE_GEN_DENORM)
        return h & 0x81FFFFFFu;
    if ((h & 0x7F800000u) == 0x7F800000u) /* Inf/NaN -> largest finite binade */
        h &= ~0x00800000u;
    return h;
}

uint16_t oracle_gen_bf16_bits(uint32_t key, uint64_t index, int mode)
{
    uint32_t h = oracle_hash32(key, index);
    if (mode == ORACLE_GEN_UNIT) {
        /* k * 2^-8, k < 2^8: 8 significant bits, exact in bf16 */
        float f = (float)(h >> 24) * 0x1p-8f;
        uint32_t bits;
        memcpy(&bits, &f, 4);
        return (uint16_t)(bits >> 16);
    }
    uint16_t b = (uint16_t)(h >> 16);
    if (mode == ORACLE_GEN_DENORM)
        return b & 0x81FFu;
    if ((b & 0x7F80u) == 0x7F80u)
        b &= (uint16_t)~0x0080u;
    return b;
}

void oracle_fill_f32(float *dst, uint64_t start, size_t n, uint32_t seed, uint32_t stream, int mode)
{
    uint32_t key = oracle_stream_key(seed, stream);
    uint32_t *out = (uint32_t *)dst;
#pragma omp parallel for schedule(static)
    for (size_t j = 0; j < n; ++j)
        out[j] = oracle_gen_f32_bits(key, start + j, mode);
}

void oracle_fill_bf16(uint16_t *dst, uint64_t start, size_t n, uint32_t seed, uint32_t stream, int mode)
{
    uint32_t key = oracle_stream_key(seed, stream);
#pragma omp parallel for schedule(static)
    for (size_t j = 0; j < n; ++j)
        dst[j] = oracle_gen_bf16_bits(key, start + j, mode);
}

/* ------------------------------------------------------------------------ */
/* bf16                                                                      */

float oracle_bf16_to_f32(uint16_t h)
{
    uint32_t bits = (uint32_t)h << 16;
    float f;
    memcpy(&f, &bits, 4);
    return f;
}

uint16_t oracle_f32_to_bf16_rne(float f)
{
    uint32_t bits;
    memcpy(&bits, &f, 4);
    if ((bits & 0x7FFFFFFFu) > 0x7F800000u) /* NaN: canonical, as cvt.rn.bf16.f32 gives */
        return 0x7FFFu;
    /* add 0x7FFF + lsb of the kept part, then truncate: round-to-nearest-even;
     * a carry out of the mantissa correctly bumps the exponent (up to Inf). */
    uint32_t lsb = (bits >> 16) & 1u;
    bits += 0x7FFFu + lsb;
    return (uint16_t)(bits >> 16);
}

/* ------------------------------------------------------------------------ */
/* loops                                                                     */

static int resolve_threads(int threads)
{
#ifdef _OPENMP
    if (threads == 0)
        return omp_get_max_threads();
    return threads < 1 ? 1 : threads;
#else
    (void)threads;
    return 1;
#endif
}

int oracle_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

#define PAR_FOR _Pragma("omp parallel for schedule(static) num_threads(nt) if (nt > 1)")

void oracle_add_f32(float *c, const float *a, const float *b, size_t n, int threads)
{
    int nt = resolve_threads(threads);
    (void)nt;
    PAR_FOR
    for (size_t i = 0; i < n; ++i)
        c[i] = a[i] + b[i];
}

void oracle_copy_f32(float *c, const float *a, size_t n, int threads)
{
    int nt = resolve_threads(threads);
    (void)nt;
    PAR_FOR
    for (size_t i = 0; i < n; ++i)
        c[i] = a[i];
}

void oracle_scale_f32(float *b, float s, const float *c, size_t n, int threads)
{
    int nt = resolve_threads(threads);
    (void)nt;
    PAR_FOR
    for (size_t i = 0; i < n; ++i)
        b[i] = s * c[i];
}

void oracle_triad_f32(float *a, float s, const float *c, const float *b, size_t n, int threads)
{
    int nt = resolve_threads(threads);
    (void)nt;
    PAR_FOR
    for (size_t i = 0; i < n; ++i)
        a[i] = fmaf(s, c[i], b[i]);
}

void oracle_add_bf16(uint16_t *c, const uint16_t *a, const uint16_t *b, size_t n, int threads)
{
    int nt = resolve_threads(threads);
    (void)nt;
    PAR_FOR
    for (size_t i = 0; i < n; ++i)
        c[i] = oracle_f32_to_bf16_rne(oracle_bf16_to_f32(a[i]) + oracle_bf16_to_f32(b[i]));
}

void oracle_copy_bf16(uint16_t *c, const uint16_t *a, size_t n, int threads)
{
    int nt = resolve_threads(threads);
    (void)nt;
    PAR_FOR
    for (size_t i = 0; i < n; ++i)
        c[i] = a[i];
}

void oracle_scale_bf16(uint16_t *b, float s, const uint16_t *c, size_t n, int threads)
{
    int nt = resolve_threads(threads);
    (void)nt;
    PAR_FOR
    for (size_t i = 0; i < n; ++i)
        b[i] = oracle_f32_to_bf16_rne(s * oracle_bf16_to_f32(c[i]));
}

void oracle_triad_bf16(uint16_t *a, float s, const uint16_t *c, const uint16_t *b, size_t n, int threads)
{
    int nt = resolve_threads(threads);
    (void)nt;
    PAR_FOR
    for (size_t i = 0; i < n; ++i)
        a[i] = oracle_f32_to_bf16_rne(fmaf(s, oracle_bf16_to_f32(c[i]), oracle_bf16_to_f32(b[i])));
}

/* ------------------------------------------------------------------------ */
/* checksum                                                                  */

void oracle_checksum(const void *buf, size_t nbytes, uint64_t word_offset, uint64_t out[2], int threads)
{
    int nt = resolve_threads(threads);
    (void)nt;
    const unsigned char *p = (const unsigned char *)buf;
    size_t nwords = nbytes / 4;
    uint64_t s0 = 0, s1 = 0;
#pragma omp parallel for schedule(static) num_threads(nt) if (nt > 1) reduction(+ : s0, s1)
    for (size_t j = 0; j < nwords; ++j) {
        uint32_t w;
        memcpy(&w, p + 4 * j, 4);
        s0 += w;
        s1 += (uint64_t)w * (uint64_t)((uint32_t)(word_offset + j) | 1u);
    }
    size_t rem = nbytes - 4 * nwords;
    if (rem) {
        uint32_t w = 0;
        memcpy(&w, p + 4 * nwords, rem);
        s0 += w;
        s1 += (uint64_t)w * (uint64_t)((uint32_t)(word_offset + nwords) | 1u);
    }
    out[0] = s0;
    out[1] = s1;
}

/* ------------------------------------------------------------------------ */
/* config 0: the vectorAdd PASSED/FAILED contract on the host                */

int oracle_vector_add_selftest(size_t n, uint32_t seed, long long corrupt_index)
{
    float *a = (float *)malloc(n * sizeof(float) + 4);
    float *b = (float *)malloc(n * sizeof(float) + 4);
    float *c = (float *)malloc(n * sizeof(float) + 4);
    if (!a || !b || !c) {
        free(a); free(b); free(c);
        return 2;
    }
    oracle_fill_f32(a, 0, n, seed, 0, ORACLE_GEN_UNIT);
    oracle_fill_f32(b, 0, n, seed, 1, ORACLE_GEN_UNIT);
    oracle_add_f32(c, a, b, n, 1);
    if (corrupt_index >= 0 && (size_t)corrupt_index < n) {
        uint32_t bits;
        memcpy(&bits, &c[corrupt_index], 4);
        bits ^= 1u;
        memcpy(&c[corrupt_index], &bits, 4);
    }
    int failed = 0;
    for (size_t i = 0; i < n; ++i) {
        float want = a[i] + b[i];
        if (memcmp(&want, &c[i], 4) != 0) {
            failed = 1;
            break;
        }
    }
    free(a); free(b); free(c);
    return failed;
}

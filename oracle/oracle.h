/*
 * oracle.h — CPU restatement of the GPU validation workload's arithmetic.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load or call this library; the product
 * (libb200v.so, gpu-validate) never links it and has no CPU fallback.
 *
 * PARITY UNPINNED at the reference level.  The reference
 * (mysticrenji/kubernetes-with-nvidia-gpu) is a README and a LICENSE: its
 * Step 9 pod is *named* cuda-vector-add (/root/reference/README.md:307,311)
 * but runs `nvidia-smi` (README:313-314) and computes nothing, so it holds no
 * golden vector, fixture or test for this path.  The
 * arithmetic comes from BASELINE.json's north_star: NVIDIA cuda-samples
 * vectorAdd semantics (third-party, not vendored, no version pinned by the
 * reference; published algorithm: C[i] = A[i] + B[i] in IEEE-754 binary32,
 * N = 50,000) plus STREAM copy/scale/add/triad, "bit-exact against a host C
 * reference loop, with triad using explicit fmaf on both sides".
 * What pins this oracle instead: tests/golden/kat_*.json, known-answer
 * vectors produced by exact rational arithmetic (tests/golden/gen_kat.py,
 * Python fractions — no C, no fmaf, no numpy rounding involved).
 *
 * Semantics (all loops element-wise over contiguous 1-D arrays):
 *   add   : c[i] = a[i] + b[i]            one RNE rounding
 *   copy  : c[i] = a[i]                   bit copy
 *   scale : b[i] = s * c[i]               one RNE rounding
 *   triad : a[i] = fmaf(s, c[i], b[i])    ONE rounding (fused), never mul+add
 *   bf16  : widen each operand to binary32 (exact), compute as above in
 *           binary32, round the binary32 result to bf16 with RNE
 *           (NaN -> 0x7FFF).  Denormals are preserved everywhere (no FTZ).
 * Build with -ffp-contract=off so the compiler fuses nothing on its own.
 */
#ifndef B200V_ORACLE_H
#define B200V_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- deterministic index-derived inputs (SURVEY.md §8d "Inputs") -------- */

enum {
    ORACLE_GEN_UNIT   = 0, /* fp32: 24-bit k * 2^-24 in [0,1); bf16: 8-bit k * 2^-8 */
    ORACLE_GEN_BITS   = 1, /* any finite bit pattern (sign, exponent 0..0xFE, mantissa) */
    ORACLE_GEN_DENORM = 2  /* exponent field 0..3: denormals and their neighbours */
};

uint32_t oracle_stream_key(uint32_t seed, uint32_t stream);
uint32_t oracle_hash32(uint32_t key, uint64_t index);
uint32_t oracle_gen_f32_bits(uint32_t key, uint64_t index, int mode);
uint16_t oracle_gen_bf16_bits(uint32_t key, uint64_t index, int mode);

/* dst[j] = value at global index start + j, for j in [0, n) */
void oracle_fill_f32(float *dst, uint64_t start, size_t n, uint32_t seed, uint32_t stream, int mode);
void oracle_fill_bf16(uint16_t *dst, uint64_t start, size_t n, uint32_t seed, uint32_t stream, int mode);

/* ---- bf16 <-> fp32, integer bit-ops only -------------------------------- */
float    oracle_bf16_to_f32(uint16_t h);
uint16_t oracle_f32_to_bf16_rne(float f);

/* ---- the four loops, fp32.  threads<=1: serial; else OpenMP with that many
 *      threads (0 = omp_get_max_threads()).  Serial and OpenMP agree bit for
 *      bit: each element is independent.                                    */
void oracle_add_f32  (float *c, const float *a, const float *b, size_t n, int threads);
void oracle_copy_f32 (float *c, const float *a, size_t n, int threads);
void oracle_scale_f32(float *b, float s, const float *c, size_t n, int threads);
void oracle_triad_f32(float *a, float s, const float *c, const float *b, size_t n, int threads);

/* ---- the four loops, bf16 (uint16_t bit patterns) ----------------------- */
void oracle_add_bf16  (uint16_t *c, const uint16_t *a, const uint16_t *b, size_t n, int threads);
void oracle_copy_bf16 (uint16_t *c, const uint16_t *a, size_t n, int threads);
void oracle_scale_bf16(uint16_t *b, float s, const uint16_t *c, size_t n, int threads);
void oracle_triad_bf16(uint16_t *a, float s, const uint16_t *c, const uint16_t *b, size_t n, int threads);

/* ---- order-independent position-sensitive checksum over a byte buffer ----
 * The buffer is read as little-endian 32-bit words w_j (a trailing 1..3 bytes
 * are zero-extended into one last word).  j counts words from word_offset.
 *   out[0] = sum_j w_j                      (mod 2^64)
 *   out[1] = sum_j w_j * ((uint32_t)j | 1)  (mod 2^64, 32x32->64 product)
 * Sums of shards' checksums (each with its own word_offset) equal the
 * checksum of the whole: this is what verifies 2^32-element runs in chunks. */
void oracle_checksum(const void *buf, size_t nbytes, uint64_t word_offset, uint64_t out[2], int threads);

/* ---- cuda-samples vectorAdd contract, config 0 (BASELINE.json configs[0]):
 * fills A,B (UNIT mode, streams 0 and 1), computes C=A+B with the serial
 * loop, re-checks it, returns 0 (PASSED) or 1 (FAILED).  If corrupt_index is
 * >= 0 that element of C is bit-flipped before the check (FAILED path).     */
int oracle_vector_add_selftest(size_t n, uint32_t seed, long long corrupt_index);

int oracle_max_threads(void);

#ifdef __cplusplus
}
#endif
#endif

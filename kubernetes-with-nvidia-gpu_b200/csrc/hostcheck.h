// hostcheck.h — the launcher's own host reference loops ("results must be bit-exact against a host
// C reference loop, with triad using explicit fmaf on both sides", BASELINE.json north_star) and
// the OpenMP CPU baseline it times in the same run.  This is the verification half of the
// workload, as in cuda-samples vectorAdd where the host re-computes and compares; it is never a
// substitute for the GPU path.  Built with -ffp-contract=off.
#pragma once

#include <cstddef>
#include <cstdint>

namespace hostcheck {

enum Op { ADD = 0, COPY = 1, SCALE = 2, TRIAD = 3 };
enum Dtype { F32 = 0, BF16 = 1 };
enum Gen { UNIT = 0, BITS = 1, DENORM = 2 };

uint32_t stream_key(uint32_t seed, uint32_t stream_id);
uint32_t gen_f32(uint32_t key, uint64_t index, int mode);
uint16_t gen_bf16(uint32_t key, uint64_t index, int mode);
void fill(int dtype, void* dst, uint64_t start, size_t n, uint32_t seed, uint32_t stream_id, int mode);

uint16_t f32_to_bf16(float f);
float bf16_to_f32(uint16_t h);

// dst = op(x, y, s), element-wise; `parallel` selects the OpenMP loop (same results bit for bit)
void apply(int op, int dtype, void* dst, const void* x, const void* y, float s, size_t n, bool parallel);

// expected checksum of op(gen(stream 0), gen(stream 1), s) over global indices [start, start+n),
// computed in chunks without materialising the arrays; word indices count from word_offset
void expected_checksum(int op, int dtype, uint64_t start, size_t n, uint32_t seed, int mode, float s,
                       uint64_t word_offset, uint64_t out[2]);
// checksum of the generator's own output for one stream (what a copied buffer must hold)
void generator_checksum(int dtype, uint64_t start, size_t n, uint32_t seed, uint32_t stream_id, int mode,
                        uint64_t word_offset, uint64_t out[2]);
void checksum(const void* buf, size_t nbytes, uint64_t word_offset, uint64_t out[2]);

// index of the first differing element, or n if the arrays match bit for bit
size_t first_mismatch(int dtype, const void* got, const void* want, size_t n);

int max_threads();

}  // namespace hostcheck

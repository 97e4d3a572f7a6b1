// hostcheck.cpp — see hostcheck.h.
#include "hostcheck.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#ifdef _OPENMP
#include <omp.h>
#endif

namespace hostcheck {

namespace {

inline uint32_t scramble(uint32_t v)
{
    v = (v ^ (v >> 16)) * 0x7FEB352Du;
    v = (v ^ (v >> 15)) * 0x846CA68Bu;
    return v ^ (v >> 16);
}

inline uint32_t index_hash(uint32_t key, uint64_t index)
{
    const uint32_t low = static_cast<uint32_t>(index);
    const uint32_t high = static_cast<uint32_t>(index >> 32);
    return scramble(low ^ key ^ (high * 0x9E3779B9u));
}

inline uint32_t bits_of(float f)
{
    uint32_t u;
    std::memcpy(&u, &f, sizeof u);
    return u;
}

inline float float_of(uint32_t u)
{
    float f;
    std::memcpy(&f, &u, sizeof f);
    return f;
}

// The loops are compiled twice: a baseline x86-64 clone and an AVX2+FMA clone picked at load time.
// With FMA in the ISA std::fmaf becomes one vfmadd instead of a libm call per element; both are
// correctly rounded, so the clones agree bit for bit (tests/test_launcher_cpu.py cross-checks).
#if defined(__x86_64__) && defined(__GNUC__)
#define HOSTCHECK_CLONES __attribute__((target_clones("default", "arch=haswell")))
#else
#define HOSTCHECK_CLONES
#endif

template <int OP> inline __attribute__((always_inline)) float op1(float x, float y, float s)
{
    if (OP == ADD) return x + y;
    if (OP == COPY) return x;
    if (OP == SCALE) return s * x;
    return std::fmaf(s, x, y);  // correctly rounded whether or not the CPU has FMA
}

template <int OP> HOSTCHECK_CLONES void loop_f32(float* d, const float* x, const float* y, float s, size_t n, bool par)
{
#pragma omp parallel for schedule(static) if (par)
    for (size_t i = 0; i < n; ++i) d[i] = op1<OP>(x[i], OP == ADD || OP == TRIAD ? y[i] : 0.0f, s);
}

template <int OP> HOSTCHECK_CLONES void loop_bf16(uint16_t* d, const uint16_t* x, const uint16_t* y, float s, size_t n, bool par)
{
#pragma omp parallel for schedule(static) if (par)
    for (size_t i = 0; i < n; ++i) {
        if (OP == COPY) {
            d[i] = x[i];
        } else {
            const float yy = (OP == ADD || OP == TRIAD) ? bf16_to_f32(y[i]) : 0.0f;
            d[i] = f32_to_bf16(op1<OP>(bf16_to_f32(x[i]), yy, s));
        }
    }
}

struct Sum {
    uint64_t plain = 0, weighted = 0;
    void word(uint32_t w, uint64_t j)
    {
        plain += w;
        weighted += static_cast<uint64_t>(w) * static_cast<uint64_t>(static_cast<uint32_t>(j) | 1u);
    }
};

}  // namespace

uint32_t stream_key(uint32_t seed, uint32_t stream_id) { return scramble(seed ^ (0x9E3779B9u * (stream_id + 1u))); }

uint32_t gen_f32(uint32_t key, uint64_t index, int mode)
{
    uint32_t h = index_hash(key, index);
    switch (mode) {
    case UNIT: return bits_of(static_cast<float>(h >> 8) * 0x1p-24f);
    case DENORM: return h & 0x81FFFFFFu;
    default:
        if ((h & 0x7F800000u) == 0x7F800000u) h &= ~0x00800000u;
        return h;
    }
}

uint16_t gen_bf16(uint32_t key, uint64_t index, int mode)
{
    const uint32_t h = index_hash(key, index);
    if (mode == UNIT) return static_cast<uint16_t>(bits_of(static_cast<float>(h >> 24) * 0x1p-8f) >> 16);
    uint16_t b = static_cast<uint16_t>(h >> 16);
    if (mode == DENORM) return b & 0x81FFu;
    if ((b & 0x7F80u) == 0x7F80u) b = static_cast<uint16_t>(b & ~0x0080u);
    return b;
}

void fill(int dtype, void* dst, uint64_t start, size_t n, uint32_t seed, uint32_t stream_id, int mode)
{
    const uint32_t key = stream_key(seed, stream_id);
    if (dtype == F32) {
        uint32_t* out = static_cast<uint32_t*>(dst);
#pragma omp parallel for schedule(static)
        for (size_t i = 0; i < n; ++i) out[i] = gen_f32(key, start + i, mode);
    } else {
        uint16_t* out = static_cast<uint16_t*>(dst);
#pragma omp parallel for schedule(static)
        for (size_t i = 0; i < n; ++i) out[i] = gen_bf16(key, start + i, mode);
    }
}

uint16_t f32_to_bf16(float f)
{
    uint32_t u = bits_of(f);
    if ((u << 1) > 0xFF000000u) return 0x7FFFu;  // NaN, as cvt.rn.bf16.f32 produces
    u += 0x7FFFu + ((u >> 16) & 1u);             // round to nearest, ties to even
    return static_cast<uint16_t>(u >> 16);
}

float bf16_to_f32(uint16_t h) { return float_of(static_cast<uint32_t>(h) << 16); }

void apply(int op, int dtype, void* dst, const void* x, const void* y, float s, size_t n, bool parallel)
{
    if (dtype == F32) {
        float* d = static_cast<float*>(dst);
        const float* a = static_cast<const float*>(x);
        const float* b = static_cast<const float*>(y);
        switch (op) {
        case ADD: loop_f32<ADD>(d, a, b, s, n, parallel); break;
        case COPY: loop_f32<COPY>(d, a, b, s, n, parallel); break;
        case SCALE: loop_f32<SCALE>(d, a, b, s, n, parallel); break;
        default: loop_f32<TRIAD>(d, a, b, s, n, parallel); break;
        }
    } else {
        uint16_t* d = static_cast<uint16_t*>(dst);
        const uint16_t* a = static_cast<const uint16_t*>(x);
        const uint16_t* b = static_cast<const uint16_t*>(y);
        switch (op) {
        case ADD: loop_bf16<ADD>(d, a, b, s, n, parallel); break;
        case COPY: loop_bf16<COPY>(d, a, b, s, n, parallel); break;
        case SCALE: loop_bf16<SCALE>(d, a, b, s, n, parallel); break;
        default: loop_bf16<TRIAD>(d, a, b, s, n, parallel); break;
        }
    }
}

void checksum(const void* buf, size_t nbytes, uint64_t word_offset, uint64_t out[2])
{
    const unsigned char* p = static_cast<const unsigned char*>(buf);
    const size_t whole = nbytes / 4;
    uint64_t plain = 0, weighted = 0;
#pragma omp parallel for schedule(static) reduction(+ : plain, weighted)
    for (size_t j = 0; j < whole; ++j) {
        uint32_t w;
        std::memcpy(&w, p + 4 * j, 4);
        plain += w;
        weighted += static_cast<uint64_t>(w) * static_cast<uint64_t>(static_cast<uint32_t>(word_offset + j) | 1u);
    }
    Sum tail;
    if (nbytes % 4) {
        uint32_t w = 0;
        std::memcpy(&w, p + 4 * whole, nbytes % 4);
        tail.word(w, word_offset + whole);
    }
    out[0] = plain + tail.plain;
    out[1] = weighted + tail.weighted;
}

namespace {

// chunked: generate inputs, apply, checksum; chunk boundaries fall on even element counts so bf16
// pairs never straddle two chunks
template <typename Produce>
void chunked_checksum(int dtype, size_t n, uint64_t word_offset, uint64_t out[2], Produce produce)
{
    const size_t esz = dtype == F32 ? 4 : 2;
    const size_t per_word = 4 / esz;
    const size_t chunk = size_t(1) << 22;
    std::vector<unsigned char> buf(chunk * esz);
    uint64_t plain = 0, weighted = 0;
    for (size_t off = 0; off < n; off += chunk) {
        const size_t m = std::min(chunk, n - off);
        produce(off, m, buf.data());
        uint64_t c[2];
        checksum(buf.data(), m * esz, word_offset + off / per_word, c);
        plain += c[0];
        weighted += c[1];
    }
    out[0] = plain;
    out[1] = weighted;
}

}  // namespace

void expected_checksum(int op, int dtype, uint64_t start, size_t n, uint32_t seed, int mode, float s,
                       uint64_t word_offset, uint64_t out[2])
{
    const size_t esz = dtype == F32 ? 4 : 2;
    const size_t chunk = size_t(1) << 22;
    std::vector<unsigned char> x(chunk * esz), y(chunk * esz);
    chunked_checksum(dtype, n, word_offset, out, [&](size_t off, size_t m, unsigned char* dst) {
        fill(dtype, x.data(), start + off, m, seed, 0, mode);
        if (op == ADD || op == TRIAD) fill(dtype, y.data(), start + off, m, seed, 1, mode);
        apply(op, dtype, dst, x.data(), y.data(), s, m, true);
    });
}

void generator_checksum(int dtype, uint64_t start, size_t n, uint32_t seed, uint32_t stream_id, int mode,
                        uint64_t word_offset, uint64_t out[2])
{
    chunked_checksum(dtype, n, word_offset, out, [&](size_t off, size_t m, unsigned char* dst) {
        fill(dtype, dst, start + off, m, seed, stream_id, mode);
    });
}

size_t first_mismatch(int dtype, const void* got, const void* want, size_t n)
{
    const size_t esz = dtype == F32 ? 4 : 2;
    if (std::memcmp(got, want, n * esz) == 0) return n;
    const unsigned char* g = static_cast<const unsigned char*>(got);
    const unsigned char* w = static_cast<const unsigned char*>(want);
    for (size_t i = 0; i < n; ++i)
        if (std::memcmp(g + i * esz, w + i * esz, esz) != 0) return i;
    return n;
}

int max_threads()
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

}  // namespace hostcheck

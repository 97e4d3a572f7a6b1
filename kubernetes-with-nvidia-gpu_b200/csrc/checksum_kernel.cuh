// checksum_kernel.cuh — the checksum reduction (see include/b200v.h: b200v_checksum).
//   out[0] = sum w_j, out[1] = sum w_j * ((uint32_t)(word_offset + j) | 1)  (mod 2^64)
//
// Scratch layout (caller-owned workspace, or the library's static array): a uint32 arrival counter
// (zeroed by the entry point before every launch), then one {s0, s1} uint64 pair per CTA.  Every CTA
// reduces its share, publishes its pair, and the last CTA to arrive adds the pairs up.
//
// UNROLL = 128-bit loads in flight per thread.  TILED = each CTA iteration reads one contiguous tile
// of BLOCK x UNROLL vectors (tiles walked grid-stride) instead of UNROLL separate grid-strided
// streams; tools/tune.cu `cksum` measures both.
#pragma once

#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

namespace b200v {

constexpr int kChecksumBlock = 256;
constexpr int kChecksumMaxBlocks = 2048;

__device__ __forceinline__ uint4 ld_stream_u4(const uint4* p)
{
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.b32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p));
    return v;
}

__device__ __forceinline__ void acc_word(uint64_t& s0, uint64_t& s1, uint32_t w, uint64_t j)
{
    s0 += w;
    s1 += (uint64_t)w * (uint64_t)((uint32_t)j | 1u);
}

template <int UNROLL, bool TILED>
__global__ void __launch_bounds__(kChecksumBlock)
checksum_kernel(const unsigned char* __restrict__ buf, size_t nbytes, uint64_t word_offset,
                unsigned int* __restrict__ counter, uint64_t* __restrict__ partials, uint64_t* __restrict__ out)
{
    const size_t nwords = nbytes / 4;
    const uint32_t* words = reinterpret_cast<const uint32_t*>(buf);
    size_t head = ((16 - ((uintptr_t)buf & 15)) & 15) / 4;   // words before 16-byte alignment
    if (head > nwords) head = nwords;
    const size_t nvec = (nwords - head) / 4;
    const uint4* vec = reinterpret_cast<const uint4*>(words + head);
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;

    uint64_t s0 = 0, s1 = 0;
    size_t v = tid;
    if constexpr (TILED) {
        constexpr size_t TILE = (size_t)kChecksumBlock * UNROLL;
        const size_t tiles = nvec / TILE;
        for (size_t t = blockIdx.x; t < tiles; t += gridDim.x) {
            const size_t base = t * TILE + threadIdx.x;
            uint4 q[UNROLL];
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) q[u] = ld_stream_u4(vec + base + (size_t)u * kChecksumBlock);
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                const uint64_t j = word_offset + head + (base + (size_t)u * kChecksumBlock) * 4;
                acc_word(s0, s1, q[u].x, j + 0);
                acc_word(s0, s1, q[u].y, j + 1);
                acc_word(s0, s1, q[u].z, j + 2);
                acc_word(s0, s1, q[u].w, j + 3);
            }
        }
        v = tiles * TILE + tid;       // the < 1 tile remainder goes through the plain loop below
    } else {
        for (; v + (UNROLL - 1) * stride < nvec; v += UNROLL * stride) {
            uint4 q[UNROLL];
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) q[u] = ld_stream_u4(vec + v + u * stride);
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                const uint64_t j = word_offset + head + (v + u * stride) * 4;
                acc_word(s0, s1, q[u].x, j + 0);
                acc_word(s0, s1, q[u].y, j + 1);
                acc_word(s0, s1, q[u].z, j + 2);
                acc_word(s0, s1, q[u].w, j + 3);
            }
        }
    }
    for (; v < nvec; v += stride) {
        uint4 q = ld_stream_u4(vec + v);
        const uint64_t j = word_offset + head + v * 4;
        acc_word(s0, s1, q.x, j + 0);
        acc_word(s0, s1, q.y, j + 1);
        acc_word(s0, s1, q.z, j + 2);
        acc_word(s0, s1, q.w, j + 3);
    }
    // ragged words: [0, head) and [head + 4*nvec, nwords), then the 1..3 trailing bytes
    const size_t tail0 = head + nvec * 4;
    const size_t ragged = head + (nwords - tail0);
    for (size_t r = tid; r < ragged; r += stride) {
        const size_t j = r < head ? r : tail0 + (r - head);
        acc_word(s0, s1, words[j], word_offset + j);
    }
    if (tid == 0 && (nbytes & 3)) {
        uint32_t w = 0;
        for (size_t k = 0; k < (nbytes & 3); ++k) w |= (uint32_t)buf[nwords * 4 + k] << (8 * k);
        acc_word(s0, s1, w, word_offset + nwords);
    }

    // CTA reduction
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        s0 += __shfl_down_sync(0xFFFFFFFFu, s0, d);
        s1 += __shfl_down_sync(0xFFFFFFFFu, s1, d);
    }
    __shared__ uint64_t sh0[kChecksumBlock / 32], sh1[kChecksumBlock / 32];
    __shared__ bool last;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) { sh0[warp] = s0; sh1[warp] = s1; }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint64_t t0 = 0, t1 = 0;
#pragma unroll
        for (int w = 0; w < kChecksumBlock / 32; ++w) { t0 += sh0[w]; t1 += sh1[w]; }
        partials[2 * blockIdx.x + 0] = t0;
        partials[2 * blockIdx.x + 1] = t1;
        __threadfence();
        last = (atomicAdd(counter, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    uint64_t t0 = 0, t1 = 0;
    for (unsigned b = threadIdx.x; b < gridDim.x; b += kChecksumBlock) {
        t0 += __ldcg(&partials[2 * b + 0]);
        t1 += __ldcg(&partials[2 * b + 1]);
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        t0 += __shfl_down_sync(0xFFFFFFFFu, t0, d);
        t1 += __shfl_down_sync(0xFFFFFFFFu, t1, d);
    }
    __syncthreads();
    if (lane == 0) { sh0[warp] = t0; sh1[warp] = t1; }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint64_t r0 = 0, r1 = 0;
#pragma unroll
        for (int w = 0; w < kChecksumBlock / 32; ++w) { r0 += sh0[w]; r1 += sh1[w]; }
        out[0] = r0;
        out[1] = r1;
        *counter = 0;
    }
}


}  // namespace b200v

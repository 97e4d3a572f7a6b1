// stream_tma.cuh — the same streaming ops staged through shared memory with the bulk async-copy
// engine (TMA 1-D: cp.async.bulk + mbarrier; SASS: UBLKCP), as an alternative to the register
// path in stream_kernels.cuh.  Persistent CTAs; each owns a ring of STAGES shared-memory slots.
// One elected thread drives the ring:
//
//   load(tile t+STAGES-1) -> [mbarrier full]  -> all threads: LDS, op, STS in place
//        -> fence.proxy.async + __syncthreads -> bulk store(tile t) -> commit_group
//   the slot of tile t-1 is refilled only after cp.async.bulk.wait_group.read says its store has
//   drained the slot.
//
// Bytes in flight per SM are set by STAGES x TILE_BYTES, not by registers.  Whether that beats
// 256-bit LDG/STG is an empirical question: tools/tune.cu measures both (profiles/).
#pragma once

#include "stream_kernels.cuh"

namespace b200v {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE;\n"
        "bra WAIT_LOOP;\n"
        "DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void bulk_store(void* gmem_dst, const void* smem_src, uint32_t bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst), "r"(smem_u32(smem_src)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait_read()
{
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// n counts ELEMENTS; only whole tiles are handled here (the host sends the remainder to the
// register kernel).  Dynamic shared memory: STAGES x (1 or 2) x TILE_BYTES, 128-byte aligned.
template <int OP, bool BF16, int TILE_BYTES, int STAGES, int BLOCK>
__global__ void __launch_bounds__(BLOCK)
stream_kernel_tma(void* __restrict__ dst, const void* __restrict__ x, const void* __restrict__ y, float s, size_t tiles)
{
    constexpr bool TWO = kReadsY<OP>;
    constexpr int SLOT_BYTES = TILE_BYTES * (TWO ? 2 : 1);
    constexpr int VECS = TILE_BYTES / 16;            // 16-byte vectors per array per tile
    static_assert(VECS % BLOCK == 0, "tile must be a whole number of 16-byte vectors per thread");
    constexpr int PER_THREAD = VECS / BLOCK;

    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ uint64_t full[STAGES];

    const char* gx = static_cast<const char*>(x);
    const char* gy = static_cast<const char*>(y);
    char* gd = static_cast<char*>(dst);

    if (threadIdx.x == 0) {
        for (int i = 0; i < STAGES; ++i) mbar_init(&full[i], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    // number of tiles this CTA owns: blockIdx.x, blockIdx.x + grid, ...
    const size_t mine = tiles > blockIdx.x ? (tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;

    auto issue_load = [&](size_t k) {   // k-th tile of this CTA into slot k % STAGES
        const int slot = (int)(k % STAGES);
        const size_t off = (blockIdx.x + k * gridDim.x) * (size_t)TILE_BYTES;
        unsigned char* sx = smem + slot * SLOT_BYTES;
        mbar_expect_tx(&full[slot], SLOT_BYTES);
        bulk_load(sx, gx + off, TILE_BYTES, &full[slot]);
        if constexpr (TWO) bulk_load(sx + TILE_BYTES, gy + off, TILE_BYTES, &full[slot]);
    };

    if (threadIdx.x == 0)
        for (size_t k = 0; k < (size_t)(STAGES - 1) && k < mine; ++k) issue_load(k);

    for (size_t k = 0; k < mine; ++k) {
        const int slot = (int)(k % STAGES);
        const uint32_t parity = (uint32_t)((k / STAGES) & 1);
        uint4* sx = reinterpret_cast<uint4*>(smem + slot * SLOT_BYTES);
        const uint4* sy = reinterpret_cast<const uint4*>(smem + slot * SLOT_BYTES + TILE_BYTES);
        mbar_wait(&full[slot], parity);
#pragma unroll
        for (int u = 0; u < PER_THREAD; ++u) {
            const int v = u * BLOCK + threadIdx.x;
            uint4 a = sx[v];
            uint4 b = TWO ? sy[v] : make_uint4(0, 0, 0, 0);
            uint4 r;
            r.x = op_word<OP, BF16>(a.x, b.x, s);
            r.y = op_word<OP, BF16>(a.y, b.y, s);
            r.z = op_word<OP, BF16>(a.z, b.z, s);
            r.w = op_word<OP, BF16>(a.w, b.w, s);
            sx[v] = r;
        }
        fence_proxy_async();          // generic-proxy STS must be visible to the bulk store
        __syncthreads();
        if (threadIdx.x == 0) {
            const size_t off = (blockIdx.x + k * gridDim.x) * (size_t)TILE_BYTES;
            bulk_store(gd + off, sx, TILE_BYTES);
            bulk_commit();
            // refill the slot that tile k-1 used: its store (the group before this one) must have
            // finished reading shared memory
            if (k + STAGES - 1 < mine) {
                bulk_wait_read<1>();
                issue_load(k + STAGES - 1);
            }
        }
    }
    if (threadIdx.x == 0) bulk_wait_read<0>();
}

}  // namespace b200v

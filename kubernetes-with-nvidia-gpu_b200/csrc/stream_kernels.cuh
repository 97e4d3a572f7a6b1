// stream_kernels.cuh — HBM streaming kernels for sm_100a: vectorAdd / STREAM add, copy, scale, triad
// in fp32 and bf16.  These take the place of the `nvidia-smi` payload of the reference's Step 9 pod
// (/root/reference/README.md:313-314) with the cuda-samples vectorAdd semantics BASELINE.json names.
//
// Bit-exactness rules: explicit __fadd_rn/__fmul_rn/__fmaf_rn only, so the
// compiler never contracts or reorders; build with -fmad=false and without -use_fast_math (no FTZ);
// bf16 = widen (exact) -> fp32 op -> cvt.rn.bf16x2.f32 (RNE, NaN -> 0x7FFF).
//
// The kernels are HBM-bound (<= 1/6 flop per byte): no tensor cores, no shared-memory reuse.  The
// levers are bytes in flight per SM (vector width x unroll x resident warps), full-line coalescing,
// and cache hints that keep a read-once / write-once stream from being retained.
#pragma once

#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

namespace b200v {

enum Op : int { OP_ADD = 0, OP_COPY = 1, OP_SCALE = 2, OP_TRIAD = 3 };

// Load / store cache hints (SASS shown is what ptxas emits for sm_100a).
enum LdHint : int {
    LD_DEFAULT = 0,   // LDG.E.ENL2
    LD_NC_NA = 1,     // ld.global.nc.L1::no_allocate            -> LDG.E.NA.ENL2.CONSTANT
    LD_NC_NA_EF = 2,  // ... + L2::evict_first                   -> LDG.E.NA.EFL2.CONSTANT
    LD_CS = 3,        // ld.global.cs (evict-first in L1 and L2)  -> LDG.E.EF
    LD_NA = 4         // ld.global.L1::no_allocate (coherent path: ptxas keeps it ahead of the stores)
};
enum StHint : int {
    ST_DEFAULT = 0,   // STG.E.ENL2
    ST_CS = 1,        // st.global.cs                             -> STG.E.EF
    ST_NA_EF = 2      // st.global.L1::no_allocate.L2::evict_first -> STG.E.NA.EFL2
};

// ---------------------------------------------------------------------------------------------
// 128-bit and 256-bit global access.  256-bit (`.v8.b32`, LDG.E.256 / STG.E.256) exists only on
// sm_100+: one instruction moves 1 KiB per warp.

template <int BYTES> struct Vec;
template <> struct __align__(16) Vec<16> { uint32_t w[4]; };
template <> struct __align__(32) Vec<32> { uint32_t w[8]; };

#define B200V_LD4(STR)                                                                             \
    asm volatile(STR " {%0,%1,%2,%3}, [%4];"                                                       \
                 : "=r"(v.w[0]), "=r"(v.w[1]), "=r"(v.w[2]), "=r"(v.w[3])                          \
                 : "l"(p))
#define B200V_LD8(STR)                                                                             \
    asm volatile(STR " {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"                                           \
                 : "=r"(v.w[0]), "=r"(v.w[1]), "=r"(v.w[2]), "=r"(v.w[3]), "=r"(v.w[4]),           \
                   "=r"(v.w[5]), "=r"(v.w[6]), "=r"(v.w[7])                                        \
                 : "l"(p))
#define B200V_ST4(STR)                                                                             \
    asm volatile(STR " [%4], {%0,%1,%2,%3};" ::"r"(v.w[0]), "r"(v.w[1]), "r"(v.w[2]), "r"(v.w[3]), \
                 "l"(p)                                                                            \
                 : "memory")
#define B200V_ST8(STR)                                                                             \
    asm volatile(STR " [%8], {%0,%1,%2,%3,%4,%5,%6,%7};" ::"r"(v.w[0]), "r"(v.w[1]), "r"(v.w[2]),  \
                 "r"(v.w[3]), "r"(v.w[4]), "r"(v.w[5]), "r"(v.w[6]), "r"(v.w[7]), "l"(p)           \
                 : "memory")

template <int H> __device__ __forceinline__ Vec<16> ldg(const Vec<16>* p)
{
    Vec<16> v;
    if constexpr (H == LD_DEFAULT) B200V_LD4("ld.global.v4.b32");
    else if constexpr (H == LD_NC_NA) B200V_LD4("ld.global.nc.L1::no_allocate.v4.b32");
    else if constexpr (H == LD_NC_NA_EF) B200V_LD4("ld.global.nc.L1::no_allocate.v4.b32");  // L2 hints: 256-bit only
    else if constexpr (H == LD_NA) B200V_LD4("ld.global.L1::no_allocate.v4.b32");
    else B200V_LD4("ld.global.cs.v4.b32");
    return v;
}
template <int H> __device__ __forceinline__ Vec<32> ldg(const Vec<32>* p)
{
    Vec<32> v;
    if constexpr (H == LD_DEFAULT) B200V_LD8("ld.global.v8.b32");
    else if constexpr (H == LD_NC_NA) B200V_LD8("ld.global.nc.L1::no_allocate.v8.b32");
    else if constexpr (H == LD_NC_NA_EF) B200V_LD8("ld.global.nc.L1::no_allocate.L2::evict_first.v8.b32");
    else if constexpr (H == LD_NA) B200V_LD8("ld.global.L1::no_allocate.v8.b32");
    else B200V_LD8("ld.global.cs.v8.b32");
    return v;
}
template <int H> __device__ __forceinline__ void stg(Vec<16>* p, const Vec<16>& v)
{
    if constexpr (H == ST_DEFAULT) B200V_ST4("st.global.v4.b32");
    else if constexpr (H == ST_CS) B200V_ST4("st.global.cs.v4.b32");
    else B200V_ST4("st.global.cs.v4.b32");  // ptxas takes .L2::evict_first on st only at 256 bits
}
template <int H> __device__ __forceinline__ void stg(Vec<32>* p, const Vec<32>& v)
{
    if constexpr (H == ST_DEFAULT) B200V_ST8("st.global.v8.b32");
    else if constexpr (H == ST_CS) B200V_ST8("st.global.cs.v8.b32");
    else B200V_ST8("st.global.L1::no_allocate.L2::evict_first.v8.b32");
}

// ---------------------------------------------------------------------------------------------
// Arithmetic on one 32-bit word: one fp32 element, or two bf16 elements (low half = lower index).
//   add: x + y     copy: x     scale: s * x     triad: fma(s, x, y)

template <int OP> __device__ __forceinline__ float op_f32(float x, float y, float s)
{
    if constexpr (OP == OP_ADD) return __fadd_rn(x, y);
    else if constexpr (OP == OP_COPY) return x;
    else if constexpr (OP == OP_SCALE) return __fmul_rn(s, x);
    else return __fmaf_rn(s, x, y);
}

__device__ __forceinline__ uint32_t pack_bf16x2_rn(float lo, float hi)
{
    uint32_t r;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}

template <int OP, bool BF16> __device__ __forceinline__ uint32_t op_word(uint32_t x, uint32_t y, float s)
{
    if constexpr (OP == OP_COPY) {
        return x;
    } else if constexpr (!BF16) {
        return __float_as_uint(op_f32<OP>(__uint_as_float(x), __uint_as_float(y), s));
    } else {
        float lo = op_f32<OP>(__uint_as_float(x << 16), __uint_as_float(y << 16), s);
        float hi = op_f32<OP>(__uint_as_float(x & 0xFFFF0000u), __uint_as_float(y & 0xFFFF0000u), s);
        return pack_bf16x2_rn(lo, hi);
    }
}

template <int OP> __device__ __forceinline__ uint16_t op_bf16_scalar(uint16_t x, uint16_t y, float s)
{
    if constexpr (OP == OP_COPY) return x;
    float r = op_f32<OP>(__uint_as_float((uint32_t)x << 16), __uint_as_float((uint32_t)y << 16), s);
    return (uint16_t)(pack_bf16x2_rn(r, 0.0f) & 0xFFFFu);
}

template <int OP> inline constexpr bool kReadsY = (OP == OP_ADD || OP == OP_TRIAD);

// ---------------------------------------------------------------------------------------------
// The streaming kernel.  A tile is BLOCK x UNROLL vectors per array; inside a tile, unroll step u
// of thread t touches vector u*BLOCK + t, so every warp instruction covers 32 x VECB contiguous
// bytes.  All loads of a tile are issued before the first use: UNROLL x (1 or 2) x VECB bytes in
// flight per thread.  CTAs walk tiles grid-stride (grid == tiles gives one tile per CTA; grid ==
// k x SMs gives persistent CTAs).  The < 1 tile remainder, and a trailing odd bf16 element, are
// done element-wise by the CTA that would have taken the next tile.
//
// n counts ELEMENTS.  dst/x/y must be VECB-aligned (the host dispatcher guarantees it).

//
// PDL = true adds the programmatic-dependent-launch handshake (griddepcontrol): the kernel lets
// its stream successor become resident as soon as this grid's last CTAs are placed, and itself
// touches no memory until its predecessor has completed and flushed.  Results are unchanged; what
// it removes is the drain/launch bubble between back-to-back launches.

template <int OP, bool BF16, int VECB, int UNROLL, int BLOCK, int LDH, int STH, bool PDL = false>
__global__ void __launch_bounds__(BLOCK)
stream_kernel(void* __restrict__ dst, const void* __restrict__ x, const void* __restrict__ y, float s, size_t n)
{
    if constexpr (PDL) {
        asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
        asm volatile("griddepcontrol.wait;" ::: "memory");
    }
    using V = Vec<VECB>;
    constexpr int WORDS = VECB / 4;
    constexpr size_t EPW = BF16 ? 2 : 1;                         // elements per 32-bit word
    constexpr size_t TILE_VECS = (size_t)BLOCK * UNROLL;
    constexpr size_t TILE_ELEMS = TILE_VECS * WORDS * EPW;

    const size_t full_tiles = n / TILE_ELEMS;
    V* __restrict__ vd = reinterpret_cast<V*>(dst);
    const V* __restrict__ vx = reinterpret_cast<const V*>(x);
    const V* __restrict__ vy = reinterpret_cast<const V*>(y);

    // (Left to the compiler's default unrolling on purpose.  `#pragma unroll 1` removes the 64-bit
    // trip-count division nvcc puts in front of the loop, but ptxas then allocates ~58 instead of 80
    // registers, sinks loads between the stores again, and the kernel loses 1.5 %:
    // profiles/r01_tune_negative_results.md.)
    for (size_t tile = blockIdx.x; tile < full_tiles; tile += gridDim.x) {
        const size_t base = tile * TILE_VECS + threadIdx.x;
        V a[UNROLL];
        V b[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) a[u] = ldg<LDH>(vx + base + (size_t)u * BLOCK);
        if constexpr (kReadsY<OP>) {
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) b[u] = ldg<LDH>(vy + base + (size_t)u * BLOCK);
        }
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            V r;
#pragma unroll
            for (int k = 0; k < WORDS; ++k)
                r.w[k] = op_word<OP, BF16>(a[u].w[k], kReadsY<OP> ? b[u].w[k] : 0u, s);
            stg<STH>(vd + base + (size_t)u * BLOCK, r);
        }
    }

    // The remainder belongs to the CTA that would take the next tile.  With one tile per CTA that is
    // CTA 0 (or the only CTA when there is no full tile); the 64-bit modulo is only evaluated for a
    // clamped or persistent grid, and nothing at all runs here when n is a whole number of tiles —
    // a CTA that lives for a few microseconds should not end with a ~100-cycle division.
    const size_t done = full_tiles * TILE_ELEMS;
    if (done == n) return;
    const size_t owner = gridDim.x >= full_tiles ? (gridDim.x == full_tiles ? 0 : full_tiles) : full_tiles % gridDim.x;
    if (blockIdx.x == owner) {
        if constexpr (!BF16) {
            float* d = reinterpret_cast<float*>(dst);
            const float* fx = reinterpret_cast<const float*>(x);
            const float* fy = reinterpret_cast<const float*>(y);
            for (size_t i = done + threadIdx.x; i < n; i += BLOCK)
                d[i] = op_f32<OP>(fx[i], kReadsY<OP> ? fy[i] : 0.0f, s);
        } else {
            uint16_t* d = reinterpret_cast<uint16_t*>(dst);
            const uint16_t* hx = reinterpret_cast<const uint16_t*>(x);
            const uint16_t* hy = reinterpret_cast<const uint16_t*>(y);
            for (size_t i = done + threadIdx.x; i < n; i += BLOCK)
                d[i] = op_bf16_scalar<OP>(hx[i], kReadsY<OP> ? hy[i] : (uint16_t)0, s);
        }
    }
}

// Element-wise kernel for pointers that are not even 16-byte aligned (sub-views handed in by a
// caller).  Correct for any element-aligned pointer; not a performance path.
template <int OP, bool BF16>
__global__ void __launch_bounds__(256)
stream_kernel_unaligned(void* __restrict__ dst, const void* __restrict__ x, const void* __restrict__ y, float s, size_t n)
{
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        if constexpr (!BF16) {
            reinterpret_cast<float*>(dst)[i] =
                op_f32<OP>(reinterpret_cast<const float*>(x)[i],
                           kReadsY<OP> ? reinterpret_cast<const float*>(y)[i] : 0.0f, s);
        } else {
            reinterpret_cast<uint16_t*>(dst)[i] =
                op_bf16_scalar<OP>(reinterpret_cast<const uint16_t*>(x)[i],
                                   kReadsY<OP> ? reinterpret_cast<const uint16_t*>(y)[i] : (uint16_t)0, s);
        }
    }
}

}  // namespace b200v

// kernels.cu — device code of libb200v and the C-ABI entry points that launch it:
// the streaming kernels (add/copy/scale/triad, fp32 + bf16), the index-hash fill, the
// checksum reduction, and the peer pull / peer push copies.  See include/b200v.h for the contract and the
// reference lines each group stands in for.
#include <cstdint>
#include <cstring>

#include "checksum_kernel.cuh"
#include "internal.h"
#include "stream_kernels.cuh"
#include "tuning.h"

namespace b200v {

// ------------------------------------------------------------------------------------------
// device properties the launch shapes need, cached per device

struct DevCache {
    bool known = false;
    int sm_count = 0;
};
static DevCache g_dev[64];

static int current_sm_count(int* sms)
{
    int dev = 0;
    B200V_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return set_error(B200V_ERR_INVALID, "device index %d out of range", dev);
    if (!g_dev[dev].known) {
        int n = 0, major = 0;
        B200V_CUDA(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
        B200V_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
        if (major != 10)
            return set_error(B200V_ERR_ARCH, "device %d is compute capability %d.x; libb200v carries sm_100a code only", dev, major);
        g_dev[dev].sm_count = n;
        g_dev[dev].known = true;
    }
    *sms = g_dev[dev].sm_count;
    return B200V_OK;
}

static bool overlaps(const void* a, size_t abytes, const void* b, size_t bbytes)
{
    uintptr_t a0 = (uintptr_t)a, b0 = (uintptr_t)b;
    return a0 < b0 + bbytes && b0 < a0 + abytes;
}

// ------------------------------------------------------------------------------------------
// streaming ops

template <int OP, bool BF16, int VECB, int UNROLL, int BLOCK>
static int launch_vec(void* dst, const void* x, const void* y, float s, size_t n, int sms, cudaStream_t st)
{
    constexpr size_t EPW = BF16 ? 2 : 1;
    constexpr size_t TILE_ELEMS = (size_t)BLOCK * UNROLL * (VECB / 4) * EPW;
    size_t tiles = n / TILE_ELEMS;
    size_t grid = tiles ? tiles : 1;
    if (Tuning::kCtasPerSm > 0) {
        size_t cap = (size_t)sms * Tuning::kCtasPerSm;
        if (grid > cap) grid = cap;
    }
    if (grid > 0x7FFFFFFFull) grid = 0x7FFFFFFFull;
    // Programmatic dependent launch: back-to-back launches on one stream overlap the drain of one
    // grid with the placement of the next (the kernel waits at griddepcontrol.wait before it touches
    // memory, so stream order of the data is unchanged).  profiles/r01_tune_pdl.csv: +0.6 % at 2^28
    // elements, +6..12 % at 2^24, +30 % at 2^22.
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(BLOCK);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    B200V_CUDA(cudaLaunchKernelEx(&cfg, stream_kernel<OP, BF16, VECB, UNROLL, BLOCK, Tuning::kLd, Tuning::kSt, true>, dst, x,
                                  y, s, n));
    return B200V_OK;
}

template <int OP, bool BF16>
static int launch_stream(void* dst, const void* x, const void* y, float s, size_t n, cudaStream_t st)
{
    constexpr size_t ESZ = BF16 ? 2 : 4;
    constexpr bool TWO = kReadsY<OP>;
    B200V_REQUIRE(dst && x && (!TWO || y), B200V_ERR_INVALID, "NULL array pointer");
    if (n == 0) return B200V_OK;
    uintptr_t bits = (uintptr_t)dst | (uintptr_t)x | (TWO ? (uintptr_t)y : 0);
    B200V_REQUIRE(bits % ESZ == 0, B200V_ERR_ALIGN, "array pointer not aligned to its %zu-byte element", ESZ);
    B200V_REQUIRE(!overlaps(dst, n * ESZ, x, n * ESZ) && (!TWO || !overlaps(dst, n * ESZ, y, n * ESZ)),
                  B200V_ERR_INVALID, "output array overlaps an input array (STREAM arrays are distinct)");
    int sms = 0;
    if (int rc = current_sm_count(&sms)) return rc;

    const size_t bytes = n * ESZ;
    if (bits % 32 == 0) {
        if (bytes < Tuning::kMediumBytes)
            return launch_vec<OP, BF16, 32, ShapeSmall::unroll, ShapeSmall::block>(dst, x, y, s, n, sms, st);
        if (bytes < Tuning::kLargeBytes || OP == OP_COPY)
            return launch_vec<OP, BF16, 32, ShapeMedium::unroll, ShapeMedium::block>(dst, x, y, s, n, sms, st);
        if constexpr (TWO && !BF16)
            return launch_vec<OP, BF16, 32, ShapeLarge2::unroll, ShapeLarge2::block>(dst, x, y, s, n, sms, st);
        else   // scale, and the bf16 two-input kernels (profiles/r01_tune_bf16.csv)
            return launch_vec<OP, BF16, 32, ShapeLarge1024::unroll, ShapeLarge1024::block>(dst, x, y, s, n, sms, st);
    }
    if (bits % 16 == 0) {
        if constexpr (TWO)
            return launch_vec<OP, BF16, 16, ShapeVec16Two::unroll, ShapeVec16Two::block>(dst, x, y, s, n, sms, st);
        else
            return launch_vec<OP, BF16, 16, ShapeVec16One::unroll, ShapeVec16One::block>(dst, x, y, s, n, sms, st);
    }
    size_t grid = (n + 255) / 256;
    size_t cap = (size_t)sms * 32;
    if (grid > cap) grid = cap;
    stream_kernel_unaligned<OP, BF16><<<(unsigned)grid, 256, 0, st>>>(dst, x, y, s, n);
    B200V_CUDA(cudaGetLastError());
    return B200V_OK;
}

// ------------------------------------------------------------------------------------------
// fill: the index-hash generator (same integer spec as oracle/oracle.c and csrc/hostcheck.cpp)

__host__ __device__ __forceinline__ uint32_t mix32(uint32_t x)
{
    x ^= x >> 16;
    x *= 0x7FEB352Du;
    x ^= x >> 15;
    x *= 0x846CA68Bu;
    x ^= x >> 16;
    return x;
}

__host__ __device__ __forceinline__ uint32_t hash32(uint32_t key, uint64_t index)
{
    return mix32((uint32_t)index ^ key ^ ((uint32_t)(index >> 32) * 0x9E3779B9u));
}

__device__ __forceinline__ uint32_t gen_f32_bits(uint32_t key, uint64_t index, int mode)
{
    uint32_t h = hash32(key, index);
    if (mode == B200V_GEN_UNIT) return __float_as_uint(__fmul_rn((float)(h >> 8), 0x1p-24f));
    if (mode == B200V_GEN_DENORM) return h & 0x81FFFFFFu;
    if ((h & 0x7F800000u) == 0x7F800000u) h &= ~0x00800000u;
    return h;
}

__device__ __forceinline__ uint32_t gen_bf16_bits(uint32_t key, uint64_t index, int mode)
{
    uint32_t h = hash32(key, index);
    if (mode == B200V_GEN_UNIT) return __float_as_uint(__fmul_rn((float)(h >> 24), 0x1p-8f)) >> 16;
    uint32_t b = h >> 16;
    if (mode == B200V_GEN_DENORM) return b & 0x81FFu;
    if ((b & 0x7F80u) == 0x7F80u) b &= ~0x0080u;
    return b;
}

// The fill was integer-ALU bound (ncu on the round-1 shape: ALU pipe 94 % busy, DRAM 70 % for fp32), so the vector body
// spends as few ALU-pipe instructions per element as the integer spec allows (now DRAM 83 % for fp32, 68 % for bf16:
// profiles/r02_ncu_all_kernels.txt):
//  * the generator mode is a template parameter;
//  * the hash's high-half term is hoisted out of the per-element work (32-bit index arithmetic; a vector that
//    straddles a 2^32 index boundary takes the plain path);
//  * when the first index of every vector is a multiple of the vector length (ALIGNED; true whenever start + head is,
//    which the host checks once), lo + k == lo ^ k, so (lo + k) ^ kh is ONE xor with a hoisted lo ^ kh;
//  * the last xor-shift of mix32 is folded into what consumes it: for fp32 UNIT, (float)(h >> 8) * 2^-24 is computed
//    as (float)(h & 0xFFFFFF00) * 2^-32 (same value: 24 significant bits either way), and (x ^ t) & mask is a single
//    LOP3; bf16 only ever reads the top 16 bits of h, which the last xor-shift (x ^= x >> 16) does not change, so it
//    is skipped there.
// First version 4.2 / 3.6 TB/s (fp32 / bf16 UNIT), round 1 5.9 / 4.3, now 6.9 / 6.1 (cudaMemset: 7.1).  Doing the shifts as
// multiply-high on the FMA pipe was tried and is slower (IMAD.HI is half rate).
// Same integer function as mix32/hash32 above, bit for bit (tests/test_gpu_parity.py: fill vs oracle).

// mix32 up to and including the second multiply; h = x ^ (x >> 16) is left to the consumer
__device__ __forceinline__ uint32_t mix32_head(uint32_t x)
{
    x ^= x >> 16;
    x *= 0x7FEB352Du;
    x ^= x >> 15;
    x *= 0x846CA68Bu;
    return x;
}

template <int MODE> __device__ __forceinline__ uint32_t f32_from_head(uint32_t x)
{
    const uint32_t t = x >> 16;
    if constexpr (MODE == B200V_GEN_UNIT) {
        return __float_as_uint(__fmul_rn((float)((x ^ t) & 0xFFFFFF00u), 0x1p-32f));
    } else if constexpr (MODE == B200V_GEN_DENORM) {
        return (x ^ t) & 0x81FFFFFFu;
    } else {
        const uint32_t h = x ^ t;
        return (~h & 0x7F800000u) == 0u ? h & ~0x00800000u : h;
    }
}

template <int MODE> __device__ __forceinline__ uint32_t bf16_from_head(uint32_t x)
{
    // h >> 16 == x >> 16: the last xor-shift only touches the low half
    if constexpr (MODE == B200V_GEN_UNIT) {
        return __float_as_uint(__fmul_rn((float)(x >> 24), 0x1p-8f)) >> 16;
    } else if constexpr (MODE == B200V_GEN_DENORM) {
        return (x >> 16) & 0x81FFu;
    } else {
        const uint32_t b = x >> 16;
        return (~b & 0x7F80u) == 0u ? b & ~0x0080u : b;
    }
}

// One thread produces 16 bytes (4 fp32 / 8 bf16) where dst is 16-byte aligned; the ragged
// head and tail are written element-wise.
template <bool BF16, int MODE, bool ALIGNED>
__global__ void __launch_bounds__(256)
fill_kernel(void* __restrict__ dst, uint64_t start, size_t n, uint32_t key)
{
    constexpr size_t EPV = BF16 ? 8 : 4;
    constexpr size_t ESZ = BF16 ? 2 : 4;
    size_t head = ((16 - ((uintptr_t)dst & 15)) & 15) / ESZ;
    if (head > n) head = n;
    const size_t nvec = (n - head) / EPV;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint4* vdst = reinterpret_cast<uint4*>(reinterpret_cast<char*>(dst) + head * ESZ);
    for (size_t v = tid; v < nvec; v += stride) {
        const uint64_t i0 = start + head + v * EPV;
        const uint32_t lo0 = (uint32_t)i0;
        uint32_t w[EPV];
        if (ALIGNED || lo0 <= 0xFFFFFFFFu - (uint32_t)EPV) {
            // all EPV indices share the high half: one hoisted term, 32-bit index arithmetic
            // (ALIGNED: lo0 is a multiple of EPV, so the vector can never straddle a 2^32 boundary)
            const uint32_t kh = key ^ ((uint32_t)(i0 >> 32) * 0x9E3779B9u);
            const uint32_t base = lo0 ^ kh;
#pragma unroll
            for (int k = 0; k < (int)EPV; ++k) {
                const uint32_t x = mix32_head(ALIGNED ? base ^ (uint32_t)k : (lo0 + (uint32_t)k) ^ kh);
                w[k] = BF16 ? bf16_from_head<MODE>(x) : f32_from_head<MODE>(x);
            }
        } else {
#pragma unroll
            for (int k = 0; k < (int)EPV; ++k)
                w[k] = BF16 ? gen_bf16_bits(key, i0 + k, MODE) : gen_f32_bits(key, i0 + k, MODE);
        }
        uint4 o;
        if constexpr (!BF16) {
            o = make_uint4(w[0], w[1], w[2], w[3]);
        } else {
            o = make_uint4(w[0] | (w[1] << 16), w[2] | (w[3] << 16), w[4] | (w[5] << 16), w[6] | (w[7] << 16));
        }
        vdst[v] = o;
    }
    // ragged ends: [0, head) and [head + nvec*EPV, n)
    const size_t tail0 = head + nvec * EPV;
    const size_t ragged = head + (n - tail0);
    for (size_t r = tid; r < ragged; r += stride) {
        const size_t j = r < head ? r : tail0 + (r - head);
        if constexpr (!BF16) reinterpret_cast<uint32_t*>(dst)[j] = gen_f32_bits(key, start + j, MODE);
        else reinterpret_cast<uint16_t*>(dst)[j] = (uint16_t)gen_bf16_bits(key, start + j, MODE);
    }
}

// ------------------------------------------------------------------------------------------
// checksum: out[0] = sum w_j, out[1] = sum w_j * ((uint32_t)(word_offset + j) | 1)  (mod 2^64)
//
// Workspace layout (caller-owned scratch): bytes [0,16): a uint32 arrival counter (zeroed by the
// entry point before every launch); then one {s0, s1} uint64 pair per CTA.  Every CTA reduces its
// grid-stride share, publishes its pair, and the last CTA to arrive adds the pairs up.

// shipped shape of the reduction (tools/tune.cu `cksum`, profiles/)
constexpr int kChecksumUnroll = 4;
constexpr bool kChecksumTiled = true;   // 6870 vs 6683 GB/s read at 1 GiB (profiles/r01_tune_cksum.csv)

// ------------------------------------------------------------------------------------------
// peer pull / peer push: one GPU's SMs move a buffer across NVLink with 256-bit accesses.
//   pull: the DESTINATION GPU runs the kernel: loads from the peer's address space, stores to local HBM.
//         Every 32 bytes cost a read request going out and the data coming back, so under bidirectional
//         load the requests share the link direction the other flow's data uses.
//   push: the SOURCE GPU runs the kernel: loads from local HBM, stores into the peer's address space.
//         Stores are posted: nothing travels back but acknowledgements, which is why the copy engine
//         (which also pushes) keeps its rate when every GPU sends and receives at once.
// Peer addresses bypass the local L2, so no cache hint is asked for.  Same body for both directions;
// two entry kernels so a profile shows which one ran.
//
// Measured (tools/tune_p2p.cu, profiles/r02_tune_p2p_2gpu.csv, 1 GiB, GPU0 <-> GPU1), one way / both ways at once:
//   copy engine 779 / 773   pull 784 / 672   push 714 / 705 GB/s per direction.
// Every SM-issued store variant lands on the same 714 / 704: 128- or 256-bit, any unroll / block / grid, .cs and
// no-allocate / evict-first hints, and the bulk-copy engine (cp.async.bulk shared -> peer global, 4..64 KiB tiles,
// 2..4 stages, 1..8 CTAs per SM).  The ceiling is the write packets SMs (and TMA) emit on NVLink, not the kernel:
// the copy engine's writes carry more payload per packet.  So: push is the better KERNEL when traffic is
// bidirectional (+5.5 % over pull), pull the better one for a one-way transfer, and cudaMemcpyPeer beats both
// when both directions are busy.

constexpr int kPullBlock = 512;
constexpr int kPullUnroll = 4;

template <int BLOCK, int UNROLL>
__device__ __forceinline__ void peer_move_body(Vec<32>* __restrict__ dst, const Vec<32>* __restrict__ src, size_t nvec)
{
    constexpr size_t TILE = (size_t)BLOCK * UNROLL;
    const size_t tiles = nvec / TILE;
    for (size_t t = blockIdx.x; t < tiles; t += gridDim.x) {
        const size_t base = t * TILE + threadIdx.x;
        Vec<32> v[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) v[u] = ldg<LD_DEFAULT>(src + base + (size_t)u * BLOCK);
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) stg<ST_DEFAULT>(dst + base + (size_t)u * BLOCK, v[u]);
    }
    if (blockIdx.x == tiles % gridDim.x)
        for (size_t i = tiles * TILE + threadIdx.x; i < nvec; i += BLOCK)
            stg<ST_DEFAULT>(dst + i, ldg<LD_DEFAULT>(src + i));
}

__global__ void __launch_bounds__(kPullBlock)
peer_pull_kernel(Vec<32>* __restrict__ dst, const Vec<32>* __restrict__ src, size_t nvec)
{
    peer_move_body<kPullBlock, kPullUnroll>(dst, src, nvec);
}

__global__ void __launch_bounds__(kPullBlock)
peer_push_kernel(Vec<32>* __restrict__ peer_dst, const Vec<32>* __restrict__ local_src, size_t nvec)
{
    peer_move_body<kPullBlock, kPullUnroll>(peer_dst, local_src, nvec);
}

__global__ void __launch_bounds__(256)
byte_copy_kernel(unsigned char* __restrict__ dst, const unsigned char* __restrict__ src, size_t n)
{
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) dst[i] = src[i];
}

}  // namespace b200v

// ==========================================================================================
// C ABI
// ==========================================================================================
using namespace b200v;

extern "C" {

int b200v_stream_op(int op, int dtype, void* dst, const void* x, const void* y, float s, size_t n, void* stream)
{
    cudaStream_t st = as_stream(stream);
    B200V_REQUIRE(dtype == B200V_F32 || dtype == B200V_BF16, B200V_ERR_INVALID, "unknown dtype %d", dtype);
    const bool bf = dtype == B200V_BF16;
    switch (op) {
    case B200V_OP_ADD:
        return bf ? launch_stream<OP_ADD, true>(dst, x, y, s, n, st) : launch_stream<OP_ADD, false>(dst, x, y, s, n, st);
    case B200V_OP_COPY:
        return bf ? launch_stream<OP_COPY, true>(dst, x, y, s, n, st) : launch_stream<OP_COPY, false>(dst, x, y, s, n, st);
    case B200V_OP_SCALE:
        return bf ? launch_stream<OP_SCALE, true>(dst, x, y, s, n, st) : launch_stream<OP_SCALE, false>(dst, x, y, s, n, st);
    case B200V_OP_TRIAD:
        return bf ? launch_stream<OP_TRIAD, true>(dst, x, y, s, n, st) : launch_stream<OP_TRIAD, false>(dst, x, y, s, n, st);
    default:
        return set_error(B200V_ERR_INVALID, "unknown op %d", op);
    }
}

int b200v_add_f32(float* c, const float* a, const float* b, size_t n, void* stream)
{
    return launch_stream<OP_ADD, false>(c, a, b, 0.0f, n, as_stream(stream));
}
int b200v_copy_f32(float* c, const float* a, size_t n, void* stream)
{
    return launch_stream<OP_COPY, false>(c, a, nullptr, 0.0f, n, as_stream(stream));
}
int b200v_scale_f32(float* b, float s, const float* c, size_t n, void* stream)
{
    return launch_stream<OP_SCALE, false>(b, c, nullptr, s, n, as_stream(stream));
}
int b200v_triad_f32(float* a, float s, const float* c, const float* b, size_t n, void* stream)
{
    return launch_stream<OP_TRIAD, false>(a, c, b, s, n, as_stream(stream));
}
int b200v_add_bf16(uint16_t* c, const uint16_t* a, const uint16_t* b, size_t n, void* stream)
{
    return launch_stream<OP_ADD, true>(c, a, b, 0.0f, n, as_stream(stream));
}
int b200v_copy_bf16(uint16_t* c, const uint16_t* a, size_t n, void* stream)
{
    return launch_stream<OP_COPY, true>(c, a, nullptr, 0.0f, n, as_stream(stream));
}
int b200v_scale_bf16(uint16_t* b, float s, const uint16_t* c, size_t n, void* stream)
{
    return launch_stream<OP_SCALE, true>(b, c, nullptr, s, n, as_stream(stream));
}
int b200v_triad_bf16(uint16_t* a, float s, const uint16_t* c, const uint16_t* b, size_t n, void* stream)
{
    return launch_stream<OP_TRIAD, true>(a, c, b, s, n, as_stream(stream));
}

uint64_t b200v_op_bytes(int op, int dtype, size_t n)
{
    uint64_t esz = dtype == B200V_BF16 ? 2 : 4;
    uint64_t arrays = (op == B200V_OP_ADD || op == B200V_OP_TRIAD) ? 3 : 2;
    return arrays * esz * (uint64_t)n;
}

int b200v_fill(int dtype, void* d_dst, uint64_t start, size_t n, uint32_t seed, uint32_t stream_id, int mode, void* stream)
{
    B200V_REQUIRE(dtype == B200V_F32 || dtype == B200V_BF16, B200V_ERR_INVALID, "unknown dtype %d", dtype);
    B200V_REQUIRE(mode >= B200V_GEN_UNIT && mode <= B200V_GEN_DENORM, B200V_ERR_INVALID, "unknown generator mode %d", mode);
    B200V_REQUIRE(d_dst != nullptr, B200V_ERR_INVALID, "NULL destination");
    if (n == 0) return B200V_OK;
    const size_t esz = dtype == B200V_BF16 ? 2 : 4;
    B200V_REQUIRE((uintptr_t)d_dst % esz == 0, B200V_ERR_ALIGN, "destination not aligned to its %zu-byte element", esz);
    int sms = 0;
    if (int rc = current_sm_count(&sms)) return rc;
    const uint32_t key = mix32(seed ^ (0x9E3779B9u * (stream_id + 1u)));
    const size_t epv = dtype == B200V_BF16 ? 8 : 4;
    size_t grid = (n / epv + 255) / 256 + 1;
    // 64 CTAs per SM (8 waves of 8 resident CTAs): enough CTAs for the block scheduler to even out the SMs, few enough
    // that their prologues do not show.  1 GiB fp32 UNIT: 16/SM 6.40, 32/SM 6.73, 64/SM 6.93, 128/SM 7.05 (but bf16 and
    // BITS drop), 256/SM 6.40, one vector per thread (262144 CTAs) 2.8 TB/s (profiles/r02_time_fill.txt).
    const size_t cap = (size_t)sms * 64;
    if (grid > cap) grid = cap;
    cudaStream_t st = as_stream(stream);
    const unsigned g = (unsigned)grid;
    const bool bf = dtype == B200V_BF16;
    // is the first index of every vector a multiple of the vector length?  (head = the kernel's: elements before
    // 16-byte alignment of the destination)
    size_t head = ((16 - ((uintptr_t)d_dst & 15)) & 15) / esz;
    if (head > n) head = n;
    const bool aligned = ((start + head) % epv) == 0;
#define B200V_FILL(BF, MODE)                                                                  \
    do {                                                                                      \
        if (aligned) fill_kernel<BF, MODE, true><<<g, 256, 0, st>>>(d_dst, start, n, key);    \
        else fill_kernel<BF, MODE, false><<<g, 256, 0, st>>>(d_dst, start, n, key);           \
    } while (0)
    switch (mode) {
    case B200V_GEN_UNIT:
        if (bf) B200V_FILL(true, B200V_GEN_UNIT);
        else B200V_FILL(false, B200V_GEN_UNIT);
        break;
    case B200V_GEN_BITS:
        if (bf) B200V_FILL(true, B200V_GEN_BITS);
        else B200V_FILL(false, B200V_GEN_BITS);
        break;
    default:
        if (bf) B200V_FILL(true, B200V_GEN_DENORM);
        else B200V_FILL(false, B200V_GEN_DENORM);
        break;
    }
#undef B200V_FILL
    B200V_CUDA(cudaGetLastError());
    return B200V_OK;
}

size_t b200v_checksum_workspace_bytes(void) { return 16 + (size_t)kChecksumMaxBlocks * 16; }

int b200v_checksum(const void* d_buf, size_t nbytes, uint64_t word_offset, void* d_workspace, size_t workspace_bytes,
                   uint64_t* d_out, void* stream)
{
    B200V_REQUIRE(d_workspace && d_out, B200V_ERR_INVALID, "NULL workspace or output");
    B200V_REQUIRE(d_buf || nbytes == 0, B200V_ERR_INVALID, "NULL buffer");
    B200V_REQUIRE(workspace_bytes >= b200v_checksum_workspace_bytes(), B200V_ERR_INVALID,
                  "workspace is %zu bytes, need %zu", workspace_bytes, b200v_checksum_workspace_bytes());
    B200V_REQUIRE((uintptr_t)d_workspace % 16 == 0 && (uintptr_t)d_out % 8 == 0, B200V_ERR_ALIGN,
                  "workspace must be 16-byte and output 8-byte aligned");
    B200V_REQUIRE((uintptr_t)d_buf % 4 == 0, B200V_ERR_ALIGN, "buffer must be 4-byte aligned");
    int sms = 0;
    if (int rc = current_sm_count(&sms)) return rc;
    cudaStream_t st = as_stream(stream);
    size_t grid = (nbytes / 16 + kChecksumBlock * kChecksumUnroll - 1) / (kChecksumBlock * kChecksumUnroll);
    size_t cap = (size_t)sms * 8;
    if (cap > kChecksumMaxBlocks) cap = kChecksumMaxBlocks;
    if (grid > cap) grid = cap;
    if (grid == 0) grid = 1;
    unsigned int* counter = reinterpret_cast<unsigned int*>(d_workspace);
    uint64_t* partials = reinterpret_cast<uint64_t*>(reinterpret_cast<char*>(d_workspace) + 16);
    B200V_CUDA(cudaMemsetAsync(counter, 0, 16, st));
    checksum_kernel<kChecksumUnroll, kChecksumTiled><<<(unsigned)grid, kChecksumBlock, 0, st>>>(reinterpret_cast<const unsigned char*>(d_buf), nbytes,
                                                              word_offset, counter, partials, d_out);
    B200V_CUDA(cudaGetLastError());
    return B200V_OK;
}

// ------------------------------------------------------------------------------------------
// b200v_checksum_static: the checksum for callers that have no way to hand over scratch memory.
//
// Some users of this library cannot allocate device memory at the moment they want a checksum:
// a burn-in that has deliberately filled every free byte of HBM with its arrays (a later
// cudaMalloc simply fails), a liveness probe living inside a process whose framework allocator
// has already reserved the whole GPU, or a monitoring hook that must not call cudaMalloc/cudaFree
// because those calls synchronise the device and would stall the work it is watching.  For them
// the reduction's scratch lives in the library image itself: the file-scope array below, present
// on every GPU that loads libb200v (about 16 KiB per GPU, held until the process exits).
//
// Same kernel, same result as b200v_checksum; only the origin of the scratch differs.
//  * Limit: buffers up to B200V_CHECKSUM_STATIC_MAX_BYTES (4 GiB).  The array has room for one
//    partial pair per CTA for at most kStaticMaxBlocks CTAs, and the limit budgets 4 MiB of input
//    per slot: since every caller on the device shares the array, one call should not keep it
//    busy for long (4 GiB is about 0.6 ms of HBM reading).  Anything larger is refused with
//    B200V_ERR_INVALID — checksum it in pieces (pieces add up, see b200v_checksum) or use
//    b200v_checksum with your own workspace.
//  * The FIRST call on each device resolves the array's device address (cudaGetSymbolAddress,
//    with that device current), which may make the driver load the module.  Make that first call
//    during start-up, before any stream capture is begun.
//  * Every call on a device shares the one array and finds the previous call's partials in it.
//    The arrival counter is cleared on the stream before each launch; the partials need no
//    clearing (each CTA overwrites its own slot before the combine reads it).  Two calls that
//    overlap on the GPU (different streams, no ordering between them) race on the array and
//    give garbage: keep calls on one stream, or order them with events.
//  * Nothing here calls cudaMalloc, cudaFree or any allocating library routine.

constexpr int kStaticMaxBlocks = 1024;
constexpr size_t kStaticHeaderBytes = 256;   // arrival counter, padded so the partials stay 256-byte aligned
constexpr size_t kStaticScratchBytes = kStaticHeaderBytes + (size_t)kStaticMaxBlocks * 16;
constexpr size_t kStaticBytesPerBlock = (size_t)4 << 20;
static_assert((uint64_t)kStaticMaxBlocks * kStaticBytesPerBlock == B200V_CHECKSUM_STATIC_MAX_BYTES,
              "the documented limit and the scratch array disagree");

// In host code this name is only a handle for cudaGetSymbolAddress, never a device pointer.
static __device__ __align__(256) unsigned char g_checksum_static_scratch[kStaticScratchBytes];

// device address of the array, per device ordinal, resolved on first use
static void* g_checksum_static_addr[64] = {};

int b200v_checksum_static(const void* d_buf, size_t nbytes, uint64_t word_offset, uint64_t* d_out, void* stream)
{
    B200V_REQUIRE(d_out != nullptr, B200V_ERR_INVALID, "NULL output");
    B200V_REQUIRE(d_buf || nbytes == 0, B200V_ERR_INVALID, "NULL buffer");
    B200V_REQUIRE((uint64_t)nbytes <= B200V_CHECKSUM_STATIC_MAX_BYTES, B200V_ERR_INVALID,
                  "b200v_checksum_static takes at most %llu bytes per call (got %zu): split the buffer or use "
                  "b200v_checksum with a caller-owned workspace",
                  (unsigned long long)B200V_CHECKSUM_STATIC_MAX_BYTES, nbytes);
    B200V_REQUIRE((uintptr_t)d_out % 8 == 0, B200V_ERR_ALIGN, "output must be 8-byte aligned");
    B200V_REQUIRE((uintptr_t)d_buf % 4 == 0, B200V_ERR_ALIGN, "buffer must be 4-byte aligned");
    int sms = 0;
    if (int rc = current_sm_count(&sms)) return rc;
    int dev = 0;
    B200V_CUDA(cudaGetDevice(&dev));
    if (g_checksum_static_addr[dev] == nullptr) {   // once per device, with that device current
        void* p = nullptr;
        B200V_CUDA(cudaGetSymbolAddress(&p, g_checksum_static_scratch));
        B200V_REQUIRE(p != nullptr, B200V_ERR_CUDA, "cudaGetSymbolAddress returned a NULL scratch address");
        g_checksum_static_addr[dev] = p;
    }
    unsigned char* scratch = static_cast<unsigned char*>(g_checksum_static_addr[dev]);
    unsigned int* counter = reinterpret_cast<unsigned int*>(scratch);
    uint64_t* partials = reinterpret_cast<uint64_t*>(scratch + kStaticHeaderBytes);

    cudaStream_t st = as_stream(stream);
    size_t grid = (nbytes / 16 + kChecksumBlock * kChecksumUnroll - 1) / (kChecksumBlock * kChecksumUnroll);
    size_t cap = (size_t)sms * 4;
    if (cap > kStaticMaxBlocks) cap = kStaticMaxBlocks;
    if (grid > cap) grid = cap;
    if (grid == 0) grid = 1;
    B200V_CUDA(cudaMemsetAsync(counter, 0, 16, st));   // the counter must start at zero; a memset allocates nothing
    checksum_kernel<kChecksumUnroll, kChecksumTiled><<<(unsigned)grid, kChecksumBlock, 0, st>>>(reinterpret_cast<const unsigned char*>(d_buf), nbytes,
                                                              word_offset, counter, partials, d_out);
    B200V_CUDA(cudaGetLastError());
    return B200V_OK;
}

// Shared by pull and push: the 32-byte-aligned body goes through the 256-bit kernel, whatever is left
// (an unaligned pair of pointers, or the last < 32 bytes) through the byte kernel.
static int launch_peer_move(bool push, void* dst, const void* src, size_t bytes, void* stream)
{
    B200V_REQUIRE(dst && src, B200V_ERR_INVALID, "NULL pointer");
    if (bytes == 0) return B200V_OK;
    int sms = 0;
    if (int rc = current_sm_count(&sms)) return rc;
    cudaStream_t st = as_stream(stream);
    const bool aligned = (((uintptr_t)dst | (uintptr_t)src) % 32) == 0;
    size_t body = aligned ? (bytes / 32) * 32 : 0;
    if (body) {
        size_t nvec = body / 32;
        // one 64 KiB tile per CTA, like the streaming kernels (profiles/r02_tune_p2p_2gpu.csv: pull 784 / 672 GB/s
        // one way / both ways against 777 / 667 for 4 CTAs per SM; push is 713 / 704 either way)
        size_t grid = nvec / ((size_t)kPullBlock * kPullUnroll);
        if (grid > 0x7FFFFFFFull) grid = 0x7FFFFFFFull;
        if (grid == 0) grid = 1;
        if (push)
            peer_push_kernel<<<(unsigned)grid, kPullBlock, 0, st>>>(reinterpret_cast<Vec<32>*>(dst),
                                                                   reinterpret_cast<const Vec<32>*>(src), nvec);
        else
            peer_pull_kernel<<<(unsigned)grid, kPullBlock, 0, st>>>(reinterpret_cast<Vec<32>*>(dst),
                                                                   reinterpret_cast<const Vec<32>*>(src), nvec);
        B200V_CUDA(cudaGetLastError());
    }
    if (body < bytes) {
        size_t rest = bytes - body;
        size_t grid = (rest + 255) / 256;
        size_t cap = (size_t)sms * 8;
        if (grid > cap) grid = cap;
        byte_copy_kernel<<<(unsigned)grid, 256, 0, st>>>(reinterpret_cast<unsigned char*>(dst) + body,
                                                        reinterpret_cast<const unsigned char*>(src) + body, rest);
        B200V_CUDA(cudaGetLastError());
    }
    return B200V_OK;
}

int b200v_peer_pull(void* d_local_dst, const void* d_peer_src, size_t bytes, void* stream)
{
    return launch_peer_move(false, d_local_dst, d_peer_src, bytes, stream);
}

int b200v_peer_push(void* d_peer_dst, const void* d_local_src, size_t bytes, void* stream)
{
    return launch_peer_move(true, d_peer_dst, d_local_src, bytes, stream);
}

}  // extern "C"

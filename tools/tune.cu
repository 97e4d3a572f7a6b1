// tune.cu — sweep the launch shapes / cache hints of the streaming kernels on a B200 and print a
// CSV table (one row per configuration).  The product's csrc/tuning.h is set from these tables
// (kept under profiles/).  Not part of the product; builds against the same stream_kernels.cuh.
//
//   tools/tune [log2_n=28] [iters=10] [stage=all|shape|hint|tma|tma1|scale|small|pdl|bf16|lockstep|ref]
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <string>
#include <vector>

#include "stream_kernels.cuh"
#include "stream_tma.cuh"
#include "checksum_kernel.cuh"

using namespace b200v;

#define CK(x)                                                                                   \
    do {                                                                                        \
        cudaError_t e_ = (x);                                                                   \
        if (e_ != cudaSuccess) {                                                                \
            fprintf(stderr, "%s: %s at %s:%d\n", #x, cudaGetErrorString(e_), __FILE__, __LINE__); \
            exit(2);                                                                            \
        }                                                                                       \
    } while (0)

__global__ void count_mismatch(const uint32_t* a, const uint32_t* b, size_t n, unsigned long long* out)
{
    unsigned long long bad = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        bad += a[i] != b[i];
    if (bad) atomicAdd(out, bad);
}

__global__ void init_values(float* p, size_t n, uint32_t salt)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        p[i] = (float)((((uint32_t)i ^ salt) * 2654435761u) >> 8) * 0x1p-24f;
}

// Experiment for DESIGN.md's "persistent CTAs run in lock-step" reading: the product's main loop as a
// persistent grid, optionally with every CTA starting after its own short delay (0 .. 6.3 us).
// Occupancy probe: the one-tile body with PAD registers held live across it (read from %clock at entry,
// "used" at exit), so the resident CTAs per SM can be dialled through the register file alone — no shared
// memory request (which would shrink L1), no change to the memory instructions.
template <int OP, int UNROLL, int BLOCK, int PAD>
__global__ void __launch_bounds__(BLOCK)
occupancy_probe_kernel(void* __restrict__ dst, const void* __restrict__ x, const void* __restrict__ y, float s, size_t n)
{
    using V = Vec<32>;
    constexpr size_t TILE_VECS = (size_t)BLOCK * UNROLL;
    const size_t full_tiles = n / (TILE_VECS * 8);
    uint32_t pad[PAD > 0 ? PAD : 1];
#pragma unroll
    for (int i = 0; i < PAD; ++i) asm volatile("mov.u32 %0, %%clock;" : "=r"(pad[i]));
    V* vd = reinterpret_cast<V*>(dst);
    const V* vx = reinterpret_cast<const V*>(x);
    const V* vy = reinterpret_cast<const V*>(y);
    uint32_t last = 0;
#pragma unroll 1
    for (size_t tile = blockIdx.x; tile < full_tiles; tile += gridDim.x) {
        const size_t base = tile * TILE_VECS + threadIdx.x;
        V a[UNROLL], b[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) a[u] = ldg<LD_DEFAULT>(vx + base + (size_t)u * BLOCK);
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) b[u] = ldg<LD_DEFAULT>(vy + base + (size_t)u * BLOCK);
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            V r;
#pragma unroll
            for (int k = 0; k < 8; ++k) r.w[k] = op_word<OP, false>(a[u].w[k], b[u].w[k], s);
            stg<ST_DEFAULT>(vd + base + (size_t)u * BLOCK, r);
            last = r.w[0];
        }
    }
    // a chain that needs every pad register after the last loaded value is known keeps them all live
    if (PAD > 0) {
        uint32_t acc = last;
#pragma unroll
        for (int i = 0; i < PAD; ++i) acc = acc * pad[i] + 1u;
        if (acc == 0xDEADBEEFu) reinterpret_cast<uint32_t*>(dst)[0] = acc;
    }
}

template <int OP, int UNROLL, int BLOCK, bool STAGGER>
__global__ void __launch_bounds__(BLOCK)
persistent_probe_kernel(void* __restrict__ dst, const void* __restrict__ x, const void* __restrict__ y, float s, size_t n)
{
    using V = Vec<32>;
    constexpr size_t TILE_VECS = (size_t)BLOCK * UNROLL;
    const size_t full_tiles = n / (TILE_VECS * 8);
    if (STAGGER) __nanosleep(((blockIdx.x * 37u) % 64u) * 100u);
    V* vd = reinterpret_cast<V*>(dst);
    const V* vx = reinterpret_cast<const V*>(x);
    const V* vy = reinterpret_cast<const V*>(y);
    for (size_t tile = blockIdx.x; tile < full_tiles; tile += gridDim.x) {
        const size_t base = tile * TILE_VECS + threadIdx.x;
        V a[UNROLL], b[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) a[u] = ldg<LD_DEFAULT>(vx + base + (size_t)u * BLOCK);
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) b[u] = ldg<LD_DEFAULT>(vy + base + (size_t)u * BLOCK);
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            V r;
#pragma unroll
            for (int k = 0; k < 8; ++k) r.w[k] = op_word<OP, false>(a[u].w[k], b[u].w[k], s);
            stg<ST_DEFAULT>(vd + base + (size_t)u * BLOCK, r);
        }
    }
}

// Mechanism probes for "why does a persistent grid lose to one tile per CTA?" (profiles/r02_persist_vs_tiles.txt).
//   MODE 0: plain persistent loop (as persistent_probe_kernel)
//   MODE 1: the same with __syncthreads() at the top of every tile: the CTA's 16 warps stay in phase, so every tile is
//           requested as one 2 x 64 KiB burst, as it is when a fresh CTA starts (tests: request-phase coherence)
//   MODE 2: persistent CTAs that FETCH their next tile from a global counter instead of owning a fixed stride (tests:
//           dynamic load balancing over SMs); the fetch needs a CTA-wide broadcast, so it implies MODE 1's barrier
//   MODE 3: static stride, but the CTA's whole tile range is contiguous (CTA b owns tiles [b*T/G, (b+1)*T/G)) — each SM
//           streams its own private region (tests: does it matter that concurrently active tiles are neighbours?)
__device__ unsigned long long g_tile_counter;

template <int OP, int UNROLL, int BLOCK, int MODE>
__global__ void __launch_bounds__(BLOCK)
mechanism_probe_kernel(void* __restrict__ dst, const void* __restrict__ x, const void* __restrict__ y, float s, size_t n)
{
    using V = Vec<32>;
    constexpr size_t TILE_VECS = (size_t)BLOCK * UNROLL;
    const size_t full_tiles = n / (TILE_VECS * 8);
    V* vd = reinterpret_cast<V*>(dst);
    const V* vx = reinterpret_cast<const V*>(x);
    const V* vy = reinterpret_cast<const V*>(y);
    __shared__ unsigned long long next_tile;
    size_t tile = blockIdx.x, end = full_tiles, step = gridDim.x;
    if (MODE == 3) {
        tile = full_tiles * blockIdx.x / gridDim.x;
        end = full_tiles * (blockIdx.x + 1) / gridDim.x;
        step = 1;
    }
    while (true) {
        if (MODE == 2) {
            __syncthreads();
            if (threadIdx.x == 0) next_tile = atomicAdd(&g_tile_counter, 1ull);
            __syncthreads();
            tile = next_tile;
        } else if (MODE == 1) {
            __syncthreads();
        }
        if (tile >= end) break;
        const size_t base = tile * TILE_VECS + threadIdx.x;
        V a[UNROLL], b[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) a[u] = ldg<LD_DEFAULT>(vx + base + (size_t)u * BLOCK);
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) b[u] = ldg<LD_DEFAULT>(vy + base + (size_t)u * BLOCK);
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            V r;
#pragma unroll
            for (int k = 0; k < 8; ++k) r.w[k] = op_word<OP, false>(a[u].w[k], b[u].w[k], s);
            stg<ST_DEFAULT>(vd + base + (size_t)u * BLOCK, r);
        }
        tile += step;
    }
}

struct Config {
    std::string stage;
    int op, vecb, unroll, block, ldh, sth;
    int smem = 0;          // > 0: TMA variant (vecb = tile KiB, unroll = stages); grid modes are CTAs per SM

    size_t tile_elems;
    std::function<void(unsigned, void*, const void*, const void*, float, size_t)> launch;
    int regs;
};

static std::vector<Config> g_cfg;

// BF16 = true: the arrays are read as bf16 (2 elements per word); n passed to the kernel is doubled so
// the bytes moved are the same as in the fp32 rows
template <int OP, int VECB, int UNROLL, int BLOCK, int LDH, int STH, bool BF16 = false>
static void add_cfg(const char* stage)
{
    Config c;
    c.stage = stage;
    c.op = OP; c.vecb = VECB; c.unroll = UNROLL; c.block = BLOCK; c.ldh = LDH; c.sth = STH;
    c.tile_elems = (size_t)BLOCK * UNROLL * (VECB / 4);
    auto k = stream_kernel<OP, BF16, VECB, UNROLL, BLOCK, LDH, STH>;
    cudaFuncAttributes fa;
    CK(cudaFuncGetAttributes(&fa, k));
    c.regs = fa.numRegs;
    c.launch = [k](unsigned grid, void* d, const void* x, const void* y, float s, size_t n) {
        k<<<grid, BLOCK>>>(d, x, y, s, BF16 ? 2 * n : n);
    };
    g_cfg.push_back(c);
}

template <int OP> static void add_bf16_shapes(const char* stage)
{
    add_cfg<OP, 32, 1, 256, LD_DEFAULT, ST_DEFAULT, true>(stage);
    add_cfg<OP, 32, 2, 256, LD_DEFAULT, ST_DEFAULT, true>(stage);
    add_cfg<OP, 32, 4, 256, LD_DEFAULT, ST_DEFAULT, true>(stage);
    add_cfg<OP, 32, 1, 512, LD_DEFAULT, ST_DEFAULT, true>(stage);
    add_cfg<OP, 32, 2, 512, LD_DEFAULT, ST_DEFAULT, true>(stage);
    add_cfg<OP, 32, 4, 512, LD_DEFAULT, ST_DEFAULT, true>(stage);
    add_cfg<OP, 32, 1, 1024, LD_DEFAULT, ST_DEFAULT, true>(stage);
    add_cfg<OP, 32, 2, 1024, LD_DEFAULT, ST_DEFAULT, true>(stage);
    add_cfg<OP, 32, 2, 128, LD_DEFAULT, ST_DEFAULT, true>(stage);
    add_cfg<OP, 32, 4, 128, LD_DEFAULT, ST_DEFAULT, true>(stage);
    add_cfg<OP, 16, 4, 512, LD_DEFAULT, ST_DEFAULT, true>(stage);
    add_cfg<OP, 16, 4, 1024, LD_DEFAULT, ST_DEFAULT, true>(stage);
    add_cfg<OP, 32, 4, 512, LD_NC_NA, ST_DEFAULT, true>(stage);
    add_cfg<OP, 32, 2, 512, LD_NC_NA, ST_DEFAULT, true>(stage);
}

template <int OP, int UNROLL, int BLOCK, bool STAGGER> static void add_probe(const char* stage)
{
    Config c;
    c.stage = stage;
    c.op = OP; c.vecb = 32; c.unroll = UNROLL; c.block = BLOCK; c.ldh = STAGGER ? 8 : 7; c.sth = 0;   // ld 7 = plain, 8 = staggered
    c.tile_elems = (size_t)BLOCK * UNROLL * 8;
    auto k = persistent_probe_kernel<OP, UNROLL, BLOCK, STAGGER>;
    cudaFuncAttributes fa;
    CK(cudaFuncGetAttributes(&fa, k));
    c.regs = fa.numRegs;
    c.launch = [k](unsigned grid, void* d, const void* x, const void* y, float s, size_t n) { k<<<grid, BLOCK>>>(d, x, y, s, n); };
    g_cfg.push_back(c);
}

template <int OP, int UNROLL, int BLOCK, int MODE> static void add_mechanism(const char* stage)
{
    Config c;
    c.stage = stage;
    c.op = OP; c.vecb = 32; c.unroll = UNROLL; c.block = BLOCK; c.ldh = 10 + MODE; c.sth = 0;   // ld 10..13 = mechanism probe MODE 0..3
    c.tile_elems = (size_t)BLOCK * UNROLL * 8;
    auto k = mechanism_probe_kernel<OP, UNROLL, BLOCK, MODE>;
    cudaFuncAttributes fa;
    CK(cudaFuncGetAttributes(&fa, k));
    c.regs = fa.numRegs;
    c.launch = [k](unsigned grid, void* d, const void* x, const void* y, float s, size_t n) {
        if (MODE == 2) {
            unsigned long long zero = 0;
            CK(cudaMemcpyToSymbolAsync(g_tile_counter, &zero, sizeof zero, 0, cudaMemcpyHostToDevice, 0));
        }
        k<<<grid, BLOCK>>>(d, x, y, s, n);
    };
    g_cfg.push_back(c);
}

// the product kernel with its residency capped by a dummy dynamic shared-memory request: st column = CTAs per SM allowed
template <int OP, int VECB, int UNROLL, int BLOCK> static void add_occ(const char* stage, int ctas)
{
    Config c;
    c.stage = stage;
    c.op = OP; c.vecb = VECB; c.unroll = UNROLL; c.block = BLOCK; c.ldh = 6; c.sth = ctas;
    c.tile_elems = (size_t)BLOCK * UNROLL * (VECB / 4);
    auto k = stream_kernel<OP, false, VECB, UNROLL, BLOCK, LD_DEFAULT, ST_DEFAULT>;
    const int smem = ctas > 0 ? (227 * 1024 / ctas - 1024) & ~127 : 0;
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    cudaFuncAttributes fa;
    CK(cudaFuncGetAttributes(&fa, k));
    c.regs = fa.numRegs;
    c.launch = [k, smem](unsigned grid, void* d, const void* x, const void* y, float s, size_t n) {
        k<<<grid, BLOCK, smem>>>(d, x, y, s, n);
    };
    g_cfg.push_back(c);
}

// the product kernel with the L1/shared split moved (no shared memory is used): st column = carve-out percent
template <int OP, int VECB, int UNROLL, int BLOCK> static void add_carve(const char* stage, int percent)
{
    Config c;
    c.stage = stage;
    c.op = OP; c.vecb = VECB; c.unroll = UNROLL; c.block = BLOCK; c.ldh = 5; c.sth = percent;
    c.tile_elems = (size_t)BLOCK * UNROLL * (VECB / 4);
    auto k = stream_kernel<OP, false, VECB, UNROLL, BLOCK, LD_DEFAULT, ST_DEFAULT>;
    cudaFuncAttributes fa;
    CK(cudaFuncGetAttributes(&fa, k));
    c.regs = fa.numRegs;
    c.launch = [k, percent](unsigned grid, void* d, const void* x, const void* y, float s, size_t n) {
        CK(cudaFuncSetAttribute(k, cudaFuncAttributePreferredSharedMemoryCarveout, percent));
        k<<<grid, BLOCK>>>(d, x, y, s, n);
    };
    g_cfg.push_back(c);
}

template <int OP, int UNROLL, int BLOCK, int PAD> static void add_occprobe(const char* stage)
{
    Config c;
    c.stage = stage;
    c.op = OP; c.vecb = 32; c.unroll = UNROLL; c.block = BLOCK; c.ldh = 4; c.sth = PAD;   // st column = PAD
    c.tile_elems = (size_t)BLOCK * UNROLL * 8;
    auto k = occupancy_probe_kernel<OP, UNROLL, BLOCK, PAD>;
    cudaFuncAttributes fa;
    CK(cudaFuncGetAttributes(&fa, k));
    c.regs = fa.numRegs;
    c.launch = [k](unsigned grid, void* d, const void* x, const void* y, float s, size_t n) { k<<<grid, BLOCK>>>(d, x, y, s, n); };
    g_cfg.push_back(c);
}

template <int OP, int UNROLL, int BLOCK> static void add_occprobe_all(const char* stage)
{
    add_occprobe<OP, UNROLL, BLOCK, 0>(stage);
    add_occprobe<OP, UNROLL, BLOCK, 8>(stage);
    add_occprobe<OP, UNROLL, BLOCK, 16>(stage);
    add_occprobe<OP, UNROLL, BLOCK, 24>(stage);
    add_occprobe<OP, UNROLL, BLOCK, 32>(stage);
    add_occprobe<OP, UNROLL, BLOCK, 40>(stage);
    add_occprobe<OP, UNROLL, BLOCK, 56>(stage);
    add_occprobe<OP, UNROLL, BLOCK, 72>(stage);
    add_occprobe<OP, UNROLL, BLOCK, 100>(stage);
    add_occprobe<OP, UNROLL, BLOCK, 140>(stage);
    add_occprobe<OP, UNROLL, BLOCK, 200>(stage);
}

// checksum reduction variants: vecb column = 16, unroll = loads in flight, ld column = 1 if tiled, ctas_per_sm modes apply
static unsigned int* g_ck_counter = nullptr;
static uint64_t* g_ck_partials = nullptr;
static uint64_t* g_ck_out = nullptr;
template <int UNROLL, bool TILED> static void add_cksum(const char* stage)
{
    Config c;
    c.stage = stage;
    c.op = OP_COPY;   // 4 bytes read per word: reported through the 8 B/word copy accounting, halved below
    c.vecb = 16; c.unroll = UNROLL; c.block = kChecksumBlock; c.ldh = TILED ? 1 : 0; c.sth = 0;
    c.tile_elems = (size_t)kChecksumBlock * UNROLL * 4;
    auto k = checksum_kernel<UNROLL, TILED>;
    cudaFuncAttributes fa;
    CK(cudaFuncGetAttributes(&fa, k));
    c.regs = fa.numRegs;
    c.smem = 0;
    c.launch = [k](unsigned grid, void*, const void* x, const void*, float, size_t n) {
        if (grid > 2048) grid = 2048;
        CK(cudaMemsetAsync(g_ck_counter, 0, 16));
        k<<<grid, kChecksumBlock>>>((const unsigned char*)x, n * 4, 0, g_ck_counter, g_ck_partials, g_ck_out);
    };
    g_cfg.push_back(c);
}

// the same register kernel launched with programmatic stream serialization (PDL); ld column = 9
template <int OP, int VECB, int UNROLL, int BLOCK> static void add_pdl(const char* stage)
{
    Config c;
    c.stage = stage;
    c.op = OP; c.vecb = VECB; c.unroll = UNROLL; c.block = BLOCK; c.ldh = 9; c.sth = 0;
    c.tile_elems = (size_t)BLOCK * UNROLL * (VECB / 4);
    auto k = stream_kernel<OP, false, VECB, UNROLL, BLOCK, LD_DEFAULT, ST_DEFAULT, true>;
    cudaFuncAttributes fa;
    CK(cudaFuncGetAttributes(&fa, k));
    c.regs = fa.numRegs;
    c.launch = [k](unsigned grid, void* d, const void* x, const void* y, float s, size_t n) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(grid);
        cfg.blockDim = dim3(BLOCK);
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at;
        cfg.numAttrs = 1;
        CK(cudaLaunchKernelEx(&cfg, k, d, x, y, s, n));
    };
    g_cfg.push_back(c);
}

template <int OP, int TILE_BYTES, int STAGES, int BLOCK> static void add_tma(const char* stage)
{
    Config c;
    c.stage = stage;
    c.op = OP; c.vecb = TILE_BYTES / 1024; c.unroll = STAGES; c.block = BLOCK; c.ldh = -1; c.sth = -1;
    c.tile_elems = TILE_BYTES / 4;
    c.smem = STAGES * TILE_BYTES * (kReadsY<OP> ? 2 : 1);
    auto k = stream_kernel_tma<OP, false, TILE_BYTES, STAGES, BLOCK>;
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, c.smem));
    cudaFuncAttributes fa;
    CK(cudaFuncGetAttributes(&fa, k));
    c.regs = fa.numRegs;
    const int smem = c.smem;
    c.launch = [k, smem](unsigned grid, void* d, const void* x, const void* y, float s, size_t n) {
        k<<<grid, BLOCK, smem>>>(d, x, y, s, n / (TILE_BYTES / 4));
    };
    g_cfg.push_back(c);
}

// one tile per CTA (grid = tiles), one shared-memory slot: the register kernel's structure, staged by TMA
template <int OP, int TILE_BYTES, int BLOCK> static void add_tma_oneshot(const char* stage)
{
    Config c;
    c.stage = stage;
    c.op = OP; c.vecb = TILE_BYTES / 1024; c.unroll = 1; c.block = BLOCK; c.ldh = -2; c.sth = -2;
    c.tile_elems = TILE_BYTES / 4;
    c.smem = -TILE_BYTES * (kReadsY<OP> ? 2 : 1);      // negative: marks the one-shot variant (only grid = tiles is run)
    auto k = stream_kernel_tma<OP, false, TILE_BYTES, 2, BLOCK>;
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, -c.smem));
    cudaFuncAttributes fa;
    CK(cudaFuncGetAttributes(&fa, k));
    c.regs = fa.numRegs;
    const int smem = -c.smem;
    c.launch = [k, smem](unsigned grid, void* d, const void* x, const void* y, float s, size_t n) {
        k<<<grid, BLOCK, smem>>>(d, x, y, s, n / (TILE_BYTES / 4));
    };
    g_cfg.push_back(c);
}

template <int OP> static void add_tma_oneshot_all(const char* stage)
{
    add_tma_oneshot<OP, 4096, 128>(stage);
    add_tma_oneshot<OP, 4096, 256>(stage);
    add_tma_oneshot<OP, 8192, 128>(stage);
    add_tma_oneshot<OP, 8192, 256>(stage);
    add_tma_oneshot<OP, 8192, 512>(stage);
    add_tma_oneshot<OP, 16384, 256>(stage);
    add_tma_oneshot<OP, 16384, 512>(stage);
    add_tma_oneshot<OP, 16384, 1024>(stage);
    add_tma_oneshot<OP, 32768, 256>(stage);
    add_tma_oneshot<OP, 32768, 512>(stage);
    add_tma_oneshot<OP, 32768, 1024>(stage);
    add_tma_oneshot<OP, 65536, 512>(stage);
    add_tma_oneshot<OP, 65536, 1024>(stage);
}

template <int OP> static void add_tma_all(const char* stage)
{
    add_tma<OP, 8192, 2, 256>(stage);
    add_tma<OP, 8192, 3, 256>(stage);
    add_tma<OP, 8192, 4, 256>(stage);
    add_tma<OP, 8192, 6, 256>(stage);
    add_tma<OP, 16384, 2, 256>(stage);
    add_tma<OP, 16384, 3, 256>(stage);
    add_tma<OP, 16384, 4, 256>(stage);
    add_tma<OP, 16384, 3, 512>(stage);
    add_tma<OP, 16384, 4, 512>(stage);
    add_tma<OP, 32768, 2, 512>(stage);
    add_tma<OP, 32768, 3, 512>(stage);
    add_tma<OP, 32768, 3, 1024>(stage);
    add_tma<OP, 4096, 4, 128>(stage);
    add_tma<OP, 4096, 8, 128>(stage);
}

template <int OP, int LDH, int STH> static void add_shapes(const char* stage)
{
    add_cfg<OP, 16, 1, 256, LDH, STH>(stage);
    add_cfg<OP, 16, 2, 256, LDH, STH>(stage);
    add_cfg<OP, 16, 4, 256, LDH, STH>(stage);
    add_cfg<OP, 16, 8, 256, LDH, STH>(stage);
    add_cfg<OP, 16, 4, 512, LDH, STH>(stage);
    add_cfg<OP, 16, 4, 1024, LDH, STH>(stage);
    add_cfg<OP, 32, 1, 128, LDH, STH>(stage);
    add_cfg<OP, 32, 2, 128, LDH, STH>(stage);
    add_cfg<OP, 32, 4, 128, LDH, STH>(stage);
    add_cfg<OP, 32, 1, 256, LDH, STH>(stage);
    add_cfg<OP, 32, 2, 256, LDH, STH>(stage);
    add_cfg<OP, 32, 4, 256, LDH, STH>(stage);
    add_cfg<OP, 32, 8, 256, LDH, STH>(stage);
    add_cfg<OP, 32, 1, 512, LDH, STH>(stage);
    add_cfg<OP, 32, 2, 512, LDH, STH>(stage);
    add_cfg<OP, 32, 4, 512, LDH, STH>(stage);
    add_cfg<OP, 32, 1, 1024, LDH, STH>(stage);
    add_cfg<OP, 32, 2, 1024, LDH, STH>(stage);
}

template <int OP, int VECB, int UNROLL, int BLOCK> static void add_hints(const char* stage)
{
    add_cfg<OP, VECB, UNROLL, BLOCK, LD_DEFAULT, ST_CS>(stage);
    add_cfg<OP, VECB, UNROLL, BLOCK, LD_DEFAULT, ST_NA_EF>(stage);
    add_cfg<OP, VECB, UNROLL, BLOCK, LD_NC_NA, ST_CS>(stage);
    add_cfg<OP, VECB, UNROLL, BLOCK, LD_NC_NA, ST_NA_EF>(stage);
    add_cfg<OP, VECB, UNROLL, BLOCK, LD_NC_NA_EF, ST_DEFAULT>(stage);
    add_cfg<OP, VECB, UNROLL, BLOCK, LD_NC_NA_EF, ST_CS>(stage);
    add_cfg<OP, VECB, UNROLL, BLOCK, LD_NC_NA_EF, ST_NA_EF>(stage);
    add_cfg<OP, VECB, UNROLL, BLOCK, LD_CS, ST_DEFAULT>(stage);
    add_cfg<OP, VECB, UNROLL, BLOCK, LD_CS, ST_CS>(stage);
    add_cfg<OP, VECB, UNROLL, BLOCK, LD_NA, ST_DEFAULT>(stage);
    add_cfg<OP, VECB, UNROLL, BLOCK, LD_NA, ST_CS>(stage);
    add_cfg<OP, VECB, UNROLL, BLOCK, LD_NA, ST_NA_EF>(stage);
}

int main(int argc, char** argv)
{
    const int log2n = argc > 1 ? atoi(argv[1]) : 28;
    const int iters = argc > 2 ? atoi(argv[2]) : 10;
    const char* stage = argc > 3 ? argv[3] : "all";
    const size_t n = (size_t)1 << log2n;

    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    const int sms = prop.multiProcessorCount;
    fprintf(stderr, "# %s, %d SMs, L2 %d MiB, n=2^%d fp32\n", prop.name, sms, prop.l2CacheSize >> 20, log2n);

    float *a, *b, *c;
    CK(cudaMalloc(&a, n * 4));
    CK(cudaMalloc(&b, n * 4));
    CK(cudaMalloc(&c, n * 4));
    init_values<<<2048, 256>>>(a, n, 0x1234567u);
    init_values<<<2048, 256>>>(b, n, 0x89abcdeu);
    CK(cudaDeviceSynchronize());
    CK(cudaMemset(c, 0, n * 4));

    const bool all = !strcmp(stage, "all");
    if (all || !strcmp(stage, "shape")) {
        add_shapes<OP_ADD, LD_NC_NA, ST_DEFAULT>("shape");
        add_shapes<OP_ADD, LD_DEFAULT, ST_DEFAULT>("shape");
        add_shapes<OP_TRIAD, LD_NC_NA, ST_DEFAULT>("shape");
        add_shapes<OP_TRIAD, LD_DEFAULT, ST_DEFAULT>("shape");
        add_shapes<OP_COPY, LD_NC_NA, ST_DEFAULT>("shape");
        add_shapes<OP_COPY, LD_DEFAULT, ST_DEFAULT>("shape");
    }
    if (!strcmp(stage, "persist")) {   // the shipped add shape alone, grid = tiles and 1/2/4 CTAs per SM: for an ncu side-by-side
        add_cfg<OP_ADD, 32, 4, 512, LD_DEFAULT, ST_DEFAULT>("persist");
    }
    if (!strcmp(stage, "mechanism")) {   // ld column: 0 = product kernel, 10 = plain persistent, 11 = + CTA barrier per tile, 12 = dynamic tile fetch, 13 = contiguous range per CTA
        add_cfg<OP_ADD, 32, 4, 512, LD_DEFAULT, ST_DEFAULT>("mechanism");
        add_mechanism<OP_ADD, 4, 512, 0>("mechanism");
        add_mechanism<OP_ADD, 4, 512, 1>("mechanism");
        add_mechanism<OP_ADD, 4, 512, 2>("mechanism");
        add_mechanism<OP_ADD, 4, 512, 3>("mechanism");
        add_cfg<OP_ADD, 32, 1, 256, LD_DEFAULT, ST_DEFAULT>("mechanism");
        add_mechanism<OP_ADD, 1, 256, 0>("mechanism");
        add_mechanism<OP_ADD, 1, 256, 1>("mechanism");
        add_mechanism<OP_ADD, 1, 256, 2>("mechanism");
        add_mechanism<OP_ADD, 1, 256, 3>("mechanism");
    }
    if (!strcmp(stage, "scale")) {
        add_shapes<OP_SCALE, LD_NC_NA, ST_DEFAULT>("scale");
        add_shapes<OP_SCALE, LD_DEFAULT, ST_DEFAULT>("scale");
        add_shapes<OP_COPY, LD_DEFAULT, ST_DEFAULT>("scale");
    }
    if (!strcmp(stage, "bf16")) {   // rows are bf16 kernels over the same bytes; fp32 rows of the shipped shapes for comparison
        add_bf16_shapes<OP_ADD>("bf16");
        add_bf16_shapes<OP_TRIAD>("bf16");
        add_bf16_shapes<OP_SCALE>("bf16");
        add_cfg<OP_ADD, 32, 4, 512, LD_DEFAULT, ST_DEFAULT>("f32ref");
        add_cfg<OP_TRIAD, 32, 4, 512, LD_DEFAULT, ST_DEFAULT>("f32ref");
        add_cfg<OP_SCALE, 32, 1, 1024, LD_DEFAULT, ST_DEFAULT>("f32ref");
    }
    if (!strcmp(stage, "recheck")) {   // the shipped shapes against their neighbours, three interleaved rounds (box noise)
        for (int round = 0; round < 3; ++round) {
            add_cfg<OP_ADD, 32, 4, 512, LD_DEFAULT, ST_DEFAULT>("recheck");
            add_cfg<OP_ADD, 32, 1, 256, LD_DEFAULT, ST_DEFAULT>("recheck");
            add_cfg<OP_ADD, 32, 2, 256, LD_DEFAULT, ST_DEFAULT>("recheck");
            add_cfg<OP_ADD, 32, 1, 128, LD_DEFAULT, ST_DEFAULT>("recheck");
            add_cfg<OP_ADD, 32, 1, 512, LD_DEFAULT, ST_DEFAULT>("recheck");
            add_cfg<OP_ADD, 32, 2, 512, LD_DEFAULT, ST_DEFAULT>("recheck");
            add_probe<OP_ADD, 1, 256, false>("probe");
            add_probe<OP_ADD, 4, 512, false>("probe");
            add_pdl<OP_ADD, 32, 4, 512>("recheck-pdl");
            add_pdl<OP_ADD, 32, 1, 256>("recheck-pdl");
        }
    }
    if (!strcmp(stage, "occ")) {
        for (int ctas : {0, 8, 6, 5, 4, 3, 2, 1}) {
            add_occ<OP_ADD, 32, 1, 256>("occ", ctas);
            add_occ<OP_ADD, 32, 2, 256>("occ", ctas);
            add_occ<OP_ADD, 32, 1, 128>("occ", ctas);
            add_occ<OP_ADD, 32, 1, 512>("occ", ctas);
            add_occ<OP_ADD, 32, 2, 512>("occ", ctas);
            add_occ<OP_ADD, 32, 4, 256>("occ", ctas);
            add_occ<OP_ADD, 32, 2, 128>("occ", ctas);
            add_occ<OP_ADD, 32, 4, 128>("occ", ctas);
            add_occ<OP_COPY, 32, 1, 256>("occ", ctas);
            add_occ<OP_COPY, 32, 2, 256>("occ", ctas);
            add_occ<OP_COPY, 32, 4, 256>("occ", ctas);
        }
        add_cfg<OP_ADD, 32, 4, 512, LD_DEFAULT, ST_DEFAULT>("occref");
        add_probe<OP_ADD, 1, 256, false>("probe");
    }
    if (!strcmp(stage, "carve")) {
        for (int pct : {-1, 0, 25, 50, 75, 90, 100}) {
            add_carve<OP_ADD, 32, 4, 512>("carve", pct);
            add_carve<OP_ADD, 32, 1, 256>("carve", pct);
            add_carve<OP_ADD, 32, 2, 512>("carve", pct);
            add_carve<OP_COPY, 32, 1, 256>("carve", pct);
            add_carve<OP_SCALE, 32, 1, 1024>("carve", pct);
        }
        add_probe<OP_ADD, 1, 256, false>("probe");
    }
    if (!strcmp(stage, "occprobe")) {
        add_occprobe_all<OP_ADD, 1, 256>("occprobe");
        add_occprobe_all<OP_ADD, 1, 128>("occprobe");
        add_occprobe_all<OP_ADD, 2, 256>("occprobe");
        add_occprobe_all<OP_ADD, 2, 128>("occprobe");
        add_occprobe_all<OP_ADD, 4, 128>("occprobe");
        add_occprobe_all<OP_ADD, 1, 512>("occprobe");
        add_occprobe_all<OP_COPY, 1, 256>("occprobe");
        add_occprobe_all<OP_COPY, 2, 256>("occprobe");
    }
    if (!strcmp(stage, "cksum")) {   // GB/s columns count 8 B per word (copy accounting): halve them for the 4 B actually read
        CK(cudaMalloc(&g_ck_counter, 256));
        CK(cudaMalloc(&g_ck_partials, 2048 * 16));
        CK(cudaMalloc(&g_ck_out, 16));
        add_cksum<4, false>("cksum");
        add_cksum<8, false>("cksum");
        add_cksum<2, true>("cksum");
        add_cksum<4, true>("cksum");
        add_cksum<8, true>("cksum");
        add_cksum<16, true>("cksum");
    }
    if (!strcmp(stage, "lockstep")) {
        add_probe<OP_ADD, 4, 512, false>("lockstep");
        add_probe<OP_ADD, 4, 512, true>("lockstep");
        add_probe<OP_ADD, 2, 256, false>("lockstep");
        add_probe<OP_ADD, 2, 256, true>("lockstep");
        add_probe<OP_ADD, 1, 256, false>("lockstep");
        add_probe<OP_ADD, 1, 256, true>("lockstep");
    }
    if (!strcmp(stage, "pdl")) {
        add_cfg<OP_ADD, 32, 4, 512, LD_DEFAULT, ST_DEFAULT>("pdl");
        add_pdl<OP_ADD, 32, 4, 512>("pdl");
        add_cfg<OP_ADD, 32, 1, 256, LD_DEFAULT, ST_DEFAULT>("pdl");
        add_pdl<OP_ADD, 32, 1, 256>("pdl");
        add_cfg<OP_ADD, 32, 2, 256, LD_DEFAULT, ST_DEFAULT>("pdl");
        add_pdl<OP_ADD, 32, 2, 256>("pdl");
        add_cfg<OP_COPY, 32, 1, 256, LD_DEFAULT, ST_DEFAULT>("pdl");
        add_pdl<OP_COPY, 32, 1, 256>("pdl");
        add_cfg<OP_COPY, 32, 2, 256, LD_DEFAULT, ST_DEFAULT>("pdl");
        add_pdl<OP_COPY, 32, 2, 256>("pdl");
    }
    if (!strcmp(stage, "small")) {   // sizes where the grid is only a few waves: which tile shape fills the SMs best
        add_shapes<OP_ADD, LD_DEFAULT, ST_DEFAULT>("small");
        add_shapes<OP_COPY, LD_DEFAULT, ST_DEFAULT>("small");
    }
    if (all || !strcmp(stage, "hint")) {
        add_hints<OP_ADD, 32, 4, 256>("hint");
        add_hints<OP_ADD, 32, 2, 512>("hint");
        add_hints<OP_ADD, 16, 4, 256>("hint");
        add_hints<OP_TRIAD, 32, 4, 256>("hint");
        add_hints<OP_COPY, 32, 4, 256>("hint");
        add_hints<OP_COPY, 32, 8, 256>("hint");
    }

    if (!strcmp(stage, "tma1")) {
        add_tma_oneshot_all<OP_ADD>("tma1");
        add_tma_oneshot_all<OP_TRIAD>("tma1");
        add_tma_oneshot_all<OP_COPY>("tma1");
        add_cfg<OP_ADD, 32, 4, 512, LD_DEFAULT, ST_DEFAULT>("regref");
        add_cfg<OP_COPY, 32, 1, 256, LD_DEFAULT, ST_DEFAULT>("regref");
    }
    if (all || !strcmp(stage, "tma")) {
        add_tma_all<OP_ADD>("tma");
        add_tma_all<OP_TRIAD>("tma");
        add_tma_all<OP_COPY>("tma");
    }
    float* cref = nullptr;
    unsigned long long* dbad = nullptr;
    CK(cudaMalloc(&dbad, 8));

    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));

    printf("stage,op,vecb,unroll,block,ld,st,regs,ctas_per_sm,grid,best_ms,mean_ms,best_gbs,mean_gbs\n");

    // reference points: the driver's own device-to-device copy and memset
    if (all || !strcmp(stage, "ref")) {
        for (int which = 0; which < 2; ++which) {
            float best = 1e30f, tot = 0;
            for (int i = 0; i < iters + 2; ++i) {
                CK(cudaEventRecord(e0));
                if (which == 0) CK(cudaMemcpyAsync(c, a, n * 4, cudaMemcpyDeviceToDevice));
                else CK(cudaMemsetAsync(c, 1, n * 4));
                CK(cudaEventRecord(e1));
                CK(cudaEventSynchronize(e1));
                float ms;
                CK(cudaEventElapsedTime(&ms, e0, e1));
                if (i >= 2) { best = std::min(best, ms); tot += ms; }
            }
            const double bytes = which == 0 ? 8.0 * n : 4.0 * n;
            printf("ref,%s,0,0,0,0,0,0,0,0,%.4f,%.4f,%.1f,%.1f\n", which == 0 ? "cudaMemcpyD2D" : "cudaMemset", best,
                   tot / iters, bytes / best * 1e-6, bytes / (tot / iters) * 1e-6);
        }
    }

    const int ctas_modes[] = {0, 1, 2, 4, 8};
    for (const Config& cf : g_cfg) {
        const double bytes = (cf.op == OP_COPY || cf.op == OP_SCALE ? 8.0 : 12.0) * n;
        if (cf.smem != 0) {   // correctness of the TMA variant against the register kernel, once per config
            if (!cref) CK(cudaMalloc(&cref, n * 4));
            CK(cudaMemset(c, 0, n * 4));
            CK(cudaMemset(dbad, 0, 8));
            cf.launch(cf.smem < 0 ? (unsigned)(n / cf.tile_elems) : (unsigned)sms, c, a, b, 3.0f, n);
            if (cf.op == OP_ADD) stream_kernel<OP_ADD, false, 32, 4, 512, 0, 0><<<(unsigned)(n / 16384), 512>>>(cref, a, b, 3.0f, n);
            else if (cf.op == OP_TRIAD) stream_kernel<OP_TRIAD, false, 32, 4, 512, 0, 0><<<(unsigned)(n / 16384), 512>>>(cref, a, b, 3.0f, n);
            else stream_kernel<OP_COPY, false, 32, 4, 512, 0, 0><<<(unsigned)(n / 16384), 512>>>(cref, a, b, 3.0f, n);
            count_mismatch<<<1024, 256>>>((const uint32_t*)c, (const uint32_t*)cref, n, dbad);
            unsigned long long bad = 0;
            CK(cudaMemcpy(&bad, dbad, 8, cudaMemcpyDeviceToHost));
            if (bad) {
                fprintf(stderr, "TMA variant tile=%dK stages=%d block=%d op=%d: %llu mismatches\n", cf.vecb, cf.unroll, cf.block, cf.op, bad);
                return 3;
            }
        }
        for (int cps : ctas_modes) {
            size_t tiles = n / cf.tile_elems;
            if (tiles == 0) continue;
            size_t grid = tiles;
            if (cps > 0) grid = std::min(tiles, (size_t)sms * cps);
            if (cps * cf.block > 2048) continue;
            if (cf.smem > 0 && (cps == 0 || (size_t)cps * (cf.smem + 1024) > 227 * 1024)) continue;
            if (cf.smem < 0 && cps != 0) continue;
            for (int i = 0; i < 2; ++i) cf.launch((unsigned)grid, c, a, b, 3.0f, n);
            CK(cudaGetLastError());
            float best = 1e30f;
            for (int i = 0; i < 3; ++i) {
                CK(cudaEventRecord(e0));
                cf.launch((unsigned)grid, c, a, b, 3.0f, n);
                CK(cudaEventRecord(e1));
                CK(cudaEventSynchronize(e1));
                float ms;
                CK(cudaEventElapsedTime(&ms, e0, e1));
                best = std::min(best, ms);
            }
            CK(cudaEventRecord(e0));
            for (int i = 0; i < iters; ++i) cf.launch((unsigned)grid, c, a, b, 3.0f, n);
            CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1));
            float tot;
            CK(cudaEventElapsedTime(&tot, e0, e1));
            const float mean = tot / iters;
            static const char* opn[] = {"add", "copy", "scale", "triad"};
            printf("%s,%s,%d,%d,%d,%d,%d,%d,%d,%zu,%.4f,%.4f,%.1f,%.1f\n", cf.stage.c_str(), opn[cf.op], cf.vecb,
                   cf.unroll, cf.block, cf.ldh, cf.sth, cf.regs, cps, grid, best, mean, bytes / best * 1e-6,
                   bytes / mean * 1e-6);
        }
        fflush(stdout);
    }
    return 0;
}

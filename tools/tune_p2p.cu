// tune_p2p.cu — developer tool: which kernel shape moves a buffer to a peer GPU fastest over NVLink?
//
//   tools/tune_p2p [csv]        needs >= 2 GPUs; uses GPU 0 and GPU 1
//
// For every variant: one direction alone (0 -> 1) and both directions at once (0 -> 1 while 1 -> 0),
// 1 GiB per transfer, device-timed, best of 5; the destination is compared word for word with the source.
// Variants: cudaMemcpyPeerAsync (the copy engine), register kernels that LOAD from the peer (pull) or
// STORE to the peer (push) with 128/256-bit accesses in several launch shapes and store hints, and a
// bulk-copy (TMA, cp.async.bulk) push: local HBM -> shared memory ring -> peer memory, one elected thread
// per CTA issuing both copies.  Output feeds csrc/kernels.cu's peer_push_kernel (profiles/r02_tune_p2p*.csv).
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "stream_kernels.cuh"
#include "stream_tma.cuh"

using namespace b200v;

#define CK(call)                                                                               \
    do {                                                                                       \
        cudaError_t e__ = (call);                                                              \
        if (e__ != cudaSuccess) {                                                              \
            fprintf(stderr, "%s failed: %s (%s:%d)\n", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
            exit(2);                                                                           \
        }                                                                                      \
    } while (0)

// ---- register kernels ------------------------------------------------------------------------
template <int VECB, int UNROLL, int BLOCK, int STH>
__global__ void __launch_bounds__(BLOCK) move_kernel(Vec<VECB>* __restrict__ dst, const Vec<VECB>* __restrict__ src, size_t nvec)
{
    constexpr size_t TILE = (size_t)BLOCK * UNROLL;
    const size_t tiles = nvec / TILE;
    for (size_t t = blockIdx.x; t < tiles; t += gridDim.x) {
        const size_t base = t * TILE + threadIdx.x;
        Vec<VECB> v[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) v[u] = ldg<LD_DEFAULT>(src + base + (size_t)u * BLOCK);
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) stg<STH>(dst + base + (size_t)u * BLOCK, v[u]);
    }
}

// ---- bulk-copy push: persistent CTAs, a ring of STAGES slots of TILE_BYTES, one thread does everything ----
template <int TILE_BYTES, int STAGES>
__global__ void __launch_bounds__(32) tma_move_kernel(char* __restrict__ dst, const char* __restrict__ src, size_t tiles)
{
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ uint64_t full[STAGES];
    if (threadIdx.x != 0) return;
    for (int i = 0; i < STAGES; ++i) mbar_init(&full[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    const size_t mine = tiles > blockIdx.x ? (tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    auto issue_load = [&](size_t k) {
        const int slot = (int)(k % STAGES);
        const size_t off = (blockIdx.x + k * gridDim.x) * (size_t)TILE_BYTES;
        mbar_expect_tx(&full[slot], TILE_BYTES);
        bulk_load(smem + slot * TILE_BYTES, src + off, TILE_BYTES, &full[slot]);
    };
    for (size_t k = 0; k < (size_t)(STAGES - 1) && k < mine; ++k) issue_load(k);
    for (size_t k = 0; k < mine; ++k) {
        const int slot = (int)(k % STAGES);
        mbar_wait(&full[slot], (uint32_t)((k / STAGES) & 1));
        const size_t off = (blockIdx.x + k * gridDim.x) * (size_t)TILE_BYTES;
        bulk_store(dst + off, smem + slot * TILE_BYTES, TILE_BYTES);
        bulk_commit();
        if (k + STAGES - 1 < mine) {
            bulk_wait_read<1>();      // the store of tile k-1 has finished reading its slot
            issue_load(k + STAGES - 1);
        }
    }
    bulk_wait_read<0>();
}

__global__ void compare_kernel(const uint32_t* a, const uint32_t* b, size_t n, unsigned long long* bad)
{
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    unsigned long long mine = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) mine += a[i] != b[i];
    if (mine) atomicAdd(bad, mine);
}

__global__ void pattern_kernel(uint32_t* a, size_t n, uint32_t salt)
{
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) a[i] = (uint32_t)i * 2654435761u ^ salt;
}

// ---------------------------------------------------------------------------------------------------
struct Variant {
    std::string name;
    // launches the transfer dst <- src on the CURRENT device's stream st
    void (*launch)(void* dst, const void* src, size_t bytes, int sms, int arg, cudaStream_t st);
    int arg;          // grid multiplier (CTAs per SM; 0 = one tile per CTA)
    bool pull;        // the DESTINATION device runs the kernel
    bool copy_engine;
};

template <int VECB, int UNROLL, int BLOCK, int STH>
static void launch_move(void* dst, const void* src, size_t bytes, int sms, int arg, cudaStream_t st)
{
    const size_t nvec = bytes / VECB;
    size_t grid = nvec / ((size_t)BLOCK * UNROLL);
    if (arg > 0) grid = std::min(grid, (size_t)sms * arg);
    move_kernel<VECB, UNROLL, BLOCK, STH><<<(unsigned)grid, BLOCK, 0, st>>>((Vec<VECB>*)dst, (const Vec<VECB>*)src, nvec);
}

template <int TILE_BYTES, int STAGES>
static void launch_tma(void* dst, const void* src, size_t bytes, int sms, int arg, cudaStream_t st)
{
    const size_t tiles = bytes / TILE_BYTES;
    static bool attr_set[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (!attr_set[dev]) {
        CK(cudaFuncSetAttribute(tma_move_kernel<TILE_BYTES, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, TILE_BYTES * STAGES));
        attr_set[dev] = true;
    }
    const size_t grid = std::min(tiles, (size_t)sms * arg);
    tma_move_kernel<TILE_BYTES, STAGES><<<(unsigned)grid, 32, TILE_BYTES * STAGES, st>>>((char*)dst, (const char*)src, tiles);
}

static void launch_copy_engine(void*, const void*, size_t, int, int, cudaStream_t) {}

static std::vector<Variant> g_v;

template <int VECB, int UNROLL, int BLOCK, int STH> static void add_push(const char* hint, std::initializer_list<int> grids)
{
    for (int g : grids) {
        char nm[128];
        snprintf(nm, sizeof nm, "push reg v%d u%d b%d st=%s grid=%s", VECB * 8, UNROLL, BLOCK, hint, g ? (std::to_string(g) + "/SM").c_str() : "tiles");
        g_v.push_back({nm, launch_move<VECB, UNROLL, BLOCK, STH>, g, false, false});
    }
}
template <int VECB, int UNROLL, int BLOCK> static void add_pull(std::initializer_list<int> grids)
{
    for (int g : grids) {
        char nm[128];
        snprintf(nm, sizeof nm, "pull reg v%d u%d b%d grid=%s", VECB * 8, UNROLL, BLOCK, g ? (std::to_string(g) + "/SM").c_str() : "tiles");
        g_v.push_back({nm, launch_move<VECB, UNROLL, BLOCK, ST_DEFAULT>, g, true, false});
    }
}
template <int TILE_BYTES, int STAGES> static void add_tma(std::initializer_list<int> grids)
{
    for (int g : grids) {
        if ((size_t)TILE_BYTES * STAGES * g > 200 * 1024) continue;     // shared memory per SM
        char nm[128];
        snprintf(nm, sizeof nm, "push tma tile=%dK stages=%d grid=%d/SM", TILE_BYTES / 1024, STAGES, g);
        g_v.push_back({nm, launch_tma<TILE_BYTES, STAGES>, g, false, false});
    }
}

int main(int argc, char** argv)
{
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    if (ndev < 2) {
        printf("tune_p2p: needs 2 GPUs, %d visible\n", ndev);
        return 0;
    }
    const size_t bytes = (size_t)1 << 30, n = bytes / 4;
    void *src[2], *dst[2];
    unsigned long long* bad[2];
    cudaStream_t st[2];
    cudaEvent_t e0[2], e1[2];
    int sms = 0;
    for (int d = 0; d < 2; ++d) {
        CK(cudaSetDevice(d));
        CK(cudaDeviceEnablePeerAccess(1 - d, 0));
        CK(cudaMalloc(&src[d], bytes));
        CK(cudaMalloc(&dst[d], bytes));
        CK(cudaMalloc(&bad[d], 8));
        CK(cudaStreamCreateWithFlags(&st[d], cudaStreamNonBlocking));
        CK(cudaEventCreate(&e0[d]));
        CK(cudaEventCreate(&e1[d]));
        pattern_kernel<<<1024, 256>>>((uint32_t*)src[d], n, 0xA5A50000u + d);
        CK(cudaDeviceSynchronize());
        CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, d));
    }

    g_v.push_back({"cudaMemcpyPeerAsync (copy engine)", launch_copy_engine, 0, false, true});
    add_pull<32, 4, 512>({4});
    add_pull<32, 4, 512>({0, 1, 2, 8});
    add_pull<32, 8, 256>({2, 4});
    add_pull<16, 4, 512>({4});
    add_push<32, 4, 512, ST_DEFAULT>("default", {4});      // the round-2 first shape
    add_push<32, 4, 512, ST_DEFAULT>("default", {0, 1, 2, 8});
    add_push<32, 1, 256, ST_DEFAULT>("default", {0, 4, 8, 16});
    add_push<32, 2, 256, ST_DEFAULT>("default", {0, 4, 8});
    add_push<32, 8, 256, ST_DEFAULT>("default", {2, 4});
    add_push<32, 1, 1024, ST_DEFAULT>("default", {0, 1, 2});
    add_push<32, 4, 512, ST_CS>("cs", {0, 2, 4});
    add_push<32, 4, 512, ST_NA_EF>("na_ef", {0, 2, 4});
    add_push<16, 4, 512, ST_DEFAULT>("default", {0, 2, 4});
    add_push<16, 8, 256, ST_DEFAULT>("default", {0, 4});
    add_push<16, 4, 512, ST_CS>("cs", {4});
    add_tma<4096, 4>({1, 2, 4, 8});
    add_tma<8192, 2>({1, 2, 4, 8});
    add_tma<8192, 4>({1, 2, 4});
    add_tma<16384, 2>({1, 2, 4});
    add_tma<16384, 4>({1, 2});
    add_tma<32768, 2>({1, 2});
    add_tma<32768, 3>({1, 2});
    add_tma<65536, 2>({1});
    add_tma<65536, 3>({1});

    FILE* csv = argc > 1 ? fopen(argv[1], "w") : nullptr;
    if (csv) fprintf(csv, "variant,one_way_gbs,both_ways_gbs_per_direction,verified\n");
    printf("%-52s %12s %22s  verify\n", "variant (1 GiB, GPU0 <-> GPU1)", "one way GB/s", "both ways GB/s per dir");

    for (const Variant& v : g_v) {
        // transfer "dir": source GPU dir, destination GPU 1-dir; the issuing GPU is the destination for a pull
        auto issue = [&](int dir) {
            const int s = dir, d = 1 - dir, owner = v.pull ? d : s;
            CK(cudaSetDevice(owner));
            CK(cudaEventRecord(e0[owner], st[owner]));
            if (v.copy_engine) CK(cudaMemcpyPeerAsync(dst[d], d, src[s], s, bytes, st[owner]));
            else v.launch(dst[d], src[s], bytes, sms, v.arg, st[owner]);
            CK(cudaGetLastError());
            CK(cudaEventRecord(e1[owner], st[owner]));
            return owner;
        };
        auto elapsed = [&](int owner) {
            CK(cudaSetDevice(owner));
            CK(cudaEventSynchronize(e1[owner]));
            float ms = 0;
            CK(cudaEventElapsedTime(&ms, e0[owner], e1[owner]));
            return (double)ms;
        };
        for (int d = 0; d < 2; ++d) {
            CK(cudaSetDevice(d));
            CK(cudaMemsetAsync(dst[d], 0xEE, bytes, st[d]));
            CK(cudaMemsetAsync(bad[d], 0, 8, st[d]));
            CK(cudaStreamSynchronize(st[d]));
        }
        double one = 1e30, both = 1e30;
        for (int rep = 0; rep < 6; ++rep) {
            const double ms = elapsed(issue(0));
            if (rep) one = std::min(one, ms);
        }
        for (int rep = 0; rep < 6; ++rep) {
            const int o0 = issue(0), o1 = issue(1);
            const double ms = std::max(elapsed(o0), elapsed(o1));
            if (rep) both = std::min(both, ms);
        }
        // verify both destinations: dst[1] must equal src[0], dst[0] must equal src[1]
        unsigned long long nbad = 0;
        for (int d = 0; d < 2; ++d) {
            CK(cudaSetDevice(d));
            compare_kernel<<<1024, 256, 0, st[d]>>>((const uint32_t*)dst[d], (const uint32_t*)src[1 - d], n, bad[d]);
            unsigned long long h = 0;
            CK(cudaMemcpyAsync(&h, bad[d], 8, cudaMemcpyDeviceToHost, st[d]));
            CK(cudaStreamSynchronize(st[d]));
            nbad += h;
        }
        const double g1 = bytes / one * 1e-6, g2 = bytes / both * 1e-6;
        printf("%-52s %12.1f %22.1f  %s\n", v.name.c_str(), g1, g2, nbad ? "MISMATCH" : "ok");
        fflush(stdout);
        if (csv) fprintf(csv, "\"%s\",%.1f,%.1f,%d\n", v.name.c_str(), g1, g2, nbad ? 0 : 1);
    }
    if (csv) fclose(csv);
    return 0;
}

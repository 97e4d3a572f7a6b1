// tune.cu — sweep the launch shapes / cache hints of the streaming kernels on a B200 and print a
// CSV table (one row per configuration).  The product's csrc/tuning.h is set from these tables
// (kept under profiles/).  Not part of the product; builds against the same stream_kernels.cuh.
//
//   tools/tune [log2_n=28] [iters=10] [stage=all|shape|hint|ref]
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <string>
#include <vector>

#include "stream_kernels.cuh"

using namespace b200v;

#define CK(x)                                                                                   \
    do {                                                                                        \
        cudaError_t e_ = (x);                                                                   \
        if (e_ != cudaSuccess) {                                                                \
            fprintf(stderr, "%s: %s at %s:%d\n", #x, cudaGetErrorString(e_), __FILE__, __LINE__); \
            exit(2);                                                                            \
        }                                                                                       \
    } while (0)

struct Config {
    std::string stage;
    int op, vecb, unroll, block, ldh, sth;
    size_t tile_elems;
    std::function<void(unsigned, void*, const void*, const void*, float, size_t)> launch;
    int regs;
};

static std::vector<Config> g_cfg;

template <int OP, int VECB, int UNROLL, int BLOCK, int LDH, int STH>
static void add_cfg(const char* stage)
{
    Config c;
    c.stage = stage;
    c.op = OP; c.vecb = VECB; c.unroll = UNROLL; c.block = BLOCK; c.ldh = LDH; c.sth = STH;
    c.tile_elems = (size_t)BLOCK * UNROLL * (VECB / 4);
    auto k = stream_kernel<OP, false, VECB, UNROLL, BLOCK, LDH, STH>;
    cudaFuncAttributes fa;
    CK(cudaFuncGetAttributes(&fa, k));
    c.regs = fa.numRegs;
    c.launch = [k](unsigned grid, void* d, const void* x, const void* y, float s, size_t n) {
        k<<<grid, BLOCK>>>(d, x, y, s, n);
    };
    g_cfg.push_back(c);
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
    CK(cudaMemset(a, 0x3c, n * 4));
    CK(cudaMemset(b, 0x3d, n * 4));
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
    if (all || !strcmp(stage, "hint")) {
        add_hints<OP_ADD, 32, 4, 256>("hint");
        add_hints<OP_ADD, 32, 2, 512>("hint");
        add_hints<OP_ADD, 16, 4, 256>("hint");
        add_hints<OP_TRIAD, 32, 4, 256>("hint");
        add_hints<OP_COPY, 32, 4, 256>("hint");
        add_hints<OP_COPY, 32, 8, 256>("hint");
    }

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
        const double bytes = (cf.op == OP_COPY ? 8.0 : 12.0) * n;
        for (int cps : ctas_modes) {
            size_t tiles = n / cf.tile_elems;
            size_t grid = tiles;
            if (cps > 0) grid = std::min(tiles, (size_t)sms * cps);
            if (cps * cf.block > 2048) continue;
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

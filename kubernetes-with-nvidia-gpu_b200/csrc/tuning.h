// tuning.h — launch shapes of the streaming kernels, chosen by tools/tune.cu sweeps on a B200
// (profiles/ holds the sweep tables).  One place, compile-time, so the product and the tuner
// instantiate the same kernels.
#pragma once

#include "stream_kernels.cuh"

namespace b200v {

// profiles/r01_tune_sweep_2p28.csv (B200, N = 2^28 fp32, mean of 10 back-to-back launches):
//   add/triad  VECB=32 UNROLL=4 BLOCK=512 default ld/st, one tile per CTA : 7113 GB/s
//              same shape as a persistent grid of 1/2/4 CTAs per SM        : 6420 / 6591 / 6764
//              ld.global.nc variants: ptxas sinks the loads between the stores (40 regs instead
//              of 80) and lands ~1-2 % lower (6982)
//   copy       VECB=32 UNROLL=1 BLOCK=256, one tile per CTA                : 7033 GB/s
//   for scale: cudaMemcpy D2D 6450 GB/s, cudaMemset 7061 GB/s on the same box
// One tile per CTA wins everywhere: the block scheduler hands out tiles in address order, so the
// set of tiles in flight is one contiguous window of each array.
struct Tuning {
    // two-input kernels (add, triad): 2 x UNROLL x 32 B in flight per thread
    static constexpr int kBlock2 = 512;
    static constexpr int kUnroll2 = 4;
    // one-input kernels (copy, scale)
    static constexpr int kBlock1 = 256;
    static constexpr int kUnroll1 = 1;
    // plain (coherent) loads: program order is kept, every load of a tile is issued before the
    // first store
    static constexpr int kLd = LD_DEFAULT;
    static constexpr int kSt = ST_DEFAULT;
    // 0 = one tile per CTA (grid == tiles); k > 0 = persistent grid of k CTAs per SM
    static constexpr int kCtasPerSm = 0;
};

}  // namespace b200v

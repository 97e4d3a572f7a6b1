// tuning.h — launch shapes of the streaming kernels, chosen by tools/tune.cu sweeps on a B200
// (profiles/ holds the sweep tables).  One place, compile-time, so the product and the tuner
// instantiate the same kernels.
#pragma once

#include "stream_kernels.cuh"

namespace b200v {

struct Tuning {
    // two-input kernels (add, triad): 2 x UNROLL x 32 B in flight per thread
    static constexpr int kBlock2 = 256;
    static constexpr int kUnroll2 = 4;
    // one-input kernels (copy, scale)
    static constexpr int kBlock1 = 256;
    static constexpr int kUnroll1 = 8;
    static constexpr int kLd = LD_NC_NA;
    static constexpr int kSt = ST_DEFAULT;
    // 0 = one tile per CTA (grid == tiles); k > 0 = persistent grid of k CTAs per SM
    static constexpr int kCtasPerSm = 0;
};

}  // namespace b200v

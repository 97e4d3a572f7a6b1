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
//   scale      VECB=32 UNROLL=1 BLOCK=1024                                 : 7028 GB/s (BLOCK=256: 6877)
//   for scale: cudaMemcpy D2D 6450 GB/s, cudaMemset 7061 GB/s on the same box
// One tile per CTA wins everywhere: the block scheduler hands out tiles in address order, so the
// set of tiles in flight is one contiguous window of each array.
//
// profiles/r01_tune_small_sizes.csv (2^20 .. 2^27 elements): the 64 KiB-per-array tile leaves too
// few CTAs below 2^27 elements (one CTA per SM at 80 registers), so smaller tiles take over:
//   add, 2^25 elements: <1,256> 6716  <2,256> 6553  <4,512> 6429
//   add, 2^24 elements: <2,256> 6548  <1,256> 6438  <4,512> 5855
//   add, 2^22 elements: <2,256> 8177 (L2-resident)  <4,512> 6858
//   copy, 2^24 elements: <2,256> 6551  <1,256> 5954
struct Tuning {
    // plain (coherent) loads: program order is kept, every load of a tile is issued before the
    // first store
    static constexpr int kLd = LD_DEFAULT;
    static constexpr int kSt = ST_DEFAULT;
    // 0 = one tile per CTA (grid == tiles); k > 0 = persistent grid of k CTAs per SM
    static constexpr int kCtasPerSm = 0;
    // bytes per array at which the launch shape changes
    static constexpr size_t kLargeBytes = (size_t)512 << 20;
    static constexpr size_t kMediumBytes = (size_t)128 << 20;
};

// <UNROLL, BLOCK> of the 256-bit kernel by op class and size class
struct ShapeLarge2 { static constexpr int unroll = 4, block = 512; };    // fp32 add/triad, >= 512 MiB per array
struct ShapeMedium { static constexpr int unroll = 1, block = 256; };    // every op, 128..512 MiB; copy above
struct ShapeSmall { static constexpr int unroll = 2, block = 256; };     // every op, < 128 MiB
// scale, and bf16 add/triad (the unpack/convert code lets ptxas interleave the <4,512> loads anyway:
// 60 registers, 6999 GB/s; <1,1024> reaches 7088 at the same bytes), >= 512 MiB
struct ShapeLarge1024 { static constexpr int unroll = 1, block = 1024; };
// 128-bit fallback for pointers that are 16- but not 32-byte aligned
struct ShapeVec16Two { static constexpr int unroll = 4, block = 1024; };
struct ShapeVec16One { static constexpr int unroll = 2, block = 256; };

}  // namespace b200v

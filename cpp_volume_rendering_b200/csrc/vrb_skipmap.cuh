// Empty-space skip map (see k_skipmap.cu).
#pragma once
#include <cstdint>
#include <algorithm>
struct vrb_ctx;

namespace vrb {

struct SkipMap {
  static constexpr int MAX_DIST = 64;
  uint8_t* dist = nullptr;   // per block: 0 = may contribute; r >= 1 = the (2r-1)^3 cube of blocks around it is empty
  int B = 0, nbx = 0, nby = 0, nbz = 0;
  uint64_t vol_gen = 0, tf_gen = 0;
  int ensure(vrb_ctx* ctx, int block);   // (re)builds on the context stream when the volume, TF or block size changed
  void release();
};

}  // namespace vrb

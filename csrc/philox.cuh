// Philox4x32-10, the CUDA twin of utils/philox.py (same counter/key layout, same keep rule).
#pragma once
#include <stdint.h>

namespace dfm {

struct DropoutCfg {
  uint32_t threshold;  // keep iff u32 < threshold
  uint32_t enabled;    // 0 -> keep everything, scale 1
  float scale;         // 1 / keep_prob
  uint32_t seed_lo, seed_hi;
  uint32_t step_lo, step_hi;
  uint32_t layer;
  uint32_t row_offset;  // first global row of this rank's batch slice
};

__device__ __forceinline__ uint4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                               uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return make_uint4(c0, c1, c2, c3);
}

// Random words for the 4 consecutive elements [4*q, 4*q+4) of a layer's output (elem = row*width+col).
__device__ __forceinline__ uint4 dropout_words(const DropoutCfg& d, uint64_t quad) {
  return philox4x32_10((uint32_t)quad, d.layer, d.step_lo, d.step_hi, d.seed_lo, d.seed_hi);
}

}  // namespace dfm

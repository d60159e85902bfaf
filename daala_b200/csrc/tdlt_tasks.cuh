// The 4x4 leaf transforms of a lane that holds two stacked 4x4 blocks (an 8-row by
// 4-column strip of an 8x8 unit split into four 4x4 blocks).
#pragma once
#include "tdlt_device.cuh"

namespace tdlt {

// ---------------------------------------------------------------- 4x4 leaves
// v[8*q + r] = sample (row r, column q) of the 8x4 strip: two stacked blocks.
// Column DCTs first, then row DCTs (reference src/dct.c:137-142).
__device__ __forceinline__ void fdct4x4_pair(int *v) {
#pragma unroll
  for (int b = 0; b < 2; b++) {
#pragma unroll
    for (int q = 0; q < 4; q++) tdlt_fdct4(v + 8*q + 4*b);   // column q, rows 4b..4b+3
#pragma unroll
    for (int r = 0; r < 4; r++) {
      int t[4] = {v[4*b + r], v[8 + 4*b + r], v[16 + 4*b + r], v[24 + 4*b + r]};
      tdlt_fdct4(t);
      v[4*b + r] = t[0]; v[8 + 4*b + r] = t[1]; v[16 + 4*b + r] = t[2]; v[24 + 4*b + r] = t[3];
    }
  }
}
// Row inverses first, then column inverses (reference src/dct.c:144-149).
__device__ __forceinline__ void idct4x4_pair(int *v) {
#pragma unroll
  for (int b = 0; b < 2; b++) {
#pragma unroll
    for (int r = 0; r < 4; r++) {
      int t[4] = {v[4*b + r], v[8 + 4*b + r], v[16 + 4*b + r], v[24 + 4*b + r]};
      tdlt_idct4(t);
      v[4*b + r] = t[0]; v[8 + 4*b + r] = t[1]; v[16 + 4*b + r] = t[2]; v[24 + 4*b + r] = t[3];
    }
#pragma unroll
    for (int q = 0; q < 4; q++) tdlt_idct4(v + 8*q + 4*b);
  }
}

}  // namespace tdlt

// Host-visible declarations of the batch kernels (internal to the library).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include "tdlt_device.cuh"
#include "tdlt_region.cuh"

namespace tdlt {

// TMA descriptors, one per plane: `in` is read by tiled loads, `out` written
// by tiled stores (forward only).
struct TmapSet {
  CUtensorMap in[kMaxPlanes];
  CUtensorMap out[kMaxPlanes];
};

struct InvOut {
  int *c[kMaxPlanes];
};

__global__ void tdlt_fwd_region_kernel(const __grid_constant__ TmapSet maps,
                                       const __grid_constant__ RegionBatch bd,
                                       const unsigned char *__restrict__ bsize);
__global__ void tdlt_inv_region_kernel(const __grid_constant__ TmapSet maps,
                                       const __grid_constant__ RegionBatch bd,
                                       const unsigned char *__restrict__ bsize, InvOut out,
                                       int2 *__restrict__ strips, int *__restrict__ ticket,
                                       int epoch);
constexpr int kRegionStripInts = 2*kRegW + 2*kRegH;

}  // namespace tdlt

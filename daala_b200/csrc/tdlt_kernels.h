// Host-visible declarations of the batch kernels (internal to the library).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include "tdlt_device.cuh"
#include "tdlt_region.cuh"

namespace tdlt {

// TMA descriptors of a kernel's input planes (3-D: x, y, frame).
struct TmapSet {
  CUtensorMap in[kMaxPlanes];
};

// One device pointer per plane (nframes planes each, 16-byte aligned).
struct PlanePtrs {
  int *p[kMaxPlanes];
};

// plan: one kPlanBytes record per region, index frame*regs_per_frame + region
__global__ void tdlt_plan_kernel(const __grid_constant__ RegionBatch bd,
                                 const unsigned char *__restrict__ bsize,
                                 unsigned char *__restrict__ plan);
__global__ void tdlt_fwd_region_kernel(const __grid_constant__ TmapSet maps,
                                       const __grid_constant__ RegionBatch bd,
                                       const unsigned char *__restrict__ plan, PlanePtrs out);
__global__ void tdlt_fwd_region_pix8_kernel(const __grid_constant__ TmapSet maps,
                                            const __grid_constant__ RegionBatch bd,
                                            const unsigned char *__restrict__ plan, PlanePtrs out);
__global__ void tdlt_fwd_region_pix16_kernel(const __grid_constant__ TmapSet maps,
                                             const __grid_constant__ RegionBatch bd,
                                             const unsigned char *__restrict__ plan, PlanePtrs out);
__global__ void tdlt_fwd_haar_kernel(const __grid_constant__ TmapSet maps,
                                     const __grid_constant__ RegionBatch bd,
                                     const unsigned char *__restrict__ plan, PlanePtrs out);
__global__ void tdlt_inv_region_kernel(const __grid_constant__ TmapSet maps,
                                       const __grid_constant__ RegionBatch bd,
                                       const unsigned char *__restrict__ plan,
                                       PlanePtrs out, int2 *__restrict__ strips,
                                       int *__restrict__ ticket, int epoch,
                                       const unsigned short *__restrict__ order, PlanePtrs in);
__global__ void tdlt_inv_region_pix8_kernel(const __grid_constant__ TmapSet maps,
                                            const __grid_constant__ RegionBatch bd,
                                            const unsigned char *__restrict__ plan,
                                            PlanePtrs out, int2 *__restrict__ strips,
                                            int *__restrict__ ticket, int epoch,
                                            const unsigned short *__restrict__ order, PlanePtrs in);
__global__ void tdlt_inv_region_pix16_kernel(const __grid_constant__ TmapSet maps,
                                             const __grid_constant__ RegionBatch bd,
                                             const unsigned char *__restrict__ plan,
                                             PlanePtrs out, int2 *__restrict__ strips,
                                             int *__restrict__ ticket, int epoch,
                                             const unsigned short *__restrict__ order, PlanePtrs in);
#define TDLT_DECL_INV_CO(name) \
  __global__ void name(const __grid_constant__ TmapSet maps, const __grid_constant__ RegionBatch bd, \
                       const unsigned char *__restrict__ plan, PlanePtrs out, int2 *__restrict__ strips, \
                       int *__restrict__ ticket, int epoch, const unsigned short *__restrict__ order, \
                       PlanePtrs in)
TDLT_DECL_INV_CO(tdlt_inv_region_co_kernel);
TDLT_DECL_INV_CO(tdlt_inv_region_co_pix8_kernel);
TDLT_DECL_INV_CO(tdlt_inv_region_co_pix16_kernel);
__global__ void tdlt_inv_haar_kernel(const __grid_constant__ TmapSet maps,
                                     const __grid_constant__ RegionBatch bd,
                                     const unsigned char *__restrict__ plan,
                                     PlanePtrs out, int2 *__restrict__ strips,
                                     int *__restrict__ ticket, int epoch,
                                     const unsigned short *__restrict__ order, PlanePtrs in);
constexpr int kRegionStripInts = 2*kRegW + 2*kRegH;

}  // namespace tdlt

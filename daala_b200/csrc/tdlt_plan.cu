// The per-region PLAN of a batch: what the reference's recursive quadtree walk
// (od_compute_dcts src/encode.c:1286-1343, od_decode_recursive src/decode.c:842-866)
// decides block by block -- which transform size covers each 8x8 unit, which edges
// are lapped at which level -- flattened into one 1 KB record per 128x64 region:
// unit tables, four class-sorted task lists with their counts, edge bit masks
// (layout: tdlt_region.cuh, kOffLs .. kOffBar).  It depends on the block-size map and
// the geometry only, so the forward and the inverse kernel of the same frames read
// the same records, each with one bulk copy while its samples are in flight.
// One CTA builds kPlanPerCta consecutive regions of one frame: every thread first
// issues the map-byte loads of its unit in all of them (the kernel is bound by that
// latency), then the records are built one after the other.  Grid
// (ceil(regions per frame / kPlanPerCta), frames).
#include "tdlt_kernels.h"

namespace tdlt {

__global__ void __launch_bounds__(192)
tdlt_plan_kernel(const __grid_constant__ RegionBatch bd, const unsigned char *__restrict__ bsize,
                 unsigned char *__restrict__ plan) {
  __shared__ __align__(16) unsigned char rec[kPlanPerCta][kPlanBytes];
  const int tid = threadIdx.x, lane = tid & 31;
  const int frame = blockIdx.y, r0 = blockIdx.x*kPlanPerCta;
  const int nreg = min(kPlanPerCta, bd.regs_per_frame - r0);
  const unsigned char *map = bsize + (size_t)frame*bd.bw*bd.bh;
  const int ur = tid/18, uc = tid - ur*18;

  int raw[kPlanPerCta];
  bool inside[kPlanPerCta];
  RegionId ids[kPlanPerCta];
#pragma unroll
  for (int k = 0; k < kPlanPerCta; k++) {
    raw[k] = 0;
    inside[k] = false;
    ids[k] = decode_region(bd, frame, min(r0 + k, bd.regs_per_frame - 1));
    if (k < nreg && tid < 18*10) {
      const RegionId id = ids[k];
      int idx = region_unit_map_index(bd, bd.pl[id.pli], id, ur, uc);
      inside[k] = idx >= 0;
      if (inside[k]) raw[k] = map[idx];
    }
  }
  for (int i = tid; i < kPlanPerCta*kPlanBytes/16; i += 192) {
    reinterpret_cast<int4 *>(&rec[0][0])[i] = make_int4(0, 0, 0, 0);
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < kPlanPerCta; k++) {
    if (k < nreg) {
      RegionSmem s;
      s.A = nullptr;
      s.ls = rec[k];
      s.ub = rec[k] + (kOffUb - kOffLs);
      s.tasks = rec[k] + (kOffTasks - kOffLs);
      s.cnt = reinterpret_cast<int *>(rec[k] + (kOffCnt - kOffLs));
      s.msk = reinterpret_cast<unsigned *>(rec[k] + (kOffMsk - kOffLs));
      s.bar = nullptr;
      s.dyn = nullptr;
      const RegionId id = ids[k];
      region_setup_units(s, bd, bd.pl[id.pli], id, tid, lane, inside[k], raw[k]);
    }
  }
  __syncthreads();
  int4 *dst = reinterpret_cast<int4 *>(plan + ((size_t)frame*bd.regs_per_frame + r0)*kPlanBytes);
  for (int i = tid; i < nreg*(kPlanBytes/16); i += 192) {
    dst[i] = reinterpret_cast<const int4 *>(&rec[0][0])[i];
  }
}

}  // namespace tdlt

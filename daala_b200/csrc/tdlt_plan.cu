// The per-region PLAN of a batch: what the reference's recursive quadtree walk
// (od_compute_dcts src/encode.c:1286-1343, od_decode_recursive src/decode.c:842-866)
// decides block by block -- which transform size covers each 8x8 unit, which edges
// are lapped at which level -- flattened into one 1 KB record per 128x64 region:
// unit tables, four class-sorted task lists with their counts, edge bit masks
// (layout: tdlt_region.cuh, kOffLs .. kOffBar).  It depends on the block-size map and
// the geometry only, so the forward and the inverse kernel of the same frames read
// the same records, each with one bulk copy while its samples are in flight.
//
// One thread per unit of the 18 x 10 grid around a region (the region's 16 x 8 units
// and a ring of one):
//  - leaf size: OD_BLOCK_SIZE4x4 (reference src/block_size.h:32-35), bs =
//    OD_MAXI(obs, xdec) - xdec (src/encode.c:1297-1301); lossless mode forces
//    32x32 (src/encode.c:3090); 0xFF outside the plane;
//  - which of the unit's edges are filtered: superblock edges always, over the
//    whole padded plane (src/filter.c:1567-1584); split mid-edges gated by
//    hfilter / vfilter = the split block's right / bottom side lies inside the
//    LUMA picture, compared in plane-local coordinates (src/encode.c:1318-1319,
//    src/filter.c:1494-1508);
//  - the unit's transform goes to the task list of its class (4x4-unit, 8x8, 16x16,
//    32x32; blocks larger than a unit are entered by their first unit).
//
// One CTA builds kPlanPerCta horizontally adjacent regions (same plane, same region
// row, so everything that depends on the row is computed once): every thread first
// issues the map-byte loads of its unit in all of them, then the records are built.
// Grid (groups per frame, frames).
#include "tdlt_kernels.h"

namespace tdlt {

__global__ void __launch_bounds__(192)
tdlt_plan_kernel(const __grid_constant__ RegionBatch bd, const unsigned char *__restrict__ bsize,
                 unsigned char *__restrict__ plan) {
  __shared__ __align__(16) unsigned char rec[kPlanPerCta][kPlanBytes];
  const int tid = threadIdx.x, lane = tid & 31;
  const int frame = blockIdx.y;
  // group -> (plane, region row, first region column); uniform over the CTA
  int g = blockIdx.x, p = 0;
#pragma unroll
  for (int q = 1; q < kMaxPlanes; q++) {
    if (q < bd.nplanes && g >= bd.pl[q].grp0) p = q;
  }
  const RegionPlane &pd = bd.pl[p];
  g -= pd.grp0;
  const int ry = g/pd.ngx, rx0 = (g - ry*pd.ngx)*kPlanPerCta;
  const int nreg = min(kPlanPerCta, pd.nrx - rx0);
  const int dec = pd.dec, sb = 32 >> dec;
  const bool haar = (bd.flags & 2) != 0;
  const unsigned char *map = bsize + (size_t)frame*bd.bw*bd.bh;

  // this thread's grid position (r, c): plane unit row uy, unit columns ux0 + 16 k
  const int r = tid/18, c = tid - r*18;
  const bool active = tid < 18*10;
  const int uy = ry*kRegUY + r - 1, ux0 = rx0*kRegUX + c - 1;
  const int Y = 8*uy;
  const bool iny = active && uy >= 0 && Y < pd.h;
  const unsigned char *maprow = map + min(uy << dec, bd.bh - 1)*bd.bw;

  int raw[kPlanPerCta];
  bool inside[kPlanPerCta];
#pragma unroll
  for (int k = 0; k < kPlanPerCta; k++) {
    int ux = ux0 + kRegUX*k;
    inside[k] = iny && k < nreg && ux >= 0 && 8*ux < pd.w;
    raw[k] = inside[k] ? maprow[min(ux << dec, bd.bw - 1)] : 0;
  }
  for (int i = tid; i < kPlanPerCta*kPlanBytes/16; i += 192) {
    reinterpret_cast<int4 *>(&rec[0][0])[i] = make_int4(0, 0, 0, 0);
  }
  __syncthreads();

  // what depends on the row only: is the unit's top edge a superblock edge, and the
  // size of the block whose mid-edge it would otherwise be
  const bool top_sb = (Y & (sb - 1)) == 0;
  const int nbY = 2*(Y & -Y);
  const bool midv_in = Y + 8 <= bd.pic_h;
  const int UR = r - 1, UC = c - 1;
  const bool owned = active && UR >= 0 && UR < kRegUY && UC >= 0 && UC < kRegUX && !haar;
  if (tid < 192) {
#pragma unroll
    for (int k = 0; k < kPlanPerCta; k++) {
      if (k >= nreg) break;                     // uniform
      unsigned char *ls = rec[k];
      unsigned char *ub = rec[k] + (kOffUb - kOffLs);
      unsigned char *tasks = rec[k] + (kOffTasks - kOffLs);
      int *cnt = reinterpret_cast<int *>(rec[k] + (kOffCnt - kOffLs));
      // Branch-free (bitwise and / or, selects): the lanes of a warp differ in every one of
      // these conditions.
      const bool in = inside[k];
      const int Lin = haar ? 3 : max(raw[k], dec) - dec;
      const int X = 8*(ux0 + kRegUX*k);
      const int nm = (4 << Lin) - 1;          // block size - 1
      const int nbX = 2*(X & -X);
      // an edge inside the unit's block is no edge; the mid-edge of a split block is filtered
      // along the whole block iff the block's right (bottom) side is inside the picture
      const bool eT = (Y > 0) & (top_sb | (((Y & nm) == 0) & ((X & -nbY) + nbY <= bd.pic_w)));
      const bool eL = (X > 0) & (((X & (sb - 1)) == 0) | (((X & nm) == 0) & ((Y & -nbX) + nbX <= bd.pic_h)));
      const bool l0 = Lin == 0;
      int bits = (eT ? kUbTop : 0) | (eL ? kUbLeft : 0) | ((l0 & (X + 8 <= bd.pic_w)) ? kUbMidH : 0) |
                 ((l0 & midv_in) ? kUbMidV : 0);
      bits = (in & !haar) ? bits : 0;
      const int L = in ? Lin : 0xFF;
      const bool first2 = ((UR | UC) & 1) == 0, first3 = ((UR | UC) & 3) == 0;
      int which = Lin <= 1 ? Lin : ((Lin == 2) & first2) ? 2 : ((Lin == 3) & first3) ? 3 : -1;
      which = (in & owned) ? which : -1;
      if (active) {
        ls[r*kUS + c] = (unsigned char)L;
        ub[r*kUS + c] = (unsigned char)bits;
      }
      list_push4(tasks, kTaskListStride, cnt + 4, which, (unsigned char)((UR << 4) | UC), lane);
    }
  }
  // the order of the two filter directions at the crossing of row line 8*iu and column line
  // 8*ju (region_lap_pass): bit ju of word iu = the row line is at least as coarse
  if (tid < kRegUY + 1) {
    const int li = lap_level8(tid, sb);
    unsigned m = 0;
#pragma unroll
    for (int ju = 0; ju <= kRegUX; ju++) m |= (unsigned)(li >= lap_level8(ju, sb)) << ju;
    for (int k = 0; k < nreg; k++) reinterpret_cast<unsigned *>(rec[k] + (kOffMsk - kOffLs))[tid] = m;
  }
  __syncthreads();
  int4 *dst = reinterpret_cast<int4 *>(
      plan + ((size_t)frame*bd.regs_per_frame + pd.reg0 + ry*pd.nrx + rx0)*kPlanBytes);
  for (int i = tid; i < nreg*(kPlanBytes/16); i += 192) {
    dst[i] = reinterpret_cast<const int4 *>(&rec[0][0])[i];
  }
}

}  // namespace tdlt

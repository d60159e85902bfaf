// Forward TDLT over a batch of frames: c -> d.
//   SB-edge prefilter (reference src/filter.c:1556-1586), split prefilters
//   (:1486-1510), leaf DCTs (src/dct.c), keyframe Haar-DC merge
//   (src/encode.c:1286-1343); lossless mode: od_haar (src/dct.c:1960-1997).
// One CTA per 128x64 region of one plane; see tdlt_region.cuh for the layout
// and the task vocabulary.  The input tile needs only INPUT samples from the
// neighbouring regions (a 2-sample ring), so regions are independent.
#include <cuda.h>
#include "tdlt_tasks.cuh"
#include "tdlt_region.cuh"
#include "tdlt_kernels.h"

namespace tdlt {

#ifdef TDLT_TRACE
__device__ long long *g_tdlt_trace = nullptr;
#endif

// ---- lossless mode: multi-level Haar wavelet of each superblock of a tile
// (reference src/dct.c:1960-1997).  T = the tile's part of A, Y = its B tile.
__device__ __forceinline__ void region_haar_fwd_tile(int *At, int *Bt, int lane, int sb) {
  int ln = (sb == 32) ? 5 : 4;
  int nb = 32/sb;                         // superblocks per tile side
  for (int level = 0; level < ln; level++) {
    int np = sb >> level >> 1;
    int items = nb*nb*np*np;
    int a[8], b[8], c[8], d[8];
#pragma unroll
    for (int k = 0; k < 8; k++) {
      int it = lane + 32*k;
      if (it < items) {
        int blk = it/(np*np), e = it - blk*np*np;
        int i = e/np, j = e - i*np;
        int r0 = (blk/nb)*sb, c0 = (blk%nb)*sb;
        a[k] = At[(r0 + 2*i)*kAP + c0 + 2*j];
        b[k] = At[(r0 + 2*i + 1)*kAP + c0 + 2*j];
        c[k] = At[(r0 + 2*i)*kAP + c0 + 2*j + 1];
        d[k] = At[(r0 + 2*i + 1)*kAP + c0 + 2*j + 1];
      }
    }
    __syncwarp();
#pragma unroll
    for (int k = 0; k < 8; k++) {
      int it = lane + 32*k;
      if (it < items) {
        int blk = it/(np*np), e = it - blk*np*np;
        int i = e/np, j = e - i*np;
        int r0 = (blk/nb)*sb, c0 = (blk%nb)*sb;
        haar2x2(a[k], b[k], c[k], d[k]);
        At[(r0 + i)*kAP + c0 + j] = a[k];
        Bt[b_off(r0 + i, c0 + j + np)] = b[k];
        Bt[b_off(r0 + i + np, c0 + j)] = c[k];
        Bt[b_off(r0 + i + np, c0 + j + np)] = d[k];
        if (np == 1) Bt[b_off(r0, c0)] = a[k];
      }
    }
    __syncwarp();
  }
}

__global__ void __launch_bounds__(kRegThreads, TDLT_MINB)
tdlt_fwd_region_kernel(const __grid_constant__ TmapSet maps,
                       const __grid_constant__ RegionBatch bd,
                       const unsigned char *__restrict__ bsize) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  TDLT_MARK(blockIdx.y*gridDim.x + blockIdx.x, 0);
  const RegionSmem s = carve_region(smem);
  const RegionId id = decode_region(bd, blockIdx.y, blockIdx.x);
  const RegionPlane &pd = bd.pl[id.pli];
  const int X0 = id.rx*kRegW, Y0 = id.ry*kRegH;
  const bool haar = (bd.flags & 2) != 0;

  if (tid == 0 && !(TDLT_ABLATE & 16)) {
    mbar_init(s.bar, 1);
    fence_mbar_init();
    mbar_expect_tx(s.bar, kARegBytes);
    tma_load_3d(s.A, &maps.in[id.pli], s.bar, X0 - 4, Y0 - 2, id.frame);
  }
  if (tid >= 228) s.cnt[tid - 228] = 0;       // counters and edge masks
  __syncthreads();
  if (!(TDLT_ABLATE & 8)) region_setup_units(s, bsize, bd, pd, id, tid, lane);
  __syncthreads();
  TDLT_MARK(blockIdx.y*gridDim.x + blockIdx.x, 1);
  if (!(TDLT_ABLATE & 16)) mbar_wait(s.bar, 0);
  TDLT_MARK(blockIdx.y*gridDim.x + blockIdx.x, 2);

  if (!haar) {
    // ---- lapping, level by level, coarse to fine (level 4 is done in registers
    // by the transform tasks of the units it concerns) ---------------------------
    if (!(TDLT_ABLATE & 1)) {
#ifndef TDLT_LAP_BLOCKS
      const int sb = 32 >> pd.dec;
#pragma unroll 1
      for (int lv = sb; lv >= 8; lv >>= 1) {
        region_lap_cols<false>(s.A, s.msk, lv, sb, warp, lane);
        __syncthreads();
        region_lap_rows<false>(s.A, s.msk, lv, sb, warp, lane, false);
        __syncthreads();
      }
#else
      region_lap_blocks<false>(s.A, s.ub, 32 >> pd.dec, warp, lane, 9, 17, 9, 17);
      __syncthreads();
#endif
    }
    TDLT_MARK(blockIdx.y*gridDim.x + blockIdx.x, 4);
    // ---- transforms ---------------------------------------------------------
    if (!(TDLT_ABLATE & 2)) {
      const int n0 = s.cnt[4], n1 = s.cnt[5], n2 = s.cnt[6], n3 = s.cnt[7];
      const int q0 = (n0 + 15) >> 4, q1 = (n1 + 15) >> 4, q2 = (n2 + 3) >> 2;
      const int ntask = q0 + q1 + q2 + n3;
      unsigned char *Bb = reinterpret_cast<unsigned char *>(s.B);
      int v[32];
#pragma unroll 1
      for (;;) {
        // longest tasks first; a warp that is done draws the next one
        int g = 0;
        if (lane == 0) g = atomicAdd(s.cnt + 2, 1);
        g = __shfl_sync(0xffffffffu, g, 0);
        if (g >= ntask) break;
        if (g < n3) {
          // 32x32: lane = column, then lane = row
          int e = s.tasks[3*kTaskListStride + g];
          int UR = e >> 4, UC = e & 15;
          unsigned char *tile = Bb + ((UR >> 2)*4 + (UC >> 2))*kBBytes;
#pragma unroll 1
          for (int pass = 0; pass < 2; pass++) {
            if (pass == 0) a_col32<false>(s.A + (8*UR + 2)*kAP + 8*UC + 4 + lane, v);
            else b_row32<false>(tile, lane, v);
            tdlt_fdct32(v);
            if (pass == 0) b_col32<true>(tile, lane, v);
            else b_row32<true>(tile, lane, v);
            __syncwarp();
          }
        }
        else if (g < n3 + q2) {
          // 16x16: 8 lanes per block
          int idx = (g - n3)*4 + (lane >> 3), k = lane & 7;
          bool on = idx < n2;
          int e = s.tasks[2*kTaskListStride + (on ? idx : 0)];
          int UR = e >> 4, UC = e & 15;
          unsigned char *tile = Bb + ((UR >> 2)*4 + (UC >> 2))*kBBytes;
          int r16 = (UR & 2)*8, c16 = (UC & 2)*8;
#pragma unroll 1
          for (int pass = 0; pass < 2; pass++) {
            if (on) {
              if (pass == 0) a_cols16<false>(s.A + (8*UR + 2)*kAP + 8*UC + 4 + 2*k, v);
              else b_rows16<false>(tile, r16, c16, k, v);
              tdlt_fdct16(v);
              tdlt_fdct16(v + 16);
              if (pass == 0) b_cols16<true>(tile, r16, c16, k, v);
              else b_rows16<true>(tile, r16, c16, k, v);
            }
            __syncwarp();
          }
        }
        else if (g < n3 + q2 + q1) {
          // 8x8: 2 lanes per block
          int idx = (g - n3 - q2)*16 + (lane >> 1), h = lane & 1;
          bool on = idx < n1;
          int e = s.tasks[1*kTaskListStride + (on ? idx : 0)];
          int UR = e >> 4, UC = e & 15;
          unsigned char *tile = Bb + ((UR >> 2)*4 + (UC >> 2))*kBBytes;
          int ur = UR & 3, uc = UC & 3;
          int mu = (uc ^ (lane >> 1)) & 3;
#pragma unroll 1
          for (int pass = 0; pass < 2; pass++) {
            if (on) {
              if (pass == 0) ld_a_8x4(s.A + (8*UR + 2)*kAP + 8*UC + 4 + 4*h, v);
              else b_rows8<false>(tile, ur, uc, h, mu, v);
#pragma unroll
              for (int q = 0; q < 4; q++) tdlt_fdct8(v + 8*q);
              if (pass == 0) b_8x4<true>(tile, ur, uc, h, v);
              else b_rows8<true>(tile, ur, uc, h, mu, v);
            }
            __syncwarp();
          }
        }
        else {
          // 8x8 unit of four 4x4 blocks: each lane holds two stacked blocks; the
          // unit's own mid-edges are lapped here, in registers
          int idx = (g - n3 - q2 - q1)*16 + (lane >> 1), h = lane & 1;
          bool on = idx < n0;
          int e = s.tasks[on ? idx : 0];
          int UR = e >> 4, UC = e & 15;
          unsigned char *tile = Bb + ((UR >> 2)*4 + (UC >> 2))*kBBytes;
          int bits = on ? s.ub[ut(UR, UC)] : 0;
          if (on) ld_a_8x4(s.A + (8*UR + 2)*kAP + 8*UC + 4 + 4*h, v);
          lap4_unit<false>(v, h, bits);
          if (on) {
            fdct4x4_pair(v);
            b_8x4<true>(tile, UR & 3, UC & 3, h, v);
          }
        }
      }
    }
    fence_proxy_async_smem();   // B is read by the TMA store (async proxy) below
    __syncthreads();
    TDLT_MARK(blockIdx.y*gridDim.x + blockIdx.x, 5);
  }

  // ---- per tile: DC pyramid (or the Haar wavelet), then one TMA store ------
  {
    const int tx = warp & 3, ty = warp >> 2;
    const int TX = X0 + 32*tx, TY = Y0 + 32*ty;
    if (TX < pd.w && TY < pd.h) {
      int *Bt = s.B + warp*(kBBytes/4);
      if (haar) {
        region_haar_fwd_tile(s.A + (32*ty + 2)*kAP + 32*tx + 4, Bt, lane, 32 >> pd.dec);
      }
      else if ((bd.flags & 1) && !(TDLT_ABLATE & 4)) {
        const unsigned char *ls0 = s.ls + ut(4*ty, 4*tx);
        region_dc_level<false>(Bt, ls0, lane, 8);
        region_dc_level<false>(Bt, ls0, lane, 16);
        if (pd.dec == 0) region_dc_level<false>(Bt, ls0, lane, 32);
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0 && !(TDLT_ABLATE & 16)) {
        tma_store_3d(&maps.out[id.pli], Bt, TX, TY, id.frame);
        tma_store_commit_and_wait_read();
      }
    }
  }
  TDLT_MARK(blockIdx.y*gridDim.x + blockIdx.x, 6);
}

}  // namespace tdlt

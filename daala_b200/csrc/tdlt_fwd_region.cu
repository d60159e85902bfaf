// Forward TDLT over a batch of frames: c -> d.
//   SB-edge prefilter (reference src/filter.c:1556-1586), split prefilters
//   (:1486-1510), leaf DCTs (src/dct.c), keyframe Haar-DC merge
//   (src/encode.c:1286-1343); lossless mode: od_haar (src/dct.c:1960-1997).
// One CTA per 128x64 region of one plane; see tdlt_region.cuh for the layout
// and the task vocabulary.  A region needs only INPUT samples from its
// neighbours (a 2-sample ring), so regions are independent.  Everything happens
// in place in one shared-memory buffer: TMA load with halo -> lapping (one pass) ->
// transforms -> DC pyramid -> 16-byte stores.
#include <cuda.h>
#include "tdlt_tasks.cuh"
#include "tdlt_region.cuh"
#include "tdlt_kernels.h"

namespace tdlt {

#ifdef TDLT_TRACE
__device__ long long *g_tdlt_trace = nullptr;
#endif

template <bool kHaar, int kPix>
__device__ __forceinline__ void tdlt_fwd_region_body(const TmapSet &maps,
                       const RegionBatch &bd,
                       const unsigned char *__restrict__ plan, PlanePtrs out) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  TDLT_MARK(blockIdx.y*gridDim.x + blockIdx.x, 0);
  const RegionSmem s = carve_region(smem);
  const RegionId id = decode_region(bd, blockIdx.y, blockIdx.x);
  const RegionPlane &pd = bd.pl[id.pli];
  const int X0 = id.rx*kRegW, Y0 = id.ry*kRegH;
  constexpr bool haar = kHaar;      // lossless mode is a kernel of its own (less code in either)

  // One thread starts both loads: the samples (TMA box with the halo) and the region's
  // plan record (tdlt_plan.cu), on one mbarrier.
  if (tid == 0) {
    mbar_init(s.bar, 1);
    fence_mbar_init();
    if (kPix == 0) {
      mbar_expect_tx(s.bar, kARegBytes + kPlanBytes);
      tma_load_3d(s.A, &maps.in[id.pli], s.bar, X0 - 4, Y0 - 2, id.frame);
    }
    else {
      // the region as 8- / 16-bit reference-buffer samples, into the tail of A's storage
      constexpr int kP = kPix ? kPix : 1;
      mbar_expect_tx(s.bar, PixBox<kP>::kBytes + kPlanBytes);
      tma_load_3d(smem + kOffA + PixBox<kP>::kOff, &maps.in[id.pli], s.bar, X0 - 4 - PixBox<kP>::kSkip, Y0 - 2, id.frame);
    }
    bulk_g2s(s.ls, plan + ((size_t)blockIdx.y*bd.regs_per_frame + blockIdx.x)*kPlanBytes,
             kPlanBytes, s.bar);
    s.dyn[1] = 0;                             // task draw counter
  }
  __syncthreads();
  TDLT_MARK(blockIdx.y*gridDim.x + blockIdx.x, 1);
  mbar_wait(s.bar, 0);
  if (kPix) {
    region_widen_box<(kPix ? kPix : 1)>(smem, bd.pix_shift, tid);
    __syncthreads();
  }
  TDLT_MARK(blockIdx.y*gridDim.x + blockIdx.x, 2);

  if (!haar) {
    // ---- lapping of the levels superblock .. 8, one pass (level 4 is done in registers
    // by the transform tasks of the units it concerns) ---------------------------
    if (!(TDLT_ABLATE & 1)) {
      region_lap_pass<false>(s.A, s.ub, s.msk, kCellsFwd, warp, lane);
      __syncthreads();
    }
    TDLT_MARK(blockIdx.y*gridDim.x + blockIdx.x, 4);
    // ---- transforms, in place: columns first, then rows (src/dct.c:137-142) ----
    if (!(TDLT_ABLATE & 2)) {
      const int n0 = s.cnt[4], n1 = s.cnt[5], n2 = s.cnt[6], n3 = s.cnt[7];
      const int q0 = (n0 + 15) >> 4, q1 = (n1 + 15) >> 4, q2 = (n2 + 3) >> 2;
      const int ntask = q0 + q1 + q2 + n3;
      int v[32];
#pragma unroll 1
      for (;;) {
        // longest tasks first; a warp that is done draws the next one
        int g = 0;
        if (lane == 0) g = atomicAdd(s.dyn + 1, 1);
        g = __shfl_sync(0xffffffffu, g, 0);
        if (g >= ntask) break;
        if (g < n3 && !(TDLT_ABLATE & 256)) {
          // 32x32: lane = column, then lane = row
          int e = s.tasks[3*kTaskListStride + g];
          int *blk = s.A + (8*(e >> 4) + 2)*kAP + 8*(e & 15) + 4;
#pragma unroll 1
          for (int pass = 0; pass < 2; pass++) {
            int *line = pass == 0 ? blk + lane : blk + lane*kAP;
            if (pass == 0) a_col32<false>(line, v); else a_row32<false>(line, v);
            tdlt_fdct32(v);
            if (pass == 0) a_col32<true>(line, v); else a_row32<true>(line, v);
            __syncwarp();
          }
        }
        else if (g < n3 + q2) {
          // 16x16: 8 lanes per block
          int idx = (g - n3)*4 + (lane >> 3), k = lane & 7;
          bool on = idx < n2;
          int e = s.tasks[2*kTaskListStride + (on ? idx : 0)];
          int *blk = s.A + (8*(e >> 4) + 2)*kAP + 8*(e & 15) + 4;
#pragma unroll 1
          for (int pass = 0; pass < 2; pass++) {
            if (on) {
              if (pass == 0) a_cols16<false>(blk + 2*k, v); else a_rows16<false>(blk, k, v);
              tdlt_fdct16(v);
              tdlt_fdct16(v + 16);
              if (pass == 0) a_cols16<true>(blk + 2*k, v); else a_rows16<true>(blk, k, v);
            }
            __syncwarp();
          }
        }
        else if (g < n3 + q2 + q1) {
          // 8x8: 2 lanes per block
          int idx = (g - n3 - q2)*16 + (lane >> 1), h = lane & 1;
          bool on = idx < n1;
          int e = s.tasks[1*kTaskListStride + (on ? idx : 0)];
          int *blk = s.A + (8*(e >> 4) + 2)*kAP + 8*(e & 15) + 4;
          int mu = ((e & 15) - (lane >> 1)) & 3;
#pragma unroll 1
          for (int pass = 0; pass < 2; pass++) {
            if (on) {
              if (pass == 0) ld_a_8x4(blk + 4*h, v); else a_rows8<false>(blk, h, mu, v);
#pragma unroll
              for (int q = 0; q < 4; q++) tdlt_fdct8(v + 8*q);
              if (pass == 0) st_a_8x4(blk + 4*h, v); else a_rows8<true>(blk, h, mu, v);
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
          int *blk = s.A + (8*UR + 2)*kAP + 8*UC + 4 + 4*h;
          int bits = on ? s.ub[ut(UR, UC)] : 0;
          if (on) ld_a_8x4(blk, v);
          lap4_unit<false>(v, h, bits);
          if (on) {
            fdct4x4_pair(v);
            st_a_8x4(blk, v);
          }
        }
      }
    }
    __syncthreads();
    TDLT_MARK(blockIdx.y*gridDim.x + blockIdx.x, 5);
  }

  // ---- keyframe DC pyramid of the tiles: the warps 0..3 take two tiles each, 16 lanes per
  // tile (half the instructions of one tile per warp); lossless mode: the Haar wavelet of
  // its tile by every warp.  Then the region goes out.
  const bool coding = !haar && (bd.flags & kFlagCodingOrder);
  if (haar) {
    const int tx = warp & 3, ty = warp >> 2;
    if (X0 + 32*tx < pd.w && Y0 + 32*ty < pd.h) {
      region_haar_tile<false>(s.A + (32*ty + 2)*kAP + 32*tx + 4, lane, 32 >> pd.dec);
    }
  }
  else if ((bd.flags & 1) && warp < kRegTiles/2 && !(TDLT_ABLATE & 4)) {
    const int tl = 2*warp + (lane >> 4), ttx = tl & 3, tty = tl >> 2;
    const bool on = X0 + 32*ttx < pd.w && Y0 + 32*tty < pd.h;
    region_dc_merge_tile(s.A + (32*tty + 2)*kAP + 32*ttx + 4, s.ls + ut(4*tty, 4*ttx), lane, on,
                         pd.dec == 0);
  }
  if (coding || !TDLT_BULK_STORE) {
    if (!haar) __syncthreads();           // the pyramid of a tile is another warp's work
    const int tx = warp & 3, ty = warp >> 2;
    const int TX = X0 + 32*tx, TY = Y0 + 32*ty;
    if (TX < pd.w && TY < pd.h) {
      int *At = s.A + (32*ty + 2)*kAP + 32*tx + 4;
      if (coding) {
        // the tile leaves as the coding-order vectors of its blocks (block-linear plane)
        const unsigned char *ls0 = s.ls + ut(4*ty, 4*tx);
        const int lS = 5 - pd.dec, sbw = pd.w >> lS;
        int *gtile = out.p[id.pli] + (size_t)id.frame*pd.w*pd.h +
                     ((size_t)(TY >> lS)*sbw + (TX >> lS))*(size_t)(1 << 2*lS);
        region_store_tile_coding(At, ls0, bd.cotab, pd.dec, gtile, sbw << 2*lS, lane);
      }
      else if (!(TDLT_ABLATE & 16)) {
        int *g = out.p[id.pli] + ((size_t)id.frame*pd.h + TY + (lane >> 3))*pd.w + TX + 4*(lane & 7);
        const int *a = At + (lane >> 3)*kAP + 4*(lane & 7);
#pragma unroll
        for (int it = 0; it < 8; it++) {
          *reinterpret_cast<int4 *>(g + (size_t)(4*it)*pd.w) =
              *reinterpret_cast<const int4 *>(a + 4*it*kAP);
        }
      }
    }
  }
  if (TDLT_BULK_STORE && !(TDLT_ABLATE & 16) && (haar || !(bd.flags & kFlagCodingOrder))) {
    // the region's rows go out by the copy engine, one bulk copy of up to 512 bytes per row
    fence_proxy_async_smem();
    __syncthreads();
    // (warp w: rows 8w .. 8w+7, issued by one elected lane from warp-uniform addresses -- a
    // copy per lane compiles to a loop over the lanes, the copy engine's operands being
    // uniform registers)
    const int wu = __shfl_sync(0xffffffffu, warp, 0);
    constexpr int kRows = kRegH/kRegWarps;
    const int bytes = 4*min(kRegW, pd.w - X0);
    const int n = min(kRegH, pd.h - Y0) - kRows*wu;     // rows of this warp that exist
    if (elect_one()) {
      unsigned long long ga = reinterpret_cast<unsigned long long>(
          out.p[id.pli] + ((size_t)id.frame*pd.h + Y0 + kRows*wu)*pd.w + X0);
      unsigned sa = smem_u32(s.A + (kRows*wu + 2)*kAP + 4);
      const unsigned long long gstep = 4ull*pd.w;
#pragma unroll
      for (int i = 0; i < kRows; i++) {
        if (i < n) bulk_s2g_addr(ga, sa, bytes);
        ga += gstep;
        sa += 4*kAP;
      }
      bulk_commit_wait_read();
    }
  }
  TDLT_MARK(blockIdx.y*gridDim.x + blockIdx.x, 6);
}

__global__ void __launch_bounds__(kRegThreads, TDLT_MINB)
tdlt_fwd_region_kernel(const __grid_constant__ TmapSet maps,
                       const __grid_constant__ RegionBatch bd,
                       const unsigned char *__restrict__ plan, PlanePtrs out) {
  tdlt_fwd_region_body<false, 0>(maps, bd, plan, out);
}

// The same with od_ref_buf_to_coeff fused into the load: maps describe planes of 8-bit
// (pix8) or 16-bit (pix16) reference-buffer samples.
__global__ void __launch_bounds__(kRegThreads, TDLT_MINB)
tdlt_fwd_region_pix8_kernel(const __grid_constant__ TmapSet maps,
                       const __grid_constant__ RegionBatch bd,
                       const unsigned char *__restrict__ plan, PlanePtrs out) {
  tdlt_fwd_region_body<false, 1>(maps, bd, plan, out);
}
__global__ void __launch_bounds__(kRegThreads, TDLT_MINB)
tdlt_fwd_region_pix16_kernel(const __grid_constant__ TmapSet maps,
                       const __grid_constant__ RegionBatch bd,
                       const unsigned char *__restrict__ plan, PlanePtrs out) {
  tdlt_fwd_region_body<false, 2>(maps, bd, plan, out);
}

// Lossless mode (OD_TDLT_HAAR): the Haar wavelet of every superblock instead of lapping + DCTs.
__global__ void __launch_bounds__(kRegThreads, TDLT_MINB)
tdlt_fwd_haar_kernel(const __grid_constant__ TmapSet maps,
                       const __grid_constant__ RegionBatch bd,
                       const unsigned char *__restrict__ plan, PlanePtrs out) {
  tdlt_fwd_region_body<true, 0>(maps, bd, plan, out);
}


}  // namespace tdlt

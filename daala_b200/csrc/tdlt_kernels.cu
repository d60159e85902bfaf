// The two frame-batch kernels of the TDLT stage, hand-written for sm_100a.
//
//   tdlt_fwd_kernel : c -> d   SB-edge + split prefilters, leaf DCTs, Haar-DC merge
//   tdlt_inv_kernel : d -> c   Haar-DC split, leaf iDCTs, split + SB-edge postfilters
//
// One WARP owns one 32x32 tile of one plane of one frame (a luma superblock,
// or 2x2 chroma superblocks in 4:2:0) from load to store; warps never
// synchronise with each other inside a kernel except through the strip
// exchange of the inverse.  Tiles are staged by TMA (cp.async.bulk.tensor, 3-D
// maps x/y/frame) into shared memory; see tdlt_device.cuh for the two layouts
// and tdlt_tasks.cuh for how the lanes share a tile.
//
// Reference behaviour reproduced (KyleSiefring/daala): src/filter.c:1486-1646,
// src/encode.c:1286-1343, src/decode.c:528-639,767-868, src/dct.c:74-2042.
#include <cuda.h>
#include "tdlt_tasks.cuh"
#include "tdlt_kernels.h"

namespace tdlt {

struct WarpSmem {
  int *A;
  int *B;
  unsigned char *ls;
  unsigned long long *bar;
};

__device__ __forceinline__ WarpSmem carve(unsigned char *base, int warp) {
  WarpSmem s;
  s.B = reinterpret_cast<int *>(base + warp*kBBytes);
  s.A = reinterpret_cast<int *>(base + kWarpsPerCta*kBBytes + warp*kAAlloc);
  s.ls = base + kWarpsPerCta*(kBBytes + kAAlloc) + warp*kLsBytes;
  s.bar = reinterpret_cast<unsigned long long *>(
      base + kWarpsPerCta*(kBBytes + kAAlloc + kLsBytes) + warp*8);
  return s;
}

// ---- keyframe DC pyramid (reference src/encode.c:1328-1341, src/decode.c:733)
// One quadtree level: nodes of size n (8, 16 or 32) whose block is split.
template <bool kInverse>
__device__ __forceinline__ void dc_level(int *B, const unsigned char *ls, int lane, int n) {
  int per = 32/n;                 // nodes per tile side
  if (lane < per*per) {
    int r0 = (lane/per)*n, c0 = (lane%per)*n;
    int lsz = ls[(1 + (r0 >> 3))*kLsStride + 1 + (c0 >> 3)];
    if ((4 << lsz) < n) {
      int hn = n >> 1;
      int o0 = b_off(r0, c0), o1 = b_off(r0, c0 + hn);
      int o2 = b_off(r0 + hn, c0), o3 = b_off(r0 + hn, c0 + hn);
      int x0 = B[o0], x1 = B[o1], x2 = B[o2], x3 = B[o3];
      if (kInverse) haar2x2(x0, x1, x2, x3); else haar2x2(x0, x2, x1, x3);
      B[o0] = x0; B[o1] = x1; B[o2] = x2; B[o3] = x3;
    }
  }
  __syncwarp();
}

// ---- lossless mode: multi-level Haar wavelet of each superblock of the tile
// (reference src/dct.c:1960-2026).  T = interior of A (work plane), Y = B.
__device__ __forceinline__ void haar_fwd_tile(int *A, int *B, int lane, int sb) {
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
        a[k] = A[a_at(r0 + 2*i, c0 + 2*j)];
        b[k] = A[a_at(r0 + 2*i + 1, c0 + 2*j)];
        c[k] = A[a_at(r0 + 2*i, c0 + 2*j + 1)];
        d[k] = A[a_at(r0 + 2*i + 1, c0 + 2*j + 1)];
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
        A[a_at(r0 + i, c0 + j)] = a[k];
        B[b_off(r0 + i, c0 + j + np)] = b[k];
        B[b_off(r0 + i + np, c0 + j)] = c[k];
        B[b_off(r0 + i + np, c0 + j + np)] = d[k];
        if (np == 1) B[b_off(r0, c0)] = a[k];
      }
    }
    __syncwarp();
  }
}

__device__ __forceinline__ void haar_inv_tile(int *A, int *B, int lane, int sb) {
  int ln = (sb == 32) ? 5 : 4;
  int nb = 32/sb;
  for (int level = ln - 1; level >= 0; level--) {
    int np = 1 << (ln - 1 - level);
    int items = nb*nb*np*np;
    int a[8], b[8], c[8], d[8];
#pragma unroll
    for (int k = 0; k < 8; k++) {
      int it = lane + 32*k;
      if (it < items) {
        int blk = it/(np*np), e = it - blk*np*np;
        int i = e/np, j = e - i*np;
        int r0 = (blk/nb)*sb, c0 = (blk%nb)*sb;
        a[k] = (np == 1) ? B[b_off(r0, c0)] : A[a_at(r0 + i, c0 + j)];
        b[k] = B[b_off(r0 + i, c0 + j + np)];
        c[k] = B[b_off(r0 + i + np, c0 + j)];
        d[k] = B[b_off(r0 + i + np, c0 + j + np)];
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
        A[a_at(r0 + 2*i, c0 + 2*j)] = a[k];
        A[a_at(r0 + 2*i + 1, c0 + 2*j)] = b[k];
        A[a_at(r0 + 2*i, c0 + 2*j + 1)] = c[k];
        A[a_at(r0 + 2*i + 1, c0 + 2*j + 1)] = d[k];
      }
    }
    __syncwarp();
  }
}

// ===========================================================================
// Forward
// ===========================================================================
__global__ void __launch_bounds__(kWarpsPerCta*32, 6)
tdlt_fwd_kernel(const __grid_constant__ TmapSet maps, const __grid_constant__ BatchDesc bd,
                const unsigned char *__restrict__ bsize) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int t = blockIdx.x*kWarpsPerCta + warp;
  if (t >= bd.tiles_per_frame*bd.nframes) return;
  const WarpSmem s = carve(smem, warp);
  const TileId id = decode_tile(bd, t);
  const PlaneDesc &pd = bd.pl[id.pli];
  const int X0 = id.tx*kTile, Y0 = id.ty*kTile;
  const bool haar = (bd.flags & 2) != 0;

  if (lane == 0) {
    mbar_init(s.bar, 1);
    fence_mbar_init();
    mbar_expect_tx(s.bar, kABytes);
    tma_load_3d(s.A, &maps.in[id.pli], s.bar, X0 - kAColOff, Y0 - kARowOff, id.frame);
  }
  load_leaf_sizes(s.ls, bsize, bd, pd, id, lane);
  __syncwarp();
  mbar_wait(s.bar, 0);

  if (haar) {
    haar_fwd_tile(s.A, s.B, lane, 32 >> pd.dec);
  }
  else {
    // Lapping: 9x9 cells centred on the multiples of 4 from 0 to 32; the
    // outer ring is shared with (and recomputed by) the neighbouring tiles.
#pragma unroll 1
    for (int cell = lane; cell < 81; cell += 32) {
      int i = cell/9, j = cell - i*9;
      int fl = cell_flags(s.ls, 4*i, 4*j, Y0, X0, pd, bd.pic_w, bd.pic_h);
      lap_cell<false>(s.A, i, j, fl);
    }
    __syncwarp();

    const Lane L = make_lane(lane);
    const int lsz = s.ls[(1 + L.ur)*kLsStride + 1 + L.uc];
    int v[32];
#pragma unroll 1
    for (int pass = 0; pass < 2; pass++) {
      if (lsz == 0) {
        if (pass == 0) {
          colA_8<false>(s.A, L, v);
          fdct4x4_pair(v);
          colB_8<true>(s.B, L, v);
        }
      }
      else {
        if (pass == 0) {
          if (lsz == 3) colA_32<false>(s.A, lane, v);
          else if (lsz == 2) colA_16<false>(s.A, L, v);
          else colA_8<false>(s.A, L, v);
        }
        else {
          if (lsz == 3) rowB_32<false>(s.B, lane, v);
          else if (lsz == 2) rowB_16<false>(s.B, L, v);
          else rowB_8<false>(s.B, L, v);
        }
        lines_dct<false>(lsz, v);
        if (pass == 0) {
          if (lsz == 3) colB_32<true>(s.B, lane, v);
          else if (lsz == 2) colB_16<true>(s.B, L, v);
          else colB_8<true>(s.B, L, v);
        }
        else {
          if (lsz == 3) rowB_32<true>(s.B, lane, v);
          else if (lsz == 2) rowB_16<true>(s.B, L, v);
          else rowB_8<true>(s.B, L, v);
        }
      }
      __syncwarp();
    }
    if (bd.flags & 1) {
      dc_level<false>(s.B, s.ls, lane, 8);
      dc_level<false>(s.B, s.ls, lane, 16);
      if (pd.dec == 0) dc_level<false>(s.B, s.ls, lane, 32);
    }
  }
  fence_proxy_async_smem();
  __syncwarp();
  if (lane == 0) {
    tma_store_3d(&maps.out[id.pli], s.B, X0, Y0, id.frame);
    tma_store_commit_and_wait_read();
  }
}

// ===========================================================================
// Inverse
//
// The post-filter of a block edge needs reconstructed samples from both sides
// of it, so a tile cannot finish alone: tile (tx, ty) owns the OUTPUT region
// shifted by (-2, -2) -- plane rows Y0-2..Y0+29, columns X0-2..X0+29, which is
// a whole number of lapping cells -- and needs the two bottom rows / right
// columns of raw iDCT output of its upper / left / upper-left neighbours.
// Every tile publishes those strips (512 bytes) to global memory and raises a
// flag; consumers wait on the flags of tiles that precede them in a ticket
// order (CTAs draw tickets in the order they start running, so a tile only ever
// waits on work that is already resident: no deadlock, and on average the
// producer is ahead because it started first).
// ===========================================================================
__device__ __forceinline__ int ld_acquire(const int *p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release(int *p, int v) {
  asm volatile("st.release.gpu.global.s32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}

__global__ void __launch_bounds__(kWarpsPerCta*32, 6)
tdlt_inv_kernel(const __grid_constant__ TmapSet maps, const __grid_constant__ BatchDesc bd,
                const unsigned char *__restrict__ bsize, InvOut out, int *__restrict__ strips,
                int *__restrict__ flags, int *__restrict__ ticket, int epoch) {
  extern __shared__ __align__(1024) unsigned char smem[];
  int *s_ticket = reinterpret_cast<int *>(smem + kCtaSmemBytes);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) *s_ticket = atomicAdd(ticket, 1);
  __syncthreads();
  const int t = *s_ticket*kWarpsPerCta + warp;
  if (t >= bd.tiles_per_frame*bd.nframes) return;
  const WarpSmem s = carve(smem, warp);
  const TileId id = decode_tile(bd, t);
  const PlaneDesc &pd = bd.pl[id.pli];
  const int X0 = id.tx*kTile, Y0 = id.ty*kTile;
  const bool haar = (bd.flags & 2) != 0;

  if (lane == 0) {
    mbar_init(s.bar, 1);
    fence_mbar_init();
    mbar_expect_tx(s.bar, kBBytes);
    tma_load_3d(s.B, &maps.in[id.pli], s.bar, X0, Y0, id.frame);
  }
  load_leaf_sizes(s.ls, bsize, bd, pd, id, lane);
  __syncwarp();
  mbar_wait(s.bar, 0);

  int *gout = out.c[id.pli] + (size_t)id.frame*pd.w*pd.h;

  if (haar) {
    haar_inv_tile(s.A, s.B, lane, 32 >> pd.dec);
    // No lapping in lossless mode: the tile is stored where it lies.
#pragma unroll 1
    for (int it = 0; it < 16; it++) {
      int r = 2*it + (lane >> 4), c = 2*(lane & 15);
      int2 val = *reinterpret_cast<int2 *>(s.A + a_at(r, c));
      *reinterpret_cast<int2 *>(gout + (size_t)(Y0 + r)*pd.w + X0 + c) = val;
    }
    return;
  }

  if (bd.flags & 1) {
    if (pd.dec == 0) dc_level<true>(s.B, s.ls, lane, 32);
    dc_level<true>(s.B, s.ls, lane, 16);
    dc_level<true>(s.B, s.ls, lane, 8);
  }

  const Lane L = make_lane(lane);
  const int lsz = s.ls[(1 + L.ur)*kLsStride + 1 + L.uc];
  {
    int v[32];
#pragma unroll 1
    for (int pass = 0; pass < 2; pass++) {
      if (lsz == 0) {
        if (pass == 0) {
          colB_8<false>(s.B, L, v);
          idct4x4_pair(v);
          colA_8<true>(s.A, L, v);
        }
      }
      else {
        if (pass == 0) {
          if (lsz == 3) rowB_32<false>(s.B, lane, v);
          else if (lsz == 2) rowB_16<false>(s.B, L, v);
          else rowB_8<false>(s.B, L, v);
        }
        else {
          if (lsz == 3) colB_32<false>(s.B, lane, v);
          else if (lsz == 2) colB_16<false>(s.B, L, v);
          else colB_8<false>(s.B, L, v);
        }
        lines_dct<true>(lsz, v);
        if (pass == 0) {
          if (lsz == 3) rowB_32<true>(s.B, lane, v);
          else if (lsz == 2) rowB_16<true>(s.B, L, v);
          else rowB_8<true>(s.B, L, v);
        }
        else {
          if (lsz == 3) colA_32<true>(s.A, lane, v);
          else if (lsz == 2) colA_16<true>(s.A, L, v);
          else colA_8<true>(s.A, L, v);
        }
      }
      __syncwarp();
    }
  }

  // ---- publish my bottom rows / right columns (raw iDCT output) -----------
  {
    int *mine = strips + (size_t)t*128;
    int e = 2*lane;                       // bottom rows 30, 31: 64 words
    int a = e >> 5, c = e & 31;
    *reinterpret_cast<int2 *>(mine + e) =
        *reinterpret_cast<int2 *>(s.A + a_at(30 + a, c));
    *reinterpret_cast<int2 *>(mine + 64 + 2*lane) =   // right columns 30, 31
        *reinterpret_cast<int2 *>(s.A + a_at(lane, 30));
    __syncwarp();
    if (lane == 0) {
      __threadfence();
      st_release(flags + t, epoch);
    }
  }

  // ---- fetch the ring from the tiles above / left / above-left -----------
  {
    const bool has_l = id.tx > 0, has_u = id.ty > 0;
    if (lane < 3) {
      int nb = (lane == 0) ? t - 1 : (lane == 1) ? t - pd.ntx : t - pd.ntx - 1;
      bool need = (lane == 0) ? has_l : (lane == 1) ? has_u : (has_l && has_u);
      if (need) {
        while (ld_acquire(flags + nb) != epoch) __nanosleep(64);
      }
    }
    __syncwarp();
    if (has_u) {
      const int *up = strips + (size_t)(t - pd.ntx)*128;
      int e = 2*lane, a = e >> 5, c = e & 31;
      int2 val = __ldcg(reinterpret_cast<const int2 *>(up + e));
      *reinterpret_cast<int2 *>(s.A + a_at(a - 2, c)) = val;
    }
    if (has_l) {
      const int *left = strips + (size_t)(t - 1)*128;
      int2 val = __ldcg(reinterpret_cast<const int2 *>(left + 64 + 2*lane));
      *reinterpret_cast<int2 *>(s.A + a_at(lane, -2)) = val;
    }
    if (has_u && has_l && lane < 2) {
      const int *ul = strips + (size_t)(t - pd.ntx - 1)*128;
      int2 val = __ldcg(reinterpret_cast<const int2 *>(ul + lane*32 + 30));
      *reinterpret_cast<int2 *>(s.A + a_at(lane - 2, -2)) = val;
    }
    __syncwarp();
  }

  // ---- un-lap the cells of my output region -------------------------------
  const bool last_r = id.ty == pd.nty - 1, last_c = id.tx == pd.ntx - 1;
  {
    const int nj = 8 + (last_c ? 1 : 0), ni = 8 + (last_r ? 1 : 0);
#pragma unroll 1
    for (int cell = lane; cell < ni*nj; cell += 32) {
      int i = cell/nj, j = cell - i*nj;
      int fl = cell_flags(s.ls, 4*i, 4*j, Y0, X0, pd, bd.pic_w, bd.pic_h);
      lap_cell<true>(s.A, i, j, fl);
    }
    __syncwarp();
  }

  // ---- store: region row a <-> plane row Y0-2+a, region column b <-> X0-2+b -
  {
    const int a0 = (id.ty == 0) ? 2 : 0, a1 = last_r ? 34 : 32;
    const int b0 = (id.tx == 0) ? 2 : 0;
    int *g = gout + (ptrdiff_t)(Y0 - 2)*pd.w + (X0 - 2);
#pragma unroll 1
    for (int it = 0; it < 17; it++) {
      int a = 2*it + (lane >> 4), b = 2*(lane & 15);
      if (a >= a0 && a < a1 && b >= b0) {
        *reinterpret_cast<int2 *>(g + (ptrdiff_t)a*pd.w + b) =
            *reinterpret_cast<int2 *>(s.A + a_at(a - 2, b - 2));
      }
    }
    if (last_c) {
#pragma unroll 1
      for (int a = lane; a < 34; a += 32) {
        if (a >= a0 && a < a1) {
          *reinterpret_cast<int2 *>(g + (ptrdiff_t)a*pd.w + 32) =
              *reinterpret_cast<int2 *>(s.A + a_at(a - 2, 30));
        }
      }
    }
  }
}

}  // namespace tdlt

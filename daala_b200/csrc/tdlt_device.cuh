// Device-side building blocks of the TDLT kernels: the 4-tap lapping filters,
// the Haar butterfly, the shared-memory tile layouts and the block-edge logic.
// Hand-written for sm_100a; no dependency beyond the CUDA runtime headers.
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>
#include "tdlt_lift_gen.cuh"

namespace tdlt {

// ---------------------------------------------------------------------------
// Geometry shared by host and device
// ---------------------------------------------------------------------------
constexpr int kTile = 32;          // a work tile is 32x32 samples of one plane
constexpr int kAStride = 40;       // words per row of the lapping buffer
constexpr int kARows = 36;
constexpr int kAColOff = 4;        // tile column 0 sits at buffer column 4 ...
constexpr int kARowOff = 2;        // ... and tile row 0 at buffer row 2
constexpr int kABytes = kAStride*kARows*4;   // 5760, what the TMA load delivers (45 x 128)
constexpr int kAAlloc = kABytes;
constexpr int kBBytes = kTile*kTile*4;       // 4096, TMA SWIZZLE_128B layout
constexpr int kLsStride = 8;       // leaf-size ring: 6x6 units, row stride 8
constexpr int kLsBytes = 64;
constexpr int kWarpsPerCta = 4;
constexpr int kMaxPlanes = 4;

struct PlaneDesc {
  int w, h;          // plane size in samples (frame size >> dec)
  int dec;           // 0 or 1
  int ntx, nty;      // tiles across / down
  int tile0;         // index of this plane's first tile within a frame
};

struct BatchDesc {
  PlaneDesc pl[kMaxPlanes];
  int nplanes;
  int tiles_per_frame;
  int nframes;
  int pic_w, pic_h;      // LUMA picture size (gates the split filters)
  int bw, bh;            // block-size map: entries across / down (8x8 luma units)
  int flags;             // OD_TDLT_KEYFRAME | OD_TDLT_HAAR
};

// ---------------------------------------------------------------------------
// 4-tap lapping filters (reference src/filter.c:174-249; parameters 85, 75,
// -15, 33 in 1/64ths, :164-167).  In place on four values.
// ---------------------------------------------------------------------------
__device__ __forceinline__ void pre4(int &x0, int &x1, int &x2, int &x3) {
  int t3 = x0 - x3;
  int t2 = x1 - x2;
  int t1 = x1 - (t2 >> 1);
  int t0 = x0 - (t3 >> 1);
  t2 = (t2*85) >> 6;
  t2 += (t2 > 0);
  t3 = (t3*75) >> 6;
  t3 += (t3 > 0);
  t3 += (t2*-15 + 32) >> 6;
  t2 += (t3*33 + 32) >> 6;
  t0 += t3 >> 1;
  t1 += t2 >> 1;
  x0 = t0;
  x1 = t1;
  x2 = t1 - t2;
  x3 = t0 - t3;
}

constexpr int kPostM75 = (int)(3665038760u);   // ceil(2^38/75) - 2^32
constexpr int kPostM85 = (int)(3233857729u);   // ceil(2^38/85) - 2^32
__device__ __forceinline__ void post4(int &y0, int &y1, int &y2, int &y3) {
  int t3 = y0 - y3;
  int t2 = y1 - y2;
  int t1 = y1 - (t2 >> 1);
  int t0 = y0 - (t3 >> 1);
  t2 -= (t3*33 + 32) >> 6;
  t3 -= (t2*-15 + 32) >> 6;
  // (t*64)/75 and (t*64)/85, C division (truncating), as one multiply-high each:
  // trunc(t*64/d) = floor(t*M/2^32) + (t < 0) with M = ceil(2^38/d), exact for
  // |t| < 2^25, the range in which the reference's own t*64 does not overflow
  // (checked exhaustively on the host, tests/test_lifting_header.py).
  t3 = __mulhi(t3, kPostM75) + t3 + (int)((unsigned)t3 >> 31);
  t2 = __mulhi(t2, kPostM85) + t2 + (int)((unsigned)t2 >> 31);
  t0 += t3 >> 1;
  t1 += t2 >> 1;
  y0 = t0;
  y1 = t1;
  y2 = t1 - t2;
  y3 = t0 - t3;
}

// Reversible 2x2 Haar butterfly (reference src/tf.h:34-45).
__device__ __forceinline__ void haar2x2(int &ll, int &lh, int &hl, int &hh) {
  ll += hl;
  hh -= lh;
  int m = (ll - hh) >> 1;
  lh = m - lh;
  hl = m - hl;
  ll -= lh;
  hh += hl;
}

// ---------------------------------------------------------------------------
// Shared-memory layouts
//  A: 36 rows x 40 words.  Holds plane rows Y0-2..Y0+33, columns X0-4..X0+35
//     of a tile whose origin is (X0, Y0).  The innermost TMA coordinate must be
//     16-byte aligned (measured on B200: a tiled load at x = 30 raises "illegal
//     instruction", x = 28 or 32 is fine), hence the 4-column halo although the
//     lapping only needs 2.  Block aligned: 16-byte chunks start on multiples
//     of 4 columns; the 4x4 lapping cells (centred on multiples of 4) straddle
//     two chunks and are accessed as 8-byte halves.
//  B: 32x32 words in the TMA SWIZZLE_128B pattern (16-byte chunk index XOR
//     row&7; the buffer is 1024-byte aligned): "block aligned" -- conflict
//     free for 128-bit row accesses by 8 consecutive rows AND for column-wise
//     accesses, and it is exactly what a TMA tile load/store produces/consumes.
// ---------------------------------------------------------------------------
__device__ __forceinline__ int a_off(int r, int c) { return r*kAStride + c; }
// tile sample (r, c), r and c in [-2, 34)
__device__ __forceinline__ int a_at(int r, int c) { return (r + kARowOff)*kAStride + c + kAColOff; }
__device__ __forceinline__ int b_off(int r, int c) {
  return r*32 + ((((c >> 2) ^ (r & 7)) << 2) | (c & 3));
}

// ---------------------------------------------------------------------------
// mbarrier / TMA (cp.async.bulk.tensor) wrappers
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(void *bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(void *bar, int bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
               :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(void *bar, int parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" :: "r"(smem_u32(bar)), "r"(parity) : "memory");
}
// 3-D tiled load: coordinates (x, y, frame), innermost first.
__device__ __forceinline__ void tma_load_3d(void *dst, const void *tmap, void *bar,
                                            int x, int y, int z) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%2, %3, %4}], [%5];"
      :: "r"(smem_u32(dst)), "l"(tmap), "r"(x), "r"(y), "r"(z), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void tma_store_3d(const void *tmap, const void *src,
                                             int x, int y, int z) {
  asm volatile(
      "cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
      :: "l"(tmap), "r"(smem_u32(src)), "r"(x), "r"(y), "r"(z) : "memory");
}
__device__ __forceinline__ void tma_store_commit_and_wait_read() {
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void prefetch_tmap(const void *tmap) {
  asm volatile("prefetch.tensormap [%0];" :: "l"(tmap) : "memory");
}

// ---------------------------------------------------------------------------
// Which tile is this?
// ---------------------------------------------------------------------------
struct TileId {
  int frame, pli, tx, ty;
};

__device__ __forceinline__ TileId decode_tile(const BatchDesc &bd, int t) {
  TileId id;
  id.frame = t / bd.tiles_per_frame;
  int r = t - id.frame*bd.tiles_per_frame;
  int p = 0;
#pragma unroll
  for (int q = 1; q < kMaxPlanes; q++) {
    if (q < bd.nplanes && r >= bd.pl[q].tile0) p = q;
  }
  id.pli = p;
  r -= bd.pl[p].tile0;
  id.ty = r / bd.pl[p].ntx;
  id.tx = r - id.ty*bd.pl[p].ntx;
  return id;
}

// ---------------------------------------------------------------------------
// Leaf sizes.  ls[] is a 6x6 ring (row stride 8) of log2(leaf size)-2 per 8x8
// PLANE unit around the tile: entry (1+ur, 1+uc) is unit (ur, uc) of the tile.
// reference: OD_BLOCK_SIZE4x4 (src/block_size.h:32-35), bs = OD_MAXI(obs, xdec)
// - xdec (src/encode.c:1297-1301).
// ---------------------------------------------------------------------------
__device__ __forceinline__ void load_leaf_sizes(unsigned char *ls, const unsigned char *bsize,
                                                const BatchDesc &bd, const PlaneDesc &pd,
                                                const TileId &id, int lane) {
  const unsigned char *map = bsize + (size_t)id.frame*bd.bw*bd.bh;
#pragma unroll
  for (int e = lane; e < 36; e += 32) {
    int r = e/6, c = e - r*6;
    int uy = id.ty*4 + r - 1;   // plane unit (8x8 plane samples)
    int ux = id.tx*4 + c - 1;
    int my = min(max(uy << pd.dec, 0), bd.bh - 1);  // map entry (8x8 luma)
    int mx = min(max(ux << pd.dec, 0), bd.bw - 1);
    int v = map[my*bd.bw + mx];
    v = (bd.flags & 2) ? 3 : v;               // lossless Haar: all 32x32
    v = max(v, pd.dec) - pd.dec;
    ls[r*kLsStride + c] = (unsigned char)v;
  }
}

// Leaf log-size at tile-local sample (y, x), each in [-8, 40).
__device__ __forceinline__ int leaf_at(const unsigned char *ls, int y, int x) {
  return ls[((y + 8) >> 3)*kLsStride + ((x + 8) >> 3)];
}

// Does a filtered block edge run along row yc (a multiple of 4, tile-local,
// plane row Y0+yc) at tile-local column x?  reference: SB edges
// src/filter.c:1567-1576 (always, whole padded width); split mid-edges
// src/filter.c:1494-1502 gated by `hfilter` = block right edge <= LUMA
// pic_width (src/encode.c:1318).
__device__ __forceinline__ bool edge_across_row(const unsigned char *ls, int yc, int x,
                                                int Y0, int X0, int sb, int pic_w) {
  int Y = Y0 + yc;
  if ((Y & (sb - 1)) == 0) return true;
  int n = 4 << leaf_at(ls, yc, x);
  if (Y & (n - 1)) return false;
  int nb = 2*(Y & -Y);                 // the block this row splits
  int right = ((X0 + x) & ~(nb - 1)) + nb;
  return right <= pic_w;
}
__device__ __forceinline__ bool edge_across_col(const unsigned char *ls, int xc, int y,
                                                int Y0, int X0, int sb, int pic_h) {
  int X = X0 + xc;
  if ((X & (sb - 1)) == 0) return true;
  int n = 4 << leaf_at(ls, y, xc);
  if (X & (n - 1)) return false;
  int nb = 2*(X & -X);
  int bottom = ((Y0 + y) & ~(nb - 1)) + nb;
  return bottom <= pic_h;
}

// Flags of the 4x4 lapping cell centred on tile-local (yc, xc):
//  bit0/1: filter down the left / right column pair (edge along row yc)
//  bit2/3: filter along the top / bottom row pair (edge along column xc)
//  bit4  : the vertical-direction filter comes first in the FORWARD order
//          (coarser level first; on a tie the column-direction pass first:
//          src/filter.c:1493-1508, 1567-1584 and the recursion order of
//          src/encode.c:1311-1327).
__device__ __forceinline__ int cell_flags(const unsigned char *ls, int yc, int xc,
                                          int Y0, int X0, const PlaneDesc &pd,
                                          int pic_w, int pic_h) {
  int sb = 32 >> pd.dec;
  int Y = Y0 + yc, X = X0 + xc;
  int f = 0;
  if (Y > 0 && Y < pd.h) {
    if (X > 0 && edge_across_row(ls, yc, xc - 1, Y0, X0, sb, pic_w)) f |= 1;
    if (X < pd.w && edge_across_row(ls, yc, xc, Y0, X0, sb, pic_w)) f |= 2;
  }
  if (X > 0 && X < pd.w) {
    if (Y > 0 && edge_across_col(ls, xc, yc - 1, Y0, X0, sb, pic_h)) f |= 4;
    if (Y < pd.h && edge_across_col(ls, xc, yc, Y0, X0, sb, pic_h)) f |= 8;
  }
  int lv = min(Y & -Y, sb), lh = min(X & -X, sb);
  if (Y == 0) lv = sb;
  if (X == 0) lh = sb;
  if (lv >= lh) f |= 16;
  return f;
}

}  // namespace tdlt

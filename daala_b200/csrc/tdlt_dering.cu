// Deringing filter of the reconstructed planes: the step right after
// the inverse stage.  Reference: od_dering and its helpers, src/filter.c:1648-1941, driven
// per 32x32 superblock by the loop of src/decode.c:1044-1122 (encoder twin in
// src/encode.c).  Every superblock reads only the UNFILTERED frame (the reference's etmp
// copy), so superblocks are independent: one CTA per superblock and frame, all planes.
//
//   per plane (luma first; chroma reuses luma's directions):
//     the (n+6) x (n+8) tile arrives as ONE TMA box -> A (the next plane's box is asked
//       for before this plane is worked on); samples outside the frame = 30000; the
//       reference's int16 etmp copy is the low half of every word, read as such
//     luma: direction search of the 16 8x8 blocks (thread = block x direction), variances,
//           per-block thresholds; all planes: skip test of every block
//     pass 1: 6 taps along the block's direction          A -> B (interior)
//     pass 2: 4 taps across it, threshold tightened by |pass-1 change|   A,B -> out
//   superblocks whose flag is clear are copied through.
//
// HBM-bound by design: 4 bytes read + 4 written per sample; the 3-sample ring around a
// superblock is re-read from L2 (neighbouring CTAs are resident at the same time).
#include <cuda.h>
#include <cuda_runtime.h>
#include "tdlt_device.cuh"
#include "tdlt_aux_kernels.h"

namespace tdlt {

namespace {

constexpr int kSB = 32;               // luma superblock (reference OD_BSIZE_MAX)
constexpr int kBorder = 3;            // reference OD_FILT_BORDER
constexpr int kPadL = 4;              // left / right pad of the tile: interior 16-byte aligned
constexpr int kFar = 30000;           // reference OD_DERING_VERY_LARGE
constexpr int kThreads = 128;

// The tile of a plane of one superblock, ring included: n + 6 rows of n + 8 words, dense (it
// is a TMA box).  With row pitches of 40 / 24 words the 4 x 8 pixels of a warp access (8x8
// block) and the 2 x (4 x 4) of a chroma warp access fall into 32 different banks.
template <int DEC> struct Tile {
  static constexpr int n = kSB >> DEC;
  static constexpr int pitch = n + 2*kPadL;
  static constexpr int rows = n + 2*kBorder;
  static constexpr int words = rows*pitch;
  static constexpr int origin = kBorder*pitch + kPadL;      // the superblock's first sample
};
constexpr int kTileWords = 1536;      // Tile<0>::words = 1520, rounded up to whole 128-byte lines
static_assert(Tile<0>::words <= kTileWords, "tile buffer");

// The reference filters its int16_t copy of the frame: the low half of the od_coeff word.
__device__ __forceinline__ int ld16(const int *p) { return *reinterpret_cast<const short *>(p); }

// Line of direction D that pixel (i, j) of an 8x8 block lies on (reference
// src/filter.c:1666-1673).
template <int D> __host__ __device__ constexpr int dir_line(int i, int j) {
  return D == 0 ? i + j : D == 1 ? i + j/2 : D == 2 ? i : D == 3 ? 3 + i - j/2 :
         D == 4 ? 7 + i - j : D == 5 ? 3 - i/2 + j : D == 6 ? j : i/2 + j;
}
template <int D> struct DirLines {   // pixels on each line, counted at compile time
  int n[15];
  constexpr DirLines() : n{} {
    for (int i = 0; i < 8; i++) {
      for (int j = 0; j < 8; j++) n[dir_line<D>(i, j)]++;
    }
  }
};

// Costs of the directions D0 and D1 for the 8x8 block at blk (pixels already shifted down to
// pixel scale, the reference's x >> OD_COEFF_SHIFT; reference src/filter.c:1655-1697): the
// sum over the direction's lines of (line sum)^2 / (pixels on the line).  Fully unrolled: the
// line sums of both directions live in registers, every row is loaded once for the two.
template <int D> __device__ __forceinline__ int lines_cost(const int *s) {
  constexpr DirLines<D> lines{};
  int cost = 0;
#pragma unroll
  for (int k = 0; k < 15; k++) {
    const int np = lines.n[k];
    if (np == 8) cost += (s[k]*s[k]) >> 3;
    else if (np > 0) cost += (int)((unsigned)(s[k]*s[k])/(unsigned)np);
  }
  return cost;
}
template <int D0, int D1, int kPitch>
__device__ __forceinline__ void dir_cost2(const int *blk, int &c0, int &c1) {
  int s0[15], s1[15];
#pragma unroll
  for (int k = 0; k < 15; k++) s0[k] = s1[k] = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    int4 a = *reinterpret_cast<const int4 *>(blk + i*kPitch);
    int4 b = *reinterpret_cast<const int4 *>(blk + i*kPitch + 4);
    int v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
    for (int j = 0; j < 8; j++) {
      s0[dir_line<D0>(i, j)] += v[j];
      s1[dir_line<D1>(i, j)] += v[j];
    }
  }
  c0 = lines_cost<D0>(s0);
  c1 = lines_cost<D1>(s1);
}

__constant__ int kThreshQ8[18] = {  // reference src/filter.c:1803-1807 (formula in its comment)
    128, 134, 150, 168, 188, 210, 234, 262, 292, 327, 365, 408, 455, 509, 569, 635, 710, 768};
// (dy, dx) of the three taps of each direction (reference direction_offsets_table,
// src/filter.c:142-151)
__constant__ signed char kDirStep[8][3][2] = {
    {{-1, 1}, {-2, 2}, {-3, 3}}, {{0, 1}, {-1, 2}, {-1, 3}}, {{0, 1}, {0, 2}, {0, 3}},
    {{0, 1}, {1, 2}, {1, 3}},    {{1, 1}, {2, 2}, {3, 3}},   {{1, 0}, {2, 1}, {3, 1}},
    {{1, 0}, {2, 0}, {3, 0}},    {{1, 0}, {2, -1}, {3, -1}}};

// p if |p| < thr, else 0, with one unsigned compare: |p| < thr  <=>  (unsigned)(p + thr - 1) <
// 2*thr - 1; the bound is clamped at 0 so that a threshold of 0 accepts nothing.
struct Gate { int c; unsigned lim; };
__device__ __forceinline__ Gate gate(int thr) {
  Gate g;
  g.c = thr - 1;
  g.lim = (unsigned)max(2*thr - 1, 0);
  return g;
}
__device__ __forceinline__ int tap(int p, Gate g) { return (unsigned)(p + g.c) < g.lim ? p : 0; }
// The same shifted by c, so that the subtraction of the centre pixel folds into one constant
// per pixel: q = the neighbour, cx = c - centre; returns tap(q - centre) + c (a rejected tap
// counts as c; the caller takes the c's off the weighted sum in one go).
__device__ __forceinline__ int tapc(int q, int cx, Gate g) {
  const int t = q + cx;
  return (unsigned)t < g.lim ? t : g.c;
}

}  // namespace

// Tile offset of tap K of direction DIR (the table above, as a compile-time value).
template <int DIR, int K, int PITCH> struct DirOff {
  static constexpr int dy = DIR == 0 ? -(K + 1) : DIR == 1 ? (K == 0 ? 0 : -1) : DIR == 2 ? 0 :
                            DIR == 3 ? (K == 0 ? 0 : 1) : K + 1;
  static constexpr int dx = DIR <= 3 ? K + 1 : DIR == 4 ? K + 1 : DIR == 5 ? (K == 0 ? 0 : 1) :
                            DIR == 6 ? 0 : (K == 0 ? 0 : -1);
  static constexpr int value = dy*PITCH + dx;
};

// Pass 1 of one 8x8 block by one warp: this lane's two pixels (a, a + 4 rows).
template <int DIR>
__device__ __forceinline__ void pass1_block(const int *a, int *b, int thr) {
  constexpr int kPitch = Tile<0>::pitch;
  constexpr int o0 = DirOff<DIR, 0, kPitch>::value, o1 = DirOff<DIR, 1, kPitch>::value,
                o2 = DirOff<DIR, 2, kPitch>::value;
  const Gate g = gate(thr);
  const int k1 = 8 - 14*g.c;            // rounding offset less the c's of the 3*2 + 2*4 tap weights
#pragma unroll
  for (int h = 0; h < 2; h++) {
    const int *q = a + h*4*kPitch;
    const int xx = ld16(q);
    const int cx = g.c - xx;
    const int s0 = tapc(ld16(q + o0), cx, g) + tapc(ld16(q - o0), cx, g);
    const int s1 = tapc(ld16(q + o1), cx, g) + tapc(ld16(q - o1), cx, g) +
                   tapc(ld16(q + o2), cx, g) + tapc(ld16(q - o2), cx, g);
    b[h*4*kPitch] = (short)(xx + ((3*s0 + 2*s1 + k1) >> 4));
  }
}

// Pass 2 of one 8x8 block by one warp; O = tile offset across the direction.
template <int O>
__device__ __forceinline__ void pass2_block(const int *a, const int *b, int *g, int w, int thr) {
  constexpr int kPitch = Tile<0>::pitch;
  const int third = thr/3;
#pragma unroll
  for (int h = 0; h < 2; h++) {
    const int *bq = b + h*4*kPitch;
    const int yy = ld16(bq);
    const Gate ga = gate(min(thr, third + abs(yy - ld16(a + h*4*kPitch))));
    const int cy = ga.c - yy;
    const int sum = tapc(ld16(bq + O), cy, ga) + tapc(ld16(bq - O), cy, ga) +
                    tapc(ld16(bq + 2*O), cy, ga) + tapc(ld16(bq - 2*O), cy, ga);
    g[h*4*(long long)w] = (short)(yy + ((3*sum + 8 - 12*ga.c) >> 4));
  }
}

// One plane of one filtered superblock.  DEC is the plane's decimation: n = 32 >> DEC
// samples a side, filter blocks of 8 >> DEC.  The plane's tile is on its way into A (bar,
// parity); B is free.  s_dir[], s_lthr[]: directions and thresholds of the 16 luma blocks.
template <int DEC>
__device__ __forceinline__ void dering_plane(const DeringParams &P, int p, int frame, int sbx,
                                             int sby, int *A, int *B, void *bar, int parity,
                                             const int *s_dir, const int *s_lthr) {
  using T = Tile<DEC>;
  constexpr int n = T::n, kPitch = T::pitch;
  const int t = threadIdx.x;
  const int w = (P.nhsb*kSB) >> DEC;
  int *dst = P.out[p] + frame*P.plane_size[p] + (long long)(sby*n)*w + sbx*n;
  int *Ai = A + T::origin;                              // interior origin
  int *Bi = B + T::origin;
  // Skip test first, so that its loads are in flight under the tile's: a block whose own
  // 4x4 units and the ring of units around them (lapping reaches there) are all skipped is
  // left alone (reference src/filter.c:1899-1920; the ring is cut at the frame border only).
  // Eight lanes per block, two flags of the 4x4 window each.  Warp w tests the blocks 4w ..
  // 4w + 3, the blocks it filters below: byte k of its ballot = block 4w + k.
  unsigned coded_mask;
  {
    constexpr int u = 2 >> DEC;
    const int b = t >> 3;
    const unsigned char *sk = P.bskip[p] + (long long)frame*P.skip_frame +
                              (long long)((sby << (3 - DEC)) + (b >> 2)*u)*P.skip_stride +
                              (sbx << (3 - DEC)) + (b & 3)*u;
    const int lo_i = sby ? -1 : 0, hi_i = u + (sby != P.nvsb - 1);
    const int lo_j = sbx ? -1 : 0, hi_j = u + (sbx != P.nhsb - 1);
    const int wi = ((t & 7) >> 1) - 1;
    int coded = 0;
#pragma unroll
    for (int e = 0; e < 2; e++) {
      const int wj = (t & 1)*2 + e - 1;
      if (wi >= lo_i && wi < hi_i && wj >= lo_j && wj < hi_j) coded |= !__ldg(sk + wi*P.skip_stride + wj);
    }
    coded_mask = __ballot_sync(0xffffffffu, coded);
  }
  mbar_wait(bar, parity);
  // What lies outside the frame arrives as zeros and becomes "far away" (reference
  // OD_DERING_VERY_LARGE): whole rows and 4-word column groups of border superblocks only.
  {
    const bool top = sby == 0, bottom = sby == P.nvsb - 1;
    const bool left = sbx == 0, right = sbx == P.nhsb - 1;
    if (top || bottom || left || right) {               // uniform over the CTA
      for (int i = t; i < T::words; i += kThreads) {
        const int r = i/kPitch, c = i - r*kPitch;
        const bool outside = (top && r < kBorder) || (bottom && r >= n + kBorder) ||
                             (left && c < kPadL) || (right && c >= n + kPadL);
        if (outside) A[i] = kFar;
      }
      __syncthreads();
    }
  }
  // B's ring (pass 2 reaches two samples out of the superblock, straight up / down / left /
  // right) = the unfiltered samples
  if (t < n) {                                          // rows -2, -1, n, n + 1
    const int rs = t/(n/4), cq = t - rs*(n/4);
    const int rr = rs < 2 ? rs - 2 : n + rs - 2;
    const int o = (rr + kBorder)*kPitch + kPadL + 4*cq;
    *reinterpret_cast<int4 *>(B + o) = *reinterpret_cast<const int4 *>(A + o);
  }
  else if (t < 3*n) {                                   // columns -2, -1 and n, n + 1
    const int k = t - n, rr = k >> 1;
    const int o = (rr + kBorder)*kPitch + ((k & 1) ? n + kPadL : kPadL - 2);
    *reinterpret_cast<int2 *>(B + o) = *reinterpret_cast<const int2 *>(A + o);
  }
  __syncthreads();
  // thresholds: luma's come from the direction search (dering_dir_kernel, s_lthr[]), chroma's
  // are flat (reference src/filter.c:1877-1897); a block that is all skipped passes through
  const int flat_thr = P.thr[p];
  auto blk_thr = [&](int b) {
    return ((coded_mask >> (8*(b & 3))) & 0xffu) ? (p == 0 ? s_lthr[b] : flat_thr) : 0;
  };
  // ---- the two passes.  A block with threshold 0 passes through both unchanged.
  // 8x8 blocks: one warp per block at a time (two pixels per lane, four rows apart), so the
  // block's direction and threshold are warp-uniform and the tap offsets compile-time
  // constants.  Chroma of 4:2:0 (4x4 blocks): half a warp per block, offsets from the table.
  const int warp = t >> 5, lane = t & 31;
  if (DEC == 0) {
    for (int k = 0; k < 4; k++) {
      const int b = warp*4 + k;
      const int thr = blk_thr(b);
      const int off = ((b >> 2)*8 + (lane >> 3))*kPitch + (b & 3)*8 + (lane & 7);
      if (thr == 0) {
        Bi[off] = ld16(Ai + off);
        Bi[off + 4*kPitch] = ld16(Ai + off + 4*kPitch);
        continue;
      }
      switch (s_dir[b]) {
        case 0: pass1_block<0>(Ai + off, Bi + off, thr); break;
        case 1: pass1_block<1>(Ai + off, Bi + off, thr); break;
        case 2: pass1_block<2>(Ai + off, Bi + off, thr); break;
        case 3: pass1_block<3>(Ai + off, Bi + off, thr); break;
        case 4: pass1_block<4>(Ai + off, Bi + off, thr); break;
        case 5: pass1_block<5>(Ai + off, Bi + off, thr); break;
        case 6: pass1_block<6>(Ai + off, Bi + off, thr); break;
        default: pass1_block<7>(Ai + off, Bi + off, thr); break;
      }
    }
    __syncthreads();
    // (a warp's four blocks lie side by side: block row `warp`)
    int *grow = dst + (long long)(warp*8 + (lane >> 3))*w + (lane & 7);
    for (int k = 0; k < 4; k++) {
      const int b = warp*4 + k;
      const int thr = blk_thr(b);
      const int off = (warp*8 + (lane >> 3))*kPitch + k*8 + (lane & 7);
      int *g = grow + k*8;
      if (thr == 0) {
        g[0] = Bi[off];
        g[4*(long long)w] = Bi[off + 4*kPitch];
      }
      else if (s_dir[b] <= 4) pass2_block<kPitch>(Ai + off, Bi + off, g, w, thr);
      else pass2_block<1>(Ai + off, Bi + off, g, w, thr);
    }
  }
  else {
    const int half = lane >> 4, l16 = lane & 15;
    for (int k = 0; k < 2; k++) {
      const int b = (warp*2 + k)*2 + half;              // two horizontally adjacent blocks per warp
      const int thr = blk_thr(b);
      const int off = ((b >> 2)*4 + (l16 >> 2))*kPitch + (b & 3)*4 + (l16 & 3);
      const int xx = ld16(Ai + off);
      int yy = xx;
      if (thr) {
        const int *a = Ai + off;
        const int dir = s_dir[b];
        const int o0 = kDirStep[dir][0][0]*kPitch + kDirStep[dir][0][1];
        const int o1 = kDirStep[dir][1][0]*kPitch + kDirStep[dir][1][1];
        const int o2 = kDirStep[dir][2][0]*kPitch + kDirStep[dir][2][1];
        const Gate g = gate(thr);
        const int cx = g.c - xx;
        const int s0 = tapc(ld16(a + o0), cx, g) + tapc(ld16(a - o0), cx, g);
        const int s1 = tapc(ld16(a + o1), cx, g) + tapc(ld16(a - o1), cx, g) +
                       tapc(ld16(a + o2), cx, g) + tapc(ld16(a - o2), cx, g);
        yy = (short)(xx + ((3*s0 + 2*s1 + 8 - 14*g.c) >> 4));
      }
      Bi[off] = yy;
    }
    __syncthreads();
    for (int k = 0; k < 2; k++) {
      const int b = (warp*2 + k)*2 + half;
      const int thr = blk_thr(b);
      const int r = (b >> 2)*4 + (l16 >> 2), c = (b & 3)*4 + (l16 & 3);
      const int *bq = Bi + r*kPitch + c;
      const int yy = ld16(bq);
      int out = yy;
      if (thr) {
        const int o = s_dir[b] <= 4 ? kPitch : 1;
        const Gate ga = gate(min(thr, thr/3 + abs(yy - ld16(Ai + r*kPitch + c))));
        const int cy = ga.c - yy;
        const int sum = tapc(ld16(bq + o), cy, ga) + tapc(ld16(bq - o), cy, ga) +
                        tapc(ld16(bq + 2*o), cy, ga) + tapc(ld16(bq - 2*o), cy, ga);
        out = (short)(yy + ((3*sum + 8 - 12*ga.c) >> 4));
      }
      dst[(long long)r*w + c] = out;
    }
  }
}

#ifndef TDLT_DERING_MINB
#define TDLT_DERING_MINB 10
#endif
__global__ void __launch_bounds__(kThreads, TDLT_DERING_MINB)
dering_kernel(const __grid_constant__ DeringMaps maps, const __grid_constant__ DeringParams P,
              const int *__restrict__ dirthr) {
  __shared__ __align__(128) int A[2][kTileWords];
  __shared__ __align__(16) int B[kTileWords];
  __shared__ __align__(8) unsigned long long bars[2];
  __shared__ int s_dir[16];
  __shared__ int s_lthr[16];
  const int t = threadIdx.x;
  // grid (superblocks of a frame, frames); the row of a superblock by multiply-high
  const int sb = blockIdx.x, frame = blockIdx.y;
  const int sby = div_recip(sb, P.recip_nhsb), sbx = sb - sby*P.nhsb;
  const bool filtered = P.flags[(long long)frame*gridDim.x + sb] != 0;
  if (!filtered) {
    for (int p = 0; p < P.nplanes; p++) {
      // copy through: n rows of n words as 16-byte vectors
      const int n = kSB >> P.dec[p];
      const int w = (P.nhsb*kSB) >> P.dec[p];
      const long long o = frame*P.plane_size[p] + (long long)(sby*n)*w + sbx*n;
      const int lv = 3 - P.dec[p];                      // log2 of the vectors per row
      for (int i = t; i < n << lv; i += kThreads) {
        const int r = i >> lv, c = i & ((1 << lv) - 1);
        reinterpret_cast<int4 *>(P.out[p] + o + (long long)r*w)[c] =
            __ldg(reinterpret_cast<const int4 *>(P.in[p] + o + (long long)r*w) + c);
      }
    }
    return;
  }
  // One thread asks for the planes' tiles: plane 0 now, plane p + 1 when plane p - 1 is done
  // with its buffer.  The box starts 3 rows above and 4 words left of the superblock (a
  // 16-byte boundary); what it takes from outside the plane reads as zero.
  auto fetch = [&](int p) {
    const int dec = P.dec[p], n = kSB >> dec;
    const int bytes = (dec ? Tile<1>::words : Tile<0>::words)*4;
    mbar_expect_tx(&bars[p & 1], bytes);
    tma_load_3d(A[p & 1], &maps.in[p], &bars[p & 1], sbx*n - kPadL, sby*n - kBorder, frame);
  };
  if (t == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    fence_mbar_init();
    fetch(0);
  }
  // the luma blocks' directions and thresholds (dering_dir_kernel): dir | thr << 3
  if (t < 16) {
    const int dt = __ldg(dirthr + ((long long)frame*gridDim.x + sb)*16 + t);
    s_dir[t] = dt & 7;
    s_lthr[t] = dt >> 3;
  }
  __syncthreads();
  for (int p = 0; p < P.nplanes; p++) {
    if (p > 0) {
      // plane p - 1 is done with its A and with B; its border fill wrote A through the
      // generic proxy, the next box arrives through the async one
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncthreads();
    }
    if (t == 0 && p + 1 < P.nplanes) fetch(p + 1);
    const int parity = (p >> 1) & 1;
    if (P.dec[p] == 0) dering_plane<0>(P, p, frame, sbx, sby, A[p & 1], B, &bars[p & 1], parity, s_dir, s_lthr);
    else dering_plane<1>(P, p, frame, sbx, sby, A[p & 1], B, &bars[p & 1], parity, s_dir, s_lthr);
  }
}

// Direction search of the luma 8x8 blocks (reference od_dir_find8 + od_compute_thresh,
// src/filter.c:1655-1697, 1816-1833, called from od_dering :1865-1897): one CTA per PAIR of
// horizontally adjacent superblocks, so that a warp runs one direction for 32 blocks at a
// time (warp w: directions 2w and 2w + 1) with every lane busy.  Result per block: dir |
// thr << 3 (threshold before the skip test) -> dirthr[(frame*nsb + sb)*16 + block].
constexpr int kDirPitch = 2*kSB;
constexpr int kBlkPitch = 68;          // words between the 8x8 blocks of the regrouped copy
__global__ void __launch_bounds__(kThreads)
dering_dir_kernel(const __grid_constant__ DeringMaps maps, const __grid_constant__ DeringParams P,
                  int *__restrict__ dirthr) {
  __shared__ __align__(128) int T[kSB*kDirPitch];       // the pair's luma samples, row by row
  __shared__ __align__(16) int S[32*kBlkPitch];         // the same as pixels, block by block
  __shared__ __align__(8) unsigned long long bar;
  __shared__ int s_cost[32][9];
  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  const int npair = (P.nhsb + 1) >> 1;
  const int sby = div_recip((int)blockIdx.x, P.recip_npair), px = (int)blockIdx.x - sby*npair;
  const int frame = blockIdx.y;
  const int sb0 = sby*P.nhsb + 2*px;
  const bool two = 2*px + 1 < P.nhsb;
  const unsigned char *fl = P.flags + (long long)frame*P.nhsb*P.nvsb + sb0;
  if (!fl[0] && !(two && fl[1])) return;                // uniform
  if (t == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
    mbar_expect_tx(&bar, kSB*kDirPitch*4);
    tma_load_3d(T, &maps.dir, &bar, 2*px*kSB, sby*kSB, frame);
  }
  __syncthreads();
  mbar_wait(&bar, 0);
  // pixel scale (the reference searches (int16_t)x >> OD_COEFF_SHIFT), regrouped block by
  // block: block (superblock s, row by, column bx) = 64 words at S + (16 s + 4 by + bx)*68.
  // With 17 16-byte chunks between blocks the eight lanes of a quarter warp, each reading
  // the same chunk of its own block, hit eight different bank groups.
#pragma unroll
  for (int it = 0; it < kSB*kDirPitch/4/kThreads; it++) {
    const int i = t + it*kThreads, r = i >> 4, c = i & 15;          // row, 16-byte chunk of the row
    int4 v = reinterpret_cast<const int4 *>(T)[i];
    v.x = (short)v.x >> 4; v.y = (short)v.y >> 4; v.z = (short)v.z >> 4; v.w = (short)v.w >> 4;
    const int blk = ((c >> 3) << 4) + ((r >> 3) << 2) + ((c >> 1) & 3);
    *reinterpret_cast<int4 *>(S + blk*kBlkPitch + (r & 7)*8 + (c & 1)*4) = v;
  }
  __syncthreads();
  {
    const int *bp = S + lane*kBlkPitch;                 // lane = block: 16*superblock + 4*by + bx
    int c0, c1;
    switch (warp) {                                     // uniform
      case 0: dir_cost2<0, 1, 8>(bp, c0, c1); break;
      case 1: dir_cost2<2, 3, 8>(bp, c0, c1); break;
      case 2: dir_cost2<4, 5, 8>(bp, c0, c1); break;
      default: dir_cost2<6, 7, 8>(bp, c0, c1); break;
    }
    s_cost[lane][2*warp] = c0;
    s_cost[lane][2*warp + 1] = c1;
  }
  __syncthreads();
  if (t < 32) {
    int best = 0, best_cost = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
      int ck = s_cost[t][k];
      if (ck > best_cost) { best_cost = ck; best = k; }
    }
    const int var = best_cost - s_cost[t][(best + 4) & 7];
    int varsum = var;                                   // over the 16 blocks of the lane's superblock
#pragma unroll
    for (int o = 8; o >= 1; o >>= 1) varsum += __shfl_xor_sync(0xffffffffu, varsum, o);
    // reference od_compute_thresh, src/filter.c:1816-1833
    int v1 = min(32767, var >> 6);
    int v2 = min(32767, varsum/(kSB*kSB));
    unsigned prod = (unsigned)(v1*v2);
    int e = (32 - __clz(prod)) - 9;
    e = max(0, min(17, e));
    const int thr = (P.thr[0]*kThreshQ8[e]) >> 8;
    if (t < 16 || two) {
      dirthr[((long long)frame*P.nhsb*P.nvsb + sb0 + (t >> 4))*16 + (t & 15)] = best | (thr << 3);
    }
  }
}

void launch_dering(const DeringMaps &maps, const DeringParams &P, int *dirthr, int nframes,
                   cudaStream_t st) {
  const int npair = (P.nhsb + 1) >> 1;
  dering_dir_kernel<<<dim3((unsigned)(npair*P.nvsb), (unsigned)nframes), kThreads, 0, st>>>(maps, P, dirthr);
  dering_kernel<<<dim3((unsigned)(P.nhsb*P.nvsb), (unsigned)nframes), kThreads, 0, st>>>(maps, P, dirthr);
}

}  // namespace tdlt

// ---------------------------------------------------------------------------
// Keyframe bilinear smoothing (reference od_bilinear_smooth, src/filter.c:1952-2008, applied
// by od_smooth_recursive with min_bs = 32x32 from src/decode.c:1134-1146): every superblock
// coded as ONE 32x32 block is pulled towards the bilinear interpolation of its four
// corners, by a weight that falls with the distortion this causes relative to the
// quantization noise.  In place; work unit = (superblock, frame, plane), one warp per unit; a
// warp takes whole rows (the distortion is truncated per row before it is summed).
// ---------------------------------------------------------------------------
namespace tdlt {

namespace {
__constant__ int kBilinearStrength[4] = {5, 20, 20, 5};   // reference src/filter.c:154

// One unit by ONE warp, two passes over its rows (distortion, then update; the second pass
// re-reads the block from L1/L2): no barrier and no shared memory, so the four warps of a
// CTA smooth four units side by side.
template <int DEC>
__device__ __forceinline__ void smooth_block(int *x, int w, int q, int pli) {
  constexpr int ln = 5 - DEC, n = 1 << ln;
  constexpr int shift = 2*4 + 2*ln - 16 > 0 ? 2*4 + 2*ln - 16 : 0;   // OD_COEFF_SHIFT = 4
  constexpr int rows_per_step = 32/n;                   // one row per warp step (luma) or two (chroma)
  constexpr int steps = n/rows_per_step;
  const int lane = threadIdx.x & 31;
  const int j = lane & (n - 1), i0 = lane >> ln;
  const int x00 = x[0], x01 = x[n - 1], x10 = x[(long long)(n - 1)*w],
            x11 = x[(long long)(n - 1)*w + n - 1];
  int a01 = x01 - x00, a10 = x10 - x00, a11 = x11 + x00 - x10 - x01;
  a01 += (a01 + n/2) >> ln;
  a10 += (a10 + n/2) >> ln;
  a11 += (2*a10 + n/2) >> ln;
  int dist = 0;
#pragma unroll 8
  for (int s = 0; s < steps; s++) {
    const int i = i0 + s*rows_per_step;
    const int xv = x[(long long)i*w + j];
    const int yv = x00 + ((j*a01 + i*a10 + ((j*i*a11) >> ln) + n/2) >> ln);
    const int d = yv - xv;
    int row = d*d;
#pragma unroll
    for (int o = n/2; o >= 1; o >>= 1) row += __shfl_xor_sync(0xffffffffu, row, o);
    dist += row >> shift;                               // truncated per row, as the reference does
  }
  int total = dist;                                     // every lane holds its row group's sum
  if (rows_per_step == 2) total += __shfl_xor_sync(0xffffffffu, dist, 16);
  total += n/2;
  total >>= 2*ln - shift;
  int wgt = min(1024, kBilinearStrength[pli]*q*q/(1 + 12*total));
  wgt = (wgt*wgt) >> 12;
#pragma unroll 8
  for (int s = 0; s < steps; s++) {
    const int i = i0 + s*rows_per_step;
    int *px = x + (long long)i*w + j;
    const int xv = *px;
    const int yv = x00 + ((j*a01 + i*a10 + ((j*i*a11) >> ln) + n/2) >> ln);
    *px = xv - ((wgt*(xv - yv) + 128) >> 8);
  }
}
}  // namespace

__global__ void __launch_bounds__(kThreads) smooth_kernel(SmoothParams P, int nunits) {
  __shared__ int s_list[kThreads];
  __shared__ int s_count;
  // Work unit = (superblock, frame, plane).  Most superblocks are split and have nothing to
  // do, so a fixed-size grid walks the units 128 at a time: every thread looks one unit's map
  // entry up (the loads of a chunk are in flight together), the units that are one 32x32
  // block are collected in shared memory, then each warp smooths one of them at a time.
  const int t = threadIdx.x;
  const int per = P.nhsb*P.nvsb;
  for (int base = blockIdx.x*kThreads; base < nunits; base += gridDim.x*kThreads) {
    if (t == 0) s_count = 0;
    __syncthreads();
    const int unit = base + t;
    if (unit < nunits) {
      const int fs = unit/P.nplanes;
      const int frame = fs/per;
      const int sb = fs - frame*per;
      const int sby = sb/P.nhsb, sbx = sb - sby*P.nhsb;
      // the superblock is one 32x32 block when its first map entry says so
      if (__ldg(P.bsize + (long long)frame*per*16 + (long long)(4*sby)*(4*P.nhsb) + 4*sbx) == 3) {
        s_list[atomicAdd(&s_count, 1)] = unit;
      }
    }
    __syncthreads();
    const int count = s_count;
    for (int k = t >> 5; k < count; k += kThreads/32) {   // one unit per warp
      const int u = s_list[k];
      const int p = u % P.nplanes;
      const int fs = u/P.nplanes;
      const int frame = fs/per;
      const int sb = fs - frame*per;
      const int sby = sb/P.nhsb, sbx = sb - sby*P.nhsb;
      const int dec = P.dec[p];
      const int n = kSB >> dec;
      const int w = (P.nhsb*kSB) >> dec;
      int *x = P.c[p] + frame*P.plane_size[p] + (long long)(sby*n)*w + sbx*n;
      if (dec == 0) smooth_block<0>(x, w, P.q[p], p);
      else smooth_block<1>(x, w, P.q[p], p);
    }
    __syncthreads();                                    // s_list, s_count free for the next chunk
  }
}

void launch_smooth(const SmoothParams &P, int nframes, cudaStream_t st) {
  const long long nunits = (long long)P.nhsb*P.nvsb*nframes*P.nplanes;
  const long long chunks = (nunits + kThreads - 1)/kThreads;
  const int grid = (int)(chunks < 148*16 ? chunks : 148*16);    // up to 16 CTAs of 128 threads per SM
  smooth_kernel<<<grid, kThreads, 0, st>>>(P, (int)nunits);
}

}  // namespace tdlt

// ---------------------------------------------------------------------------
// The two filters with the reference's own signatures, one block (the entries of its
// dispatch table, od_filter_dering_direction / _orthogonal, src/filter.h:46-65): parity
// surface behind the *_cuda calls, one thread per pixel.  `in`: the block with its 3-sample
// ring, dense, pitch n + 6.
// ---------------------------------------------------------------------------
namespace tdlt {

__global__ void dering_block_kernel(short *y, const short *in, const short *x, int n, int thr,
                                    int dir, int orthogonal) {
  const int i = threadIdx.x/n, j = threadIdx.x - i*n;
  if (i >= n) return;
  const int pitch = n + 2*kBorder;
  const short *a = in + (i + kBorder)*pitch + j + kBorder;
  const int xx = a[0];
  if (!orthogonal) {
    const Gate g = gate(thr);
    int s[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
      const int oo = kDirStep[dir][k][0]*pitch + kDirStep[dir][k][1];
      s[k] = tap(a[oo] - xx, g) + tap(a[-oo] - xx, g);
    }
    y[i*n + j] = (short)(xx + ((3*s[0] + 2*(s[1] + s[2]) + 8) >> 4));
  }
  else {
    const int o = dir <= 4 ? pitch : 1;
    const Gate g = gate(min(thr, thr/3 + abs(xx - x[i*n + j])));
    const int sum = tap(a[o] - xx, g) + tap(a[-o] - xx, g) + tap(a[2*o] - xx, g) + tap(a[-2*o] - xx, g);
    y[i*n + j] = (short)(xx + ((3*sum + 8) >> 4));
  }
}

void launch_dering_block(short *y, const short *in, const short *x, int n, int thr, int dir,
                         bool orthogonal, cudaStream_t st) {
  dering_block_kernel<<<1, 64, 0, st>>>(y, in, x, n, thr, dir, orthogonal ? 1 : 0);
}

}  // namespace tdlt

// ---------------------------------------------------------------------------
// Raster <-> coding order (reference od_raster_to_coding_order /
// od_coding_order_to_raster, src/partition.c:95-200): every transform block's coefficients
// permuted into the order the PVQ coder takes them in.  The scan tables are data of the
// caller (od_tdlt_set_scan_tables); the vectors go to a block-linear plane: superblocks in
// raster order, inside a superblock the 4x4-sample units in Z order, 16 entries per unit, so
// that a quadtree-aligned n x n block owns n*n consecutive entries.  One thread per entry of
// that plane: to_coding gathers (coalesced writes), the way back scatters (coalesced reads).
// Entries the scans do not cover (the upper half of a 32x32 block's vector) are left alone,
// as the reference leaves them.  A fixed grid walks the superblocks; the tables sit in
// shared memory.
// ---------------------------------------------------------------------------
namespace tdlt {

namespace {
constexpr int kScanEntries = 15 + 48 + 192 + 768;
constexpr int kCoThreads = 256;
}  // namespace

// Where entry e of a superblock's block-linear range comes from when the block it lies in
// has a side of 4 << ln4 samples: position inside the superblock (row*S + column), or kNoPos
// when the codec does not order that entry.  Depends on nothing but (e, ln4) and the scan
// tables, so every CTA tabulates it once in shared memory.
constexpr unsigned short kNoPos = 0xffff;
template <int DEC>
__device__ __forceinline__ unsigned short coding_source_pos(int e, int ln4, const unsigned short *s_scan,
                                                            const int *scan_len) {
  constexpr int S = kSB >> DEC;
  const int z = e >> 4;
  // Z order -> unit coordinates (4x4 samples of this plane)
  const int ux = (z & 1) | ((z >> 1) & 2) | ((z >> 2) & 4);
  const int uy = ((z >> 1) & 1) | ((z >> 2) & 2) | ((z >> 3) & 4);
  const int zb = z & ~((1 << 2*ln4) - 1);               // the block's first unit in Z order
  const int i = ((z - zb) << 4) | (e & 15);             // index in the block's vector
  int x = 0, y = 0;
  if (i) {
    const int k = i < 16 ? 0 : i < 64 ? 1 : i < 256 ? 2 : 3;
    const int first = k == 0 ? 1 : 4 << 2*k;             // 1, 16, 64, 256
    const int j = i - first;
    if (j >= scan_len[k]) return kNoPos;
    const int toff = k == 0 ? 0 : k == 1 ? 15 : k == 2 ? 63 : 255;
    const unsigned short xy = s_scan[toff + j];
    x = xy & 255;
    y = xy >> 8;
  }
  const int bx = (ux & ~((1 << ln4) - 1)) << 2, by = (uy & ~((1 << ln4) - 1)) << 2;
  return (unsigned short)((by + y)*S + bx + x);
}

// One plane of NSB superblocks at once: S*S entries each, S*S/256 per thread and superblock,
// all of a thread's loads in flight before its first store (the kernel is bound by the
// bytes in flight per SM).  mo[]: this thread's map offsets (fixed per thread).
template <int DEC, int NSB>
__device__ __forceinline__ void coding_order_sb(int to_coding, const unsigned short *tab, const int *mo,
                                                const unsigned char *const map[], int *const v[],
                                                int *const d[], int w) {
  constexpr int S = kSB >> DEC;
  constexpr int lS = 5 - DEC;
  constexpr int iters = S*S/kCoThreads;
  int off[NSB][iters];
  int val[NSB][iters];
#pragma unroll
  for (int u = 0; u < NSB; u++) {
#pragma unroll
    for (int it = 0; it < iters; it++) {
      const int e = threadIdx.x + it*kCoThreads;
      const int ln4 = max((int)__ldg(map[u] + mo[it]) - DEC, 0);   // log2 of the block side in 4-sample units
      const int pos = tab[ln4*S*S + e];
      off[u][it] = pos == kNoPos ? -1 : (pos >> lS)*w + (pos & (S - 1));
      if (to_coding) val[u][it] = off[u][it] >= 0 ? d[u][off[u][it]] : 0;
      else val[u][it] = v[u][e];
    }
  }
#pragma unroll
  for (int u = 0; u < NSB; u++) {
#pragma unroll
    for (int it = 0; it < iters; it++) {
      const int e = threadIdx.x + it*kCoThreads;
      if (off[u][it] < 0) continue;                     // left alone, as the reference leaves it
      if (to_coding) v[u][e] = val[u][it];
      else d[u][off[u][it]] = val[u][it];
    }
  }
}

__global__ void __launch_bounds__(kCoThreads) coding_order_kernel(CodingParams P, int nunits) {
  __shared__ unsigned short s_scan[kScanEntries];       // (x, y) byte pairs
  __shared__ unsigned short s_tab0[4*kSB*kSB];          // planes at full size: block sides 4..32
  __shared__ unsigned short s_tab1[3*(kSB/2)*(kSB/2)];  // 4:2:0 chroma: block sides 4..16
  __shared__ int s_len[4];
  const int t = threadIdx.x;
  for (int i = t; i < kScanEntries; i += kCoThreads) {
    s_scan[i] = reinterpret_cast<const unsigned short *>(P.scan)[i];
  }
  if (t < 4) s_len[t] = P.scan_len[t];
  __syncthreads();
  for (int i = t; i < 4*kSB*kSB; i += kCoThreads) {
    s_tab0[i] = coding_source_pos<0>(i & (kSB*kSB - 1), i >> 10, s_scan, s_len);
  }
  for (int i = t; i < 3*(kSB/2)*(kSB/2); i += kCoThreads) {
    s_tab1[i] = coding_source_pos<1>(i & 255, i >> 8, s_scan, s_len);
  }
  // this thread's entries and the map entries they depend on, the same for every superblock
  int mo0[kSB*kSB/kCoThreads], mo1[1];
#pragma unroll
  for (int it = 0; it < kSB*kSB/kCoThreads; it++) {
    const int z = (t + it*kCoThreads) >> 4;
    const int ux = (z & 1) | ((z >> 1) & 2) | ((z >> 2) & 4);
    const int uy = ((z >> 1) & 1) | ((z >> 2) & 2) | ((z >> 3) & 4);
    mo0[it] = (uy >> 1)*(4*P.nhsb) + (ux >> 1);
  }
  {
    const int z = t >> 4;
    const int ux = (z & 1) | ((z >> 1) & 2), uy = ((z >> 1) & 1) | ((z >> 2) & 2);
    mo1[0] = uy*(4*P.nhsb) + ux;
  }
  __syncthreads();
  // walk the superblocks, two horizontal neighbours at a time (the frame is a multiple of 64
  // samples wide, so nhsb is even), without dividing: (frame, pair in frame) advance by the
  // grid size, the row by a multiply-high with the host's reciprocal (exact below 65536)
  const int per = P.nhsb*P.nvsb;
  const int per2 = per >> 1;
  int frame = blockIdx.x/per2;
  int pr = blockIdx.x - frame*per2;
  const int step_f = gridDim.x/per2, step_pr = gridDim.x - step_f*per2;
  for (int unit = blockIdx.x; unit < nunits; unit += gridDim.x) {
    const int sb = 2*pr;
    const int sby = div_recip(sb, P.recip_nhsb), sbx = sb - sby*P.nhsb;
    const unsigned char *map0 = P.bsize + (long long)frame*per*16 + (long long)(4*sby)*(4*P.nhsb) + 4*sbx;
    const unsigned char *const map[2] = {map0, map0 + 4};
    for (int p = 0; p < P.nplanes; p++) {
      const int dec = P.dec[p];
      const int S = kSB >> dec;
      const int w = (P.nhsb*kSB) >> dec;
      const long long fo = frame*P.plane_size[p];
      int *v0 = P.v[p] + fo + (long long)sb*S*S;
      int *d0 = P.d[p] + fo + (long long)(sby*S)*w + sbx*S;
      int *const v[2] = {v0, v0 + S*S};
      int *const d[2] = {d0, d0 + S};
      if (dec == 0) coding_order_sb<0, 2>(P.to_coding, s_tab0, mo0, map, v, d, w);
      else coding_order_sb<1, 2>(P.to_coding, s_tab1, mo1, map, v, d, w);
    }
    frame += step_f;
    pr += step_pr;
    if (pr >= per2) {
      pr -= per2;
      frame++;
    }
  }
}

// The same tabulation in global memory, once per od_tdlt_set_scan_tables, for the transform
// kernels' fused coding-order store / load (OD_TDLT_CODING_ORDER): 4 x 1024 entries for
// planes at full size, then 3 x 256 for 4:2:0 chroma (kCoTabEntries in tdlt_region.cuh).
__global__ void cotab_build_kernel(unsigned short *tab, const unsigned char *scan, int l0, int l1,
                                   int l2, int l3) {
  __shared__ unsigned short s_scan[kScanEntries];
  __shared__ int s_len[4];
  for (int i = threadIdx.x; i < kScanEntries; i += blockDim.x) {
    s_scan[i] = reinterpret_cast<const unsigned short *>(scan)[i];
  }
  if (threadIdx.x == 0) { s_len[0] = l0; s_len[1] = l1; s_len[2] = l2; s_len[3] = l3; }
  __syncthreads();
  for (int i = threadIdx.x; i < 4*kSB*kSB; i += blockDim.x) {
    tab[i] = coding_source_pos<0>(i & (kSB*kSB - 1), i >> 10, s_scan, s_len);
  }
  for (int i = threadIdx.x; i < 3*(kSB/2)*(kSB/2); i += blockDim.x) {
    tab[4*kSB*kSB + i] = coding_source_pos<1>(i & 255, i >> 8, s_scan, s_len);
  }
}
void launch_cotab_build(unsigned short *tab, const unsigned char *scan, const int lens[4],
                        cudaStream_t st) {
  cotab_build_kernel<<<1, 256, 0, st>>>(tab, scan, lens[0], lens[1], lens[2], lens[3]);
}

void launch_coding_order(const CodingParams &P, int nframes, cudaStream_t st) {
  const long long nunits = (long long)(P.nhsb/2)*P.nvsb*nframes;   // pairs of superblocks
  const long long cap = 148ll*12;                       // a fixed grid walks the superblocks (6 and 12
                                                        // CTAs per SM measure the same, 16+ is 8 % slower)
  const int grid = (int)(nunits < cap ? nunits : cap);
  coding_order_kernel<<<grid, kCoThreads, 0, st>>>(P, (int)nunits);
}

}  // namespace tdlt

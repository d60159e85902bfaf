// Deringing filter of the reconstructed planes: the step right after
// the inverse stage.  Reference: od_dering and its helpers, src/filter.c:1648-1941, driven
// per 32x32 superblock by the loop of src/decode.c:1044-1122 (encoder twin in
// src/encode.c).  Every superblock reads only the UNFILTERED frame (the reference's etmp
// copy), so superblocks are independent: one CTA per superblock and frame, all planes.
//
//   per plane (luma first; chroma reuses luma's directions):
//     load the (n+6)^2 tile -> A (int16 semantics; samples outside the frame = 30000)
//     luma: direction search of the 16 8x8 blocks (thread = block x direction), variances,
//           per-block thresholds; all planes: skip test of every block
//     pass 1: 6 taps along the block's direction          A -> B (interior)
//     pass 2: 4 taps across it, threshold tightened by |pass-1 change|   A,B -> out
//   superblocks whose flag is clear are copied through.
//
// HBM-bound by design: 4 bytes read + 4 written per sample; the 3-sample ring around a
// superblock is re-read from L2 (neighbouring CTAs are resident at the same time).
#include <cuda_runtime.h>
#include "tdlt_aux_kernels.h"

namespace tdlt {

namespace {

constexpr int kSB = 32;               // luma superblock (reference OD_BSIZE_MAX)
constexpr int kBorder = 3;            // reference OD_FILT_BORDER
constexpr int kPadL = 4;              // left pad of the tile: interior 16-byte aligned
constexpr int kPitch = 40;            // words per tile row
constexpr int kRows = kSB + 2*kBorder;
constexpr int kTile = kRows*kPitch;
constexpr int kFar = 30000;           // reference OD_DERING_VERY_LARGE
constexpr int kThreads = 128;

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

// Cost of direction D for the 8x8 block at blk (pixels already shifted down to pixel scale,
// the reference's x >> OD_COEFF_SHIFT; reference src/filter.c:1655-1697): the sum
// over the direction's lines of (line sum)^2 / (pixels on the line).  Fully unrolled: the
// line sums live in registers.
template <int D> __device__ __forceinline__ int dir_cost(const int *blk) {
  int s[15];
#pragma unroll
  for (int k = 0; k < 15; k++) s[k] = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    int4 a = *reinterpret_cast<const int4 *>(blk + i*kPitch);
    int4 b = *reinterpret_cast<const int4 *>(blk + i*kPitch + 4);
    int v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
    for (int j = 0; j < 8; j++) s[dir_line<D>(i, j)] += v[j];
  }
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

__constant__ int kThreshQ8[18] = {  // reference src/filter.c:1803-1807 (formula in its comment)
    128, 134, 150, 168, 188, 210, 234, 262, 292, 327, 365, 408, 455, 509, 569, 635, 710, 768};
// (dy, dx) of the three taps of each direction as tile offsets (reference
// direction_offsets_table, src/filter.c:142-151)
__constant__ int kDirOff[8][3] = {
    {-1*kPitch + 1, -2*kPitch + 2, -3*kPitch + 3}, {1, -1*kPitch + 2, -1*kPitch + 3},
    {1, 2, 3}, {1, 1*kPitch + 2, 1*kPitch + 3},
    {1*kPitch + 1, 2*kPitch + 2, 3*kPitch + 3}, {1*kPitch, 2*kPitch + 1, 3*kPitch + 1},
    {1*kPitch, 2*kPitch, 3*kPitch}, {1*kPitch, 2*kPitch - 1, 3*kPitch - 1}};

// the same taps as (dy, dx) steps, for the block calls whose buffers have another pitch
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

}  // namespace

// Tile offset of tap K of direction DIR (the table above, as a compile-time value).
template <int DIR, int K> struct DirOff {
  static constexpr int dy = DIR == 0 ? -(K + 1) : DIR == 1 ? (K == 0 ? 0 : -1) : DIR == 2 ? 0 :
                            DIR == 3 ? (K == 0 ? 0 : 1) : K + 1;
  static constexpr int dx = DIR <= 3 ? K + 1 : DIR == 4 ? K + 1 : DIR == 5 ? (K == 0 ? 0 : 1) :
                            DIR == 6 ? 0 : (K == 0 ? 0 : -1);
  static constexpr int value = dy*kPitch + dx;
};

// Pass 1 of one 8x8 luma block by one warp: this lane's two pixels (a, a + 4 rows).
template <int DIR>
__device__ __forceinline__ void pass1_block(const int *a, int *b, int thr) {
  constexpr int o0 = DirOff<DIR, 0>::value, o1 = DirOff<DIR, 1>::value, o2 = DirOff<DIR, 2>::value;
  const Gate g = gate(thr);
#pragma unroll
  for (int h = 0; h < 2; h++) {
    const int *q = a + h*4*kPitch;
    const int xx = q[0];
    int s0 = tap(q[o0] - xx, g) + tap(q[-o0] - xx, g);
    int s1 = tap(q[o1] - xx, g) + tap(q[-o1] - xx, g) + tap(q[o2] - xx, g) + tap(q[-o2] - xx, g);
    b[h*4*kPitch] = (short)(xx + ((3*s0 + 2*s1 + 8) >> 4));
  }
}

// Pass 2 of one 8x8 luma block by one warp; O = tile offset across the direction.
template <int O>
__device__ __forceinline__ void pass2_block(const int *a, const int *b, int *g, int w, int thr) {
  const int third = thr/3;
#pragma unroll
  for (int h = 0; h < 2; h++) {
    const int *bq = b + h*4*kPitch;
    const int yy = bq[0];
    const Gate ga = gate(min(thr, third + abs(yy - a[h*4*kPitch])));
    int sum = tap(bq[O] - yy, ga) + tap(bq[-O] - yy, ga) + tap(bq[2*O] - yy, ga) +
              tap(bq[-2*O] - yy, ga);
    g[h*4*(long long)w] = (short)(yy + ((3*sum + 8) >> 4));
  }
}

// One plane of one filtered superblock.  DEC is the plane's decimation: n = 32 >> DEC
// samples a side, filter blocks of 8 >> DEC.
template <int DEC>
__device__ __forceinline__ void dering_plane(const DeringParams &P, int p, int frame, int sbx,
                                             int sby, int *A, int *B, int (*s_cost)[8],
                                             int *s_dir, int *s_thr, int *s_coded) {
  constexpr int n = kSB >> DEC;
  constexpr int span = n + 2*kBorder;
  constexpr int lb = 3 - DEC;
  const int t = threadIdx.x;
  const int w = (P.nhsb*kSB) >> DEC;
  const int *src = P.in[p] + frame*P.plane_size[p] + (long long)(sby*n)*w + sbx*n;
  int *dst = P.out[p] + frame*P.plane_size[p] + (long long)(sby*n)*w + sbx*n;
  int *Ai = A + kBorder*kPitch + kPadL;                 // interior origin
  int *Bi = B + kBorder*kPitch + kPadL;
  // ---- load the tile, ring included, as 16-byte vectors: rows -3 .. n+2, columns -4 .. n+3
  // (the superblock starts on a 16-byte boundary).  What lies outside the frame does so in
  // whole vectors and rows and becomes "far away" (reference OD_DERING_VERY_LARGE).
  __syncthreads();                                      // previous plane done with A, B
  // Skip test first, so that its loads are in flight under the tile's: a block whose own
  // 4x4 units and the ring of units around them (lapping reaches there) are all skipped is
  // left alone (reference src/filter.c:1899-1920; the ring is cut at the frame border only).
  // Eight lanes per block, two flags of the 4x4 window each.
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
    const unsigned m = __ballot_sync(0xffffffffu, coded);
    if ((t & 7) == 0) s_coded[b] = (m >> (t & 24)) & 0xffu;
  }
  {
    constexpr int vrow = n/4 + 2;                       // vectors per tile row
    const bool top = sby == 0, bottom = sby == P.nvsb - 1;
    const bool left = sbx == 0, right = sbx == P.nhsb - 1;
    for (int i = t; i < span*vrow; i += kThreads) {
      const int r = i/vrow, c = i - r*vrow;
      const int rr = r - kBorder;
      const bool ring = rr < 0 || rr >= n || c == 0 || c == vrow - 1;
      const bool outside = (top && rr < 0) || (bottom && rr >= n) || (left && c == 0) ||
                           (right && c == vrow - 1);
      int4 v = make_int4(kFar, kFar, kFar, kFar);
      if (!outside) {
        v = __ldg(reinterpret_cast<const int4 *>(src + (long long)rr*w - 4) + c);
        v.x = (short)v.x; v.y = (short)v.y; v.z = (short)v.z; v.w = (short)v.w;  // the int16_t etmp copy
      }
      *reinterpret_cast<int4 *>(A + r*kPitch + 4*c) = v;
      if (p == 0 && !ring) {  // luma: pixel-scale copy for the direction search (B's interior is free until pass 1)
        v.x >>= 4; v.y >>= 4; v.z >>= 4; v.w >>= 4;
      }
      if (p == 0 || ring) *reinterpret_cast<int4 *>(B + r*kPitch + 4*c) = v;
    }
  }
  __syncthreads();
  // ---- luma: direction search -------------------------------------------------------------
  if (DEC == 0 && p == 0) {
    const int blk = t & 15, d = t >> 4;
    const int *bp = Bi + (blk >> 2)*8*kPitch + (blk & 3)*8;
    int c;
    switch (d) {
      case 0: c = dir_cost<0>(bp); break;
      case 1: c = dir_cost<1>(bp); break;
      case 2: c = dir_cost<2>(bp); break;
      case 3: c = dir_cost<3>(bp); break;
      case 4: c = dir_cost<4>(bp); break;
      case 5: c = dir_cost<5>(bp); break;
      case 6: c = dir_cost<6>(bp); break;
      default: c = dir_cost<7>(bp); break;
    }
    s_cost[blk][d] = c;
    __syncthreads();
    if (t < 32) {
      int var = 0;
      if (t < 16) {
        int best = 0, best_cost = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) {
          int ck = s_cost[t][k];
          if (ck > best_cost) { best_cost = ck; best = k; }
        }
        var = best_cost - s_cost[t][(best + 4) & 7];
        s_dir[t] = best;
      }
      int varsum = var;
#pragma unroll
      for (int o = 8; o >= 1; o >>= 1) varsum += __shfl_xor_sync(0xffffffffu, varsum, o);
      if (t < 16) {
        // reference od_compute_thresh, src/filter.c:1816-1833
        int v1 = min(32767, var >> 6);
        int v2 = min(32767, varsum/(kSB*kSB));
        unsigned prod = (unsigned)(v1*v2);
        int e = (32 - __clz(prod)) - 9;
        e = max(0, min(17, e));
        s_thr[t] = (P.thr[0]*kThreshQ8[e]) >> 8;
      }
    }
  }
  else if (t < 16) {
    s_thr[t] = P.thr[p];
  }
  if (t < 16 && !s_coded[t]) s_thr[t] = 0;            // all skipped: pass through
  __syncthreads();
  // ---- the two passes.  A block with threshold 0 passes through both unchanged.
  // Luma: one warp per 8x8 block at a time (two pixels per lane, four rows apart), so the
  // block's direction and threshold are warp-uniform and the tap offsets compile-time
  // constants; with a row pitch of 40 words the 4 x 8 pixels of a warp access fall into 32
  // different banks.  Chroma (4x4 blocks): half a warp per block, offsets from the table.
  const int warp = t >> 5, lane = t & 31;
  if (DEC == 0) {
    for (int k = 0; k < 4; k++) {
      const int b = warp*4 + k;
      const int thr = s_thr[b];
      const int off = ((b >> 2)*8 + (lane >> 3))*kPitch + (b & 3)*8 + (lane & 7);
      if (thr == 0) {
        Bi[off] = Ai[off];
        Bi[off + 4*kPitch] = Ai[off + 4*kPitch];
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
    for (int k = 0; k < 4; k++) {
      const int b = warp*4 + k;
      const int thr = s_thr[b];
      const int r = (b >> 2)*8 + (lane >> 3), c = (b & 3)*8 + (lane & 7);
      const int off = r*kPitch + c;
      int *g = dst + (long long)r*w + c;
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
      const int thr = s_thr[b];
      const int off = ((b >> 2)*4 + (l16 >> 2))*kPitch + (b & 3)*4 + (l16 & 3);
      const int xx = Ai[off];
      int yy = xx;
      if (thr) {
        const int *a = Ai + off;
        const int dir = s_dir[b];
        const int o0 = kDirOff[dir][0], o1 = kDirOff[dir][1], o2 = kDirOff[dir][2];
        const Gate g = gate(thr);
        int s0 = tap(a[o0] - xx, g) + tap(a[-o0] - xx, g);
        int s1 = tap(a[o1] - xx, g) + tap(a[-o1] - xx, g) + tap(a[o2] - xx, g) + tap(a[-o2] - xx, g);
        yy = (short)(xx + ((3*s0 + 2*s1 + 8) >> 4));
      }
      Bi[off] = yy;
    }
    __syncthreads();
    for (int k = 0; k < 2; k++) {
      const int b = (warp*2 + k)*2 + half;
      const int thr = s_thr[b];
      const int r = (b >> 2)*4 + (l16 >> 2), c = (b & 3)*4 + (l16 & 3);
      const int *bq = Bi + r*kPitch + c;
      const int yy = bq[0];
      int out = yy;
      if (thr) {
        const int o = s_dir[b] <= 4 ? kPitch : 1;
        const Gate ga = gate(min(thr, thr/3 + abs(yy - Ai[r*kPitch + c])));
        int sum = tap(bq[o] - yy, ga) + tap(bq[-o] - yy, ga) + tap(bq[2*o] - yy, ga) +
                  tap(bq[-2*o] - yy, ga);
        out = (short)(yy + ((3*sum + 8) >> 4));
      }
      dst[(long long)r*w + c] = out;
    }
  }
}

__global__ void __launch_bounds__(kThreads) dering_kernel(DeringParams P) {
  __shared__ __align__(16) int A[kTile];
  __shared__ __align__(16) int B[kTile];
  __shared__ int s_cost[16][8];
  __shared__ int s_dir[16];
  __shared__ int s_thr[16];
  __shared__ int s_coded[16];
  const int t = threadIdx.x;
  // (a 3-D grid that spares these divisions measures 1.5 % slower)
  const int per = P.nhsb*P.nvsb;
  const int frame = blockIdx.x/per;
  const int sb = blockIdx.x - frame*per;
  const int sby = sb/P.nhsb, sbx = sb - sby*P.nhsb;
  const bool filtered = P.flags[(long long)frame*per + sb] != 0;
  for (int p = 0; p < P.nplanes; p++) {
    const int dec = P.dec[p];
    if (!filtered) {
      // copy through: n rows of n words as 16-byte vectors
      const int n = kSB >> dec;
      const int w = (P.nhsb*kSB) >> dec;
      const long long o = frame*P.plane_size[p] + (long long)(sby*n)*w + sbx*n;
      const int vec_per_row = n >> 2;
      for (int i = t; i < n*vec_per_row; i += kThreads) {
        int r = i/vec_per_row, c = i - r*vec_per_row;
        reinterpret_cast<int4 *>(P.out[p] + o + (long long)r*w)[c] =
            __ldg(reinterpret_cast<const int4 *>(P.in[p] + o + (long long)r*w) + c);
      }
      continue;
    }
    if (dec == 0) dering_plane<0>(P, p, frame, sbx, sby, A, B, s_cost, s_dir, s_thr, s_coded);
    else dering_plane<1>(P, p, frame, sbx, sby, A, B, s_cost, s_dir, s_thr, s_coded);
  }
}

void launch_dering(const DeringParams &P, int nframes, cudaStream_t st) {
  dering_kernel<<<(unsigned)(P.nhsb*P.nvsb*nframes), kThreads, 0, st>>>(P);
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
    const int sby = (int)__umulhi((unsigned)sb, P.recip_nhsb), sbx = sb - sby*P.nhsb;
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

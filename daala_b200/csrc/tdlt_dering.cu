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

// Cost of direction D for the 8x8 block at blk (reference src/filter.c:1655-1697): the sum
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
    for (int j = 0; j < 8; j++) s[dir_line<D>(i, j)] += v[j] >> 4;
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

__device__ __forceinline__ int tap(int p, int thr) { return abs(p) < thr ? p : 0; }

}  // namespace

__global__ void __launch_bounds__(kThreads) dering_kernel(DeringParams P) {
  __shared__ __align__(16) int A[kTile];
  __shared__ __align__(16) int B[kTile];
  __shared__ int s_cost[16][8];
  __shared__ int s_dir[16];
  __shared__ int s_thr[16];
  const int t = threadIdx.x;
  const int per = P.nhsb*P.nvsb;
  const int frame = blockIdx.x/per;
  const int sb = blockIdx.x - frame*per;
  const int sby = sb/P.nhsb, sbx = sb - sby*P.nhsb;
  const bool filtered = P.flags[(long long)frame*per + sb] != 0;

  for (int p = 0; p < P.nplanes; p++) {
    const int dec = P.dec[p];
    const int n = kSB >> dec;
    const int w = (P.nhsb*kSB) >> dec;
    const int *src = P.in[p] + frame*P.plane_size[p] + (long long)(sby*n)*w + sbx*n;
    int *dst = P.out[p] + frame*P.plane_size[p] + (long long)(sby*n)*w + sbx*n;
    if (!filtered) {
      // copy through: n rows of n words as 16-byte vectors
      const int vec_per_row = n >> 2;
      for (int i = t; i < n*vec_per_row; i += kThreads) {
        int r = i/vec_per_row, c = i - r*vec_per_row;
        reinterpret_cast<int4 *>(dst + (long long)r*w)[c] =
            __ldg(reinterpret_cast<const int4 *>(src + (long long)r*w) + c);
      }
      continue;
    }
    // ---- load the tile with its ring; outside the frame: far away ------------------
    const int lo_r = sby ? -kBorder : 0, hi_r = n + (sby != P.nvsb - 1 ? kBorder : 0);
    const int lo_c = sbx ? -kBorder : 0, hi_c = n + (sbx != P.nhsb - 1 ? kBorder : 0);
    const int span = n + 2*kBorder;
    __syncthreads();                                    // previous plane done with A, B
    for (int i = t; i < span*span; i += kThreads) {
      int r = i/span, c = i - r*span;
      int rr = r - kBorder, cc = c - kBorder;
      int v = kFar;
      if (rr >= lo_r && rr < hi_r && cc >= lo_c && cc < hi_c) {
        v = (short)__ldg(src + (long long)rr*w + cc);   // the reference's int16_t etmp copy
      }
      A[r*kPitch + kPadL - kBorder + c] = v;
      B[r*kPitch + kPadL - kBorder + c] = v;
    }
    __syncthreads();
    const int *Ai = A + kBorder*kPitch + kPadL;         // interior origin
    int *Bi = B + kBorder*kPitch + kPadL;
    // ---- luma: direction search ----------------------------------------------------
    if (p == 0) {
      const int blk = t & 15, d = t >> 4;
      const int *bp = Ai + (blk >> 2)*8*kPitch + (blk & 3)*8;
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
    __syncthreads();
    // ---- blocks whose units and the ring of units around them are all skipped -----------
    if (t < 16) {
      const int u = 2 >> dec;
      const int by = t >> 2, bx = t & 3;
      const unsigned char *sk = P.bskip[p] + (long long)frame*P.skip_frame +
                                (long long)((sby << (3 - dec)) + by*u)*P.skip_stride +
                                (sbx << (3 - dec)) + bx*u;
      int all = 1;
      for (int i = sby ? -1 : 0; i < u + (sby != P.nvsb - 1); i++) {
        for (int j = sbx ? -1 : 0; j < u + (sbx != P.nhsb - 1); j++) {
          all &= sk[i*P.skip_stride + j] != 0;
        }
      }
      if (all) s_thr[t] = 0;
    }
    __syncthreads();
    // ---- the two passes: thread = one column, rows strided --------------------------------
    const int lb = 3 - dec;
    const int col = t & (n - 1);
    const int row0 = t >> (5 - dec);
    const int rstep = kThreads >> (5 - dec);
    for (int r = row0; r < n; r += rstep) {
      const int b = (r >> lb)*4 + (col >> lb);
      const int thr = s_thr[b], dir = s_dir[b];
      const int *a = Ai + r*kPitch + col;
      const int xx = a[0];
      int sum = 0;
#pragma unroll
      for (int k = 0; k < 3; k++) {
        const int o = kDirOff[dir][k];
        const int tp = k == 0 ? 3 : 2;
        sum += tp*(tap(a[o] - xx, thr) + tap(a[-o] - xx, thr));
      }
      Bi[r*kPitch + col] = (short)(xx + ((sum + 8) >> 4));
    }
    __syncthreads();
    for (int r = row0; r < n; r += rstep) {
      const int b = (r >> lb)*4 + (col >> lb);
      const int thr = s_thr[b], dir = s_dir[b];
      const int o = dir <= 4 ? kPitch : 1;
      const int *bq = Bi + r*kPitch + col;
      const int yy = bq[0];
      const int ath = min(thr, thr/3 + abs(yy - Ai[r*kPitch + col]));
      int sum = tap(bq[o] - yy, ath) + tap(bq[-o] - yy, ath) + tap(bq[2*o] - yy, ath) +
                tap(bq[-2*o] - yy, ath);
      dst[(long long)r*w + col] = (short)(yy + ((3*sum + 8) >> 4));
    }
  }
}

void launch_dering(const DeringParams &P, int nframes, cudaStream_t st) {
  dering_kernel<<<(unsigned)(P.nhsb*P.nvsb*nframes), kThreads, 0, st>>>(P);
}

}  // namespace tdlt

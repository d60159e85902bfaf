// Auxiliary kernels: pixel <-> coefficient conversion, and the standalone
// per-block / per-vector forms of the transforms behind the reference-signature
// entry points (layer 3 of include/daala_tdlt.h).  Hand-written CUDA for
// sm_100a; streaming or tiny, so no TMA here.
#include "tdlt_device.cuh"
#include "tdlt_aux_kernels.h"

namespace tdlt {

// ---------------------------------------------------------------------------
// Pixel <-> coefficient (reference src/state.c:1209-1316; clamps
// src/odintrin.h:104-113).  One thread per 4 samples.
//   xstride 1: c = (p - 128) << shift            p = clamp255(((c + rnd) >> shift) + 128)
//   xstride 2: c = (p - 2048 + rnd) >> shift     p = clampFPR((c << shift) + 2048)
// ---------------------------------------------------------------------------
__global__ void pix_to_coeff_kernel(int4 *__restrict__ c, const void *__restrict__ pix,
                                    long long n4, int xstride, int shift) {
  long long i = (long long)blockIdx.x*blockDim.x + threadIdx.x;
  if (i >= n4) return;
  int4 o;
  if (xstride == 1) {
    uchar4 p = reinterpret_cast<const uchar4 *>(pix)[i];
    o.x = ((int)p.x - 128)*(1 << shift);
    o.y = ((int)p.y - 128)*(1 << shift);
    o.z = ((int)p.z - 128)*(1 << shift);
    o.w = ((int)p.w - 128)*(1 << shift);
  }
  else {
    short4 p = reinterpret_cast<const short4 *>(pix)[i];
    int rnd = (1 << shift) >> 1;
    o.x = ((int)p.x - 2048 + rnd) >> shift;
    o.y = ((int)p.y - 2048 + rnd) >> shift;
    o.z = ((int)p.z - 2048 + rnd) >> shift;
    o.w = ((int)p.w - 2048 + rnd) >> shift;
  }
  c[i] = o;
}

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return min(max(v, lo), hi); }

__global__ void coeff_to_pix_kernel(void *__restrict__ pix, const int4 *__restrict__ c,
                                    long long n4, int xstride, int shift) {
  long long i = (long long)blockIdx.x*blockDim.x + threadIdx.x;
  if (i >= n4) return;
  int4 v = c[i];
  if (xstride == 1) {
    int rnd = (1 << shift) >> 1;
    uchar4 o;
    o.x = (unsigned char)clampi(((v.x + rnd) >> shift) + 128, 0, 255);
    o.y = (unsigned char)clampi(((v.y + rnd) >> shift) + 128, 0, 255);
    o.z = (unsigned char)clampi(((v.z + rnd) >> shift) + 128, 0, 255);
    o.w = (unsigned char)clampi(((v.w + rnd) >> shift) + 128, 0, 255);
    reinterpret_cast<uchar4 *>(pix)[i] = o;
  }
  else {
    short4 o;
    o.x = (short)clampi(v.x*(1 << shift) + 2048, 0, 4095);
    o.y = (short)clampi(v.y*(1 << shift) + 2048, 0, 4095);
    o.z = (short)clampi(v.z*(1 << shift) + 2048, 0, 4095);
    o.w = (short)clampi(v.w*(1 << shift) + 2048, 0, 4095);
    reinterpret_cast<short4 *>(pix)[i] = o;
  }
}

// ---------------------------------------------------------------------------
// Standalone 2-D DCT of many blocks (reference src/dct.c:137-149, 335-347,
// 774-788, 2028-2042).  One CTA of N threads per block; thread = line.
// ---------------------------------------------------------------------------
template <int N> struct LineDct;
template <> struct LineDct<4> {
  static __device__ __forceinline__ void f(int *v) { tdlt_fdct4(v); }
  static __device__ __forceinline__ void i(int *v) { tdlt_idct4(v); }
};
template <> struct LineDct<8> {
  static __device__ __forceinline__ void f(int *v) { tdlt_fdct8(v); }
  static __device__ __forceinline__ void i(int *v) { tdlt_idct8(v); }
};
template <> struct LineDct<16> {
  static __device__ __forceinline__ void f(int *v) { tdlt_fdct16(v); }
  static __device__ __forceinline__ void i(int *v) { tdlt_idct16(v); }
};
template <> struct LineDct<32> {
  static __device__ __forceinline__ void f(int *v) { tdlt_fdct32(v); }
  static __device__ __forceinline__ void i(int *v) { tdlt_idct32(v); }
};

template <int N, bool kInverse>
__global__ void __launch_bounds__(N) dct2d_blocks_kernel(int *__restrict__ out, int ostride,
                                                         long long oblock,
                                                         const int *__restrict__ in, int istride,
                                                         long long iblock) {
  __shared__ int tile[N][N + 1];
  const int *src = in + (long long)blockIdx.x*iblock;
  int *dst = out + (long long)blockIdx.x*oblock;
  const int l = threadIdx.x;
  int v[N];
  if (!kInverse) {
    // column l -> vertical DCT -> column l of the tile
#pragma unroll
    for (int r = 0; r < N; r++) v[r] = src[r*istride + l];
    LineDct<N>::f(v);
#pragma unroll
    for (int r = 0; r < N; r++) tile[r][l] = v[r];
    __syncthreads();
    // row l -> horizontal DCT -> row l of the output
#pragma unroll
    for (int c = 0; c < N; c++) v[c] = tile[l][c];
    LineDct<N>::f(v);
#pragma unroll
    for (int c = 0; c < N; c++) dst[l*ostride + c] = v[c];
  }
  else {
#pragma unroll
    for (int c = 0; c < N; c++) v[c] = src[l*istride + c];
    LineDct<N>::i(v);
#pragma unroll
    for (int c = 0; c < N; c++) tile[l][c] = v[c];
    __syncthreads();
#pragma unroll
    for (int r = 0; r < N; r++) v[r] = tile[r][l];
    LineDct<N>::i(v);
#pragma unroll
    for (int r = 0; r < N; r++) dst[r*ostride + l] = v[r];
  }
}

void launch_dct2d_blocks(int bs, bool inverse, int *out, int ostride, long long oblock,
                         const int *in, int istride, long long iblock, int nblocks,
                         cudaStream_t st) {
#define TDLT_CASE(N) \
  if (inverse) dct2d_blocks_kernel<N, true><<<nblocks, N, 0, st>>>(out, ostride, oblock, in, istride, iblock); \
  else dct2d_blocks_kernel<N, false><<<nblocks, N, 0, st>>>(out, ostride, oblock, in, istride, iblock)
  switch (bs) {
    case 0: TDLT_CASE(4); break;
    case 1: TDLT_CASE(8); break;
    case 2: TDLT_CASE(16); break;
    default: TDLT_CASE(32); break;
  }
#undef TDLT_CASE
}

// ---------------------------------------------------------------------------
// General N-tap lapping filters, N = 4 << f (reference src/filter.c:174-1373,
// TYPE3 parameter sets :164-167, 285-294, 485-507, 812-858).  Only the 4-tap
// one is on the codec's path (OD_FILT_SIZE == 0, src/filter.h:99); the larger
// ones exist for the reference's tools and self-test.  Parameters: N/2 scales,
// N/2-1 "p", N/2-1 "u", in 1/64ths.
// ---------------------------------------------------------------------------
__constant__ int kFilt4[4] = {85, 75, -15, 33};
__constant__ int kFilt8[10] = {93, 72, 73, 78, -28, -23, -10, 50, 37, 23};
__constant__ int kFilt16[22] = {94, 71, 68, 68, 68, 69, 70, 73,
                                -32, -37, -36, -32, -26, -17, -7,
                                56, 49, 45, 40, 34, 26, 15};
__constant__ int kFilt32[46] = {91, 70, 68, 67, 67, 67, 67, 66, 66, 67, 67, 66, 67, 67, 67, 70,
                                -32, -41, -42, -41, -40, -38, -36, -34, -32, -29, -24, -19,
                                -14, -9, -5,
                                58, 52, 50, 48, 45, 43, 40, 38, 35, 32, 29, 24, 18, 13, 8};

template <int N, bool kInverse>
__device__ void lap_filter_n(int *t, const int *par) {
  constexpr int H = N/2;
  const int *s = par, *p = par + H, *u = par + 2*H - 1;
  int x[N];
#pragma unroll
  for (int i = 0; i < N; i++) x[i] = t[i];
#pragma unroll
  for (int i = 0; i < H; i++) t[N - 1 - i] = x[i] - x[N - 1 - i];
#pragma unroll
  for (int i = 0; i < H; i++) t[i] = x[i] - (t[N - 1 - i] >> 1);
  if (!kInverse) {
#pragma unroll
    for (int i = 0; i < H; i++) {
      if (s[i] != 64) {
        int v = (t[H + i]*s[i]) >> 6;
        t[H + i] = v + (v > 0);
      }
    }
#pragma unroll
    for (int i = N - 1; i > H; i--) {
      t[i] += (t[i - 1]*p[i - 1 - H] + 32) >> 6;
      t[i - 1] += (t[i]*u[i - 1 - H] + 32) >> 6;
    }
  }
  else {
#pragma unroll
    for (int i = H + 1; i < N; i++) {
      t[i - 1] -= (t[i]*u[i - 1 - H] + 32) >> 6;
      t[i] -= (t[i - 1]*p[i - 1 - H] + 32) >> 6;
    }
#pragma unroll
    for (int i = 0; i < H; i++) {
      if (s[i] != 64) t[H + i] = (t[H + i]*64)/s[i];
    }
  }
#pragma unroll
  for (int i = 0; i < H; i++) t[i] += t[N - 1 - i] >> 1;
#pragma unroll
  for (int i = 0; i < H; i++) t[N - 1 - i] = t[i] - t[N - 1 - i];
}

template <int N, bool kInverse>
__global__ void lap_vectors_kernel(int *__restrict__ out, const int *__restrict__ in, int nvec) {
  int i = blockIdx.x*blockDim.x + threadIdx.x;
  if (i >= nvec) return;
  int t[N];
#pragma unroll
  for (int k = 0; k < N; k++) t[k] = in[(long long)i*N + k];
  const int *par = N == 4 ? kFilt4 : N == 8 ? kFilt8 : N == 16 ? kFilt16 : kFilt32;
  lap_filter_n<N, kInverse>(t, par);
#pragma unroll
  for (int k = 0; k < N; k++) out[(long long)i*N + k] = t[k];
}

void launch_lap_vectors(int f, bool inverse, int *out, const int *in, int nvec, cudaStream_t st) {
  int grid = (nvec + 127)/128;
#define TDLT_CASE(N) \
  if (inverse) lap_vectors_kernel<N, true><<<grid, 128, 0, st>>>(out, in, nvec); \
  else lap_vectors_kernel<N, false><<<grid, 128, 0, st>>>(out, in, nvec)
  switch (f) {
    case 0: TDLT_CASE(4); break;
    case 1: TDLT_CASE(8); break;
    case 2: TDLT_CASE(16); break;
    default: TDLT_CASE(32); break;
  }
#undef TDLT_CASE
}

// 1-D DCTs of many vectors (reference OD_FDCT_1D / OD_IDCT_1D, src/dct.c:58-72), natural
// order in and out: what the self-test's check_bias streams through (src/dct.c:3565-3598).
template <int N, bool kInverse>
__global__ void dct1d_vectors_kernel(int *__restrict__ out, const int *__restrict__ in, int nvec) {
  int i = blockIdx.x*blockDim.x + threadIdx.x;
  if (i >= nvec) return;
  int t[N];
  const int4 *src = reinterpret_cast<const int4 *>(in + (long long)i*N);
#pragma unroll
  for (int k = 0; k < N/4; k++) {
    int4 q = src[k];
    t[4*k] = q.x; t[4*k + 1] = q.y; t[4*k + 2] = q.z; t[4*k + 3] = q.w;
  }
  if (kInverse) LineDct<N>::i(t); else LineDct<N>::f(t);
  int4 *dst = reinterpret_cast<int4 *>(out + (long long)i*N);
#pragma unroll
  for (int k = 0; k < N/4; k++) dst[k] = make_int4(t[4*k], t[4*k + 1], t[4*k + 2], t[4*k + 3]);
}

void launch_dct1d_vectors(int bs, bool inverse, int *out, const int *in, int nvec, cudaStream_t st) {
  int grid = (nvec + 127)/128;
#define TDLT_CASE(N) \
  if (inverse) dct1d_vectors_kernel<N, true><<<grid, 128, 0, st>>>(out, in, nvec); \
  else dct1d_vectors_kernel<N, false><<<grid, 128, 0, st>>>(out, in, nvec)
  switch (bs) {
    case 0: TDLT_CASE(4); break;
    case 1: TDLT_CASE(8); break;
    case 2: TDLT_CASE(16); break;
    default: TDLT_CASE(32); break;
  }
#undef TDLT_CASE
}

// Strided line filters used by the host-pointer forms of od_prefilter_split /
// od_apply_prefilter_frame_sbs and their inverses: `count` lines, line i
// starts at base + i*line_step and its 4 taps are tap_step apart.
template <bool kInverse>
__global__ void lap4_lines_kernel(int *__restrict__ base, long long line_step, long long tap_step,
                                  int count) {
  int i = blockIdx.x*blockDim.x + threadIdx.x;
  if (i >= count) return;
  int *p = base + i*line_step;
  int a = p[0], b = p[tap_step], c = p[2*tap_step], d = p[3*tap_step];
  if (kInverse) post4(a, b, c, d); else pre4(a, b, c, d);
  p[0] = a; p[tap_step] = b; p[2*tap_step] = c; p[3*tap_step] = d;
}

void launch_lap4_lines(bool inverse, int *base, long long line_step, long long tap_step,
                       int count, cudaStream_t st) {
  int grid = (count + 127)/128;
  if (inverse) lap4_lines_kernel<true><<<grid, 128, 0, st>>>(base, line_step, tap_step, count);
  else lap4_lines_kernel<false><<<grid, 128, 0, st>>>(base, line_step, tap_step, count);
}

// ---------------------------------------------------------------------------
// Standalone Haar wavelet of one 2^ln block (reference src/dct.c:1960-2026).
// One CTA; x and y are device buffers with the given strides.
// ---------------------------------------------------------------------------
__global__ void haar_block_kernel(int *__restrict__ y, int ystride, const int *__restrict__ x,
                                  int xstride, int ln, int inverse) {
  __shared__ int T[32*32];
  const int n = 1 << ln;
  const int tid = threadIdx.x, nt = blockDim.x;
  if (!inverse) {
    for (int e = tid; e < n*n; e += nt) T[(e/n)*32 + e%n] = x[(e/n)*xstride + e%n];
    __syncthreads();
    for (int level = 0; level < ln; level++) {
      int np = n >> level >> 1;
      // np*np <= 256 items; each thread owns at most one (blockDim = 256)
      int a = 0, b = 0, c = 0, d = 0, i = tid/np, j = tid%np;
      bool act = tid < np*np;
      if (act) {
        a = T[2*i*32 + 2*j]; b = T[(2*i + 1)*32 + 2*j];
        c = T[2*i*32 + 2*j + 1]; d = T[(2*i + 1)*32 + 2*j + 1];
      }
      __syncthreads();
      if (act) {
        haar2x2(a, b, c, d);
        T[i*32 + j] = a;
        y[i*ystride + j + np] = b;
        y[(i + np)*ystride + j] = c;
        y[(i + np)*ystride + j + np] = d;
      }
      __syncthreads();
    }
    if (tid == 0) y[0] = T[0];
  }
  else {
    if (tid == 0) T[0] = x[0];
    __syncthreads();
    for (int level = ln - 1; level >= 0; level--) {
      int np = 1 << (ln - 1 - level);
      int a = 0, b = 0, c = 0, d = 0, i = tid/np, j = tid%np;
      bool act = tid < np*np;
      if (act) {
        a = T[i*32 + j];
        b = x[i*xstride + j + np];
        c = x[(i + np)*xstride + j];
        d = x[(i + np)*xstride + j + np];
      }
      __syncthreads();
      if (act) {
        haar2x2(a, b, c, d);
        T[2*i*32 + 2*j] = a; T[(2*i + 1)*32 + 2*j] = b;
        T[2*i*32 + 2*j + 1] = c; T[(2*i + 1)*32 + 2*j + 1] = d;
      }
      __syncthreads();
    }
    for (int e = tid; e < n*n; e += nt) y[(e/n)*ystride + e%n] = T[(e/n)*32 + e%n];
  }
}

void launch_haar_block(bool inverse, int *y, int ystride, const int *x, int xstride, int ln,
                       cudaStream_t st) {
  haar_block_kernel<<<1, 256, 0, st>>>(y, ystride, x, xstride, ln, inverse ? 1 : 0);
}

// ---------------------------------------------------------------------------
// Time-frequency merge / split (reference src/tf.c:112-169).
// up:   2x2 group of nxn blocks (src, 2n x 2n) -> one 2nx2n block
// down: one nxn block -> 2x2 group of n/2 x n/2 blocks
// One thread per 2x2 butterfly.
// ---------------------------------------------------------------------------
__global__ void tf_hv_kernel(int *__restrict__ dst, int dstride, const int *__restrict__ src,
                             int sstride, int n, int down) {
  int x = blockIdx.x*blockDim.x + threadIdx.x;
  int y = blockIdx.y*blockDim.y + threadIdx.y;
  if (x >= n || y >= n) return;
  int hs = x & 1, vs = y & 1;
  if (!down) {
    int ll = src[y*sstride + x], lh = src[y*sstride + x + n];
    int hl = src[(y + n)*sstride + x], hh = src[(y + n)*sstride + x + n];
    haar2x2(ll, hl, lh, hh);
    dst[(2*y + vs)*dstride + 2*x + hs] = ll;
    dst[(2*y + vs)*dstride + 2*x + 1 - hs] = lh;
    dst[(2*y + 1 - vs)*dstride + 2*x + hs] = hl;
    dst[(2*y + 1 - vs)*dstride + 2*x + 1 - hs] = hh;
  }
  else {
    int ll = src[(2*y + vs)*sstride + 2*x + hs], lh = src[(2*y + vs)*sstride + 2*x + 1 - hs];
    int hl = src[(2*y + 1 - vs)*sstride + 2*x + hs];
    int hh = src[(2*y + 1 - vs)*sstride + 2*x + 1 - hs];
    haar2x2(ll, lh, hl, hh);
    dst[y*dstride + x] = ll;
    dst[y*dstride + x + n] = lh;
    dst[(y + n)*dstride + x] = hl;
    dst[(y + n)*dstride + x + n] = hh;
  }
}

void launch_tf_hv(bool down, int *dst, int dstride, const int *src, int sstride, int n,
                  cudaStream_t st) {
  dim3 blk(16, 16), grid((n + 15)/16, (n + 15)/16);
  tf_hv_kernel<<<grid, blk, 0, st>>>(dst, dstride, src, sstride, n, down ? 1 : 0);
}

// ---------------------------------------------------------------------------
// The remaining time-frequency helpers of reference src/tf.c, one CTA per call
// (a block is at most 32x32): the low-pass-only merges od_tf_up_{h,v,hv}_lp
// (:38-108), the TF smoothing filters od_tf_filter_2d / _inv_2d (:171-235) and
// od_convert_block_down (:237-277).  Parity surface; nothing here is hot.
// ---------------------------------------------------------------------------
enum { kTfUpH = 0, kTfUpV = 1, kTfUpHV = 2, kTfFilt = 3, kTfFiltInv = 4, kTfDown = 5 };

// One line of the TF filter: a chain over the odd positions (src/tf.c:171-199).
__device__ __forceinline__ void tf_filter_line(int *v, int stride, int n, bool inverse) {
  int m = (n >> 1) - 1;
  if (!inverse) {
    int *q = v + stride;
    for (int i = 0; i < m; i++) {
      int *u = q;
      q += 2*stride;
      *u += *q >> 1;
      *q -= *u >> 1;
    }
  }
  else {
    int *u = v + stride*(n - 1);
    for (int i = 0; i < m; i++) {
      int *q = u;
      u -= 2*stride;
      *q += *u >> 1;
      *u -= *q >> 1;
    }
  }
}

__global__ void tf_ops_kernel(int op, int *__restrict__ dst, int dstride,
                              const int *__restrict__ src, int sstride, int n, int p0, int p1,
                              int p2) {
  __shared__ int buf[2][32*32];
  const int tid = threadIdx.x, nt = blockDim.x;
  if (op == kTfUpH) {                // p0 = dx
    for (int e = tid; e < n*(n >> 1); e += nt) {
      int y = e/(n >> 1), x = e - y*(n >> 1);
      int ll = src[y*sstride + x];
      int lh = ll - src[y*sstride + x + p0];
      int hs = x & 1;
      ll -= tdlt_half(lh);
      dst[y*dstride + 2*x + hs] = ll;
      dst[y*dstride + 2*x + 1 - hs] = lh;
    }
  }
  else if (op == kTfUpV) {           // p0 = dy
    for (int e = tid; e < n*(n >> 1); e += nt) {
      int y = e/n, x = e - y*n;
      int vs = y & 1;
      int ll = src[y*sstride + x];
      int hl = ll - src[(y + p0)*sstride + x];
      ll -= tdlt_half(hl);
      dst[(2*y + vs)*dstride + x] = ll;
      dst[(2*y + 1 - vs)*dstride + x] = hl;
    }
  }
  else if (op == kTfUpHV) {          // p0 = dx, p1 = dy
    int h = n >> 1;
    for (int e = tid; e < h*h; e += nt) {
      int y = e/h, x = e - y*h;
      int ll = src[y*sstride + x], lh = src[y*sstride + x + p0];
      int hl = src[(y + p1)*sstride + x], hh = src[(y + p1)*sstride + x + p0];
      haar2x2(ll, hl, lh, hh);
      int hs = x & 1, vs = y & 1;
      dst[(2*y + vs)*dstride + 2*x + hs] = ll;
      dst[(2*y + vs)*dstride + 2*x + 1 - hs] = lh;
      dst[(2*y + 1 - vs)*dstride + 2*x + hs] = hl;
      dst[(2*y + 1 - vs)*dstride + 2*x + 1 - hs] = hh;
    }
  }
  else {
    // in shared memory: buf[cur] holds the n x n block, row stride 32
    int cur = 0;
    for (int e = tid; e < n*n; e += nt) buf[0][(e/n)*32 + e%n] = src[(e/n)*sstride + e%n];
    __syncthreads();
    if (op == kTfFilt || op == kTfFiltInv) {
      bool inv = op == kTfFiltInv;
      // forward: rows then columns; inverse: columns then rows
      for (int step = 0; step < 2; step++) {
        bool rows = (step == 0) != inv;
        if (tid < n) tf_filter_line(rows ? buf[0] + 32*tid : buf[0] + tid, rows ? 1 : 32, n, inv);
        __syncthreads();
      }
    }
    else {                           // od_convert_block_down: p0 = curr, p1 = dest, p2 = filter
      int s = n;                     // current block size
      for (int lvl = p0; lvl > p1; lvl--, s >>= 1) {
        int nb = n/s;                // blocks per side at this level
        if (p2) {
          for (int step = 0; step < 2; step++) {      // inverse filter: columns, then rows
            for (int e = tid; e < nb*nb*s; e += nt) {
              int blk = e/s, l = e - blk*s;
              int *b0 = buf[cur] + (blk/nb)*s*32 + (blk%nb)*s;
              tf_filter_line(step == 0 ? b0 + l : b0 + 32*l, step == 0 ? 32 : 1, s, true);
            }
            __syncthreads();
          }
        }
        int h = s >> 1;
        for (int e = tid; e < nb*nb*h*h; e += nt) {
          int blk = e/(h*h), r = e - blk*h*h;
          int y = r/h, x = r - y*h;
          const int *a = buf[cur] + (blk/nb)*s*32 + (blk%nb)*s;
          int *o = buf[cur ^ 1] + (blk/nb)*s*32 + (blk%nb)*s;
          int hs = x & 1, vs = y & 1;
          int ll = a[(2*y + vs)*32 + 2*x + hs], lh = a[(2*y + vs)*32 + 2*x + 1 - hs];
          int hl = a[(2*y + 1 - vs)*32 + 2*x + hs], hh = a[(2*y + 1 - vs)*32 + 2*x + 1 - hs];
          haar2x2(ll, lh, hl, hh);
          o[y*32 + x] = ll;
          o[y*32 + x + h] = lh;
          o[(y + h)*32 + x] = hl;
          o[(y + h)*32 + x + h] = hh;
        }
        __syncthreads();
        cur ^= 1;
      }
    }
    for (int e = tid; e < n*n; e += nt) dst[(e/n)*dstride + e%n] = buf[cur][(e/n)*32 + e%n];
  }
}

void launch_tf_op(int op, int *dst, int dstride, const int *src, int sstride, int n, int p0,
                  int p1, int p2, cudaStream_t st) {
  tf_ops_kernel<<<1, 256, 0, st>>>(op, dst, dstride, src, sstride, n, p0, p1, p2);
}

// od_init_skipped_coeffs (reference src/state.c:1340-1357), one block: d is n x n with row
// stride w; keyframe: everything but the DC cleared; else d = pred (dense n x n).
__global__ void init_skipped_kernel(int *d, const int *pred, int is_keyframe, int n, int w) {
  for (int e = threadIdx.x; e < n*n; e += blockDim.x) {
    const int i = e/n, j = e - i*n;
    if (is_keyframe) { if (e) d[i*w + j] = 0; }
    else d[i*w + j] = pred[e];
  }
}
void launch_init_skipped(int *d, const int *pred, int is_keyframe, int n, int w, cudaStream_t st) {
  init_skipped_kernel<<<1, 256, 0, st>>>(d, pred, is_keyframe, n, w);
}

void launch_pix_to_coeff(int *c, const void *pix, long long nsamples, int xstride, int shift,
                         cudaStream_t st) {
  long long n4 = nsamples/4;
  pix_to_coeff_kernel<<<(unsigned)((n4 + 255)/256), 256, 0, st>>>(
      reinterpret_cast<int4 *>(c), pix, n4, xstride, shift);
}
void launch_coeff_to_pix(void *pix, const int *c, long long nsamples, int xstride, int shift,
                         cudaStream_t st) {
  long long n4 = nsamples/4;
  coeff_to_pix_kernel<<<(unsigned)((n4 + 255)/256), 256, 0, st>>>(
      pix, reinterpret_cast<const int4 *>(c), n4, xstride, shift);
}


// ---------------------------------------------------------------------------
// Chroma-from-luma predictor planes (reference od_resample_luma_coeffs,
// src/intra.c:72-109, as the leaf visits of src/encode.c:1491-1517 call it for
// 4:2:0): for every chroma leaf block the low-frequency coefficients of the luma
// transform covering it.  One thread per 8x8 luma unit = 4x4 samples of the
// half-resolution plane `l`:
//   unit size 0 (four 4x4 luma transforms over one 4x4 chroma block): the 2x2 low
//   corners of the four blocks merged by one Haar level, then the trained 4x4 scaling;
//   otherwise: the 4x4 patch of the luma block's upper-left quarter this unit covers.
// Memory bound: per `l` sample 4 bytes read + 4 written (+ 1/16 byte of map).
// Threads of a warp cover 32 consecutive units of a unit row: their four 16-byte
// stores per thread tile four full 512-byte row segments.
// ---------------------------------------------------------------------------
__constant__ int kCflScaling4[16] = {   // reference src/intra.c:65-70, [x][y] flattened
    128, 128, 100, 36, 128, 80, 71, 35, 100, 71, 35, 31, 36, 35, 31, 18};

__global__ void __launch_bounds__(256) cfl_pred_kernel(
    int *__restrict__ l, int lstride, long long lframe, const int *__restrict__ d_luma,
    int dstride, long long dframe, const unsigned char *__restrict__ bsize, int bw, int bh,
    long long nunits) {
  long long i = (long long)blockIdx.x*blockDim.x + threadIdx.x;
  if (i >= nunits) return;
  int per = bw*bh;
  int frame = (int)(i/per);
  int u = (int)(i - (long long)frame*per);
  int uy = u/bw, ux = u - uy*bw;
  int obs = __ldg(bsize + (long long)frame*per + u);
  const int *src = d_luma + frame*dframe;
  int *dst = l + frame*lframe + (long long)(4*uy)*lstride + 4*ux;
  if (obs == 0) {
    const int *s = src + (long long)(8*uy)*dstride + 8*ux;
    int o[4][4];
#pragma unroll
    for (int y = 0; y < 2; y++) {
      int2 a = *reinterpret_cast<const int2 *>(s + y*dstride);
      int2 b = *reinterpret_cast<const int2 *>(s + y*dstride + 4);
      int2 c = *reinterpret_cast<const int2 *>(s + (y + 4)*dstride);
      int2 e = *reinterpret_cast<const int2 *>(s + (y + 4)*dstride + 4);
      int ll[2] = {a.x, a.y}, lh[2] = {b.x, b.y}, hl[2] = {c.x, c.y}, hh[2] = {e.x, e.y};
#pragma unroll
      for (int x = 0; x < 2; x++) {
        haar2x2(ll[x], hl[x], lh[x], hh[x]);
        // low band at the position of its parity, as the reference's merge interleaves them
        int hs = x, vs = y;
        o[2*y + vs][2*x + hs] = ll[x];
        o[2*y + vs][2*x + 1 - hs] = lh[x];
        o[2*y + 1 - vs][2*x + hs] = hl[x];
        o[2*y + 1 - vs][2*x + 1 - hs] = hh[x];
      }
    }
#pragma unroll
    for (int y = 0; y < 4; y++) {
      int4 v;
      v.x = (kCflScaling4[0 + y]*o[y][0] + 64) >> 7;
      v.y = (kCflScaling4[4 + y]*o[y][1] + 64) >> 7;
      v.z = (kCflScaling4[8 + y]*o[y][2] + 64) >> 7;
      v.w = (kCflScaling4[12 + y]*o[y][3] + 64) >> 7;
      *reinterpret_cast<int4 *>(dst + (long long)y*lstride) = v;
    }
    return;
  }
  // The luma block of 2^(obs-1) x 2^(obs-1) units this unit belongs to; the chroma block
  // has the same origin at half scale and reads the block's first rows / columns.
  int m = (1 << (obs - 1)) - 1;
  int bx = ux & ~m, by = uy & ~m;
  const int *s = src + (long long)(8*by + 4*(uy - by))*dstride + 8*bx + 4*(ux - bx);
  int4 v[4];
#pragma unroll
  for (int y = 0; y < 4; y++) v[y] = __ldg(reinterpret_cast<const int4 *>(s + (long long)y*dstride));
#pragma unroll
  for (int y = 0; y < 4; y++) *reinterpret_cast<int4 *>(dst + (long long)y*lstride) = v[y];
}

void launch_cfl_pred(int *l, int lstride, long long lframe, const int *d_luma, int dstride,
                     long long dframe, const unsigned char *bsize, int bw, int bh, int nframes,
                     cudaStream_t st) {
  long long nunits = (long long)bw*bh*nframes;
  cfl_pred_kernel<<<(unsigned)((nunits + 255)/256), 256, 0, st>>>(
      l, lstride, lframe, d_luma, dstride, dframe, bsize, bw, bh, nunits);
}

}  // namespace tdlt

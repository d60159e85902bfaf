// The encoder's input side on the device: od_img_copy_pad
// (reference src/encode.c:1728-1750) = per plane od_img_plane_copy with its bit-depth
// conversion (src/state.c:92-217), the low-pass extension of the picture into the padded
// frame (od_img_plane_copy_pad, src/encode.c:610-700) and, where the destination has a
// border, od_img_plane_edge_ext (src/state.c:1100-1159).  The output is the plane the
// forward kernel's fused pixel load reads (od_tdlt_fwd_batch_pix), so a picture goes from
// the caller's 8/16-bit samples to coefficients without a host pass.
//   img_copy_kernel   one thread per picture sample: convert and store (streaming)
//   img_pad_kernel    one CTA per (plane, frame): the extension is a recurrence -- column x
//                     is a [1 2 1]/4 filter of column x-1 with rounding at every step, then
//                     row y of row y-1 -- so the previous line lives in shared memory and
//                     the CTA walks at most 63 columns and 63 rows
//   img_edge_kernel   one thread per border sample: the nearest plane sample
#include <cuda_runtime.h>
#include <stdint.h>
#include "tdlt_aux_kernels.h"

namespace tdlt {

__device__ __forceinline__ int img_get(const unsigned char *p, int xs) {
  return xs == 1 ? (int)*p : (int)*reinterpret_cast<const short *>(p);
}
__device__ __forceinline__ void img_put(unsigned char *p, int xs, int v) {
  if (xs == 1) *p = (unsigned char)v;
  else *reinterpret_cast<short *>(p) = (short)v;
}
// src/state.c:141-212
__device__ __forceinline__ int img_convert(int v, int sb, int db) {
  if (sb > 8) {
    if (db > 8) {
      if (db >= sb) return min(max(v << (db - sb), 0), (1 << db) - 1);
      const int dn = sb - db;
      return min(max((v + ((1 << dn) >> 1)) >> dn, 0), (1 << db) - 1);
    }
    const int dn = sb - 8;
    return min(max((v + ((1 << dn) >> 1)) >> dn, 0), 255);
  }
  if (db > 8) return min(max(v << (db - 8), 0), (1 << db) - 1);
  return v;
}

__global__ void __launch_bounds__(256) img_copy_kernel(const __grid_constant__ ImgPadParams P) {
  const int pli = blockIdx.z % P.nplanes, frame = blockIdx.z / P.nplanes;
  const ImgPadPlane &pl = P.pl[pli];
  const unsigned char *src = pl.src + (long long)frame*pl.src_frame_stride;
  unsigned char *dst = pl.dst + (long long)frame*pl.dst_frame_stride;
  for (int y = blockIdx.y; y < pl.pic_h; y += gridDim.y) {
    for (int x = blockIdx.x*blockDim.x + threadIdx.x; x < pl.pic_w; x += gridDim.x*blockDim.x) {
      const int v = img_get(src + (long long)y*pl.src_ystride + (long long)x*P.src_xs, P.src_xs);
      img_put(dst + (long long)y*pl.dst_ystride + (long long)x*P.dst_xs, P.dst_xs,
              img_convert(v, P.src_bits, P.dst_bits));
    }
  }
}

__global__ void __launch_bounds__(1024) img_pad_kernel(const __grid_constant__ ImgPadParams P) {
  extern __shared__ unsigned short line[];     // two lines of max(pic_h, plane_w) samples
  const int pli = blockIdx.x % P.nplanes, frame = blockIdx.x / P.nplanes;
  const ImgPadPlane &pl = P.pl[pli];
  unsigned char *dst = pl.dst + (long long)frame*pl.dst_frame_stride;
  const int xs = P.dst_xs, n = max(pl.pic_h, pl.plane_w);
  unsigned short *prev = line, *next = line + n;
  const int tid = threadIdx.x, nt = blockDim.x;
  if (pl.pic_w == 0 || pl.pic_h == 0) {
    // no data at all: the reference clears the plane (src/encode.c:622-628)
    for (int i = tid; i < pl.plane_h*pl.plane_w; i += nt) {
      img_put(dst + (long long)(i/pl.plane_w)*pl.dst_ystride + (i%pl.plane_w)*xs, xs, 0);
    }
    return;
  }
  // ---- to the right: column x from column x - 1, rows of the picture (src/encode.c:634-664)
  if (pl.plane_w > pl.pic_w) {
    for (int y = tid; y < pl.pic_h; y += nt) {
      prev[y] = (unsigned short)img_get(dst + (long long)y*pl.dst_ystride + (pl.pic_w - 1)*xs, xs);
    }
    __syncthreads();
    for (int x = pl.pic_w; x < pl.plane_w; x++) {
      for (int y = tid; y < pl.pic_h; y += nt) {
        const int c = prev[y], u = prev[max(y - 1, 0)], d = prev[min(y + 1, pl.pic_h - 1)];
        const int v = (2*c + u + d + 2) >> 2;
        next[y] = (unsigned short)v;
        img_put(dst + (long long)y*pl.dst_ystride + x*xs, xs, v);
      }
      __syncthreads();
      unsigned short *t = prev; prev = next; next = t;
    }
  }
  // ---- downwards: row y from row y - 1, all columns of the plane (:665-695)
  if (pl.plane_h > pl.pic_h) {
    __syncthreads();      // the last picture row's right part was written above by this CTA
    for (int x = tid; x < pl.plane_w; x += nt) {
      prev[x] = (unsigned short)img_get(dst + (long long)(pl.pic_h - 1)*pl.dst_ystride + x*xs, xs);
    }
    __syncthreads();
    for (int y = pl.pic_h; y < pl.plane_h; y++) {
      for (int x = tid; x < pl.plane_w; x += nt) {
        const int c = prev[x], l = prev[max(x - 1, 0)], r = prev[min(x + 1, pl.plane_w - 1)];
        const int v = (2*c + l + r + 2) >> 2;
        next[x] = (unsigned short)v;
        img_put(dst + (long long)y*pl.dst_ystride + x*xs, xs, v);
      }
      __syncthreads();
      unsigned short *t = prev; prev = next; next = t;
    }
  }
}

// Border samples of a plane: first the 2*border samples beside every plane row, then the
// border rows above and below over the widened width (od_img_plane_edge_ext's two steps
// give every border sample the nearest plane sample).
__global__ void __launch_bounds__(256) img_edge_kernel(const __grid_constant__ ImgPadParams P) {
  const int pli = blockIdx.y % P.nplanes, frame = blockIdx.y / P.nplanes;
  const ImgPadPlane &pl = P.pl[pli];
  const int b = pl.border, w = pl.plane_w, h = pl.plane_h, xs = P.dst_xs;
  if (b <= 0) return;
  unsigned char *dst = pl.dst + (long long)frame*pl.dst_frame_stride;
  const long long nside = (long long)h*2*b, ntop = (long long)2*b*(w + 2*b);
  for (long long i = (long long)blockIdx.x*blockDim.x + threadIdx.x; i < nside + ntop;
       i += (long long)gridDim.x*blockDim.x) {
    int x, y;
    if (i < nside) {
      y = (int)(i/(2*b));
      const int k = (int)(i - (long long)y*2*b);
      x = k < b ? k - b : w + (k - b);
    }
    else {
      const long long j = i - nside;
      const int r = (int)(j/(w + 2*b));
      x = (int)(j - (long long)r*(w + 2*b)) - b;
      y = r < b ? r - b : h + (r - b);
    }
    const int sx = min(max(x, 0), w - 1), sy = min(max(y, 0), h - 1);
    img_put(dst + (long long)y*pl.dst_ystride + (long long)x*xs, xs,
            img_get(dst + (long long)sy*pl.dst_ystride + (long long)sx*xs, xs));
  }
}

void launch_img_copy_pad(const ImgPadParams &P, cudaStream_t st) {
  int maxw = 1, maxh = 1, maxn = 1, maxb = 0;
  for (int p = 0; p < P.nplanes; p++) {
    maxw = max(maxw, P.pl[p].pic_w);
    maxh = max(maxh, P.pl[p].pic_h);
    maxn = max(maxn, max(P.pl[p].pic_h, P.pl[p].plane_w));
    maxb = max(maxb, P.pl[p].border);
  }
  dim3 g1((unsigned)min((maxw + 255)/256, 64), (unsigned)min(maxh, 512), (unsigned)(P.nplanes*P.nframes));
  img_copy_kernel<<<g1, 256, 0, st>>>(P);
  img_pad_kernel<<<(unsigned)(P.nplanes*P.nframes), 1024, 2*maxn*sizeof(unsigned short), st>>>(P);
  if (maxb > 0) {
    dim3 g3(64, (unsigned)(P.nplanes*P.nframes));
    img_edge_kernel<<<g3, 256, 0, st>>>(P);
  }
}

}  // namespace tdlt

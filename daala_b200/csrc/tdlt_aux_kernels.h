// Launchers of the auxiliary kernels (internal to the library).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

namespace tdlt {

// x / d for x < 2^16 by multiply-high with recip = ceil(2^32 / d) (tdlt_capi.cu recip32; 0 = d is 1).
__device__ __forceinline__ int div_recip(int x, unsigned recip) {
  return recip ? (int)__umulhi((unsigned)x, recip) : x;
}

void launch_pix_to_coeff(int *c, const void *pix, long long nsamples, int xstride, int shift,
                         cudaStream_t st);
void launch_coeff_to_pix(void *pix, const int *c, long long nsamples, int xstride, int shift,
                         cudaStream_t st);
void launch_dct2d_blocks(int bs, bool inverse, int *out, int ostride, long long oblock,
                         const int *in, int istride, long long iblock, int nblocks,
                         cudaStream_t st);
void launch_lap_vectors(int f, bool inverse, int *out, const int *in, int nvec, cudaStream_t st);
void launch_dct1d_vectors(int bs, bool inverse, int *out, const int *in, int nvec, cudaStream_t st);
void launch_lap4_lines(bool inverse, int *base, long long line_step, long long tap_step,
                       int count, cudaStream_t st);
void launch_haar_block(bool inverse, int *y, int ystride, const int *x, int xstride, int ln,
                       cudaStream_t st);
void launch_tf_hv(bool down, int *dst, int dstride, const int *src, int sstride, int n,
                  cudaStream_t st);
// op: 0 up_h_lp(p0 = dx), 1 up_v_lp(p0 = dy), 2 up_hv_lp(p0 = dx, p1 = dy), 3 filter_2d,
// 4 filter_inv_2d, 5 convert_block_down(p0 = curr_size, p1 = dest_size, p2 = filter)
void launch_tf_op(int op, int *dst, int dstride, const int *src, int sstride, int n, int p0,
                  int p1, int p2, cudaStream_t st);

// Chroma-from-luma predictor planes: bw x bh 8x8 luma units per frame, planes with row
// strides lstride / dstride (words, multiples of 4) and frame strides lframe / dframe.
void launch_cfl_pred(int *l, int lstride, long long lframe, const int *d_luma, int dstride,
                     long long dframe, const unsigned char *bsize, int bw, int bh, int nframes,
                     cudaStream_t st);

// Deringing (tdlt_dering.cu): one CTA per 32x32 luma superblock and frame, all planes.
struct DeringParams {
  const int *in[4];
  int *out[4];
  const unsigned char *bskip[4];   // per plane: one flag per 4x4 samples, row stride skip_stride
  const unsigned char *flags;      // one per superblock, nhsb*nvsb per frame
  long long plane_size[4];
  long long skip_frame;            // bytes per frame of every bskip plane
  int dec[4];
  int thr[4];                      // (int)pow(q, 0.84182) per plane, computed on the host
  int nplanes, nhsb, nvsb, skip_stride;
  unsigned recip_nhsb;             // ceil(2^32 / nhsb): row of a superblock by multiply-high
  unsigned recip_npair;            // the same for (nhsb + 1) / 2 (pairs of superblocks)
};
// One 3-D map (x, y, frame) per input plane; box = the superblock's tile with its ring:
// (32 >> dec) + 8 words by (32 >> dec) + 6 rows.
struct DeringMaps {
  CUtensorMap in[4];
  CUtensorMap dir;                 // luma, box 64 x 32: a pair of superblocks without ring
};
// dirthr: 16 ints per superblock and frame (direction | threshold << 3 of every luma 8x8 block)
void launch_dering(const DeringMaps &maps, const DeringParams &P, int *dirthr, int nframes,
                   cudaStream_t st);

// Keyframe bilinear smoothing of the superblocks coded as one 32x32 block, in place.
struct SmoothParams {
  int *c[4];
  const unsigned char *bsize;      // block-size maps, (nvsb*4) x (nhsb*4) per frame
  long long plane_size[4];
  int dec[4];
  int q[4];
  int nplanes, nhsb, nvsb;
};
void launch_smooth(const SmoothParams &P, int nframes, cudaStream_t st);

// Raster <-> coding order of every transform block (tdlt_dering.cu).
struct CodingParams {
  int *v[4];                       // block-linear planes of coding-order vectors
  int *d[4];                       // raster coefficient planes
  const unsigned char *bsize;
  const unsigned char *scan;       // device copy of the four scan tables, (x, y) byte pairs
  long long plane_size[4];
  int scan_len[4];
  int dec[4];
  int nplanes, nhsb, nvsb, to_coding;
  unsigned recip_nhsb;             // ceil(2^32 / nhsb): row of a superblock by multiply-high
  unsigned recip_npair;            // the same for (nhsb + 1) / 2 (pairs of superblocks)
};
void launch_coding_order(const CodingParams &P, int nframes, cudaStream_t st);
// tab: 4*1024 + 3*256 positions (tdlt_region.cuh: kCoTabEntries), see cotab_build_kernel
void launch_cotab_build(unsigned short *tab, const unsigned char *scan, const int lens[4],
                        cudaStream_t st);

// One block (n = 4 or 8) through one of the two deringing filters; `in` dense with its
// 3-sample ring (pitch n + 6), x: the unfiltered block (orthogonal filter only).
void launch_dering_block(short *y, const short *in, const short *x, int n, int thr, int dir,
                         bool orthogonal, cudaStream_t st);

void launch_init_skipped(int *d, const int *pred, int is_keyframe, int n, int w, cudaStream_t st);

// The encoder's input side (tdlt_input.cu): od_img_copy_pad per plane.
struct ImgPadPlane {
  const unsigned char *src;        // first picture sample of frame 0
  unsigned char *dst;              // first visible sample of frame 0's plane buffer
  long long src_frame_stride, dst_frame_stride;    // bytes
  int src_ystride, dst_ystride;    // bytes
  int pic_w, pic_h;                // picture size in this plane
  int plane_w, plane_h;            // padded frame size in this plane
  int border;                      // samples of border around the plane in dst (0: none)
};
struct ImgPadParams {
  ImgPadPlane pl[4];
  int nplanes, nframes;
  int src_xs, dst_xs;              // bytes per sample (1 or 2)
  int src_bits, dst_bits;
};
void launch_img_copy_pad(const ImgPadParams &P, cudaStream_t st);

}  // namespace tdlt

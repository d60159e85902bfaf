/* TEST INFRASTRUCTURE -- not part of the product.
 *
 * Glue that exposes the UNMODIFIED reference (KyleSiefring/daala, compiled from
 * the sources where they lie under /root/reference by oracle/Makefile) to the
 * tests through a flat C interface.  Nothing here re-implements arithmetic:
 * every sample goes through the reference's own functions.
 *
 * This translation unit #includes the reference's src/encode.c (the idiom the
 * reference itself uses in src/tests/test_coef_coder.c:25-34) so that the
 * `static` driver od_compute_dcts (src/encode.c:1286-1343) is reachable.
 * The inverse walk below mirrors the transform slices of od_decode_recursive
 * (src/decode.c:842-866) and od_block_decode (src/decode.c:632-638), which are
 * interleaved with entropy decoding in the reference and so cannot be called
 * as they are.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load the library built from this file.
 */
#include <stdlib.h>
#include <string.h>
#include "encode.c"
#include "tf.h"

typedef struct ref_ctx {
  daala_enc_ctx *enc;
  int pic_width;
  int pic_height;
} ref_ctx;

/* impl: 0 = generic C tables (OD_FDCT_2D_C), 1 = whatever the reference's
   runtime CPU dispatch installs (src/x86/x86state.c:39-97). */
static void ref_pick_vtbl(ref_ctx *r, int impl) {
  int i;
  if (impl == 0) {
    for (i = 0; i <= OD_NBSIZES; i++) {
      r->enc->state.opt_vtbl.fdct_2d[i] = OD_FDCT_2D_C[i];
      r->enc->state.opt_vtbl.idct_2d[i] = OD_IDCT_2D_C[i];
    }
  }
  else {
#if defined(OD_X86ASM)
    od_state_opt_vtbl_init_x86(&r->enc->state);
#endif
  }
}

void *ref_ctx_create(int pic_width, int pic_height) {
  daala_info info;
  ref_ctx *r;
  daala_info_init(&info);
  info.pic_width = pic_width;
  info.pic_height = pic_height;
  info.nplanes = 3;
  info.plane_info[0].xdec = info.plane_info[0].ydec = 0;
  info.plane_info[1].xdec = info.plane_info[1].ydec = 1;
  info.plane_info[2].xdec = info.plane_info[2].ydec = 1;
  info.timebase_numerator = 30;
  info.timebase_denominator = 1;
  info.frame_duration = 1;
  info.pixel_aspect_numerator = 1;
  info.pixel_aspect_denominator = 1;
  info.keyframe_rate = 1;
  r = (ref_ctx *)calloc(1, sizeof(*r));
  if (!r) return NULL;
  r->enc = daala_encode_create(&info);
  if (!r->enc) {
    free(r);
    return NULL;
  }
  r->pic_width = pic_width;
  r->pic_height = pic_height;
  return r;
}

void ref_ctx_free(void *h) {
  ref_ctx *r = (ref_ctx *)h;
  if (!r) return;
  daala_encode_free(r->enc);
  free(r);
}

int ref_ctx_frame_width(void *h) { return ((ref_ctx *)h)->enc->state.frame_width; }
int ref_ctx_frame_height(void *h) { return ((ref_ctx *)h)->enc->state.frame_height; }
int ref_ctx_nhsb(void *h) { return ((ref_ctx *)h)->enc->state.nhsb; }
int ref_ctx_nvsb(void *h) { return ((ref_ctx *)h)->enc->state.nvsb; }

/* Copies a caller map (one byte per 8x8 luma block, nhsb*4 x nvsb*4 entries,
   row stride `stride`) into the state's padded map. */
static void ref_load_bsize(ref_ctx *r, const unsigned char *bsize, int stride) {
  od_state *s = &r->enc->state;
  int i;
  int j;
  for (i = 0; i < s->nvsb*OD_BSIZE_GRID; i++) {
    for (j = 0; j < s->nhsb*OD_BSIZE_GRID; j++) {
      s->bsize[i*s->bstride + j] = bsize[i*stride + j];
    }
  }
}

/* Forward stage of one plane, exactly as od_encode_coefficients runs it
   (src/encode.c:2425-2429, 2492): SB-edge prefilter over the plane, then
   od_compute_dcts on every superblock.  c is lapped in place, d receives the
   coefficients.  Both are (frame_width>>dec) x (frame_height>>dec). */
void ref_tdlt_fwd_plane(void *h, od_coeff *d, od_coeff *c, int dec,
 const unsigned char *bsize, int bsize_stride, int is_keyframe, int use_haar,
 int impl) {
  ref_ctx *r = (ref_ctx *)h;
  od_state *s = &r->enc->state;
  od_mb_enc_ctx ctx;
  od_coeff *dplanes[OD_NPLANES_MAX];
  int sbx;
  int sby;
  ref_pick_vtbl(r, impl);
  ref_load_bsize(r, bsize, bsize_stride);
  memset(&ctx, 0, sizeof(ctx));
  memset(dplanes, 0, sizeof(dplanes));
  dplanes[0] = d;
  ctx.d = dplanes;
  ctx.c = c;
  ctx.is_keyframe = is_keyframe;
  ctx.use_haar_wavelet = use_haar;
  if (!use_haar) {
    od_apply_prefilter_frame_sbs(c, s->frame_width >> dec, s->nhsb, s->nvsb,
     dec, dec);
  }
  for (sby = 0; sby < s->nvsb; sby++) {
    for (sbx = 0; sbx < s->nhsb; sbx++) {
      od_compute_dcts(r->enc, &ctx, 0, sbx, sby, OD_NBSIZES - 1, dec, dec,
       use_haar);
    }
  }
}

static void ref_inv_recursive(ref_ctx *r, od_coeff *c, od_coeff *d, int bx,
 int by, int bsi, int dec, int dc_split, int use_haar) {
  od_state *s = &r->enc->state;
  int obs;
  int bs;
  int w;
  int bo;
  w = s->frame_width >> dec;
  obs = OD_BLOCK_SIZE4x4(s->bsize, s->bstride, bx << bsi, by << bsi);
  bs = OD_MAXI(obs, dec);
  if (bs == bsi) {
    bs -= dec;
    bo = (by << (OD_LOG_BSIZE0 + bs))*w + (bx << (OD_LOG_BSIZE0 + bs));
    if (use_haar) od_haar_inv(c + bo, w, d + bo, w, bs + 2);
    else (*s->opt_vtbl.idct_2d[bs])(c + bo, w, d + bo, w);
  }
  else {
    int f;
    int hfilter;
    int vfilter;
    bs = bsi - dec;
    f = OD_FILT_SIZE(bs - 1, dec);
    bo = (by << (OD_LOG_BSIZE0 + bs))*w + (bx << (OD_LOG_BSIZE0 + bs));
    hfilter = (bx + 1) << (OD_LOG_BSIZE0 + bs) <= s->info.pic_width;
    vfilter = (by + 1) << (OD_LOG_BSIZE0 + bs) <= s->info.pic_height;
    if (dc_split) {
      /* The decoder's top-down DC reconstruction, src/decode.c:722-738, with
         no quantisation or gradient prediction in between. */
      od_coeff x[4];
      int ln;
      int cx;
      int cy;
      ln = bsi - 1 - dec + 2;
      cx = 2*bx;
      cy = 2*by;
      x[0] = d[(cy << ln)*w + (cx << ln)];
      x[1] = d[(cy << ln)*w + ((cx + 1) << ln)];
      x[2] = d[((cy + 1) << ln)*w + (cx << ln)];
      x[3] = d[((cy + 1) << ln)*w + ((cx + 1) << ln)];
      OD_HAAR_KERNEL(x[0], x[1], x[2], x[3]);
      d[(cy << ln)*w + (cx << ln)] = x[0];
      d[(cy << ln)*w + ((cx + 1) << ln)] = x[1];
      d[((cy + 1) << ln)*w + (cx << ln)] = x[2];
      d[((cy + 1) << ln)*w + ((cx + 1) << ln)] = x[3];
    }
    ref_inv_recursive(r, c, d, 2*bx + 0, 2*by + 0, bsi - 1, dec, dc_split,
     use_haar);
    ref_inv_recursive(r, c, d, 2*bx + 1, 2*by + 0, bsi - 1, dec, dc_split,
     use_haar);
    ref_inv_recursive(r, c, d, 2*bx + 0, 2*by + 1, bsi - 1, dec, dc_split,
     use_haar);
    ref_inv_recursive(r, c, d, 2*bx + 1, 2*by + 1, bsi - 1, dec, dc_split,
     use_haar);
    od_postfilter_split(c + bo, w, bs, f, 0, NULL, 0, hfilter, vfilter);
  }
}

/* Inverse stage of one plane: per superblock the decoder's recursion (DC
   split, leaf inverse transforms, split postfilters), then the SB-edge
   postfilter over the plane (src/decode.c:1010-1041).  d is modified when
   dc_split is set, as it is in the decoder. */
void ref_tdlt_inv_plane(void *h, od_coeff *c, od_coeff *d, int dec,
 const unsigned char *bsize, int bsize_stride, int dc_split, int use_haar,
 int impl) {
  ref_ctx *r = (ref_ctx *)h;
  od_state *s = &r->enc->state;
  int sbx;
  int sby;
  ref_pick_vtbl(r, impl);
  ref_load_bsize(r, bsize, bsize_stride);
  for (sby = 0; sby < s->nvsb; sby++) {
    for (sbx = 0; sbx < s->nhsb; sbx++) {
      ref_inv_recursive(r, c, d, sbx, sby, OD_NBSIZES - 1, dec, dc_split,
       use_haar);
    }
  }
  if (!use_haar) {
    od_apply_postfilter_frame_sbs(c, s->frame_width >> dec, s->nhsb, s->nvsb,
     dec, dec, 0, NULL, 0);
  }
}

/* Pixel <-> coefficient conversion (src/state.c:1209-1315) on raw buffers.
   xstride is 1 (8-bit references) or 2 (16-bit references). */
void ref_buf_to_coeff(int bitdepth_mode, od_coeff *dst, int dst_ystride,
 int lossless_p, unsigned char *src, int src_xstride, int src_ystride, int w,
 int h) {
  od_state *s = (od_state *)calloc(1, sizeof(*s));
  s->info.bitdepth_mode = bitdepth_mode;
  od_ref_buf_to_coeff(s, dst, dst_ystride, lossless_p, src, src_xstride,
   src_ystride, w, h);
  free(s);
}

void ref_coeff_to_buf(int bitdepth_mode, unsigned char *dst, int dst_xstride,
 int dst_ystride, od_coeff *src, int src_ystride, int lossless_p, int w,
 int h) {
  od_state *s = (od_state *)calloc(1, sizeof(*s));
  s->info.bitdepth_mode = bitdepth_mode;
  od_coeff_to_ref_buf(s, dst, dst_xstride, dst_ystride, src, src_ystride,
   lossless_p, w, h);
  free(s);
}

/* 1-D transforms through the reference's tables (src/dct.c:58-72). */
void ref_fdct1d(int bs, od_coeff *y, const od_coeff *x, int xstride) {
  (*OD_FDCT_1D[bs])(y, x, xstride);
}

void ref_idct1d(int bs, od_coeff *x, int xstride, const od_coeff *y) {
  (*OD_IDCT_1D[bs])(x, xstride, y);
}

/* Runs the real encoder on one 4:2:0 8-bit frame (y, u, v are pic-sized /
   half-sized planes) as a keyframe, and returns the block-size map its RDO
   chose (nhsb*4 x nvsb*4 bytes, tightly packed) -- used to obtain realistic
   maps for BASELINE config 1/3.  Returns 0 on success. */
int ref_encode_keyframe_bsize(void *h, const unsigned char *y,
 const unsigned char *u, const unsigned char *v, int quality, int complexity,
 unsigned char *bsize_out) {
  ref_ctx *r = (ref_ctx *)h;
  od_state *s = &r->enc->state;
  od_img img;
  daala_packet pkt;
  int i;
  int j;
  int ret;
  int left;
  int bframes;
  memset(&img, 0, sizeof(img));
  img.nplanes = 3;
  img.width = r->pic_width;
  img.height = r->pic_height;
  img.planes[0].data = (unsigned char *)y;
  img.planes[0].xstride = 1;
  img.planes[0].ystride = r->pic_width;
  img.planes[1].data = (unsigned char *)u;
  img.planes[1].xdec = img.planes[1].ydec = 1;
  img.planes[1].xstride = 1;
  img.planes[1].ystride = (r->pic_width + 1) >> 1;
  img.planes[2].data = (unsigned char *)v;
  img.planes[2].xdec = img.planes[2].ydec = 1;
  img.planes[2].xstride = 1;
  img.planes[2].ystride = (r->pic_width + 1) >> 1;
  for (i = 0; i < 3; i++) img.planes[i].bitdepth = 8;
  daala_encode_ctl(r->enc, OD_SET_QUANT, &quality, sizeof(quality));
  daala_encode_ctl(r->enc, OD_SET_COMPLEXITY, &complexity, sizeof(complexity));
  bframes = 0;
  daala_encode_ctl(r->enc, OD_SET_B_FRAMES, &bframes, sizeof(bframes));
  left = 0;
  ret = daala_encode_img_in(r->enc, &img, 0, 0, &left);
  if (ret < 0) return ret;
  while (daala_encode_packet_out(r->enc, 0, &pkt) > 0);
  for (i = 0; i < s->nvsb*OD_BSIZE_GRID; i++) {
    for (j = 0; j < s->nhsb*OD_BSIZE_GRID; j++) {
      bsize_out[i*s->nhsb*OD_BSIZE_GRID + j] = s->bsize[i*s->bstride + j];
    }
  }
  return 0;
}

/* ---- deringing (src/filter.c:1835, the step after the post-filter) ---------------- */

/* od_dering as the decoder calls it (src/decode.c:1104-1110), on caller-provided planes.
   impl: 0 = the C filters, 1 = whatever the runtime dispatch installed (SSE2 twins). */
void ref_dering(void *h, int16_t *y, int ystride, int16_t *x, int xstride, int ln, int sbx,
 int sby, int nhsb, int nvsb, int q, int xdec, int *dir, int pli, unsigned char *bskip,
 int skip_stride, int impl) {
  ref_ctx *r = (ref_ctx *)h;
  od_state *state = &r->enc->state;
  if (impl == 0) {
    state->opt_vtbl.filter_dering_direction[0] = od_filter_dering_direction_4x4_c;
    state->opt_vtbl.filter_dering_direction[1] = od_filter_dering_direction_8x8_c;
    state->opt_vtbl.filter_dering_orthogonal[0] = od_filter_dering_orthogonal_4x4_c;
    state->opt_vtbl.filter_dering_orthogonal[1] = od_filter_dering_orthogonal_8x8_c;
  }
  else {
#if defined(OD_X86ASM)
    od_state_opt_vtbl_init_x86(state);
#endif
  }
  od_dering(state, y, ystride, x, xstride, ln, sbx, sby, nhsb, nvsb, q, xdec,
   (int (*)[OD_DERING_NBLOCKS])dir, pli, bskip, skip_stride);
}

/* The two filters through the dispatch table, one block: bsi 0 = 4x4, 1 = 8x8; `in` has the
   reference's row pitch OD_FILT_BSTRIDE. */
void ref_dering_direction(void *h, int bsi, int16_t *y, int ystride, int16_t *in, int threshold,
 int dir) {
  ref_ctx *r = (ref_ctx *)h;
  (*r->enc->state.opt_vtbl.filter_dering_direction[bsi])(y, ystride, in, threshold, dir);
}

void ref_dering_orthogonal(void *h, int bsi, int16_t *y, int ystride, int16_t *in, int16_t *x,
 int xstride, int threshold, int dir) {
  ref_ctx *r = (ref_ctx *)h;
  (*r->enc->state.opt_vtbl.filter_dering_orthogonal[bsi])(y, ystride, in, x, xstride, threshold,
   dir);
}

/* ---- batched transforms for the self-test streams (src/dct.c:3439-3469, 3565-3598) ----
   nblocks n x n blocks, tightly packed, through the 2-D tables: impl 0 = OD_FDCT_2D_C /
   OD_IDCT_2D_C, 1 = the tables the runtime dispatch installs (SSE2 / SSE4.1 / AVX2). */
void ref_dct2d_blocks(void *h, int bs, int inverse, int impl, od_coeff *out,
 const od_coeff *in, int nblocks) {
  ref_ctx *r = (ref_ctx *)h;
  int n;
  int k;
  od_dct_func_2d f;
  n = 4 << bs;
  ref_pick_vtbl(r, impl);
  f = inverse ? r->enc->state.opt_vtbl.idct_2d[bs] : r->enc->state.opt_vtbl.fdct_2d[bs];
  for (k = 0; k < nblocks; k++) (*f)(out + (size_t)k*n*n, n, in + (size_t)k*n*n, n);
}

/* nvec vectors of n = 4 << bs values through the 1-D tables (src/dct.c:58-72). */
void ref_dct1d_vectors(int bs, int inverse, od_coeff *out, const od_coeff *in, int nvec) {
  int n;
  int k;
  n = 4 << bs;
  for (k = 0; k < nvec; k++) {
    if (inverse) (*OD_IDCT_1D[bs])(out + (size_t)k*n, 1, in + (size_t)k*n);
    else (*OD_FDCT_1D[bs])(out + (size_t)k*n, in + (size_t)k*n, 1);
  }
}

/* ---- frame-level drivers of the neighbouring stages ---------------------------------
   The same loops tests/oracle_libs.py runs from Python (one reference call per superblock
   or block), in C, so that bench.py and tools/bench_*.py can TIME the reference on these
   stages without the interpreter in the way.  Every sample still goes through the
   reference's own functions; tests/test_oracle_vs_ref.py holds them to the Python loops. */

/* od_dering over a frame in the decoder's loop order (src/decode.c:1055-1122): in plays the
   unfiltered copy etmp (narrowed to int16 as that copy is), out the reconstruction ctmp. */
void ref_dering_frame(void *h, od_coeff *const out[], od_coeff *const in[], int nplanes,
 const int dec[], const unsigned char *flags, unsigned char *const bskip[], int skip_stride,
 const int q[], int impl) {
  ref_ctx *r = (ref_ctx *)h;
  od_state *s = &r->enc->state;
  int16_t *e[OD_NPLANES_MAX];
  int16_t buf[32*32];
  int W;
  int H;
  int pli;
  int sbx;
  int sby;
  W = s->frame_width;
  H = s->frame_height;
  for (pli = 0; pli < nplanes; pli++) {
    size_t n;
    size_t i;
    n = (size_t)(W >> dec[pli])*(H >> dec[pli]);
    e[pli] = (int16_t *)malloc(n*sizeof(int16_t));
    for (i = 0; i < n; i++) {
      e[pli][i] = (int16_t)in[pli][i];
      out[pli][i] = in[pli][i];
    }
  }
  for (sby = 0; sby < s->nvsb; sby++) {
    for (sbx = 0; sbx < s->nhsb; sbx++) {
      int dir[OD_DERING_NBLOCKS][OD_DERING_NBLOCKS];
      if (!flags[sby*s->nhsb + sbx]) continue;
      memset(dir, 0, sizeof(dir));
      for (pli = 0; pli < nplanes; pli++) {
        int ln;
        int n;
        int w;
        int y;
        int x;
        ln = 5 - dec[pli];
        n = 1 << ln;
        w = W >> dec[pli];
        ref_dering(h, buf, n, e[pli] + (size_t)(sby << ln)*w + (sbx << ln), w, ln, sbx, sby,
         s->nhsb, s->nvsb, q[pli], dec[pli], &dir[0][0], pli,
         bskip[pli] + (sby << (3 - dec[pli]))*skip_stride + (sbx << (3 - dec[pli])),
         skip_stride, impl);
        for (y = 0; y < n; y++) {
          for (x = 0; x < n; x++) {
            out[pli][(size_t)((sby << ln) + y)*w + (sbx << ln) + x] = buf[y*n + x];
          }
        }
      }
    }
  }
  for (pli = 0; pli < nplanes; pli++) free(e[pli]);
}

/* od_smooth_recursive per superblock and plane with the decoder's arguments
   (src/decode.c:1139-1143), in place. */
void ref_smooth_frame(void *h, od_coeff *const c[], int nplanes, const int dec[],
 unsigned char *bsize, int bstride, const int q[]) {
  ref_ctx *r = (ref_ctx *)h;
  od_state *s = &r->enc->state;
  int pli;
  int sbx;
  int sby;
  for (pli = 0; pli < nplanes; pli++) {
    for (sby = 0; sby < s->nvsb; sby++) {
      for (sbx = 0; sbx < s->nhsb; sbx++) {
        od_smooth_recursive(c[pli], bsize, bstride, sbx, sby, 3, s->frame_width >> dec[pli],
         dec[pli], dec[pli], 3, q[pli], pli);
      }
    }
  }
}

/* od_resample_luma_coeffs once per chroma leaf block of a 4:2:0 frame, with the arguments
   of its call sites (src/encode.c:1512-1517): l is the half-resolution predictor plane. */
void ref_cfl_pred_frame(od_coeff *l, const od_coeff *d_luma, int W, int H,
 const unsigned char *bsize, int bstride) {
  int uy;
  int ux;
  int cw;
  cw = W >> 1;
  for (uy = 0; uy < H >> 3; uy++) {
    for (ux = 0; ux < W >> 3; ux++) {
      int obs;
      int bsi;
      int m;
      obs = bsize[uy*bstride + ux];
      bsi = obs > 1 ? obs : 1;
      m = 1 << (bsi - 1);
      if ((ux & (m - 1)) || (uy & (m - 1))) continue;
      od_resample_luma_coeffs(l + (size_t)(uy*4)*cw + ux*4, cw,
       d_luma + (size_t)(uy*8)*W + ux*8, W, 1, 1, bsi - 1, obs);
    }
  }
}

/* od_raster_to_coding_order / od_coding_order_to_raster (src/partition.c:144-200) block by
   block over the map, vectors placed in the product's block-linear plane
   (include/daala_tdlt.h: superblocks in raster order, 4x4 units in Z order inside). */
static size_t ref_block_linear_offset(int x0, int y0, int w, int S) {
  int ux;
  int uy;
  int z;
  int b;
  ux = (x0 % S) >> 2;
  uy = (y0 % S) >> 2;
  z = 0;
  for (b = 0; b < 3; b++) z |= (((ux >> b) & 1) << (2*b)) | (((uy >> b) & 1) << (2*b + 1));
  return ((size_t)(y0/S)*(w/S) + x0/S)*S*S + 16*(size_t)z;
}

void ref_coding_order_plane(od_coeff *v, od_coeff *d, int W, int H, int dec,
 const unsigned char *bsize, int bstride, int to_coding) {
  od_coeff tmp[32*32];
  int uy;
  int ux;
  for (uy = 0; uy < (H << dec) >> 3; uy++) {
    for (ux = 0; ux < (W << dec) >> 3; ux++) {
      int obs;
      int lu;
      int n;
      int nsub;
      int k;
      obs = bsize[uy*bstride + ux];
      lu = obs > 1 ? obs - 1 : 0;
      if ((ux & ((1 << lu) - 1)) || (uy & ((1 << lu) - 1))) continue;
      n = (4 << obs) >> dec;
      if (n < 4) n = 4;
      nsub = (obs == 0 && dec == 0) ? 4 : 1;
      for (k = 0; k < nsub; k++) {
        int x0;
        int y0;
        int ncoded;
        od_coeff *vec;
        x0 = ((ux*8) >> dec) + 4*(k & 1);
        y0 = ((uy*8) >> dec) + 4*(k >> 1);
        vec = v + ref_block_linear_offset(x0, y0, W, 32 >> dec);
        ncoded = n*n < 512 ? n*n : 512;
        if (to_coding) {
          od_raster_to_coding_order(tmp, n, d + (size_t)y0*W + x0, W);
          memcpy(vec, tmp, ncoded*sizeof(od_coeff));
        }
        else {
          memcpy(tmp, vec, n*n*sizeof(od_coeff));
          od_coding_order_to_raster(d + (size_t)y0*W + x0, W, tmp, n);
        }
      }
    }
  }
}

/* ---- the whole codec through the reference's own call sites ------------------------------
   daala_encode_img_in -> daala_encode_packet_out -> daala_decode_packet_in ->
   daala_decode_img_out over a sequence of 8-bit 4:2:0 frames (src/encode.c:2876, :3246,
   src/decode.c:1159), i.e. the loop of examples/encoder_example.c + dump_video.c without the
   Ogg container.  impl selects what the dispatch tables of BOTH states hold: 0 = the C
   functions, 1 = the runtime x86 dispatch, 2 = the tables a caller installed with
   ref_set_external_vtbl (the CUDA library's od_tdlt_opt_vtbl_init_cuda /
   od_tdlt_dering_vtbl_init_cuda).  In _ref/libdaala_ref_cuda.so (cuda_redirect.h) the direct
   calls -- split and superblock-edge filters, Haar wavelet -- go to the CUDA library as
   well, whatever impl says.  Every packet is decoded as it is produced, as
   OD_ENCODER_CHECK's in-loop decoder does (src/encode.c:3203-3243). */
#include "decint.h"

typedef void (*ref_dct_vtbl_init)(od_dct_func_2d fdct_2d[], od_dct_func_2d idct_2d[]);
typedef void (*ref_dering_vtbl_init)(od_filter_dering_direction_func direction[],
 od_filter_dering_orthogonal_func orthogonal[]);
static ref_dct_vtbl_init ref_ext_dct_init;
static ref_dering_vtbl_init ref_ext_dering_init;

void ref_set_external_vtbl(void *dct_init, void *dering_init) {
  ref_ext_dct_init = (ref_dct_vtbl_init)dct_init;
  ref_ext_dering_init = (ref_dering_vtbl_init)dering_init;
}

int ref_direct_calls_redirected(void) {
#if defined(REF_GLUE_CUDA_REDIRECT)
  return 1;
#else
  return 0;
#endif
}

#if defined(REF_GLUE_CUDA_REDIRECT)
void ref_chk_install_vtbl(od_state *s, void *dct_init, void *dering_init);
#endif

static int ref_state_vtbl(od_state *s, int impl) {
  od_state_opt_vtbl_init_c(s);
  if (impl == 1) {
#if defined(OD_X86ASM)
    od_state_opt_vtbl_init_x86(s);
#endif
  }
  else if (impl == 2) {
    if (!ref_ext_dct_init) return -1;
    (*ref_ext_dct_init)(s->opt_vtbl.fdct_2d, s->opt_vtbl.idct_2d);
    if (ref_ext_dering_init) {
      (*ref_ext_dering_init)(s->opt_vtbl.filter_dering_direction,
       s->opt_vtbl.filter_dering_orthogonal);
    }
  }
#if defined(REF_GLUE_CUDA_REDIRECT)
  else if (impl == 3) {
    /* OD_CHECKASM style: every table entry runs the C function and the external one and
       compares (oracle/cuda_check.c); the direct calls do the same while ref_chk_mode is set */
    if (!ref_ext_dct_init) return -1;
    ref_chk_install_vtbl(s, (void *)ref_ext_dct_init, (void *)ref_ext_dering_init);
  }
#endif
  return 0;
}

/* yuv: nframes x (y: w*h, u: cw*ch, v: cw*ch) bytes, cw = (w+1)/2.  packets: the data packets
   back to back (capacity packets_cap), packet_bytes[f] = size of frame f's packet.  recon:
   the decoder's output pictures in the layout of yuv.  Returns the number of frames the
   decoder put out, or a negative error. */
/* Rows in which the decoder's reconstruction differed from the encoder's during the last
   ref_codec_roundtrip (OD_ENCODER_CHECK's "Pixel mismatch in row" count; 0 = in step). */
static long long ref_encoder_check_rows;
long long ref_encoder_check_mismatched_rows(void) { return ref_encoder_check_rows; }

int ref_codec_roundtrip(int pic_width, int pic_height, int nframes, const unsigned char *yuv,
 int quality, int complexity, int keyframe_rate, int impl, unsigned char *packets,
 long long packets_cap, int *packet_bytes, unsigned char *recon) {
  daala_info info;
  daala_info dinfo;
  daala_comment dc;
  daala_setup_info *setup;
  daala_enc_ctx *enc;
  daala_dec_ctx *dec;
  daala_packet pkt;
  od_img img;
  od_img out;
  long long used;
  size_t ysz;
  size_t csz;
  int cw;
  int ch;
  int f;
  int nout;
  int ret;
  int zero;
  daala_info_init(&info);
  info.pic_width = pic_width;
  info.pic_height = pic_height;
  info.nplanes = 3;
  info.plane_info[0].xdec = info.plane_info[0].ydec = 0;
  info.plane_info[1].xdec = info.plane_info[1].ydec = 1;
  info.plane_info[2].xdec = info.plane_info[2].ydec = 1;
  info.timebase_numerator = 30;
  info.timebase_denominator = 1;
  info.frame_duration = 1;
  info.pixel_aspect_numerator = 1;
  info.pixel_aspect_denominator = 1;
  info.keyframe_rate = keyframe_rate;
  enc = daala_encode_create(&info);
  if (!enc) return -100;
  zero = 0;
  daala_encode_ctl(enc, OD_SET_QUANT, &quality, sizeof(quality));
  daala_encode_ctl(enc, OD_SET_COMPLEXITY, &complexity, sizeof(complexity));
  daala_encode_ctl(enc, OD_SET_B_FRAMES, &zero, sizeof(zero));
  if (ref_state_vtbl(&enc->state, impl)) { daala_encode_free(enc); return -101; }
  /* headers -> decoder */
  daala_comment_init(&dc);
  daala_info_init(&dinfo);
  setup = NULL;
  {
    daala_comment ec;
    daala_comment_init(&ec);
    while ((ret = daala_encode_flush_header(enc, &ec, &pkt)) > 0) {
      ret = daala_decode_header_in(&dinfo, &dc, &setup, &pkt);
      if (ret < 0) { daala_encode_free(enc); return -102; }
    }
    daala_comment_clear(&ec);
  }
  dec = daala_decode_create(&dinfo, setup);
  daala_setup_free(setup);
  if (!dec) { daala_encode_free(enc); return -103; }
  if (ref_state_vtbl(&dec->state, impl)) { daala_decode_free(dec); daala_encode_free(enc); return -101; }
  cw = (pic_width + 1) >> 1;
  ch = (pic_height + 1) >> 1;
  ysz = (size_t)pic_width*pic_height;
  csz = (size_t)cw*ch;
  used = 0;
  nout = 0;
  ref_encoder_check_rows = 0;
  for (f = 0; f < nframes; f++) {
    const unsigned char *fr;
    int last;
    int left;
    int i;
    fr = yuv + (size_t)f*(ysz + 2*csz);
    last = f == nframes - 1;
    memset(&img, 0, sizeof(img));
    img.nplanes = 3;
    img.width = pic_width;
    img.height = pic_height;
    img.planes[0].data = (unsigned char *)fr;
    img.planes[0].xstride = 1;
    img.planes[0].ystride = pic_width;
    img.planes[1].data = (unsigned char *)fr + ysz;
    img.planes[2].data = (unsigned char *)fr + ysz + csz;
    for (i = 1; i < 3; i++) {
      img.planes[i].xdec = img.planes[i].ydec = 1;
      img.planes[i].xstride = 1;
      img.planes[i].ystride = cw;
    }
    for (i = 0; i < 3; i++) img.planes[i].bitdepth = 8;
    left = 0;
    /* (end_of_input means "this call brings no frame", src/encode.c:2907-2923) */
    ret = daala_encode_img_in(enc, &img, 0, 0, &left);
    if (ret < 0) { daala_decode_free(dec); daala_encode_free(enc); return ret; }
    packet_bytes[f] = 0;
    while (daala_encode_packet_out(enc, last, &pkt) > 0) {
      int pli;
      if (used + pkt.bytes > packets_cap) { daala_decode_free(dec); daala_encode_free(enc); return -104; }
      memcpy(packets + used, pkt.packet, pkt.bytes);
      used += pkt.bytes;
      packet_bytes[f] += (int)pkt.bytes;
      ret = daala_decode_packet_in(dec, &pkt);
      if (ret < 0) { daala_decode_free(dec); daala_encode_free(enc); return -105; }
      /* OD_ENCODER_CHECK (src/encode.c:3203-3243): the decoder's reconstruction of this
         packet against the encoder's own, row by row over the padded frame */
      {
        od_img *ei;
        od_img *di;
        ei = enc->state.ref_imgs + enc->state.ref_imgi[OD_FRAME_SELF];
        di = dec->state.ref_imgs + dec->state.ref_imgi[OD_FRAME_SELF];
        for (pli = 0; pli < ei->nplanes; pli++) {
          int pw;
          int ph;
          int y;
          pw = dec->state.frame_width >> di->planes[pli].xdec;
          ph = dec->state.frame_height >> di->planes[pli].ydec;
          for (y = 0; y < ph; y++) {
            if (memcmp(ei->planes[pli].data + (size_t)ei->planes[pli].ystride*y,
             di->planes[pli].data + (size_t)di->planes[pli].ystride*y, pw)) {
              ref_encoder_check_rows++;
            }
          }
        }
      }
      /* one picture per packet at most (examples/dump_video.c:414-436); the pictures the
         decoder still holds come out after the last packet, when it has seen e_o_s */
      while (daala_decode_img_out(dec, &out) > 0) {
        unsigned char *dst;
        if (nout >= nframes) break;
        dst = recon + (size_t)nout*(ysz + 2*csz);
        for (pli = 0; pli < 3; pli++) {
          int w;
          int h;
          int y;
          w = pli ? cw : pic_width;
          h = pli ? ch : pic_height;
          for (y = 0; y < h; y++) {
            memcpy(dst + (size_t)y*w, out.planes[pli].data + (size_t)y*out.planes[pli].ystride, w);
          }
          dst += (size_t)w*h;
        }
        nout++;
        if (!last) break;
      }
    }
  }
  daala_comment_clear(&dc);
  daala_decode_free(dec);
  daala_encode_free(enc);
  return nout;
}

/* ---- the encoder's input side: od_img_copy_pad (src/encode.c:1728-1750) on one plane ----
   od_img_plane_copy_pad (static, src/encode.c:610) + od_img_edge_ext (src/state.c:1161) on
   caller buffers: dst -> first visible sample of a plane buffer with a border of
   OD_BUFFER_PADDING >> dec samples all round. */
void ref_img_copy_pad_plane(unsigned char *dst, int dst_xstride, int dst_ystride, int dst_bits,
 int plane_w, int plane_h, int dec, unsigned char *src, int src_xstride, int src_ystride,
 int src_bits, int pic_w, int pic_h, int edge_ext) {
  od_img d;
  od_img s;
  memset(&d, 0, sizeof(d));
  memset(&s, 0, sizeof(s));
  d.nplanes = 1;
  d.width = plane_w << dec;
  d.height = plane_h << dec;
  d.planes[0].data = dst;
  d.planes[0].xstride = dst_xstride;
  d.planes[0].ystride = dst_ystride;
  d.planes[0].xdec = d.planes[0].ydec = dec;
  d.planes[0].bitdepth = dst_bits;
  s.nplanes = 1;
  s.width = pic_w << dec;
  s.height = pic_h << dec;
  s.planes[0].data = src;
  s.planes[0].xstride = src_xstride;
  s.planes[0].ystride = src_ystride;
  s.planes[0].xdec = s.planes[0].ydec = dec;
  s.planes[0].bitdepth = src_bits;
  od_img_plane_copy_pad(&d, plane_w, plane_h, &s, pic_w, pic_h, 0);
  if (edge_ext) od_img_edge_ext(&d);
}
int ref_buffer_padding(void) { return OD_BUFFER_PADDING; }

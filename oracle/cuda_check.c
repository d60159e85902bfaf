/* TEST INFRASTRUCTURE -- the reference's OD_CHECKASM idea (src/x86/x86dct.h:305-315,
 * od_dct_check src/dct.c:2167-2189) for the CUDA library: every call the reference's
 * encoder and decoder make into the path runs BOTH the reference's C function and the CUDA
 * entry point on the same input and the two results are compared element by element.
 *
 * Built only into oracle/_ref/libdaala_ref_cuda.so.  cuda_redirect.h sends the direct calls
 * of src/encode.c / src/decode.c to the ref_chk_* functions below; this file is compiled
 * WITHOUT that header, so it sees both the reference's own functions (src/filter.c,
 * src/dct.c) and the product's *_cuda symbols (resolved at load time from
 * libdaala_tdlt_cuda.so).  With ref_chk_mode == 0 the functions just forward to CUDA.
 * The dispatch-table entries (DCTs, deringing filters) are checked the same way through
 * ref_chk_install_vtbl.  Counters are read back by tests/test_zy_dropin_codec.py. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "filter.h"
#include "dct.h"
#include "state.h"

void od_prefilter_split_cuda(od_coeff *c0, int stride, int bs, int f, int hfilter, int vfilter);
void od_postfilter_split_cuda(od_coeff *c0, int stride, int bs, int f, int q,
 unsigned char *skip, int skip_stride, int hfilter, int vfilter);
void od_apply_prefilter_frame_sbs_cuda(od_coeff *c, int stride, int nhsb, int nvsb, int xdec,
 int ydec);
void od_apply_postfilter_frame_sbs_cuda(od_coeff *c, int stride, int nhsb, int nvsb, int xdec,
 int ydec, int q, unsigned char *skip, int skip_stride);
void od_haar_cuda(od_coeff *y, int ystride, const od_coeff *x, int xstride, int ln);
void od_haar_inv_cuda(od_coeff *x, int xstride, const od_coeff *y, int ystride, int ln);

int ref_chk_mode;
long long ref_chk_calls;
long long ref_chk_mismatches;
char ref_chk_first[256];

void ref_chk_reset(int mode) {
  ref_chk_mode = mode;
  ref_chk_calls = 0;
  ref_chk_mismatches = 0;
  ref_chk_first[0] = 0;
}
long long ref_chk_get_calls(void) { return ref_chk_calls; }
long long ref_chk_get_mismatches(void) { return ref_chk_mismatches; }
const char *ref_chk_get_first(void) { return ref_chk_first; }

/* Compares an h x w rectangle of two strided buffers (elements of esize bytes). */
static void chk_cmp(const char *what, const void *a, int astride, const void *b, int bstride,
 int w, int h, int esize, int p0, int p1, int p2) {
  int y;
  ref_chk_calls++;
  for (y = 0; y < h; y++) {
    const char *ra = (const char *)a + (size_t)y*astride*esize;
    const char *rb = (const char *)b + (size_t)y*bstride*esize;
    if (memcmp(ra, rb, (size_t)w*esize)) {
      if (!ref_chk_mismatches) {
        int x;
        for (x = 0; x < w && !memcmp(ra + x*esize, rb + x*esize, esize); x++);
        snprintf(ref_chk_first, sizeof(ref_chk_first),
         "%s(%d,%d,%d): %dx%d, first difference at row %d col %d (call %lld)", what, p0, p1, p2, w,
         h, y, x, ref_chk_calls);
      }
      ref_chk_mismatches++;
      return;
    }
  }
}

void ref_chk_prefilter_split(od_coeff *c0, int stride, int bs, int f, int hfilter,
 int vfilter) {
  int n;
  int y;
  od_coeff *t;
  if (!ref_chk_mode) { od_prefilter_split_cuda(c0, stride, bs, f, hfilter, vfilter); return; }
  n = 4 << bs;
  t = (od_coeff *)malloc(sizeof(*t)*n*n);
  for (y = 0; y < n; y++) memcpy(t + y*n, c0 + (size_t)y*stride, sizeof(*t)*n);
  od_prefilter_split(t, n, bs, f, hfilter, vfilter);
  od_prefilter_split_cuda(c0, stride, bs, f, hfilter, vfilter);
  chk_cmp("od_prefilter_split", c0, stride, t, n, n, n, sizeof(*t), bs, hfilter, vfilter);
  free(t);
}

void ref_chk_postfilter_split(od_coeff *c0, int stride, int bs, int f, int q,
 unsigned char *skip, int skip_stride, int hfilter, int vfilter) {
  int n;
  int y;
  od_coeff *t;
  if (!ref_chk_mode) {
    od_postfilter_split_cuda(c0, stride, bs, f, q, skip, skip_stride, hfilter, vfilter);
    return;
  }
  n = 4 << bs;
  t = (od_coeff *)malloc(sizeof(*t)*n*n);
  for (y = 0; y < n; y++) memcpy(t + y*n, c0 + (size_t)y*stride, sizeof(*t)*n);
  od_postfilter_split(t, n, bs, f, q, skip, skip_stride, hfilter, vfilter);
  od_postfilter_split_cuda(c0, stride, bs, f, q, skip, skip_stride, hfilter, vfilter);
  chk_cmp("od_postfilter_split", c0, stride, t, n, n, n, sizeof(*t), bs, hfilter, vfilter);
  free(t);
}

void ref_chk_apply_prefilter_frame_sbs(od_coeff *c, int stride, int nhsb, int nvsb, int xdec,
 int ydec) {
  int w;
  int h;
  int y;
  od_coeff *t;
  if (!ref_chk_mode) { od_apply_prefilter_frame_sbs_cuda(c, stride, nhsb, nvsb, xdec, ydec); return; }
  w = (nhsb*32) >> xdec;
  h = (nvsb*32) >> ydec;
  t = (od_coeff *)malloc(sizeof(*t)*w*h);
  for (y = 0; y < h; y++) memcpy(t + (size_t)y*w, c + (size_t)y*stride, sizeof(*t)*w);
  od_apply_prefilter_frame_sbs(t, w, nhsb, nvsb, xdec, ydec);
  od_apply_prefilter_frame_sbs_cuda(c, stride, nhsb, nvsb, xdec, ydec);
  chk_cmp("od_apply_prefilter_frame_sbs", c, stride, t, w, w, h, sizeof(*t), nhsb, nvsb, xdec);
  free(t);
}

void ref_chk_apply_postfilter_frame_sbs(od_coeff *c, int stride, int nhsb, int nvsb, int xdec,
 int ydec, int q, unsigned char *skip, int skip_stride) {
  int w;
  int h;
  int y;
  od_coeff *t;
  if (!ref_chk_mode) {
    od_apply_postfilter_frame_sbs_cuda(c, stride, nhsb, nvsb, xdec, ydec, q, skip, skip_stride);
    return;
  }
  w = (nhsb*32) >> xdec;
  h = (nvsb*32) >> ydec;
  t = (od_coeff *)malloc(sizeof(*t)*w*h);
  for (y = 0; y < h; y++) memcpy(t + (size_t)y*w, c + (size_t)y*stride, sizeof(*t)*w);
  od_apply_postfilter_frame_sbs(t, w, nhsb, nvsb, xdec, ydec, q, skip, skip_stride);
  od_apply_postfilter_frame_sbs_cuda(c, stride, nhsb, nvsb, xdec, ydec, q, skip, skip_stride);
  chk_cmp("od_apply_postfilter_frame_sbs", c, stride, t, w, w, h, sizeof(*t), nhsb, nvsb, xdec);
  free(t);
}

void ref_chk_haar(od_coeff *y, int ystride, const od_coeff *x, int xstride, int ln) {
  if (ref_chk_mode) {
    int n;
    od_coeff *t;
    n = 1 << ln;
    t = (od_coeff *)malloc(sizeof(*t)*n*n);
    od_haar(t, n, x, xstride, ln);
    od_haar_cuda(y, ystride, x, xstride, ln);
    chk_cmp("od_haar", y, ystride, t, n, n, n, sizeof(*t), ln, 0, 0);
    free(t);
  }
  else od_haar_cuda(y, ystride, x, xstride, ln);
}

void ref_chk_haar_inv(od_coeff *x, int xstride, const od_coeff *y, int ystride, int ln) {
  if (ref_chk_mode) {
    int n;
    od_coeff *t;
    n = 1 << ln;
    t = (od_coeff *)malloc(sizeof(*t)*n*n);
    od_haar_inv(t, n, y, ystride, ln);
    od_haar_inv_cuda(x, xstride, y, ystride, ln);
    chk_cmp("od_haar_inv", x, xstride, t, n, n, n, sizeof(*t), ln, 0, 0);
    free(t);
  }
  else od_haar_inv_cuda(x, xstride, y, ystride, ln);
}

/* ---- dispatch-table entries: the external (CUDA) tables are kept here and the state gets
   checking trampolines, one per table slot ------------------------------------------- */
static od_dct_func_2d chk_cuda_fdct[OD_NBSIZES + 1];
static od_dct_func_2d chk_cuda_idct[OD_NBSIZES + 1];
static od_filter_dering_direction_func chk_cuda_dir[OD_DERINGSIZES];
static od_filter_dering_orthogonal_func chk_cuda_orth[OD_DERINGSIZES];

static void chk_dct(int bs, int inverse, od_coeff *out, int out_stride, const od_coeff *in,
 int in_stride) {
  int n;
  od_coeff *t;
  n = 4 << bs;
  t = (od_coeff *)malloc(sizeof(*t)*n*n);
  (*(inverse ? OD_IDCT_2D_C : OD_FDCT_2D_C)[bs])(t, n, in, in_stride);
  (*(inverse ? chk_cuda_idct : chk_cuda_fdct)[bs])(out, out_stride, in, in_stride);
  chk_cmp(inverse ? "idct_2d" : "fdct_2d", out, out_stride, t, n, n, n, sizeof(*t), bs, 0, 0);
  free(t);
}
#define CHK_DCT(bs) \
  static void chk_fdct##bs(od_coeff *o, int os, const od_coeff *i, int is) { chk_dct(bs, 0, o, os, i, is); } \
  static void chk_idct##bs(od_coeff *o, int os, const od_coeff *i, int is) { chk_dct(bs, 1, o, os, i, is); }
CHK_DCT(0)
CHK_DCT(1)
CHK_DCT(2)
CHK_DCT(3)

static void chk_dering_direction(int bsi, int16_t *y, int ystride, int16_t *in, int threshold,
 int dir) {
  int n;
  int16_t t[8*8];
  n = 4 << bsi;
  if (bsi) od_filter_dering_direction_8x8_c(t, n, in, threshold, dir);
  else od_filter_dering_direction_4x4_c(t, n, in, threshold, dir);
  (*chk_cuda_dir[bsi])(y, ystride, in, threshold, dir);
  chk_cmp("filter_dering_direction", y, ystride, t, n, n, n, sizeof(*t), bsi, threshold, dir);
}
static void chk_dering_orthogonal(int bsi, int16_t *y, int ystride, int16_t *in, int16_t *x,
 int xstride, int threshold, int dir) {
  int n;
  int16_t t[8*8];
  n = 4 << bsi;
  if (bsi) od_filter_dering_orthogonal_8x8_c(t, n, in, x, xstride, threshold, dir);
  else od_filter_dering_orthogonal_4x4_c(t, n, in, x, xstride, threshold, dir);
  (*chk_cuda_orth[bsi])(y, ystride, in, x, xstride, threshold, dir);
  chk_cmp("filter_dering_orthogonal", y, ystride, t, n, n, n, sizeof(*t), bsi, threshold, dir);
}
static void chk_dir0(int16_t *y, int ys, int16_t *in, int th, int d) { chk_dering_direction(0, y, ys, in, th, d); }
static void chk_dir1(int16_t *y, int ys, int16_t *in, int th, int d) { chk_dering_direction(1, y, ys, in, th, d); }
static void chk_orth0(int16_t *y, int ys, int16_t *in, int16_t *x, int xs, int th, int d) {
  chk_dering_orthogonal(0, y, ys, in, x, xs, th, d);
}
static void chk_orth1(int16_t *y, int ys, int16_t *in, int16_t *x, int xs, int th, int d) {
  chk_dering_orthogonal(1, y, ys, in, x, xs, th, d);
}

typedef void (*chk_dct_init)(od_dct_func_2d fdct_2d[], od_dct_func_2d idct_2d[]);
typedef void (*chk_dering_init)(od_filter_dering_direction_func direction[],
 od_filter_dering_orthogonal_func orthogonal[]);

/* Fills the state's tables with the checking trampolines; the CUDA functions they call come
   from the two init calls of the product (od_tdlt_opt_vtbl_init_cuda,
   od_tdlt_dering_vtbl_init_cuda). */
void ref_chk_install_vtbl(od_state *s, void *dct_init, void *dering_init) {
  static const od_dct_func_2d F[4] = {chk_fdct0, chk_fdct1, chk_fdct2, chk_fdct3};
  static const od_dct_func_2d I[4] = {chk_idct0, chk_idct1, chk_idct2, chk_idct3};
  int i;
  memcpy(chk_cuda_fdct, OD_FDCT_2D_C, sizeof(chk_cuda_fdct));
  memcpy(chk_cuda_idct, OD_IDCT_2D_C, sizeof(chk_cuda_idct));
  (*(chk_dct_init)dct_init)(chk_cuda_fdct, chk_cuda_idct);
  for (i = 0; i < 4; i++) {
    s->opt_vtbl.fdct_2d[i] = F[i];
    s->opt_vtbl.idct_2d[i] = I[i];
  }
  if (dering_init) {
    (*(chk_dering_init)dering_init)(chk_cuda_dir, chk_cuda_orth);
    s->opt_vtbl.filter_dering_direction[0] = chk_dir0;
    s->opt_vtbl.filter_dering_direction[1] = chk_dir1;
    s->opt_vtbl.filter_dering_orthogonal[0] = chk_orth0;
    s->opt_vtbl.filter_dering_orthogonal[1] = chk_orth1;
  }
}

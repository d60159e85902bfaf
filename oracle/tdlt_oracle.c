/* TEST INFRASTRUCTURE -- CPU restatement of the reference's TDLT path.
 * See tdlt_oracle.h for the rules on who may use it and how it is pinned.
 *
 * Written for clarity, not speed: the DCTs run through a small interpreter
 * over step tables (lift_tables_gen.h, produced from tools/gen_lifting.py),
 * the lapping filters through one parametric routine.
 */
#include <string.h>
#include "tdlt_oracle.h"
#include "lift_tables_gen.h"

#define OC_BSIZE_MAX 32
#define OC_COEFF_SHIFT 4 /* reference src/internal.h:119 */

/* Division by two rounding toward zero: reference OD_DCT_RSHIFT(a, 1),
   src/filter.h:40-43. */
static int oc_half(int a) {
  return (int)(((uint32_t)a >> 31) + (uint32_t)a) >> 1;
}

/* ---------------------------------------------------------------- DCTs -- */

typedef struct {
  int n;
  const unsigned char *in;
  const tl_step *steps;
  int nsteps;
} oc_prog;

#define OC_PROG(N, NAME) \
  {N, TL_##NAME##_IN, TL_##NAME##_STEPS, \
   (int)(sizeof(TL_##NAME##_STEPS)/sizeof(TL_##NAME##_STEPS[0]))}

static const oc_prog OC_PROGS[4] = {
  OC_PROG(4, DCT4), OC_PROG(8, DCT8), OC_PROG(16, DCT16), OC_PROG(32, DCT32)
};

static int oc_lift(int s, int k, int shift) {
  return (s*k + (1 << (shift - 1))) >> shift;
}

static void oc_step(int *r, int op, int d, int s, int k, int shift) {
  switch (op) {
    case TL_ADD: r[d] += r[s]; break;
    case TL_SUB: r[d] -= r[s]; break;
    case TL_RSUB: r[d] = r[s] - r[d]; break;
    case TL_ADDH: r[d] += oc_half(r[s]); break;
    case TL_SUBH: r[d] -= oc_half(r[s]); break;
    case TL_RSUBH: r[d] = oc_half(r[s]) - r[d]; break;
    case TL_NEG: r[d] = -r[d]; break;
    case TL_LADD: r[d] += oc_lift(r[s], k, shift); break;
    case TL_LSUB: r[d] -= oc_lift(r[s], k, shift); break;
  }
}

static int oc_inverse_op(int op) {
  switch (op) {
    case TL_ADD: return TL_SUB;
    case TL_SUB: return TL_ADD;
    case TL_ADDH: return TL_SUBH;
    case TL_SUBH: return TL_ADDH;
    case TL_LADD: return TL_LSUB;
    case TL_LSUB: return TL_LADD;
    default: return op; /* RSUB, RSUBH, NEG are involutions */
  }
}

/* reference src/dct.c:74 (4), :151 (8), :349 (16), :1757 (32) */
void oc_fdct1d(int bs, oc_coeff *y, const oc_coeff *x, int xstride) {
  const oc_prog *p = &OC_PROGS[bs];
  int r[32];
  int i;
  for (i = 0; i < p->n; i++) r[p->in[i]] = x[i*xstride];
  for (i = 0; i < p->nsteps; i++) {
    const tl_step *s = &p->steps[i];
    oc_step(r, s->op, s->dst, s->src, s->k, s->shift);
  }
  for (i = 0; i < p->n; i++) y[i] = r[i];
}

/* reference src/dct.c:114 (4), :271 (8), :642 (16), :1859 (32) */
void oc_idct1d(int bs, oc_coeff *x, int xstride, const oc_coeff *y) {
  const oc_prog *p = &OC_PROGS[bs];
  int r[32];
  int i;
  for (i = 0; i < p->n; i++) r[i] = y[i];
  for (i = p->nsteps - 1; i >= 0; i--) {
    const tl_step *s = &p->steps[i];
    oc_step(r, oc_inverse_op(s->op), s->dst, s->src, s->k, s->shift);
  }
  for (i = 0; i < p->n; i++) x[i*xstride] = r[p->in[i]];
}

/* Columns first, then rows: reference src/dct.c:137-142, 335-340, 774-780,
   2028-2034. */
void oc_fdct2d(int bs, oc_coeff *y, int ystride, const oc_coeff *x, int xstride) {
  oc_coeff z[OC_BSIZE_MAX*OC_BSIZE_MAX];
  int n = 4 << bs;
  int i;
  for (i = 0; i < n; i++) oc_fdct1d(bs, z + n*i, x + i, xstride);
  for (i = 0; i < n; i++) oc_fdct1d(bs, y + ystride*i, z + i, n);
}

/* Rows first, then columns: reference src/dct.c:144-149, 342-347, 782-788,
   2036-2042. */
void oc_idct2d(int bs, oc_coeff *x, int xstride, const oc_coeff *y, int ystride) {
  oc_coeff z[OC_BSIZE_MAX*OC_BSIZE_MAX];
  int n = 4 << bs;
  int i;
  for (i = 0; i < n; i++) oc_idct1d(bs, z + i, n, y + ystride*i);
  for (i = 0; i < n; i++) oc_idct1d(bs, x + i, xstride, z + n*i);
}

/* Batched forms for the self-test streams (src/dct.c:3439-3469, 3565-3598): nblocks
   tightly packed n x n blocks / nvec tightly packed vectors. */
void oc_dct2d_blocks(int bs, int inverse, oc_coeff *out, const oc_coeff *in, int nblocks) {
  int n = 4 << bs;
  int k;
  for (k = 0; k < nblocks; k++) {
    if (inverse) oc_idct2d(bs, out + (size_t)k*n*n, n, in + (size_t)k*n*n, n);
    else oc_fdct2d(bs, out + (size_t)k*n*n, n, in + (size_t)k*n*n, n);
  }
}

void oc_dct1d_vectors(int bs, int inverse, oc_coeff *out, const oc_coeff *in, int nvec) {
  int n = 4 << bs;
  int k;
  for (k = 0; k < nvec; k++) {
    if (inverse) oc_idct1d(bs, out + (size_t)k*n, 1, in + (size_t)k*n);
    else oc_fdct1d(bs, out + (size_t)k*n, in + (size_t)k*n, 1);
  }
}

/* ------------------------------------------------------ lapping filters -- */

/* Parameter tables: n/2 scale factors s[], then n/2-1 "p" and n/2-1 "u"
   rotation parameters, all in 1/64ths (reference src/filter.c:164-167,
   285-294, 485-507, 812-858; the TYPE3 variants are the ones compiled in). */
static const int OC_FP4[4] = {85, 75, -15, 33};
static const int OC_FP8[10] = {93, 72, 73, 78, -28, -23, -10, 50, 37, 23};
static const int OC_FP16[22] = {
  94, 71, 68, 68, 68, 69, 70, 73,
  -32, -37, -36, -32, -26, -17, -7,
  56, 49, 45, 40, 34, 26, 15
};
static const int OC_FP32[46] = {
  91, 70, 68, 67, 67, 67, 67, 66, 66, 67, 67, 66, 67, 67, 67, 70,
  -32, -41, -42, -41, -40, -38, -36, -34, -32, -29, -24, -19, -14, -9, -5,
  58, 52, 50, 48, 45, 43, 40, 38, 35, 32, 29, 24, 18, 13, 8
};
static const int *const OC_FP[4] = {OC_FP4, OC_FP8, OC_FP16, OC_FP32};

/* reference src/filter.c:174-220 (4), 306-391 (8), 546-703 (16), 879-1171 (32) */
void oc_pre_filter(int f, oc_coeff *y, const oc_coeff *x) {
  int n = 4 << f;
  int h = n >> 1;
  const int *s = OC_FP[f];
  const int *p = s + h;
  const int *u = p + h - 1;
  int t[32];
  int i;
  for (i = 0; i < h; i++) t[n - 1 - i] = x[i] - x[n - 1 - i];
  for (i = 0; i < h; i++) t[i] = x[i] - (t[n - 1 - i] >> 1);
  for (i = 0; i < h; i++) {
    if (s[i] != 64) {
      int v = (t[h + i]*s[i]) >> 6;
      v += (-v >> 31) & 1; /* +1 when v > 0 */
      t[h + i] = v;
    }
  }
  for (i = n - 1; i > h; i--) {
    t[i] += (t[i - 1]*p[i - 1 - h] + 32) >> 6;
    t[i - 1] += (t[i]*u[i - 1 - h] + 32) >> 6;
  }
  for (i = 0; i < h; i++) t[i] += t[n - 1 - i] >> 1;
  for (i = 0; i < h; i++) {
    y[i] = t[i];
    y[n - 1 - i] = t[i] - t[n - 1 - i];
  }
}

/* reference src/filter.c:222-249 (4), 393-449 (8), 705-809 (16), 1173-1373 (32) */
void oc_post_filter(int f, oc_coeff *x, const oc_coeff *y) {
  int n = 4 << f;
  int h = n >> 1;
  const int *s = OC_FP[f];
  const int *p = s + h;
  const int *u = p + h - 1;
  int t[32];
  int i;
  for (i = 0; i < h; i++) t[n - 1 - i] = y[i] - y[n - 1 - i];
  for (i = 0; i < h; i++) t[i] = y[i] - (t[n - 1 - i] >> 1);
  for (i = h + 1; i < n; i++) {
    t[i - 1] -= (t[i]*u[i - 1 - h] + 32) >> 6;
    t[i] -= (t[i - 1]*p[i - 1 - h] + 32) >> 6;
  }
  for (i = 0; i < h; i++) {
    /* C division: truncation toward zero. */
    if (s[i] != 64) t[h + i] = (t[h + i]*64)/s[i];
  }
  for (i = 0; i < h; i++) t[i] += t[n - 1 - i] >> 1;
  for (i = 0; i < h; i++) {
    x[i] = t[i];
    x[n - 1 - i] = t[i] - t[n - 1 - i];
  }
}

/* ------------------------------------------------------------------ Haar -- */

/* The reversible 2x2 Haar butterfly, reference src/tf.h:34-45. */
static void oc_haar_kernel(oc_coeff *ll, oc_coeff *lh, oc_coeff *hl, oc_coeff *hh) {
  oc_coeff m;
  *ll += *hl;
  *hh -= *lh;
  m = (*ll - *hh) >> 1;
  *lh = m - *lh;
  *hl = m - *hl;
  *ll -= *lh;
  *hh += *hl;
}

/* reference src/dct.c:1960-1997 */
void oc_haar(oc_coeff *y, int ystride, const oc_coeff *x, int xstride, int ln) {
  oc_coeff tmp[OC_BSIZE_MAX*OC_BSIZE_MAX];
  int n = 1 << ln;
  int level;
  int i;
  int j;
  for (i = 0; i < n; i++) {
    for (j = 0; j < n; j++) tmp[i*n + j] = x[i*xstride + j];
  }
  for (level = 0; level < ln; level++) {
    int np = n >> level >> 1;
    for (i = 0; i < np; i++) {
      for (j = 0; j < np; j++) {
        oc_coeff a = tmp[2*i*n + 2*j];
        oc_coeff b = tmp[(2*i + 1)*n + 2*j];
        oc_coeff c = tmp[2*i*n + 2*j + 1];
        oc_coeff d = tmp[(2*i + 1)*n + 2*j + 1];
        oc_haar_kernel(&a, &b, &c, &d);
        tmp[i*n + j] = a;
        y[i*ystride + j + np] = b;
        y[(i + np)*ystride + j] = c;
        y[(i + np)*ystride + j + np] = d;
      }
    }
  }
  y[0] = tmp[0];
}

/* reference src/dct.c:1999-2026 */
void oc_haar_inv(oc_coeff *x, int xstride, const oc_coeff *y, int ystride, int ln) {
  int level;
  int i;
  int j;
  x[0] = y[0];
  for (level = ln - 1; level >= 0; level--) {
    int np = 1 << (ln - 1 - level);
    for (i = np - 1; i >= 0; i--) {
      for (j = np - 1; j >= 0; j--) {
        oc_coeff a = x[i*xstride + j];
        oc_coeff b = y[i*ystride + j + np];
        oc_coeff c = y[(i + np)*ystride + j];
        oc_coeff d = y[(i + np)*ystride + j + np];
        oc_haar_kernel(&a, &b, &c, &d);
        x[2*i*xstride + 2*j] = a;
        x[(2*i + 1)*xstride + 2*j] = b;
        x[2*i*xstride + 2*j + 1] = c;
        x[(2*i + 1)*xstride + 2*j + 1] = d;
      }
    }
  }
}

/* ------------------------------------------------- time-frequency (tf.c) -- */

/* reference src/tf.c:38-56 */
void oc_tf_up_h_lp(oc_coeff *dst, int dstride, const oc_coeff *src, int sstride,
 int dx, int n) {
  int x;
  int y;
  for (y = 0; y < n; y++) {
    for (x = 0; x < n >> 1; x++) {
      oc_coeff ll = src[y*sstride + x];
      oc_coeff lh = ll - src[y*sstride + x + dx];
      int hs = x & 1;
      ll -= oc_half(lh);
      dst[y*dstride + 2*x + hs] = ll;
      dst[y*dstride + 2*x + 1 - hs] = lh;
    }
  }
}

/* reference src/tf.c:60-78 */
void oc_tf_up_v_lp(oc_coeff *dst, int dstride, const oc_coeff *src, int sstride,
 int dy, int n) {
  int x;
  int y;
  for (y = 0; y < n >> 1; y++) {
    int vs = y & 1;
    for (x = 0; x < n; x++) {
      oc_coeff ll = src[y*sstride + x];
      oc_coeff hl = ll - src[(y + dy)*sstride + x];
      ll -= oc_half(hl);
      dst[(2*y + vs)*dstride + x] = ll;
      dst[(2*y + 1 - vs)*dstride + x] = hl;
    }
  }
}

static void oc_tf_store4(oc_coeff *dst, int dstride, int x, int y,
 oc_coeff ll, oc_coeff lh, oc_coeff hl, oc_coeff hh) {
  int hs = x & 1;
  int vs = y & 1;
  dst[(2*y + vs)*dstride + 2*x + hs] = ll;
  dst[(2*y + vs)*dstride + 2*x + 1 - hs] = lh;
  dst[(2*y + 1 - vs)*dstride + 2*x + hs] = hl;
  dst[(2*y + 1 - vs)*dstride + 2*x + 1 - hs] = hh;
}

/* reference src/tf.c:82-108 */
void oc_tf_up_hv_lp(oc_coeff *dst, int dstride, const oc_coeff *src, int sstride,
 int dx, int dy, int n) {
  int x;
  int y;
  for (y = 0; y < n >> 1; y++) {
    for (x = 0; x < n >> 1; x++) {
      oc_coeff ll = src[y*sstride + x];
      oc_coeff lh = src[y*sstride + x + dx];
      oc_coeff hl = src[(y + dy)*sstride + x];
      oc_coeff hh = src[(y + dy)*sstride + x + dx];
      oc_haar_kernel(&ll, &hl, &lh, &hh);
      oc_tf_store4(dst, dstride, x, y, ll, lh, hl, hh);
    }
  }
}

/* reference src/tf.c:112-138 */
void oc_tf_up_hv(oc_coeff *dst, int dstride, const oc_coeff *src, int sstride, int n) {
  int x;
  int y;
  for (y = 0; y < n; y++) {
    for (x = 0; x < n; x++) {
      oc_coeff ll = src[y*sstride + x];
      oc_coeff lh = src[y*sstride + x + n];
      oc_coeff hl = src[(y + n)*sstride + x];
      oc_coeff hh = src[(y + n)*sstride + x + n];
      oc_haar_kernel(&ll, &hl, &lh, &hh);
      oc_tf_store4(dst, dstride, x, y, ll, lh, hl, hh);
    }
  }
}

/* reference src/tf.c:142-169 */
void oc_tf_down_hv(oc_coeff *dst, int dstride, const oc_coeff *src, int sstride, int n) {
  int x;
  int y;
  n >>= 1;
  for (y = 0; y < n; y++) {
    int vs = y & 1;
    for (x = 0; x < n; x++) {
      int hs = x & 1;
      oc_coeff ll = src[(2*y + vs)*sstride + 2*x + hs];
      oc_coeff lh = src[(2*y + vs)*sstride + 2*x + 1 - hs];
      oc_coeff hl = src[(2*y + 1 - vs)*sstride + 2*x + hs];
      oc_coeff hh = src[(2*y + 1 - vs)*sstride + 2*x + 1 - hs];
      oc_haar_kernel(&ll, &lh, &hl, &hh);
      dst[y*dstride + x] = ll;
      dst[y*dstride + x + n] = lh;
      dst[(y + n)*dstride + x] = hl;
      dst[(y + n)*dstride + x + n] = hh;
    }
  }
}

/* reference src/tf.c:171-199 */
static void oc_tf_filter(oc_coeff *v, int stride, int n) {
  int m = (n >> 1) - 1;
  int i;
  v += stride;
  for (i = 0; i < m; i++) {
    oc_coeff *u = v;
    v += 2*stride;
    *u += *v >> 1;
    *v -= *u >> 1;
  }
}

static void oc_tf_filter_inv(oc_coeff *base, int stride, int n) {
  oc_coeff *u = base + stride*(n - 1);
  int m = (n >> 1) - 1;
  int i;
  for (i = 0; i < m; i++) {
    oc_coeff *v = u;
    u -= 2*stride;
    *v += *u >> 1;
    *u -= *v >> 1;
  }
}

static void oc_copy_block(oc_coeff *dst, int dstride, const oc_coeff *src,
 int sstride, int n) {
  int i;
  int j;
  if (dst == src) return;
  for (j = 0; j < n; j++) {
    for (i = 0; i < n; i++) dst[dstride*j + i] = src[sstride*j + i];
  }
}

/* reference src/tf.c:201-217 */
void oc_tf_filter_2d(oc_coeff *dst, int dstride, const oc_coeff *src, int sstride, int n) {
  int i;
  oc_copy_block(dst, dstride, src, sstride, n);
  for (i = 0; i < n; i++) oc_tf_filter(dst + dstride*i, 1, n);
  for (i = 0; i < n; i++) oc_tf_filter(dst + i, dstride, n);
}

/* reference src/tf.c:219-235 */
void oc_tf_filter_inv_2d(oc_coeff *dst, int dstride, const oc_coeff *src, int sstride, int n) {
  int i;
  oc_copy_block(dst, dstride, src, sstride, n);
  for (i = 0; i < n; i++) oc_tf_filter_inv(dst + i, dstride, n);
  for (i = 0; i < n; i++) oc_tf_filter_inv(dst + dstride*i, 1, n);
}

/* reference src/tf.c:237-277 */
void oc_convert_block_down(oc_coeff *dst, int dstride, const oc_coeff *src, int sstride,
 int curr_size, int dest_size, int filter) {
  oc_coeff a[OC_BSIZE_MAX*OC_BSIZE_MAX];
  int n = 4 << curr_size;
  if (curr_size == dest_size) {
    oc_copy_block(dst, dstride, src, sstride, n);
    return;
  }
  oc_copy_block(a, OC_BSIZE_MAX, src, sstride, n);
  if (filter) oc_tf_filter_inv_2d(a, OC_BSIZE_MAX, a, OC_BSIZE_MAX, n);
  if (curr_size - 1 == dest_size) {
    oc_tf_down_hv(dst, dstride, a, OC_BSIZE_MAX, n);
  }
  else {
    oc_coeff b[OC_BSIZE_MAX*OC_BSIZE_MAX];
    int i;
    int j;
    oc_tf_down_hv(b, OC_BSIZE_MAX, a, OC_BSIZE_MAX, n);
    n >>= 1;
    for (j = 0; j < 2; j++) {
      for (i = 0; i < 2; i++) {
        oc_convert_block_down(dst + dstride*n*j + n*i, dstride,
         b + OC_BSIZE_MAX*n*j + n*i, OC_BSIZE_MAX, curr_size - 1, dest_size,
         filter);
      }
    }
  }
}

/* -------------------------------------------------------- lapping drivers -- */

static void oc_filter_cols(oc_coeff *c, int stride, int ncols, int f, int inv) {
  /* Filters down each of ncols columns, 4<<f samples starting at c. */
  oc_coeff t[32];
  int n = 4 << f;
  int j;
  int k;
  for (j = 0; j < ncols; j++) {
    for (k = 0; k < n; k++) t[k] = c[stride*k + j];
    if (inv) oc_post_filter(f, t, t);
    else oc_pre_filter(f, t, t);
    for (k = 0; k < n; k++) c[stride*k + j] = t[k];
  }
}

static void oc_filter_rows(oc_coeff *c, int stride, int nrows, int f, int inv) {
  int i;
  for (i = 0; i < nrows; i++) {
    if (inv) oc_post_filter(f, c + i*stride, c + i*stride);
    else oc_pre_filter(f, c + i*stride, c + i*stride);
  }
}

/* reference src/filter.c:1486-1510.  The flag named hfilter gates the filter
   ACROSS THE HORIZONTAL mid-edge (it runs down the columns); vfilter gates
   the one across the vertical mid-edge. */
void oc_prefilter_split(oc_coeff *c0, int stride, int bs, int f, int hfilter, int vfilter) {
  if (hfilter) oc_filter_cols(c0 + ((2 << bs) - (2 << f))*stride, stride, 4 << bs, f, 0);
  if (vfilter) oc_filter_rows(c0 + (2 << bs) - (2 << f), stride, 4 << bs, f, 0);
}

/* reference src/filter.c:1512-1554 (the OD_DEBLOCKING == 0 branch) */
void oc_postfilter_split(oc_coeff *c0, int stride, int bs, int f, int hfilter, int vfilter) {
  if (vfilter) oc_filter_rows(c0 + (2 << bs) - (2 << f), stride, 4 << bs, f, 1);
  if (hfilter) oc_filter_cols(c0 + ((2 << bs) - (2 << f))*stride, stride, 4 << bs, f, 1);
}

/* reference src/filter.c:1556-1586; f is always 0 (src/filter.h:99) */
void oc_apply_prefilter_frame_sbs(oc_coeff *c0, int stride, int nhsb, int nvsb,
 int xdec, int ydec) {
  int sbw = OC_BSIZE_MAX >> xdec;
  int sbh = OC_BSIZE_MAX >> ydec;
  int sbx;
  int sby;
  for (sby = 1; sby < nvsb; sby++) {
    oc_filter_cols(c0 + (sby*sbh - 2)*stride, stride, nhsb*sbw, 0, 0);
  }
  for (sbx = 1; sbx < nhsb; sbx++) {
    oc_filter_rows(c0 + sbx*sbw - 2, stride, nvsb*sbh, 0, 0);
  }
}

/* reference src/filter.c:1588-1646 (the OD_DEBLOCKING == 0 branch) */
void oc_apply_postfilter_frame_sbs(oc_coeff *c0, int stride, int nhsb, int nvsb,
 int xdec, int ydec) {
  int sbw = OC_BSIZE_MAX >> xdec;
  int sbh = OC_BSIZE_MAX >> ydec;
  int sbx;
  int sby;
  for (sbx = 1; sbx < nhsb; sbx++) {
    oc_filter_rows(c0 + sbx*sbw - 2, stride, nvsb*sbh, 0, 1);
  }
  for (sby = 1; sby < nvsb; sby++) {
    oc_filter_cols(c0 + (sby*sbh - 2)*stride, stride, nhsb*sbw, 0, 1);
  }
}

/* ---------------------------------------------------- whole-plane stages -- */

typedef struct {
  oc_coeff *c;
  oc_coeff *d;
  int w;
  int dec;
  const unsigned char *bsize;
  int bstride;
  int pic_width;
  int pic_height;
  int dc;
  int use_haar;
} oc_walk;

static int oc_bsize_at(const oc_walk *k, int bx4, int by4) {
  /* bx4, by4 in 4x4 units: reference OD_BLOCK_SIZE4x4, src/block_size.h:32-35 */
  return k->bsize[(by4 >> 1)*k->bstride + (bx4 >> 1)];
}

/* reference src/encode.c:1286-1343 (od_compute_dcts) */
static void oc_fwd_recursive(const oc_walk *k, int bx, int by, int bsi) {
  int obs = oc_bsize_at(k, bx << bsi, by << bsi);
  int bs = obs > k->dec ? obs : k->dec;
  int w = k->w;
  if (bs == bsi) {
    int bo;
    bs -= k->dec;
    bo = (by << (2 + bs))*w + (bx << (2 + bs));
    if (k->use_haar) oc_haar(k->d + bo, w, k->c + bo, w, bs + 2);
    else oc_fdct2d(bs, k->d + bo, w, k->c + bo, w);
  }
  else {
    int hfilter;
    int vfilter;
    int bo;
    bs = bsi - k->dec;
    bo = (by << (2 + bs))*w + (bx << (2 + bs));
    hfilter = ((bx + 1) << (2 + bs)) <= k->pic_width;
    vfilter = ((by + 1) << (2 + bs)) <= k->pic_height;
    oc_prefilter_split(k->c + bo, w, bs, 0, hfilter, vfilter);
    bsi--;
    bx <<= 1;
    by <<= 1;
    oc_fwd_recursive(k, bx + 0, by + 0, bsi);
    oc_fwd_recursive(k, bx + 1, by + 0, bsi);
    oc_fwd_recursive(k, bx + 0, by + 1, bsi);
    oc_fwd_recursive(k, bx + 1, by + 1, bsi);
    if (k->dc) {
      int ln = bsi - k->dec + 2;
      oc_coeff *d = k->d;
      oc_coeff *x0 = d + (by << ln)*w + (bx << ln);
      oc_coeff *x1 = d + (by << ln)*w + ((bx + 1) << ln);
      oc_coeff *x2 = d + ((by + 1) << ln)*w + (bx << ln);
      oc_coeff *x3 = d + ((by + 1) << ln)*w + ((bx + 1) << ln);
      oc_haar_kernel(x0, x2, x1, x3);
    }
  }
}

/* Transform slices of reference src/decode.c:767-868 (od_decode_recursive),
   :722-738 (DC split) and :632-638 (leaf inverse). */
static void oc_inv_recursive(const oc_walk *k, int bx, int by, int bsi) {
  int obs = oc_bsize_at(k, bx << bsi, by << bsi);
  int bs = obs > k->dec ? obs : k->dec;
  int w = k->w;
  if (bs == bsi) {
    int bo;
    bs -= k->dec;
    bo = (by << (2 + bs))*w + (bx << (2 + bs));
    if (k->use_haar) oc_haar_inv(k->c + bo, w, k->d + bo, w, bs + 2);
    else oc_idct2d(bs, k->c + bo, w, k->d + bo, w);
  }
  else {
    int hfilter;
    int vfilter;
    int bo;
    bs = bsi - k->dec;
    bo = (by << (2 + bs))*w + (bx << (2 + bs));
    hfilter = ((bx + 1) << (2 + bs)) <= k->pic_width;
    vfilter = ((by + 1) << (2 + bs)) <= k->pic_height;
    if (k->dc) {
      int ln = bsi - 1 - k->dec + 2;
      int cx = 2*bx;
      int cy = 2*by;
      oc_coeff *d = k->d;
      oc_coeff *x0 = d + (cy << ln)*w + (cx << ln);
      oc_coeff *x1 = d + (cy << ln)*w + ((cx + 1) << ln);
      oc_coeff *x2 = d + ((cy + 1) << ln)*w + (cx << ln);
      oc_coeff *x3 = d + ((cy + 1) << ln)*w + ((cx + 1) << ln);
      oc_haar_kernel(x0, x1, x2, x3);
    }
    oc_inv_recursive(k, 2*bx + 0, 2*by + 0, bsi - 1);
    oc_inv_recursive(k, 2*bx + 1, 2*by + 0, bsi - 1);
    oc_inv_recursive(k, 2*bx + 0, 2*by + 1, bsi - 1);
    oc_inv_recursive(k, 2*bx + 1, 2*by + 1, bsi - 1);
    oc_postfilter_split(k->c + bo, w, bs, 0, hfilter, vfilter);
  }
}

/* reference src/encode.c:2425-2429 + :2460-2492 (forward stage of a plane) */
void oc_tdlt_fwd_plane(oc_coeff *d, oc_coeff *c, int nhsb, int nvsb, int dec,
 const unsigned char *bsize, int bstride, int pic_width, int pic_height,
 int is_keyframe, int use_haar) {
  oc_walk k;
  int sbx;
  int sby;
  k.c = c; k.d = d; k.w = (nhsb*OC_BSIZE_MAX) >> dec; k.dec = dec;
  k.bsize = bsize; k.bstride = bstride;
  k.pic_width = pic_width; k.pic_height = pic_height;
  k.dc = is_keyframe; k.use_haar = use_haar;
  if (!use_haar) oc_apply_prefilter_frame_sbs(c, k.w, nhsb, nvsb, dec, dec);
  for (sby = 0; sby < nvsb; sby++) {
    for (sbx = 0; sbx < nhsb; sbx++) oc_fwd_recursive(&k, sbx, sby, 3);
  }
}

/* reference src/decode.c:1010-1041 (inverse stage of a plane); d is updated
   in place by the DC split, as in the decoder. */
void oc_tdlt_inv_plane(oc_coeff *c, oc_coeff *d, int nhsb, int nvsb, int dec,
 const unsigned char *bsize, int bstride, int pic_width, int pic_height,
 int dc_split, int use_haar) {
  oc_walk k;
  int sbx;
  int sby;
  k.c = c; k.d = d; k.w = (nhsb*OC_BSIZE_MAX) >> dec; k.dec = dec;
  k.bsize = bsize; k.bstride = bstride;
  k.pic_width = pic_width; k.pic_height = pic_height;
  k.dc = dc_split; k.use_haar = use_haar;
  for (sby = 0; sby < nvsb; sby++) {
    for (sbx = 0; sbx < nhsb; sbx++) oc_inv_recursive(&k, sbx, sby, 3);
  }
  if (!use_haar) oc_apply_postfilter_frame_sbs(c, k.w, nhsb, nvsb, dec, dec);
}

/* ---------------------------------------------- chroma-from-luma predictor -- */

/* reference src/intra.c:65-70 (a table of trained constants; the only data of this step) */
static const int OC_CFL_SCALING4[4][4] = {
  {128, 128, 100, 36},
  {128, 80, 71, 35},
  {100, 71, 35, 31},
  {36, 35, 31, 18},
};

/* reference src/intra.c:72-109.  bs: chroma transform size (n = 4 << bs samples);
   chroma_bs: the LUMA block size at this position (the call sites pass obs,
   src/encode.c:1514-1516, src/decode.c:818-820).  4:2:0 and 4:4:4 only, as the
   callers assert (src/encode.c:1499). */
void oc_resample_luma_coeffs(oc_coeff *chroma_pred, int cpstride, const oc_coeff *decoded_luma,
 int dlstride, int xdec, int ydec, int bs, int chroma_bs) {
  int n = 4 << bs;
  int x;
  int y;
  if (chroma_bs == 0 && xdec && ydec) {
    /* four 4x4 luma transforms merged into one 4x4: their low 2x2 corners */
    oc_tf_up_hv_lp(chroma_pred, cpstride, decoded_luma, dlstride, n, n, n);
    for (y = 0; y < 4; y++) {
      for (x = 0; x < 4; x++) {
        chroma_pred[y*cpstride + x] = (OC_CFL_SCALING4[x][y]*chroma_pred[y*cpstride + x] + 64) >> 7;
      }
    }
    return;
  }
  if (chroma_bs == 0 && xdec) {
    oc_tf_up_h_lp(chroma_pred, cpstride, decoded_luma, dlstride, n, n);
    return;
  }
  if (chroma_bs == 0 && ydec) {
    oc_tf_up_v_lp(chroma_pred, cpstride, decoded_luma, dlstride, n, n);
    return;
  }
  /* the luma transform is at least as large: its upper-left n x n coefficients */
  for (y = 0; y < n; y++) {
    for (x = 0; x < n; x++) chroma_pred[y*cpstride + x] = decoded_luma[y*dlstride + x];
  }
}

/* The predictor of every 4:2:0 chroma leaf block of one frame, each written at the
   block's own position of a half-resolution plane `l` (stride nhsb*16): what the
   recursion of src/encode.c:1491-1517 / src/decode.c:780-821 hands od_resample_luma_coeffs,
   leaf by leaf, with ctx->l repointed into a plane. */
void oc_cfl_pred_plane(oc_coeff *l, const oc_coeff *d_luma, int nhsb, int nvsb,
 const unsigned char *bsize, int bstride) {
  int fw = nhsb*32;
  int cw = fw >> 1;
  int ux;
  int uy;
  for (uy = 0; uy < nvsb*4; uy++) {
    for (ux = 0; ux < nhsb*4; ux++) {
      int obs = bsize[uy*bstride + ux];
      int bsi = obs > 1 ? obs : 1;
      int m = 1 << (bsi - 1);
      if ((ux & (m - 1)) || (uy & (m - 1))) continue;   /* not the block's first unit */
      oc_resample_luma_coeffs(l + (uy*4)*cw + ux*4, cw, d_luma + (uy*8)*fw + ux*8, fw,
       1, 1, bsi - 1, obs);
    }
  }
}

/* ------------------------------------------------- pixel <-> coefficient -- */

/* reference src/state.c:1209-1248.  bitdepth_mode: 1/2/3 for 8/10/12 bits
   (include/daala/codec.h:196-200). */
void oc_buf_to_coeff(int bitdepth_mode, oc_coeff *dst, int dst_ystride, int lossless_p,
 const unsigned char *src, int src_xstride, int src_ystride, int w, int h) {
  int x;
  int y;
  if (src_xstride == 1) {
    int shift = lossless_p ? (bitdepth_mode - 1)*2 : OC_COEFF_SHIFT;
    for (y = 0; y < h; y++) {
      for (x = 0; x < w; x++) dst[x] = (src[x] - 128)*(1 << shift);
      dst += dst_ystride;
      src += src_ystride;
    }
  }
  else {
    int shift = lossless_p ? OC_COEFF_SHIFT - (bitdepth_mode - 1)*2 : 0;
    for (y = 0; y < h; y++) {
      const int16_t *s = (const int16_t *)src;
      for (x = 0; x < w; x++) {
        dst[x] = (s[x] - (1 << (8 + OC_COEFF_SHIFT) >> 1) + (1 << shift >> 1)) >> shift;
      }
      dst += dst_ystride;
      src += src_ystride;
    }
  }
}

/* reference src/state.c:1274-1316, clamps src/odintrin.h:104-113 */
void oc_coeff_to_buf(int bitdepth_mode, unsigned char *dst, int dst_xstride, int dst_ystride,
 const oc_coeff *src, int src_ystride, int lossless_p, int w, int h) {
  int x;
  int y;
  if (dst_xstride == 1) {
    int shift = lossless_p ? (bitdepth_mode - 1)*2 : OC_COEFF_SHIFT;
    for (y = 0; y < h; y++) {
      for (x = 0; x < w; x++) {
        int v = ((src[x] + (1 << shift >> 1)) >> shift) + 128;
        dst[x] = (unsigned char)(v < 0 ? 0 : v > 255 ? 255 : v);
      }
      dst += dst_ystride;
      src += src_ystride;
    }
  }
  else {
    int shift = lossless_p ? OC_COEFF_SHIFT - (bitdepth_mode - 1)*2 : 0;
    int top = (1 << (8 + OC_COEFF_SHIFT)) - 1;
    for (y = 0; y < h; y++) {
      int16_t *o = (int16_t *)dst;
      for (x = 0; x < w; x++) {
        int v = src[x]*(1 << shift) + (128 << OC_COEFF_SHIFT);
        o[x] = (int16_t)(v < 0 ? 0 : v > top ? top : v);
      }
      dst += dst_ystride;
      src += src_ystride;
    }
  }
}

/* TEST INFRASTRUCTURE -- CPU restatement of the reference's deringing filter
 * (the step right after the post-filter).
 * See tdlt_oracle.h for the rules on who may use it and how it is pinned:
 * tests/test_oracle_vs_ref.py compares every function here with the compiled,
 * unmodified reference (od_dering and the od_filter_dering_* entry points of its
 * dispatch table, C and SSE twins).
 *
 * Reference: src/filter.c:1648-1950 (direction search, the two filters, thresholds,
 * od_dering) and the frame loop of src/decode.c:1044-1122.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "tdlt_oracle.h"

#define OC_DERING_BORDER 3        /* reference OD_FILT_BORDER, src/filter.h:102 */
#define OC_DERING_SB 32           /* reference OD_BSIZE_MAX: one deringing superblock */
#define OC_DERING_PITCH (OC_DERING_SB + 2*OC_DERING_BORDER)
#define OC_DERING_FAR 30000       /* reference OD_DERING_VERY_LARGE, src/filter.c:1710 */

/* The three taps of each direction as (dy, dx) steps away from the pixel; the same
   steps are taken the other way (reference direction_offsets_table, src/filter.c:142-151:
   0 = 45 degrees up-right, 2 = horizontal, 4 = 45 degrees down-right, 6 = vertical). */
static const signed char OC_DIR_STEPS[8][3][2] = {
  {{-1, 1}, {-2, 2}, {-3, 3}},
  {{0, 1}, {-1, 2}, {-1, 3}},
  {{0, 1}, {0, 2}, {0, 3}},
  {{0, 1}, {1, 2}, {1, 3}},
  {{1, 1}, {2, 2}, {3, 3}},
  {{1, 0}, {2, 1}, {3, 1}},
  {{1, 0}, {2, 0}, {3, 0}},
  {{1, 0}, {2, -1}, {3, -1}},
};

/* Which of the lines of direction `d` pixel (i, j) of an 8x8 block lies on
   (reference src/filter.c:1666-1673). */
static int oc_dir_line(int d, int i, int j) {
  switch (d) {
    case 0: return i + j;
    case 1: return i + j/2;
    case 2: return i;
    case 3: return 3 + i - j/2;
    case 4: return 7 + i - j;
    case 5: return 3 - i/2 + j;
    case 6: return j;
    default: return i/2 + j;
  }
}

/* reference src/filter.c:1655-1707.  The cost of a direction is the sum over its lines
   of (sum of the line)^2 / (pixels on the line); the largest wins.  The reference
   divides with OD_DIVU_SMALL, an exact unsigned division by a small constant. */
int oc_dir_find8(const int16_t *img, int stride, int32_t *var) {
  int cost[8];
  int d;
  int best = 0;
  int best_cost = 0;
  for (d = 0; d < 8; d++) {
    int sum[15];
    int npix[15];
    int i;
    int j;
    memset(sum, 0, sizeof(sum));
    memset(npix, 0, sizeof(npix));
    for (i = 0; i < 8; i++) {
      for (j = 0; j < 8; j++) {
        int k = oc_dir_line(d, i, j);
        sum[k] += img[i*stride + j] >> 4;      /* OD_COEFF_SHIFT */
        npix[k]++;
      }
    }
    cost[d] = 0;
    for (i = 0; i < 15; i++) {
      if (npix[i] == 8) cost[d] += sum[i]*sum[i] >> 3;
      else if (npix[i]) cost[d] += (int)((uint32_t)(sum[i]*sum[i])/(uint32_t)npix[i]);
    }
  }
  for (d = 0; d < 8; d++) {
    if (cost[d] > best_cost) {
      best_cost = cost[d];
      best = d;
    }
  }
  *var = best_cost - cost[(best + 4) & 7];
  return best;
}

/* reference src/filter.c:1715-1740: taps 3, 2, 2 along the direction, both ways, each
   neighbour counted only if it differs from the pixel by less than the threshold.
   `in` has row pitch OC_DERING_PITCH. */
void oc_dering_direction(int16_t *y, int ystride, const int16_t *in, int ln, int threshold,
 int dir) {
  static const int taps[3] = {3, 2, 2};
  int i;
  int j;
  int k;
  for (i = 0; i < 1 << ln; i++) {
    for (j = 0; j < 1 << ln; j++) {
      int xx = in[i*OC_DERING_PITCH + j];
      int sum = 0;
      for (k = 0; k < 3; k++) {
        int o = OC_DIR_STEPS[dir][k][0]*OC_DERING_PITCH + OC_DIR_STEPS[dir][k][1];
        int p0 = in[i*OC_DERING_PITCH + j + o] - xx;
        int p1 = in[i*OC_DERING_PITCH + j - o] - xx;
        if (abs(p0) < threshold) sum += taps[k]*p0;
        if (abs(p1) < threshold) sum += taps[k]*p1;
      }
      y[i*ystride + j] = (int16_t)(xx + ((sum + 8) >> 4));
    }
  }
}

/* reference src/filter.c:1753-1789: two neighbours each way across the direction
   (vertically for dir <= 4, else horizontally), threshold tightened by how much the
   first pass changed the pixel. */
void oc_dering_orthogonal(int16_t *y, int ystride, const int16_t *in, const int16_t *x,
 int xstride, int ln, int threshold, int dir) {
  int o = dir <= 4 ? OC_DERING_PITCH : 1;
  int i;
  int j;
  int k;
  for (i = 0; i < 1 << ln; i++) {
    for (j = 0; j < 1 << ln; j++) {
      int yy = in[i*OC_DERING_PITCH + j];
      int a = threshold/3 + abs(yy - x[i*xstride + j]);
      int sum = 0;
      if (a > threshold) a = threshold;
      for (k = -2; k <= 2; k++) {
        int p;
        if (k == 0) continue;
        p = in[i*OC_DERING_PITCH + j + k*o] - yy;
        if (abs(p) < a) sum += p;
      }
      y[i*ystride + j] = (int16_t)(yy + ((3*sum + 8) >> 4));
    }
  }
}

/* reference src/filter.c:1803-1807; the comment there gives the formula
   round(256*min(3, max(.5, 1.08*(sqrt(2)*2^(i+8)/65536)^.16))), which
   tests/test_oracle_vs_ref.py evaluates against these values. */
const int16_t OC_DERING_THRESH_Q8[18] = {
  128, 134, 150, 168, 188, 210, 234, 262, 292, 327, 365, 408, 455, 509, 569, 635, 710, 768,
};

static int oc_ilog(uint32_t v) {   /* reference od_ilog, src/internal.c:401 */
  int n = 0;
  while (v) {
    n++;
    v >>= 1;
  }
  return n;
}

/* reference src/filter.c:1816-1833 */
int oc_dering_block_thresh(int threshold, int32_t var, int32_t sb_var) {
  int v1 = var >> 6;
  int v2 = sb_var/(OC_DERING_SB*OC_DERING_SB);
  int e;
  if (v1 > 32767) v1 = 32767;
  if (v2 > 32767) v2 = 32767;
  e = oc_ilog((uint32_t)(v1*v2)) - 9;
  if (e < 0) e = 0;
  if (e > 17) e = 17;
  return threshold*OC_DERING_THRESH_Q8[e] >> 8;
}

/* reference src/filter.c:1877: the ringing expected of quantizer q */
int oc_dering_threshold(int q) {
  return (int)(1.0*pow(q, 0.84182));
}

/* reference od_dering, src/filter.c:1835-1941: one deringing superblock of one plane.
   y: n x n output; x: the unfiltered plane at the superblock's first sample; dir: found on
   luma (pli == 0), reused by the chroma planes; bskip: this plane's skip flags (one per 4x4
   samples) at the superblock's first block. */
void oc_dering(int16_t *y, int ystride, const int16_t *x, int xstride, int ln, int sbx, int sby,
 int nhsb, int nvsb, int q, int xdec, int dir[4][4], int pli, const unsigned char *bskip,
 int skip_stride) {
  int16_t buf[OC_DERING_PITCH*OC_DERING_PITCH];
  int16_t *in = buf + OC_DERING_BORDER*OC_DERING_PITCH + OC_DERING_BORDER;
  int n = 1 << ln;
  int lb = 3 - xdec;                      /* log2 of the filter block: 8x8 luma, 4x4 chroma */
  int nb = n >> lb;
  int threshold = oc_dering_threshold(q);
  int thresh[4][4];
  int32_t var[4][4];
  int32_t varsum = 0;
  int i;
  int j;
  int bx;
  int by;
  /* samples outside the frame are "far away": no tap ever accepts them */
  for (i = 0; i < OC_DERING_PITCH*OC_DERING_PITCH; i++) buf[i] = OC_DERING_FAR;
  for (i = sby ? -OC_DERING_BORDER : 0; i < n + (sby != nvsb - 1 ? OC_DERING_BORDER : 0); i++) {
    for (j = sbx ? -OC_DERING_BORDER : 0; j < n + (sbx != nhsb - 1 ? OC_DERING_BORDER : 0); j++) {
      in[i*OC_DERING_PITCH + j] = x[i*xstride + j];
    }
  }
  for (by = 0; by < nb; by++) {
    for (bx = 0; bx < nb; bx++) {
      if (pli == 0) {
        dir[by][bx] = oc_dir_find8(x + 8*by*xstride + 8*bx, xstride, &var[by][bx]);
        varsum += var[by][bx];
      }
    }
  }
  for (by = 0; by < nb; by++) {
    for (bx = 0; bx < nb; bx++) {
      int all_skipped = 1;
      int u = 2 >> xdec;                 /* 4x4 units per filter block side */
      thresh[by][bx] = pli == 0 ? oc_dering_block_thresh(threshold, var[by][bx], varsum) : threshold;
      /* a block whose own units and the ring of units around them (lapping reaches there)
         are all skipped is left alone */
      for (i = sby ? -1 : 0; i < u + (sby != nvsb - 1); i++) {
        for (j = sbx ? -1 : 0; j < u + (sbx != nhsb - 1); j++) {
          if (!bskip[(by*u + i)*skip_stride + bx*u + j]) all_skipped = 0;
        }
      }
      if (all_skipped) thresh[by][bx] = 0;
    }
  }
  for (by = 0; by < nb; by++) {
    for (bx = 0; bx < nb; bx++) {
      oc_dering_direction(y + (by << lb)*ystride + (bx << lb), ystride,
       in + (by << lb)*OC_DERING_PITCH + (bx << lb), lb, thresh[by][bx], dir[by][bx]);
    }
  }
  /* second pass: the superblock's own samples are the first pass's output, the border
     stays unfiltered */
  for (i = 0; i < n; i++) {
    for (j = 0; j < n; j++) in[i*OC_DERING_PITCH + j] = y[i*ystride + j];
  }
  for (by = 0; by < nb; by++) {
    for (bx = 0; bx < nb; bx++) {
      oc_dering_orthogonal(y + (by << lb)*ystride + (bx << lb), ystride,
       in + (by << lb)*OC_DERING_PITCH + (bx << lb), x + (by << lb)*xstride + (bx << lb), xstride,
       lb, thresh[by][bx], dir[by][bx]);
    }
  }
}

/* The frame loop of src/decode.c:1044-1122 (and src/encode.c's twin) for given
   superblock decisions: out = in, then every superblock whose flag is set is filtered,
   plane by plane, always reading the UNFILTERED frame (the reference's etmp copy,
   int16_t) and writing the reconstruction (ctmp).  Planes are nhsb*32 >> dec wide;
   bskip[p]: one flag per 4x4 samples of plane p, row stride skip_stride. */
void oc_dering_frame(oc_coeff *const out[], const oc_coeff *const in[], int nplanes,
 const int dec[], int nhsb, int nvsb, const unsigned char *flags,
 const unsigned char *const bskip[], int skip_stride, const int q[]) {
  int16_t *e[4];
  int p;
  int sbx;
  int sby;
  for (p = 0; p < nplanes; p++) {
    int w = nhsb*OC_DERING_SB >> dec[p];
    int h = nvsb*OC_DERING_SB >> dec[p];
    int i;
    e[p] = (int16_t *)malloc(sizeof(int16_t)*w*h);
    for (i = 0; i < w*h; i++) {
      e[p][i] = (int16_t)in[p][i];
      out[p][i] = in[p][i];
    }
  }
  for (sby = 0; sby < nvsb; sby++) {
    for (sbx = 0; sbx < nhsb; sbx++) {
      int dir[4][4];
      if (!flags[sby*nhsb + sbx]) continue;
      for (p = 0; p < nplanes; p++) {
        int16_t buf[OC_DERING_SB*OC_DERING_SB];
        int w = nhsb*OC_DERING_SB >> dec[p];
        int ln = 5 - dec[p];
        int n = 1 << ln;
        int x;
        int y;
        oc_dering(buf, n, e[p] + (sby << ln)*w + (sbx << ln), w, ln, sbx, sby, nhsb, nvsb, q[p],
         dec[p], dir, p, bskip[p] + (sby << (3 - dec[p]))*skip_stride + (sbx << (3 - dec[p])),
         skip_stride);
        for (y = 0; y < n; y++) {
          for (x = 0; x < n; x++) out[p][((sby << ln) + y)*w + (sbx << ln) + x] = buf[y*n + x];
        }
      }
    }
  }
  for (p = 0; p < nplanes; p++) free(e[p]);
}

/* ---------------------------------------------------------- bilinear smoothing -- */

/* reference OD_BILINEAR_STRENGTH, src/filter.c:154 */
static const int OC_BILINEAR_STRENGTH[4] = {5, 20, 20, 5};

/* reference od_bilinear_smooth, src/filter.c:1952-2008: the block is pulled towards the
   bilinear interpolation of its four corners, by a Wiener-style weight that falls with the
   distortion this would cause relative to the quantization noise.  Plain int arithmetic
   as in the reference (including a11's correction taking the already corrected a10). */
void oc_bilinear_smooth(oc_coeff *x, int ln, int stride, int q, int pli) {
  int n = 1 << ln;
  int a00 = x[0];
  int a01 = x[n - 1] - x[0];
  int a10 = x[(n - 1)*stride] - x[0];
  int a11 = x[(n - 1)*stride + n - 1] + x[0] - x[(n - 1)*stride] - x[n - 1];
  int shift = 2*4 + 2*ln - 16 > 0 ? 2*4 + 2*ln - 16 : 0;   /* OD_COEFF_SHIFT = 4 */
  int32_t dist = 0;
  int w;
  int i;
  int j;
  oc_coeff y[32][32];
  /* times 1 + 1/n, about n/(n - 1): the loop divides by n */
  a01 += (a01 + n/2) >> ln;
  a10 += (a10 + n/2) >> ln;
  a11 += (2*a10 + n/2) >> ln;
  for (i = 0; i < n; i++) {
    int32_t row = 0;
    for (j = 0; j < n; j++) {
      int d;
      y[i][j] = a00 + ((j*a01 + i*a10 + (j*i*a11 >> ln) + n/2) >> ln);
      d = y[i][j] - x[i*stride + j];
      row += d*d;
    }
    dist += row >> shift;
  }
  dist += n/2;
  dist >>= 2*ln - shift;
  w = OC_BILINEAR_STRENGTH[pli]*q*q/(1 + 12*dist);
  if (w > 1024) w = 1024;
  w = w*w >> 12;
  for (i = 0; i < n; i++) {
    for (j = 0; j < n; j++) x[i*stride + j] -= (w*(x[i*stride + j] - y[i][j]) + 128) >> 8;
  }
}

/* od_smooth_recursive with min_bs = 32x32 over every superblock, as the decoder calls it on
   keyframes (src/decode.c:1134-1146, src/filter.c:2010-2039): only superblocks coded as one
   32x32 block are smoothed, in every plane (4:2:0 chroma: the 16x16 block). */
void oc_smooth_frame(oc_coeff *const c[], int nplanes, const int dec[], int nhsb, int nvsb,
 const unsigned char *bsize, int bstride, const int q[]) {
  int p;
  int sbx;
  int sby;
  for (p = 0; p < nplanes; p++) {
    int w = nhsb*32 >> dec[p];
    int ln = 5 - dec[p];
    for (sby = 0; sby < nvsb; sby++) {
      for (sbx = 0; sbx < nhsb; sbx++) {
        if (bsize[4*sby*bstride + 4*sbx] != 3) continue;
        oc_bilinear_smooth(c[p] + (sby << ln)*w + (sbx << ln), ln, w, q[p], p);
      }
    }
  }
}

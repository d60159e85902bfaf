/* TEST INFRASTRUCTURE -- CPU restatement of the reference's raster <-> coding-order
 * permutation of a transform block.  See tdlt_oracle.h for who may
 * use it.  Reference: od_raster_to_coding_order / od_coding_order_to_raster and the band
 * copies under them, src/partition.c:95-200.
 *
 * The permutation is driven by four scan tables (the reference's OD_ZIGZAG4/8/16/32,
 * generated from measured coefficient statistics): they are DATA of the reference and are
 * passed in by the caller, never stored here.  Table k lists, as (x, y) byte pairs, the
 * coefficients of the (4 << k)-point block that the (2 << k)-point block does not contain
 * (15, 48, 192 and 768 of them; DC is handled apart), and scan_len[k] says how many of
 * them the codec actually orders: the reference's OD_LAYOUT32 stops after 256 (two bands
 * of 128, src/partition.c:43-49), so a 32x32 block has 512 coded coefficients; the others
 * are neither read nor written, here as there.
 */
#include "tdlt_oracle.h"

/* reference src/partition.c:143-163: DC first, then the 4x4 scan (without DC) at 1, the
   8x8 scan of what 4x4 left out at 16, 16x16 at 64, 32x32 at 256 */
void oc_raster_to_coding_order(oc_coeff *dst, int n, const oc_coeff *src, int stride,
 const unsigned char *const scan[4], const int scan_len[4]) {
  int k;
  int i;
  for (k = 0; k < 4 && n >= 4 << k; k++) {
    oc_coeff *out = dst + (k ? 1 << 2*(k + 1) : 1);
    for (i = 0; i < scan_len[k]; i++) out[i] = src[scan[k][2*i + 1]*stride + scan[k][2*i]];
  }
  dst[0] = src[0];
}

/* reference src/partition.c:177-199 */
void oc_coding_order_to_raster(oc_coeff *dst, int stride, const oc_coeff *src, int n,
 const unsigned char *const scan[4], const int scan_len[4]) {
  int k;
  int i;
  for (k = 0; k < 4 && n >= 4 << k; k++) {
    const oc_coeff *in = src + (k ? 1 << 2*(k + 1) : 1);
    for (i = 0; i < scan_len[k]; i++) dst[scan[k][2*i + 1]*stride + scan[k][2*i]] = in[i];
  }
  dst[0] = src[0];
}

/* Position of a block's vector in the block-linear plane the GPU call produces: superblocks
   in raster order, inside a superblock the 4x4-sample units in Z order (x bits even, y bits
   odd), 16 samples per unit; a quadtree-aligned n x n block owns n*n consecutive entries. */
long long oc_block_linear_offset(int x0, int y0, int w, int sb) {
  int sbx = x0/sb;
  int sby = y0/sb;
  int ux = (x0 - sbx*sb) >> 2;
  int uy = (y0 - sby*sb) >> 2;
  int z = 0;
  int b;
  for (b = 0; b < 3; b++) z |= ((ux >> b) & 1) << (2*b) | ((uy >> b) & 1) << (2*b + 1);
  return ((long long)sby*(w/sb) + sbx)*sb*sb + z*16;
}

/* Every transform block of one plane, as the block walk of src/encode.c:1491-1517 visits
   them (block size from the map; 4:2:0 chroma halves it, not below 4x4): to_coding != 0
   writes each block's vector into the block-linear plane v, else the inverse. */
void oc_coding_order_plane(oc_coeff *v, oc_coeff *d, int nhsb, int nvsb, int dec,
 const unsigned char *bsize, int bstride, const unsigned char *const scan[4],
 const int scan_len[4], int to_coding) {
  int w = nhsb*32 >> dec;
  int sb = 32 >> dec;
  int ux;
  int uy;
  for (uy = 0; uy < nvsb*4; uy++) {
    for (ux = 0; ux < nhsb*4; ux++) {
      int obs = bsize[uy*bstride + ux];
      int lu = obs > 1 ? obs - 1 : 0;             /* log2 of the block side in 8x8 luma units */
      int n;
      int x0;
      int y0;
      int i;
      int j;
      if ((ux & ((1 << lu) - 1)) || (uy & ((1 << lu) - 1))) continue;   /* not the block's first unit */
      n = (4 << obs) >> dec;
      if (n < 4) n = 4;
      /* a unit of 4x4 luma blocks holds four of them (one 4x4 chroma block with 4:2:0) */
      for (j = 0; j < ((8 << lu) >> dec)/n; j++) {
        for (i = 0; i < ((8 << lu) >> dec)/n; i++) {
          x0 = ((ux*8) >> dec) + i*n;
          y0 = ((uy*8) >> dec) + j*n;
          if (to_coding) {
            oc_raster_to_coding_order(v + oc_block_linear_offset(x0, y0, w, sb), n, d + y0*w + x0, w, scan,
             scan_len);
          }
          else {
            oc_coding_order_to_raster(d + y0*w + x0, w, v + oc_block_linear_offset(x0, y0, w, sb), n, scan,
             scan_len);
          }
        }
      }
    }
  }
}

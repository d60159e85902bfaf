/* TEST INFRASTRUCTURE -- CPU restatement of the encoder's input side for the parity tests:
 * od_img_plane_copy with its bit-depth conversion (reference src/state.c:92-217), the low-pass
 * extension of the picture into the padded frame (od_img_plane_copy_pad,
 * src/encode.c:610-700) and the border replication od_img_plane_edge_ext
 * (src/state.c:1100-1159), i.e. what od_img_copy_pad (src/encode.c:1728-1750) does to one
 * plane.  Pinned against the compiled reference by tests/test_oracle_vs_ref.py. */
#include <stdint.h>
#include <stddef.h>
#include <string.h>

static int oc_clampi(int lo, int v, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

static int oc_get(const unsigned char *p, int xstride) {
  return xstride == 1 ? *p : *(const int16_t *)p;
}
static void oc_put(unsigned char *p, int xstride, int v) {
  if (xstride == 1) *p = (unsigned char)v;
  else *(int16_t *)p = (int16_t)v;
}

/* One sample through the conversion of src/state.c:141-212. */
static int oc_convert_sample(int v, int src_bits, int dst_bits) {
  if (src_bits > 8) {
    if (dst_bits > 8) {
      if (dst_bits >= src_bits) return oc_clampi(0, v << (dst_bits - src_bits), (1 << dst_bits) - 1);
      else {
        int dn = src_bits - dst_bits;
        return oc_clampi(0, (v + (1 << dn >> 1)) >> dn, (1 << dst_bits) - 1);
      }
    }
    else {
      int dn = src_bits - 8;
      return oc_clampi(0, (v + (1 << dn >> 1)) >> dn, 255);
    }
  }
  if (dst_bits > 8) return oc_clampi(0, v << (dst_bits - 8), (1 << dst_bits) - 1);
  return v;
}

/* dst -> the first visible sample of a plane buffer with `border` samples of padding all
   round (rows dst_ystride BYTES apart, samples dst_xstride bytes: 1 with 8 bits, 2 above);
   src -> pic_w x pic_h picture samples (src_xstride 1 or 2 bytes, src_bits deep).  The plane
   is plane_w x plane_h (the frame padded to a multiple of 64, decimated). */
void oc_img_copy_pad_plane(unsigned char *dst, int dst_xstride, int dst_ystride, int dst_bits,
 int plane_w, int plane_h, int border, const unsigned char *src, int src_xstride, int src_ystride,
 int src_bits, int pic_w, int pic_h) {
  int x;
  int y;
  if (pic_w == 0 || pic_h == 0) {
    for (y = 0; y < plane_h; y++) memset(dst + (size_t)y*dst_ystride, 0, (size_t)plane_w*dst_xstride);
  }
  else {
    /* step 1: the picture (src/state.c:92-217) */
    for (y = 0; y < pic_h; y++) {
      for (x = 0; x < pic_w; x++) {
        oc_put(dst + (size_t)y*dst_ystride + x*dst_xstride, dst_xstride,
         oc_convert_sample(oc_get(src + (size_t)y*src_ystride + x*src_xstride, src_xstride), src_bits, dst_bits));
      }
    }
    /* step 2: low-pass extension to the right, column after column (src/encode.c:634-664) */
    for (x = pic_w; x < plane_w; x++) {
      for (y = 0; y < pic_h; y++) {
        const unsigned char *c = dst + (size_t)y*dst_ystride + (x - 1)*dst_xstride;
        int vc = oc_get(c, dst_xstride) & 0xffff;
        int vu = oc_get(c - (y > 0 ? dst_ystride : 0), dst_xstride) & 0xffff;
        int vd = oc_get(c + (y + 1 < pic_h ? dst_ystride : 0), dst_xstride) & 0xffff;
        oc_put(dst + (size_t)y*dst_ystride + x*dst_xstride, dst_xstride, (2*vc + vu + vd + 2) >> 2);
      }
    }
    /* ... and downwards, row after row (:665-695) */
    for (y = pic_h; y < plane_h; y++) {
      const unsigned char *up = dst + (size_t)(y - 1)*dst_ystride;
      for (x = 0; x < plane_w; x++) {
        int vc = oc_get(up + x*dst_xstride, dst_xstride) & 0xffff;
        int vl = oc_get(up + (x - (x > 0))*dst_xstride, dst_xstride) & 0xffff;
        int vr = oc_get(up + (x + (x + 1 < plane_w))*dst_xstride, dst_xstride) & 0xffff;
        oc_put(dst + (size_t)y*dst_ystride + x*dst_xstride, dst_xstride, (2*vc + vl + vr + 2) >> 2);
      }
    }
  }
  /* step 3: od_img_plane_edge_ext (src/state.c:1100-1159): left / right, then top / bottom
     over the widened rows = every border sample takes the nearest plane sample */
  for (y = -border; y < plane_h + border; y++) {
    int sy = oc_clampi(0, y, plane_h - 1);
    for (x = -border; x < plane_w + border; x++) {
      int sx;
      if (y >= 0 && y < plane_h && x >= 0 && x < plane_w) continue;
      sx = oc_clampi(0, x, plane_w - 1);
      memcpy(dst + (ptrdiff_t)y*dst_ystride + (ptrdiff_t)x*dst_xstride,
       dst + (ptrdiff_t)sy*dst_ystride + (ptrdiff_t)sx*dst_xstride, dst_xstride);
    }
  }
}

// Lane tasks: how the 32 lanes of a warp share one 32x32 tile.
//
// The tile is 16 units of 8x8 samples, numbered in raster order; lane l owns
// half h = l&1 of unit u = l>>1 (unit row ur = l>>3, unit column uc = (l>>1)&3).
// What "half a unit" means depends on the size of the leaf block that covers
// the unit, so that EVERY lane always holds exactly 32 samples in registers:
//
//   leaf   column pass (vertical 1-D)        row pass (horizontal 1-D)
//   32x32  1 column  x 32 rows (lane = col)   1 row x 32 cols (lane = row)
//   16x16  2 columns x 16 rows                rows k and k+8 of the leaf
//   8x8    4 columns x 8 rows                 rows h, h+2, h+4, h+6 of the unit
//   4x4    4 columns x 8 rows = two stacked 4x4 blocks, both passes in registers
//
// (k = lane's index 0..7 among the 8 lanes of a 16x16 leaf.)  The assignments
// were chosen so that every shared-memory access below is bank-conflict free
// for the transforms; the lapping cells on buffer A see at most 2-way conflicts.
#pragma once
#include "tdlt_device.cuh"

namespace tdlt {

struct Lane {
  int ur, uc, h;     // unit row/col, half
  int k16;           // index within a 16x16 leaf
  int r16, c16;      // origin of the 16x16 quadrant containing the unit
};

__device__ __forceinline__ Lane make_lane(int lane) {
  Lane L;
  L.ur = lane >> 3;
  L.uc = (lane >> 1) & 3;
  L.h = lane & 1;
  L.k16 = ((L.ur & 1) << 2) | ((L.uc & 1) << 1) | L.h;
  L.r16 = (L.ur >> 1)*16;
  L.c16 = (L.uc >> 1)*16;
  return L;
}

// ------------------------------------------------------------------ buffer A
// Column tasks on A (tile sample (r, c) lives at A[a_at(r, c)]).
template <bool kStore>
__device__ __forceinline__ void colA_32(int *A, int lane, int *v) {
  int *p = A + a_at(0, lane);
#pragma unroll
  for (int r = 0; r < 32; r++) {
    if (kStore) p[r*kAStride] = v[r]; else v[r] = p[r*kAStride];
  }
}
template <bool kStore>
__device__ __forceinline__ void colA_16(int *A, const Lane &L, int *v) {
  int2 *p = reinterpret_cast<int2 *>(A + a_at(L.r16, L.c16 + 2*L.k16));
#pragma unroll
  for (int r = 0; r < 16; r++) {
    if (kStore) p[r*(kAStride/2)] = make_int2(v[r], v[16 + r]);
    else { int2 t = p[r*(kAStride/2)]; v[r] = t.x; v[16 + r] = t.y; }
  }
}
// 8 rows x 4 columns, element (r, q) <-> v[8*q + r]  (q-th column contiguous)
template <bool kStore>
__device__ __forceinline__ void colA_8(int *A, const Lane &L, int *v) {
  int4 *p = reinterpret_cast<int4 *>(A + a_at(L.ur*8, L.uc*8 + 4*L.h));
#pragma unroll
  for (int r = 0; r < 8; r++) {
    if (kStore) p[r*(kAStride/4)] = make_int4(v[r], v[8 + r], v[16 + r], v[24 + r]);
    else { int4 t = p[r*(kAStride/4)]; v[r] = t.x; v[8 + r] = t.y; v[16 + r] = t.z; v[24 + r] = t.w; }
  }
}

// ------------------------------------------------------------------ buffer B
template <bool kStore>
__device__ __forceinline__ void colB_32(int *B, int lane, int *v) {
#pragma unroll
  for (int r = 0; r < 32; r++) {
    int o = b_off(r, lane);
    if (kStore) B[o] = v[r]; else v[r] = B[o];
  }
}
template <bool kStore>
__device__ __forceinline__ void colB_16(int *B, const Lane &L, int *v) {
#pragma unroll
  for (int r = 0; r < 16; r++) {
    int2 *p = reinterpret_cast<int2 *>(B + b_off(L.r16 + r, L.c16 + 2*L.k16));
    if (kStore) *p = make_int2(v[r], v[16 + r]);
    else { int2 t = *p; v[r] = t.x; v[16 + r] = t.y; }
  }
}
template <bool kStore>
__device__ __forceinline__ void colB_8(int *B, const Lane &L, int *v) {
#pragma unroll
  for (int r = 0; r < 8; r++) {
    int4 *p = reinterpret_cast<int4 *>(B + b_off(L.ur*8 + r, L.uc*8 + 4*L.h));
    if (kStore) *p = make_int4(v[r], v[8 + r], v[16 + r], v[24 + r]);
    else { int4 t = *p; v[r] = t.x; v[8 + r] = t.y; v[16 + r] = t.z; v[24 + r] = t.w; }
  }
}
template <bool kStore>
__device__ __forceinline__ void rowB_32(int *B, int lane, int *v) {
#pragma unroll
  for (int j = 0; j < 8; j++) {
    int4 *p = reinterpret_cast<int4 *>(B + b_off(lane, 4*j));
    if (kStore) *p = make_int4(v[4*j], v[4*j + 1], v[4*j + 2], v[4*j + 3]);
    else { int4 t = *p; v[4*j] = t.x; v[4*j + 1] = t.y; v[4*j + 2] = t.z; v[4*j + 3] = t.w; }
  }
}
template <bool kStore>
__device__ __forceinline__ void rowB_16(int *B, const Lane &L, int *v) {
#pragma unroll
  for (int a = 0; a < 2; a++) {
#pragma unroll
    for (int j = 0; j < 4; j++) {
      int4 *p = reinterpret_cast<int4 *>(B + b_off(L.r16 + L.k16 + 8*a, L.c16 + 4*j));
      int *w = v + 16*a + 4*j;
      if (kStore) *p = make_int4(w[0], w[1], w[2], w[3]);
      else { int4 t = *p; w[0] = t.x; w[1] = t.y; w[2] = t.z; w[3] = t.w; }
    }
  }
}
template <bool kStore>
__device__ __forceinline__ void rowB_8(int *B, const Lane &L, int *v) {
#pragma unroll
  for (int a = 0; a < 4; a++) {
#pragma unroll
    for (int j = 0; j < 2; j++) {
      int4 *p = reinterpret_cast<int4 *>(B + b_off(L.ur*8 + 2*a + L.h, L.uc*8 + 4*j));
      int *w = v + 8*a + 4*j;
      if (kStore) *p = make_int4(w[0], w[1], w[2], w[3]);
      else { int4 t = *p; w[0] = t.x; w[1] = t.y; w[2] = t.z; w[3] = t.w; }
    }
  }
}

// ---------------------------------------------------------------- 4x4 leaves
// v[8*q + r] = sample (row r, column q) of the 8x4 strip: two stacked blocks.
// Column DCTs first, then row DCTs (reference src/dct.c:137-142).
__device__ __forceinline__ void fdct4x4_pair(int *v) {
#pragma unroll
  for (int b = 0; b < 2; b++) {
#pragma unroll
    for (int q = 0; q < 4; q++) tdlt_fdct4(v + 8*q + 4*b);   // column q, rows 4b..4b+3
#pragma unroll
    for (int r = 0; r < 4; r++) {
      int t[4] = {v[4*b + r], v[8 + 4*b + r], v[16 + 4*b + r], v[24 + 4*b + r]};
      tdlt_fdct4(t);
      v[4*b + r] = t[0]; v[8 + 4*b + r] = t[1]; v[16 + 4*b + r] = t[2]; v[24 + 4*b + r] = t[3];
    }
  }
}
// Row inverses first, then column inverses (reference src/dct.c:144-149).
__device__ __forceinline__ void idct4x4_pair(int *v) {
#pragma unroll
  for (int b = 0; b < 2; b++) {
#pragma unroll
    for (int r = 0; r < 4; r++) {
      int t[4] = {v[4*b + r], v[8 + 4*b + r], v[16 + 4*b + r], v[24 + 4*b + r]};
      tdlt_idct4(t);
      v[4*b + r] = t[0]; v[8 + 4*b + r] = t[1]; v[16 + 4*b + r] = t[2]; v[24 + 4*b + r] = t[3];
    }
#pragma unroll
    for (int q = 0; q < 4; q++) tdlt_idct4(v + 8*q + 4*b);
  }
}

// 1-D transforms of whatever lines a lane holds, by leaf size.
template <bool kInverse>
__device__ __forceinline__ void lines_dct(int lsz, int *v) {
  if (lsz == 3) {
    if (kInverse) tdlt_idct32(v); else tdlt_fdct32(v);
  }
  else if (lsz == 2) {
#pragma unroll
    for (int q = 0; q < 2; q++) { if (kInverse) tdlt_idct16(v + 16*q); else tdlt_fdct16(v + 16*q); }
  }
  else {
#pragma unroll
    for (int q = 0; q < 4; q++) { if (kInverse) tdlt_idct8(v + 8*q); else tdlt_fdct8(v + 8*q); }
  }
}

// --------------------------------------------------------------- lapping cells
// One 4x4 cell held as four rows.  vert(): the filter runs DOWN the columns
// (across a horizontal edge); horz(): ALONG the rows (across a vertical edge).
template <bool kInverse>
__device__ __forceinline__ void cell_vert(int4 &r0, int4 &r1, int4 &r2, int4 &r3, int fl) {
  if (fl & 1) {
    if (kInverse) { post4(r0.x, r1.x, r2.x, r3.x); post4(r0.y, r1.y, r2.y, r3.y); }
    else { pre4(r0.x, r1.x, r2.x, r3.x); pre4(r0.y, r1.y, r2.y, r3.y); }
  }
  if (fl & 2) {
    if (kInverse) { post4(r0.z, r1.z, r2.z, r3.z); post4(r0.w, r1.w, r2.w, r3.w); }
    else { pre4(r0.z, r1.z, r2.z, r3.z); pre4(r0.w, r1.w, r2.w, r3.w); }
  }
}
template <bool kInverse>
__device__ __forceinline__ void cell_horz(int4 &r0, int4 &r1, int4 &r2, int4 &r3, int fl) {
  if (fl & 4) {
    if (kInverse) { post4(r0.x, r0.y, r0.z, r0.w); post4(r1.x, r1.y, r1.z, r1.w); }
    else { pre4(r0.x, r0.y, r0.z, r0.w); pre4(r1.x, r1.y, r1.z, r1.w); }
  }
  if (fl & 8) {
    if (kInverse) { post4(r2.x, r2.y, r2.z, r2.w); post4(r3.x, r3.y, r3.z, r3.w); }
    else { pre4(r2.x, r2.y, r2.z, r2.w); pre4(r3.x, r3.y, r3.z, r3.w); }
  }
}

// Processes cell (i, j) in place: rows 4i..4i+3, columns 4j+2..4j+5 of A, i.e.
// the cell centred on tile-local (4i, 4j).  Forward order: coarser edge first,
// vertical-direction first on ties; the inverse undoes that back to front.
template <bool kInverse>
__device__ __forceinline__ void lap_cell(int *A, int i, int j, int fl) {
  if ((fl & 15) == 0) return;
  int2 *p = reinterpret_cast<int2 *>(A + a_off(4*i, 4*j + 2));
  constexpr int S = kAStride/2;
  int2 lo, hi;
  int4 r0, r1, r2, r3;
  lo = p[0]; hi = p[1]; r0 = make_int4(lo.x, lo.y, hi.x, hi.y);
  lo = p[S]; hi = p[S + 1]; r1 = make_int4(lo.x, lo.y, hi.x, hi.y);
  lo = p[2*S]; hi = p[2*S + 1]; r2 = make_int4(lo.x, lo.y, hi.x, hi.y);
  lo = p[3*S]; hi = p[3*S + 1]; r3 = make_int4(lo.x, lo.y, hi.x, hi.y);
  bool vfirst = ((fl & 16) != 0) != kInverse;
  if (!vfirst) cell_horz<kInverse>(r0, r1, r2, r3, fl);
  cell_vert<kInverse>(r0, r1, r2, r3, fl);
  if (vfirst) cell_horz<kInverse>(r0, r1, r2, r3, fl);
  p[0] = make_int2(r0.x, r0.y); p[1] = make_int2(r0.z, r0.w);
  p[S] = make_int2(r1.x, r1.y); p[S + 1] = make_int2(r1.z, r1.w);
  p[2*S] = make_int2(r2.x, r2.y); p[2*S + 1] = make_int2(r2.z, r2.w);
  p[3*S] = make_int2(r3.x, r3.y); p[3*S + 1] = make_int2(r3.z, r3.w);
}

}  // namespace tdlt

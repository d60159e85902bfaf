// Region kernels: shared geometry, shared-memory layout and the task code the
// forward and inverse TDLT kernels are assembled from.  sm_100a only.
//
// A CTA (8 warps) owns one REGION of one plane of one frame: 128 x 64 samples =
// 4 x 2 tiles of 32 x 32 = 16 x 8 UNITS of 8 x 8.  All the work of a region is
// first sorted into lists of equal-shaped tasks, so that every warp instruction
// runs the same code on 32 busy lanes whatever the block-size map looks like:
//
//   lapping    the levels superblock / 16 / 8 in ONE barrier-free pass over the 4x4 blocks
//              of the grid offset by (2, 2) (region_lap_pass below): cell tasks (lane =
//              a crossing of two 8-lines, its block and the half blocks either side on
//              the row line) and edge tasks (lane = one row of a column line between
//              crossings).  Level 4 (mid-edges of a unit split into 4x4) is filtered in
//              registers by the transform task of that unit.
//   transforms class 0: an 8x8 unit split into four 4x4 blocks  (2 lanes)
//              class 1: an 8x8 block                            (2 lanes)
//              class 2: a 16x16 block                           (8 lanes)
//              class 3: a 32x32 block                           (32 lanes)
//              every lane always carries 32 samples in registers.
//
// Shared memory of a CTA (39 KB: five CTAs per SM together with 48 registers):
//   A   68 rows x 140 words: region rows -2..65, columns -4..135 (one TMA box;
//       the innermost TMA coordinate must be 16-byte aligned, hence 4 columns
//       of halo where the lapping needs 2; 140 words = 35 16-byte chunks, an
//       odd number, so that 8 consecutive rows of one chunk column fall in 8
//       different bank groups -- for 16-byte accesses along a row AND down a
//       column, so the two passes of every transform run in place in A)
//   unit tables (leaf size, edge bits), crossing-order masks, task lists, counters.
#pragma once
#include "tdlt_device.cuh"
#include "tdlt_tasks.cuh"

namespace tdlt {

// Phase timestamps for tools/trace_phases.py (debug builds only: -DTDLT_TRACE).
#ifdef TDLT_TRACE
extern __device__ long long *g_tdlt_trace;
#define TDLT_MARK(cta, k) \
  do { \
    if (threadIdx.x == 0 && g_tdlt_trace && (cta) < 4096) g_tdlt_trace[(cta)*16 + (k)] = clock64(); \
  } while (0)
#else
#define TDLT_MARK(cta, k) do { } while (0)
#endif

// Experiment switches (tools/variants.py builds libraries with them; the product
// build leaves them at their defaults): minimum CTAs per SM the kernels are compiled
// for, and phases to skip when measuring what a phase costs (results are then wrong;
// 512: inverse takes the regions of a frame in raster order -- still right).
#ifndef TDLT_MINB
#define TDLT_MINB 5
#endif
#ifndef TDLT_BULK_STORE
#define TDLT_BULK_STORE 1   // forward: the coefficient rows leave shared memory as bulk copies (measured -2.5 %)
#endif
#ifndef TDLT_BULK_STORE_INV
#define TDLT_BULK_STORE_INV 0   // inverse: the same for the owned rectangle (measured +1 %: off; 2 .. 5 are the experiments of DESIGN section 4)
#endif
#ifndef TDLT_ABLATE
#define TDLT_ABLATE 0    // 1: no lapping, 2: no transforms, 4: no DC pyramid, 8: no set-up, 16: no global I/O (forward); inverse: 32 no strip exchange, 64 ticket = blockIdx, 128 no output stores, 1024 no polling for the ring (what is not there after the transforms is left out), 2048 the left neighbour's columns are not fetched at all, 4096 the inverse stores columns 0 .. 127 (whole sectors, wrong at the seams)
#endif

constexpr int kRegW = 128, kRegH = 64;       // region size in samples
constexpr int kRegTX = 4, kRegTY = 2;        // tiles across / down
constexpr int kRegTiles = kRegTX*kRegTY;
constexpr int kRegUX = 16, kRegUY = 8;       // units across / down
constexpr int kAP = 140;                     // words per row of the region buffer A
constexpr int kARegRows = kRegH + 4;
constexpr int kARegBytes = kAP*kARegRows*4;  // 38080, what the TMA load delivers
constexpr int kRegThreads = 256;
constexpr int kRegWarps = kRegThreads/32;
constexpr int kUS = 20;                      // row stride of the unit tables (18 used)
constexpr int kUTab = kUS*(kRegUY + 2);      // 200 entries: units -1..16 x -1..8

// Shared-memory carve-up (bytes from a 1024-aligned base).  The kPlanBytes from
// kOffLs on are the region's PLAN -- unit tables, task lists, class counts, edge masks
// -- which depends on the block-size map and the geometry only: tdlt_plan_kernel
// writes one such record per region to global memory and the transform kernels
// fetch it with one bulk copy next to the samples.
constexpr int kOffA = 0;
constexpr int kOffLs = kOffA + kARegBytes;               // 38080
constexpr int kOffUb = kOffLs + kUTab;                   // unit edge bits
constexpr int kOffTasks = kOffUb + kUTab;
constexpr int kTaskListStride = 128;                     // bytes per class list
constexpr int kOffCnt = kOffTasks + 4*kTaskListStride;   // 8 ints, [4..7] = class counts
constexpr int kOffMsk = kOffCnt + 32;                    // 9 crossing-order masks (20 words reserved)
constexpr int kOffBar = kOffMsk + 80;
constexpr int kPlanBytes = kOffBar - kOffLs;
#ifndef TDLT_PLAN_PER_CTA
#define TDLT_PLAN_PER_CTA 8
#endif
constexpr int kPlanPerCta = TDLT_PLAN_PER_CTA;           // regions per CTA of tdlt_plan_kernel
constexpr int kOffDyn = kOffBar + 16;                    // [0] ticket, [1] task draw counter
constexpr int kRegSmemBytes = kOffDyn + 16;
static_assert(kOffCnt % 4 == 0 && kOffBar % 8 == 0 && kOffLs % 16 == 0, "alignment");
static_assert(kPlanBytes == 1024, "one plan record = 1 KB");
static_assert(5*(kRegSmemBytes + 1024) <= 232448, "five CTAs per SM");

struct RegionPlane {
  int w, h;          // plane size in samples
  int dec;           // 0 or 1
  int nrx, nry;      // regions across / down
  int reg0;          // index of this plane's first region within a frame
  float inv_nrx;     // 1 / nrx
  int ngx, grp0;     // plan kernel: groups of kPlanPerCta regions across; first group of the plane
};

struct RegionBatch {
  RegionPlane pl[kMaxPlanes];
  int nplanes;
  int regs_per_frame;
  float inv_rpf;          // 1 / regs_per_frame
  int groups_per_frame;   // plan kernel grid
  int nframes;
  float inv_nframes;      // 1 / nframes
  int pic_w, pic_h;      // LUMA picture size (gates the split filters)
  int bw, bh;            // block-size map: entries across / down (8x8 luma units)
  int flags;             // OD_TDLT_KEYFRAME | OD_TDLT_HAAR | kFlagCodingOrder
  int pix_shift;         // fused pixel conversion: the shift of src/state.c:1221-1223 / 1236-1238
  const unsigned short *cotab;   // fused coding order: source positions (kCoTabEntries), or null
};

struct RegionId {
  int frame, pli, rx, ry;
};

// r = region index within its frame.
__device__ __forceinline__ RegionId decode_region(const RegionBatch &bd, int frame, int r) {
  RegionId id;
  id.frame = frame;
  int p = 0;
#pragma unroll
  for (int q = 1; q < kMaxPlanes; q++) {
    if (q < bd.nplanes && r >= bd.pl[q].reg0) p = q;
  }
  id.pli = p;
  r -= bd.pl[p].reg0;
  // r / nrx without an integer division (r < 2^16, nrx <= 2^9: the float quotient of
  // r + 0.5 is never within rounding distance of an integer)
  id.ry = __float2int_rz((r + 0.5f)*bd.pl[p].inv_nrx);
  id.rx = r - id.ry*bd.pl[p].nrx;
  return id;
}

struct RegionSmem {
  int *A;
  unsigned char *ls;      // leaf log-size per unit, 0xFF outside the plane
  unsigned char *ub;      // edge bits per unit (kUb*)
  unsigned char *tasks;   // 4 lists
  int *cnt;               // [4..7] task classes
  unsigned *msk;          // [iu]: bit ju = row line 8*iu is at least as coarse as column line 8*ju
  unsigned long long *bar;
  int *dyn;               // [0] ticket (inverse), [1] task draw counter
};

__device__ __forceinline__ RegionSmem carve_region(unsigned char *base) {
  RegionSmem s;
  s.A = reinterpret_cast<int *>(base + kOffA);
  s.ls = base + kOffLs;
  s.ub = base + kOffUb;
  s.tasks = base + kOffTasks;
  s.cnt = reinterpret_cast<int *>(base + kOffCnt);
  s.msk = reinterpret_cast<unsigned *>(base + kOffMsk);
  s.bar = reinterpret_cast<unsigned long long *>(base + kOffBar);
  s.dyn = reinterpret_cast<int *>(base + kOffDyn);
  return s;
}

// ---------------------------------------------------------------------------
// Fused pixel <-> coefficient conversion (reference od_ref_buf_to_coeff /
// od_coeff_to_ref_buf, src/state.c:1209-1316; clamps src/odintrin.h:104-113): the
// forward kernel takes its region as a TMA box of 8- or 16-bit reference-buffer samples
// and widens it into A; the inverse kernel narrows its owned rectangle on the way out.
//   kPix 1 (xstride 1):  c = (p - 128) << shift           p = clamp255(((c + rnd) >> shift) + 128)
//   kPix 2 (xstride 2):  c = (p - 2048 + rnd) >> shift    p = clamp(c << shift + 2048, 0, 4095)
// The sample box (PixBox below) lands in the tail of A's own storage, at a 128-byte aligned offset; all threads take
// their part of it into registers, a barrier, then they write the widened rows.
// ---------------------------------------------------------------------------
// The box starts kSkip samples left of A's first column because the innermost TMA coordinate
// must be a multiple of 16 BYTES (16 samples of one byte, 8 of two), as must the box row.
template <int kPix> struct PixBox {
  static constexpr int kSkip = 16/kPix - 4;                       // box column of A's column 0: 12 / 4
  static constexpr int kW = kPix == 1 ? 160 : 144;                // samples per box row (>= kSkip + kAP)
  static constexpr int kBytes = kW*kARegRows*kPix;
  static constexpr int kOff = ((kARegBytes - kBytes)/128)*128;
  static_assert(kW >= kSkip + kAP && (kW*kPix) % 16 == 0, "box row");
  static_assert(kOff >= 0 && kOff + kBytes <= kARegBytes, "the sample box fits in A");
};
template <int kPix>
__device__ __forceinline__ int pix_to_coeff1(int p, int shift) {
  return kPix == 1 ? (p - 128)*(1 << shift) : (p - 2048 + ((1 << shift) >> 1)) >> shift;
}
template <int kPix>
__device__ __forceinline__ int coeff_to_pix1(int c, int shift) {
  return kPix == 1 ? min(max(((c + ((1 << shift) >> 1)) >> shift) + 128, 0), 255)
                   : min(max(c*(1 << shift) + 2048, 0), 4095);
}
// Called by all threads after the box has landed; ends with the widened region in A (the
// caller's next barrier orders it).  Work item = 4 samples = one 16-byte chunk of A.
template <int kPix>
__device__ __forceinline__ void region_widen_box(unsigned char *smem, int shift, int tid) {
  constexpr int kItems = kARegRows*(kAP/4);                       // 2380
  constexpr int kIter = (kItems + kRegThreads - 1)/kRegThreads;   // 10
  const unsigned char *box = smem + kOffA + PixBox<kPix>::kOff;
  int4 *A4 = reinterpret_cast<int4 *>(smem + kOffA);
  unsigned lo[kIter], hi[kIter];
#pragma unroll
  for (int it = 0; it < kIter; it++) {
    const int j = tid + kRegThreads*it;
    lo[it] = hi[it] = 0;
    if (j < kItems) {
      const int row = (j*1873) >> 16, q = j - row*(kAP/4);        // j / 35, exact below 2^15
      const unsigned char *src = box + (row*PixBox<kPix>::kW + PixBox<kPix>::kSkip + 4*q)*kPix;
      if (kPix == 1) lo[it] = *reinterpret_cast<const unsigned *>(src);
      else { uint2 t = *reinterpret_cast<const uint2 *>(src); lo[it] = t.x; hi[it] = t.y; }
    }
  }
  __syncthreads();
#pragma unroll
  for (int it = 0; it < kIter; it++) {
    const int j = tid + kRegThreads*it;
    if (j < kItems) {
      int4 o;
      if (kPix == 1) {
        o.x = pix_to_coeff1<1>(lo[it] & 255, shift);         o.y = pix_to_coeff1<1>((lo[it] >> 8) & 255, shift);
        o.z = pix_to_coeff1<1>((lo[it] >> 16) & 255, shift); o.w = pix_to_coeff1<1>(lo[it] >> 24, shift);
      }
      else {
        o.x = pix_to_coeff1<2>((short)(lo[it] & 0xffff), shift); o.y = pix_to_coeff1<2>((short)(lo[it] >> 16), shift);
        o.z = pix_to_coeff1<2>((short)(hi[it] & 0xffff), shift); o.w = pix_to_coeff1<2>((short)(hi[it] >> 16), shift);
      }
      A4[j] = o;
    }
  }
}

// Unit-table index of region unit (uy, ux), uy in [-1, 8], ux in [-1, 16].
__device__ __forceinline__ int ut(int uy, int ux) { return (uy + 1)*kUS + ux + 1; }

// Edge bits of a unit (all zero outside the plane).
//  kUbTop   the horizontal edge along the unit's top side is filtered
//  kUbLeft  the vertical edge along its left side is filtered
//  kUbMidH  the unit is split into 4x4 blocks and its horizontal mid-edge is filtered
//  kUbMidV  ... and its vertical mid-edge
constexpr int kUbTop = 1, kUbLeft = 2, kUbMidH = 4, kUbMidV = 8;

// One lane appends `item` to the list `which` (0..3, or -1 for none) of a
// family of 4 lists; warp-aggregated (one shared-memory atomic per list and call).
template <typename T>
__device__ __forceinline__ void list_push4(T *lists, int stride, int *cnt, int which, T item,
                                           int lane) {
  unsigned peers = __match_any_sync(0xffffffffu, which);
  int leader = __ffs(peers) - 1;
  int base = 0;
  if (lane == leader && which >= 0) base = atomicAdd(cnt + which, __popc(peers));
  base = __shfl_sync(0xffffffffu, base, leader);
  if (which >= 0) lists[which*stride + base + __popc(peers & ((1u << lane) - 1))] = item;
}

// ---------------------------------------------------------------------------
// Lapping.  The filtered edges of a plane lie on lines at multiples of 4; the
// LEVEL of a line is the largest power of two dividing its coordinate, capped
// at the superblock size (multiples of 32 -> the SB edges / 32x32 splits, 16,
// 8, and 4 = the mid-edges of 8x8 units split into 4x4).  The reference filters
// level by level, coarse to fine, and inside a level first across all
// horizontal edges (down the columns), then across all vertical edges
// (src/filter.c:1493-1508, 1567-1584; recursion src/encode.c:1311-1327); the
// inverse runs the same sequence backwards.  One phase = one (level, direction):
//   down the columns:  a task is a 16-byte chunk (4 columns) of one line: 4 rows
//                      x int4, four filters side by side.  8 consecutive lanes
//                      read 8 consecutive chunks: no bank conflicts.
//   along the rows:    a task is one row of one line: the last two words of the
//                      chunk left of the edge and the first two of the chunk
//                      right of it.  8 consecutive lanes read 8 consecutive rows
//                      (35 chunks per row, odd): no bank conflicts.
// Region lines: rows / columns 0, 8, .., 64 / 128 (levels >= 8) and 4, 12, ..
// (level 4).  Regions start at multiples of 64 / 128, so the level of a line
// follows from its region-local coordinate.
// ---------------------------------------------------------------------------
// Forward: the region needs its 128 x 64 samples lapped; the lines on its far
// sides (row 64, column 128) and the halo are recomputed from input samples.
// Inverse: the region OWNS the cells centred on rows 0..60 and columns 0..124,
// i.e. output rows -2..61 and columns -2..125 (plus the last two rows / columns
// of the plane where it ends there).
// Level 4 (the mid-edges of units split into 4x4 blocks) never leaves its unit
// and is filtered in registers by the transform task of that unit (lap4_unit
// below), so the phases here run over the levels 8 .. superblock only.
// Level of the line at coordinate 8k: the largest power of two dividing it, capped at
// the superblock size.
__device__ __forceinline__ int lap_level8(int k, int sb) {
  int c = 8*k;
  return (c & (sb - 1)) == 0 ? sb : (c & -c);
}
// Transposes a 4x4 register block when t (12 selects): used where a lane needs the two
// filter directions in the other order than the code runs them.
__device__ __forceinline__ void transpose4_if(int (&x)[4][4], bool t) {
#pragma unroll
  for (int r = 0; r < 4; r++) {
#pragma unroll
    for (int c = r + 1; c < 4; c++) {
      int a = x[r][c], b = x[c][r];
      x[r][c] = t ? b : a;
      x[c][r] = t ? a : b;
    }
  }
}
// ---------------------------------------------------------------------------
// One-pass lapping of the levels 8 .. superblock: every 4-tap
// filter of those levels lies inside ONE 4x4 block of the grid offset by (2, 2), so all
// of them can run in a single barrier-free pass, each block taken into registers, its
// filters applied in the reference's order, and written back.  Two kinds of warp task,
// both with whole-warp conflict-poor shared-memory access:
//   cell task  lane = the crossing (iu, ju) of row line 8*iu and column line 8*ju (iu in
//              0..8, ju in 0..16, 17 per row line).  The lane takes the two 16-byte chunks
//              around the crossing, buffer rows 8*iu .. +3 (region rows 8*iu-2 .. +1),
//              region columns 8*ju-4 .. 8*ju+3: the crossing's own 4x4 block in the middle
//              (filters of both directions, order by level) and, either side, half of the
//              blocks that only the horizontal edge crosses (filters down the columns).
//   edge task  lane = one row of one column line between two crossings (region rows
//              8*iu+2 .. +5, columns 8*ju-2 .. +1): one filter along the row; two column
//              lines and two units per lane pair up so that a half warp spreads over all
//              the 8-byte bank pairs its rows can reach.
// The forward pass needs the crossings 0..8 x 0..16 (its far sides are recomputed from the
// halo); the inverse owns 0..7 x 0..15, plus column-line 16's left half (columns 124, 125
// belong to it) and row line 8 where the plane ends.
// Order inside a crossing (src/filter.c:1493-1508, 1567-1584; src/encode.c:1311-1327):
// forward, the coarser line first and on a tie the filters down the columns (across the
// horizontal edge); the inverse backwards.  trm[iu] bit ju = level(row line iu) >=
// level(column line ju) (tdlt_plan.cu).  The code always runs "down the columns, then
// along the rows"; a lane that needs the other order works on the transposed block.
// ---------------------------------------------------------------------------
template <bool kInverse>
__device__ __forceinline__ void filt4(int &a, int &b, int &c, int &d) {
  if (kInverse) post4(a, b, c, d); else pre4(a, b, c, d);
}
constexpr int kCellsPerLine = kRegUX + 1;              // 17
constexpr int kCellsFwd = (kRegUY + 1)*kCellsPerLine;  // 153
constexpr int kCellsInv = kRegUY*kCellsPerLine;        // 136 (153 where the plane ends below)
constexpr int kCellTasks = (kCellsFwd + 31)/32;        // 5
constexpr int kEdgeTasksFwd = 9, kEdgeTasksInv = 8;

template <bool kInverse>
__device__ __forceinline__ void lap_cell_task(int *A, const unsigned char *ub, const unsigned *trm,
                                              int idx, int ncells) {
  if (idx >= ncells) return;
  const int iu = (idx*241) >> 12, ju = idx - kCellsPerLine*iu;     // idx / 17, exact below 2^11
  const unsigned char *u = ub + ut(iu, ju);
  const unsigned u0 = u[0], ul = u[-1], ua = u[-kUS];
  static_assert(kUbTop == 1 && kUbLeft == 2, "bit positions used below");
  // bit 0: filters down the columns 0..3 of the eight (the unit left of the crossing has its
  // top edge filtered), 1: columns 4..7, 2: along the rows 0, 1 of the crossing's block (the
  // unit above has its left edge filtered), 3: rows 2, 3
  unsigned f = (ul & 1) | ((u0 & 1) << 1) | ((ua & 2) << 1) | ((u0 & 2) << 2);
  if (kInverse && ju == kRegUX) f &= 1;     // column line 16 is the next region's: only 124, 125
  if (!f) return;
  const bool both = (f & 3) && (f & 12);
  const bool tr = both && ((((trm[iu] >> ju) & 1) != 0) == kInverse);
#ifdef TDLT_FAKE_CELLS
  int4 *p = reinterpret_cast<int4 *>(A + 8*iu*kAP + 4*(idx & 31));   // timing experiment: conflict-free, wrong
#else
  int4 *p = reinterpret_cast<int4 *>(A + 8*iu*kAP + 8*ju);
#endif
  constexpr int S = kAP/4;
  int4 L0 = p[0], R0 = p[1], L1 = p[S], R1 = p[S + 1];
  int4 L2 = p[2*S], R2 = p[2*S + 1], L3 = p[3*S], R3 = p[3*S + 1];
  if (f & 1) { filt4<kInverse>(L0.x, L1.x, L2.x, L3.x); filt4<kInverse>(L0.y, L1.y, L2.y, L3.y); }
  if (f & 2) { filt4<kInverse>(R0.z, R1.z, R2.z, R3.z); filt4<kInverse>(R0.w, R1.w, R2.w, R3.w); }
  int x[4][4] = {{L0.z, L0.w, R0.x, R0.y}, {L1.z, L1.w, R1.x, R1.y},
                 {L2.z, L2.w, R2.x, R2.y}, {L3.z, L3.w, R3.x, R3.y}};
  const unsigned g = tr ? ((f & 3) << 2) | (f >> 2) : f;   // on the transposed block rows and columns swap
  transpose4_if(x, tr);
  if (g & 1) { filt4<kInverse>(x[0][0], x[1][0], x[2][0], x[3][0]); filt4<kInverse>(x[0][1], x[1][1], x[2][1], x[3][1]); }
  if (g & 2) { filt4<kInverse>(x[0][2], x[1][2], x[2][2], x[3][2]); filt4<kInverse>(x[0][3], x[1][3], x[2][3], x[3][3]); }
  if (g & 4) { filt4<kInverse>(x[0][0], x[0][1], x[0][2], x[0][3]); filt4<kInverse>(x[1][0], x[1][1], x[1][2], x[1][3]); }
  if (g & 8) { filt4<kInverse>(x[2][0], x[2][1], x[2][2], x[2][3]); filt4<kInverse>(x[3][0], x[3][1], x[3][2], x[3][3]); }
  transpose4_if(x, tr);
  p[0] = make_int4(L0.x, L0.y, x[0][0], x[0][1]);       p[1] = make_int4(x[0][2], x[0][3], R0.z, R0.w);
  p[S] = make_int4(L1.x, L1.y, x[1][0], x[1][1]);       p[S + 1] = make_int4(x[1][2], x[1][3], R1.z, R1.w);
  p[2*S] = make_int4(L2.x, L2.y, x[2][0], x[2][1]);     p[2*S + 1] = make_int4(x[2][2], x[2][3], R2.z, R2.w);
  p[3*S] = make_int4(L3.x, L3.y, x[3][0], x[3][1]);     p[3*S + 1] = make_int4(x[3][2], x[3][3], R3.z, R3.w);
}

// Edge task v: v < 8 -> column lines j0 and j0 + 2, j0 = 4*(v >> 1) + (v & 1), lane = (row k of
// the four, which line, unit iu in 0..3) and the same for unit iu + 4; v == 8 (forward only)
// -> column line 16, lane = (k, iu in 0..7).
template <bool kInverse>
__device__ __forceinline__ void lap_edge_task(int *A, const unsigned char *ub, int v, int lane) {
  const int k = lane & 3;
  const bool two = v < 8;
  const int ju = two ? ((v >> 1) << 2) + (v & 1) + ((lane >> 1) & 2) : kRegUX;
  const int iua = two ? lane >> 3 : lane >> 2;
  const bool ona = (ub[ut(iua, ju)] & kUbLeft) != 0;
  const bool onb = two && (ub[ut(iua + 4, ju)] & kUbLeft) != 0;
  int2 *pa = reinterpret_cast<int2 *>(A + (8*iua + 4 + k)*kAP + 8*ju + 2);   // region columns 8ju-2 .. +1
  int2 *pb = pa + 32*kAP/2;                                                  // the unit four rows of units below
  int2 la, ha, lb, hb;
  if (ona) { la = pa[0]; ha = pa[1]; }
  if (onb) { lb = pb[0]; hb = pb[1]; }
  if (ona) filt4<kInverse>(la.x, la.y, ha.x, ha.y);
  if (onb) filt4<kInverse>(lb.x, lb.y, hb.x, hb.y);
  if (ona) { pa[0] = la; pa[1] = ha; }
  if (onb) { pb[0] = lb; pb[1] = hb; }
}

// The whole pass, called by all warps of the CTA; the caller puts a block barrier either
// side.  ncells: kCellsFwd, or kCellsInv (inverse; kCellsFwd where the plane ends below).
template <bool kInverse>
__device__ __forceinline__ void region_lap_pass(int *A, const unsigned char *ub, const unsigned *trm,
                                                int ncells, int warp, int lane) {
  if (warp < kCellTasks) {
    lap_cell_task<kInverse>(A, ub, trm, 32*warp + lane, ncells);
  }
  else {
    constexpr int nv = kInverse ? kEdgeTasksInv : kEdgeTasksFwd;
#pragma unroll 1
    for (int v = warp - kCellTasks; v < nv; v += kRegWarps - kCellTasks) {
      lap_edge_task<kInverse>(A, ub, v, lane);
    }
  }
}

// Inverse: the owned rectangle goes out after the pass, buffer rows b0 .. b1-1, region
// columns cs .. ce-1 (cs = -2 or 0; ce = 126, or a multiple of 32 where the plane ends);
// g -> the plane sample at buffer row 0, region column -2.  A warp takes a row at a time:
// 16-byte vectors for the columns 0 .. 123 (whole sectors), 8-byte pieces at either end.
__device__ __forceinline__ void region_store_owned(const int *A, int *g, size_t pitch, int b0, int b1,
                                                   int cs, int ce, int warp, int lane) {
  const int c = 4*lane;
  const bool full = c + 3 < ce, half = !full && c + 1 < ce, left = lane == 0 && cs < 0;
  const int *arow = A + (b0 + warp)*kAP + 4 + c;       // -> region column c
  int *grow = g + (size_t)(b0 + warp)*pitch + 2 + c;
#pragma unroll 2
  for (int b = b0 + warp; b < b1; b += kRegWarps, arow += kRegWarps*kAP, grow += kRegWarps*pitch) {
    // (one 16-byte load: left to itself the compiler splits it in two because of the 8-byte
    // store below, and 8-byte loads at a 16-byte lane stride conflict two ways)
    int4 v;
    asm volatile("ld.shared.v4.s32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(smem_u32(arow)));
    if (full) *reinterpret_cast<int4 *>(grow) = v;
    if (half) *reinterpret_cast<int2 *>(grow) = make_int2(v.x, v.y);
    if (left) *reinterpret_cast<int2 *>(grow - 2) = *reinterpret_cast<const int2 *>(arow - 2);
  }
}

// The same by the copy engine: one thread per row issues a bulk copy of the row's 16-byte
// aligned part (region columns 0 .. 123, or up to where the plane ends) and stores the 8-byte
// pieces at either end itself.  The caller has fenced the lapping pass's writes
// (fence_proxy_async_smem) before its barrier.
__device__ __forceinline__ void region_store_owned_bulk(const int *A, int *g, size_t pitch, int b0,
                                                        int b1, int cs, int ce, int tid) {
  const int b = b0 + tid;
  if (b >= b1) return;
  const int nb = ce & ~3;                               // columns that go out in bulk (124, or all)
  const int *arow = A + b*kAP + 4;                      // -> region column 0
  int *grow = g + (size_t)b*pitch + 2;
  // (TDLT_BULK_STORE_INV 2: no end pieces, 3: 384 bytes only -- timing experiments, wrong results)
  if (TDLT_BULK_STORE_INV == 4 && nb > 96) {            // whole 128-byte lines, then the rest
    bulk_s2g(grow, arow, 384);
    bulk_s2g(grow + 96, arow + 96, 4*(nb - 96));
  }
  else bulk_s2g(grow, arow, TDLT_BULK_STORE_INV == 3 ? 384 : 4*nb);
  if (TDLT_BULK_STORE_INV == 1 || TDLT_BULK_STORE_INV == 4) {
    if (cs < 0) *reinterpret_cast<int2 *>(grow - 2) = *reinterpret_cast<const int2 *>(arow - 2);
    if (ce > nb) *reinterpret_cast<int2 *>(grow + nb) = *reinterpret_cast<const int2 *>(arow + nb);
  }
  bulk_commit_wait_read();
}

// Mixed form: the first 96 columns of every owned row (three whole 128-byte lines of the
// plane) leave as a bulk copy issued by one elected lane per warp, the columns 96 .. ce-1 and
// the two left of the region by ordinary stores, four rows per warp instruction (lane = row
// of the four, 16-byte chunk of the eight).  The caller has fenced the lapping pass's writes.
__device__ __forceinline__ void region_store_owned_mixed(const int *A, int *g, size_t pitch, int b0,
                                                         int b1, int cs, int ce, int warp, int lane) {
  constexpr int kBulkCols = 96, kRowsPerWarp = 9;       // 8 x 9 rows cover the 66 a region can own
  const int wu = __shfl_sync(0xffffffffu, warp, 0);
  const int nbulk = min(ce, kBulkCols);
  if (elect_one()) {
    const int bfirst = max(b0, kRowsPerWarp*wu);
    unsigned long long ga = reinterpret_cast<unsigned long long>(g + (size_t)bfirst*pitch + 2);
    unsigned sa = smem_u32(A + bfirst*kAP + 4);
    const int n = min(b1, kRowsPerWarp*(wu + 1)) - bfirst;
#pragma unroll
    for (int i = 0; i < kRowsPerWarp; i++) {
      if (i < n) bulk_s2g_addr(ga, sa, 4*nbulk);
      ga += 4ull*pitch;
      sa += 4*kAP;
    }
  }
  // the rest of the rows
  const int c = kBulkCols + 4*(lane & 7);
  const bool full = c + 3 < ce, half = !full && c + 1 < ce, left = (lane & 7) == 0 && cs < 0;
#pragma unroll 1
  for (int b = b0 + 4*warp + (lane >> 3); b < b1; b += 4*kRegWarps) {
    const int *arow = A + b*kAP + 4;
    int *grow = g + (size_t)b*pitch + 2;
    if (full) *reinterpret_cast<int4 *>(grow + c) = *reinterpret_cast<const int4 *>(arow + c);
    if (half) *reinterpret_cast<int2 *>(grow + c) = *reinterpret_cast<const int2 *>(arow + c);
    if (left) *reinterpret_cast<int2 *>(grow - 2) = *reinterpret_cast<const int2 *>(arow - 2);
  }
  bulk_commit_wait_read();
}

// The same with the samples narrowed to kPix-byte reference-buffer pixels on the way out
// (od_coeff_to_ref_buf): g -> the pixel at buffer row 0, region column -2; pitch in bytes.
template <int kPix>
__device__ __forceinline__ void region_store_owned_pix(const int *A, unsigned char *g, size_t pitch,
                                                       int b0, int b1, int cs, int ce, int shift,
                                                       int warp, int lane) {
  const int c = 4*lane;
  const bool full = c + 3 < ce, half = !full && c + 1 < ce, left = lane == 0 && cs < 0;
  const int *arow = A + (b0 + warp)*kAP + 4 + c;
  unsigned char *grow = g + (size_t)(b0 + warp)*pitch + (2 + c)*kPix;     // -> region column c
#pragma unroll 2
  for (int b = b0 + warp; b < b1; b += kRegWarps, arow += kRegWarps*kAP, grow += kRegWarps*pitch) {
    int4 v;
    asm volatile("ld.shared.v4.s32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(smem_u32(arow)));
    const unsigned p0 = coeff_to_pix1<kPix>(v.x, shift), p1 = coeff_to_pix1<kPix>(v.y, shift);
    const unsigned p2 = coeff_to_pix1<kPix>(v.z, shift), p3 = coeff_to_pix1<kPix>(v.w, shift);
    if (kPix == 1) {
      if (full) *reinterpret_cast<unsigned *>(grow) = p0 | (p1 << 8) | (p2 << 16) | (p3 << 24);
      if (half) *reinterpret_cast<unsigned short *>(grow) = (unsigned short)(p0 | (p1 << 8));
    }
    else {
      if (full) *reinterpret_cast<uint2 *>(grow) = make_uint2(p0 | (p1 << 16), p2 | (p3 << 16));
      if (half) *reinterpret_cast<unsigned *>(grow) = p0 | (p1 << 16);
    }
    if (left) {
      const int2 l = *reinterpret_cast<const int2 *>(arow - 2);
      const unsigned q0 = coeff_to_pix1<kPix>(l.x, shift), q1 = coeff_to_pix1<kPix>(l.y, shift);
      if (kPix == 1) *reinterpret_cast<unsigned short *>(grow - 2) = (unsigned short)(q0 | (q1 << 8));
      else *reinterpret_cast<unsigned *>(grow - 4) = q0 | (q1 << 16);
    }
  }
}

// ---------------------------------------------------------------------------
// Fused coding order (OD_TDLT_CODING_ORDER; reference od_raster_to_coding_order /
// od_coding_order_to_raster, src/partition.c:143-199, as od_block_encode / od_block_decode
// call them per block, src/encode.c:1160-1161, 1219-1220, src/decode.c:580-581, 630): the
// forward kernel stores, and the inverse kernel loads, the BLOCK-LINEAR plane of coding-order
// vectors (layout: include/daala_tdlt.h, od_tdlt_coding_order_batch) instead of the raster
// plane, so the permutation never costs a pass of its own.  Where entry e of a superblock's
// range comes from depends on e and the size of the block it lies in only; cotab (built once
// per od_tdlt_set_scan_tables, tdlt_dering.cu) holds that position, row*S + column inside the
// superblock or kCoNoPos where the scans order nothing (the second half of a 32x32 block).
// A warp takes its tile: one 32x32 superblock of a full-size plane, or 2x2 superblocks of 16x16
// of a 4:2:0 chroma plane -- 1024 entries either way, four consecutive entries per lane and
// step (they lie in one 4x4 unit, hence in one block), eight steps.
// ---------------------------------------------------------------------------
constexpr int kFlagCodingOrder = 8;
constexpr int kCoTabEntries = 4*1024 + 3*256;
constexpr unsigned kCoNoPos = 0xffff;

struct CoSite {
  int4 *vec;            // the four entries in the block-linear plane
  int a[4];             // word offsets from the tile's origin in A, or -1 (entry not ordered)
};
// gtile -> the block-linear plane entry of the tile's first superblock; sbstep = entries from
// a superblock to the one below it (4:2:0 chroma only).
__device__ __forceinline__ CoSite co_site(const unsigned short *__restrict__ cotab,
                                          const unsigned char *ls0, int dec, int it, int lane,
                                          int *gtile, int sbstep) {
  CoSite c;
  const int idx = 128*it + 4*lane;
  int e, rowoff, coloff, L;
  uint2 pk;
  if (dec == 0) {
    e = idx;
    const int z = e >> 4;      // Z order of the 4x4 unit: x bit 0, y bit 0, x bit 1, ...
    const int ux = (z & 1) | ((z >> 1) & 2) | ((z >> 2) & 4), uy = ((z >> 1) & 1) | ((z >> 2) & 2) | ((z >> 3) & 4);
    L = ls0[(uy >> 1)*kUS + (ux >> 1)];
    pk = __ldg(reinterpret_cast<const uint2 *>(cotab + (L << 10) + e));
    rowoff = 0; coloff = 0;
    c.vec = reinterpret_cast<int4 *>(gtile + e);
  }
  else {
    const int sb = idx >> 8, sx = sb & 1, sy = sb >> 1;
    e = idx & 255;
    const int z = e >> 4;
    const int ux = (z & 1) | ((z >> 1) & 2), uy = ((z >> 1) & 1) | ((z >> 2) & 2);
    L = ls0[(2*sy + (uy >> 1))*kUS + 2*sx + (ux >> 1)];
    pk = __ldg(reinterpret_cast<const uint2 *>(cotab + 4*1024 + (L << 8) + e));
    rowoff = 16*sy; coloff = 16*sx;
    c.vec = reinterpret_cast<int4 *>(gtile + sy*sbstep + sx*256 + e);
  }
  const unsigned p[4] = {pk.x & 0xffff, pk.x >> 16, pk.y & 0xffff, pk.y >> 16};
  const int sh = 5 - dec, msk = (32 >> dec) - 1;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    c.a[k] = p[k] == kCoNoPos ? -1 : (rowoff + (int)(p[k] >> sh))*kAP + coloff + (int)(p[k] & msk);
  }
  return c;
}

// Forward: the tile leaves as coding-order vectors.  Entries the scans do not order are left
// as they are in the destination, as the reference leaves them.
__device__ __forceinline__ void region_store_tile_coding(const int *At, const unsigned char *ls0,
                                                         const unsigned short *cotab, int dec,
                                                         int *gtile, int sbstep, int lane) {
#pragma unroll 2
  for (int it = 0; it < 8; it++) {
    const CoSite c = co_site(cotab, ls0, dec, it, lane, gtile, sbstep);
    int v[4];
#pragma unroll
    for (int k = 0; k < 4; k++) v[k] = c.a[k] >= 0 ? At[c.a[k]] : 0;
    if ((c.a[0] | c.a[1] | c.a[2] | c.a[3]) >= 0) *c.vec = make_int4(v[0], v[1], v[2], v[3]);
    else {
      int *g = reinterpret_cast<int *>(c.vec);
#pragma unroll
      for (int k = 0; k < 4; k++) if (c.a[k] >= 0) g[k] = v[k];
    }
  }
}
// Inverse: the tile arrives as coding-order vectors and is scattered to its raster place in
// A; what the scans do not order starts at zero (od_init_skipped_coeffs on a keyframe,
// src/state.c:1340-1349: everything but the vector's entries is cleared first).
__device__ __forceinline__ void region_load_tile_coding(int *At, const unsigned char *ls0,
                                                        const unsigned short *cotab, int dec,
                                                        int *gtile, int sbstep, int lane) {
  {
    int4 *row = reinterpret_cast<int4 *>(At + (lane >> 3)*kAP + 4*(lane & 7));
#pragma unroll
    for (int it = 0; it < 8; it++) row[it*kAP] = make_int4(0, 0, 0, 0);    // 4 rows per step
  }
  __syncwarp();
  // four vectors in flight per lane before the first scatter (the loads are plain global
  // loads here, not a TMA box: their latency is this warp's to hide)
#pragma unroll 1
  for (int h = 0; h < 2; h++) {
    int4 v[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int idx = 128*(4*h + k) + 4*lane;
      const int *src = dec == 0 ? gtile + idx
                                : gtile + (idx >> 9)*sbstep + ((idx >> 8) & 1)*256 + (idx & 255);
      v[k] = __ldg(reinterpret_cast<const int4 *>(src));
    }
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const CoSite c = co_site(cotab, ls0, dec, 4*h + k, lane, gtile, sbstep);
      if (c.a[0] >= 0) At[c.a[0]] = v[k].x;
      if (c.a[1] >= 0) At[c.a[1]] = v[k].y;
      if (c.a[2] >= 0) At[c.a[2]] = v[k].z;
      if (c.a[3] >= 0) At[c.a[3]] = v[k].w;
    }
  }
}

// Level 4, in registers.  The two lanes of a unit hold it as v[8*q + r] = sample
// (row r, column 4*h + q).  The filter across the horizontal mid-edge (rows
// 2..5 of every column) is local; the one across the vertical mid-edge (columns
// 2..5 of every row) needs the neighbour lane's two adjacent columns: lane 0
// filters rows 0..3, lane 1 rows 4..7, and the halves travel by shuffle.
// Forward: columns first, then rows (src/filter.c:1493-1508); inverse: rows,
// then columns (:1537-1552).
template <bool kInverse>
__device__ __forceinline__ void lap4_rows(int *v, int h, unsigned mask) {
  int ra[4], rb[4];
#pragma unroll
  for (int j = 0; j < 4; j++) {
    int sa = h ? v[j] : v[20 + j];          // lane 1 sends columns 4, 5 of rows 0..3,
    int sc = h ? v[8 + j] : v[28 + j];      // lane 0 columns 2, 3 of rows 4..7
    ra[j] = __shfl_xor_sync(mask, sa, 1);
    rb[j] = __shfl_xor_sync(mask, sc, 1);
  }
#pragma unroll
  for (int j = 0; j < 4; j++) {
    int x0 = h ? ra[j] : v[16 + j];
    int x1 = h ? rb[j] : v[24 + j];
    int x2 = h ? v[4 + j] : ra[j];
    int x3 = h ? v[12 + j] : rb[j];
    if (kInverse) post4(x0, x1, x2, x3); else pre4(x0, x1, x2, x3);
    int ta = h ? x0 : x2, tb = h ? x1 : x3;   // the neighbour's half goes back
    int ua = __shfl_xor_sync(mask, ta, 1);
    int uc = __shfl_xor_sync(mask, tb, 1);
    v[4 + j] = h ? x2 : v[4 + j];
    v[12 + j] = h ? x3 : v[12 + j];
    v[16 + j] = h ? v[16 + j] : x0;
    v[24 + j] = h ? v[24 + j] : x1;
    v[j] = h ? ua : v[j];
    v[8 + j] = h ? uc : v[8 + j];
    v[20 + j] = h ? v[20 + j] : ua;
    v[28 + j] = h ? v[28 + j] : uc;
  }
}
template <bool kInverse>
__device__ __forceinline__ void lap4_cols(int *v) {
#pragma unroll
  for (int q = 0; q < 4; q++) {
    if (kInverse) post4(v[8*q + 2], v[8*q + 3], v[8*q + 4], v[8*q + 5]);
    else pre4(v[8*q + 2], v[8*q + 3], v[8*q + 4], v[8*q + 5]);
  }
}
// bits: the unit's edge bits (0 on idle lanes).  Called by the whole warp.
template <bool kInverse>
__device__ __forceinline__ void lap4_unit(int *v, int h, int bits) {
  unsigned mask = __ballot_sync(0xffffffffu, (bits & kUbMidV) != 0);
  if (kInverse) {
    if (bits & kUbMidV) lap4_rows<true>(v, h, mask);
    if (bits & kUbMidH) lap4_cols<true>(v);
  }
  else {
    if (bits & kUbMidH) lap4_cols<false>(v);
    if (bits & kUbMidV) lap4_rows<false>(v, h, mask);
  }
}

// ---------------------------------------------------------------------------
// Transform tasks.  A unit is named by its region unit coordinates (UR, UC),
// UR in 0..7, UC in 0..15; a task list entry is (UR << 4) | UC.  Sample (r, c) of
// the region lies at A[(r + 2)*kAP + c + 4].  Both passes of a transform run IN
// PLACE in A: every lane reads the lines it owns into registers, transforms
// them and writes them back; a __syncwarp separates the passes.  The row pitch
// of 35 16-byte chunks makes the bank group of chunk (row, q) equal to
// (3*row + q) mod 8, so eight lanes that read the same chunk column of eight
// consecutive rows, or eight consecutive chunks of one row, never collide.
// ---------------------------------------------------------------------------

// -- 8 rows x 4 columns: element (r, q) <-> v[8*q + r] -----------------------
__device__ __forceinline__ void ld_a_8x4(const int *a, int *v) {
#pragma unroll
  for (int r = 0; r < 8; r++) {
    int4 t = *reinterpret_cast<const int4 *>(a + r*kAP);
    v[r] = t.x; v[8 + r] = t.y; v[16 + r] = t.z; v[24 + r] = t.w;
  }
}
__device__ __forceinline__ void st_a_8x4(int *a, const int *v) {
#pragma unroll
  for (int r = 0; r < 8; r++) {
    *reinterpret_cast<int4 *>(a + r*kAP) = make_int4(v[r], v[8 + r], v[16 + r], v[24 + r]);
  }
}

// -- rows of an 8x8 block (a -> its origin): register group g (8 words) holds unit row
//    2*((g + mu) & 3) + h, all 8 columns.  mu = (UC - slot) & 3 rotates the four
//    independent rows per lane so that the four units of a quarter warp (slot = their
//    index in it) fall on four different bank groups whatever columns they lie in; the two
//    lanes of a unit (h) differ in the parity of the group.
template <bool kStore>
__device__ __forceinline__ void a_rows8(int *a, int h, int mu, int *v) {
#pragma unroll
  for (int g = 0; g < 4; g++) {
    int *row = a + (2*((g + mu) & 3) + h)*kAP;
#pragma unroll
    for (int j = 0; j < 2; j++) {
      int4 *p = reinterpret_cast<int4 *>(row + 4*j);
      int *w = v + 8*g + 4*j;
      if (kStore) *p = make_int4(w[0], w[1], w[2], w[3]);
      else { int4 t = *p; w[0] = t.x; w[1] = t.y; w[2] = t.z; w[3] = t.w; }
    }
  }
}

// -- 16x16 block, 8 lanes: lane k holds columns 2k, 2k+1 (v[r], v[16 + r]) ---
template <bool kStore>
__device__ __forceinline__ void a_cols16(int *a, int *v) {   // a -> (row 0, column 2k)
#pragma unroll
  for (int r = 0; r < 16; r++) {
    int2 *p = reinterpret_cast<int2 *>(a + r*kAP);
    if (kStore) *p = make_int2(v[r], v[16 + r]);
    else { int2 t = *p; v[r] = t.x; v[16 + r] = t.y; }
  }
}
// lane k holds rows k and k + 8 (v[0..15], v[16..31]); a -> the block's origin
template <bool kStore>
__device__ __forceinline__ void a_rows16(int *a, int k, int *v) {
#pragma unroll
  for (int g = 0; g < 2; g++) {
#pragma unroll
    for (int j = 0; j < 4; j++) {
      int4 *p = reinterpret_cast<int4 *>(a + (k + 8*g)*kAP + 4*j);
      int *w = v + 16*g + 4*j;
      if (kStore) *p = make_int4(w[0], w[1], w[2], w[3]);
      else { int4 t = *p; w[0] = t.x; w[1] = t.y; w[2] = t.z; w[3] = t.w; }
    }
  }
}

// -- 32x32 block, one warp: lane = column, then lane = row -------------------
template <bool kStore>
__device__ __forceinline__ void a_col32(int *a, int *v) {     // a -> (row 0, column lane)
#pragma unroll
  for (int r = 0; r < 32; r++) {
    if (kStore) a[r*kAP] = v[r]; else v[r] = a[r*kAP];
  }
}
template <bool kStore>
__device__ __forceinline__ void a_row32(int *a, int *v) {     // a -> (row lane, column 0)
#pragma unroll
  for (int j = 0; j < 8; j++) {
    int4 *p = reinterpret_cast<int4 *>(a + 4*j);
    if (kStore) *p = make_int4(v[4*j], v[4*j + 1], v[4*j + 2], v[4*j + 3]);
    else { int4 t = *p; v[4*j] = t.x; v[4*j + 1] = t.y; v[4*j + 2] = t.z; v[4*j + 3] = t.w; }
  }
}

// ---- keyframe DC pyramid on one tile (reference src/encode.c:1328-1341,
// src/decode.c:733).  At -> the tile's origin in A, ls0 -> unit table entry of the
// tile's first unit.
// The forward direction (bottom-up merge) of a whole tile in ONE step: lane j of a half warp
// takes unit j of the tile -- the DCs of its 4x4 blocks if it is split -- and the levels
// above travel by shuffle: the four lanes of a 16x16 quadrant exchange their unit DCs,
// the lanes 0, 2, 8, 10 the quadrant DCs (luma only: in 4:2:0 chroma the tile is four
// superblocks).  Every lane of a group computes the group's butterfly redundantly and
// keeps its own output.  Called by the whole warp.
__device__ __forceinline__ void region_dc_merge_tile(int *At, const unsigned char *ls0, int lane,
                                                     bool act, bool luma) {
  // lane = 0..31: the unit (lane & 15) of the tile of this half warp (At / ls0 / act may
  // differ between the halves: two tiles per warp); the shuffles stay inside the half
  const int hb = lane & 16;
  const int uy = (lane >> 2) & 3, ux = lane & 3;
  int *U = At + 8*uy*kAP + 8*ux;
  int vu = 0, L = 3, Lq = 3;
  if (act) {
    L = ls0[uy*kUS + ux];
    Lq = ls0[(uy & 2)*kUS + (ux & 2)];            // leaf size at the quadrant's origin
    vu = U[0];
    if (L == 0) {
      int c1 = U[4], c2 = U[4*kAP], c3 = U[4*kAP + 4];
      haar2x2(vu, c2, c1, c3);
      U[4] = c1; U[4*kAP] = c2; U[4*kAP + 4] = c3;
    }
  }
  {
    const int base = lane & ~5;
    int a0 = __shfl_sync(0xffffffffu, vu, base), a1 = __shfl_sync(0xffffffffu, vu, base + 1);
    int a2 = __shfl_sync(0xffffffffu, vu, base + 4), a3 = __shfl_sync(0xffffffffu, vu, base + 5);
    if (Lq < 2) {
      haar2x2(a0, a2, a1, a3);
      vu = (lane & 4) ? ((lane & 1) ? a3 : a2) : ((lane & 1) ? a1 : a0);
    }
  }
  if (luma) {
    const int L0 = __shfl_sync(0xffffffffu, L, hb);
    int q0 = __shfl_sync(0xffffffffu, vu, hb), q1 = __shfl_sync(0xffffffffu, vu, hb + 2);
    int q2 = __shfl_sync(0xffffffffu, vu, hb + 8), q3 = __shfl_sync(0xffffffffu, vu, hb + 10);
    if (L0 < 3 && !(lane & 5)) {
      haar2x2(q0, q2, q1, q3);
      vu = (lane & 8) ? ((lane & 2) ? q3 : q2) : ((lane & 2) ? q1 : q0);
    }
  }
  if (act) U[0] = vu;
  __syncwarp();
}

// The inverse direction (top-down split) of a whole tile in ONE step: lane j < 16 takes
// unit j of the tile and follows the chain superblock -> 16x16 quadrant -> its unit
// -> (if the unit is split) its 4x4 blocks in registers, from the values as they are
// before any split; lanes that share ancestors compute them redundantly.  Every lane
// reads first, then writes the DC(s) at positions inside its own unit: no ordering
// between levels through shared memory.  luma: the tile is one 32x32 superblock;
// chroma (4:2:0): four 16x16 superblocks.  lane = 0..15 (the unit), act = the tile is
// inside the plane; At / ls0 may differ between the two halves of a warp (two tiles per
// warp).  Called by the whole warp.
__device__ __forceinline__ void region_dc_split_tile(int *At, const unsigned char *ls0, int lane,
                                                     bool act, bool luma) {
  const int uy = (lane >> 2) & 3, ux = lane & 3, qy = uy >> 1, qx = ux >> 1;
  int *Q = At + 16*qy*kAP + 16*qx;            // my quadrant
  int *U = At + 8*uy*kAP + 8*ux;              // my unit
  int vq, b1, b2, b3, c1 = 0, c2 = 0, c3 = 0;
  const int L = act ? ls0[uy*kUS + ux] : 3;
  if (act) {
    vq = Q[0];
    if (luma && ls0[0] < 3) {
      int a0 = At[0], a1 = At[16], a2 = At[16*kAP], a3 = At[16*kAP + 16];
      haar2x2(a0, a1, a2, a3);
      vq = qy ? (qx ? a3 : a2) : (qx ? a1 : a0);
    }
    b1 = Q[8]; b2 = Q[8*kAP]; b3 = Q[8*kAP + 8];
    if (L == 0) { c1 = U[4]; c2 = U[4*kAP]; c3 = U[4*kAP + 4]; }
  }
  __syncwarp();
  if (act) {
    int b0 = vq;
    if (ls0[2*qy*kUS + 2*qx] < 2) haar2x2(b0, b1, b2, b3);
    int vu = (uy & 1) ? ((ux & 1) ? b3 : b2) : ((ux & 1) ? b1 : b0);
    if (L == 0) {
      haar2x2(vu, c1, c2, c3);
      U[4] = c1; U[4*kAP] = c2; U[4*kAP + 4] = c3;
    }
    U[0] = vu;
  }
  __syncwarp();
}

// ---- lossless mode: the multi-level Haar wavelet of each superblock of a tile, in
// place (reference src/dct.c:1960-2026).  At -> the tile's origin in A.  Every level
// reads its 2x2 groups into registers before it writes anything.
template <bool kInverse>
__device__ __forceinline__ void region_haar_tile(int *At, int lane, int sb) {
  const int ln = (sb == 32) ? 5 : 4;
  const int nb = 32/sb;                   // superblocks per tile side
#pragma unroll 1
  for (int step = 0; step < ln; step++) {
    const int level = kInverse ? ln - 1 - step : step;
    const int np = sb >> level >> 1;      // half the side of this level's input
    const int items = nb*nb*np*np;
    int a[8], b[8], c[8], d[8];
#pragma unroll
    for (int k = 0; k < 8; k++) {
      int it = lane + 32*k;
      if (it < items) {
        int blk = it/(np*np), e = it - blk*np*np;
        int i = e/np, j = e - i*np;
        int *q = At + (blk/nb)*sb*kAP + (blk%nb)*sb;
        if (!kInverse) {
          a[k] = q[2*i*kAP + 2*j];       b[k] = q[(2*i + 1)*kAP + 2*j];
          c[k] = q[2*i*kAP + 2*j + 1];   d[k] = q[(2*i + 1)*kAP + 2*j + 1];
        }
        else {
          a[k] = q[i*kAP + j];           b[k] = q[i*kAP + j + np];
          c[k] = q[(i + np)*kAP + j];    d[k] = q[(i + np)*kAP + j + np];
        }
      }
    }
    __syncwarp();
#pragma unroll
    for (int k = 0; k < 8; k++) {
      int it = lane + 32*k;
      if (it < items) {
        int blk = it/(np*np), e = it - blk*np*np;
        int i = e/np, j = e - i*np;
        int *q = At + (blk/nb)*sb*kAP + (blk%nb)*sb;
        haar2x2(a[k], b[k], c[k], d[k]);
        if (!kInverse) {
          q[i*kAP + j] = a[k];           q[i*kAP + j + np] = b[k];
          q[(i + np)*kAP + j] = c[k];    q[(i + np)*kAP + j + np] = d[k];
        }
        else {
          q[2*i*kAP + 2*j] = a[k];       q[(2*i + 1)*kAP + 2*j] = b[k];
          q[2*i*kAP + 2*j + 1] = c[k];   q[(2*i + 1)*kAP + 2*j + 1] = d[k];
        }
      }
    }
    __syncwarp();
  }
}

}  // namespace tdlt

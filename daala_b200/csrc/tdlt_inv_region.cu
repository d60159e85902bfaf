// Inverse TDLT over a batch of frames: d -> c.
//   keyframe Haar-DC split (reference src/decode.c:697-738), leaf iDCTs
//   (src/dct.c), split postfilters bottom-up (src/filter.c:1512-1554) and the
//   SB-edge postfilter (:1588-1646); lossless mode: od_haar_inv (src/dct.c:1999).
// One CTA per 128x64 region of one plane (tdlt_region.cuh); everything happens in
// place in one shared-memory buffer.
//
// The post-filter of a block edge needs reconstructed samples from both sides
// of it, so a region cannot finish alone: region (rx, ry) owns the OUTPUT
// rectangle shifted by (-2, -2) -- plane rows Y0-2..Y0+61, columns
// X0-2..X0+125, a whole number of lapping cells -- and needs the two bottom
// rows / right columns of raw iDCT output of its upper / left / upper-left
// neighbours.  Every region publishes those strips to global memory as 8-byte
// (sample, epoch) pairs -- one aligned 64-bit scalar store each, which is single-copy
// atomic, so a reader that sees the epoch of this launch has the sample too: no flag,
// no fence --
// and consumers take the pairs of regions that precede them in a TICKET order
// (CTAs draw tickets in the order they start running, so a region only ever
// waits on work that is already resident: no deadlock).  The tickets walk the
// batch frame-interleaved, a few skewed rows of regions of every frame at a
// time, so that a region's producers are hundreds of tickets ahead and their
// strips are there when it starts (see the ticket -> region mapping below).
//
// Phases: ticket, TMA box + plan record, ring fetch | DC pyramid top-down |
// inverse transforms in place (class-sorted warp tasks) | publish | un-lapping of
// the levels 8 .. superblock in one pass (region_lap_pass) | store of the owned rectangle.
#include "tdlt_tasks.cuh"
#include "tdlt_region.cuh"
#include "tdlt_kernels.h"

namespace tdlt {

constexpr int kStripInts = 2*kRegW + 2*kRegH;   // bottom rows 62, 63; right columns 126, 127

// One (sample, epoch) pair travels as ONE 64-bit scalar access: a naturally aligned
// b64 access is single-copy atomic in the PTX memory model (a v2.b32 access is formally two).
__device__ __forceinline__ int2 ld_pair(const int2 *p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.b64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return make_int2((int)(unsigned)v, (int)(unsigned)(v >> 32));
}
__device__ __forceinline__ void st_pair(int2 *p, int sample, int epoch) {
  const unsigned long long v = (unsigned long long)(unsigned)sample | ((unsigned long long)(unsigned)epoch << 32);
  asm volatile("st.relaxed.gpu.global.b64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}

// The two samples of the ring this thread is responsible for (threads 0..193):
// src -> two (sample, epoch) pairs of a neighbour's strip, dst -> their place in A.
struct RingSlot {
  const int2 *src;
  int2 *dst;
  bool need;
};
__device__ __forceinline__ RingSlot ring_slot(int *A, const int2 *strips, int t, int nrx,
                                              bool has_l, bool has_u, int tid) {
  RingSlot r;
  r.src = nullptr; r.dst = nullptr; r.need = false;
  if (tid < 128) {                    // rows -2, -1 <- bottom rows of the region above
    int a = tid >> 6, c = 2*(tid & 63);
    r.src = strips + (size_t)(t - nrx)*kStripInts + a*kRegW + c;
    r.dst = reinterpret_cast<int2 *>(A + a*kAP + c + 4);
    r.need = has_u;
  }
  else if (tid < 192) {               // columns -2, -1 <- right columns of the left region
    int row = tid - 128;
    r.src = strips + (size_t)(t - 1)*kStripInts + 2*kRegW + 2*row;
    r.dst = reinterpret_cast<int2 *>(A + (row + 2)*kAP + 2);
    r.need = has_l;
  }
  else if (tid < 194) {               // the corner
    int a = tid - 192;
    r.src = strips + (size_t)(t - nrx - 1)*kStripInts + a*kRegW + 126;
    r.dst = reinterpret_cast<int2 *>(A + a*kAP + 2);
    r.need = has_u && has_l;
  }
  return r;
}
// One look; true (and the samples stored into A) when both pairs carry this epoch.
__device__ __forceinline__ bool ring_try(const RingSlot &r, int epoch) {
  int2 p0 = ld_pair(r.src), p1 = ld_pair(r.src + 1);
  bool ok = p0.y == epoch && p1.y == epoch;
  if (ok) *r.dst = make_int2(p0.x, p1.x);
  return ok;
}

#ifdef TDLT_RINGSTAT
// measurement build only (tools/ablate.py): how often the ring is not there yet
__device__ unsigned long long g_ringstat[8];
#define RINGSTAT(i, n) atomicAdd(&g_ringstat[i], (unsigned long long)(n))
#else
#define RINGSTAT(i, n) ((void)0)
#endif

template <bool kHaar, int kPix, bool kCoding>
__device__ __forceinline__ void tdlt_inv_region_body(const TmapSet &maps,
                       const RegionBatch &bd,
                       const unsigned char *__restrict__ plan, PlanePtrs out,
                       int2 *__restrict__ strips, int *__restrict__ ticket, int epoch,
                       const unsigned short *__restrict__ order, PlanePtrs in) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const RegionSmem s = carve_region(smem);
  if (tid == 0) {
    s.dyn[0] = (TDLT_ABLATE & 64) ? (int)blockIdx.x : atomicAdd(ticket, 1);
    s.dyn[1] = 0;                             // task draw counter
    mbar_init(s.bar, 1);
    fence_mbar_init();
  }
  __syncthreads();
  const int t = s.dyn[0];
  TDLT_MARK(t, 0);
  // Ticket -> region.  Consecutive tickets go to the SAME region position of consecutive
  // frames, and the positions of a frame are taken in the order of order[] (built by the
  // host, tdlt_capi.cu: a few rows of regions at a time, skewed): the regions a region
  // waits for -- left, above, above-left -- drew their tickets at least a few times
  // nframes earlier, so their strips are there when it starts and nobody polls, while
  // the regions in flight still touch only a few rows of every frame.  (With a small
  // batch some may poll: the dependencies still have smaller tickets, which is all the
  // protocol needs.)
  // t / nframes by a float estimate and one correction step (exact: the estimate is
  // within 1 of the quotient for every t < 2^31)
  int k = __float2int_rz(__int2float_rz(t)*bd.inv_nframes);
  int frame = t - k*bd.nframes;
  if (frame < 0) { k--; frame += bd.nframes; }
  else if (frame >= bd.nframes) { k++; frame -= bd.nframes; }
  const int rloc = (TDLT_ABLATE & 512) ? k : (int)__ldg(order + k);
  const int lin = frame*bd.regs_per_frame + rloc;     // index of the plan record / strip
  const RegionId id = decode_region(bd, frame, rloc);
  const RegionPlane &pd = bd.pl[id.pli];
  const int X0 = id.rx*kRegW, Y0 = id.ry*kRegH;
  constexpr bool haar = kHaar;      // lossless mode is a kernel of its own (less code in either)
  const int sb = 32 >> pd.dec;

  // ---- the region's coefficients: the same TMA box as the forward kernel's (the halo
  // part of it is neighbours' coefficients, of no use here: the ring of raw samples
  // replaces it below), and the region's plan record (tdlt_plan.cu), on one mbarrier
  constexpr bool coding = kCoding;      // a kernel of its own: the raster kernels carry none of it
  if (tid == 0) {
    if (coding) mbar_expect_tx(s.bar, kPlanBytes);      // the tiles are gathered below
    else {
      mbar_expect_tx(s.bar, kARegBytes + kPlanBytes);
      tma_load_3d(s.A, &maps.in[id.pli], s.bar, X0 - 4, Y0 - 2, id.frame);
    }
    bulk_g2s(s.ls, plan + (size_t)lin*kPlanBytes, kPlanBytes, s.bar);
  }
  // ---- the ring of raw samples from the regions above / left / above-left: one
  // int2 per thread.  With the ticket order above they finished long ago as a rule:
  // the ring is read now and put into the halo of A once the box has landed; what is
  // not there yet (small batches) is waited for after the transforms.
  bool ring_done;
  int2 ring_val = make_int2(0, 0);
  RingSlot rs = ring_slot(s.A, strips, lin, pd.nrx, id.rx > 0, id.ry > 0, tid);
  bool ring_early = false;
  {
    ring_done = !rs.need || (TDLT_ABLATE & 32) || ((TDLT_ABLATE & 2048) && tid >= 128 && tid < 192);
    if (!ring_done) {
      int2 p0 = ld_pair(rs.src), p1 = ld_pair(rs.src + 1);
      ring_early = p0.y == epoch && p1.y == epoch;
      ring_val = make_int2(p0.x, p1.x);
      ring_done = ring_early;
      RINGSTAT(1, 1);
      if (!ring_early) RINGSTAT(2, 1);
    }
  }
  TDLT_MARK(t, 1);
  mbar_wait(s.bar, 0);
  if (ring_early) *rs.dst = ring_val;
  TDLT_MARK(t, 2);

  const int tx = warp & 3, ty = warp >> 2;
  const bool tile_on = X0 + 32*tx < pd.w && Y0 + 32*ty < pd.h;
  int *At = s.A + (32*ty + 2)*kAP + 32*tx + 4;      // warp w <-> tile w
  if (coding) {
    // the region's coefficients arrive as coding-order vectors (block-linear plane): every
    // warp scatters its tile to its raster place in A
    if (tile_on) {
      const int lS = 5 - pd.dec, sbw = pd.w >> lS;
      int *gtile = in.p[id.pli] + (size_t)id.frame*pd.w*pd.h +
                   ((size_t)((Y0 + 32*ty) >> lS)*sbw + ((X0 + 32*tx) >> lS))*(size_t)(1 << 2*lS);
      region_load_tile_coding(At, s.ls + ut(4*ty, 4*tx), bd.cotab, pd.dec, gtile, sbw << 2*lS, lane);
    }
    __syncthreads();
  }

  if (haar) {
    // No lapping in lossless mode: every tile is stored where it lies.
    if (tile_on) {
      region_haar_tile<true>(At, lane, sb);
      int *g = out.p[id.pli] + ((size_t)id.frame*pd.h + Y0 + 32*ty)*pd.w + X0 + 32*tx;
      int *gq = g + (size_t)(lane >> 3)*pd.w + 4*(lane & 7);
      const int *aq = At + (lane >> 3)*kAP + 4*(lane & 7);
#pragma unroll
      for (int it = 0; it < 8; it++) {      // 4 rows of 128 bytes per warp instruction
        *reinterpret_cast<int4 *>(gq + (size_t)(4*it)*pd.w) = *reinterpret_cast<const int4 *>(aq + 4*it*kAP);
      }
    }
    return;
  }

  // ---- keyframe: undo the DC pyramid, top-down --------------------------------
  // (16 lanes per tile: the warps 0..3 take two tiles each, half the instructions of one
  // tile per warp)
  if ((bd.flags & 1) && warp < kRegTiles/2 && !(TDLT_ABLATE & 4)) {
    const int tl = 2*warp + (lane >> 4), ttx = tl & 3, tty = tl >> 2;
    const bool on = X0 + 32*ttx < pd.w && Y0 + 32*tty < pd.h;
    region_dc_split_tile(s.A + (32*tty + 2)*kAP + 32*ttx + 4, s.ls + ut(4*tty, 4*ttx), lane & 15,
                         on, pd.dec == 0);
  }
  __syncthreads();
  TDLT_MARK(t, 3);

  // ---- inverse transforms, in place: rows first, then columns (src/dct.c:144-149)
  if (!(TDLT_ABLATE & 2)) {
    const int n0 = s.cnt[4], n1 = s.cnt[5], n2 = s.cnt[6], n3 = s.cnt[7];
    const int q0 = (n0 + 15) >> 4, q1 = (n1 + 15) >> 4, q2 = (n2 + 3) >> 2;
    const int ntask = q0 + q1 + q2 + n3;
    int v[32];
#pragma unroll 1
    for (;;) {
      // longest tasks first; a warp that is done draws the next one
      int g = 0;
      if (lane == 0) g = atomicAdd(s.dyn + 1, 1);
      g = __shfl_sync(0xffffffffu, g, 0);
      if (g >= ntask) break;
      if (g < n3 && !(TDLT_ABLATE & 256)) {
        int e = s.tasks[3*kTaskListStride + g];
        int *blk = s.A + (8*(e >> 4) + 2)*kAP + 8*(e & 15) + 4;
#pragma unroll 1
        for (int pass = 0; pass < 2; pass++) {
          int *line = pass == 0 ? blk + lane*kAP : blk + lane;
          if (pass == 0) a_row32<false>(line, v); else a_col32<false>(line, v);
          tdlt_idct32(v);
          if (pass == 0) a_row32<true>(line, v); else a_col32<true>(line, v);
          __syncwarp();
        }
      }
      else if (g < n3 + q2) {
        int idx = (g - n3)*4 + (lane >> 3), k = lane & 7;
        bool on = idx < n2;
        int e = s.tasks[2*kTaskListStride + (on ? idx : 0)];
        int *blk = s.A + (8*(e >> 4) + 2)*kAP + 8*(e & 15) + 4;
#pragma unroll 1
        for (int pass = 0; pass < 2; pass++) {
          if (on) {
            if (pass == 0) a_rows16<false>(blk, k, v); else a_cols16<false>(blk + 2*k, v);
            tdlt_idct16(v);
            tdlt_idct16(v + 16);
            if (pass == 0) a_rows16<true>(blk, k, v); else a_cols16<true>(blk + 2*k, v);
          }
          __syncwarp();
        }
      }
      else if (g < n3 + q2 + q1) {
        int idx = (g - n3 - q2)*16 + (lane >> 1), h = lane & 1;
        bool on = idx < n1;
        int e = s.tasks[1*kTaskListStride + (on ? idx : 0)];
        int *blk = s.A + (8*(e >> 4) + 2)*kAP + 8*(e & 15) + 4;
        int mu = ((e & 15) - (lane >> 1)) & 3;
#pragma unroll 1
        for (int pass = 0; pass < 2; pass++) {
          if (on) {
            if (pass == 0) a_rows8<false>(blk, h, mu, v); else ld_a_8x4(blk + 4*h, v);
#pragma unroll
            for (int q = 0; q < 4; q++) tdlt_idct8(v + 8*q);
            if (pass == 0) a_rows8<true>(blk, h, mu, v); else st_a_8x4(blk + 4*h, v);
          }
          __syncwarp();
        }
      }
      else {
        // 4x4 blocks: inverse transforms, then the unit's own mid-edges un-lapped
        // in registers
        int idx = (g - n3 - q2 - q1)*16 + (lane >> 1), h = lane & 1;
        bool on = idx < n0;
        int e = s.tasks[on ? idx : 0];
        int UR = e >> 4, UC = e & 15;
        int *blk = s.A + (8*UR + 2)*kAP + 8*UC + 4 + 4*h;
        int bits = on ? s.ub[ut(UR, UC)] : 0;
        if (on) {
          ld_a_8x4(blk, v);
          idct4x4_pair(v);
        }
        lap4_unit<true>(v, h, bits);
        if (on) st_a_8x4(blk, v);
      }
    }
  }
  // the ring samples still missing (the left neighbour's columns, as a rule) are asked
  // for now, so that the round trip to L2 overlaps the barrier and the publication
  int2 pf0 = make_int2(0, 0), pf1 = make_int2(0, 0);
  if (!ring_done) {
    pf0 = ld_pair(rs.src);
    pf1 = ld_pair(rs.src + 1);
  }
  __syncthreads();
  TDLT_MARK(t, 4);

  // ---- publish my bottom rows / right columns (raw iDCT output) ---------------
  if (tid < 192) {
    int2 *mine = strips + (size_t)lin*kStripInts;
    int2 val;
    int at;
    if (tid < 128) {          // rows 62, 63: 2 x 128 samples
      int a = tid >> 6, c = 2*(tid & 63);
      val = *reinterpret_cast<int2 *>(s.A + (62 + a + 2)*kAP + c + 4);
      at = a*kRegW + c;
    }
    else {                    // columns 126, 127 of rows 0..63
      int r = tid - 128;
      val = *reinterpret_cast<int2 *>(s.A + (r + 2)*kAP + 126 + 4);
      at = 2*kRegW + 2*r;
    }
    if (!(TDLT_ABLATE & 32)) {
      st_pair(mine + at, val.x, epoch);
      st_pair(mine + at + 1, val.y, epoch);
    }
  }
  TDLT_MARK(t, 5);

  // ---- the ring: whatever was not there yet when the region started --------------
  if (!ring_done) {
    if (pf0.y == epoch && pf1.y == epoch) *rs.dst = make_int2(pf0.x, pf1.x);
    else if (!(TDLT_ABLATE & 1024)) {
      RINGSTAT(3, 1);
      while (!ring_try(rs, epoch)) { RINGSTAT(4, 1); __nanosleep(20); }
    }
  }
  __syncthreads();      // the ring is in A, and the strips were read before anyone laps them
  TDLT_MARK(t, 6);

  // ---- un-lap (levels 8 .. superblock) in one pass, in place (region_lap_pass), then the
  // owned rectangle goes out row by row (region_store_owned).  buffer row b <-> plane row
  // Y0 - 2 + b.  Owned: buffer rows b0 .. b1-1 and region columns cs .. ce-1 (cs = -2, or 0
  // in the first region column; ce = 126, or where the plane ends).
  const bool last_r = Y0 + kRegH >= pd.h, last_c = X0 + kRegW >= pd.w;
  TDLT_MARK(t, 7);
  {
    const int b0 = (id.ry == 0) ? 2 : 0;
    const int b1 = last_r ? min(kRegH, pd.h - Y0) + 2 : kRegH;
    // (TDLT_ABLATE 4096: whole sectors only -- columns 0 .. 127, wrong at the region seams)
    const int cs = (id.rx > 0 && !(TDLT_ABLATE & 4096)) ? -2 : 0;
    const int ce = (last_c || (TDLT_ABLATE & 4096)) ? min(kRegW, pd.w - X0) : kRegW - 2;
    int *g = out.p[id.pli] + ((ptrdiff_t)id.frame*pd.h + Y0 - 2)*(ptrdiff_t)pd.w + (X0 - 2);
    if (!(TDLT_ABLATE & 1)) region_lap_pass<true>(s.A, s.ub, s.msk, last_r ? kCellsFwd : kCellsInv, warp, lane);
    if (TDLT_BULK_STORE_INV && !kPix) fence_proxy_async_smem();   // the copy engine reads A next
    __syncthreads();
    if (kPix) {
      // od_coeff_to_ref_buf fused into the store: out.p[] are planes of kPix-byte samples
      unsigned char *gp = reinterpret_cast<unsigned char *>(out.p[id.pli]) +
                          (((ptrdiff_t)id.frame*pd.h + Y0 - 2)*(ptrdiff_t)pd.w + (X0 - 2))*kPix;
      region_store_owned_pix<(kPix ? kPix : 1)>(s.A, gp, (size_t)pd.w*kPix, b0, b1, cs, ce, bd.pix_shift, warp, lane);
    }
    else if (TDLT_ABLATE & 128) { }
    else if (TDLT_BULK_STORE_INV == 5) region_store_owned_mixed(s.A, g, (size_t)pd.w, b0, b1, cs, ce, warp, lane);
    else if (TDLT_BULK_STORE_INV) region_store_owned_bulk(s.A, g, (size_t)pd.w, b0, b1, cs, ce, tid);
    else region_store_owned(s.A, g, (size_t)pd.w, b0, b1, cs, ce, warp, lane);
  }
  TDLT_MARK(t, 8);
}

__global__ void __launch_bounds__(kRegThreads, TDLT_MINB)
tdlt_inv_region_kernel(const __grid_constant__ TmapSet maps,
                       const __grid_constant__ RegionBatch bd,
                       const unsigned char *__restrict__ plan, PlanePtrs out,
                       int2 *__restrict__ strips, int *__restrict__ ticket, int epoch,
                       const unsigned short *__restrict__ order, PlanePtrs in) {
  tdlt_inv_region_body<false, 0, false>(maps, bd, plan, out, strips, ticket, epoch, order, in);
}

// The same with od_coeff_to_ref_buf fused into the store: out.p[] are planes of 8-bit (pix8)
// or 16-bit (pix16) reference-buffer samples.
__global__ void __launch_bounds__(kRegThreads, TDLT_MINB)
tdlt_inv_region_pix8_kernel(const __grid_constant__ TmapSet maps,
                       const __grid_constant__ RegionBatch bd,
                       const unsigned char *__restrict__ plan, PlanePtrs out,
                       int2 *__restrict__ strips, int *__restrict__ ticket, int epoch,
                       const unsigned short *__restrict__ order, PlanePtrs in) {
  tdlt_inv_region_body<false, 1, false>(maps, bd, plan, out, strips, ticket, epoch, order, in);
}
__global__ void __launch_bounds__(kRegThreads, TDLT_MINB)
tdlt_inv_region_pix16_kernel(const __grid_constant__ TmapSet maps,
                       const __grid_constant__ RegionBatch bd,
                       const unsigned char *__restrict__ plan, PlanePtrs out,
                       int2 *__restrict__ strips, int *__restrict__ ticket, int epoch,
                       const unsigned short *__restrict__ order, PlanePtrs in) {
  tdlt_inv_region_body<false, 2, false>(maps, bd, plan, out, strips, ticket, epoch, order, in);
}

// OD_TDLT_CODING_ORDER: the coefficients arrive as block-linear coding-order vectors (in.p[]).
#define TDLT_INV_CODING_KERNEL(name, pix) \
  __global__ void __launch_bounds__(kRegThreads, TDLT_MINB) \
  name(const __grid_constant__ TmapSet maps, const __grid_constant__ RegionBatch bd, \
       const unsigned char *__restrict__ plan, PlanePtrs out, int2 *__restrict__ strips, \
       int *__restrict__ ticket, int epoch, const unsigned short *__restrict__ order, PlanePtrs in) { \
    tdlt_inv_region_body<false, pix, true>(maps, bd, plan, out, strips, ticket, epoch, order, in); \
  }
TDLT_INV_CODING_KERNEL(tdlt_inv_region_co_kernel, 0)
TDLT_INV_CODING_KERNEL(tdlt_inv_region_co_pix8_kernel, 1)
TDLT_INV_CODING_KERNEL(tdlt_inv_region_co_pix16_kernel, 2)

// Lossless mode (OD_TDLT_HAAR): the Haar wavelet of every superblock instead of lapping + DCTs.
__global__ void __launch_bounds__(kRegThreads, TDLT_MINB)
tdlt_inv_haar_kernel(const __grid_constant__ TmapSet maps,
                       const __grid_constant__ RegionBatch bd,
                       const unsigned char *__restrict__ plan, PlanePtrs out,
                       int2 *__restrict__ strips, int *__restrict__ ticket, int epoch,
                       const unsigned short *__restrict__ order, PlanePtrs in) {
  tdlt_inv_region_body<true, 0, false>(maps, bd, plan, out, strips, ticket, epoch, order, in);
}


}  // namespace tdlt

#ifdef TDLT_RINGSTAT
extern "C" int od_tdlt_debug_ringstat(unsigned long long *out, int reset) {
  if (cudaMemcpyFromSymbol(out, tdlt::g_ringstat, sizeof(tdlt::g_ringstat)) != cudaSuccess) return -1;
  if (reset) {
    unsigned long long z[8] = {0};
    cudaMemcpyToSymbol(tdlt::g_ringstat, z, sizeof(z));
  }
  return 0;
}
#endif

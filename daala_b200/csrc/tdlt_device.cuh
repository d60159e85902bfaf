// Device-side building blocks of the TDLT kernels: the 4-tap lapping filters,
// the Haar butterfly and the mbarrier / TMA wrappers.
// Hand-written for sm_100a; no dependency beyond the CUDA runtime headers.
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>
#include "tdlt_lift_gen.cuh"

namespace tdlt {

// ---------------------------------------------------------------------------
// Geometry shared by host and device
// ---------------------------------------------------------------------------
constexpr int kTile = 32;                    // a tile is 32x32 samples of one plane
constexpr int kMaxPlanes = 4;

// ---------------------------------------------------------------------------
// 4-tap lapping filters (reference src/filter.c:174-249; parameters 85, 75,
// -15, 33 in 1/64ths, :164-167).  In place on four values.
// ---------------------------------------------------------------------------
__device__ __forceinline__ void pre4(int &x0, int &x1, int &x2, int &x3) {
  int t3 = x0 - x3;
  int t2 = x1 - x2;
  int t1 = x1 - (t2 >> 1);
  int t0 = x0 - (t3 >> 1);
  // (t*85 >> 6) + (that > 0): the product keeps the sign of t and is >= 64 whenever
  // t >= 1, so "> 0" is the sign bit of -85 t (one multiply-add pipe instruction and
  // one shift-add instead of compare / add / select)
  t2 = ((t2*85) >> 6) + (int)((unsigned)(t2*-85) >> 31);
  t3 = ((t3*75) >> 6) + (int)((unsigned)(t3*-75) >> 31);
  t3 += (t2*-15 + 32) >> 6;
  t2 += (t3*33 + 32) >> 6;
  t0 += t3 >> 1;
  t1 += t2 >> 1;
  x0 = t0;
  x1 = t1;
  x2 = t1 - t2;
  x3 = t0 - t3;
}

constexpr int kPostM75 = (int)(3665038760u);   // ceil(2^38/75) - 2^32
constexpr int kPostM85 = (int)(3233857729u);   // ceil(2^38/85) - 2^32
__device__ __forceinline__ void post4(int &y0, int &y1, int &y2, int &y3) {
  int t3 = y0 - y3;
  int t2 = y1 - y2;
  int t1 = y1 - (t2 >> 1);
  int t0 = y0 - (t3 >> 1);
  // x -= (a + 32) >> 6  ==  x += (-a + 31) >> 6  (one shift-add instead of shift, subtract)
  t2 += (t3*-33 + 31) >> 6;
  t3 += (t2*15 + 31) >> 6;
  // (t*64)/75 and (t*64)/85, C division (truncating), as one multiply-high each:
  // trunc(t*64/d) = floor(t*M/2^32) + (t < 0) with M = ceil(2^38/d), exact for
  // |t| < 2^25, the range in which the reference's own t*64 does not overflow
  // (checked exhaustively on the host, tests/test_cabi_and_host.py::test_post_filter_division_is_exact).
  t3 = __mulhi(t3, kPostM75) + t3 + (int)((unsigned)t3 >> 31);
  t2 = __mulhi(t2, kPostM85) + t2 + (int)((unsigned)t2 >> 31);
  t0 += t3 >> 1;
  t1 += t2 >> 1;
  y0 = t0;
  y1 = t1;
  y2 = t1 - t2;
  y3 = t0 - t3;
}

// Reversible 2x2 Haar butterfly (reference src/tf.h:34-45).
__device__ __forceinline__ void haar2x2(int &ll, int &lh, int &hl, int &hh) {
  ll += hl;
  hh -= lh;
  int m = (ll - hh) >> 1;
  lh = m - lh;
  hl = m - hl;
  ll -= lh;
  hh += hl;
}

// ---------------------------------------------------------------------------
// mbarrier / TMA (cp.async.bulk.tensor) wrappers
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(void *bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(void *bar, int bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
               :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(void *bar, int parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" :: "r"(smem_u32(bar)), "r"(parity) : "memory");
}
// 1-D bulk copy global -> shared (16-byte aligned, size a multiple of 16),
// completing on `bar`.
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, int bytes, void *bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
      :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// 1-D bulk copy shared -> global (16-byte aligned, size a multiple of 16) in the calling
// thread's bulk group; the thread must wait for its group before the CTA's shared memory
// may go away.  Shared memory written with ordinary stores needs fence_proxy_async_smem()
// by the writers (and a barrier) before the copy engine reads it.
__device__ __forceinline__ void bulk_s2g(void *dst, const void *src, int bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
               :: "l"(dst), "r"(smem_u32(src)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_s2g_addr(unsigned long long dst, unsigned src, int bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
               :: "l"(dst), "r"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit_wait_read() {
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
// One lane of the (converged) warp: true for the elected lane only.
__device__ __forceinline__ bool elect_one() {
  unsigned pred;
  asm volatile("{\n.reg .pred P;\nelect.sync _|P, 0xffffffff;\nselp.u32 %0, 1, 0, P;\n}" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
// 3-D tiled load: coordinates (x, y, frame), innermost first.
__device__ __forceinline__ void tma_load_3d(void *dst, const void *tmap, void *bar,
                                            int x, int y, int z) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%2, %3, %4}], [%5];"
      :: "r"(smem_u32(dst)), "l"(tmap), "r"(x), "r"(y), "r"(z), "r"(smem_u32(bar))
      : "memory");
}
}  // namespace tdlt

// Debug probe: isolates TMA load / store / mbarrier pieces of the TDLT kernels.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <cuda.h>
#include <cuda_runtime.h>
#include "../../daala_b200/csrc/tdlt_device.cuh"
using namespace tdlt;

typedef CUresult (*enc_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

struct Maps { CUtensorMap in, out; };

__global__ void k_load(const __grid_constant__ Maps m, int *dst, int x, int y, int stage, int nbytes) {
  extern __shared__ __align__(1024) unsigned char smem[];
  int *B = (int *)smem;
  int *A = (int *)(smem + 4096);
  unsigned long long *bar = (unsigned long long *)(smem + 4096 + 5248);
  int lane = threadIdx.x;
  if (lane == 0) {
    mbar_init(bar, 1);
    if (stage >= 1) fence_mbar_init();
    if (stage >= 2) mbar_expect_tx(bar, nbytes);
    if (stage >= 3) tma_load_3d(A, &m.in, bar, x, y, 0);
  }
  __syncwarp();
  if (stage >= 4) mbar_wait(bar, 0);
  if (stage >= 5) {
    for (int i = lane; i < 36*36; i += 32) dst[i] = A[i];
  }
  if (stage >= 6) {
    for (int r = 0; r < 32; r++) B[b_off(r, lane)] = A[a_off(r + 2, lane + 2)];
    fence_proxy_async_smem();
    __syncwarp();
    if (lane == 0) {
      tma_store_3d(&m.out, B, x + 2, y + 2, 0);
      tma_store_commit_and_wait_read();
    }
  }
}

int main(int argc, char **argv) {
  int bx = argc > 1 ? atoi(argv[1]) : 36; int rank = argc > 2 ? atoi(argv[2]) : 3; int x0 = argc > 3 ? atoi(argv[3]) : 30;
  const int W = 128, H = 96;
  std::vector<int> h(W*H);
  for (int i = 0; i < W*H; i++) h[i] = i;
  int *din, *dout, *ddst;
  cudaMalloc(&din, W*H*4); cudaMalloc(&dout, W*H*4); cudaMalloc(&ddst, 36*36*4);
  cudaMemcpy(din, h.data(), W*H*4, cudaMemcpyHostToDevice);
  cudaMemset(dout, 0, W*H*4);
  void *p = nullptr; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
  enc_fn enc = (enc_fn)p;
  Maps m; memset(&m, 0, sizeof(m));
  cuuint64_t dims[3] = {W, H, 1}; cuuint64_t str[2] = {W*4, (cuuint64_t)W*H*4};
  cuuint32_t box1[3] = {(cuuint32_t)bx, 36, 1}, box2[3] = {32, 32, 1}, es[3] = {1, 1, 1};
  CUresult r1 = enc(&m.in, CU_TENSOR_MAP_DATA_TYPE_INT32, rank, din, dims, str, box1, es,
    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  CUresult r2 = enc(&m.out, CU_TENSOR_MAP_DATA_TYPE_INT32, 3, dout, dims, str, box2, es,
    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  printf("encode: %d %d\n", (int)r1, (int)r2);
  for (int stage = 0; stage <= 6; stage++) {
    k_load<<<1, 32, 16384>>>(m, ddst, x0, 30, stage, bx*36*4);
    cudaError_t e = cudaDeviceSynchronize();
    printf("stage %d: %s\n", stage, cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
  }
  std::vector<int> a(36*36), o(W*H);
  cudaMemcpy(a.data(), ddst, 36*36*4, cudaMemcpyDeviceToHost);
  cudaMemcpy(o.data(), dout, W*H*4, cudaMemcpyDeviceToHost);
  printf("A[0]=%d (want %d) A[37]=%d (want %d)\n", a[0], 30*W + 30, a[37], 31*W + 31);
  int bad = 0;
  for (int r = 0; r < 32; r++) for (int c = 0; c < 32; c++) bad += o[(32 + r)*W + 32 + c] != h[(32 + r)*W + 32 + c];
  printf("store mismatches: %d\n", bad);
  // negative coordinates
  k_load<<<1, 32, 16384>>>(m, ddst, -2, -2, 5, bx*36*4);
  printf("neg coords: %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  cudaMemcpy(a.data(), ddst, 36*36*4, cudaMemcpyDeviceToHost);
  printf("A[0]=%d (want 0) A[2*36+2]=%d (want 0) A[2*36+3]=%d (want 1)\n", a[0], a[74], a[75]);
  return 0;
}

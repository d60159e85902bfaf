// Probe: TMA tiled load with elementStrides (4, 4): what lands in shared memory?
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
__global__ void k(const __grid_constant__ CUtensorMap m, int *dst, int x, int y, int nbytes, int nwords) {
  extern __shared__ __align__(1024) unsigned char smem[];
  int *A = (int *)smem;
  unsigned long long *bar = (unsigned long long *)(smem + 40960);
  for (int i = threadIdx.x; i < 10240; i += blockDim.x) A[i] = -777;
  __syncthreads();
  if (threadIdx.x == 0) { mbar_init(bar, 1); fence_mbar_init(); mbar_expect_tx(bar, nbytes); tma_load_3d(A, &m, bar, x, y, 0); }
  __syncthreads();
  for (int i = 0; i < 2000; i++) __nanosleep(1000);      // no wait on the barrier: just look
  __syncthreads();
  for (int i = threadIdx.x; i < nwords; i += blockDim.x) dst[i] = A[i];
}
int main(int argc, char **argv) {
  int bx = atoi(argv[1]), by = atoi(argv[2]), es = atoi(argv[3]), nbytes = atoi(argv[4]);
  const int W = 256, H = 128;
  std::vector<int> h(W*H);
  for (int i = 0; i < W*H; i++) h[i] = i;
  int *din, *ddst; cudaMalloc(&din, W*H*4); cudaMalloc(&ddst, 10240*4);
  cudaMemcpy(din, h.data(), W*H*4, cudaMemcpyHostToDevice);
  void *p = nullptr; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
  CUtensorMap m; memset(&m, 0, sizeof(m));
  cuuint64_t dims[3] = {W, H, 1}; cuuint64_t str[2] = {W*4, (cuuint64_t)W*H*4};
  cuuint32_t box[3] = {(cuuint32_t)bx, (cuuint32_t)by, 1}, e[3] = {(cuuint32_t)es, (cuuint32_t)es, 1};
  CUresult r = ((enc_fn)p)(&m, CU_TENSOR_MAP_DATA_TYPE_INT32, 3, din, dims, str, box, e, CU_TENSOR_MAP_INTERLEAVE_NONE,
    CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  printf("encode box %dx%d stride %d: %d\n", bx, by, es, (int)r); fflush(stdout);
  if (r) return 1;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 41984);
  k<<<1, 128, 41984>>>(m, ddst, 128, 64, nbytes, 10240);
  cudaError_t err = cudaDeviceSynchronize();
  printf("run: %s\n", cudaGetErrorString(err)); fflush(stdout);
  if (err) return 1;
  std::vector<int> a(10240); cudaMemcpy(a.data(), ddst, 10240*4, cudaMemcpyDeviceToHost);
  int n = 0; for (int i = 0; i < 10240; i++) n += a[i] != -777;
  int last = -1; for (int i = 0; i < 10240; i++) if (a[i] != -777) last = i;
  printf("words written: %d (last index %d); first 6: %d %d %d %d %d %d; word 32: %d word 128: %d (row 64 col 128 = %d, row 68 = %d)\n", n, last, a[0], a[1], a[2], a[3], a[4], a[5], a[32], a[128], 64*W + 128, 68*W + 128);
  return 0;
}

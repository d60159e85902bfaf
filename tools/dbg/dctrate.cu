// How fast can the SM issue the lifting code itself?  Registers only, no memory, no barriers:
// 256 threads x 3 CTAs per SM, each thread runs fdct/idct of one size on v[32] in a loop.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I daala_b200/csrc -o tools/dbg/dctrate tools/dbg/dctrate.cu
#include <cstdio>
#include <cuda_runtime.h>
#include "tdlt_lift_gen.cuh"

template <int MODE>
__global__ void __launch_bounds__(256, 3) k(int *out, int n, long long *cyc) {
  int v[32];
#pragma unroll
  for (int i = 0; i < 32; i++) v[i] = (threadIdx.x*31 + i*7) & 1023;
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < n; it++) {
    if (MODE == 8) {
#pragma unroll
      for (int q = 0; q < 4; q++) tdlt_fdct8(v + 8*q);
#pragma unroll
      for (int q = 0; q < 4; q++) tdlt_idct8(v + 8*q);
    }
    if (MODE == 16) { tdlt_fdct16(v); tdlt_fdct16(v + 16); tdlt_idct16(v); tdlt_idct16(v + 16); }
    if (MODE == 32) { tdlt_fdct32(v); tdlt_idct32(v); }
    if (MODE == 4) {
#pragma unroll
      for (int q = 0; q < 8; q++) tdlt_fdct4(v + 4*q);
#pragma unroll
      for (int q = 0; q < 8; q++) tdlt_idct4(v + 4*q);
    }
  }
  long long t1 = clock64();
  int s = 0;
#pragma unroll
  for (int i = 0; i < 32; i++) s ^= v[i];
  out[blockIdx.x*blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(const char *name) {
  int nsm; cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, 0);
  int grid = nsm*3, n = 2000;
  int *out; long long *cyc;
  cudaMalloc(&out, grid*256*4); cudaMalloc(&cyc, grid*8);
  k<MODE><<<grid, 256>>>(out, 10, cyc);
  k<MODE><<<grid, 256>>>(out, n, cyc);
  cudaDeviceSynchronize();
  long long *h = new long long[grid];
  cudaMemcpy(h, cyc, grid*8, cudaMemcpyDeviceToHost);
  double avg = 0; for (int i = 0; i < grid; i++) avg += h[i]; avg /= grid;
  printf("%-10s %10.0f cycles  %8.1f cycles per iteration (6 warps per scheduler)\n", name, avg, avg/n);
  cudaFree(out); cudaFree(cyc); delete[] h;
}

int main() {
  run<4>("dct4 x8");
  run<8>("dct8 x4");
  run<16>("dct16 x2");
  run<32>("dct32 x1");
  return 0;
}

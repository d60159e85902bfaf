// Issue-rate microbenchmark for the integer pipes of sm_100a (tools only).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/dbg/pipes tools/dbg/pipes.cu
// Each kernel runs N iterations of 16 independent operations per thread (8 chains x 2),
// 256 threads x 3 CTAs per SM; prints warp-instructions per cycle per SM sub-partition.
#include <cstdio>
#include <cuda_runtime.h>

#define CHAINS 8
template <int MODE>
__global__ void __launch_bounds__(256, 3) k(int *out, int n, int a, int b, long long *cyc) {
  int x[CHAINS];
#pragma unroll
  for (int i = 0; i < CHAINS; i++) x[i] = threadIdx.x + i;
  long long t0 = clock64();
  for (int it = 0; it < n; it++) {
#pragma unroll
    for (int rep = 0; rep < 4; rep++) {
#pragma unroll
      for (int i = 0; i < CHAINS; i++) {
        if (MODE == 0) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(a), "r"(b));
        if (MODE == 1) asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(a), "r"(b));
        if (MODE == 2) {
          if (i & 1) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(a), "r"(b));
          else asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(a), "r"(b));
        }
        if (MODE == 3) asm volatile("shf.r.clamp.b32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(a), "r"(b));
        if (MODE == 4) {   // one lifting step: t = x*K + R (IMAD); x = (t >> 13) + x (LEA.HI.SX32)
          int t;
          asm volatile("mad.lo.s32 %0, %1, %2, 4096;" : "=r"(t) : "r"(x[i]), "r"(a));
          asm volatile("{.reg .s32 u; shr.s32 u, %1, 13; add.s32 %0, %0, u;}" : "+r"(x[i]) : "r"(t));
        }
        if (MODE == 5) {   // two ALU ops per IMAD (the measured mix: ~40% alu, ~33% fma-pipe)
          int t;
          asm volatile("mad.lo.s32 %0, %1, %2, 4096;" : "=r"(t) : "r"(x[i]), "r"(a));
          asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(t), "r"(b));
          asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(a), "r"(b));
        }
      }
    }
  }
  long long t1 = clock64();
  int s = 0;
#pragma unroll
  for (int i = 0; i < CHAINS; i++) s ^= x[i];
  out[blockIdx.x*blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(const char *name, int per_iter) {
  int nsm; cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, 0);
  int grid = nsm*3, n = 4000;
  int *out; long long *cyc;
  cudaMalloc(&out, grid*256*4); cudaMalloc(&cyc, grid*8);
  k<MODE><<<grid, 256>>>(out, 10, 3, 5, cyc);
  k<MODE><<<grid, 256>>>(out, n, 3, 5, cyc);
  cudaDeviceSynchronize();
  long long *h = new long long[grid];
  cudaMemcpy(h, cyc, grid*8, cudaMemcpyDeviceToHost);
  double avg = 0; for (int i = 0; i < grid; i++) avg += h[i]; avg /= grid;
  // per SMSP: 3 CTAs x 8 warps / 4 = 6 warps, each n*per_iter instructions
  double ipc = 6.0*n*per_iter/avg;
  printf("%-34s %8.0f cycles  %.3f warp-instr/cycle/SMSP\n", name, avg, ipc);
  cudaFree(out); cudaFree(cyc); delete[] h;
}

int main() {
  run<0>("lop3 (alu pipe)", 32);
  run<1>("imad (fma pipe)", 32);
  run<2>("imad / lop3 alternating", 32);
  run<3>("shf (alu pipe)", 32);
  run<4>("lift step: imad + lea.hi.sx32", 64);
  run<5>("imad + 2 lop3", 96);
  return 0;
}

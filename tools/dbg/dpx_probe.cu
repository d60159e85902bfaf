// Throughput probe: packed 16-bit add / min-max instructions of sm_100a against plain
// 32-bit integer ones (the deringing taps could be written with either).
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dpx_probe dpx_probe.cu && ./dpx_probe
#include <cstdio>
#include <cuda_runtime.h>
template <int OP> __global__ void k(unsigned *out, unsigned a0, unsigned b0, int iters) {
  unsigned x[8];
#pragma unroll
  for (int i = 0; i < 8; i++) x[i] = a0 + threadIdx.x*(i + 1);
  unsigned b = b0, c = b0 ^ 0x00050003;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int r = 0; r < 8; r++) {
#pragma unroll
      for (int i = 0; i < 8; i++) {
        if (OP == 0) x[i] = x[i] + b + c;                                   // IADD3
        if (OP == 1) x[i] = __viaddmax_s16x2(x[i], b, c);                   // VIADDMNMX.S16x2
        if (OP == 2) x[i] = __vadd2(x[i], b);                               // VIADD.16x2
        if (OP == 3) x[i] = __vmaxs2(x[i], b);                              // VIMNMX.S16x2
        if (OP == 4) x[i] = ((unsigned)(x[i] + b) < c) ? x[i] + b : c;      // IADD + ISETP + SEL
        if (OP == 5) x[i] = max((int)x[i] + (int)b, (int)c);                // VIADDMNMX (32 bit)
      }
    }
  }
  unsigned s = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) s ^= x[i];
  out[blockIdx.x*blockDim.x + threadIdx.x] = s;
}
template <int OP> void run(const char *name, unsigned *out) {
  const int iters = 2000, grid = 148*8, block = 256;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<OP><<<grid, block>>>(out, 3, 0x00010002, 10); cudaDeviceSynchronize();
  cudaEventRecord(e0); k<OP><<<grid, block>>>(out, 3, 0x00010002, iters); cudaEventRecord(e1);
  cudaDeviceSynchronize();
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  double ops = (double)grid*block*iters*64;
  printf("%-28s %.3f ms  %.1f Gop/s (thread-level)  %.1f warp-instr/clk/SM at 1.965 GHz\n", name, ms,
         ops/ms/1e6, ops/32/(ms*1e-3)/148/1.965e9);
}
int main() {
  unsigned *out; cudaMalloc(&out, 148*8*256*4);
  run<0>("IADD3", out); run<1>("VIADDMNMX.S16x2", out); run<2>("VIADD.16x2", out);
  run<3>("VIMNMX.S16x2", out); run<4>("IADD+ISETP+SEL", out); run<5>("VIADDMNMX 32", out);
  return 0;
}

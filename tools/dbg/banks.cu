// Shared-memory wavefront microbenchmark for 16-byte accesses on sm_100a (tools only).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/dbg/banks tools/dbg/banks.cu
// One warp issues a stream of independent LDS.128 at per-lane word offsets given by
// the host; prints cycles per instruction (= wavefronts per instruction when the data
// pipe is the limit).  Patterns are the region kernels' access shapes on the 140-word
// pitch.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

constexpr int kAP = 140;

__global__ void k(const int *offs, int n, long long *cyc, int *sink) {
  extern __shared__ int4 sm[];
  for (int i = threadIdx.x; i < 68*kAP/4; i += 32) sm[i] = make_int4(i, i, i, i);
  __syncwarp();
  const int4 *p = reinterpret_cast<const int4 *>(reinterpret_cast<const int *>(sm) + offs[threadIdx.x]);
  int acc = 0;
  long long t0 = clock64();
  for (int it = 0; it < n; it++) {
#pragma unroll
    for (int u = 0; u < 16; u++) {
      int4 t;
      asm volatile("ld.volatile.shared.v4.s32 {%0,%1,%2,%3}, [%4];"
                   : "=r"(t.x), "=r"(t.y), "=r"(t.z), "=r"(t.w)
                   : "r"((unsigned)__cvta_generic_to_shared(p)));
      acc += t.x ^ t.y ^ t.z ^ t.w;
    }
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) *cyc = t1 - t0;
  sink[threadIdx.x] = acc;
}

static int *d_off; static long long *d_cyc; static int *d_sink;
static void run(const char *name, const int *off) {
  cudaMemcpy(d_off, off, 32*sizeof(int), cudaMemcpyHostToDevice);
  const int n = 4096;
  k<<<1, 32, 68*kAP*4>>>(d_off, n, d_cyc, d_sink);
  k<<<1, 32, 68*kAP*4>>>(d_off, n, d_cyc, d_sink);
  long long c;
  cudaMemcpy(&c, d_cyc, sizeof(c), cudaMemcpyDeviceToHost);
  // model: fixed quarters of 8 lanes, wavefronts = sum over quarters of the max
  // multiplicity of a 16-byte bank group with distinct addresses
  int wf = 0;
  for (int q = 0; q < 4; q++) {
    int mx = 0;
    for (int g = 0; g < 8; g++) {
      int cnt = 0; int seen[8]; int ns = 0;
      for (int l = 8*q; l < 8*q + 8; l++) {
        int chunk = off[l] >> 2;
        if ((chunk & 7) != g) continue;
        bool dup = false;
        for (int i = 0; i < ns; i++) dup |= seen[i] == chunk;
        if (!dup) { seen[ns++] = chunk; cnt++; }
      }
      if (cnt > mx) mx = cnt;
    }
    wf += mx;
  }
  printf("%-44s %6.2f cycles/LDS.128   quarter-warp model %d wavefronts\n", name,
         (double)c/(n*16.0), wf);
}

int main() {
  cudaMalloc(&d_off, 32*sizeof(int)); cudaMalloc(&d_cyc, 8); cudaMalloc(&d_sink, 32*sizeof(int));
  int off[32];
  for (int l = 0; l < 32; l++) off[l] = 4*l;
  run("contiguous chunks", off);
  for (int l = 0; l < 32; l++) off[l] = 32*l % (60*kAP);
  run("all lanes one bank group (8-way per quarter)", off);
  for (int l = 0; l < 32; l++) off[l] = (l & 7)*kAP + 4*(l >> 3);
  run("8 rows x 4 chunk columns", off);
  // ld_a_8x4 of 16 units, 2 lanes each (column halves), row r = 0: unit (UR, UC)
  srand(7);
  for (int trial = 0; trial < 3; trial++) {
    int UR[16], UC[16];
    for (int u = 0; u < 16; u++) {
      if (trial == 0) { UR[u] = 0; UC[u] = u; }
      else { UR[u] = rand() % 8; UC[u] = rand() % 16; }
    }
    for (int l = 0; l < 32; l++) off[l] = (8*UR[l >> 1] + 2)*kAP + 8*UC[l >> 1] + 4 + 4*(l & 1);
    run(trial == 0 ? "ld_a_8x4, 16 consecutive units" : "ld_a_8x4, random units", off);
    for (int g = 0; g < 4; g += 3) {
      for (int l = 0; l < 32; l++) {
        int u = l >> 1, h = l & 1, mu = (UC[u] - u) & 3;
        off[l] = (8*UR[u] + 2 + 2*((g + mu) & 3) + h)*kAP + 8*UC[u] + 4;
      }
      run(trial == 0 ? "a_rows8, 16 consecutive units" : "a_rows8, random units", off);
    }
  }
  // different groupings: conflict free only if wavefronts are lanes {0-7},{8-15},..
  for (int l = 0; l < 32; l++) off[l] = 4*(l & 7) + 32*(l >> 3);
  run("each quarter covers all groups, quarters collide", off);
  // conflict free only if lanes l and l+16 may share a wavefront (half-warp pairs)
  for (int l = 0; l < 32; l++) off[l] = 4*((l & 3) + 4*((l >> 4) & 1)) + 32*((l >> 2) & 3);
  run("lanes {0-3,16-19} cover all groups", off);
  printf("done\n");
  return 0;
}

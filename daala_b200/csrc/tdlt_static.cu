// od_tdlt_fwd_batch_static: the forward batch call for callers that must not allocate
// (code under CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).
// No counterpart in the reference.
//
// The forward stage needs one piece of scratch memory: the per-region PLAN records that
// tdlt_plan_kernel builds from the block-size maps and tdlt_fwd_region_kernel reads
// (tdlt_plan.cu).  od_tdlt_fwd_batch keeps them in a cudaMalloc'ed buffer of its context;
// here they live in a file-scope __device__ array of a fixed, modest size
// (OD_TDLT_STATIC_MAX_REGIONS records of 1 KB = 4 MiB, held by every GPU that loads the
// library until the process exits).  The entry point needs no context and calls no
// cudaMalloc / cudaFree / allocating library routine: geometry on the host stack, TMA
// descriptors as kernel parameters, two kernel launches on the caller's stream.
//
// Every call shares the one array.  Nothing in it has to start at zero (the plan kernel
// writes every byte of every record the forward kernel reads), but calls that overlap on
// the GPU -- two streams, or two host threads -- race on it: serialise them.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <cuda.h>
#include <cuda_runtime.h>
#include "../../include/daala_tdlt.h"
#include "tdlt_kernels.h"

using namespace tdlt;

static_assert(OD_TDLT_STATIC_MAX_REGIONS*(long long)kPlanBytes == 4ll << 20, "4 MiB of plan records");
static __device__ __align__(256) unsigned char g_static_plan[(size_t)OD_TDLT_STATIC_MAX_REGIONS*kPlanBytes];

namespace {

constexpr int kMaxDevices = 64;
std::mutex g_setup_mutex;
unsigned char *g_plan_addr[kMaxDevices];   // DEVICE address of g_static_plan, per device

int complain(int code, const char *what) {
  fprintf(stderr, "od_tdlt_fwd_batch_static: %s\n", what);
  return code;
}

// Once per device, with that device current: in host code `g_static_plan` is only a
// stand-in for the symbol, the device address comes from cudaGetSymbolAddress.
int plan_address(unsigned char **out) {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    return complain(OD_TDLT_EFAULT, "no current CUDA device");
  }
  std::lock_guard<std::mutex> lock(g_setup_mutex);
  if (!g_plan_addr[dev]) {
    void *p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_static_plan);
    if (e != cudaSuccess || !p) {
      fprintf(stderr, "od_tdlt_fwd_batch_static: cudaGetSymbolAddress: %s\n", cudaGetErrorString(e));
      return OD_TDLT_EFAULT;
    }
    e = cudaFuncSetAttribute(tdlt_fwd_region_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kRegSmemBytes);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(tdlt_fwd_haar_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kRegSmemBytes);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(tdlt_fwd_region_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(tdlt_fwd_haar_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
    if (e != cudaSuccess) {
      fprintf(stderr, "od_tdlt_fwd_batch_static: cudaFuncSetAttribute: %s\n", cudaGetErrorString(e));
      return OD_TDLT_EFAULT;
    }
    g_plan_addr[dev] = static_cast<unsigned char *>(p);
  }
  *out = g_plan_addr[dev];
  return OD_TDLT_SUCCESS;
}

typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *,
                                    const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                                    const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
encode_tiled_fn encode_tiled() {
  static encode_tiled_fn fn = nullptr;
  if (!fn) {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess) {
      fn = reinterpret_cast<encode_tiled_fn>(p);
    }
  }
  return fn;
}

}  // namespace

extern "C" int od_tdlt_fwd_batch_static(const od_tdlt_geometry *g, od_coeff *const d[],
                                        const od_coeff *const c[], const unsigned char *bsize,
                                        int nframes, int flags, void *stream) {
  if (!g || !d || !c || !bsize) return complain(OD_TDLT_EFAULT, "null argument");
  if (g->nplanes < 1 || g->nplanes > kMaxPlanes || g->pic_width < 1 || g->pic_height < 1 ||
      nframes < 1 || nframes > 65535) {
    return complain(OD_TDLT_EINVAL, "bad geometry or frame count");
  }
  // the same geometry od_tdlt_ctx_create derives (reference src/state.c:372-375)
  const int fw = (g->pic_width + 63) & ~63, fh = (g->pic_height + 63) & ~63;
  RegionBatch rb;
  memset(&rb, 0, sizeof(rb));
  rb.nplanes = g->nplanes;
  rb.pic_w = g->pic_width;
  rb.pic_h = g->pic_height;
  rb.bw = (fw >> 5)*4;
  rb.bh = (fh >> 5)*4;
  int reg0 = 0, grp0 = 0;
  for (int p = 0; p < g->nplanes; p++) {
    if (g->xdec[p] != g->ydec[p] || g->xdec[p] < 0 || g->xdec[p] > 1) {
      return complain(OD_TDLT_EINVAL, "decimation must be 0 or 1, equal in x and y");
    }
    RegionPlane &rp = rb.pl[p];
    rp.dec = g->xdec[p];
    rp.w = fw >> rp.dec;
    rp.h = fh >> rp.dec;
    rp.nrx = (rp.w + kRegW - 1)/kRegW;
    rp.nry = (rp.h + kRegH - 1)/kRegH;
    rp.reg0 = reg0;
    rp.inv_nrx = 1.0f/(float)rp.nrx;
    reg0 += rp.nrx*rp.nry;
    rp.ngx = (rp.nrx + kPlanPerCta - 1)/kPlanPerCta;
    rp.grp0 = grp0;
    grp0 += rp.ngx*rp.nry;
  }
  rb.regs_per_frame = reg0;
  rb.inv_rpf = 1.0f/(float)reg0;
  rb.groups_per_frame = grp0;
  rb.nframes = nframes;
  rb.inv_nframes = 1.0f/(float)nframes;
  rb.flags = flags & (OD_TDLT_KEYFRAME | OD_TDLT_HAAR);
  const long long total = (long long)reg0*nframes;
  if (total > OD_TDLT_STATIC_MAX_REGIONS) {
    fprintf(stderr, "od_tdlt_fwd_batch_static: %d frames x %d regions of 128x64 = %lld regions exceed "
            "OD_TDLT_STATIC_MAX_REGIONS (%d): split the batch or use od_tdlt_fwd_batch\n",
            nframes, reg0, total, OD_TDLT_STATIC_MAX_REGIONS);
    return OD_TDLT_EINVAL;
  }
  unsigned char *plan = nullptr;
  int rc = plan_address(&plan);
  if (rc) return rc;
  encode_tiled_fn enc = encode_tiled();
  if (!enc) return complain(OD_TDLT_EFAULT, "cuTensorMapEncodeTiled unavailable");
  TmapSet maps;
  memset(&maps, 0, sizeof(maps));
  PlanePtrs out;
  memset(&out, 0, sizeof(out));
  for (int p = 0; p < rb.nplanes; p++) {
    if (!c[p] || !d[p]) return complain(OD_TDLT_EFAULT, "null plane pointer");
    if (((uintptr_t)c[p] | (uintptr_t)d[p]) & 15) {
      return complain(OD_TDLT_EINVAL, "plane pointers must be 16-byte aligned");
    }
    cuuint64_t dims[3] = {(cuuint64_t)rb.pl[p].w, (cuuint64_t)rb.pl[p].h, (cuuint64_t)nframes};
    cuuint64_t strides[2] = {(cuuint64_t)rb.pl[p].w*4, (cuuint64_t)rb.pl[p].w*rb.pl[p].h*4};
    cuuint32_t boxd[3] = {(cuuint32_t)kAP, (cuuint32_t)kARegRows, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    if (enc(&maps.in[p], CU_TENSOR_MAP_DATA_TYPE_INT32, 3, const_cast<od_coeff *>(c[p]), dims, strides,
            boxd, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
            CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
      return complain(OD_TDLT_EFAULT, "cuTensorMapEncodeTiled failed");
    }
    out.p[p] = d[p];
  }
  cudaStream_t st = (cudaStream_t)stream;
  tdlt_plan_kernel<<<dim3((unsigned)rb.groups_per_frame, (unsigned)nframes), 192, 0, st>>>(rb, bsize, plan);
  dim3 grid((unsigned)rb.regs_per_frame, (unsigned)nframes);
  if (rb.flags & OD_TDLT_HAAR) {
    tdlt_fwd_haar_kernel<<<grid, kRegThreads, kRegSmemBytes, st>>>(maps, rb, plan, out);
  }
  else {
    tdlt_fwd_region_kernel<<<grid, kRegThreads, kRegSmemBytes, st>>>(maps, rb, plan, out);
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    fprintf(stderr, "od_tdlt_fwd_batch_static: launch failed: %s\n", cudaGetErrorString(e));
    return OD_TDLT_EFAULT;
  }
  return OD_TDLT_SUCCESS;
}

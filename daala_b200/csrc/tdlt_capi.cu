// C-ABI of the library (include/daala_tdlt.h): context management, TMA
// descriptor construction, kernel launches and the host-pointer wrappers.
// No CPU arithmetic path exists here: every entry point launches CUDA kernels
// or fails.
#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <vector>
#include <cuda.h>
#include <cuda_runtime.h>
#include "../../include/daala_tdlt.h"
#include "tdlt_kernels.h"
#include "tdlt_aux_kernels.h"

using namespace tdlt;

static std::atomic<long long> g_launches{0};
constexpr int kMaxChunks = 64;   // pipeline depth of one host-pointer call

// Region orders of the inverse kernel: rows of regions in flight per frame, by batch size.
constexpr int kOrderTables = 3;
#ifndef TDLT_BANDS0
#define TDLT_BANDS0 3      // (a build flag for tools/ablate.py experiments)
#endif
constexpr int kOrderBands[kOrderTables] = {TDLT_BANDS0, 8, 24};

struct od_tdlt_ctx {
  od_tdlt_geometry g;
  int device;
  int max_frames;
  int fw, fh, nhsb, nvsb;
  RegionBatch rb;               // template: nframes / flags filled per call
  long long plane_size[kMaxPlanes];
  // strip exchange of the inverse kernel
  int2 *strips;                 // (sample, epoch) pairs, kRegionStripInts per region
  int *ticket;
  unsigned short *order;        // inverse kernel: kOrderTables orders of the region positions of a frame
  unsigned char *plan;          // kPlanBytes per region of the largest batch so far
  long long plan_regions;
  const unsigned char *plan_maps;   // what the records in `plan` were built from
  int plan_nframes, plan_flags;
  cudaStream_t plan_stream;
  long long scratch_tiles;
  int epoch;
  // device staging for the host-pointer calls
  od_coeff *dev_c[kMaxPlanes];
  od_coeff *dev_d[kMaxPlanes];
  void *dev_pix[kMaxPlanes];
  unsigned char *dev_bsize;
  unsigned char *dev_aux;       // od_tdlt_dering_frames: superblock flags + skip maps
  unsigned char *dev_scan;      // the caller's scan tables (od_tdlt_set_scan_tables)
  unsigned short *dev_cotab;    // source positions for the fused coding-order store / load
  int *dering_dirs;             // deringing: 16 ints per superblock and frame (search -> filter kernel)
  long long dering_sbs;         // ... capacity in superblocks
  int scan_len[4];
  size_t dev_aux_bytes;
  cudaStream_t stream;          // kernels
  cudaStream_t s_in, s_out;     // uploads / downloads of the host-pointer calls
  cudaEvent_t ev_in[kMaxChunks], ev_run[kMaxChunks];
  char err[256];
};

static int fail(od_tdlt_ctx *ctx, int code, const char *fmt, ...) {
  if (ctx) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(ctx->err, sizeof(ctx->err), fmt, ap);
    va_end(ap);
  }
  return code;
}

#define CK(ctx, call) \
  do { \
    cudaError_t e_ = (call); \
    if (e_ != cudaSuccess) { \
      return fail(ctx, OD_TDLT_EFAULT, "%s:%d %s: %s", __FILE__, __LINE__, #call, \
                  cudaGetErrorString(e_)); \
    } \
  } while (0)

// Every context entry point makes the context's device current for the duration of the
// call and puts the caller's device back on return (a process may hold one context per GPU).
struct DeviceGuard {
  int prev = -1;
  bool ok = false, switched = false;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    if (prev == dev) { ok = true; return; }
    ok = cudaSetDevice(dev) == cudaSuccess;
    switched = ok;
  }
  ~DeviceGuard() { if (switched && prev >= 0) cudaSetDevice(prev); }
  DeviceGuard(const DeviceGuard &) = delete;
  DeviceGuard &operator=(const DeviceGuard &) = delete;
};

// cuTensorMapEncodeTiled is a driver-API symbol; fetch it through the runtime
// so the library does not link against libcuda directly.
typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *,
                                    const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                                    const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static encode_tiled_fn get_encode_tiled() {
  static encode_tiled_fn fn = nullptr;
  if (!fn) {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess && q == cudaDriverEntryPointSuccess) {
      fn = reinterpret_cast<encode_tiled_fn>(p);
    }
  }
  return fn;
}

static bool ranges_overlap(const void *a, size_t na, const void *b, size_t nb) {
  const uintptr_t x = (uintptr_t)a, y = (uintptr_t)b;
  return x < y + nb && y < x + na;
}

// ceil(2^32 / d) for the kernels' divisions by multiply-high (tdlt_aux_kernels.h div_recip);
// 0 stands for d = 1, whose reciprocal does not fit.
static unsigned recip32(int d) { return d <= 1 ? 0u : (unsigned)(((1ull << 32) + d - 1)/d); }

// 3-D map (x, y, frame) over nframes planes of w x h int32; box bx x by x 1.
static int make_tmap(od_tdlt_ctx *ctx, CUtensorMap *m, const void *base, int w, int h,
                     int nframes, int boxw, int boxh, int elem_bytes = 4) {
  encode_tiled_fn enc = get_encode_tiled();
  if (!enc) return fail(ctx, OD_TDLT_EFAULT, "cuTensorMapEncodeTiled unavailable");
  cuuint64_t dims[3] = {(cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)nframes};
  cuuint64_t strides[2] = {(cuuint64_t)w*elem_bytes, (cuuint64_t)w*h*elem_bytes};
  cuuint32_t boxd[3] = {(cuuint32_t)boxw, (cuuint32_t)boxh, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  const CUtensorMapDataType dt = elem_bytes == 1 ? CU_TENSOR_MAP_DATA_TYPE_UINT8
                                 : (elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_UINT16
                                                    : CU_TENSOR_MAP_DATA_TYPE_INT32);
  CUresult r = enc(m, dt, 3, const_cast<void *>(base), dims, strides,
                   boxd, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(ctx, OD_TDLT_EFAULT, "cuTensorMapEncodeTiled failed: %d", (int)r);
  return OD_TDLT_SUCCESS;
}

// The order in which the inverse kernel's tickets walk the region positions of a frame
// (tdlt_inv_region.cu): plane by plane, by rx + S*ry -- every region comes after the
// regions left of, above and above-left of it for any S >= 1.  S = nrx is raster order
// (one row of regions in flight per frame: the smallest working set of pages); S =
// nrx / b puts b rows in flight and the left neighbour b * nframes tickets back.
// Host only; order[] gets regs_per_frame region indices (reg0 + ry*nrx + rx).
struct OrderPlane { int nrx, nry, reg0; };
static void build_region_order(const OrderPlane *pl, int nplanes, int bands,
                               std::vector<unsigned short> &order) {
  std::vector<unsigned long long> key;
  for (int p = 0; p < nplanes; p++) {
    const int S = std::max(1, (pl[p].nrx + bands - 1)/bands);
    for (int ry = 0; ry < pl[p].nry; ry++) {
      for (int rx = 0; rx < pl[p].nrx; rx++) {
        key.push_back(((unsigned long long)p << 48) | ((unsigned long long)(rx + S*ry) << 16) |
                      (unsigned long long)(pl[p].reg0 + ry*pl[p].nrx + rx));
      }
    }
  }
  std::sort(key.begin(), key.end());
  order.resize(key.size());
  for (size_t i = 0; i < key.size(); i++) order[i] = (unsigned short)(key[i] & 0xffff);
}

extern "C" {

int od_tdlt_region_order(const od_tdlt_geometry *g, int rows_in_flight, unsigned short *order,
                         int capacity) {
  if (!g || !order) return OD_TDLT_EFAULT;
  if (g->nplanes < 1 || g->nplanes > kMaxPlanes || g->pic_width < 1 || g->pic_height < 1 ||
      rows_in_flight < 1) {
    return OD_TDLT_EINVAL;
  }
  OrderPlane pl[kMaxPlanes];
  int reg0 = 0;
  for (int p = 0; p < g->nplanes; p++) {
    if (g->xdec[p] != g->ydec[p] || g->xdec[p] < 0 || g->xdec[p] > 1) return OD_TDLT_EINVAL;
    const int w = ((g->pic_width + 63) & ~63) >> g->xdec[p], h = ((g->pic_height + 63) & ~63) >> g->xdec[p];
    pl[p].nrx = (w + kRegW - 1)/kRegW;
    pl[p].nry = (h + kRegH - 1)/kRegH;
    pl[p].reg0 = reg0;
    reg0 += pl[p].nrx*pl[p].nry;
  }
  if (reg0 > 65535 || reg0 > capacity) return OD_TDLT_EINVAL;
  std::vector<unsigned short> v;
  build_region_order(pl, g->nplanes, rows_in_flight, v);
  memcpy(order, v.data(), v.size()*sizeof(unsigned short));
  return reg0;
}

int od_tdlt_ctx_create(od_tdlt_ctx **out, const od_tdlt_geometry *g, int device,
                       int max_frames) {
  if (!out || !g) return OD_TDLT_EFAULT;
  *out = nullptr;
  if (g->nplanes < 1 || g->nplanes > kMaxPlanes || g->pic_width < 1 || g->pic_height < 1 ||
      max_frames < 0) {
    return OD_TDLT_EINVAL;
  }
  for (int p = 0; p < g->nplanes; p++) {
    if (g->xdec[p] != g->ydec[p] || g->xdec[p] < 0 || g->xdec[p] > 1) return OD_TDLT_EINVAL;
  }
  od_tdlt_ctx *ctx = (od_tdlt_ctx *)calloc(1, sizeof(*ctx));
  if (!ctx) return OD_TDLT_EFAULT;
  ctx->g = *g;
  ctx->device = device;
  ctx->max_frames = max_frames;
  // reference src/state.c:372-375: frame = picture padded to a multiple of 64
  ctx->fw = (g->pic_width + 63) & ~63;
  ctx->fh = (g->pic_height + 63) & ~63;
  ctx->nhsb = ctx->fw >> 5;
  ctx->nvsb = ctx->fh >> 5;
  RegionBatch &rb = ctx->rb;
  rb.nplanes = g->nplanes;
  rb.pic_w = g->pic_width;
  rb.pic_h = g->pic_height;
  rb.bw = ctx->nhsb*4;
  rb.bh = ctx->nvsb*4;
  int reg0 = 0, grp0 = 0;
  for (int p = 0; p < g->nplanes; p++) {
    RegionPlane &rp = rb.pl[p];
    rp.dec = g->xdec[p];
    rp.w = ctx->fw >> rp.dec;
    rp.h = ctx->fh >> rp.dec;
    rp.nrx = (rp.w + kRegW - 1)/kRegW;
    rp.nry = (rp.h + kRegH - 1)/kRegH;
    rp.reg0 = reg0;
    rp.inv_nrx = 1.0f/(float)rp.nrx;
    reg0 += rp.nrx*rp.nry;
    rp.ngx = (rp.nrx + kPlanPerCta - 1)/kPlanPerCta;
    rp.grp0 = grp0;
    grp0 += rp.ngx*rp.nry;
    ctx->plane_size[p] = (long long)rp.w*rp.h;
  }
  rb.regs_per_frame = reg0;
  rb.inv_rpf = 1.0f/(float)reg0;
  rb.groups_per_frame = grp0;
  *out = ctx;
  int rc = OD_TDLT_SUCCESS;
  do {
    if (cudaSetDevice(device) != cudaSuccess) { rc = OD_TDLT_EFAULT; break; }
    if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) {
      rc = OD_TDLT_EFAULT; break;
    }
    if (cudaStreamCreateWithFlags(&ctx->s_in, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&ctx->s_out, cudaStreamNonBlocking) != cudaSuccess) {
      rc = OD_TDLT_EFAULT; break;
    }
    {
      bool ok = true;
      for (int k = 0; k < kMaxChunks && ok; k++) {
        ok = cudaEventCreateWithFlags(&ctx->ev_in[k], cudaEventDisableTiming) == cudaSuccess &&
             cudaEventCreateWithFlags(&ctx->ev_run[k], cudaEventDisableTiming) == cudaSuccess;
      }
      if (!ok) { rc = OD_TDLT_EFAULT; break; }
    }
    if (cudaMalloc(&ctx->ticket, 256) != cudaSuccess) { rc = OD_TDLT_EFAULT; break; }
    {
      // Region orders of the inverse kernel (build_region_order above).  Measured (64 4K
      // frames, ms per launch incl. plan): frame-major raster 2.23; frame-interleaved with
      // b = 1 / 2 / 3 / 5 / 8 rows of regions in flight 2.12 / 2.09 / 2.08 / 2.08 / 2.09;
      // whole diagonals 2.14 - 2.25.  Three tables, the batch size picks one: about 200
      // tickets between a region and its left neighbour.
      if (rb.regs_per_frame > 65535) { rc = OD_TDLT_EINVAL; break; }
      const int nreg = rb.regs_per_frame;
      std::vector<unsigned short> order((size_t)kOrderTables*nreg), one;
      OrderPlane opl[kMaxPlanes];
      for (int p = 0; p < g->nplanes; p++) opl[p] = OrderPlane{rb.pl[p].nrx, rb.pl[p].nry, rb.pl[p].reg0};
      for (int tb = 0; tb < kOrderTables; tb++) {
        int bands = kOrderBands[tb];
        build_region_order(opl, g->nplanes, bands, one);
        memcpy(&order[(size_t)tb*nreg], one.data(), (size_t)nreg*sizeof(unsigned short));
      }
      if (cudaMalloc(&ctx->order, order.size()*sizeof(unsigned short)) != cudaSuccess ||
          cudaMemcpy(ctx->order, order.data(), order.size()*sizeof(unsigned short),
                     cudaMemcpyHostToDevice) != cudaSuccess) { rc = OD_TDLT_EFAULT; break; }
    }
    cudaFuncSetAttribute(tdlt_fwd_region_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kRegSmemBytes);
    cudaFuncSetAttribute(tdlt_inv_region_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kRegSmemBytes);
    cudaFuncSetAttribute(tdlt_fwd_haar_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kRegSmemBytes);
    cudaFuncSetAttribute(tdlt_inv_haar_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kRegSmemBytes);
    {
      const void *pixk[7] = {(const void *)tdlt_fwd_region_pix8_kernel, (const void *)tdlt_fwd_region_pix16_kernel,
                             (const void *)tdlt_inv_region_pix8_kernel, (const void *)tdlt_inv_region_pix16_kernel,
                             (const void *)tdlt_inv_region_co_kernel, (const void *)tdlt_inv_region_co_pix8_kernel,
                             (const void *)tdlt_inv_region_co_pix16_kernel};
      for (int k = 0; k < 7; k++) {
        cudaFuncSetAttribute(pixk[k], cudaFuncAttributeMaxDynamicSharedMemorySize, kRegSmemBytes);
        cudaFuncSetAttribute(pixk[k], cudaFuncAttributePreferredSharedMemoryCarveout, 100);
      }
    }
    cudaFuncSetAttribute(tdlt_fwd_region_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
    cudaFuncSetAttribute(tdlt_inv_region_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
    cudaFuncSetAttribute(tdlt_fwd_haar_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
    cudaFuncSetAttribute(tdlt_inv_haar_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
    if (max_frames > 0) {
      bool ok = true;
      for (int p = 0; p < g->nplanes && ok; p++) {
        size_t bytes = (size_t)ctx->plane_size[p]*max_frames*sizeof(od_coeff);
        ok = cudaMalloc(&ctx->dev_c[p], bytes) == cudaSuccess &&
             cudaMalloc(&ctx->dev_d[p], bytes) == cudaSuccess &&
             cudaMalloc(&ctx->dev_pix[p], (size_t)ctx->plane_size[p]*max_frames*2) == cudaSuccess;
      }
      ok = ok && cudaMalloc(&ctx->dev_bsize, (size_t)rb.bw*rb.bh*max_frames) == cudaSuccess;
      if (!ok) { rc = OD_TDLT_EFAULT; break; }
    }
  } while (0);
  if (rc != OD_TDLT_SUCCESS) {
    fprintf(stderr, "od_tdlt_ctx_create: CUDA setup failed: %s\n",
            cudaGetErrorString(cudaGetLastError()));
    od_tdlt_ctx_destroy(ctx);
    *out = nullptr;
  }
  return rc;
}

void od_tdlt_ctx_destroy(od_tdlt_ctx *ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  if (ctx->stream) cudaStreamSynchronize(ctx->stream);
  for (int p = 0; p < kMaxPlanes; p++) {
    cudaFree(ctx->dev_c[p]);
    cudaFree(ctx->dev_d[p]);
    cudaFree(ctx->dev_pix[p]);
  }
  cudaFree(ctx->dev_bsize);
  cudaFree(ctx->dev_aux);
  cudaFree(ctx->dev_scan);
  cudaFree(ctx->dev_cotab);
  cudaFree(ctx->strips);
  cudaFree(ctx->ticket);
  cudaFree(ctx->order);
  cudaFree(ctx->dering_dirs);
  cudaFree(ctx->plan);
  for (int k = 0; k < kMaxChunks; k++) {
    if (ctx->ev_in[k]) cudaEventDestroy(ctx->ev_in[k]);
    if (ctx->ev_run[k]) cudaEventDestroy(ctx->ev_run[k]);
  }
  if (ctx->s_in) cudaStreamDestroy(ctx->s_in);
  if (ctx->s_out) cudaStreamDestroy(ctx->s_out);
  if (ctx->stream) cudaStreamDestroy(ctx->stream);
  free(ctx);
}

int od_tdlt_frame_width(const od_tdlt_ctx *ctx) { return ctx->fw; }
int od_tdlt_frame_height(const od_tdlt_ctx *ctx) { return ctx->fh; }
int od_tdlt_nhsb(const od_tdlt_ctx *ctx) { return ctx->nhsb; }
int od_tdlt_nvsb(const od_tdlt_ctx *ctx) { return ctx->nvsb; }
long long od_tdlt_plane_size(const od_tdlt_ctx *ctx, int pli) {
  return (pli >= 0 && pli < ctx->g.nplanes) ? ctx->plane_size[pli] : 0;
}
long long od_tdlt_bsize_size(const od_tdlt_ctx *ctx) { return (long long)ctx->rb.bw*ctx->rb.bh; }
const char *od_tdlt_last_error(const od_tdlt_ctx *ctx) { return ctx ? ctx->err : "null context"; }
long long od_tdlt_launch_count(void) { return g_launches.load(); }
#ifdef TDLT_TRACE
int od_tdlt_debug_set_trace(long long *dev_buf) {
  return cudaMemcpyToSymbol(tdlt::g_tdlt_trace, &dev_buf, sizeof(dev_buf)) == cudaSuccess ? 0 : -1;
}
#endif

// Builds the plan records of a batch (one small kernel ahead of either transform
// kernel, same stream).
static int launch_plan(od_tdlt_ctx *ctx, const RegionBatch &rb, const unsigned char *bsize,
                       int nframes, int flags, cudaStream_t st) {
  if (flags & OD_TDLT_SAME_MAPS) {
    // (the plan depends on the maps and on lossless mode only, not on the coding-order flag)
    if (!ctx->plan || ctx->plan_maps != bsize || ctx->plan_nframes != nframes ||
        ctx->plan_flags != (rb.flags & (OD_TDLT_KEYFRAME | OD_TDLT_HAAR)) || ctx->plan_stream != st) {
      return fail(ctx, OD_TDLT_EINVAL, "OD_TDLT_SAME_MAPS without a matching previous call");
    }
    return OD_TDLT_SUCCESS;
  }
  long long total = (long long)rb.regs_per_frame*nframes;
  if (total > ctx->plan_regions) {
    CK(ctx, cudaDeviceSynchronize());
    cudaFree(ctx->plan);
    ctx->plan = nullptr;
    ctx->plan_regions = 0;
    CK(ctx, cudaMalloc(&ctx->plan, (size_t)total*kPlanBytes));
    ctx->plan_regions = total;
  }
  dim3 grid((unsigned)rb.groups_per_frame, (unsigned)nframes);
  tdlt_plan_kernel<<<grid, 192, 0, st>>>(rb, bsize, ctx->plan);
  g_launches++;
  ctx->plan_maps = nullptr;
  CK(ctx, cudaGetLastError());
  ctx->plan_maps = bsize;
  ctx->plan_nframes = nframes;
  ctx->plan_flags = rb.flags & (OD_TDLT_KEYFRAME | OD_TDLT_HAAR);
  ctx->plan_stream = st;
  return OD_TDLT_SUCCESS;
}

// ---------------------------------------------------------------- batches --
static int conv_shift(int xstride, int bitdepth_mode, int lossless_p) {
  // reference src/state.c:1221-1223, 1236-1238 (OD_COEFF_SHIFT == 4)
  if (xstride == 1) return lossless_p ? (bitdepth_mode - 1)*2 : 4;
  return lossless_p ? 4 - (bitdepth_mode - 1)*2 : 0;
}

// OD_TDLT_CODING_ORDER: the coefficient side is the block-linear plane of coding-order vectors.
static int check_coding_order(od_tdlt_ctx *ctx, RegionBatch &rb) {
  rb.cotab = nullptr;
  if (!(rb.flags & OD_TDLT_CODING_ORDER)) return OD_TDLT_SUCCESS;
  if (rb.flags & OD_TDLT_HAAR) {
    return fail(ctx, OD_TDLT_EINVAL, "OD_TDLT_CODING_ORDER does not apply to lossless mode (the "
                "reference codes the wavelet's coefficients in raster order, src/encode.c:1150-1157)");
  }
  if (!ctx->dev_cotab) return fail(ctx, OD_TDLT_EINVAL, "od_tdlt_set_scan_tables has not been called");
  rb.cotab = ctx->dev_cotab;
  return OD_TDLT_SUCCESS;
}

// The forward batch: in[] are od_coeff planes (pix_bytes 0) or planes of pix_bytes-byte
// reference-buffer samples converted on the fly (od_ref_buf_to_coeff fused into the load).
static int fwd_batch_any(od_tdlt_ctx *ctx, od_coeff *const d[], const void *const in[],
                         const unsigned char *bsize, int nframes, int flags, int pix_bytes,
                         int pix_shift, void *stream) {
  if (!ctx || !d || !in || !bsize) return OD_TDLT_EFAULT;
  if (nframes < 1) return fail(ctx, OD_TDLT_EINVAL, "nframes < 1");
  DeviceGuard dev_guard(ctx->device);
  if (!dev_guard.ok) return fail(ctx, OD_TDLT_EFAULT, "cudaSetDevice(%d) failed", ctx->device);
  TmapSet maps;
  memset(&maps, 0, sizeof(maps));
  RegionBatch rb = ctx->rb;
  rb.nframes = nframes;
  rb.inv_nframes = 1.0f/(float)nframes;
  rb.flags = flags & (OD_TDLT_KEYFRAME | OD_TDLT_HAAR | OD_TDLT_CODING_ORDER);
  rb.pix_shift = pix_shift;
  if (pix_bytes && (rb.flags & OD_TDLT_HAAR)) {
    return fail(ctx, OD_TDLT_EINVAL, "the fused pixel path has no lossless (OD_TDLT_HAAR) form");
  }
  if (int crc = check_coding_order(ctx, rb)) return crc;
  PlanePtrs out;
  memset(&out, 0, sizeof(out));
  for (int p = 0; p < rb.nplanes; p++) {
    if (((uintptr_t)in[p] | (uintptr_t)d[p]) & 15) {
      return fail(ctx, OD_TDLT_EINVAL, "plane pointers must be 16-byte aligned");
    }
    const size_t nsamp = (size_t)ctx->plane_size[p]*nframes;
    if (ranges_overlap(in[p], nsamp*(pix_bytes ? pix_bytes : sizeof(od_coeff)), d[p], nsamp*sizeof(od_coeff))) {
      return fail(ctx, OD_TDLT_EINVAL, "plane %d: input and output overlap (the stage is out of place)", p);
    }
    int rc = pix_bytes ? make_tmap(ctx, &maps.in[p], in[p], rb.pl[p].w, rb.pl[p].h, nframes,
                                   pix_bytes == 1 ? PixBox<1>::kW : PixBox<2>::kW, kARegRows, pix_bytes)
                       : make_tmap(ctx, &maps.in[p], in[p], rb.pl[p].w, rb.pl[p].h, nframes, kAP, kARegRows);
    if (rc) return rc;
    out.p[p] = d[p];
  }
  if (nframes > 65535) return fail(ctx, OD_TDLT_EINVAL, "nframes > 65535");
  int prc = launch_plan(ctx, rb, bsize, nframes, flags, (cudaStream_t)stream);
  if (prc) return prc;
  dim3 grid((unsigned)rb.regs_per_frame, (unsigned)nframes);
  cudaStream_t st = (cudaStream_t)stream;
  if (rb.flags & OD_TDLT_HAAR) tdlt_fwd_haar_kernel<<<grid, kRegThreads, kRegSmemBytes, st>>>(maps, rb, ctx->plan, out);
  else if (pix_bytes == 1) tdlt_fwd_region_pix8_kernel<<<grid, kRegThreads, kRegSmemBytes, st>>>(maps, rb, ctx->plan, out);
  else if (pix_bytes == 2) tdlt_fwd_region_pix16_kernel<<<grid, kRegThreads, kRegSmemBytes, st>>>(maps, rb, ctx->plan, out);
  else tdlt_fwd_region_kernel<<<grid, kRegThreads, kRegSmemBytes, st>>>(maps, rb, ctx->plan, out);
  g_launches++;
  CK(ctx, cudaGetLastError());
  return OD_TDLT_SUCCESS;
}

int od_tdlt_fwd_batch(od_tdlt_ctx *ctx, od_coeff *const d[], const od_coeff *const c[],
                      const unsigned char *bsize, int nframes, int flags, void *stream) {
  return fwd_batch_any(ctx, d, reinterpret_cast<const void *const *>(c), bsize, nframes, flags, 0, 0,
                       stream);
}

// od_img_copy_pad (reference src/encode.c:1728-1750) for a batch of pictures.
int od_tdlt_img_copy_pad_batch(od_tdlt_ctx *ctx, void *const dst[], const int dst_ystride[],
                               const long long dst_frame_stride[], int dst_xstride, int dst_bitdepth,
                               int border, const void *const src[], const int src_ystride[],
                               const long long src_frame_stride[], int src_xstride, int src_bitdepth,
                               int nframes, void *stream) {
  if (!ctx || !dst || !dst_ystride || !dst_frame_stride || !src || !src_ystride || !src_frame_stride) {
    return OD_TDLT_EFAULT;
  }
  if (nframes < 1 || border < 0 || (dst_xstride != 1 && dst_xstride != 2) ||
      (src_xstride != 1 && src_xstride != 2) || src_bitdepth < 8 || src_bitdepth > 16 ||
      dst_bitdepth < 8 || dst_bitdepth > 16 || (src_bitdepth > 8) != (src_xstride == 2) ||
      (dst_bitdepth > 8) != (dst_xstride == 2)) {
    return fail(ctx, OD_TDLT_EINVAL, "od_tdlt_img_copy_pad_batch: 8-bit samples are one byte, deeper ones two");
  }
  DeviceGuard dev_guard(ctx->device);
  if (!dev_guard.ok) return fail(ctx, OD_TDLT_EFAULT, "cudaSetDevice(%d) failed", ctx->device);
  ImgPadParams P;
  memset(&P, 0, sizeof(P));
  P.nplanes = ctx->g.nplanes;
  P.nframes = nframes;
  P.src_xs = src_xstride; P.dst_xs = dst_xstride;
  P.src_bits = src_bitdepth; P.dst_bits = dst_bitdepth;
  for (int p = 0; p < P.nplanes; p++) {
    if (!dst[p] || !src[p]) return OD_TDLT_EFAULT;
    const int dec = ctx->g.xdec[p];
    ImgPadPlane &pl = P.pl[p];
    pl.src = (const unsigned char *)src[p];
    pl.dst = (unsigned char *)dst[p];
    pl.src_ystride = src_ystride[p]; pl.dst_ystride = dst_ystride[p];
    pl.src_frame_stride = src_frame_stride[p]; pl.dst_frame_stride = dst_frame_stride[p];
    // src/encode.c:1743-1744, src/state.c:134-139
    pl.pic_w = (ctx->g.pic_width + (1 << dec) - 1) >> dec;
    pl.pic_h = (ctx->g.pic_height + (1 << dec) - 1) >> dec;
    pl.plane_w = ctx->fw >> dec;
    pl.plane_h = ctx->fh >> dec;
    pl.border = border >> dec;
    if (dst_xstride == 2 && ((uintptr_t)dst[p] | (uintptr_t)dst_ystride[p] | (uintptr_t)dst_frame_stride[p]) & 1) {
      return fail(ctx, OD_TDLT_EINVAL, "16-bit destination must be 2-byte aligned");
    }
    if (src_xstride == 2 && ((uintptr_t)src[p] | (uintptr_t)src_ystride[p] | (uintptr_t)src_frame_stride[p]) & 1) {
      return fail(ctx, OD_TDLT_EINVAL, "16-bit source must be 2-byte aligned");
    }
  }
  launch_img_copy_pad(P, (cudaStream_t)stream);
  g_launches += border > 0 ? 3 : 2;
  CK(ctx, cudaGetLastError());
  return OD_TDLT_SUCCESS;
}

static int check_pix_args(od_tdlt_ctx *ctx, int xstride, int bitdepth_mode) {
  if ((xstride != 1 && xstride != 2) || bitdepth_mode < 1 || bitdepth_mode > 3) {
    return fail(ctx, OD_TDLT_EINVAL, "bad conversion arguments");
  }
  return OD_TDLT_SUCCESS;
}

int od_tdlt_fwd_batch_pix(od_tdlt_ctx *ctx, od_coeff *const d[], const void *const pix[],
                          const unsigned char *bsize, int nframes, int flags, int xstride,
                          int bitdepth_mode, int lossless_p, void *stream) {
  if (!ctx || !pix) return OD_TDLT_EFAULT;
  int rc = check_pix_args(ctx, xstride, bitdepth_mode);
  if (rc) return rc;
  return fwd_batch_any(ctx, d, pix, bsize, nframes, flags, xstride,
                       conv_shift(xstride, bitdepth_mode, lossless_p), stream);
}

static int ensure_scratch(od_tdlt_ctx *ctx, long long tiles) {
  if (tiles <= ctx->scratch_tiles) return OD_TDLT_SUCCESS;
  CK(ctx, cudaDeviceSynchronize());
  cudaFree(ctx->strips);
  ctx->strips = nullptr;
  ctx->scratch_tiles = 0;
  size_t bytes = (size_t)tiles*kRegionStripInts*sizeof(int2);
  CK(ctx, cudaMalloc(&ctx->strips, bytes));
  CK(ctx, cudaMemset(ctx->strips, 0, bytes));    // epoch 0 = nothing published yet
  ctx->scratch_tiles = tiles;
  ctx->epoch = 0;
  return OD_TDLT_SUCCESS;
}

// The inverse batch: out[] are od_coeff planes (pix_bytes 0) or planes of pix_bytes-byte
// reference-buffer samples (od_coeff_to_ref_buf fused into the store).
static int inv_batch_any(od_tdlt_ctx *ctx, void *const outp[], const od_coeff *const d[],
                         const unsigned char *bsize, int nframes, int flags, int pix_bytes,
                         int pix_shift, void *stream) {
  if (!ctx || !d || !outp || !bsize) return OD_TDLT_EFAULT;
  if (nframes < 1) return fail(ctx, OD_TDLT_EINVAL, "nframes < 1");
  DeviceGuard dev_guard(ctx->device);
  if (!dev_guard.ok) return fail(ctx, OD_TDLT_EFAULT, "cudaSetDevice(%d) failed", ctx->device);
  RegionBatch rb = ctx->rb;
  rb.nframes = nframes;
  rb.inv_nframes = 1.0f/(float)nframes;
  rb.flags = flags & (OD_TDLT_KEYFRAME | OD_TDLT_HAAR | OD_TDLT_CODING_ORDER);
  rb.pix_shift = pix_shift;
  if (pix_bytes && (rb.flags & OD_TDLT_HAAR)) {
    return fail(ctx, OD_TDLT_EINVAL, "the fused pixel path has no lossless (OD_TDLT_HAAR) form");
  }
  if (int crc = check_coding_order(ctx, rb)) return crc;
  TmapSet maps;
  memset(&maps, 0, sizeof(maps));
  PlanePtrs out, inp;
  memset(&out, 0, sizeof(out));
  memset(&inp, 0, sizeof(inp));
  for (int p = 0; p < rb.nplanes; p++) {
    if (((uintptr_t)outp[p] | (uintptr_t)d[p]) & 15) {
      return fail(ctx, OD_TDLT_EINVAL, "plane pointers must be 16-byte aligned");
    }
    const size_t nsamp = (size_t)ctx->plane_size[p]*nframes;
    if (ranges_overlap(outp[p], nsamp*(pix_bytes ? pix_bytes : sizeof(od_coeff)), d[p], nsamp*sizeof(od_coeff))) {
      return fail(ctx, OD_TDLT_EINVAL, "plane %d: input and output overlap (the stage is out of place)", p);
    }
    int trc = make_tmap(ctx, &maps.in[p], d[p], rb.pl[p].w, rb.pl[p].h, nframes, kAP, kARegRows);
    if (trc) return trc;
    out.p[p] = reinterpret_cast<int *>(outp[p]);
    inp.p[p] = const_cast<int *>(d[p]);
  }
  long long total = (long long)rb.regs_per_frame*nframes;
  int rc = ensure_scratch(ctx, total);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  if (nframes > 65535) return fail(ctx, OD_TDLT_EINVAL, "nframes > 65535");
  rc = launch_plan(ctx, rb, bsize, nframes, flags, st);
  if (rc) return rc;
  CK(ctx, cudaMemsetAsync(ctx->ticket, 0, sizeof(int), st));
  ctx->epoch++;
  const unsigned short *order =
      ctx->order + (size_t)(nframes >= 48 ? 0 : (nframes >= 16 ? 1 : 2))*rb.regs_per_frame;
  const unsigned grid = (unsigned)total;
  if (rb.flags & OD_TDLT_CODING_ORDER) {
    auto k = pix_bytes == 1 ? tdlt_inv_region_co_pix8_kernel
             : (pix_bytes == 2 ? tdlt_inv_region_co_pix16_kernel : tdlt_inv_region_co_kernel);
    k<<<grid, kRegThreads, kRegSmemBytes, st>>>(maps, rb, ctx->plan, out, ctx->strips, ctx->ticket,
                                                ctx->epoch, order, inp);
  }
  else if (rb.flags & OD_TDLT_HAAR) {
    tdlt_inv_haar_kernel<<<grid, kRegThreads, kRegSmemBytes, st>>>(
        maps, rb, ctx->plan, out, ctx->strips, ctx->ticket, ctx->epoch, order, inp);
  }
  else if (pix_bytes == 1) {
    tdlt_inv_region_pix8_kernel<<<grid, kRegThreads, kRegSmemBytes, st>>>(
        maps, rb, ctx->plan, out, ctx->strips, ctx->ticket, ctx->epoch, order, inp);
  }
  else if (pix_bytes == 2) {
    tdlt_inv_region_pix16_kernel<<<grid, kRegThreads, kRegSmemBytes, st>>>(
        maps, rb, ctx->plan, out, ctx->strips, ctx->ticket, ctx->epoch, order, inp);
  }
  else {
    tdlt_inv_region_kernel<<<grid, kRegThreads, kRegSmemBytes, st>>>(
        maps, rb, ctx->plan, out, ctx->strips, ctx->ticket, ctx->epoch, order, inp);
  }
  g_launches++;
  CK(ctx, cudaGetLastError());
  return OD_TDLT_SUCCESS;
}

int od_tdlt_inv_batch(od_tdlt_ctx *ctx, od_coeff *const c[], const od_coeff *const d[],
                      const unsigned char *bsize, int nframes, int flags, void *stream) {
  return inv_batch_any(ctx, reinterpret_cast<void *const *>(c), d, bsize, nframes, flags, 0, 0, stream);
}

int od_tdlt_inv_batch_pix(od_tdlt_ctx *ctx, void *const pix[], const od_coeff *const d[],
                          const unsigned char *bsize, int nframes, int flags, int xstride,
                          int bitdepth_mode, int lossless_p, void *stream) {
  if (!ctx || !pix) return OD_TDLT_EFAULT;
  int rc = check_pix_args(ctx, xstride, bitdepth_mode);
  if (rc) return rc;
  return inv_batch_any(ctx, pix, d, bsize, nframes, flags, xstride,
                       conv_shift(xstride, bitdepth_mode, lossless_p), stream);
}

int od_tdlt_pix_to_coeff_batch(od_tdlt_ctx *ctx, od_coeff *c, const void *pix, int pli,
                               int nframes, int xstride, int bitdepth_mode, int lossless_p,
                               void *stream) {
  if (!ctx || !c || !pix) return OD_TDLT_EFAULT;
  if (pli < 0 || pli >= ctx->g.nplanes || (xstride != 1 && xstride != 2) || bitdepth_mode < 1 ||
      bitdepth_mode > 3) {
    return fail(ctx, OD_TDLT_EINVAL, "bad conversion arguments");
  }
  DeviceGuard dev_guard(ctx->device);
  if (!dev_guard.ok) return fail(ctx, OD_TDLT_EFAULT, "cudaSetDevice(%d) failed", ctx->device);
  launch_pix_to_coeff(c, pix, ctx->plane_size[pli]*nframes, xstride,
                      conv_shift(xstride, bitdepth_mode, lossless_p), (cudaStream_t)stream);
  g_launches++;
  CK(ctx, cudaGetLastError());
  return OD_TDLT_SUCCESS;
}

int od_tdlt_coeff_to_pix_batch(od_tdlt_ctx *ctx, void *pix, const od_coeff *c, int pli,
                               int nframes, int xstride, int bitdepth_mode, int lossless_p,
                               void *stream) {
  if (!ctx || !c || !pix) return OD_TDLT_EFAULT;
  if (pli < 0 || pli >= ctx->g.nplanes || (xstride != 1 && xstride != 2) || bitdepth_mode < 1 ||
      bitdepth_mode > 3) {
    return fail(ctx, OD_TDLT_EINVAL, "bad conversion arguments");
  }
  DeviceGuard dev_guard(ctx->device);
  if (!dev_guard.ok) return fail(ctx, OD_TDLT_EFAULT, "cudaSetDevice(%d) failed", ctx->device);
  launch_coeff_to_pix(pix, c, ctx->plane_size[pli]*nframes, xstride,
                      conv_shift(xstride, bitdepth_mode, lossless_p), (cudaStream_t)stream);
  g_launches++;
  CK(ctx, cudaGetLastError());
  return OD_TDLT_SUCCESS;
}

// Chroma-from-luma predictor planes (reference src/intra.c:72-109 over the leaf visits of
// src/encode.c:1491-1517 / src/decode.c:780-821): l is a plane of the size of plane 1.
int od_tdlt_cfl_pred_batch(od_tdlt_ctx *ctx, od_coeff *l, const od_coeff *d_luma,
                           const unsigned char *bsize, int nframes, void *stream) {
  if (!ctx || !l || !d_luma || !bsize) return OD_TDLT_EFAULT;
  if (ctx->g.nplanes < 2 || ctx->g.xdec[0] || ctx->g.ydec[0] || ctx->g.xdec[1] != 1 ||
      ctx->g.ydec[1] != 1) {
    return fail(ctx, OD_TDLT_EINVAL, "chroma-from-luma planes need 4:2:0 (plane 0 full size, "
                "plane 1 halved both ways); with 4:4:4 the predictor is the luma plane itself");
  }
  if (nframes < 1) return fail(ctx, OD_TDLT_EINVAL, "nframes %d", nframes);
  DeviceGuard dev_guard(ctx->device);
  if (!dev_guard.ok) return fail(ctx, OD_TDLT_EFAULT, "cudaSetDevice(%d) failed", ctx->device);
  launch_cfl_pred(l, ctx->fw >> 1, ctx->plane_size[1], d_luma, ctx->fw, ctx->plane_size[0], bsize,
                  ctx->nhsb*4, ctx->nvsb*4, nframes, (cudaStream_t)stream);
  g_launches++;
  CK(ctx, cudaGetLastError());
  return OD_TDLT_SUCCESS;
}

// Deringing of the reconstructed planes (reference od_dering, src/filter.c:1835-1941, over
// the superblock loop of src/decode.c:1044-1122): out = in, superblocks whose flag is set
// filtered.  The thresholds follow the reference's double-precision expression, evaluated
// on the host exactly as the reference does (src/filter.c:1877).
int od_tdlt_dering_batch(od_tdlt_ctx *ctx, od_coeff *const out[], const od_coeff *const in[],
                         const unsigned char *flags, const unsigned char *const bskip[],
                         const int q[], int nframes, void *stream) {
  if (!ctx || !out || !in || !flags || !bskip || !q) return OD_TDLT_EFAULT;
  if (nframes < 1) return fail(ctx, OD_TDLT_EINVAL, "nframes %d", nframes);
  DeringParams P = {};
  P.nplanes = ctx->g.nplanes;
  for (int p = 0; p < P.nplanes; p++) {
    if (!out[p] || !in[p] || !bskip[p]) return OD_TDLT_EFAULT;
    if (out[p] == in[p]) return fail(ctx, OD_TDLT_EINVAL, "deringing is out of place (plane %d)", p);
    if (ctx->g.xdec[p] != ctx->g.ydec[p] || ctx->g.xdec[p] > 1 || (p == 0 && ctx->g.xdec[p])) {
      return fail(ctx, OD_TDLT_EINVAL, "deringing needs 4:4:4 or 4:2:0 planes (the reference "
                  "asserts xdec == ydec, src/decode.c:1099)");
    }
    if (q[p] < 0) return fail(ctx, OD_TDLT_EINVAL, "quantizer %d", q[p]);
    P.in[p] = in[p];
    P.out[p] = out[p];
    P.bskip[p] = bskip[p];
    P.plane_size[p] = ctx->plane_size[p];
    P.dec[p] = ctx->g.xdec[p];
    P.thr[p] = (int)(1.0*pow((double)q[p], 0.84182));
  }
  P.flags = flags;
  P.nhsb = ctx->nhsb;
  P.nvsb = ctx->nvsb;
  P.skip_stride = ctx->nhsb*8;
  P.skip_frame = (long long)ctx->nhsb*8*ctx->nvsb*8;
  P.recip_nhsb = recip32(ctx->nhsb);
  if (nframes > 65535) return fail(ctx, OD_TDLT_EINVAL, "nframes %d (at most 65535 per call)", nframes);
  // the superblock tiles arrive as TMA boxes: (32 >> dec) + 8 words by (32 >> dec) + 6 rows
  DeringMaps maps;
  for (int p = 0; p < P.nplanes; p++) {
    if ((uintptr_t)in[p] & 15) return fail(ctx, OD_TDLT_EINVAL, "plane %d is not 16-byte aligned", p);
    const int dec = P.dec[p], n = 32 >> dec;
    if (int rc = make_tmap(ctx, &maps.in[p], in[p], ctx->fw >> dec, ctx->fh >> dec, nframes, n + 8, n + 6)) return rc;
  }
  if (int rc = make_tmap(ctx, &maps.dir, in[0], ctx->fw, ctx->fh, nframes, 64, 32)) return rc;
  const int npair = (ctx->nhsb + 1)/2;
  P.recip_npair = recip32(npair);
  DeviceGuard dev_guard(ctx->device);
  if (!dev_guard.ok) return fail(ctx, OD_TDLT_EFAULT, "cudaSetDevice(%d) failed", ctx->device);
  // directions and thresholds of the luma blocks, between the search and the filter kernel
  const long long nsb = (long long)ctx->nhsb*ctx->nvsb*nframes;
  if (nsb > ctx->dering_sbs) {
    CK(ctx, cudaDeviceSynchronize());
    cudaFree(ctx->dering_dirs);
    ctx->dering_dirs = nullptr;
    ctx->dering_sbs = 0;
    CK(ctx, cudaMalloc(&ctx->dering_dirs, (size_t)nsb*16*sizeof(int)));
    ctx->dering_sbs = nsb;
  }
  launch_dering(maps, P, ctx->dering_dirs, nframes, (cudaStream_t)stream);
  g_launches++;
  g_launches++;
  CK(ctx, cudaGetLastError());
  return OD_TDLT_SUCCESS;
}

// Keyframe bilinear smoothing (reference od_smooth_recursive with min_bs = 32x32,
// src/filter.c:2010-2039 / :1952-2008, called per superblock and plane from
// src/decode.c:1134-1146): in place on the reconstructed planes.
int od_tdlt_smooth_batch(od_tdlt_ctx *ctx, od_coeff *const c[], const unsigned char *bsize,
                         const int q[], int nframes, void *stream) {
  if (!ctx || !c || !bsize || !q) return OD_TDLT_EFAULT;
  if (nframes < 1) return fail(ctx, OD_TDLT_EINVAL, "nframes %d", nframes);
  SmoothParams P = {};
  P.nplanes = ctx->g.nplanes;
  for (int p = 0; p < P.nplanes; p++) {
    if (!c[p]) return OD_TDLT_EFAULT;
    if (ctx->g.xdec[p] != ctx->g.ydec[p] || ctx->g.xdec[p] > 1) {
      return fail(ctx, OD_TDLT_EINVAL, "smoothing needs 4:4:4 or 4:2:0 planes (the reference "
                  "asserts xdec == ydec, src/filter.c:2016)");
    }
    if (q[p] < 0 || q[p] > 10000) return fail(ctx, OD_TDLT_EINVAL, "quantizer %d", q[p]);
    P.c[p] = c[p];
    P.plane_size[p] = ctx->plane_size[p];
    P.dec[p] = ctx->g.xdec[p];
    P.q[p] = q[p];
  }
  P.bsize = bsize;
  P.nhsb = ctx->nhsb;
  P.nvsb = ctx->nvsb;
  DeviceGuard dev_guard(ctx->device);
  if (!dev_guard.ok) return fail(ctx, OD_TDLT_EFAULT, "cudaSetDevice(%d) failed", ctx->device);
  launch_smooth(P, nframes, (cudaStream_t)stream);
  g_launches++;
  CK(ctx, cudaGetLastError());
  return OD_TDLT_SUCCESS;
}

// Raster <-> coding order (reference src/partition.c:95-200).  The scan tables are the
// caller's: OD_ZIGZAG4/8/16/32 as they lie in the reference (src/zigzag*.c; 15, 48, 192 and
// 768 (x, y) byte pairs) and how many entries of each its layouts order
// (OD_LAYOUT*.band_offsets[nb_bands], src/partition.c:43-73: 15, 48, 192, 256).
int od_tdlt_set_scan_tables(od_tdlt_ctx *ctx, const unsigned char *const tables[4],
                            const int lens[4]) {
  static const int kTableLen[4] = {15, 48, 192, 768};
  if (!ctx || !tables || !lens) return OD_TDLT_EFAULT;
  unsigned char host[2*(15 + 48 + 192 + 768)];
  int off = 0;
  for (int k = 0; k < 4; k++) {
    if (!tables[k]) return OD_TDLT_EFAULT;
    if (lens[k] < 0 || lens[k] > kTableLen[k]) {
      return fail(ctx, OD_TDLT_EINVAL, "scan length %d of table %d", lens[k], k);
    }
    for (int i = 0; i < lens[k]; i++) {
      if (tables[k][2*i] >= (4 << k) || tables[k][2*i + 1] >= (4 << k)) {
        return fail(ctx, OD_TDLT_EINVAL, "scan table %d, entry %d outside the block", k, i);
      }
    }
    memcpy(host + off, tables[k], 2*kTableLen[k]);
    off += 2*kTableLen[k];
  }
  DeviceGuard dev_guard(ctx->device);
  if (!dev_guard.ok) return fail(ctx, OD_TDLT_EFAULT, "cudaSetDevice(%d) failed", ctx->device);
  if (!ctx->dev_scan) CK(ctx, cudaMalloc(&ctx->dev_scan, sizeof(host)));
  CK(ctx, cudaMemcpy(ctx->dev_scan, host, sizeof(host), cudaMemcpyHostToDevice));
  for (int k = 0; k < 4; k++) ctx->scan_len[k] = lens[k];
  // where every entry of a superblock's block-linear range comes from, per block size: what
  // the transform kernels' fused coding-order store / load read (OD_TDLT_CODING_ORDER)
  if (!ctx->dev_cotab) CK(ctx, cudaMalloc(&ctx->dev_cotab, kCoTabEntries*sizeof(unsigned short)));
  launch_cotab_build(ctx->dev_cotab, ctx->dev_scan, ctx->scan_len, 0);
  g_launches++;
  CK(ctx, cudaGetLastError());
  CK(ctx, cudaDeviceSynchronize());
  return OD_TDLT_SUCCESS;
}

int od_tdlt_coding_order_batch(od_tdlt_ctx *ctx, od_coeff *const v[], od_coeff *const d[],
                               const unsigned char *bsize, int nframes, int to_coding,
                               void *stream) {
  if (!ctx || !v || !d || !bsize) return OD_TDLT_EFAULT;
  if (!ctx->dev_scan) return fail(ctx, OD_TDLT_EINVAL, "od_tdlt_set_scan_tables has not been called");
  if (nframes < 1) return fail(ctx, OD_TDLT_EINVAL, "nframes %d", nframes);
  CodingParams P = {};
  P.nplanes = ctx->g.nplanes;
  for (int p = 0; p < P.nplanes; p++) {
    if (!v[p] || !d[p]) return OD_TDLT_EFAULT;
    if (v[p] == d[p]) return fail(ctx, OD_TDLT_EINVAL, "coding order is out of place (plane %d)", p);
    if (ctx->g.xdec[p] != ctx->g.ydec[p] || ctx->g.xdec[p] > 1) {
      return fail(ctx, OD_TDLT_EINVAL, "coding order needs 4:4:4 or 4:2:0 planes");
    }
    P.v[p] = v[p];
    P.d[p] = d[p];
    P.plane_size[p] = ctx->plane_size[p];
    P.dec[p] = ctx->g.xdec[p];
  }
  for (int k = 0; k < 4; k++) P.scan_len[k] = ctx->scan_len[k];
  P.bsize = bsize;
  P.scan = ctx->dev_scan;
  P.nhsb = ctx->nhsb;
  P.nvsb = ctx->nvsb;
  P.to_coding = to_coding != 0;
  if ((long long)ctx->nhsb*ctx->nvsb >= 65536) return fail(ctx, OD_TDLT_EINVAL, "frame too large for the coding-order walk");
  P.recip_nhsb = recip32(ctx->nhsb);
  DeviceGuard dev_guard(ctx->device);
  if (!dev_guard.ok) return fail(ctx, OD_TDLT_EFAULT, "cudaSetDevice(%d) failed", ctx->device);
  launch_coding_order(P, nframes, (cudaStream_t)stream);
  g_launches++;
  CK(ctx, cudaGetLastError());
  return OD_TDLT_SUCCESS;
}

}  // extern "C"

// ------------------------------------------------- whole-frame host calls --
static int check_frames(od_tdlt_ctx *ctx, int nframes) {
  if (nframes < 1 || nframes > ctx->max_frames) {
    return fail(ctx, OD_TDLT_EINVAL, "nframes %d outside 1..max_frames %d", nframes,
                ctx->max_frames);
  }
  return OD_TDLT_SUCCESS;
}

// One host call = a pipeline over chunks of frames on three streams: the upload
// of chunk k+1 (host -> device), the kernels of chunk k and the download of chunk
// k-1 (device -> host) overlap, so a call costs about max(upload, download) of
// the PCIe link instead of their sum.  Chunks are independent (frames are), and
// every chunk has its own slice of the staging buffers.
struct FramesJob {
  bool inverse;
  bool pix;                       // the spatial side is 8/16-bit reference-buffer samples
  od_coeff *const *coef_out;      // host planes written   (forward: d, inverse without pix: c)
  const od_coeff *const *coef_in; // host planes read      (forward without pix: c, inverse: d)
  void *const *pix_out;           // inverse + pix
  const void *const *pix_in;      // forward + pix
  const int *ystride;
  const long long *frame_stride;
  const unsigned char *bsize;
  int nframes, flags, xstride, bitdepth_mode, lossless_p;
};

// Queues the whole pipeline; any error returns at once -- run_frames below drains the
// streams whatever happened, so no copy is still touching the caller's buffers when
// the call returns.
static int run_frames_queue(od_tdlt_ctx *ctx, const FramesJob &j) {
  int rc = OD_TDLT_SUCCESS;
  const int np = ctx->g.nplanes;
  long long frame_samples = 0;
  for (int p = 0; p < np; p++) frame_samples += ctx->plane_size[p];
  // chunks of about 384 MB of coefficients, at most kMaxChunks of them (measured, 64 4K frames
  // per call, Mpixel/s end to end with 1 / 2 / 4 host threads: 96 MB 4447 / 5098 / 5385,
  // 200 MB 4434 / 5157 / 5495, 400 MB 4376 / 5153 / 5558, 800 MB - / - / 5564)
#ifndef TDLT_CHUNK_MB
#define TDLT_CHUNK_MB 384
#endif
  long long per = ((long long)TDLT_CHUNK_MB << 20)/(frame_samples*(long long)sizeof(od_coeff));
  int chunk = (int)(per < 1 ? 1 : per);
  if ((j.nframes + chunk - 1)/chunk > kMaxChunks) chunk = (j.nframes + kMaxChunks - 1)/kMaxChunks;
  const long long bsz = od_tdlt_bsize_size(ctx);
  cudaStream_t s_in = ctx->s_in, s_run = ctx->stream, s_out = ctx->s_out;
  int k = 0;
  for (int f0 = 0; f0 < j.nframes; f0 += chunk, k++) {
    const int n = (j.nframes - f0 < chunk) ? j.nframes - f0 : chunk;
    od_coeff *dc[kMaxPlanes], *dd[kMaxPlanes];
    for (int p = 0; p < np; p++) {
      dc[p] = ctx->dev_c[p] + (size_t)f0*ctx->plane_size[p];
      dd[p] = ctx->dev_d[p] + (size_t)f0*ctx->plane_size[p];
    }
    unsigned char *db = ctx->dev_bsize + (size_t)f0*bsz;
    // ---- upload -----------------------------------------------------------
    for (int p = 0; p < np; p++) {
      const size_t bytes = (size_t)ctx->plane_size[p]*n*sizeof(od_coeff);
      if (j.inverse) {
        CK(ctx, cudaMemcpyAsync(dd[p], j.coef_in[p] + (size_t)f0*ctx->plane_size[p], bytes,
                                cudaMemcpyHostToDevice, s_in));
      }
      else if (!j.pix) {
        CK(ctx, cudaMemcpyAsync(dc[p], j.coef_in[p] + (size_t)f0*ctx->plane_size[p], bytes,
                                cudaMemcpyHostToDevice, s_in));
      }
      else {
        const RegionPlane &pd = ctx->rb.pl[p];
        const size_t row = (size_t)pd.w*j.xstride;
        if ((size_t)j.ystride[p] == row && (size_t)j.frame_stride[p] == row*pd.h) {
          // dense frames one after the other: the chunk's pixels in one copy
          CK(ctx, cudaMemcpyAsync((char *)ctx->dev_pix[p] + (size_t)f0*row*pd.h,
                                  (const char *)j.pix_in[p] + (size_t)f0*row*pd.h, row*pd.h*n,
                                  cudaMemcpyHostToDevice, s_in));
        }
        else for (int f = f0; f < f0 + n; f++) {
          CK(ctx, cudaMemcpy2DAsync((char *)ctx->dev_pix[p] + (size_t)f*row*pd.h, row,
                                    (const char *)j.pix_in[p] + f*j.frame_stride[p], j.ystride[p],
                                    row, pd.h, cudaMemcpyHostToDevice, s_in));
        }
      }
    }
    CK(ctx, cudaMemcpyAsync(db, j.bsize + (size_t)f0*bsz, (size_t)bsz*n, cudaMemcpyHostToDevice, s_in));
    CK(ctx, cudaEventRecord(ctx->ev_in[k], s_in));
    // ---- kernels ------------------------------------------------------------
    CK(ctx, cudaStreamWaitEvent(s_run, ctx->ev_in[k], 0));
    // the pixel conversion is fused into the transform kernels' load / store; lossless mode
    // (another kernel pair) keeps the separate conversion passes
    const bool fused = j.pix && !(j.flags & OD_TDLT_HAAR);
    void *dpix[kMaxPlanes];
    for (int p = 0; p < np; p++) dpix[p] = (char *)ctx->dev_pix[p] + (size_t)f0*ctx->plane_size[p]*j.xstride;
    if (fused) {
      rc = j.inverse ? od_tdlt_inv_batch_pix(ctx, dpix, dd, db, n, j.flags, j.xstride, j.bitdepth_mode,
                                             j.lossless_p, s_run)
                     : od_tdlt_fwd_batch_pix(ctx, dd, dpix, db, n, j.flags, j.xstride, j.bitdepth_mode,
                                             j.lossless_p, s_run);
      if (rc) return rc;
    }
    else if (!j.inverse) {
      if (j.pix) {
        for (int p = 0; p < np; p++) {
          rc = od_tdlt_pix_to_coeff_batch(ctx, dc[p],
              (char *)ctx->dev_pix[p] + (size_t)f0*ctx->plane_size[p]*j.xstride, p, n, j.xstride,
              j.bitdepth_mode, j.lossless_p, s_run);
          if (rc) return rc;
        }
      }
      rc = od_tdlt_fwd_batch(ctx, dd, dc, db, n, j.flags, s_run);
      if (rc) return rc;
    }
    else {
      rc = od_tdlt_inv_batch(ctx, dc, dd, db, n, j.flags, s_run);
      if (rc) return rc;
      if (j.pix) {
        for (int p = 0; p < np; p++) {
          rc = od_tdlt_coeff_to_pix_batch(ctx,
              (char *)ctx->dev_pix[p] + (size_t)f0*ctx->plane_size[p]*j.xstride, dc[p], p, n,
              j.xstride, j.bitdepth_mode, j.lossless_p, s_run);
          if (rc) return rc;
        }
      }
    }
    CK(ctx, cudaEventRecord(ctx->ev_run[k], s_run));
    // ---- download -----------------------------------------------------------
    CK(ctx, cudaStreamWaitEvent(s_out, ctx->ev_run[k], 0));
    for (int p = 0; p < np; p++) {
      const size_t bytes = (size_t)ctx->plane_size[p]*n*sizeof(od_coeff);
      if (!j.inverse) {
        CK(ctx, cudaMemcpyAsync(j.coef_out[p] + (size_t)f0*ctx->plane_size[p], dd[p], bytes,
                                cudaMemcpyDeviceToHost, s_out));
      }
      else if (!j.pix) {
        CK(ctx, cudaMemcpyAsync(j.coef_out[p] + (size_t)f0*ctx->plane_size[p], dc[p], bytes,
                                cudaMemcpyDeviceToHost, s_out));
      }
      else {
        const RegionPlane &pd = ctx->rb.pl[p];
        const size_t row = (size_t)pd.w*j.xstride;
        if ((size_t)j.ystride[p] == row && (size_t)j.frame_stride[p] == row*pd.h) {
          CK(ctx, cudaMemcpyAsync((char *)j.pix_out[p] + (size_t)f0*row*pd.h,
                                  (const char *)ctx->dev_pix[p] + (size_t)f0*row*pd.h, row*pd.h*n,
                                  cudaMemcpyDeviceToHost, s_out));
        }
        else for (int f = f0; f < f0 + n; f++) {
          CK(ctx, cudaMemcpy2DAsync((char *)j.pix_out[p] + f*j.frame_stride[p], j.ystride[p],
                                    (const char *)ctx->dev_pix[p] + (size_t)f*row*pd.h, row, row,
                                    pd.h, cudaMemcpyDeviceToHost, s_out));
        }
      }
    }
  }
  return OD_TDLT_SUCCESS;
}

static int run_frames(od_tdlt_ctx *ctx, const FramesJob &j) {
  int rc = check_frames(ctx, j.nframes);
  if (rc) return rc;
  DeviceGuard dev_guard(ctx->device);
  if (!dev_guard.ok) return fail(ctx, OD_TDLT_EFAULT, "cudaSetDevice(%d) failed", ctx->device);
  rc = run_frames_queue(ctx, j);
  // drain all three streams on every path (an error above must not leave copies in flight)
  const cudaError_t e0 = cudaStreamSynchronize(ctx->s_in);
  const cudaError_t e1 = cudaStreamSynchronize(ctx->stream);
  const cudaError_t e2 = cudaStreamSynchronize(ctx->s_out);
  if (rc) return rc;
  const cudaError_t e = e0 != cudaSuccess ? e0 : (e1 != cudaSuccess ? e1 : e2);
  if (e != cudaSuccess) return fail(ctx, OD_TDLT_EFAULT, "host-pointer call: %s", cudaGetErrorString(e));
  return OD_TDLT_SUCCESS;
}

extern "C" {

// Page-locks caller memory so that the host-pointer calls can overlap their copies with the
// kernels (the reference's planes come from plain malloc, src/state.c:409-428: from pageable
// memory every cudaMemcpyAsync is staged by the driver and the three-stream pipeline of
// run_frames degenerates to one copy at a time).
int od_tdlt_host_register(void *ptr, size_t bytes) {
  if (!ptr || !bytes) return OD_TDLT_EFAULT;
  cudaError_t e = cudaHostRegister(ptr, bytes, cudaHostRegisterPortable);
  if (e == cudaErrorHostMemoryAlreadyRegistered) { cudaGetLastError(); return OD_TDLT_SUCCESS; }
  if (e != cudaSuccess) { cudaGetLastError(); return e == cudaErrorInvalidValue ? OD_TDLT_EINVAL : OD_TDLT_EFAULT; }
  return OD_TDLT_SUCCESS;
}
int od_tdlt_host_unregister(void *ptr) {
  if (!ptr) return OD_TDLT_EFAULT;
  cudaError_t e = cudaHostUnregister(ptr);
  if (e != cudaSuccess) { cudaGetLastError(); return OD_TDLT_EINVAL; }
  return OD_TDLT_SUCCESS;
}

int od_tdlt_fwd_frames(od_tdlt_ctx *ctx, od_coeff *const d[], const od_coeff *const c[],
                       const unsigned char *bsize, int nframes, int flags) {
  if (!ctx || !d || !c || !bsize) return OD_TDLT_EFAULT;
  FramesJob j = {};
  j.coef_out = d; j.coef_in = c; j.bsize = bsize; j.nframes = nframes; j.flags = flags;
  return run_frames(ctx, j);
}

int od_tdlt_inv_frames(od_tdlt_ctx *ctx, od_coeff *const c[], const od_coeff *const d[],
                       const unsigned char *bsize, int nframes, int flags) {
  if (!ctx || !d || !c || !bsize) return OD_TDLT_EFAULT;
  FramesJob j = {};
  j.inverse = true;
  j.coef_out = c; j.coef_in = d; j.bsize = bsize; j.nframes = nframes; j.flags = flags;
  return run_frames(ctx, j);
}

int od_tdlt_cfl_pred_frames(od_tdlt_ctx *ctx, od_coeff *l, const od_coeff *d_luma,
                            const unsigned char *bsize, int nframes) {
  if (!ctx || !l || !d_luma || !bsize) return OD_TDLT_EFAULT;
  int rc = check_frames(ctx, nframes);
  if (rc) return rc;
  if (ctx->g.nplanes < 2) return fail(ctx, OD_TDLT_EINVAL, "no chroma plane");
  DeviceGuard dev_guard(ctx->device);
  if (!dev_guard.ok) return fail(ctx, OD_TDLT_EFAULT, "cudaSetDevice(%d) failed", ctx->device);
  // staging: the luma coefficient plane and the predictor plane in the slots the
  // transform calls use for them (plane 0 coefficients, plane 1 samples)
  cudaStream_t st = ctx->stream;
  CK(ctx, cudaMemcpyAsync(ctx->dev_d[0], d_luma, (size_t)ctx->plane_size[0]*nframes*sizeof(od_coeff),
                          cudaMemcpyHostToDevice, st));
  CK(ctx, cudaMemcpyAsync(ctx->dev_bsize, bsize, (size_t)od_tdlt_bsize_size(ctx)*nframes,
                          cudaMemcpyHostToDevice, st));
  rc = od_tdlt_cfl_pred_batch(ctx, ctx->dev_c[1], ctx->dev_d[0], ctx->dev_bsize, nframes, st);
  if (rc) return rc;
  CK(ctx, cudaMemcpyAsync(l, ctx->dev_c[1], (size_t)ctx->plane_size[1]*nframes*sizeof(od_coeff),
                          cudaMemcpyDeviceToHost, st));
  CK(ctx, cudaStreamSynchronize(st));
  return OD_TDLT_SUCCESS;
}

int od_tdlt_dering_frames(od_tdlt_ctx *ctx, od_coeff *const out[], const od_coeff *const in[],
                          const unsigned char *flags, const unsigned char *const bskip[],
                          const int q[], int nframes) {
  if (!ctx || !out || !in || !flags || !bskip || !q) return OD_TDLT_EFAULT;
  int rc = check_frames(ctx, nframes);
  if (rc) return rc;
  DeviceGuard dev_guard(ctx->device);
  if (!dev_guard.ok) return fail(ctx, OD_TDLT_EFAULT, "cudaSetDevice(%d) failed", ctx->device);
  const int np = ctx->g.nplanes;
  const size_t nsb = (size_t)ctx->nhsb*ctx->nvsb*nframes;
  const size_t skip = (size_t)ctx->nhsb*8*ctx->nvsb*8*nframes;
  const size_t want = nsb + skip*np;
  if (want > ctx->dev_aux_bytes) {
    cudaFree(ctx->dev_aux);
    ctx->dev_aux = nullptr;
    ctx->dev_aux_bytes = 0;
    CK(ctx, cudaMalloc(&ctx->dev_aux, want));
    ctx->dev_aux_bytes = want;
  }
  cudaStream_t st = ctx->stream;
  const unsigned char *dskip[kMaxPlanes];
  CK(ctx, cudaMemcpyAsync(ctx->dev_aux, flags, nsb, cudaMemcpyHostToDevice, st));
  for (int p = 0; p < np; p++) {
    if (!out[p] || !in[p] || !bskip[p]) return OD_TDLT_EFAULT;
    unsigned char *ds = ctx->dev_aux + nsb + skip*p;
    dskip[p] = ds;
    CK(ctx, cudaMemcpyAsync(ds, bskip[p], skip, cudaMemcpyHostToDevice, st));
    CK(ctx, cudaMemcpyAsync(ctx->dev_c[p], in[p], (size_t)ctx->plane_size[p]*nframes*sizeof(od_coeff),
                            cudaMemcpyHostToDevice, st));
  }
  rc = od_tdlt_dering_batch(ctx, ctx->dev_d, ctx->dev_c, ctx->dev_aux, dskip, q, nframes, st);
  if (rc) return rc;
  for (int p = 0; p < np; p++) {
    CK(ctx, cudaMemcpyAsync(out[p], ctx->dev_d[p], (size_t)ctx->plane_size[p]*nframes*sizeof(od_coeff),
                            cudaMemcpyDeviceToHost, st));
  }
  CK(ctx, cudaStreamSynchronize(st));
  return OD_TDLT_SUCCESS;
}

int od_tdlt_smooth_frames(od_tdlt_ctx *ctx, od_coeff *const c[], const unsigned char *bsize,
                          const int q[], int nframes) {
  if (!ctx || !c || !bsize || !q) return OD_TDLT_EFAULT;
  int rc = check_frames(ctx, nframes);
  if (rc) return rc;
  DeviceGuard dev_guard(ctx->device);
  if (!dev_guard.ok) return fail(ctx, OD_TDLT_EFAULT, "cudaSetDevice(%d) failed", ctx->device);
  cudaStream_t st = ctx->stream;
  const int np = ctx->g.nplanes;
  for (int p = 0; p < np; p++) {
    if (!c[p]) return OD_TDLT_EFAULT;
    CK(ctx, cudaMemcpyAsync(ctx->dev_c[p], c[p], (size_t)ctx->plane_size[p]*nframes*sizeof(od_coeff),
                            cudaMemcpyHostToDevice, st));
  }
  CK(ctx, cudaMemcpyAsync(ctx->dev_bsize, bsize, (size_t)od_tdlt_bsize_size(ctx)*nframes,
                          cudaMemcpyHostToDevice, st));
  rc = od_tdlt_smooth_batch(ctx, ctx->dev_c, ctx->dev_bsize, q, nframes, st);
  if (rc) return rc;
  for (int p = 0; p < np; p++) {
    CK(ctx, cudaMemcpyAsync(c[p], ctx->dev_c[p], (size_t)ctx->plane_size[p]*nframes*sizeof(od_coeff),
                            cudaMemcpyDeviceToHost, st));
  }
  CK(ctx, cudaStreamSynchronize(st));
  return OD_TDLT_SUCCESS;
}

int od_tdlt_coding_order_frames(od_tdlt_ctx *ctx, od_coeff *const v[], od_coeff *const d[],
                                const unsigned char *bsize, int nframes, int to_coding) {
  if (!ctx || !v || !d || !bsize) return OD_TDLT_EFAULT;
  int rc = check_frames(ctx, nframes);
  if (rc) return rc;
  DeviceGuard dev_guard(ctx->device);
  if (!dev_guard.ok) return fail(ctx, OD_TDLT_EFAULT, "cudaSetDevice(%d) failed", ctx->device);
  cudaStream_t st = ctx->stream;
  const int np = ctx->g.nplanes;
  // both sides go up: entries the scans do not cover keep what the destination held
  for (int p = 0; p < np; p++) {
    if (!v[p] || !d[p]) return OD_TDLT_EFAULT;
    const size_t bytes = (size_t)ctx->plane_size[p]*nframes*sizeof(od_coeff);
    CK(ctx, cudaMemcpyAsync(ctx->dev_c[p], v[p], bytes, cudaMemcpyHostToDevice, st));
    CK(ctx, cudaMemcpyAsync(ctx->dev_d[p], d[p], bytes, cudaMemcpyHostToDevice, st));
  }
  CK(ctx, cudaMemcpyAsync(ctx->dev_bsize, bsize, (size_t)od_tdlt_bsize_size(ctx)*nframes,
                          cudaMemcpyHostToDevice, st));
  rc = od_tdlt_coding_order_batch(ctx, ctx->dev_c, ctx->dev_d, ctx->dev_bsize, nframes, to_coding, st);
  if (rc) return rc;
  for (int p = 0; p < np; p++) {
    const size_t bytes = (size_t)ctx->plane_size[p]*nframes*sizeof(od_coeff);
    CK(ctx, cudaMemcpyAsync(to_coding ? v[p] : d[p], to_coding ? ctx->dev_c[p] : ctx->dev_d[p], bytes,
                            cudaMemcpyDeviceToHost, st));
  }
  CK(ctx, cudaStreamSynchronize(st));
  return OD_TDLT_SUCCESS;
}

int od_tdlt_fwd_frames_pix(od_tdlt_ctx *ctx, od_coeff *const d[], const void *const pix[],
                           const int ystride[], const long long frame_stride[],
                           const unsigned char *bsize, int nframes, int flags, int xstride,
                           int bitdepth_mode, int lossless_p) {
  if (!ctx || !d || !pix || !ystride || !frame_stride || !bsize) return OD_TDLT_EFAULT;
  FramesJob j = {};
  j.pix = true;
  j.coef_out = d; j.pix_in = pix; j.ystride = ystride; j.frame_stride = frame_stride;
  j.bsize = bsize; j.nframes = nframes; j.flags = flags; j.xstride = xstride;
  j.bitdepth_mode = bitdepth_mode; j.lossless_p = lossless_p;
  return run_frames(ctx, j);
}

int od_tdlt_inv_frames_pix(od_tdlt_ctx *ctx, void *const pix[], const int ystride[],
                           const long long frame_stride[], const od_coeff *const d[],
                           const unsigned char *bsize, int nframes, int flags, int xstride,
                           int bitdepth_mode, int lossless_p) {
  if (!ctx || !d || !pix || !ystride || !frame_stride || !bsize) return OD_TDLT_EFAULT;
  FramesJob j = {};
  j.inverse = true;
  j.pix = true;
  j.pix_out = pix; j.coef_in = d; j.ystride = ystride; j.frame_stride = frame_stride;
  j.bsize = bsize; j.nframes = nframes; j.flags = flags; j.xstride = xstride;
  j.bitdepth_mode = bitdepth_mode; j.lossless_p = lossless_p;
  return run_frames(ctx, j);
}

}  // extern "C"

// ------------------------------------------- reference-signature wrappers --
// Host pointers in, host pointers out, synchronous, on the current device.
// A failure here cannot be reported through the reference's void signatures,
// so it aborts (the reference's OD_FATAL convention, src/internal.h:128-158).
static void die(const char *what, cudaError_t e) {
  fprintf(stderr, "daala_tdlt (CUDA): fatal: %s: %s\n", what, cudaGetErrorString(e));
  abort();
}
#define CKF(call) \
  do { cudaError_t e_ = (call); if (e_ != cudaSuccess) die(#call, e_); } while (0)

namespace {
// Scratch of the reference-signature wrappers: one pair of buffers per host thread AND per
// device (the wrappers run on whatever device is current, as the reference's functions run
// on whatever core calls them), so two threads or two GPUs never share or free each other's
// memory.  Buffers live until the process ends.
struct Scratch {
  static constexpr int kMaxDev = 64;
  int *a[kMaxDev] = {}, *b[kMaxDev] = {};
  size_t na[kMaxDev] = {}, nb[kMaxDev] = {};
  static int dev() {
    int d = 0;
    CKF(cudaGetDevice(&d));
    if (d < 0 || d >= kMaxDev) die("device index outside the scratch table", cudaErrorInvalidDevice);
    return d;
  }
  int *get(int *&p, size_t &n, size_t want) {
    if (want > n) {
      if (p) CKF(cudaFree(p));
      p = nullptr;
      n = 0;
      CKF(cudaMalloc(&p, want*sizeof(int)));
      n = want;
    }
    return p;
  }
  int *A(size_t want) { const int d = dev(); return get(a[d], na[d], want); }
  int *B(size_t want) { const int d = dev(); return get(b[d], nb[d], want); }
};
thread_local Scratch g_scratch;

// Copies a strided host rectangle (rows x cols ints) to a dense device buffer.
void h2d_rect(int *dev, const od_coeff *host, int stride, int rows, int cols) {
  CKF(cudaMemcpy2D(dev, (size_t)cols*4, host, (size_t)stride*4, (size_t)cols*4, rows,
                   cudaMemcpyHostToDevice));
}
void d2h_rect(od_coeff *host, int stride, const int *dev, int rows, int cols) {
  CKF(cudaMemcpy2D(host, (size_t)stride*4, dev, (size_t)cols*4, (size_t)cols*4, rows,
                   cudaMemcpyDeviceToHost));
}

void dct_one(int bs, bool inverse, od_coeff *out, int ostride, const od_coeff *in, int istride) {
  int n = 4 << bs;
  int *din = g_scratch.A((size_t)n*n), *dout = g_scratch.B((size_t)n*n);
  h2d_rect(din, in, istride, n, n);
  launch_dct2d_blocks(bs, inverse, dout, n, 0, din, n, 0, 1, 0);
  g_launches++;
  CKF(cudaGetLastError());
  d2h_rect(out, ostride, dout, n, n);
}
}  // namespace

extern "C" {

void od_init_skipped_coeffs_cuda(od_coeff *d, od_coeff *pred, int is_keyframe, int bo, int n, int w) {
  if (n < 1 || n > 64 || w < n) die("od_init_skipped_coeffs: bad block", cudaErrorInvalidValue);
  int *dd = g_scratch.A((size_t)n*n), *dp = g_scratch.B((size_t)n*n);
  h2d_rect(dd, d + bo, w, n, n);
  if (!is_keyframe) CKF(cudaMemcpy(dp, pred, sizeof(int)*n*n, cudaMemcpyHostToDevice));
  launch_init_skipped(dd, dp, is_keyframe, n, n, 0);
  g_launches++;
  CKF(cudaGetLastError());
  d2h_rect(d + bo, w, dd, n, n);
}

void od_bin_fdct4x4_cuda(od_coeff *y, int ys, const od_coeff *x, int xs) { dct_one(0, false, y, ys, x, xs); }
void od_bin_idct4x4_cuda(od_coeff *x, int xs, const od_coeff *y, int ys) { dct_one(0, true, x, xs, y, ys); }
void od_bin_fdct8x8_cuda(od_coeff *y, int ys, const od_coeff *x, int xs) { dct_one(1, false, y, ys, x, xs); }
void od_bin_idct8x8_cuda(od_coeff *x, int xs, const od_coeff *y, int ys) { dct_one(1, true, x, xs, y, ys); }
void od_bin_fdct16x16_cuda(od_coeff *y, int ys, const od_coeff *x, int xs) { dct_one(2, false, y, ys, x, xs); }
void od_bin_idct16x16_cuda(od_coeff *x, int xs, const od_coeff *y, int ys) { dct_one(2, true, x, xs, y, ys); }
void od_bin_fdct32x32_cuda(od_coeff *y, int ys, const od_coeff *x, int xs) { dct_one(3, false, y, ys, x, xs); }
void od_bin_idct32x32_cuda(od_coeff *x, int xs, const od_coeff *y, int ys) { dct_one(3, true, x, xs, y, ys); }

// The deringing filters of the reference's dispatch table (src/filter.h:46-65,
// src/filter.c:1715-1800), one block, host pointers, synchronous.  `in` points at the
// block inside a buffer of row pitch OD_FILT_BSTRIDE = 38 with at least three samples
// around it, exactly as od_dering passes it (src/filter.c:1922-1940).
static void dering_block(int n, bool orthogonal, int16_t *y, int ystride, const int16_t *in,
                         const int16_t *x, int xstride, int threshold, int dir) {
  if (dir < 0 || dir > 7) die("od_filter_dering_*: direction outside 0..7", cudaErrorInvalidValue);
  const int kInPitch = 38, span = n + 6;
  short *d = reinterpret_cast<short *>(g_scratch.A(256));
  short *din = d, *dx = d + 200, *dy = d + 300;       // 14*14 + 8*8 + 8*8 shorts
  CKF(cudaMemcpy2D(din, (size_t)span*2, in - 3*kInPitch - 3, (size_t)kInPitch*2, (size_t)span*2, span,
                   cudaMemcpyHostToDevice));
  if (orthogonal) {
    CKF(cudaMemcpy2D(dx, (size_t)n*2, x, (size_t)xstride*2, (size_t)n*2, n, cudaMemcpyHostToDevice));
  }
  launch_dering_block(dy, din, dx, n, threshold, dir, orthogonal, 0);
  g_launches++;
  CKF(cudaGetLastError());
  CKF(cudaMemcpy2D(y, (size_t)ystride*2, dy, (size_t)n*2, (size_t)n*2, n, cudaMemcpyDeviceToHost));
}
void od_filter_dering_direction_4x4_cuda(int16_t *y, int ystride, int16_t *in, int threshold, int dir) {
  dering_block(4, false, y, ystride, in, nullptr, 0, threshold, dir);
}
void od_filter_dering_direction_8x8_cuda(int16_t *y, int ystride, int16_t *in, int threshold, int dir) {
  dering_block(8, false, y, ystride, in, nullptr, 0, threshold, dir);
}
void od_filter_dering_orthogonal_4x4_cuda(int16_t *y, int ystride, int16_t *in, int16_t *x, int xstride,
                                          int threshold, int dir) {
  dering_block(4, true, y, ystride, in, x, xstride, threshold, dir);
}
void od_filter_dering_orthogonal_8x8_cuda(int16_t *y, int ystride, int16_t *in, int16_t *x, int xstride,
                                          int threshold, int dir) {
  dering_block(8, true, y, ystride, in, x, xstride, threshold, dir);
}
void od_tdlt_dering_vtbl_init_cuda(od_tdlt_dering_direction_func direction[],
                                   od_tdlt_dering_orthogonal_func orthogonal[]) {
  direction[0] = od_filter_dering_direction_4x4_cuda;
  direction[1] = od_filter_dering_direction_8x8_cuda;
  orthogonal[0] = od_filter_dering_orthogonal_4x4_cuda;
  orthogonal[1] = od_filter_dering_orthogonal_8x8_cuda;
}

void od_tdlt_opt_vtbl_init_cuda(od_tdlt_dct_func_2d fdct_2d[], od_tdlt_dct_func_2d idct_2d[]) {
  fdct_2d[0] = od_bin_fdct4x4_cuda;   idct_2d[0] = od_bin_idct4x4_cuda;
  fdct_2d[1] = od_bin_fdct8x8_cuda;   idct_2d[1] = od_bin_idct8x8_cuda;
  fdct_2d[2] = od_bin_fdct16x16_cuda; idct_2d[2] = od_bin_idct16x16_cuda;
  fdct_2d[3] = od_bin_fdct32x32_cuda; idct_2d[3] = od_bin_idct32x32_cuda;
}

int od_tdlt_dct2d_blocks(int bs, int inverse, od_coeff *out, int ostride, long long oblock,
                         const od_coeff *in, int istride, long long iblock, int nblocks) {
  if (bs < 0 || bs > 3 || nblocks < 1 || !out || !in) return OD_TDLT_EINVAL;
  int n = 4 << bs;
  // Stage densely: block i at i*n*n.
  size_t tot = (size_t)nblocks*n*n;
  int *din = g_scratch.A(tot), *dout = g_scratch.B(tot);
  // tightly packed blocks travel in one copy
  if (istride == n && iblock == (long long)n*n) CKF(cudaMemcpy(din, in, tot*4, cudaMemcpyHostToDevice));
  else for (int i = 0; i < nblocks; i++) h2d_rect(din + (size_t)i*n*n, in + i*iblock, istride, n, n);
  launch_dct2d_blocks(bs, inverse != 0, dout, n, (long long)n*n, din, n, (long long)n*n, nblocks, 0);
  g_launches++;
  if (cudaGetLastError() != cudaSuccess) return OD_TDLT_EFAULT;
  if (ostride == n && oblock == (long long)n*n) CKF(cudaMemcpy(out, dout, tot*4, cudaMemcpyDeviceToHost));
  else for (int i = 0; i < nblocks; i++) d2h_rect(out + i*oblock, ostride, dout + (size_t)i*n*n, n, n);
  return OD_TDLT_SUCCESS;
}

int od_tdlt_dct1d_vectors(int bs, int inverse, od_coeff *out, const od_coeff *in, int nvec) {
  if (bs < 0 || bs > 3 || nvec < 1 || !out || !in) return OD_TDLT_EINVAL;
  size_t tot = (size_t)nvec*(4 << bs);
  int *din = g_scratch.A(tot), *dout = g_scratch.B(tot);
  CKF(cudaMemcpy(din, in, tot*4, cudaMemcpyHostToDevice));
  launch_dct1d_vectors(bs, inverse != 0, dout, din, nvec, 0);
  g_launches++;
  if (cudaGetLastError() != cudaSuccess) return OD_TDLT_EFAULT;
  CKF(cudaMemcpy(out, dout, tot*4, cudaMemcpyDeviceToHost));
  return OD_TDLT_SUCCESS;
}

int od_tdlt_lap_filter_vectors(int f, int inverse, od_coeff *out, const od_coeff *in, int nvec) {
  if (f < 0 || f > 3 || nvec < 1 || !out || !in) return OD_TDLT_EINVAL;
  size_t tot = (size_t)nvec*(4 << f);
  int *din = g_scratch.A(tot), *dout = g_scratch.B(tot);
  CKF(cudaMemcpy(din, in, tot*4, cudaMemcpyHostToDevice));
  launch_lap_vectors(f, inverse != 0, dout, din, nvec, 0);
  g_launches++;
  if (cudaGetLastError() != cudaSuccess) return OD_TDLT_EFAULT;
  CKF(cudaMemcpy(out, dout, tot*4, cudaMemcpyDeviceToHost));
  return OD_TDLT_SUCCESS;
}

static void haar_one(bool inverse, od_coeff *out, int ostride, const od_coeff *in, int istride,
                     int ln) {
  int n = 1 << ln;
  if (ln < 1 || ln > 5) die("od_haar: ln out of range", cudaErrorInvalidValue);
  int *din = g_scratch.A((size_t)n*n), *dout = g_scratch.B((size_t)n*n);
  h2d_rect(din, in, istride, n, n);
  launch_haar_block(inverse, dout, n, din, n, ln, 0);
  g_launches++;
  CKF(cudaGetLastError());
  d2h_rect(out, ostride, dout, n, n);
}
void od_haar_cuda(od_coeff *y, int ys, const od_coeff *x, int xs, int ln) { haar_one(false, y, ys, x, xs, ln); }
void od_haar_inv_cuda(od_coeff *x, int xs, const od_coeff *y, int ys, int ln) { haar_one(true, x, xs, y, ys, ln); }

static void tf_one(bool down, od_coeff *dst, int dstride, const od_coeff *src, int sstride,
                   int n) {
  // up: n = size of each of the 2x2 source blocks; down: n = size of the source block.
  int full = down ? n : 2*n;
  int half = full/2;
  int *din = g_scratch.A((size_t)full*full), *dout = g_scratch.B((size_t)full*full);
  h2d_rect(din, src, sstride, full, full);
  launch_tf_hv(down, dout, full, din, full, half, 0);
  g_launches++;
  CKF(cudaGetLastError());
  d2h_rect(dst, dstride, dout, full, full);
}
void od_tf_up_hv_cuda(od_coeff *dst, int ds, const od_coeff *src, int ss, int n) { tf_one(false, dst, ds, src, ss, n); }
void od_tf_down_hv_cuda(od_coeff *dst, int ds, const od_coeff *src, int ss, int n) { tf_one(true, dst, ds, src, ss, n); }

// The other helpers of src/tf.c: source extent sw x sh, result n x n.
static void tf_op(int op, od_coeff *dst, int dstride, const od_coeff *src, int sstride, int n,
                  int sw, int sh, int p0, int p1, int p2) {
  if (n < 2 || n > 32 || sw > 64 || sh > 64) die("od_tf_*: block too large", cudaErrorInvalidValue);
  int *din = g_scratch.A((size_t)sw*sh), *dout = g_scratch.B((size_t)n*n);
  h2d_rect(din, src, sstride, sh, sw);
  launch_tf_op(op, dout, n, din, sw, n, p0, p1, p2, 0);
  g_launches++;
  CKF(cudaGetLastError());
  d2h_rect(dst, dstride, dout, n, n);
}
void od_tf_up_h_lp_cuda(od_coeff *dst, int ds, const od_coeff *src, int ss, int dx, int n) {
  tf_op(0, dst, ds, src, ss, n, (n >> 1) + dx, n, dx, 0, 0);
}
void od_tf_up_v_lp_cuda(od_coeff *dst, int ds, const od_coeff *src, int ss, int dy, int n) {
  tf_op(1, dst, ds, src, ss, n, n, (n >> 1) + dy, dy, 0, 0);
}
void od_tf_up_hv_lp_cuda(od_coeff *dst, int ds, const od_coeff *src, int ss, int dx, int dy,
                         int n) {
  tf_op(2, dst, ds, src, ss, n, (n >> 1) + dx, (n >> 1) + dy, dx, dy, 0);
}
// reference src/intra.c:72-109, one block.  The block goes through the frame kernel as a
// frame of its own: n/4 x n/4 units whose map entries make the kernel take the same branch.
void od_resample_luma_coeffs_cuda(od_coeff *chroma_pred, int cpstride, const od_coeff *decoded_luma,
                                  int dlstride, int xdec, int ydec, int bs, int chroma_bs) {
  if (xdec != ydec || bs < 0 || bs > 3) {
    die("od_resample_luma_coeffs: 4:2:0 / 4:4:4 and bs 0..3 only (the reference's callers "
        "assert the same, src/encode.c:1499)", cudaErrorInvalidValue);
  }
  int n = 4 << bs;
  bool merge = chroma_bs == 0 && xdec;
  if (merge && bs != 0) die("od_resample_luma_coeffs: merged block larger than 4x4", cudaErrorInvalidValue);
  int units = merge ? 1 : n >> 2;
  int ln = merge ? 8 : n;                     // luma samples read per row / column
  int *din = g_scratch.A((size_t)ln*ln + 16), *dout = g_scratch.B((size_t)n*n);
  unsigned char map[64];
  memset(map, merge ? 0 : bs + 1, sizeof(map));
  unsigned char *dmap = reinterpret_cast<unsigned char *>(din + (size_t)ln*ln);
  h2d_rect(din, decoded_luma, dlstride, ln, ln);
  CKF(cudaMemcpy(dmap, map, (size_t)units*units, cudaMemcpyHostToDevice));
  launch_cfl_pred(dout, n, 0, din, ln, 0, dmap, units, units, 1, 0);
  g_launches++;
  CKF(cudaGetLastError());
  d2h_rect(chroma_pred, cpstride, dout, n, n);
}
void od_tf_filter_2d_cuda(od_coeff *dst, int ds, const od_coeff *src, int ss, int n) {
  tf_op(3, dst, ds, src, ss, n, n, n, 0, 0, 0);
}
void od_tf_filter_inv_2d_cuda(od_coeff *dst, int ds, const od_coeff *src, int ss, int n) {
  tf_op(4, dst, ds, src, ss, n, n, n, 0, 0, 0);
}
void od_convert_block_down_cuda(od_coeff *dst, int ds, const od_coeff *src, int ss, int curr_size,
                                int dest_size, int filter) {
  if (curr_size < dest_size || curr_size > 3 || dest_size < 0) {
    die("od_convert_block_down: bad sizes", cudaErrorInvalidValue);
  }
  int n = 4 << curr_size;
  tf_op(5, dst, ds, src, ss, n, n, n, curr_size, dest_size, filter);
}

// General f is supported through the N-tap vector kernel only for f == 0 here:
// the codec never uses another size on this path (OD_FILT_SIZE == 0,
// reference src/filter.h:99), and the reference asserts nothing else either.
static void split_filter(od_coeff *c0, int stride, int bs, int f, int hfilter, int vfilter,
                         bool inverse) {
  if (f != 0) die("od_*filter_split: only f == 0 (4-tap lapping) is on the path", cudaErrorInvalidValue);
  int n = 4 << bs;
  int *dev = g_scratch.A((size_t)n*n);
  h2d_rect(dev, c0, stride, n, n);
  int mid = (n >> 1) - 2;
  // forward: across the horizontal mid-edge first (down the columns), then across the
  // vertical one (reference src/filter.c:1493-1508); inverse: the other way round (:1537-1552).
  for (int step = 0; step < 2; step++) {
    bool cols = (step == 0) != inverse;
    if (cols && hfilter) {
      launch_lap4_lines(inverse, dev + mid*n, 1, n, n, 0);
      g_launches++;
    }
    if (!cols && vfilter) {
      launch_lap4_lines(inverse, dev + mid, n, 1, n, 0);
      g_launches++;
    }
  }
  CKF(cudaGetLastError());
  d2h_rect(c0, stride, dev, n, n);
}
void od_prefilter_split_cuda(od_coeff *c0, int stride, int bs, int f, int hfilter, int vfilter) {
  split_filter(c0, stride, bs, f, hfilter, vfilter, false);
}
void od_postfilter_split_cuda(od_coeff *c0, int stride, int bs, int f, int q, unsigned char *skip,
                              int skip_stride, int hfilter, int vfilter) {
  (void)q; (void)skip; (void)skip_stride;
  split_filter(c0, stride, bs, f, hfilter, vfilter, true);
}

static void frame_sbs(od_coeff *c, int stride, int nhsb, int nvsb, int xdec, int ydec,
                      bool inverse) {
  int w = (nhsb*32) >> xdec, h = (nvsb*32) >> ydec;
  int sbw = 32 >> xdec, sbh = 32 >> ydec;
  int *dev = g_scratch.A((size_t)w*h);
  h2d_rect(dev, c, stride, h, w);
  // forward: horizontal SB edges (down the columns) then vertical ones (reference
  // src/filter.c:1567-1584); inverse: vertical then horizontal (:1627-1644).
  for (int step = 0; step < 2; step++) {
    bool cols = (step == 0) != inverse;
    if (cols) {
      for (int sby = 1; sby < nvsb; sby++) {
        launch_lap4_lines(inverse, dev + (size_t)(sby*sbh - 2)*w, 1, w, w, 0);
        g_launches++;
      }
    }
    else {
      for (int sbx = 1; sbx < nhsb; sbx++) {
        launch_lap4_lines(inverse, dev + sbx*sbw - 2, w, 1, h, 0);
        g_launches++;
      }
    }
  }
  CKF(cudaGetLastError());
  d2h_rect(c, stride, dev, h, w);
}
void od_apply_prefilter_frame_sbs_cuda(od_coeff *c, int stride, int nhsb, int nvsb, int xdec,
                                       int ydec) {
  frame_sbs(c, stride, nhsb, nvsb, xdec, ydec, false);
}
void od_apply_postfilter_frame_sbs_cuda(od_coeff *c, int stride, int nhsb, int nvsb, int xdec,
                                        int ydec, int q, unsigned char *skip, int skip_stride) {
  (void)q; (void)skip; (void)skip_stride;
  frame_sbs(c, stride, nhsb, nvsb, xdec, ydec, true);
}

}  // extern "C"

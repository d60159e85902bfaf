/* daala_tdlt.h -- C-ABI of the B200 (sm_100a) implementation of Daala's
 * time-domain lapped transform stage.
 *
 * Everything here is plain C: pointers, ints, no CUDA or torch types (a CUDA
 * stream crosses the boundary as a void*).  Each entry point cites the
 * reference interface it replaces (KyleSiefring/daala, file:line).
 *
 * Results equal the reference's bit for bit for every input whose intermediate
 * products stay inside 32 bits in the reference's own transforms (what its
 * OD_DCT_OVERFLOW_CHECK asserts; 8- to 16-bit video is far inside).  Where the
 * reference's int arithmetic overflows its answer is that of signed overflow in
 * the compiled C and is not reproduced.
 *
 * Three layers, coarsest first:
 *   1. device-resident batch calls   (od_tdlt_fwd_batch / od_tdlt_inv_batch)
 *      -- what bench.py times as `value`; pointers are DEVICE pointers.
 *   2. whole-frame host calls        (od_tdlt_fwd_frames / od_tdlt_inv_frames,
 *      od_tdlt_fwd_frames_pix / od_tdlt_inv_frames_pix)
 *      -- the real drop-in for encode.c / decode.c; pointers are HOST pointers;
 *      uploads, kernels and downloads are pipelined over chunks of frames.
 *   3. reference-signature calls     (od_bin_fdct8x8_cuda, od_prefilter_split_cuda ...)
 *      -- same prototypes as the reference functions, host pointers,
 *      synchronous; they exist so the reference's own call sites and
 *      OD_CHECKASM-style cross-checks can run against the GPU.
 *
 * The same three forms exist for the neighbours of the path built so far:
 *   chroma-from-luma predictor planes  od_tdlt_cfl_pred_batch / _frames, od_resample_luma_coeffs_cuda
 *   deringing                          od_tdlt_dering_batch / _frames, od_filter_dering_*_cuda
 *   keyframe bilinear smoothing        od_tdlt_smooth_batch / _frames
 *   raster <-> coding order            od_tdlt_set_scan_tables, od_tdlt_coding_order_batch / _frames
 *
 * There is no CPU fallback anywhere in this library: a call either runs on
 * the GPU or fails (int-returning calls return OD_EFAULT; void calls abort,
 * mirroring OD_FATAL in reference src/internal.h:128-158).
 */
#ifndef DAALA_TDLT_H
#define DAALA_TDLT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* reference src/filter.h:31 */
typedef int32_t od_coeff;

/* reference include/daala/codec.h:89-103 */
#define OD_TDLT_SUCCESS (0)
#define OD_TDLT_EFAULT (-1)
#define OD_TDLT_EINVAL (-10)
#define OD_TDLT_EIMPL (-23)

#define OD_TDLT_NPLANES_MAX (4)

/* Flags for the frame / batch calls. */
#define OD_TDLT_KEYFRAME (1)   /* forward: Haar-merge child DCs (encode.c:1328-1341);
                                  inverse: undo it top-down (decode.c:733) */
#define OD_TDLT_HAAR     (2)   /* lossless mode: od_haar / od_haar_inv instead of
                                  lapping + DCT (encode.c:1304-1306, decode.c:632-634) */
#define OD_TDLT_SAME_MAPS (4)  /* device batch calls only: bsize, nframes and the mode
                                  flags are those of the previous batch call on this
                                  context and stream (an encoder's forward transform
                                  followed by its reconstruction, encode.c:2492 then
                                  :2506-2533): the per-region plan built from the maps
                                  by that call is reused instead of rebuilt.
                                  OD_TDLT_EINVAL if there is no such plan. */
#define OD_TDLT_CODING_ORDER (8) /* the coefficient side of the call is the BLOCK-LINEAR plane
                                  of coding-order vectors (layout: od_tdlt_coding_order_batch
                                  below) instead of the raster plane: the forward stage
                                  stores every block as od_raster_to_coding_order would
                                  (src/partition.c:143-170; call sites src/encode.c:1160-1161),
                                  the inverse stage loads it as od_coding_order_to_raster
                                  does after od_init_skipped_coeffs on a keyframe
                                  (src/state.c:1340-1349, src/encode.c:1219-1220,
                                  src/decode.c:580-581, 630): entries the scans do not
                                  order (half of a 32x32 block) are left alone by the
                                  forward store and read as zero by the inverse load.
                                  Needs od_tdlt_set_scan_tables; not with OD_TDLT_HAAR.
                                  The permutation then costs no pass of its own. */

/* Geometry of a stream; mirrors the fields of daala_info / od_state the path
   reads (reference src/state.c:366-375, 393-394). */
typedef struct od_tdlt_geometry {
  int pic_width;                      /* luma picture width  (info.pic_width)  */
  int pic_height;                     /* luma picture height (info.pic_height) */
  int nplanes;                        /* 1..4 */
  int xdec[OD_TDLT_NPLANES_MAX];      /* 0 or 1; xdec == ydec (encode.c:1296) */
  int ydec[OD_TDLT_NPLANES_MAX];
} od_tdlt_geometry;

typedef struct od_tdlt_ctx od_tdlt_ctx;

/* ---- context -----------------------------------------------------------
   Created next to od_state_init (reference src/state.c:355-496), destroyed
   next to od_state_clear (:507-...).  max_frames bounds the batch size of the
   host-pointer calls (device staging for that many frames is allocated). */
int od_tdlt_ctx_create(od_tdlt_ctx **ctx, const od_tdlt_geometry *geom,
 int device, int max_frames);
void od_tdlt_ctx_destroy(od_tdlt_ctx *ctx);

/* Padded frame geometry (multiple of 64, reference src/state.c:372-375). */
int od_tdlt_frame_width(const od_tdlt_ctx *ctx);
int od_tdlt_frame_height(const od_tdlt_ctx *ctx);
int od_tdlt_nhsb(const od_tdlt_ctx *ctx);
int od_tdlt_nvsb(const od_tdlt_ctx *ctx);
/* Number of od_coeff in one plane of one frame, (fw>>xdec)*(fh>>ydec). */
long long od_tdlt_plane_size(const od_tdlt_ctx *ctx, int pli);
/* Bytes of one frame's block-size map: (nvsb*4)*(nhsb*4), one byte per 8x8
   luma block holding 0..3 (reference src/state.h:207-224, without the border). */
long long od_tdlt_bsize_size(const od_tdlt_ctx *ctx);
const char *od_tdlt_last_error(const od_tdlt_ctx *ctx);

/* ---- 1. device-resident batches ---------------------------------------
   c[pli], d[pli]: DEVICE pointers to nframes contiguous planes
   ([frame][row][col], row stride = frame_width>>xdec, no padding -- the
   reference's ctmp/dtmp layout, src/state.c:409-428, repeated per frame).
   Every plane pointer must be 16-byte aligned (cudaMalloc gives 256), else
   OD_TDLT_EINVAL; c and d must not overlap.
   bsize: DEVICE pointer, nframes maps of od_tdlt_bsize_size() bytes.
   stream: a cudaStream_t (NULL = default stream).  Asynchronous.

   Forward  = od_apply_prefilter_frame_sbs (src/filter.c:1556) + od_compute_dcts
              on every superblock (src/encode.c:1286), c -> d.  c is NOT modified.
   Inverse  = per superblock: [DC split], od_bin_idct* / od_haar_inv at the
              leaves, od_postfilter_split bottom-up (src/decode.c:767-868,
              528-639), then od_apply_postfilter_frame_sbs (src/filter.c:1588),
              d -> c.  d is NOT modified. */
int od_tdlt_fwd_batch(od_tdlt_ctx *ctx, od_coeff *const d[], const od_coeff *const c[],
 const unsigned char *bsize, int nframes, int flags, void *stream);
int od_tdlt_inv_batch(od_tdlt_ctx *ctx, od_coeff *const c[], const od_coeff *const d[],
 const unsigned char *bsize, int nframes, int flags, void *stream);
/* The same two stages with the pixel conversion fused into the kernels' load / store, so
   that the od_coeff copy of the spatial side never exists in memory:
     forward = od_ref_plane_to_coeff (src/state.c:1252-1272, od_ref_buf_to_coeff :1209-1250)
               + the forward stage, pix -> d;
     inverse = the inverse stage + od_coeff_to_ref_plane (src/state.c:1320-1338,
               od_coeff_to_ref_buf :1274-1316), d -> pix.
   pix[pli]: DEVICE pointers to nframes contiguous planes of (fw>>xdec) x (fh>>ydec)
   reference-buffer samples, xstride bytes each (1: 8-bit, 2: 16-bit), no padding, 16-byte
   aligned; xstride / bitdepth_mode / lossless_p as in the conversion calls below.  Not
   available with OD_TDLT_HAAR (OD_TDLT_EINVAL): lossless mode uses the separate conversion
   calls.  Per sample and direction these calls move 4 + xstride bytes instead of 12 + xstride
   (conversion pass + transform pass). */
int od_tdlt_fwd_batch_pix(od_tdlt_ctx *ctx, od_coeff *const d[], const void *const pix[],
 const unsigned char *bsize, int nframes, int flags, int xstride, int bitdepth_mode,
 int lossless_p, void *stream);
int od_tdlt_inv_batch_pix(od_tdlt_ctx *ctx, void *const pix[], const od_coeff *const d[],
 const unsigned char *bsize, int nframes, int flags, int xstride, int bitdepth_mode,
 int lossless_p, void *stream);

/* The forward batch call for callers that must not allocate (code under CUDA graph
   capture, serving loops, harnesses that forbid cudaMalloc).  It has NO counterpart in
   the reference.
   Same arguments and results as od_tdlt_fwd_batch, but no context: the geometry is passed
   directly and the CURRENT CUDA device is used.  The only scratch memory of the forward
   stage (1 KB of plan per 128x64 region of a plane) lives in a file-scope __device__ array
   of OD_TDLT_STATIC_MAX_REGIONS records (4 MiB, held by every GPU that loads the library
   until the process exits); the call makes no cudaMalloc / cudaFree / allocating library
   call.  regions = nframes * sum over planes of ceil(w/128)*ceil(h/64) of the padded
   planes (1080p 4:2:0: 399 per frame, 4K 4:2:0: 1530); more than the maximum returns
   OD_TDLT_EINVAL with a message on stderr.  All calls share the one array: calls that
   overlap on the GPU (two streams, two host threads) race on it and must be serialised by
   the caller.  The first call on a device looks the array's address up
   (cudaGetSymbolAddress) and sets the kernels' attributes: make it outside graph capture.
   Errors: OD_TDLT_EFAULT / OD_TDLT_EINVAL, message on stderr (there is no context to hold
   it).  Asynchronous on `stream`. */
#define OD_TDLT_STATIC_MAX_REGIONS (4096)
int od_tdlt_fwd_batch_static(const od_tdlt_geometry *geom, od_coeff *const d[],
 const od_coeff *const c[], const unsigned char *bsize, int nframes, int flags,
 void *stream);

/* Pixel <-> coefficient conversion on device batches
   (od_ref_buf_to_coeff / od_coeff_to_ref_buf, src/state.c:1209-1316).
   pix: DEVICE pointer to nframes tightly packed planes of (fw>>xdec) x (fh>>ydec)
   samples, xstride = 1 (uint8) or 2 (int16) bytes. */
int od_tdlt_pix_to_coeff_batch(od_tdlt_ctx *ctx, od_coeff *c, const void *pix,
 int pli, int nframes, int xstride, int bitdepth_mode, int lossless_p, void *stream);
int od_tdlt_coeff_to_pix_batch(od_tdlt_ctx *ctx, void *pix, const od_coeff *c,
 int pli, int nframes, int xstride, int bitdepth_mode, int lossless_p, void *stream);

/* Chroma-from-luma predictor planes for 4:2:0 (reference od_resample_luma_coeffs,
   src/intra.c:72-109, as called for every chroma leaf block at src/encode.c:1512-1517 and
   src/decode.c:816-821).  d_luma: the decoded luma coefficient planes (plane 0 layout);
   l: planes of the size of plane 1, every chroma leaf block's predictor at the block's own
   position (the reference's ctx->l, one block at a time, laid out as a plane).  DEVICE
   pointers, asynchronous on `stream`; nframes is not limited by max_frames. */
int od_tdlt_cfl_pred_batch(od_tdlt_ctx *ctx, od_coeff *l, const od_coeff *d_luma,
 const unsigned char *bsize, int nframes, void *stream);

/* Deringing of the reconstructed planes, the step after the inverse stage (reference
   od_dering, src/filter.c:1835-1941, called per 32x32 luma superblock and plane by the loops
   at src/decode.c:1055-1122 and src/encode.c; includes od_dir_find8, od_compute_thresh and
   the od_filter_dering_direction / _orthogonal 4x4 and 8x8 filters of the dispatch table).
   out[p] = in[p] with every superblock whose flag is set filtered; out != in (the reference
   filters from its unfiltered int16_t copy etmp into ctmp; `in` plays etmp, `out` ctmp, and
   samples are taken as int16_t exactly as that copy does).
     flags: nframes x nvsb x nhsb bytes, state->dering_flags after the decision
            (src/decode.c:1093);
     bskip[p]: per plane nframes x (nvsb*8) x (nhsb*8) bytes, one skip flag per 4x4 samples of
            plane p, row stride nhsb*8 for every plane (state->bskip[p], skip_stride,
            src/state.c:450-463; 4:2:0 chroma uses the upper-left quarter);
     q[p]:  the plane quantizers (state->quantizer[p]) of this batch, HOST array.
   4:4:4 or 4:2:0.  DEVICE pointers in out/in/flags/bskip (16-byte aligned planes),
   asynchronous on `stream`, at most 65535 frames per call.  Two kernels: the direction
   search of the luma blocks, then the filters; 64 bytes per superblock of context-owned
   device scratch pass between them (grown on demand), so calls on one context must be
   ordered by their stream like every other call here. */
int od_tdlt_dering_batch(od_tdlt_ctx *ctx, od_coeff *const out[], const od_coeff *const in[],
 const unsigned char *flags, const unsigned char *const bskip[], const int q[], int nframes,
 void *stream);

/* Keyframe bilinear smoothing, the step after deringing (reference od_smooth_recursive with
   min_bs = OD_BLOCK_32X32 -> od_bilinear_smooth, src/filter.c:1952-2039, called per
   superblock and plane at src/decode.c:1134-1146 when quantizer[0] > 0 on keyframes): every
   superblock the map codes as one 32x32 block is smoothed IN PLACE in every plane.
   bsize: the block-size maps (layout above); q[p]: plane quantizers, HOST array.  4:4:4 or
   4:2:0.  DEVICE pointers in c/bsize, asynchronous on `stream`. */
int od_tdlt_smooth_batch(od_tdlt_ctx *ctx, od_coeff *const c[], const unsigned char *bsize,
 const int q[], int nframes, void *stream);

/* Raster <-> coding order of every transform block (reference od_raster_to_coding_order /
   od_coding_order_to_raster, src/partition.c:143-199; call sites src/encode.c:1160-1161,
   1219-1220, src/decode.c:580-581, 630): the permutation the PVQ coder wants next.
   The scan tables are data of the reference and are handed over by the caller once per
   context: tables[k] = OD_ZIGZAG4/8/16/32 ((x, y) byte pairs: 15, 48, 192, 768 of them,
   src/zigzag*.c), lens[k] = how many entries layout k orders
   (OD_LAYOUT*.band_offsets[nb_bands], src/partition.c:43-73: 15, 48, 192, 256). */
int od_tdlt_set_scan_tables(od_tdlt_ctx *ctx, const unsigned char *const tables[4],
 const int lens[4]);
/* d[p]: raster coefficient planes (layout above).  v[p]: planes of the same size holding
   every block's coding-order vector: superblocks (32x32 luma samples, 16x16 with 4:2:0) in
   raster order, inside a superblock the 4x4-sample units in Z order (x bit 0, y bit 0,
   x bit 1, ...), 16 entries per unit -- a quadtree-aligned n x n block at plane position
   (x0, y0) owns the n*n consecutive entries starting at
     ((y0/S)*(w/S) + x0/S)*S*S + 16*zorder((x0%S)/4, (y0%S)/4),   S = 32 >> dec.
   to_coding != 0: v <- d (od_raster_to_coding_order per block); else d <- v.  Entries the
   scans do not cover (a 32x32 block orders 512 of its 1024 coefficients) are left as they
   were in the destination, as the reference leaves them.  Block sizes follow the map as in
   the transform calls (4:2:0 chroma halves them, not below 4x4).  DEVICE pointers,
   asynchronous on `stream`; v != d. */
int od_tdlt_coding_order_batch(od_tdlt_ctx *ctx, od_coeff *const v[], od_coeff *const d[],
 const unsigned char *bsize, int nframes, int to_coding, void *stream);

/* The encoder's input side, the step before od_ref_plane_to_coeff (reference
   od_img_copy_pad, src/encode.c:1728-1750, called from daala_encode_img_in at :2840): per
   plane od_img_plane_copy with its bit-depth conversion and clamps (src/state.c:92-217: 8 ->
   8 copy, deeper -> 8 rounding, 8 -> deeper and deeper -> deeper shifts), the low-pass
   extension of the picture to the padded frame (od_img_plane_copy_pad, src/encode.c:610-700)
   and, when border > 0, od_img_plane_edge_ext (src/state.c:1100-1159) into a border of
   border >> xdec samples all round (OD_BUFFER_PADDING = 64 in the reference's buffers).
     src[p]: picture samples of plane p, ((pic_width + (1 << xdec) - 1) >> xdec) x (... height),
             src_xstride bytes per sample (1 for 8 bits, 2 above), rows src_ystride[p] BYTES
             apart, frames src_frame_stride[p] bytes apart;
     dst[p]: first visible sample of the (fw >> xdec) x (fh >> ydec) plane, same conventions.
   With border = 0, dst_ystride[p] = (fw >> xdec)*dst_xstride and frames back to back, dst is
   exactly what od_tdlt_fwd_batch_pix reads.  DEVICE pointers, asynchronous on `stream`. */
int od_tdlt_img_copy_pad_batch(od_tdlt_ctx *ctx, void *const dst[], const int dst_ystride[],
 const long long dst_frame_stride[], int dst_xstride, int dst_bitdepth, int border,
 const void *const src[], const int src_ystride[], const long long src_frame_stride[],
 int src_xstride, int src_bitdepth, int nframes, void *stream);

/* ---- 2. whole-frame host calls ----------------------------------------
   HOST pointers, synchronous, nframes <= max_frames.  Planes and maps are laid
   out as above but in host memory.  Pageable memory works, but every copy is then staged
   by the driver and nothing overlaps (measured: bench.py's e2e.pageable); the reference's
   planes are plain malloc (src/state.c:409-428), so page-lock them once after
   od_state_init with
     od_tdlt_host_register(state->ctmp[pli], plane_bytes);   ... dtmp, the od_img planes
   and release them before od_state_clear frees them (od_tdlt_host_unregister).  Both wrap
   cudaHostRegister / cudaHostUnregister (portable across devices); registering a range
   twice is not an error.  Returns OD_TDLT_SUCCESS, OD_TDLT_EINVAL (bad range) or
   OD_TDLT_EFAULT.
   These replace, in od_encode_coefficients / od_decode_coefficients, the loops
   at src/encode.c:2425-2437 + :2492 and src/decode.c:632-638, 854-866, 1032-1041. */
int od_tdlt_host_register(void *ptr, size_t bytes);
int od_tdlt_host_unregister(void *ptr);
int od_tdlt_fwd_frames(od_tdlt_ctx *ctx, od_coeff *const d[], const od_coeff *const c[],
 const unsigned char *bsize, int nframes, int flags);
int od_tdlt_inv_frames(od_tdlt_ctx *ctx, od_coeff *const c[], const od_coeff *const d[],
 const unsigned char *bsize, int nframes, int flags);
/* Host-pointer form of od_tdlt_cfl_pred_batch (synchronous, nframes <= max_frames). */
int od_tdlt_cfl_pred_frames(od_tdlt_ctx *ctx, od_coeff *l, const od_coeff *d_luma,
 const unsigned char *bsize, int nframes);
/* Host-pointer form of od_tdlt_dering_batch (synchronous, nframes <= max_frames). */
int od_tdlt_dering_frames(od_tdlt_ctx *ctx, od_coeff *const out[], const od_coeff *const in[],
 const unsigned char *flags, const unsigned char *const bskip[], const int q[], int nframes);
/* Host-pointer form of od_tdlt_smooth_batch (synchronous, in place, nframes <= max_frames). */
int od_tdlt_smooth_frames(od_tdlt_ctx *ctx, od_coeff *const c[], const unsigned char *bsize,
 const int q[], int nframes);
/* Host-pointer form of od_tdlt_coding_order_batch (synchronous, nframes <= max_frames; the
   destination's previous contents are kept where the scans do not write). */
int od_tdlt_coding_order_frames(od_tdlt_ctx *ctx, od_coeff *const v[], od_coeff *const d[],
 const unsigned char *bsize, int nframes, int to_coding);
/* Same with the pixel conversion done on the device: the host ships 8/16-bit
   reference-buffer samples instead of od_coeff (od_ref_plane_to_coeff,
   src/state.c:1252; od_coeff_to_ref_plane, :1320).  pix[pli] points at the
   first visible sample of frame 0 (od_img_plane.data); rows are ystride[pli]
   BYTES apart (od_img_plane.ystride, which includes the reference's border
   padding) and frames frame_stride[pli] bytes apart; xstride = 1 or 2 bytes
   per sample (od_img_plane.xstride); lossless_p as OD_LOSSLESS(). */
int od_tdlt_fwd_frames_pix(od_tdlt_ctx *ctx, od_coeff *const d[], const void *const pix[],
 const int ystride[], const long long frame_stride[], const unsigned char *bsize,
 int nframes, int flags, int xstride, int bitdepth_mode, int lossless_p);
int od_tdlt_inv_frames_pix(od_tdlt_ctx *ctx, void *const pix[], const int ystride[],
 const long long frame_stride[], const od_coeff *const d[], const unsigned char *bsize,
 int nframes, int flags, int xstride, int bitdepth_mode, int lossless_p);

/* ---- 3. reference-signature calls (host pointers, synchronous) --------- */
/* od_dct_func_2d, reference src/dct.h:61-62; tables src/dct.c:42-56 */
void od_bin_fdct4x4_cuda(od_coeff *y, int ystride, const od_coeff *x, int xstride);
void od_bin_idct4x4_cuda(od_coeff *x, int xstride, const od_coeff *y, int ystride);
void od_bin_fdct8x8_cuda(od_coeff *y, int ystride, const od_coeff *x, int xstride);
void od_bin_idct8x8_cuda(od_coeff *x, int xstride, const od_coeff *y, int ystride);
void od_bin_fdct16x16_cuda(od_coeff *y, int ystride, const od_coeff *x, int xstride);
void od_bin_idct16x16_cuda(od_coeff *x, int xstride, const od_coeff *y, int ystride);
void od_bin_fdct32x32_cuda(od_coeff *y, int ystride, const od_coeff *x, int xstride);
void od_bin_idct32x32_cuda(od_coeff *x, int xstride, const od_coeff *y, int ystride);
/* Installs the eight calls above into the reference's dispatch tables
   (od_state.opt_vtbl.fdct_2d[] / idct_2d[], reference src/state.h:123-124), the way
   od_state_opt_vtbl_init_x86 overrides them per CPU feature
   (src/x86/x86state.c:39-97); entry OD_NBSIZES (the 64-point placeholder) is left
   alone.  Call as
     od_tdlt_opt_vtbl_init_cuda(state->opt_vtbl.fdct_2d, state->opt_vtbl.idct_2d);  */
typedef void (*od_tdlt_dct_func_2d)(od_coeff *out, int out_stride, const od_coeff *in,
 int in_stride);
void od_tdlt_opt_vtbl_init_cuda(od_tdlt_dct_func_2d fdct_2d[], od_tdlt_dct_func_2d idct_2d[]);
/* od_init_skipped_coeffs, reference src/state.c:1340-1357 (call sites src/encode.c:1219,
   src/decode.c:580): the block at d + bo (n x n, row stride w) is prepared for
   od_coding_order_to_raster -- on a keyframe everything but the DC is cleared, else the block
   takes the prediction pred (n x n, dense). */
void od_init_skipped_coeffs_cuda(od_coeff *d, od_coeff *pred, int is_keyframe, int bo, int n, int w);

/* The deringing filters of the same table (od_filter_dering_direction_func /
   od_filter_dering_orthogonal_func, src/filter.h:46-65; entries [0] = 4x4, [1] = 8x8,
   src/state.h:120-121; C versions src/filter.c:1715-1800, SSE2 twins in
   src/x86/sse2filter.c).  Host pointers, synchronous, one launch per block: for checking
   through od_dering's own call sites (src/filter.c:1922-1940), not for speed.  `in` points
   at the block inside a buffer of row pitch OD_FILT_BSTRIDE (38) with three samples around it.
     od_tdlt_dering_vtbl_init_cuda(state->opt_vtbl.filter_dering_direction,
                                   state->opt_vtbl.filter_dering_orthogonal);  */
typedef void (*od_tdlt_dering_direction_func)(int16_t *y, int ystride, int16_t *in,
 int threshold, int dir);
typedef void (*od_tdlt_dering_orthogonal_func)(int16_t *y, int ystride, int16_t *in,
 int16_t *x, int xstride, int threshold, int dir);
void od_filter_dering_direction_4x4_cuda(int16_t *y, int ystride, int16_t *in, int threshold,
 int dir);
void od_filter_dering_direction_8x8_cuda(int16_t *y, int ystride, int16_t *in, int threshold,
 int dir);
void od_filter_dering_orthogonal_4x4_cuda(int16_t *y, int ystride, int16_t *in, int16_t *x,
 int xstride, int threshold, int dir);
void od_filter_dering_orthogonal_8x8_cuda(int16_t *y, int ystride, int16_t *in, int16_t *x,
 int xstride, int threshold, int dir);
void od_tdlt_dering_vtbl_init_cuda(od_tdlt_dering_direction_func direction[],
 od_tdlt_dering_orthogonal_func orthogonal[]);
/* Many blocks in one launch: nblocks blocks of (4<<bs)^2, block i at x + i*xblock
   (resp. y + i*yblock); the useful form of the above for parity sweeps. */
int od_tdlt_dct2d_blocks(int bs, int inverse, od_coeff *out, int ostride, long long oblock,
 const od_coeff *in, int istride, long long iblock, int nblocks);
/* The 1-D transforms od_bin_fdct{4,8,16,32} / od_bin_idct* behind the reference's tables
   OD_FDCT_1D / OD_IDCT_1D (src/dct.c:58-72; src/dct.c:74-135, 151-333, 349-772, 1757-1958):
   nvec vectors of 4<<bs values, tightly packed, natural order in and out -- what the
   reference's check_bias streams through (src/dct.c:3565-3598). */
int od_tdlt_dct1d_vectors(int bs, int inverse, od_coeff *out, const od_coeff *in, int nvec);
/* od_pre_filter{4,8,16,32} / od_post_filter*, reference src/filter.c:174-1373;
   nvec vectors of 4<<f samples, tightly packed. */
int od_tdlt_lap_filter_vectors(int f, int inverse, od_coeff *out, const od_coeff *in, int nvec);
/* reference src/dct.h:181-183, src/dct.c:1960-2026 */
void od_haar_cuda(od_coeff *y, int ystride, const od_coeff *x, int xstride, int ln);
void od_haar_inv_cuda(od_coeff *x, int xstride, const od_coeff *y, int ystride, int ln);
/* reference src/tf.h:53-56, src/tf.c:112-169 */
void od_tf_up_hv_cuda(od_coeff *dst, int dstride, const od_coeff *src, int sstride, int n);
void od_tf_down_hv_cuda(od_coeff *dst, int dstride, const od_coeff *src, int sstride, int n);
/* reference src/tf.h:46-53, src/tf.c:38-108 (low-pass-only merges; the live caller is
   od_resample_luma_coeffs, src/intra.c:81-94) */
void od_tf_up_h_lp_cuda(od_coeff *dst, int dstride, const od_coeff *src, int sstride, int dx,
 int n);
void od_tf_up_v_lp_cuda(od_coeff *dst, int dstride, const od_coeff *src, int sstride, int dy,
 int n);
void od_tf_up_hv_lp_cuda(od_coeff *dst, int dstride, const od_coeff *src, int sstride, int dx,
 int dy, int n);
/* reference src/intra.h / src/intra.c:72-109, one block (host pointers, synchronous; the
   4:2:0 and 4:4:4 branches: xdec == ydec) */
void od_resample_luma_coeffs_cuda(od_coeff *chroma_pred, int cpstride,
 const od_coeff *decoded_luma, int dlstride, int xdec, int ydec, int bs, int chroma_bs);
/* reference src/tf.h:60-66, src/tf.c:171-277 */
void od_tf_filter_2d_cuda(od_coeff *dst, int dstride, const od_coeff *src, int sstride, int n);
void od_tf_filter_inv_2d_cuda(od_coeff *dst, int dstride, const od_coeff *src, int sstride,
 int n);
void od_convert_block_down_cuda(od_coeff *dst, int dstride, const od_coeff *src, int sstride,
 int curr_size, int dest_size, int filter);
/* reference src/filter.h:121-128, src/filter.c:1486-1646 (OD_DEBLOCKING == 0:
   q, skip and skip_stride are ignored, as in the reference) */
void od_prefilter_split_cuda(od_coeff *c0, int stride, int bs, int f, int hfilter, int vfilter);
void od_postfilter_split_cuda(od_coeff *c0, int stride, int bs, int f, int q,
 unsigned char *skip, int skip_stride, int hfilter, int vfilter);
void od_apply_prefilter_frame_sbs_cuda(od_coeff *c, int stride, int nhsb, int nvsb,
 int xdec, int ydec);
void od_apply_postfilter_frame_sbs_cuda(od_coeff *c, int stride, int nhsb, int nvsb,
 int xdec, int ydec, int q, unsigned char *skip, int skip_stride);

/* Introspection (host only, no device needed): the order in which the inverse
   kernel walks the 128x64 regions of one frame when `rows_in_flight` rows of
   regions are in flight per frame -- region indices plane by plane, row-major
   within a plane.  A region's strip exchange (tdlt_inv_region.cu) only needs its
   left / upper / upper-left neighbours to come earlier in this order; the tests
   check exactly that.  Returns the number of regions of a frame (<= capacity), or
   a negative OD_TDLT_* code. */
int od_tdlt_region_order(const od_tdlt_geometry *g, int rows_in_flight, unsigned short *order,
                         int capacity);

/* Number of kernel launches issued by this library since load (bench.py's
   gpu_launches claim is read from here). */
long long od_tdlt_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif

/* TEST INFRASTRUCTURE.  Force-included (gcc -include) in front of the reference's encode.c
   and decode.c when oracle/Makefile builds _ref/libdaala_ref_cuda.so: the direct calls the
   reference's encoder and decoder make into the path (src/encode.c:1129-1132, 1225, 1305,
   1320, 1591-1593, 1620, 2428-2435, 2530; src/decode.c:562, 633, 848, 864, 1005, 1037) then
   bind to the product's reference-signature entry points (include/daala_tdlt.h section 3),
   exactly the substitution INTEGRATION.md section 2 describes for a maintainer.  The header
   declarations (src/filter.h:121-128, src/dct.h:181-183) are renamed with them, so the
   prototypes are checked by the compiler against the reference's own.  The symbols are
   resolved when the library is loaded, from libdaala_tdlt_cuda.so (loaded RTLD_GLOBAL first). */
#define od_prefilter_split od_prefilter_split_cuda
#define od_postfilter_split od_postfilter_split_cuda
#define od_apply_prefilter_frame_sbs od_apply_prefilter_frame_sbs_cuda
#define od_apply_postfilter_frame_sbs od_apply_postfilter_frame_sbs_cuda
#define od_haar od_haar_cuda
#define od_haar_inv od_haar_inv_cuda
#define REF_GLUE_CUDA_REDIRECT 1

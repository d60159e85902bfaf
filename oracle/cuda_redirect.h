/* TEST INFRASTRUCTURE.  Force-included (gcc -include) in front of the reference's encode.c
   and decode.c when oracle/Makefile builds _ref/libdaala_ref_cuda.so: the direct calls the
   reference's encoder and decoder make into the path (src/encode.c:1129-1132, 1225, 1305,
   1320, 1591-1593, 1620, 2428-2435, 2530; src/decode.c:562, 633, 848, 864, 1005, 1037) then
   bind to the forwarding functions of oracle/cuda_check.c, which call the product's
   reference-signature entry points (include/daala_tdlt.h section 3: od_prefilter_split_cuda
   ...) -- the substitution INTEGRATION.md section 2 describes for a maintainer -- and, in
   check mode, the reference's own function beside them (OD_CHECKASM style).  The header
   declarations (src/filter.h:121-128, src/dct.h:181-183) are renamed with them, so the
   prototypes are checked by the compiler against the reference's own.  The *_cuda symbols are
   resolved when the library is loaded, from libdaala_tdlt_cuda.so (loaded RTLD_GLOBAL first). */
#define od_prefilter_split ref_chk_prefilter_split
#define od_postfilter_split ref_chk_postfilter_split
#define od_apply_prefilter_frame_sbs ref_chk_apply_prefilter_frame_sbs
#define od_apply_postfilter_frame_sbs ref_chk_apply_postfilter_frame_sbs
#define od_haar ref_chk_haar
#define od_haar_inv ref_chk_haar_inv
#define REF_GLUE_CUDA_REDIRECT 1

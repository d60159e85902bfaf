import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if len(sys.argv) > 1 and sys.argv[1] == "--one":
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import ctypes as C, numpy as np
    import oracle_libs as ol, make_golden as mg, daala_b200
    prelude = sys.argv[2]
    L = daala_b200.lib()
    if "refload" in prelude:
        ol.ref()
    if "refcall" in prelude:
        R = ol.ref(); R.od_prefilter_split.restype = None
        a = np.zeros((16, 16), np.int32); R.od_prefilter_split(ol.ptr(a), 16, 2, 0, 1, 1)
    if "split" in prelude:
        a = np.arange(256, dtype=np.int32).reshape(16, 16).copy()
        L.od_prefilter_split_cuda(a.ctypes.data, 16, 2, 0, 1, 1)
    if "sbs" in prelude:
        a = np.arange(160*96, dtype=np.int32).reshape(96, 160).copy()
        L.od_apply_prefilter_frame_sbs_cuda(a.ctypes.data, 160, 5, 3, 0, 0)
    if "torch" in prelude:
        import torch; torch.zeros(10, device="cuda")
    if "ctx" in prelude:
        c = daala_b200.TdltContext(352, 288, device=0, max_frames=2); c.close()
    product = C.CDLL(daala_b200.LIB_PATH, mode=C.RTLD_GLOBAL)
    refc = C.CDLL(ol.REF_CUDA_SO)
    refc.ref_set_external_vtbl.argtypes = [C.c_void_p, C.c_void_p]
    impl = 2
    if "novtbl" in prelude:
        impl = 1
    refc.ref_set_external_vtbl(C.cast(product.od_tdlt_opt_vtbl_init_cuda, C.c_void_p),
                               None if "nodering" in prelude else C.cast(product.od_tdlt_dering_vtbl_init_cuda, C.c_void_p))
    frames = [mg.cif_frame(t) for t in range(4)]
    if "check" in prelude:
        impl = 3
        refc.ref_chk_reset(1)
    pk, sz, out = ol.codec_roundtrip(refc, frames, 30, impl, keyframe_rate=4)
    refc.ref_chk_get_calls.restype = C.c_longlong
    refc.ref_chk_get_mismatches.restype = C.c_longlong
    refc.ref_chk_get_first.restype = C.c_char_p
    print(prelude, sz.tolist(), "checked calls", refc.ref_chk_get_calls(), "mismatches", refc.ref_chk_get_mismatches(),
          refc.ref_chk_get_first().decode())
    sys.exit(0)
for prelude in ["none", "refload", "none,check", "refload,check", "split,check"]:
    r = subprocess.run([sys.executable, __file__, "--one", prelude], capture_output=True, text=True, timeout=300)
    print((r.stdout.strip() or r.stderr.strip()[-300:]), flush=True)
# the unmodified reference alone under the same preludes: is ITS inter bitstream reproducible?
if len(sys.argv) == 1:
    code = r'''
import sys, os
ROOT = %r
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import numpy as np, oracle_libs as ol, make_golden as mg
pre = sys.argv[1]
if "cuda" in pre:
    import daala_b200
    L = daala_b200.lib(); a = np.arange(256, dtype=np.int32).reshape(16, 16).copy(); L.od_prefilter_split_cuda(a.ctypes.data, 16, 2, 0, 1, 1)
if "alloc" in pre:
    junk = [np.full(n, 7, np.uint8) for n in (1000, 50000, 3000000, 123456)]; del junk
frames = [mg.cif_frame(t) for t in range(4)]
pk, sz, out = ol.codec_roundtrip(ol.ref(), frames, 30, 1, keyframe_rate=4)
print("reference alone,", pre, sz.tolist())
''' % ROOT
    for pre in ["none", "cuda", "alloc", "cuda,alloc"]:
        r = subprocess.run([sys.executable, "-c", code, pre], capture_output=True, text=True, timeout=300)
        print((r.stdout.strip() or r.stderr.strip()[-300:]), flush=True)

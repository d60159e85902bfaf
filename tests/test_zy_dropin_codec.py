"""The drop-in, dropped in: the reference's own encoder and decoder run with every call they
make into the path bound to the CUDA library, and must produce the same bitstream and the
same pictures as the unmodified reference.

  * oracle/_ref/libdaala_ref_cuda.so is the reference compiled with oracle/cuda_redirect.h
    force-included in front of src/encode.c and src/decode.c: od_prefilter_split,
    od_postfilter_split, od_apply_prefilter_frame_sbs, od_apply_postfilter_frame_sbs, od_haar and
    od_haar_inv become the product's *_cuda entry points;
  * impl = 2 installs od_tdlt_opt_vtbl_init_cuda / od_tdlt_dering_vtbl_init_cuda into the
    opt_vtbl of the encoder's AND the decoder's od_state (src/state.h:116-124), the way
    od_state_opt_vtbl_init_x86 does (src/x86/x86state.c:39-97);
  * ref_codec_roundtrip (oracle/ref_glue.c) is daala_encode_img_in -> daala_encode_packet_out
    -> daala_decode_packet_in -> daala_decode_img_out (src/encode.c:2876, src/decode.c:1159).

Equality of packets and decoded pictures with the plain reference is OD_ENCODER_CHECK's
criterion (src/encode.c:3203-3243) extended to the bitstream; at quality 0 the output must be
the input (tools/check_lossless.sh).  One kernel launch per block: slow by design, this is a
correctness test.  Collected after the parity tests and before test_zz_no_alloc.py.
"""
import ctypes as C
import hashlib
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(HERE, "golden"))
import oracle_libs as ol  # noqa: E402

pytestmark = pytest.mark.gpu


def _cuda_routed_reference():
    import daala_b200
    if not (ol.have_ref() and os.path.exists(ol.REF_CUDA_SO)):
        pytest.skip("oracle/_ref/libdaala_ref_cuda.so is not built (needs /root/reference at build time)")
    daala_b200.lib()
    product = C.CDLL(daala_b200.LIB_PATH, mode=C.RTLD_GLOBAL)     # the *_cuda symbols, globally visible
    refc = C.CDLL(ol.REF_CUDA_SO)
    assert refc.ref_direct_calls_redirected() == 1 and ol.ref().ref_direct_calls_redirected() == 0
    refc.ref_set_external_vtbl.argtypes = [C.c_void_p, C.c_void_p]
    refc.ref_set_external_vtbl(C.cast(product.od_tdlt_opt_vtbl_init_cuda, C.c_void_p),
                               C.cast(product.od_tdlt_dering_vtbl_init_cuda, C.c_void_p))
    return refc


@pytest.mark.parametrize("quality,keyframe_rate,nframes", [(20, 1, 10), (0, 1, 10), (30, 4, 4)])
def test_reference_codec_through_the_cuda_library(quality, keyframe_rate, nframes):
    """BASELINE config 1's ten CIF frames (intra-only) at quality 20 and lossless, and a short
    inter-coded run (the split filters and DCTs of the motion-compensated planes as well)."""
    import daala_b200
    import make_golden as mg
    refc = _cuda_routed_reference()
    frames = [mg.cif_frame(t) for t in range(nframes)]
    n0 = daala_b200.launch_count()
    pk_gpu, sz_gpu, out_gpu = ol.codec_roundtrip(refc, frames, quality, 2, keyframe_rate=keyframe_rate)
    launches = daala_b200.launch_count() - n0
    assert launches > 1000 * nframes, "the codec did not go through the CUDA library (%d launches)" % launches
    pk_ref, sz_ref, out_ref = ol.codec_roundtrip(ol.ref(), frames, quality, 1, keyframe_rate=keyframe_rate)
    assert np.array_equal(sz_gpu, sz_ref), "packet sizes differ: %s vs %s" % (sz_gpu, sz_ref)
    assert np.array_equal(pk_gpu, pk_ref), "bitstreams differ"
    for f, (a, b) in enumerate(zip(out_gpu, out_ref)):
        for p in range(3):
            assert np.array_equal(a[p], b[p]), "decoded frame %d plane %d differs" % (f, p)
    if quality == 0:
        md5 = lambda fr: hashlib.md5(b"".join(p.tobytes() for f in fr for p in f)).hexdigest()  # noqa: E731
        assert md5(out_gpu) == md5(frames), "lossless mode must return the input (check_lossless.sh)"

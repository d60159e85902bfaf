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


def _counters(lib):
    for f in (lib.ref_chk_get_calls, lib.ref_chk_get_mismatches, lib.ref_encoder_check_mismatched_rows):
        f.restype = C.c_longlong
    lib.ref_chk_get_first.restype = C.c_char_p
    return lib.ref_chk_get_calls(), lib.ref_chk_get_mismatches(), lib.ref_chk_get_first().decode()


@pytest.mark.parametrize("quality", [20, 0])
def test_intra_codec_through_the_cuda_library_same_bitstream(quality):
    """BASELINE config 1's ten CIF frames, intra-only, at quality 20 and lossless: the
    bitstream and the decoded pictures of the CUDA-routed codec equal the plain reference's,
    and the decoder stays in step with the encoder (OD_ENCODER_CHECK)."""
    import daala_b200
    import make_golden as mg
    refc = _cuda_routed_reference()
    frames = [mg.cif_frame(t) for t in range(10)]
    refc.ref_chk_reset(0)
    n0 = daala_b200.launch_count()
    pk_gpu, sz_gpu, out_gpu = ol.codec_roundtrip(refc, frames, quality, 2)
    launches = daala_b200.launch_count() - n0
    assert launches > 10000, "the codec did not go through the CUDA library (%d launches)" % launches
    refc.ref_encoder_check_mismatched_rows.restype = C.c_longlong
    assert refc.ref_encoder_check_mismatched_rows() == 0
    pk_ref, sz_ref, out_ref = ol.codec_roundtrip(ol.ref(), frames, quality, 1)
    assert np.array_equal(sz_gpu, sz_ref), "packet sizes differ: %s vs %s" % (sz_gpu, sz_ref)
    assert np.array_equal(pk_gpu, pk_ref), "bitstreams differ"
    for f, (a, b) in enumerate(zip(out_gpu, out_ref)):
        for p in range(3):
            assert np.array_equal(a[p], b[p]), "decoded frame %d plane %d differs" % (f, p)
    if quality == 0:
        md5 = lambda fr: hashlib.md5(b"".join(p.tobytes() for f in fr for p in f)).hexdigest()  # noqa: E731
        assert md5(out_gpu) == md5(frames), "lossless mode must return the input (check_lossless.sh)"


@pytest.mark.parametrize("quality,keyframe_rate,nframes", [(20, 1, 3), (0, 1, 2), (30, 4, 4)])
def test_codec_checkasm_style(quality, keyframe_rate, nframes):
    """OD_CHECKASM's criterion (src/x86/x86dct.h:305-315) for the CUDA library, over every call
    the encoder and the decoder make while coding: each DCT, deringing filter, split filter,
    superblock-edge filter and Haar call runs the reference's C function and the CUDA entry
    point on the same input (oracle/cuda_check.c) -- no call may differ.  Includes an
    inter-coded run (motion-compensated planes through od_prefilter_split and the DCTs).
    The inter-coded bitstream itself is NOT compared with a second run of the reference: the
    reference's inter encoder is not reproducible from one process history to another -- the
    UNMODIFIED library alone produces packet sizes [2331, 96, 94, 123] or [2331, 96, 97, 121]
    for these four frames depending on what the process allocated before
    (tools/dbg/dropin_probe.py, profiles/dropin_probe_r2.txt), while every transform call
    still matches -- so equality of every call is the criterion there."""
    import make_golden as mg
    refc = _cuda_routed_reference()
    frames = [mg.cif_frame(t) for t in range(nframes)]
    refc.ref_chk_reset(1)
    try:
        pk, sz, out = ol.codec_roundtrip(refc, frames, quality, 3, keyframe_rate=keyframe_rate)
        calls, bad, first = _counters(refc)
        rows = refc.ref_encoder_check_mismatched_rows()
    finally:
        refc.ref_chk_reset(0)
    assert calls > 1000 * nframes, "only %d calls were checked" % calls
    assert bad == 0, "%d of %d calls differ from the reference's C functions; first: %s" % (bad, calls, first)
    assert rows == 0, "decoder out of step with the encoder in %d rows (OD_ENCODER_CHECK)" % rows
    if quality == 0:
        for a, b in zip(out, frames):
            assert all(np.array_equal(x, y) for x, y in zip(a, b))

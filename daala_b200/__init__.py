"""daala_b200 -- B200 (sm_100a) implementation of Daala's time-domain lapped
transform stage behind a C-ABI shared library.

This module is the thin Python binding used by the tests and bench.py: it loads
``libdaala_tdlt_cuda.so`` (built in-tree by ``daala_b200/build.py``) with
ctypes and mirrors include/daala_tdlt.h one to one.  There is NO fallback: if
the library is missing or a CUDA call fails, an exception is raised.  Nothing
here imports or calls anything under ``oracle/``.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("DAALA_TDLT_LIB", os.path.join(_HERE, "libdaala_tdlt_cuda.so"))

OD_TDLT_KEYFRAME = 1
OD_TDLT_HAAR = 2
OD_TDLT_SAME_MAPS = 4
OD_TDLT_CODING_ORDER = 8
NPLANES_MAX = 4


class TdltError(RuntimeError):
    pass


class Geometry(C.Structure):
    _fields_ = [("pic_width", C.c_int), ("pic_height", C.c_int), ("nplanes", C.c_int),
                ("xdec", C.c_int * NPLANES_MAX), ("ydec", C.c_int * NPLANES_MAX)]


_lib = None


def lib():
    """The loaded C-ABI library (raises if it has not been built)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise TdltError("%s is missing: run `python daala_b200/build.py` "
                            "(there is no CPU fallback)" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        vp, ci, ll = C.c_void_p, C.c_int, C.c_longlong
        L.od_tdlt_ctx_create.argtypes = [C.POINTER(vp), C.POINTER(Geometry), ci, ci]
        L.od_tdlt_ctx_destroy.argtypes = [vp]
        L.od_tdlt_ctx_destroy.restype = None
        for f in ("od_tdlt_frame_width", "od_tdlt_frame_height", "od_tdlt_nhsb", "od_tdlt_nvsb"):
            getattr(L, f).argtypes = [vp]
        L.od_tdlt_plane_size.argtypes = [vp, ci]
        L.od_tdlt_plane_size.restype = ll
        L.od_tdlt_bsize_size.argtypes = [vp]
        L.od_tdlt_bsize_size.restype = ll
        L.od_tdlt_last_error.argtypes = [vp]
        L.od_tdlt_last_error.restype = C.c_char_p
        L.od_tdlt_launch_count.restype = ll
        L.od_tdlt_region_order.argtypes = [C.POINTER(Geometry), ci, vp, ci]
        arr = C.POINTER(vp)
        L.od_tdlt_fwd_batch.argtypes = [vp, arr, arr, vp, ci, ci, vp]
        L.od_tdlt_inv_batch.argtypes = [vp, arr, arr, vp, ci, ci, vp]
        L.od_tdlt_fwd_batch_static.argtypes = [C.POINTER(Geometry), arr, arr, vp, ci, ci, vp]
        L.od_tdlt_fwd_batch_pix.argtypes = [vp, arr, arr, vp, ci, ci, ci, ci, ci, vp]
        L.od_tdlt_inv_batch_pix.argtypes = [vp, arr, arr, vp, ci, ci, ci, ci, ci, vp]
        L.od_tdlt_pix_to_coeff_batch.argtypes = [vp, vp, vp, ci, ci, ci, ci, ci, vp]
        L.od_tdlt_img_copy_pad_batch.argtypes = [vp, arr, C.POINTER(ci), C.POINTER(ll), ci, ci, ci,
                                                 arr, C.POINTER(ci), C.POINTER(ll), ci, ci, ci, vp]
        L.od_tdlt_coeff_to_pix_batch.argtypes = [vp, vp, vp, ci, ci, ci, ci, ci, vp]
        L.od_tdlt_cfl_pred_batch.argtypes = [vp, vp, vp, vp, ci, vp]
        L.od_tdlt_cfl_pred_frames.argtypes = [vp, vp, vp, vp, ci]
        L.od_resample_luma_coeffs_cuda.argtypes = [vp, ci, vp, ci, ci, ci, ci, ci]
        L.od_resample_luma_coeffs_cuda.restype = None
        L.od_tdlt_dering_batch.argtypes = [vp, arr, arr, vp, arr, C.POINTER(ci), ci, vp]
        L.od_tdlt_dering_frames.argtypes = [vp, arr, arr, vp, arr, C.POINTER(ci), ci]
        L.od_tdlt_smooth_batch.argtypes = [vp, arr, vp, C.POINTER(ci), ci, vp]
        L.od_tdlt_smooth_frames.argtypes = [vp, arr, vp, C.POINTER(ci), ci]
        L.od_tdlt_set_scan_tables.argtypes = [vp, arr, C.POINTER(ci)]
        L.od_tdlt_coding_order_batch.argtypes = [vp, arr, arr, vp, ci, ci, vp]
        L.od_tdlt_coding_order_frames.argtypes = [vp, arr, arr, vp, ci, ci]
        L.od_tdlt_host_register.argtypes = [vp, C.c_size_t]
        L.od_tdlt_host_unregister.argtypes = [vp]
        L.od_tdlt_fwd_frames.argtypes = [vp, arr, arr, vp, ci, ci]
        L.od_tdlt_inv_frames.argtypes = [vp, arr, arr, vp, ci, ci]
        iarr, larr = C.POINTER(ci), C.POINTER(ll)
        L.od_tdlt_fwd_frames_pix.argtypes = [vp, arr, arr, iarr, larr, vp, ci, ci, ci, ci, ci]
        L.od_tdlt_inv_frames_pix.argtypes = [vp, arr, iarr, larr, arr, vp, ci, ci, ci, ci, ci]
        for n in (4, 8, 16, 32):
            for d in ("f", "i"):
                fn = getattr(L, "od_bin_%sdct%dx%d_cuda" % (d, n, n))
                fn.argtypes = [vp, ci, vp, ci]
                fn.restype = None
        L.od_tdlt_dct2d_blocks.argtypes = [ci, ci, vp, ci, ll, vp, ci, ll, ci]
        L.od_tdlt_lap_filter_vectors.argtypes = [ci, ci, vp, vp, ci]
        L.od_tdlt_dct1d_vectors.argtypes = [ci, ci, vp, vp, ci]
        for f in ("od_haar_cuda", "od_haar_inv_cuda"):
            getattr(L, f).argtypes = [vp, ci, vp, ci, ci]
            getattr(L, f).restype = None
        for f in ("od_tf_up_hv_cuda", "od_tf_down_hv_cuda"):
            getattr(L, f).argtypes = [vp, ci, vp, ci, ci]
            getattr(L, f).restype = None
        for f, na in (("od_tf_up_h_lp_cuda", 2), ("od_tf_up_v_lp_cuda", 2), ("od_tf_up_hv_lp_cuda", 3),
                      ("od_tf_filter_2d_cuda", 1), ("od_tf_filter_inv_2d_cuda", 1),
                      ("od_convert_block_down_cuda", 3)):
            getattr(L, f).argtypes = [vp, ci, vp, ci] + [ci] * na
            getattr(L, f).restype = None
        L.od_tdlt_opt_vtbl_init_cuda.argtypes = [vp, vp]
        L.od_init_skipped_coeffs_cuda.argtypes = [vp, vp, ci, ci, ci, ci]
        L.od_init_skipped_coeffs_cuda.restype = None
        L.od_tdlt_opt_vtbl_init_cuda.restype = None
        L.od_prefilter_split_cuda.argtypes = [vp, ci, ci, ci, ci, ci]
        L.od_prefilter_split_cuda.restype = None
        L.od_postfilter_split_cuda.argtypes = [vp, ci, ci, ci, ci, vp, ci, ci, ci]
        L.od_postfilter_split_cuda.restype = None
        L.od_apply_prefilter_frame_sbs_cuda.argtypes = [vp, ci, ci, ci, ci, ci]
        L.od_apply_prefilter_frame_sbs_cuda.restype = None
        L.od_apply_postfilter_frame_sbs_cuda.argtypes = [vp, ci, ci, ci, ci, ci, ci, vp, ci]
        L.od_apply_postfilter_frame_sbs_cuda.restype = None
        _lib = L
    return _lib


def _ptr_array(ptrs):
    a = (C.c_void_p * NPLANES_MAX)()
    for i, p in enumerate(ptrs):
        a[i] = p
    return a


class TdltContext:
    """One od_tdlt_ctx (one per od_state / per GPU).  Mirrors the lifetime the
    reference gives od_state (src/state.c:355-496, 507-...)."""

    def __init__(self, pic_width, pic_height, decs=(0, 1, 1), device=0, max_frames=0):
        L = lib()
        g = Geometry()
        g.pic_width, g.pic_height, g.nplanes = pic_width, pic_height, len(decs)
        for i, d in enumerate(decs):
            g.xdec[i] = d
            g.ydec[i] = d
        self._h = C.c_void_p()
        rc = L.od_tdlt_ctx_create(C.byref(self._h), C.byref(g), device, max_frames)
        if rc != 0:
            raise TdltError("od_tdlt_ctx_create failed: %d" % rc)
        self.L = L
        self.decs = tuple(decs)
        self.nplanes = len(decs)
        self.device = device
        self.max_frames = max_frames
        self.pic_width, self.pic_height = pic_width, pic_height
        self.frame_width = L.od_tdlt_frame_width(self._h)
        self.frame_height = L.od_tdlt_frame_height(self._h)
        self.nhsb = L.od_tdlt_nhsb(self._h)
        self.nvsb = L.od_tdlt_nvsb(self._h)
        self.plane_shapes = [(self.frame_height >> d, self.frame_width >> d) for d in decs]
        self.bsize_shape = (self.nvsb * 4, self.nhsb * 4)

    def close(self):
        if self._h:
            self.L.od_tdlt_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, rc, what):
        if rc != 0:
            raise TdltError("%s failed (%d): %s" % (what, rc, self.L.od_tdlt_last_error(self._h).decode()))

    # ---- layer 1: device pointers (ints), asynchronous on `stream` ----------
    def fwd_batch(self, d_ptrs, c_ptrs, bsize_ptr, nframes, flags=0, stream=0):
        self._check(self.L.od_tdlt_fwd_batch(self._h, _ptr_array(d_ptrs), _ptr_array(c_ptrs),
                                             bsize_ptr, nframes, flags, stream), "od_tdlt_fwd_batch")

    def inv_batch(self, c_ptrs, d_ptrs, bsize_ptr, nframes, flags=0, stream=0):
        self._check(self.L.od_tdlt_inv_batch(self._h, _ptr_array(c_ptrs), _ptr_array(d_ptrs),
                                             bsize_ptr, nframes, flags, stream), "od_tdlt_inv_batch")

    def fwd_batch_pix(self, d_ptrs, pix_ptrs, bsize_ptr, nframes, flags=0, xstride=1, bitdepth_mode=1,
                      lossless=0, stream=0):
        """Device pointers: the forward stage straight from planes of 8/16-bit reference-buffer
        samples (od_ref_buf_to_coeff fused into the load)."""
        self._check(self.L.od_tdlt_fwd_batch_pix(self._h, _ptr_array(d_ptrs), _ptr_array(pix_ptrs),
                                                 bsize_ptr, nframes, flags, xstride, bitdepth_mode,
                                                 lossless, stream), "od_tdlt_fwd_batch_pix")

    def inv_batch_pix(self, pix_ptrs, d_ptrs, bsize_ptr, nframes, flags=0, xstride=1, bitdepth_mode=1,
                      lossless=0, stream=0):
        """Device pointers: the inverse stage straight to planes of 8/16-bit reference-buffer
        samples (od_coeff_to_ref_buf fused into the store)."""
        self._check(self.L.od_tdlt_inv_batch_pix(self._h, _ptr_array(pix_ptrs), _ptr_array(d_ptrs),
                                                 bsize_ptr, nframes, flags, xstride, bitdepth_mode,
                                                 lossless, stream), "od_tdlt_inv_batch_pix")

    def img_copy_pad_batch(self, dst_ptrs, dst_ystride, dst_frame_stride, dst_xstride, dst_bitdepth, border,
                           src_ptrs, src_ystride, src_frame_stride, src_xstride, src_bitdepth, nframes,
                           stream=0):
        """Device pointers: the encoder's input side (od_img_copy_pad) for a batch of pictures."""
        I4, L4 = C.c_int * NPLANES_MAX, C.c_longlong * NPLANES_MAX
        pad = lambda v: list(v) + [0] * (NPLANES_MAX - len(v))  # noqa: E731
        self._check(self.L.od_tdlt_img_copy_pad_batch(
            self._h, _ptr_array(dst_ptrs), I4(*pad(dst_ystride)), L4(*pad(dst_frame_stride)), dst_xstride,
            dst_bitdepth, border, _ptr_array(src_ptrs), I4(*pad(src_ystride)), L4(*pad(src_frame_stride)),
            src_xstride, src_bitdepth, nframes, stream), "od_tdlt_img_copy_pad_batch")

    def cfl_pred_batch(self, l_ptr, d_luma_ptr, bsize_ptr, nframes, stream=0):
        """Device pointers: chroma-from-luma predictor planes (od_tdlt_cfl_pred_batch)."""
        self._check(self.L.od_tdlt_cfl_pred_batch(self._h, l_ptr, d_luma_ptr, bsize_ptr, nframes,
                                                  stream), "od_tdlt_cfl_pred_batch")

    def dering_batch(self, out_ptrs, in_ptrs, flags_ptr, bskip_ptrs, q, nframes, stream=0):
        """Device pointers: deringing of the reconstructed planes (od_tdlt_dering_batch)."""
        qa = (C.c_int * NPLANES_MAX)(*q)
        self._check(self.L.od_tdlt_dering_batch(self._h, _ptr_array(out_ptrs), _ptr_array(in_ptrs),
                                                flags_ptr, _ptr_array(bskip_ptrs), qa, nframes, stream),
                    "od_tdlt_dering_batch")

    def smooth_batch(self, c_ptrs, bsize_ptr, q, nframes, stream=0):
        """Device pointers: keyframe bilinear smoothing in place (od_tdlt_smooth_batch)."""
        qa = (C.c_int * NPLANES_MAX)(*q)
        self._check(self.L.od_tdlt_smooth_batch(self._h, _ptr_array(c_ptrs), bsize_ptr, qa, nframes,
                                                stream), "od_tdlt_smooth_batch")

    def set_scan_tables(self, tables, lens):
        """tables: four contiguous uint8 numpy arrays of (x, y) pairs (the reference's
        OD_ZIGZAG4/8/16/32); lens: how many entries of each the reference's layouts order."""
        self._check(self.L.od_tdlt_set_scan_tables(self._h, _ptr_array([t.ctypes.data for t in tables]),
                                                   (C.c_int * 4)(*lens)), "od_tdlt_set_scan_tables")

    def coding_order_batch(self, v_ptrs, d_ptrs, bsize_ptr, nframes, to_coding=True, stream=0):
        self._check(self.L.od_tdlt_coding_order_batch(self._h, _ptr_array(v_ptrs), _ptr_array(d_ptrs),
                                                      bsize_ptr, nframes, 1 if to_coding else 0, stream),
                    "od_tdlt_coding_order_batch")

    def pix_to_coeff_batch(self, c_ptr, pix_ptr, pli, nframes, xstride=1, bitdepth_mode=1,
                           lossless=0, stream=0):
        self._check(self.L.od_tdlt_pix_to_coeff_batch(self._h, c_ptr, pix_ptr, pli, nframes, xstride,
                                                      bitdepth_mode, lossless, stream),
                    "od_tdlt_pix_to_coeff_batch")

    def coeff_to_pix_batch(self, pix_ptr, c_ptr, pli, nframes, xstride=1, bitdepth_mode=1,
                           lossless=0, stream=0):
        self._check(self.L.od_tdlt_coeff_to_pix_batch(self._h, pix_ptr, c_ptr, pli, nframes, xstride,
                                                      bitdepth_mode, lossless, stream),
                    "od_tdlt_coeff_to_pix_batch")

    # ---- layer 2: host pointers (numpy arrays), synchronous -----------------
    def fwd_frames(self, d_planes, c_planes, bsize, flags=0):
        n = bsize.shape[0]
        self._check(self.L.od_tdlt_fwd_frames(
            self._h, _ptr_array([a.ctypes.data for a in d_planes]),
            _ptr_array([a.ctypes.data for a in c_planes]), bsize.ctypes.data, n, flags),
            "od_tdlt_fwd_frames")

    def inv_frames(self, c_planes, d_planes, bsize, flags=0):
        n = bsize.shape[0]
        self._check(self.L.od_tdlt_inv_frames(
            self._h, _ptr_array([a.ctypes.data for a in c_planes]),
            _ptr_array([a.ctypes.data for a in d_planes]), bsize.ctypes.data, n, flags),
            "od_tdlt_inv_frames")

    def cfl_pred_frames(self, l, d_luma, bsize):
        """Host arrays: l (nframes, fh/2, fw/2) <- d_luma (nframes, fh, fw), bsize maps."""
        n = bsize.shape[0]
        self._check(self.L.od_tdlt_cfl_pred_frames(self._h, l.ctypes.data, d_luma.ctypes.data,
                                                   bsize.ctypes.data, n), "od_tdlt_cfl_pred_frames")

    def dering_frames(self, out_planes, in_planes, flags, bskip, q):
        """Host arrays: planes (nframes, h, w) int32; flags (nframes, nvsb, nhsb) uint8; bskip[p]
        (nframes, nvsb*8, nhsb*8) uint8; q: one quantizer per plane."""
        qa = (C.c_int * NPLANES_MAX)(*q)
        self._check(self.L.od_tdlt_dering_frames(
            self._h, _ptr_array([a.ctypes.data for a in out_planes]),
            _ptr_array([a.ctypes.data for a in in_planes]), flags.ctypes.data,
            _ptr_array([a.ctypes.data for a in bskip]), qa, flags.shape[0]), "od_tdlt_dering_frames")

    def smooth_frames(self, c_planes, bsize, q):
        """Host arrays, in place: planes (nframes, h, w) int32, bsize maps, one quantizer per plane."""
        qa = (C.c_int * NPLANES_MAX)(*q)
        self._check(self.L.od_tdlt_smooth_frames(
            self._h, _ptr_array([a.ctypes.data for a in c_planes]), bsize.ctypes.data, qa,
            bsize.shape[0]), "od_tdlt_smooth_frames")

    def coding_order_frames(self, v_planes, d_planes, bsize, to_coding=True):
        """Host arrays: to_coding fills v_planes from d_planes, else the other way round."""
        self._check(self.L.od_tdlt_coding_order_frames(
            self._h, _ptr_array([a.ctypes.data for a in v_planes]),
            _ptr_array([a.ctypes.data for a in d_planes]), bsize.ctypes.data, bsize.shape[0],
            1 if to_coding else 0), "od_tdlt_coding_order_frames")

    def _pix_args(self, pix_planes, xstride):
        ys = (C.c_int * NPLANES_MAX)()
        fs = (C.c_longlong * NPLANES_MAX)()
        for i, a in enumerate(pix_planes):
            ys[i] = a.strides[1]
            fs[i] = a.strides[0]
        return ys, fs

    def fwd_frames_pix(self, d_planes, pix_planes, bsize, flags=0, xstride=1, bitdepth_mode=1,
                       lossless=0):
        """pix_planes[p]: numpy (nframes, h, w_padded...) uint8/int16 arrays; row and frame
        strides are taken from the arrays (they may be views into padded buffers)."""
        n = bsize.shape[0]
        ys, fs = self._pix_args(pix_planes, xstride)
        self._check(self.L.od_tdlt_fwd_frames_pix(
            self._h, _ptr_array([a.ctypes.data for a in d_planes]),
            _ptr_array([a.ctypes.data for a in pix_planes]), ys, fs, bsize.ctypes.data, n, flags,
            xstride, bitdepth_mode, lossless), "od_tdlt_fwd_frames_pix")

    def inv_frames_pix(self, pix_planes, d_planes, bsize, flags=0, xstride=1, bitdepth_mode=1,
                       lossless=0):
        n = bsize.shape[0]
        ys, fs = self._pix_args(pix_planes, xstride)
        self._check(self.L.od_tdlt_inv_frames_pix(
            self._h, _ptr_array([a.ctypes.data for a in pix_planes]), ys, fs,
            _ptr_array([a.ctypes.data for a in d_planes]), bsize.ctypes.data, n, flags,
            xstride, bitdepth_mode, lossless), "od_tdlt_inv_frames_pix")


def region_order(pic_width, pic_height, decs=(0, 1, 1), rows_in_flight=3):
    """Host-only: the order in which the inverse kernel walks the regions of a frame
    (od_tdlt_region_order).  Returns a numpy uint16 array of region indices."""
    import numpy as np
    g = Geometry()
    g.pic_width, g.pic_height, g.nplanes = pic_width, pic_height, len(decs)
    for i, d in enumerate(decs):
        g.xdec[i] = d
        g.ydec[i] = d
    out = np.zeros(65536, np.uint16)
    n = lib().od_tdlt_region_order(C.byref(g), rows_in_flight, out.ctypes.data, out.size)
    if n < 0:
        raise TdltError("od_tdlt_region_order failed: %d" % n)
    return out[:n].copy()


def fwd_batch_static(pic_width, pic_height, decs, d_ptrs, c_ptrs, bsize_ptr, nframes, flags=0, stream=0):
    """od_tdlt_fwd_batch_static: the forward batch call without a context and without any
    allocation (plan records in a static device array); current CUDA device.  Returns the
    library's error code instead of raising, so that callers can test the size limit."""
    g = Geometry()
    g.pic_width, g.pic_height, g.nplanes = pic_width, pic_height, len(decs)
    for i, d in enumerate(decs):
        g.xdec[i] = d
        g.ydec[i] = d
    return lib().od_tdlt_fwd_batch_static(C.byref(g), _ptr_array(d_ptrs), _ptr_array(c_ptrs),
                                          bsize_ptr, nframes, flags, stream)


def host_register(array):
    """Page-locks a numpy array's memory (od_tdlt_host_register)."""
    rc = lib().od_tdlt_host_register(array.ctypes.data, array.nbytes)
    if rc != 0:
        raise TdltError("od_tdlt_host_register failed: %d" % rc)


def host_unregister(array):
    rc = lib().od_tdlt_host_unregister(array.ctypes.data)
    if rc != 0:
        raise TdltError("od_tdlt_host_unregister failed: %d" % rc)


def launch_count():
    return lib().od_tdlt_launch_count()

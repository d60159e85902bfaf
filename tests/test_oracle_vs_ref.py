"""Pins the C restatement (oracle/tdlt_oracle.c) against the UNMODIFIED reference
compiled into oracle/_ref/libdaala_ref.so -- element by element, on the kind of
vectors the reference's own self-test uses (src/dct.c:3322-3391, 3439-3469:
uniform ranges +-256/255, +-5, +-300 scaled by OD_COEFF_SCALE; impulses and
extremes :3759-3830; all sign patterns of +-676 for the filters,
src/filter.c:2041-2116).  CPU only.
"""
import ctypes as C
import os
import sys

import numpy as np
import pytest

import oracle_libs as ol
from oracle_libs import ptr

SIZES = [0, 1, 2, 3]


def _vectors(n, rng, count=400):
    """Rows of length n: self-test style ranges, impulses, extremes."""
    rows = []
    for lo, hi in ((-256, 255), (-5, 5), (-300, 300)):
        rows.append(rng.integers(lo, hi + 1, size=(count, n)) * 16)
    rows.append(rng.integers(-(1 << 15), 1 << 15, size=(count, n)))
    imp = np.zeros((2 * n, n), dtype=np.int64)
    for i in range(n):
        imp[2 * i, i] = 255 * 16
        imp[2 * i + 1, i] = -256 * 16
    rows.append(imp)
    ext = np.where(rng.integers(0, 2, size=(count, n)) > 0, 255 * 16, -256 * 16)
    rows.append(ext)
    return np.ascontiguousarray(np.concatenate(rows, 0), dtype=np.int32)


@pytest.mark.parametrize("bs", SIZES)
def test_dct_1d_matches_reference(reflib, bs):
    n = 4 << bs
    rng = np.random.default_rng(100 + bs)
    X = _vectors(n, rng)
    port = ol.port()
    for x in X:
        y_p = np.zeros(n, np.int32)
        y_r = np.zeros(n, np.int32)
        port.oc_fdct1d(bs, ptr(y_p), ptr(x), 1)
        reflib.ref_fdct1d(bs, ptr(y_r), ptr(x), 1)
        assert np.array_equal(y_p, y_r)
        # inverse on RAW vectors as coefficients (not only on fdct outputs)
        x_p = np.zeros(n, np.int32)
        x_r = np.zeros(n, np.int32)
        port.oc_idct1d(bs, ptr(x_p), 1, ptr(x))
        reflib.ref_idct1d(bs, ptr(x_r), 1, ptr(x))
        assert np.array_equal(x_p, x_r)
        # perfect reconstruction
        port.oc_idct1d(bs, ptr(x_p), 1, ptr(y_p))
        assert np.array_equal(x_p, x)


def _ref_2d(reflib, name):
    f = getattr(reflib, name)
    f.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
    return f


@pytest.mark.parametrize("bs", SIZES)
def test_dct_2d_matches_reference_c_and_simd(reflib, bs):
    n = 4 << bs
    rng = np.random.default_rng(200 + bs)
    port = ol.port()
    nm = "%dx%d" % (n, n)
    fwd = [_ref_2d(reflib, "od_bin_fdct" + nm)]
    inv = [_ref_2d(reflib, "od_bin_idct" + nm)]
    if bs == 0:
        fwd += [_ref_2d(reflib, "od_bin_fdct4x4_sse2"), _ref_2d(reflib, "od_bin_fdct4x4_sse41")]
        inv += [_ref_2d(reflib, "od_bin_idct4x4_sse2"), _ref_2d(reflib, "od_bin_idct4x4_sse41")]
    if bs == 1:
        fwd += [_ref_2d(reflib, "od_bin_fdct8x8_" + s) for s in ("sse2", "sse41", "avx2")]
        inv += [_ref_2d(reflib, "od_bin_idct8x8_" + s) for s in ("sse2", "sse41", "avx2")]
    blocks = _vectors(n * n, rng, count=60 if bs < 3 else 20).reshape(-1, n, n)
    for x in blocks:
        x = np.ascontiguousarray(x)
        y_p = np.zeros((n, n), np.int32)
        port.oc_fdct2d(bs, ptr(y_p), n, ptr(x), n)
        for f in fwd:
            y_r = np.zeros((n, n), np.int32)
            f(ptr(y_r), n, ptr(x), n)
            assert np.array_equal(y_p, y_r)
        x_p = np.zeros((n, n), np.int32)
        port.oc_idct2d(bs, ptr(x_p), n, ptr(x), n)
        for f in inv:
            x_r = np.zeros((n, n), np.int32)
            f(ptr(x_r), n, ptr(x), n)
            assert np.array_equal(x_p, x_r)
        port.oc_idct2d(bs, ptr(x_p), n, ptr(y_p), n)
        assert np.array_equal(x_p, x)


@pytest.mark.parametrize("f", SIZES)
def test_lapping_filters_match_reference(reflib, f):
    n = 4 << f
    rng = np.random.default_rng(300 + f)
    port = ol.port()
    pre = getattr(reflib, "od_pre_filter%d" % n)
    post = getattr(reflib, "od_post_filter%d" % n)
    X = [_vectors(n, rng)]
    if n <= 8:
        # every sign pattern of +-676 (reference src/filter.c:2041-2116)
        pats = np.array([[676 if (m >> k) & 1 else -676 for k in range(n)] for m in range(1 << n)])
        X.append(pats.astype(np.int32))
    else:
        X.append(np.where(rng.integers(0, 2, size=(4096, n)) > 0, 676, -676).astype(np.int32))
    for x in np.concatenate(X, 0):
        x = np.ascontiguousarray(x)
        a = np.zeros(n, np.int32)
        b = np.zeros(n, np.int32)
        port.oc_pre_filter(f, ptr(a), ptr(x))
        pre(ptr(b), ptr(x))
        assert np.array_equal(a, b)
        port.oc_post_filter(f, ptr(a), ptr(x))
        post(ptr(b), ptr(x))
        assert np.array_equal(a, b)
        port.oc_pre_filter(f, ptr(a), ptr(x))
        port.oc_post_filter(f, ptr(b), ptr(a))
        assert np.array_equal(b, x)


@pytest.mark.parametrize("ln", [2, 3, 4, 5])
def test_haar_matches_reference(reflib, ln):
    n = 1 << ln
    rng = np.random.default_rng(400 + ln)
    port = ol.port()
    for _ in range(20):
        x = rng.integers(-4096, 4096, size=(n, n)).astype(np.int32)
        a = np.zeros((n, n), np.int32)
        b = np.zeros((n, n), np.int32)
        port.oc_haar(ptr(a), n, ptr(x), n, ln)
        reflib.od_haar(ptr(b), n, ptr(x), n, ln)
        assert np.array_equal(a, b)
        xa = np.zeros((n, n), np.int32)
        xb = np.zeros((n, n), np.int32)
        port.oc_haar_inv(ptr(xa), n, ptr(x), n, ln)
        reflib.od_haar_inv(ptr(xb), n, ptr(x), n, ln)
        assert np.array_equal(xa, xb)
        port.oc_haar_inv(ptr(xa), n, ptr(a), n, ln)
        assert np.array_equal(xa, x)


@pytest.mark.parametrize("n", [4, 8, 16])
def test_tf_matches_reference(reflib, n):
    rng = np.random.default_rng(500 + n)
    port = ol.port()
    m = 2 * n
    for _ in range(10):
        src = rng.integers(-30000, 30000, size=(m, m)).astype(np.int32)
        for name, args in (("tf_up_hv", (n,)), ("tf_down_hv", (m,)),
                           ("tf_up_hv_lp", (n, n, m)), ("tf_up_h_lp", (n, m)),
                           ("tf_up_v_lp", (n, m)), ("tf_filter_2d", (m,)),
                           ("tf_filter_inv_2d", (m,))):
            a = np.zeros((m, m), np.int32)
            b = np.zeros((m, m), np.int32)
            getattr(port, "oc_" + name)(ptr(a), m, ptr(src), m, *args)
            getattr(reflib, "od_" + name)(ptr(b), m, ptr(src), m, *args)
            assert np.array_equal(a, b), name
        # up then down is the identity
        a = np.zeros((m, m), np.int32)
        b = np.zeros((m, m), np.int32)
        port.oc_tf_up_hv(ptr(a), m, ptr(src), m, n)
        port.oc_tf_down_hv(ptr(b), m, ptr(a), m, m)
        assert np.array_equal(b, src)


@pytest.mark.parametrize("curr,dest,filt", [(1, 0, 0), (2, 0, 1), (3, 1, 0), (3, 0, 1), (2, 2, 0)])
def test_convert_block_down_matches_reference(reflib, curr, dest, filt):
    rng = np.random.default_rng(600 + curr * 10 + dest)
    port = ol.port()
    n = 4 << curr
    src = rng.integers(-30000, 30000, size=(n, n)).astype(np.int32)
    a = np.zeros((n, n), np.int32)
    b = np.zeros((n, n), np.int32)
    port.oc_convert_block_down(ptr(a), n, ptr(src), n, curr, dest, filt)
    reflib.od_convert_block_down(ptr(b), n, ptr(src), n, curr, dest, filt)
    assert np.array_equal(a, b)


PLANE_CASES = [
    # pic_w, pic_h, dec, keyframe
    (352, 288, 0, 1), (352, 288, 1, 1), (352, 288, 0, 0), (352, 288, 1, 0),
    (200, 136, 0, 1), (200, 136, 1, 1), (64, 64, 0, 1), (64, 64, 1, 0),
]


@pytest.mark.parametrize("pic_w,pic_h,dec,key", PLANE_CASES)
def test_plane_stages_match_reference(reflib, pic_w, pic_h, dec, key):
    rng = np.random.default_rng(700 + pic_w + dec * 7 + key)
    with ol.RefCtx(pic_w, pic_h) as rc:
        H, W = rc.fh >> dec, rc.fw >> dec
        for trial in range(3):
            if trial == 0:
                bsize = ol.random_quadtree_bsize(rc.nhsb, rc.nvsb, rng)
            elif trial == 1:
                bsize = np.full((rc.nvsb * 4, rc.nhsb * 4), rng.integers(0, 4), np.uint8)
            else:
                bsize = ol.random_quadtree_bsize(rc.nhsb, rc.nvsb, rng, 1.0, 1.0, 0.5)
            c = rng.integers(-2048, 2048, size=(H, W)).astype(np.int32)
            d_r, lap_r = rc.fwd_plane(c, bsize, dec, key, impl=0)
            d_s, _ = rc.fwd_plane(c, bsize, dec, key, impl=1)
            d_p, lap_p = ol.port_fwd_plane(c, bsize, dec, pic_w, pic_h, key)
            assert np.array_equal(lap_p, lap_r)
            assert np.array_equal(d_p, d_r)
            assert np.array_equal(d_s, d_r)  # C == SIMD dispatch
            # inverse on the forward output (round trip) and on raw coefficients
            for dd in (d_r, rng.integers(-30000, 30000, size=(H, W)).astype(np.int32)):
                c_r = rc.inv_plane(dd, bsize, dec, key, impl=0)
                c_s = rc.inv_plane(dd, bsize, dec, key, impl=1)
                c_p = ol.port_inv_plane(dd, bsize, dec, pic_w, pic_h, key)
                assert np.array_equal(c_p, c_r)
                assert np.array_equal(c_s, c_r)
            assert np.array_equal(rc.inv_plane(d_r, bsize, dec, key), c)


@pytest.mark.parametrize("dec", [0, 1])
def test_haar_plane_matches_reference(reflib, dec):
    rng = np.random.default_rng(800 + dec)
    pic_w, pic_h = 176, 144
    with ol.RefCtx(pic_w, pic_h) as rc:
        H, W = rc.fh >> dec, rc.fw >> dec
        bsize = np.full((rc.nvsb * 4, rc.nhsb * 4), 3, np.uint8)
        c = rng.integers(-128, 128, size=(H, W)).astype(np.int32)
        d_r, _ = rc.fwd_plane(c, bsize, dec, 1, use_haar=1)
        d_p, _ = ol.port_fwd_plane(c, bsize, dec, pic_w, pic_h, 1, use_haar=1)
        assert np.array_equal(d_p, d_r)
        c_r = rc.inv_plane(d_r, bsize, dec, 0, use_haar=1)
        c_p = ol.port_inv_plane(d_r, bsize, dec, pic_w, pic_h, 0, use_haar=1)
        assert np.array_equal(c_p, c_r)
        assert np.array_equal(c_r, c)


@pytest.mark.parametrize("mode,lossless,xstride", [
    (1, 0, 1), (1, 1, 1), (2, 1, 1), (3, 1, 1), (1, 0, 2), (2, 1, 2), (3, 1, 2), (2, 0, 2)])
def test_pixel_conversion_matches_reference(reflib, mode, lossless, xstride):
    rng = np.random.default_rng(900 + mode * 10 + lossless + xstride)
    port = ol.port()
    w, h = 48, 20
    ystride = (w + 16) * xstride
    if xstride == 1:
        src = rng.integers(0, 256, size=(h, ystride)).astype(np.uint8)
    else:
        src = rng.integers(0, 4096, size=(h, ystride // 2)).astype(np.int16).view(np.uint8).reshape(h, ystride)
    src = np.ascontiguousarray(src)
    a = np.zeros((h, w), np.int32)
    b = np.zeros((h, w), np.int32)
    port.oc_buf_to_coeff(mode, ptr(a), w, lossless, ptr(src), xstride, ystride, w, h)
    reflib.ref_buf_to_coeff(mode, ptr(b), w, lossless, ptr(src), xstride, ystride, w, h)
    assert np.array_equal(a, b)
    co = rng.integers(-5000, 5000, size=(h, w)).astype(np.int32)
    pa = np.zeros((h, ystride), np.uint8)
    pb = np.zeros((h, ystride), np.uint8)
    port.oc_coeff_to_buf(mode, ptr(pa), xstride, ystride, ptr(co), w, lossless, w, h)
    reflib.ref_coeff_to_buf(mode, ptr(pb), xstride, ystride, ptr(co), w, lossless, w, h)
    assert np.array_equal(pa, pb)


@pytest.mark.parametrize("chroma_bs,dec", [(0, 1), (1, 1), (2, 1), (3, 1), (0, 0), (1, 0), (2, 0), (3, 0)])
def test_resample_luma_block_matches_reference(reflib, chroma_bs, dec):
    """od_resample_luma_coeffs (src/intra.c:72-109), one block: the 4:2:0 branches (4x4 chroma
    from four 4x4 luma blocks; upper-left quarter otherwise) and the 4:4:4 copy."""
    rng = np.random.default_rng(1000 + chroma_bs * 2 + dec)
    port = ol.port()
    # the call sites' arguments (src/encode.c:1512-1517): chroma_bs is the LUMA block size,
    # bs the chroma transform size
    bs = max(chroma_bs, 1) - 1 if dec else chroma_bs
    n = 4 << bs
    ln = n << dec
    f = reflib.od_resample_luma_coeffs
    f.restype = None
    f.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
    for trial in range(4):
        luma = rng.integers(-30000, 30000, size=(ln, ln + 5)).astype(np.int32)
        a = np.full((n, n + 3), 77, np.int32)
        b = a.copy()
        f(ptr(a), n + 3, ptr(luma), ln + 5, dec, dec, bs, chroma_bs)
        port.oc_resample_luma_coeffs(ptr(b), n + 3, ptr(luma), ln + 5, dec, dec, bs, chroma_bs)
        assert np.array_equal(a, b)
        assert np.all(a[:, n:] == 77)


@pytest.mark.parametrize("w,h", [(64, 64), (352, 288), (200, 136)])
def test_cfl_pred_plane_matches_reference(reflib, w, h):
    """Whole-frame driver of the restatement against the reference called per chroma leaf."""
    rng = np.random.default_rng(1100 + w)
    fw, fh = ol.padded_dims(w, h)
    nhsb, nvsb = fw // 32, fh // 32
    maps = [ol.random_quadtree_bsize(nhsb, nvsb, rng),
            ol.random_quadtree_bsize(nhsb, nvsb, rng, 1.0, 0.9, 0.5)]
    maps += [np.full((nvsb * 4, nhsb * 4), k, np.uint8) for k in range(4)]
    for bsize in maps:
        d = rng.integers(-30000, 30000, size=(fh, fw)).astype(np.int32)
        assert np.array_equal(ol.port_cfl_pred_plane(d, bsize), ol.ref_cfl_pred_plane(d, bsize))


# ---- deringing ----------------------------------------------------
def test_dering_threshold_table_follows_its_formula():
    port = ol.port()
    tab = (C.c_int16 * 18).in_dll(port, "OC_DERING_THRESH_Q8")
    want = [round(256 * min(3, max(.5, 1.08 * (np.sqrt(2) * 2.0 ** (i + 8) / 65536) ** .16))) for i in range(18)]
    assert list(tab) == want


@pytest.mark.parametrize("bsi", [0, 1])
def test_dering_filters_match_reference(reflib, bsi):
    """The direction and orthogonal filters, one block, all 8 directions, against the reference's
    C functions and whatever its runtime dispatch installs (the SSE2 twins)."""
    rng = np.random.default_rng(1400 + bsi)
    port = ol.port()
    n = 4 << bsi
    P = 38
    cfun = [(reflib.od_filter_dering_direction_4x4_c, reflib.od_filter_dering_orthogonal_4x4_c),
            (reflib.od_filter_dering_direction_8x8_c, reflib.od_filter_dering_orthogonal_8x8_c)][bsi]
    with ol.RefCtx(64, 64) as rc:
        for trial in range(40):
            d = trial % 8
            thr = int(rng.choice([0, 1, 7, 40, 300, 2000]))   # 14*thr fits 16 bits (SSE2 sums)
            # 8-bit video range (|x| <= 2048 + filter overshoot): the SSE2 twins work in 16 bits, so
            # 30000 - x must not wrap
            buf = rng.integers(-400, 400, size=(n + 6, P)).astype(np.int16) * int(rng.choice([1, 3, 6]))
            if trial % 5 == 0:
                buf[:, :4] = 30000       # outside the frame
            org = rng.integers(-3000, 3000, size=(n, n + 2)).astype(np.int16)
            inp = buf.ctypes.data + 2 * (3 * P + 3)
            outs = []
            for who in ("c", "dispatch", "port"):
                y1 = np.zeros((n, n + 1), np.int16)
                y2 = np.zeros((n, n + 1), np.int16)
                if who == "c":
                    cfun[0](ptr(y1), n + 1, C.c_void_p(inp), thr, d)
                    cfun[1](ptr(y2), n + 1, C.c_void_p(inp), ptr(org), n + 2, thr, d)
                elif who == "dispatch":
                    reflib.ref_dering_direction(C.c_void_p(rc.h), bsi, ptr(y1), n + 1, C.c_void_p(inp), thr, d)
                    reflib.ref_dering_orthogonal(C.c_void_p(rc.h), bsi, ptr(y2), n + 1, C.c_void_p(inp),
                                                 ptr(org), n + 2, thr, d)
                else:
                    port.oc_dering_direction(ptr(y1), n + 1, C.c_void_p(inp), 2 + bsi, thr, d)
                    port.oc_dering_orthogonal(ptr(y2), n + 1, C.c_void_p(inp), ptr(org), n + 2, 2 + bsi, thr, d)
                outs.append((y1, y2))
            for y1, y2 in outs[1:]:
                assert np.array_equal(y1, outs[0][0]) and np.array_equal(y2, outs[0][1])


@pytest.mark.parametrize("nhsb,nvsb,q", [(3, 3, 60), (1, 1, 200), (4, 2, 9), (2, 5, 1000), (1, 3, 33)])
def test_dering_frame_matches_reference(reflib, nhsb, nvsb, q):
    """Whole frames: every superblock position (corners, edges, interior), random skip maps and
    decisions, luma directions reused by chroma; the restatement's frame loop against od_dering
    called the way the decoder calls it, C filters and dispatched filters."""
    rng = np.random.default_rng(1500 + nhsb * 10 + nvsb)
    decs = (0, 1, 1)
    qs = (q, max(1, q * 3 // 4), max(1, q // 2))
    with ol.RefCtx(64, 64) as rc:
        for trial in range(3):
            planes, flags, bskip = ol.dering_test_frame(rng, nhsb, nvsb, decs, skip_prob=(0.2, 0.6, 0.95)[trial])
            if trial == 2:
                planes = [rng.integers(-2600, 2600, size=p.shape).astype(np.int32) for p in planes]
            a = ol.ref_dering_frame(rc, planes, decs, flags, bskip, qs, impl=0)
            b = ol.ref_dering_frame(rc, planes, decs, flags, bskip, qs, impl=1)
            c = ol.port_dering_frame(planes, decs, flags, bskip, qs)
            changed = 0
            for p in range(3):
                assert np.array_equal(a[p], b[p]), "reference C vs SIMD, plane %d" % p
                assert np.array_equal(a[p], c[p]), "restatement, plane %d" % p
                changed += int(np.count_nonzero(a[p] != planes[p]))
            if trial == 0 and flags.any():
                assert changed > 0          # the test frame does exercise the filter


@pytest.mark.parametrize("w,h,q", [(64, 64, 30), (352, 288, 200), (200, 136, 1000)])
def test_smooth_frame_matches_reference(reflib, w, h, q):
    """Keyframe bilinear smoothing (od_smooth_recursive / od_bilinear_smooth) of every
    superblock coded as one 32x32 block, luma and 4:2:0 chroma."""
    rng = np.random.default_rng(1600 + w)
    fw, fh = ol.padded_dims(w, h)
    nhsb, nvsb = fw // 32, fh // 32
    decs = (0, 1, 1)
    qs = (q, max(1, 3 * q // 4), max(1, q // 2))
    for trial in range(3):
        bsize = ol.random_quadtree_bsize(nhsb, nvsb, rng, 0.4, 0.6, 0.5) if trial < 2 else \
            np.full((nvsb * 4, nhsb * 4), 3, np.uint8)
        yy, xx = np.mgrid[0:fh, 0:fw]
        luma = (xx * rng.uniform(-3, 3) + yy * rng.uniform(-3, 3) + rng.normal(0, (3, 40, 400)[trial], size=(fh, fw)))
        planes = [np.clip(luma, -2040, 2040).astype(np.int32)]
        planes += [np.ascontiguousarray(planes[0][::2, ::2] // 2 + rng.integers(-20, 20, size=(fh // 2, fw // 2))).astype(np.int32)
                   for _ in range(2)]
        a = ol.ref_smooth_frame(planes, decs, bsize, qs)
        b = ol.port_smooth_frame(planes, decs, bsize, qs)
        for p in range(3):
            assert np.array_equal(a[p], b[p]), "plane %d" % p
        if (bsize[::4, ::4] == 3).any() and trial == 0 and q >= 200:   # smooth block, coarse quantizer
            assert any(np.count_nonzero(a[p] != planes[p]) for p in range(3))


@pytest.mark.parametrize("bs", [0, 1, 2, 3])
def test_coding_order_block_matches_reference(reflib, bs):
    """od_raster_to_coding_order / od_coding_order_to_raster (src/partition.c:143-199), the
    restatement driven by the reference's own scan tables."""
    rng = np.random.default_rng(1900 + bs)
    port = ol.port()
    n = 4 << bs
    keep, sa, la = ol._scan_arg(ol.scan_tables())
    ncoded = min(n * n, 512)              # a 32x32 block has 512 coded coefficients
    for f in (reflib.od_raster_to_coding_order, reflib.od_coding_order_to_raster):
        f.restype = None
    for trial in range(3):
        src = rng.integers(-30000, 30000, size=(n, n + 3)).astype(np.int32)
        a = np.zeros(n * n, np.int32)
        b = np.zeros(n * n, np.int32)
        reflib.od_raster_to_coding_order(ptr(a), n, ptr(src), n + 3)
        port.oc_raster_to_coding_order(ptr(b), n, ptr(src), n + 3, sa, la)
        assert np.array_equal(a, b)
        assert not a[ncoded:].any()                                          # left alone
        if n < 32:
            assert sorted(a.tolist()) == sorted(src[:, :n].ravel().tolist())     # a permutation
        ra = np.full((n, n + 2), 9, np.int32)
        rb = ra.copy()
        reflib.od_coding_order_to_raster(ptr(ra), n + 2, ptr(a), n)
        port.oc_coding_order_to_raster(ptr(rb), n + 2, ptr(a), n, sa, la)
        assert np.array_equal(ra, rb)
        assert np.all(ra[:, n:] == 9)
        covered = ra[:, :n] != 9
        assert np.count_nonzero(covered) >= ncoded - 3 and np.array_equal(ra[:, :n][covered], src[:, :n][covered])


@pytest.mark.parametrize("dec", [0, 1])
def test_coding_order_plane_round_trip_and_layout(reflib, dec):
    """The whole-plane driver: every block's vector equals the reference's for that block, at the
    block-linear position; the way back restores the plane."""
    rng = np.random.default_rng(1950 + dec)
    tables = ol.scan_tables()
    port = ol.port()
    port.oc_block_linear_offset.restype = C.c_longlong
    nhsb, nvsb = 3, 2
    bsize = ol.random_quadtree_bsize(nhsb, nvsb, rng)
    H, W = (nvsb * 32) >> dec, (nhsb * 32) >> dec
    d = rng.integers(-30000, 30000, size=(H, W)).astype(np.int32)
    v = ol.port_coding_order_plane(d, bsize, dec, tables, True)
    back = ol.port_coding_order_plane(v, bsize, dec, tables, False, out=np.full_like(d, 123456))
    covered = back != 123456
    assert np.array_equal(back[covered], d[covered])
    # everything is covered except the uncoded half of the 32x32 blocks (16x16 with 4:2:0: all coded)
    n32 = int(np.count_nonzero(bsize[::4, ::4] == 3)) if dec == 0 else 0
    assert np.count_nonzero(~covered) == 512 * n32
    reflib.od_raster_to_coding_order.restype = None
    flat = v.ravel()
    for uy in range(nvsb * 4):
        for ux in range(nhsb * 4):
            obs = int(bsize[uy, ux])
            lu = max(obs - 1, 0)
            if (ux & ((1 << lu) - 1)) or (uy & ((1 << lu) - 1)):
                continue
            n = max(4, (4 << obs) >> dec)
            x0, y0 = (ux * 8) >> dec, (uy * 8) >> dec
            want = np.zeros(n * n, np.int32)
            blk = np.ascontiguousarray(d[y0:y0 + n, x0:x0 + n])
            reflib.od_raster_to_coding_order(ptr(want), n, ptr(blk), n)
            off = port.oc_block_linear_offset(x0, y0, W, 32 >> dec)
            assert np.array_equal(flat[off:off + n * n], want)


def test_c_frame_drivers_equal_the_python_loops(reflib):
    """The C frame drivers of oracle/ref_glue.c that the benches TIME (ref_dering_frame,
    ref_smooth_frame, ref_cfl_pred_frame, ref_coding_order_plane) make the same reference
    calls as the Python loops above: identical planes."""
    rng = np.random.default_rng(2100)
    decs = (0, 1, 1)
    with ol.RefCtx(96, 64) as rc:
        nhsb, nvsb = rc.nhsb, rc.nvsb
        planes, flags, bskip = ol.dering_test_frame(rng, nhsb, nvsb, decs, skip_prob=0.4)
        q = (60, 45, 30)
        for impl in (0, 1):
            a = ol.ref_dering_frame(rc, planes, decs, flags, bskip, q, impl=impl)
            b = ol.ref_dering_frame_c(rc, planes, decs, flags, bskip, q, impl=impl)
            for p in range(3):
                assert np.array_equal(a[p], b[p]), "deringing plane %d impl %d" % (p, impl)
        bsize = ol.random_quadtree_bsize(nhsb, nvsb, rng, 0.4, 0.6, 0.5)
        a = ol.ref_smooth_frame(planes, decs, bsize, (200, 150, 100))
        b = ol.ref_smooth_frame_c(rc, planes, decs, bsize, (200, 150, 100))
        for p in range(3):
            assert np.array_equal(a[p], b[p]), "smoothing plane %d" % p
        d_luma = rng.integers(-30000, 30000, size=planes[0].shape).astype(np.int32)
        bsize = ol.random_quadtree_bsize(nhsb, nvsb, rng)
        assert np.array_equal(ol.ref_cfl_pred_plane(d_luma, bsize), ol.ref_cfl_pred_plane_c(d_luma, bsize))
        for dec in (0, 1):
            d = rng.integers(-30000, 30000, size=planes[dec].shape).astype(np.int32)
            v1 = ol.ref_coding_order_plane(d, bsize, dec, True)
            v2 = ol.ref_coding_order_plane_c(d, bsize, dec, True)
            assert np.array_equal(v1, v2), "to coding order, dec %d" % dec
            b1 = ol.ref_coding_order_plane(v1, bsize, dec, False, out=np.full_like(d, 77))
            b2 = ol.ref_coding_order_plane_c(v1, bsize, dec, False, out=np.full_like(d, 77))
            assert np.array_equal(b1, b2), "to raster, dec %d" % dec


@pytest.mark.parametrize("quality,keyframe_rate", [(20, 1), (0, 1), (30, 4)])
def test_codec_roundtrip_harness_c_vs_simd(reflib, quality, keyframe_rate):
    """The whole-codec harness (encoder -> packets -> decoder through the reference's own API)
    with the C tables and with the runtime x86 dispatch: same packets, same pictures -- the
    equality the CUDA-routed run is held to on the GPU (tests/test_zy_dropin_codec.py); at
    quality 0 the decoder returns the input (tools/check_lossless.sh)."""
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
    import make_golden as mg
    frames = [mg.cif_frame(t) for t in range(4)]      # (the decoder puts pictures out one packet late)
    pk0, sz0, out0 = ol.codec_roundtrip(reflib, frames, quality, 0, keyframe_rate=keyframe_rate)
    pk1, sz1, out1 = ol.codec_roundtrip(reflib, frames, quality, 1, keyframe_rate=keyframe_rate)
    assert len(pk0) > 100 and np.array_equal(sz0, sz1) and np.array_equal(pk0, pk1)
    for a, b in zip(out0, out1):
        assert all(np.array_equal(x, y) for x, y in zip(a, b))
    if quality == 0:
        for a, b in zip(out0, frames):
            assert all(np.array_equal(x, y) for x, y in zip(a, b)), "lossless mode must return the input"
    else:
        err = np.abs(out0[0][0].astype(int) - frames[0][0].astype(int)).mean()
        assert err < 8, "decoded picture is far from the input (%.1f)" % err


IMG_PAD_CASES = [(352, 288, 8, 8), (352, 288, 10, 8), (200, 136, 10, 12), (33, 17, 8, 10),
                 (64, 64, 12, 10), (130, 70, 16, 8), (1, 1, 8, 8)]


@pytest.mark.parametrize("pic_w,pic_h,src_bits,dst_bits", IMG_PAD_CASES)
def test_img_copy_pad_matches_reference(reflib, pic_w, pic_h, src_bits, dst_bits):
    """The encoder's input side: od_img_plane_copy's bit-depth conversions
    with their clamps, the low-pass extension to the padded frame and od_img_edge_ext, luma
    and 4:2:0 chroma planes, pictures that are no multiple of 64 (or of 2)."""
    rng = np.random.default_rng(2300 + pic_w + src_bits)
    fw, fh = ol.padded_dims(pic_w, pic_h)
    for dec in (0, 1):
        pw, ph = (pic_w + (1 << dec) - 1) >> dec, (pic_h + (1 << dec) - 1) >> dec
        if src_bits == 8:
            pic = rng.integers(0, 256, size=(ph, pw)).astype(np.uint8)
        else:      # out-of-range samples too: the copy clamps
            pic = rng.integers(-40, (1 << src_bits) + 40, size=(ph, pw)).clip(-32768, 32767).astype(np.int16)
        a = ol.ref_img_copy_pad_plane(pic, src_bits, fw >> dec, fh >> dec, dec, dst_bits)
        b = ol.port_img_copy_pad_plane(pic, src_bits, fw >> dec, fh >> dec, 64 >> dec, dst_bits)
        assert a.shape == b.shape and np.array_equal(a, b), "dec %d" % dec
        assert not (a == 77).all()

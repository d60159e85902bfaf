"""Known-answer tests against tests/golden/tdlt_golden.npz -- outputs of the
UNMODIFIED reference (tests/golden/make_golden.py, run where /root/reference is
mounted).  The CPU half holds the C restatement to them; the GPU half holds the
CUDA library to them, through the C-ABI, on a box where the reference does not
exist.  Bit-exact everywhere (np.array_equal on integer arrays)."""
import ctypes as C
import hashlib
import os
import sys

import numpy as np
import pytest

import oracle_libs as ol
from oracle_libs import ptr

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))


@pytest.fixture(scope="module")
def G():
    return np.load(os.path.join(HERE, "golden", "tdlt_golden.npz"))


def _cif_planes(fw=384, fh=320):
    import make_golden as mg
    y, u, v = mg.cif_frame(0)
    out = []
    for p, dec in ((y, 0), (u, 1), (v, 1)):
        out.append((mg.pad_plane(p, fh >> dec, fw >> dec).astype(np.int32) - 128) << 4)
    return out


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


# ------------------------------------------------------------------ CPU: port --
@pytest.mark.parametrize("bs", [0, 1, 2, 3])
def test_port_dcts(G, bs):
    P = ol.port()
    n = 4 << bs
    for x, f, i in zip(G["dct1d_%d_x" % bs], G["dct1d_%d_fwd" % bs], G["dct1d_%d_inv" % bs]):
        x = np.ascontiguousarray(x)
        y = np.zeros(n, np.int32)
        P.oc_fdct1d(bs, ptr(y), ptr(x), 1)
        assert np.array_equal(y, f)
        P.oc_idct1d(bs, ptr(y), 1, ptr(x))
        assert np.array_equal(y, i)
    for x, f, i in zip(G["dct2d_%d_x" % bs], G["dct2d_%d_fwd" % bs], G["dct2d_%d_inv" % bs]):
        x = np.ascontiguousarray(x)
        y = np.zeros((n, n), np.int32)
        P.oc_fdct2d(bs, ptr(y), n, ptr(x), n)
        assert np.array_equal(y, f)
        P.oc_idct2d(bs, ptr(y), n, ptr(x), n)
        assert np.array_equal(y, i)


@pytest.mark.parametrize("f", [0, 1, 2, 3])
def test_port_filters(G, f):
    P = ol.port()
    n = 4 << f
    for x, a, b in zip(G["filt_%d_x" % f], G["filt_%d_pre" % f], G["filt_%d_post" % f]):
        x = np.ascontiguousarray(x)
        y = np.zeros(n, np.int32)
        P.oc_pre_filter(f, ptr(y), ptr(x))
        assert np.array_equal(y, a)
        P.oc_post_filter(f, ptr(y), ptr(x))
        assert np.array_equal(y, b)


def test_port_haar_and_tf(G):
    P = ol.port()
    for ln in (2, 3, 4, 5):
        n = 1 << ln
        x = np.ascontiguousarray(G["haar_%d_x" % ln])
        y = np.zeros((n, n), np.int32)
        P.oc_haar(ptr(y), n, ptr(x), n, ln)
        assert np.array_equal(y, G["haar_%d_fwd" % ln])
        P.oc_haar_inv(ptr(y), n, ptr(x), n, ln)
        assert np.array_equal(y, G["haar_%d_inv" % ln])
    for n in (4, 8, 16):
        m = 2 * n
        src = np.ascontiguousarray(G["tf_%d_src" % n])
        for name, args in (("tf_up_hv", (n,)), ("tf_down_hv", (m,)), ("tf_up_hv_lp", (n, n, m)),
                           ("tf_up_h_lp", (n, m)), ("tf_up_v_lp", (n, m)), ("tf_filter_2d", (m,)),
                           ("tf_filter_inv_2d", (m,))):
            out = np.zeros((m, m), np.int32)
            getattr(P, "oc_" + name)(ptr(out), m, ptr(src), m, *args)
            assert np.array_equal(out, G["tf_%d_%s" % (n, name)]), name
    for curr, dest, filt in CBD_CASES:
        n = 4 << curr
        src = np.ascontiguousarray(G["cbd_%d_%d_%d_src" % (curr, dest, filt)])
        out = np.zeros((n, n), np.int32)
        P.oc_convert_block_down(ptr(out), n, ptr(src), n, curr, dest, filt)
        assert np.array_equal(out, G["cbd_%d_%d_%d_out" % (curr, dest, filt)])


CBD_CASES = [(1, 0, 0), (2, 0, 1), (3, 1, 0), (3, 0, 1), (2, 2, 0), (3, 2, 1)]


def test_port_planes(G):
    pic_w, pic_h = (int(v) for v in G["plane_pic"])
    bsize = G["plane_bsize"]
    for pli, dec in enumerate((0, 1, 1)):
        c, dq = G["plane_%d_c" % pli], G["plane_%d_dq" % pli]
        for key in (0, 1):
            d, _ = ol.port_fwd_plane(c, bsize, dec, pic_w, pic_h, key)
            assert np.array_equal(d, G["plane_%d_d_key%d" % (pli, key)])
            cq = ol.port_inv_plane(dq, bsize, dec, pic_w, pic_h, key)
            assert np.array_equal(cq, G["plane_%d_cq_key%d" % (pli, key)])
        d, _ = ol.port_fwd_plane(G["plane_%d_cl" % pli], np.full_like(bsize, 3), dec, pic_w, pic_h, 1,
                                 use_haar=1)
        assert np.array_equal(d, G["plane_%d_d_haar" % pli])


def test_port_cfl_pred_plane(G):
    assert np.array_equal(ol.port_cfl_pred_plane(G["plane_0_dq"], G["plane_bsize"]), G["plane_cfl"])


def test_port_coding_order_blocks(G):
    port = ol.port()
    keep, sa, la = ol._scan_arg([G["scan_table_%d" % (4 << k)] for k in range(4)])
    src = np.ascontiguousarray(G["scan_block_src"])
    for bs in range(4):
        n = 4 << bs
        v = np.zeros(n * n, np.int32)
        port.oc_raster_to_coding_order(ol.ptr(v), n, ol.ptr(src), 32, sa, la)
        assert np.array_equal(v, G["scan_block_%d" % n])


def test_port_cif_keyframe(G):
    """BASELINE config 1: CIF keyframe with the block sizes the reference encoder's
    RDO chose; the coefficient planes are pinned by SHA-256."""
    bs = G["cif_bsize"]
    for pli, (c, dec) in enumerate(zip(_cif_planes(), (0, 1, 1))):
        d, _ = ol.port_fwd_plane(c, bs, dec, 352, 288, 1)
        assert _sha(d) == str(G["cif_d_sha256"][pli])
        if pli == 0:
            assert np.array_equal(d[96:104], G["cif_d0_rows"])
        assert np.array_equal(ol.port_inv_plane(d, bs, dec, 352, 288, 1), c)


PIX_CASES = [(1, 0, 1), (1, 1, 1), (2, 0, 2), (2, 1, 2), (3, 1, 2), (1, 0, 2)]


@pytest.mark.parametrize("mode,lossless,xs", PIX_CASES)
def test_port_pixels(G, mode, lossless, xs):
    P = ol.port()
    tag = "pix_%d_%d_%d" % (mode, lossless, xs)
    src, co = np.ascontiguousarray(G[tag + "_src"]), np.ascontiguousarray(G[tag + "_co"])
    h, w = co.shape
    ystride = src.shape[1]
    a = np.zeros((h, w), np.int32)
    P.oc_buf_to_coeff(mode, ptr(a), w, lossless, ptr(src), xs, ystride, w, h)
    assert np.array_equal(a, G[tag + "_coeff"])
    pb = np.zeros((h, ystride), np.uint8)
    P.oc_coeff_to_buf(mode, ptr(pb), xs, ystride, ptr(co), w, lossless, w, h)
    assert np.array_equal(pb, G[tag + "_buf"])


# ------------------------------------------------------------------ GPU: CUDA --
@pytest.mark.gpu
@pytest.mark.parametrize("bs", [0, 1, 2, 3])
def test_cuda_dct_blocks(G, bs):
    import daala_b200
    L = daala_b200.lib()
    n = 4 << bs
    X = np.ascontiguousarray(G["dct2d_%d_x" % bs])
    nb = X.shape[0]
    out = np.zeros_like(X)
    assert L.od_tdlt_dct2d_blocks(bs, 0, out.ctypes.data, n, n * n, X.ctypes.data, n, n * n, nb) == 0
    assert np.array_equal(out, G["dct2d_%d_fwd" % bs])
    assert L.od_tdlt_dct2d_blocks(bs, 1, out.ctypes.data, n, n * n, X.ctypes.data, n, n * n, nb) == 0
    assert np.array_equal(out, G["dct2d_%d_inv" % bs])
    # the reference-signature single-block calls (od_dct_func_2d)
    fw = getattr(L, "od_bin_fdct%dx%d_cuda" % (n, n))
    iv = getattr(L, "od_bin_idct%dx%d_cuda" % (n, n))
    y = np.zeros((n, n), np.int32)
    fw(y.ctypes.data, n, X[0].ctypes.data, n)
    assert np.array_equal(y, G["dct2d_%d_fwd" % bs][0])
    iv(y.ctypes.data, n, X[0].ctypes.data, n)
    assert np.array_equal(y, G["dct2d_%d_inv" % bs][0])


@pytest.mark.gpu
@pytest.mark.parametrize("f", [0, 1, 2, 3])
def test_cuda_filters(G, f):
    import daala_b200
    L = daala_b200.lib()
    X = np.ascontiguousarray(G["filt_%d_x" % f])
    out = np.zeros_like(X)
    assert L.od_tdlt_lap_filter_vectors(f, 0, out.ctypes.data, X.ctypes.data, X.shape[0]) == 0
    assert np.array_equal(out, G["filt_%d_pre" % f])
    assert L.od_tdlt_lap_filter_vectors(f, 1, out.ctypes.data, X.ctypes.data, X.shape[0]) == 0
    assert np.array_equal(out, G["filt_%d_post" % f])


@pytest.mark.gpu
def test_cuda_haar_and_tf(G):
    import daala_b200
    L = daala_b200.lib()
    for ln in (2, 3, 4, 5):
        n = 1 << ln
        x = np.ascontiguousarray(G["haar_%d_x" % ln])
        y = np.zeros((n, n), np.int32)
        L.od_haar_cuda(y.ctypes.data, n, x.ctypes.data, n, ln)
        assert np.array_equal(y, G["haar_%d_fwd" % ln])
        L.od_haar_inv_cuda(y.ctypes.data, n, x.ctypes.data, n, ln)
        assert np.array_equal(y, G["haar_%d_inv" % ln])
    for n in (4, 8, 16):
        m = 2 * n
        src = np.ascontiguousarray(G["tf_%d_src" % n])
        out = np.zeros((m, m), np.int32)
        L.od_tf_up_hv_cuda(out.ctypes.data, m, src.ctypes.data, m, n)
        assert np.array_equal(out, G["tf_%d_tf_up_hv" % n])
        L.od_tf_down_hv_cuda(out.ctypes.data, m, src.ctypes.data, m, m)
        assert np.array_equal(out, G["tf_%d_tf_down_hv" % n])
        for name, args in (("tf_up_hv_lp", (n, n, m)), ("tf_up_h_lp", (n, m)), ("tf_up_v_lp", (n, m)),
                           ("tf_filter_2d", (m,)), ("tf_filter_inv_2d", (m,))):
            out = np.full((m, m), 7, np.int32)
            getattr(L, "od_%s_cuda" % name)(out.ctypes.data, m, src.ctypes.data, m, *args)
            assert np.array_equal(out, G["tf_%d_%s" % (n, name)]), name
    for curr, dest, filt in CBD_CASES:
        n = 4 << curr
        src = np.ascontiguousarray(G["cbd_%d_%d_%d_src" % (curr, dest, filt)])
        out = np.zeros((n, n), np.int32)
        L.od_convert_block_down_cuda(out.ctypes.data, n, src.ctypes.data, n, curr, dest, filt)
        assert np.array_equal(out, G["cbd_%d_%d_%d_out" % (curr, dest, filt)])


@pytest.mark.gpu
def test_cuda_planes(G):
    import daala_b200
    pic_w, pic_h = (int(v) for v in G["plane_pic"])
    bsize = np.ascontiguousarray(G["plane_bsize"])[None]
    with daala_b200.TdltContext(pic_w, pic_h, decs=(0, 1, 1), device=0, max_frames=1) as ctx:
        c = [np.ascontiguousarray(G["plane_%d_c" % p])[None] for p in range(3)]
        dq = [np.ascontiguousarray(G["plane_%d_dq" % p])[None] for p in range(3)]
        for key in (0, 1):
            flags = daala_b200.OD_TDLT_KEYFRAME if key else 0
            d = [np.zeros_like(a) for a in c]
            ctx.fwd_frames(d, c, bsize, flags)
            cq = [np.zeros_like(a) for a in c]
            ctx.inv_frames(cq, dq, bsize, flags)
            for p in range(3):
                assert np.array_equal(d[p][0], G["plane_%d_d_key%d" % (p, key)])
                assert np.array_equal(cq[p][0], G["plane_%d_cq_key%d" % (p, key)])
        cl = [np.ascontiguousarray(G["plane_%d_cl" % p])[None] for p in range(3)]
        d = [np.zeros_like(a) for a in cl]
        hb = np.full_like(bsize, 3)
        ctx.fwd_frames(d, cl, hb, daala_b200.OD_TDLT_KEYFRAME | daala_b200.OD_TDLT_HAAR)
        for p in range(3):
            assert np.array_equal(d[p][0], G["plane_%d_d_haar" % p])
        l = np.zeros((1,) + G["plane_cfl"].shape, np.int32)
        ctx.cfl_pred_frames(l, dq[0], bsize)
        assert np.array_equal(l[0], G["plane_cfl"])


@pytest.mark.gpu
def test_cuda_cif_keyframe(G):
    import daala_b200
    bs = np.ascontiguousarray(G["cif_bsize"])[None]
    c = [np.ascontiguousarray(p)[None] for p in _cif_planes()]
    with daala_b200.TdltContext(352, 288, decs=(0, 1, 1), device=0, max_frames=1) as ctx:
        d = [np.zeros_like(a) for a in c]
        ctx.fwd_frames(d, c, bs, daala_b200.OD_TDLT_KEYFRAME)
        for p in range(3):
            assert _sha(d[p][0]) == str(G["cif_d_sha256"][p])
        assert np.array_equal(d[0][0][96:104], G["cif_d0_rows"])
        c2 = [np.zeros_like(a) for a in c]
        ctx.inv_frames(c2, d, bs, daala_b200.OD_TDLT_KEYFRAME)
        for p in range(3):
            assert np.array_equal(c2[p], c[p])


# ---------------------------------------------- BASELINE configs at their stated size --
@pytest.fixture(scope="module")
def GC():
    return np.load(os.path.join(HERE, "golden", "tdlt_golden_configs.npz"))


def test_port_cif_ten_frames(GC):
    """Config 1 as BASELINE.json states it: all ten frames, each with the map the reference
    encoder's RDO chose for it (the moving square makes them differ)."""
    import make_golden as mg
    assert len({GC["cif10_bsize"][t].tobytes() for t in range(10)}) > 1
    for t in range(10):
        bs = GC["cif10_bsize"][t]
        for pli, (c, dec) in enumerate(zip(mg.planes_to_coeff(mg.cif_frame(t), 384, 320), (0, 1, 1))):
            d, _ = ol.port_fwd_plane(c, bs, dec, 352, 288, 1)
            assert _sha(d) == str(GC["cif10_d_sha256"][t][pli]), (t, pli)


def test_port_hd_frame_rdo_map_and_uniform_8x8(GC):
    """Configs 2 and 3(i) at 1080p: fixed 8x8 blocks, and the RDO map of the textured frame."""
    import make_golden as mg
    c = mg.planes_to_coeff(mg.hd_frame(0), 1920, 1088)
    for key, bs in (("hd_d_sha256", GC["hd_bsize"]), ("hd_d_sha256_8x8", np.full_like(GC["hd_bsize"], 1))):
        for pli, dec in enumerate((0, 1, 1)):
            d, _ = ol.port_fwd_plane(c[pli], bs, dec, 1920, 1080, 1)
            assert _sha(d) == str(GC[key][pli]), (key, pli)


@pytest.mark.gpu
def test_cuda_cif_ten_frames(GC):
    import daala_b200
    import make_golden as mg
    frames = [mg.planes_to_coeff(mg.cif_frame(t), 384, 320) for t in range(10)]
    c = [np.ascontiguousarray(np.stack([f[p] for f in frames])) for p in range(3)]
    bs = np.ascontiguousarray(GC["cif10_bsize"])
    with daala_b200.TdltContext(352, 288, decs=(0, 1, 1), device=0, max_frames=10) as ctx:
        d = [np.zeros_like(a) for a in c]
        ctx.fwd_frames(d, c, bs, daala_b200.OD_TDLT_KEYFRAME)
        for t in range(10):
            for p in range(3):
                assert _sha(d[p][t]) == str(GC["cif10_d_sha256"][t][p]), (t, p)
        c2 = [np.zeros_like(a) for a in c]
        ctx.inv_frames(c2, d, bs, daala_b200.OD_TDLT_KEYFRAME)
        for p in range(3):
            assert np.array_equal(c2[p], c[p])


@pytest.mark.gpu
def test_cuda_hd_frame_rdo_map_and_uniform_8x8(GC):
    import daala_b200
    import make_golden as mg
    c = [np.ascontiguousarray(p)[None] for p in mg.planes_to_coeff(mg.hd_frame(0), 1920, 1088)]
    with daala_b200.TdltContext(1920, 1080, decs=(0, 1, 1), device=0, max_frames=1) as ctx:
        for key, bs in (("hd_d_sha256", GC["hd_bsize"]), ("hd_d_sha256_8x8", np.full_like(GC["hd_bsize"], 1))):
            bs = np.ascontiguousarray(bs)[None]
            d = [np.zeros_like(a) for a in c]
            ctx.fwd_frames(d, c, bs, daala_b200.OD_TDLT_KEYFRAME)
            for p in range(3):
                assert _sha(d[p][0]) == str(GC[key][p]), (key, p)
            c2 = [np.zeros_like(a) for a in c]
            ctx.inv_frames(c2, d, bs, daala_b200.OD_TDLT_KEYFRAME)
            for p in range(3):
                assert np.array_equal(c2[p], c[p])
            if ol.have_ref():
                # and the decoder side on arbitrary coefficients, against the reference itself
                from checker import Checker
                chk = Checker(1920, 1080)
                rng = np.random.default_rng(5)
                dq = [rng.integers(-30000, 30000, size=a.shape).astype(np.int32) for a in c]
                ctx.inv_frames(c2, dq, bs, daala_b200.OD_TDLT_KEYFRAME)
                c_ref = chk.inv_frames(dq, bs, (0, 1, 1), 1)
                for p in range(3):
                    assert np.array_equal(c2[p], c_ref[p]), (key, p)
                chk.close()

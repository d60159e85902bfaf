"""GPU parity tests (run with -m gpu on the B200 box): the CUDA path, called
through the C-ABI, against the CPU checker on identical seeded inputs.
Bit-exact: every comparison is np.array_equal on int32 / uint8 / int16."""
import numpy as np
import pytest

import oracle_libs as ol
from checker import Checker, make_maps

pytestmark = pytest.mark.gpu

DECS = (0, 1, 1)


def _ctx(pic_w, pic_h, max_frames, decs=DECS):
    import daala_b200
    return daala_b200.TdltContext(pic_w, pic_h, decs=decs, device=0, max_frames=max_frames)


def _rand_planes(shapes, nframes, rng, lo=-2048, hi=2048):
    return [rng.integers(lo, hi, size=(nframes,) + s).astype(np.int32) for s in shapes]


FRAME_CASES = [
    # pic_w, pic_h, map kind, keyframe
    (352, 288, "quad", 1), (352, 288, "quad", 0), (352, 288, "deep", 1),
    (352, 288, "uniform0", 1), (352, 288, "uniform1", 0), (352, 288, "uniform2", 1),
    (352, 288, "uniform3", 1), (200, 136, "quad", 1), (200, 136, "deep", 0),
    (64, 64, "quad", 1), (33, 17, "deep", 1), (1920, 1080, "quad", 1),
    (1920, 1080, "uniform1", 1),      # BASELINE config 2 at its stated size
]


@pytest.mark.parametrize("pic_w,pic_h,kind,key", FRAME_CASES)
def test_frames_match_checker(pic_w, pic_h, kind, key):
    import daala_b200
    rng = np.random.default_rng(pic_w * 7 + pic_h * 3 + key + sum(map(ord, kind)))
    nframes = 2 if pic_w < 1000 else 1
    chk = Checker(pic_w, pic_h)
    with _ctx(pic_w, pic_h, nframes) as ctx:
        assert (ctx.frame_width, ctx.frame_height) == (chk.fw, chk.fh)
        bsize = make_maps(kind, nframes, ctx.nhsb, ctx.nvsb, rng)
        c = _rand_planes(ctx.plane_shapes, nframes, rng)
        flags = daala_b200.OD_TDLT_KEYFRAME if key else 0
        d = [np.zeros_like(a) for a in c]
        ctx.fwd_frames(d, c, bsize, flags)
        d_ref = chk.fwd_frames(c, bsize, DECS, key)
        for p in range(3):
            assert np.array_equal(d[p], d_ref[p]), "forward plane %d" % p
        # inverse of the forward output: exact reconstruction of c
        c2 = [np.zeros_like(a) for a in c]
        ctx.inv_frames(c2, d, bsize, flags)
        for p in range(3):
            assert np.array_equal(c2[p], c[p]), "round trip plane %d" % p
        # inverse of arbitrary coefficients (what a decoder feeds it)
        dq = _rand_planes(ctx.plane_shapes, nframes, rng, -30000, 30000)
        ctx.inv_frames(c2, dq, bsize, flags)
        c_ref = chk.inv_frames(dq, bsize, DECS, key)
        for p in range(3):
            assert np.array_equal(c2[p], c_ref[p]), "inverse plane %d" % p
    chk.close()


@pytest.mark.parametrize("nframes", [5, 20, 50])
def test_inverse_batch_sizes_pick_every_region_order(nframes):
    """The inverse kernel walks the batch frame-interleaved and picks one of three region
    orders by batch size (< 16, < 48, >= 48 frames): each of them, on a plane that is several
    regions wide and high, must give the checker's samples for arbitrary coefficients."""
    import daala_b200
    rng = np.random.default_rng(900 + nframes)
    pic_w, pic_h = 400, 250            # 448 x 256 padded: luma 4 x 4 regions, chroma 2 x 2
    chk = Checker(pic_w, pic_h)
    with _ctx(pic_w, pic_h, nframes) as ctx:
        bsize = make_maps("quad", nframes, ctx.nhsb, ctx.nvsb, rng)
        dq = _rand_planes(ctx.plane_shapes, nframes, rng, -20000, 20000)
        c2 = [np.zeros_like(a) for a in dq]
        for rep in range(2):           # twice: the strip buffer is reused with a new epoch
            ctx.inv_frames(c2, dq, bsize, daala_b200.OD_TDLT_KEYFRAME)
            c_ref = chk.inv_frames(dq, bsize, DECS, 1)
            for p in range(3):
                assert np.array_equal(c2[p], c_ref[p]), "inverse plane %d" % p
    chk.close()


@pytest.mark.parametrize("decs", [(0,), (0, 0, 0), (0, 1, 1, 0)])
def test_other_plane_layouts(decs):
    import daala_b200
    rng = np.random.default_rng(77 + len(decs))
    pic_w, pic_h = 176, 144
    chk = Checker(pic_w, pic_h)
    with _ctx(pic_w, pic_h, 1, decs) as ctx:
        bsize = make_maps("deep", 1, ctx.nhsb, ctx.nvsb, rng)
        c = _rand_planes(ctx.plane_shapes, 1, rng)
        d = [np.zeros_like(a) for a in c]
        ctx.fwd_frames(d, c, bsize, daala_b200.OD_TDLT_KEYFRAME)
        d_ref = chk.fwd_frames(c, bsize, decs, 1)
        c2 = [np.zeros_like(a) for a in c]
        ctx.inv_frames(c2, d, bsize, daala_b200.OD_TDLT_KEYFRAME)
        for p in range(len(decs)):
            assert np.array_equal(d[p], d_ref[p])
            assert np.array_equal(c2[p], c[p])
    chk.close()


@pytest.mark.parametrize("pic_w,pic_h", [(352, 288), (1280, 720)])
def test_lossless_haar_mode(pic_w, pic_h):
    """BASELINE config 5 (ci/lossless.sh equivalent): 8-bit pixels -> coeff_shift 0
    -> od_haar -> od_haar_inv -> pixels must be the identity, and the
    coefficients must equal the checker's."""
    import daala_b200
    rng = np.random.default_rng(5)
    chk = Checker(pic_w, pic_h)
    with _ctx(pic_w, pic_h, 1) as ctx:
        bsize = make_maps("uniform3", 1, ctx.nhsb, ctx.nvsb, rng)
        pix = [rng.integers(0, 256, size=(1,) + s).astype(np.uint8) for s in ctx.plane_shapes]
        d = [np.zeros((1,) + s, np.int32) for s in ctx.plane_shapes]
        fl = daala_b200.OD_TDLT_KEYFRAME | daala_b200.OD_TDLT_HAAR
        ctx.fwd_frames_pix(d, pix, bsize, fl, xstride=1, bitdepth_mode=1, lossless=1)
        c = [(a.astype(np.int32) - 128) for a in pix]
        d_ref = chk.fwd_frames(c, bsize, DECS, 1, haar=1)
        for p in range(3):
            assert np.array_equal(d[p], d_ref[p])
        out = [np.zeros_like(a) for a in pix]
        ctx.inv_frames_pix(out, d, bsize, fl, xstride=1, bitdepth_mode=1, lossless=1)
        for p in range(3):
            assert np.array_equal(out[p], pix[p])
    chk.close()


@pytest.mark.parametrize("xstride,mode,lossless", [(1, 1, 0), (2, 2, 0), (2, 3, 1), (1, 2, 1), (2, 1, 1)])
def test_pixel_paths_match_checker(xstride, mode, lossless):
    """coeff_shift paths: 8-bit and 16-bit reference buffers with row padding."""
    import daala_b200
    rng = np.random.default_rng(600 + xstride * 10 + mode)
    pic_w, pic_h = 352, 288
    port = ol.port()
    chk = Checker(pic_w, pic_h)
    with _ctx(pic_w, pic_h, 2) as ctx:
        bsize = make_maps("quad", 2, ctx.nhsb, ctx.nvsb, rng)
        pad = 32
        pix, views = [], []
        for (h, w) in ctx.plane_shapes:
            if xstride == 1:
                buf = rng.integers(0, 256, size=(2, h + 2 * pad, w + 2 * pad)).astype(np.uint8)
            else:
                buf = rng.integers(0, 4096, size=(2, h + 2 * pad, w + 2 * pad)).astype(np.int16)
            pix.append(buf)
            views.append(buf[:, pad:pad + h, pad:pad + w])
        d = [np.zeros((2,) + s, np.int32) for s in ctx.plane_shapes]
        ctx.fwd_frames_pix(d, views, bsize, 1, xstride=xstride, bitdepth_mode=mode, lossless=lossless)
        c = []
        for v in views:
            cc = np.zeros(v.shape, np.int32)
            for f in range(2):
                src = np.ascontiguousarray(v[f])
                port.oc_buf_to_coeff(mode, ol.ptr(cc[f]), v.shape[2], lossless,
                                     ol.ptr(src.view(np.uint8)), xstride, src.strides[0],
                                     v.shape[2], v.shape[1])
                if ol.have_ref():     # od_ref_buf_to_coeff itself (src/state.c:1209)
                    cr = np.zeros_like(cc[f])
                    ol.ref().ref_buf_to_coeff(mode, ol.ptr(cr), v.shape[2], lossless,
                                              ol.ptr(src.view(np.uint8)), xstride, src.strides[0],
                                              v.shape[2], v.shape[1])
                    assert np.array_equal(cr, cc[f])
            c.append(cc)
        d_ref = chk.fwd_frames(c, bsize, DECS, 1)
        for p in range(3):
            assert np.array_equal(d[p], d_ref[p])
        # inverse into fresh padded buffers; only the visible window may change
        outs = [np.full_like(b, 7) for b in pix]
        oviews = [b[:, pad:pad + s[0], pad:pad + s[1]] for b, s in zip(outs, ctx.plane_shapes)]
        dq = [(a // 3) * 3 for a in d]   # a crude "quantiser"
        ctx.inv_frames_pix(oviews, dq, bsize, 1, xstride=xstride, bitdepth_mode=mode, lossless=lossless)
        c_ref = chk.inv_frames(dq, bsize, DECS, 1)
        for p in range(3):
            h, w = ctx.plane_shapes[p]
            exp = np.zeros((2, h, w), pix[p].dtype)
            for f in range(2):
                tmp = np.zeros((h, w * xstride), np.uint8)
                port.oc_coeff_to_buf(mode, ol.ptr(tmp), xstride, w * xstride, ol.ptr(c_ref[p][f]), w,
                                     lossless, w, h)
                exp[f] = tmp.view(pix[p].dtype).reshape(h, w)
                if ol.have_ref():     # od_coeff_to_ref_buf itself (src/state.c:1274)
                    tr = np.zeros_like(tmp)
                    ol.ref().ref_coeff_to_buf(mode, ol.ptr(tr), xstride, w * xstride,
                                              ol.ptr(c_ref[p][f]), w, lossless, w, h)
                    assert np.array_equal(tr, tmp)
            assert np.array_equal(oviews[p], exp)
            border = outs[p].copy()
            border[:, pad:pad + h, pad:pad + w] = 7
            assert (border == 7).all()
    chk.close()


@pytest.mark.parametrize("xstride,mode,lossless", [(1, 1, 0), (2, 2, 0), (2, 3, 1)])
def test_fused_pixel_batches_equal_the_two_pass_path(xstride, mode, lossless):
    """od_tdlt_fwd_batch_pix / od_tdlt_inv_batch_pix (conversion fused into the transform
    kernels' load / store) against the conversion kernel + transform kernel pair, bit for
    bit, on an 8-frame 1080p batch (partial regions at the bottom of every plane); the
    lossless kernels refuse the fused form."""
    import torch
    import daala_b200
    pic_w, pic_h, nframes = 1920, 1080, 8
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(4100 + xstride)
    with _ctx(pic_w, pic_h, 0) as ctx:
        bsize = make_maps("quad", nframes, ctx.nhsb, ctx.nvsb, rng)
        maps = torch.from_numpy(bsize).to(dev)
        if xstride == 1:
            pix_np = [rng.integers(0, 256, size=(nframes,) + s).astype(np.uint8) for s in ctx.plane_shapes]
        else:
            pix_np = [rng.integers(0, 4096, size=(nframes,) + s).astype(np.int16) for s in ctx.plane_shapes]
        pix = [torch.from_numpy(a).to(dev) for a in pix_np]
        c = [torch.empty(a.shape, dtype=torch.int32, device=dev) for a in pix_np]
        d1 = [torch.empty_like(t) for t in c]
        d2 = [torch.zeros_like(t) for t in c]
        st = torch.cuda.current_stream().cuda_stream
        KEY = daala_b200.OD_TDLT_KEYFRAME
        for p in range(3):
            ctx.pix_to_coeff_batch(c[p].data_ptr(), pix[p].data_ptr(), p, nframes, xstride, mode, lossless, st)
        ctx.fwd_batch([t.data_ptr() for t in d1], [t.data_ptr() for t in c], maps.data_ptr(), nframes, KEY, st)
        ctx.fwd_batch_pix([t.data_ptr() for t in d2], [t.data_ptr() for t in pix], maps.data_ptr(), nframes,
                          KEY, xstride, mode, lossless, st)
        torch.cuda.synchronize()
        for p in range(3):
            assert torch.equal(d1[p], d2[p]), "forward plane %d" % p
        dq = [(t // 5) * 5 for t in d1]          # a crude quantiser: the inverse must clamp
        c2 = [torch.empty_like(t) for t in c]
        o1 = [torch.full_like(t, 9) for t in pix]
        o2 = [torch.full_like(t, 9) for t in pix]
        ctx.inv_batch([t.data_ptr() for t in c2], [t.data_ptr() for t in dq], maps.data_ptr(), nframes, KEY, st)
        for p in range(3):
            ctx.coeff_to_pix_batch(o1[p].data_ptr(), c2[p].data_ptr(), p, nframes, xstride, mode, lossless, st)
        ctx.inv_batch_pix([t.data_ptr() for t in o2], [t.data_ptr() for t in dq], maps.data_ptr(), nframes,
                          KEY, xstride, mode, lossless, st)
        torch.cuda.synchronize()
        for p in range(3):
            assert torch.equal(o1[p], o2[p]), "inverse plane %d" % p
        with pytest.raises(daala_b200.TdltError):
            ctx.fwd_batch_pix([t.data_ptr() for t in d2], [t.data_ptr() for t in pix], maps.data_ptr(),
                              nframes, KEY | daala_b200.OD_TDLT_HAAR, xstride, mode, lossless, st)


@pytest.mark.parametrize("pic_w,pic_h,src_bits,dst_bits", [(352, 288, 8, 8), (352, 288, 10, 8), (200, 136, 10, 12),
                                                            (33, 17, 8, 10), (1920, 1080, 8, 8), (130, 70, 16, 8)])
def test_img_copy_pad_matches_checker(pic_w, pic_h, src_bits, dst_bits):
    """The encoder's input side on the device (od_tdlt_img_copy_pad_batch = od_img_copy_pad):
    three pictures per call, all planes, into bordered buffers (OD_BUFFER_PADDING = 64) --
    against the compiled reference's od_img_plane_copy_pad + od_img_edge_ext when it is here,
    else the restatement; nothing outside the bordered planes may change."""
    import torch
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(4300 + pic_w + src_bits)
    nframes, border, guard = 3, 64, 5
    with _ctx(pic_w, pic_h, 0) as ctx:
        fw, fh = ctx.frame_width, ctx.frame_height
        sdt = np.uint8 if src_bits == 8 else np.int16
        ddt = np.uint8 if dst_bits == 8 else np.int16
        srcs, dsts, want = [], [], []
        for dec in DECS:
            pw, ph = (pic_w + (1 << dec) - 1) >> dec, (pic_h + (1 << dec) - 1) >> dec
            if src_bits == 8:
                pic = rng.integers(0, 256, size=(nframes, ph, pw + 3)).astype(np.uint8)     # padded rows
            else:
                pic = rng.integers(-40, (1 << src_bits) + 40, size=(nframes, ph, pw + 3)).clip(-32768, 32767).astype(np.int16)
            srcs.append(pic)
            b = border >> dec
            dsts.append(np.full((nframes, (fh >> dec) + 2 * b + 2 * guard, (fw >> dec) + 2 * b + 2 * guard), 55, ddt))
            per = []
            for f in range(nframes):
                p = np.ascontiguousarray(pic[f, :, :pw])
                if ol.have_ref():
                    per.append(ol.ref_img_copy_pad_plane(p, src_bits, fw >> dec, fh >> dec, dec, dst_bits))
                else:
                    per.append(ol.port_img_copy_pad_plane(p, src_bits, fw >> dec, fh >> dec, b, dst_bits))
            want.append(per)
        ts = [torch.from_numpy(a).to(dev) for a in srcs]
        td = [torch.from_numpy(a).to(dev) for a in dsts]
        isz = np.dtype(ddt).itemsize
        dptr = [t.data_ptr() + ((guard + (border >> dec)) * t.shape[2] + guard + (border >> dec)) * isz
                for t, dec in zip(td, DECS)]
        ctx.img_copy_pad_batch(dptr, [t.shape[2] * isz for t in td], [t.shape[1] * t.shape[2] * isz for t in td],
                               isz, dst_bits, border,
                               [t.data_ptr() for t in ts], [a.strides[1] for a in srcs], [a.strides[0] for a in srcs],
                               np.dtype(sdt).itemsize, src_bits, nframes, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        for p, dec in enumerate(DECS):
            got = td[p].cpu().numpy()
            for f in range(nframes):
                inner = got[f, guard:-guard, guard:-guard]
                assert np.array_equal(inner, want[p][f]), "plane %d frame %d" % (p, f)
                outer = got[f].copy()
                outer[guard:-guard, guard:-guard] = 55
                assert (outer == 55).all(), "wrote outside the bordered plane"


def test_picture_to_coefficients_on_the_device():
    """od_tdlt_img_copy_pad_batch (border 0, packed planes) feeding od_tdlt_fwd_batch_pix: an
    8-bit 200x136 picture goes from the caller's samples to coefficients without a host pass
    and equals the checker's forward stage of the padded picture."""
    import torch
    import daala_b200
    dev = torch.device("cuda", 0)
    pic_w, pic_h, nframes = 200, 136, 2
    rng = np.random.default_rng(4400)
    chk = Checker(pic_w, pic_h)
    with _ctx(pic_w, pic_h, 0) as ctx:
        fw, fh = ctx.frame_width, ctx.frame_height
        bsize = make_maps("quad", nframes, ctx.nhsb, ctx.nvsb, rng)
        pics = [rng.integers(0, 256, size=(nframes, (pic_h + d) >> d, (pic_w + d) >> d)).astype(np.uint8) for d in DECS]
        tp = [torch.from_numpy(a).to(dev) for a in pics]
        planes = [torch.zeros((nframes,) + s, dtype=torch.uint8, device=dev) for s in ctx.plane_shapes]
        d = [torch.zeros((nframes,) + s, dtype=torch.int32, device=dev) for s in ctx.plane_shapes]
        st = torch.cuda.current_stream().cuda_stream
        ctx.img_copy_pad_batch([t.data_ptr() for t in planes], [s[1] for s in ctx.plane_shapes],
                               [s[0] * s[1] for s in ctx.plane_shapes], 1, 8, 0,
                               [t.data_ptr() for t in tp], [a.strides[1] for a in pics], [a.strides[0] for a in pics],
                               1, 8, nframes, st)
        maps = torch.from_numpy(bsize).to(dev)
        ctx.fwd_batch_pix([t.data_ptr() for t in d], [t.data_ptr() for t in planes], maps.data_ptr(), nframes,
                          daala_b200.OD_TDLT_KEYFRAME, 1, 1, 0, st)
        torch.cuda.synchronize()
        c = []
        for p, dec in enumerate(DECS):
            cc = np.zeros((nframes,) + ctx.plane_shapes[p], np.int32)
            for f in range(nframes):
                padded = ol.port_img_copy_pad_plane(np.ascontiguousarray(pics[p][f]), 8, fw >> dec, fh >> dec, 0, 8)
                cc[f] = (padded.astype(np.int32) - 128) << 4
            c.append(cc)
        d_ref = chk.fwd_frames(c, bsize, DECS, 1)
        for p in range(3):
            assert np.array_equal(d[p].cpu().numpy(), d_ref[p]), "plane %d" % p
    chk.close()


# ---- reference-signature entry points (include/daala_tdlt.h section 3) --------
def _selftest_vectors(n, rng, count):
    """The ranges of the reference's self-test (src/dct.c:3322-3391) plus impulses
    and the +-extreme patterns (:3759-3830)."""
    rows = [rng.integers(lo, hi + 1, size=(count, n)) * 16 for lo, hi in ((-256, 255), (-5, 5), (-300, 300))]
    rows.append(rng.integers(-(1 << 15), 1 << 15, size=(count, n)))
    rows.append(np.where(rng.integers(0, 2, size=(count, n)) > 0, 255 * 16, -256 * 16))
    return np.ascontiguousarray(np.concatenate(rows, 0), dtype=np.int32)


@pytest.mark.parametrize("bs", [0, 1, 2, 3])
def test_dct_block_sweep(bs):
    import daala_b200
    L, port = daala_b200.lib(), ol.port()
    n = 4 << bs
    rng = np.random.default_rng(1000 + bs)
    X = _selftest_vectors(n * n, rng, 40 if bs < 3 else 12).reshape(-1, n, n)
    nb = X.shape[0]
    F = np.zeros_like(X)
    I = np.zeros_like(X)
    assert L.od_tdlt_dct2d_blocks(bs, 0, F.ctypes.data, n, n * n, X.ctypes.data, n, n * n, nb) == 0
    assert L.od_tdlt_dct2d_blocks(bs, 1, I.ctypes.data, n, n * n, X.ctypes.data, n, n * n, nb) == 0
    for k in range(nb):
        y = np.zeros((n, n), np.int32)
        port.oc_fdct2d(bs, ol.ptr(y), n, ol.ptr(X[k]), n)
        assert np.array_equal(F[k], y)
        port.oc_idct2d(bs, ol.ptr(y), n, ol.ptr(X[k]), n)
        assert np.array_equal(I[k], y)
    if ol.have_ref():     # the compiled reference itself: C tables and the x86 dispatch
        with ol.RefCtx(64, 64) as rc:
            for impl in (0, 1):
                assert np.array_equal(F, ol.dct2d_blocks("ref", bs, 0, X, rc, impl))
                assert np.array_equal(I, ol.dct2d_blocks("ref", bs, 1, X, rc, impl))
    # perfect reconstruction, and strided single-block calls inside a larger plane
    R = np.zeros_like(X)
    assert L.od_tdlt_dct2d_blocks(bs, 1, R.ctypes.data, n, n * n, F.ctypes.data, n, n * n, nb) == 0
    assert np.array_equal(R, X)
    plane = rng.integers(-2048, 2048, size=(2 * n, 3 * n)).astype(np.int32)
    out = np.full_like(plane, 7)
    fw = getattr(L, "od_bin_fdct%dx%d_cuda" % (n, n))
    off = n * plane.shape[1] + n
    fw(out.ctypes.data + 4 * off, plane.shape[1], plane.ctypes.data + 4 * off, plane.shape[1])
    y = np.zeros((n, n), np.int32)
    src = np.ascontiguousarray(plane[n:2 * n, n:2 * n])
    port.oc_fdct2d(bs, ol.ptr(y), n, ol.ptr(src), n)
    assert np.array_equal(out[n:2 * n, n:2 * n], y)
    out[n:2 * n, n:2 * n] = 7
    assert (out == 7).all()


@pytest.mark.parametrize("f", [0, 1, 2, 3])
def test_lap_filter_sweep(f):
    import daala_b200
    L, port = daala_b200.lib(), ol.port()
    n = 4 << f
    rng = np.random.default_rng(1100 + f)
    X = [_selftest_vectors(n, rng, 300)]
    if n <= 8:   # every sign pattern of +-676 (reference src/filter.c:2041-2116)
        X.append(np.array([[676 if (m >> k) & 1 else -676 for k in range(n)] for m in range(1 << n)], np.int32))
    else:
        X.append(np.where(rng.integers(0, 2, size=(2048, n)) > 0, 676, -676).astype(np.int32))
    X = np.ascontiguousarray(np.concatenate(X, 0))
    P = np.zeros_like(X)
    Q = np.zeros_like(X)
    assert L.od_tdlt_lap_filter_vectors(f, 0, P.ctypes.data, X.ctypes.data, X.shape[0]) == 0
    assert L.od_tdlt_lap_filter_vectors(f, 1, Q.ctypes.data, X.ctypes.data, X.shape[0]) == 0
    y = np.zeros(n, np.int32)
    for k in range(X.shape[0]):
        port.oc_pre_filter(f, ol.ptr(y), ol.ptr(X[k]))
        assert np.array_equal(P[k], y)
        port.oc_post_filter(f, ol.ptr(y), ol.ptr(X[k]))
        assert np.array_equal(Q[k], y)
    if ol.have_ref():
        pre, post = getattr(ol.ref(), "od_pre_filter%d" % n), getattr(ol.ref(), "od_post_filter%d" % n)
        for k in range(X.shape[0]):
            pre(ol.ptr(y), ol.ptr(X[k]))
            assert np.array_equal(P[k], y)
            post(ol.ptr(y), ol.ptr(X[k]))
            assert np.array_equal(Q[k], y)
    R = np.zeros_like(X)
    assert L.od_tdlt_lap_filter_vectors(f, 1, R.ctypes.data, P.ctypes.data, X.shape[0]) == 0
    assert np.array_equal(R, X)


def test_split_and_sb_edge_filters():
    """od_prefilter_split / od_postfilter_split / od_apply_*filter_frame_sbs with the
    reference's own signatures (host pointers, in place)."""
    import daala_b200
    L, port = daala_b200.lib(), ol.port()
    R = ol.ref() if ol.have_ref() else None
    if R:
        for fn in (R.od_prefilter_split, R.od_postfilter_split, R.od_apply_prefilter_frame_sbs,
                   R.od_apply_postfilter_frame_sbs):
            fn.restype = None
    rng = np.random.default_rng(1200)
    for bs in (1, 2, 3):
        n = 4 << bs
        for hf, vf in ((1, 1), (1, 0), (0, 1), (0, 0)):
            stride = n + 9
            a = rng.integers(-2048, 2048, size=(n, stride)).astype(np.int32)
            b = a.copy()
            r = a.copy()
            L.od_prefilter_split_cuda(a.ctypes.data, stride, bs, 0, hf, vf)
            port.oc_prefilter_split(ol.ptr(b), stride, bs, 0, hf, vf)
            assert np.array_equal(a, b)
            if R:
                R.od_prefilter_split(ol.ptr(r), stride, bs, 0, hf, vf)
                assert np.array_equal(a, r)
            L.od_postfilter_split_cuda(a.ctypes.data, stride, bs, 0, 0, None, 0, hf, vf)
            port.oc_postfilter_split(ol.ptr(b), stride, bs, 0, hf, vf)
            assert np.array_equal(a, b)
            if R:
                R.od_postfilter_split(ol.ptr(r), stride, bs, 0, 0, None, 0, hf, vf)
                assert np.array_equal(a, r)
    for nhsb, nvsb, dec in ((5, 3, 0), (6, 4, 1), (1, 1, 0), (1, 3, 1)):
        w, h = (nhsb * 32) >> dec, (nvsb * 32) >> dec
        a = rng.integers(-2048, 2048, size=(h, w)).astype(np.int32)
        b = a.copy()
        r = a.copy()
        L.od_apply_prefilter_frame_sbs_cuda(a.ctypes.data, w, nhsb, nvsb, dec, dec)
        port.oc_apply_prefilter_frame_sbs(ol.ptr(b), w, nhsb, nvsb, dec, dec)
        assert np.array_equal(a, b)
        if R:
            R.od_apply_prefilter_frame_sbs(ol.ptr(r), w, nhsb, nvsb, dec, dec)
            assert np.array_equal(a, r)
        L.od_apply_postfilter_frame_sbs_cuda(a.ctypes.data, w, nhsb, nvsb, dec, dec, 0, None, 0)
        port.oc_apply_postfilter_frame_sbs(ol.ptr(b), w, nhsb, nvsb, dec, dec)
        assert np.array_equal(a, b)
        if R:
            R.od_apply_postfilter_frame_sbs(ol.ptr(r), w, nhsb, nvsb, dec, dec, 0, None, 0)
            assert np.array_equal(a, r)


CFL_CASES = [(352, 288, "quad"), (352, 288, "deep"), (200, 136, "quad"), (33, 17, "deep"),
             (64, 64, "uniform0"), (352, 288, "uniform1"), (352, 288, "uniform2"),
             (352, 288, "uniform3"), (1920, 1080, "quad")]


@pytest.mark.parametrize("pic_w,pic_h,kind", CFL_CASES)
def test_cfl_predictor_planes_match_checker(pic_w, pic_h, kind):
    """Chroma-from-luma predictor planes (od_resample_luma_coeffs over every chroma leaf
    block) against the reference called leaf by leaf."""
    rng = np.random.default_rng(pic_w + 5 * pic_h + sum(map(ord, kind)))
    nframes = 3 if pic_w < 1000 else 1
    chk = Checker(pic_w, pic_h)
    with _ctx(pic_w, pic_h, nframes) as ctx:
        bsize = make_maps(kind, nframes, ctx.nhsb, ctx.nvsb, rng)
        d_luma = rng.integers(-30000, 30000, size=(nframes, chk.fh, chk.fw)).astype(np.int32)
        l = np.full((nframes, chk.fh // 2, chk.fw // 2), 12345, np.int32)
        ctx.cfl_pred_frames(l, d_luma, bsize)
        assert np.array_equal(l, chk.cfl_pred_frames(d_luma, bsize))
    chk.close()


def test_cfl_predictor_needs_420():
    import daala_b200
    with _ctx(64, 64, 1, decs=(0, 0, 0)) as ctx:
        l = np.zeros((1, 64, 64), np.int32)
        with pytest.raises(daala_b200.TdltError):
            ctx.cfl_pred_frames(l, l.copy(), np.zeros((1, 8, 8), np.uint8))


@pytest.mark.parametrize("obs,dec", [(0, 1), (1, 1), (2, 1), (3, 1), (0, 0), (1, 0), (2, 0), (3, 0)])
def test_resample_luma_block(obs, dec):
    """od_resample_luma_coeffs with the reference's signature, one block, the arguments of
    its call sites: the 4:2:0 merge, the 4:2:0 upper-left quarter and the 4:4:4 copy."""
    import daala_b200
    L, port = daala_b200.lib(), ol.port()
    rng = np.random.default_rng(1300 + 2 * obs + dec)
    bs = max(obs, 1) - 1 if dec else obs
    n = 4 << bs
    ln = n << dec
    for trial in range(3):
        luma = rng.integers(-30000, 30000, size=(ln, ln + 7)).astype(np.int32)
        a = np.full((n, n + 5), 99, np.int32)
        b = a.copy()
        L.od_resample_luma_coeffs_cuda(a.ctypes.data, n + 5, luma.ctypes.data, ln + 7, dec, dec, bs, obs)
        port.oc_resample_luma_coeffs(ol.ptr(b), n + 5, ol.ptr(luma), ln + 7, dec, dec, bs, obs)
        assert np.array_equal(a, b)
        if ol.have_ref():         # od_resample_luma_coeffs itself (src/intra.c:72)
            r = np.full((n, n + 5), 99, np.int32)
            ol.ref().od_resample_luma_coeffs.restype = None
            ol.ref().od_resample_luma_coeffs(ol.ptr(r), n + 5, ol.ptr(luma), ln + 7, dec, dec, bs, obs)
            assert np.array_equal(a, r)


def _dering_batch(rng, nframes, nhsb, nvsb, decs, skip_prob):
    fr = [ol.dering_test_frame(rng, nhsb, nvsb, decs, skip_prob=skip_prob) for _ in range(nframes)]
    planes = [np.stack([f[0][p] for f in fr]) for p in range(len(decs))]
    flags = np.stack([f[1] for f in fr])
    bskip = [np.stack([f[2][p] for f in fr]) for p in range(len(decs))]
    return planes, flags, bskip


@pytest.mark.parametrize("pic_w,pic_h,q,skip_prob", [
    (352, 288, 60, 0.3), (352, 288, 400, 0.7), (200, 136, 12, 0.0), (64, 64, 100, 0.5),
    (33, 17, 33, 0.2), (20, 70, 90, 0.3), (90, 20, 50, 0.1), (1920, 1080, 80, 0.5)])
def test_dering_frames_match_checker(pic_w, pic_h, q, skip_prob):
    """The deringing filter over whole frames (every superblock position, random decisions and
    skip maps, chroma reusing luma's directions) against the reference's od_dering."""
    rng = np.random.default_rng(pic_w + 3 * pic_h + q)
    nframes = 2 if pic_w < 1000 else 1
    chk = Checker(pic_w, pic_h)
    qs = (q, max(1, 3 * q // 4), max(1, q // 2))
    with _ctx(pic_w, pic_h, nframes) as ctx:
        planes, flags, bskip = _dering_batch(rng, nframes, ctx.nhsb, ctx.nvsb, DECS, skip_prob)
        out = [np.full_like(p, 77777) for p in planes]
        ctx.dering_frames(out, planes, flags, bskip, qs)
        want = chk.dering_frames(planes, DECS, flags, bskip, qs)
        for p in range(3):
            assert np.array_equal(out[p], want[p]), "plane %d" % p
        assert any(np.count_nonzero(out[p] != planes[p]) for p in range(3))
    chk.close()


def test_dering_444_and_all_clear():
    """4:4:4 planes (8x8 filter blocks on every plane) and a batch whose decisions are all
    clear (pure copy)."""
    rng = np.random.default_rng(1700)
    decs = (0, 0, 0)
    chk = Checker(96, 64)
    with _ctx(96, 64, 2, decs=decs) as ctx:
        planes, flags, bskip = _dering_batch(rng, 2, ctx.nhsb, ctx.nvsb, decs, 0.3)
        out = [np.zeros_like(p) for p in planes]
        ctx.dering_frames(out, planes, flags, bskip, (90, 70, 70))
        want = chk.dering_frames(planes, decs, flags, bskip, (90, 70, 70))
        for p in range(3):
            assert np.array_equal(out[p], want[p]), "plane %d" % p
        flags[:] = 0
        ctx.dering_frames(out, planes, flags, bskip, (90, 70, 70))
        for p in range(3):
            assert np.array_equal(out[p], planes[p])
    chk.close()


@pytest.mark.parametrize("pic_w,pic_h,q,decs", [
    (352, 288, 200, (0, 1, 1)), (200, 136, 1000, (0, 1, 1)), (64, 64, 30, (0, 1, 1)),
    (96, 64, 300, (0, 0, 0)), (1920, 1080, 150, (0, 1, 1))])
def test_smooth_frames_match_checker(pic_w, pic_h, q, decs):
    """Keyframe bilinear smoothing of the 32x32-coded superblocks, in place, against the
    reference's od_smooth_recursive."""
    rng = np.random.default_rng(pic_w + pic_h + q)
    nframes = 2 if pic_w < 1000 else 1
    chk = Checker(pic_w, pic_h)
    qs = (q, max(1, 3 * q // 4), max(1, q // 2))
    with _ctx(pic_w, pic_h, nframes, decs=decs) as ctx:
        bsize = np.stack([ol.random_quadtree_bsize(ctx.nhsb, ctx.nvsb, rng, 0.4, 0.6, 0.5)
                          for _ in range(nframes)])
        yy, xx = np.mgrid[0:chk.fh, 0:chk.fw]
        luma = np.stack([xx * rng.uniform(-3, 3) + yy * rng.uniform(-3, 3) +
                         rng.normal(0, 3 + 30 * f, size=(chk.fh, chk.fw)) for f in range(nframes)])
        planes = [np.clip(luma, -2040, 2040).astype(np.int32)]
        for dec in decs[1:]:
            planes.append(np.ascontiguousarray(planes[0][:, ::1 << dec, ::1 << dec] // 2).astype(np.int32))
        want = chk.smooth_frames(planes, decs, bsize, qs)
        got = [p.copy() for p in planes]
        ctx.smooth_frames(got, bsize, qs)
        for p in range(len(decs)):
            assert np.array_equal(got[p], want[p]), "plane %d" % p
        if q >= 150:
            assert any(np.count_nonzero(got[p] != planes[p]) for p in range(len(decs)))
    chk.close()


@pytest.mark.parametrize("bsi", [0, 1])
def test_dering_block_filters(bsi):
    """The deringing filters with the reference's dispatch-table signatures, one block, all
    eight directions, installed through od_tdlt_dering_vtbl_init_cuda."""
    import ctypes as C
    import daala_b200
    L, port = daala_b200.lib(), ol.port()
    DIRF = C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int)
    ORTF = C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int)
    dtab, otab = (DIRF * 2)(), (ORTF * 2)()
    L.od_tdlt_dering_vtbl_init_cuda.argtypes = [C.c_void_p, C.c_void_p]
    L.od_tdlt_dering_vtbl_init_cuda.restype = None
    L.od_tdlt_dering_vtbl_init_cuda(dtab, otab)
    rng = np.random.default_rng(1800 + bsi)
    n, P = 4 << bsi, 38
    ref_pitch = 38                      # OD_FILT_BSTRIDE, the row pitch the reference's `in` has
    rc = ol.RefCtx(64, 64) if ol.have_ref() else None
    for trial in range(24):
        d = trial % 8
        thr = int(rng.choice([0, 1, 7, 40, 300, 2000]))
        buf = rng.integers(-400, 400, size=(n + 6, P)).astype(np.int16) * int(rng.choice([1, 3, 6]))
        if trial % 5 == 0:
            buf[:, :4] = 30000
        org = rng.integers(-3000, 3000, size=(n, n + 2)).astype(np.int16)
        inp = buf.ctypes.data + 2 * (3 * P + 3)
        y1, y2 = np.full((n, n + 1), 5, np.int16), np.full((n, n + 1), 5, np.int16)
        w1, w2 = y1.copy(), y2.copy()
        dtab[bsi](y1.ctypes.data, n + 1, inp, thr, d)
        otab[bsi](y2.ctypes.data, n + 1, inp, org.ctypes.data, n + 2, thr, d)
        port.oc_dering_direction(ol.ptr(w1), n + 1, C.c_void_p(inp), 2 + bsi, thr, d)
        port.oc_dering_orthogonal(ol.ptr(w2), n + 1, C.c_void_p(inp), ol.ptr(org), n + 2, 2 + bsi, thr, d)
        assert np.array_equal(y1, w1) and np.array_equal(y2, w2)
        if rc is not None and P == ref_pitch:
            # the reference's own filters through its dispatch table (C and the SSE2 twins;
            # the SSE2 ones need 30000 - x and 14*threshold in 16 bits: thresholds <= 2000 do)
            for impl in (0, 1):
                r1, r2 = np.full((n, n + 1), 5, np.int16), np.full((n, n + 1), 5, np.int16)
                ol.ref_dering_block(rc, impl, bsi, r1, r2, n + 1, inp, org, n + 2, thr, d)
                assert np.array_equal(y1, r1) and np.array_equal(y2, r2), (impl, trial)


@pytest.mark.parametrize("pic_w,pic_h,kind,decs", [
    (352, 288, "quad", (0, 1, 1)), (200, 136, "deep", (0, 1, 1)), (64, 64, "uniform3", (0, 1, 1)),
    (96, 64, "quad", (0, 0, 0)), (352, 288, "uniform0", (0, 1, 1)), (20, 70, "quad", (0, 1, 1)),
    (1920, 1080, "quad", (0, 1, 1))])
def test_coding_order_frames_match_checker(pic_w, pic_h, kind, decs):
    """Raster <-> coding order of every transform block with the reference's scan tables: the
    block-linear vectors against the restatement (pinned block by block against the reference's
    od_raster_to_coding_order on the CPU), the uncovered entries left alone, and the way back."""
    rng = np.random.default_rng(pic_w * 3 + pic_h + sum(map(ord, kind)))
    nframes = 2 if pic_w < 1000 else 1
    tables = ol.scan_tables()
    with _ctx(pic_w, pic_h, nframes, decs=decs) as ctx:
        ctx.set_scan_tables(tables, ol.SCAN_LEN)
        bsize = make_maps(kind, nframes, ctx.nhsb, ctx.nvsb, rng)
        d = _rand_planes(ctx.plane_shapes, nframes, rng, -30000, 30000)
        v = [np.full_like(a, 424242) for a in d]
        ctx.coding_order_frames(v, d, bsize, True)
        for p, dec in enumerate(decs):
            for f in range(nframes):
                want = ol.port_coding_order_plane(d[p][f], bsize[f], dec, tables, True,
                                                  out=np.full_like(d[p][f], 424242))
                assert np.array_equal(v[p][f], want), "to coding order, plane %d frame %d" % (p, f)
                if ol.have_ref() and pic_w < 1000:
                    want = ol.ref_coding_order_plane(d[p][f], bsize[f], dec, True,
                                                     out=np.full_like(d[p][f], 424242))
                    assert np.array_equal(v[p][f], want), "vs reference, plane %d frame %d" % (p, f)
        back = [np.full_like(a, -7) for a in d]
        ctx.coding_order_frames(v, back, bsize, False)
        for p, dec in enumerate(decs):
            for f in range(nframes):
                want = ol.port_coding_order_plane(v[p][f], bsize[f], dec, tables, False,
                                                  out=np.full_like(d[p][f], -7))
                assert np.array_equal(back[p][f], want), "to raster, plane %d frame %d" % (p, f)
                if ol.have_ref() and pic_w < 1000:
                    want = ol.ref_coding_order_plane(v[p][f], bsize[f], dec, False,
                                                     out=np.full_like(d[p][f], -7))
                    assert np.array_equal(back[p][f], want), "vs reference, plane %d frame %d" % (p, f)
            covered = back[p] != -7
            assert np.array_equal(back[p][covered], d[p][covered])


@pytest.mark.parametrize("pic_w,pic_h,kind,decs,pix", [
    (352, 288, "quad", (0, 1, 1), 0), (352, 288, "uniform3", (0, 1, 1), 0), (200, 136, "deep", (0, 1, 1), 1),
    (352, 288, "quad", (0, 0, 0), 0), (1920, 1080, "quad", (0, 1, 1), 1), (64, 64, "uniform0", (0, 1, 1), 0)])
def test_fused_coding_order_store_and_load(pic_w, pic_h, kind, decs, pix):
    """OD_TDLT_CODING_ORDER: the forward kernel's store writes every block's coding-order
    vector (block-linear plane) and the inverse kernel's load reads it, with no permutation
    pass.  Forward: equal to the checker's forward stage followed by the REFERENCE's
    od_raster_to_coding_order block by block (the restatement where the reference is absent),
    entries the scans do not order untouched.  Inverse: equal to od_init_skipped_coeffs
    (keyframe: cleared) + od_coding_order_to_raster + the checker's inverse stage.  Also from
    / to 8-bit pixels (pix)."""
    import torch
    import daala_b200
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(4500 + pic_w + len(kind))
    nframes = 2
    tables = ol.scan_tables()
    chk = Checker(pic_w, pic_h) if decs == DECS else None
    KEY, CO = daala_b200.OD_TDLT_KEYFRAME, daala_b200.OD_TDLT_CODING_ORDER
    with _ctx(pic_w, pic_h, 0, decs=decs) as ctx:
        bsize = make_maps(kind, nframes, ctx.nhsb, ctx.nvsb, rng)
        maps = torch.from_numpy(bsize).to(dev)
        st = torch.cuda.current_stream().cuda_stream
        if pix:
            px_np = [rng.integers(0, 256, size=(nframes,) + s).astype(np.uint8) for s in ctx.plane_shapes]
            c_np = [(a.astype(np.int32) - 128) << 4 for a in px_np]
            px = [torch.from_numpy(a).to(dev) for a in px_np]
        else:
            c_np = _rand_planes(ctx.plane_shapes, nframes, rng)
        c = [torch.from_numpy(a).to(dev) for a in c_np]
        d = [torch.empty_like(t) for t in c]
        v = [torch.full_like(t, 424242) for t in c]
        ptrs = lambda ts: [t.data_ptr() for t in ts]  # noqa: E731
        with pytest.raises(daala_b200.TdltError):       # no scan tables yet
            ctx.fwd_batch(ptrs(v), ptrs(c), maps.data_ptr(), nframes, KEY | CO, st)
        ctx.set_scan_tables(tables, ol.SCAN_LEN)
        ctx.fwd_batch(ptrs(d), ptrs(c), maps.data_ptr(), nframes, KEY, st)
        if pix:
            ctx.fwd_batch_pix(ptrs(v), ptrs(px), maps.data_ptr(), nframes, KEY | CO, 1, 1, 0, st)
        else:
            ctx.fwd_batch(ptrs(v), ptrs(c), maps.data_ptr(), nframes, KEY | CO, st)
        torch.cuda.synchronize()
        d_np = [t.cpu().numpy() for t in d]
        v_np = [t.cpu().numpy() for t in v]
        if chk is not None and pic_w < 1000:
            d_ref = chk.fwd_frames(c_np, bsize, decs, 1)
            for p in range(3):
                assert np.array_equal(d_np[p], d_ref[p])
        d0 = []
        for p, dec in enumerate(decs):
            zeroed = np.zeros_like(d_np[p])
            for f in range(nframes):
                if ol.have_ref() and pic_w < 1000:
                    want = ol.ref_coding_order_plane_c(d_np[p][f], bsize[f], dec, True,
                                                       out=np.full_like(d_np[p][f], 424242))
                else:
                    want = ol.port_coding_order_plane(d_np[p][f], bsize[f], dec, tables, True,
                                                      out=np.full_like(d_np[p][f], 424242))
                assert np.array_equal(v_np[p][f], want), "fused store, plane %d frame %d" % (p, f)
                # what the decoder side holds before its inverse transforms: cleared block
                # (od_init_skipped_coeffs, keyframe) + od_coding_order_to_raster
                zeroed[f] = ol.port_coding_order_plane(want, bsize[f], dec, tables, False,
                                                       out=np.zeros_like(d_np[p][f]))
            d0.append(zeroed)
        # ---- inverse, straight from the vectors (the sentinel entries must read as zero) ----
        td0 = [torch.from_numpy(a).to(dev) for a in d0]
        c_a = [torch.empty_like(t) for t in c]
        c_b = [torch.empty_like(t) for t in c]
        ctx.inv_batch(ptrs(c_a), ptrs(td0), maps.data_ptr(), nframes, KEY, st)
        if pix:
            o_a = [torch.empty_like(t) for t in px]
            o_b = [torch.empty_like(t) for t in px]
            for p in range(len(decs)):
                ctx.coeff_to_pix_batch(o_a[p].data_ptr(), c_a[p].data_ptr(), p, nframes, 1, 1, 0, st)
            ctx.inv_batch_pix(ptrs(o_b), ptrs(v), maps.data_ptr(), nframes, KEY | CO, 1, 1, 0, st)
            torch.cuda.synchronize()
            for p in range(len(decs)):
                assert torch.equal(o_a[p], o_b[p]), "fused load to pixels, plane %d" % p
        else:
            ctx.inv_batch(ptrs(c_b), ptrs(v), maps.data_ptr(), nframes, KEY | CO, st)
            torch.cuda.synchronize()
            for p in range(len(decs)):
                assert torch.equal(c_a[p], c_b[p]), "fused load, plane %d" % p
        if chk is not None and pic_w < 1000:
            c_ref = chk.inv_frames(d0, bsize, decs, 1)
            for p in range(3):
                assert np.array_equal(c_a[p].cpu().numpy(), c_ref[p])
        with pytest.raises(daala_b200.TdltError):       # lossless mode codes in raster order
            ctx.fwd_batch(ptrs(v), ptrs(c), maps.data_ptr(), nframes, KEY | CO | daala_b200.OD_TDLT_HAAR, st)
    if chk is not None:
        chk.close()


def test_init_skipped_coeffs_wrapper():
    """od_init_skipped_coeffs with the reference's signature (src/state.c:1340-1357)."""
    import daala_b200
    L = daala_b200.lib()
    rng = np.random.default_rng(4600)
    for n in (4, 8, 16, 32):
        for key in (1, 0):
            w, bo = n + 11, 3 * (n + 11) + 5
            d = rng.integers(-30000, 30000, size=(n + 8, w)).astype(np.int32)
            pred = rng.integers(-30000, 30000, size=(n, n)).astype(np.int32)
            want = d.copy()
            blk = want.reshape(-1)[bo:].reshape(-1)
            if ol.have_ref():
                R = ol.ref()
                R.od_init_skipped_coeffs.restype = None
                R.od_init_skipped_coeffs(ol.ptr(want), ol.ptr(pred), key, bo, n, w)
            else:
                for i in range(n):
                    for j in range(n):
                        if key:
                            if i or j:
                                blk[i * w + j] = 0
                        else:
                            blk[i * w + j] = pred[i, j]
            got = d.copy()
            L.od_init_skipped_coeffs_cuda(got.ctypes.data, pred.ctypes.data, key, bo, n, w)
            assert np.array_equal(got, want), "n %d keyframe %d" % (n, key)


def test_coding_order_needs_tables():
    import daala_b200
    with _ctx(64, 64, 1) as ctx:
        a = [np.zeros((1,) + s, np.int32) for s in ctx.plane_shapes]
        b = [np.zeros((1,) + s, np.int32) for s in ctx.plane_shapes]
        with pytest.raises(daala_b200.TdltError):
            ctx.coding_order_frames(a, b, np.zeros((1, 8, 8), np.uint8), True)


def test_full_size_properties():
    """BASELINE config 4 shape (4K 4:2:0, 10-bit samples in 16-bit words) at full
    size, where the CPU checker is too slow to be run on every frame: (1) the round
    trip is the identity, (2) frame 0 matches the checker, (3) identical inputs at
    two batch positions give identical coefficients (frames are independent)."""
    import daala_b200
    rng = np.random.default_rng(4000)
    pic_w, pic_h, nframes = 3840, 2160, 3
    chk = Checker(pic_w, pic_h)
    with _ctx(pic_w, pic_h, nframes) as ctx:
        bsize = make_maps("quad", 1, ctx.nhsb, ctx.nvsb, rng)
        bsize = np.ascontiguousarray(np.repeat(bsize, nframes, 0))
        bsize[1] = make_maps("deep", 1, ctx.nhsb, ctx.nvsb, rng)[0]
        pix = [rng.integers(0, 1024, size=(nframes,) + s).astype(np.int16) for s in ctx.plane_shapes]
        for p in range(3):
            pix[p][2] = pix[p][0]
        d = [np.zeros((nframes,) + s, np.int32) for s in ctx.plane_shapes]
        # 16-bit reference buffers, lossy: (src - 2048 + rnd) >> shift with mode 2 -> coeff
        ctx.fwd_frames_pix(d, pix, bsize, 1, xstride=2, bitdepth_mode=2, lossless=0)
        port = ol.port()
        for p, dec in enumerate(DECS):
            h, w = ctx.plane_shapes[p]
            c0 = np.zeros((h, w), np.int32)
            src = np.ascontiguousarray(pix[p][0])
            port.oc_buf_to_coeff(2, ol.ptr(c0), w, 0, ol.ptr(src.view(np.uint8)), 2, 2 * w, w, h)
            assert np.array_equal(d[p][0], chk.fwd_plane(c0, bsize[0], dec, 1))
            assert np.array_equal(d[p][2], d[p][0])
        out = [np.zeros_like(a) for a in pix]
        ctx.inv_frames_pix(out, d, bsize, 1, xstride=2, bitdepth_mode=2, lossless=0)
        for p in range(3):
            assert np.array_equal(out[p], pix[p])
    chk.close()


def test_8k_frame_matches_checker():
    """The largest frame the region index allows comfortably (7680x4320 4:2:0: 60 x 68 luma
    regions, 6120 regions per frame): forward against the checker, inverse of arbitrary
    coefficients against the checker, exact round trip -- non-keyframe and keyframe."""
    import daala_b200
    rng = np.random.default_rng(8000)
    pic_w, pic_h = 7680, 4320
    chk = Checker(pic_w, pic_h)
    with _ctx(pic_w, pic_h, 1) as ctx:
        bsize = make_maps("quad", 1, ctx.nhsb, ctx.nvsb, rng)
        c = _rand_planes(ctx.plane_shapes, 1, rng)
        for key in (0, 1):
            flags = daala_b200.OD_TDLT_KEYFRAME if key else 0
            d = [np.zeros_like(a) for a in c]
            ctx.fwd_frames(d, c, bsize, flags)
            d_ref = chk.fwd_frames(c, bsize, DECS, key)
            c2 = [np.zeros_like(a) for a in c]
            ctx.inv_frames(c2, d, bsize, flags)
            for p in range(3):
                assert np.array_equal(d[p], d_ref[p]), "forward plane %d key %d" % (p, key)
                assert np.array_equal(c2[p], c[p]), "round trip plane %d key %d" % (p, key)
        dq = _rand_planes(ctx.plane_shapes, 1, rng, -30000, 30000)
        ctx.inv_frames(c2, dq, bsize, daala_b200.OD_TDLT_KEYFRAME)
        c_ref = chk.inv_frames(dq, bsize, DECS, 1)
        for p in range(3):
            assert np.array_equal(c2[p], c_ref[p]), "inverse plane %d" % p
    chk.close()


def test_device_batches_plan_reuse_and_stress():
    """Device-pointer batch calls on a 16-frame 1080p batch, run three times on the same
    buffers (the inverse kernel's strip exchange reuses its buffer with a new epoch each
    launch): identical coefficients every time, exact round trip, frame 0 equal to the
    checker; OD_TDLT_SAME_MAPS reuses the previous call's plan and is refused when
    there is none to reuse."""
    import torch
    import daala_b200
    pic_w, pic_h, nframes = 1920, 1080, 16
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(1616)
    chk = Checker(pic_w, pic_h)
    with _ctx(pic_w, pic_h, nframes) as ctx:
        bsize = make_maps("quad", nframes, ctx.nhsb, ctx.nvsb, rng)
        c_np = _rand_planes(ctx.plane_shapes, nframes, rng)
        c = [torch.from_numpy(a).to(dev) for a in c_np]
        d = [torch.empty_like(t) for t in c]
        c2 = [torch.empty_like(t) for t in c]
        maps = torch.from_numpy(bsize).to(dev)
        st = torch.cuda.current_stream().cuda_stream
        cp, dp, c2p = ([t.data_ptr() for t in x] for x in (c, d, c2))
        KEY, SAME = daala_b200.OD_TDLT_KEYFRAME, daala_b200.OD_TDLT_SAME_MAPS
        with pytest.raises(daala_b200.TdltError):
            ctx.inv_batch(c2p, dp, maps.data_ptr(), nframes, KEY | SAME, st)   # nothing to reuse
        first = None
        for rep in range(3):
            ctx.fwd_batch(dp, cp, maps.data_ptr(), nframes, KEY, st)
            ctx.inv_batch(c2p, dp, maps.data_ptr(), nframes, KEY | SAME, st)
            torch.cuda.synchronize()
            assert all(torch.equal(a, b) for a, b in zip(c, c2)), "round trip, repetition %d" % rep
            if first is None:
                first = [t.clone() for t in d]
            assert all(torch.equal(a, b) for a, b in zip(first, d))
        with pytest.raises(daala_b200.TdltError):
            ctx.inv_batch(c2p, dp, maps.data_ptr(), nframes - 1, KEY | SAME, st)  # other batch
        # without the flag the inverse builds its own plan
        for t in c2:
            t.zero_()
        ctx.inv_batch(c2p, dp, maps.data_ptr(), nframes, KEY, st)
        torch.cuda.synchronize()
        assert all(torch.equal(a, b) for a, b in zip(c, c2))
        d_ref = chk.fwd_frames([a[:1] for a in c_np], bsize[:1], DECS, 1)
        for p in range(3):
            assert np.array_equal(d[p][0].cpu().numpy(), d_ref[p][0])
        # the stage is out of place (the reference's callers pass ctmp and dtmp): an in-place
        # or overlapping call is refused in both directions instead of corrupting silently
        with pytest.raises(daala_b200.TdltError):
            ctx.fwd_batch(cp, cp, maps.data_ptr(), nframes, KEY, st)
        with pytest.raises(daala_b200.TdltError):
            ctx.inv_batch([x + 64 for x in dp], dp, maps.data_ptr(), nframes, KEY, st)
        # a context call leaves the caller's current device as it found it
        assert torch.cuda.current_device() == 0
    chk.close()

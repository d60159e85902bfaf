#!/usr/bin/env python3
"""Generates tests/golden/tdlt_golden.npz from the UNMODIFIED reference.

Run in the build container (where /root/reference is mounted):

    python tests/golden/make_golden.py

The outputs come from oracle/_ref/libdaala_ref.so, i.e. the reference's own C
(and x86 SIMD) code compiled by oracle/Makefile -- never from the restatement
and never from the CUDA path.  The reference stores no coefficient vectors of
its own, so these are the known answers the oracle and the CUDA
library are both held to wherever the reference itself cannot travel.

Inputs are seeded with the LCG of the reference's self-test
(x = x*1103515245 + 12345, src/dct.c:2385) so the file is reproducible bit for
bit; tests/test_golden.py re-derives nothing but the CIF frame from it.
"""
import ctypes as C
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_libs as ol  # noqa: E402
from oracle_libs import ptr  # noqa: E402

OUT = os.path.join(HERE, "tdlt_golden.npz")


def lcg_ints(seed, n, lo, hi):
    """n integers in [lo, hi) from the self-test LCG (high bits)."""
    s = ol.lcg_stream(seed, n)
    return (lo + ((s >> 8) % (hi - lo))).astype(np.int32)


def cif_frame(t=0, w=352, h=288):
    """smooth gradient + moving 64x64 square + LCG noise +-8."""
    yy, xx = np.mgrid[0:h, 0:w]
    img = 40 + (xx * 120) // w + (yy * 60) // h
    x0, y0 = 40 + 8 * t, 60 + 4 * t
    img[y0:y0 + 64, x0:x0 + 64] += 50
    noise = lcg_ints(1 + t, w * h, -8, 9).reshape(h, w)
    y = np.clip(img + noise, 0, 255).astype(np.uint8)
    u = np.clip(128 + (xx[::2, ::2] - w // 2) // 8 + lcg_ints(101 + t, (w // 2) * (h // 2), -3, 4)
                .reshape(h // 2, w // 2), 0, 255).astype(np.uint8)
    v = np.clip(128 - (yy[::2, ::2] - h // 2) // 8 + lcg_ints(201 + t, (w // 2) * (h // 2), -3, 4)
                .reshape(h // 2, w // 2), 0, 255).astype(np.uint8)
    return y, u, v


def lcg_ints_fast(seed, n, lo, hi):
    """lcg_ints without the Python loop (same values)."""
    s, _ = ol.lcg_stream_fast(seed, n)
    return (lo + ((s.astype(np.int64) >> 8) % (hi - lo))).astype(np.int32)


def hd_frame(t=0, w=1920, h=1080):
    """SURVEY 8d C2/C3 stand-in for "textured 1080p frames", integer-only so that it is
    reproducible bit for bit anywhere: a gradient plus noise at three scales whose amplitude
    changes from one 64x64 area to the next (flat areas next to busy ones, so that the
    reference encoder's RDO picks every block size)."""
    yy, xx = np.mgrid[0:h, 0:w]
    base = 60 + (xx * 90) // w + (yy * 50) // h
    nbx, nby = (w + 63) // 64, (h + 63) // 64
    amps = np.array([0, 1, 3, 8, 20, 40], np.int32)
    amp = amps[lcg_ints_fast(500 + t, nbx * nby, 0, 6)].reshape(nby, nbx)
    amp = np.repeat(np.repeat(amp, 64, 0), 64, 1)[:h, :w]
    fine = lcg_ints_fast(600 + t, w * h, -8, 9).reshape(h, w)
    mid = lcg_ints_fast(700 + t, ((h + 3) // 4) * ((w + 3) // 4), -8, 9).reshape((h + 3) // 4, (w + 3) // 4)
    mid = np.repeat(np.repeat(mid, 4, 0), 4, 1)[:h, :w]
    coarse = lcg_ints_fast(800 + t, ((h + 15) // 16) * ((w + 15) // 16), -8, 9).reshape((h + 15) // 16, (w + 15) // 16)
    coarse = np.repeat(np.repeat(coarse, 16, 0), 16, 1)[:h, :w]
    y = np.clip(base + (amp * (fine + 2 * mid + 3 * coarse)) // 16, 0, 255).astype(np.uint8)
    hw, hh = w // 2, h // 2
    u = np.clip(128 + (xx[::2, ::2] - hw) // 24 + (amp[::2, ::2] * coarse[::2, ::2]) // 32
                + lcg_ints_fast(900 + t, hw * hh, -2, 3).reshape(hh, hw), 0, 255).astype(np.uint8)
    v = np.clip(128 - (yy[::2, ::2] - hh) // 24 + (amp[::2, ::2] * mid[::2, ::2]) // 32
                + lcg_ints_fast(1000 + t, hw * hh, -2, 3).reshape(hh, hw), 0, 255).astype(np.uint8)
    return y, u, v


def planes_to_coeff(yuv, fw, fh):
    """8-bit picture planes -> padded int32 coefficient planes, (pix - 128) << 4
    (od_ref_buf_to_coeff, src/state.c:1221-1230, on edge-extended planes)."""
    return [(pad_plane(p, fh >> dec, fw >> dec).astype(np.int32) - 128) << 4
            for p, dec in zip(yuv, (0, 1, 1))]


def pad_plane(p, H, W):
    """Edge-extend a picture plane to the padded frame (what od_img_edge_ext leaves
    in the coded area, src/state.c:170-217, is not on the TDLT path; any fill does)."""
    out = np.empty((H, W), p.dtype)
    h, w = p.shape
    out[:h, :w] = p
    out[:h, w:] = p[:, -1:]
    out[h:, :] = out[h - 1:h, :]
    return out


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    assert ol.have_ref(), "needs oracle/_ref/libdaala_ref.so (make -C oracle ref)"
    R = ol.ref()
    G = {}

    def f2d(name):
        f = getattr(R, name)
        f.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        return f

    # ---- 1-D and 2-D DCTs ------------------------------------------------------
    for bs in range(4):
        n = 4 << bs
        X = lcg_ints(10 + bs, 48 * n, -4096, 4096).reshape(48, n)
        X[0] = 255 * 16
        X[1] = -256 * 16
        X[2] = 0
        X[2, n - 1] = 255 * 16
        F = np.zeros_like(X)
        I = np.zeros_like(X)
        for k in range(48):
            x = np.ascontiguousarray(X[k])
            y = np.zeros(n, np.int32)
            R.ref_fdct1d(bs, ptr(y), ptr(x), 1)
            F[k] = y
            R.ref_idct1d(bs, ptr(y), 1, ptr(x))
            I[k] = y
        G["dct1d_%d_x" % bs], G["dct1d_%d_fwd" % bs], G["dct1d_%d_inv" % bs] = X, F, I
        nb = 4 if bs < 3 else 2
        B = lcg_ints(20 + bs, nb * n * n, -4096, 4096).reshape(nb, n, n)
        FB = np.zeros_like(B)
        IB = np.zeros_like(B)
        fw, iv = f2d("od_bin_fdct%dx%d" % (n, n)), f2d("od_bin_idct%dx%d" % (n, n))
        for k in range(nb):
            x = np.ascontiguousarray(B[k])
            y = np.zeros((n, n), np.int32)
            fw(ptr(y), n, ptr(x), n)
            FB[k] = y
            iv(ptr(y), n, ptr(x), n)
            IB[k] = y
        G["dct2d_%d_x" % bs], G["dct2d_%d_fwd" % bs], G["dct2d_%d_inv" % bs] = B, FB, IB

    # ---- lapping filters, 4..32 taps -------------------------------------------
    for f in range(4):
        n = 4 << f
        X = lcg_ints(30 + f, 64 * n, -4096, 4096).reshape(64, n)
        X[0] = 676
        X[1] = -676
        X[2] = np.where(np.arange(n) & 1, 676, -676)
        P = np.zeros_like(X)
        Q = np.zeros_like(X)
        pre, post = getattr(R, "od_pre_filter%d" % n), getattr(R, "od_post_filter%d" % n)
        for k in range(64):
            x = np.ascontiguousarray(X[k])
            y = np.zeros(n, np.int32)
            pre(ptr(y), ptr(x))
            P[k] = y
            post(ptr(y), ptr(x))
            Q[k] = y
        G["filt_%d_x" % f], G["filt_%d_pre" % f], G["filt_%d_post" % f] = X, P, Q

    # ---- Haar wavelet blocks, TF merges / splits -----------------------------------
    for ln in (2, 3, 4, 5):
        n = 1 << ln
        x = np.ascontiguousarray(lcg_ints(40 + ln, n * n, -4096, 4096).reshape(n, n))
        a = np.zeros((n, n), np.int32)
        b = np.zeros((n, n), np.int32)
        R.od_haar(ptr(a), n, ptr(x), n, ln)
        R.od_haar_inv(ptr(b), n, ptr(x), n, ln)
        G["haar_%d_x" % ln], G["haar_%d_fwd" % ln], G["haar_%d_inv" % ln] = x, a, b
    for n in (4, 8, 16):
        m = 2 * n
        src = np.ascontiguousarray(lcg_ints(50 + n, m * m, -30000, 30000).reshape(m, m))
        G["tf_%d_src" % n] = src
        for name, args in (("tf_up_hv", (n,)), ("tf_down_hv", (m,)), ("tf_up_hv_lp", (n, n, m)),
                           ("tf_up_h_lp", (n, m)), ("tf_up_v_lp", (n, m)), ("tf_filter_2d", (m,)),
                           ("tf_filter_inv_2d", (m,))):
            out = np.zeros((m, m), np.int32)
            getattr(R, "od_" + name)(ptr(out), m, ptr(src), m, *args)
            G["tf_%d_%s" % (n, name)] = out

    for curr, dest, filt in ((1, 0, 0), (2, 0, 1), (3, 1, 0), (3, 0, 1), (2, 2, 0), (3, 2, 1)):
        n = 4 << curr
        src = np.ascontiguousarray(lcg_ints(55 + curr * 8 + dest * 2 + filt, n * n, -30000, 30000).reshape(n, n))
        out = np.zeros((n, n), np.int32)
        R.od_convert_block_down(ptr(out), n, ptr(src), n, curr, dest, filt)
        G["cbd_%d_%d_%d_src" % (curr, dest, filt)] = src
        G["cbd_%d_%d_%d_out" % (curr, dest, filt)] = out

    # ---- whole planes: 100x60 picture -> 128x64 frame, 4:2:0 ------------------------
    pic_w, pic_h = 100, 60
    with ol.RefCtx(pic_w, pic_h) as rc:
        st = np.random.RandomState(7)     # legacy generator: stable stream across numpy versions
        bsize = np.full((rc.nvsb * 4, rc.nhsb * 4), 3, np.uint8)
        pick = st.randint(0, 100, size=(rc.nvsb, rc.nhsb, 21))
        for sy in range(rc.nvsb):
            for sx in range(rc.nhsb):
                p = pick[sy, sx]
                if p[0] >= 80:
                    continue
                for q in range(4):
                    y0, x0 = sy * 4 + (q >> 1) * 2, sx * 4 + (q & 1) * 2
                    if p[1 + q] >= 60:
                        bsize[y0:y0 + 2, x0:x0 + 2] = 2
                        continue
                    for r in range(4):
                        bsize[y0 + (r >> 1), x0 + (r & 1)] = 0 if p[5 + 4 * q + r] < 50 else 1
        G["plane_bsize"] = bsize
        G["plane_pic"] = np.array([pic_w, pic_h], np.int32)
        for pli, dec in enumerate((0, 1, 1)):
            H, W = rc.fh >> dec, rc.fw >> dec
            c = np.ascontiguousarray(lcg_ints(60 + pli, H * W, -2048, 2048).reshape(H, W))
            dq = np.ascontiguousarray(lcg_ints(70 + pli, H * W, -30000, 30000).reshape(H, W))
            G["plane_%d_c" % pli] = c
            G["plane_%d_dq" % pli] = dq
            for key in (0, 1):
                d, _ = rc.fwd_plane(c, bsize, dec, key, impl=0)
                G["plane_%d_d_key%d" % (pli, key)] = d
                G["plane_%d_cq_key%d" % (pli, key)] = rc.inv_plane(dq, bsize, dec, key, impl=0)
            hb = np.full_like(bsize, 3)
            cl = np.ascontiguousarray(lcg_ints(80 + pli, H * W, -128, 128).reshape(H, W))
            G["plane_%d_cl" % pli] = cl
            G["plane_%d_d_haar" % pli] = rc.fwd_plane(cl, hb, dec, 1, use_haar=1, impl=0)[0]
        # chroma-from-luma predictor plane of the decoded luma coefficients above
        # (od_resample_luma_coeffs, src/intra.c:72, called per chroma leaf block)
        G["plane_cfl"] = ol.ref_cfl_pred_plane(G["plane_0_dq"], bsize)

    # ---- C1: CIF keyframe, block sizes chosen by the reference encoder's RDO ------
    with ol.RefCtx(352, 288) as rc:
        y, u, v = cif_frame(0)
        bs = rc.encode_keyframe_bsize(y, u, v, quality=20, complexity=7)
        G["cif_bsize"] = bs
        hashes = []
        for pli, (p, dec) in enumerate(((y, 0), (u, 1), (v, 1))):
            H, W = rc.fh >> dec, rc.fw >> dec
            c = ((pad_plane(p, H, W).astype(np.int32) - 128) << 4)
            d, _ = rc.fwd_plane(c, bs, dec, 1, impl=0)
            hashes.append(sha(d))
            if pli == 0:
                G["cif_d0_rows"] = d[96:104].copy()      # a readable slice next to the hashes
        G["cif_d_sha256"] = np.array(hashes)

    # ---- pixel <-> coefficient conversions -----------------------------------------
    w, h = 48, 20
    for mode, lossless, xs in ((1, 0, 1), (1, 1, 1), (2, 0, 2), (2, 1, 2), (3, 1, 2), (1, 0, 2)):
        ystride = (w + 16) * xs
        if xs == 1:
            src = lcg_ints(90 + mode, h * ystride, 0, 256).astype(np.uint8).reshape(h, ystride)
        else:
            src = lcg_ints(90 + mode, h * ystride // 2, 0, 4096).astype(np.int16).view(np.uint8).reshape(h, ystride)
        src = np.ascontiguousarray(src)
        co = np.ascontiguousarray(lcg_ints(95 + mode, h * w, -5000, 5000).reshape(h, w))
        a = np.zeros((h, w), np.int32)
        R.ref_buf_to_coeff(mode, ptr(a), w, lossless, ptr(src), xs, ystride, w, h)
        pb = np.zeros((h, ystride), np.uint8)
        R.ref_coeff_to_buf(mode, ptr(pb), xs, ystride, ptr(co), w, lossless, w, h)
        tag = "pix_%d_%d_%d" % (mode, lossless, xs)
        G[tag + "_src"], G[tag + "_coeff"], G[tag + "_co"], G[tag + "_buf"] = src, a, co, pb

    # ---- the reference's scan tables (OD_ZIGZAG4/8/16/32, src/zigzag*.c): inputs of the
    # coding-order calls, which take them from the caller; stored so that the GPU box, where
    # the reference may be absent, can still drive those calls with the real tables
    for k, t in enumerate(ol.scan_tables()):
        G["scan_table_%d" % (4 << k)] = t
    rng_d = np.ascontiguousarray(lcg_ints(99, 32 * 32, -30000, 30000).reshape(32, 32))
    for bs in range(4):
        n = 4 << bs
        v = np.zeros(n * n, np.int32)
        R.od_raster_to_coding_order.restype = None
        R.od_raster_to_coding_order(ptr(v), n, ptr(rng_d), 32)
        G["scan_block_%d" % n] = v
    G["scan_block_src"] = rng_d

    np.savez_compressed(OUT, **G)
    print("wrote %s: %d arrays, %d bytes" % (OUT, len(G), os.path.getsize(OUT)))


OUT_CONFIGS = os.path.join(HERE, "tdlt_golden_configs.npz")


def main_configs():
    """BASELINE configs at their stated size (second file, so that the first keeps its bytes):
    config 1 -- ALL ten CIF frames, each a keyframe (keyframe_rate = 1) with the block sizes the
    reference encoder's RDO chose for it; config 3(i) -- a 1080p frame with its RDO map.
    Stored: the maps and the SHA-256 of every coefficient plane the reference's forward stage
    produces from them."""
    assert ol.have_ref(), "needs oracle/_ref/libdaala_ref.so (make -C oracle ref)"
    G = {}
    maps, hashes = [], []
    for t in range(10):
        with ol.RefCtx(352, 288) as rc:
            yuv = cif_frame(t)
            bs = rc.encode_keyframe_bsize(*yuv, quality=20, complexity=7)
            maps.append(bs)
            row = []
            for pli, (c, dec) in enumerate(zip(planes_to_coeff(yuv, rc.fw, rc.fh), (0, 1, 1))):
                row.append(sha(rc.fwd_plane(c, bs, dec, 1, impl=0)[0]))
            hashes.append(row)
    G["cif10_bsize"] = np.stack(maps)
    G["cif10_d_sha256"] = np.array(hashes)
    with ol.RefCtx(1920, 1080) as rc:
        yuv = hd_frame(0)
        bs = rc.encode_keyframe_bsize(*yuv, quality=20, complexity=7)
        G["hd_bsize"] = bs
        G["hd_d_sha256"] = np.array([sha(rc.fwd_plane(c, bs, dec, 1, impl=0)[0])
                                     for c, dec in zip(planes_to_coeff(yuv, rc.fw, rc.fh), (0, 1, 1))])
        c8 = planes_to_coeff(yuv, rc.fw, rc.fh)
        u8 = np.full_like(bs, 1)
        G["hd_d_sha256_8x8"] = np.array([sha(rc.fwd_plane(c, u8, dec, 1, impl=0)[0])
                                         for c, dec in zip(c8, (0, 1, 1))])
    np.savez_compressed(OUT_CONFIGS, **G)
    print("wrote %s: %d arrays, %d bytes" % (OUT_CONFIGS, len(G), os.path.getsize(OUT_CONFIGS)))
    print("hd map histogram:", np.bincount(G["hd_bsize"].ravel(), minlength=4))
    print("cif maps differ:", len({m.tobytes() for m in maps}))


if __name__ == "__main__":
    if "--configs" in sys.argv:
        main_configs()
    else:
        main()
        main_configs()

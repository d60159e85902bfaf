#!/usr/bin/env python3
"""Randomised parity sweep (test infrastructure, like tests/): random picture sizes, map
kinds, plane layouts, frame counts and flags through the whole-frame C-ABI calls, against the
CPU checker (the compiled reference when it is built, else the restatement).  Bit-exact or
it stops.  One summary line on stdout.

  python tools/fuzz_parity.py [--seconds 120] [--seed 1]
"""
import argparse
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=120.0)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--what", default="tdlt", choices=["tdlt", "dering", "pix", "coding"],
                    help="tdlt: forward / inverse; dering: the deringing filter over whole frames")
    ap.add_argument("--coeff", type=int, default=8000,
                    help="range of the arbitrary coefficients fed to the inverse (inside the domain in "
                         "which the reference's own int32 products do not overflow: DESIGN.md section 5)")
    ap.add_argument("--max-frames", type=int, default=3,
                    help="tdlt: frames per case are drawn from 1 .. this (the inverse kernel picks one of "
                         "three region orders by batch size: < 16, < 48, >= 48)")
    ap.add_argument("--start", type=int, default=0, help="draw the first N cases without running them")
    ap.add_argument("--dump", default="", help="on a mismatch, write the case to this .npz")
    args = ap.parse_args()
    import daala_b200
    import oracle_libs as ol
    from checker import Checker, make_maps
    rng = np.random.default_rng(args.seed)
    if args.what == "dering":
        return fuzz_dering(args, rng)
    if args.what == "pix":
        return fuzz_pix(args, rng)
    if args.what == "coding":
        return fuzz_coding(args, rng)
    kinds = ["quad", "deep", "uniform0", "uniform1", "uniform2", "uniform3"]
    layouts = [(0, 1, 1), (0, 1, 1), (0, 1, 1), (0,), (0, 0, 0)]
    t0 = time.time()
    cases = 0
    samples = 0
    t_run = None
    while True:
        if cases >= args.start:
            if t_run is None:
                t_run = time.time()
            elif time.time() - t_run > args.seconds:
                break
        pic_w = int(rng.integers(8, 700))
        pic_h = int(rng.integers(8, 420))
        decs = layouts[int(rng.integers(len(layouts)))]
        kind = kinds[int(rng.integers(len(kinds)))]
        nframes = int(rng.integers(1, args.max_frames + 1))
        key = int(rng.integers(2))
        haar = int(rng.integers(8) == 0)
        if haar:
            kind = "uniform3"       # lossless mode codes whole superblocks (src/encode.c:3090)
        flags = (daala_b200.OD_TDLT_KEYFRAME if key else 0) | (daala_b200.OD_TDLT_HAAR if haar else 0)
        desc = "%dx%d decs %s %s frames %d key %d haar %d" % (pic_w, pic_h, decs, kind, nframes, key, haar)
        if cases < args.start:
            # keep the random stream aligned: the same draws, no GPU or checker work
            fw, fh = (pic_w + 63) & ~63, (pic_h + 63) & ~63
            make_maps(kind, nframes, fw // 32, fh // 32, rng)
            for lo_, hi_ in ((-2048, 2048), (-args.coeff, args.coeff)):
                for dec in decs:
                    rng.integers(lo_, hi_, size=(nframes, fh >> dec, fw >> dec))
            cases += 1
            continue
        if cases == args.start and args.start:
            print("first case run:", desc, flush=True)
        chk = Checker(pic_w, pic_h)
        with daala_b200.TdltContext(pic_w, pic_h, decs=decs, device=0, max_frames=nframes) as ctx:
            bsize = make_maps(kind, nframes, ctx.nhsb, ctx.nvsb, rng)
            c = [rng.integers(-2048, 2048, size=(nframes,) + s).astype(np.int32) for s in ctx.plane_shapes]
            d = [np.zeros_like(a) for a in c]
            ctx.fwd_frames(d, c, bsize, flags)
            d_ref = chk.fwd_frames(c, bsize, decs, key, haar=haar)
            for p in range(len(decs)):
                assert np.array_equal(d[p], d_ref[p]), "forward plane %d: %s" % (p, desc)
            c2 = [np.zeros_like(a) for a in c]
            ctx.inv_frames(c2, d, bsize, flags)
            for p in range(len(decs)):
                assert np.array_equal(c2[p], c[p]), "round trip plane %d: %s" % (p, desc)
            dq = [rng.integers(-args.coeff, args.coeff, size=a.shape).astype(np.int32) for a in c]
            ctx.inv_frames(c2, dq, bsize, flags)
            c_ref = chk.inv_frames(dq, bsize, decs, key, haar=haar)
            for p in range(len(decs)):
                if not np.array_equal(c2[p], c_ref[p]):
                    bad = np.argwhere(c2[p] != c_ref[p])
                    print("MISMATCH inverse plane %d: %s: case %d, %d samples, frames %s rows %s cols %s" % (
                        p, desc, cases, len(bad), sorted(set(bad[:, 0].tolist())),
                        sorted(set(bad[:, 1].tolist()))[:24], sorted(set(bad[:, 2].tolist()))[:24]))
                    again = [np.zeros_like(a) for a in c]
                    ctx.inv_frames(again, dq, bsize, flags)
                    print("  the same call again: %s" % ("equal to the checker" if np.array_equal(again[p], c_ref[p])
                                                        else "differs again"))
                    if args.dump:
                        np.savez_compressed(args.dump, pic=np.array([pic_w, pic_h]), decs=np.array(decs),
                                            key=key, bsize=bsize, **{"dq%d" % i: a for i, a in enumerate(dq)},
                                            **{"gpu%d" % i: a for i, a in enumerate(c2)},
                                            **{"ref%d" % i: a for i, a in enumerate(c_ref)})
                    raise SystemExit(1)
            samples += sum(a.size for a in c)
        chk.close()
        cases += 1
    print("fuzz_parity: %d random cases, %.1f Msamples, all bit-exact against %s (seed %d, %.0f s)"
          % (cases, samples / 1e6, "the compiled reference" if ol.have_ref() else "the restatement",
             args.seed, time.time() - t0))


def fuzz_pix(args, rng):
    """The pixel-facing whole-frame calls (conversion fused into the transform kernels): random
    sizes, 8- and 16-bit reference buffers, dense or padded rows, quadtree maps; forward
    against od_ref_buf_to_coeff + the checker's forward, inverse against the checker's inverse
    + od_coeff_to_ref_buf, and nothing outside the picture window may change."""
    import daala_b200
    import oracle_libs as ol
    from checker import Checker, make_maps
    port = ol.port()
    decs = (0, 1, 1)
    t0 = time.time()
    cases = 0
    samples = 0
    while time.time() - t0 < args.seconds:
        pic_w = int(rng.integers(8, 700))
        pic_h = int(rng.integers(8, 420))
        xstride, mode, lossless = [(1, 1, 0), (2, 2, 0), (2, 3, 0), (1, 2, 0), (2, 1, 0)][int(rng.integers(5))]
        nframes = int(rng.integers(1, 4))
        pad = int(rng.integers(3)) * 16
        kind = ["quad", "deep", "uniform1"][int(rng.integers(3))]
        desc = "%dx%d xstride %d mode %d frames %d pad %d %s" % (pic_w, pic_h, xstride, mode, nframes, pad, kind)
        chk = Checker(pic_w, pic_h)
        with daala_b200.TdltContext(pic_w, pic_h, decs=decs, device=0, max_frames=nframes) as ctx:
            bsize = make_maps(kind, nframes, ctx.nhsb, ctx.nvsb, rng)
            pix, views = [], []
            for (h, w) in ctx.plane_shapes:
                if xstride == 1:
                    buf = rng.integers(0, 256, size=(nframes, h + 2 * pad, w + 2 * pad)).astype(np.uint8)
                else:
                    buf = rng.integers(0, 4096, size=(nframes, h + 2 * pad, w + 2 * pad)).astype(np.int16)
                pix.append(buf)
                views.append(buf[:, pad:pad + h, pad:pad + w])
            d = [np.zeros((nframes,) + sh, np.int32) for sh in ctx.plane_shapes]
            ctx.fwd_frames_pix(d, views, bsize, 1, xstride=xstride, bitdepth_mode=mode, lossless=lossless)
            c = []
            for v in views:
                cc = np.zeros(v.shape, np.int32)
                for f in range(nframes):
                    src = np.ascontiguousarray(v[f])
                    port.oc_buf_to_coeff(mode, ol.ptr(cc[f]), v.shape[2], lossless, ol.ptr(src.view(np.uint8)),
                                         xstride, src.strides[0], v.shape[2], v.shape[1])
                c.append(cc)
            d_ref = chk.fwd_frames(c, bsize, decs, 1)
            for p in range(3):
                assert np.array_equal(d[p], d_ref[p]), "forward plane %d: %s" % (p, desc)
            outs = [np.full_like(b, 7) for b in pix]
            oviews = [b[:, pad:pad + sh[0], pad:pad + sh[1]] for b, sh in zip(outs, ctx.plane_shapes)]
            dq = [(a // 3) * 3 for a in d]
            ctx.inv_frames_pix(oviews, dq, bsize, 1, xstride=xstride, bitdepth_mode=mode, lossless=lossless)
            c_ref = chk.inv_frames(dq, bsize, decs, 1)
            for p in range(3):
                h, w = ctx.plane_shapes[p]
                exp = np.zeros((nframes, h, w), pix[p].dtype)
                for f in range(nframes):
                    tmp = np.zeros((h, w * xstride), np.uint8)
                    port.oc_coeff_to_buf(mode, ol.ptr(tmp), xstride, w * xstride, ol.ptr(c_ref[p][f]), w,
                                         lossless, w, h)
                    exp[f] = tmp.view(pix[p].dtype).reshape(h, w)
                assert np.array_equal(oviews[p], exp), "inverse plane %d: %s" % (p, desc)
                border = outs[p].copy()
                border[:, pad:pad + h, pad:pad + w] = 7
                assert (border == 7).all(), "outside the window, plane %d: %s" % (p, desc)
            samples += sum(a.size for a in d)
        chk.close()
        cases += 1
    print("fuzz_parity pix: %d random cases, %.1f Msamples, all bit-exact (seed %d, %.0f s)"
          % (cases, samples / 1e6, args.seed, time.time() - t0))


def fuzz_coding(args, rng):
    """OD_TDLT_CODING_ORDER through the device-pointer batch calls: the forward kernel's
    block-linear coding-order vectors against forward + od_raster_to_coding_order, the inverse
    kernel's load of them against od_coding_order_to_raster (into a cleared block) + inverse."""
    import torch
    import daala_b200
    import oracle_libs as ol
    from checker import Checker, make_maps
    dev = torch.device("cuda", 0)
    tables = ol.scan_tables()
    KEY, CO = daala_b200.OD_TDLT_KEYFRAME, daala_b200.OD_TDLT_CODING_ORDER
    t0 = time.time()
    cases = 0
    samples = 0
    ptrs = lambda ts: [t.data_ptr() for t in ts]  # noqa: E731
    while time.time() - t0 < args.seconds:
        pic_w = int(rng.integers(8, 600))
        pic_h = int(rng.integers(8, 360))
        decs = (0, 1, 1) if rng.integers(4) else (0, 0, 0)
        kind = ["quad", "deep", "uniform0", "uniform1", "uniform2", "uniform3"][int(rng.integers(6))]
        nframes = int(rng.integers(1, 4))
        desc = "%dx%d decs %s %s frames %d" % (pic_w, pic_h, decs, kind, nframes)
        chk = Checker(pic_w, pic_h)
        with daala_b200.TdltContext(pic_w, pic_h, decs=decs, device=0, max_frames=0) as ctx:
            ctx.set_scan_tables(tables, ol.SCAN_LEN)
            bsize = make_maps(kind, nframes, ctx.nhsb, ctx.nvsb, rng)
            maps = torch.from_numpy(bsize).to(dev)
            st = torch.cuda.current_stream().cuda_stream
            c_np = [rng.integers(-2048, 2048, size=(nframes,) + sh).astype(np.int32) for sh in ctx.plane_shapes]
            c = [torch.from_numpy(a).to(dev) for a in c_np]
            v = [torch.full_like(t, 424242) for t in c]
            ctx.fwd_batch(ptrs(v), ptrs(c), maps.data_ptr(), nframes, KEY | CO, st)
            torch.cuda.synchronize()
            d_ref = chk.fwd_frames(c_np, bsize, decs, 1)
            d0 = []
            for p, dec in enumerate(decs):
                zeroed = np.zeros_like(d_ref[p])
                vp = v[p].cpu().numpy()
                for f in range(nframes):
                    want = ol.port_coding_order_plane(d_ref[p][f], bsize[f], dec, tables, True,
                                                      out=np.full_like(d_ref[p][f], 424242))
                    assert np.array_equal(vp[f], want), "fused store plane %d frame %d: %s" % (p, f, desc)
                    zeroed[f] = ol.port_coding_order_plane(want, bsize[f], dec, tables, False,
                                                           out=np.zeros_like(d_ref[p][f]))
                d0.append(zeroed)
            c_b = [torch.empty_like(t) for t in c]
            ctx.inv_batch(ptrs(c_b), ptrs(v), maps.data_ptr(), nframes, KEY | CO, st)
            torch.cuda.synchronize()
            c_ref = chk.inv_frames(d0, bsize, decs, 1)
            for p in range(3):
                assert np.array_equal(c_b[p].cpu().numpy(), c_ref[p]), "fused load plane %d: %s" % (p, desc)
            samples += sum(a.size for a in c_np)
        chk.close()
        cases += 1
    print("fuzz_parity coding: %d random cases, %.1f Msamples, all bit-exact (seed %d, %.0f s)"
          % (cases, samples / 1e6, args.seed, time.time() - t0))


def fuzz_dering(args, rng):
    """Random sizes, 4:2:0 or 4:4:4, quantizers and skip densities through od_tdlt_dering_frames
    against the reference's od_dering (the loop of src/decode.c:1044-1122)."""
    import daala_b200
    import oracle_libs as ol
    from checker import Checker
    t0 = time.time()
    cases = 0
    samples = 0
    while time.time() - t0 < args.seconds:
        pic_w = int(rng.integers(8, 500))
        pic_h = int(rng.integers(8, 300))
        decs = (0, 1, 1) if rng.integers(4) else (0, 0, 0)
        nframes = int(rng.integers(1, 4))
        q = int(rng.integers(1, 600))
        skip_prob = float(rng.random())
        qs = (q, max(1, 3 * q // 4), max(1, q // 2))
        desc = "%dx%d decs %s frames %d q %d skip %.2f" % (pic_w, pic_h, decs, nframes, q, skip_prob)
        chk = Checker(pic_w, pic_h)
        with daala_b200.TdltContext(pic_w, pic_h, decs=decs, device=0, max_frames=nframes) as ctx:
            fr = [ol.dering_test_frame(rng, ctx.nhsb, ctx.nvsb, decs, skip_prob=skip_prob) for _ in range(nframes)]
            planes = [np.stack([f[0][p] for f in fr]) for p in range(3)]
            flags = np.stack([f[1] for f in fr])
            bskip = [np.stack([f[2][p] for f in fr]) for p in range(3)]
            out = [np.full_like(p, 77777) for p in planes]
            ctx.dering_frames(out, planes, flags, bskip, qs)
            want = chk.dering_frames(planes, decs, flags, bskip, qs)
            for p in range(3):
                if not np.array_equal(out[p], want[p]):
                    bad = np.argwhere(out[p] != want[p])
                    print("MISMATCH dering plane %d: %s: %d samples, first %s" % (p, desc, len(bad), bad[0]))
                    raise SystemExit(1)
            samples += sum(a.size for a in planes)
        chk.close()
        cases += 1
    print("fuzz_parity dering: %d random cases, %.1f Msamples, all bit-exact against %s (seed %d, %.0f s)"
          % (cases, samples / 1e6, "the compiled reference" if ol.have_ref() else "the restatement",
             args.seed, time.time() - t0))


if __name__ == "__main__":
    main()

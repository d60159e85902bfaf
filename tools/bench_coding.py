#!/usr/bin/env python3
"""Raster <-> coding order over whole frames: kernel time against the HBM roofline, the CPU
restatement timed beside it on a bounded sample.  One JSON line on stdout.

  python tools/bench_coding.py [--workload 4k] [--frames 64] [--steps 20]

Algorithmic bytes: 8 per ordered coefficient (read int32, write int32); a 32x32 block orders
512 of its 1024 coefficients, so the figure depends on the map.  The scan tables come from the
compiled reference when it is there, else from the golden fixture taken from it.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="4k", choices=sorted(bench.WORKLOADS))
    ap.add_argument("--frames", type=int, default=0)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    import torch
    import daala_b200
    import oracle_libs as ol
    pic_w, pic_h, nframes, kind = bench.WORKLOADS[args.workload]
    nframes = args.frames or nframes
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    ctx = daala_b200.TdltContext(pic_w, pic_h, decs=bench.DECS, device=0, max_frames=0)
    tables = ol.scan_tables()
    ctx.set_scan_tables(tables, ol.SCAN_LEN)
    maps_np = bench.make_maps_np(kind, nframes, ctx.nhsb, ctx.nvsb, 79)
    maps = torch.from_numpy(maps_np).to(dev)
    g = torch.Generator(device=dev)
    g.manual_seed(5)
    d = [torch.randint(-30000, 30000, (nframes,) + s, dtype=torch.int32, device=dev, generator=g)
         for s in ctx.plane_shapes]
    v = [torch.zeros_like(t) for t in d]
    back = [torch.zeros_like(t) for t in d]
    st = torch.cuda.current_stream().cuda_stream
    dp, vp, bp = ([t.data_ptr() for t in x] for x in (d, v, back))
    res = {}
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for name, fn in (("to_coding", lambda: ctx.coding_order_batch(vp, dp, maps.data_ptr(), nframes, True, st)),
                     ("to_raster", lambda: ctx.coding_order_batch(vp, bp, maps.data_ptr(), nframes, False, st))):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(args.steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        res[name] = e0.elapsed_time(e1) / args.steps
    nsamples = sum(t.numel() for t in d)
    # luma 32x32 blocks order half of their coefficients (4:2:0 chroma blocks are at most 16x16)
    frac32 = float((maps_np[:, ::4, ::4] == 3).mean())
    ordered = nsamples - d[0].numel() * frac32 / 2
    alg_bytes = 8 * ordered
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    peak = float(peaks.get("hbm_gbs", 0)) or 6500.0
    ms = res["to_coding"] + res["to_raster"]
    line = {"metric": "raster <-> coding order, whole frames", "workload": "%dx%d 4:2:0, %s maps, %d frames"
            % (pic_w, pic_h, kind, nframes), "ms_per_step": round(ms, 4),
            "to_coding_ms": round(res["to_coding"], 4), "to_raster_ms": round(res["to_raster"], 4),
            "value": round(pic_w * pic_h * nframes / (ms * 1e-3) / 1e6, 1), "unit": "Mpixel/s",
            "gpu_launches": 2 * args.steps,
            "roofline": {"bound": "hbm", "achieved": round(2 * alg_bytes / (ms * 1e-3) / 1e9, 1), "peak": peak,
                         "unit": "GB/s", "frac": round(2 * alg_bytes / (ms * 1e-3) / 1e9 / peak, 3),
                         "to_coding_frac": round(alg_bytes / (res["to_coding"] * 1e-3) / 1e9 / peak, 3),
                         "to_raster_frac": round(alg_bytes / (res["to_raster"] * 1e-3) / 1e9 / peak, 3),
                         "traffic": None, "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback"},
            "algorithmic_bytes_per_direction": int(alg_bytes)}
    # parity of what was just timed: frame 0 against the restatement, both directions
    for p, dec in enumerate(bench.DECS):
        want = ol.port_coding_order_plane(d[p][0].cpu().numpy(), maps_np[0], dec, tables, True)
        assert np.array_equal(v[p][0].cpu().numpy(), want), "plane %d differs from the oracle" % p
        covered = ol.port_coding_order_plane(want, maps_np[0], dec, tables, False,
                                             out=np.full(want.shape, 1 << 30, np.int32)) != (1 << 30)
        assert np.array_equal(back[p][0].cpu().numpy()[covered], d[p][0].cpu().numpy()[covered])
    line["parity"] = "frame 0 equal to the oracle, round trip exact on the ordered coefficients"
    if not args.no_cpu and ol.have_ref():
        # od_raster_to_coding_order itself, block by block, against the GPU's planes (before the
        # fused runs below reuse the buffers)
        for p, dec in enumerate(bench.DECS):
            want = ol.ref_coding_order_plane_c(d[p][0].cpu().numpy(), maps_np[0], dec, True)
            assert np.array_equal(v[p][0].cpu().numpy(), want), "plane %d differs from the reference" % p
        line["parity"] += "; frame 0 equal to the compiled reference"
    # ---- the permutation fused into the transform kernels (OD_TDLT_CODING_ORDER) -------------
    # forward: coefficients leave as vectors; inverse: they arrive as vectors.  Timed against
    # the same kernels writing / reading the raster plane (plan kernel included in the forward).
    KEY, CO = daala_b200.OD_TDLT_KEYFRAME, daala_b200.OD_TDLT_CODING_ORDER
    SAME = daala_b200.OD_TDLT_SAME_MAPS
    c = [(t >> 4) for t in d]                      # spatial-domain stand-in, 12-bit range
    fused = {}
    for name, fn in (("fwd_raster", lambda: ctx.fwd_batch(bp, [t.data_ptr() for t in c], maps.data_ptr(), nframes, KEY, st)),
                     ("fwd_coding", lambda: ctx.fwd_batch(vp, [t.data_ptr() for t in c], maps.data_ptr(), nframes, KEY | CO, st)),
                     ("inv_raster", lambda: ctx.inv_batch(dp, bp, maps.data_ptr(), nframes, KEY | SAME, st)),
                     ("inv_coding", lambda: ctx.inv_batch(dp, vp, maps.data_ptr(), nframes, KEY | CO | SAME, st))):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(args.steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        fused[name + "_ms"] = round(e0.elapsed_time(e1) / args.steps, 4)
    fused["fwd_ratio"] = round(fused["fwd_coding_ms"] / fused["fwd_raster_ms"], 3)
    fused["inv_ratio"] = round(fused["inv_coding_ms"] / fused["inv_raster_ms"], 3)
    fused["note"] = ("OD_TDLT_CODING_ORDER: the forward kernel stores, the inverse kernel loads the "
                     "block-linear vectors; against the standalone pass above the permutation costs "
                     "(coding - raster) ms instead of to_coding_ms / to_raster_ms")
    line["fused"] = fused
    if not args.no_cpu:
        a = [t[0].cpu().numpy() for t in d]
        use_ref = ol.have_ref()
        t0 = time.perf_counter()
        reps = 0
        while time.perf_counter() - t0 < 5.0:
            for p, dec in enumerate(bench.DECS):
                if use_ref:
                    w = ol.ref_coding_order_plane_c(a[p], maps_np[0], dec, True)
                    ol.ref_coding_order_plane_c(w, maps_np[0], dec, False)
                else:
                    w = ol.port_coding_order_plane(a[p], maps_np[0], dec, tables, True)
                    ol.port_coding_order_plane(w, maps_np[0], dec, tables, False)
            reps += 1
        dt = time.perf_counter() - t0
        line["cpu_baseline"] = {"value": round(pic_w * pic_h * reps / dt / 1e6, 1), "unit": "Mpixel/s",
                                "cores": 1, "kind": "reference" if use_ref else "port",
                                "sample": "1 frame x %d passes, both directions" % reps}
    print(json.dumps(line))
    ctx.close()


if __name__ == "__main__":
    main()

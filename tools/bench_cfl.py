#!/usr/bin/env python3
"""Chroma-from-luma predictor planes: kernel time against the HBM roofline, with the CPU
restatement timed beside it on a bounded sample.  One JSON line on stdout.

  python tools/bench_cfl.py [--workload 4k] [--frames 64] [--steps 20] [--warmup 3]

Algorithmic bytes: 8 per sample of `l` (4 read from the luma plane, 4 written) plus one map
byte per 16 samples.  Inputs (2.1 GB of luma coefficients at 4K x 64) are far larger than L2.
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
import bench  # noqa: E402  (workload table, map generator, measured peaks)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="4k", choices=sorted(bench.WORKLOADS))
    ap.add_argument("--frames", type=int, default=0)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--maps", default="", help="override the map kind (quad, uniform0..3)")
    args = ap.parse_args()
    import torch
    import daala_b200
    pic_w, pic_h, nframes, kind = bench.WORKLOADS[args.workload]
    nframes = args.frames or nframes
    kind = args.maps or kind
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    ctx = daala_b200.TdltContext(pic_w, pic_h, decs=bench.DECS, device=0, max_frames=0)
    fw, fh = ctx.frame_width, ctx.frame_height
    maps_np = bench.make_maps_np(kind, nframes, ctx.nhsb, ctx.nvsb, 77)
    maps = torch.from_numpy(maps_np).to(dev)
    g = torch.Generator(device=dev)
    g.manual_seed(99)
    d_luma = torch.randint(-30000, 30000, (nframes, fh, fw), dtype=torch.int32, device=dev, generator=g)
    l = torch.zeros((nframes, fh // 2, fw // 2), dtype=torch.int32, device=dev)
    st = torch.cuda.current_stream().cuda_stream

    def step():
        ctx.cfl_pred_batch(l.data_ptr(), d_luma.data_ptr(), maps.data_ptr(), nframes, st)

    for _ in range(max(args.warmup, 3)):
        step()
    torch.cuda.synchronize()
    n0 = daala_b200.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    launches = daala_b200.launch_count() - n0
    ms = e0.elapsed_time(e1) / args.steps
    nl = l.numel()
    alg_bytes = 8 * nl + maps.numel()
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    peak = float(peaks.get("hbm_gbs", 0)) or 6500.0
    achieved = alg_bytes / (ms * 1e-3) / 1e9
    line = {"metric": "chroma-from-luma predictor planes", "workload": "%dx%d 4:2:0, %s maps, %d frames"
            % (pic_w, pic_h, kind, nframes), "ms_per_step": round(ms, 4),
            "value": round(pic_w * pic_h * nframes / (ms * 1e-3) / 1e6, 1), "unit": "Mpixel/s",
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                         "frac": round(achieved / peak, 3), "traffic": None,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback"},
            "algorithmic_bytes_per_step": int(alg_bytes)}
    # parity of what was just timed, frame 0 and the last frame, against the restatement
    import oracle_libs as ol
    for f in (0, nframes - 1):
        want = ol.port_cfl_pred_plane(d_luma[f].cpu().numpy(), maps_np[f])
        assert np.array_equal(l[f].cpu().numpy(), want), "frame %d differs from the oracle" % f
    line["parity"] = "frames 0 and %d equal to the oracle" % (nframes - 1)
    if not args.no_cpu:
        # CPU baseline: the restatement, one thread, on a few frames of the same workload
        dn = d_luma[:4].cpu().numpy()
        use_ref = ol.have_ref()
        fn = ol.ref_cfl_pred_plane_c if use_ref else ol.port_cfl_pred_plane
        if use_ref:      # od_resample_luma_coeffs itself, leaf by leaf, against the GPU's plane
            assert np.array_equal(l[0].cpu().numpy(), fn(dn[0], maps_np[0])), "frame 0 differs from the reference"
            line["parity"] += " and (frame 0) to the compiled reference"
        t0 = time.perf_counter()
        reps = 0
        while time.perf_counter() - t0 < 5.0:
            for f in range(dn.shape[0]):
                fn(dn[f], maps_np[f])
            reps += 1
        dt = time.perf_counter() - t0
        line["cpu_baseline"] = {"value": round(pic_w * pic_h * dn.shape[0] * reps / dt / 1e6, 1),
                                "unit": "Mpixel/s", "cores": 1, "kind": "reference" if use_ref else "port",
                                "sample": "%d frames x %d passes%s" % (dn.shape[0], reps,
                                          " (od_resample_luma_coeffs per chroma leaf)" if use_ref else "")}
    print(json.dumps(line))
    ctx.close()


if __name__ == "__main__":
    main()

#!/usr/bin/env python3
"""Deringing filter over whole frames: kernel time against the HBM roofline, the CPU
restatement timed beside it on a bounded sample.  One JSON line on stdout.

  python tools/bench_dering.py [--workload 4k] [--frames 64] [--steps 20] [--flag-prob 0.7]

Algorithmic bytes: 8 per plane sample (read int32, write int32); flags and skip maps
(1/16 byte per sample) are ignored.  4K x 64 frames = 6.4 GB per launch, far above L2.
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
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--flag-prob", type=float, default=0.7)
    ap.add_argument("--skip-prob", type=float, default=0.4)
    ap.add_argument("--q", type=int, default=80)
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    import torch
    import daala_b200
    pic_w, pic_h, nframes, _ = bench.WORKLOADS[args.workload]
    nframes = args.frames or nframes
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    ctx = daala_b200.TdltContext(pic_w, pic_h, decs=bench.DECS, device=0, max_frames=0)
    fw, fh, nhsb, nvsb = ctx.frame_width, ctx.frame_height, ctx.nhsb, ctx.nvsb
    g = torch.Generator(device=dev)
    g.manual_seed(4242)
    # oriented stripes per superblock + noise: edges for the direction search to find
    yy = torch.arange(fh, device=dev, dtype=torch.float32)[:, None]
    xx = torch.arange(fw, device=dev, dtype=torch.float32)[None, :]
    planes = []
    luma = torch.empty((nframes, fh, fw), dtype=torch.int32, device=dev)
    for f in range(nframes):
        ang = torch.rand((nvsb, nhsb), device=dev, generator=g) * 3.14159
        ang = ang.repeat_interleave(32, 0).repeat_interleave(32, 1)
        s = torch.sign(torch.sin((xx * torch.cos(ang) + yy * torch.sin(ang)) * (6.2832 / 17.0)))
        v = 900 * s + 40 * torch.randn((fh, fw), device=dev, generator=g)
        luma[f] = torch.clamp(v, -2040, 2040).to(torch.int32)
    planes.append(luma)
    for _ in range(2):
        c = luma[:, ::2, ::2] // 2 + torch.randint(-300, 300, (nframes, fh // 2, fw // 2), device=dev,
                                                   dtype=torch.int32, generator=g)
        planes.append(c.contiguous())
    out = [torch.zeros_like(p) for p in planes]
    flags = (torch.rand((nframes, nvsb, nhsb), device=dev, generator=g) < args.flag_prob).to(torch.uint8)
    bskip = [(torch.rand((nframes, nvsb * 2, nhsb * 2), device=dev, generator=g) < args.skip_prob)
             .repeat_interleave(4, 1).repeat_interleave(4, 2).to(torch.uint8).contiguous() for _ in range(3)]
    q = (args.q, max(1, 3 * args.q // 4), max(1, args.q // 2))
    st = torch.cuda.current_stream().cuda_stream
    op, ip, bp = ([t.data_ptr() for t in x] for x in (out, planes, bskip))

    def step():
        ctx.dering_batch(op, ip, flags.data_ptr(), bp, q, nframes, st)

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
    nsamples = sum(p.numel() for p in planes)
    alg_bytes = 8 * nsamples
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    peak = float(peaks.get("hbm_gbs", 0)) or 6500.0
    achieved = alg_bytes / (ms * 1e-3) / 1e9
    changed = sum(int((a != b).sum()) for a, b in zip(out, planes))
    line = {"metric": "deringing filter, whole frames", "workload":
            "%dx%d 4:2:0, %d frames, %.0f %% of the superblocks filtered, %.0f %% of the 16x16 areas skipped, q %d"
            % (pic_w, pic_h, nframes, 100 * args.flag_prob, 100 * args.skip_prob, args.q),
            "ms_per_step": round(ms, 4),
            "value": round(pic_w * pic_h * nframes / (ms * 1e-3) / 1e6, 1), "unit": "Mpixel/s",
            "gpu_launches": int(launches), "samples_changed_frac": round(changed / nsamples, 4),
            "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                         "frac": round(achieved / peak, 3), "traffic": None,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback"},
            "algorithmic_bytes_per_step": int(alg_bytes)}
    # parity of what was just timed (frame 0) against the restatement
    import oracle_libs as ol
    want = ol.port_dering_frame([p[0].cpu().numpy() for p in planes], bench.DECS, flags[0].cpu().numpy(),
                                [b[0].cpu().numpy() for b in bskip], q)
    for p in range(3):
        assert np.array_equal(out[p][0].cpu().numpy(), want[p]), "plane %d differs from the oracle" % p
    line["parity"] = "frame 0 equal to the oracle"
    dering_out0 = [o[0].cpu().numpy().copy() for o in out]
    # ---- keyframe bilinear smoothing of the 32x32-coded superblocks, in place on the output ----
    maps_np = bench.make_maps_np("quad", nframes, nhsb, nvsb, 78)
    maps = torch.from_numpy(maps_np).to(dev)
    before = [o[0].cpu().numpy().copy() for o in out]
    ctx.smooth_batch(op, maps.data_ptr(), q, nframes, st)       # once, for the parity check
    torch.cuda.synchronize()
    want = ol.port_smooth_frame(before, bench.DECS, maps_np[0], q)
    for p in range(3):
        assert np.array_equal(out[p][0].cpu().numpy(), want[p]), "smoothing: plane %d differs from the oracle" % p
    e0.record()
    for _ in range(args.steps):
        ctx.smooth_batch(op, maps.data_ptr(), q, nframes, st)
    e1.record()
    torch.cuda.synchronize()
    sms = e0.elapsed_time(e1) / args.steps
    frac32 = float((maps_np[:, ::4, ::4] == 3).mean())
    sbytes = 8 * nsamples * frac32
    line["smoothing"] = {"ms_per_step": round(sms, 4), "superblocks_32x32_frac": round(frac32, 3),
                         "algorithmic_bytes_per_step": int(sbytes),
                         "achieved_GBps": round(sbytes / (sms * 1e-3) / 1e9, 1),
                         "frac_of_peak": round(sbytes / (sms * 1e-3) / 1e9 / peak, 3),
                         "parity": "frame 0 equal to the oracle"}
    if not args.no_cpu:
        a = ([p[0].cpu().numpy() for p in planes], bench.DECS, flags[0].cpu().numpy(),
             [b[0].cpu().numpy() for b in bskip], q)
        use_ref = ol.have_ref()
        rc = ol.RefCtx(pic_w, pic_h) if use_ref else None
        if use_ref:      # the reference itself (od_dering, SSE2 filters), frame 0, against the GPU's
            want = ol.ref_dering_frame_c(rc, *a, impl=1)
            for p in range(3):
                assert np.array_equal(dering_out0[p], want[p]), "plane %d differs from the reference" % p
            line["parity"] = "frame 0 equal to the oracle and to the compiled reference"
        t0 = time.perf_counter()
        reps = 0
        while time.perf_counter() - t0 < 5.0:
            if use_ref:
                ol.ref_dering_frame_c(rc, *a, impl=1)
            else:
                ol.port_dering_frame(*a)
            reps += 1
        dt = time.perf_counter() - t0
        line["cpu_baseline"] = {"value": round(pic_w * pic_h * reps / dt / 1e6, 1), "unit": "Mpixel/s",
                                "cores": 1, "kind": "reference" if use_ref else "port",
                                "sample": "1 frame x %d passes (od_dering per superblock and plane, "
                                          "dispatched SSE2 filters)" % reps}
        if use_ref:
            sm_in = [o[0].cpu().numpy() for o in planes]
            t0 = time.perf_counter()
            reps = 0
            while time.perf_counter() - t0 < 3.0:
                ol.ref_smooth_frame_c(rc, sm_in, bench.DECS, maps_np[0], q)
                reps += 1
            dt = time.perf_counter() - t0
            line["smoothing"]["cpu_baseline"] = {"value": round(pic_w * pic_h * reps / dt / 1e6, 1),
                                                 "unit": "Mpixel/s", "cores": 1, "kind": "reference",
                                                 "sample": "1 frame x %d passes (od_smooth_recursive)" % reps}
            rc.close()
    print(json.dumps(line))
    ctx.close()


if __name__ == "__main__":
    main()

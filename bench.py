#!/usr/bin/env python3
"""bench.py -- forward+inverse TDLT throughput (Mpixel/s) on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload 4k|1080p|1080p8x8|720p]
  python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...
  python bench.py --impl reference ...      (the reference's CPU path, host cores)

One "step" = forward stage then inverse stage over a batch of synthetic 4:2:0
frames that is already resident in HBM.  Default workload = the one BASELINE.json's
metric and target are quoted on: 3840x2160 4:2:0, variable block sizes (random
quadtree maps), 64 frames per GPU, keyframe mode.  Frames are
independent, so N GPUs each take their own batch (weak scaling, no collective on
the data path; torch.distributed is used only for the timing barrier / max).

"pixel" = one luma picture pixel (pic_width*pic_height*frames); its luma sample
and 0.5 chroma samples go through both stages.  Algorithmic bytes: 8 per padded
plane sample per direction (read int32, write int32), SURVEY 8d.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (pic_w, pic_h, frames per GPU, map kind)
    "4k": (3840, 2160, 64, "quad"),
    "1080p": (1920, 1080, 64, "quad"),
    "1080p8x8": (1920, 1080, 64, "uniform1"),
    "720p": (1280, 720, 64, "quad"),
    "cif": (352, 288, 10, "quad"),
}
DECS = (0, 1, 1)


_REAL_STDOUT = None


def make_maps_np(kind, nframes, nhsb, nvsb, seed):
    """Vectorised random quadtree (split 32 w.p. .8, 16 w.p. .6, 8 w.p. .5)."""
    rng = np.random.default_rng(seed)
    if kind.startswith("uniform"):
        return np.full((nframes, nvsb * 4, nhsb * 4), int(kind[-1]), np.uint8)
    s32 = rng.random((nframes, nvsb, nhsb)) < 0.8
    s16 = rng.random((nframes, nvsb * 2, nhsb * 2)) < 0.6
    s8 = rng.random((nframes, nvsb * 4, nhsb * 4)) < 0.5
    up = lambda a, k: np.repeat(np.repeat(a, k, axis=1), k, axis=2)
    m = np.where(s8, 0, 1)
    m = np.where(up(s16, 2), m, 2)
    m = np.where(up(s32, 4), m, 3)
    return m.astype(np.uint8)


def frame_dims(pic_w, pic_h):
    return (pic_w + 63) & ~63, (pic_h + 63) & ~63


# ----------------------------------------------------------------- clocks ----
class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1]))
                mx = float(f[2])
            except ValueError:
                continue
            for n, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx,
                "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------- reference (CPU) ----
def cpu_reference_run(pic_w, pic_h, kind, steps, warmup, budget_s=20.0, nthreads=0, opt="-O2"):
    """Times the reference's own CPU implementation of the path (oracle/_ref when it
    is built, else the C restatement) on the host cores: frames spread over threads
    (the functions are re-entrant; ctypes releases the GIL).  nthreads 0 = all cores (at
    most 32).  Returns a dict with the whole fwd+inv figure and the two halves."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_libs as ol
    use_ref = ol.have_ref()
    fw, fh = frame_dims(pic_w, pic_h)
    nhsb, nvsb = fw // 32, fh // 32
    cores = os.cpu_count() or 1
    nthreads = nthreads or max(1, min(cores, 32))
    rng = np.random.default_rng(11)
    maps = make_maps_np(kind, 1, nhsb, nvsb, 7)[0]
    pix = [rng.integers(0, 256, size=(fh >> d, fw >> d)).astype(np.uint8) for d in DECS]
    lib = ol.ref() if use_ref else ol.port()
    to_coeff = lib.ref_buf_to_coeff if use_ref else lib.oc_buf_to_coeff
    to_buf = lib.ref_coeff_to_buf if use_ref else lib.oc_coeff_to_buf
    phase = [[0.0, 0.0] for _ in range(nthreads)]      # seconds in the forward / inverse half

    def make_worker(i):
        ctx = ol.RefCtx(pic_w, pic_h) if use_ref else None
        cbuf = [np.empty(p.shape, np.int32) for p in pix]
        obuf = [np.empty_like(p) for p in pix]

        def one_frame():
            # pixels -> od_coeff (src/state.c:1209), forward stage, inverse stage,
            # od_coeff -> pixels (:1274): SURVEY 8d's CPU baseline, one frame
            for p, dec in enumerate(DECS):
                h, w = pix[p].shape
                t0 = time.perf_counter()
                to_coeff(1, ol.ptr(cbuf[p]), w, 0, ol.ptr(pix[p]), 1, w, w, h)
                if ctx:
                    d, _ = ctx.fwd_plane(cbuf[p], maps, dec, 1, impl=1)
                    t1 = time.perf_counter()
                    c2 = ctx.inv_plane(d, maps, dec, 1, impl=1)
                else:
                    d, _ = ol.port_fwd_plane(cbuf[p], maps, dec, pic_w, pic_h, 1)
                    t1 = time.perf_counter()
                    c2 = ol.port_inv_plane(d, maps, dec, pic_w, pic_h, 1)
                to_buf(1, ol.ptr(obuf[p]), 1, w, ol.ptr(c2), w, 0, w, h)
                t2 = time.perf_counter()
                phase[i][0] += t1 - t0
                phase[i][1] += t2 - t1
        return one_frame

    workers = [make_worker(i) for i in range(nthreads)]
    # size the per-step sample from one single-thread frame
    t0 = time.perf_counter()
    workers[0]()
    t_frame = time.perf_counter() - t0
    total_steps = steps + warmup
    per_thread = max(1, int(budget_s / max(t_frame, 1e-6) / max(total_steps, 1)))
    per_thread = min(per_thread, 64)

    def run_step():
        ths = [threading.Thread(target=lambda w=w: [w() for _ in range(per_thread)]) for w in workers]
        t0 = time.perf_counter()
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        return time.perf_counter() - t0

    for _ in range(warmup):
        run_step()
    for ph in phase:
        ph[0] = ph[1] = 0.0
    times = [run_step() for _ in range(steps)]
    frames = nthreads * per_thread
    el = float(np.mean(times))
    mpix = pic_w * pic_h * frames / el / 1e6
    # the two halves: frames / (mean per-thread time spent in that half per step)
    tf = float(np.mean([ph[0] for ph in phase])) / max(steps, 1)
    ti = float(np.mean([ph[1] for ph in phase])) / max(steps, 1)
    return {"value": mpix, "unit": "Mpixel/s", "cores": nthreads,
            "kind": "reference" if use_ref else "port",
            "sample": "%d frames of %dx%d per step (%d threads x %d), 8-bit pixels -> coeff -> fwd -> inv -> "
                      "pixels, gcc %s, x86 SIMD dispatch for 4x4/8x8" % (frames, pic_w, pic_h, nthreads,
                                                                            per_thread, opt),
            "fwd_only": pic_w * pic_h * frames / max(tf, 1e-9) / 1e6,
            "inv_only": pic_w * pic_h * frames / max(ti, 1e-9) / 1e6,
            "opt": opt, "ms_per_step": el * 1e3, "host_cores": cores}


def cpu_baseline_variants(pic_w, pic_h, kind, workload):
    """the reference at gcc -O2 and -O3 -march=native, on one thread and
    on all host threads, whole stage and the two halves.  The -O3 build runs in a subprocess
    (another library of the same symbols; -march=native was chosen on the build machine, so
    an illegal instruction on this host is reported, not fatal).  ~35 s in all."""
    out = []
    r1 = cpu_reference_run(pic_w, pic_h, kind, 1, 0, budget_s=4.0, nthreads=1)
    out.append({"opt": "-O2", "threads": 1, "value": r1["value"], "fwd_only": r1["fwd_only"],
                "inv_only": r1["inv_only"], "kind": r1["kind"]})
    o3 = os.path.join(ROOT, "oracle", "_ref", "libdaala_ref_o3.so")
    if os.path.exists(o3):
        for nthr, budget in ((1, 4.0), (0, 8.0)):
            try:
                env = dict(os.environ, DAALA_REF_LIB=o3)
                pr = subprocess.run([sys.executable, os.path.abspath(__file__), "--impl", "reference",
                                     "--workload", workload, "--steps", "1", "--warmup", "1" if nthr == 0 else "0",
                                     "--cpu-threads", str(nthr), "--cpu-budget", str(budget),
                                     "--cpu-opt", "-O3 -march=native"],
                                    env=env, capture_output=True, text=True, timeout=120)
                ln = json.loads(pr.stdout.strip().splitlines()[-1])
                cb = ln["cpu_baseline"]
                out.append({"opt": "-O3 -march=native", "threads": cb["cores"], "value": cb["value"],
                            "fwd_only": cb.get("fwd_only"), "inv_only": cb.get("inv_only"), "kind": cb["kind"]})
            except Exception as e:      # noqa: BLE001 -- reported in the line
                out.append({"opt": "-O3 -march=native", "threads": nthr or "all",
                            "unavailable": "%s: %s" % (type(e).__name__, str(e)[:80])})
    return out


# ------------------------------------------------------------------- main ----
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="4k", choices=sorted(WORKLOADS))
    ap.add_argument("--frames", type=int, default=0, help="frames per GPU (0 = workload default)")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-threads", type=int, default=0,
                    help="host threads of the e2e figure, one context and one batch each "
                         "(0 = 6 on one GPU, 2 per rank under torchrun)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the secondary lines (fused pixel kernels, 16-bit / pageable e2e, neighbours)")
    ap.add_argument("--cpu-threads", type=int, default=0, help="--impl reference: host threads (0 = all, <= 32)")
    ap.add_argument("--cpu-budget", type=float, default=20.0, help="--impl reference: seconds of CPU work")
    ap.add_argument("--cpu-opt", default="-O2", help="--impl reference: label of the library build timed")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    pic_w, pic_h, nframes, kind = WORKLOADS[args.workload]
    if args.frames:
        nframes = args.frames
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    config = {"workload": "%dx%d 4:2:0, %s block-size maps, %d frames per GPU, keyframe DC pyramid, "
                          "od_coeff int32 planes" % (pic_w, pic_h,
                                                     "random-quadtree 4x4..32x32" if kind == "quad" else kind,
                                                     nframes),
              "frames_per_gpu": nframes, "l2": "inputs larger than L2 (no flush needed)",
              "parallelism": "frames sharded over %d GPU(s), no collective" % world}

    if args.impl == "reference":
        if rank != 0:
            return
        r = cpu_reference_run(pic_w, pic_h, kind, args.steps, args.warmup, budget_s=args.cpu_budget,
                              nthreads=args.cpu_threads, opt=args.cpu_opt)
        line = {"impl": "reference", "metric": "Mpixel/s fwd+inv TDLT (4:2:0, var bsize)",
                "value": r["value"], "unit": "Mpixel/s", "n_gpus": args.gpus, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
                "config": config,
                "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample", "fwd_only",
                                                   "inv_only", "opt")},
                "e2e": {"value": r["value"], "unit": "Mpixel/s", "h2d_bytes_per_step": 0,
                        "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line))
        return

    import torch
    import daala_b200
    from daala_b200.sharding import mpixels_per_second, reduce_timing
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback for the product path)")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        # stdout carries exactly one JSON line: NCCL writes its version banner to fd 1 from
        # C (at WARN level too), so fd 1 points at stderr while the job runs and the line is
        # written to the saved descriptor at the end
        os.environ.setdefault("NCCL_DEBUG", "WARN")
        sys.stdout.flush()
        global _REAL_STDOUT
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)

    ctx = daala_b200.TdltContext(pic_w, pic_h, decs=DECS, device=local_rank,
                                 max_frames=0 if args.no_e2e else nframes)
    fw, fh = ctx.frame_width, ctx.frame_height
    shapes = ctx.plane_shapes
    samples_per_frame = sum(h * w for h, w in shapes)
    g = torch.Generator(device=dev)
    g.manual_seed(1234 + rank)

    def synth(h, w):
        # smooth field + texture, 12-bit coefficient scale (8-bit << 4), on the GPU
        low = torch.randn((nframes, 1, h // 8 + 1, w // 8 + 1), device=dev, generator=g)
        img = torch.nn.functional.interpolate(low, size=(h, w), mode="bilinear", align_corners=False)
        img = img.squeeze(1) * 600.0 + torch.randn((nframes, h, w), device=dev, generator=g) * 60.0
        return img.round().clamp_(-2048, 2047).to(torch.int32).contiguous()

    c = [synth(h, w) for h, w in shapes]
    d = [torch.empty_like(t) for t in c]
    c2 = [torch.empty_like(t) for t in c]
    maps_np = make_maps_np(kind, nframes, ctx.nhsb, ctx.nvsb, 99 + rank)
    maps = torch.from_numpy(maps_np).to(dev)
    stream = torch.cuda.current_stream().cuda_stream
    cp = [t.data_ptr() for t in c]
    dp = [t.data_ptr() for t in d]
    c2p = [t.data_ptr() for t in c2]
    FL = daala_b200.OD_TDLT_KEYFRAME

    def fwd():
        ctx.fwd_batch(dp, cp, maps.data_ptr(), nframes, FL, stream)

    def inv():
        # same maps as the forward call just before: its per-region plan is reused
        ctx.inv_batch(c2p, dp, maps.data_ptr(), nframes, FL | daala_b200.OD_TDLT_SAME_MAPS, stream)

    def barrier():
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
            torch.cuda.synchronize()

    sampler = ClockSampler(local_rank)
    sampler.start()          # samples through warm-up and the timed region: always under load
    for _ in range(args.warmup):
        fwd()
        inv()
    torch.cuda.synchronize()
    ok = all(torch.equal(a, b) for a, b in zip(c, c2))   # lossless round trip at full size
    if not ok:
        raise SystemExit("bench.py: round trip is not the identity -- refusing to report a number")

    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2 * args.steps + 1)]
    barrier()
    l0 = daala_b200.launch_count()
    ev[0].record()
    for k in range(args.steps):
        fwd()
        ev[2 * k + 1].record()
        inv()
        ev[2 * k + 2].record()
    torch.cuda.synchronize()
    launches = daala_b200.launch_count() - l0
    # a short timed region may end before nvidia-smi's first sample: keep the same load
    # running (untimed) until a few samples are in
    t_end = time.perf_counter() + 1.5
    while len(sampler.lines) < 4 and time.perf_counter() < t_end:
        fwd()
        inv()
        torch.cuda.synchronize()
    clocks = sampler.stop()
    barrier()
    total_ms = ev[0].elapsed_time(ev[-1])
    fwd_ms = float(np.mean([ev[2 * k].elapsed_time(ev[2 * k + 1]) for k in range(args.steps)]))
    inv_ms = float(np.mean([ev[2 * k + 1].elapsed_time(ev[2 * k + 2]) for k in range(args.steps)]))
    # whole job: max elapsed over ranks, frames summed over ranks (no data-path collective)
    total_ms, frames_all = reduce_timing(dist, total_ms, nframes, dev)
    ms_per_step = total_ms / args.steps
    value = mpixels_per_second(pic_w, pic_h, frames_all, ms_per_step)

    # roofline of the dominant kernel (8 algorithmic bytes per padded plane sample per launch)
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = json.load(open(peaks_path))["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    alg_bytes = 8.0 * samples_per_frame * nframes
    dom, dom_ms = (("tdlt_inv_region_kernel", inv_ms) if inv_ms >= fwd_ms
                   else ("tdlt_fwd_region_kernel", fwd_ms))
    ach = alg_bytes / (dom_ms * 1e-3) / 1e9
    roof = {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak, "unit": "GB/s",
            "frac": ach / peak, "traffic": None, "peak_source": peak_src,
            "algorithmic_bytes_per_launch": alg_bytes,
            "fwd_ms": fwd_ms, "inv_ms": inv_ms,
            "fwd_GBps": alg_bytes / (fwd_ms * 1e-3) / 1e9, "inv_GBps": alg_bytes / (inv_ms * 1e-3) / 1e9,
            "step_frac_of_peak": 2 * alg_bytes / (ms_per_step * 1e-3) / 1e9 / peak}
    prof = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(prof):
        try:
            tr = json.load(open(prof))
            # dram bytes read + written per launch from one `ncu --set full` capture of this
            # workload (tools/ncu_traffic.py); only meaningful for the same frames per launch
            if tr.get("frames_per_launch") == nframes and tr.get("workload") == args.workload:
                roof["traffic"] = tr.get(dom)
                roof["traffic_source"] = tr.get("source")
        except Exception:
            pass

    # a decoder never runs the forward stage first: the inverse step that builds its own plan
    def inv_own_plan():
        ctx.inv_batch(c2p, dp, maps.data_ptr(), nframes, FL, stream)
    for _ in range(2):
        inv_own_plan()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(max(5, args.steps // 2)):
        inv_own_plan()
    e1.record()
    torch.cuda.synchronize()
    inv_plan_ms = e0.elapsed_time(e1) / max(5, args.steps // 2)
    roof["inv_with_plan_ms"] = inv_plan_ms
    roof["inv_with_plan_frac"] = alg_bytes / (inv_plan_ms * 1e-3) / 1e9 / peak

    # ---- the same stages with the pixel conversion fused into load / store ------------------
    # (od_tdlt_fwd_batch_pix / od_tdlt_inv_batch_pix, device-resident): SURVEY 8d accounting of
    # the bytes this variant must move, (xstride + 4) per padded sample and direction
    pix_lines = None
    if not args.no_extras:
        pix_lines = {}
        for xs, mode, name in ((1, 1, "8bit"), (2, 2, "16bit")):
            if xs == 1:
                px = [((t >> 4) + 128).clamp_(0, 255).to(torch.uint8) for t in c]
            else:
                px = [(t + 2048).clamp_(0, 4095).to(torch.int16) for t in c]
            po = [torch.empty_like(t) for t in px]
            pxp, pop = [t.data_ptr() for t in px], [t.data_ptr() for t in po]

            def fwd_pix():
                ctx.fwd_batch_pix(dp, pxp, maps.data_ptr(), nframes, FL, xs, mode, 0, stream)

            def inv_pix():
                ctx.inv_batch_pix(pop, dp, maps.data_ptr(), nframes, FL | daala_b200.OD_TDLT_SAME_MAPS,
                                  xs, mode, 0, stream)
            for _ in range(3):
                fwd_pix()
                inv_pix()
            torch.cuda.synchronize()
            if not all(torch.equal(a_, b_) for a_, b_ in zip(px, po)):
                raise SystemExit("bench.py: fused pixel round trip (%s) is not the identity" % name)
            evp = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            k2 = max(5, args.steps // 2)
            evp[0].record()
            for _ in range(k2):
                fwd_pix()
            evp[1].record()
            for _ in range(k2):
                inv_pix()
            evp[2].record()
            torch.cuda.synchronize()
            fms, ims = evp[0].elapsed_time(evp[1]) / k2, evp[1].elapsed_time(evp[2]) / k2
            byts = float(xs + 4) * samples_per_frame * nframes
            pix_lines[name] = {
                "fwd_ms": fms, "inv_ms": ims,
                "value": mpixels_per_second(pic_w, pic_h, nframes, fms + ims), "unit": "Mpixel/s",
                "bytes_per_launch": byts, "accounting": "(%d + 4) B per padded sample and direction "
                "(pixels one side, od_coeff the other; SURVEY 8d fused variant)" % xs,
                "fwd_frac_of_peak": byts / (fms * 1e-3) / 1e9 / peak,
                "inv_frac_of_peak": byts / (ims * 1e-3) / 1e9 / peak,
                "two_pass_equivalent_bytes": float(xs + 12) * samples_per_frame * nframes}
            del px, po

    # ---- the host link: plain cudaMemcpyAsync both ways at once, pinned memory ---------------
    def link_ceiling(nbytes=1 << 30, reps=3):
        hu = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
        hd = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
        du = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        dd_ = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

        def run(up, down):
            barrier()
            t0 = time.perf_counter()
            for _ in range(reps):
                if up:
                    with torch.cuda.stream(s1):
                        du.copy_(hu, non_blocking=True)
                if down:
                    with torch.cuda.stream(s2):
                        hd.copy_(dd_, non_blocking=True)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            dt_ms, _ = reduce_timing(dist, dt * 1e3, 1, dev)
            return world * reps * nbytes * (int(up) + int(down)) / (dt_ms * 1e-3) / 1e9
        run(True, True)
        return {"h2d_alone_GBps": run(True, False), "d2h_alone_GBps": run(False, True),
                "both_GBps": run(True, True), "processes": world,
                "how": "cudaMemcpyAsync of %d MiB pinned buffers, %d each way, two streams; all "
                       "ranks at once, whole-node total" % (nbytes >> 20, reps)}

    # end to end: host buffers through the frame-level C-ABI calls the codec would make:
    # reference-buffer pixels in -> od_coeff d out (encoder side), d in -> pixels out (decoder
    # side); every byte crosses PCIe inside the timed region
    e2e = None
    if not args.no_e2e:
        import ctypes as C
        link = link_ceiling()
        # Several host threads, one context each (contexts are
        # independent), each running whole steps on its own batch, started a fraction of a
        # step apart: the forward call is download-heavy (pixels up, od_coeff down), the
        # inverse call upload-heavy, so the two directions of the link are busy at the same
        # time.  Measured on one GPU (profiles/e2e_threads_r2w.txt): 1 / 2 / 4 / 6 / 8 threads
        # reach 0.67 / 0.78 / 0.89 / 0.93-0.95 / 1.0 of the link (6 pin 48 GB of host memory);
        # under torchrun the ranks already share the host's link, two per rank reach 0.96 at
        # N = 2.
        nthr = args.e2e_threads if args.e2e_threads > 0 else (6 if world == 1 else 2)
        if args.e2e_threads <= 0:
            # every thread holds its own host planes (od_coeff + 16-bit pixels of a batch):
            # stay well inside the host's free memory, two threads at least
            try:
                import psutil
                per_thread = 6 * samples_per_frame * nframes * world
                while nthr > 2 and 2 * nthr * per_thread > psutil.virtual_memory().available:
                    nthr -= 1
            except ImportError:
                pass
        L = ctx.L
        arr = daala_b200._ptr_array
        ctxs = [ctx] + [daala_b200.TdltContext(pic_w, pic_h, device=local_rank, max_frames=nframes)
                        for _ in range(nthr - 1)]
        hmaps = torch.from_numpy(maps_np).pin_memory()

        def e2e_variant(xs, mode, memory):
            """memory: "pinned" (cudaHostAlloc), "pageable" (plain malloc, what the reference's
            od_state holds) or "registered" (plain malloc + od_tdlt_host_register)."""
            pdt, ndt = (torch.uint8, np.uint8) if xs == 1 else (torch.int16, np.int16)

            def host(shape, tdt, ndt_):
                if memory == "pinned":
                    return torch.empty(shape, dtype=tdt, pin_memory=True)
                a_ = np.empty(shape, ndt_)
                a_[...] = 0                       # touch the pages
                if memory == "registered":
                    daala_b200.host_register(a_)
                    regd.append(a_)
                return torch.from_numpy(a_)
            regd = []
            hp = [host(t.shape, pdt, ndt) for t in c]
            for a_, b_ in zip(hp, c):
                if xs == 1:
                    a_.copy_(((b_ >> 4) + 128).clamp_(0, 255).to(torch.uint8))
                else:
                    a_.copy_((b_ + 2048).clamp_(0, 4095).to(torch.int16))
            ys = (C.c_int * 4)(*[t.shape[2] * xs for t in hp] + [0] * (4 - len(hp)))
            fs = (C.c_longlong * 4)(*[t.shape[1] * t.shape[2] * xs for t in hp] + [0] * (4 - len(hp)))
            hds = [[host(t.shape, torch.int32, np.int32) for t in c] for _ in range(nthr)]
            hp2s = [[host(t.shape, pdt, ndt) for t in c] for _ in range(nthr)]
            errs = []

            def e2e_fwd(i):
                rc = L.od_tdlt_fwd_frames_pix(ctxs[i]._h, arr([x.data_ptr() for x in hds[i]]),
                                              arr([x.data_ptr() for x in hp]), ys, fs, hmaps.data_ptr(),
                                              nframes, FL, xs, mode, 0)
                if rc:
                    errs.append("fwd: " + L.od_tdlt_last_error(ctxs[i]._h).decode())

            def e2e_inv(i):
                rc = L.od_tdlt_inv_frames_pix(ctxs[i]._h, arr([x.data_ptr() for x in hp2s[i]]), ys, fs,
                                              arr([x.data_ptr() for x in hds[i]]), hmaps.data_ptr(),
                                              nframes, FL, xs, mode, 0)
                if rc:
                    errs.append("inv: " + L.od_tdlt_last_error(ctxs[i]._h).decode())

            def worker(i, nsteps, go, started):
                go.wait()
                for k in range(nsteps):
                    e2e_fwd(i)
                    if k == 0:
                        started.set()          # the next thread starts now: half a step behind this one
                    e2e_inv(i)

            def run_steps(nsteps):
                evs = [threading.Event() for _ in range(nthr + 1)]
                ths = [threading.Thread(target=worker, args=(i, nsteps, evs[i], evs[i + 1]))
                       for i in range(nthr)]
                for th in ths:
                    th.start()
                t0 = time.perf_counter()
                evs[0].set()
                for th in ths:
                    th.join()
                torch.cuda.synchronize()
                return time.perf_counter() - t0

            e2e_steps = max(2, min(args.steps, 3))      # steps per thread
            run_steps(1)
            barrier()
            el = run_steps(e2e_steps)
            if errs:
                raise SystemExit("e2e call failed: %s" % errs[0])
            el_ms, frames_e2e = reduce_timing(dist, el * 1e3, nframes * nthr, dev)
            el = el_ms * 1e-3
            for hp2 in hp2s:
                if not all(torch.equal(a_, b_) for a_, b_ in zip(hp, hp2)):
                    raise SystemExit("bench.py: e2e round trip mismatch")
            for a_ in regd:
                daala_b200.host_unregister(a_)
            coeff_bytes = 4 * samples_per_frame * nframes
            pix_bytes = xs * samples_per_frame * nframes
            up, down = pix_bytes + coeff_bytes + 2 * maps_np.nbytes, coeff_bytes + pix_bytes
            gbps = (up + down) * e2e_steps * nthr * world / el / 1e9
            return {"value": pic_w * pic_h * frames_e2e * e2e_steps / el / 1e6, "unit": "Mpixel/s",
                    "h2d_bytes_per_step": up, "d2h_bytes_per_step": down, "steps": e2e_steps * nthr,
                    "host_threads": nthr, "host_memory": memory, "sample_bytes": xs,
                    "link_GBps": gbps, "link_frac": gbps / link["both_GBps"]}

        e2e = e2e_variant(1, 1, "pinned")
        e2e["api"] = ("od_tdlt_fwd_frames_pix (8-bit pixels -> od_coeff) + od_tdlt_inv_frames_pix "
                      "(od_coeff -> 8-bit pixels), pixel conversion fused into the transform kernels, "
                      "synchronous calls; %d host thread(s), one context and one batch each, each "
                      "started when the one before it has finished its first forward call" % nthr)
        e2e["link"] = link
        e2e["link_frac_means"] = ("bytes moved both ways per second / what plain cudaMemcpyAsync moves "
                                  "both ways at once on this box (link.both_GBps)")
        if world == 1 and not args.no_extras:
            # config 4's 16-bit reference buffers, and what the call costs from the plain-malloc
            # planes the reference's od_state holds, before and after od_tdlt_host_register
            e2e["variants"] = {"16bit_pinned": e2e_variant(2, 2, "pinned"),
                               "8bit_pageable": e2e_variant(1, 1, "pageable"),
                               "8bit_registered": e2e_variant(1, 1, "registered")}
        for cx in ctxs[1:]:
            cx.close()

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        r = cpu_reference_run(pic_w, pic_h, kind, 2, 1, budget_s=15.0)
        cpu = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample", "fwd_only", "inv_only", "opt")}
        variants = cpu_baseline_variants(pic_w, pic_h, kind, args.workload)
        variants.insert(1, {"opt": "-O2", "threads": r["cores"], "value": r["value"],
                            "fwd_only": r["fwd_only"], "inv_only": r["inv_only"], "kind": r["kind"]})
        cpu["variants"] = variants

    # ---- the neighbouring stages (SURVEY 8f rows built so far), each by its own tool --------
    neighbours = None
    if rank == 0 and world == 1 and not args.no_extras:
        ctx.close()
        ctx = None
        del c, d, c2
        torch.cuda.empty_cache()
        neighbours = {}
        for name, tool, extra in (("deringing_and_smoothing", "bench_dering.py", []),
                                  ("chroma_from_luma", "bench_cfl.py", []),
                                  ("coding_order", "bench_coding.py", [])):
            try:
                pr = subprocess.run([sys.executable, os.path.join(ROOT, "tools", tool), "--workload",
                                     args.workload, "--steps", "10"] + extra,
                                    capture_output=True, text=True, timeout=300)
                neighbours[name] = json.loads(pr.stdout.strip().splitlines()[-1])
            except Exception as e:      # noqa: BLE001 -- reported in the line
                neighbours[name] = {"unavailable": "%s: %s" % (type(e).__name__, str(e)[:120])}

    if rank == 0:
        line = {"metric": "Mpixel/s fwd+inv TDLT (4:2:0, var bsize)", "value": value, "unit": "Mpixel/s",
                "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32",
                "data": "synthetic", "config": config, "clocks": clocks, "roofline": roof,
                "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches),
                "round_trip_exact": True, "fused_pixel_kernels": pix_lines, "neighbours": neighbours}
        if _REAL_STDOUT is not None:
            sys.stdout.flush()
            os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())
        else:
            print(json.dumps(line))
    if ctx is not None:
        ctx.close()
    if dist:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

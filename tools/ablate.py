#!/usr/bin/env python3
"""Times forward and inverse launches of every library under daala_b200/variants/
(plus the product library) on the bench workload.  Results of ablated variants are wrong
by construction; only the product library is checked.  usage: ablate.py [frames] [kind]"""
import glob, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if len(sys.argv) > 1 and sys.argv[1] == "--one":
    sys.path.insert(0, ROOT)
    import numpy as np, torch
    import daala_b200
    from bench import make_maps_np
    nframes, kind, flags = int(sys.argv[2]), sys.argv[3], int(sys.argv[4])
    dev = torch.device("cuda", 0)
    ctx = daala_b200.TdltContext(3840, 2160, device=0)
    g = torch.Generator(device=dev); g.manual_seed(1)
    c = [torch.randint(-2048, 2048, (nframes,) + s, dtype=torch.int32, device=dev, generator=g) for s in ctx.plane_shapes]
    d = [torch.empty_like(t) for t in c]; c2 = [torch.empty_like(t) for t in c]
    if kind in ("half38", "half31", "quarters"):
        # regions are 128x64 luma samples = 16 x 8 map entries: classes laid out so that every
        # region holds whole warp-tasks only (no partially filled task)
        m = np.zeros((nframes, ctx.nvsb*4, ctx.nhsb*4), np.uint8)
        xx = np.arange(ctx.nhsb*4)[None, None, :] % 16
        yy = np.arange(ctx.nvsb*4)[None, :, None] % 8
        if kind == "half38":
            m[:] = np.where(xx < 8, 3, 1)
        elif kind == "half31":
            m[:] = np.where(xx < 8, 3, 0)
        else:
            m[:] = np.where(xx < 4, 3, np.where(xx < 8, 2, np.where(xx < 12, 1, 0)))
        maps = torch.from_numpy(m).to(dev)
    else:
        maps = torch.from_numpy(make_maps_np(kind, nframes, ctx.nhsb, ctx.nvsb, 5)).to(dev)
    st = torch.cuda.current_stream().cuda_stream
    def fwd(): ctx.fwd_batch([t.data_ptr() for t in d], [t.data_ptr() for t in c], maps.data_ptr(), nframes, flags, st)
    def inv(): ctx.inv_batch([t.data_ptr() for t in c2], [t.data_ptr() for t in d], maps.data_ptr(), nframes, flags, st)
    res = []
    for f in (fwd, inv):
        for _ in range(3): f()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10): f()
        e1.record(); torch.cuda.synchronize()
        res.append(e0.elapsed_time(e1)/10)
    ok = all(torch.equal(a, b) for a, b in zip(c, c2))
    extra = ""
    try:      # measurement builds (-DTDLT_RINGSTAT) count how often the inverse waits for its ring
        import ctypes
        fn = daala_b200._lib.od_tdlt_debug_ringstat
        out = (ctypes.c_ulonglong*8)()
        fn(out, 1); inv(); torch.cuda.synchronize(); fn(out, 0)
        extra = "  ring: need %d, not there at start %d, not there after the transforms %d, polls %d" % (out[1], out[2], out[3], out[4])
    except AttributeError:
        pass
    print("fwd %.3f ms  inv %.3f ms  roundtrip %s%s" % (res[0], res[1], "ok" if ok else "WRONG", extra))
    sys.exit(0)
nframes = sys.argv[1] if len(sys.argv) > 1 else "64"
kind = sys.argv[2] if len(sys.argv) > 2 else "quad"
flags = sys.argv[3] if len(sys.argv) > 3 else "1"
libs = [os.path.join(ROOT, "daala_b200", "libdaala_tdlt_cuda.so")] + sorted(glob.glob(os.path.join(ROOT, "daala_b200", "variants", "lib_*.so")))
for lib in libs:
    env = dict(os.environ, DAALA_TDLT_LIB=lib)
    try:
        r = subprocess.run([sys.executable, __file__, "--one", nframes, kind, flags], env=env, capture_output=True,
                           text=True, timeout=120)
        print("%-28s %s" % (os.path.basename(lib), (r.stdout.strip() or r.stderr.strip()[-200:])), flush=True)
    except subprocess.TimeoutExpired:
        print("%-28s TIMEOUT" % os.path.basename(lib), flush=True)

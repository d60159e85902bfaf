#!/usr/bin/env python3
"""Per-phase clock64() timeline of the region kernels (debug build only).
  python daala_b200/build.py --trace
  DAALA_TDLT_LIB=daala_b200/libdaala_tdlt_cuda_trace.so python tools/trace_phases.py
Prints the mean / median cycles between consecutive marks over the first 4096 CTAs."""
import ctypes, os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import daala_b200
from bench import make_maps_np

pic_w, pic_h, nframes = 3840, 2160, 8
dev = torch.device("cuda", 0)
ctx = daala_b200.TdltContext(pic_w, pic_h, device=0)
L = ctx.L
L.od_tdlt_debug_set_trace.argtypes = [ctypes.c_void_p]
trace = torch.zeros((4096, 16), dtype=torch.int64, device=dev)
c = [torch.randint(-2048, 2048, (nframes,) + s, dtype=torch.int32, device=dev) for s in ctx.plane_shapes]
d = [torch.empty_like(t) for t in c]
c2 = [torch.empty_like(t) for t in c]
maps = torch.from_numpy(make_maps_np("quad", nframes, ctx.nhsb, ctx.nvsb, 5)).to(dev)
st = torch.cuda.current_stream().cuda_stream
names = {"fwd": ["start", "setup", "tma wait", "-", "lapping", "transforms", "dc+store"],
         "inv": ["start", "setup", "tma wait", "dc split", "transforms", "publish", "fetch", "lapping", "store"]}
for which in ("fwd", "inv"):
    for rep in range(3):
        trace.zero_()
        assert L.od_tdlt_debug_set_trace(trace.data_ptr()) == 0
        if which == "fwd":
            ctx.fwd_batch([t.data_ptr() for t in d], [t.data_ptr() for t in c], maps.data_ptr(), nframes, 1, st)
        else:
            ctx.inv_batch([t.data_ptr() for t in c2], [t.data_ptr() for t in d], maps.data_ptr(), nframes, 1, st)
        torch.cuda.synchronize()
    t = trace.cpu().numpy()
    nm = names[which]
    # skip the first wave (cold) and use CTAs 1000..4095
    t = t[1000:]
    prev = 0
    print(which, "cycles per phase (mean / median):")
    tot = t[:, len(nm) - 1] - t[:, 0]
    for k in range(1, len(nm)):
        if nm[k] == "-":
            continue
        dlt = t[:, k] - t[:, prev]
        print("  %-12s %8.0f %8.0f" % (nm[k], dlt.mean(), np.median(dlt)))
        prev = k
    print("  %-12s %8.0f %8.0f" % ("lifetime", tot.mean(), np.median(tot)))
assert all(torch.equal(a, b) for a, b in zip(c, c2))

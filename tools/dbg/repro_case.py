#!/usr/bin/env python3
"""Replays tools/dbg/fuzz_case.npz (a tools/fuzz_parity.py dump) with variations."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import daala_b200
from checker import Checker
z = np.load(os.path.join(ROOT, "tools", "dbg", "fuzz_case.npz"))
W, H = [int(v) for v in z["pic"]]; decs = tuple(int(v) for v in z["decs"]); bsize = z["bsize"]
dq0 = [z["dq%d" % i] for i in range(3)]
chk = Checker(W, H)
def run(name, dq, key, frames=None):
    if frames is not None:
        dq = [a[frames] for a in dq]; bs = bsize[frames]
    else:
        bs = bsize
    n = dq[0].shape[0]
    flags = daala_b200.OD_TDLT_KEYFRAME if key else 0
    with daala_b200.TdltContext(W, H, decs=decs, device=0, max_frames=n) as ctx:
        out = [np.zeros_like(a) for a in dq]
        ctx.inv_frames(out, dq, bs, flags)
    ref = chk.inv_frames(dq, bs, decs, key)
    for p in range(3):
        bad = np.argwhere(out[p] != ref[p])
        if len(bad):
            print("%-28s plane %d: %d differ, frames %s rows %s cols %s" % (name, p, len(bad), sorted(set(bad[:, 0].tolist())),
                  sorted(set(bad[:, 1].tolist()))[:12], sorted(set(bad[:, 2].tolist()))[:12]))
            return
    print("%-28s equal" % name)
run("as dumped", dq0, 1)
run("not a keyframe", dq0, 0)
run("frame 1 alone", dq0, 1, slice(1, 2))
run("halved coefficients", [a // 2 for a in dq0], 1)
run("quartered coefficients", [a // 4 for a in dq0], 1)
# only the neighbourhood of the mismatch kept (chroma rows 32..63, cols 96..127 = luma 64..127, 192..255)
m = [np.zeros_like(a) for a in dq0]
for p, dec in enumerate(decs):
    ys, xs = (64 >> dec, 128 >> dec), (192 >> dec, 256 >> dec)
    m[p][:, ys[0]:ys[1], xs[0]:xs[1]] = dq0[p][:, ys[0]:ys[1], xs[0]:xs[1]]
run("only the neighbourhood", m, 1)

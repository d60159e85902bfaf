#!/usr/bin/env python3
"""Repro helper: inverse of arbitrary coefficients for one geometry, several seeds; prints
where the CUDA planes differ from the checker.  usage: repro_inv.py W H kind nframes key [lo hi]"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import daala_b200
from checker import Checker, make_maps
W, H, kind, nframes, key = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3], int(sys.argv[4]), int(sys.argv[5])
lo, hi = (int(sys.argv[6]), int(sys.argv[7])) if len(sys.argv) > 7 else (-30000, 30000)
decs = (0, 1, 1)
flags = daala_b200.OD_TDLT_KEYFRAME if key else 0
chk = Checker(W, H)
with daala_b200.TdltContext(W, H, decs=decs, device=0, max_frames=nframes) as ctx:
    for seed in range(int(os.environ.get('SEEDS', '6'))):
        rng = np.random.default_rng(seed)
        bsize = make_maps(kind, nframes, ctx.nhsb, ctx.nvsb, rng)
        dq = [rng.integers(lo, hi, size=(nframes,) + s).astype(np.int32) for s in ctx.plane_shapes]
        ref = chk.inv_frames(dq, bsize, decs, key)
        c0 = [rng.integers(-2048, 2048, size=a.shape).astype(np.int32) for a in dq]
        d0 = [np.zeros_like(a) for a in dq]
        ctx.fwd_frames(d0, c0, bsize, flags)
        ctx.inv_frames([np.zeros_like(a) for a in dq], d0, bsize, flags)
        for rep in range(2):
            c2 = [np.zeros_like(a) for a in dq]
            ctx.inv_frames(c2, dq, bsize, flags)
            for p in range(3):
                bad = np.argwhere(c2[p] != ref[p])
                if len(bad):
                    f, y, x = bad.min(0); F, Y, X = bad.max(0)
                    print("seed %d rep %d plane %d: %d samples differ, frames %d..%d rows %d..%d cols %d..%d; first %s got %d want %d"
                          % (seed, rep, p, len(bad), f, F, y, Y, x, X, tuple(bad[0]), c2[p][tuple(bad[0])], ref[p][tuple(bad[0])]))
                    ys = sorted(set(bad[:, 1].tolist())); xs = sorted(set(bad[:, 2].tolist()))
                    print("   rows", ys[:40], "cols", xs[:40])
chk.close()
print('repro done')

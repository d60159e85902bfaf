import os, sys
os.environ["CUDA_LAUNCH_BLOCKING"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import daala_b200
dev = torch.device("cuda", 0)
for xs in (1, 2):
    ctx = daala_b200.TdltContext(352, 288, device=0)
    n = 2
    maps = torch.ones((n,) + ctx.bsize_shape, dtype=torch.uint8, device=dev)
    dt = torch.uint8 if xs == 1 else torch.int16
    pix = [torch.randint(0, 200, (n,) + s, device=dev).to(dt) for s in ctx.plane_shapes]
    d = [torch.zeros((n,) + s, dtype=torch.int32, device=dev) for s in ctx.plane_shapes]
    try:
        ctx.fwd_batch_pix([t.data_ptr() for t in d], [t.data_ptr() for t in pix], maps.data_ptr(), n, 1, xs, 1, 0, 0)
        torch.cuda.synchronize()
        print("xstride", xs, "fwd ok")
    except Exception as e:
        print("xstride", xs, "fwd FAILED", e)
        break

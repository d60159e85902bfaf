#!/usr/bin/env python3
"""The host link of this box: plain cudaMemcpyAsync host -> device and device -> host from
pinned memory, each direction alone and both at once -- the ceiling of any end-to-end number
that ships bytes both ways (bench.py's e2e.link_frac divides by `both_GBps`).

  python tools/bench_hostlink.py [--mib 1024] [--reps 5]
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/bench_hostlink.py
      (N processes, one per GPU, all copying at once: the node's aggregate)

One JSON line on stdout (rank 0).  --pageable repeats the measurement from plain malloc memory.
"""
import argparse
import json
import os
import time

import numpy as np
import torch


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mib", type=int, default=1024)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--pageable", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    n = args.mib << 20

    def host():
        if args.pageable:
            a = np.empty(n, np.uint8)
            a[...] = 1
            return torch.from_numpy(a)
        return torch.empty(n, dtype=torch.uint8, pin_memory=True)
    hu, hd = host(), host()
    du = torch.empty(n, dtype=torch.uint8, device=dev)
    dd = torch.ones(n, dtype=torch.uint8, device=dev)
    s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

    def run(up, down):
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
            torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(args.reps):
            if up:
                with torch.cuda.stream(s1):
                    du.copy_(hu, non_blocking=True)
            if down:
                with torch.cuda.stream(s2):
                    hd.copy_(dd, non_blocking=True)
        torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if dist:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        return world * args.reps * n * (int(up) + int(down)) / float(dt.item()) / 1e9
    run(True, True)
    line = {"metric": "host link, plain cudaMemcpyAsync", "processes": world,
            "memory": "pageable (malloc)" if args.pageable else "pinned (cudaHostAlloc)",
            "MiB_per_copy": args.mib, "reps": args.reps,
            "h2d_alone_GBps": round(run(True, False), 2), "d2h_alone_GBps": round(run(False, True), 2),
            "both_GBps": round(run(True, True), 2), "unit": "GB/s, whole node (sum over processes)"}
    if rank == 0:
        print(json.dumps(line))
    if dist:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

"""The N>1 host logic on CPU: two processes under the gloo backend shard a batch
of frames, each runs ITS frames (through the CPU checker here -- there is no GPU
in this container; on the B200 box the same shard goes through the CUDA library,
tests/test_gpu_parity.py), and the pieces must tile the single-process answer
with no exchange on the data path."""
import hashlib
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_frame_shard_tiles_the_batch():
    from daala_b200.sharding import frame_owner, frame_shard
    for total in (0, 1, 5, 64, 67):
        for world in (1, 2, 3, 4, 8):
            seen = []
            for r in range(world):
                first, count = frame_shard(total, r, world)
                seen += list(range(first, first + count))
                for f in range(first, first + count):
                    assert frame_owner(f, total, world) == r
            assert seen == list(range(total))
            sizes = [frame_shard(total, r, world)[1] for r in range(world)]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        frame_shard(4, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, total, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    import oracle_libs as ol
    from daala_b200.sharding import frame_shard, reduce_timing
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        pic_w, pic_h = 100, 60
        fw, fh = ol.padded_dims(pic_w, pic_h)
        first, count = frame_shard(total, rank, world)
        digests = []
        for f in range(first, first + count):
            rng = np.random.default_rng(500 + f)           # the frame's content depends on f only
            bsize = ol.random_quadtree_bsize(fw // 32, fh // 32, rng)
            h = hashlib.sha256()
            for dec in (0, 1, 1):
                c = rng.integers(-2048, 2048, size=(fh >> dec, fw >> dec)).astype(np.int32)
                d, _ = ol.port_fwd_plane(c, bsize, dec, pic_w, pic_h, 1)
                assert np.array_equal(ol.port_inv_plane(d, bsize, dec, pic_w, pic_h, 1), c)
                h.update(d.tobytes())
            digests.append((f, h.hexdigest()))
        dist.barrier()
        ms, frames = reduce_timing(dist, 10.0 * (rank + 1), count)
        gathered = [None] * world
        dist.all_gather_object(gathered, digests)       # test bookkeeping, not the data path
        if rank == 0:
            q.put((ms, frames, sum(gathered, [])))
    finally:
        dist.destroy_process_group()


def test_two_ranks_tile_the_single_process_answer():
    import torch.multiprocessing as mp
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_libs as ol
    ol.port()                                           # build once, before forking workers
    ctx = mp.get_context("spawn")
    total, world = 5, 2
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, total, q)) for r in range(world)]
    for p in procs:
        p.start()
    ms, frames, digests = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ms == 20.0 and frames == total               # max over ranks, sum over ranks
    assert [f for f, _ in digests] == list(range(total))
    # single process, all frames
    pic_w, pic_h = 100, 60
    fw, fh = ol.padded_dims(pic_w, pic_h)
    for f, dig in digests:
        rng = np.random.default_rng(500 + f)
        bsize = ol.random_quadtree_bsize(fw // 32, fh // 32, rng)
        h = hashlib.sha256()
        for dec in (0, 1, 1):
            c = rng.integers(-2048, 2048, size=(fh >> dec, fw >> dec)).astype(np.int32)
            h.update(ol.port_fwd_plane(c, bsize, dec, pic_w, pic_h, 1)[0].tobytes())
        assert h.hexdigest() == dig

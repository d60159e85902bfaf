"""Frame sharding over the GPUs of one box.

Frames are independent on this path (reference: one od_state per picture,
src/state.c:355-496; nothing is carried from frame to frame inside the
transform stage), so a batch is split into contiguous runs of frames, one run
per rank, and there is no data-path collective.  torch.distributed is used for
exactly two things: the barrier around a timed region and the max-over-ranks /
sum-over-ranks of the numbers each rank measured.

Nothing here touches CUDA, so the same code runs under the gloo backend in the
CPU tests (tests/test_sharding_gloo.py).
"""


def frame_shard(total_frames, rank, world):
    """(first, count) of the frames rank `rank` of `world` owns: contiguous,
    sizes differ by at most one, earlier ranks take the remainder."""
    if world < 1 or not 0 <= rank < world or total_frames < 0:
        raise ValueError("bad shard request: total=%r rank=%r world=%r" % (total_frames, rank, world))
    base, rem = divmod(total_frames, world)
    count = base + (1 if rank < rem else 0)
    first = rank * base + min(rank, rem)
    return first, count


def frame_owner(frame, total_frames, world):
    """Inverse of frame_shard: the rank that owns `frame`."""
    base, rem = divmod(total_frames, world)
    edge = rem * (base + 1)
    if frame < edge:
        return frame // (base + 1)
    return rem + (frame - edge) // base


def reduce_timing(dist, elapsed_ms, frames_done, device=None):
    """Whole-job numbers from per-rank ones: (max elapsed over ranks, frames summed
    over ranks).  `dist` is torch.distributed, or None for a single process."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return float(elapsed_ms), int(frames_done)
    import torch
    t = torch.tensor([float(elapsed_ms)], dtype=torch.float64, device=device)
    n = torch.tensor([int(frames_done)], dtype=torch.int64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.all_reduce(n, op=dist.ReduceOp.SUM)
    return float(t.item()), int(n.item())


def mpixels_per_second(pic_width, pic_height, frames, elapsed_ms):
    """BASELINE.json's unit: luma picture pixels through the stage(s) per second / 1e6."""
    return pic_width * pic_height * frames / (elapsed_ms * 1e-3) / 1e6

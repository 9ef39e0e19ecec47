"""Frame-parallel sharding of a video stream over the GPUs of one box.

Output frames are independent (the reference re-creates the canvas per frame,
/root/reference/tools/stitch/main.cpp:144), so frame f goes to rank f mod G; every rank holds a
full replica of the calibration and there is no data-path collective.  torch.distributed is
used only to agree on timing (max over ranks) and, for tests and small runs, to put per-frame
results back in stream order on rank 0.
"""


def frames_for_rank(n_frames, rank, world):
    """Indices of the frames rank `rank` stitches: f = rank, rank + world, ..."""
    return list(range(rank, n_frames, world))


def owner_of_frame(f, world):
    return f % world


def max_over_ranks(value, dist=None, device=None):
    """Largest `value` (a float, e.g. elapsed ms) over all ranks; identity without a process group."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return float(value)
    import torch
    t = torch.tensor([float(value)], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def resequence(local_items, n_frames, dist=None):
    """local_items: {frame_index: picklable}.  Returns the list ordered by frame index on rank 0
    (None on other ranks).  Host-side only: meant for digests, timings and small results."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return [local_items[f] for f in range(n_frames)]
    world, rank = dist.get_world_size(), dist.get_rank()
    gathered = [None] * world if rank == 0 else None
    dist.gather_object(local_items, gathered, dst=0)
    if rank != 0:
        return None
    merged = {}
    for r, part in enumerate(gathered):
        for f, item in part.items():
            if owner_of_frame(f, world) != r:
                raise ValueError(f"frame {f} came from rank {r}, owner is {owner_of_frame(f, world)}")
            merged[f] = item
    missing = [f for f in range(n_frames) if f not in merged]
    if missing:
        raise ValueError(f"frames missing after gather: {missing}")
    return [merged[f] for f in range(n_frames)]

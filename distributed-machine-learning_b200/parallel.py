"""Image sharding across the GPUs of one box: one process per GPU, no data-path collective.

The reference spreads a job's batches over 8 worker VMs (/root/reference/worker.py:52,
:270-298); here the 8 workers are the 8 GPUs.  Each rank runs its contiguous slice of the file
list on its own engine; the only collectives are one broadcast of the packed weights at load
and one gather of the (tiny) per-file results.  Works with the NCCL backend on GPUs and with
gloo on the CPU (the host logic is tested that way).
"""
from __future__ import annotations


def shard_range(n_items: int, world: int, rank: int):
    """Contiguous slice [start, end) of rank: sizes differ by at most one, earlier ranks get the extra."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad world/rank")
    base, extra = divmod(n_items, world)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def replicate_weights(packed, src: int = 0):
    """Broadcast rank src's packed weight blob (a uint8 tensor view) to every rank; returns it."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.broadcast(packed, src=src)
    return packed


def run_sharded(predict_fn, files):
    """predict_fn(list_of_files) -> {basename: result} on this rank's slice; returns the merged dict in
    the order of ``files`` on every rank (duplicates: the later file wins, like dict.update)."""
    import torch.distributed as dist
    files = list(files)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return predict_fn(files)
    world, rank = dist.get_world_size(), dist.get_rank()
    a, b = shard_range(len(files), world, rank)
    mine = predict_fn(files[a:b]) if b > a else {}
    parts = [None] * world
    dist.all_gather_object(parts, mine)
    merged = {}
    for part in parts:
        merged.update(part)
    return {f.split("/")[-1]: merged[f.split("/")[-1]] for f in files}

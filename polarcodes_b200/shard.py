"""Frame sharding and counter reduction for N>1 GPUs (one process per GPU).

Frames are independent: rank r decodes the global frame ids
[total*r/world, total*(r+1)/world) — the same split the C++ driver uses for PCB_GPUS — and the
only collective of the path is a SUM all-reduce of {frames, frame_errors, bit_errors}
(NCCL on GPUs, gloo in the CPU tests)."""


def shard_range(total, rank, world):
    return total * rank // world, total * (rank + 1) // world


def reduce_counters(counters, device=None):
    """SUM all-reduce of the three uint64 error counters over the default process group."""
    import torch
    import torch.distributed as dist
    keys = ("frames", "frame_errors", "bit_errors")
    t = torch.tensor([int(counters[k]) for k in keys], dtype=torch.int64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return {k: int(v) for k, v in zip(keys, t.tolist())}

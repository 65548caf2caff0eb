"""Multi-GPU plumbing for bench.py: the image stream shards by frame with no data-path collective;
torch.distributed is used only for the barrier and for the max-over-ranks of the elapsed time."""
import os
import torch
import torch.distributed as dist


def env():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


def shard(n_items, rank, world):
    """Contiguous, balanced [begin, end) slice of n_items for this rank (weak scaling uses n_items per rank instead)."""
    base, rem = divmod(n_items, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


_group = None          # the group barrier() / max_over_ranks() / sum_over_ranks() run on (None: the default group)


def use_group(group):
    """bench.py starts on a gloo default group (CPU-only waits) and switches its timing plumbing to an NCCL group here"""
    global _group
    _group = group


def _backend():
    return dist.get_backend(_group) if _group is not None else dist.get_backend()


def _device():
    return torch.device("cuda", torch.cuda.current_device()) if _backend() == "nccl" else torch.device("cpu")


def barrier():
    if dist.is_available() and dist.is_initialized():
        if _backend() == "nccl":
            dist.barrier(group=_group, device_ids=[torch.cuda.current_device()])
        else:
            dist.barrier(group=_group)
    if torch.cuda.is_available():
        torch.cuda.synchronize()


_cpu_group = None


def cpu_barrier():
    """Barrier that waits on the CPU (a gloo group next to the NCCL one): for the legs in which rank 0 alone drives every
    GPU of the box - an NCCL barrier would leave a spinning kernel on the waiting ranks' GPUs."""
    global _cpu_group
    if not (dist.is_available() and dist.is_initialized()):
        return
    if dist.get_backend() == "gloo":
        dist.barrier(); return
    if _cpu_group is None:
        _cpu_group = dist.new_group(backend="gloo")
    dist.barrier(group=_cpu_group)


def max_over_ranks(x):
    if not (dist.is_available() and dist.is_initialized()):
        return float(x)
    t = torch.tensor([float(x)], dtype=torch.float64, device=_device())
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=_group)
    return float(t.item())


def sum_over_ranks(x):
    if not (dist.is_available() and dist.is_initialized()):
        return float(x)
    t = torch.tensor([float(x)], dtype=torch.float64, device=_device())
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=_group)
    return float(t.item())

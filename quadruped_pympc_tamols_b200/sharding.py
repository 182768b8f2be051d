"""Robot-level sharding of the batched entry points over the GPUs of one box.

The path partitions by independent robot instances -- upstream `simulation/batched_simulations.py:43-58` runs them as
separate OS processes that exchange nothing -- so there is NO collective on the control path: every rank solves its
own contiguous block of robots with `qpmpc_solve_batched` / `tamols_score_device`.  Robot r always draws its noise
from Philox subsequence r (`robot_offset` below is added to the local index), hence the results do not depend on how
many ranks there are.  torch.distributed (nccl on GPUs, gloo in the CPU tests) is only used to collect the small
per-robot outputs on rank 0 when a caller wants them in one place.
"""
import numpy as np


def shard_bounds(num_robots, world_size, rank):
    """[lo, hi) of the robots rank `rank` owns: contiguous blocks, the first `num_robots % world_size` ranks get one more."""
    if not (0 <= rank < world_size):
        raise ValueError(f"rank {rank} outside world of {world_size}")
    base, extra = divmod(int(num_robots), int(world_size))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard(array, world_size, rank, axis=0):
    """The slice of a per-robot array this rank owns."""
    lo, hi = shard_bounds(np.shape(array)[axis], world_size, rank)
    return np.take(array, np.arange(lo, hi), axis=axis)


def gather_to_rank0(local, num_robots, dist=None):
    """Collect per-robot rows (numpy [n_local, ...]) on rank 0 in robot order; other ranks get None.
    `dist` = an initialised torch.distributed module (any backend); None = single process."""
    local = np.ascontiguousarray(local)
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return local
    import torch
    world, rank = dist.get_world_size(), dist.get_rank()
    lo, hi = shard_bounds(num_robots, world, rank)
    assert local.shape[0] == hi - lo, (local.shape, lo, hi)
    backend = dist.get_backend()
    dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    rows = max(shard_bounds(num_robots, world, k)[1] - shard_bounds(num_robots, world, k)[0] for k in range(world))
    pad = np.zeros((rows,) + local.shape[1:], local.dtype)
    pad[:local.shape[0]] = local
    mine = torch.from_numpy(pad).to(dev)
    bufs = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(bufs, mine)
    if rank != 0:
        return None
    out = []
    for k in range(world):
        klo, khi = shard_bounds(num_robots, world, k)
        out.append(bufs[k][:khi - klo].cpu().numpy())
    return np.concatenate(out, axis=0)

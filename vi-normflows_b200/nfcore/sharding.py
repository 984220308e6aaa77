"""Sample sharding across ranks: who owns which rows, and the one exchange per step.

Monte-Carlo samples (or minibatch rows) are independent, parameters are replicated, and
every kernel returns SUMS over its local rows.  A step therefore needs exactly one
all-reduce(sum) of the small result vector [F parts | gradient partials] plus the local row
count; each rank then divides by the global count and applies the same update.
Works with any torch.distributed backend (NCCL on the GPUs, gloo in the CPU tests).
"""
import torch


def shard_bounds(n_rows, rank, world):
    """Contiguous [lo, hi) slice of `n_rows` rows owned by `rank` (sizes differ by at most one)."""
    base, rem = divmod(int(n_rows), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def reduce_sums(sums, n_local, group=None):
    """In-place all-reduce(sum) of `sums` (a tensor of per-rank sums) and of the row count.
    Returns the global row count.  With group=None and no initialised process group it is a no-op."""
    import torch.distributed as dist
    if group is None and not (dist.is_available() and dist.is_initialized()):
        return int(n_local)
    cnt = torch.tensor([float(n_local)], dtype=torch.float64, device=sums.device)
    dist.all_reduce(sums, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(cnt, op=dist.ReduceOp.SUM, group=group)
    return int(round(float(cnt.item())))


def reduce_with_count(sums, n_local, group=None):
    """One all-reduce(sum) that carries the per-rank sums AND this rank's row count (appended to the vector, exact in
    float64): every rank learns the global count of THIS call -- nothing is cached between calls, so callers may change
    the number of rows from one call to the next.  `sums` is overwritten with the global sums (in its own dtype);
    returns the global row count."""
    import torch.distributed as dist
    if group is None and not (dist.is_available() and dist.is_initialized()):
        return int(n_local)
    buf = torch.empty(sums.numel() + 1, dtype=torch.float64, device=sums.device)
    buf[:-1] = sums.reshape(-1)
    buf[-1] = float(n_local)
    dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group)
    sums.reshape(-1).copy_(buf[:-1].to(sums.dtype))
    return int(round(float(buf[-1].item())))

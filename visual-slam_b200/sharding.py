"""Multi-GPU plumbing for the hot path: units (stereo pairs, keyframe-pair jobs) are independent, so
the only "parallel strategy" is a contiguous split by rank and a gather of per-rank results.
No collective runs on the data path (BASELINE.json north_star)."""
import numpy as np


def shard_range(n_units, rank, world):
    """Contiguous block [lo, hi) of rank `rank`; sizes differ by at most one."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    base, rem = divmod(int(n_units), world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def gather_counts(local_counts, dist=None):
    """Result gather: every rank's per-unit counts concatenated in rank order (all ranks get it).
    `dist` is torch.distributed (or None for a single process)."""
    local = np.ascontiguousarray(local_counts, np.int32)
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return local.copy()
    parts = [None] * dist.get_world_size()
    dist.all_gather_object(parts, local)
    return np.concatenate(parts)


def max_over_ranks(value, dist=None):
    """Timing rule: a multi-GPU time is the max over ranks."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return float(value)
    import torch
    t = torch.tensor([float(value)], dtype=torch.float64)
    if dist.get_backend() == "nccl":
        t = t.cuda()
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def gather_shard_parts(part, dist=None):
    """Config 4 with desc2 sharded over ranks: gather every rank's [n1][4] int32 part (16 bytes per query --
    the only exchange of the sharded matcher) into [world][n1][4] on every rank.  With NCCL the parts travel as
    device tensors; with gloo (CPU tests) as host tensors."""
    part = np.ascontiguousarray(part, np.int32)
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return part[None].copy()
    import torch
    t = torch.from_numpy(part)
    if dist.get_backend() == "nccl":
        t = t.cuda()
    out = [torch.empty_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t)
    return torch.stack(out).cpu().numpy()

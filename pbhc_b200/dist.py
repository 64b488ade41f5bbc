"""Env sharding and the tiny cross-rank reductions of the path.  The reference is single-process
(no torch.distributed anywhere); these helpers are what a multi-GPU caller of TrackingStep / compute_returns uses.
Pure torch: works with any backend (NCCL on the GPUs, gloo in the CPU tests)."""
import torch
import torch.distributed as dist


def shard_range(n_total, rank, world):
    """contiguous block partition of env ids: rank r owns [lo, hi); sizes differ by at most one."""
    base, rem = divmod(int(n_total), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def all_reduce_sum_(t, group=None):
    """in-place SUM all-reduce when a process group is up; no-op for a single process."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return t


def moments_to_mean_std(stats3):
    """(count, sum, sum of squares) fp64 -> (mean, unbiased std) exactly as pbhc_adv_normalize uses them
    (csrc/gae.cu; the reference's `advantages.std()` is unbiased, agents/ppo/ppo.py:354)."""
    n, s, ss = (float(v) for v in stats3)
    mean = s / n
    var = (ss - s * mean) / (n - 1.0) if n > 1 else float("nan")
    return mean, max(var, 0.0) ** 0.5


def global_env_means(acc, n_slot):
    """per-step sums over the local envs (PBHC_A_* layout) -> global means after the all-reduce."""
    return acc / acc[n_slot]

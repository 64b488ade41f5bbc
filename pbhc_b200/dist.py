"""Env sharding and the tiny cross-rank reductions of the path.  The reference is single-process
(no torch.distributed anywhere); these helpers are what a multi-GPU caller of TrackingStep / compute_returns uses.

all_reduce_sum_ is the one exchange primitive of the product (TrackingStep.finalize, gae.compute_returns).  For the
fp64 CUDA vectors of the path (<= 256 doubles) on an NCCL group it runs pbhc_peer_allreduce_sum - one single-block kernel
per rank over NVLink peer memory (csrc/peer.cu; buffers and signal pads come from torch symmetric memory) - and falls
back to the library collective when peer memory cannot be set up (PBHC_PEER_ALLREDUCE=0 forces that).  Anything else
(gloo in the CPU tests, other dtypes) goes through torch.distributed."""
import ctypes as C
import os

import torch
import torch.distributed as dist


def shard_range(n_total, rank, world):
    """contiguous block partition of env ids: rank r owns [lo, hi); sizes differ by at most one."""
    base, rem = divmod(int(n_total), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


class PeerAllReduce:
    """pbhc_peer_allreduce_sum on one process group: symmetric buffer + signal pads, the device call counter."""

    def __init__(self, group, device):
        import torch.distributed._symmetric_memory as symm
        from . import _lib
        self._lib = _lib
        self.group = group if group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(self.group), dist.get_rank(self.group)
        self.device = torch.device(device)
        self.buf = symm.empty(2 * self.world * _lib.PEER_MAX_DOUBLES, dtype=torch.float64, device=self.device)
        self.buf.zero_()
        self.handle = symm.rendezvous(self.buf, self.group)
        entries = int(self.handle.signal_pad_size) // 4
        if entries < 2 * self.world:
            raise RuntimeError("signal pad too small")
        self.signal_base = entries - self.world  # the tail of the pad: torch's own barriers use the head
        self.counter = torch.zeros(1, device=self.device, dtype=torch.int32)
        torch.cuda.synchronize(self.device)
        dist.barrier(self.group)  # every rank's buffer is zeroed and mapped before anybody's first call

    def __call__(self, t):
        L = self._lib
        L.check(L.lib().pbhc_peer_allreduce_sum(L.ptr(t, torch.float64), t.numel(), C.c_void_p(self.handle.buffer_ptrs_dev),
                                                C.c_void_p(self.handle.signal_pad_ptrs_dev), L.ptr(self.counter), self.rank, self.world,
                                                self.signal_base, L.stream_ptr(self.device)))
        return t


_peer = {}      # (group id, device index) -> PeerAllReduce, or None when the set-up failed (library collective from then on)
peer_status = {}  # the same keys -> "peer memory" / the reason for the fallback (benchmarks report it)


def _peer_for(t, group):
    if os.environ.get("PBHC_PEER_ALLREDUCE", "1") == "0":
        return None
    g = group if group is not None else dist.group.WORLD
    if dist.get_backend(g) != "nccl":
        return None
    key = (id(g), t.device.index)
    if key not in _peer:
        try:
            _peer[key] = PeerAllReduce(g, t.device)
            peer_status[key] = "peer memory (pbhc_peer_allreduce_sum)"
        except Exception as e:  # noqa: BLE001 - no NVLink peer mapping on this box / torch build: the library collective still works
            _peer[key] = None
            peer_status[key] = f"library collective ({type(e).__name__}: {e})"[:200]
    return _peer[key]


def all_reduce_sum_(t, group=None):
    """in-place SUM all-reduce when a process group is up; no-op for a single process."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        if t.is_cuda and t.dtype == torch.float64 and t.is_contiguous() and t.numel() <= 256:
            peer = _peer_for(t, group)
            if peer is not None:
                return peer(t)
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

"""World-size-2 gloo tests (CPU) of the N>1 host logic: env sharding and the two tiny reductions of the path -
(count, sum, sum of squares) of the advantages and the per-step statistics vector - give the single-process result."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from pbhc_b200.dist import shard_range, all_reduce_sum_, moments_to_mean_std


def test_shard_range_partitions_everything():
    for n in (0, 1, 7, 4096, 65537):
        for w in (1, 2, 3, 8):
            spans = [shard_range(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(0)
        adv = torch.randn(24, 1001, generator=g, dtype=torch.float32)  # the whole job's raw advantages [T, N]
        err = torch.rand(1001, 10, generator=g, dtype=torch.float32)   # per-env adaptive-sigma error terms
        lo, hi = shard_range(1001, rank, world)
        mine = adv[:, lo:hi].double()
        stats = torch.tensor([mine.numel(), mine.sum(), (mine * mine).sum()], dtype=torch.float64)
        all_reduce_sum_(stats)
        acc = torch.cat([err[lo:hi].double().sum(0), torch.tensor([float(hi - lo)], dtype=torch.float64)])
        all_reduce_sum_(acc)
        q.put((rank, stats.tolist(), acc.tolist()))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_two_rank_reductions_match_single_process():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = [q.get(timeout=100) for _ in procs]
    for p in procs:
        p.join(timeout=30)
        assert p.exitcode == 0
    g = torch.Generator().manual_seed(0)
    adv = torch.randn(24, 1001, generator=g, dtype=torch.float32)
    err = torch.rand(1001, 10, generator=g, dtype=torch.float32)
    for _, stats, acc in out:
        mean, std = moments_to_mean_std(stats)
        assert stats[0] == adv.numel()
        assert abs(mean - adv.double().mean().item()) < 1e-12
        assert abs(std - adv.double().std().item()) < 1e-10  # unbiased, like advantages.std() (ppo.py:354)
        means = torch.tensor(acc[:10], dtype=torch.float64) / acc[10]
        assert torch.allclose(means, err.double().mean(0), rtol=1e-12, atol=0)
    assert out[0][1] == out[1][1] and out[0][2] == out[1][2]  # every rank holds the same global statistics

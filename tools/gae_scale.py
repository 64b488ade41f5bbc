#!/usr/bin/env python
"""BASELINE.json configs[4]: PPO rollout post-processing - GAE reverse scan + advantage normalisation over a
24-step x (--envs-total) buffer sharded over the ranks (one NCCL all-reduce of the three fp64 advantage moments).

    python tools/gae_scale.py --envs-total 1048576                       # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 tools/gae_scale.py

Each rank times `compute_returns` (pbhc_gae -> all-reduce -> pbhc_adv_normalize) with CUDA events between barriers,
the max over ranks is taken, rank 0 prints one JSON line per head count R (1 = PPO, 21 = MHPPO with config A's reward
heads).  Buffers are larger than L2 or rotated so that no timed call finds its inputs in L2.  Development aid;
bench.py is the contract."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs-total", type=int, default=1 << 20)
    ap.add_argument("--steps", type=int, default=24)
    ap.add_argument("--iters", type=int, default=30)
    ap.add_argument("--heads", type=int, nargs="*", default=[1, 21])
    a = ap.parse_args()
    from pbhc_b200.dist import shard_range
    from pbhc_b200.gae import compute_returns
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    peak = 6447.8
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    lo, hi = shard_range(a.envs_total, rank, world)
    N, T = hi - lo, a.steps

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for R in a.heads:
        byt = T * R * 12 + T + R * 4 + T * 4 + T * 8  # per env: rewards/values read, returns written, dones, last values, adv written + normalised in place
        n_sets = max(2, int(2 * 126e6 // (byt * N)) + 1)  # rotate input sets so each call's inputs were evicted from the 126-MB L2
        g = torch.Generator(device=dev).manual_seed(100 + rank)
        sets = [(torch.randn(T, N, R, device=dev, generator=g), torch.randn(T, N, R, device=dev, generator=g),
                 torch.rand(T, N, 1, device=dev, generator=g) < 0.02, torch.randn(N, R, device=dev, generator=g)) for _ in range(n_sets)]
        for i in range(3):
            compute_returns(*sets[i % n_sets], 0.99, 0.95)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(a.iters):
            compute_returns(*sets[i % n_sets], 0.99, 0.95)
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1) / a.iters], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        if rank == 0:
            m = float(ms)
            print(json.dumps({"kernel": f"gae_R{R}+allreduce+normalize", "n_gpus": world, "T": T, "envs_total": a.envs_total, "envs_per_gpu": N,
                              "ms": m, "env_rollouts_per_s": a.envs_total / m * 1e3, "bytes_per_env": byt,
                              "gbs_per_gpu": byt * N / m / 1e6, "frac_of_hbm_per_gpu": byt * N / m / 1e6 / peak,
                              "l2": f"{n_sets} input sets rotated"}), flush=True)
        del sets
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        os._exit(0)


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Timeline of the sharded step, per rank (run under torchrun): CUDA events around the pieces of Workload.step -
fused kernel + block reduction | cross-rank exchange of acc (waits for the slowest rank) | finalize kernel, and the K2 FK
branch that runs beside the last two - averaged over --steps eager steps after warm-up.  Rank 0 prints one JSON line:
per-rank microseconds of each piece and of the whole step.  PBHC_PEER_ALLREDUCE=0 times the NCCL exchange instead."""
import argparse
import ctypes as C
import json
import os
import sys
import tempfile

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pbhc_b200 import _lib, dist as pd  # noqa: E402
from benchkit.workload import Workload  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs", type=int, default=65536)
    ap.add_argument("--steps", type=int, default=200)
    a = ap.parse_args()
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    wl = Workload(a.envs, dev, tempfile.mkdtemp(prefix=f"tl{rank}_"), seed=1000 + rank)
    ts = wl.ts
    for _ in range(5):
        wl.step(True)
    torch.cuda.synchronize()
    cur, side = torch.cuda.current_stream(dev), torch.cuda.Stream(dev)
    L = _lib.lib()
    names = ("track+reduce", "exchange", "finalize", "fk_branch", "step")
    acc = [0.0] * len(names)
    if world > 1:
        dist.barrier()
    for _ in range(a.steps):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
        ev[0].record(cur)
        ts.launch()
        ev[1].record(cur)
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            wl.fk()
            ev[5].record(side)
        pd.all_reduce_sum_(ts.acc)
        ev[2].record(cur)
        _lib.check(L.pbhc_track_finalize(C.byref(ts._cfg), _lib.ptr(ts.acc), _lib.ptr(ts.scalars), _lib.stream_ptr(dev)))
        ev[3].record(cur)
        cur.wait_stream(side)
        ev[4].record(cur)
        torch.cuda.synchronize()
        for i, (x, y) in enumerate(((0, 1), (1, 2), (2, 3), (1, 5), (0, 4))):
            acc[i] += ev[x].elapsed_time(ev[y]) * 1e3 / a.steps
    mine = torch.tensor(acc, device=dev, dtype=torch.float64)
    allv = [torch.empty_like(mine) for _ in range(world)]
    if world > 1:
        dist.all_gather(allv, mine)
    else:
        allv = [mine]
    if rank == 0:
        print(json.dumps({"what": "eager step pieces, microseconds, mean over %d steps (CUDA events; the exchange includes waiting for the slowest rank's kernel)" % a.steps,
                          "world": world, "envs_per_gpu": a.envs, "exchange": list(pd.peer_status.values()) or ["none / single process"],
                          "per_rank_us": {n: [round(float(v[i]), 1) for v in allv] for i, n in enumerate(names)}}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

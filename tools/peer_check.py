#!/usr/bin/env python
"""Multi-GPU check of the path's cross-rank exchange (run under torchrun, one rank per GPU):
  * pbhc_peer_allreduce_sum == dist.all_reduce (NCCL) bit for bit on random fp64 vectors, eager and replayed from a CUDA graph,
  * a sharded TrackingStep.finalize (each rank its block of envs) gives every rank the scalars of the single-process step,
  * timing of the 64-double exchange: peer kernel vs NCCL (CUDA events, graph replay).
Prints one JSON line on rank 0; exits non-zero on a mismatch."""
import json
import os
import sys
import tempfile

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pbhc_b200 import dist as pd  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    out = {"world": world}
    gen = torch.Generator(device="cpu").manual_seed(100 + rank)
    # ---- 1. bit-exactness against NCCL's result is not required (summation order may differ); compare with the rank-ordered fp64 sum
    for n in (3, 64, 256):
        for it in range(20):
            mine = torch.randn(n, generator=gen, dtype=torch.float64)
            allv = [torch.empty(n, dtype=torch.float64, device=dev) for _ in range(world)]
            x = mine.to(dev)
            dist.all_gather(allv, x)
            want = torch.zeros(n, dtype=torch.float64, device=dev)
            for v in allv:  # rank order
                want = want + v
            got = pd.all_reduce_sum_(x.clone())
            torch.cuda.synchronize()
            if not torch.equal(got, want):
                print(f"rank {rank}: peer all-reduce mismatch n={n} it={it}: max |d| = {(got - want).abs().max().item():.3e}", flush=True)
                sys.exit(1)
    out["status"] = list(pd.peer_status.values())
    # ---- 2. graph capture + replay
    x = torch.ones(64, dtype=torch.float64, device=dev) * (rank + 1)
    y = x.clone()
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        y.copy_(x); pd.all_reduce_sum_(y)
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    with torch.cuda.graph(g):
        y.copy_(x)
        pd.all_reduce_sum_(y)
    for _ in range(50):
        g.replay()
    torch.cuda.synchronize()
    want = float(sum(range(1, world + 1)))
    if not bool((y == want).all()):
        print(f"rank {rank}: graph replay mismatch {y[:4].tolist()} != {want}", flush=True)
        sys.exit(1)
    # ---- 3. timing: peer kernel vs NCCL, 64 doubles, graph replay of 100 exchanges
    def timed(fn):
        z = torch.ones(64, dtype=torch.float64, device=dev)
        for _ in range(5):
            fn(z)
        torch.cuda.synchronize()
        gg = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gg):
            for _ in range(100):
                fn(z)
        gg.replay()
        dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            gg.replay()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / 500.0], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item() * 1e3
    out["peer_us"] = timed(lambda z: pd.all_reduce_sum_(z))
    out["nccl_us"] = timed(lambda z: dist.all_reduce(z))
    # ---- 4. sharded finalize == single-process finalize
    from benchkit.workload import Workload
    N = 64 * world
    full = Workload(N, dev, tempfile.mkdtemp(prefix=f"pc_full{rank}_"), seed=5)
    lo, hi = pd.shard_range(N, rank, world)
    part = Workload(hi - lo, dev, tempfile.mkdtemp(prefix=f"pc_part{rank}_"), seed=5)
    # the shard holds rows [lo, hi) of the full workload's state
    for name in ("_rigid_body_state", "all_root_states", "dof_state", "contact_forces", "_base_com_bias", "_link_mass_scale", "friction_coeffs"):
        src = getattr(full.sim, name)
        getattr(part.sim, name).copy_(src.view(N, -1)[lo:hi].reshape(getattr(part.sim, name).shape))
    for a, b in ((part.actions, full.actions), (part.torques, full.torques), (part.env_origins, full.env_origins)):
        a.copy_(b[lo:hi])
    for a, b in ((part.ts.episode_length_buf, full.ts.episode_length_buf), (part.ts.env_f32, full.ts.env_f32), (part.ts._kp_scale, full.ts._kp_scale),
                 (part.ts._kd_scale, full.ts._kd_scale), (part.ts.last_actions, full.ts.last_actions), (part.ts.last_dof_vel, full.ts.last_dof_vel),
                 (part.ts.last_dof_pos, full.ts.last_dof_pos)):
        a.copy_(b[lo:hi])
    for _ in range(3):
        full.ts.bind(full.sim, full.actions, full.torques, full.env_origins, full.motion_ids); full.ts.launch()
        from pbhc_b200 import _lib
        _lib.check(_lib.lib().pbhc_track_finalize(__import__("ctypes").byref(full.ts._cfg), _lib.ptr(full.ts.acc), _lib.ptr(full.ts.scalars), _lib.stream_ptr(dev)))
        part.ts.bind(part.sim, part.actions, part.torques, part.env_origins, part.motion_ids)
        part.ts._io.env_offset = lo  # the in-kernel reset draws are keyed by the global env id
        part.ts.launch()
        part.ts.finalize()  # all-reduce of acc over the ranks, then pbhc_track_finalize
        torch.cuda.synchronize()
        a, b = part.ts.scalars, full.ts.scalars
        bad = ~(torch.isclose(a, b, rtol=1e-12, atol=1e-300) | (a.isnan() & b.isnan()))
        if bad.any():
            i = int(bad.nonzero()[0])
            print(f"rank {rank}: sharded finalize differs at scalar {i}: {a[i].item()!r} vs {b[i].item()!r}", flush=True)
            sys.exit(1)
    out["sharded_finalize"] = "equal to the single-process step (rtol 1e-12) on every rank"
    dist.barrier()
    if rank == 0:
        print(json.dumps(out), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

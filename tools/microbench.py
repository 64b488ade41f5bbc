#!/usr/bin/env python
"""Per-kernel timing on one GPU (CUDA events, L2 flushed by working sets larger than L2):
python tools/microbench.py [--envs 65536] [--iters 20].  Prints one JSON line per kernel with achieved
algorithmic GB/s (SURVEY.md 8d byte figures).  Development aid; bench.py is the contract."""
import argparse
import json
import os
import sys
import tempfile

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def timeit(fn, iters, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(iters + 1)]
    ev[0].record()
    for i in range(iters):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    ts = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(iters))
    return ts[len(ts) // 2]


def timeit_graph(fn, iters, warmup=3):
    """the same call captured once in a CUDA graph and replayed: device time without the host-side call overhead"""
    st = torch.cuda.Stream()
    st.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(st):
        for _ in range(warmup):
            fn()
    torch.cuda.current_stream().wait_stream(st)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        keep = fn()  # noqa: F841  (outputs stay alive in the graph's pool)
    return timeit(g.replay, iters, warmup)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs", type=int, default=65536)
    ap.add_argument("--iters", type=int, default=20)
    a = ap.parse_args()
    from tests._util import product_cfg
    from pbhc_b200.motion_lib import MotionLibRobot
    from pbhc_b200.gae import compute_returns
    dev = "cuda:0"
    n = a.envs
    peak = 6447.8
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    for tag, names in (("23", ["Horse-stance_pose", "Bruce_Lee_pose", "Charleston_dance", "Hooks_punch",
                               "Horse-stance_punch", "Roundhouse_kick", "Side_kick"]), ("29", ["g1_29dof_clip"])):
        ml = MotionLibRobot(product_cfg(tag, tempfile.mkdtemp(), names), n, dev)
        ml.load_motions(random_sample=False)
        J, Jb, D, Cm = ml._dims
        ids = torch.randint(0, n, (n,), device=dev)
        t = torch.rand(n, device=dev) * ml._motion_lengths[ids]
        off = torch.randn(n, 3, device=dev)
        ms = timeit(lambda: ml.get_motion_state(ids, t, off), a.iters)
        rec = 13 * J + 2 * D + Cm
        outf = 3 + 4 + D + 3 + 3 + D + 3 * Jb + 17 + 13 * Jb + 13 * J + Cm
        byt = 24 + 2 * rec * 4 + outf * 4
        print(json.dumps(dict(kernel=f"lookup_{tag}", n=n, ms=ms, bytes_per_unit=byt, gbs=byt * n / ms / 1e6,
                              frac=byt * n / ms / 1e6 / peak)))
        ms = timeit_graph(lambda: ml.get_motion_state(ids, t, off), a.iters)
        print(json.dumps(dict(kernel=f"lookup_{tag}_graph_replay", n=n, ms=ms, bytes_per_unit=byt, gbs=byt * n / ms / 1e6,
                              frac=byt * n / ms / 1e6 / peak)))
        pose = torch.randn(n, J, 3, device=dev) * 0.5
        trans = torch.randn(n, 3, device=dev)
        hb = ml.mesh_parsers
        ms = timeit(lambda: hb.fk(pose, trans), a.iters)
        byt = J * 12 + 12 + J * (12 + 16 + 36)
        print(json.dumps(dict(kernel=f"fk_{tag}", n=n, ms=ms, bytes_per_unit=byt, gbs=byt * n / ms / 1e6,
                              frac=byt * n / ms / 1e6 / peak)))
        ms = timeit_graph(lambda: hb.fk(pose, trans), a.iters)
        print(json.dumps(dict(kernel=f"fk_{tag}_graph_replay", n=n, ms=ms, bytes_per_unit=byt, gbs=byt * n / ms / 1e6,
                              frac=byt * n / ms / 1e6 / peak)))
    for R, N in ((1, n * 2), (1, n * 4), (1, n * 16), (21, n)):
        T = 24
        v, r = torch.randn(T, N, R, device=dev), torch.randn(T, N, R, device=dev)
        d = torch.rand(T, N, 1, device=dev) < 0.02
        lv = torch.randn(N, R, device=dev)
        ms = timeit(lambda: compute_returns(v, r, d, lv, 0.99, 0.95), a.iters)
        byt = T * R * 12 + T + R * 4 + T * 4 + T * 8
        print(json.dumps(dict(kernel=f"gae_R{R}", n=N, ms=ms, bytes_per_unit=byt, gbs=byt * N / ms / 1e6,
                              frac=byt * N / ms / 1e6 / peak)))
        ms = timeit_graph(lambda: compute_returns(v, r, d, lv, 0.99, 0.95), a.iters)
        print(json.dumps(dict(kernel=f"gae_R{R}_graph_replay", n=N, ms=ms, bytes_per_unit=byt, gbs=byt * N / ms / 1e6,
                              frac=byt * N / ms / 1e6 / peak)))
        del v, r, d, lv
    # RolloutStorage shuffle-gather: the keys PPO._setup_storage registers (ppo.py:124-135) at config A's
    # widths, T = 24 steps; bytes = every stored row read once + written once + the int64 permutation.  The eager-torch
    # line is the reference expression (one index_select + contiguous per key) on the same buffers.
    from pbhc_b200.rollout import RolloutStorage
    T = 24
    for N in sorted({4096, min(n, 32768)}):
        st = RolloutStorage(N, T, device=dev)
        dims = dict(actor_obs=380, critic_obs=463, actions=23, rewards=1, dones=1, values=1, returns=1, advantages=1,
                    actions_log_prob=1, action_mean=23, action_sigma=23)
        for k, w in dims.items():
            st.register_key(k, shape=(w,), dtype=torch.bool if k == "dones" else torch.float)
            if k != "dones":
                st.batch_update_data(k, torch.randn(T, N, w, device=dev))
        row = sum(getattr(st, k)[0, 0].numel() * getattr(st, k).element_size() for k in st.stored_keys)
        byt = 2 * row + 8
        idx = torch.randperm(T * N, device=dev)
        out = st.gather(idx)
        ms = timeit(lambda: st.gather(idx, out=out, check=False), a.iters)
        print(json.dumps(dict(kernel="rollout_gather", n=T * N, ms=ms, bytes_per_unit=byt, gbs=byt * T * N / ms / 1e6,
                              frac=byt * T * N / ms / 1e6 / peak, launches=1)))
        ms = timeit(lambda: {k: getattr(st, k)[:].flatten(0, 1)[idx].contiguous() for k in st.stored_keys}, a.iters)
        print(json.dumps(dict(kernel="rollout_gather_eager_torch", n=T * N, ms=ms, bytes_per_unit=byt,
                              gbs=byt * T * N / ms / 1e6, frac=byt * T * N / ms / 1e6 / peak, launches=len(dims))))
        del st, out


if __name__ == "__main__":
    main()

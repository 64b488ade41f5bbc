#!/usr/bin/env python
"""Generates tests/golden/ref_rollout.npz by running the UNMODIFIED reference RolloutStorage
(/root/reference/humanoidverse/agents/modules/data_utils.py) on CPU and the time-out bootstrap expression of
agents/ppo/ppo.py:266-270 in eager torch.  Test infrastructure; run in the build container from the repo root:

    python -m tools.make_golden_rollout

Stored: the per-step inputs of every key, the permutation the reference drew (recovered through a `row_id` key
holding arange(T*N)), every yielded minibatch (epoch-major) and the bootstrap inputs/outputs.
"""
import importlib.util
import os

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("PBHC_REFERENCE_ROOT", "/root/reference")
GOLD = os.path.join(ROOT, "tests", "golden")

# key -> (shape, dtype name): the keys PPO._setup_storage registers (ppo.py:124-135) at toy widths, plus odd row
# widths (1, 3, 5, 11 elements; bool rows) so that every vector width of the gather kernel is exercised
KEYS = {"actor_obs": ((11,), "float"), "critic_obs": ((16,), "float"), "actions": ((3,), "float"), "rewards": ((1,), "float"),
        "dones": ((1,), "bool"), "values": ((1,), "float"), "action_mean": ((5,), "float"), "row_id": ((1,), "long"),
        "flags3": ((3,), "bool"), "grid": ((2, 4), "float")}
T, N, MB, EPOCHS = 5, 7, 3, 2


def main():
    spec = importlib.util.spec_from_file_location("ref_data_utils", os.path.join(REF, "humanoidverse/agents/modules/data_utils.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    torch.manual_seed(20240611)
    st = mod.RolloutStorage(N, T, device="cpu")
    tdt = {"float": torch.float, "bool": torch.bool, "long": torch.long}
    for k, (shape, dt) in KEYS.items():
        st.register_key(k, shape=shape, dtype=tdt[dt])
    out = {}
    g = torch.Generator().manual_seed(7)
    for step in range(T):
        for k, (shape, dt) in KEYS.items():
            if k == "row_id":
                data = (torch.arange(N) + step * N).reshape(N, 1)
            elif dt == "bool":
                data = torch.rand((N,) + shape, generator=g) < 0.4
            else:
                data = torch.randn((N,) + shape, generator=g)
            st.update_key(k, data)
            out.setdefault(f"in_{k}", []).append(data.numpy())
        st.increment_step()
    ret = torch.randn(T, N, 1, generator=g)
    st.register_key("returns", shape=(1,), dtype=torch.float)
    st.batch_update_data("returns", ret)
    out = {k: np.stack(v) for k, v in out.items()}
    out["in_returns"] = ret.numpy()
    batches = list(st.mini_batch_generator(MB, EPOCHS))
    assert len(batches) == MB * EPOCHS
    for k in st.stored_keys:
        out[f"mb_{k}"] = np.stack([b[k].numpy() for b in batches])
    perm = torch.cat([b["row_id"] for b in batches[:MB]]).reshape(-1).numpy()
    out["perm_head"] = perm  # the first MB * (T*N // MB) entries of the permutation (the rest is never yielded)
    # time-out bootstrap (ppo.py:266-270), scalar and vector (MHPPO-like R=3) value heads
    for tag, R in (("r1", 1), ("r3", 3)):
        rewards = torch.randn(64, generator=g) if R == 1 else torch.randn(64, R, generator=g)
        values = torch.randn(64, R, generator=g) * 3
        time_outs = torch.rand(64, generator=g) < 0.3
        gamma = 0.99
        rs = rewards.clone().unsqueeze(1) if R == 1 else rewards.clone()
        rs += gamma * values * time_outs.unsqueeze(1)
        out[f"bs_{tag}_rewards"], out[f"bs_{tag}_values"] = rewards.numpy(), values.numpy()
        out[f"bs_{tag}_time_outs"], out[f"bs_{tag}_out"] = time_outs.numpy(), rs.numpy()
    out["meta"] = np.array([T, N, MB, EPOCHS], np.int64)
    np.savez_compressed(os.path.join(GOLD, "ref_rollout.npz"), **out)
    print("wrote ref_rollout.npz", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()

"""env / gae stages of tools/make_golden.py: run the UNMODIFIED reference classes on seeded inputs."""
import os
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


def _np(x):
    return x.detach().cpu().numpy() if torch.is_tensor(x) else np.asarray(x)


def gae_inputs(T, N, R, seed):
    g = torch.Generator().manual_seed(seed)
    values = torch.randn(T, N, R, generator=g)
    rewards = torch.randn(T, N, R, generator=g)
    dones = torch.rand(T, N, 1, generator=g) < 0.1
    dones[-1, ::7] = True   # dones at the last step
    dones[0, ::5] = True    # and at the first
    dones[:, 3] = True      # an env that is done at every position
    last_values = torch.randn(N, R, generator=g)
    return values, rewards, dones, last_values


def stage_gae(ns):
    from humanoidverse.agents.ppo.ppo import PPO
    from humanoidverse.agents.mh_ppo.mh_ppo import MHPPO
    for tag, cls, (T, N, R) in (("ppo", PPO, (24, 96, 1)), ("mhppo", MHPPO, (24, 64, 21))):
        values, rewards, dones, last_values = gae_inputs(T, N, R, 77)
        stub = types.SimpleNamespace(device="cpu", gamma=0.99, lam=0.95,
                                     critic=types.SimpleNamespace(evaluate=lambda obs: last_values))
        ret, adv = cls._compute_returns(stub, {"critic_obs": None},
                                        {"values": values.clone(), "dones": dones.clone(), "rewards": rewards.clone()})
        np.savez_compressed(os.path.join(GOLD, f"ref_gae_{tag}.npz"), values=_np(values), rewards=_np(rewards),
                            dones=_np(dones), last_values=_np(last_values), returns=_np(ret), advantages=_np(adv),
                            gamma=0.99, lam=0.95)


def run(ns, stages):
    if "gae" in stages:
        stage_gae(ns)
    if "env" in stages:
        from tools import make_golden_env_step
        make_golden_env_step.run(ns)

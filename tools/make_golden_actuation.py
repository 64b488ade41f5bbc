#!/usr/bin/env python
"""Generates tests/golden/ref_actuation.npz by driving the UNMODIFIED reference env's _pre_physics_step and
_compute_torques (humanoidverse/envs/legged_base_task/legged_robot_base.py:268-286, 795-838) on CPU, with the stub
simulator of oracle/ref_env.py.  Test infrastructure; run in the build container:  python -m tools.make_golden_actuation

torch.rand_like inside _compute_torques (torque RFI) is captured by re-seeding torch before the call and drawing the
same shape again afterwards; the values are stored as `rfi_u` so that a kernel fed with them must match bit for bit.
"""
import copy
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_loader as R  # noqa: E402
from oracle import ref_env as E  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def main(N=40, steps=5, seed=11):
    ov = {"domain_rand.push_robots": False, "robot.control.action_clip_value": 1.5}
    cfg = E.build_config(N, R.example_motion("Horse-stance_pose"), obs="motion_tracking/main", overrides=ov)
    cj = copy.deepcopy(cfg)
    cj["robot"]["motion"]["motion_file"] = "Horse-stance_pose"
    cj["robot"]["motion"]["asset"]["assetRoot"] = "description/robots/g1/"
    env, ns = E.make_env(cfg, env_cls="motion_tracking", seed=seed)
    sim = env.simulator
    gen = torch.Generator().manual_seed(seed)
    D = env.num_dof
    dr = cfg["domain_rand"]
    env._kp_scale[:] = 0.9 + 0.2 * torch.rand(N, D, generator=gen)
    env._kd_scale[:] = 0.9 + 0.2 * torch.rand(N, D, generator=gen)
    env._rfi_lim_scale[:] = 0.5 + torch.rand(N, D, generator=gen)
    env._rao_scale[:] = (torch.rand(N, D, generator=gen) * 2 - 1) * dr["rao_lim"]
    env.action_delay_idx[:] = torch.randint(dr["ctrl_delay_step_range"][0], dr["ctrl_delay_step_range"][1] + 1, (N,), generator=gen)
    out = {"meta_config_json": json.dumps(cj), "meta": np.array([N, D, steps, env.action_queue.shape[1]], np.int64),
           "kp_scale": env._kp_scale.numpy().copy(), "kd_scale": env._kd_scale.numpy().copy(),
           "rfi_lim_scale": env._rfi_lim_scale.numpy().copy(), "rao_scale": env._rao_scale.numpy().copy(),
           "action_delay_idx": env.action_delay_idx.numpy().copy(), "default_dof_pos": env.default_dof_pos.numpy().copy(),
           "p_gains": env.p_gains.numpy().copy(), "d_gains": env.d_gains.numpy().copy(), "torque_limits": env.torque_limits.numpy().copy()}
    for s in range(steps):
        actions = torch.randn(N, D, generator=gen) * 1.2
        actions[s % N, :3] = torch.tensor([1.5, -1.5, 7.0])  # exactly at / beyond the clip value
        env._pre_physics_step(actions.clone())
        p = f"s{s}_"
        out.update({p + "actions_in": actions.numpy(), p + "actions": env.actions.numpy().copy(),
                    p + "action_queue": env.action_queue.numpy().copy(), p + "actions_after_delay": env.actions_after_delay.numpy().copy(),
                    p + "action_clip_frac": np.float64(env.log_dict["action_clip_frac"])})
        for k in range(cfg["simulator"]["config"]["sim"]["control_decimation"]):
            dof = torch.stack([torch.randn(N, D, generator=gen) * 0.4, torch.randn(N, D, generator=gen) * 3.0], dim=-1)
            sim.dof_state.copy_(dof.reshape(-1, 2))
            torch.manual_seed(1000 * s + k)
            tq = env._compute_torques(env.actions_after_delay)
            torch.manual_seed(1000 * s + k)
            u = torch.rand_like(tq)
            out.update({p + f"k{k}_dof_state": dof.numpy(), p + f"k{k}_rfi_u": u.numpy(), p + f"k{k}_torques": tq.numpy().copy()})
    np.savez_compressed(os.path.join(GOLD, "ref_actuation.npz"), **out)
    clipped = float(np.mean([np.abs(out[f"s{s}_k0_torques"]) == out["torque_limits"] for s in range(steps)]))
    print("wrote ref_actuation.npz", len(out), "arrays; control", cfg["robot"]["control"]["control_type"], "rfi", dr["randomize_torque_rfi"],
          "rao", dr["use_rao"], "clip", cfg["robot"]["control"]["clip_torques"], "frac torques at limit", clipped,
          "clip_frac", [float(out[f"s{s}_action_clip_frac"]) for s in range(steps)])


if __name__ == "__main__":
    main()

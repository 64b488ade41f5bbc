"""Pins oracle/pbhc_oracle_env.py (the CPU restatement of the motion_tracking env step) against golden
fixtures produced by the UNMODIFIED reference env (tools/make_golden_env_step.py).  CPU only."""
import json

import numpy as np
import pytest
import torch

from oracle import pbhc_oracle as O, pbhc_oracle_env as OE
from pbhc_b200.spec import TrackSpec
from tests._util import gold, oracle_clip

INJECT = ("motion_start_times", "kp_scale", "kd_scale", "rfi_lim_scale", "rao_scale", "action_delay_idx")
SCALARS = ("average_episode_length", "reward_penalty_scale", "motion_far_threshold", "soft_dof_pos", "soft_dof_vel", "soft_torque",
           "current_noise_curriculum_value")


def load_case(name):
    g = gold(f"ref_env_{name}.npz")
    spec = TrackSpec.from_config(json.loads(str(g["meta_config_json"])))
    N = int(g["meta_n_envs"])
    from benchkit.fixtures import clips_of_case
    tag, names = clips_of_case(json.loads(str(g["meta_config_json"])))
    clips = [oracle_clip(tag, nm) for nm in names]
    tab = O.MotionTable(clips, spec.Jb, motion_to_clip=[i % len(clips) for i in range(N)])  # load_motions(random_sample=False)
    return g, spec, N, tab


def state_from(g, prefix, spec):
    S = {}
    for k in g:
        if k.startswith(prefix):
            v = g[k]
            S[k[len(prefix):]] = torch.from_numpy(v) if v.ndim else float(v)
    for k in ("dof_state", "root_state"):
        S.pop(k, None)
    return S


@pytest.mark.parametrize("case", ["mt23_main", "mt23_bench", "gt29_teacher", "gt29_student", "mt23_edge", "gt29_edge", "mt23_term", "gt29_term", "mt23_p4", "gt29_p4", "mt23_push", "gt29_push", "mt23_eval", "gt29_eval"])
def test_env_step_oracle_matches_reference(case):
    g, spec, N, tab = load_case(case)
    S = state_from(g, "init_", spec)
    for step in range(int(g["meta_n_steps"])):
        p = f"s{step}_"
        inp = {k: torch.from_numpy(g[p + "in_" + k].copy()) for k in
               ("body_state", "root_state", "dof_state", "contact_forces", "actions", "torques")}
        inject = {k: torch.from_numpy(g[p + "out_" + k]) for k in INJECT}
        if case.endswith("_eval"):  # set_is_evaluating(): resets start at time 0 by themselves (motion_tracking.py:374-375)
            S["is_evaluating"] = True
            del inject["motion_start_times"]
        if p + "in_term_gates" in g:  # the reference's `torch.rand(1) < p` draws of this step (legged_robot_base.py:455-473)
            inject["term_gates"] = int(g[p + "in_term_gates"])
        if p + "in_push_robot_counter" in g:  # the fixture moves a few envs next to their push before every step
            S["push_robot_counter"] = torch.from_numpy(g[p + "in_push_robot_counter"].copy())
        if p + "out_push_interval_s" in g:  # the reference's torch.randint / torch_rand_float draws of the envs it pushed (legged_robot_base.py:387, 1097)
            inject["push_interval_s"] = torch.from_numpy(g[p + "out_push_interval_s"])
            inject["push_vel"] = torch.from_numpy(g[p + "out_push_robot_vel_buf"])
        obs, log, extras = OE.step(spec, tab, S, inp, inject)
        for gname, v in obs.items():
            assert np.array_equal(v.numpy(), g[p + "obs_" + gname]), (step, gname)
        for k in g:
            if not k.startswith(p + "out_"):
                continue
            name = k[len(p) + 4:]
            if name == "dof_state":
                assert np.array_equal(inp["dof_state"].reshape(-1, 2).numpy(), g[k]), (step, name)
            elif name == "root_state":
                assert np.array_equal(inp["root_state"].numpy(), g[k]), (step, name)
            elif name in SCALARS or name.startswith("sigma_") or name.startswith("ema_"):
                assert float(S[name]) == float(g[k]), (step, name, float(S[name]), float(g[k]))
            else:
                mine = S[name].numpy() if name != "last_contacts" else S[name].float().numpy()
                assert np.array_equal(mine.astype(g[k].dtype), g[k]), (step, name)
        for k, v in log.items():
            assert np.float64(np.float32(float(v))) == np.float64(np.float32(g[p + "log_" + k])), (step, k, float(v), float(g[p + "log_" + k]))
        assert np.array_equal(extras["ref_body_pos_extend"].numpy(), g[p + "extras_ref_body_pos_extend"])
        assert np.array_equal(extras["ref_body_rot_extend"].numpy(), g[p + "extras_ref_body_rot_extend"])

"""env stage of tools/make_golden.py: drive the UNMODIFIED reference LeggedRobotMotionTracking (on
oracle/ref_env.FakeSim) for a few steps on seeded synthetic simulator state and dump, per step, every
input of `_post_physics_step` and every output / mutated buffer -> tests/golden/ref_env_<case>.npz.
"""
import os

import numpy as np
import torch

from oracle import ref_env as E, ref_loader as R
from benchkit import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


def _np(x):
    if torch.is_tensor(x):
        return x.detach().cpu().numpy().copy()
    return np.asarray(x)


def env_state(env):
    """every persistent per-env / scalar buffer the hot path reads or writes."""
    s = dict(
        episode_length_buf=env.episode_length_buf, motion_start_times=env.motion_start_times, motion_len=env.motion_len,
        motion_ids=env.motion_ids, actions=env.actions, last_actions=env.last_actions, last_dof_pos=env.last_dof_pos,
        last_dof_vel=env.last_dof_vel, last_contacts=env.last_contacts.float(), feet_air_time=env.feet_air_time,
        kp_scale=env._kp_scale, kd_scale=env._kd_scale, rfi_lim_scale=env._rfi_lim_scale, rao_scale=env._rao_scale,
        action_delay_idx=env.action_delay_idx, default_dof_pos=env.default_dof_pos,
        base_com_bias=env.simulator._base_com_bias, link_mass_scale=env.simulator._link_mass_scale,
        friction_coeffs=env.simulator.friction_coeffs, env_origins=env.env_origins,
        average_episode_length=np.float64(env.average_episode_length), reward_penalty_scale=np.float64(getattr(env, "reward_penalty_scale", 1.0)),
        motion_far_threshold=np.float64(env.terminate_when_motion_far_threshold),
        soft_dof_pos=np.float64(getattr(env, "soft_dof_pos_curriculum_value", 1.0)), soft_dof_vel=np.float64(getattr(env, "soft_dof_vel_curriculum_value", 1.0)),
        soft_torque=np.float64(getattr(env, "soft_torque_curriculum_value", 1.0)),
        current_noise_curriculum_value=np.float64(env.current_noise_curriculum_value), rew_buf=env.rew_buf, reset_buf=env.reset_buf,
        time_out_buf=env.time_out_buf, dof_state=env.simulator.dof_state, root_state=env.simulator.all_root_states,
        end_time_ratio_buf=env.end_time_ratio_buf)
    if hasattr(env.simulator, "_base_mass_scale"):
        s["base_mass_scale"] = env.simulator._base_mass_scale
    if env.config.domain_rand.push_robots:  # legged_robot_base.py:135-147, 1090-1106
        s.update(push_robot_counter=env.push_robot_counter, push_interval_s=env.push_interval_s, push_robot_vel_buf=env.push_robot_vel_buf)
    for k, v in env.history_handler.history.items():
        s["hist_" + k] = v
    for k, v in env.episode_sums.items():
        s["epsum_" + k] = v
    for k, v in env.config.rewards.reward_tracking_sigma.items():
        s["sigma_" + k] = np.float64(v)
    for k, v in getattr(env, "_reward_error_ema", {}).items():
        s["ema_" + k] = np.float64(v)
    return {k: _np(v) for k, v in s.items()}


def _library29(tmp="/tmp/pbhc_golden"):
    """the 3-clip 29-DoF library of benchkit.fixtures.CLIPS29 as a joblib file the reference loads."""
    import joblib
    from benchkit.fixtures import CLIPS29, motion_npz
    os.makedirs(tmp, exist_ok=True)
    data = {}
    for nm in CLIPS29:
        m = motion_npz(nm)
        data[nm] = dict(root_trans_offset=m["root_trans_offset"], pose_aa=m["pose_aa"], dof=m["dof"], root_rot=m["root_rot"], fps=int(m["fps"]))
    path = os.path.join(tmp, "motions29_3.pkl")
    joblib.dump(data, path)
    return path


def run_case(name, n_envs, n_steps, obs, seed, overrides=None, general=False, motion_file=None, motion_name=None, start_times=None, tweak=None,
             tweak_env=None, evaluating=False):
    """motion_file / motion_name: a joblib motion file instead of the default clip(s) and the fixture name recorded for it;
    start_times(env) -> tensor: motion start times instead of the seeded spread; tweak(step, body, root, dof, cf, ref): edits
    the synthetic simulator state in place (edge-case fixtures, tools/make_golden_edge.py); tweak_env(step, env): edits env
    tensors (torques) before the step.  The `torch.rand(1)` draws of the step (the gates of the dof-limit terminations,
    legged_robot_base.py:455-473) are recorded as s<k>_in_term_gates (bit i = i-th enabled check applies)."""
    ov = {"domain_rand.push_robots": False}
    ov.update(overrides or {})
    if general:
        cfg = E.build_config(n_envs, motion_file or _library29(), robot="g1/g1_29dof_general", rewards="motion_tracking/general_main", obs=obs,
                             env="general_tracking", overrides=ov)
    else:
        cfg = E.build_config(n_envs, motion_file or R.example_motion("Horse-stance_pose"), obs=obs, overrides=ov)
    import copy, json
    cj = copy.deepcopy(cfg)  # before the env mutates it (reward_scales *= dt, adaptive sigmas)
    cj["robot"]["motion"]["motion_file"] = motion_name or ("clips29" if general else "Horse-stance_pose")
    cj["robot"]["motion"]["asset"]["assetRoot"] = "description/robots/g1/"
    cj["env_class"] = "general_tracking" if general else "motion_tracking"
    cfg_json = json.dumps(cj)
    env, ns = E.make_env(cfg, env_cls="general" if general else "motion_tracking", seed=seed)
    sim = env.simulator
    if evaluating:
        env.set_is_evaluating()
    gen = torch.Generator().manual_seed(seed)
    N, D, Jb = n_envs, env.num_dof, env.num_bodies
    # spread the envs over the clip and put a few near its end so motion_end fires
    env.motion_start_times[:] = torch.rand(N, generator=gen) * env.motion_len
    env.motion_start_times[:3] = env.motion_len[:3] - torch.tensor([0.01, 0.03, 0.05])
    env.episode_length_buf[:] = torch.randint(0, 50, (N,), generator=gen)
    env.episode_length_buf[:3] = 0
    if start_times is not None:
        env.motion_start_times[:] = start_times(env)
        env.episode_length_buf[:] = 0
    # privileged DR tensors that the critic observes
    sim._base_com_bias[:] = (torch.rand(N, 3, generator=gen) - 0.5) * 0.1
    sim._link_mass_scale[:] = 0.9 + 0.2 * torch.rand(N, sim._link_mass_scale.shape[1], generator=gen)
    sim.friction_coeffs[:] = 0.2 + torch.rand(N, 1, 1, generator=gen)
    if general:
        sim._base_mass_scale[:] = 0.9 + 0.2 * torch.rand(N, 1, generator=gen)
    env._kp_scale[:] = 0.9 + 0.2 * torch.rand(N, D, generator=gen)
    env._kd_scale[:] = 0.9 + 0.2 * torch.rand(N, D, generator=gen)
    out = {"meta_config_json": cfg_json, "meta_n_envs": N, "meta_n_steps": n_steps, "meta_obs": obs, "meta_dt": env.dt,
           "meta_reward_names": np.array(env.reward_names), "meta_reward_scales": np.array([env.reward_scales[k] for k in env.reward_names]),
           "meta_termination_scale": np.float64(env.reward_scales["termination"]),
           "meta_history_keys": np.array(list(env.history_handler.history.keys()))}
    for k, v in env_state(env).items():
        out["init_" + k] = v
    out["init_ref_init_yaw"] = np.float64(float(env.ref_init_rpy[0, 2]))  # motion_tracking.py:186-187
    for step in range(n_steps):
        actions = synth.synth_actions(N, D, gen)
        env._pre_physics_step(actions)
        env._physics_step()
        t = (env.episode_length_buf + 2) * env.dt + env.motion_start_times
        ref = env._motion_lib.get_motion_state(env.motion_ids, t, offset=env.env_origins)
        body, root, dof, cf = synth.synth_sim_state(ref, Jb, gen, feet=tuple(int(i) for i in env.feet_indices))
        if tweak is not None:
            tweak(step, body, root, dof, cf, ref)
        sim.set_state(body, root, dof, cf)
        if tweak_env is not None:
            tweak_env(step, env)
        p = f"s{step}_"
        if cfg["domain_rand"]["push_robots"]:  # (tweak_env may have moved envs next to their push)
            out[p + "in_push_robot_counter"] = _np(env.push_robot_counter)
        out.update({p + "in_body_state": _np(body), p + "in_root_state": _np(root), p + "in_dof_state": _np(dof),
                    p + "in_contact_forces": _np(cf), p + "in_actions": _np(env.actions), p + "in_torques": _np(env.torques)})
        draws, orig_rand = [], torch.rand

        def rec_rand(*a_, **k_):
            r_ = orig_rand(*a_, **k_)
            if a_ == (1,) and not k_:
                draws.append(float(r_))
            return r_
        torch.rand = rec_rand
        try:
            env._post_physics_step()
        finally:
            torch.rand = orig_rand
        tcfg, tprob = cfg["termination"], cfg.get("termination_probality", {})
        bits, di = 0, 0
        for bit, key in enumerate(("terminate_when_close_to_dof_pos_limit", "terminate_when_close_to_dof_vel_limit", "terminate_when_close_to_torque_limit")):
            if tcfg.get(key, False):
                bits |= (1 << bit) if draws[di] < tprob[key] else 0
                di += 1
        out[p + "in_term_gates"] = np.int64(bits)
        for k, v in env.obs_buf_dict.items():
            out[p + "obs_" + k] = _np(v)
        for k, v in env_state(env).items():
            out[p + "out_" + k] = v
        for k, v in env.log_dict.items():
            out[p + "log_" + k] = _np(v).astype(np.float64)
        out[p + "extras_ref_body_pos_extend"] = _np(env.extras["ref_body_pos_extend"])
        out[p + "extras_ref_body_rot_extend"] = _np(env.extras["ref_body_rot_extend"])
        out[p + "dbg_dif_global_body_pos"] = _np(env.dif_global_body_pos)
        out[p + "dbg_ref_motion_phase"] = _np(env._ref_motion_phase)
        out[p + "dbg_base_ang_vel"] = _np(env.base_ang_vel)
        out[p + "dbg_projected_gravity"] = _np(env.projected_gravity)
        if general:
            for k in ("dif_local_body_pos", "dif_local_body_rot", "dif_global_body_rot", "dif_root_velocity", "dif_anchor_pos_z", "dif_anchor_ori",
                      "body_pos_relative_w", "_obs_local_body_pos", "_obs_local_body_rot", "_obs_anchor_ref_pos", "_obs_anchor_ref_rot",
                      "obs_next_step_mimic_buf", "rpy"):
                out[p + "dbg_" + k.lstrip("_")] = _np(getattr(env, k))
    np.savez_compressed(os.path.join(GOLD, f"ref_env_{name}.npz"), **out)
    return env


ZERO_NOISE = {f"obs.noise_scales.{k}": 0.0 for k in ("base_ang_vel", "projected_gravity", "dof_pos", "dof_vel")}


def run_term():
    """mt23_term: the base-class terminations switched on (contact, low height, close-to-dof-pos / dof-vel / torque limit with
    their per-step `torch.rand(1) < p` gates, legged_robot_base.py:433-476), a few envs driven into each of them"""
    ov = dict(ZERO_NOISE)
    ov.update({"termination.terminate_by_contact": True, "termination.terminate_by_low_height": True,
               "termination.terminate_when_close_to_dof_pos_limit": True, "termination.terminate_when_close_to_dof_vel_limit": True,
               "termination.terminate_when_close_to_torque_limit": True,
               "termination_probality.terminate_when_close_to_dof_pos_limit": 0.5, "termination_probality.terminate_when_close_to_dof_vel_limit": 0.5,
               "termination_probality.terminate_when_close_to_torque_limit": 0.5})
    state = {}

    def tweak(step, body, root, dof, cf, ref):
        cf[5, 0, 2] = 50.0                      # pelvis contact
        cf[6, 0, :] = torch.tensor([0.6, 0.6, 0.53])   # |f| just above 1 (0.36 + 0.36 + 0.2809 = 1.0009)
        cf[11, 0, :] = torch.tensor([0.6, 0.6, 0.52])  # |f| just below 1 (0.9904)
        root[7, 2] = 0.1; body[7, 0, 2] = 0.1   # low base
        lim = state["lim"]
        dof[8, 3, 0] = lim[3, 1] + 0.05         # beyond the upper termination limit
        dof[12, 4, 0] = lim[4, 0] - 0.05        # beyond the lower one
        dof[9, 2, 1] = 100.0                    # far beyond the velocity limit

    def tweak_env(step, env):
        state["lim"] = env.simulator.dof_pos_limits_termination.clone()
        env.torques[10, 1] = env.torque_limits[1]   # at the limit: above 0.98 x limit

    def tweak_first(step, body, root, dof, cf, ref):
        if "lim" in state:
            tweak(step, body, root, dof, cf, ref)

    # the limits are only known once the env exists: tweak_env runs before the step's state is recorded, so read them there first
    run_case("mt23_term", 32, 6, "motion_tracking/main", 13, ov, tweak=tweak_first, tweak_env=tweak_env)
    # the same five checks under LeggedRobotGeneralTracking (G1 29-DoF, teacher observations)
    state.clear()
    ovc = dict(ZERO_NOISE_C)
    ovc.update({k: v for k, v in ov.items() if k.startswith("termination")})
    run_case("gt29_term", 32, 4, "motion_tracking/obs_ppo_teacher", 23, ovc, general=True, tweak=tweak_first, tweak_env=tweak_env)


def run_eval():
    """mt23_eval / gt29_eval: the env after set_is_evaluating() (resets start their clip at time 0, motion_tracking.py:374-377) with the
    observation-noise curriculum on (obs.add_noise_currculum; noise scales 0 so that the fixture stays deterministic: it pins the
    curriculum value, its log and the thresholds it borrows from the reward-penalty curriculum, legged_robot_base.py:637-645)"""
    ov = dict(ZERO_NOISE)
    ov.update({"obs.add_noise_currculum": True, "obs.noise_initial_value": 0.5, "obs.soft_dof_pos_curriculum_degree": 0.01})
    if "general" not in sys.argv:
        run_case("mt23_eval", 16, 4, "motion_tracking/main", 51, ov, evaluating=True)
    ovc = dict(ZERO_NOISE_C)
    ovc.update({"obs.add_noise_currculum": True, "obs.noise_initial_value": 0.5, "obs.soft_dof_pos_curriculum_degree": 0.01,
                "obs.noise_value_max": 1.0, "obs.noise_value_min": 0.00001, "obs.soft_dof_pos_curriculum_level_down_threshold": 40})
    run_case("gt29_eval", 16, 3, "motion_tracking/obs_ppo_teacher", 52, ovc, general=True, evaluating=True)


def run_push():
    """mt23_push / gt29_push: domain_rand.push_robots as shipped in domain_rand/main.yaml (legged_robot_base.py:382-388, 1090-1106):
    a few envs are put one step short of their push interval in every step - among them env 0, which also resets at the end of
    its clip (the reset wins over the push) - so that pushes, new intervals and need_to_refresh_envs are exercised."""
    def tweak_env(step, env):
        thr = (env.push_interval_s / env.dt).int()
        ids = torch.tensor([0, 3 + step, 9, 14 - step])
        env.push_robot_counter[ids] = thr[ids] - 1
    ov = dict(ZERO_NOISE)
    ov["domain_rand.push_robots"] = True
    if "general" not in sys.argv:
        run_case("mt23_push", 16, 4, "motion_tracking/main", 41, ov, tweak_env=tweak_env)
    ovc = dict(ZERO_NOISE_C)
    ovc["domain_rand.push_robots"] = True
    ovc["domain_rand._push_fixed"] = False  # (main.yaml ships True: the push is added; False: it replaces the root velocity)
    run_case("gt29_push", 16, 3, "motion_tracking/obs_ppo_teacher", 42, ovc, general=True, tweak_env=tweak_env)


def run_p4():
    """mt23_p4 / gt29_p4: the observation sources and reward terms no shipped config uses: relyaw,
    local_ref_rigid_body_pos_relyaw, vr_3point_pos, future_ref_dof_pos / _vel (obs.future_ref_steps = 3; motion_tracking only);
    teleop_radial_body_velocity_extend / teleop_radial_joint_velocity (motion_tracking only), feet_heading_alignment(_contact),
    penalty_feet_ori(_contact)"""
    import yaml
    def obs_lists(path):
        with open(os.path.join(R.REF_ROOT, "humanoidverse/config/obs", path + ".yaml")) as f:
            d = yaml.safe_load(f)["obs"]["obs_dict"]
        second = "critic_obs" if "critic_obs" in d else "priv_obs"
        return list(d["actor_obs"]), list(d[second]), second
    extra = ["relyaw", "local_ref_rigid_body_pos_relyaw", "vr_3point_pos"]
    fut = ["future_ref_dof_pos", "future_ref_dof_vel"]
    feet = {"feet_heading_alignment": -0.3, "penalty_feet_ori": -0.4, "feet_heading_alignment_contact": -0.2, "penalty_feet_ori_contact": -0.6}

    def overrides(base, path, J, D, keys, rewards):
        a, c, second = obs_lists(path)
        ov = dict(base)
        ov["obs.obs_dict.actor_obs"] = a + ["relyaw", "vr_3point_pos"]
        ov[f"obs.obs_dict.{second}"] = c + keys
        dims = {"relyaw": 1, "local_ref_rigid_body_pos_relyaw": 3 * J, "vr_3point_pos": 9, "future_ref_dof_pos": 3 * D, "future_ref_dof_vel": 3 * D}
        scales = {"relyaw": 0.5, "local_ref_rigid_body_pos_relyaw": 1.0, "vr_3point_pos": 2.0, "future_ref_dof_pos": 1.0, "future_ref_dof_vel": 0.05}
        for k in keys:
            ov[f"obs.obs_dims.{k}"] = dims[k]
            ov[f"obs.obs_scales.{k}"] = scales[k]
            ov[f"obs.noise_scales.{k}"] = 0.0
        for k, v in rewards.items():
            ov[f"rewards.reward_scales.{k}"] = v
        return ov
    ov = overrides(ZERO_NOISE, "motion_tracking/main", 27, 23, extra + fut,
                   dict(feet, teleop_radial_body_velocity_extend=0.7, teleop_radial_joint_velocity=0.9))
    ov["obs.future_ref_steps"] = 3
    if "general" not in sys.argv:
        run_case("mt23_p4", 40, 5, "motion_tracking/main", 31, ov)
    ovc = overrides(ZERO_NOISE_C, "motion_tracking/obs_ppo_teacher", 32, 29, extra, feet)
    run_case("gt29_p4", 24, 3, "motion_tracking/obs_ppo_teacher", 32, ovc, general=True)


def run(ns=None):
    # obs=main ships uniform observation noise (helpers.py:152, torch.rand_like): parity runs set the four
    # non-zero noise scales to 0 exactly like the reference's own obs/motion_tracking/benchmark.yaml does.
    run_case("mt23_main", 48, 6, "motion_tracking/main", 11, ZERO_NOISE)
    run_case("mt23_bench", 32, 3, "motion_tracking/benchmark", 12)
    run_general()


ZERO_NOISE_C = {f"obs.noise_scales.{k}": 0.0 for k in ("base_ang_vel", "dof_pos", "dof_vel", "roll_pitch")}


def run_general():
    # config C: LeggedRobotGeneralTracking, G1 29-DoF, obs_ppo_teacher, rewards general_main, 3-clip library
    run_case("gt29_teacher", 48, 6, "motion_tracking/obs_ppo_teacher", 21, ZERO_NOISE_C, general=True)
    # scripts/train/student_ue_29dof.sh: the student observations (history inside actor_obs, no key-body targets)
    run_case("gt29_student", 32, 4, "motion_tracking/obs_ppo_student", 22, ZERO_NOISE_C, general=True)


if __name__ == "__main__":
    import sys
    run_eval() if "eval" in sys.argv else run_push() if "push" in sys.argv else run_p4() if "p4" in sys.argv else (run_general() if "general" in sys.argv else (run_term() if "term" in sys.argv else run()))

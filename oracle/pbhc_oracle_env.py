"""CPU oracle for the motion_tracking / general_tracking env step (K3).  TEST INFRASTRUCTURE ONLY.

A torch-fp32 (CPU) restatement of what one `_post_physics_step` of the reference's
LeggedRobotMotionTracking does (humanoidverse/envs/legged_base_task/legged_robot_base.py:297-338 and the
callbacks it dispatches to; envs/motion_tracking/motion_tracking.py) or LeggedRobotGeneralTracking
(envs/motion_tracking/general_tracking.py, spec.general) does, driven by a pbhc_b200.spec.TrackSpec
and the oracle motion table (oracle/pbhc_oracle.py).  Each block cites the reference lines it follows.

Parity pin: tests/test_oracle_env_golden.py replays tests/golden/ref_env_*.npz - inputs and outputs of the
UNMODIFIED reference env run on a fake simulator by tools/make_golden_env_step.py - through this file and
requires bit-exact obs / rewards / masks / state.  Random draws the reference makes at resets
(sample_time, episodic domain randomisation) are injected from the fixture (`inject`).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
import numpy as np
import torch

from . import pbhc_oracle as O


def yaw_zero(n):
    return torch.zeros(n, 3)


def new_state(spec, N, env_origins=None):
    """buffers as LeggedRobotBase._init_buffers / _prepare_reward_function / HistoryHandler create them."""
    D = spec.D
    S = dict(
        episode_length_buf=torch.zeros(N, dtype=torch.long), motion_start_times=torch.zeros(N), motion_len=torch.zeros(N),
        motion_ids=torch.arange(N), actions=torch.zeros(N, D), last_actions=torch.zeros(N, D), last_dof_pos=torch.zeros(N, D),
        last_dof_vel=torch.zeros(N, D), last_contacts=torch.zeros(N, len(spec.feet_ids)), feet_air_time=torch.zeros(N, len(spec.feet_ids)),
        kp_scale=torch.ones(N, D), kd_scale=torch.ones(N, D), rfi_lim_scale=torch.ones(N, D), rao_scale=torch.ones(N, D),
        action_delay_idx=torch.zeros(N, dtype=torch.long), default_dof_pos=torch.tensor(spec.default_joint_angles).repeat(N, 1),
        base_com_bias=torch.zeros(N, 3), link_mass_scale=torch.ones(N, spec.L), friction_coeffs=torch.ones(N, 1, 1),
        base_mass_scale=torch.ones(N, 1),
        env_origins=torch.zeros(N, 3) if env_origins is None else env_origins,
        rew_buf=torch.zeros(N, len(spec.reward_names) + 1) if spec.use_vec_reward else torch.zeros(N),
        reset_buf=torch.ones(N, dtype=torch.long), time_out_buf=torch.zeros(N, dtype=torch.bool),
        end_time_ratio_buf=torch.zeros(N),
        average_episode_length=0.0, reward_penalty_scale=spec.penalty["initial"], motion_far_threshold=spec.motion_far_threshold0,
        soft_dof_pos=spec.limit_cur["dof_pos"]["initial"], soft_dof_vel=spec.limit_cur["dof_vel"]["initial"],
        soft_torque=spec.limit_cur["torque"]["initial"], current_noise_curriculum_value=spec.noise_cur["initial"])
    if spec.dr.get("push_robots"):  # legged_robot_base.py:135, 139
        S["push_robot_counter"] = torch.zeros(N, dtype=torch.int)
        S["push_interval_s"] = torch.randint(int(spec.dr["push_interval_s"][0]), int(spec.dr["push_interval_s"][1]), (N,))
        S["push_robot_vel_buf"] = torch.zeros(N, 2)
    for k in spec.hist_keys:
        S["hist_" + k] = torch.zeros(N, spec.hist_len[k], spec.hist_dim[k])
    for k in spec.episode_sum_names:
        S["epsum_" + k] = torch.zeros(N)
    for k, v in spec.sigma0.items():
        S["sigma_" + k] = float(v)
        S["ema_" + k] = float(v)
    return S


def _update_sigma(spec, S, err, term):
    """_update_adaptive_sigma, motion_tracking.py:1030-1048 (python float64 state, .mean().item())."""
    a = spec.adaptive
    if not a["enable"]:
        return
    S["ema_" + term] = S["ema_" + term] * (1 - a["alpha"]) + err.mean().item() * a["alpha"]
    ema, sig = S["ema_" + term], S["sigma_" + term]
    if a["type"] == "scale":
        S["sigma_" + term] = min(ema * a["scale"], sig)
    elif a["type"] == "mean":  # general_tracking.py:985 keeps only the EMA
        S["sigma_" + term] = ema if spec.general else (min(ema, sig) + ema) / 2
    elif a["type"] == "origin":
        S["sigma_" + term] = min(ema, sig)


def _pre_compute(spec, tab, S, inp):
    """LeggedRobotBase._pre_compute_observations_callback (legged_robot_base.py:346-380) followed by
    LeggedRobotMotionTracking._pre_compute_observations_callback (motion_tracking.py:583-747)."""
    N, Jb, J = inp["root_state"].shape[0], spec.Jb, spec.J
    W = {}
    root = inp["root_state"]
    q = root[:, 3:7]
    gravity = torch.tensor([0.0, 0.0, -1.0]).repeat(N, 1)
    W["base_lin_vel"] = O.quat_rotate_inverse(q, root[:, 7:10])
    W["base_ang_vel"] = O.quat_rotate_inverse(q, root[:, 10:13])
    W["projected_gravity"] = O.quat_rotate_inverse(q, gravity)
    cf = inp["contact_forces"]
    W["contacts"] = (cf[:, spec.feet_ids, :].norm(dim=-1) > 1.0).float()
    W["contacts_filt"] = torch.logical_or(W["contacts"], S["last_contacts"]).float()
    # kick_motion_res, motion_tracking.py:548-558
    times = (S["episode_length_buf"] + 1) * spec.dt + S["motion_start_times"]
    ref = O.get_motion_state(tab, S["motion_ids"], times, offset=S["env_origins"])
    W["ref"], W["motion_times"] = ref, times
    body = inp["body_state"]
    pos, rot, vel, ang = body[..., 0:3], body[..., 3:7], body[..., 7:10], body[..., 10:13]
    # extended bodies, motion_tracking.py:619-643
    par = spec.ext_parent
    E = len(par)
    ext_pos = torch.tensor(spec.ext_pos).repeat(N, 1, 1)
    ext_rot_xyzw = torch.tensor(spec.ext_rot_wxyz).repeat(N, 1, 1)[:, :, [1, 2, 3, 0]]
    rotated = O.quat_rotate(rot[:, par].reshape(-1, 4), ext_pos.reshape(-1, 3))
    e_pos = O.quat_rotate(ext_rot_xyzw.reshape(-1, 4), rotated).view(N, -1, 3) + pos[:, par]
    e_rot = O.quat_mul(rot[:, par].reshape(-1, 4), ext_rot_xyzw.reshape(-1, 4)).view(N, -1, 4)
    e_vel = vel[:, par] + torch.cross(ang[:, par], ext_pos.view(N, -1, 3), dim=2).view(N, -1, 3)
    W["pos_ext"] = torch.cat([pos, e_pos], dim=1)
    W["rot_ext"] = torch.cat([rot, e_rot], dim=1)
    W["vel_ext"] = torch.cat([vel, e_vel], dim=1)
    W["ang_ext"] = torch.cat([ang, ang[:, par]], dim=1)
    # differences, motion_tracking.py:645-665
    W["dif_pos"] = ref["rg_pos_t"] - W["pos_ext"]
    W["dif_rot"] = ref["rg_rot_t"] - W["rot_ext"]
    W["dif_vel"] = ref["body_vel_t"] - W["vel_ext"]
    W["dif_ang"] = ref["body_ang_vel_t"] - W["ang_ext"]
    dof_pos, dof_vel = inp["dof_state"][..., 0], inp["dof_state"][..., 1]
    W["dif_joint_angles"] = ref["dof_pos"] - dof_pos
    W["dif_joint_velocities"] = ref["dof_vel"] - dof_vel
    # heading-frame observation tensors, motion_tracking.py:678-721
    h_inv = O.calc_heading_quat_inv(q.clone())
    h_exp = h_inv.unsqueeze(1).expand(-1, J, -1).reshape(-1, 4)
    W["dif_local_rigid_body_pos"] = O.quat_rotate(h_exp, W["dif_pos"].reshape(-1, 3)).view(N, -1)
    g_ref = ref["rg_pos_t"] - root[:, :3].view(N, 1, 3)
    W["local_ref_rigid_body_pos"] = O.quat_rotate(h_exp, g_ref.reshape(-1, 3)).view(N, -1)
    W["local_ref_rigid_body_vel"] = O.quat_rotate(h_exp, ref["body_vel_t"].reshape(-1, 3)).view(N, -1)
    W["global_ref_rigid_body_vel"] = ref["body_vel_t"].reshape(N, -1)
    # relyaw / vr 3-point observations, motion_tracking.py:686-687, 720-731 (general_tracking.py:682-725)
    W["rpy"] = O.get_euler_xyz_in_tensor(q)                      # legged_robot_base.py:349
    W["relyaw"] = W["rpy"][:, 2:3] - float(S.get("ref_init_yaw", 0.0))
    ry_exp = O.calc_yaw_heading_quat_inv(W["relyaw"]).unsqueeze(1).expand(-1, J, -1).reshape(-1, 4)
    W["local_ref_rigid_body_pos_relyaw"] = O.quat_rotate(ry_exp, ref["body_vel_t"].reshape(-1, 3)).view(N, -1)  # velocities (sic)
    if len(spec.vr3_ids) == 3:
        vr = ref["rg_pos_t"][:, spec.vr3_ids, :] - root[:, 0:3].view(N, 1, 3)
        W["vr_3point_pos"] = O.quat_rotate(h_inv.repeat(3, 1).view(-1, 4), vr.reshape(-1, 3)).view(N, -1)  # `repeat`, not repeat_interleave (sic)
    if not spec.general and spec.future_num_steps > 0:  # kick_motion_res_multiplestep, motion_tracking.py:562-579, 594-599
        fp, fv = [], []
        for i in range(spec.future_num_steps):
            t_i = (S["episode_length_buf"] + 1 + i) * spec.dt + S["motion_start_times"]
            m = O.get_motion_state(tab, S["motion_ids"], t_i, offset=S["env_origins"])
            fp.append(m["dof_pos"].unsqueeze(1)); fv.append(m["dof_vel"].unsqueeze(1))
        W["future_ref_dof_pos"] = torch.cat(fp, dim=1).view(N, -1)
        W["future_ref_dof_vel"] = torch.cat(fv, dim=1).view(N, -1)
    # phase, motion_tracking.py:738-744
    W["ref_motion_phase"] = (times / tab.length[S["motion_ids"]]).unsqueeze(1)
    if spec.general:
        _pre_compute_general(spec, tab, S, inp, W)
    return W


def _future_targets(spec, tab, S, W):
    """_get_future_motion_targets, general_tracking.py:500-565: F future reference frames per env (2-D lookup)."""
    N, F, a = S["motion_ids"].shape[0], spec.future_num_steps, spec.anchor_index
    steps = torch.tensor(spec.future_steps, dtype=torch.long)
    times = S["episode_length_buf"] * spec.dt + S["motion_start_times"]
    obs_times = steps * spec.dt + times[:, None]
    ids = S["motion_ids"][:, None].expand(-1, F)
    off = S["env_origins"][:, None, :].expand(-1, F, -1)
    r = O.get_motion_state(tab, ids, obs_times, off)
    flat_rot = r["root_rot"].reshape(N * F, 4)
    rpy = O.get_euler_xyz_in_tensor(flat_rot)
    roll_pitch = rpy[:, :2].reshape(N, F, 2)
    root_vel = O.quat_rotate_inverse(flat_rot, r["root_vel"].reshape(N * F, 3)).view(N, F, 3)
    root_ang = O.quat_rotate_inverse(flat_rot, r["root_ang_vel"].reshape(N * F, 3)).view(N, F, 3)
    pos_t, rot_t = r["rg_pos_t"], r["rg_rot_t"]
    J = pos_t.shape[2]
    a_pos = pos_t[..., a, :][..., None, :].repeat(1, 1, J, 1)
    a_rot = rot_t[..., a, :][..., None, :].repeat(1, 1, J, 1)
    key = O.quat_apply(O.quat_conjugate(a_rot), pos_t - a_pos)[..., spec.key_ids, :].reshape(N, F, -1)
    W["future_motion_root_height"] = r["root_pos"][..., 2:3].reshape(N, -1)
    W["future_motion_roll_pitch"] = roll_pitch.reshape(N, -1)
    W["future_motion_base_lin_vel"] = root_vel.reshape(N, -1)
    W["future_motion_base_yaw_vel"] = root_ang[..., 2:3].reshape(N, -1)
    W["future_motion_dof_pos"] = r["dof_pos"].reshape(N, -1)
    W["future_motion_local_ref_key_body_pos"] = key.reshape(N, -1)
    W["next_step_ref_motion"] = torch.cat((r["root_pos"][:, 0, 2:3], roll_pitch[:, 0, :], root_vel[:, 0, :], root_ang[:, 0, 2:3],
                                           r["dof_pos"][:, 0, :], key[:, 0, :]), dim=-1)


def _pre_compute_general(spec, tab, S, inp, W):
    """what LeggedRobotGeneralTracking._pre_compute_observations_callback adds or changes
    (general_tracking.py:568-805)."""
    N, J, a = inp["root_state"].shape[0], spec.J, spec.anchor_index
    ref, root = W["ref"], inp["root_state"]
    if spec.future_num_steps > 0:
        _future_targets(spec, tab, S, W)
    W["roll_pitch"] = W["rpy"][:, :2]
    # differences (:638-666): quaternion product instead of the component-wise subtraction of motion_tracking
    W["dif_rot"] = O.quat_mul(ref["rg_rot_t"], O.quat_conjugate(W["rot_ext"]))
    ref_root_vel_local = O.quat_rotate_inverse(ref["root_rot"].view(N, 4), ref["root_vel"].view(N, 3)).view(N, 3)
    W["dif_root_velocity"] = ref_root_vel_local - W["base_lin_vel"]
    # beyondmimic block (:738-803)
    pos_ext, rot_ext = W["pos_ext"], W["rot_ext"]
    a_pos = ref["rg_pos_t"][:, a, :][:, None, :].repeat(1, J, 1)
    a_rot = ref["rg_rot_t"][:, a, :][:, None, :].repeat(1, J, 1)
    ra_pos = pos_ext[:, a, :][:, None, :].repeat(1, J, 1)
    ra_rot = rot_ext[:, a, :][:, None, :].repeat(1, J, 1)
    delta_pos = ra_pos                      # the reference aliases: the z overwrite below also changes ra_pos (:752-753)
    delta_pos[..., 2] = a_pos[..., 2]
    delta_ori = O.yaw_quat(O.quat_mul(ra_rot, O.quat_conjugate(a_rot)))
    rel_pos = delta_pos + O.quat_apply(delta_ori, ref["rg_pos_t"] - a_pos)
    rel_rot = O.quat_mul(delta_ori, ref["rg_rot_t"])
    W["dif_local_body_pos"] = rel_pos - pos_ext
    W["dif_local_body_rot"] = O.quat_mul(rel_rot, O.quat_conjugate(rot_ext))
    W["local_body_rot"] = O.matrix_from_quat(O.quat_mul(O.quat_conjugate(ra_rot), rot_ext))[..., :2]
    W["local_body_pos"] = O.quat_apply(O.quat_conjugate(ra_rot), pos_ext - ra_pos)
    W["anchor_ref_rot"] = O.matrix_from_quat(O.quat_mul(O.quat_conjugate(rot_ext[:, a, :]), ref["rg_rot_t"][:, a, :]))[..., :2]
    W["anchor_ref_pos"] = O.quat_apply(O.quat_conjugate(rot_ext[:, a, :]), ref["rg_pos_t"][:, a, :] - pos_ext[:, a, :])
    W["dif_anchor_body_pos"] = W["dif_pos"][:, a, :]
    W["dif_anchor_pos_z"] = ref["rg_pos_t"][:, a, -1] - pos_ext[:, a, -1]
    grav = torch.tensor([0.0, 0.0, -1.0]).repeat(N, 1)
    W["dif_anchor_ori"] = (O.quat_rotate_inverse(ref["rg_rot_t"][:, a, :], grav)[:, 2]
                           - O.quat_rotate_inverse(rot_ext[:, a, :], grav)[:, 2])


def _termination(spec, S, W, log, inp=None, inject=None):
    """_check_termination (legged_robot_base.py:408-431) with the motion_tracking overrides (:330-358)."""
    N = S["reset_buf"].shape[0]
    S["reset_buf"][:] = 0
    S["time_out_buf"][:] = 0
    by = {}
    # base class (legged_robot_base.py:433-476); the three `torch.rand(1) < p` gates come in as inject["term_gates"] (bits 0..2)
    gates = int((inject or {}).get("term_gates", 7))
    if getattr(spec, "term_contact", False):
        by["contact"] = torch.any(torch.norm(inp["contact_forces"][:, spec.term_contact_ids, :], dim=-1) > 1.0, dim=1)
        S["reset_buf"] |= by["contact"]
    if spec.term_gravity:
        by["gravity"] = torch.norm(W["projected_gravity"][:, 0:2], dim=-1) > spec.term_gravity_thr
        S["reset_buf"] |= by["gravity"]
    if getattr(spec, "term_low_height", False):
        by["low_height"] = torch.any(inp["root_state"][:, 2:3] < spec.term_min_height, dim=1)
        S["reset_buf"] |= by["low_height"]
    if getattr(spec, "term_dof_pos", False):
        lo_h, hi_h = torch.tensor(spec.hard_dof_lo), torch.tensor(spec.hard_dof_hi)
        m, r = (lo_h + hi_h) / 2, hi_h - lo_h
        lo, hi = m - 0.5 * r * spec.term_dof_pos_scale, m + 0.5 * r * spec.term_dof_pos_scale  # isaacgym.py:387-388
        dof_pos = inp["dof_state"][..., 0]
        out = -(dof_pos - lo).clip(max=0.0)
        out += (dof_pos - hi).clip(min=0.0)
        out = torch.sum(out, dim=1)
        by["dof_pos_limit"] = (out > 0.0) if gates & 1 else torch.zeros_like(out)
        if gates & 1:
            S["reset_buf"] |= by["dof_pos_limit"]
    if getattr(spec, "term_dof_vel", False):
        out = torch.sum((torch.abs(inp["dof_state"][..., 1]) - torch.tensor(spec.dof_vel_limit) * spec.term_dof_vel_scale).clip(min=0.0, max=1.0), dim=1)
        by["dof_vel_limit"] = (out > 0.0) if gates & 2 else torch.zeros_like(out)
        if gates & 2:
            S["reset_buf"] |= by["dof_vel_limit"]
    if getattr(spec, "term_torque", False):
        out = torch.sum((torch.abs(inp["torques"]) - torch.tensor(spec.torque_limit) * spec.term_torque_scale).clip(min=0.0, max=1.0), dim=1)
        by["torque_limit"] = (out > 0.0) if gates & 4 else torch.zeros_like(out)
        if gates & 4:
            S["reset_buf"] |= by["torque_limit"]
    if spec.term_motion_far:
        by["motion_far"] = torch.any(torch.norm(W["dif_pos"], dim=-1) > S["motion_far_threshold"], dim=-1)
        S["reset_buf"] |= by["motion_far"]
        if spec.far_cur["enable"]:
            log["terminate_when_motion_far_threshold"] = torch.tensor(S["motion_far_threshold"], dtype=torch.float)
    if spec.general:  # general_tracking.py:241-255
        if spec.term_ref_pos_z:
            by["ref_pos_z"] = torch.abs(W["dif_anchor_pos_z"]) > spec.term_ref_pos_z_thr
            S["reset_buf"] |= by["ref_pos_z"]
        if spec.term_ref_ori:
            by["ref_ori"] = W["dif_anchor_ori"].abs() > spec.term_ref_ori_thr
            S["reset_buf"] |= by["ref_ori"]
        if spec.term_body_z:
            by["body_z"] = torch.any(W["dif_local_body_pos"][:, spec.body_z_ids, -1].abs() > spec.term_body_z_thr, dim=-1)
            S["reset_buf"] |= by["body_z"]
    S["time_out_buf"] |= S["episode_length_buf"] > spec.max_episode_length
    by["time_out"] = S["time_out_buf"]
    if spec.term_motion_end:
        cur = S["episode_length_buf"] * spec.dt + S["motion_start_times"]
        by["motion_end"] = cur > S["motion_len"]
        S["time_out_buf"] |= by["motion_end"]
    S["reset_buf"] |= S["time_out_buf"]
    for k, v in by.items():
        log["terminate_by_" + k] = v.float().mean()
    for k in list(log.keys()):
        if k.startswith("terminate_by_"):
            log[k] = log[k] / (S["reset_buf"].float().mean() + 1e-15)


def _rewards(spec, S, W, inp, log):
    """_compute_reward (legged_robot_base.py:715-761) and the reward terms of rewards/motion_tracking/main.yaml
    (motion_tracking.py:1154-1328, legged_robot_base.py:944-1087)."""
    dt = spec.dt
    cf = inp["contact_forces"]
    feet = spec.feet_ids
    dof_pos, dof_vel = inp["dof_state"][..., 0], inp["dof_state"][..., 1]
    torques, actions = inp["torques"], inp["actions"]
    tl = torch.tensor(spec.torque_limit)
    vl = torch.tensor(spec.dof_vel_limit)

    def msq(x):
        return (x ** 2).mean(dim=-1).mean(dim=-1)

    def track(err, key):
        r = torch.exp(-err / S["sigma_" + key])
        _update_sigma(spec, S, err, key)
        return r

    def r_body_pos():
        up, lo = msq(W["dif_pos"][:, spec.upper_ids, :]), msq(W["dif_pos"][:, spec.lower_ids, :])
        r_up = torch.exp(-up / S["sigma_teleop_upper_body_pos"])
        r_lo = torch.exp(-lo / S["sigma_teleop_lower_body_pos"])
        r = r_lo * spec.w_lower + r_up * spec.w_upper
        _update_sigma(spec, S, up, "teleop_upper_body_pos")
        _update_sigma(spec, S, lo, "teleop_lower_body_pos")
        return r

    def r_feet_air_time():
        contact = cf[:, feet, 2] > 1.0
        filt = torch.logical_or(contact, S["last_contacts"] > 0)
        first = (S["feet_air_time"] > 0.0) * filt
        S["feet_air_time"] += dt
        r = torch.sum((S["feet_air_time"] - spec.desired_feet_air_time) * first, dim=1)
        S["feet_air_time"] *= ~filt
        return r

    def r_limits_dof_pos():
        if spec.limit_cur["dof_pos"]["enable"]:
            lo_h, hi_h = torch.tensor(spec.hard_dof_lo), torch.tensor(spec.hard_dof_hi)
            m, r = (lo_h + hi_h) / 2, hi_h - lo_h
            lo, hi = m - 0.5 * r * S["soft_dof_pos"], m + 0.5 * r * S["soft_dof_pos"]
        else:
            lo, hi = soft_dof_limits(spec)
        out = -(dof_pos - lo).clip(max=0.0)
        out += (dof_pos - hi).clip(min=0.0)
        return torch.sum(out, dim=1)

    def radial_velocity_potential(cur_vel, ref_vel, sigma_theta=0.75, alpha_norm=0.4):  # motion_tracking.py:78-94
        cosine_error = 1 - torch.nn.functional.cosine_similarity(cur_vel, ref_vel, dim=-1)
        potential_cosine = torch.exp(-cosine_error / sigma_theta)
        norm_ratio = cur_vel.norm(dim=-1) / ref_vel.norm(dim=-1)
        return potential_cosine * (norm_ratio * torch.exp(alpha_norm * (1 - norm_ratio ** (1 / alpha_norm))))

    def r_radial_body():  # :1238-1244
        cur = W["vel_ext"].reshape(N_, -1)
        return radial_velocity_potential(cur, W["dif_vel"].reshape(N_, -1) + cur)

    def feet_heading_diffs():  # legged_robot_base.py:1031-1047
        fwd = torch.tensor([1.0, 0.0, 0.0]).repeat(N_, 1)
        out = []
        hr = O.quat_apply(inp["root_state"][:, 3:7], fwd)
        heading_root = torch.atan2(hr[:, 1], hr[:, 0])
        for f in feet[:2]:
            ff = O.quat_apply(inp["body_state"][:, f, 3:7], fwd)
            out.append(torch.abs(O.wrap_to_pi(torch.atan2(ff[:, 1], ff[:, 0]) - heading_root)))
        return out

    def feet_gravity():  # :1049-1054
        grav = torch.tensor([0.0, 0.0, -1.0]).repeat(N_, 1)
        return [O.quat_rotate_inverse(inp["body_state"][:, f, 3:7], grav) for f in feet[:2]]

    N_ = dof_pos.shape[0]
    cfilt = W["contacts_filt"]
    fn = {
        "teleop_radial_body_velocity_extend": r_radial_body,
        "teleop_radial_joint_velocity": lambda: radial_velocity_potential(dof_vel, W["dif_joint_velocities"] + dof_vel),
        "feet_heading_alignment": lambda: (lambda d: d[0] + d[1])(feet_heading_diffs()),
        "feet_heading_alignment_contact": lambda: (lambda d: d[0] * cfilt[:, 0] + d[1] * cfilt[:, 1])(feet_heading_diffs()),
        "penalty_feet_ori": lambda: (lambda g: torch.sum(torch.square(g[0][:, :2]), dim=1) ** 0.5 + torch.sum(torch.square(g[1][:, :2]), dim=1) ** 0.5)(feet_gravity()),
        "penalty_feet_ori_contact": lambda: (lambda g: torch.norm(g[0][:, :2], dim=-1) * cfilt[:, 0] + torch.norm(g[1][:, :2], dim=-1) * cfilt[:, 1])(feet_gravity()),
        "teleop_contact_mask": lambda: 1 - (W["contacts_filt"] - W["ref"]["contact_mask"]).abs().mean(dim=-1),
        "teleop_contact_mask_v2": lambda: 0.5 - (W["contacts_filt"] - W["ref"]["contact_mask"]).abs().mean(dim=-1),
        "teleop_max_joint_position": lambda: track(W["dif_joint_angles"].abs().max(dim=-1)[0], "teleop_max_joint_pos"),
        "teleop_body_position_extend": r_body_pos,
        "teleop_vr_3point": lambda: track(msq(W["dif_pos"][:, spec.vr3_ids, :]), "teleop_vr_3point_pos"),
        "teleop_body_position_feet": lambda: track(msq(W["dif_pos"][:, feet, :]), "teleop_feet_pos"),
        "teleop_body_rotation_extend": lambda: track(msq(W["dif_rot"]), "teleop_body_rot"),
        "teleop_body_ang_velocity_extend": lambda: track(msq(W["dif_ang"]), "teleop_body_ang_vel"),
        "teleop_body_velocity_extend": lambda: track(msq(W["dif_vel"]), "teleop_body_vel"),
        "teleop_joint_position": lambda: track((W["dif_joint_angles"] ** 2).mean(dim=-1), "teleop_joint_pos"),
        "teleop_joint_velocity": lambda: track((W["dif_joint_velocities"] ** 2).mean(dim=-1), "teleop_joint_vel"),
        "penalty_torques": lambda: torch.sum(torch.square(torques), dim=1),
        "penalty_action_rate": lambda: torch.sum(torch.square(S["last_actions"] - actions), dim=1),
        "penalty_dof_vel": lambda: torch.sum(torch.square(dof_vel), dim=1),
        "penalty_dof_acc": lambda: torch.sum(torch.square((S["last_dof_vel"] - dof_vel) / dt), dim=1),
        "penalty_orientation": lambda: torch.sum(torch.square(W["projected_gravity"][:, :2]), dim=1),
        "alive": lambda: 1.0 * torch.ones_like(S["reset_buf"]),
        "feet_air_time": r_feet_air_time,
        "penalty_feet_contact_forces": lambda: torch.sum((torch.norm(cf[:, feet, :], dim=-1) - spec.max_contact_force).clip(min=0.0), dim=1),
        "penalty_stumble": lambda: torch.any(torch.norm(cf[:, feet, :2], dim=2) > 5 * torch.abs(cf[:, feet, 2]), dim=1),
        "penalty_slippage": lambda: torch.sum(torch.norm(inp["body_state"][:, feet, 7:10], dim=-1) * (torch.norm(cf[:, feet, :], dim=-1) > 1.0), dim=1),
        "foot_slip_penalty": lambda: torch.sum((torch.norm(cf[:, feet, :], dim=-1) > 1.0) * torch.linalg.norm(inp["body_state"][:, feet, 7:9], dim=-1), dim=1),
        "limits_dof_pos": r_limits_dof_pos,
        "limits_dof_vel": lambda: torch.sum((torch.abs(dof_vel) - vl * (S["soft_dof_vel"] if spec.limit_cur["dof_vel"]["enable"] else spec.soft_dof_vel_limit)).clip(min=0.0, max=1.0), dim=1),
        "limits_torque": lambda: (torch.sum((torch.abs(torques) - tl * S["soft_torque"]).clip(min=0.0, max=1.0), dim=1) if spec.limit_cur["torque"]["enable"]
                                  else torch.sum((torch.abs(torques) - tl * spec.soft_torque_limit).clip(min=0.0), dim=1)),
        "collision": lambda: torch.sum(1.0 * (torch.norm(cf[:, spec.penalised_ids, :], dim=-1) > 0.1), dim=1),
    }
    if spec.general:  # general_tracking.py:1116-1270
        K, a = spec.key_ids, spec.anchor_index
        fn.update({
            "teleop_root_vel": lambda: track((W["dif_root_velocity"] ** 2).mean(dim=-1), "teleop_root_vel"),
            "teleop_anchor_body_position": lambda: track((W["dif_anchor_body_pos"] ** 2).mean(dim=-1), "teleop_anchor_body_pos"),
            "teleop_anchor_body_rotation": lambda: track(O.quat_to_angle(W["dif_rot"])[:, a] ** 2, "teleop_anchor_body_rot"),
            "local_key_body_position": lambda: track(msq(W["dif_local_body_pos"][:, K, :]), "local_key_body_pos"),
            "local_key_body_rotation": lambda: track((O.quat_to_angle(W["dif_local_body_rot"])[:, K] ** 2).mean(dim=-1), "local_key_body_rot"),
            "key_body_velocity": lambda: track(msq(W["dif_vel"][:, K, :]), "key_body_vel"),
            "key_body_ang_velocity": lambda: track(msq(W["dif_ang"][:, K, :]), "key_body_ang_vel"),
        })
    S["rew_buf"][:] = 0.0
    i = 0
    for i, name in enumerate(spec.reward_names):
        rew = fn[name]() * spec.reward_scales[i]
        if spec.reward_is_penalty[i]:
            rew *= S["reward_penalty_scale"]
        if spec.use_vec_reward:
            S["rew_buf"][:, i] += rew
        else:
            S["rew_buf"] += rew
        S["epsum_" + name] += rew
    if spec.only_positive_rewards:
        S["rew_buf"][:] = torch.clip(S["rew_buf"][:], min=0.0)
    if spec.termination_scale is not None:
        rew = (S["reset_buf"] * ~S["time_out_buf"]) * spec.termination_scale
        if spec.use_vec_reward:
            S["rew_buf"][:, i] += rew  # the stale loop index: last function's column (legged_robot_base.py:744)
        else:
            S["rew_buf"] += rew
        S["epsum_termination"] += rew
    if spec.penalty_curriculum:
        log["penalty_scale"] = torch.tensor(S["reward_penalty_scale"], dtype=torch.float)
        log["average_episode_length"] = torch.as_tensor(S["average_episode_length"], dtype=torch.float)
    for k in ("dof_pos", "dof_vel", "torque"):
        if spec.limit_cur[k]["enable"]:
            log[f"soft_{k}_curriculum_value"] = torch.tensor(S["soft_" + k], dtype=torch.float)
    if spec.add_noise_curriculum:  # legged_robot_base.py:760-761
        log["current_noise_curriculum_value"] = torch.tensor(S["current_noise_curriculum_value"], dtype=torch.float)


def soft_dof_limits(spec):
    """FakeSim/IsaacGym soft limits (isaacgym.py:381-385)."""
    lo, hi = torch.tensor(spec.hard_dof_lo), torch.tensor(spec.hard_dof_hi)
    m, r = (lo + hi) / 2, hi - lo
    return m - 0.5 * r * spec.soft_dof_pos_limit, m + 0.5 * r * spec.soft_dof_pos_limit


def _curriculum(value, avg, c, sign_down):
    """the shared shape of the reference's scalar curricula (legged_robot_base.py:883-939,
    motion_tracking.py:306-317): below `down` multiply by (1 + sign_down*degree), above `up` by (1 - sign_down*degree)."""
    if avg < c["down"]:
        value *= (1 + sign_down * c["degree"])
    elif avg > c["up"]:
        value *= (1 - sign_down * c["degree"])
    return float(np.clip(value, c["lo"], c["hi"]))


def _reset(spec, tab, S, W, inp, inject, log, last_ep_len):
    """reset_envs_idx (legged_robot_base.py:491-516) = _reset_buffers_callback (:647-686) +
    _reset_tasks_callback (:587-598 and motion_tracking.py:274-296) + _reset_dofs/_reset_root_states
    (motion_tracking.py:445-543).  Returns the episode extras."""
    ids = S["reset_buf"].nonzero(as_tuple=False).flatten()
    if len(ids) == 0:
        return None
    for k in ("actions", "last_actions", "last_dof_pos", "last_dof_vel", "feet_air_time"):
        S[k][ids] = 0.0
    S["episode_length_buf"][ids] = 0
    W["contacts"][ids] = 0.0
    W["contacts_filt"][ids] = 0.0
    S["last_contacts"][ids] = 0.0
    S["reset_buf"][ids] = 1
    # _update_average_episode_length (:875-879)
    num = len(ids)
    cur = torch.mean(last_ep_len[ids], dtype=torch.float)
    ncomp = spec.num_compute_average_epl
    S["average_episode_length"] = S["average_episode_length"] * (1 - num / ncomp) + cur * (num / ncomp)
    for k in spec.hist_keys:
        S["hist_" + k][ids] *= 0.0
    # _reset_tasks_callback: episodic domain randomisation (values injected), curricula
    for k in ("kp_scale", "kd_scale", "rfi_lim_scale", "rao_scale", "action_delay_idx"):
        if inject is not None and k in inject:
            S[k][ids] = inject[k][ids]
    avg = S["average_episode_length"]
    if spec.penalty_curriculum:
        S["reward_penalty_scale"] = _curriculum(S["reward_penalty_scale"], avg, spec.penalty, -1)
    for k in ("dof_pos", "dof_vel", "torque"):
        if spec.limit_cur[k]["enable"]:
            S["soft_" + k] = _curriculum(S["soft_" + k], avg, spec.limit_cur[k], +1)
    if spec.add_noise_curriculum:  # _update_obs_noise_curriculum, legged_robot_base.py:637-645
        S["current_noise_curriculum_value"] = _curriculum(S["current_noise_curriculum_value"], avg, spec.noise_cur, -1)
    end_time = last_ep_len[ids] * spec.dt + S["motion_start_times"][ids]
    S["end_time_ratio_buf"][ids] = end_time / S["motion_len"][ids]
    log["end_time_ratio"] = S["end_time_ratio_buf"].mean()
    log["end_time_ratio_std"] = S["end_time_ratio_buf"].std()
    # _resample_motion_times (motion_tracking.py:369-384); sample_time draws injected
    S["motion_len"][ids] = tab.length[S["motion_ids"][ids]]
    if "motion_start_times" in inject:
        S["motion_start_times"][ids] = inject["motion_start_times"][ids]
    else:  # is_evaluating and not enforce_randomize_motion_start_eval (motion_tracking.py:374-375)
        assert S.get("is_evaluating"), "oracle: start times are injected unless the env is evaluating"
        S["motion_start_times"][ids] = torch.zeros(len(ids))
    if spec.far_cur["enable"]:
        S["motion_far_threshold"] = _curriculum(S["motion_far_threshold"], avg, spec.far_cur, +1)
    # _reset_dofs / _reset_root_states: kick_motion_res() is recomputed at the new times for all envs
    times = (S["episode_length_buf"] + 1) * spec.dt + S["motion_start_times"]
    ref = O.get_motion_state(tab, S["motion_ids"], times, offset=S["env_origins"])
    if spec.noise_to_initial_level != 0:
        raise NotImplementedError("oracle: reset noise needs injected normals (reference default noise_to_initial_level = 0)")
    dof = inp["dof_state"]
    # general_tracking._reset_dofs (:464-483) queries the motion at episode_length*dt + start (no +1); motion_tracking's
    # (:524-543) shares kick_motion_res() with the root reset
    ref_d = (O.get_motion_state(tab, S["motion_ids"], S["episode_length_buf"] * spec.dt + S["motion_start_times"], offset=S["env_origins"])
             if spec.general else ref)
    dof[ids, :, 0] = ref_d["dof_pos"][ids]
    dof[ids, :, 1] = ref_d["dof_vel"][ids]
    root = inp["root_state"]
    root[ids, :3] = ref["root_pos"][ids]
    ident = torch.zeros(len(ids), 4)
    ident[:, 3] = 1.0  # small_random_quaternions(n, 0): sin(0)*axis, cos(0)
    root[ids, 3:7] = O.quat_mul(ident, ref["root_rot"][ids])
    root[ids, 7:10] = ref["root_vel"][ids]
    root[ids, 10:13] = ref["root_ang_vel"][ids]
    extras = {"rew_" + k: S["epsum_" + k][ids] / spec.max_episode_length_s for k in spec.episode_sum_names}
    for k in spec.episode_sum_names:
        S["epsum_" + k][ids] = 0.0
    extras["end_epis_length"] = last_ep_len[ids]
    return extras


def _observations(spec, S, W, inp):
    """_compute_observations + parse_observation (legged_robot_base.py:763-793, helpers.py:128-152, noise
    scales 0), the history getters (motion_tracking.py:989-1013), clip (:326-328) and history push (:330-331)."""
    N = S["actions"].shape[0]
    dof_pos, dof_vel = inp["dof_state"][..., 0], inp["dof_state"][..., 1]
    src = {
        "base_lin_vel": W["base_lin_vel"], "base_ang_vel": W["base_ang_vel"], "projected_gravity": W["projected_gravity"],
        "dof_pos": dof_pos - S["default_dof_pos"], "dof_vel": dof_vel, "actions": S["actions"],
        "ref_motion_phase": W["ref_motion_phase"], "dif_local_rigid_body_pos": W["dif_local_rigid_body_pos"],
        "local_ref_rigid_body_pos": W["local_ref_rigid_body_pos"], "dr_base_com": S["base_com_bias"],
        "dr_link_mass": S["link_mass_scale"], "dr_kp": S["kp_scale"], "dr_kd": S["kd_scale"],
        "dr_friction": S["friction_coeffs"].reshape(N, -1), "dr_ctrl_delay": S["action_delay_idx"].reshape(N, -1).float(),
        "local_ref_rigid_body_vel": W["local_ref_rigid_body_vel"], "global_ref_rigid_body_vel": W["global_ref_rigid_body_vel"],
        "dif_joint_angles": W["dif_joint_angles"], "dif_joint_velocities": W["dif_joint_velocities"],
        "base_pos_z": inp["root_state"][:, 2:3],
        "relyaw": W["relyaw"], "local_ref_rigid_body_pos_relyaw": W["local_ref_rigid_body_pos_relyaw"],
    }
    for k in ("vr_3point_pos", "future_ref_dof_pos", "future_ref_dof_vel"):
        if k in W:
            src[k] = W[k]
    if spec.general:  # getters of general_tracking.py:879-927; root_height is a live view of the root state (:608)
        K = spec.key_ids
        src.update({
            "roll_pitch": W["roll_pitch"], "local_key_body_pos": W["local_body_pos"][:, K, :].reshape(N, -1),
            "local_key_body_rot": W["local_body_rot"][:, K, ...].reshape(N, -1), "anchor_ref_pos": W["anchor_ref_pos"].reshape(N, -1),
            "anchor_ref_rot": W["anchor_ref_rot"].reshape(N, -1), "contact_mask": W["contacts_filt"],
            "root_height": inp["root_state"][:, 2:3], "dr_base_mass": S["base_mass_scale"],
        })
        for k in ("next_step_ref_motion", "future_motion_root_height", "future_motion_roll_pitch", "future_motion_base_lin_vel",
                  "future_motion_base_yaw_vel", "future_motion_dof_pos", "future_motion_local_ref_key_body_pos"):
            if k in W:
                src[k] = W[k]
    obs = {}
    for g, grp in spec.obs_groups.items():
        parts = []
        for seg in grp["segs"]:
            if seg["kind"] == "src":
                parts.append((src[seg["key"]].clone() + 0.0) * seg["scale"])
            else:
                h = S["hist_" + seg["key"]].clone()[:, :seg["len"]]
                parts.append(h.reshape(N, -1) * seg["scale"])
        obs[g] = torch.clip(torch.cat(parts, dim=-1), -spec.clip_observations, spec.clip_observations)
    hist_obs = {k: src[k].clone() * spec.obs_scales[k] for k in spec.hist_keys}
    # _post_compute_observations_callback (:398-405)
    S["last_actions"][:] = S["actions"]
    S["last_dof_pos"][:] = dof_pos
    S["last_dof_vel"][:] = dof_vel
    S["last_contacts"][:] = W["contacts"]
    for k in spec.hist_keys:  # HistoryHandler.add (history_handler.py:40-44)
        val = S["hist_" + k].clone()
        S["hist_" + k][:, 1:] = val[:, :-1]
        S["hist_" + k][:, 0] = hist_obs[k]
    return obs


def step(spec, tab, S, inp, inject=None):
    """one `_post_physics_step` (legged_robot_base.py:297-338).  `inp` tensors are modified in place where the
    reference writes simulator state at resets (dof_state, root_state).  Returns (obs, log, extras)."""
    log = {}
    S["episode_length_buf"] += 1
    last_ep_len = S["episode_length_buf"].clone()
    S["actions"] = inp["actions"].clone()
    W = _pre_compute(spec, tab, S, inp)
    dp = W["dif_pos"]
    log["upper_body_diff_norm"] = dp[:, spec.upper_ids, :].norm(dim=-1).mean()
    log["lower_body_diff_norm"] = dp[:, spec.lower_ids, :].norm(dim=-1).mean()
    log["vr_3point_diff_norm"] = dp[:, spec.vr3_ids, :].norm(dim=-1).mean()
    log["joint_pos_diff_norm"] = W["dif_joint_angles"].norm(dim=-1).mean()
    if spec.general:  # general_tracking.py:303-335
        dl = W["dif_local_body_pos"]
        log["key_body_diff_norm"] = dp[:, spec.key_ids, :].norm(dim=-1).mean()
        log["local_upper_body_diff_norm"] = dl[:, spec.upper_ids, :].norm(dim=-1).mean()
        log["local_lower_body_diff_norm"] = dl[:, spec.lower_ids, :].norm(dim=-1).mean()
        log["local_vr_3point_diff_norm"] = dl[:, spec.vr3_ids, :].norm(dim=-1).mean()
        log["local_key_body_diff_norm"] = dl[:, spec.key_ids, :].norm(dim=-1).mean()
    pushed = torch.zeros(S["episode_length_buf"].shape[0], dtype=torch.bool)
    if spec.dr.get("push_robots"):
        # _update_counters_each_step (legged_robot_base.py:143-148), the push part of _update_tasks_callback (:382-388) and
        # _push_robots (:1090-1106); the torch.randint / torch_rand_float draws come from `inject`
        S["push_robot_counter"] += 1
        ids = (S["push_robot_counter"] == (S["push_interval_s"] / spec.dt).int()).nonzero(as_tuple=False).flatten()
        S["push_robot_counter"][ids] = 0
        if inject is not None and "push_interval_s" in inject:
            S["push_interval_s"][ids] = inject["push_interval_s"][ids]
            vel = inject["push_vel"][ids]
        else:  # the reference's own draws (torch.randint, torch_rand_float)
            lo_s, hi_s = spec.dr["push_interval_s"]
            S["push_interval_s"][ids] = torch.randint(int(lo_s), int(hi_s), (len(ids),))
            mv = spec.dr["max_push_vel_xy"]
            vel = (mv - (-mv)) * torch.rand(len(ids), 2) + (-mv)
        S["push_robot_vel_buf"][ids] = vel
        if spec.dr.get("push_fixed"):
            inp["root_state"][ids, 7:9] += vel
        else:
            inp["root_state"][ids, 7:9] = vel
        pushed[ids] = True
    _termination(spec, S, W, log, inp, inject)
    _rewards(spec, S, W, inp, log)
    extras = {"ref_body_pos_extend": W["ref"]["rg_pos_t"].clone(), "ref_body_rot_extend": W["ref"]["rg_rot_t"].clone()}
    ep = _reset(spec, tab, S, W, inp, inject, log, last_ep_len)
    if ep is not None:
        extras["episode"] = ep
    extras["time_outs"] = S["time_out_buf"]
    extras["refresh_env_ids"] = torch.logical_or(S["reset_buf"] != 0, pushed).nonzero(as_tuple=False).flatten()  # need_to_refresh_envs, :314-315
    obs = _observations(spec, S, W, inp)
    return obs, log, extras

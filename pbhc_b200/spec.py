"""TrackSpec: everything the fused tracking step needs from the reference's composed config
(`env.config` of `+exp=motion_tracking`, humanoidverse/config/**), flattened into plain Python values.

It is the host-side equivalent of what LeggedRobotBase.__init__/_prepare_reward_function/_init_buffers,
BaseTask._setup_robot_body_indices, LeggedRobotMotionTracking._init_tracking_config/_init_motion_extend
and HistoryHandler.__init__ derive from the config (legged_robot_base.py:28-233, base_task.py:172-205,
motion_tracking.py:203-247, history_handler.py:12-27).  Pure Python, no torch, no CUDA.
"""
import math

# reward terms implemented by csrc/track.cu, in enum order (ids are part of the C ABI)
REWARD_IDS = {name: i for i, name in enumerate([
    "teleop_contact_mask", "teleop_max_joint_position", "teleop_body_position_extend", "teleop_vr_3point",
    "teleop_body_position_feet", "teleop_body_rotation_extend", "teleop_body_ang_velocity_extend",
    "teleop_body_velocity_extend", "teleop_joint_position", "teleop_joint_velocity", "penalty_torques",
    "penalty_action_rate", "feet_air_time", "penalty_feet_contact_forces", "penalty_stumble", "penalty_slippage",
    "limits_dof_pos", "limits_dof_vel", "limits_torque", "collision", "penalty_dof_vel", "penalty_dof_acc",
    "foot_slip_penalty", "alive", "penalty_orientation", "teleop_contact_mask_v2",
    # general_tracking (general_tracking.py:1116-1270)
    "teleop_root_vel", "teleop_anchor_body_position", "teleop_anchor_body_rotation", "local_key_body_position",
    "local_key_body_rotation", "key_body_velocity", "key_body_ang_velocity",
    # zero-scaled in every shipped config; the radial terms are LeggedRobotMotionTracking's
    # (motion_tracking.py:1238-1244, 1286-1293), the feet terms the base class's (legged_robot_base.py:1031-1082)
    "teleop_radial_body_velocity_extend", "teleop_radial_joint_velocity", "feet_heading_alignment", "penalty_feet_ori",
    "feet_heading_alignment_contact", "penalty_feet_ori_contact"])}
MT_ONLY_REWARDS = {"teleop_radial_body_velocity_extend", "teleop_radial_joint_velocity"}
# reward functions of the reference that cannot run there (so there is nothing to be equal to): what they trip over
BROKEN_IN_REFERENCE = {
    "feet_max_height_for_this_air": "`~self.last_contacts_filt` on a float tensor raises (legged_robot_base.py:68, 1022)",
    "teleop_key_body_position": "no config defines reward_tracking_sigma.teleop_key_body_pos (general_tracking.py:1112)",
    "teleop_root_pose": "no config defines reward_tracking_sigma.teleop_root_pose (general_tracking.py:1276)"}

# adaptive-sigma slots (config.rewards.reward_tracking_sigma keys), in enum order
SIGMA_IDS = {name: i for i, name in enumerate([
    "teleop_max_joint_pos", "teleop_upper_body_pos", "teleop_lower_body_pos", "teleop_vr_3point_pos",
    "teleop_feet_pos", "teleop_body_rot", "teleop_body_vel", "teleop_body_ang_vel", "teleop_joint_pos",
    "teleop_joint_vel",
    "teleop_root_vel", "teleop_anchor_body_pos", "teleop_anchor_body_rot", "local_key_body_pos", "local_key_body_rot",
    "key_body_vel", "key_body_ang_vel"])}

# observation sources (getters `_get_obs_<key>`), enum order; dims as functions of (D, J, L)
OBS_SRC = ["base_lin_vel", "base_ang_vel", "projected_gravity", "dof_pos", "dof_vel", "actions", "ref_motion_phase",
           "dif_local_rigid_body_pos", "local_ref_rigid_body_pos", "dr_base_com", "dr_link_mass", "dr_kp", "dr_kd",
           "dr_friction", "dr_ctrl_delay", "local_ref_rigid_body_vel", "global_ref_rigid_body_vel", "dif_joint_angles",
           "dif_joint_velocities", "base_pos_z",
           # general_tracking getters (general_tracking.py:879-927, legged_robot_base.py:1193)
           "roll_pitch", "local_key_body_pos", "local_key_body_rot", "anchor_ref_pos", "anchor_ref_rot", "next_step_ref_motion",
           "contact_mask", "root_height", "dr_base_mass", "future_motion_root_height", "future_motion_roll_pitch",
           "future_motion_base_lin_vel", "future_motion_base_yaw_vel", "future_motion_dof_pos",
           "future_motion_local_ref_key_body_pos",
           # both env classes (motion_tracking.py:950-961 / general_tracking.py:837-850); warp-per-env kernel
           "relyaw", "local_ref_rigid_body_pos_relyaw", "vr_3point_pos",
           # LeggedRobotMotionTracking with obs.future_ref_steps > 0 (motion_tracking.py:594-599, 986-990)
           "future_ref_dof_pos", "future_ref_dof_vel"]
GENERAL_ONLY_SRC = set(OBS_SRC[OBS_SRC.index("roll_pitch"):OBS_SRC.index("relyaw")])
MT_ONLY_SRC = {"future_ref_dof_pos", "future_ref_dof_vel"}
OBS_SRC_IDS = {k: i for i, k in enumerate(OBS_SRC)}


def obs_src_dim(key, D, J, L, K=0, F=0):
    """K = number of key bodies, F = future_num_steps (general_tracking only)."""
    return {"roll_pitch": 2, "local_key_body_pos": 3 * K, "local_key_body_rot": 6 * K, "anchor_ref_pos": 3, "anchor_ref_rot": 6,
            "next_step_ref_motion": 7 + D + 3 * K, "contact_mask": 2, "root_height": 1, "dr_base_mass": 1,
            "relyaw": 1, "local_ref_rigid_body_pos_relyaw": 3 * J, "vr_3point_pos": 9, "future_ref_dof_pos": F * D, "future_ref_dof_vel": F * D,
            "future_motion_root_height": F, "future_motion_roll_pitch": 2 * F, "future_motion_base_lin_vel": 3 * F,
            "future_motion_base_yaw_vel": F, "future_motion_dof_pos": F * D, "future_motion_local_ref_key_body_pos": F * 3 * K,
            "base_lin_vel": 3, "base_ang_vel": 3, "projected_gravity": 3, "dof_pos": D, "dof_vel": D, "actions": D,
            "ref_motion_phase": 1, "dif_local_rigid_body_pos": 3 * J, "local_ref_rigid_body_pos": 3 * J, "dr_base_com": 3,
            "dr_link_mass": L, "dr_kp": D, "dr_kd": D, "dr_friction": 1, "dr_ctrl_delay": 1,
            "local_ref_rigid_body_vel": 3 * J, "global_ref_rigid_body_vel": 3 * J, "dif_joint_angles": D,
            "dif_joint_velocities": D, "base_pos_z": 1}[key]


def _get(cfg, path, default=None):
    node = cfg
    for k in path.split("."):
        if node is None:
            return default
        if isinstance(node, dict):
            node = node.get(k, None)
        else:
            node = getattr(node, k, None)
    return default if node is None else node


class TrackSpec:
    """see module docstring.  Build with TrackSpec.from_config(env_config)."""

    @classmethod
    def from_config(cls, cfg, env_class=None):
        """cfg: the composed `env.config`.  env_class: "motion_tracking" (LeggedRobotMotionTracking) or "general_tracking"
        (LeggedRobotGeneralTracking) - what hydra's `env._target_` selects; defaults to cfg["env_class"] if present."""
        s = cls()
        s.env_class = env_class or _get(cfg, "env_class", "motion_tracking")
        if s.env_class not in ("motion_tracking", "general_tracking"):
            raise ValueError(f"unknown env class {s.env_class}")
        s.general = s.env_class == "general_tracking"
        robot, rw, obs, dr = _get(cfg, "robot"), _get(cfg, "rewards"), _get(cfg, "obs"), _get(cfg, "domain_rand")
        s.body_names = list(_get(robot, "body_names"))
        s.dof_names = list(_get(robot, "dof_names"))
        s.Jb, s.D = len(s.body_names), len(s.dof_names)
        ext = list(_get(robot, "motion.extend_config", []) or [])
        s.ext_parent = [s.body_names.index(_get(e, "parent_name")) for e in ext]
        s.ext_pos = [[float(v) for v in _get(e, "pos")] for e in ext]
        s.ext_rot_wxyz = [[float(v) for v in _get(e, "rot")] for e in ext]
        s.body_list = s.body_names + [_get(e, "joint_name") for e in ext]  # simulator._body_list after _init_motion_extend
        s.J = len(s.body_list)
        s.dt = _get(cfg, "simulator.config.sim.control_decimation") * (1.0 / _get(cfg, "simulator.config.sim.fps"))
        s.max_episode_length = math.ceil(_get(cfg, "max_episode_length_s") / s.dt)  # base_task.py:39
        s.max_episode_length_s = float(_get(cfg, "max_episode_length_s"))
        # body index sets (base_task.py:172-184, motion_tracking.py:203-211)
        s.feet_ids = [i for i, n in enumerate(s.body_names) if _get(robot, "foot_name") in n]
        pen = []
        for name in _get(robot, "penalize_contacts_on", []) or []:
            pen.extend(i for i, n in enumerate(s.body_names) if name in n)
        s.penalised_ids = pen
        s.vr3_ids = [s.body_list.index(n) for n in _get(robot, "motion.motion_tracking_link", [])]
        s.lower_ids = [s.body_list.index(n) for n in _get(robot, "motion.lower_body_link", [])]
        s.upper_ids = [s.body_list.index(n) for n in _get(robot, "motion.upper_body_link", [])]
        # general_tracking.py:86-98: key bodies; anchor = find_rigid_body_indice(anchor_link) + 1 (not found -> -1 -> body 0)
        s.key_ids = [s.body_list.index(n) for n in (_get(robot, "key_bodies", []) or [])] if s.general else []
        anchor_link = _get(robot, "motion.anchor_link", "pelvis_link")
        s.anchor_index = (s.body_names.index(anchor_link) if anchor_link in s.body_names else -1) + 1
        s.default_joint_angles = [float(_get(robot, "init_state.default_joint_angles")[n]) for n in s.dof_names]
        s.hard_dof_lo = [float(v) for v in _get(robot, "dof_pos_lower_limit_list")]
        s.hard_dof_hi = [float(v) for v in _get(robot, "dof_pos_upper_limit_list")]
        s.dof_vel_limit = [float(v) for v in _get(robot, "dof_vel_limit_list")]
        s.torque_limit = [float(v) for v in _get(robot, "dof_effort_limit_list")]
        s._from_rewards(cfg, rw)
        s._from_termination(cfg)
        s._from_obs(cfg, obs, dr)
        s.dr = dict(randomize_pd_gain=bool(_get(dr, "randomize_pd_gain", False)), kp_range=list(_get(dr, "kp_range", [1, 1])),
                    kd_range=list(_get(dr, "kd_range", [1, 1])), randomize_rfi_lim=bool(_get(dr, "randomize_rfi_lim", False)),
                    rfi_lim_range=list(_get(dr, "rfi_lim_range", [1, 1])), use_rao=bool(_get(dr, "use_rao", False)),
                    rao_lim=float(_get(dr, "rao_lim", 0.0)), randomize_ctrl_delay=bool(_get(dr, "randomize_ctrl_delay", False)),
                    ctrl_delay_step_range=list(_get(dr, "ctrl_delay_step_range", [0, 0])),
                    # legged_robot_base.py:134-147, 382-388, 1090-1106 (on in the shipped domain_rand/main.yaml)
                    push_robots=bool(_get(dr, "push_robots", False)), push_interval_s=list(_get(dr, "push_interval_s", [5, 10])),
                    max_push_vel_xy=float(_get(dr, "max_push_vel_xy", 0.0)), push_fixed=bool(_get(dr, "_push_fixed", False)))
        s.noise_to_initial_level = float(_get(cfg, "noise_to_initial_level", 0))
        s.init_noise_scale = {k: float(v) for k, v in dict(_get(cfg, "init_noise_scale", {}) or {}).items()}
        s.use_vec_reward = bool(_get(cfg, "use_vec_reward", False))
        # options that change the step in the reference and are not built here: refuse them instead of ignoring them
        for path, what in (("use_teleop_control", "teleoperation markers instead of the motion's (motion_tracking.py:112, 725-728)"),
                           ("soft_dynamic_correction.enable", "soft dynamic correction (motion_tracking.py:133, 840-870)"),
                           ("domain_rand.parallel_serial_pd.enable", "parallel / serial PD gain randomisation (legged_robot_base.py:607-623)"),
                           ("domain_rand.randomize_default_dof_pos", "default_dof_pos randomisation at resets (legged_robot_base.py:632-635)"),
                           ("domain_rand.born_offset", "born offset"), ("domain_rand.born_heading_randomization", "born heading randomisation"),
                           ("domain_rand.motion_package_loss", "motion package loss"), ("domain_rand.randomize_motion_ref_xyz", "reference xyz randomisation")):
            if _get(cfg, path, False):
                raise NotImplementedError(f"{path}: {what} is not implemented by the B200 tracking step")
        if float(_get(dr, "reinit_epis_rand", -1) or -1) > 0:
            raise NotImplementedError("domain_rand.reinit_epis_rand > 0 is not implemented by the B200 tracking step")
        s.clip_observations = float(_get(cfg, "normalization.clip_observations"))
        return s

    # ---- rewards (legged_robot_base.py:167-233) --------------------------------------------------------
    def _from_rewards(self, cfg, rw):
        self.reward_names, self.reward_scales = [], []
        self.termination_scale = None
        for name, scale in dict(_get(rw, "reward_scales")).items():
            if scale == 0:
                continue
            scale = scale * self.dt
            if name == "termination":
                self.termination_scale = scale
                continue
            if name in BROKEN_IN_REFERENCE:
                raise NotImplementedError(f"reward term '{name}' cannot run in the reference: {BROKEN_IN_REFERENCE[name]}")
            if name not in REWARD_IDS:
                raise NotImplementedError(f"reward term '{name}' is not implemented by the B200 tracking step")
            if name in MT_ONLY_REWARDS and self.general:
                raise NotImplementedError(f"reward term '{name}' exists only in motion_tracking")
            general_only = REWARD_IDS["teleop_root_vel"] <= REWARD_IDS[name] <= REWARD_IDS["key_body_ang_velocity"]
            if general_only and not self.general:
                raise NotImplementedError(f"reward term '{name}' exists only in general_tracking")
            if self.general and name in ("teleop_body_rotation_extend", "teleop_body_velocity_extend", "teleop_body_ang_velocity_extend"):
                raise NotImplementedError(f"reward term '{name}' is not implemented for general_tracking")
            self.reward_names.append(name)
            self.reward_scales.append(scale)
        self.episode_sum_names = [n for n, sc in dict(_get(rw, "reward_scales")).items() if sc != 0]
        self.only_positive_rewards = bool(_get(rw, "only_positive_rewards", False))
        pen_names = set(_get(rw, "reward_penalty_reward_names", []) or [])
        self.penalty_curriculum = bool(_get(rw, "reward_penalty_curriculum", False))
        self.reward_is_penalty = [bool(self.penalty_curriculum and n in pen_names) for n in self.reward_names]
        self.penalty = dict(initial=float(_get(rw, "reward_initial_penalty_scale", 1.0)), lo=float(_get(rw, "reward_min_penalty_scale", 0.0)),
                            hi=float(_get(rw, "reward_max_penalty_scale", 1.0)), down=float(_get(rw, "reward_penalty_level_down_threshold", 0)),
                            up=float(_get(rw, "reward_penalty_level_up_threshold", 0)), degree=float(_get(rw, "reward_penalty_degree", 0)))
        self.num_compute_average_epl = float(_get(rw, "num_compute_average_epl"))
        self.sigma0 = {k: float(v) for k, v in dict(_get(rw, "reward_tracking_sigma", {}) or {}).items()}
        ad = _get(rw, "adaptive_tracking_sigma")
        self.adaptive = dict(enable=bool(_get(ad, "enable", False)), alpha=float(_get(ad, "alpha", 0.0)),
                             scale=float(_get(ad, "scale", 1.0)), type=str(_get(ad, "type", "origin")))
        self.w_lower = float(_get(rw, "teleop_body_pos_lowerbody_weight", 1.0))
        self.w_upper = float(_get(rw, "teleop_body_pos_upperbody_weight", 1.0))
        self.desired_feet_air_time = float(_get(rw, "desired_feet_air_time", 0.0))
        self.max_contact_force = float(_get(rw, "locomotion_max_contact_force", 0.0))
        lim = _get(rw, "reward_limit")
        cur = _get(lim, "reward_limits_curriculum")
        self.soft_dof_pos_limit = float(_get(lim, "soft_dof_pos_limit"))
        self.soft_dof_vel_limit = float(_get(lim, "soft_dof_vel_limit"))
        self.soft_torque_limit = float(_get(lim, "soft_torque_limit"))
        self.limit_cur = {}
        for k in ("dof_pos", "dof_vel", "torque"):
            self.limit_cur[k] = dict(enable=bool(_get(cur, f"soft_{k}_curriculum", False)),
                                     initial=float(_get(cur, f"soft_{k}_initial_limit", 1.0)), lo=float(_get(cur, f"soft_{k}_min_limit", 0)),
                                     hi=float(_get(cur, f"soft_{k}_max_limit", 1)), degree=float(_get(cur, f"soft_{k}_curriculum_degree", 0)),
                                     down=float(_get(cur, f"soft_{k}_curriculum_level_down_threshold", 0)),
                                     up=float(_get(cur, f"soft_{k}_curriculum_level_up_threshold", 0)))

    # ---- termination (legged_robot_base.py:408-489, motion_tracking.py:330-358) -------------------------
    def _from_termination(self, cfg):
        t = _get(cfg, "termination")
        if _get(t, "terminate_when_dof_far", False):
            raise NotImplementedError("termination 'terminate_when_dof_far' is not implemented by the B200 tracking step")
        # base-class terminations (legged_robot_base.py:433-476)
        ts_, tp = _get(cfg, "termination_scales"), _get(cfg, "termination_probality")
        self.term_contact = bool(_get(t, "terminate_by_contact", False))
        self.term_low_height = bool(_get(t, "terminate_by_low_height", False))
        self.term_dof_pos = bool(_get(t, "terminate_when_close_to_dof_pos_limit", False))
        self.term_dof_vel = bool(_get(t, "terminate_when_close_to_dof_vel_limit", False))
        self.term_torque = bool(_get(t, "terminate_when_close_to_torque_limit", False))
        self.term_min_height = float(_get(ts_, "termination_min_base_height", 0.0))
        self.term_dof_pos_scale = float(_get(ts_, "termination_close_to_dof_pos_limit", 1.0))
        self.term_dof_vel_scale = float(_get(ts_, "termination_close_to_dof_vel_limit", 1.0))
        self.term_torque_scale = float(_get(ts_, "termination_close_to_torque_limit", 1.0))
        self.term_dof_pos_prob = float(_get(tp, "terminate_when_close_to_dof_pos_limit", 1.0))
        self.term_dof_vel_prob = float(_get(tp, "terminate_when_close_to_dof_vel_limit", 1.0))
        self.term_torque_prob = float(_get(tp, "terminate_when_close_to_torque_limit", 1.0))
        robot = _get(cfg, "robot")
        ids = []
        for name in _get(robot, "terminate_after_contacts_on", []) or []:  # base_task.py:179-197: every body whose name contains it
            ids.extend(i for i, n in enumerate(self.body_names) if name in n)
        self.term_contact_ids = ids
        self.term_gravity = bool(_get(t, "terminate_by_gravity", False))
        self.term_motion_far = bool(_get(t, "terminate_when_motion_far", False))
        self.term_motion_end = bool(_get(t, "terminate_when_motion_end", False))
        self.term_gravity_thr = float(_get(cfg, "termination_scales.termination_gravity", 0.0))
        # general_tracking.py:241-255
        g = self.general
        self.term_ref_pos_z = bool(g and _get(t, "terminate_by_ref_pos_z", False))
        self.term_ref_ori = bool(g and _get(t, "terminate_by_ref_ori", False))
        self.term_body_z = bool(g and _get(t, "terminate_by_body_z", False))
        self.term_ref_pos_z_thr = float(_get(cfg, "termination_scales.terminate_by_ref_pos_z_threshold", 0.25))
        self.term_ref_ori_thr = float(_get(cfg, "termination_scales.terminate_by_ref_ori_threshold", 0.8))
        self.term_body_z_thr = float(_get(cfg, "termination_scales.terminate_by_body_z_threshold", 0.25))
        self.body_z_ids = [4, 10, 24, 25, 26]  # hard-coded in the reference (:253)
        tc = _get(cfg, "termination_curriculum")
        self.far_cur = dict(enable=bool(self.term_motion_far and _get(tc, "terminate_when_motion_far_curriculum", False)),
                            initial=float(_get(tc, "terminate_when_motion_far_initial_threshold", 0)),
                            hi=float(_get(tc, "terminate_when_motion_far_threshold_max", 0)),
                            lo=float(_get(tc, "terminate_when_motion_far_threshold_min", 0)),
                            degree=float(_get(tc, "terminate_when_motion_far_curriculum_degree", 0)),
                            down=float(_get(tc, "terminate_when_motion_far_curriculum_level_down_threshold", 0)),
                            up=float(_get(tc, "terminate_when_motion_far_curriculum_level_up_threshold", 0)))
        self.motion_far_threshold0 = (self.far_cur["initial"] if self.far_cur["enable"]
                                      else float(_get(cfg, "termination_scales.termination_motion_far_threshold", 0)))

    # ---- observations + history (legged_robot_base.py:763-793, helpers.py:128-152, history_handler.py) ---
    def _from_obs(self, cfg, obs, dr):
        D, J = self.D, self.J
        self.L = len(_get(dr, "randomize_link_body_names", []) or [])
        K = len(self.key_ids)
        # general_tracking: obs.future_num_steps targets; motion_tracking: obs.future_ref_steps reference dof frames (:594-599)
        self.future_num_steps = int(_get(obs, "future_num_steps", 0) or 0) if self.general else int(_get(obs, "future_ref_steps", 0) or 0)
        self.future_max_steps = int(_get(obs, "future_max_steps", 0) or 0)
        F = self.future_num_steps
        if F > 0 and self.general:  # torch.linspace(1, max, F, dtype=long) (general_tracking.py:501-507), evaluated by torch itself
            import torch
            self.future_steps = [int(v) for v in torch.linspace(start=1, end=self.future_max_steps, steps=F, dtype=torch.long)]
        else:
            self.future_steps = []
        if bool(_get(obs, "masked_dof_vel", False)):
            raise NotImplementedError("obs.masked_dof_vel is not implemented by the B200 tracking step")
        self.obs_scales = {k: float(v) for k, v in dict(_get(obs, "obs_scales")).items()}
        self.noise_scales = {k: float(v) for k, v in dict(_get(obs, "noise_scales")).items()}
        aux = {a: dict(c) for a, c in dict(_get(obs, "obs_auxiliary", {}) or {}).items()}
        self.obs_aux = aux
        self.hist_len = {}
        for a, c in aux.items():  # history_handler.py:18-24 (insertion order of first appearance)
            for k, n in c.items():
                self.hist_len[k] = max(self.hist_len.get(k, 0), int(n))
        self.hist_keys = list(self.hist_len.keys())
        self.hist_dim = {k: obs_src_dim(k, D, J, self.L, K, F) for k in self.hist_keys}
        self.obs_groups = {}
        for g, keys in dict(_get(obs, "obs_dict")).items():
            segs, off = [], 0
            for key in sorted(keys):
                if key.endswith("_raw"):  # parse_observation stores the value under the stripped key and _post_config_observation_callback
                    # then looks up the "_raw" name (helpers.py:138-152, legged_robot_base.py:788-793): KeyError in the reference
                    raise NotImplementedError("'_raw' observation keys cannot run in the reference (KeyError in _post_config_observation_callback)")
                if key in aux:
                    for hk in sorted(aux[key].keys()):
                        n = int(aux[key][hk]) * self.hist_dim[hk]
                        segs.append(dict(kind="hist", key=hk, len=int(aux[key][hk]), dim=self.hist_dim[hk], off=off,
                                         scale=self.obs_scales[key], group_key=key))
                        off += n
                else:
                    if key in GENERAL_ONLY_SRC and not self.general:
                        raise NotImplementedError(f"observation '{key}' exists only in general_tracking")
                    if key in MT_ONLY_SRC and (self.general or F <= 0):
                        raise NotImplementedError(f"observation '{key}' needs motion_tracking with obs.future_ref_steps > 0")
                    if key == "vr_3point_pos" and len(self.vr3_ids) != 3:
                        raise ValueError("vr_3point_pos: the reference pairs 3 points per env with `heading_inv_rot.repeat(3, 1)`")
                    n = obs_src_dim(key, D, J, self.L, K, F)
                    segs.append(dict(kind="src", key=key, dim=n, off=off, scale=self.obs_scales[key],
                                     noise=self.noise_scales[key]))
                    off += n
            self.obs_groups[g] = dict(segs=segs, dim=off)
        # observation-noise curriculum (legged_robot_base.py:119-120, 637-645; the thresholds are the reference's: the level-down
        # one from obs.*, the level-up one from rewards.reward_penalty_level_up_threshold)
        self.add_noise_curriculum = bool(_get(obs, "add_noise_currculum", False))
        self.noise_cur = dict(enable=self.add_noise_curriculum, initial=float(_get(obs, "noise_initial_value", 1.0)),
                              lo=float(_get(obs, "noise_value_min", 0.0)), hi=float(_get(obs, "noise_value_max", 1.0)),
                              degree=float(_get(obs, "soft_dof_pos_curriculum_degree", 0.0)),
                              down=float(_get(obs, "soft_dof_pos_curriculum_level_down_threshold", 0.0)),
                              up=float(_get(cfg, "rewards.reward_penalty_level_up_threshold", 0.0)))
        if _get(obs, "noise_process.enable", False):
            raise NotImplementedError("obs.noise_process is not implemented by the B200 tracking step")

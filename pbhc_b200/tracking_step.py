"""TrackingStep: the fused motion-tracking env step behind the reference's env attributes.

It replaces, for LeggedRobotMotionTracking, everything `_post_physics_step` does between
`_refresh_sim_tensors()` and `extras["to_log"]` (legged_robot_base.py:297-338): one launch of pbhc_track_step
(csrc/track.cu) + pbhc_track_finalize, no host synchronisation.  State the reference keeps as env attributes
is exposed under the same names as views of the packed device buffers (episode_length_buf,
motion_start_times, motion_len, feet_air_time, last_contacts, episode_sums[name], rew_buf, reset_buf,
time_out_buf, obs_buf_dict, log_dict, history_handler.query(key), ...).  No CPU fallback.
"""
import ctypes as C
import os

import torch

from . import _lib
from .spec import REWARD_IDS, SIGMA_IDS, OBS_SRC, OBS_SRC_IDS, obs_src_dim


def _mask(ids):
    m = 0
    for i in ids:
        m |= 1 << int(i)
    return m


class _History:
    """HistoryHandler view (env_utils/history_handler.py): query(key) -> [N, H, dim] newest-first."""

    def __init__(self, step):
        self._s = step

    @property
    def history(self):
        return {k: self.query(k) for k in self._s.spec.hist_keys}

    def query(self, key):
        s = self._s
        k = s.spec.hist_keys.index(key)
        H, dim, base = s.spec.hist_len[key], s.spec.hist_dim[key], s._hist_base[k]
        raw = s.hist[:, base:base + H * dim].view(-1, H, dim)
        pushes = int(s.scalars[_lib.S_PUSH_COUNT].item())
        head = (H - pushes % H) % H
        order = [(head + i) % H for i in range(H)]
        return raw[:, order].clone()


class TrackingStep:
    def __init__(self, spec, motion_lib, num_envs, device, seed=0, env_offset=0):
        self.spec, self.lib, self.N = spec, motion_lib, int(num_envs)
        self.device = dev = torch.device(device)
        if dev.type != "cuda":
            raise _lib.PbhcError("TrackingStep needs a CUDA device (there is no CPU fallback)")
        self._L = _lib.lib()
        s, N, D = spec, self.N, spec.D
        J, Jb, Cm = motion_lib._dims[0], motion_lib._dims[1], motion_lib._dims[3]
        if (J, Jb, motion_lib._dims[2]) != (s.J, s.Jb, D):
            raise _lib.PbhcError("motion library and config disagree on the skeleton")
        if len(s.feet_ids) != 2:
            raise NotImplementedError("the B200 tracking step supports exactly two feet")
        f32 = dict(device=dev, dtype=torch.float32)
        self.n_sums = len(s.reward_names) + (1 if s.termination_scale is not None else 0)
        self._sum_names = list(s.reward_names) + (["termination"] if s.termination_scale is not None else [])
        # persistent state
        self.episode_length_buf = torch.zeros(N, device=dev, dtype=torch.int64)
        self.env_f32 = torch.zeros(N, _lib.ENVF_STRIDE, **f32)
        self.last_actions, self.last_dof_pos, self.last_dof_vel = (torch.zeros(N, D, **f32) for _ in range(3))
        self._kp_scale, self._kd_scale, self._rfi_lim_scale, self._rao_scale = (torch.ones(N, D, **f32) for _ in range(4))
        self.action_delay_idx = torch.zeros(N, device=dev, dtype=torch.int64)
        self.default_dof_pos = torch.tensor(s.default_joint_angles, **f32).repeat(N, 1)
        self._hist_base, off = [], 0
        for k in s.hist_keys:
            self._hist_base.append(off)
            off += s.hist_len[k] * s.hist_dim[k]
        self.hist_row = off
        self.hist = torch.zeros(N, max(off, 1), **f32)
        self._episode_sums = torch.zeros(N, self.n_sums, **f32)
        self.last_episode_sums = torch.zeros(N, self.n_sums, **f32)
        self.scalars = torch.zeros(_lib.SCALARS, device=dev, dtype=torch.float64)
        init = [0.0] * _lib.SCALARS
        for name, v in s.sigma0.items():
            if name in SIGMA_IDS:
                init[_lib.S_SIGMA + SIGMA_IDS[name]] = v
                init[_lib.S_EMA + SIGMA_IDS[name]] = v
        init[_lib.S_PENALTY] = s.penalty["initial"] if s.penalty_curriculum else 1.0  # unused (never applied) without the curriculum
        init[_lib.S_FAR_THR] = s.motion_far_threshold0
        init[_lib.S_SOFT_DOF_POS] = s.limit_cur["dof_pos"]["initial"]
        init[_lib.S_SOFT_DOF_VEL] = s.limit_cur["dof_vel"]["initial"]
        init[_lib.S_SOFT_TORQUE] = s.limit_cur["torque"]["initial"]
        init[_lib.S_NOISE_CUR] = s.noise_cur["initial"]
        self.scalars.copy_(torch.tensor(init, dtype=torch.float64))
        # outputs
        self.obs_buf_dict = {g: torch.zeros(N, v["dim"], **f32) for g, v in s.obs_groups.items()}
        self.rew_buf = torch.zeros(N, len(s.reward_names) + 1, **f32) if s.use_vec_reward else torch.zeros(N, **f32)
        self.reset_buf = torch.ones(N, device=dev, dtype=torch.int64)
        self.time_out_buf = torch.zeros(N, device=dev, dtype=torch.bool)
        self.ref_body_pos_extend = torch.zeros(N, J, 3, **f32)
        self.ref_body_rot_extend = torch.zeros(N, J, 4, **f32)
        self.partials = torch.zeros(self._L.pbhc_track_max_blocks(), _lib.ACC, device=dev, dtype=torch.float64)
        self.acc = torch.zeros(_lib.ACC, device=dev, dtype=torch.float64)
        self.history_handler = _History(self)
        self.extras = {}
        # boundary state of _post_physics_step (legged_robot_base.py:302, 311-319, 509-515)
        self.last_episode_length_buf = torch.zeros(N, device=dev, dtype=torch.int64)
        self._force_reset = torch.zeros(N, device=dev, dtype=torch.uint8)
        self._force_pending = False
        self.inject_next = None  # values for the reset draws of the next step() only (tests: the reference's own draws)
        self._reset_ids = torch.zeros(N, device=dev, dtype=torch.int32)
        self._reset_count = torch.zeros(1, device=dev, dtype=torch.int32)
        self._compact_scratch = torch.zeros(max(int(self._L.pbhc_compact_scratch_ints(N)), 1), device=dev, dtype=torch.int32)
        self._build_cfg(seed)
        # push_robots: per-env counter and interval live in the spare slot of the env_f32 row (counter | interval_s << 16 as int32 bits)
        self._refresh_buf = self.push_robot_vel_buf = None
        if s.dr["push_robots"]:
            lo, hi = s.dr["push_interval_s"]
            self.set_push_state(torch.zeros(N, device=dev, dtype=torch.int32), torch.randint(int(lo), int(hi), (N,), device=dev))  # legged_robot_base.py:135, 139
            self._refresh_buf = torch.zeros(N, device=dev, dtype=torch.int64)    # step stamps (PbhcTrackIO.refresh_buf)
            self._refresh_flags = torch.zeros(N, device=dev, dtype=torch.int64)  # the mask of the last step, for pbhc_compact_ids
            self.push_robot_vel_buf = torch.zeros(N, 2, **f32)
            self._refresh_ids = torch.zeros(N, device=dev, dtype=torch.int32)
            self._refresh_count = torch.zeros(1, device=dev, dtype=torch.int32)
        used = int(self._cfg.src_used_mask)
        self._heading_scratch = torch.zeros(N, 2, **f32) if used >> OBS_SRC_IDS["vr_3point_pos"] & 1 else None
        if used >> OBS_SRC_IDS["relyaw"] & 3:
            # ref_init_rpy (motion_tracking.py:186-187): the env is built with episode_length_buf = 0 and motion_start_times = 0,
            # so kick_motion_res() reads motion 0 at time 1 * dt; euler yaw of that root rotation (rotations.py:381-385)
            q = motion_lib.get_motion_state(torch.zeros(1, device=dev, dtype=torch.int64), torch.full((1,), float(s.dt), **f32))["root_rot"][0].float()
            siny = 2.0 * (q[3] * q[2] + q[0] * q[1])
            cosy = q[3] * q[3] + q[0] * q[0] - q[1] * q[1] - q[2] * q[2]
            self.set_ref_init_yaw(float(torch.atan2(siny, cosy)))
        self._io = _lib.TrackIO()
        self._io.env_offset = int(env_offset)  # global id of this rank's env 0 (pbhc_b200.dist.shard_range): keys the in-kernel random streams
        self._keep = None

    def set_is_evaluating(self, enforce_randomize_motion_start_eval=False):
        """LeggedRobotMotionTracking.set_is_evaluating (motion_tracking.py:327-328): from now on a reset starts its clip at time 0
        instead of a sampled time (`_resample_motion_times`, :374-377) unless config.enforce_randomize_motion_start_eval."""
        self.is_evaluating = True
        self._cfg.start_time_zero = 0 if enforce_randomize_motion_start_eval else 1

    def set_push_state(self, counter, interval_s):
        """push_robot_counter [N] and push_interval_s [N] (legged_robot_base.py:135, 139) into the packed per-env state."""
        thr = (interval_s.to(self.device).long() / self.spec.dt).int()  # (self.push_interval_s / self.dt).int(), legged_robot_base.py:384
        word = (counter.to(self.device).to(torch.int32) & 0xffff) | (thr << 16)
        self.env_f32[:, 7] = word.view(torch.float32)

    push_robot_counter = property(lambda self: self.env_f32[:, 7].contiguous().view(torch.int32) & 0xffff)

    @property
    def push_interval_s(self):
        """[N] int64: the interval whose step count (interval / dt).int() the env's push word holds"""
        lo, hi = self.spec.dr["push_interval_s"]
        cand = torch.arange(int(lo), int(hi), device=self.device)
        thr_of = (cand / self.spec.dt).int()
        thr = self.env_f32[:, 7].contiguous().view(torch.int32) >> 16
        return cand[(thr[:, None] == thr_of[None, :]).int().argmax(dim=1)]

    def set_ref_init_yaw(self, yaw):
        """ref_init_rpy[0, 2] of the reference env (`relyaw = rpy[:, 2:3] - ref_init_rpy[0, 2]`, motion_tracking.py:686)."""
        self.ref_init_yaw = float(yaw)
        self._cfg.ref_init_yaw = float(yaw)

    # ---- reference attribute names as views ------------------------------------------------------------
    motion_start_times = property(lambda self: self.env_f32[:, 0])
    motion_len = property(lambda self: self.env_f32[:, 1])
    feet_air_time = property(lambda self: self.env_f32[:, 2:4])
    last_contacts = property(lambda self: self.env_f32[:, 4:6])
    end_time_ratio_buf = property(lambda self: self.env_f32[:, 6])

    @property
    def episode_sums(self):
        return {n: self._episode_sums[:, i] for i, n in enumerate(self._sum_names)}

    @property
    def log_dict(self):
        vals = self.scalars[_lib.S_LOG:_lib.S_LOG + len(_lib.LOG_NAMES)].float()
        d = {n: vals[i] for i, n in enumerate(_lib.LOG_NAMES)}
        b = _lib.S_LOG
        sig = self.scalars[b + _lib.L_ADP_SIGMA:b + _lib.L_ADP_SIGMA + _lib.NUM_SIGMA].float()
        ema = self.scalars[b + _lib.L_ERROR_EMA:b + _lib.L_ERROR_EMA + _lib.NUM_SIGMA].float()
        for name, i in SIGMA_IDS.items():  # logged before the step's update, like _log_motion_tracking_info
            d["adp_sigma_" + name], d["error_ema_" + name] = sig[i], ema[i]
        if self.spec.general:  # general_tracking.py:303-335, 241-255
            for name, i in _lib.LOG_NAMES_GENERAL.items():
                d[name] = self.scalars[b + i].float()
        sp = self.spec
        for on, name in ((sp.term_contact, "terminate_by_contact"), (sp.term_low_height, "terminate_by_low_height"),
                         (sp.term_dof_pos, "terminate_by_dof_pos_limit"), (sp.term_dof_vel, "terminate_by_dof_vel_limit"),
                         (sp.term_torque, "terminate_by_torque_limit")):
            if on:  # legged_robot_base.py:425-429
                d[name] = self.scalars[b + _lib.LOG_NAMES_TERM[name]].float()
        if sp.add_noise_curriculum:  # legged_robot_base.py:760-761
            d["current_noise_curriculum_value"] = self.scalars[b + _lib.L_NOISE_CUR].float()
        return d

    def reward_tracking_sigma(self):
        """config.rewards.reward_tracking_sigma as the reference mutates it (synchronises)."""
        sig = self.scalars[_lib.S_SIGMA:_lib.S_SIGMA + _lib.NUM_SIGMA].cpu().tolist()
        return {name: sig[i] for name, i in SIGMA_IDS.items()}

    # ---- PbhcTrackCfg from the spec ---------------------------------------------------------------------
    def _build_cfg(self, seed):
        s, dev = self.spec, self.device
        c = _lib.TrackCfg()
        c.J, c.Jb, c.D, c.L, c.n_ext, c.n_feet = s.J, s.Jb, s.D, s.L, s.J - s.Jb, len(s.feet_ids)
        for i, p in enumerate(s.ext_parent):
            c.ext_parent[i] = p
            for k in range(3):
                c.ext_pos[i][k] = s.ext_pos[i][k]
            w, x, y, z = s.ext_rot_wxyz[i]
            for k, v in enumerate((x, y, z, w)):
                c.ext_rot_xyzw[i][k] = v
        for i, f in enumerate(s.feet_ids):
            c.feet_ids[i] = f
        c.upper_mask, c.lower_mask, c.vr3_mask = _mask(s.upper_ids), _mask(s.lower_ids), _mask(s.vr3_ids)
        c.penalised_mask = _mask(s.penalised_ids)
        c.n_upper, c.n_lower, c.n_vr3 = len(s.upper_ids), len(s.lower_ids), len(s.vr3_ids)
        c.dt, c.max_episode_length = s.dt, int(s.max_episode_length)
        c.n_rewards, c.use_vec_reward = len(s.reward_names), int(s.use_vec_reward)
        c.only_positive_rewards, c.has_termination = int(s.only_positive_rewards), int(s.termination_scale is not None)
        for i, n in enumerate(s.reward_names):
            c.reward_id[i], c.reward_scale[i], c.reward_is_penalty[i] = REWARD_IDS[n], s.reward_scales[i], int(s.reward_is_penalty[i])
        c.termination_scale = s.termination_scale or 0.0
        c.w_lower, c.w_upper = s.w_lower, s.w_upper
        c.desired_feet_air_time, c.max_contact_force = s.desired_feet_air_time, s.max_contact_force
        # soft limits exactly as the simulator derives them in fp32 (isaacgym.py:373-385)
        lo, hi = torch.tensor(s.hard_dof_lo), torch.tensor(s.hard_dof_hi)
        m, r = (lo + hi) / 2, hi - lo
        slo, shi = m - 0.5 * r * s.soft_dof_pos_limit, m + 0.5 * r * s.soft_dof_pos_limit
        for d in range(s.D):
            c.hard_lo[d], c.hard_hi[d], c.soft_lo[d], c.soft_hi[d] = float(lo[d]), float(hi[d]), float(slo[d]), float(shi[d])
            c.vel_limit[d], c.torque_limit[d] = s.dof_vel_limit[d], s.torque_limit[d]
        c.soft_dof_vel_limit, c.soft_torque_limit = s.soft_dof_vel_limit, s.soft_torque_limit
        c.term_gravity, c.term_motion_far, c.term_motion_end = int(s.term_gravity), int(s.term_motion_far), int(s.term_motion_end)
        c.term_gravity_thr = s.term_gravity_thr
        a = s.adaptive
        c.adaptive_enable, c.adaptive_type = int(a["enable"]), {"origin": 0, "mean": 1, "scale": 2}[a["type"]]
        c.adaptive_alpha, c.adaptive_scale = a["alpha"], a["scale"]

        def cur(d, enable):
            return _lib.Curriculum(int(enable), d["lo"], d["hi"], d["degree"], d["down"], d["up"])
        c.cur_penalty = cur(s.penalty, s.penalty_curriculum)
        c.cur_far = cur(s.far_cur, s.far_cur["enable"])
        c.cur_dof_pos = cur(s.limit_cur["dof_pos"], s.limit_cur["dof_pos"]["enable"])
        c.cur_dof_vel = cur(s.limit_cur["dof_vel"], s.limit_cur["dof_vel"]["enable"])
        c.cur_torque = cur(s.limit_cur["torque"], s.limit_cur["torque"]["enable"])
        c.num_compute_average_epl = s.num_compute_average_epl
        # observation rows: where each source / history key sits in each group (include/pbhc_b200.h)
        used = 0
        c.n_groups = len(s.obs_groups)
        for k in range(len(OBS_SRC)):
            for g in range(_lib.MAX_OBS_GROUPS):
                c.src_dst[k][g] = -1
        n_hd = 0
        for g, (name, grp) in enumerate(s.obs_groups.items()):
            c.group_dim[g] = grp["dim"]
            for seg in grp["segs"]:
                if seg["kind"] == "src":
                    k = OBS_SRC_IDS[seg["key"]]
                    used |= 1 << k
                    c.src_dst[k][g] = seg["off"]  # the source scale is applied by the producer
                else:
                    if s.noise_scales.get(seg.get("group_key", ""), 0.0) != 0:
                        raise NotImplementedError("noise on a whole history block (noise_scales.history_*) is not implemented")
                    if n_hd >= _lib.MAX_HIST_DST:
                        raise NotImplementedError("too many history segments in the observation groups")
                    c.hist_dst[n_hd] = _lib.HistDst(s.hist_keys.index(seg["key"]), g, seg["off"], seg["len"], seg["scale"])
                    n_hd += 1
        c.n_hist_dst = n_hd
        for i, k in enumerate(OBS_SRC):
            c.obs_scale[i] = s.obs_scales.get(k, 1.0)
            # helpers.py:136-152: noise_scales[key] (x1: the noise curriculum is rejected by TrackSpec); it applies to the
            # key wherever it is shown, the pushed history element included (legged_robot_base.py:781-783)
            c.obs_noise[i] = s.noise_scales.get(k, 0.0) if (used >> i) & 1 or k in s.hist_keys else 0.0
        c.clip_obs = s.clip_observations
        c.n_hist_keys, c.hist_row = len(s.hist_keys), self.hist_row
        for i, k in enumerate(s.hist_keys):
            used |= 1 << OBS_SRC_IDS[k]
            c.hist_src[i], c.hist_len[i], c.hist_dim[i] = OBS_SRC_IDS[k], s.hist_len[k], s.hist_dim[k]
        c.src_used_mask = used
        dr = s.dr
        c.randomize_pd_gain, c.randomize_rfi_lim = int(dr["randomize_pd_gain"]), int(dr["randomize_rfi_lim"])
        c.use_rao, c.randomize_ctrl_delay = int(dr["use_rao"]), int(dr["randomize_ctrl_delay"])
        c.delay_lo, c.delay_hi = int(dr["ctrl_delay_step_range"][0]), int(dr["ctrl_delay_step_range"][1])
        c.kp_lo, c.kp_hi, c.kd_lo, c.kd_hi = dr["kp_range"][0], dr["kp_range"][1], dr["kd_range"][0], dr["kd_range"][1]
        c.rfi_lo, c.rfi_hi, c.rao_lim = dr["rfi_lim_range"][0], dr["rfi_lim_range"][1], dr["rao_lim"]
        c.seed = int(seed)
        lvl, ns = s.noise_to_initial_level, s.init_noise_scale
        vals = [ns.get("dof_pos", 0) * lvl, ns.get("dof_vel", 0) * lvl, ns.get("root_pos", 0) * lvl,
                ns.get("root_rot", 0) * 3.14 / 180 * lvl, ns.get("root_vel", 0) * lvl, ns.get("root_ang_vel", 0) * lvl]
        for i, v in enumerate(vals):
            c.reset_noise[i] = v
        c.n_episode_sums = self.n_sums
        # LeggedRobotGeneralTracking extras (general_tracking.py)
        c.general = int(s.general)
        if s.general:
            if len(s.key_ids) > _lib.MAX_KEY or s.future_num_steps > _lib.MAX_FUTURE:
                raise NotImplementedError("too many key bodies / future steps for the B200 tracking step")
            if s.anchor_index >= s.Jb:
                raise NotImplementedError("the anchor must be a simulator body")
            c.anchor_index, c.n_key = s.anchor_index, len(s.key_ids)
            for i, k in enumerate(s.key_ids):
                c.key_ids[i] = k
            c.key_mask, c.body_z_mask = _mask(s.key_ids), _mask(s.body_z_ids)
            c.term_ref_pos_z, c.term_ref_ori, c.term_body_z = int(s.term_ref_pos_z), int(s.term_ref_ori), int(s.term_body_z)
            c.term_ref_pos_z_thr, c.term_ref_ori_thr, c.term_body_z_thr = s.term_ref_pos_z_thr, s.term_ref_ori_thr, s.term_body_z_thr
            c.n_future = s.future_num_steps
            for i, v in enumerate(s.future_steps):
                c.future_steps[i] = v
        else:
            if s.future_num_steps > _lib.MAX_FUTURE:
                raise NotImplementedError("too many future_ref_steps for the B200 tracking step")
            c.n_future = s.future_num_steps  # obs.future_ref_steps (motion_tracking.py:594-599)
        for i, v in enumerate(s.vr3_ids[:4]):
            c.vr3_ids[i] = v
        c.ref_init_yaw = 0.0  # set_ref_init_yaw / __init__
        # domain_rand.push_robots (legged_robot_base.py:134-147, 382-388, 1090-1106)
        c.push_robots, c.push_fixed = int(dr["push_robots"]), int(dr["push_fixed"])
        c.push_lo, c.push_hi, c.max_push_vel = int(dr["push_interval_s"][0]), int(dr["push_interval_s"][1]), dr["max_push_vel_xy"]
        c.start_time_zero = 0  # set_is_evaluating()
        nc = s.noise_cur
        c.cur_noise.enable, c.cur_noise.lo, c.cur_noise.hi = int(nc["enable"]), nc["lo"], nc["hi"]
        c.cur_noise.degree, c.cur_noise.down, c.cur_noise.up = nc["degree"], nc["down"], nc["up"]
        # base-class terminations (legged_robot_base.py:433-476)
        c.term_contact, c.term_contact_mask = int(s.term_contact), _mask(s.term_contact_ids)
        c.term_low_height, c.term_min_height = int(s.term_low_height), s.term_min_height
        c.term_dof_pos, c.term_dof_vel, c.term_torque = int(s.term_dof_pos), int(s.term_dof_vel), int(s.term_torque)
        c.term_dof_pos_prob, c.term_dof_vel_prob, c.term_torque_prob = s.term_dof_pos_prob, s.term_dof_vel_prob, s.term_torque_prob
        c.term_dof_vel_scale, c.term_torque_scale = s.term_dof_vel_scale, s.term_torque_scale
        # simulator.dof_pos_limits_termination exactly as isaacgym.py:383-388 derives it in fp32
        tlo, thi = m - 0.5 * r * s.term_dof_pos_scale, m + 0.5 * r * s.term_dof_pos_scale
        for d in range(s.D):
            c.term_pos_lo[d], c.term_pos_hi[d] = float(tlo[d]), float(thi[d])
        self._cfg = c

    @classmethod
    def specialize(cls, spec, dims, rec_stride, seed=0):
        """Compiles the kernel specialised for `spec` into the on-disk cache (pbhc_track_specialize) without a GPU:
        dims = (J, Jb, D, contact_size) of the motion library.  Returns True when the specialised kernel is cached."""
        self = object.__new__(cls)
        self.spec, self.device = spec, None
        self.n_sums = len(spec.reward_names) + (1 if spec.termination_scale is not None else 0)
        self.hist_row = sum(spec.hist_len[k] * spec.hist_dim[k] for k in spec.hist_keys)
        self._build_cfg(seed)
        J, Jb, D, Cm = dims
        desc = _lib.TableDesc(None, None, None, None, None, 0, 1, J, Jb, D, Cm, rec_stride)
        io = _lib.TrackIO()
        io.default_dof_stride = spec.D
        io.ref_body_pos_extend, io.ref_body_rot_extend = 16, 16  # (only their NULL-ness is read)
        io.fut_table = 16 if self._uses_future_table() else None    # (likewise: selects the staged or the direct-read future window)
        return _lib.lib().pbhc_track_specialize(C.byref(self._cfg), C.byref(desc), C.byref(io)) == 0

    def launch_info(self):
        """what the last pbhc_track_step launched: dict(kernel, blocks, smem_bytes, blocks_per_sm)."""
        out = (C.c_int32 * 4)()
        self._L.pbhc_track_launch_info(C.byref(out))
        return dict(kernel={0: "warp-per-env", 1: "tile (generic)", 2: "tile (specialised for the config)", 3: "warp-per-env (specialised for the config)"}.get(out[0], "?"),
                    blocks=out[1], smem_bytes=out[2], blocks_per_sm=out[3])

    # ---- one env step -----------------------------------------------------------------------------------
    def bind(self, sim, actions, torques, env_origins, motion_ids, inject=None):
        """Records the device pointers of this step's borrowed tensors in the PbhcTrackIO block (no launch).
        sim: object with Isaac-Gym-layout CUDA tensors (simulator/isaacgym/isaacgym.py:596-618):
        _rigid_body_state [N*bodies,13] or [N,bodies,13], all_root_states [N,13], dof_state [N*D,2],
        contact_forces [N,Jb,3], _base_com_bias [N,3], _link_mass_scale [N,L], friction_coeffs [N,1,1].
        actions [N,D] (zeroed in place for envs that reset) and torques [N,D] are this step's clipped actions /
        last applied torques."""
        io, p, s = self._io, _lib.ptr, self.spec
        body = sim._rigid_body_state
        io.body_state, io.body_env_stride = p(body, torch.float32), body.numel() // self.N
        io.root_state, io.dof_state = p(sim.all_root_states, torch.float32), p(sim.dof_state, torch.float32)
        io.contact_forces = p(sim.contact_forces, torch.float32)
        io.actions, io.torques = p(actions, torch.float32), p(torques, torch.float32)
        io.env_origins, io.motion_ids = p(env_origins, torch.float32), p(motion_ids, torch.int64)
        io.clip_of_motion = self.lib._clip_of_motion.data_ptr()
        io.default_dof_pos, io.default_dof_stride = p(self.default_dof_pos), s.D
        io.base_com_bias = p(sim._base_com_bias, torch.float32)
        io.link_mass_scale = p(sim._link_mass_scale, torch.float32) if s.L else None
        io.friction = p(sim.friction_coeffs, torch.float32)
        bm = getattr(sim, "_base_mass_scale", None)
        io.base_mass_scale = p(bm, torch.float32) if bm is not None else None
        io.episode_length_buf, io.env_f32 = p(self.episode_length_buf), p(self.env_f32)
        io.last_actions, io.last_dof_pos, io.last_dof_vel = p(self.last_actions), p(self.last_dof_pos), p(self.last_dof_vel)
        io.kp_scale, io.kd_scale = p(self._kp_scale), p(self._kd_scale)
        io.rfi_lim_scale, io.rao_scale = p(self._rfi_lim_scale), p(self._rao_scale)
        io.action_delay_idx, io.hist, io.episode_sums = p(self.action_delay_idx), p(self.hist), p(self._episode_sums)
        io.scalars = p(self.scalars)
        for g, name in enumerate(s.obs_groups):
            io.obs[g] = p(self.obs_buf_dict[name])
        io.rew_buf, io.reset_buf = p(self.rew_buf), p(self.reset_buf)
        io.time_out_buf = self.time_out_buf.data_ptr()
        io.ref_body_pos_extend, io.ref_body_rot_extend = p(self.ref_body_pos_extend), p(self.ref_body_rot_extend)
        io.last_episode_sums, io.partials, io.acc = p(self.last_episode_sums), p(self.partials), p(self.acc)
        io.last_episode_length_buf = p(self.last_episode_length_buf)
        io.force_reset = p(self._force_reset) if self._force_pending else None
        inj = inject or {}
        io.inject_start_times = p(inj.get("motion_start_times"))
        io.inject_kp, io.inject_kd = p(inj.get("kp_scale")), p(inj.get("kd_scale"))
        io.inject_rfi, io.inject_rao = p(inj.get("rfi_lim_scale")), p(inj.get("rao_scale"))
        io.inject_delay = p(inj.get("action_delay_idx"))
        io.inject_term_gates = int(inj.get("term_gates", -1))
        ft = self._future_table()
        io.fut_table, io.fut_stride = (p(ft), ft.shape[1]) if ft is not None else (None, 0)
        io.heading_scratch = p(self._heading_scratch)
        io.refresh_buf, io.push_vel = p(self._refresh_buf), p(self.push_robot_vel_buf)
        io.inject_push_interval, io.inject_push_vel = p(inj.get("push_interval_s"), torch.int64), p(inj.get("push_vel"), torch.float32)
        self._keep = (sim, actions, torques, env_origins, motion_ids, inj)

    def _uses_future_table(self):
        """a compact future-window table is bound to the launches of this step (one predicate for _future_table() and for the
        warm-up in specialize(): the kernel is specialised on it, csrc/track.cu fut_staged_ok)"""
        return bool(self.spec.general and self.spec.future_num_steps > 0 and os.environ.get("PBHC_FUTURE_TABLE", "1") != "0")

    def _future_table(self):
        """general_tracking: compact per-frame records for the 20-step future window (pbhc_future_table_build), rebuilt when
        the motion library's table changes (load_motions / resample_motion); None when the config has no future window."""
        if not self._uses_future_table():
            return None  # (PBHC_FUTURE_TABLE=0: the window reads the packed table itself - same results, for A/B runs)
        ver = (getattr(self.lib, "_table_version", 0), self.lib._records.data_ptr())
        if getattr(self, "_fut_ver", None) != ver:
            stride = int(self._L.pbhc_future_table_stride(C.byref(self._cfg)))
            frames = int(self.lib._desc.total_frames)
            self._fut = torch.empty(frames, stride, device=self.device, dtype=torch.float32)
            _lib.check(self._L.pbhc_future_table_build(C.byref(self._cfg), C.byref(self.lib._desc), _lib.ptr(self._fut), stride,
                                                       _lib.stream_ptr(self.device)))
            self._fut_ver = ver
        return self._fut

    def write_obs_into(self, buffers):
        """Redirects observation groups to caller-owned [N, dim] fp32 CUDA tensors from the next bind() on, e.g.
        `{g: storage.slot(g, step + 1)}` so RolloutStorage.update_key (agents/ppo/ppo.py:251-252) finds the rows
        already in place.  Groups not named keep their buffer.  Takes effect for the next launch() (the pointer block
        of the last bind() is patched too); a HostStep / CUDA graph built earlier keeps the pointers it captured."""
        names = list(self.spec.obs_groups)
        for g, t in buffers.items():
            cur = self.obs_buf_dict[g]
            if t.shape != cur.shape or t.dtype != cur.dtype or t.device != cur.device or not t.is_contiguous():
                raise _lib.PbhcError(f"write_obs_into: '{g}' must be a contiguous {tuple(cur.shape)} {cur.dtype} tensor on {cur.device}")
            self.obs_buf_dict[g] = t
            self._io.obs[names.index(g)] = _lib.ptr(t)

    def launch(self):
        """pbhc_track_step on the current stream with the pointers of the last bind(): the fused kernel + the
        deterministic reduction of its per-block sums into `acc`."""
        _lib.check(self._L.pbhc_track_step(C.byref(self._cfg), C.byref(self.lib._desc), C.byref(self._io), self.N,
                                           _lib.stream_ptr(self.device)))

    def finalize(self, group=None):
        """all-reduce of `acc` when envs are sharded over ranks (the only cross-rank exchange of the step),
        then pbhc_track_finalize advances the device scalars (adaptive sigma, curricula, log values)."""
        from .dist import all_reduce_sum_
        all_reduce_sum_(self.acc, group)  # (the world-size-2 gloo test drives this very call, tests/test_dist_cpu.py)
        _lib.check(self._L.pbhc_track_finalize(C.byref(self._cfg), _lib.ptr(self.acc), _lib.ptr(self.scalars),
                                               _lib.stream_ptr(self.device)))

    def step(self, sim, actions, torques, env_origins, motion_ids, inject=None, group=None):
        """Returns (obs_buf_dict, rew_buf, reset_buf, extras) like LeggedRobotBase.step (see bind() for the inputs)."""
        if inject is None and self.inject_next is not None:
            inject, self.inject_next = self.inject_next, None
        self.bind(sim, actions, torques, env_origins, motion_ids, inject)
        self.launch()
        if self._force_pending:  # the forced resets (resample_motion) are consumed by this launch
            self._force_reset.zero_()
            self._force_pending = False
        self.finalize(group)
        self.extras = {"time_outs": self.time_out_buf, "ref_body_pos_extend": self.ref_body_pos_extend,
                       "ref_body_rot_extend": self.ref_body_rot_extend, "to_log": None}
        return self.obs_buf_dict, self.rew_buf, self.reset_buf, self.extras

    # ---- the rest of _post_physics_step's contract: env ids of the resets, extras["episode"], resample_motion -----------
    @property
    def need_to_refresh_envs(self):
        """legged_robot_base.py:315: after step() the envs of reset_buf wait for the simulator setters - and, with
        domain_rand.push_robots, the envs whose root velocity the step just pushed (:1095)."""
        if self._refresh_buf is None:
            return self.reset_buf != 0
        # the kernel stamps the envs it resets or pushes with the step count finalize() then reaches (no per-env store otherwise)
        return self._refresh_buf == self.scalars[_lib.S_PUSH_COUNT].long()

    def refresh_env_ids(self):
        """`self.need_to_refresh_envs.nonzero(as_tuple=False).flatten()` (legged_robot_base.py:315) on the device: (ids, count) like
        reset_env_ids().  Without push_robots these ARE the reset ids."""
        if self._refresh_buf is None:
            return self.reset_env_ids()
        self._refresh_flags.copy_(self.need_to_refresh_envs)
        _lib.check(self._L.pbhc_compact_ids(_lib.ptr(self._refresh_flags), self.N, _lib.ptr(self._refresh_ids), _lib.ptr(self._refresh_count),
                                            _lib.ptr(self._compact_scratch), _lib.stream_ptr(self.device)))
        return self._refresh_ids, self._refresh_count

    def refresh_env_ids_host(self):
        ids, count = self.refresh_env_ids()
        return ids[:int(count.item())]

    def reset_env_ids(self):
        """`self.reset_buf.nonzero(as_tuple=False).flatten()` (legged_robot_base.py:311) without leaving the device:
        (ids int32 [N], count int32 [1]) - ids[:count] are the envs reset by the last step, ascending (pbhc_compact_ids)."""
        _lib.check(self._L.pbhc_compact_ids(_lib.ptr(self.reset_buf), self.N, _lib.ptr(self._reset_ids), _lib.ptr(self._reset_count),
                                            _lib.ptr(self._compact_scratch), _lib.stream_ptr(self.device)))
        return self._reset_ids, self._reset_count

    def reset_env_ids_host(self):
        """The same as an int32 device tensor of exactly `count` ids, for setters that need len(env_ids) on the host
        (simulator/isaacgym/isaacgym.py:642-660).  One 4-byte read-back - the only synchronisation of the step."""
        ids, count = self.reset_env_ids()
        return ids[:int(count.item())]

    def episode_extras(self, env_ids):
        """extras["episode"] of reset_envs_idx (legged_robot_base.py:508-515) for the envs reset by the last step:
        rew_<name> = episode sums / max_episode_length_s, end_epis_length = last_episode_length_buf[env_ids]."""
        idx = env_ids.long()
        sums = self.last_episode_sums[idx] / self.spec.max_episode_length_s
        ep = {"rew_" + n: sums[:, i] for i, n in enumerate(self._sum_names)}
        ep["end_epis_length"] = self.last_episode_length_buf[idx]
        return ep

    def resample_motion(self, **load_kwargs):
        """LeggedRobotMotionTracking.resample_motion (motion_tracking.py:385-389): a new draw of clips for the envs, then
        every env is reset.  The reference resets inside _update_tasks_callback of the running step; here the reset of all
        envs is part of the next step() (PbhcTrackIO.force_reset), which samples start times from the new clips."""
        load_kwargs.setdefault("random_sample", True)
        self.lib.load_motions(**load_kwargs)
        self._force_reset.fill_(1)
        self._force_pending = True

    # ---- chunked launches and the host-buffer pipeline --------------------------------------------------------
    def _row_bytes(self):
        """bytes per env of every per-env pointer of PbhcTrackIO (what a launch on envs [lo, hi) offsets)."""
        s, io = self.spec, self._io
        D, J, Jb, L = s.D, s.J, s.Jb, s.L
        rew = (len(s.reward_names) + 1) * 4 if s.use_vec_reward else 4
        rb = dict(body_state=io.body_env_stride * 4, root_state=52, dof_state=D * 8, contact_forces=Jb * 12, actions=D * 4, torques=D * 4,
                  env_origins=12, motion_ids=8, default_dof_pos=io.default_dof_stride * 4, base_com_bias=12, link_mass_scale=L * 4,
                  friction=4, base_mass_scale=4, episode_length_buf=8, env_f32=_lib.ENVF_STRIDE * 4, last_actions=D * 4, last_dof_pos=D * 4,
                  last_dof_vel=D * 4, kp_scale=D * 4, kd_scale=D * 4, rfi_lim_scale=D * 4, rao_scale=D * 4, action_delay_idx=8,
                  hist=self.hist_row * 4, episode_sums=self.n_sums * 4, rew_buf=rew, reset_buf=8, time_out_buf=1,
                  ref_body_pos_extend=J * 12, ref_body_rot_extend=J * 16, last_episode_sums=self.n_sums * 4,
                  last_episode_length_buf=8, force_reset=1,
                  inject_start_times=4, inject_kp=D * 4, inject_kd=D * 4, inject_rfi=D * 4, inject_rao=D * 4, inject_delay=8,
                  refresh_buf=8, push_vel=8, inject_push_interval=8, inject_push_vel=8)
        return rb

    def launch_range(self, lo, hi, acc):
        """pbhc_track_step on envs [lo, hi) of the last bind(), per-step sums into `acc` (a [PBHC_ACC] fp64 tensor).
        Used to pipeline a step in chunks; sum the chunk `acc`s into self.acc before finalize()."""
        if self._heading_scratch is not None:
            raise NotImplementedError("vr_3point_pos pairs env i with the heading of env (3 i + k) mod N (motion_tracking.py:730-731): not defined for a launch on a sub-range")
        io = _lib.TrackIO.from_buffer_copy(self._io)
        for name, nbytes in self._row_bytes().items():
            v = getattr(io, name)
            if v:
                setattr(io, name, v + lo * nbytes)
        for g, name in enumerate(self.spec.obs_groups):
            io.obs[g] = self._io.obs[g] + lo * self.obs_buf_dict[name].shape[1] * 4
        io.acc, io.env_offset = _lib.ptr(acc, torch.float64), self._io.env_offset + lo
        _lib.check(self._L.pbhc_track_step(C.byref(self._cfg), C.byref(self.lib._desc), C.byref(io), hi - lo,
                                           _lib.stream_ptr(self.device)))


class HostStep:
    """The step for callers whose simulator state lives in HOST memory (pinned): `step()` copies the inputs of this step
    to the device, runs the fused step and copies observations / rewards / reset flags back, pipelined over `chunks`
    ranges of envs on three streams so that the host->device copy of chunk k+1, the kernels of chunk k and the
    device->host copy of chunk k-1 overlap (PCIe is full duplex).  Returns when the results are on the host.
    Measured on the B200 box (65536 envs, 150 MB up / 271 MB down per step, 50 GB/s each way when both directions are
    busy): every copy costs ~8 us of fixed overhead, so few large chunks win - 4 chunks 6.2 ms, 8: 6.6, 16: 7.3."""

    def __init__(self, tracking_step, sim, actions, torques, env_origins, motion_ids, host_in, extra=None, chunks=4, graph=True):
        """host_in: dict name -> pinned CPU tensor for body_state / root_state / dof_state / contact_forces / actions /
        torques, shaped like the device tensors; extra: optional (fn(lo, hi), [(host, device), ...]) run per chunk before
        the tracking kernel (the benchmark uses it for K2 on the same envs); graph: after one eager step the whole
        three-stream pipeline (copies, kernels, exchange, finalize) is captured into a CUDA graph and replayed - the
        ~15 x chunks host-side launches per step otherwise sit on the critical path of a PCIe-bound step."""
        ts = self.ts = tracking_step
        self._graph, self._want_graph, self._steps = None, bool(graph), 0
        self.N, self.dev = ts.N, ts.device
        ts.bind(sim, actions, torques, env_origins, motion_ids)
        self.pairs = [(host_in["body_state"], sim._rigid_body_state.view(self.N, -1)), (host_in["root_state"], sim.all_root_states),
                      (host_in["dof_state"], sim.dof_state.view(self.N, -1)), (host_in["contact_forces"], sim.contact_forces.view(self.N, -1)),
                      (host_in["actions"], actions), (host_in["torques"], torques)]
        self.pairs = [(h.view(self.N, -1), d) for h, d in self.pairs]
        self.extra_fn = None
        if extra is not None:
            self.extra_fn = extra[0]
            self.pairs += [(h.view(self.N, -1), d.view(self.N, -1)) for h, d in extra[1]]
        outs = list(ts.obs_buf_dict.values()) + [ts.rew_buf.view(self.N, -1), ts.reset_buf.view(self.N, 1), ts.time_out_buf.view(self.N, 1)]
        self.outs = [(o, torch.empty(o.shape, dtype=o.dtype).pin_memory()) for o in outs]
        for h, _ in self.pairs:
            if not h.is_pinned():
                raise _lib.PbhcError("HostStep needs pinned host tensors")
        c = max(1, min(int(chunks), self.N // 4 or 1))
        step = -(-self.N // c)
        step = (step + 3) // 4 * 4  # chunk starts on multiples of 4 envs keep every row 16-byte aligned
        self.ranges = [(lo, min(lo + step, self.N)) for lo in range(0, self.N, step)]
        self.s_in, self.s_k, self.s_out = (torch.cuda.Stream(self.dev) for _ in range(3))
        self.acc = torch.zeros(len(self.ranges), _lib.ACC, device=self.dev, dtype=torch.float64)
        self.h2d_bytes = sum(h.numel() * h.element_size() for h, _ in self.pairs)
        self.d2h_bytes = sum(o.numel() * o.element_size() for o, _ in self.outs)

    @property
    def host_out(self):
        names = list(self.ts.obs_buf_dict.keys()) + ["rew_buf", "reset_buf", "time_out_buf"]
        return {n: h for n, (_, h) in zip(names, self.outs)}

    def step(self):
        if self._graph is not None:
            self._graph.replay()
            torch.cuda.current_stream(self.dev).synchronize()
            return self.host_out
        if self._want_graph and self._steps >= 2:  # the first steps ran eagerly (lazy one-time set-up done, incl. the peer-memory exchange)
            try:  # sharded steps too: the exchange of finalize() is a capturable kernel (csrc/peer.cu) or a capturable NCCL call
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    self._enqueue()
                self._graph = g
                return self.step()
            except Exception:  # noqa: BLE001 - a collective that cannot be captured here: stay eager
                self._want_graph = False
                torch.cuda.synchronize(self.dev)
        self._steps += 1
        self._enqueue()
        torch.cuda.current_stream(self.dev).synchronize()  # the caller owns the results on the host when step() returns
        return self.host_out

    def _enqueue(self):
        ts = self.ts
        cur = torch.cuda.current_stream(self.dev)
        for s in (self.s_in, self.s_k, self.s_out):
            s.wait_stream(cur)
        for k, (lo, hi) in enumerate(self.ranges):
            with torch.cuda.stream(self.s_in):
                for h, d in self.pairs:
                    d[lo:hi].copy_(h[lo:hi], non_blocking=True)
                ev_in = self.s_in.record_event()
            with torch.cuda.stream(self.s_k):
                self.s_k.wait_event(ev_in)
                if self.extra_fn is not None:
                    self.extra_fn(lo, hi)
                ts.launch_range(lo, hi, self.acc[k])
                ev_k = self.s_k.record_event()
            with torch.cuda.stream(self.s_out):
                self.s_out.wait_event(ev_k)
                for o, h in self.outs:
                    h[lo:hi].copy_(o[lo:hi], non_blocking=True)
        with torch.cuda.stream(self.s_k):
            torch.sum(self.acc, dim=0, out=ts.acc)
            ts.finalize()
        cur.wait_stream(self.s_in)
        cur.wait_stream(self.s_k)
        cur.wait_stream(self.s_out)

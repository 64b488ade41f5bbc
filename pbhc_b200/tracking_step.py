"""TrackingStep: the fused motion-tracking env step behind the reference's env attributes.

It replaces, for LeggedRobotMotionTracking, everything `_post_physics_step` does between
`_refresh_sim_tensors()` and `extras["to_log"]` (legged_robot_base.py:297-338): one launch of pbhc_track_step
(csrc/track.cu) + pbhc_track_finalize, no host synchronisation.  State the reference keeps as env attributes
is exposed under the same names as views of the packed device buffers (episode_length_buf,
motion_start_times, motion_len, feet_air_time, last_contacts, episode_sums[name], rew_buf, reset_buf,
time_out_buf, obs_buf_dict, log_dict, history_handler.query(key), ...).  No CPU fallback.
"""
import ctypes as C

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
    def __init__(self, spec, motion_lib, num_envs, device, seed=0):
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
        init[_lib.S_PENALTY] = s.penalty["initial"]
        init[_lib.S_FAR_THR] = s.motion_far_threshold0
        init[_lib.S_SOFT_DOF_POS] = s.limit_cur["dof_pos"]["initial"]
        init[_lib.S_SOFT_DOF_VEL] = s.limit_cur["dof_vel"]["initial"]
        init[_lib.S_SOFT_TORQUE] = s.limit_cur["torque"]["initial"]
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
        self._build_cfg(seed)
        self._io = _lib.TrackIO()
        self._keep = None

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
        sig = self.scalars[_lib.S_SIGMA:_lib.S_SIGMA + _lib.NUM_SIGMA].float()
        ema = self.scalars[_lib.S_EMA:_lib.S_EMA + _lib.NUM_SIGMA].float()
        for name, i in SIGMA_IDS.items():
            d["adp_sigma_" + name], d["error_ema_" + name] = sig[i], ema[i]
        return d

    def reward_tracking_sigma(self):
        """config.rewards.reward_tracking_sigma as the reference mutates it (synchronises)."""
        sig = self.scalars[_lib.S_SIGMA:_lib.S_SIGMA + _lib.NUM_SIGMA].cpu().tolist()
        return {name: sig[i] for name, i in SIGMA_IDS.items()}

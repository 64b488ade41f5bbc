"""Actuation half of the env step on the sm_100a kernels of csrc/actuate.cu:

  * `_pre_physics_step(actions)`  = LeggedRobotBase._pre_physics_step (envs/legged_base_task/legged_robot_base.py:268-286):
    action clip, `log_dict["action_clip_frac"]`, control-delay queue shift + per-env delayed pick;
  * `_compute_torques(actions)`   = LeggedRobotBase._compute_torques (:795-838): PD law ("P" / "V" / "T") with the
    per-env kp / kd scales, torque RFI noise, RAO offset and torque clipping; the reference runs it
    control_decimation (4) times per env step, between physics substeps (`_apply_force_in_physics_step`, :292-294).

Same attribute names as the reference env (actions, actions_after_delay, action_queue, action_delay_idx, torques,
p_gains, d_gains, torque_limits, _kp_scale, _kd_scale, _rfi_lim_scale, _rao_scale, log_dict).  Built next to a
TrackingStep it shares that object's per-env randomisation tensors (K3 resamples them at resets) and its
default_dof_pos.  fp32 operations in the reference's order: bit-exact except for the RFI draw, which comes from an
in-kernel Philox stream (or from `rfi_uniform`, the torch.rand_like values, in tests).  No CPU fallback.
"""
import ctypes as C

import torch

from . import _lib
from .spec import _get

CONTROL_TYPES = {"P": 0, "V": 1, "T": 2}


class Actuation:
    def __init__(self, cfg, num_envs, device, tracking_step=None, seed=0, env_offset=0):
        dev = torch.device(device)
        if dev.type != "cuda":
            raise _lib.PbhcError("Actuation needs a CUDA device (there is no CPU fallback)")
        self.device, self.num_envs = dev, num_envs
        robot, dr, ctl = _get(cfg, "robot"), _get(cfg, "domain_rand"), _get(cfg, "robot.control")
        self.dof_names = list(_get(robot, "dof_names"))
        D = self.num_dof = len(self.dof_names)
        if D > _lib.MAX_DOF:
            raise _lib.PbhcError(f"{D} dofs > PBHC_MAX_DOF")
        control_type = _get(ctl, "control_type")
        if control_type not in CONTROL_TYPES:
            raise NameError(f"Unknown controller type: {control_type}")  # legged_robot_base.py:817
        if _get(dr, "parallel_serial_tau.enable", False):
            raise NotImplementedError("domain_rand.parallel_serial_tau (no shipped config enables it)")
        # legged_robot_base.py:89-108: gains by substring match of the stiffness keys
        stiff, damp, scale = _get(ctl, "stiffness"), _get(ctl, "damping"), _get(ctl, "action_scale")
        per_dof = not isinstance(scale, (int, float))
        p_gains, d_gains, scales = [0.0] * D, [0.0] * D, [0.0] * D
        for i, name in enumerate(self.dof_names):
            found = False
            for key in stiff.keys():
                if key in name:
                    p_gains[i], d_gains[i] = float(stiff[key]), float(damp[key])
                    if per_dof:
                        scales[i] = float(scale[key])
                    found = True
            if not found and control_type in ("P", "V"):
                raise ValueError(f"PD gain of joint {name} were not defined. Should be defined in the yaml file.")
        f32 = dict(device=dev, dtype=torch.float32)
        self.p_gains, self.d_gains = torch.tensor(p_gains, **f32), torch.tensor(d_gains, **f32)
        self.torque_limits = torch.tensor([float(v) for v in _get(robot, "dof_effort_limit_list")], **f32)
        self.action_scales = torch.tensor(scales, **f32) if per_dof else None
        delay = bool(_get(dr, "randomize_ctrl_delay", False))
        rng = list(_get(dr, "ctrl_delay_step_range", [0, 0]))
        Q = int(rng[1]) + 1 if delay else 0
        c = self._cfg = _lib.ActuationCfg()
        c.D, c.Q, c.action_clip, c.control_type = D, Q, float(_get(ctl, "action_clip_value")), CONTROL_TYPES[control_type]
        c.per_dof_action_scale, c.action_scale = int(per_dof), 0.0 if per_dof else float(scale)
        for i in range(D):
            c.action_scales[i], c.p_gains[i], c.d_gains[i] = scales[i], p_gains[i], d_gains[i]
            c.torque_limits[i] = float(self.torque_limits[i])
        c.rfi, c.rfi_lim = int(bool(_get(dr, "randomize_torque_rfi", False))), float(_get(dr, "rfi_lim", 0.0))
        c.rao, c.clip_torques = int(bool(_get(dr, "use_rao", False))), int(bool(_get(ctl, "clip_torques", False)))
        c.sim_dt = 1.0 / float(_get(cfg, "simulator.config.sim.fps"))
        c.seed = int(seed) & 0xFFFFFFFFFFFFFFFF
        N = num_envs
        ts = self._ts = tracking_step
        if ts is not None:
            self._kp_scale, self._kd_scale, self._rfi_lim_scale, self._rao_scale = ts._kp_scale, ts._kd_scale, ts._rfi_lim_scale, ts._rao_scale
            self.action_delay_idx, self.default_dof_pos, self.last_dof_vel = ts.action_delay_idx, ts.default_dof_pos, ts.last_dof_vel
        else:
            self._kp_scale, self._kd_scale, self._rfi_lim_scale, self._rao_scale = (torch.ones(N, D, **f32) for _ in range(4))
            # legged_robot_base.py:154-155
            self.action_delay_idx = torch.randint(int(rng[0]), int(rng[1]) + 1, (N,), device=dev) if delay else torch.zeros(N, device=dev, dtype=torch.int64)
            self.default_dof_pos = torch.tensor([float(_get(robot, "init_state.default_joint_angles")[n]) for n in self.dof_names], **f32).repeat(N, 1)
            self.last_dof_vel = torch.zeros(N, D, **f32)
        self.actions, self.actions_after_delay, self.torques = (torch.zeros(N, D, **f32) for _ in range(3))
        self.action_queue = torch.zeros(N, Q, D, **f32) if Q else None
        self._clip_count = torch.zeros(1, device=dev, dtype=torch.int64)
        self._numel = torch.tensor(float(N * D), **f32)
        # the RFI noise stream is keyed by a DEVICE call counter (advanced by pbhc_compute_torques) and the global env id, so a
        # replayed CUDA graph draws fresh noise and ranks that share a seed draw different streams
        self._call_counter = torch.zeros(1, device=dev, dtype=torch.int32)
        c.call_counter = self._call_counter.data_ptr()
        c.env_offset = int(ts._io.env_offset) if ts is not None else int(env_offset)
        self.log_dict = {}
        self._L = _lib.lib()

    # ---- legged_robot_base.py:268-286 ---------------------------------------------------------------------------------
    def _pre_physics_step(self, actions):
        a = actions.to(device=self.device, dtype=torch.float32).contiguous()
        if a.shape != self.actions.shape:
            raise _lib.PbhcError(f"actions {tuple(a.shape)}, expected {tuple(self.actions.shape)}")
        self._clip_count.zero_()
        p = lambda t: None if t is None else t.data_ptr()
        _lib.check(self._L.pbhc_pre_physics(C.byref(self._cfg), p(a), p(self.actions), p(self.action_queue), p(self.action_delay_idx),
                                            p(self.actions_after_delay), p(self._clip_count), self.num_envs, _lib.stream_ptr(self.device)))
        # count / numel as an fp32 tensor-by-tensor division (eager torch divides by a python scalar through a reciprocal
        # on CUDA, which is 1 ulp off the CPU result the reference logs)
        self.log_dict["action_clip_frac"] = self._clip_count[0].to(torch.float32) / self._numel
        return self.actions_after_delay

    pre_physics_step = _pre_physics_step

    # ---- legged_robot_base.py:795-838 ---------------------------------------------------------------------------------
    def _compute_torques(self, actions, dof_state, rfi_uniform=None, out=None):
        """actions: the delayed actions [N,D]; dof_state: Isaac Gym dof_state [N*D,2] or [N,D,2] (pos, vel);
        rfi_uniform [N,D]: U[0,1) values to use instead of the in-kernel stream.  Returns torques [N,D] (self.torques)."""
        a = actions.to(torch.float32).contiguous()
        ds = dof_state.contiguous()
        if ds.dtype != torch.float32 or ds.numel() != self.num_envs * self.num_dof * 2 or not ds.is_cuda:
            raise _lib.PbhcError("dof_state must be a CUDA fp32 tensor of N*D*2 elements")
        out = self.torques if out is None else out
        u = None if rfi_uniform is None else rfi_uniform.to(torch.float32).contiguous()
        p = lambda t: None if t is None else t.data_ptr()
        stride = self.default_dof_pos.shape[1] if self.default_dof_pos.dim() == 2 and self.default_dof_pos.shape[0] > 1 else 0
        _lib.check(self._L.pbhc_compute_torques(C.byref(self._cfg), p(a), p(ds), p(self.default_dof_pos), stride, p(self.last_dof_vel),
                                                p(self._kp_scale), p(self._kd_scale), p(self._rfi_lim_scale), p(self._rao_scale), p(u),
                                                0, p(out), self.num_envs, _lib.stream_ptr(self.device)))
        return out

    compute_torques = _compute_torques

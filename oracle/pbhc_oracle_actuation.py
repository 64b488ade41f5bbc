"""TEST INFRASTRUCTURE ONLY - torch-fp32 CPU restatement of the reference's actuation path.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this; pbhc_b200/ never does.

Follows humanoidverse/envs/legged_base_task/legged_robot_base.py:268-286 (_pre_physics_step) and :795-838
(_compute_torques).  Pinned against the unmodified reference env by tests/golden/ref_actuation.npz
(tools/make_golden_actuation.py) in tests/test_oracle_golden.py.
"""
import torch


class Actuation:
    def __init__(self, N, D, Q, action_clip, control_type, action_scale, p_gains, d_gains, torque_limits, default_dof_pos,
                 rfi, rfi_lim, rao, clip_torques, sim_dt=0.005):
        self.N, self.D, self.Q = N, D, Q
        self.action_clip, self.control_type, self.action_scale = action_clip, control_type, action_scale
        self.p_gains, self.d_gains, self.torque_limits = (torch.as_tensor(x, dtype=torch.float32) for x in (p_gains, d_gains, torque_limits))
        self.default_dof_pos = torch.as_tensor(default_dof_pos, dtype=torch.float32)
        self.rfi, self.rfi_lim, self.rao, self.clip_torques, self.sim_dt = rfi, rfi_lim, rao, clip_torques, sim_dt
        self.kp_scale, self.kd_scale, self.rfi_lim_scale, self.rao_scale = (torch.ones(N, D) for _ in range(4))
        self.action_delay_idx = torch.zeros(N, dtype=torch.int64)
        self.action_queue = torch.zeros(N, max(Q, 1), D)
        self.last_dof_vel = torch.zeros(N, D)
        self.log_dict = {}

    def pre_physics_step(self, actions):  # legged_robot_base.py:268-286
        self.actions = torch.clip(actions, -self.action_clip, self.action_clip)
        self.log_dict["action_clip_frac"] = (self.actions.abs() == self.action_clip).sum() / self.actions.numel()
        if self.Q > 0:
            self.action_queue[:, 1:] = self.action_queue[:, :-1].clone()
            self.action_queue[:, 0] = self.actions.clone()
            self.actions_after_delay = self.action_queue[torch.arange(self.N), self.action_delay_idx].clone()
        else:
            self.actions_after_delay = self.actions.clone()
        return self.actions_after_delay

    def compute_torques(self, actions, dof_pos, dof_vel, rfi_uniform=None):  # legged_robot_base.py:795-838
        a = actions * self.action_scale
        if self.control_type == "P":
            t = self.kp_scale * self.p_gains * (a + self.default_dof_pos - dof_pos) - self.kd_scale * self.d_gains * dof_vel
        elif self.control_type == "V":
            t = self.kp_scale * self.p_gains * (a - dof_vel) - self.kd_scale * self.d_gains * (dof_vel - self.last_dof_vel) / self.sim_dt
        elif self.control_type == "T":
            t = a
        else:
            raise NameError(f"Unknown controller type: {self.control_type}")
        if self.rfi:
            u = torch.rand_like(t) if rfi_uniform is None else rfi_uniform
            t = t + (u * 2. - 1.) * self.rfi_lim * self.rfi_lim_scale * self.torque_limits
        if self.rao:
            t = t + self.rao_scale * self.torque_limits
        if self.clip_torques:
            return torch.clip(t, -self.torque_limits, self.torque_limits)
        return t

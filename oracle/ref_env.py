"""TEST INFRASTRUCTURE ONLY - runs the UNMODIFIED reference env classes (LeggedRobotMotionTracking /
LeggedRobotGeneralTracking) headless on a fake simulator made of synthetic tensors with Isaac Gym's
layouts (simulator/isaacgym/isaacgym.py:596-618).  Used by tools/make_golden_env_step.py to dump golden
fixtures that pin oracle/pbhc_oracle_env.py, and by tools/calibrate_cpu_port.py to time the reference's own
step beside the port that bench.py times (bench.py itself never imports this module: /root/reference does not
exist on the GPU box).  Needs /root/reference: never imported on the GPU box by tests, smoke() or the product.
"""
import os
import re
import types

import numpy as np
import torch
import yaml

from . import ref_loader as R

CFG_ROOT = os.path.join(R.REF_ROOT, "humanoidverse", "config")


class _Loader(yaml.SafeLoader):
    pass


# YAML 1.2 floats ("1e-3" is a float for OmegaConf, a string for PyYAML)
_Loader.add_implicit_resolver(
    "tag:yaml.org,2002:float",
    re.compile(r"^[-+]?(?:[0-9][0-9_]*\.[0-9_]*(?:[eE][-+]?[0-9]+)?|\.[0-9_]+(?:[eE][-+]?[0-9]+)?|[0-9][0-9_]*[eE][-+]?[0-9]+|\.(?:inf|Inf|INF)|\.(?:nan|NaN|NAN))$"),
    list("-+0123456789."))


def yaml_load(rel):
    with open(os.path.join(CFG_ROOT, rel)) as f:
        return yaml.load(f, Loader=_Loader)


def _merge(a, b):
    for k, v in b.items():
        if isinstance(v, dict) and isinstance(a.get(k), dict):
            _merge(a[k], v)
        else:
            a[k] = v
    return a


def _group(rel, key):
    """one config group file with its `defaults:` parents merged underneath."""
    d = yaml_load(rel)
    out = {}
    for parent in d.get("defaults", []) or []:
        if isinstance(parent, str) and not parent.startswith("/"):
            d_ = os.path.dirname(rel)
            while d_ and not os.path.exists(os.path.join(CFG_ROOT, d_, parent + ".yaml")):
                d_ = os.path.dirname(d_)
            _merge(out, _group(os.path.join(d_, parent + ".yaml"), key))
    node = d
    for k in key.split("."):
        node = node.get(k, {}) if isinstance(node, dict) else {}
    return _merge(out, node or {})


def build_config(num_envs, motion_file, robot="g1/g1_23dof_lock_wrist", rewards="motion_tracking/main",
                 obs="motion_tracking/main", env="motion_tracking", domain_rand="main", overrides=None):
    """What hydra composes for `+exp=motion_tracking +robot=.. +rewards=.. +obs=.. +domain_rand=..` (env.config)."""
    robot_c = _group(f"robot/{robot}.yaml", "robot")
    rewards_c = _group(f"rewards/{rewards}.yaml", "rewards")
    obs_c = _group(f"obs/{obs}.yaml", "obs")
    dr_c = _group(f"domain_rand/{domain_rand}.yaml", "domain_rand")
    terrain_c = _group("terrain/terrain_locomotion_plane.yaml", "terrain")
    sim_c = _group("simulator/isaacgym.yaml", "simulator")
    cfg = _group(f"env/{env}.yaml", "env.config")
    cfg.update(robot=robot_c, rewards=rewards_c, obs=obs_c, domain_rand=dr_c, terrain=terrain_c, simulator=sim_c,
               num_envs=num_envs, headless=True, experiment_name="oracle")
    sim_c["_target_"] = "oracle.ref_env.FakeSim"
    robot_c["motion"]["motion_file"] = motion_file
    robot_c["motion"]["asset"]["assetRoot"] = os.path.join(R.REF_ROOT, robot_c["motion"]["asset"]["assetRoot"])
    # ${robot.dof_obs_size} / ${eval:'3 * ${robot.num_bodies} + 9'} in obs_dims; list of 1-key dicts -> dict
    # (utils/helpers.py:47-56 determine_obs_dim)
    dims = {}
    for d in obs_c["obs_dims"]:
        for k, v in d.items():
            if isinstance(v, str):
                v = re.sub(r"\$\{robot\.([a-z_]+)\}", lambda m: str(robot_c[m.group(1)]), v)
                m = re.match(r"^\$\{eval:'(.*)'\}$", v)
                v = int(eval(m.group(1))) if m else int(v)
            dims[k] = v
    obs_c["obs_dims"] = dims

    def _interp(x):  # ${obs.history_length} etc. inside the obs group (OmegaConf interpolation)
        if isinstance(x, dict):
            return {k: _interp(v) for k, v in x.items()}
        if isinstance(x, list):
            return [_interp(v) for v in x]
        if isinstance(x, str):
            m = re.match(r"^\$\{obs\.([A-Za-z_]+)\}$", x)
            if m:
                return obs_c[m.group(1)]
        return x
    for k in list(obs_c.keys()):
        obs_c[k] = _interp(obs_c[k])
    for k, v in (overrides or {}).items():
        node = cfg
        ks = k.split(".")
        for kk in ks[:-1]:
            node = node[kk]
        node[ks[-1]] = v
    return R.to_attr(_drop_missing(cfg))


def _drop_missing(x):
    """OmegaConf semantics: a '???' value is MISSING and `key in cfg` is False for it."""
    if isinstance(x, dict):
        return {k: _drop_missing(v) for k, v in x.items() if not (isinstance(v, str) and v == "???")}
    if isinstance(x, list):
        return [_drop_missing(v) for v in x]
    return x


class FakeSim:
    """Stand-in for simulator.isaacgym.IsaacGym: same attribute names and tensor layouts
    (isaacgym.py:234-239, 366-388, 596-618), no physics.  Tensors are filled by set_state()."""

    def __init__(self, config, device):
        self.config = self.env_config = config
        self.robot_config = config.robot
        self.sim_device = self.device = device
        self.sim_dt = 1.0 / config.simulator.config.sim.fps
        self.viewer = None
        self.headless = True

    def set_headless(self, h):
        self.headless = h

    def setup(self):
        pass

    def setup_terrain(self, mesh_type):
        pass

    def load_assets(self):
        rc = self.robot_config
        self.dof_names, self.body_names = list(rc.dof_names), list(rc.body_names)
        self.num_dof, self.num_bodies = len(self.dof_names), len(self.body_names)

    def create_envs(self, num_envs, env_origins, base_init_state):
        self.num_envs = num_envs
        self._body_list = list(self.body_names)
        dr = self.env_config.domain_rand
        dev = self.device
        self._base_com_bias = torch.zeros(num_envs, 3, device=dev)
        self.friction_coeffs = torch.ones(num_envs, 1, 1, device=dev)
        self._link_mass_scale = torch.ones(num_envs, len(dr.randomize_link_body_names), device=dev)
        self._base_mass_scale = torch.ones(num_envs, 1, device=dev)

    def get_dof_limits_properties(self):
        rc, dev, D = self.robot_config, self.device, self.num_dof
        self.hard_dof_pos_limits = torch.zeros(D, 2, device=dev)
        self.dof_pos_limits = torch.zeros(D, 2, device=dev)
        self.dof_vel_limits = torch.zeros(D, device=dev)
        self.torque_limits = torch.zeros(D, device=dev)
        self.dof_pos_limits_termination = torch.zeros(D, 2, device=dev)
        for i in range(D):  # isaacgym.py:373-388, same op order
            lo, hi = np.float32(rc.dof_pos_lower_limit_list[i]).item(), np.float32(rc.dof_pos_upper_limit_list[i]).item()
            self.hard_dof_pos_limits[i, 0] = lo
            self.hard_dof_pos_limits[i, 1] = hi
            self.dof_pos_limits[i, 0] = lo
            self.dof_pos_limits[i, 1] = hi
            self.dof_vel_limits[i] = rc.dof_vel_limit_list[i]
            self.torque_limits[i] = rc.dof_effort_limit_list[i]
            m = (self.dof_pos_limits[i, 0] + self.dof_pos_limits[i, 1]) / 2
            r = self.dof_pos_limits[i, 1] - self.dof_pos_limits[i, 0]
            self.dof_pos_limits[i, 0] = m - 0.5 * r * self.env_config.rewards.reward_limit.soft_dof_pos_limit
            self.dof_pos_limits[i, 1] = m + 0.5 * r * self.env_config.rewards.reward_limit.soft_dof_pos_limit
            s = self.env_config.termination_scales.termination_close_to_dof_pos_limit
            self.dof_pos_limits_termination[i, 0] = m - 0.5 * r * s
            self.dof_pos_limits_termination[i, 1] = m + 0.5 * r * s
        return self.dof_pos_limits, self.dof_vel_limits, self.torque_limits

    def find_rigid_body_indice(self, name):
        return self._body_list.index(name) if name in self._body_list else -1

    def prepare_sim(self):
        N, Jb, D, dev = self.num_envs, self.num_bodies, self.num_dof, self.device
        self._rigid_body_state = torch.zeros(N * Jb, 13, device=dev)
        self._rigid_body_state[:, 6] = 1.0
        v = self._rigid_body_state.view(N, Jb, 13)
        self._rigid_body_state_reshaped = v
        self._rigid_body_pos, self._rigid_body_rot = v[..., :Jb, 0:3], v[..., :Jb, 3:7]
        self._rigid_body_vel, self._rigid_body_ang_vel = v[..., :Jb, 7:10], v[..., :Jb, 10:13]
        self.all_root_states = torch.zeros(N, 13, device=dev)
        self.all_root_states[:, 6] = 1.0
        self.robot_root_states = self.all_root_states.view(N, 1, 13)[..., 0, :]
        self.base_quat = self.robot_root_states[..., 3:7]
        self.dof_state = torch.zeros(N * D, 2, device=dev)
        self.dof_pos = self.dof_state.view(N, -1, 2)[..., 0]
        self.dof_vel = self.dof_state.view(N, -1, 2)[..., 1]
        self.contact_forces = torch.zeros(N, Jb, 3, device=dev)

    def refresh_sim_tensors(self):
        pass

    def set_actor_root_state_tensor(self, ids, states):
        pass

    def set_dof_state_tensor(self, ids, states):
        pass

    def apply_torques_at_dof(self, torques):
        pass

    def simulate_at_each_physics_step(self):
        pass

    def render(self, *a, **k):
        pass

    # ---- harness side ------------------------------------------------------------------------
    def set_state(self, body_state, root_state, dof_state, contact_forces):
        """copy synthetic state INTO the persistent buffers (the env holds views of them)."""
        self._rigid_body_state.copy_(body_state.reshape(-1, 13))
        self.all_root_states.copy_(root_state)
        self.dof_state.copy_(dof_state.reshape(-1, 2))
        self.contact_forces.copy_(contact_forces)


def make_env(cfg, env_cls="motion_tracking", seed=0):
    """Instantiates the reference env on the FakeSim through its own __init__."""
    os.chdir(R.REF_ROOT)
    ns = R.load_modules()
    import hydra.utils as hu
    hu.get_class = lambda target: FakeSim
    torch.manual_seed(seed)
    np.random.seed(seed)
    if env_cls == "motion_tracking":
        from humanoidverse.envs.motion_tracking.motion_tracking import LeggedRobotMotionTracking as Cls
    else:
        from humanoidverse.envs.motion_tracking.general_tracking import LeggedRobotGeneralTracking as Cls
    env = Cls(cfg, "cpu")
    return env, ns

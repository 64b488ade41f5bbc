"""The benchmark workload: config A/B of SURVEY.md 8 (G1 23-DoF `+exp=motion_tracking +obs=motion_tracking/main
+rewards=motion_tracking/main`, single clip Horse-stance_pose, synthetic simulator state) assembled from the
committed fixtures only.  Shared by bench.py, __graft_entry__.smoke() and the GPU tests so they all drive the
product through the same calls.  Nothing here imports oracle/ or /root/reference.
"""
import types

import torch

from . import synth
from .fixtures import env_config, product_cfg
from pbhc_b200.spec import TrackSpec

CLIP = "Horse-stance_pose"
CASE = "mt23_main"  # obs=motion_tracking/main with the four observation-noise scales set to 0 (= benchmark.yaml's choice)
CASE_C = "gt29_teacher"  # config C of SURVEY.md 8: general_tracking, G1 29-DoF, obs_ppo_teacher, rewards general_main, 3-clip library


def step_bytes(spec, contact_size=2):
    """Algorithmic HBM bytes of ONE fused env step (SURVEY.md 8d row "Fused tracked env-step"), recomputed from
    the live config.  The motion table is excluded when it is a single L2-resident clip."""
    D, J, Jb, L = spec.D, spec.J, spec.Jb, spec.L
    n_sums = len(spec.reward_names) + (1 if spec.termination_scale is not None else 0)
    hist_step = sum(spec.hist_dim[k] for k in spec.hist_keys)
    hist_all = sum(spec.hist_dim[k] * spec.hist_len[k] for k in spec.hist_keys)
    reads = dict(body_state=Jb * 13 * 4, root_state=13 * 4, dof_state=D * 2 * 4, contact_forces=Jb * 3 * 4,
                 actions_last_actions_last_dof_vel_torques=4 * D * 4,
                 per_env_scalars=8 + 8 + 4 + 4 + 12,            # episode_length, motion_id, start time, motion_len, origin
                 stateful_reward_state=4 * contact_size * 2,    # feet_air_time, last_contacts
                 privileged_dr=12 + L * 4 + 2 * D * 4 + 4 + 8,  # com bias, link mass, kp, kd, friction, ctrl delay
                 default_dof_pos=D * 4, history=hist_all * 4, episode_sums=n_sums * 4)
    rew_w = (len(spec.reward_names) + 1) * 4 if spec.use_vec_reward else 4
    writes = dict(obs=sum(g["dim"] for g in spec.obs_groups.values()) * 4, rew_buf=rew_w, reset_timeout=8 + 1,
                  episode_sums=n_sums * 4, history_push=hist_step * 4,
                  last_state=3 * D * 4 + 8 + 4 * 4 * 2 + 24,     # last_actions/dof_pos/dof_vel, episode_length, air/contacts, times
                  extras_ref=J * 7 * 4)
    return sum(reads.values()) + sum(writes.values()), reads, writes


def fk_bytes(J):
    """K2 per pose: pose_aa + trans in; positions, xyzw quaternions and 3x3 matrices out."""
    return J * 12 + 12 + J * (12 + 16 + 36)


class Workload:
    """Everything one rank needs to run env steps on `device` for `n_envs` environments."""

    def __init__(self, n_envs, device, tmpdir, seed=0, case=CASE, clip=CLIP, noise=None, library_clips=0, push_robots=None, env_offset=0):
        """library_clips > 0 (config C only): the SURVEY 8d motion library of that many 29-DoF clips (fixtures.library29)
        instead of the 3-clip one, so that the table is served from HBM."""
        from .fixtures import CLIPS29, library29
        from pbhc_b200.motion_lib import MotionLibRobot
        from pbhc_b200.tracking_step import TrackingStep
        self.N, self.device = int(n_envs), torch.device(device)
        self.spec = spec = TrackSpec.from_config(env_config(case, noise, push_robots))
        self.tag = "29" if spec.general else "23"
        self.clips = list(CLIPS29) if spec.general else [clip]
        self.library_clips = int(library_clips) if spec.general else 0
        data = library29(self.library_clips, seed=0) if self.library_clips else None
        self.library = data
        if data is not None:
            self.clips = list(data)
        self.lib = MotionLibRobot(product_cfg(self.tag, tmpdir, self.clips, motion_data=data), self.N, self.device)
        self.lib.load_motions(random_sample=False)
        self.ts = TrackingStep(spec, self.lib, self.N, self.device, seed=seed, env_offset=env_offset)  # env_offset: this rank's first global env (dist.shard_range)
        self.host = self.make_inputs(seed)
        dev = self.device
        h = self.host
        self.sim = types.SimpleNamespace(
            _rigid_body_state=h["body_state"].reshape(-1, 13).to(dev), all_root_states=h["root_state"].to(dev),
            dof_state=h["dof_state"].reshape(-1, 2).to(dev), contact_forces=h["contact_forces"].to(dev),
            _base_com_bias=h["base_com_bias"].to(dev), _link_mass_scale=h["link_mass_scale"].to(dev),
            friction_coeffs=h["friction_coeffs"].to(dev), _base_mass_scale=h["base_mass_scale"].to(dev))
        self.actions, self.torques = h["actions"].to(dev), h["torques"].to(dev)
        self.env_origins, self.motion_ids = h["env_origins"].to(dev), h["motion_ids"].to(dev)
        self.pose_aa, self.trans = h["pose_aa"].to(dev), h["trans"].to(dev)
        self.load_env_state(h)
        self._fk_stream = None  # side stream of step(): K2 overlaps the all-reduce + finalize tail
        self.ts.bind(self.sim, self.actions, self.torques, self.env_origins, self.motion_ids)

    # ---- seeded synthetic inputs (CPU tensors: the CPU oracle consumes the very same bits) -------------------
    def make_inputs(self, seed):
        N, spec, dev = self.N, self.spec, self.device
        D, Jb, J = spec.D, spec.Jb, spec.J
        gen = torch.Generator().manual_seed(seed)
        h = {}
        h["motion_ids"] = torch.arange(N, dtype=torch.int64)  # per-env motion ids; load_motions maps id -> clip (id % num clips)
        h["motion_len"] = self.lib.get_motion_length(h["motion_ids"].to(dev)).cpu()
        h["motion_start_times"] = torch.rand(N, generator=gen) * h["motion_len"]
        h["episode_length_buf"] = torch.randint(0, 50, (N,), generator=gen)
        side = int(N ** 0.5) + 1  # env origins on a grid, spacing 5 m (base_task.py:136-142)
        idx = torch.arange(N)
        h["env_origins"] = torch.stack([(idx // side).float() * 5.0, (idx % side).float() * 5.0, torch.zeros(N)], dim=1)
        t = (h["episode_length_buf"] + 2) * spec.dt + h["motion_start_times"]
        ref = self.lib.get_motion_state(h["motion_ids"].to(dev), t.to(dev), h["env_origins"].to(dev))
        ref = {k: v.cpu() for k, v in ref.items()}
        body, root, dof, cf = synth.synth_sim_state(ref, Jb, gen, feet=tuple(spec.feet_ids))
        h.update(body_state=body, root_state=root, dof_state=dof, contact_forces=cf)
        h["actions"] = synth.synth_actions(N, D, gen)
        lim = torch.tensor(spec.torque_limit)
        h["torques"] = torch.maximum(torch.minimum(torch.randn(N, D, generator=gen) * 20.0, lim), -lim)
        h["base_com_bias"] = (torch.rand(N, 3, generator=gen) - 0.5) * 0.1
        h["link_mass_scale"] = 0.9 + 0.2 * torch.rand(N, spec.L, generator=gen)
        h["friction_coeffs"] = 0.2 + torch.rand(N, 1, 1, generator=gen)
        h["base_mass_scale"] = 0.9 + 0.2 * torch.rand(N, 1, generator=gen) if spec.general else torch.ones(N, 1)
        h["kp_scale"] = 0.9 + 0.2 * torch.rand(N, D, generator=gen)
        h["kd_scale"] = 0.9 + 0.2 * torch.rand(N, D, generator=gen)
        h["last_actions"] = synth.synth_actions(N, D, gen)
        h["last_dof_vel"] = dof[..., 1] + torch.randn(N, D, generator=gen) * 0.1
        # FK input of the step: the reference pose of each env's current frame (fk_batch(B=N, T=1), BASELINE configs[0])
        h["pose_aa"] = torch.randn(N, J, 3, generator=gen) * 0.3
        h["trans"] = root[:, :3].clone()
        return h

    def load_env_state(self, h):
        ts = self.ts
        ts.episode_length_buf.copy_(h["episode_length_buf"])
        push_word = ts.env_f32[:, 7].clone()  # push_robot_counter | push_interval_s << 16 (drawn at construction)
        ts.env_f32.zero_()
        ts.env_f32[:, 7] = push_word
        ts.env_f32[:, 0].copy_(h["motion_start_times"])
        ts.env_f32[:, 1].copy_(h["motion_len"])
        ts._kp_scale.copy_(h["kp_scale"])
        ts._kd_scale.copy_(h["kd_scale"])
        ts.last_actions.copy_(h["last_actions"])
        ts.last_dof_vel.copy_(h["last_dof_vel"])

    # ---- the step ------------------------------------------------------------------------------------------
    def fk(self):
        """K2 on the N per-env poses: positions + xyzw quaternions + rotation matrices."""
        self.fk_out = self.lib.mesh_parsers.fk(self.pose_aa, self.trans, want_mat=True, want_quat=True, out=getattr(self, "fk_out", None))
        return self.fk_out

    def step(self, with_fk=True):
        """one tracked env step: [K2 FK] + fused lookup/diff/termination/reward/reset/obs/history kernel + finalize.
        K2 does not depend on the fused kernel's results, so it runs on a side stream forked AFTER the fused kernel and
        its block reduction (the persistent kernel keeps every SM to itself) and joined at the end of the step: the
        serial tail of the step - the 64-double all-reduce across ranks and the one-warp finalize kernel - hides behind
        K2 instead of adding to it.  Captured in a CUDA graph this is a fork / join of two branches."""
        if not with_fk:
            self.ts.launch()
            self.ts.finalize()
            return self.ts.obs_buf_dict, self.ts.rew_buf, self.ts.reset_buf
        if self._fk_stream is None:
            self._fk_stream = torch.cuda.Stream(self.device)
        cur = torch.cuda.current_stream(self.device)
        self.ts.launch()
        self._fk_stream.wait_stream(cur)  # forking BEFORE the fused kernel was measured: K2's blocks take SM residency from the persistent grid, 0.399 -> 0.604 ms
        with torch.cuda.stream(self._fk_stream):
            self.fk()
        self.ts.finalize()
        cur.wait_stream(self._fk_stream)
        return self.ts.obs_buf_dict, self.ts.rew_buf, self.ts.reset_buf

    LAUNCHES_PER_STEP = {"fk": 1, "track": 2, "finalize": 1}

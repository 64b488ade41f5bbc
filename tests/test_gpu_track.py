"""GPU parity of the fused tracking step (K3, csrc/track.cu through the C ABI) against golden fixtures of the
UNMODIFIED reference env (tests/golden/ref_env_*.npz) and, at larger sizes, against the CPU oracle
(oracle/pbhc_oracle_env.py).  Masks / indices bit-exact; fp32 within 1e-5 rel / 1e-6 abs."""
import json
import types

import numpy as np
import pytest
import torch

from oracle import pbhc_oracle as O, pbhc_oracle_env as OE
from benchkit import synth
from pbhc_b200 import _lib
from pbhc_b200.spec import TrackSpec, SIGMA_IDS
from tests._util import gold, oracle_clip, product_cfg, assert_close, SLERP_FLIP, ATOL

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
INJECT = ("motion_start_times", "kp_scale", "kd_scale", "rfi_lim_scale", "rao_scale", "action_delay_idx")
NAMES = ["Horse-stance_pose"]
# observation keys that contain reference (table) velocities
VEL_OBS = {"next_step_ref_motion", "future_motion_base_lin_vel", "future_motion_base_yaw_vel", "local_ref_rigid_body_vel",
           "global_ref_rigid_body_vel", "dif_joint_velocities"}


def make_step(spec, N, tmp_path, clip_names=None, exact=True):
    from benchkit.fixtures import CLIPS29
    from pbhc_b200.motion_lib import MotionLibRobot
    from pbhc_b200.tracking_step import TrackingStep
    from tests.test_gpu_parity import _install_exact_table
    tag = "29" if spec.general else "23"   # config C: G1 29-DoF with the 3-clip library, config A/B: 23-DoF single clip
    clip_names = clip_names or (list(CLIPS29) if spec.general else NAMES)
    ml = MotionLibRobot(product_cfg(tag, str(tmp_path), clip_names), N, DEV)
    ml.load_motions(random_sample=False)
    if exact:
        _install_exact_table(ml, [oracle_clip(tag, n) for n in clip_names])
    return TrackingStep(spec, ml, N, DEV), ml


def make_sim(N, Jb, D, L):
    z = lambda *s: torch.zeros(*s, device=DEV)
    return types.SimpleNamespace(_rigid_body_state=z(N * Jb, 13), all_root_states=z(N, 13), dof_state=z(N * D, 2),
                                 contact_forces=z(N, Jb, 3), _base_com_bias=z(N, 3), _link_mass_scale=z(N, L),
                                 friction_coeffs=z(N, 1, 1), _base_mass_scale=torch.ones(N, 1, device=DEV))


def load_state(ts, S, spec):
    """oracle / golden state dict -> TrackingStep buffers (history is key-major, logical order at push count 0)."""
    t = lambda k: torch.as_tensor(np.asarray(S[k])).to(DEV)
    ts.episode_length_buf.copy_(t("episode_length_buf"))
    ts.env_f32[:, 0], ts.env_f32[:, 1] = t("motion_start_times"), t("motion_len")
    ts.env_f32[:, 2:4], ts.env_f32[:, 4:6] = t("feet_air_time"), t("last_contacts").float()
    ts.env_f32[:, 6] = t("end_time_ratio_buf")
    for a, k in ((ts.last_actions, "last_actions"), (ts.last_dof_pos, "last_dof_pos"), (ts.last_dof_vel, "last_dof_vel"),
                 (ts._kp_scale, "kp_scale"), (ts._kd_scale, "kd_scale"), (ts._rfi_lim_scale, "rfi_lim_scale"),
                 (ts._rao_scale, "rao_scale"), (ts.action_delay_idx, "action_delay_idx"), (ts.default_dof_pos, "default_dof_pos")):
        a.copy_(t(k))
    for k, base in zip(spec.hist_keys, ts._hist_base):
        n = spec.hist_len[k] * spec.hist_dim[k]
        ts.hist[:, base:base + n] = t("hist_" + k).reshape(ts.N, -1)
    for i, name in enumerate(ts._sum_names):
        ts._episode_sums[:, i] = t("epsum_" + name)
    if "push_interval_s" in S:
        ts.set_push_state(t("push_robot_counter"), t("push_interval_s"))


def compare_state(ts, G, spec, tag, loose=None):
    """G: dict of expected (reference / oracle) post-step buffers.  loose: see _run_golden_case."""
    rew_tol = (loose or {}).get("rew", {})
    sig_tol = (loose or {}).get("sigma", 1e-6)
    e = lambda k: np.asarray(G[k])
    assert np.array_equal(ts.episode_length_buf.cpu().numpy(), e("episode_length_buf")), tag
    assert np.array_equal(ts.reset_buf.cpu().numpy(), e("reset_buf")), tag
    assert np.array_equal(ts.time_out_buf.cpu().numpy(), e("time_out_buf")), tag
    assert np.array_equal(ts.action_delay_idx.cpu().numpy(), e("action_delay_idx")), tag
    assert_close(ts.rew_buf.cpu(), e("rew_buf"), what=f"{tag} rew_buf", **rew_tol)
    for name, mine in (("motion_start_times", ts.motion_start_times), ("motion_len", ts.motion_len),
                       ("feet_air_time", ts.feet_air_time), ("end_time_ratio_buf", ts.end_time_ratio_buf),
                       ("last_actions", ts.last_actions), ("last_dof_pos", ts.last_dof_pos), ("last_dof_vel", ts.last_dof_vel),
                       ("kp_scale", ts._kp_scale), ("kd_scale", ts._kd_scale), ("rfi_lim_scale", ts._rfi_lim_scale),
                       ("rao_scale", ts._rao_scale)):
        assert_close(mine.cpu(), e(name), what=f"{tag} {name}")
    assert np.array_equal(ts.last_contacts.cpu().numpy(), e("last_contacts").astype(np.float32)), tag
    for k in spec.hist_keys:
        assert_close(ts.history_handler.query(k).cpu(), e("hist_" + k), what=f"{tag} hist_{k}")
    for name, v in ts.episode_sums.items():
        assert_close(v.cpu(), e("epsum_" + name), what=f"{tag} epsum_{name}", **rew_tol)
    sc = ts.scalars.cpu().numpy()
    for name, i in SIGMA_IDS.items():
        if "sigma_" + name not in G:
            continue
        assert abs(sc[_lib.S_SIGMA + i] - float(G["sigma_" + name])) <= sig_tol * abs(float(G["sigma_" + name])), (tag, name)
        if "ema_" + name in G:
            assert abs(sc[_lib.S_EMA + i] - float(G["ema_" + name])) <= sig_tol * abs(float(G["ema_" + name])), (tag, name)
    for name, i in (("average_episode_length", _lib.S_AVG_EPL), ("reward_penalty_scale", _lib.S_PENALTY),
                    ("motion_far_threshold", _lib.S_FAR_THR), ("soft_dof_pos", _lib.S_SOFT_DOF_POS),
                    ("soft_dof_vel", _lib.S_SOFT_DOF_VEL), ("soft_torque", _lib.S_SOFT_TORQUE), ("current_noise_curriculum_value", _lib.S_NOISE_CUR)):
        if name not in G:
            continue
        assert abs(sc[i] - float(G[name])) <= 1e-6 * abs(float(G[name])) + 1e-12, (tag, name, sc[i], float(G[name]))


def _run_golden_case(case, tmp_path, exact, loose=None):
    """loose: dict of looser assert_close tolerances for the quantities that inherit the velocity error of a
    product-built table (None: the north_star bar everywhere)."""
    g = gold(f"ref_env_{case}.npz")
    spec = TrackSpec.from_config(json.loads(str(g["meta_config_json"])))
    N = int(g["meta_n_envs"])
    from benchkit.fixtures import clips_of_case
    ts, ml = make_step(spec, N, tmp_path, clip_names=clips_of_case(json.loads(str(g["meta_config_json"])))[1], exact=exact)
    sim = make_sim(N, spec.Jb, spec.D, spec.L)
    init = {k[5:]: g[k] for k in g if k.startswith("init_")}
    if hasattr(ts, "ref_init_yaw"):  # ref_init_rpy[0, 2]: the step derived it itself from the motion library (K1 at construction)
        assert abs(ts.ref_init_yaw - float(init["ref_init_yaw"])) < (1e-6 if exact else 1e-5), (ts.ref_init_yaw, float(init["ref_init_yaw"]))
    load_state(ts, init, spec)
    if case.endswith("_eval"):
        ts.set_is_evaluating()
    sim._base_com_bias.copy_(torch.from_numpy(init["base_com_bias"]))
    sim._link_mass_scale.copy_(torch.from_numpy(init["link_mass_scale"]))
    sim.friction_coeffs.copy_(torch.from_numpy(init["friction_coeffs"]))
    if "base_mass_scale" in init:
        sim._base_mass_scale.copy_(torch.from_numpy(init["base_mass_scale"]))
    origins = torch.from_numpy(init["env_origins"]).to(DEV)
    mids = torch.from_numpy(init["motion_ids"]).to(DEV)
    for step in range(int(g["meta_n_steps"])):
        p = f"s{step}_"
        sim._rigid_body_state.copy_(torch.from_numpy(g[p + "in_body_state"]).reshape(-1, 13))
        sim.all_root_states.copy_(torch.from_numpy(g[p + "in_root_state"]))
        sim.dof_state.copy_(torch.from_numpy(g[p + "in_dof_state"]).reshape(-1, 2))
        sim.contact_forces.copy_(torch.from_numpy(g[p + "in_contact_forces"]))
        actions = torch.from_numpy(g[p + "in_actions"]).to(DEV).contiguous()
        torques = torch.from_numpy(g[p + "in_torques"]).to(DEV).contiguous()
        inject = {k: torch.from_numpy(g[p + "out_" + k]).to(DEV).contiguous() for k in INJECT}
        if case.endswith("_eval"):
            del inject["motion_start_times"]
        if p + "in_term_gates" in g:
            inject["term_gates"] = int(g[p + "in_term_gates"])
        if p + "in_push_robot_counter" in g:  # domain_rand.push_robots: the fixture moves a few envs next to their push before every step
            ts.set_push_state(torch.from_numpy(g[p + "in_push_robot_counter"]), ts.push_interval_s)
            inject["push_interval_s"] = torch.from_numpy(g[p + "out_push_interval_s"]).to(DEV).contiguous()
            inject["push_vel"] = torch.from_numpy(g[p + "out_push_robot_vel_buf"]).to(DEV).contiguous()
        obs, rew, reset, extras = ts.step(sim, actions, torques, origins, mids, inject=inject)
        torch.cuda.synchronize()
        tag = f"{case} step {step}"
        for name, v in obs.items():
            otol = {}
            if loose:  # columns that show reference velocities inherit the table's velocity error, the rest its position error
                at = torch.full((v.shape[1],), loose["obs"]["atol"], dtype=torch.float64)
                for seg in spec.obs_groups[name]["segs"]:
                    if seg["key"] in VEL_OBS:
                        at[seg["off"]:seg["off"] + seg["dim"] * (seg.get("len") or 1)] = loose["state"]["atol"]
                otol = dict(atol=at, rtol=loose["obs"]["rtol"])
            if not loose and any(seg["key"] == "local_ref_rigid_body_pos_relyaw" for seg in spec.obs_groups[name]["segs"]):
                # reference velocities of several m/s (the 29-DoF clips) rotated by the relyaw quaternion: the angle goes through
                # atan2f - (ref_init yaw, itself an atan2f of a K1 lookup) - sinf / cosf, ~3e-7 rad of fp32 rounding between CUDA
                # and glibc, which shows as that fraction OF THE VECTOR NORM in its small components (observed 1.7e-6)
                at = torch.full((v.shape[1],), ATOL, dtype=torch.float64)
                for seg in spec.obs_groups[name]["segs"]:
                    if seg["key"] == "local_ref_rigid_body_pos_relyaw":
                        at[seg["off"]:seg["off"] + seg["dim"]] = 4e-6
                otol = dict(atol=at)
            assert_close(v.cpu(), g[p + "obs_" + name], what=f"{tag} obs {name}", **otol)
        G = {k[len(p) + 4:]: g[k] for k in g if k.startswith(p + "out_")}
        compare_state(ts, G, spec, tag, loose)
        if "push_interval_s" in G:  # counters, new intervals, velocities of the pushed envs, need_to_refresh_envs = reset | pushed
            assert np.array_equal(ts.push_robot_counter.cpu().numpy(), G["push_robot_counter"]), tag
            assert np.array_equal(ts.push_interval_s.cpu().numpy(), G["push_interval_s"]), tag
            pushed = G["push_robot_counter"] == 0
            assert pushed.sum() >= 3
            assert np.array_equal(ts.push_robot_vel_buf.cpu().numpy()[pushed], G["push_robot_vel_buf"][pushed]), tag
            want = np.flatnonzero(pushed | (G["reset_buf"] != 0))
            assert np.array_equal(ts.refresh_env_ids_host().cpu().numpy(), want), tag
            assert np.array_equal(np.flatnonzero(ts.need_to_refresh_envs.cpu().numpy()), want), tag
        assert_close(actions.cpu(), g[p + "out_actions"], what=f"{tag} actions")
        assert_close(sim.dof_state.cpu(), g[p + "out_dof_state"], what=f"{tag} dof_state")
        assert_close(sim.all_root_states.cpu(), g[p + "out_root_state"], what=f"{tag} root_state", **(loose or {}).get("state", {}))
        assert_close(extras["ref_body_pos_extend"].cpu(), g[p + "extras_ref_body_pos_extend"], what=f"{tag} ref pos")
        assert_close(extras["ref_body_rot_extend"].cpu(), g[p + "extras_ref_body_rot_extend"], what=f"{tag} ref rot", allow=SLERP_FLIP if loose else None)
        log = ts.log_dict
        for k in g:
            if k.startswith(p + "log_") and k[len(p) + 4:] in log:
                assert_close(log[k[len(p) + 4:]].cpu(), np.float32(g[k]), what=f"{tag} log {k}", **(loose or {}).get("log", dict(rtol=1e-5, atol=1e-6)))


@pytest.mark.parametrize("case", ["mt23_main", "mt23_bench", "gt29_teacher", "gt29_student", "mt23_edge", "gt29_edge", "mt23_term", "gt29_term", "mt23_p4", "gt29_p4", "mt23_push", "gt29_push", "mt23_eval", "gt29_eval"])
def test_track_step_matches_reference_golden(case, tmp_path):
    """*_eval: the env after set_is_evaluating() (a reset starts its clip at time 0 - not injected here) with the observation-noise
    curriculum on (its value, log and borrowed thresholds; legged_robot_base.py:637-645).
    *_push: domain_rand.push_robots as the shipped domain_rand/main.yaml has it (legged_robot_base.py:382-388, 1090-1106): pushes
    added to / replacing the root velocity, an env that is pushed and reset in the same step, need_to_refresh_envs.
    *_p4: the observation sources and reward terms no shipped config uses (tools/make_golden_env_step.py run_p4): relyaw,
    local_ref_rigid_body_pos_relyaw, vr_3point_pos (with the reference's cross-env heading pairing), future_ref_dof_pos / _vel,
    the radial velocity rewards, feet_heading_alignment(_contact), penalty_feet_ori(_contact).
    mt23_term: the base-class terminations on (contact with |f| just above / below 1, low base, dof position / velocity /
    torque limits behind the reference's own per-step gates; legged_robot_base.py:433-476).  *_edge: the synthetic clips of tools/make_golden_edge.py - envs sitting on identical frames, sign-flipped frame
    quaternions and sin(half angle) on both sides of slerp's 1e-3; in gt29_edge also body rotations within 0 .. 40 urad of
    the reference's, i.e. quat_to_angle_axis on both sides of its 1e-5 threshold (rotations.py:185-232)"""
    _run_golden_case(case, tmp_path, exact=True)


def test_general_tracking_without_the_compact_future_table(tmp_path, monkeypatch):
    """the future window read from the packed motion table itself (PbhcTrackIO.fut_table = NULL) gives the same golden
    results as through the compact per-frame records every other general_tracking test uses"""
    monkeypatch.setenv("PBHC_FUTURE_TABLE", "0")
    _run_golden_case("gt29_teacher", tmp_path, exact=True)


@pytest.mark.parametrize("case", ["mt23_main", "gt29_teacher"])
def test_track_step_on_product_built_table_matches_reference_golden(case, tmp_path):
    """the fused step on the motion table load_motions builds on the device (no oracle table installed) against the
    goldens of the unmodified reference env: masks, episode lengths and frame-dependent integers bit-exact; observations
    at the north_star bar (none of these configs observes reference velocities); the quantities that contain the table's
    body / angular velocities - the velocity rewards and their episode sums, adaptive sigma, root / dof state written at
    resets - at the error measured for the table (tools/table_error.py, profiles/r02_table_error.json)."""
    loose = dict(rew=dict(atol=5e-5, rtol=1e-4), state=dict(atol=8e-3, rtol=0.0), sigma=2e-3, log=dict(rtol=1e-4, atol=1e-5),
                 obs=dict(atol=5e-6, rtol=1e-5))
    _run_golden_case(case, tmp_path, exact=False, loose=loose)


@pytest.mark.parametrize("n_envs", [1, 37, 4096])
def test_workload_step_vs_oracle(n_envs, tmp_path):
    """the bench workload itself (ragged and default-training sizes) against the CPU oracle, incl. resets."""
    from benchkit.workload import Workload
    from tests._util import check_workload_step_against_oracle
    wl = Workload(n_envs, DEV, str(tmp_path), seed=3)
    n_reset = check_workload_step_against_oracle(wl, n_steps=3)
    if n_envs >= 4096:
        assert n_reset > 0


@pytest.mark.parametrize("n_envs", [37, 4096])
def test_workload_step_vs_oracle_general_tracking(n_envs, tmp_path):
    """config C (general_tracking, G1 29-DoF, 3-clip library with per-env motion ids and random start times) against the
    CPU oracle at ragged and training sizes, incl. resets and the 20-step future window."""
    from benchkit.workload import Workload, CASE_C
    from tests._util import check_workload_step_against_oracle
    wl = Workload(n_envs, DEV, str(tmp_path), seed=4, case=CASE_C)
    n_reset = check_workload_step_against_oracle(wl, n_steps=3)
    if n_envs >= 4096:
        assert n_reset > 0


def test_workload_step_vs_oracle_general_tracking_large_library(tmp_path):
    """config C on the SURVEY 8d style library (fixtures.library29: the 29-DoF clip + the zero-padded, time-warped 23-DoF
    clips; 96 clips here so that the CPU oracle builds it in seconds - bench.py runs 1024): per-env motion ids over all
    clips, the table built by ONE batched FK + pbhc_clips_prepare launch"""
    from benchkit.workload import Workload, CASE_C
    from tests._util import check_workload_step_against_oracle
    wl = Workload(4096, DEV, str(tmp_path), seed=6, case=CASE_C, library_clips=96)
    assert wl.lib._desc.num_clips == 96
    n_reset = check_workload_step_against_oracle(wl, n_steps=2)
    assert n_reset > 0


def test_host_step_chunked_pipeline_equals_single_launch(tmp_path):
    """HostStep (host buffers, 5 ragged env chunks on 3 streams) gives bit-identical per-env results to one launch;
    in-kernel reset draws are keyed by the global env id, so they do not depend on the chunking either."""
    from benchkit.workload import Workload
    from pbhc_b200.tracking_step import HostStep
    n = 1003
    a = Workload(n, DEV, str(tmp_path / "a"), seed=7)
    b = Workload(n, DEV, str(tmp_path / "b"), seed=7)
    pin = lambda t: t.contiguous().pin_memory()
    host_in = {k: pin(b.host[k]) for k in ("body_state", "root_state", "dof_state", "contact_forces", "actions", "torques")}
    hs = HostStep(b.ts, b.sim, b.actions, b.torques, b.env_origins, b.motion_ids, host_in, chunks=5)
    assert len(hs.ranges) >= 4 and hs.ranges[-1][1] == n
    for step in range(3):
        a.sim._rigid_body_state.copy_(a.host["body_state"].reshape(-1, 13))  # what the host step re-uploads every step
        a.sim.all_root_states.copy_(a.host["root_state"])
        a.sim.dof_state.copy_(a.host["dof_state"].reshape(-1, 2))
        a.actions.copy_(a.host["actions"])
        a.step(with_fk=False)
        out = hs.step()
        torch.cuda.synchronize()
        for name, v in a.ts.obs_buf_dict.items():
            assert torch.equal(v.cpu(), out[name]), (step, name)
        assert torch.equal(a.ts.rew_buf.cpu().view(n, -1), out["rew_buf"])
        assert torch.equal(a.ts.reset_buf.cpu().view(n, 1), out["reset_buf"])
        assert torch.equal(a.ts.time_out_buf.cpu().view(n, 1), out["time_out_buf"])
        assert torch.equal(a.ts.env_f32.cpu(), b.ts.env_f32.cpu())
        assert torch.equal(a.ts.hist.cpu(), b.ts.hist.cpu())
        assert torch.allclose(a.ts.scalars.cpu(), b.ts.scalars.cpu(), rtol=1e-12, atol=0)
    assert int(a.ts.reset_buf.sum()) > 0


def test_observation_noise_statistics(tmp_path):
    """obs/motion_tracking/main.yaml noise: (value + U(-1,1)*noise_scale) * obs_scale, an independent draw per element,
    per observation group and for the pushed history element (helpers.py:128-152).  torch.rand_like cannot be replayed,
    so parity here is distributional: bounds, mean 0, variance 1/3, independence between groups, exact equality of the
    noise-free columns and of everything that is not an observation, reproducibility under the seed."""
    from benchkit.fixtures import MAIN_NOISE
    from benchkit.workload import Workload
    n = 8192
    clean = Workload(n, DEV, str(tmp_path / "c"), seed=5)
    noisy = Workload(n, DEV, str(tmp_path / "n"), seed=5, noise=MAIN_NOISE)
    again = Workload(n, DEV, str(tmp_path / "a"), seed=5, noise=MAIN_NOISE)
    other = Workload(n, DEV, str(tmp_path / "o"), seed=6, noise=MAIN_NOISE)
    other.load_env_state(noisy.host)  # same inputs, different seed (Workload draws its inputs from the seed too)
    for name in ("_rigid_body_state", "all_root_states", "dof_state", "contact_forces", "_base_com_bias", "_link_mass_scale", "friction_coeffs"):
        getattr(other.sim, name).copy_(getattr(noisy.sim, name))
    other.actions.copy_(noisy.actions); other.torques.copy_(noisy.torques); other.env_origins.copy_(noisy.env_origins)
    for wl in (clean, noisy, again, other):
        wl.step(with_fk=False)
    torch.cuda.synchronize()
    spec = noisy.spec
    assert torch.equal(clean.ts.rew_buf, noisy.ts.rew_buf) and torch.equal(clean.ts.reset_buf, noisy.ts.reset_buf)
    keep = ~clean.ts.reset_buf.bool()
    units = {}
    for g, grp in spec.obs_groups.items():
        a, b = clean.ts.obs_buf_dict[g], noisy.ts.obs_buf_dict[g]
        assert torch.equal(b, again.ts.obs_buf_dict[g])
        for seg in grp["segs"]:
            cols = slice(seg["off"], seg["off"] + (seg["dim"] if seg["kind"] == "src" else seg["len"] * seg["dim"]))
            ns = seg.get("noise", 0.0)
            if seg["kind"] == "src" and ns != 0:
                u = ((b[:, cols] - a[:, cols]) / (ns * seg["scale"]))[keep].double()
                assert u.abs().max() <= 1.0 + 1e-4
                assert abs(u.mean().item()) < 0.02 and abs(u.var().item() - 1.0 / 3.0) < 0.02
                lo, hi = (u < -0.5).double().mean().item(), (u > 0.5).double().mean().item()
                assert abs(lo - 0.25) < 0.02 and abs(hi - 0.25) < 0.02
                if u.shape[1] > 1:  # neighbouring elements are independent
                    assert abs((u[:, 0] * u[:, 1]).mean().item()) < 0.02
                units[(g, seg["key"])] = u
                o = ((other.ts.obs_buf_dict[g][:, cols] - a[:, cols]) / (ns * seg["scale"]))[keep].double()
                assert abs((o * u).mean().item()) < 0.02 and not torch.equal(o, u)
            elif seg["kind"] == "src":
                assert torch.equal(a[:, cols], b[:, cols]), seg["key"]
    groups = list(spec.obs_groups)
    shared = [k for (g, k) in units if g == groups[0] and (groups[1], k) in units]
    assert shared
    for k in shared:  # actor and critic see different draws
        assert abs((units[(groups[0], k)] * units[(groups[1], k)]).mean().item()) < 0.02
    for k in spec.hist_keys:  # the pushed history element carries its own draw
        ns = spec.noise_scales.get(k, 0.0)
        new_c, new_n = clean.ts.history_handler.query(k)[:, 0], noisy.ts.history_handler.query(k)[:, 0]
        if ns == 0:
            assert torch.equal(new_c, new_n), k
        else:
            u = ((new_n - new_c) / (ns * spec.obs_scales[k]))[keep].double()
            assert u.abs().max() <= 1.0 + 1e-4 and abs(u.var().item() - 1.0 / 3.0) < 0.02
            for g in groups:
                if (g, k) in units:
                    assert abs((u * units[(g, k)]).mean().item()) < 0.02


@pytest.mark.parametrize("case", ["mt23_push", "gt29_push"])
def test_push_robots_in_kernel_draws(case, tmp_path):
    """without injected draws the pushes come from the kernel's Philox stream: every env one step short of its interval is
    pushed, gets a new interval in [lo, hi) and a velocity in U(-max, max) that is added to (mt23: _push_fixed) or replaces
    (gt29) the root lin vel x / y of envs that do not reset; envs not at their interval only count up"""
    g = gold(f"ref_env_{case}.npz")
    cfg = json.loads(str(g["meta_config_json"]))
    spec = TrackSpec.from_config(cfg)
    N = 2048
    from benchkit.fixtures import clips_of_case
    ts, ml = make_step(spec, N, tmp_path, clip_names=clips_of_case(cfg)[1], exact=False)
    sim = make_sim(N, spec.Jb, spec.D, spec.L)
    mids = torch.zeros(N, device=DEV, dtype=torch.int64)
    # the robot sits exactly on its reference (nothing terminates), except for the root velocity the pushes act on
    ref = ml.get_motion_state(mids, torch.full((N,), 2 * spec.dt, device=DEV))
    body = sim._rigid_body_state.view(N, -1, 13)
    body[:, :, 0:3], body[:, :, 3:7] = ref["rg_pos_t"][:, :spec.Jb], ref["rg_rot_t"][:, :spec.Jb]
    body[:, :, 7:10], body[:, :, 10:13] = ref["body_vel_t"][:, :spec.Jb], ref["body_ang_vel_t"][:, :spec.Jb]
    sim.all_root_states[:, 0:3], sim.all_root_states[:, 3:7] = ref["root_pos"], ref["root_rot"]
    sim.all_root_states[:, 10:13] = ref["root_ang_vel"]
    sim.all_root_states[:, 7:9] = 0.25
    sim.dof_state.view(N, -1, 2)[:, :, 0], sim.dof_state.view(N, -1, 2)[:, :, 1] = ref["dof_pos"], ref["dof_vel"]
    lo, hi = spec.dr["push_interval_s"]
    iv = ts.push_interval_s.clone()
    assert int(iv.min()) >= lo and int(iv.max()) < hi and ts.push_robot_vel_buf.shape == (N, 2)
    thr = (iv / spec.dt).int()
    due = torch.arange(N, device=DEV) % 2 == 0
    ts.set_push_state(torch.where(due, thr - 1, torch.full_like(thr, 7)), iv)
    ts.env_f32[:, 1] = 1e6  # (no motion-end resets)
    z = torch.zeros(N, spec.D, device=DEV)
    ts.step(sim, z.clone(), z.clone(), torch.zeros(N, 3, device=DEV), mids)
    torch.cuda.synchronize()
    cnt, iv2, vel, reset = ts.push_robot_counter, ts.push_interval_s, ts.push_robot_vel_buf, ts.reset_buf != 0
    assert torch.equal(cnt[due], torch.zeros_like(cnt[due])) and bool((cnt[~due] == 8).all())
    assert torch.equal(iv2[~due], iv[~due]) and int(iv2[due].min()) >= lo and int(iv2[due].max()) < hi
    assert len(torch.unique(iv2[due])) == hi - lo
    m = spec.dr["max_push_vel_xy"]
    v = vel[due]
    assert float(v.abs().max()) <= m and abs(float(v.mean())) < 0.06 * m and abs(float(v.var()) - m * m / 3) < 0.1 * m * m
    assert bool((vel[~due] == 0).all())
    keep = due & ~reset
    assert int(keep.sum()) > N // 4
    got = sim.all_root_states[keep, 7:9]
    want = vel[keep] + 0.25 if spec.dr["push_fixed"] else vel[keep]
    assert torch.equal(got, want)
    assert bool((sim.all_root_states[~due & ~reset, 7:9] == 0.25).all())
    assert torch.equal(ts.need_to_refresh_envs, due | reset)


def test_observation_noise_follows_the_noise_curriculum(tmp_path):
    """obs.add_noise_currculum: every noise scale is multiplied by current_noise_curriculum_value (helpers.py:142) - with
    dof_vel noise 1.0 and the value at 0.5 the dof_vel observations deviate from the noise-free ones by U(-0.5, 0.5) x obs scale"""
    g = gold("ref_env_mt23_eval.npz")
    outs = []
    for noise in (0.0, 1.0):
        cfg = json.loads(str(g["meta_config_json"]))
        cfg["obs"]["noise_scales"]["dof_vel"] = noise
        spec = TrackSpec.from_config(cfg)
        N = int(g["meta_n_envs"])
        ts, ml = make_step(spec, N, tmp_path / str(noise), clip_names=NAMES, exact=True)
        sim = make_sim(N, spec.Jb, spec.D, spec.L)
        init = {k[5:]: g[k] for k in g if k.startswith("init_")}
        load_state(ts, init, spec)
        assert float(ts.scalars[_lib.S_NOISE_CUR]) == 0.5
        sim._rigid_body_state.copy_(torch.from_numpy(g["s0_in_body_state"]).reshape(-1, 13))
        sim.all_root_states.copy_(torch.from_numpy(g["s0_in_root_state"]))
        sim.dof_state.copy_(torch.from_numpy(g["s0_in_dof_state"]).reshape(-1, 2))
        sim.contact_forces.copy_(torch.from_numpy(g["s0_in_contact_forces"]))
        a = torch.from_numpy(g["s0_in_actions"]).to(DEV).contiguous()
        t = torch.from_numpy(g["s0_in_torques"]).to(DEV).contiguous()
        inject = {k: torch.from_numpy(g["s0_out_" + k]).to(DEV).contiguous() for k in INJECT}
        obs, *_ = ts.step(sim, a, t, torch.from_numpy(init["env_origins"]).to(DEV), torch.from_numpy(init["motion_ids"]).to(DEV), inject=inject)
        torch.cuda.synchronize()
        seg = next(s for s in spec.obs_groups["actor_obs"]["segs"] if s["key"] == "dof_vel")
        outs.append((obs["actor_obs"][:, seg["off"]:seg["off"] + seg["dim"]].cpu().double(), seg["scale"], obs["actor_obs"].cpu().clone(), seg))
    (clean, sc, full0, seg), (noisy, _, full1, _) = outs
    u = (noisy - clean) / sc
    assert float(u.abs().max()) <= 0.5 + 1e-6 and float(u.abs().max()) > 0.4
    assert abs(float(u.mean())) < 0.05 and abs(float(u.var()) - 0.25 / 3) < 0.02
    full0[:, seg["off"]:seg["off"] + seg["dim"]] = 0
    full1[:, seg["off"]:seg["off"] + seg["dim"]] = 0
    hist_cols = [s for s in spec.obs_groups["actor_obs"]["segs"] if s["kind"] == "hist"]
    assert hist_cols  # (the history rows hold earlier pushes: untouched by this step's noise)
    assert torch.equal(full0, full1)

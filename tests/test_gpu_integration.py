"""The drop-in boundary as INTEGRATION.md shows it: the env binding block of that file is executed VERBATIM on a shim of
LeggedRobotMotionTracking whose simulator records the setter calls (simulator/isaacgym/isaacgym.py:642-660), and every
step is compared with the golden fixture of the unmodified reference env - observations, masks, per-env state, the
env_ids handed to the simulator, extras["episode"] (against the oracle, pinned bit-exact to the same fixture)."""
import json
import os
import types

import numpy as np
import pytest
import torch

from oracle import pbhc_oracle_env as OE
from pbhc_b200 import _lib
from pbhc_b200.spec import TrackSpec
from tests._util import gold, assert_close
from tests.test_gpu_track import DEV, INJECT, make_step, make_sim, load_state, compare_state
from tests.test_oracle_env_golden import load_case, state_from

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def binding_source():
    text = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    a, b = text.index("# --- begin env binding ---"), text.index("# --- end env binding ---")
    return text[a:b]


class Cfg(dict):
    """dict with attribute access, standing in for the reference's OmegaConf DictConfig"""

    def __getattr__(self, k):
        try:
            v = self[k]
        except KeyError:
            raise AttributeError(k)
        return Cfg(v) if isinstance(v, dict) else v


class RecordingSim(types.SimpleNamespace):
    """Isaac Gym tensors + the two indexed setters of the reference simulator; a setter records (env_ids, rows it was given)"""

    def set_actor_root_state_tensor(self, set_env_ids, root_states):
        assert root_states.data_ptr() == self.all_root_states.data_ptr()
        ids = set_env_ids.to(torch.int32)  # what the reference setter does first (isaacgym.py:643)
        self.calls.append(("root", ids.cpu().numpy().copy(), len(ids)))

    def set_dof_state_tensor(self, set_env_ids, dof_states):
        assert dof_states.data_ptr() == self.dof_state.data_ptr()
        ids = set_env_ids.to(torch.int32)
        self.calls.append(("dof", ids.cpu().numpy().copy(), len(ids)))


def make_env(case, tmp_path, resample=False, interval=1 << 30):
    g = gold(f"ref_env_{case}.npz")
    cfg = json.loads(str(g["meta_config_json"]))
    cfg["resample_motion_when_training"] = resample
    spec = TrackSpec.from_config(cfg)
    N = int(g["meta_n_envs"])
    from benchkit.fixtures import clips_of_case
    _, ml = make_step(spec, N, tmp_path, clip_names=clips_of_case(cfg)[1], exact=True)
    sim = make_sim(N, spec.Jb, spec.D, spec.L)
    sim = RecordingSim(**vars(sim), calls=[])
    init = {k[5:]: g[k] for k in g if k.startswith("init_")}
    env = types.SimpleNamespace(config=Cfg(cfg), _motion_lib=ml, num_envs=N, device=DEV, simulator=sim, extras={}, log_dict={},
                                common_step_counter=0, resample_time_interval=interval,
                                env_origins=torch.from_numpy(init["env_origins"]).to(DEV),
                                motion_ids=torch.from_numpy(init["motion_ids"]).to(DEV), actions=None, torques=None, _step=0)
    ns = {}
    exec(compile(binding_source(), "INTEGRATION.md", "exec"), ns)
    env._init_b200 = types.MethodType(ns["_init_b200"], env)
    env._post_physics_step = types.MethodType(ns["_post_physics_step"], env)

    def refresh(self=env):  # stands in for the physics step + _refresh_sim_tensors: this step's golden simulator state
        p = f"s{self._step}_"
        sim._rigid_body_state.copy_(torch.from_numpy(g[p + "in_body_state"]).reshape(-1, 13))
        sim.all_root_states.copy_(torch.from_numpy(g[p + "in_root_state"]))
        sim.dof_state.copy_(torch.from_numpy(g[p + "in_dof_state"]).reshape(-1, 2))
        sim.contact_forces.copy_(torch.from_numpy(g[p + "in_contact_forces"]))
    env._refresh_sim_tensors = refresh
    env._init_b200()
    load_state(env._b200, init, spec)
    sim._base_com_bias.copy_(torch.from_numpy(init["base_com_bias"]))
    sim._link_mass_scale.copy_(torch.from_numpy(init["link_mass_scale"]))
    sim.friction_coeffs.copy_(torch.from_numpy(init["friction_coeffs"]))
    return env, g, spec, N


@pytest.mark.parametrize("case", ["mt23_main", "mt23_push"])
def test_integration_snippet_runs_verbatim_against_reference_golden(case, tmp_path):
    """mt23_push: domain_rand.push_robots on - the setters receive need_to_refresh_envs = reset | pushed (legged_robot_base.py:314-319)"""
    env, g, spec, N = make_env(case, tmp_path)
    sim, ts = env.simulator, env._b200
    # the attributes the binding re-exports are the very buffers of the step, not copies
    assert env.reset_buf.data_ptr() == ts.reset_buf.data_ptr() and env.obs_buf_dict is ts.obs_buf_dict
    _, ospec, _, otab = load_case(case)
    S = state_from(g, "init_", ospec)
    n_calls = 0
    for step in range(int(g["meta_n_steps"])):
        p = f"s{step}_"
        env._step = step
        env.actions = torch.from_numpy(g[p + "in_actions"]).to(DEV).contiguous()
        env.torques = torch.from_numpy(g[p + "in_torques"]).to(DEV).contiguous()
        ts.inject_next = {k: torch.from_numpy(g[p + "out_" + k]).to(DEV).contiguous() for k in INJECT}  # the reference's own reset draws
        oinj = {k: torch.from_numpy(g[p + "out_" + k]) for k in INJECT}
        if p + "in_push_robot_counter" in g:
            ts.set_push_state(torch.from_numpy(g[p + "in_push_robot_counter"]), ts.push_interval_s)
            S["push_robot_counter"] = torch.from_numpy(g[p + "in_push_robot_counter"].copy())
            oinj.update(push_interval_s=torch.from_numpy(g[p + "out_push_interval_s"]), push_vel=torch.from_numpy(g[p + "out_push_robot_vel_buf"]))
            ts.inject_next.update({k: v.to(DEV).contiguous() for k, v in oinj.items() if k.startswith("push")})
        prev_ep = ts.episode_length_buf.cpu().numpy().copy()
        env._post_physics_step()
        torch.cuda.synchronize()
        tag = f"integration {case} step {step}"
        assert env.common_step_counter == step + 1
        for name, v in env.obs_buf_dict.items():
            assert_close(v.cpu(), g[p + "obs_" + name], what=f"{tag} obs {name}")
        G = {k[len(p) + 4:]: g[k] for k in g if k.startswith(p + "out_")}
        compare_state(ts, G, spec, tag)
        assert_close(sim.dof_state.cpu(), g[p + "out_dof_state"], what=f"{tag} dof_state")
        assert_close(sim.all_root_states.cpu(), g[p + "out_root_state"], what=f"{tag} root_state")
        # env_ids = reset_buf.nonzero().flatten() of the REFERENCE went to both setters, root first (legged_robot_base.py:316-319)
        want_ids = np.flatnonzero(g[p + "out_reset_buf"]).astype(np.int32)
        refresh = g[p + "out_reset_buf"] != 0
        if p + "out_push_robot_counter" in g:
            refresh = refresh | (g[p + "out_push_robot_counter"] == 0)
            assert refresh.sum() > len(want_ids)
        want_refresh = np.flatnonzero(refresh).astype(np.int32)
        assert len(want_ids) > 0
        assert [c[0] for c in sim.calls[n_calls:]] == ["root", "dof"], tag
        for kind, ids, n in sim.calls[n_calls:]:
            assert n == len(want_refresh) and np.array_equal(ids, want_refresh), (tag, kind)
        n_calls = len(sim.calls)
        assert torch.equal(ts.need_to_refresh_envs.cpu(), torch.from_numpy(refresh))
        # extras: the oracle (bit-exact against this fixture, tests/test_oracle_env_golden.py) gives extras["episode"]
        inp = {k: torch.from_numpy(g[p + "in_" + k].copy()) for k in ("body_state", "root_state", "dof_state", "contact_forces", "actions", "torques")}
        _, olog, oex = OE.step(ospec, otab, S, inp, oinj)
        assert np.array_equal(oex["refresh_env_ids"].numpy(), want_refresh)
        ep = env.extras["episode"]
        assert sorted(ep) == sorted(oex["episode"])
        assert np.array_equal(ep["end_epis_length"].cpu().numpy(), prev_ep[want_ids] + 1)
        assert np.array_equal(ep["end_epis_length"].cpu().numpy(), oex["episode"]["end_epis_length"].numpy())
        for k, v in oex["episode"].items():
            if k != "end_epis_length":
                assert_close(ep[k].cpu(), v, what=f"{tag} extras episode {k}")
        assert torch.equal(env.extras["time_outs"].cpu(), torch.from_numpy(g[p + "out_time_out_buf"]))
        assert_close(env.extras["ref_body_pos_extend"].cpu(), g[p + "extras_ref_body_pos_extend"], what=f"{tag} ref pos")
        assert env.extras["to_log"] is env.log_dict
        for k in g:
            if k.startswith(p + "log_") and k[len(p) + 4:] in env.log_dict:
                assert_close(env.log_dict[k[len(p) + 4:]].cpu(), np.float32(g[k]), rtol=1e-5, atol=1e-6, what=f"{tag} log {k}")


@pytest.mark.parametrize("n", [1, 15, 16, 4095, 4097, 65536, 200003])
def test_compact_ids_matches_nonzero(n):
    """pbhc_compact_ids == flags.nonzero().flatten() (ascending), count on the device; all-zero and all-set masks"""
    L = _lib.lib()
    gen = torch.Generator().manual_seed(n)
    scratch = torch.zeros(max(int(L.pbhc_compact_scratch_ints(n)), 1), device=DEV, dtype=torch.int32)
    for density in (0.0, 0.03, 0.5, 1.0):
        flags = (torch.rand(n, generator=gen) < density).to(torch.int64) * torch.randint(1, 1 << 40, (n,), generator=gen)
        flags = flags.to(DEV)
        ids = torch.full((n,), -7, device=DEV, dtype=torch.int32)
        count = torch.full((1,), -1, device=DEV, dtype=torch.int32)
        _lib.check(L.pbhc_compact_ids(_lib.ptr(flags), n, _lib.ptr(ids), _lib.ptr(count), _lib.ptr(scratch), _lib.stream_ptr(torch.device(DEV))))
        want = flags.nonzero().flatten().to(torch.int32)
        c = int(count.item())
        assert c == want.numel(), (n, density)
        assert torch.equal(ids[:c], want), (n, density)
        assert bool((ids[c:] == -7).all()), "entries past count must stay untouched"
    assert L.pbhc_compact_ids(None, n, _lib.ptr(ids), _lib.ptr(count), _lib.ptr(scratch), None) == -1
    assert L.pbhc_compact_ids(_lib.ptr(flags), 0, _lib.ptr(ids), _lib.ptr(count), _lib.ptr(scratch), None) == -3


def test_resample_motion_resets_every_env(tmp_path):
    """_update_tasks_callback -> resample_motion (motion_tracking.py:319-328, 385-389): when the step counter hits the
    interval the library is re-drawn and every env is reset by that step - new start times inside the new clips, episode
    length 0, all env ids to the simulator; a forced reset is no time-out and earns no termination reward; the step after
    it is an ordinary one."""
    env, g, spec, N = make_env("mt23_main", tmp_path, resample=True, interval=2)
    sim, ts = env.simulator, env._b200
    loads = []
    orig = env._motion_lib.load_motions
    env._motion_lib.load_motions = lambda **kw: (loads.append(kw), orig(**kw))[1]
    for step in range(3):
        p = f"s{step}_"
        env._step = step
        env.actions = torch.from_numpy(g[p + "in_actions"]).to(DEV).contiguous()
        env.torques = torch.from_numpy(g[p + "in_torques"]).to(DEV).contiguous()
        prev_ep = ts.episode_length_buf.clone()
        st0, ln0 = ts.motion_start_times.clone(), ts.motion_len.clone()
        n_before = len(sim.calls)
        env._post_physics_step()
        torch.cuda.synchronize()
        natural = torch.from_numpy(g[p + "out_reset_buf"] != 0)
        if step == 1:  # common_step_counter == 2
            assert loads == [dict(random_sample=True)]
            assert bool((ts.reset_buf == 1).all()) and bool((ts.episode_length_buf == 0).all())
            assert [(k, n) for k, _, n in sim.calls[n_before:]] == [("root", N), ("dof", N)]
            assert np.array_equal(sim.calls[-1][1], np.arange(N, dtype=np.int32))
            assert torch.equal(env.extras["episode"]["end_epis_length"], prev_ep + 1)
            ep1 = prev_ep + 1  # forced resets are not time-outs: time_out_buf is what the termination check alone says
            want_to = (ep1 > spec.max_episode_length) | ((ep1.float() * torch.tensor(spec.dt, dtype=torch.float32) + st0 > ln0) & bool(spec.term_motion_end))
            assert torch.equal(ts.time_out_buf, want_to) and int(want_to.sum()) < N
            st, ln = ts.motion_start_times.cpu(), ts.motion_len.cpu()
            assert bool((st >= 0).all()) and bool((st <= ln).all())
            assert torch.equal(ln, env._motion_lib.get_motion_length(env.motion_ids).cpu())
            # the root / dof state the simulator receives is the reference motion at the new start times (motion_tracking.py:445-543)
            ref = env._motion_lib.get_motion_state(env.motion_ids, (ts.motion_start_times + spec.dt).contiguous(), offset=env.env_origins)
            assert_close(sim.all_root_states[:, :3].cpu(), ref["root_pos"].cpu(), what="forced reset root pos")
            assert_close(sim.dof_state.view(N, -1, 2)[..., 0].cpu(), ref["dof_pos"].cpu(), what="forced reset dof pos")
            forced_only = ~natural
            term_col = ts._sum_names.index("termination")
            assert bool((ts.last_episode_sums[:, term_col].cpu()[forced_only] == 0).all())
        else:
            assert not ts._force_pending and int(ts._force_reset.sum()) == 0
            if step == 0:
                assert torch.equal(ts.reset_buf.cpu() != 0, natural)
            else:
                assert int(ts.reset_buf.sum()) < N

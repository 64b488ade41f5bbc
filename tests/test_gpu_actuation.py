"""GPU parity of pbhc_b200.actuation.Actuation (csrc/actuate.cu through the C ABI) against the golden recorded from the
reference env's _pre_physics_step / _compute_torques (tests/golden/ref_actuation.npz) and against the torch-fp32 oracle.
Everything is bit-exact when the torch.rand_like values of the torque RFI are fed in; the in-kernel RNG is checked
statistically."""
import json

import numpy as np
import pytest
import torch

from tests._util import gold
from tests.test_oracle_golden import oracle_actuation_from_golden

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda:0")


def _product_from_golden(g, **override):
    from pbhc_b200.actuation import Actuation
    cfg = json.loads(str(g["meta_config_json"]))
    for k, v in override.items():
        node = cfg
        *path, leaf = k.split(".")
        for p in path:
            node = node[p]
        node[leaf] = v
    N = int(g["meta"][0])
    a = Actuation(cfg, N, DEV, seed=5)
    for k in ("kp_scale", "kd_scale", "rfi_lim_scale", "rao_scale"):
        getattr(a, "_" + k).copy_(torch.from_numpy(g[k]))
    a.action_delay_idx.copy_(torch.from_numpy(g["action_delay_idx"]))
    return a, cfg


def test_reference_golden_steps():
    g = gold("ref_actuation.npz")
    a, cfg = _product_from_golden(g)
    N, D, steps, Q = (int(x) for x in g["meta"])
    assert np.array_equal(a.p_gains.cpu().numpy(), g["p_gains"]) and np.array_equal(a.d_gains.cpu().numpy(), g["d_gains"])
    assert np.array_equal(a.torque_limits.cpu().numpy(), g["torque_limits"]) and np.array_equal(a.default_dof_pos.cpu().numpy(), g["default_dof_pos"])
    assert a.action_queue.shape == (N, Q, D)
    for s in range(steps):
        p = f"s{s}_"
        after = a._pre_physics_step(torch.from_numpy(g[p + "actions_in"]).to(DEV))
        assert np.array_equal(a.actions.cpu().numpy(), g[p + "actions"])
        assert np.array_equal(a.action_queue.cpu().numpy(), g[p + "action_queue"])
        assert np.array_equal(after.cpu().numpy(), g[p + "actions_after_delay"])
        assert float(a.log_dict["action_clip_frac"]) == float(g[p + "action_clip_frac"])
        for k in range(4):
            dof = torch.from_numpy(g[p + f"k{k}_dof_state"]).to(DEV)
            t = a._compute_torques(a.actions_after_delay, dof.reshape(-1, 2), rfi_uniform=torch.from_numpy(g[p + f"k{k}_rfi_u"]).to(DEV))
            assert np.array_equal(t.cpu().numpy(), g[p + f"k{k}_torques"]), (s, k)


@pytest.mark.parametrize("control_type,delay,clip", [("P", True, True), ("V", False, True), ("T", True, False), ("P", False, False)])
def test_variants_large_vs_oracle(control_type, delay, clip):
    """control types, delay on/off, clipping on/off at 5000 envs against the oracle (bit-exact with injected draws)"""
    from oracle import pbhc_oracle_actuation as OA
    g = gold("ref_actuation.npz")
    cfg = json.loads(str(g["meta_config_json"]))
    cfg["robot"]["control"]["control_type"], cfg["robot"]["control"]["clip_torques"] = control_type, clip
    cfg["domain_rand"]["randomize_ctrl_delay"] = delay
    from pbhc_b200.actuation import Actuation
    N, D = 5000, 23
    a = Actuation(cfg, N, DEV)
    Q = 3 if delay else 0
    ctl, dr = cfg["robot"]["control"], cfg["domain_rand"]
    o = OA.Actuation(N, D, Q, ctl["action_clip_value"], control_type, ctl["action_scale"], g["p_gains"], g["d_gains"], g["torque_limits"],
                     g["default_dof_pos"][:1].repeat(N, 0), dr["randomize_torque_rfi"], dr["rfi_lim"], dr["use_rao"], clip, 1.0 / 200)
    gen = torch.Generator().manual_seed(17)
    o.kp_scale, o.kd_scale = 0.9 + 0.2 * torch.rand(N, D, generator=gen), 0.9 + 0.2 * torch.rand(N, D, generator=gen)
    o.rfi_lim_scale, o.rao_scale = 0.5 + torch.rand(N, D, generator=gen), (torch.rand(N, D, generator=gen) - 0.5) * 0.1
    o.last_dof_vel = torch.randn(N, D, generator=gen)
    o.action_delay_idx = torch.randint(0, 3, (N,), generator=gen) if delay else torch.zeros(N, dtype=torch.int64)
    for k in ("kp_scale", "kd_scale", "rfi_lim_scale", "rao_scale"):
        getattr(a, "_" + k).copy_(getattr(o, k))
    a.last_dof_vel.copy_(o.last_dof_vel)
    a.action_delay_idx.copy_(o.action_delay_idx)
    for s in range(4):
        actions = torch.randn(N, D, generator=gen) * 1.2
        actions[0, 0] = float("nan")  # torch.clip propagates NaN
        o.pre_physics_step(actions)
        a._pre_physics_step(actions.to(DEV))
        assert np.array_equal(a.actions_after_delay.cpu().numpy(), o.actions_after_delay.numpy(), equal_nan=True)
        assert float(a.log_dict["action_clip_frac"]) == float(o.log_dict["action_clip_frac"])
        dof = torch.stack([torch.randn(N, D, generator=gen) * 0.4, torch.randn(N, D, generator=gen) * 3.0], dim=-1)
        u = torch.rand(N, D, generator=gen)
        want = o.compute_torques(o.actions_after_delay, dof[..., 0], dof[..., 1], u)
        got = a._compute_torques(a.actions_after_delay, dof.to(DEV), rfi_uniform=u.to(DEV))
        assert np.array_equal(got.cpu().numpy(), want.numpy(), equal_nan=True), s


def test_in_kernel_rfi_statistics():
    """without injected draws the RFI term is U(-1,1) * rfi_lim * scale * limit from the Philox stream: implied uniforms
    lie in [0,1), have the right mean / variance, differ between calls and between envs, and repeat for the same seed"""
    g = gold("ref_actuation.npz")
    from pbhc_b200.actuation import Actuation
    cfg = json.loads(str(g["meta_config_json"]))
    cfg["robot"]["control"]["clip_torques"] = False
    N, D = 20000, 23
    runs = []
    for seed in (5, 5, 6):
        a = Actuation(cfg, N, DEV, seed=seed)
        dof = torch.zeros(N, D, 2, device=DEV)
        acts = torch.zeros(N, D, device=DEV)
        base = a._compute_torques(acts, dof, rfi_uniform=torch.full((N, D), 0.5, device=DEV)).clone()  # u = 0.5 -> no RFI term
        t1 = a._compute_torques(acts, dof).clone()
        t2 = a._compute_torques(acts, dof).clone()
        scale = cfg["domain_rand"]["rfi_lim"] * a.torque_limits
        runs.append(((t1 - base) / scale, (t2 - base) / scale))
    v1, v2 = runs[0]
    assert float(v1.min()) >= -1.0 - 1e-5 and float(v1.max()) <= 1.0 + 1e-5
    assert abs(float(v1.mean())) < 0.01 and abs(float(v1.var()) - 1.0 / 3.0) < 0.01
    assert not torch.equal(v1, v2) and abs(float((v1 * v2).mean())) < 0.01  # successive calls: fresh draws
    assert abs(float((v1[:-1] * v1[1:]).mean())) < 0.01                       # neighbouring envs uncorrelated
    assert torch.equal(runs[1][0], v1) and not torch.equal(runs[2][0], v1)    # seed-reproducible


def test_rfi_stream_advances_under_graph_replay_and_follows_the_global_env_id():
    """the call index of the RFI Philox stream lives in device memory and is advanced by the launch itself: a replayed CUDA
    graph draws fresh noise every replay, identical to the same number of eager calls; a shard with env_offset = k draws
    what envs k.. of the unsharded run draw"""
    g = gold("ref_actuation.npz")
    from pbhc_b200.actuation import Actuation
    cfg = json.loads(str(g["meta_config_json"]))
    N, D = 512, 23
    dof = torch.zeros(N, D, 2, device=DEV)
    acts = torch.zeros(N, D, device=DEV)
    eager = Actuation(cfg, N, DEV, seed=9)
    want = [eager._compute_torques(acts, dof).clone() for _ in range(4)]
    rep = Actuation(cfg, N, DEV, seed=9)
    out = torch.empty(N, D, device=DEV)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        rep._compute_torques(acts, dof, out=out)  # call 0, eager (warm-up on the capture stream)
        first = out.clone()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=s):
            rep._compute_torques(acts, dof, out=out)
        got = [first]
        for _ in range(3):
            graph.replay()
            got.append(out.clone())
    torch.cuda.synchronize()
    for i in range(4):
        assert torch.equal(got[i], want[i]), f"call {i}"
    assert not torch.equal(got[1], got[2]) and not torch.equal(got[2], got[3])
    half = Actuation(cfg, N // 2, DEV, seed=9, env_offset=N // 2)
    assert torch.equal(half._compute_torques(acts[N // 2:], dof[N // 2:]), want[0][N // 2:])


def test_shares_randomisation_tensors_with_tracking_step(tmp_path):
    from pbhc_b200.actuation import Actuation
    from benchkit.fixtures import env_config
    from benchkit.workload import Workload
    wl = Workload(64, DEV, str(tmp_path), seed=1)
    a = Actuation(env_config(), 64, DEV, tracking_step=wl.ts)
    assert a._kp_scale.data_ptr() == wl.ts._kp_scale.data_ptr() and a.action_delay_idx.data_ptr() == wl.ts.action_delay_idx.data_ptr()
    after = a._pre_physics_step(torch.randn(64, 23, device=DEV))
    t = a._compute_torques(after, wl.sim.dof_state)
    assert t.shape == (64, 23) and bool(torch.isfinite(t).all())
    assert bool((t.abs() <= a.torque_limits + 1e-6).all())
    with pytest.raises(Exception):
        Actuation(env_config(), 64, "cpu")

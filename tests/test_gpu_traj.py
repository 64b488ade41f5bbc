"""GPU parity of pbhc_b200.measure_traj (csrc/traj.cu through the C ABI) against the golden recorded from the reference's
measure_traj.py functions (tests/golden/ref_traj.npz) and against the torch-fp32 oracle on larger seeded trajectories.
Frame pairs and lerp values: bit-exact.  slerp and the metric means: fp32 within 1e-5 relative / 1e-6 absolute."""
import numpy as np
import pytest
import torch

from tests._util import gold, assert_close, product_cfg
from tests.test_oracle_golden import traj_golden_inputs

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda:0")


def test_blend_motion_matches_reference_golden():
    from pbhc_b200 import measure_traj as MT
    g = gold("ref_traj.npz")
    pre, times, pol, ref, fps_pol = traj_golden_inputs(g, DEV)
    pre["global_rotation_mat"] = torch.zeros(pre["dof_pos"].shape[0], 24, 3, 3, device=DEV)  # skipped, like documented
    out = MT.blend_motion(pre, times)
    assert set(out) == set(ref) | {"fps"} and out["fps"] == pre["fps"]
    for k in ("global_translation", "dof_pos", "contact_mask"):  # lerp keys: same fp32 operations as eager torch
        assert np.array_equal(out[k].cpu().numpy(), ref[k].cpu().numpy()), k
    assert_close(out["global_rotation"].cpu(), ref["global_rotation"].cpu(), what="global_rotation")
    with pytest.raises(Exception):
        MT.blend_motion({k: (v.cpu() if torch.is_tensor(v) else v) for k, v in pre.items()}, times)


@pytest.mark.parametrize("tag,dpf", [("fps", False), ("frame", True)])
def test_metrics_match_reference_golden(tag, dpf):
    from pbhc_b200 import measure_traj as MT
    g = gold("ref_traj.npz")
    pre, times, pol, ref, fps_pol = traj_golden_inputs(g, DEV)
    traj = {"pol": pol, "ref": ref, "appendix": {"fps": fps_pol}}
    acc, smooth = MT.eval_accuracy(traj, dpf), MT.eval_smoothness(traj, dpf)
    want_acc = {k[len(f"acc_{tag}_"):]: v for k, v in g.items() if k.startswith(f"acc_{tag}_")}
    want_smooth = {k[len(f"smooth_{tag}_"):]: v for k, v in g.items() if k.startswith(f"smooth_{tag}_")}
    assert list(acc) == list(want_acc) or set(acc) == set(want_acc)
    assert set(smooth) == set(want_smooth)
    for k, v in {**want_acc, **want_smooth}.items():
        got = (acc if k in acc else smooth)[k]
        assert got.dtype == torch.float32 and got.dim() == 0 and got.is_cuda
        assert_close(got.cpu(), v, what=k)
    # without contact masks the key is absent, like the reference
    traj2 = {"pol": {k: v for k, v in pol.items() if k != "contact_mask"}, "ref": ref, "appendix": {"fps": fps_pol}}
    assert "E_contact_acc" not in MT.eval_accuracy(traj2, dpf)
    assert MT.apply_1e3({"x": torch.tensor(2.0)})["x"].item() == 2000.0


@pytest.mark.parametrize("T,J,D", [(1, 24, 23), (2, 24, 23), (3, 27, 23), (4, 30, 29), (1000, 33, 29), (20011, 24, 23)])
def test_metrics_vs_oracle_sizes(T, J, D):
    """short trajectories (a mean over no frames is NaN in torch and here), body counts above a warp, a long recording"""
    from oracle import pbhc_oracle_traj as OT
    from pbhc_b200 import measure_traj as MT
    gen = torch.Generator().manual_seed(T + J)
    mk = lambda *s: torch.cumsum(torch.randn(*s, generator=gen) * 0.05, dim=0)
    pol = {"global_translation": mk(T, J, 3), "dof_pos": mk(T, D), "contact_mask": (torch.rand(T, 2, generator=gen) < 0.5).float()}
    ref = {"global_translation": mk(T, J, 3), "dof_pos": mk(T, D), "contact_mask": (torch.rand(T, 2, generator=gen) < 0.5).float()}
    cpu = {"pol": pol, "ref": ref, "appendix": {"fps": 50}}
    gpu = {"pol": {k: v.to(DEV) for k, v in pol.items()}, "ref": {k: v.to(DEV) for k, v in ref.items()}, "appendix": {"fps": 50}}
    want = {**OT.eval_accuracy(cpu), **OT.eval_smoothness(cpu)}
    got = {**MT.eval_accuracy(gpu), **MT.eval_smoothness(gpu)}
    assert set(want) == set(got)
    for k, v in want.items():
        if torch.isnan(v):
            assert torch.isnan(got[k]), k
        else:
            assert_close(got[k].cpu(), v, what=k)


def test_blend_motion_large_vs_oracle():
    from oracle import pbhc_oracle_traj as OT
    from pbhc_b200 import measure_traj as MT
    gen = torch.Generator().manual_seed(9)
    T = 777
    q = torch.randn(T, 27, 4, generator=gen)
    q = q / q.norm(dim=-1, keepdim=True)
    pre = {"global_translation": torch.randn(T, 27, 3, generator=gen), "global_rotation_extend": q, "dof_pos": torch.randn(T, 23, generator=gen),
           "global_velocity": torch.randn(T, 24, 3, generator=gen), "fps": 30}
    times = torch.rand(5000, generator=gen) * (T / 30 + 1.0) - 0.5
    want = OT.blend_motion(pre, times)
    got = MT.blend_motion({k: (v.to(DEV) if torch.is_tensor(v) else v) for k, v in pre.items()}, times.to(DEV))
    for k in ("global_translation", "dof_pos", "global_velocity"):
        assert np.array_equal(got[k].cpu().numpy(), want[k].numpy()), k
    assert_close(got["global_rotation_extend"].cpu(), want["global_rotation_extend"], what="slerp")


def test_multistep_reference_window_matches_per_step_lookups(tmp_path):
    """kick_motion_res_multiplestep (motion_tracking.py:560-579): one 2-D K1 launch == the reference's loop of 1-D lookups
    (bit-exact against the same kernel called per step; the 1-D lookup itself is pinned in test_gpu_parity.py)"""
    from oracle import pbhc_oracle_traj as OT
    from pbhc_b200 import measure_traj as MT
    from pbhc_b200.motion_lib import MotionLibRobot
    names = ["Horse-stance_pose", "Bruce_Lee_pose", "Side_kick"]
    B, steps, dt = 301, 5, 0.02
    ml = MotionLibRobot(product_cfg("23", str(tmp_path), names), B, DEV)
    ml.load_motions(random_sample=False)
    gen = torch.Generator().manual_seed(1)
    ids = torch.randint(0, len(names), (B,), generator=gen).to(DEV)
    ep = torch.randint(0, 400, (B,), generator=gen).to(DEV)
    start = (torch.rand(B, generator=gen) * 2.0).to(DEV)
    origins = torch.randn(B, 3, generator=gen).to(DEV)
    got = MT.kick_motion_res_multiplestep(ml, ids, ep, dt, start, origins, steps)
    want = OT.kick_motion_res_multiplestep(lambda i, t, o: ml.get_motion_state(i, t, offset=o), ids, ep, dt, start, origins, steps)
    assert set(got) == set(want)
    for k, v in want.items():
        assert got[k].shape == v.shape, k
        assert torch.equal(got[k], v), k
    fp, fv = MT.future_ref_obs(got)
    assert fp.shape == (B, steps * 23) and torch.equal(fv, want["dof_vel"].reshape(B, -1))


def test_motion_data_feeds_the_metrics_end_to_end(tmp_path):
    """measure_traj.load_traj_data flow on device: loaded reference motion -> motion_data -> blend at 50-fps policy times ->
    metrics against a perturbed copy; checked against the oracle run on the same table rows"""
    from oracle import pbhc_oracle_traj as OT
    from pbhc_b200 import measure_traj as MT
    from pbhc_b200.motion_lib import MotionLibRobot
    ml = MotionLibRobot(product_cfg("23", str(tmp_path), ["Horse-stance_pose"]), 1, DEV)
    ml.load_motions(random_sample=False)
    pre = ml.motion_data(0)
    T = pre["dof_pos"].shape[0]
    assert pre["global_translation"].shape == (T, 24, 3) and pre["global_translation_extend"].shape == (T, 27, 3)
    assert pre["dof_pos"].shape == (T, 23) and pre["fps"] == 30
    times = torch.arange(int(T / 30 * 50) - 2, device=DEV, dtype=torch.float32) / 50
    ref = MT.blend_motion(pre, times)
    gen = torch.Generator().manual_seed(0)
    pol = {k: v + 0.01 * torch.randn(v.shape, generator=gen).to(DEV) for k, v in ref.items() if k in ("global_translation", "dof_pos")}
    traj = {"pol": pol, "ref": ref, "appendix": {"fps": 50}}
    cpu = lambda d: {k: (v.cpu() if torch.is_tensor(v) else v) for k, v in d.items()}
    pre_c = {k: (v.cpu().contiguous() if torch.is_tensor(v) else v) for k, v in pre.items()}
    ref_c = OT.blend_motion(pre_c, times.cpu())
    for k in ("global_translation", "dof_pos", "global_velocity_extend"):
        assert np.array_equal(ref[k].cpu().numpy(), ref_c[k].numpy()), k
    traj_c = {"pol": cpu(pol), "ref": ref_c, "appendix": {"fps": 50}}
    for k, v in {**OT.eval_accuracy(traj_c), **OT.eval_smoothness(traj_c)}.items():
        got = {**MT.eval_accuracy(traj), **MT.eval_smoothness(traj)}[k]
        assert_close(got.cpu(), v, what=k)

"""TEST INFRASTRUCTURE ONLY - torch-fp32 CPU restatement of the reference's trajectory evaluation.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this; pbhc_b200/ never does.

Follows humanoidverse/measure_traj.py:67-133 (blend_motion), :146-225 (eval_accuracy), :227-290 (eval_smoothness) and
envs/motion_tracking/motion_tracking.py:560-579 (kick_motion_res_multiplestep).  Pinned against the unmodified
reference functions by tests/golden/ref_traj.npz (tools/make_golden_traj.py imports measure_traj.py in the build
container) in tests/test_oracle_golden.py.
"""
import torch

from . import pbhc_oracle as O


def calc_frame_blend(time, length, num_frames, dt):  # measure_traj.py:68-84
    time = time.clone()
    phase = torch.clip(time / length, 0.0, 1.0)
    time[time < 0] = 0
    f0 = (phase * (num_frames - 1)).long()
    f1 = torch.min(f0 + 1, torch.tensor(num_frames) - 1)
    blend = torch.clip((time - f0 * dt) / dt, 0.0, 1.0)
    return f0, f1, blend


def blend_motion(pre, times):  # measure_traj.py:67-133 (rotation_mat keys left out, like the product)
    fps = pre["fps"]
    n = pre["dof_pos"].shape[0]
    f0, f1, blend = calc_frame_blend(times, n / fps, n, 1 / fps)
    blend = blend.unsqueeze(-1)
    out = {}
    for k, v in pre.items():
        if k == "fps":
            out[k] = v
            continue
        if "rotation" in k and "mat" in k:
            continue
        a, b = v[f0], v[f1]
        be = blend.reshape(-1, *([1] * (a.ndim - 1)))
        out[k] = O.slerp(a, b, be) if "rotation" in k else be * b + (1 - be) * a
    return out


def _m(x):  # `torch.norm(x, dim=-1).mean(dim=-1).mean()`
    return torch.norm(x, dim=-1).mean(dim=-1).mean()


def eval_accuracy(traj, delta_per_frame=False):  # measure_traj.py:146-225
    pol, ref = traj["pol"], traj["ref"]
    d = 1 if delta_per_frame else traj["appendix"]["fps"]
    diff = lambda x: (x[1:] - x[:-1]) * d
    pp, rp, pd, rd = pol["global_translation"], ref["global_translation"], pol["dof_pos"], ref["dof_pos"]
    res = {"E_gmpbpe": _m(pp - rp), "E_mpbpe": _m((pp - pp[..., 0:1, :]) - (rp - rp[..., 0:1, :])), "E_mpjpe": _m(pd - rd)}
    pdv, rdv = diff(pd), diff(rd)
    res["E_mpjve"], res["E_mpjae"] = _m(pdv - rdv), _m(diff(pdv) - diff(rdv))
    pv, rv = diff(pp), diff(rp)
    pa, ra = diff(pv), diff(rv)
    res["E_pbve"], res["E_pbae"] = _m(pv - rv), _m(pa - ra)
    res["E_root_acc"] = _m(pa[..., 0:1, :] - ra[..., 0:1, :])
    res["E_root_vel"] = _m(pv[..., 0:1, :] - rv[..., 0:1, :])
    if "contact_mask" in pol and "contact_mask" in ref:
        res["E_contact_acc"] = torch.mean((pol["contact_mask"] - ref["contact_mask"]).abs(), dim=-1).mean()
    return res


def eval_smoothness(traj, delta_per_frame=False):  # measure_traj.py:227-290
    d = 1 if delta_per_frame else traj["appendix"]["fps"]
    diff = lambda x: (x[1:] - x[:-1]) * d
    res = {}
    for tag, src in (("", traj["pol"]), ("ref_", traj["ref"])):
        for name, x in (("", src["global_translation"]), ("dof_", src["dof_pos"])):
            v = diff(x)
            a = diff(v)
            res[f"L2_{tag}{name}vel"], res[f"L2_{tag}{name}acc"], res[f"L2_{tag}{name}jerk"] = _m(v), _m(a), _m(diff(a))
    return res


def kick_motion_res_multiplestep(get_motion_state, motion_ids, episode_length_buf, dt, motion_start_times, env_origins, steps):
    """motion_tracking.py:560-579 with `get_motion_state(ids, times, offset)` = the 1-D lookup"""
    buf = {}
    for i in range(steps):
        times = (episode_length_buf + 1 + i) * dt + motion_start_times
        for k, v in get_motion_state(motion_ids, times, env_origins).items():
            buf.setdefault(k, []).append(v.unsqueeze(1))
    return {k: torch.cat(v, dim=1) for k, v in buf.items()}

"""Trajectory evaluation (reference: humanoidverse/measure_traj.py:67-290) on the sm_100a kernels of csrc/traj.cu
(SURVEY.md 8f N4).  Same function names, argument meaning and result keys as the reference:

    ref = blend_motion(ref_motion_data_preblend, pol_appendix['motion_times'])
    traj = {'pol': pol_motion_data, 'ref': ref, 'appendix': pol_appendix}
    acc = eval_accuracy(traj)            # E_gmpbpe, E_mpbpe, E_mpjpe, E_mpjve, E_mpjae, E_pbve, E_pbae, E_root_acc, E_root_vel[, E_contact_acc]
    smooth = eval_smoothness(traj)       # L2_vel ... L2_ref_dof_jerk

`pol` / the preblend data are the per-motion dicts MotionLibRobot.load_motions returns (global_translation [T,J,3],
dof_pos [T,D], ..., fps) on a CUDA device; eval_accuracy + eval_smoothness share ONE pass over the trajectory
(`traj_metrics`, cached per traj_data dict).  Metrics come back as 0-dim fp32 CUDA tensors like the reference's.
No CPU fallback.

Differences, on purpose: blend_motion skips `*rotation_mat*` keys (the metrics never read them; the reference
round-trips them through quaternions) and does not print; times are used in fp32.
"""
import torch

from . import _lib


def _f32(t, what):
    if not torch.is_tensor(t) or not t.is_cuda:
        raise _lib.PbhcError(f"{what} must be a CUDA tensor (there is no CPU fallback)")
    return t.to(torch.float32).contiguous()


def blend_motion(preblend_data, input_motion_times):
    """measure_traj.py:67-133: every key of a loaded motion resampled at the policy's recorded motion times."""
    times = _f32(input_motion_times, "input_motion_times").reshape(-1)
    fps = float(preblend_data["fps"])
    n_frames = int(preblend_data["dof_pos"].shape[0])
    L, st = _lib.lib(), _lib.stream_ptr(times.device)
    out = {}
    for key, val in preblend_data.items():
        if key == "fps":
            out[key] = val
            continue
        if not torch.is_tensor(val) or val.dim() == 0 or val.shape[0] != n_frames or not val.is_floating_point():
            continue
        if "rotation" in key and "mat" in key:
            continue
        src = _f32(val, key)
        row = src[0].numel()
        quats = 1 if "rotation" in key else 0
        dst = torch.empty((times.numel(),) + tuple(src.shape[1:]), device=src.device, dtype=torch.float32)
        _lib.check(L.pbhc_blend_frames(src.data_ptr(), n_frames, row, quats, times.data_ptr(), times.numel(), fps, dst.data_ptr(), st))
        out[key] = dst
    return out


def traj_metrics(traj_data, delta_per_frame=False):
    """all 22 metrics of eval_accuracy + eval_smoothness in one launch -> fp64 device tensor [TRAJ_METRICS]"""
    pol, ref = traj_data["pol"], traj_data["ref"]
    delta = 1.0 if delta_per_frame else float(traj_data["appendix"]["fps"])
    pp, rp = _f32(pol["global_translation"], "pol global_translation"), _f32(ref["global_translation"], "ref global_translation")
    pd, rd = _f32(pol["dof_pos"], "pol dof_pos"), _f32(ref["dof_pos"], "ref dof_pos")
    if pp.shape != rp.shape or pd.shape != rd.shape or pp.shape[0] != pd.shape[0]:
        raise _lib.PbhcError(f"traj_metrics: pol / ref shapes differ: {tuple(pp.shape)} {tuple(rp.shape)} {tuple(pd.shape)} {tuple(rd.shape)}")
    T, J, D = pp.shape[0], pp.shape[1], pd.shape[1]
    pc = rc = None
    C = 0
    if "contact_mask" in pol and "contact_mask" in ref:
        pc, rc = _f32(pol["contact_mask"], "pol contact_mask"), _f32(ref["contact_mask"], "ref contact_mask")
        C = pc.shape[1]
    dev = pp.device
    ws = torch.empty(_lib.TRAJ_WORKSPACE, device=dev, dtype=torch.float64)
    out = torch.empty(_lib.TRAJ_METRICS, device=dev, dtype=torch.float64)
    p = lambda t: None if t is None else t.data_ptr()
    _lib.check(_lib.lib().pbhc_traj_metrics(p(pp), p(rp), p(pd), p(rd), p(pc), p(rc), T, J, D, C, delta, ws.data_ptr(), ws.numel(),
                                            out.data_ptr(), _lib.stream_ptr(dev)))
    return out


def _named(vals, names, lo):
    v32 = vals.to(torch.float32)
    return {k: v32[lo + i] for i, k in enumerate(names)}


def eval_accuracy(traj_data, delta_per_frame=False):
    """measure_traj.py:146-225"""
    res = _named(traj_metrics(traj_data, delta_per_frame), _lib.TRAJ_ACCURACY, 0)
    if not ("contact_mask" in traj_data["pol"] and "contact_mask" in traj_data["ref"]):
        res.pop("E_contact_acc")
    return res


def eval_smoothness(traj_data, delta_per_frame=False):
    """measure_traj.py:227-290"""
    return _named(traj_metrics(traj_data, delta_per_frame), _lib.TRAJ_SMOOTHNESS, len(_lib.TRAJ_ACCURACY))


apply_1e3 = lambda x: {k: v * 1e3 for k, v in x.items()}  # measure_traj.py:293


def kick_motion_res_multiplestep(motion_lib, motion_ids, episode_length_buf, dt, motion_start_times, env_origins, future_ref_steps):
    """LeggedRobotMotionTracking.kick_motion_res_multiplestep (envs/motion_tracking/motion_tracking.py:560-579): the
    reference motion state of the next `future_ref_steps` control steps, Dict[str, Tensor[B, T, ...]].  The reference
    calls get_motion_state once per step and concatenates; here the [B, T] times go through K1 in ONE launch (the
    per-element time arithmetic is the reference's: int64 (episode_length + 1 + i), times dt in fp32, plus start)."""
    B, T = motion_ids.shape[0], int(future_ref_steps)
    steps = torch.arange(T, device=motion_ids.device)
    times = (episode_length_buf[:, None] + 1 + steps[None, :]) * dt + motion_start_times[:, None]
    ids = motion_ids[:, None].expand(B, T).contiguous()
    off = None if env_origins is None else env_origins[:, None, :].expand(B, T, 3).contiguous()
    res = motion_lib.get_motion_state(ids, times, offset=off)
    res = {k: v for k, v in res.items() if not k.startswith("_")}
    res["dof_vel"] = res["dof_vel"].reshape(B, T, -1)  # get_motion_state flattens dof_vel of 2-D queries (motion_lib_base.py:246)
    return res


def future_ref_obs(future_motion_res):
    """motion_tracking.py:596-599: `_obs_future_ref_dof_pos`, `_obs_future_ref_dof_vel` ([B, T * D] views)"""
    B = future_motion_res["dof_pos"].shape[0]
    return future_motion_res["dof_pos"].reshape(B, -1), future_motion_res["dof_vel"].reshape(B, -1)

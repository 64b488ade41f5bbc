#!/usr/bin/env python
"""Edge-case fixtures for slerp / quat_to_angle_axis (isaac_utils/rotations.py:185-232), generated with the UNMODIFIED
reference in the build container:  python tools/make_golden_edge.py

A synthetic clip per robot (23 / 29 DoF) whose adjacent frames contain, by construction,
  * identical frames (cos_half_theta == 1: the `abs(cos) >= 1` branch, :228),
  * joint increments of 1.8 .. 2.2 mrad per frame, i.e. sin_half_theta straddling 1e-3 (the 0.5 a + 0.5 b branch, :226),
  * a root yaw sweep through -pi/2, where matrix_to_quaternion switches its best-conditioned candidate and the frame
    quaternion changes sign (the `cos < 0 -> q1 = -q1` branch, :213-216),
  * ordinary large rotations.
Outputs (tests/golden/): motion_edge{23,29}.npz (the clip: an input fixture like the other motion_*.npz),
ref_table_edge23.npz, ref_lookup_edge{23,29}.npz (get_motion_state at every frame pair x several blends) and env-step
fixtures ref_env_mt23_edge.npz / ref_env_gt29_edge.npz whose envs sit on those frame pairs; in the general_tracking fixture
the first envs hold body rotations equal to the reference's up to 0 .. 40 urad, so quat_to_angle_axis sees sin(theta / 2)
on both sides of its 1e-5 threshold (:193-200).
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_loader as R  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
T, FPS = 48, 30
INCR = [1.8e-3, 1.98e-3, 2.0e-3, 2.02e-3, 2.2e-3, 1.0e-4, 0.0, 5.0e-3]  # joint increments per frame [rad]


def edge_clip(n_pose, seed):
    """pose_aa [T, n_pose, 3] (root, joints, 3 zero extended entries), root translation [T, 3]."""
    rng = np.random.RandomState(seed)
    nj = n_pose - 3
    axis = R_axes(nj)
    base = (rng.rand(nj) - 0.5) * 0.6
    base[0] = 0.0
    ang = np.tile(base, (T, 1))                      # hinge angle of joint j at frame t (entry 0 = root, handled below)
    # frames 0..3 identical; frames 4..20: every joint drifts, the root sweeps its yaw through -pi/2
    for t in range(4, 21):
        ang[t, 1:] = base[1:] + 0.04 * (t - 3) * np.sign(base[1:] + 1e-9)
    ang[21:] = ang[20]
    # frames 24..40: joints 3 and 10 (and the last arm joint) advance by INCR, everything else is static
    for k, j in enumerate((3, 10, nj - 1)):
        a = ang[23, j]
        for t in range(24, 41):
            a = a + INCR[(t - 24 + k) % len(INCR)]
            ang[t, j] = a
        ang[41:, j] = a
    pose = np.zeros((T, n_pose, 3), np.float32)
    for j in range(1, nj):
        pose[:, j] = (ang[:, j:j + 1] * axis[j][None]).astype(np.float32)
    yaw = np.full(T, -1.2)
    yaw[4:21] = -1.2 - 0.05 * np.arange(1, 18)       # -1.25 .. -2.05: crosses -pi/2 between frames 10 and 11
    yaw[21:] = yaw[20]
    tilt = 0.05
    for t in range(T):
        pose[t, 0] = axis_angle_of(yaw[t], tilt)
    trans = np.stack([0.02 * np.arange(T), 0.01 * np.sin(np.arange(T) * 0.3), 0.78 + 0.01 * np.cos(np.arange(T) * 0.2)], 1).astype(np.float32)
    return pose, trans


def R_axes(nj):
    ax = np.zeros((nj, 3))
    for j in range(nj):
        ax[j, j % 3] = 1.0
    return ax


def axis_angle_of(yaw, tilt):
    """axis-angle of Rz(yaw) * Rx(tilt)"""
    from scipy.spatial.transform import Rotation as Rot
    return (Rot.from_euler("z", yaw) * Rot.from_euler("x", tilt)).as_rotvec().astype(np.float32)


def _np(x):
    return x.detach().cpu().numpy().copy() if torch.is_tensor(x) else np.asarray(x)


def write_clip(tag, n_pose, seed, tmp):
    import joblib
    pose, trans = edge_clip(n_pose, seed)
    rngc = np.random.RandomState(seed + 1)
    entry = dict(root_trans_offset=trans, pose_aa=pose, dof=pose[:, 1:n_pose - 3].sum(-1), root_rot=np.zeros((T, 4), np.float32), fps=FPS)
    extra = {}
    if tag == "23":
        entry["contact_mask"] = (rngc.rand(T, 2) < 0.5).astype(np.float64)
        extra["contact_mask"] = entry["contact_mask"]
    np.savez_compressed(os.path.join(GOLD, f"motion_edge{tag}.npz"), root_trans_offset=trans, pose_aa=pose, dof=entry["dof"],
                        root_rot=entry["root_rot"], fps=np.int64(FPS), **extra)
    path = os.path.join(tmp, f"edge{tag}.pkl")
    joblib.dump({f"edge{tag}": entry}, path)
    return path


def lookup_stage(ns, tag, robot, path):
    from tools.make_golden import _table_arrays
    ml = ns.motion_lib_robot.MotionLibRobot(R.motion_cfg(robot, path), num_envs=2, device="cpu")
    ml.load_motions(random_sample=False)
    tab = _table_arrays(ml)
    one = {k: (v[:T] if v.shape[0] == 2 * T else v) for k, v in tab.items()}
    np.savez_compressed(os.path.join(GOLD, f"ref_table_edge{tag}.npz"), **one)
    dt = 1.0 / FPS
    blends = [0.0, 1e-4, 0.25, 0.5, 0.9999]
    t = torch.tensor([(i + b) * dt for i in range(T - 1) for b in blends], dtype=torch.float32)
    ids = torch.zeros(t.numel(), dtype=torch.long)
    off = torch.zeros(t.numel(), 3)
    res = ml.get_motion_state(ids, t.clone(), offset=off)
    np.savez_compressed(os.path.join(GOLD, f"ref_lookup_edge{tag}.npz"), motion_ids=_np(ids), motion_times=_np(t), offset=_np(off),
                        **{"out_" + k: _np(v) for k, v in res.items()})
    # what the fixture exercises, for the record
    q = torch.from_numpy(one["grs_t"])
    c = (q[1:] * q[:-1]).sum(-1)
    s = torch.sqrt(torch.clamp(1 - c * c, min=0))
    print(f"edge{tag}: frame pairs x bodies {c.numel()}: sign-flipped {(c < 0).sum().item()}, |cos|>=1 {(c.abs() >= 1).sum().item()}, "
          f"sin<1e-3 {(s < 1e-3).sum().item()}, sin in [0.9e-3, 1.1e-3] {((s > 0.9e-3) & (s < 1.1e-3)).sum().item()}")


def env_stage(tag, path, general):
    from tools import make_golden_env_step as G

    def start_times(env):  # env e sits on frame pair e (mod T-2) a little after the frame
        e = torch.arange(env.num_envs, dtype=torch.float32)
        return ((e % (T - 3)) + 0.3) / FPS - 2 * env.dt

    def quat_mul(a, b):
        x1, y1, z1, w1 = a.unbind(-1)
        x2, y2, z2, w2 = b.unbind(-1)
        return torch.stack([w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2, w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2,
                            w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2, w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2], dim=-1)

    def tweak(step, body, root, dof, cf, ref):
        # the first envs: body rotations = the reference's rotated by 0 .. 40 urad about x (quat_to_angle_axis threshold)
        Jb = body.shape[1]
        angs = [0.0, 1.0e-5, 1.9e-5, 2.0e-5, 2.1e-5, 3.0e-5, 4.0e-5, 1.0e-6]
        for e, a in enumerate(angs):
            tiny = torch.tensor([np.sin(a / 2), 0.0, 0.0, np.cos(a / 2)], dtype=torch.float32)
            body[e, :, 3:7] = quat_mul(tiny.expand(Jb, 4), ref["rg_rot_t"][e, :Jb])
            body[e, :, 0:3] = ref["rg_pos_t"][e, :Jb] + 0.01
        root[:len(angs)] = body[:len(angs), 0]

    if general:
        G.run_case("gt29_edge", 48, 4, "motion_tracking/obs_ppo_teacher", 31, G.ZERO_NOISE_C, general=True, motion_file=path,
                   motion_name="edge29", start_times=start_times, tweak=tweak)
    else:
        G.run_case("mt23_edge", 48, 4, "motion_tracking/main", 32, G.ZERO_NOISE, motion_file=path, motion_name="edge23",
                   start_times=start_times, tweak=tweak)


def main():
    tmp = "/tmp/pbhc_golden_edge"
    os.makedirs(tmp, exist_ok=True)
    p23, p29 = write_clip("23", 27, 7, tmp), write_clip("29", 33, 8, tmp)
    os.chdir(R.REF_ROOT)
    ns = R.load_modules()
    torch.set_num_threads(1)
    lookup_stage(ns, "23", R.G1_23, p23)
    lookup_stage(ns, "29", R.G1_29, p29)
    env_stage("23", p23, False)
    env_stage("29", p29, True)
    print("edge goldens written to", GOLD)


if __name__ == "__main__":
    main()

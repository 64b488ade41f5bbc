"""CPU oracle for the PBHC motion-tracking hot path.  TEST INFRASTRUCTURE ONLY.

A torch-fp32 (CPU) restatement of the reference algorithm for the path SURVEY.md section 8 names:
motion lookup (L1-L3), G1 forward kinematics (F1-F3), the motion_tracking env step
(P0-P5, T1, T2-A, R*, O*, H1) and GAE (G1-G3).  Each function cites the reference file:line it
follows (paths relative to /root/reference/humanoidverse).  Only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs may import this file; the product package
pbhc_b200/ never does and fails loudly without its CUDA library.

Parity pin: the reference ships no golden vectors for this path.  The oracle is
pinned instead against outputs of the unmodified reference itself, generated in the build
container by tools/make_golden.py and committed under tests/golden/ (tests/test_oracle_golden.py).

The arithmetic is written with the same torch ops in the same association order as the reference
so that fp32 rounding matches it bit for bit wherever that is cheap; where a different but
equivalent op is used (e.g. where() instead of masked assignment) the values are unchanged.
"""
import math
from types import SimpleNamespace

import numpy as np
import torch

# --------------------------------------------------------------------------------------
# quaternion helpers (XYZW unless stated)                      isaac_utils/isaac_utils/rotations.py
# --------------------------------------------------------------------------------------


def normalize(x, eps=1e-9):
    """maths.py:7-8"""
    return x / x.norm(p=2, dim=-1).clamp(min=eps).unsqueeze(-1)


def slerp(q0, q1, t):
    """rotations.py:210-232 (not renormalised; thresholds 0.001 and 1 are part of the contract)."""
    c = torch.sum(q0 * q1, dim=-1)
    q1 = torch.where((c < 0).unsqueeze(-1), -q1, q1)
    c = torch.abs(c).unsqueeze(-1)
    h = torch.acos(c)
    s = torch.sqrt(1.0 - c * c)
    ra = torch.sin((1 - t) * h) / s
    rb = torch.sin(t * h) / s
    q = ra * q0 + rb * q1
    q = torch.where(torch.abs(s) < 0.001, 0.5 * q0 + 0.5 * q1, q)
    q = torch.where(torch.abs(c) >= 1, q0, q)
    return q


def _rot_terms(q, v):
    """shared body of quat_rotate / quat_rotate_inverse / my_quat_rotate (rotations.py:83-117, 244-253)."""
    qw = q[:, 3]
    qv = q[:, :3]
    a = v * (2.0 * qw ** 2 - 1.0).unsqueeze(-1)
    b = torch.cross(qv, v, dim=-1) * qw.unsqueeze(-1) * 2.0
    n = q.shape[0]
    c = qv * torch.bmm(qv.reshape(n, 1, 3), v.reshape(n, 3, 1)).squeeze(-1) * 2.0
    return a, b, c


def quat_rotate(q, v):
    """my_quat_rotate, rotations.py:244-253; q [M,4] xyzw, v [M,3]."""
    a, b, c = _rot_terms(q, v)
    return a + b + c


def quat_rotate_inverse(q, v):
    """rotations.py:101-117 / utils/torch_utils.py:72-81."""
    a, b, c = _rot_terms(q, v)
    return a - b + c


def quat_mul(a, b):
    """rotations.py:414-441 with w_last=True (the 8-multiplication form; association order kept)."""
    shape = a.shape
    a = a.reshape(-1, 4)
    b = b.reshape(-1, 4)
    x1, y1, z1, w1 = a[:, 0], a[:, 1], a[:, 2], a[:, 3]
    x2, y2, z2, w2 = b[:, 0], b[:, 1], b[:, 2], b[:, 3]
    ww = (z1 + x1) * (x2 + y2)
    yy = (w1 - y1) * (w2 + z2)
    zz = (w1 + y1) * (w2 - z2)
    xx = ww + yy + zz
    qq = 0.5 * (xx + (z1 - x1) * (x2 - y2))
    w = qq - ww + (z1 - y1) * (y2 - z2)
    x = qq - xx + (x1 + w1) * (x2 + w2)
    y = qq - yy + (w1 - x1) * (y2 + z2)
    z = qq - zz + (z1 + y1) * (w2 - x2)
    return torch.stack([x, y, z, w], dim=-1).view(shape)


def quat_conjugate(a):
    """rotations.py:56-63 (w_last)."""
    return torch.cat((-a[..., :3], a[..., 3:]), dim=-1)


def quat_from_angle_axis(angle, axis):
    """rotations.py:138-146 (w_last), including both normalisations."""
    theta = (angle / 2).unsqueeze(-1)
    xyz = normalize(axis) * theta.sin()
    w = theta.cos()
    return normalize(torch.cat([xyz, w], dim=-1))


def calc_heading(q):
    """rotations.py:256-268."""
    ref_dir = torch.zeros_like(q[..., 0:3])
    ref_dir[..., 0] = 1
    rot_dir = quat_rotate(q, ref_dir)
    return torch.atan2(rot_dir[..., 1], rot_dir[..., 0])


def calc_heading_quat_inv(q):
    """rotations.py:296-306 (w_last=True)."""
    heading = calc_heading(q)
    axis = torch.zeros_like(q[..., 0:3])
    axis[..., 2] = 1
    return quat_from_angle_axis(-heading, axis)


def get_euler_xyz_in_tensor(q):
    """rotations.py:368-387 (no 2*pi wrap)."""
    qx, qy, qz, qw = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    sinr_cosp = 2.0 * (qw * qx + qy * qz)
    cosr_cosp = qw * qw - qx * qx - qy * qy + qz * qz
    roll = torch.atan2(sinr_cosp, cosr_cosp)
    sinp = 2.0 * (qw * qy - qz * qx)
    half_pi = torch.abs(torch.full_like(sinp, np.pi / 2.0)) * torch.sign(sinp)  # copysign, maths.py:16-19
    pitch = torch.where(torch.abs(sinp) >= 1, half_pi, torch.asin(sinp))
    siny_cosp = 2.0 * (qw * qz + qx * qy)
    cosy_cosp = qw * qw + qx * qx - qy * qy - qz * qz
    yaw = torch.atan2(siny_cosp, cosy_cosp)
    return torch.stack((roll, pitch, yaw), dim=-1)


def calc_yaw_heading_quat_inv(yaw):
    """rotations.py:309-322: (0, 0, -sin(yaw / 2), cos(yaw / 2)) for yaw [B,1]."""
    half = yaw[..., 0] * 0.5
    s, c = torch.sin(half), torch.cos(half)
    z = torch.zeros_like(s)
    return torch.stack([z, z, -s, c], dim=-1)


def wrap_to_pi(angles):
    """rotations.py:49-53 (in place there; on a copy here).  The reference function is @torch.jit.script, and TorchScript
    compiles `angles %= 2*np.pi` to aten::fmod_ (sign of the dividend), not to the remainder eager Python would take: a
    negative angle stays negative, and values in (-2 pi, -pi) are NOT wrapped."""
    angles = torch.fmod(angles, 2 * np.pi)
    angles -= 2 * np.pi * (angles > np.pi)
    return angles


def quat_apply(a, b):
    """utils/torch_utils.py:51-58 (xyzw): b + w*t + xyz x t, t = 2 (xyz x b)."""
    shape = b.shape
    a = a.reshape(-1, 4)
    b = b.reshape(-1, 3)
    xyz = a[:, :3]
    t = xyz.cross(b, dim=-1) * 2
    return (b + a[:, 3:] * t + xyz.cross(t, dim=-1)).view(shape)


def yaw_quat(quat):
    """utils/torch_utils.py:239-270 (w_last): yaw-only quaternion, renormalised."""
    shape = quat.shape
    f = quat.reshape(-1, 4)
    qx, qy, qz, qw = f[:, 0], f[:, 1], f[:, 2], f[:, 3]
    yaw = torch.atan2(2 * (qw * qz + qx * qy), 1 - 2 * (qy * qy + qz * qz))
    out = torch.zeros_like(f)
    out[:, 2] = torch.sin(yaw / 2)
    out[:, 3] = torch.cos(yaw / 2)
    return normalize(out).view(shape)


def matrix_from_quat(q):
    """utils/torch_utils.py:274-297 (w_last): rotation matrix [..., 3, 3] with 2/|q|^2."""
    i, j, k, r = torch.unbind(q, -1)
    two_s = 2.0 / (q * q).sum(-1)
    o = torch.stack((1 - two_s * (j * j + k * k), two_s * (i * j - k * r), two_s * (i * k + j * r),
                     two_s * (i * j + k * r), 1 - two_s * (i * i + k * k), two_s * (j * k - i * r),
                     two_s * (i * k - j * r), two_s * (j * k + i * r), 1 - two_s * (i * i + j * j)), -1)
    return o.reshape(q.shape[:-1] + (3, 3))


def quat_to_angle(q):
    """angle part of quat_to_angle_axis (rotations.py:185-207): 2 acos(w) wrapped by atan2(sin, cos), 0 where
    sqrt(1 - w^2) <= 1e-5 (NaN-compare false also gives 0)."""
    w = q[..., 3]
    sin_theta = torch.sqrt(1 - w * w)
    angle = 2 * torch.acos(w)
    angle = torch.atan2(torch.sin(angle), torch.cos(angle))
    return torch.where(torch.abs(sin_theta) > 1e-5, angle, torch.zeros_like(angle))


# ---- WXYZ helpers used by FK (rotations.py:519-636, "FROM PHC rotation_conversions.py") ----


def axis_angle_to_quaternion_wxyz(aa):
    """rotations.py:554-578: small-angle series below 1e-6."""
    angles = torch.norm(aa, p=2, dim=-1, keepdim=True)
    half = angles * 0.5
    small = angles.abs() < 1e-6
    big_val = torch.sin(half) / angles
    small_val = 0.5 - (angles * angles) / 48
    k = torch.where(small, small_val, big_val)
    return torch.cat([torch.cos(half), aa * k], dim=-1)


def quaternion_to_matrix_wxyz(q):
    """rotations.py:519-550 (two_s = 2/|q|^2)."""
    r, i, j, k = torch.unbind(q, -1)
    two_s = 2.0 / (q * q).sum(-1)
    o = torch.stack((
        1 - two_s * (j * j + k * k), two_s * (i * j - k * r), two_s * (i * k + j * r),
        two_s * (i * j + k * r), 1 - two_s * (i * i + k * k), two_s * (j * k - i * r),
        two_s * (i * k - j * r), two_s * (j * k + i * r), 1 - two_s * (i * i + j * j)), -1)
    return o.reshape(q.shape[:-1] + (3, 3))


def matrix_to_quaternion_wxyz(m):
    """rotations.py:589-636: best-conditioned of four candidates, denominators floored at 0.1."""
    batch = m.shape[:-2]
    m00, m01, m02, m10, m11, m12, m20, m21, m22 = torch.unbind(m.reshape(batch + (9,)), dim=-1)
    pre = torch.stack([1.0 + m00 + m11 + m22, 1.0 + m00 - m11 - m22,
                       1.0 - m00 + m11 - m22, 1.0 - m00 - m11 + m22], dim=-1)
    q_abs = torch.where(pre > 0, torch.sqrt(pre.clamp(min=0)), torch.zeros_like(pre))  # :639-647
    cand = torch.stack([
        torch.stack([q_abs[..., 0] ** 2, m21 - m12, m02 - m20, m10 - m01], dim=-1),
        torch.stack([m21 - m12, q_abs[..., 1] ** 2, m10 + m01, m02 + m20], dim=-1),
        torch.stack([m02 - m20, m10 + m01, q_abs[..., 2] ** 2, m12 + m21], dim=-1),
        torch.stack([m10 - m01, m20 + m02, m21 + m12, q_abs[..., 3] ** 2], dim=-1)], dim=-2)
    cand = cand / (2.0 * q_abs[..., None].clamp(min=0.1))
    best = q_abs.argmax(dim=-1)
    idx = best[..., None, None].expand(batch + (1, 4))
    return torch.gather(cand, -2, idx).squeeze(-2)


# --------------------------------------------------------------------------------------
# kinematic tree + forward kinematics                 utils/motion_lib/torch_humanoid_batch.py
# --------------------------------------------------------------------------------------


def make_tree(names, parents, offsets, local_rot_wxyz, num_bodies, dof_axis=None):
    """Tree constants as Humanoid_Batch.__init__ holds them (:52-101): the MJCF bodies in DFS order
    followed by the extend_config bodies."""
    t = SimpleNamespace()
    t.names = list(names)
    t.parents = [int(p) for p in parents]
    t.offsets = torch.as_tensor(np.asarray(offsets), dtype=torch.float32).reshape(-1, 3)
    t.local_rot = torch.as_tensor(np.asarray(local_rot_wxyz), dtype=torch.float32).reshape(-1, 4)
    t.local_rot_mat = quaternion_to_matrix_wxyz(t.local_rot).float()  # :101
    t.num_bodies = int(num_bodies)
    t.num_joints = len(t.parents)
    t.dof_axis = None if dof_axis is None else torch.as_tensor(np.asarray(dof_axis))
    return t


def fk(tree, pose_aa, trans):
    """fk_batch + forward_kinematics_batch (:168-207, :229-269) without velocities.

    pose_aa [B,T,J,3] axis-angle, trans [B,T,3]  ->  pos [B,T,J,3], rotmat [B,T,J,3,3],
    quat_xyzw [B,T,J,4] for all J = num_bodies + extended bodies."""
    J = tree.num_joints
    pose = pose_aa[..., :J, :]
    B, T = pose.shape[:2]
    pose_quat = axis_angle_to_quaternion_wxyz(pose.clone())
    pose_mat = quaternion_to_matrix_wxyz(pose_quat).reshape(B, T, -1, 3, 3)
    rot_w = [None] * J
    pos_w = [None] * J
    for i in range(J):
        p = tree.parents[i]
        if p == -1:
            pos_w[i] = trans
            rot_w[i] = pose_mat[:, :, 0:1]
        else:
            off = tree.offsets[i].expand(B, T, 3)
            pos_w[i] = torch.matmul(rot_w[p][:, :, 0], off[:, :, :, None]).squeeze(-1) + pos_w[p]
            rot_w[i] = torch.matmul(rot_w[p], torch.matmul(tree.local_rot_mat[None, i:i + 1], pose_mat[:, :, i:i + 1]))
    pos = torch.stack(pos_w, dim=2)
    rot = torch.cat(rot_w, dim=2)
    quat = matrix_to_quaternion_wxyz(rot)[..., [1, 2, 3, 0]]  # wxyz_to_xyzw, :188
    return pos, rot, quat, pose_quat


def _gauss_filter_time(x, axis):
    import scipy.ndimage
    return scipy.ndimage.gaussian_filter1d(x, 2, axis=axis, mode="nearest")


def body_velocity(pos, dt):
    """_compute_velocity (:272-279): central differences then Gaussian sigma=2 'nearest' along time."""
    v = np.gradient(pos.numpy(), axis=-3) / dt
    return torch.from_numpy(_gauss_filter_time(v, -3)).to(pos)


def body_angular_velocity(quat_xyzw, dt):
    """_compute_angular_velocity (:282-290): q[t+1] * conj(q[t]) normalised -> angle*axis/dt, filtered.
    The last frame's difference is the identity (zero velocity before filtering)."""
    r = quat_xyzw
    dq = torch.zeros_like(r)
    dq[..., 3] = 1.0
    prod = quat_mul(r[..., 1:, :, :].contiguous(), quat_conjugate(r[..., :-1, :, :]).contiguous())
    dq[..., :-1, :, :] = normalize(prod)  # quat_mul_norm, rotations.py:452-459 (second definition wins)
    w = dq[..., 3]
    axis = dq[..., :3].clone()
    s = 2 * (w ** 2) - 1
    angle = s.clamp(-1, 1).arccos()  # quat_angle_axis, rotations.py:120-135
    axis = axis / axis.norm(p=2, dim=-1, keepdim=True).clamp(min=1e-9)
    av = axis * angle.unsqueeze(-1) / dt
    return torch.from_numpy(_gauss_filter_time(av.numpy(), -3))


def rebase_heading(pose_aa, trans, target_heading):
    """motion_lib_base.py:447-460 (load_motion_with_skeleton, target_heading given): rotate the clip about z so that its first
    frame heads along target_heading (xyzw).  fp64 through scipy like the reference; calc_heading_quat_inv is the reference's
    torch function (rotations.py:296-306) evaluated on the fp64 start quaternion.  Returns (pose_aa fp32, trans fp64) - the
    reference then runs its FK with the fp64 translations."""
    from scipy.spatial.transform import Rotation as sRot
    pose_aa = torch.as_tensor(np.asarray(pose_aa)).float().clone()
    trans = torch.as_tensor(np.asarray(trans)).float().clone()
    start_root_rot = sRot.from_rotvec(pose_aa[0, 0])
    heading_inv_rot = sRot.from_quat(calc_heading_quat_inv(torch.from_numpy(start_root_rot.as_quat()[None,])))
    heading_delta = sRot.from_quat(target_heading) * heading_inv_rot
    pose_aa[:, 0] = torch.tensor((heading_delta * sRot.from_rotvec(pose_aa[:, 0])).as_rotvec())
    trans = torch.matmul(trans.to(torch.float64), torch.from_numpy(heading_delta.as_matrix().squeeze().T))
    return pose_aa, trans


def load_clip(tree, pose_aa, trans, fps, contact_mask=None):
    """What load_motion_with_skeleton + fk_batch(return_full=True) produce for one clip
    (motion_lib_base.py:403-472, torch_humanoid_batch.py:168-226).  Returns per-frame arrays
    for the *extended* body set (the non-extended arrays are their [:num_bodies] prefix)."""
    pose_aa = torch.as_tensor(pose_aa).clone()
    trans = torch.as_tensor(trans).clone()
    dt = 1 / fps
    pos, rot, quat, _ = fk(tree, pose_aa[None], trans[None])
    c = SimpleNamespace()
    c.num_frames = int(pose_aa.shape[0])
    c.fps = int(1 / dt)
    c.pos = pos[0].float()
    c.rot = quat[0].float()
    c.vel = body_velocity(pos, dt)[0].float()
    c.ang_vel = body_angular_velocity(quat, dt)[0].float()
    nb = tree.num_bodies
    dof_pos = pose_aa[None][..., :tree.num_joints, :].sum(dim=-1)[..., 1:nb]
    dv = (dof_pos[:, 1:] - dof_pos[:, :-1]) / dt
    c.dof_pos = dof_pos[0].float()
    c.dof_vel = torch.cat([dv, dv[:, -2:-1]], dim=1)[0].float()  # :223-224 pads with the 2nd-to-last row
    c.contact = None if contact_mask is None else torch.as_tensor(np.asarray(contact_mask)).float()
    return c


class MotionTable:
    """Concatenation of clips exactly as load_motions lays them out (motion_lib_base.py:342-384):
    per-clip num_frames / length / dt / fps and length_starts = cumsum(roll(num_frames))."""

    def __init__(self, clips, num_bodies, motion_to_clip=None):
        """motion_to_clip (optional, [num_motions] long): the reference holds one copy of its clip per sampled
        motion; with this indirection the oracle stores each unique clip once - same values, same indices
        relative to the clip start."""
        self.num_bodies = num_bodies
        self.clips = list(clips)
        cat = lambda k: torch.cat([getattr(c, k) for c in clips], dim=0).float()
        self.pos, self.rot, self.vel, self.ang_vel = cat("pos"), cat("rot"), cat("vel"), cat("ang_vel")
        self.dof_pos, self.dof_vel = cat("dof_pos"), cat("dof_vel")
        self.contact = cat("contact") if clips[0].contact is not None else None
        nf = [c.num_frames for c in clips]
        fps = [c.fps for c in clips]
        self.num_frames = torch.tensor(nf)
        self.fps = torch.tensor(fps, dtype=torch.float32)
        self.dt = torch.tensor([1.0 / f for f in fps], dtype=torch.float32)
        self.length = torch.tensor([1.0 / f * (n - 1) for f, n in zip(fps, nf)], dtype=torch.float32)
        shifted = self.num_frames.roll(1)
        shifted[0] = 0
        self.starts = shifted.cumsum(0)
        if motion_to_clip is not None:
            m = torch.as_tensor(motion_to_clip, dtype=torch.long)
            self.num_frames, self.fps, self.dt = self.num_frames[m], self.fps[m], self.dt[m]
            self.length, self.starts = self.length[m], self.starts[m]


def calc_frame_blend(time, length, num_frames, dt):
    """_calc_frame_blend, motion_lib_base.py:503-513.  idx0/idx1 are int64 and must be bit-exact."""
    time = time.clone()
    phase = torch.clip(time / length, 0.0, 1.0)
    time[time < 0] = 0
    idx0 = (phase * (num_frames - 1)).long()
    idx1 = torch.min(idx0 + 1, num_frames - 1)
    blend = torch.clip((time - idx0 * dt) / dt, 0.0, 1.0)
    return idx0, idx1, blend


def get_motion_state(tab, motion_ids, motion_times, offset=None):
    """get_motion_state, motion_lib_base.py:123-259, for robot clips (dof_pos present, extended
    arrays present, flags.real_traj off).  Also returns f0l/f1l/blend for index parity."""
    mlen = tab.length[motion_ids]
    nf = tab.num_frames[motion_ids]
    dt = tab.dt[motion_ids]
    i0, i1, blend = calc_frame_blend(motion_times, mlen, nf, dt)
    f0 = i0 + tab.starts[motion_ids]
    f1 = i1 + tab.starts[motion_ids]
    b = blend.unsqueeze(-1)
    be = b.unsqueeze(-1)
    nb = tab.num_bodies

    def lerp(a, w):
        return (1.0 - w) * a[f0] + w * a[f1]

    pos_t = lerp(tab.pos, be)
    pos = lerp(tab.pos[:, :nb], be)
    if offset is not None:
        pos_t = pos_t + offset[..., None, :]
        pos = pos + offset[..., None, :]
    vel_t, ang_t = lerp(tab.vel, be), lerp(tab.ang_vel, be)
    vel, ang = lerp(tab.vel[:, :nb], be), lerp(tab.ang_vel[:, :nb], be)
    dof_vel = lerp(tab.dof_vel, b)
    dof_pos = lerp(tab.dof_pos, b)
    rot_t = slerp(tab.rot[f0], tab.rot[f1], be)
    rot = slerp(tab.rot[:, :nb][f0], tab.rot[:, :nb][f1], be)
    out = {}
    if tab.contact is not None:
        out["contact_mask"] = lerp(tab.contact, b)
    out.update({
        "root_pos": pos[..., 0, :].clone(), "root_rot": rot[..., 0, :].clone(), "dof_pos": dof_pos.clone(),
        "root_vel": vel[..., 0, :].clone(), "root_ang_vel": ang[..., 0, :].clone(),
        "dof_vel": dof_vel.view(dof_vel.shape[0], -1),
        "motion_aa": torch.zeros(motion_ids.shape + (3 * nb,)),      # robot clips: zeros, :318-323
        "motion_bodies": torch.zeros(motion_ids.shape + (17,)),
        "rg_pos": pos, "rb_rot": rot, "body_vel": vel, "body_ang_vel": ang,
        "rg_pos_t": pos_t, "rg_rot_t": rot_t, "body_vel_t": vel_t, "body_ang_vel_t": ang_t,
    })
    out["_f0l"], out["_f1l"], out["_blend"] = f0, f1, blend
    return out


# --------------------------------------------------------------------------------------
# GAE                                            agents/ppo/ppo.py:338-355, mh_ppo/mh_ppo.py:376-395
# --------------------------------------------------------------------------------------


def gae(values, rewards, dones, last_values, gamma, lam, vec_reward):
    """values/rewards [T,N,R], dones [T,N,1] bool, last_values [N,R] -> (returns, normalised advantages [T,N,1]).
    vec_reward=True is MHPPO (advantages summed over the R heads before normalising)."""
    returns = torch.zeros_like(values)
    T = returns.shape[0]
    advantage = 0
    for step in reversed(range(T)):
        next_values = last_values if step == T - 1 else values[step + 1]
        nt = 1.0 - dones[step].float()
        delta = rewards[step] + nt * gamma * next_values - values[step]
        advantage = delta + nt * gamma * lam * advantage
        returns[step] = advantage + values[step]
    adv = returns - values
    if vec_reward:
        adv = adv.sum(dim=-1)
        adv = (adv - adv.mean()) / (adv.std() + 1e-8)
        return returns, adv.unsqueeze(-1)
    adv = (adv - adv.mean()) / (adv.std() + 1e-8)
    return returns, adv

"""Seeded synthetic simulator state (SURVEY.md 8d input table) - the stand-in for the external physics.
Pure torch, device-agnostic (tensors follow `ref`'s device and the generator's), no oracle / reference import:
used by the golden generator, the parity tests, smoke() and bench.py so that the CPU oracle and the GPU path
see identical bits."""
import math

import torch


def _quat_mul(a, b):
    x1, y1, z1, w1 = a.unbind(-1)
    x2, y2, z2, w2 = b.unbind(-1)
    return torch.stack([w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2, w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2,
                        w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2, w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2], dim=-1)


def small_rotations(shape, max_angle, gen):
    axis = torch.randn(*shape, 3, generator=gen)
    axis = axis / axis.norm(dim=-1, keepdim=True).clamp(min=1e-6)
    ang = torch.rand(*shape, 1, generator=gen) * max_angle
    return torch.cat([axis * torch.sin(ang / 2), torch.cos(ang / 2)], dim=-1)


def synth_sim_state(ref, num_bodies, gen, feet=(6, 12), far_frac=0.02, tilt_frac=0.02):
    """ref: get_motion_state() dict at (roughly) the time the env will query.  Returns Isaac-Gym-layout
    tensors: body_state [N,Jb,13], root_state [N,13], dof_state [N,D,2], contact_forces [N,Jb,3]."""
    Jb = num_bodies
    pos = ref["rg_pos_t"][:, :Jb].clone()
    rot = ref["rg_rot_t"][:, :Jb].clone()
    vel = ref["body_vel_t"][:, :Jb].clone()
    ang = ref["body_ang_vel_t"][:, :Jb].clone()
    N = pos.shape[0]
    pos += torch.randn(N, Jb, 3, generator=gen) * 0.05
    rot = _quat_mul(small_rotations((N, Jb), math.radians(10), gen), rot)
    rot = rot / rot.norm(dim=-1, keepdim=True)
    vel += torch.randn(N, Jb, 3, generator=gen) * 0.3
    ang += torch.randn(N, Jb, 3, generator=gen) * 1.0
    n_far = max(1, int(N * far_frac))
    far = torch.randperm(N, generator=gen)[:n_far]
    pos[far] += torch.tensor([2.0, 0.0, 0.0])
    n_tilt = max(1, int(N * tilt_frac))
    tilt = torch.randperm(N, generator=gen)[:n_tilt]
    tq = torch.tensor([math.sin(0.6), 0.0, 0.0, math.cos(0.6)])  # 1.2 rad roll -> gravity termination
    rot[tilt, 0] = _quat_mul(tq.expand(n_tilt, 4), rot[tilt, 0])
    body = torch.cat([pos, rot, vel, ang], dim=-1).contiguous()
    root = body[:, 0].clone()
    D = ref["dof_pos"].shape[1]
    dof = torch.stack([ref["dof_pos"] + torch.randn(N, D, generator=gen) * 0.1,
                       ref["dof_vel"] + torch.randn(N, D, generator=gen) * 1.0], dim=-1).contiguous()
    cf = torch.zeros(N, Jb, 3)
    for f in feet:
        on = torch.rand(N, generator=gen) < 0.5
        cf[:, f, 2] = torch.rand(N, generator=gen) * 600.0 * on
        cf[:, f, :2] = torch.randn(N, 2, generator=gen) * 20.0 * on[:, None]
    other = torch.rand(N, Jb, generator=gen) < 0.005
    cf += other[..., None] * torch.randn(N, Jb, 3, generator=gen) * 5.0
    return body, root, dof, cf


def synth_actions(N, D, gen):
    return (torch.randn(N, D, generator=gen)).clamp(-100, 100)

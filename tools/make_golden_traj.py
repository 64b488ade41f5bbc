#!/usr/bin/env python
"""Generates tests/golden/ref_traj.npz by calling the UNMODIFIED reference functions blend_motion / eval_accuracy /
eval_smoothness of /root/reference/humanoidverse/measure_traj.py on CPU.  Test infrastructure; run in the build
container from the repo root:   python -m tools.make_golden_traj

Inputs are synthetic but shaped like the real thing: `ref` = a 30-fps loaded motion (global_translation [T,J,3],
global_rotation [T,J,4] unit quaternions incl. sign flips, dof_pos [T,D], contact_mask [T,2]), `pol` = a 50-fps
recording (the blended reference plus noise) with its motion_times (incl. negative and past-the-end times).
"""
import contextlib
import io
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_loader as R  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def _import_measure_traj():
    R._install_paths()
    for name, attrs in (("hydra.core", {}), ("hydra.core.hydra_config", {"HydraConfig": object}),
                        ("hydra.core.config_store", {"ConfigStore": object}), ("toolz", {})):
        if name not in sys.modules:
            m = types.ModuleType(name)
            m.__dict__.update(attrs)
            sys.modules[name] = m
    import importlib
    return importlib.import_module("humanoidverse.measure_traj")


def make_inputs(seed=3, T_ref=64, J=24, D=23, fps_ref=30, fps_pol=50, T_pol=101):
    g = torch.Generator().manual_seed(seed)
    t = torch.arange(T_ref, dtype=torch.float32)[:, None, None] / fps_ref
    base = torch.randn(1, J, 3, generator=g)
    pos = base + 0.3 * torch.sin(2.0 * t + torch.randn(1, J, 3, generator=g)) + torch.tensor([0.5, 0.0, 0.0]) * t
    quat = torch.randn(T_ref, J, 4, generator=g) * 0.15 + torch.tensor([0.0, 0.0, 0.0, 1.0])
    quat = quat / quat.norm(dim=-1, keepdim=True)
    quat[::7] = -quat[::7]  # sign flips between neighbouring frames (slerp's neg_mask)
    quat[5] = quat[4]       # identical neighbours (cos_half_theta >= 1)
    dof = 0.5 * torch.sin(3.0 * t[:, :, 0] + torch.randn(1, D, generator=g))
    contact = (torch.rand(T_ref, 2, generator=g) < 0.5).float()
    pre = {"global_translation": pos.contiguous(), "global_rotation": quat.contiguous(), "dof_pos": dof.contiguous(),
           "contact_mask": contact, "fps": fps_ref}
    times = torch.arange(T_pol, dtype=torch.float32) / fps_pol
    times[0], times[-1] = -0.01, T_ref / fps_ref + 0.2  # out-of-range ends are clipped
    return pre, times, fps_pol, g


def main():
    mt = _import_measure_traj()
    pre, times, fps_pol, g = make_inputs()
    with contextlib.redirect_stdout(io.StringIO()):
        ref = mt.blend_motion(pre, times)
    pol = {k: (v + 0.02 * torch.randn(v.shape, generator=g) if k in ("global_translation", "dof_pos") else v)
           for k, v in ref.items() if k != "fps"}
    pol["contact_mask"] = (torch.rand(ref["contact_mask"].shape, generator=g) < 0.5).float()
    out = {f"pre_{k}": v.numpy() for k, v in pre.items() if k != "fps"}
    out["times"] = times.numpy()
    out["fps"] = np.array([pre["fps"], fps_pol], np.int64)
    out.update({f"ref_{k}": v.numpy() for k, v in ref.items() if k != "fps"})
    out.update({f"pol_{k}": v.numpy() for k, v in pol.items()})
    for tag, dpf in (("fps", False), ("frame", True)):
        traj = {"pol": pol, "ref": ref, "appendix": {"fps": fps_pol}}
        for k, v in mt.eval_accuracy(traj, delta_per_frame=dpf).items():
            out[f"acc_{tag}_{k}"] = v.numpy()
        for k, v in mt.eval_smoothness(traj, delta_per_frame=dpf).items():
            out[f"smooth_{tag}_{k}"] = v.numpy()
    np.savez_compressed(os.path.join(GOLD, "ref_traj.npz"), **out)
    print("wrote ref_traj.npz:", len(out), "arrays;", {k: float(v) for k, v in out.items() if k.startswith("acc_fps")})


if __name__ == "__main__":
    main()

"""Shared helpers for the test-suite: golden loading and oracle object construction."""
import json
import os

import numpy as np
import torch

from oracle import pbhc_oracle as O

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

# north_star tolerance: fp32 within 1e-5 relative / 1e-6 absolute
RTOL, ATOL = 1e-5, 1e-6


def gold(name):
    return dict(np.load(os.path.join(GOLD, name), allow_pickle=False))


def tree_json(tag):
    with open(os.path.join(GOLD, f"g1_{tag}dof_tree.json")) as f:
        return json.load(f)


def oracle_tree(tag):
    t = tree_json(tag)
    return O.make_tree(t["names"], t["parents"], t["offsets"], t["local_rot_wxyz"], t["num_bodies"], t["dof_axis"])


def motion_npz(name):
    return gold(f"motion_{name}.npz")


def oracle_clip(tag, name):
    m = motion_npz(name)
    return O.load_clip(oracle_tree(tag), m["pose_aa"], m["root_trans_offset"], int(m["fps"]),
                       m.get("contact_mask"))


def assert_close(a, b, rtol=RTOL, atol=ATOL, what=""):
    a = torch.as_tensor(np.asarray(a)).double()
    b = torch.as_tensor(np.asarray(b)).double()
    assert a.shape == b.shape, f"{what}: shape {tuple(a.shape)} vs {tuple(b.shape)}"
    err = (a - b).abs()
    tol = atol + rtol * b.abs()
    bad = err > tol
    if bad.any():
        i = torch.argmax(err - tol)
        raise AssertionError(f"{what}: {int(bad.sum())}/{bad.numel()} outside tol; worst |d|={err.flatten()[i]:.3e} "
                             f"at flat {int(i)} (got {a.flatten()[i]:.8g}, want {b.flatten()[i]:.8g})")

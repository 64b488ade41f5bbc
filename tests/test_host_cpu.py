"""Host-side logic that needs no GPU: MJCF parsing, table layout helpers, loud failure without CUDA."""
import os

import numpy as np
import pytest
import torch

from pbhc_b200 import _lib
from pbhc_b200.fk import Humanoid_Batch, parse_mjcf
from pbhc_b200.motion_lib import MotionLibRobot, gaussian_weights, record_stride
from tests._util import tree_json, product_cfg


@pytest.mark.parametrize("tag", ["23", "29"])
def test_humanoid_batch_constants_from_synthetic_mjcf(tag, tmp_path):
    cfg = product_cfg(tag, str(tmp_path), ["Horse-stance_pose"] if tag == "23" else ["g1_29dof_clip"])
    hb = Humanoid_Batch(cfg)
    t = tree_json(tag)
    assert hb.body_names_augment == t["names"]
    assert hb._parents.tolist() == t["parents"]
    assert np.array_equal(hb._offsets[0].numpy(), np.asarray(t["offsets"], dtype=np.float32))
    assert np.array_equal(hb._local_rotation[0].numpy(), np.asarray(t["local_rot_wxyz"], dtype=np.float32))
    assert hb.num_bodies == t["num_bodies"] and hb.num_dof == t["num_dof"]
    assert hb.dof_axis.tolist() == t["dof_axis"]
    assert np.allclose(hb.joints_range.numpy(), np.asarray(t["joints_range"]))


@pytest.mark.parametrize("tag,xml", [("23", "g1_23dof_lock_wrist_fitmotionONLY.xml"), ("29", "g1_29dof_rev_1_0.xml")])
def test_parser_on_reference_mjcf_when_present(tag, xml):
    path = os.path.join("/root/reference/description/robots/g1", xml)
    if not os.path.exists(path):
        pytest.skip("reference tree not present (GPU box)")
    d = parse_mjcf(path)
    t = tree_json(tag)
    nb = t["num_bodies"]
    assert d["node_names"] == t["names"][:nb]
    assert d["parent_indices"].tolist() == t["parents"][:nb]
    assert np.array_equal(d["local_translation"], np.asarray(t["offsets"][:nb], dtype=np.float32))
    assert np.array_equal(d["local_rotation"], np.asarray(t["local_rot_wxyz"][:nb], dtype=np.float32))
    assert d["dof_axis"] == t["dof_axis"]
    assert np.array_equal(d["joints_range"], np.asarray(t["joints_range"]))


def test_gaussian_weights_match_scipy():
    from scipy.ndimage import gaussian_filter1d
    r, w = gaussian_weights()
    assert r == 8 and len(w) == 17
    x = np.zeros(41)
    x[20] = 1.0
    assert np.array_equal(gaussian_filter1d(x, 2, mode="nearest")[12:29], w[::-1])


def test_record_stride():
    assert record_stride(27, 23, 2) == 400 and record_stride(33, 29, 0) == 488


def test_motion_lib_refuses_cpu(tmp_path):
    cfg = product_cfg("23", str(tmp_path), ["Horse-stance_pose"])
    with pytest.raises(_lib.PbhcError):
        MotionLibRobot(cfg, 4, "cpu")

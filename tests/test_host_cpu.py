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


def test_track_spec_parses_the_unused_sources_and_names_the_terms_the_reference_cannot_run():
    """TrackSpec (host logic, no GPU): the P4 / R15 keys of the reference resolve to source / reward ids; reward functions
    that raise inside the reference itself are refused with the reason (spec.BROKEN_IN_REFERENCE)."""
    import copy
    import json
    from pbhc_b200.spec import TrackSpec, OBS_SRC_IDS, REWARD_IDS, BROKEN_IN_REFERENCE
    from tests._util import gold
    cfg = json.loads(str(gold("ref_env_mt23_p4.npz")["meta_config_json"]))
    s = TrackSpec.from_config(cfg)
    assert s.future_num_steps == 3 and not s.general
    keys = {seg["key"]: seg["dim"] for seg in s.obs_groups["critic_obs"]["segs"] if seg["kind"] == "src"}
    assert keys["relyaw"] == 1 and keys["vr_3point_pos"] == 9 and keys["local_ref_rigid_body_pos_relyaw"] == 3 * s.J
    assert keys["future_ref_dof_pos"] == keys["future_ref_dof_vel"] == 3 * s.D
    assert OBS_SRC_IDS["future_ref_dof_vel"] == _lib.NUM_OBS_SRC - 1
    for name in ("teleop_radial_body_velocity_extend", "teleop_radial_joint_velocity", "feet_heading_alignment", "penalty_feet_ori_contact"):
        assert name in s.reward_names and name in REWARD_IDS
    for name in BROKEN_IN_REFERENCE:
        bad = copy.deepcopy(cfg)
        bad["rewards"]["reward_scales"][name] = 1.0
        with pytest.raises(NotImplementedError, match="cannot run in the reference"):
            TrackSpec.from_config(bad)
    g = json.loads(str(gold("ref_env_gt29_p4.npz")["meta_config_json"]))
    g["rewards"]["reward_scales"]["teleop_radial_joint_velocity"] = 1.0  # LeggedRobotMotionTracking only
    with pytest.raises(NotImplementedError, match="only in motion_tracking"):
        TrackSpec.from_config(g)


def test_track_spec_refuses_step_options_it_does_not_implement():
    """a config option that changes the reference's step and is not built must fail loudly, never be ignored"""
    import copy
    import json
    from pbhc_b200.spec import TrackSpec
    from tests._util import gold
    cfg = json.loads(str(gold("ref_env_mt23_push.npz")["meta_config_json"]))
    s = TrackSpec.from_config(cfg)
    assert s.dr["push_robots"] and s.dr["push_fixed"] and s.dr["push_interval_s"] == [5, 10] and not s.add_noise_curriculum

    def with_(path, value):
        c = copy.deepcopy(cfg)
        node = c
        ks = path.split(".")
        for k in ks[:-1]:
            node = node.setdefault(k, {})
        node[ks[-1]] = value
        return c
    for path, value in (("use_teleop_control", True), ("soft_dynamic_correction.enable", True), ("domain_rand.parallel_serial_pd.enable", True),
                        ("domain_rand.randomize_default_dof_pos", True), ("domain_rand.born_offset", True), ("domain_rand.reinit_epis_rand", 100),
                        ("domain_rand.motion_package_loss", True), ("obs.noise_process", {"enable": True}), ("termination.terminate_when_dof_far", True)):
        with pytest.raises(NotImplementedError):
            TrackSpec.from_config(with_(path, value))
    raw = copy.deepcopy(cfg)
    raw["obs"]["obs_dict"]["actor_obs"] = list(raw["obs"]["obs_dict"]["actor_obs"]) + ["dof_pos_raw"]
    with pytest.raises(NotImplementedError, match="cannot run in the reference"):
        TrackSpec.from_config(raw)
    e = TrackSpec.from_config(json.loads(str(gold("ref_env_mt23_eval.npz")["meta_config_json"])))
    assert e.add_noise_curriculum and e.noise_cur["initial"] == 0.5 and e.noise_cur["up"] == float(cfg["rewards"]["reward_penalty_level_up_threshold"])

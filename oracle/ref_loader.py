"""TEST INFRASTRUCTURE ONLY - imports the *unmodified* reference (kyungminn/PBHC) from
/root/reference so that (a) the CPU restatement in oracle/pbhc_oracle.py can be pinned
against it and (b) tools/make_golden.py can dump golden fixtures into tests/golden/.

/root/reference exists only in the build container; nothing that runs on the GPU box
(-m gpu tests, smoke(), bench.py) may import this module.  The product package
(pbhc_b200/) never imports anything under oracle/.

The reference needs seven third-party modules only at import time (easydict, lxml, open3d,
stl, hydra, omegaconf, termcolor, matplotlib); oracle/ref_stubs/ holds ~80 lines of stand-ins.
"""
import contextlib
import os
import sys
import types

REF_ROOT = os.environ.get("PBHC_REFERENCE_ROOT", "/root/reference")
_HERE = os.path.dirname(os.path.abspath(__file__))
_STUBS = os.path.join(_HERE, "ref_stubs")


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REF_ROOT, "humanoidverse"))


def _install_paths():
    for p in (REF_ROOT, os.path.join(REF_ROOT, "humanoidverse", "isaac_utils"), _STUBS):
        if p not in sys.path:
            sys.path.insert(0, p)


@contextlib.contextmanager
def _quiet():
    from loguru import logger
    logger.disable("humanoidverse")
    try:
        yield
    finally:
        pass


class AttrDict(dict):
    """attribute-style dict with .get - enough of OmegaConf for the hot-path code."""

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v

    def pop(self, k, *a):
        return dict.pop(self, k, *a)


def to_attr(x):
    if isinstance(x, dict):
        return AttrDict({k: to_attr(v) for k, v in x.items()})
    if isinstance(x, (list, tuple)):
        return [to_attr(v) for v in x]
    return x


def load_modules():
    """Returns a namespace of the reference modules on the hot path."""
    if not reference_available():
        raise RuntimeError(f"reference tree not found at {REF_ROOT}")
    _install_paths()
    import warnings
    warnings.filterwarnings("ignore")
    from loguru import logger
    logger.remove()
    ns = types.SimpleNamespace()
    import isaac_utils.rotations as rotations
    import isaac_utils.maths as maths
    ns.rotations, ns.maths = rotations, maths
    import humanoidverse.utils.torch_utils as torch_utils
    ns.torch_utils = torch_utils
    from humanoidverse.utils.motion_lib import torch_humanoid_batch, motion_lib_base, motion_lib_robot
    from humanoidverse.utils.motion_lib import motion_lib_robot_WJX
    ns.torch_humanoid_batch = torch_humanoid_batch
    ns.motion_lib_base = motion_lib_base
    ns.motion_lib_robot = motion_lib_robot
    ns.motion_lib_robot_WJX = motion_lib_robot_WJX
    return ns


G1_23 = dict(
    assetRoot=os.path.join(REF_ROOT, "description/robots/g1/"),
    assetFileName="g1_23dof_lock_wrist_fitmotionONLY.xml",
    extend=[
        dict(joint_name="left_hand_link", parent_name="left_elbow_link", pos=[0.25, 0.0, 0.0], rot=[1.0, 0.0, 0.0, 0.0]),
        dict(joint_name="right_hand_link", parent_name="right_elbow_link", pos=[0.25, 0.0, 0.0], rot=[1.0, 0.0, 0.0, 0.0]),
        dict(joint_name="head_link", parent_name="torso_link", pos=[0.0, 0.0, 0.42], rot=[1.0, 0.0, 0.0, 0.0]),
    ],
)
G1_29 = dict(
    assetRoot=os.path.join(REF_ROOT, "description/robots/g1/"),
    assetFileName="g1_29dof_rev_1_0.xml",
    extend=[
        dict(joint_name="left_hand_link", parent_name="left_elbow_link", pos=[0.25, 0.0, 0.0], rot=[1.0, 0.0, 0.0, 0.0]),
        dict(joint_name="right_hand_link", parent_name="right_elbow_link", pos=[0.25, 0.0, 0.0], rot=[1.0, 0.0, 0.0, 0.0]),
        dict(joint_name="head_link", parent_name="torso_link", pos=[0.0, 0.0, 0.42], rot=[1.0, 0.0, 0.0, 0.0]),
    ],
)


def motion_cfg(robot=G1_23, motion_file=None, step_dt=0.02):
    """motion_lib_cfg as config/robot/g1/*.yaml `robot.motion` would compose it."""
    return to_attr(dict(
        asset=dict(assetRoot=robot["assetRoot"], assetFileName=robot["assetFileName"]),
        motion_file=motion_file,
        extend_config=robot["extend"],
        step_dt=step_dt,
    ))


def example_motion(name="Horse-stance_pose"):
    return os.path.join(REF_ROOT, "example/motion_data", name + ".pkl")

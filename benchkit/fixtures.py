"""Product-side fixtures built only from committed files under tests/golden/ (no /root/reference, no oracle/):
a synthesised G1 MJCF, a joblib motion file and the motion_lib_cfg / env config the drop-in classes take.
Used by tests, smoke() and bench.py."""
import json
import os

import numpy as np

GOLD = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def gold(name):
    return dict(np.load(os.path.join(GOLD, name), allow_pickle=False))


def tree_json(tag):
    with open(os.path.join(GOLD, f"g1_{tag}dof_tree.json")) as f:
        return json.load(f)


# the multi-clip 29-DoF library of config C: three clips cut out of the one committed 29-DoF sequence (frame ranges)
CLIPS29 = {"clip_full": (0, 520), "clip_a": (100, 300), "clip_b": (250, 500)}


def motion_npz(name):
    if name in CLIPS29:
        lo, hi = CLIPS29[name]
        m = gold("motion_g1_29dof_clip.npz")
        return {k: (v[lo:hi] if getattr(v, "ndim", 0) else v) for k, v in m.items()}
    return gold(f"motion_{name}.npz")


CLIPS23 = ["Horse-stance_pose", "Bruce_Lee_pose", "Charleston_dance", "Hooks_punch", "Horse-stance_punch", "Roundhouse_kick", "Side_kick"]


def library29(n_clips=1024, seed=0, min_frames=150, max_frames=600):
    """The multi-clip 29-DoF motion library of SURVEY.md 8d (config C): the shipped 29-DoF clip plus the seven 23-DoF example
    clips with zero wrist joints inserted (inverse of convert_29dof_to_23dof.py:39-46: dofs 19:22 and 26:29, i.e. pose_aa
    joints 20..22 and 27..29), each time-warped (linear resampling of pose_aa / root translation) to min_frames..max_frames
    frames.  1024 clips are ~0.38 M frames = ~0.75 GB of packed records: the table streams from HBM, not from L2.
    Returns {name: entry} in the joblib layout MotionLibRobot loads (deterministic in seed)."""
    rng = np.random.RandomState(seed)
    base = []
    m = gold("motion_g1_29dof_clip.npz")
    base.append((np.asarray(m["pose_aa"], np.float32), np.asarray(m["root_trans_offset"], np.float32), int(m["fps"])))
    for nm in CLIPS23:
        m = gold(f"motion_{nm}.npz")
        p23 = np.asarray(m["pose_aa"], np.float32)
        p29 = np.zeros((p23.shape[0], 33, 3), np.float32)
        p29[:, 0:20] = p23[:, 0:20]
        p29[:, 23:27] = p23[:, 20:24]
        base.append((p29, np.asarray(m["root_trans_offset"], np.float32), int(m["fps"])))
    out = {}
    for i in range(n_clips):
        pose, trans, fps = base[i % len(base)]
        T0 = pose.shape[0]
        T = int(rng.randint(min_frames, max_frames + 1))
        x = np.linspace(0.0, T0 - 1.0, T, dtype=np.float64)
        i0 = np.floor(x).astype(np.int64)
        i1 = np.minimum(i0 + 1, T0 - 1)
        w = (x - i0).astype(np.float32)
        p = pose[i0] * (1.0 - w)[:, None, None] + pose[i1] * w[:, None, None]
        t = trans[i0] * (1.0 - w)[:, None] + trans[i1] * w[:, None]
        out[f"clip{i:04d}"] = dict(root_trans_offset=t.astype(np.float32), pose_aa=p.astype(np.float32), dof=None, root_rot=None, fps=fps)
    return out


def clips_of_case(cfg):
    """(robot tag, fixture clip names) of a golden env fixture, from the motion file name recorded in its config:
    "Horse-stance_pose", "clips29" (the 3-clip library CLIPS29), or a single named fixture clip ("edge23", "edge29")."""
    general = cfg.get("env_class") == "general_tracking"
    name = cfg["robot"]["motion"]["motion_file"]
    tag = "29" if general else "23"
    if name == "clips29":
        return tag, list(CLIPS29)
    return tag, [name]


MAIN_NOISE = dict(base_ang_vel=0.3, projected_gravity=0.2, dof_pos=0.01, dof_vel=1.0)  # config/obs/motion_tracking/main.yaml:81-99


def env_config(case="mt23_main", noise=None, push_robots=None):
    """the composed reference env config (hydra `env.config`) saved with the golden env fixture.  The fixture was
    recorded with every observation-noise scale at 0 (torch.rand_like cannot be replayed); `noise` = dict of
    noise_scales to put back (MAIN_NOISE = the values obs/motion_tracking/main.yaml ships)."""
    cfg = json.loads(str(gold(f"ref_env_{case}.npz")["meta_config_json"]))
    if noise:
        cfg["obs"]["noise_scales"].update(noise)
    if push_robots is not None:  # the fixtures were recorded with pushes off; domain_rand/main.yaml ships them on (_push_fixed, [5, 10) s, 0.1 m/s)
        cfg["domain_rand"]["push_robots"] = bool(push_robots)
    return cfg


EXTEND = [dict(joint_name="left_hand_link", parent_name="left_elbow_link", pos=[0.25, 0.0, 0.0], rot=[1.0, 0.0, 0.0, 0.0]),
          dict(joint_name="right_hand_link", parent_name="right_elbow_link", pos=[0.25, 0.0, 0.0], rot=[1.0, 0.0, 0.0, 0.0]),
          dict(joint_name="head_link", parent_name="torso_link", pos=[0.0, 0.0, 0.42], rot=[1.0, 0.0, 0.0, 0.0])]


def write_mjcf(tag, path):
    """Synthesises an MJCF with the kinematic constants of tests/golden/g1_<tag>dof_tree.json (bodies in
    DFS order, one hinge per non-root body, a free joint on the root, one motor per hinge)."""
    import xml.etree.ElementTree as ET
    t = tree_json(tag)
    nb = t["num_bodies"]
    root = ET.Element("mujoco", model=f"g1_{tag}")
    world = ET.SubElement(root, "worldbody")
    act = ET.SubElement(root, "actuator")
    nodes = []
    fmt = lambda v: " ".join(repr(float(np.float32(x))) for x in v)
    for i in range(nb):
        parent = world if t["parents"][i] < 0 else nodes[t["parents"][i]]
        b = ET.SubElement(parent, "body", name=t["names"][i], pos=fmt(t["offsets"][i]), quat=fmt(t["local_rot_wxyz"][i]))
        nodes.append(b)
        if i == 0:
            ET.SubElement(b, "joint", name="floating_base_joint", type="free")
        else:
            jn = t["names"][i].replace("_link", "_joint")
            ET.SubElement(b, "joint", name=jn, type="hinge", axis=" ".join(str(int(a)) for a in t["dof_axis"][i - 1]),
                          range=" ".join(repr(float(x)) for x in t["joints_range"][i - 1]))
            ET.SubElement(act, "motor", name=jn, joint=jn)
    ET.ElementTree(root).write(path)
    return path


def product_cfg(tag, tmpdir, motion_names, motion_data=None):
    """motion_lib_cfg for pbhc_b200.MotionLibRobot built only from committed fixtures.  motion_data: an in-memory
    {name: entry} library (library29) instead of the named fixture clips."""
    import joblib
    from pbhc_b200.fk import AttrDict
    os.makedirs(tmpdir, exist_ok=True)
    write_mjcf(tag, os.path.join(tmpdir, f"g1_{tag}.xml"))
    if motion_data is not None:
        return AttrDict(asset=AttrDict(assetRoot=str(tmpdir), assetFileName=f"g1_{tag}.xml"), motion_file=motion_data,
                        extend_config=[AttrDict(e) for e in EXTEND], step_dt=0.02)
    data = {}
    for nm in motion_names:
        m = motion_npz(nm)
        e = dict(root_trans_offset=m["root_trans_offset"], pose_aa=m["pose_aa"], dof=m["dof"], root_rot=m["root_rot"],
                 fps=int(m["fps"]))
        if "contact_mask" in m:
            e["contact_mask"] = m["contact_mask"]
        data[nm] = e
    pkl = os.path.join(tmpdir, f"motions_{tag}_{len(motion_names)}.pkl")
    joblib.dump(data, pkl)
    return AttrDict(asset=AttrDict(assetRoot=str(tmpdir), assetFileName=f"g1_{tag}.xml"), motion_file=pkl,
                    extend_config=[AttrDict(e) for e in EXTEND], step_dt=0.02)

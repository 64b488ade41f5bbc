#!/usr/bin/env python
"""Generate tests/golden/* by running the UNMODIFIED reference (kyungminn/PBHC @ /root/reference)
on CPU in the build container.  Test infrastructure; run from the repo root:

    python tools/make_golden.py [stage ...]      # stages: assets lookup fk env gae (default: all)

Outputs (all small, committed):
  g1_{23,29}dof_tree.json        kinematic tree constants parsed by the reference's Humanoid_Batch
  motion_<name>.npz              the example motion clips' raw arrays (inputs: pose_aa, trans, dof, ...)
  ref_table_<name>.npz           reference load_motions() per-frame arrays for one clip (FK + velocities)
  ref_lookup_<cfg>.npz           reference get_motion_state() on seeded queries incl. edge cases
  ref_fk_<cfg>.npz               reference fk_batch() on seeded random poses
  ref_env_*.npz / ref_gae_*.npz  written by the env / gae stages (see tools/make_golden_env.py)
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_loader as R  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
EXAMPLES = ["Horse-stance_pose", "Bruce_Lee_pose", "Charleston_dance", "Hooks_punch",
            "Horse-stance_punch", "Roundhouse_kick", "Side_kick"]
CLIP29 = os.path.join(R.REF_ROOT, "motion_data", "g1_rig_Skeleton_Sequence_converted_processed_g1_29dof_rev_1_0.pkl")


def _np(x):
    return x.detach().cpu().numpy() if torch.is_tensor(x) else np.asarray(x)


def stage_assets(ns):
    import joblib
    for tag, robot in (("23", R.G1_23), ("29", R.G1_29)):
        cfg = R.motion_cfg(robot, R.example_motion())
        hb = ns.torch_humanoid_batch.Humanoid_Batch(cfg)
        tree = dict(
            names=list(hb.body_names_augment),
            parents=[int(p) for p in hb._parents],
            offsets=_np(hb._offsets[0]).astype(np.float32).tolist(),
            local_rot_wxyz=_np(hb._local_rotation[0]).astype(np.float32).tolist(),
            num_bodies=int(hb.num_bodies),
            num_dof=int(hb.num_dof),
            dof_axis=_np(hb.dof_axis).astype(int).tolist(),
            joints_range=_np(hb.joints_range).astype(np.float64).tolist(),
            source=f"description/robots/g1/{robot['assetFileName']} + extend_config (robot yaml)",
        )
        with open(os.path.join(GOLD, f"g1_{tag}dof_tree.json"), "w") as f:
            json.dump(tree, f)
    for name in EXAMPLES:
        d = joblib.load(R.example_motion(name))
        m = d[next(iter(d))]
        np.savez_compressed(
            os.path.join(GOLD, f"motion_{name}.npz"),
            root_trans_offset=m["root_trans_offset"], pose_aa=m["pose_aa"], dof=m["dof"],
            root_rot=m["root_rot"], fps=np.int64(m["fps"]),
            contact_mask=np.asarray(m["contact_mask"]))
    d = joblib.load(CLIP29)
    m = d[next(iter(d))]
    extra = {"contact_mask": np.asarray(m["contact_mask"])} if "contact_mask" in m else {}
    np.savez_compressed(os.path.join(GOLD, "motion_g1_29dof_clip.npz"),
                        root_trans_offset=m["root_trans_offset"], pose_aa=m["pose_aa"], dof=m["dof"],
                        root_rot=m["root_rot"], fps=np.int64(m["fps"]), **extra)


def _table_arrays(ml):
    out = dict(gts_t=_np(ml.gts_t), grs_t=_np(ml.grs_t), gvs_t=_np(ml.gvs_t), gavs_t=_np(ml.gavs_t),
               dof_pos=_np(ml.dof_pos), dvs=_np(ml.dvs),
               num_frames=_np(ml._motion_num_frames), lengths=_np(ml._motion_lengths), dt=_np(ml._motion_dt),
               starts=_np(ml.length_starts))
    if ml.has_contact_mask:
        out["contact"] = _np(ml._motion_contact_masks)
    return out


def _queries(ml, n, gen, two_d=0):
    nm = ml.num_motions()
    ids = torch.randint(0, nm, (n,), generator=gen)
    lens = ml._motion_lengths[ids]
    t = torch.rand(n, generator=gen) * lens * 1.1 - 0.03 * lens
    # edge cases: t<0, t==0, t==len, t>len, t just below len, exact frame times
    edge = torch.tensor([-0.5, 0.0, 1.0, 7.5, 0.999999, 1.0 / 30, 2.0 / 30, 0.033333335])
    k = min(len(edge), n)
    t[:k] = torch.where(edge[:k] == 1.0, lens[:k], torch.where(edge[:k] == 0.999999, lens[:k] * 0.999999, edge[:k]))
    off = torch.randn(n, 3, generator=gen) * 5.0
    if two_d:
        steps = torch.arange(two_d, dtype=torch.float32) * 0.1
        ids = ids[:, None].expand(n, two_d).contiguous()
        t = t[:, None] + steps[None]
        off = off[:, None, :].expand(n, two_d, 3).contiguous()
    return ids, t, off


def stage_lookup(ns):
    gen = torch.Generator().manual_seed(1234)
    # (a) 23-DoF, one clip (Horse-stance_pose), 3 env copies -> table golden + lookup golden
    cfg = R.motion_cfg(R.G1_23, R.example_motion("Horse-stance_pose"))
    ml = ns.motion_lib_robot.MotionLibRobot(cfg, num_envs=3, device="cpu")
    ml.load_motions(random_sample=False)
    tab = _table_arrays(ml)
    one = {k: (v[:210] if v.shape[0] == 630 else v) for k, v in tab.items()}
    np.savez_compressed(os.path.join(GOLD, "ref_table_Horse-stance_pose.npz"), **one)
    for tag, two_d in (("23_1d", 0), ("23_2d", 4)):
        ids, t, off = _queries(ml, 48, gen, two_d)
        res = ml.get_motion_state(ids, t.clone(), offset=off)
        np.savez_compressed(os.path.join(GOLD, f"ref_lookup_{tag}.npz"), motion_ids=_np(ids), motion_times=_np(t),
                            offset=_np(off), **{"out_" + k: _np(v) for k, v in res.items()})
    # (b) 29-DoF clip
    cfg = R.motion_cfg(R.G1_29, CLIP29)
    ml = ns.motion_lib_robot.MotionLibRobot(cfg, num_envs=2, device="cpu")
    ml.load_motions(random_sample=False)
    tab = _table_arrays(ml)
    nf = int(tab["num_frames"][0])
    one = {k: (v[:nf] if v.shape[0] == 2 * nf else v) for k, v in tab.items()}
    np.savez_compressed(os.path.join(GOLD, "ref_table_g1_29dof_clip.npz"), **one)
    ids, t, off = _queries(ml, 32, gen)
    res = ml.get_motion_state(ids, t.clone(), offset=off)
    np.savez_compressed(os.path.join(GOLD, "ref_lookup_29_1d.npz"), motion_ids=_np(ids), motion_times=_np(t),
                        offset=_np(off), **{"out_" + k: _np(v) for k, v in res.items()})


def stage_heading(ns):
    """load_motions(target_heading=...) (motion_lib_base.py:447-460): the clip re-based so that its first frame heads along the
    target heading - root axis-angles and root translations rotated in fp64 (scipy) before FK"""
    gen = torch.Generator().manual_seed(4321)
    th = np.array([0.0, 0.0, np.sin(np.pi / 8), np.cos(np.pi / 8)])  # xyzw: 45 degrees of yaw
    cfg = R.motion_cfg(R.G1_23, R.example_motion("Horse-stance_pose"))
    ml = ns.motion_lib_robot.MotionLibRobot(cfg, num_envs=3, device="cpu")
    ml.load_motions(random_sample=False, target_heading=th)
    ids, t, off = _queries(ml, 48, gen)
    res = ml.get_motion_state(ids, t.clone(), offset=off)
    np.savez_compressed(os.path.join(GOLD, "ref_lookup_23_heading.npz"), target_heading=th, motion_ids=_np(ids), motion_times=_np(t),
                        offset=_np(off), **{"out_" + k: _np(v) for k, v in res.items()})


def stage_fk(ns):
    gen = torch.Generator().manual_seed(4321)
    for tag, robot in (("23", R.G1_23), ("29", R.G1_29)):
        cfg = R.motion_cfg(robot, R.example_motion())
        hb = ns.torch_humanoid_batch.Humanoid_Batch(cfg)
        J = len(hb._parents)
        B = 24
        pose = torch.randn(B, 1, J, 3, generator=gen) * 0.6
        pose[0] = 0.0                      # zero-angle axis-angle (small-angle branch)
        pose[1] = pose[1] * 1e-7           # below the 1e-6 threshold
        pose[2, 0, 0] = torch.tensor([0.0, 0.0, 3.14159])   # near-pi root rotation
        pose[3, 0, 0] = torch.tensor([3.0, 0.5, -0.2])
        trans = torch.randn(B, 1, 3, generator=gen)
        out = hb.fk_batch(pose, trans)
        np.savez_compressed(os.path.join(GOLD, f"ref_fk_{tag}.npz"), pose=_np(pose), trans=_np(trans),
                            pos=_np(out.global_translation_extend), rotmat=_np(out.global_rotation_mat_extend),
                            quat=_np(out.global_rotation_extend))


def main():
    stages = sys.argv[1:] or ["assets", "lookup", "fk", "env", "gae"]
    os.makedirs(GOLD, exist_ok=True)
    os.chdir(R.REF_ROOT)
    ns = R.load_modules()
    torch.set_num_threads(1)
    if "assets" in stages:
        stage_assets(ns)
    if "lookup" in stages:
        stage_lookup(ns)
    if "heading" in stages:
        stage_heading(ns)
    if "fk" in stages:
        stage_fk(ns)
    if "env" in stages or "gae" in stages:
        from tools import make_golden_env
        make_golden_env.run(ns, stages)
    print("golden written to", GOLD)


if __name__ == "__main__":
    main()

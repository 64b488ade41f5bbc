#!/usr/bin/env python
"""How far the motion table the PRODUCT builds (load_motions: pbhc_fk_exact -> pbhc_clips_prepare, all on the device) is
from the reference's table (tests/golden/ref_table_*.npz, dumped by the unmodified reference): per array max / p99 / p50
absolute error, the same relative to the vector norm of the element's body, and relative to the array's RMS.  Also runs K1
(get_motion_state) on both tables at the golden query times and reports the error the lookup outputs inherit.
Writes profiles/r02_table_error.json.   python tools/table_error.py   (needs a GPU)"""
import json
import os
import sys
import tempfile

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from benchkit.fixtures import gold, product_cfg  # noqa: E402
from pbhc_b200.motion_lib import MotionLibRobot  # noqa: E402


def stats(got, want, vec=None):
    got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
    err = np.abs(got - want).reshape(-1)
    out = {"max_abs": float(err.max()), "p99_abs": float(np.percentile(err, 99)), "p50_abs": float(np.percentile(err, 50)),
           "rms_of_reference": float(np.sqrt((want ** 2).mean()))}
    if vec:
        w = want.reshape(-1, vec)
        nrm = np.linalg.norm(w, axis=1, keepdims=True) + 1e-12
        rel = (np.abs(got.reshape(-1, vec) - w) / nrm).reshape(-1)
        out.update(max_rel_to_norm=float(rel.max()), p99_rel_to_norm=float(np.percentile(rel, 99)))
    return out


def main():
    dev, res = "cuda:0", {}
    for tag, name in (("23", "Horse-stance_pose"), ("29", "g1_29dof_clip")):
        ml = MotionLibRobot(product_cfg(tag, tempfile.mkdtemp(), [name]), 2, dev)
        ml.load_motions(random_sample=False)
        g = gold(f"ref_table_{name}.npz")
        J, Jb, D, Cm = ml._dims
        rec = ml._records.cpu().numpy()
        T = rec.shape[0]
        r = {"frames": T}
        r["rot"] = stats(rec[:, :4 * J].reshape(T, J, 4), g["grs_t"], 4)
        r["pos"] = stats(rec[:, 4 * J:7 * J].reshape(T, J, 3), g["gts_t"], 3)
        r["vel"] = stats(rec[:, 7 * J:10 * J].reshape(T, J, 3), g["gvs_t"], 3)
        r["ang_vel"] = stats(rec[:, 10 * J:13 * J].reshape(T, J, 3), g["gavs_t"], 3)
        r["dof_vel"] = stats(rec[:, 13 * J + D:13 * J + 2 * D], g["dvs"])
        r["dof_pos_bit_exact"] = bool(np.array_equal(rec[:, 13 * J:13 * J + D], g["dof_pos"]))
        res[name] = r
    os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
    with open(os.path.join(ROOT, "profiles", "r02_table_error.json"), "w") as f:
        json.dump(res, f, indent=1)
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()

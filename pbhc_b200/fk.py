"""Humanoid_Batch drop-in (reference: humanoidverse/utils/motion_lib/torch_humanoid_batch.py:42-290).

Same constructor, attributes and fk_batch() signature/keys; the arithmetic runs in the sm_100a
kernels pbhc_fk (csrc/fk.cu) and pbhc_clip_velocities (csrc/motion_prep.cu).  The MJCF is parsed
with the stdlib XML parser (the reference uses lxml + poselib's from_mjcf, :104-165).  Mesh loading
(load_mesh / mesh_fk, :292-407) is visualisation only and out of scope.
"""
import ctypes as C
import xml.etree.ElementTree as ET
from collections import OrderedDict
from pathlib import Path

import numpy as np
import torch

from . import _lib


class AttrDict(dict):
    """EasyDict stand-in: attribute access on a dict (fk_batch returns one)."""
    __getattr__ = dict.__getitem__
    __setattr__ = dict.__setitem__


def _cfg_get(cfg, key, default=None):
    if isinstance(cfg, dict):
        return cfg.get(key, default)
    return getattr(cfg, key, default)


def parse_mjcf(path):
    """Kinematic tree in DFS order (from_mjcf, torch_humanoid_batch.py:104-165)."""
    root = ET.parse(str(path)).getroot()
    world = root.find("worldbody")
    if world is None or world.find("body") is None:
        raise ValueError("MJCF parsed incorrectly please verify it.")
    names, parents, trans, rots, ranges = [], [], [], [], []
    body_to_joint = OrderedDict()

    def add(node, parent):
        idx = len(names)
        names.append(node.attrib.get("name"))
        parents.append(parent)
        trans.append(np.array(node.attrib.get("pos", "0 0 0").split(), dtype=float))
        rots.append(np.array(node.attrib.get("quat", "1 0 0 0").split(), dtype=float))
        joints = node.findall("joint")
        if len(joints) == 6:
            joints = joints[6:]
        for j in joints:
            if j.attrib.get("range") is not None:
                ranges.append(np.array(j.attrib["range"].split(), dtype=float))
            elif j.attrib.get("type") != "free":
                ranges.append(np.array([-np.pi, np.pi]))
        for j in node.findall("joint"):
            body_to_joint[names[idx]] = j.attrib.get("name")
        for child in node.findall("body"):
            add(child, idx)

    add(world.find("body"), -1)
    all_joints = world.findall(".//joint")
    if all_joints[0].attrib.get("type") == "free" or "type" not in all_joints[0].attrib:
        axis_joints = all_joints[1:] if all_joints[0].attrib.get("type") == "free" else all_joints
        has_free = True
    else:
        axis_joints, has_free = all_joints[6:], False
    dof_axis = [[int(float(v)) for v in j.attrib["axis"].split()] for j in axis_joints]
    actuator = root.find("actuator")
    motors = sorted(m.attrib["name"] for m in list(actuator)) if actuator is not None else []
    return dict(node_names=names, parent_indices=np.array(parents, dtype=np.int32),
                local_translation=np.array(trans, dtype=np.float32), local_rotation=np.array(rots, dtype=np.float32),
                joints_range=np.array(ranges), body_to_joint=body_to_joint, dof_axis=dof_axis, motors=motors,
                has_freejoint=has_free)


def quat_wxyz_to_mat(q):
    """quaternion_to_matrix (isaac_utils/rotations.py:519-550) on host constants, fp32 like the reference."""
    q = torch.as_tensor(q, dtype=torch.float32)
    r, i, j, k = torch.unbind(q, -1)
    two_s = 2.0 / (q * q).sum(-1)
    o = torch.stack((1 - two_s * (j * j + k * k), two_s * (i * j - k * r), two_s * (i * k + j * r),
                     two_s * (i * j + k * r), 1 - two_s * (i * i + k * k), two_s * (j * k - i * r),
                     two_s * (i * k - j * r), two_s * (j * k + i * r), 1 - two_s * (i * i + j * j)), -1)
    return o.reshape(q.shape[:-1] + (3, 3))


class Humanoid_Batch:
    def __init__(self, cfg, device=torch.device("cpu")):
        self.cfg = cfg
        asset = _cfg_get(cfg, "asset")
        self.asset_root = Path(_cfg_get(asset, "assetRoot"))
        self.asset_file = _cfg_get(asset, "assetFileName")
        self.mjcf_file = self.asset_root / self.asset_file
        self.mjcf_data = d = parse_mjcf(self.mjcf_file)
        assert len(d["motors"]) > 0, "No motors found in the mjcf file"
        self.num_dof = len(d["motors"])
        self.num_extend_dof = self.num_dof
        assert len(d["joints_range"]) == self.num_dof
        self.body_names = list(d["node_names"])
        self.body_names_augment = list(d["node_names"])
        parents = [int(p) for p in d["parent_indices"]]
        offsets = [list(map(float, t)) for t in d["local_translation"]]
        lrot = [list(map(float, r)) for r in d["local_rotation"]]
        self.actuated_joints_idx = np.array([self.body_names.index(k) for k in d["body_to_joint"]])
        self.has_freejoint = d["has_freejoint"]
        self.dof_axis = torch.tensor(d["dof_axis"])
        self.extend_config = list(_cfg_get(cfg, "extend_config", []) or [])
        for ext in self.extend_config:
            self.body_names_augment.append(_cfg_get(ext, "joint_name"))
            parents.append(self.body_names.index(_cfg_get(ext, "parent_name")))
            offsets.append([float(v) for v in _cfg_get(ext, "pos")])
            lrot.append([float(v) for v in _cfg_get(ext, "rot")])
            self.num_extend_dof += 1
        self.num_bodies = len(self.body_names)
        self.num_bodies_augment = len(self.body_names_augment)
        self._parents = torch.tensor(parents, dtype=torch.int32)
        self._offsets = torch.tensor(offsets, dtype=torch.float32)[None]
        self._local_rotation = torch.tensor(lrot, dtype=torch.float32)[None]
        self._local_rotation_mat = quat_wxyz_to_mat(self._local_rotation).float()
        self.joints_range = torch.from_numpy(d["joints_range"])
        # host-side constant arrays handed to the C ABI (PbhcTree)
        self._c_parents = (C.c_int32 * len(parents))(*parents)
        self._c_offsets = (C.c_float * (3 * len(parents)))(*self._offsets.reshape(-1).tolist())
        self._c_lrot = (C.c_float * (9 * len(parents)))(*self._local_rotation_mat.reshape(-1).tolist())
        self._c_tree = _lib.Tree(len(parents), self._c_parents, self._c_offsets, self._c_lrot)

    # ------------------------------------------------------------------------------------------
    def fk(self, pose_aa, trans, want_mat=True, want_quat=True, out=None, exact=False):
        """pose_aa [n,J,3], trans [n,3] CUDA fp32 -> (pos [n,J,3], rotmat [n,J,3,3] | None, quat_xyzw [n,J,4] | None).
        `out` = a previous result tuple to overwrite instead of allocating.  exact: the reference's operation sequence step
        for step (pbhc_fk_exact) - what the motion table is built with; the default is the faster kernel within the same
        1e-5 rel / 1e-6 abs tolerance."""
        J = self.num_bodies_augment
        n = pose_aa.shape[0]
        pose_aa = pose_aa.contiguous()
        trans = trans.contiguous()
        dev = pose_aa.device
        if out is not None:
            pos, mat, quat = out
        else:
            pos = torch.empty(n, J, 3, device=dev, dtype=torch.float32)
            mat = torch.empty(n, J, 3, 3, device=dev, dtype=torch.float32) if want_mat else None
            quat = torch.empty(n, J, 4, device=dev, dtype=torch.float32) if want_quat else None
        fn = _lib.lib().pbhc_fk_exact if exact else _lib.lib().pbhc_fk
        _lib.check(fn(C.byref(self._c_tree), _lib.ptr(pose_aa, torch.float32),
                      _lib.ptr(trans, torch.float32), n, _lib.ptr(pos), _lib.ptr(mat), _lib.ptr(quat), _lib.stream_ptr(dev)))
        return pos, mat, quat

    def clip_kinematics(self, pose_aa, trans, dt, contact=None):
        """FK + filtered velocities + dof_pos/dof_vel of one clip, packed as table records
        [T, rec_stride] (layout: include/pbhc_b200.h).  pose_aa [T,J,3], trans [T,3] CUDA fp32."""
        pose_aa = pose_aa.contiguous()
        pos, _, quat = self.fk(pose_aa, trans, want_mat=False, want_quat=True, exact=True)
        return self.pack_clip(pos, quat, pose_aa, dt, contact)

    def pack_clip(self, pos, quat, pose_aa, dt, contact=None):
        """pbhc_clip_prepare: world pos [T,J,3] / quat xyzw [T,J,4] (FK output) + pose_aa -> records."""
        from .motion_lib import gaussian_weights, record_stride
        J, Jb = self.num_bodies_augment, self.num_bodies
        D = Jb - 1
        T = pose_aa.shape[0]
        dev = pose_aa.device
        Cm = 0 if contact is None else contact.shape[1]
        stride = record_stride(J, D, Cm)
        radius, w = gaussian_weights()
        w_c = (C.c_double * len(w))(*w.tolist())
        scratch = torch.empty(T, J, 6, device=dev, dtype=torch.float32)
        records = torch.empty(T, stride, device=dev, dtype=torch.float32)
        _lib.check(_lib.lib().pbhc_clip_prepare(_lib.ptr(pos.contiguous()), _lib.ptr(quat.contiguous()),
                                                _lib.ptr(pose_aa.contiguous()),
                                                None if contact is None else _lib.ptr(contact.contiguous()), T, J, Jb, D, Cm,
                                                dt, w_c, radius, _lib.ptr(scratch), _lib.ptr(records), stride,
                                                _lib.stream_ptr(dev)))
        return records

    def fk_batch(self, pose, trans, convert_to_mat=True, return_full=False, dt=1 / 30, exact=False):
        """torch_humanoid_batch.py:168-226: same keys.  pose [B,T,J,3] axis-angle, trans [B,T,3] (CUDA).  exact: see fk()."""
        if not convert_to_mat:
            raise NotImplementedError("fk_batch(convert_to_mat=False) is not part of the B200 path")
        if not pose.is_cuda:
            raise _lib.PbhcError("fk_batch needs CUDA tensors (there is no CPU fallback)")
        J, Jb = self.num_bodies_augment, self.num_bodies
        B, T = pose.shape[:2]
        p = pose[..., :J, :].to(torch.float32).reshape(B * T, J, 3)
        t = trans.to(torch.float32).reshape(B * T, 3)
        pos, mat, quat = self.fk(p, t, exact=exact or return_full)
        pos, mat, quat = pos.view(B, T, J, 3), mat.view(B, T, J, 3, 3), quat.view(B, T, J, 4)
        r = AttrDict()
        ext = len(self.extend_config) > 0
        if return_full:
            recs = torch.stack([self.clip_kinematics(p.view(B, T, J, 3)[b], t.view(B, T, 3)[b], dt) for b in range(B)])
            vel = recs[..., 7 * J:10 * J].reshape(B, T, J, 3)
            ang = recs[..., 10 * J:13 * J].reshape(B, T, J, 3)
            if ext:
                r.global_velocity_extend, r.global_angular_velocity_extend = vel, ang
        if ext:
            r.global_translation_extend, r.global_rotation_mat_extend, r.global_rotation_extend = pos, mat, quat
        nb = Jb if ext else J
        r.global_translation, r.global_rotation_mat, r.global_rotation = pos[..., :nb, :], mat[..., :nb, :, :], quat[..., :nb, :]
        if return_full:
            D = Jb - 1
            r.global_velocity, r.global_angular_velocity = vel[..., :nb, :], ang[..., :nb, :]
            r.global_root_velocity, r.global_root_angular_velocity = vel[..., 0, :], ang[..., 0, :]
            r.dof_pos = recs[..., 13 * J:13 * J + D]
            r.dof_vels = recs[..., 13 * J + D:13 * J + 2 * D]
            r.fps = int(1 / dt)
        return r

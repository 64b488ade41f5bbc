"""MotionLibRobot / MotionLibRobotWJX drop-ins (reference: humanoidverse/utils/motion_lib/
motion_lib_base.py:42-522, motion_lib_robot.py:3-7, motion_lib_robot_WJX.py).

Same constructor, methods, attributes and get_motion_state() dict; underneath ONE packed device
table holds every UNIQUE sampled clip once (the reference concatenates one copy per env,
motion_lib_base.py:356-372) and the sm_100a kernels pbhc_fk / pbhc_clip_prepare / pbhc_lookup do
the work.  No CPU fallback: construction requires a CUDA device.
"""
import ctypes as C
import glob
import os
import os.path as osp
import random

import numpy as np
import torch

from . import _lib
from .fk import Humanoid_Batch, _cfg_get

GAUSS_SIGMA, GAUSS_TRUNCATE = 2.0, 4.0


def gaussian_weights(sigma=GAUSS_SIGMA, truncate=GAUSS_TRUNCATE):
    """scipy.ndimage._gaussian_kernel1d(sigma, 0, radius) with radius = int(truncate*sigma + 0.5)."""
    radius = int(truncate * float(sigma) + 0.5)
    x = np.arange(-radius, radius + 1)
    phi = np.exp(-0.5 / (sigma * sigma) * x ** 2)
    return radius, phi / phi.sum()


def record_stride(J, D, Cm):
    return (13 * J + 2 * D + Cm + 3) // 4 * 4


class MotionLibRobot:
    def __init__(self, motion_lib_cfg, num_envs, device):
        self.m_cfg = motion_lib_cfg
        self._sim_fps = 1 / _cfg_get(motion_lib_cfg, "step_dt", 1 / 50)
        self.num_envs = num_envs
        self._device = torch.device(device)
        if self._device.type != "cuda":
            raise _lib.PbhcError("pbhc_b200.MotionLibRobot needs a CUDA device (there is no CPU fallback)")
        _lib.lib()
        self.has_action = False
        self.has_contact_mask = None
        self._contact_size = 0
        self.mesh_parsers = Humanoid_Batch(motion_lib_cfg)
        self.num_joints = self.mesh_parsers.num_bodies  # len(skeleton_tree.node_names), motion_lib_base.py:286
        self.load_data(_cfg_get(motion_lib_cfg, "motion_file"))
        self.setup_constants()

    # ---- motion_lib_base.py:63-105 -----------------------------------------------------------
    def load_data(self, motion_file):
        import joblib
        if isinstance(motion_file, dict):
            self._motion_data_load = motion_file
            values, keys = list(motion_file.values()), list(motion_file.keys())
        elif osp.isfile(motion_file):
            self._motion_data_load = joblib.load(motion_file)
            values, keys = list(self._motion_data_load.values()), list(self._motion_data_load.keys())
        else:
            files = glob.glob(osp.join(motion_file, "*.pkl"))
            values, keys = files, files
            self._motion_data_load = joblib.load(files[0])
        self._motion_data_list = np.empty(len(values), dtype=object)
        for i, v in enumerate(values):
            self._motion_data_list[i] = v
        self._motion_data_keys = np.array(keys)
        self._num_unique_motions = len(values)

    def setup_constants(self):
        dev = self._device
        self._curr_motion_ids = None
        self._termination_history = torch.zeros(self._num_unique_motions, device=dev)
        self._success_rate = torch.zeros(self._num_unique_motions, device=dev)
        self._sampling_history = torch.zeros(self._num_unique_motions, device=dev)
        self._sampling_prob = torch.ones(self._num_unique_motions, device=dev) / self._num_unique_motions

    # ---- motion_lib_base.py:261-391, 403-472 --------------------------------------------------
    @staticmethod
    def _rebase_heading(pose_aa, trans, target_heading):
        """load_motion_with_skeleton with target_heading (motion_lib_base.py:447-460): the whole clip is rotated about z so that
        its FIRST frame heads along `target_heading` (xyzw).  Host-side fp64 file preprocessing like the reference's (scipy
        Rotation); the rotated root axis-angles and translations then take the usual device FK.
        heading_inv = calc_heading_quat_inv(start root rotation) (rotations.py:256-268, 296-306): the x axis rotated by the
        root, its atan2 heading, the rotation by -heading about z."""
        from scipy.spatial.transform import Rotation as sRot
        start = sRot.from_rotvec(np.asarray(pose_aa[0, 0], dtype=np.float64))
        d = start.apply(np.array([1.0, 0.0, 0.0]))
        heading = np.arctan2(d[1], d[0])
        heading_inv = sRot.from_rotvec(np.array([0.0, 0.0, -heading]))
        delta = sRot.from_quat(np.asarray(target_heading, dtype=np.float64)) * heading_inv
        pose_aa = np.array(pose_aa, dtype=np.float32, copy=True)
        pose_aa[:, 0] = (delta * sRot.from_rotvec(pose_aa[:, 0].astype(np.float64))).as_rotvec().astype(np.float32)
        trans = np.asarray(trans, dtype=np.float64) @ delta.as_matrix().squeeze().T
        return pose_aa, trans.astype(np.float32)

    def _clip_arrays(self, entry, max_len, target_heading=None):
        if not isinstance(entry, dict):
            key = entry.split("/")[-1].split(".")[0]
            import joblib
            entry = joblib.load(entry)[key]
        seq_len = entry["root_trans_offset"].shape[0]
        if max_len == -1 or seq_len < max_len:
            start, end = 0, seq_len
        else:
            start = random.randint(0, seq_len - max_len)
            end = start + max_len
        out = dict(trans=np.asarray(entry["root_trans_offset"])[start:end], pose_aa=np.asarray(entry["pose_aa"])[start:end],
                   fps=entry["fps"], contact=None)
        if "action" in entry:
            self.has_action = True
        if "contact_mask" in entry:
            cm = np.asarray(entry["contact_mask"])
            assert cm.ndim == 2 and cm.shape[0] == seq_len
            if cm.shape[1] != 2:
                raise ValueError(f"Contact mask shape {cm.shape} is not supported")
            self._contact_size = cm.shape[1]
            self.has_contact_mask = "point"
            out["contact"] = cm[start:end]
        if target_heading is not None:
            out["pose_aa"], out["trans"] = self._rebase_heading(out["pose_aa"], out["trans"], target_heading)
        return out

    def load_motions(self, random_sample=True, start_idx=0, max_len=-1, target_heading=None, sampling_prob=None):
        dev = self._device
        n_load = self.num_envs
        if random_sample:
            prob = self._sampling_prob if sampling_prob is None else sampling_prob
            sample_idxes = torch.multinomial(prob, num_samples=n_load, replacement=True).to(dev)
        else:
            sample_idxes = torch.remainder(torch.arange(n_load) + start_idx, self._num_unique_motions).to(dev)
        self._curr_motion_ids = sample_idxes
        idx_host = sample_idxes.cpu().numpy()
        self.curr_motion_keys = self._motion_data_keys[idx_host]
        # unique clips only (a random crop makes every sample its own clip)
        if max_len == -1:
            uniq, clip_of_motion = np.unique(idx_host, return_inverse=True)
        else:
            uniq, clip_of_motion = idx_host, np.arange(n_load)
        clips = [self._clip_arrays(self._motion_data_list[u], max_len, target_heading) for u in uniq]
        self._build_table(clips)
        com = torch.as_tensor(clip_of_motion, dtype=torch.int64, device=dev)
        self._clip_of_motion = com.to(torch.int32).contiguous()
        # per-loaded-motion tables, as the reference exposes them
        self._motion_lengths = self._clip_length[com]
        self._motion_fps = self._clip_fps[com]
        self._motion_dt = self._clip_dt[com]
        self._motion_num_frames = self._clip_num_frames.to(torch.int64)[com]
        self._motion_bodies = torch.zeros(n_load, 17, device=dev)
        self._num_motions = n_load
        shifted = self._motion_num_frames.roll(1)
        shifted[0] = 0
        self.length_starts = shifted.cumsum(0)  # reference layout (one copy per env); informational
        self.motion_ids = torch.arange(n_load, dtype=torch.long, device=dev)
        self.num_bodies = self.num_joints
        return clips

    def _build_table(self, clips):
        dev = self._device
        hb = self.mesh_parsers
        J, Jb, D = hb.num_bodies_augment, hb.num_bodies, hb.num_bodies - 1
        Cm = self._contact_size if self.has_contact_mask else 0
        stride = record_stride(J, D, Cm)
        nf = [int(c["pose_aa"].shape[0]) for c in clips]
        total = int(sum(nf))
        records = torch.empty(total, stride, device=dev, dtype=torch.float32)
        radius, w = gaussian_weights()
        w_c = (C.c_double * len(w))(*w.tolist())
        L, st = _lib.lib(), _lib.stream_ptr(dev)
        # every frame of every clip in ONE host->device copy, ONE FK launch and ONE velocity / pack launch pair
        # (pbhc_clips_prepare); a 1024-clip library used to cost ~5000 tiny launches and copies
        pose_h = np.concatenate([np.asarray(c["pose_aa"][:, :J], dtype=np.float32) for c in clips], axis=0)
        trans_h = np.concatenate([np.asarray(c["trans"], dtype=np.float32) for c in clips], axis=0)
        pose = torch.from_numpy(np.ascontiguousarray(pose_h)).to(dev)
        trans = torch.from_numpy(np.ascontiguousarray(trans_h)).to(dev)
        contact = None
        if Cm:
            contact = torch.from_numpy(np.ascontiguousarray(np.concatenate([np.asarray(c["contact"], dtype=np.float32) for c in clips], axis=0))).to(dev)
        fps_l = [int(1 / (1 / c["fps"])) for c in clips]  # torch_humanoid_batch.py:225
        dt_l = [1.0 / f for f in fps_l]
        len_l = [1.0 / f * (T - 1) for f, T in zip(fps_l, nf)]
        starts_h = np.concatenate([[0], np.cumsum(nf)[:-1]]).astype(np.int64)
        clip_of_frame = torch.from_numpy(np.repeat(np.arange(len(nf), dtype=np.int32), nf)).to(dev)
        clip_start = torch.from_numpy(starts_h).to(dev)
        clip_frames = torch.tensor(nf, dtype=torch.int32, device=dev)
        clip_dt_prep = torch.tensor([1 / c["fps"] for c in clips], dtype=torch.float32, device=dev)  # the dt fk_batch is called with
        pos, _, quat = hb.fk(pose, trans, want_mat=False, want_quat=True, exact=True)
        scratch = torch.empty(total, J, 6, device=dev, dtype=torch.float32)
        _lib.check(L.pbhc_clips_prepare(_lib.ptr(pos), _lib.ptr(quat), _lib.ptr(pose), _lib.ptr(contact), total, clip_of_frame.data_ptr(),
                                        clip_start.data_ptr(), clip_frames.data_ptr(), clip_dt_prep.data_ptr(), J, Jb, D, Cm, w_c, radius,
                                        _lib.ptr(scratch), records.data_ptr(), stride, st))
        self._records = records
        self._table_version = getattr(self, "_table_version", 0) + 1  # (derived tables - TrackingStep's compact future-window records - follow it)
        self._rec_stride = stride
        self._clip_num_frames = torch.tensor(nf, dtype=torch.int32, device=dev)
        starts = np.concatenate([[0], np.cumsum(nf)[:-1]]).astype(np.int64)
        self._clip_frame_start = torch.as_tensor(starts, device=dev)
        self._clip_length = torch.tensor(len_l, dtype=torch.float32, device=dev)
        self._clip_dt = torch.tensor(dt_l, dtype=torch.float32, device=dev)
        self._clip_fps = torch.tensor(fps_l, dtype=torch.float32, device=dev)
        self._desc = _lib.TableDesc(records.data_ptr(), self._clip_num_frames.data_ptr(), self._clip_frame_start.data_ptr(),
                                    self._clip_length.data_ptr(), self._clip_dt.data_ptr(), total, len(nf), J, Jb, D, Cm, stride)
        self._dims = (J, Jb, D, Cm)

    def motion_data(self, motion=0):
        """The per-motion dict the reference's load_motions returns (Humanoid_Batch.fk_batch(return_full=True) keys,
        torch_humanoid_batch.py:189-229), as views of the packed table rows of loaded motion `motion`: what
        measure_traj.get_motionlib_data hands to blend_motion / eval_accuracy (`load_motions(...)[0]` there).
        Rotations are xyzw quaternions; the 3x3 `*_rotation_mat*` keys are not materialised."""
        J, Jb, D, Cm = self._dims
        clip = int(self._clip_of_motion[motion])
        lo, T = int(self._clip_frame_start[clip]), int(self._clip_num_frames[clip])
        rec = self._records[lo:lo + T]
        rot, pos = rec[:, :4 * J].reshape(T, J, 4), rec[:, 4 * J:7 * J].reshape(T, J, 3)
        vel, ang = rec[:, 7 * J:10 * J].reshape(T, J, 3), rec[:, 10 * J:13 * J].reshape(T, J, 3)
        out = {"global_translation_extend": pos, "global_rotation_extend": rot, "global_velocity_extend": vel,
               "global_angular_velocity_extend": ang, "global_translation": pos[:, :Jb], "global_rotation": rot[:, :Jb],
               "global_velocity": vel[:, :Jb], "global_angular_velocity": ang[:, :Jb],
               "global_root_velocity": vel[:, 0], "global_root_angular_velocity": ang[:, 0],
               "dof_pos": rec[:, 13 * J:13 * J + D], "dof_vels": rec[:, 13 * J + D:13 * J + 2 * D],
               "fps": int(self._clip_fps[clip])}
        if Cm:
            out["contact_mask"] = rec[:, 13 * J + 2 * D:13 * J + 2 * D + Cm]
        return out

    # ---- small accessors (motion_lib_base.py:474-501) ------------------------------------------
    def num_motions(self):
        return self._num_motions

    def get_total_length(self):
        return sum(self._motion_lengths)

    def get_motion_num_steps(self, motion_ids=None):
        if motion_ids is None:
            return (self._motion_num_frames * self._sim_fps / self._motion_fps).ceil().int()
        return (self._motion_num_frames[motion_ids] * self._sim_fps / self._motion_fps).ceil().int()

    def sample_time(self, motion_ids, truncate_time=None):
        phase = torch.rand(motion_ids.shape, device=self._device)
        motion_len = self._motion_lengths[motion_ids]
        if truncate_time is not None:
            assert truncate_time >= 0.0
            motion_len -= truncate_time
        return phase * motion_len

    def get_motion_length(self, motion_ids=None):
        if motion_ids is None:
            return self._motion_lengths
        return self._motion_lengths[motion_ids]

    # ---- motion_lib_base.py:123-259 -------------------------------------------------------------
    def get_motion_state(self, motion_ids, motion_times, offset=None, with_indices=False):
        J, Jb, D, Cm = self._dims
        dev = self._device
        shape = tuple(motion_ids.shape)
        n = int(np.prod(shape)) if len(shape) else 1
        ids = motion_ids.to(device=dev, dtype=torch.int64).contiguous()
        times = motion_times.to(device=dev, dtype=torch.float32).contiguous()
        off = None if offset is None else offset.to(device=dev, dtype=torch.float32).contiguous()
        f = lambda *s: torch.empty(shape + s, device=dev, dtype=torch.float32)
        res = {"root_pos": f(3), "root_rot": f(4), "dof_pos": f(D), "root_vel": f(3), "root_ang_vel": f(3),
               "dof_vel": f(D), "motion_aa": f(3 * Jb), "motion_bodies": f(17), "rg_pos": f(Jb, 3), "rb_rot": f(Jb, 4),
               "body_vel": f(Jb, 3), "body_ang_vel": f(Jb, 3), "rg_pos_t": f(J, 3), "rg_rot_t": f(J, 4),
               "body_vel_t": f(J, 3), "body_ang_vel_t": f(J, 3)}
        if Cm:
            res["contact_mask"] = f(Cm)
        extra = {}
        if with_indices:
            extra = {"f0l": torch.empty(shape, device=dev, dtype=torch.int64),
                     "f1l": torch.empty(shape, device=dev, dtype=torch.int64), "blend": f()}
        out = _lib.LookupOut()
        for k in _lib.LOOKUP_FIELDS:
            t = res.get(k, extra.get(k))
            setattr(out, k, None if t is None else t.data_ptr())
        _lib.check(_lib.lib().pbhc_lookup(C.byref(self._desc), self._clip_of_motion.data_ptr(), ids.data_ptr(),
                                          times.data_ptr(), None if off is None else off.data_ptr(), n, C.byref(out),
                                          _lib.stream_ptr(dev)))
        res["dof_vel"] = res["dof_vel"].reshape(shape[0], int(np.prod(shape[1:])) * D) if len(shape) else res["dof_vel"]  # :246
        if Cm:  # the reference puts contact_mask first (:231-238); dict order is not part of the contract
            pass
        res.update({"_" + k: v for k, v in extra.items()})
        return res


MotionLibRobotWJX = MotionLibRobot  # the single-clip "FakeCat" variant (motion_lib_robot_WJX.py) needs no special case here

"""ctypes binding of include/pbhc_b200.h.  There is NO fallback: if the CUDA library is missing or a
call fails, an exception is raised (the product path never routes through oracle/ or torch math)."""
import ctypes as C
import os

import torch  # noqa: F401  (loads libcudart.so.12 so the library resolves against torch's runtime)

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PBHC_B200_LIB") or os.path.join(_HERE, "libpbhc_b200.so")  # override: development A/B builds

c_f32p = C.c_void_p
ABI_VERSION = 11
MAX_JOINTS = 40


class PbhcError(RuntimeError):
    pass


class TableDesc(C.Structure):
    _fields_ = [("records", C.c_void_p), ("clip_num_frames", C.c_void_p), ("clip_frame_start", C.c_void_p),
                ("clip_length", C.c_void_p), ("clip_dt", C.c_void_p), ("total_frames", C.c_int64),
                ("num_clips", C.c_int32), ("num_joints", C.c_int32), ("num_bodies", C.c_int32),
                ("num_dof", C.c_int32), ("contact_size", C.c_int32), ("rec_stride", C.c_int32)]


LOOKUP_FIELDS = ["root_pos", "root_rot", "dof_pos", "root_vel", "root_ang_vel", "dof_vel", "motion_aa",
                 "motion_bodies", "rg_pos", "rb_rot", "body_vel", "body_ang_vel", "rg_pos_t", "rg_rot_t",
                 "body_vel_t", "body_ang_vel_t", "contact_mask", "f0l", "f1l", "blend"]


class LookupOut(C.Structure):
    _fields_ = [(k, C.c_void_p) for k in LOOKUP_FIELDS]


class GatherKey(C.Structure):
    _fields_ = [("src", C.c_void_p), ("dst", C.c_void_p), ("row_bytes", C.c_int64)]


MAX_GATHER_KEYS = 24
# PBHC_TM_* order of pbhc_traj_metrics: eval_accuracy keys (measure_traj.py:207-221) then eval_smoothness keys (:277-290)
TRAJ_ACCURACY = ["E_gmpbpe", "E_mpbpe", "E_mpjpe", "E_mpjve", "E_mpjae", "E_pbve", "E_pbae", "E_root_acc", "E_root_vel", "E_contact_acc"]
TRAJ_SMOOTHNESS = ["L2_vel", "L2_acc", "L2_jerk", "L2_dof_vel", "L2_dof_acc", "L2_dof_jerk", "L2_ref_vel", "L2_ref_acc", "L2_ref_jerk",
                   "L2_ref_dof_vel", "L2_ref_dof_acc", "L2_ref_dof_jerk"]
TRAJ_METRICS = len(TRAJ_ACCURACY) + len(TRAJ_SMOOTHNESS)
TRAJ_WORKSPACE = 592 * TRAJ_METRICS


class Tree(C.Structure):
    _fields_ = [("num_joints", C.c_int32), ("parents", C.POINTER(C.c_int32)), ("offsets", C.POINTER(C.c_float)),
                ("local_rot_mat", C.POINTER(C.c_float))]


_lib = None

EXPORTS = ["pbhc_abi_version", "pbhc_struct_size", "pbhc_last_error", "pbhc_lookup", "pbhc_fk", "pbhc_fk_exact", "pbhc_gae", "pbhc_adv_normalize", "pbhc_clip_prepare", "pbhc_clips_prepare",
           "pbhc_track_max_blocks", "pbhc_track_launch_info", "pbhc_track_specialize", "pbhc_track_step", "pbhc_track_finalize", "pbhc_gather_rows", "pbhc_gather_keys", "pbhc_timeout_bootstrap", "pbhc_blend_frames", "pbhc_traj_metrics", "pbhc_pre_physics", "pbhc_compute_torques",
           "pbhc_compact_scratch_ints", "pbhc_compact_ids", "pbhc_compact_ids_static", "pbhc_peer_allreduce_sum",
           "pbhc_future_table_stride", "pbhc_future_table_build"]


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise PbhcError(f"{LIB_PATH} is missing: run `python -m pbhc_b200.build` (there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    L.pbhc_abi_version.restype = C.c_int
    L.pbhc_last_error.restype = C.c_char_p
    L.pbhc_lookup.restype = C.c_int
    L.pbhc_lookup.argtypes = [C.POINTER(TableDesc), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                              C.POINTER(LookupOut), C.c_void_p]
    L.pbhc_fk.restype = C.c_int
    L.pbhc_fk_exact.restype = C.c_int
    L.pbhc_fk_exact.argtypes = [C.POINTER(Tree), C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.pbhc_fk.argtypes = [C.POINTER(Tree), C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                          C.c_void_p]
    L.pbhc_gae.restype = C.c_int
    L.pbhc_gae.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int64, C.c_int32,
                           C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.pbhc_adv_normalize.restype = C.c_int
    L.pbhc_adv_normalize.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
    L.pbhc_gather_rows.restype = C.c_int
    L.pbhc_gather_rows.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
    L.pbhc_gather_keys.restype = C.c_int
    L.pbhc_gather_keys.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p]
    L.pbhc_blend_frames.restype = C.c_int
    L.pbhc_blend_frames.argtypes = [C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_void_p, C.c_int64, C.c_float, C.c_void_p, C.c_void_p]
    L.pbhc_traj_metrics.restype = C.c_int
    L.pbhc_traj_metrics.argtypes = [C.c_void_p] * 6 + [C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_float, C.c_void_p, C.c_int64,
                                                        C.c_void_p, C.c_void_p]
    L.pbhc_pre_physics.restype = C.c_int
    L.pbhc_pre_physics.argtypes = [C.c_void_p] * 7 + [C.c_int64, C.c_void_p]
    L.pbhc_compute_torques.restype = C.c_int
    L.pbhc_compute_torques.argtypes = [C.c_void_p] * 4 + [C.c_int32] + [C.c_void_p] * 6 + [C.c_uint32, C.c_void_p, C.c_int64, C.c_void_p]
    L.pbhc_timeout_bootstrap.restype = C.c_int
    L.pbhc_timeout_bootstrap.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_float, C.c_void_p]
    L.pbhc_clip_prepare.restype = C.c_int
    L.pbhc_clip_prepare.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32,
                                    C.c_int32, C.c_int32, C.c_float, C.POINTER(C.c_double), C.c_int32, C.c_void_p,
                                    C.c_void_p, C.c_int32, C.c_void_p]
    L.pbhc_clips_prepare.restype = C.c_int
    L.pbhc_clips_prepare.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_double), C.c_int32, C.c_void_p,
                                     C.c_void_p, C.c_int32, C.c_void_p]
    bind_track(L)
    if L.pbhc_abi_version() != ABI_VERSION:
        raise PbhcError("ABI version mismatch")
    L.pbhc_struct_size.restype = C.c_int64
    L.pbhc_struct_size.argtypes = [C.c_int32]
    for i, t in enumerate((TableDesc, LookupOut, Tree, TrackCfg, TrackIO, HistDst, Curriculum, GatherKey, ActuationCfg)):
        if L.pbhc_struct_size(i) != C.sizeof(t):
            raise PbhcError(f"ABI struct {t.__name__}: library {L.pbhc_struct_size(i)} B, binding {C.sizeof(t)} B")
    _lib = L
    return L


def check(rc):
    if rc != 0:
        raise PbhcError(f"pbhc error {rc}: {lib().pbhc_last_error().decode()}")


def ptr(t, dtype=None):
    """device pointer of a contiguous CUDA tensor (None -> NULL)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise PbhcError("expected a CUDA tensor (no CPU fallback)")
    if not t.is_contiguous():
        raise PbhcError("expected a contiguous tensor")
    if dtype is not None and t.dtype != dtype:
        raise PbhcError(f"expected dtype {dtype}, got {t.dtype}")
    return t.data_ptr()


def stream_ptr(device=None):
    return torch.cuda.current_stream(device).cuda_stream


# ---- K3 structs (mirror include/pbhc_b200.h exactly) ------------------------------------------------
MAX_DOF, MAX_REWARDS, MAX_EXT, MAX_FEET, MAX_OBS_GROUPS, MAX_HIST_KEYS, MAX_HIST_DST = 32, 32, 4, 4, 4, 8, 16
NUM_OBS_SRC, NUM_SIGMA, ENVF_STRIDE, SCALARS, ACC, MAX_KEY, MAX_FUTURE = 40, 17, 8, 112, 64, 16, 32
PEER_MAX_DOUBLES = 256
COMPACT_STATIC_MAX_N = 1 << 24  # PBHC_COMPACT_STATIC_MAX_N: flags per call of pbhc_compact_ids_static (its scratch is a 16 KB device array of the library)
L_ADP_SIGMA, L_ERROR_EMA = 16, 33
S_SIGMA, S_EMA, S_AVG_EPL, S_PENALTY, S_FAR_THR, S_SOFT_DOF_POS, S_SOFT_DOF_VEL, S_SOFT_TORQUE, S_PUSH_COUNT, S_LOG = 0, 17, 34, 35, 36, 37, 38, 39, 40, 48
S_NOISE_CUR, L_NOISE_CUR = 41, 63  # obs.add_noise_currculum: current_noise_curriculum_value and its log slot
LOG_NAMES = ["upper_body_diff_norm", "lower_body_diff_norm", "vr_3point_diff_norm", "joint_pos_diff_norm",
             "terminate_by_gravity", "terminate_by_motion_far", "terminate_by_time_out", "terminate_by_motion_end",
             "end_time_ratio", "end_time_ratio_std", "terminate_when_motion_far_threshold", "penalty_scale",
             "average_episode_length", "soft_dof_pos_curriculum_value", "soft_dof_vel_curriculum_value",
             "soft_torque_curriculum_value"]
# general_tracking logs: name -> PBHC_L_* index
# base-class terminations (legged_robot_base.py:433-476): name -> PBHC_L_* index, logged when enabled
LOG_NAMES_TERM = {"terminate_by_contact": 58, "terminate_by_low_height": 59, "terminate_by_dof_pos_limit": 60,
                  "terminate_by_dof_vel_limit": 61, "terminate_by_torque_limit": 62}
LOG_NAMES_GENERAL = {"key_body_diff_norm": 50, "local_upper_body_diff_norm": 51, "local_lower_body_diff_norm": 52,
                     "local_vr_3point_diff_norm": 53, "local_key_body_diff_norm": 54, "terminate_by_ref_pos_z": 55,
                     "terminate_by_ref_ori": 56, "terminate_by_body_z": 57}


class Curriculum(C.Structure):
    _fields_ = [("enable", C.c_int32), ("lo", C.c_double), ("hi", C.c_double), ("degree", C.c_double),
                ("down", C.c_double), ("up", C.c_double)]


class HistDst(C.Structure):
    _fields_ = [("key", C.c_int32), ("group", C.c_int32), ("offset", C.c_int32), ("len", C.c_int32), ("mul", C.c_float)]


class TrackCfg(C.Structure):
    _fields_ = [
        ("J", C.c_int32), ("Jb", C.c_int32), ("D", C.c_int32), ("L", C.c_int32), ("n_ext", C.c_int32), ("n_feet", C.c_int32),
        ("ext_parent", C.c_int32 * MAX_EXT), ("ext_pos", (C.c_float * 3) * MAX_EXT), ("ext_rot_xyzw", (C.c_float * 4) * MAX_EXT),
        ("feet_ids", C.c_int32 * MAX_FEET),
        ("upper_mask", C.c_uint64), ("lower_mask", C.c_uint64), ("vr3_mask", C.c_uint64), ("penalised_mask", C.c_uint64),
        ("dt", C.c_float), ("max_episode_length", C.c_int64),
        ("n_rewards", C.c_int32), ("use_vec_reward", C.c_int32), ("only_positive_rewards", C.c_int32), ("has_termination", C.c_int32),
        ("reward_id", C.c_int32 * MAX_REWARDS), ("reward_scale", C.c_float * MAX_REWARDS), ("reward_is_penalty", C.c_int32 * MAX_REWARDS),
        ("termination_scale", C.c_float),
        ("w_lower", C.c_float), ("w_upper", C.c_float), ("desired_feet_air_time", C.c_float), ("max_contact_force", C.c_float),
        ("hard_lo", C.c_float * MAX_DOF), ("hard_hi", C.c_float * MAX_DOF), ("soft_lo", C.c_float * MAX_DOF), ("soft_hi", C.c_float * MAX_DOF),
        ("vel_limit", C.c_float * MAX_DOF), ("torque_limit", C.c_float * MAX_DOF),
        ("soft_dof_vel_limit", C.c_float), ("soft_torque_limit", C.c_float),
        ("term_gravity", C.c_int32), ("term_motion_far", C.c_int32), ("term_motion_end", C.c_int32), ("term_gravity_thr", C.c_float),
        ("adaptive_enable", C.c_int32), ("adaptive_type", C.c_int32), ("adaptive_alpha", C.c_double), ("adaptive_scale", C.c_double),
        ("cur_penalty", Curriculum), ("cur_far", Curriculum), ("cur_dof_pos", Curriculum), ("cur_dof_vel", Curriculum), ("cur_torque", Curriculum),
        ("num_compute_average_epl", C.c_double),
        ("n_groups", C.c_int32), ("group_dim", C.c_int32 * MAX_OBS_GROUPS),
        ("src_dst", (C.c_int32 * MAX_OBS_GROUPS) * NUM_OBS_SRC), ("n_hist_dst", C.c_int32), ("hist_dst", HistDst * MAX_HIST_DST),
        ("obs_scale", C.c_float * NUM_OBS_SRC), ("obs_noise", C.c_float * NUM_OBS_SRC), ("clip_obs", C.c_float),
        ("n_hist_keys", C.c_int32), ("hist_row", C.c_int32),
        ("hist_src", C.c_int32 * MAX_HIST_KEYS), ("hist_len", C.c_int32 * MAX_HIST_KEYS), ("hist_dim", C.c_int32 * MAX_HIST_KEYS),
        ("randomize_pd_gain", C.c_int32), ("randomize_rfi_lim", C.c_int32), ("use_rao", C.c_int32), ("randomize_ctrl_delay", C.c_int32),
        ("delay_lo", C.c_int32), ("delay_hi", C.c_int32),
        ("kp_lo", C.c_float), ("kp_hi", C.c_float), ("kd_lo", C.c_float), ("kd_hi", C.c_float), ("rfi_lo", C.c_float), ("rfi_hi", C.c_float),
        ("rao_lim", C.c_float),
        ("seed", C.c_uint64), ("reset_noise", C.c_float * 6), ("src_used_mask", C.c_uint64),
        ("n_upper", C.c_int32), ("n_lower", C.c_int32), ("n_vr3", C.c_int32), ("n_episode_sums", C.c_int32),
        ("general", C.c_int32), ("anchor_index", C.c_int32), ("n_key", C.c_int32), ("key_ids", C.c_int32 * MAX_KEY),
        ("key_mask", C.c_uint64), ("body_z_mask", C.c_uint64),
        ("term_ref_pos_z", C.c_int32), ("term_ref_ori", C.c_int32), ("term_body_z", C.c_int32),
        ("term_ref_pos_z_thr", C.c_float), ("term_ref_ori_thr", C.c_float), ("term_body_z_thr", C.c_float),
        ("n_future", C.c_int32), ("future_steps", C.c_int32 * MAX_FUTURE),
        ("term_contact", C.c_int32), ("term_contact_mask", C.c_uint64), ("term_low_height", C.c_int32), ("term_min_height", C.c_float),
        ("term_dof_pos", C.c_int32), ("term_dof_vel", C.c_int32), ("term_torque", C.c_int32),
        ("term_dof_pos_prob", C.c_float), ("term_dof_vel_prob", C.c_float), ("term_torque_prob", C.c_float),
        ("term_dof_vel_scale", C.c_float), ("term_torque_scale", C.c_float),
        ("term_pos_lo", C.c_float * MAX_DOF), ("term_pos_hi", C.c_float * MAX_DOF),
        ("ref_init_yaw", C.c_float), ("vr3_ids", C.c_int32 * 4),
        ("push_robots", C.c_int32), ("push_fixed", C.c_int32), ("push_lo", C.c_int32), ("push_hi", C.c_int32), ("max_push_vel", C.c_float),
        ("start_time_zero", C.c_int32), ("cur_noise", Curriculum)]


TRACK_IO_FIELDS = [
    "body_state", "body_env_stride", "root_state", "dof_state", "contact_forces", "actions", "torques", "env_origins",
    "motion_ids", "clip_of_motion", "default_dof_pos", "default_dof_stride", "base_com_bias", "link_mass_scale", "friction",
    "base_mass_scale",
    "episode_length_buf", "env_f32", "last_actions", "last_dof_pos", "last_dof_vel", "kp_scale", "kd_scale",
    "rfi_lim_scale", "rao_scale", "action_delay_idx", "hist", "episode_sums", "scalars"]


class ActuationCfg(C.Structure):
    _fields_ = [("D", C.c_int32), ("Q", C.c_int32), ("action_clip", C.c_float), ("control_type", C.c_int32),
                ("per_dof_action_scale", C.c_int32), ("action_scale", C.c_float), ("action_scales", C.c_float * MAX_DOF),
                ("p_gains", C.c_float * MAX_DOF), ("d_gains", C.c_float * MAX_DOF), ("torque_limits", C.c_float * MAX_DOF),
                ("rfi", C.c_int32), ("rfi_lim", C.c_float), ("rao", C.c_int32), ("clip_torques", C.c_int32), ("sim_dt", C.c_float),
                ("seed", C.c_uint64), ("env_offset", C.c_int64), ("call_counter", C.c_void_p)]


class TrackIO(C.Structure):
    _fields_ = ([(k, C.c_int64 if k.endswith("_stride") else C.c_void_p) for k in TRACK_IO_FIELDS]
                + [("obs", C.c_void_p * MAX_OBS_GROUPS)]
                + [(k, C.c_void_p) for k in ("rew_buf", "reset_buf", "time_out_buf", "ref_body_pos_extend", "ref_body_rot_extend",
                                             "last_episode_sums", "partials", "acc", "inject_start_times", "inject_kp",
                                             "inject_kd", "inject_rfi", "inject_rao", "inject_delay")]
                + [("env_offset", C.c_int64), ("last_episode_length_buf", C.c_void_p), ("force_reset", C.c_void_p), ("inject_term_gates", C.c_int32),
                   ("fut_table", C.c_void_p), ("fut_stride", C.c_int32), ("heading_scratch", C.c_void_p),
                   ("refresh_buf", C.c_void_p), ("push_vel", C.c_void_p), ("inject_push_interval", C.c_void_p), ("inject_push_vel", C.c_void_p)])


def bind_track(L):
    L.pbhc_track_max_blocks.restype = C.c_int
    L.pbhc_track_step.restype = C.c_int
    L.pbhc_track_step.argtypes = [C.POINTER(TrackCfg), C.POINTER(TableDesc), C.POINTER(TrackIO), C.c_int64, C.c_void_p]
    L.pbhc_track_finalize.restype = C.c_int
    L.pbhc_track_finalize.argtypes = [C.POINTER(TrackCfg), C.c_void_p, C.c_void_p, C.c_void_p]
    L.pbhc_track_specialize.restype = C.c_int
    L.pbhc_track_specialize.argtypes = [C.POINTER(TrackCfg), C.POINTER(TableDesc), C.POINTER(TrackIO)]
    L.pbhc_track_launch_info.restype = C.c_int
    L.pbhc_track_launch_info.argtypes = [C.POINTER(C.c_int32 * 4)]
    L.pbhc_compact_scratch_ints.restype = C.c_int64
    L.pbhc_compact_scratch_ints.argtypes = [C.c_int64]
    L.pbhc_compact_ids.restype = C.c_int
    L.pbhc_compact_ids.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.pbhc_peer_allreduce_sum.restype = C.c_int
    L.pbhc_peer_allreduce_sum.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]
    L.pbhc_future_table_stride.restype = C.c_int32
    L.pbhc_future_table_stride.argtypes = [C.POINTER(TrackCfg)]
    L.pbhc_future_table_build.restype = C.c_int
    L.pbhc_future_table_build.argtypes = [C.POINTER(TrackCfg), C.POINTER(TableDesc), C.c_void_p, C.c_int32, C.c_void_p]
    L.pbhc_compact_ids_static.restype = C.c_int
    L.pbhc_compact_ids_static.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]

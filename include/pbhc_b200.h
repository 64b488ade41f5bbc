/* pbhc_b200.h - C ABI of the B200-native PBHC motion-tracking step.
 *
 * The reference (kyungminn/PBHC, humanoidverse) is pure Python and has no FFI layer; its boundary
 * for this path is the Python object API.  Each entry point below names the
 * reference function it replaces.  All pointers are DEVICE pointers unless stated; memory is
 * caller-owned (PyTorch tensors); no call allocates, synchronises the device or throws.
 * Return value: 0 = ok, <0 = PBHC_E_*.  pbhc_last_error() gives a thread-local message.
 * `stream` is a cudaStream_t passed as void*.
 */
#ifndef PBHC_B200_H
#define PBHC_B200_H
#ifndef __CUDACC_RTC__ /* NVRTC compiles the kernels without host headers; the fixed-width types are typedef'd by the includer */
#include <stdint.h>
#endif
#ifdef __cplusplus
extern "C" {
#endif

#define PBHC_ABI_VERSION 11
#define PBHC_OK 0
#define PBHC_E_BADARG (-1)
#define PBHC_E_ALIGN (-2)
#define PBHC_E_RANGE (-3)
#define PBHC_E_CUDA (-4)

int pbhc_abi_version(void);
/* sizeof of an ABI struct, for bindings to verify their mirror: 0 PbhcTableDesc, 1 PbhcLookupOut, 2 PbhcTree,
 * 3 PbhcTrackCfg, 4 PbhcTrackIO, 5 PbhcHistDst, 6 PbhcCurriculum, 7 PbhcGatherKey, 8 PbhcActuationCfg; -1 for an unknown id */
int64_t pbhc_struct_size(int32_t which);
const char* pbhc_last_error(void);

/* ---- packed motion table (replaces the 11 concatenated tensors of
 * utils/motion_lib/motion_lib_base.py:342-384).  One record per frame of each UNIQUE clip:
 *   [ rot_t J*4 | pos_t J*3 | vel_t J*3 | ang_vel_t J*3 | dof_pos D | dof_vel D | contact C | pad ]
 * rec_stride (floats) is a multiple of 4 so every record and every quaternion is 16-B aligned. */
typedef struct {
  const float* records;           /* [total_frames, rec_stride] */
  const int32_t* clip_num_frames; /* [num_clips]  (_motion_num_frames) */
  const int64_t* clip_frame_start;/* [num_clips]  (length_starts) */
  const float* clip_length;       /* [num_clips]  (_motion_lengths) */
  const float* clip_dt;           /* [num_clips]  (_motion_dt) */
  int64_t total_frames;
  int32_t num_clips;
  int32_t num_joints;  /* J  = bodies incl. extended */
  int32_t num_bodies;  /* Jb */
  int32_t num_dof;     /* D */
  int32_t contact_size;/* C (0 if the clips carry no contact mask) */
  int32_t rec_stride;  /* floats */
} PbhcTableDesc;

/* outputs of get_motion_state (motion_lib_base.py:239-259); every pointer may be NULL */
typedef struct {
  float* root_pos;      /* [n,3] */
  float* root_rot;      /* [n,4] xyzw */
  float* dof_pos;       /* [n,D] */
  float* root_vel;      /* [n,3] */
  float* root_ang_vel;  /* [n,3] */
  float* dof_vel;       /* [n,D] */
  float* motion_aa;     /* [n,3*Jb]  zeros for robot clips (motion_lib_base.py:318-323) */
  float* motion_bodies; /* [n,17]    zeros */
  float* rg_pos;        /* [n,Jb,3] */
  float* rb_rot;        /* [n,Jb,4] */
  float* body_vel;      /* [n,Jb,3] */
  float* body_ang_vel;  /* [n,Jb,3] */
  float* rg_pos_t;      /* [n,J,3] */
  float* rg_rot_t;      /* [n,J,4] */
  float* body_vel_t;    /* [n,J,3] */
  float* body_ang_vel_t;/* [n,J,3] */
  float* contact_mask;  /* [n,C] */
  int64_t* f0l;         /* [n] frame index in the table (idx0 + frame_start) */
  int64_t* f1l;         /* [n] */
  float* blend;         /* [n] */
} PbhcLookupOut;

/* MotionLibBase.get_motion_state + _calc_frame_blend (motion_lib_base.py:123-259, 503-513) and
 * slerp (isaac_utils/rotations.py:210-232).  clip_of_motion maps the reference's motion id (index
 * of the per-env sampled clip) to a unique clip; NULL = identity.  offset may be NULL. */
int pbhc_lookup(const PbhcTableDesc* tab, const int32_t* clip_of_motion, const int64_t* motion_ids,
                const float* motion_times, const float* offset, int64_t n_queries,
                const PbhcLookupOut* out, void* stream);

/* ---- kinematic tree constants (Humanoid_Batch.__init__, torch_humanoid_batch.py:44-102).
 * HOST pointers; copied into the launch parameters (J <= PBHC_MAX_JOINTS). */
#define PBHC_MAX_JOINTS 40
typedef struct {
  int32_t num_joints;              /* J (MJCF bodies + extended bodies, DFS order) */
  const int32_t* parents;          /* [J], -1 for the root; parents[i] < i */
  const float* offsets;            /* [J,3] local_translation */
  const float* local_rot_mat;      /* [J,9] row-major quaternion_to_matrix(local_rotation) */
} PbhcTree;

/* Humanoid_Batch.fk_batch / forward_kinematics_batch (torch_humanoid_batch.py:168-269) incl.
 * axis_angle_to_quaternion, quaternion_to_matrix, matrix_to_quaternion (rotations.py:519-647).
 * pose_aa [n,J,3], trans [n,3] -> pos [n,J,3], rotmat [n,J,9] (may be NULL), quat_xyzw [n,J,4]
 * (may be NULL). */
int pbhc_fk(const PbhcTree* tree, const float* pose_aa, const float* trans, int64_t n_poses,
            float* pos, float* rotmat, float* quat_xyzw, void* stream);
/* The same with the reference's operation sequence kept step for step (IEEE sqrt / division, separate sin and cos, all
 * four matrix_to_quaternion candidates): what load_motions uses, because the angular velocities of the motion table
 * amplify last-ulp differences of the frame quaternions.  pbhc_fk is within the same tolerance of the
 * reference (1e-5 rel / 1e-6 abs) with fewer instructions. */
int pbhc_fk_exact(const PbhcTree* tree, const float* pose_aa, const float* trans, int64_t n_poses,
                  float* pos, float* rotmat, float* quat_xyzw, void* stream);

/* ---- GAE: PPO._compute_returns (agents/ppo/ppo.py:338-355), MHPPO._compute_returns
 * (agents/mh_ppo/mh_ppo.py:376-395).  rewards/values/returns [T,N,R], dones [T,N] (bool bytes),
 * last_values [N,R].  adv_raw [T,N] = sum_R(returns - values).  stats3 (double[3], device) is
 * ACCUMULATED with (count, sum, sum of squares) of adv_raw: zero it before the call; allreduce it
 * across ranks before pbhc_adv_normalize. */
int pbhc_gae(const float* rewards, const float* values, const uint8_t* dones,
             const float* last_values, int32_t T, int64_t N, int32_t R, float gamma, float lam,
             float* returns, float* adv_raw, double* stats3, void* stream);

/* advantages = (adv - mean) / (std_unbiased + 1e-8), in place (ppo.py:354). */
int pbhc_adv_normalize(float* adv, int64_t n, const double* stats3_global, void* stream);

/* ---- RolloutStorage (agents/modules/data_utils.py:30-152) data movement, SURVEY.md 8f N3.
 * mini_batch_generator (data_utils.py:116-152) shuffles every stored key with
 * `buf.flatten(0, 1)[indices]`, indices = randperm(T*N): dst[r, :] = src[idx[r], :] for rows of
 * row_bytes bytes (any dtype; src has src_rows rows, both dense).  An index outside [0, src_rows)
 * skips the row and sets *bad_index_flag (device int32, zero it first) - torch raises IndexError
 * there; the Python mirror checks the flag. */
int pbhc_gather_rows(const void* src, int64_t src_rows, int64_t row_bytes, const int64_t* idx,
                     int64_t n_rows, void* dst, int32_t* bad_index_flag, void* stream);

/* The same gather for every stored key of the storage in ONE launch (the reference issues one
 * index_select per key, 11+ per PPO update): the index tile is read once and shared by all keys. */
#define PBHC_MAX_GATHER_KEYS 24
typedef struct PbhcGatherKey {
  const void* src;   /* [src_rows, row_bytes] dense */
  void* dst;         /* [n_rows, row_bytes] dense */
  int64_t row_bytes;
} PbhcGatherKey;
int pbhc_gather_keys(const PbhcGatherKey* keys, int32_t n_keys, const int64_t* idx, int64_t n_rows,
                     int64_t src_rows, int32_t* bad_index_flag, void* stream);

/* Time-out bootstrap of the stored rewards, PPO._rollout_step (agents/ppo/ppo.py:266-270):
 * rewards[n,R] += gamma * values[n,R] * time_outs[n] (time_outs = bool bytes), in place. */
int pbhc_timeout_bootstrap(float* rewards, const float* values, const uint8_t* time_outs,
                           int64_t n_envs, int32_t R, float gamma, void* stream);

/* ---- trajectory evaluation (humanoidverse/measure_traj.py), SURVEY.md 8f N4.
 * blend_motion (measure_traj.py:67-133): one key of a loaded motion, src [num_frames, row_floats],
 * resampled at times[n_out] (fp32) -> out [n_out, row_floats]: frame pair and blend of
 * _calc_frame_blend (:68-84, motion length = num_frames / fps), then `b*x1 + (1-b)*x0`, or slerp
 * over consecutive xyzw quaternions when quats != 0 (:121). */
int pbhc_blend_frames(const float* src, int64_t num_frames, int32_t row_floats, int32_t quats,
                      const float* times, int64_t n_out, float fps, float* out, void* stream);

/* eval_accuracy (measure_traj.py:146-225) + eval_smoothness (:227-290) in one pass: pol/ref
 * global_translation [T,J,3], dof_pos [T,D], optional contact_mask [T,C] (both or neither);
 * delta = fps, or 1 for delta_per_frame.  out[PBHC_TRAJ_METRICS] doubles on the device, indexed by
 * PBHC_TM_*; a mean over no terms is NaN like torch.  workspace: >= 592 * PBHC_TRAJ_METRICS doubles. */
#define PBHC_TRAJ_METRICS 22
enum {
  PBHC_TM_GMPBPE = 0, PBHC_TM_MPBPE, PBHC_TM_MPJPE, PBHC_TM_MPJVE, PBHC_TM_MPJAE, PBHC_TM_PBVE, PBHC_TM_PBAE,
  PBHC_TM_ROOT_ACC, PBHC_TM_ROOT_VEL, PBHC_TM_CONTACT,
  PBHC_TM_L2_VEL, PBHC_TM_L2_ACC, PBHC_TM_L2_JERK, PBHC_TM_L2_DOF_VEL, PBHC_TM_L2_DOF_ACC, PBHC_TM_L2_DOF_JERK,
  PBHC_TM_L2_REF_VEL, PBHC_TM_L2_REF_ACC, PBHC_TM_L2_REF_JERK, PBHC_TM_L2_REF_DOF_VEL, PBHC_TM_L2_REF_DOF_ACC,
  PBHC_TM_L2_REF_DOF_JERK
};
int pbhc_traj_metrics(const float* pol_pos, const float* ref_pos, const float* pol_dof, const float* ref_dof,
                      const float* pol_contact, const float* ref_contact, int64_t T, int32_t J, int32_t D,
                      int32_t C, float delta, double* workspace, int64_t workspace_doubles, double* out,
                      void* stream);

/* ---- load-time clip preparation: FK output -> packed table records.
 * Replaces Humanoid_Batch._compute_velocity / _compute_angular_velocity / dof_pos / dof_vels
 * (torch_humanoid_batch.py:216-224, 272-290; numpy.gradient + scipy gaussian_filter1d sigma=2,
 * mode="nearest") and the concatenation in MotionLibBase.load_motions (motion_lib_base.py:342-384).
 * pos [T,J,3], quat xyzw [T,J,4] (pbhc_fk output), pose_aa [T,J,3], contact [T,C] or NULL;
 * gauss_weights_host: HOST double[2*radius+1]; scratch: device float [T,J,6];
 * records: device float [T,rec_stride], written completely (padding zeroed). */
int pbhc_clip_prepare(const float* pos, const float* quat, const float* pose_aa, const float* contact,
                      int32_t T, int32_t J, int32_t Jb, int32_t D, int32_t C, float dt,
                      const double* gauss_weights_host, int32_t radius, float* scratch,
                      float* records, int32_t rec_stride, void* stream);

/* The same for the frames of MANY clips back to back (one launch pair for a whole library instead of one per clip):
 * pos / quat / pose_aa / contact / scratch / records hold total_frames frames; clip_of_frame [total_frames] int32 and the
 * per-clip first frame / frame count / dt are DEVICE arrays.  Identical results to per-clip pbhc_clip_prepare calls. */
int pbhc_clips_prepare(const float* pos, const float* quat, const float* pose_aa, const float* contact, int64_t total_frames,
                       const int32_t* clip_of_frame, const int64_t* clip_start, const int32_t* clip_frames, const float* clip_dt,
                       int32_t J, int32_t Jb, int32_t D, int32_t C, const double* gauss_weights_host, int32_t radius,
                       float* scratch, float* records, int32_t rec_stride, void* stream);

/* ---- K3: fused motion-tracking env step (lookup + diffs + termination + rewards + reset + observations +
 * history), i.e. everything LeggedRobotBase._post_physics_step does between refresh_sim_tensors and
 * extras["to_log"] for LeggedRobotMotionTracking or LeggedRobotGeneralTracking (envs/legged_base_task/legged_robot_base.py:297-338,
 * envs/motion_tracking/motion_tracking.py:330-358, 548-747, 1030-1328; envs/motion_tracking/general_tracking.py:
 * 226-262, 500-805, 1116-1270; utils/helpers.py:128-152;
 * envs/env_utils/history_handler.py:33-48).  One launch per step, no host synchronisation: scalars the
 * reference keeps on the host (adaptive sigmas, curricula, average episode length) live in `scalars`
 * (device doubles) and are advanced by pbhc_track_finalize from per-step reductions. */
#define PBHC_MAX_DOF 32
#define PBHC_MAX_REWARDS 32
#define PBHC_MAX_EXT 4
#define PBHC_MAX_FEET 4
#define PBHC_MAX_OBS_GROUPS 4
#define PBHC_MAX_HIST_KEYS 8
#define PBHC_MAX_HIST_DST 16
#define PBHC_NUM_OBS_SRC 40  /* 20 shared + 15 of LeggedRobotGeneralTracking + 5 added with ABI 9 (relyaw, local_ref_rigid_body_pos_relyaw,
                              * vr_3point_pos, future_ref_dof_pos / _vel), enum order of pbhc_b200/spec.py OBS_SRC */
#define PBHC_NUM_SIGMA 17   /* 10 + 7 of general_tracking, enum order of SIGMA_IDS */
#define PBHC_MAX_KEY 16     /* key bodies (general_tracking) */
#define PBHC_MAX_FUTURE 32  /* future reference steps (general_tracking) */
#define PBHC_ENVF_STRIDE 8   /* env_f32 row: start_time, motion_len, air_time[2], last_contact[2], end_time_ratio, push word */
/* the push word (ABI 10) is an int32 stored in the float slot: push_robot_counter | push_threshold << 16, push_threshold =
 * (push_interval_s / dt).int() - the step count the reference compares the counter with (legged_robot_base.py:135-147, 384) */

/* layout of `scalars` (device double[PBHC_SCALARS]) */
#define PBHC_S_SIGMA 0        /* [17] reward_tracking_sigma, enum order of pbhc_b200/spec.py SIGMA_IDS */
#define PBHC_S_EMA 17         /* [17] _reward_error_ema */
#define PBHC_S_AVG_EPL 34     /* average_episode_length (an fp32 value) */
#define PBHC_S_PENALTY 35     /* reward_penalty_scale */
#define PBHC_S_FAR_THR 36     /* terminate_when_motion_far_threshold */
#define PBHC_S_SOFT_DOF_POS 37
#define PBHC_S_SOFT_DOF_VEL 38
#define PBHC_S_SOFT_TORQUE 39
#define PBHC_S_PUSH_COUNT 40  /* history pushes so far (ring-buffer head = -count mod H) */
#define PBHC_S_NOISE_CUR 41   /* current_noise_curriculum_value (obs.add_noise_currculum, legged_robot_base.py:119-120, 637-645), ABI 11 */
#define PBHC_S_LOG 48         /* [64] log_dict values of the last step, see PBHC_L_* */
#define PBHC_SCALARS 112
/* layout of `acc` (device double[PBHC_ACC]): per-step sums over envs; all-reduce across ranks before finalize */
#define PBHC_A_SIGMA_ERR 0    /* [17] sum over envs of each adaptive-sigma error term */
#define PBHC_A_UPPER_NORM 17  /* sum of |dif_pos| over upper bodies and envs (then lower, vr3, joint) */
#define PBHC_A_LOWER_NORM 18
#define PBHC_A_VR3_NORM 19
#define PBHC_A_JOINT_NORM 20
#define PBHC_A_TERM_GRAVITY 21
#define PBHC_A_TERM_FAR 22
#define PBHC_A_TERM_TIMEOUT 23
#define PBHC_A_TERM_MOTION_END 24
#define PBHC_A_RESET 25       /* number of envs reset this step */
#define PBHC_A_RESET_EPLEN 26 /* sum of their episode lengths */
#define PBHC_A_ETR_SUM 27     /* sum / sum of squares of end_time_ratio_buf over all envs */
#define PBHC_A_ETR_SQ 28
#define PBHC_A_NUM_ENVS 29
/* general_tracking only (general_tracking.py:241-255, 303-335) */
#define PBHC_A_KEY_NORM 32        /* sum of |dif_pos| over key bodies */
#define PBHC_A_LOCAL_UPPER_NORM 33 /* sums of |dif_local_body_pos| over upper / lower / vr3 / key bodies */
#define PBHC_A_LOCAL_LOWER_NORM 34
#define PBHC_A_LOCAL_VR3_NORM 35
#define PBHC_A_LOCAL_KEY_NORM 36
#define PBHC_A_TERM_REF_POS_Z 37
#define PBHC_A_TERM_REF_ORI 38
#define PBHC_A_TERM_BODY_Z 39
/* base-class terminations (legged_robot_base.py:433-476), ABI 6 */
#define PBHC_A_TERM_CONTACT 40
#define PBHC_A_TERM_LOW_HEIGHT 41
#define PBHC_A_TERM_DOF_POS 42
#define PBHC_A_TERM_DOF_VEL 43
#define PBHC_A_TERM_TORQUE 44
#define PBHC_ACC 64
#define PBHC_L_UPPER 0
#define PBHC_L_LOWER 1
#define PBHC_L_VR3 2
#define PBHC_L_JOINT 3
#define PBHC_L_TERM_GRAVITY 4
#define PBHC_L_TERM_FAR 5
#define PBHC_L_TERM_TIMEOUT 6
#define PBHC_L_TERM_MOTION_END 7
#define PBHC_L_ETR_MEAN 8
#define PBHC_L_ETR_STD 9
#define PBHC_L_FAR_THR 10     /* values logged BEFORE this step's curriculum update, like the reference */
#define PBHC_L_PENALTY 11
#define PBHC_L_AVG_EPL 12
#define PBHC_L_SOFT_DOF_POS 13
#define PBHC_L_SOFT_DOF_VEL 14
#define PBHC_L_SOFT_TORQUE 15
#define PBHC_L_ADP_SIGMA 16    /* [17] sigma / [17] ema as logged by _log_motion_tracking_info: BEFORE this step's update */
#define PBHC_L_ERROR_EMA 33
#define PBHC_L_KEY 50          /* general_tracking logs */
#define PBHC_L_LOCAL_UPPER 51
#define PBHC_L_LOCAL_LOWER 52
#define PBHC_L_LOCAL_VR3 53
#define PBHC_L_LOCAL_KEY 54
#define PBHC_L_TERM_REF_POS_Z 55
#define PBHC_L_TERM_REF_ORI 56
#define PBHC_L_TERM_BODY_Z 57
#define PBHC_L_TERM_CONTACT 58
#define PBHC_L_TERM_LOW_HEIGHT 59
#define PBHC_L_TERM_DOF_POS 60
#define PBHC_L_TERM_DOF_VEL 61
#define PBHC_L_TERM_TORQUE 62
#define PBHC_L_NOISE_CUR 63     /* log_dict["current_noise_curriculum_value"] (legged_robot_base.py:760-761), logged before the update */

typedef struct { int32_t enable; double lo, hi, degree, down, up; } PbhcCurriculum;
/* history key `key` (index into hist_src) appears in obs group `group` at `offset`, newest `len` slots, scaled by mul */
typedef struct { int32_t key, group, offset, len; float mul; } PbhcHistDst;

typedef struct {
  int32_t J, Jb, D, L, n_ext, n_feet;
  int32_t ext_parent[PBHC_MAX_EXT];
  float ext_pos[PBHC_MAX_EXT][3];
  float ext_rot_xyzw[PBHC_MAX_EXT][4];
  int32_t feet_ids[PBHC_MAX_FEET];
  uint64_t upper_mask, lower_mask, vr3_mask, penalised_mask; /* bit j = body j of the extended list */
  float dt;
  int64_t max_episode_length;
  /* rewards (ids: pbhc_b200/spec.py REWARD_IDS) */
  int32_t n_rewards, use_vec_reward, only_positive_rewards, has_termination;
  int32_t reward_id[PBHC_MAX_REWARDS];
  float reward_scale[PBHC_MAX_REWARDS];
  int32_t reward_is_penalty[PBHC_MAX_REWARDS];
  float termination_scale;
  float w_lower, w_upper, desired_feet_air_time, max_contact_force;
  float hard_lo[PBHC_MAX_DOF], hard_hi[PBHC_MAX_DOF], soft_lo[PBHC_MAX_DOF], soft_hi[PBHC_MAX_DOF];
  float vel_limit[PBHC_MAX_DOF], torque_limit[PBHC_MAX_DOF];
  float soft_dof_vel_limit, soft_torque_limit;
  /* termination */
  int32_t term_gravity, term_motion_far, term_motion_end;
  float term_gravity_thr;
  /* adaptive sigma: type 0 = origin, 1 = mean, 2 = scale */
  int32_t adaptive_enable, adaptive_type;
  double adaptive_alpha, adaptive_scale;
  PbhcCurriculum cur_penalty, cur_far, cur_dof_pos, cur_dof_vel, cur_torque;
  double num_compute_average_epl;
  /* observations (legged_robot_base.py:763-793, helpers.py:128-152): every obs group is a row of contiguous
   * segments; a segment is either one observation source (scaled by obs_scale[src]) or the newest `len` slots of one
   * history key (times `mul`, the obs scale of the history key of the group).  src_dst[k][g] = offset of source k in
   * group g (-1: absent).  Rows are clipped to +-clip_obs after assembly. */
  int32_t n_groups;
  int32_t group_dim[PBHC_MAX_OBS_GROUPS];
  int32_t src_dst[PBHC_NUM_OBS_SRC][PBHC_MAX_OBS_GROUPS];
  int32_t n_hist_dst;
  PbhcHistDst hist_dst[PBHC_MAX_HIST_DST];
  float obs_scale[PBHC_NUM_OBS_SRC], obs_noise[PBHC_NUM_OBS_SRC];
  float clip_obs;
  int32_t n_hist_keys, hist_row; /* hist_row = sum_k len_k*dim_k floats per env */
  int32_t hist_src[PBHC_MAX_HIST_KEYS], hist_len[PBHC_MAX_HIST_KEYS], hist_dim[PBHC_MAX_HIST_KEYS];
  /* episodic domain randomisation at reset (legged_robot_base.py:599-635) when values are not injected */
  int32_t randomize_pd_gain, randomize_rfi_lim, use_rao, randomize_ctrl_delay, delay_lo, delay_hi;
  float kp_lo, kp_hi, kd_lo, kd_hi, rfi_lo, rfi_hi, rao_lim;
  uint64_t seed;
  float reset_noise[6];   /* dof_pos, dof_vel, root_pos, root_rot [rad], root_vel, root_ang_vel; 0 = reference default */
  uint64_t src_used_mask; /* bit k: observation source k is referenced by a group or a history key */
  int32_t n_upper, n_lower, n_vr3; /* popcounts of the masks (log means) */
  int32_t n_episode_sums; /* n_rewards (+1 with termination): row width of episode_sums */
  /* LeggedRobotGeneralTracking (envs/motion_tracking/general_tracking.py); general = 0: LeggedRobotMotionTracking */
  int32_t general;
  int32_t anchor_index;              /* general_tracking.py:97-98; must be a simulator body (< Jb) */
  int32_t n_key;                     /* key bodies, in config order (observation layout follows it) */
  int32_t key_ids[PBHC_MAX_KEY];
  uint64_t key_mask, body_z_mask;    /* body_z: the hard-coded bodies of :253 */
  int32_t term_ref_pos_z, term_ref_ori, term_body_z;  /* :241-255 */
  float term_ref_pos_z_thr, term_ref_ori_thr, term_body_z_thr;
  int32_t n_future;                  /* obs.future_num_steps; future_steps = linspace(1, future_max_steps, n).long() (:501-507) */
  int32_t future_steps[PBHC_MAX_FUTURE];
  /* ABI 6: the base-class terminations (legged_robot_base.py:433-476; LeggedRobotMotionTracking only for now) */
  int32_t term_contact;              /* terminate_by_contact: any |contact_forces[b]| > 1 over the bodies of the mask */
  uint64_t term_contact_mask;        /* bit b = simulator body b is in termination_contact_indices (base_task.py:179-197) */
  int32_t term_low_height;           /* terminate_by_low_height: root z < term_min_height */
  float term_min_height;
  int32_t term_dof_pos, term_dof_vel, term_torque;  /* terminate_when_close_to_{dof_pos,dof_vel,torque}_limit */
  float term_dof_pos_prob, term_dof_vel_prob, term_torque_prob;  /* termination_probality.*: the check applies in a step when U[0,1) < p
                                      * (ONE draw per step for all envs, `torch.rand(1)` in the reference; here Philox keyed by
                                      * (seed, step) unless PbhcTrackIO.inject_term_gates says otherwise) */
  float term_dof_vel_scale, term_torque_scale;       /* termination_scales.termination_close_to_{dof_vel,torque}_limit */
  float term_pos_lo[PBHC_MAX_DOF], term_pos_hi[PBHC_MAX_DOF];  /* simulator.dof_pos_limits_termination (isaacgym.py:387-388) */
  /* ABI 9: the remaining heading-frame observations (motion_tracking.py:686-731, 594-599) */
  float ref_init_yaw;                /* ref_init_rpy[0, 2] (motion_tracking.py:186-187): yaw of env 0's reference root rotation at construction */
  int32_t vr3_ids[4];                /* motion_tracking_id in config order (the layout of vr_3point_pos follows it); n_vr3 <= 3 entries used */
  /* LeggedRobotMotionTracking with obs.future_ref_steps > 0: n_future holds that count (future_steps is unused there) */
  /* ABI 10: domain_rand.push_robots (legged_robot_base.py:134-147, 382-388, 1090-1106; on in the shipped domain_rand/main.yaml) */
  int32_t push_robots, push_fixed;   /* _push_fixed: the drawn velocity is added to the root velocity instead of replacing it */
  int32_t push_lo, push_hi;          /* push_interval_s: the next interval is randint(lo, hi) seconds */
  float max_push_vel;                /* max_push_vel_xy: the push is U(-max, max) on root lin vel x / y */
  /* ABI 11 */
  int32_t start_time_zero;           /* is_evaluating and not enforce_randomize_motion_start_eval: a reset starts the clip at time 0
                                      * instead of sample_time (motion_tracking.py:374-377, general_tracking.py:279-282) */
  PbhcCurriculum cur_noise;          /* obs.add_noise_currculum: every noise scale is multiplied by scalars[PBHC_S_NOISE_CUR], which a step with
                                      * resets moves by (1 -+ obs.soft_dof_pos_curriculum_degree) when the average episode length is below
                                      * obs.soft_dof_pos_curriculum_level_down_threshold / above rewards.reward_penalty_level_up_threshold (sic),
                                      * clipped to [noise_value_min, noise_value_max] (legged_robot_base.py:637-645).  The observations of step t
                                      * use the value as of the end of step t-1: the reference updates it between its reset and its observation
                                      * code, i.e. one multiplication by (1 +- degree ~ 1e-5) earlier - a difference no draw can show. */
} PbhcTrackCfg;

typedef struct {
  /* simulator state (borrowed; dof_state / root_state / actions are WRITTEN for envs that reset) */
  const float* body_state; int64_t body_env_stride; /* [N, >=Jb, 13], floats between envs */
  float* root_state;            /* [N,13] */
  float* dof_state;             /* [N,D,2] */
  const float* contact_forces;  /* [N,Jb,3] */
  float* actions;               /* [N,D] clipped actions of this step */
  const float* torques;         /* [N,D] */
  const float* env_origins;     /* [N,3] */
  const int64_t* motion_ids;    /* [N] */
  const int32_t* clip_of_motion;/* nullable */
  const float* default_dof_pos; int64_t default_dof_stride; /* [N,D] (stride D) or [D] (stride 0) */
  const float* base_com_bias;   /* [N,3] */
  const float* link_mass_scale; /* [N,L] */
  const float* friction;        /* [N] */
  const float* base_mass_scale; /* [N] simulator._base_mass_scale (general_tracking privileged obs), nullable */
  /* persistent env state */
  int64_t* episode_length_buf;  /* [N] */
  float* env_f32;               /* [N,PBHC_ENVF_STRIDE] */
  float* last_actions; float* last_dof_pos; float* last_dof_vel; /* [N,D] */
  float* kp_scale; float* kd_scale; float* rfi_lim_scale; float* rao_scale; /* [N,D] */
  int64_t* action_delay_idx;    /* [N] */
  float* hist;                  /* [N,hist_row] ring buffers, key-major */
  float* episode_sums;          /* [N,n_episode_sums] */
  double* scalars;              /* [PBHC_SCALARS] */
  /* outputs */
  float* obs[PBHC_MAX_OBS_GROUPS]; /* [N,group_dim] */
  float* rew_buf;               /* [N,n_rewards+1] (vec) or [N] */
  int64_t* reset_buf;           /* [N] */
  uint8_t* time_out_buf;        /* [N] */
  float* ref_body_pos_extend;   /* [N,J,3] nullable */
  float* ref_body_rot_extend;   /* [N,J,4] nullable */
  float* last_episode_sums;     /* [N,n_episode_sums]: episode sums of envs reset this step (rows of others untouched) */
  double* partials;             /* [max_blocks,PBHC_ACC] scratch */
  double* acc;                  /* [PBHC_ACC] */
  /* values the reference draws with torch RNG at resets; NULL = in-kernel Philox */
  const float* inject_start_times; const float* inject_kp; const float* inject_kd; const float* inject_rfi;
  const float* inject_rao; const int64_t* inject_delay;
  int64_t env_offset;           /* global id of env 0 of this call (keys the in-kernel RNG when a step is launched in chunks) */
  /* ABI 5 */
  int64_t* last_episode_length_buf; /* [N] nullable: episode length (after this step's increment) of envs reset this step, i.e.
                                     * last_episode_length_buf[env_ids] of extras["episode"]["end_epis_length"] (legged_robot_base.py:302,515);
                                     * rows of envs that did not reset are untouched */
  const uint8_t* force_reset;   /* [N] nullable: nonzero = reset this env now whatever its termination says - what
                                 * resample_motion's reset_envs_idx(arange(num_envs)) does (motion_tracking.py:385-389).  A forced reset
                                 * is not a time-out and earns no termination reward; reset_buf reports it. */
  /* ABI 6 */
  int32_t inject_term_gates;    /* < 0: the per-step gates of the dof_pos / dof_vel / torque terminations are drawn in the kernel;
                                 * >= 0: bit 0 / 1 / 2 = that check applies in this step (tests: the reference's own torch.rand(1) draws) */
  /* ABI 7 */
  const float* fut_table;       /* nullable: compact per-frame records for the future window of general_tracking (pbhc_future_table_build) */
  int32_t fut_stride;           /* floats per compact record */
  /* ABI 9 */
  float* heading_scratch;       /* [N,2] scratch, required when the vr_3point_pos observation is used: the reference rotates env i's k-th
                                 * point by the heading of env (3 i + k) mod N (`heading_inv_rot.repeat(3, 1)`, motion_tracking.py:730-731), so
                                 * every env's pre-reset heading is written here by a small kernel ahead of the step kernel */
  /* ABI 10 */
  int64_t* refresh_buf;         /* [N] nullable: need_to_refresh_envs (legged_robot_base.py:314-319, 1095) as a step stamp - an env that this
                                 * step resets or pushes gets (pushes so far, scalars[PBHC_S_PUSH_COUNT]) + 1, the value the counter has after
                                 * pbhc_track_finalize; other rows are untouched (no per-env store on the hot path), so
                                 * need_to_refresh_envs = (refresh_buf == scalars[PBHC_S_PUSH_COUNT]) */
  float* push_vel;              /* [N,2] nullable: push_robot_vel_buf / record_push_robot_vel_buf rows of envs pushed this step */
  const int64_t* inject_push_interval; /* [N] nullable: the next push_interval_s of envs pushed this step (the reference's torch.randint draw) */
  const float* inject_push_vel; /* [N,2] nullable: their push velocity (torch_rand_float) */
} PbhcTrackIO;

/* The future window of LeggedRobotGeneralTracking (_get_future_motion_targets, general_tracking.py:500-565) reads, for 20 future
 * steps x 2 frames per env, a handful of values scattered over a 2-KB table record (root and anchor pose, root velocities,
 * dof positions, key-body positions): on a motion library larger than L2 that is ~400 scattered 32-byte sectors per env.
 * This builds the same values as ONE contiguous compact record per frame,
 *   [ root quat 4 | anchor quat 4 | root vel 3 | root ang vel 3 | root pos 3 | anchor pos 3 | dof_pos D | key-body pos 3 x n_key | pad ]
 * (pbhc_future_table_stride floats, a multiple of 4), copied bit for bit from the packed table; pass it as
 * PbhcTrackIO.fut_table and the kernel reads ~3 lines per frame instead.  Rebuild after every load_motions. */
int32_t pbhc_future_table_stride(const PbhcTrackCfg* cfg);
int pbhc_future_table_build(const PbhcTrackCfg* cfg, const PbhcTableDesc* tab, float* out, int32_t stride, void* stream);

/* number of rows `partials` must have */
int pbhc_track_max_blocks(void);
/* what the last pbhc_track_step of this process launched (host-side bookkeeping for benchmarks, not thread-safe):
 * out4 = {kernel: 0 warp-per-env / 1 tile (generic) / 2 tile (specialised for the config), blocks, dynamic shared memory bytes, resident blocks per SM} */
int pbhc_track_launch_info(int32_t* out4);
/* Optional warm-up: compiles the kernel specialised for this env config into the on-disk cache (NVRTC, no GPU and no
 * device memory needed: of `tab` only the dimensions and of `io` only default_dof_stride and the NULL-ness of the
 * optional outputs are read).  pbhc_track_step does the same lazily on its first launch of a config and falls back to
 * the generic kernel when NVRTC is unavailable; < 0 here just means "no specialised kernel". */
int pbhc_track_specialize(const PbhcTrackCfg* cfg, const PbhcTableDesc* tab, const PbhcTrackIO* io);
/* main kernel + deterministic reduction of the per-block partial sums into io->acc */
int pbhc_track_step(const PbhcTrackCfg* cfg, const PbhcTableDesc* tab, const PbhcTrackIO* io, int64_t n_envs, void* stream);
/* advances `scalars` from `acc` (adaptive sigma, average episode length, curricula, log values);
 * call after pbhc_track_step (and after all-reducing acc when envs are sharded over ranks) */
int pbhc_track_finalize(const PbhcTrackCfg* cfg, const double* acc, double* scalars, void* stream);

/* ---- env ids of a reset mask: `env_ids = self.reset_buf.nonzero(as_tuple=False).flatten()` and the need_to_refresh_envs
 * twin (legged_robot_base.py:311,315) without a host round trip per mask: ids [n] int32 receives the indices of the nonzero
 * flags in ascending order, *count (device int32) their number; entries of ids past *count are untouched.
 * scratch: device int32[pbhc_compact_scratch_ints(n)].  Two launches, no atomics: the order is deterministic. */
int64_t pbhc_compact_scratch_ints(int64_t n);
int pbhc_compact_ids(const int64_t* flags, int64_t n, int32_t* ids, int32_t* count, int32_t* scratch, void* stream);

/* pbhc_compact_ids for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same inputs, outputs and
 * results as pbhc_compact_ids, but the per-block counts live in a file-scope `static __device__` array of the library
 * (16 KB, held by every GPU that loads the library until exit) instead of caller scratch; the call allocates nothing
 * (no cudaMalloc / cudaFree / allocating library routine).  Limits: n <= PBHC_COMPACT_STATIC_MAX_N, else PBHC_E_RANGE.
 * Every call on a device shares the one array (it keeps the last call's contents; nothing in it has to start at zero
 * because the first pass overwrites every count the second reads): calls that overlap on the GPU - different streams
 * without an ordering between them - race; serialise them or use pbhc_compact_ids.  The first call on a device looks up
 * the array's address (cudaGetSymbolAddress) - make that call outside a stream capture. */
#define PBHC_COMPACT_STATIC_MAX_N 16777216
int pbhc_compact_ids_static(const int64_t* flags, int64_t n, int32_t* ids, int32_t* count, void* stream);

/* ---- the cross-rank exchange of the path.  The reference is single-process (no torch.distributed anywhere); when
 * the envs are sharded over the GPUs of a box the path needs exactly two tiny SUM all-reduces: the PBHC_ACC statistics of a
 * step (between pbhc_track_step and pbhc_track_finalize) and (n, sum, sum of squares) of the advantages (between pbhc_gae
 * and pbhc_adv_normalize).  This entry point does them over NVLink peer memory in ONE single-block kernel per rank -
 * peer stores, release/acquire flags, a rank-ordered (deterministic, rank-identical) sum - instead of a library
 * collective whose cost at this size is all latency.  data [n] doubles, summed in place.  peer_buffers / peer_signal_pads:
 * DEVICE arrays [world] of device pointers to every rank's buffer (>= 2 * world * PBHC_PEER_MAX_DOUBLES doubles, zero at
 * first use) and signal pad (uint32 entries; entries signal_base .. signal_base + world are used), all mapped in this
 * process (torch symmetric memory: pbhc_b200/dist.py).  counter: this rank's device uint32, zero at first use, advanced
 * by the kernel - the call can be captured in a CUDA graph.  Every rank of the group must make the same sequence of calls. */
#define PBHC_PEER_MAX_DOUBLES 256
int pbhc_peer_allreduce_sum(double* data, int32_t n, const uint64_t* peer_buffers, const uint64_t* peer_signal_pads,
                            uint32_t* counter, int32_t rank, int32_t world, int32_t signal_base, void* stream);

/* ---- actuation half of the env step: LeggedRobotBase._pre_physics_step
 * (envs/legged_base_task/legged_robot_base.py:268-286) and _compute_torques (:795-838). */
typedef struct PbhcActuationCfg {
  int32_t D;                 /* dofs */
  int32_t Q;                 /* action-queue depth = ctrl_delay_step_range[1] + 1; 0 = randomize_ctrl_delay off */
  float action_clip;         /* robot.control.action_clip_value */
  int32_t control_type;      /* 0 "P", 1 "V", 2 "T" */
  int32_t per_dof_action_scale;
  float action_scale;        /* scalar robot.control.action_scale, or */
  float action_scales[PBHC_MAX_DOF];
  float p_gains[PBHC_MAX_DOF];
  float d_gains[PBHC_MAX_DOF];
  float torque_limits[PBHC_MAX_DOF];
  int32_t rfi;               /* domain_rand.randomize_torque_rfi */
  float rfi_lim;
  int32_t rao;               /* domain_rand.use_rao */
  int32_t clip_torques;
  float sim_dt;              /* control type "V" */
  uint64_t seed;             /* in-kernel Philox for the RFI draw when rfi_uniform is NULL */
  /* ABI 8 */
  int64_t env_offset;        /* global id of env 0 of this call (keys the RFI stream when envs are sharded over ranks) */
  uint32_t* call_counter;    /* nullable DEVICE counter: when set, the RFI stream of a call is keyed by *call_counter + call_index and
                              * pbhc_compute_torques advances the counter by one after the kernel - a replayed CUDA graph then draws
                              * fresh noise on every replay (a call index passed by value is frozen at capture) */
} PbhcActuationCfg;

/* actions = clip(actions_in, +-action_clip); *clip_count += #(|actions| == action_clip) (zero it first;
 * action_clip_frac = count / (N*D)); with Q > 0 the queue [N,Q,D] is shifted by one, actions enter
 * at slot 0 and actions_after_delay[e] = queue[e, action_delay_idx[e]]; with Q == 0 it is actions. */
int pbhc_pre_physics(const PbhcActuationCfg* cfg, const float* actions_in, float* actions,
                     float* action_queue, const int64_t* action_delay_idx, float* actions_after_delay,
                     uint64_t* clip_count, int64_t n_envs, void* stream);

/* torques [N,D] from actions_after_delay and the Isaac Gym dof_state [N*D,2] = (pos, vel).
 * default_dof_pos rows have default_dof_stride floats (0 = one shared row).  rfi_uniform [N,D]:
 * the U[0,1) draws of torch.rand_like for this call (tests), NULL = Philox keyed by
 * (seed, env_offset + env, [*call_counter +] call_index, dof). */
int pbhc_compute_torques(const PbhcActuationCfg* cfg, const float* actions_after_delay,
                         const float* dof_state, const float* default_dof_pos,
                         int32_t default_dof_stride, const float* last_dof_vel, const float* kp_scale,
                         const float* kd_scale, const float* rfi_lim_scale, const float* rao_scale,
                         const float* rfi_uniform, uint32_t call_index, float* torques,
                         int64_t n_envs, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PBHC_B200_H */

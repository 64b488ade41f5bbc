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
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define PBHC_ABI_VERSION 1
#define PBHC_OK 0
#define PBHC_E_BADARG (-1)
#define PBHC_E_ALIGN (-2)
#define PBHC_E_RANGE (-3)
#define PBHC_E_CUDA (-4)

int pbhc_abi_version(void);
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

#ifdef __cplusplus
}
#endif
#endif /* PBHC_B200_H */

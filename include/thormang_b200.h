/*
 * thormang_b200.h -- C ABI of the B200-native env-step hot path.
 *
 * The reference (Yougo-robotics/thormang_isaacgym) has NO FFI/plugin layer: the
 * boundary it offers is the Python VecTask subclass contract
 * (isaacgymenvs/tasks/base/vec_task.py:254-277,313-406).  This header is the
 * C-ABI a maintainer binds underneath that contract (ctypes stub shown in
 * INTEGRATION.md); every entry point cites the reference method it replaces.
 *
 * Conventions
 *  - plain C, no torch / CUDA types in signatures; `stream` is a cudaStream_t
 *    passed as void* (NULL = legacy default stream).
 *  - every pointer in the *Buffers structs is a DEVICE pointer borrowed from the
 *    caller (torch owns the memory); the library never frees them and only
 *    dereferences them inside a call that was given them or after tk_*_bind.
 *  - all entry points are asynchronous on `stream`, never synchronise the host,
 *    never allocate after tk_*_create, never throw, never exit.
 *  - return value: 0 = ok, < 0 = TkStatus error; tk_last_error() has the text.
 *  - quaternions are xyzw; reset_buf / progress_buf are int64 like
 *    vec_task.py:269-276; timeout_buf is 1 byte per env (torch.bool), the dtype
 *    vec_task.py:345 rebinds it to.
 *  - state tensors keep Isaac Gym's AoS layout: root_states [N,13] fp32,
 *    dof_state [N*D,2] fp32 (pos,vel interleaved), dof_force [N,D],
 *    force sensors [N,12]; rows contiguous.
 *  - a handle is bound to one device and is not thread-safe.
 */
#ifndef THORMANG_B200_H
#define THORMANG_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TK_ABI_VERSION 1
#define TK_MAX_DOF 64
#define TK_STATS 8

/* The Humanoid entry points also serve Ant (isaacgymenvs/tasks/ant.py), the same path with 8 DOFs, 60 observations
 * (no dof_force block, 24 sensor floats, raw Euler angles, no potential columns; ant.py:367-408) and every reward term
 * live (ant.py:318-364).  TkHumanoidParams.variant selects the kernel instantiation; motor_efforts carries joint_gears
 * (ant.py:273-277), sensors is [N,24], dof_force may be NULL, obs_buf is [N,60]. */
#define TK_VARIANT_HUMANOID 0
#define TK_VARIANT_ANT 1

typedef enum TkStatus {
    TK_OK = 0,
    TK_ERR_INVALID = -1,   /* bad argument / params */
    TK_ERR_CUDA = -2,      /* a CUDA runtime call failed */
    TK_ERR_UNBOUND = -3,   /* a required buffer was not bound */
    TK_ERR_ALIGN = -4      /* a pointer is not 4-byte (fp32) / 8-byte (int64) aligned */
} TkStatus;

/* How the random draws of reset_idx reach the kernel (humanoid.py:258-259,
 * gogoro_new.py:474-492,554-559 call torch.rand/randn on the spot). */
typedef enum TkNoiseMode {
    TK_NOISE_NONE = 0,     /* no reset can happen (pointers ignored; a set reset bit resets with u = 0.5) */
    TK_NOISE_COMPACT = 1,  /* row k belongs to the k-th env (ascending id) being reset: the reference's order */
    TK_NOISE_DENSE = 2,    /* row n belongs to env n */
    TK_NOISE_PHILOX = 3    /* drawn in-kernel: Philox4x32-10 keyed by (seed, step, global env id) */
} TkNoiseMode;

const char* tk_last_error(void);
int tk_abi_version(void);
/* sm count / compute capability of `device` (for grid sizing reports) */
int tk_device_info(int device, int* sm_count, int* cc_major, int* cc_minor);

/* ------------------------------------------------------------------------------------------
 * Humanoid  (isaacgymenvs/tasks/humanoid.py, cfg/task/Humanoid.yaml)
 * ------------------------------------------------------------------------------------------ */

typedef struct TkHumanoidParams {
    int32_t device;
    int32_t num_envs;
    int32_t num_dof;                 /* 21 (humanoid.py:68) | 8 (ant.py:68) */
    int32_t variant;                 /* TK_VARIANT_HUMANOID (0) | TK_VARIANT_ANT */
    float dt;                        /* cfg["sim"]["dt"], humanoid.py:115 */
    float dof_vel_scale;             /* humanoid.py:48 */
    float angular_velocity_scale;    /* humanoid.py:49 */
    float contact_force_scale;       /* humanoid.py:50 */
    float power_scale;               /* humanoid.py:51 */
    float heading_weight;            /* humanoid.py:52 */
    float up_weight;                 /* humanoid.py:53 */
    float actions_cost_scale;        /* humanoid.py:54 */
    float energy_cost_scale;         /* humanoid.py:55 */
    float joints_at_limit_cost_scale;/* humanoid.py:56 */
    float death_cost;                /* humanoid.py:57 */
    float termination_height;        /* humanoid.py:58 */
    float max_episode_length;        /* humanoid.py:65 (passed to the JIT fn as float) */
    float clip_actions;              /* vec_task.py:108 */
    float max_motor_effort;          /* humanoid.py:170 */
    float reserved1;
    /* rows of the replicated [N,*] tensors of humanoid.py:106-113 (identical for every env) */
    float targets[3];
    float inv_start_rot[4];
    float basis_vec0[3];
    float basis_vec1[3];
    float reserved2[3];
    float dof_limits_lower[TK_MAX_DOF];  /* humanoid.py:205-215 */
    float dof_limits_upper[TK_MAX_DOF];
    float motor_efforts[TK_MAX_DOF];     /* humanoid.py:160-171 */
} TkHumanoidParams;

typedef struct TkHumanoidBuffers {
    /* simulator-owned, aliased (humanoid.py:78-98) */
    float* root_states;              /* [N,13] */
    float* dof_state;                /* [N*D,2] */
    const float* dof_force;          /* [N,D] */
    const float* sensors;            /* [N,12] */
    /* task-owned (humanoid.py:92-117, vec_task.py:263-277) */
    float* actions;                  /* [N,D]  self.actions */
    float* obs_buf;                  /* [N,110] */
    float* rew_buf;                  /* [N] */
    int64_t* reset_buf;              /* [N] */
    int64_t* progress_buf;           /* [N] */
    uint8_t* timeout_buf;            /* [N] bool, may be NULL */
    int64_t* randomize_buf;          /* [N], may be NULL (only domain randomisation reads it) */
    float* potentials;               /* [N] */
    float* prev_potentials;          /* [N] */
    float* up_vec;                   /* [N,3], may be NULL (debug viz only, humanoid.py:298-316) */
    float* heading_vec;              /* [N,3], may be NULL */
    const float* initial_root_states;/* [N,13] */
    const float* initial_dof_pos;    /* [N,D] */
    float* episode_return;           /* [N], may be NULL: running return for the episode statistics */
} TkHumanoidBuffers;

typedef struct TkResetNoise {
    int32_t mode;                    /* TkNoiseMode */
    int32_t reserved;
    const float* unit_a;             /* Humanoid: U(0,1) [rows,D] for positions (humanoid.py:258)   | Gogoro: [rows,10] see below */
    const float* unit_b;             /* Humanoid: U(0,1) [rows,D] for velocities (humanoid.py:259)  | Gogoro: unused */
    int64_t rows;                    /* rows available in unit_a / unit_b */
    uint64_t seed;                   /* PHILOX */
    uint64_t step;                   /* PHILOX: step counter (< 2^32); the handle's device-side noise epoch is added as
                                      * the upper 32 bits, see tk_*_advance_epoch */
    int64_t env_id_offset;           /* PHILOX: global id of local env 0 (range-sharded multi-GPU runs) */
} TkResetNoise;

typedef struct TkHumanoid TkHumanoid;

int tk_humanoid_create(const TkHumanoidParams* params, TkHumanoid** out);
int tk_humanoid_destroy(TkHumanoid* h);
int tk_humanoid_bind(TkHumanoid* h, const TkHumanoidBuffers* buffers);

/* Humanoid.pre_physics_step (humanoid.py:281-285) fused with the action clamp of VecTask.step
 * (vec_task.py:327, applied when clip != 0): bound actions[N,D] = clamp(actions_in);
 * forces_out[N,D] = actions * motor_efforts * power_scale (the tensor handed to set_dof_actuation_force_tensor). */
int tk_humanoid_pre_physics(TkHumanoid* h, const float* actions_in, float* forces_out, int32_t clip, void* stream);

/* Humanoid.compute_observations (humanoid.py:239-251) = compute_humanoid_observations (humanoid.py:379-417):
 * writes obs_buf, potentials, prev_potentials, up_vec, heading_vec. */
int tk_humanoid_compute_observations(TkHumanoid* h, void* stream);

/* Humanoid.compute_reward (humanoid.py:219-237) = compute_humanoid_reward (humanoid.py:323-376):
 * reads obs_buf, reset_buf, progress_buf; writes rew_buf, reset_buf. */
int tk_humanoid_compute_reward(TkHumanoid* h, void* stream);

/* Humanoid.reset_idx (humanoid.py:253-279): env_ids[k] int64 device pointer, unit_pos/unit_vel U(0,1) [k,D]. */
int tk_humanoid_reset_idx(TkHumanoid* h, const int64_t* env_ids, int32_t k, const float* unit_pos,
                          const float* unit_vel, void* stream);

/* Humanoid.post_physics_step (humanoid.py:287-296) + the timeout line of VecTask.step (vec_task.py:345),
 * ONE fused kernel: progress++/randomize++, reset of the envs whose reset_buf != 0 (no host sync, no nonzero()),
 * observations, reward, reset mask, timeout mask, per-CTA episode-statistics partials. */
int tk_humanoid_post_physics(TkHumanoid* h, const TkResetNoise* noise, void* stream);

/* All reward terms of humanoid.py:344-363 incl. the ones the fork leaves out of the total:
 * terms[N,7] = heading_reward, up_reward, actions_cost, dof_at_limit_cost, electricity_cost, alive, progress. */
int tk_humanoid_reward_terms(TkHumanoid* h, float* terms, void* stream);

/* World-model style reward, second compute_reward of humanoid.py:425-472 over obs[M,110] (M = B*T). */
int tk_humanoid_wm_reward(TkHumanoid* h, const float* obs, int64_t m, float* reward_out, void* stream);

/* Materialise the U(0,1) draws TK_NOISE_PHILOX uses for (seed, step): unit_pos/unit_vel [N,D] (parity checks). */
int tk_humanoid_philox_noise(TkHumanoid* h, uint64_t seed, uint64_t step, int64_t env_id_offset,
                             float* unit_pos, float* unit_vel, void* stream);

/* Bump the device-side noise epoch (one single-thread kernel).  Host-supplied `step` values are frozen inside a captured
 * CUDA graph; capturing ONE call of this per graph makes every replay draw fresh Philox noise.  Not needed otherwise.
 * Invariant the step kernels rely on (all tk_*_advance_epoch): they load the epoch word AHEAD of their programmatic
 * dependency wait.  That is sound only because this call is a plain (non-programmatic) one-thread launch that fences
 * device-wide behind the increment and never triggers its dependents early, and because NOTHING else writes the word.
 * Do not write the handle's epoch word from other kernels or copies, and do not re-implement this bump as a kernel that
 * issues griddepcontrol.launch_dependents. */
int tk_humanoid_advance_epoch(TkHumanoid* h, void* stream);

/* Episode statistics.  Every tk_humanoid_post_physics adds its per-CTA partial sums into a handle-owned
 * small fp64 array (256 rows of TK_STATS accumulators, RED.ADD.F64 per tile).  This call sums the rows in a
 * fixed order into stats_out[TK_STATS] (device, fp64) -- sum reward, #reset, #timeout, sum episode_return of reset
 * envs, sum episode length of reset envs, sum up_proj, sum heading_proj, env-steps -- and zeroes the partials, so the
 * result covers the steps since the previous call.  accumulate != 0 adds to stats_out instead of overwriting.
 * This vector is the payload of the NCCL all-reduce in multi-GPU runs. */
int tk_humanoid_stats(TkHumanoid* h, double* stats_out, int32_t accumulate, void* stream);

/* ------------------------------------------------------------------------------------------
 * Gogoro scooter, registered variant  (isaacgymenvs/tasks/gogoro_new.py, cfg/task/Gogoro.yaml)
 * ------------------------------------------------------------------------------------------ */

typedef struct TkGogoroParams {
    int32_t device;
    int32_t num_envs;
    int32_t num_dof;                 /* 39 for scooter_V13.urdf */
    int32_t steering_joint;          /* dof_name_to_id['steering_joint'], gogoro_new.py:363 */
    int32_t rear_wheel_joint;        /* dof_name_to_id['rear_wheel_joint'], gogoro_new.py:367 */
    int32_t max_episode_length;      /* cfg["env"]["max_steps"], gogoro_new.py:39 */
    int32_t speed_freq_update;       /* gogoro_new.py:56 */
    int32_t yaw_freq_update;         /* gogoro_new.py:57 */
    float max_steering;              /* gogoro_new.py:86 */
    float max_steering_change;       /* gogoro_new.py:87 */
    /* [mean,std] or [lo,hi] pairs of cfg["noises"], gogoro_new.py:46-68.  Kept in double: the reference forms
     * hi - lo in Python doubles before torch rounds the scalar to fp32 (get_randoms, gogoro_new.py:466-467). */
    double imu_filter_noise[2];
    double imu_noise[2];
    double speed_sensor_noise[2];
    double steering_action_noise[2];
    double speed_range[2];
    double seat_offset_x_range[2];
    double seat_offset_y_range[2];
    double seat_offset_z_range[2];
    double seat_offset_xr_range[2];
    double steering_offset[2];
    double speed_sensor_offset[2];
} TkGogoroParams;

typedef struct TkGogoroBuffers {
    /* simulator-owned, aliased (gogoro_new.py:125-138) */
    float* root_tensor;              /* [N,13] */
    float* state_dof;                /* [N*D,2] */
    /* task-owned */
    float* obs_buf;                  /* [N,6]  (noisy observations, gogoro_new.py:462) */
    float* buffer_obs;               /* [N,1,6] (clean observations, gogoro_new.py:105,436-437) */
    float* rew_buf;                  /* [N] */
    int64_t* reset_buf;              /* [N] */
    int64_t* progress_buf;           /* [N] */
    uint8_t* timeout_buf;            /* [N] bool, may be NULL */
    float* yaw_command;              /* [N] gogoro_new.py:81 */
    float* curent_command;           /* [N] gogoro_new.py:91 */
    float* curent_speed;             /* [N] gogoro_new.py:90 */
    float* steer_offsets;            /* [N] gogoro_new.py:71 */
    float* imu_offsets;              /* [N] gogoro_new.py:70 */
    float* curent_speed_offset;      /* [N] gogoro_new.py:72 */
    float* action_history;           /* [N,5] gogoro_new.py:92 */
    float* config_vector;            /* [N,5] gogoro_new.py:78 */
    const float* thormang_pose;      /* [N,D] gogoro_new.py:234,262 */
    const float* root_reset_tensor;  /* [N,13] gogoro_new.py:144-145 */
    float* dof_pos_command;          /* [N,D] persistent position-target tensor (gogoro_new.py:360-364) */
    float* dof_speed_command;        /* [N,D] persistent velocity-target tensor (gogoro_new.py:366-369) */
    float* episode_return;           /* [N], may be NULL */
} TkGogoroBuffers;

/* Unit draws of one Gogoro post_physics_step.  sensor_unit N(0,1) [N,5] in the order the reference draws them
 * (gogoro_new.py:451-460).  reset rows [rows,10] in TkResetNoise.unit_a, columns in draw order: 0 U speed (:479),
 * 1 N steer offset (:481, dead), 2 U speed-sensor offset (:482), 3 U initial_target (:486), 4 U initial_rot (:487),
 * 5..9 N config_vector[0..4] (:554-559).  speed_unit U(0,1): COMPACT = k-th env whose progress hits
 * speed_freq_update (:384-386), DENSE = by env id; yaw_unit U(0,1) [N] by env id (:387 draws n_envs values). */
typedef struct TkGogoroNoise {
    TkResetNoise reset;
    const float* sensor_unit;        /* [N,5]; NULL with PHILOX */
    const float* speed_unit;         /* NULL with PHILOX */
    const float* yaw_unit;           /* [N]; NULL with PHILOX */
} TkGogoroNoise;

typedef struct TkGogoro TkGogoro;

int tk_gogoro_create(const TkGogoroParams* params, TkGogoro** out);
int tk_gogoro_destroy(TkGogoro* h);
int tk_gogoro_bind(TkGogoro* h, const TkGogoroBuffers* buffers);

/* Gogoro.pre_physics_step (gogoro_new.py:347-369): history shift, incremental steer, clamps, the two live
 * columns of the position / velocity target tensors.  steer_unit N(0,1) [N] (NULL = Philox with seed/step).
 * zero_fill != 0 also rewrites the 2 x [N,D] targets with zeros first, like the fresh torch.zeros of :360,366. */
int tk_gogoro_pre_physics(TkGogoro* h, const float* actions, const float* steer_unit, uint64_t seed, uint64_t step,
                          int64_t env_id_offset, int32_t zero_fill, void* stream);

/* compute_gogoro_observations (gogoro_new.py:692-723) into obs_out[N,6] (clean). */
int tk_gogoro_compute_observations(TkGogoro* h, float* obs_out, void* stream);

/* compute_gogoro_reward (gogoro_new.py:645-684) on cur_obs[N,6]: writes rew_buf, reset_buf. */
int tk_gogoro_compute_reward(TkGogoro* h, const float* cur_obs, void* stream);

/* Gogoro.compute_obs_rwd (gogoro_new.py:424-462): observations, history (B = 1), reward, reset, sensor-noise model. */
int tk_gogoro_compute_obs_rwd(TkGogoro* h, const float* sensor_unit, uint64_t seed, uint64_t step,
                              int64_t env_id_offset, void* stream);

/* Gogoro.reset_idx (gogoro_new.py:505-590 with randomize :474-482 and generate_spawn_r :485-492), PhysX property
 * calls left out.  env_ids[k] int64, unit [k,10]. */
int tk_gogoro_reset_idx(TkGogoro* h, const int64_t* env_ids, int32_t k, const float* unit, void* stream);

/* Gogoro.post_physics_step (gogoro_new.py:373-389) + vec_task.py:345, ONE fused kernel. */
int tk_gogoro_post_physics(TkGogoro* h, const TkGogoroNoise* noise, void* stream);

/* same as tk_humanoid_advance_epoch.  Ordering inside a capture is free: a launch recorded while its stream is being
 * captured always reads the device-side epoch word, whether this call is recorded before the steps, after them, or is
 * made between replays (eager launches skip that load until this has been called once on the handle). */
int tk_gogoro_advance_epoch(TkGogoro* h, void* stream);

/* stats: sum reward, #reset, #timeout, sum episode_return of reset envs, sum episode length of reset envs,
 * sum |tilt|, sum |yaw error|, N */
int tk_gogoro_stats(TkGogoro* h, double* stats_out, int32_t accumulate, void* stream);

/* Materialise the PHILOX draws of one step for parity checks: steer_unit [N], sensor_unit [N,5], reset_unit [N,10],
 * speed_unit [N], yaw_unit [N] (any pointer may be NULL). */
int tk_gogoro_philox_noise(TkGogoro* h, uint64_t seed, uint64_t step, int64_t env_id_offset, float* steer_unit,
                           float* sensor_unit, float* reset_unit, float* speed_unit, float* yaw_unit, void* stream);

/* ------------------------------------------------------------------------------------------
 * Gogoro scooter, 8-channel x 20-frame history variant -- the configuration of the paper runs
 * (isaacgymenvs/tasks/gogoro_realistic_turning_sim_paper.py, cfg/task/Gogoro_paper.yaml); SURVEY.md 8(f) row N1.
 * The file is not registered in isaacgymenvs/tasks/__init__.py; its class is also called Gogoro.
 *
 * Storage differs from the reference's two shift-copied [N,20,8] buffers (:114-115, :502-506):
 *   obs_buf [N,160]        = buffer_obs_noisy, time-ordered (oldest frame first), shifted IN PLACE by one frame per step
 *                            (the reference's obs_buf is a flatten() view of buffer_obs_noisy as well, :545);
 *   buffer_obs_ring [N,20,8] = the clean frames as a ring: the frame of step t sits in slot t % 20 (device-side tick);
 *                            tk_gogoro_paper_clean_frames materialises the reference's time-ordered buffer_obs;
 *   command_ring [N,20]    = channel 6 of the clean frames (the reward's act_buff, :509), same ring.
 * ------------------------------------------------------------------------------------------ */

/* The sibling file isaacgymenvs/tasks/gogoro_realistic_turning.py (same class name, buffers and step skeleton; saved with
 * DEBUG = False) is served by the same entry points with variant = TK_GOGORO_PAPER_VARIANT_TURNING:
 *   pre_physics_step :345-383  steering target = delayed command + steer_offsets (:365);
 *   compute_obs_rwd  :480-547  no command-change correction of the newest noisy frame;
 *   compute_gogoro_reward :713-765  3/(1+200 yaw^2) + 1/(1+100 tilt^2) + 10 * 1 + 2/(1+100 dtilt^2) + 10 (1 - |mean(act_buff)|),
 *                    the mean taken over every env and all 20 frames (tk_gogoro_paper_diff_sums then returns
 *                    {sum of the command ring, N * 20});
 *   generate_spawn_r :572-588, reset_idx :601-686  random spawn yaw / roll / yaw target: reset_unit has 8 columns,
 *                    5 = yaw target (:573), 6 = yaw offset (:574), 7 = roll (:575).
 * Set use_steer_delay = 1 for the file as saved (USE_STEER_DELAY = not DEBUG, :31). */
#define TK_GOGORO_PAPER_VARIANT_PAPER 0
#define TK_GOGORO_PAPER_VARIANT_TURNING 1

typedef struct TkGogoroPaperParams {
    int32_t device;
    int32_t num_envs;
    int32_t num_dof;                 /* 39 for scooter_V13.urdf */
    int32_t steering_joint;          /* dof_name_to_id['steering_joint'], :360 */
    int32_t rear_wheel_joint;        /* dof_name_to_id['rear_wheel_joint'], :376 */
    int32_t max_episode_length;      /* cfg["env"]["max_steps"], :47 */
    int32_t speed_freq_update;       /* :78 */
    int32_t yaw_freq_update;         /* :79 */
    int32_t buff_size;               /* 20, :112 */
    int32_t delay_width;             /* command_delay[1]: width of command_history, :93 (3..8) */
    int32_t use_steer_delay;         /* USE_STEER_DELAY, :31: per-env tap -steer_delay (:355) instead of the fixed -3 (:357) */
    int32_t variant;                 /* TK_GOGORO_PAPER_VARIANT_PAPER (0) | TK_GOGORO_PAPER_VARIANT_TURNING */
    float max_steering;              /* 0.5, :86 */
    float start_speed;               /* 1.3 with DEBUG_START_SPEED (:26,638-641), 0 = off */
    /* [lo,hi] pairs of cfg["noises"], :55-80; double for the same reason as in TkGogoroParams */
    double imu_filter_noise[2];
    double imu_noise[2];
    double speed_sensor_noise[2];
    double speed_sensor_offset[2];
    double imu_x_offset[2];
    double steering_offset[2];
    double command_delay[2];
    double speed_range[2];
} TkGogoroPaperParams;

typedef struct TkGogoroPaperBuffers {
    float* root_tensor;              /* [N,13] simulator-owned */
    float* state_dof;                /* [N*D,2] simulator-owned */
    float* obs_buf;                  /* [N,160], 16-byte aligned */
    float* buffer_obs_ring;          /* [N,20,8], 16-byte aligned */
    float* command_ring;             /* [N,20], 16-byte aligned */
    float* rew_buf;                  /* [N] */
    int64_t* reset_buf;              /* [N] */
    int64_t* progress_buf;           /* [N] */
    uint8_t* timeout_buf;            /* [N] bool, may be NULL */
    float* yaw_command;              /* [N] :81 */
    float* curent_command;           /* [N] :90 */
    float* curent_speed;             /* [N] :88 */
    float* steer_offsets;            /* [N] :91 */
    float* curent_speed_offset;      /* [N] :96 */
    float* curent_imu_x_offset;      /* [N] :97 */
    float* speed_no_noise;           /* [N] :100 */
    int64_t* steer_delay;            /* [N] :92 */
    float* command_history;          /* [N,delay_width] :93 */
    const float* thormang_pose;      /* [N,D] */
    const float* root_reset_tensor;  /* [N,13] :155-156 */
    float* dof_pos_command;          /* [N,D] persistent position-target tensor (:359-360) */
    float* dof_speed_command;        /* [N,D] persistent velocity-target tensor (:375-376) */
    float* episode_return;           /* [N], may be NULL */
} TkGogoroPaperBuffers;

/* U(0,1) draws of one step, addressed by env id (mode TK_NOISE_DENSE), or drawn in-kernel (TK_NOISE_PHILOX), or absent
 * (TK_NOISE_NONE: every draw = 0.5).  sensor_unit [N,6] in draw order: 2 x imu_filter (:525), 2 x imu_noise (:527),
 * speed noise (:528), imu_filter (:531).  reset_unit [N,5]: speed (:562), steer_delay (:563), steer_offsets (:564),
 * speed-sensor offset (:566), imu_x offset (:655).  speed_unit [N] (:407), yaw_unit [N] (:410). */
typedef struct TkGogoroPaperNoise {
    int32_t mode;
    int32_t reserved;
    const float* sensor_unit;
    const float* reset_unit;
    const float* speed_unit;
    const float* yaw_unit;
    uint64_t seed;
    uint64_t step;
    int64_t env_id_offset;
} TkGogoroPaperNoise;

typedef struct TkGogoroPaper TkGogoroPaper;

int tk_gogoro_paper_create(const TkGogoroPaperParams* params, TkGogoroPaper** out);
int tk_gogoro_paper_destroy(TkGogoroPaper* h);
int tk_gogoro_paper_bind(TkGogoroPaper* h, const TkGogoroPaperBuffers* buffers);

/* pre_physics_step (:342-378): clamp, direct steer command, delay line, the two live target columns. */
int tk_gogoro_paper_pre_physics(TkGogoroPaper* h, const float* actions, int32_t zero_fill, void* stream);

/* compute_gogoro_observations (:770-807) on the current state -> obs_out [N,8] (clean); nothing else is written. */
int tk_gogoro_paper_compute_observations(TkGogoroPaper* h, float* obs_out, void* stream);

/* compute_obs_rwd (:490-545) except the cross-env term of the reward; MUST be followed by
 * tk_gogoro_paper_finalize_reward on the same stream. */
int tk_gogoro_paper_compute_obs_rwd(TkGogoroPaper* h, const TkGogoroPaperNoise* noise, void* stream);

/* post_physics_step (:382-412) + vec_task.py:345 as ONE fused kernel, except the cross-env term of the reward;
 * MUST be followed by tk_gogoro_paper_finalize_reward on the same stream. */
int tk_gogoro_paper_post_physics(TkGogoroPaper* h, const TkGogoroPaperNoise* noise, void* stream);

/* reward7 = 1 - mean over ALL envs and frame pairs of diff(act_buff / 0.5)^2 (:747) couples every env.
 * sums_out[2] (device, fp64) = {sum of squared differences, number of differences} of the step just computed on THIS
 * GPU: the payload of the per-step all-reduce in range-sharded multi-GPU runs. */
int tk_gogoro_paper_diff_sums(TkGogoroPaper* h, double* sums_out, void* stream);

/* Second (small) kernel of a step: rew_buf += 0.2 * reward7, clip at 0 (:760), fall penalty -1 (:761), episode return;
 * clears the accumulators and advances the frame ring.  sums = the all-reduced pair, or NULL on a single GPU. */
int tk_gogoro_paper_finalize_reward(TkGogoroPaper* h, const double* sums, void* stream);

/* reset_idx (:599-688 with randomize :557-566 and generate_spawn_r :570-587), PhysX property calls left out.
 * env_ids[k] int64, unit [k,5]. */
int tk_gogoro_paper_reset_idx(TkGogoroPaper* h, const int64_t* env_ids, int32_t k, const float* unit, void* stream);

/* The reference's buffer_obs: clean frames in time order, frames_out [N,20,8] (16-byte aligned). */
int tk_gogoro_paper_clean_frames(TkGogoroPaper* h, float* frames_out, void* stream);

int tk_gogoro_paper_advance_epoch(TkGogoroPaper* h, void* stream);

/* stats: sum reward, #reset, #timeout, sum episode_return of reset envs, sum episode length of reset envs,
 * sum |tilt|, sum |yaw error|, N */
int tk_gogoro_paper_stats(TkGogoroPaper* h, double* stats_out, int32_t accumulate, void* stream);

/* Materialise the PHILOX draws of one step: sensor_unit [N,6], reset_unit [N,5], speed_unit [N], yaw_unit [N]. */
int tk_gogoro_paper_philox_noise(TkGogoroPaper* h, uint64_t seed, uint64_t step, int64_t env_id_offset, float* sensor_unit,
                                 float* reset_unit, float* speed_unit, float* yaw_unit, void* stream);

/* No counterpart in the reference: an allocation-free twin of tk_gogoro_paper_diff_sums, added for this library's users.
 *
 * Every other entry point of this section needs a TkGogoroPaper handle, and tk_gogoro_paper_create reserves the handle's
 * work areas with cudaMalloc.  Some hosts are not in a position to do that: a plug-in running inside a simulator or
 * trainer that has already claimed the whole device heap for its own pool, code running where the implicit device-wide
 * synchronisation of cudaMalloc / cudaFree is forbidden (a real-time step loop, a host callback), or a sidecar that must not
 * disturb another process's memory accounting on the GPU.  Such a caller can still produce the per-GPU payload of the
 * reward's cross-env term (what range-sharded runs all-reduce between the two kernels of a step) with this call.  It takes
 * no handle, calls no cudaMalloc / cudaFree / allocating library routine, and keeps its scratch (1024 fp64 block partials and
 * an arrival counter, 8448 bytes) in a fixed array built into the library; every GPU that loads the library holds those
 * bytes until the process exits.
 *
 *   command_ring [num_envs,20] fp32 (device): the clean command ring (TkGogoroPaperBuffers.command_ring) as the fused step
 *                left it; newest_slot = the ring slot that step wrote (completed steps before it, modulo 20).
 *   variant      TK_GOGORO_PAPER_VARIANT_PAPER: sums_out = {sum over envs of the 19 squared differences of consecutive
 *                commands / 0.5 in time order, 19 * num_envs};  _TURNING: {sum of the ring, 20 * num_envs}.
 *   sums_out     [2] fp64 (device).  Equal to tk_gogoro_paper_diff_sums after the same step up to fp64 summation order.
 *
 * Limit: num_envs <= 1,048,576 per call; larger sizes are refused with TK_ERR_INVALID (split the env range, the sums add).
 * Runs on the CURRENT device.  The first call on each device looks up the array's device address (cudaGetSymbolAddress,
 * which may load the module); make that first call before any stream capture.  All calls on a device share the one
 * array: it still holds the previous call's partials (each call overwrites the ones it uses and clears the arrival
 * counter on `stream` before its kernel), and two calls that overlap on the GPU -- different streams, or different host
 * threads on different streams -- race on it and return garbage.  Calls on one stream are safe back to back. */
int tk_gogoro_paper_diff_sums_static(const float* command_ring, int64_t num_envs, int32_t variant, int32_t newest_slot,
                                     double* sums_out, void* stream);

/* ------------------------------------------------------------------------------------------
 * Anymal  (isaacgymenvs/tasks/anymal.py, cfg/task/Anymal.yaml) -- SURVEY.md 8(f) row N2: the canonical
 * `targets = action_scale * actions + default_dof_pos` PD-target write (:229-232) and contact-force termination
 * (:359-362).  This fork's copy: 61 observations via compute_anymal_observations_V2 (:399-431), 12 DOFs, 13 bodies.
 * ------------------------------------------------------------------------------------------ */

typedef struct TkAnymalParams {
    int32_t device;
    int32_t num_envs;
    int32_t num_dof;                 /* 12 (:84) */
    int32_t num_bodies;              /* 13: base + 4 x (HIP, THIGH, SHANK), :192,206-213 */
    int32_t max_episode_length;      /* int(episodeLength_s / dt + 0.5), :92 */
    int32_t base_index;              /* :215,226 */
    int32_t num_knee;                /* len(knee_indices), :212-213 */
    int32_t knee_indices[8];         /* rigid-body indices of the THIGH bodies, :223-224 */
    int32_t reserved;
    float dof_pos_scale;             /* :48 */
    float dof_vel_scale;             /* :49 */
    float action_scale;              /* :50 */
    float clip_actions;              /* vec_task.py:108 */
    /* python doubles: the reference multiplies the reward scales by dt in double (:96-97) before torch rounds to fp32 */
    double dt;
    double rew_lin_vel_xy;           /* :53 */
    double rew_ang_vel_z;            /* :54 */
    double rew_torque;               /* :55 */
    double command_x_range[2];       /* :62 */
    double command_y_range[2];       /* :63 */
    double command_yaw_range[2];     /* :64 */
    /* rows of replicated [N,*] tensors (identical for every env) */
    float default_dof_pos[12];       /* :131-136 */
    float base_init_state[13];       /* :72-78,140-141 */
    float gravity_vec[3];            /* :142 */
} TkAnymalParams;

typedef struct TkAnymalBuffers {
    float* root_states;              /* [N,13] simulator-owned */
    float* dof_state;                /* [N*12,2] simulator-owned */
    const float* torques;            /* [N,12] dof-force tensor, :125 */
    const float* contact_forces;     /* [N,13,3] net contact forces, :124 */
    float* commands;                 /* [N,3] :127 */
    float* actions;                  /* [N,12] self.actions */
    float* obs_buf;                  /* [N,61] */
    float* rew_buf;                  /* [N] */
    int64_t* reset_buf;              /* [N] */
    int64_t* progress_buf;           /* [N] */
    uint8_t* timeout_buf;            /* [N] bool, may be NULL */
    float* episode_return;           /* [N], may be NULL */
} TkAnymalBuffers;

/* U(0,1) draws of reset_idx (:288-289,305-307), addressed by env id (TK_NOISE_DENSE), drawn in-kernel
 * (TK_NOISE_PHILOX) or absent (TK_NOISE_NONE: 0.5): unit_pos [N,12], unit_vel [N,12], unit_cmd [N,3] = x, y, yaw. */
typedef struct TkAnymalNoise {
    int32_t mode;
    int32_t reserved;
    const float* unit_pos;
    const float* unit_vel;
    const float* unit_cmd;
    uint64_t seed;
    uint64_t step;
    int64_t env_id_offset;
} TkAnymalNoise;

typedef struct TkAnymal TkAnymal;

int tk_anymal_create(const TkAnymalParams* params, TkAnymal** out);
int tk_anymal_destroy(TkAnymal* h);
int tk_anymal_bind(TkAnymal* h, const TkAnymalBuffers* buffers);

/* Anymal.pre_physics_step (:229-232) with the action clamp of vec_task.py:327 (when clip != 0): bound actions = clamp(in);
 * targets_out[N,12] = action_scale * actions + default_dof_pos (the tensor handed to set_dof_position_target_tensor). */
int tk_anymal_pre_physics(TkAnymal* h, const float* actions_in, float* targets_out, int32_t clip, void* stream);
/* compute_observations (:262-283) = compute_anymal_observations_V2 (:399-431): writes obs_buf. */
int tk_anymal_compute_observations(TkAnymal* h, void* stream);
/* compute_reward (:244-259) = compute_anymal_reward (:323-372): writes rew_buf, reset_buf. */
int tk_anymal_compute_reward(TkAnymal* h, void* stream);
/* reset_idx (:285-311): env_ids[k] int64, unit_pos / unit_vel [k,12], unit_cmd [k,3]. */
int tk_anymal_reset_idx(TkAnymal* h, const int64_t* env_ids, int32_t k, const float* unit_pos, const float* unit_vel,
                        const float* unit_cmd, void* stream);
/* post_physics_step (:234-242) + vec_task.py:345 as ONE fused kernel. */
int tk_anymal_post_physics(TkAnymal* h, const TkAnymalNoise* noise, void* stream);
int tk_anymal_advance_epoch(TkAnymal* h, void* stream);
/* stats: sum reward, #reset, #timeout, sum episode_return / length of reset envs, sum lin-vel error, sum ang-vel error, N */
int tk_anymal_stats(TkAnymal* h, double* stats_out, int32_t accumulate, void* stream);
int tk_anymal_philox_noise(TkAnymal* h, uint64_t seed, uint64_t step, int64_t env_id_offset, float* unit_pos, float* unit_vel,
                           float* unit_cmd, void* stream);

/* ------------------------------------------------------------------------------------------
 * MA_OP3  (isaacgymenvs/tasks/MA_OP3.py, cfg/task/MA_OP3.yaml) -- SURVEY.md 8(f) row N3: two 22-DOF ROBOTIS OP3
 * humanoids and one table per env; tensors are [N, agents, ...].  compute_op3_observations (:662-708),
 * compute_op3_reward (:550-658), compute_objective_reward (:496-547), the five contact terms (:319-358),
 * compute_observations (:411-438), compute_reward (:360-409), reset_idx (:440-489).  The reference's step hooks are
 * stubs (:309-317): the fused step is the anymal.py:229-242 ordering around those methods, with prev_torques /
 * last_actions (read at :373,:384; never assigned / only zeroed in the reference) refreshed at the end of the step.
 * ------------------------------------------------------------------------------------------ */

#define TK_OP3_AGENTS 2
#define TK_OP3_DOF 22
#define TK_OP3_OBS 88
#define TK_OP3_MAX_FEET 8
#define TK_OP3_MAX_GRIPPERS 4

typedef struct TkOp3Params {
    int32_t device;
    int32_t num_envs;
    int32_t num_agents;              /* 2 (MA_OP3.yaml numAgents) */
    int32_t num_dof;                 /* 22 per robot (:208) */
    int32_t num_bodies_total;        /* rigid bodies per env = dim 1 of contact_forces.view(N, -1, 3), :114 */
    int32_t max_episode_length;      /* int(episodeLength_s / dt + 0.5), :69 */
    int32_t num_feet;                /* len(feet_names), :237-238 */
    int32_t num_grippers;            /* len(gripper_names) >= 2, :241-242; [0] = left, [1] = right (:347-348) */
    int32_t feet_indices[TK_OP3_AGENTS][TK_OP3_MAX_FEET];          /* into dim 1 of contact_forces, :294-296 */
    int32_t gripper_indices[TK_OP3_AGENTS][TK_OP3_MAX_GRIPPERS];   /* :300-302 */
    float lin_vel_scale;             /* :18 */
    float ang_vel_scale;             /* :19 */
    float dof_pos_scale;             /* :20 */
    float dof_vel_scale;             /* :21 */
    float action_scale;              /* :22 */
    float clip_actions;
    /* python doubles: the reward scales are multiplied by dt in double (:80-81) before torch rounds to fp32 */
    double dt;
    double rew_torque, rew_up, rew_air_time, rew_syns_hip, rew_no_fly, rew_stand, rew_action_rate;   /* :25-35 */
    double command_x_range[2];       /* :42 */
    double command_y_range[2];       /* :43 */
    double command_yaw_range[2];     /* :44 */
    /* rows of replicated [N,*] tensors (identical for every env) */
    float default_dof_pos[TK_OP3_DOF];   /* :126-129 */
    float inv_start_rot[4];          /* :143 */
    float gravity_vec[3];            /* :132 */
    float up_vec[3];                 /* :139 (vec1 of both reward functions) */
    float targets[3];                /* :144 */
    float goal_pos[3];               /* :100 */
} TkOp3Params;

typedef struct TkOp3Buffers {
    float* root_states;              /* [N,3,13] simulator-owned: robot 0, robot 1, table (:98) */
    const float* initial_root_states;/* [N,3,13] (:99-100) */
    float* dof_state;                /* [N,2,22,2] simulator-owned (:109-111) */
    const float* torques;            /* [N,2,22] dof-force tensor (:115) */
    float* prev_torques;             /* [N,2,22] (:373) */
    const float* contact_forces;     /* [N,bodies,3] (:114) */
    float* commands;                 /* [N,3] (:117) */
    float* actions;                  /* [N,2,22] */
    float* last_actions;             /* [N,2,22] (:137) */
    float* feet_air_time;            /* [N,2,F] (:134) */
    uint8_t* last_contacts;          /* [N,2,F] bool (:135) */
    float* potentials;               /* [N,2] (:145) */
    float* prev_potentials;          /* [N,2] (:146) */
    float* table_potentials;         /* [N] (:106) */
    float* prev_table_potentials;    /* [N] (:107) */
    float* obs_buf;                  /* [N,2,88] */
    float* rew_buf;                  /* [N,2] */
    int64_t* reset_buf;              /* [N] */
    int64_t* progress_buf;           /* [N] */
    uint8_t* timeout_buf;            /* [N] bool, may be NULL */
    float* episode_return;           /* [N] (sum over both agents), may be NULL */
    int64_t* grip_counts;            /* [agents,2] batch-global gripper counts (:351-352), see tk_ma_op3_grip_counts */
} TkOp3Buffers;

/* U(0,1) draws of reset_idx (:472-474): unit_cmd [N,3] = x, y, yaw addressed by env id (TK_NOISE_DENSE), drawn in-kernel
 * (TK_NOISE_PHILOX) or absent (TK_NOISE_NONE: 0.5). */
typedef struct TkOp3Noise {
    int32_t mode;
    int32_t reserved;
    const float* unit_cmd;
    uint64_t seed;
    uint64_t step;
    int64_t env_id_offset;
} TkOp3Noise;

typedef struct TkOp3 TkOp3;

int tk_ma_op3_create(const TkOp3Params* params, TkOp3** out);
int tk_ma_op3_destroy(TkOp3* h);
int tk_ma_op3_bind(TkOp3* h, const TkOp3Buffers* buffers);
/* Un-stubbed pre_physics_step (:309-310 is `pass`; anymal.py:229-232 + clamp vec_task.py:327): actions = clamp(in),
 * targets_out[N,2,22] = action_scale * actions + default_dof_pos. */
int tk_ma_op3_pre_physics(TkOp3* h, const float* actions_in, float* targets_out, int32_t clip, void* stream);
/* _reward_gripper_contact_hold (:346-358) reduces over ALL envs.  This call counts, into the bound grip_counts array
 * (int64 [agents,2]), the envs whose left / right gripper x-force exceeds 0.1.  A sharded run all-reduces (SUM) that
 * array over ranks before the reward kernel -- the one exchange step of this path.  tk_ma_op3_compute_reward /
 * tk_ma_op3_post_physics read it; they MUST follow this call on the same stream. */
int tk_ma_op3_grip_counts(TkOp3* h, void* stream);
/* compute_observations (:411-438): obs_buf, potentials, prev_potentials, prev_table_potentials. */
int tk_ma_op3_compute_observations(TkOp3* h, void* stream);
/* compute_reward (:360-409): rew_buf, reset_buf, feet_air_time, last_contacts. */
int tk_ma_op3_compute_reward(TkOp3* h, void* stream);
/* reset_idx (:440-489): env_ids[k] int64, unit_cmd [k,3]. */
int tk_ma_op3_reset_idx(TkOp3* h, const int64_t* env_ids, int32_t k, const float* unit_cmd, void* stream);
/* progress += 1, reset prologue, compute_observations, compute_reward, prev_torques / last_actions refresh and the
 * timeout line (multi_vec_task.py:193) as ONE fused kernel.  The kernel also counts the gripper contacts of :351-352
 * into grip_counts (no pre-pass) and writes rewards with _reward_gripper_contact_hold = 0; it MUST be followed by
 * tk_ma_op3_finalize_reward on the same stream -- after an all-reduce (SUM) of grip_counts when the envs are sharded. */
int tk_ma_op3_post_physics(TkOp3* h, const TkOp3Noise* noise, void* stream);
/* Adds the batch-global hold term (:346-358, :648-650) to the rewards of the agents whose counts are exactly (1, 1);
 * a grid that exits immediately otherwise. */
int tk_ma_op3_finalize_reward(TkOp3* h, void* stream);
int tk_ma_op3_advance_epoch(TkOp3* h, void* stream);
/* stats: sum reward (both agents), #reset, #timeout, sum episode_return / length of reset envs, sum agent-0 / agent-1
 * distance to the table, N */
int tk_ma_op3_stats(TkOp3* h, double* stats_out, int32_t accumulate, void* stream);
int tk_ma_op3_philox_noise(TkOp3* h, uint64_t seed, uint64_t step, int64_t env_id_offset, float* unit_cmd, void* stream);

/* ------------------------------------------------------------------------------------------
 * VecTask.step prologue / epilogue (isaacgymenvs/tasks/base/vec_task.py:323-327,347-353) and the non-physical
 * domain-randomisation lambdas of apply_randomizations (:576-646) -- SURVEY.md 8(f) row N4.
 * ------------------------------------------------------------------------------------------ */

/* One entry of dr_randomizations['observations' | 'actions'] AFTER the schedule scaling of :592-634 (host arithmetic).
 * gaussian: range = [mu, var], range_correlated = [mu_corr, var_corr] (:591-592);
 * uniform:  range = [lo, hi],  range_correlated = [lo_corr, hi_corr]  (:624-625). */
typedef struct TkNoiseLambda {
    int32_t distribution;            /* 0 none, 1 gaussian, 2 uniform */
    int32_t operation;               /* 0 additive (operator.add), 1 scaling (operator.mul), :582 */
    double range[2];
    double range_correlated[2];
} TkNoiseLambda;

/* ONE pass over a tensor of `count` floats: x' = op(x, (corr * a_corr + b_corr) + unit * a + b)  (noise_lambda, :612-620 /
 * :635-641), y = clamp(x', +-clip).  noisy_out (may alias `in`: what `self.obs_buf = noise_lambda(self.obs_buf)` rebinds
 * to) and clamped_out (obs_dict["obs"] / action_tensor; e.g. slot t of a [horizon, N, obs] rollout buffer) may each be
 * NULL.  corr: the persistent N(0,1) draw params['corr'] or NULL (= 0).  unit: N(0,1) / U(0,1) draws of this call, or NULL:
 * in-kernel Philox keyed by (seed, step, (elem_offset + i) / 4); elem_offset (a multiple of 4) is the global index of
 * element 0 in range-sharded runs.  lambda NULL or distribution 0: clamp copy only.  clip = INFINITY: no clamp. */
int tk_step_noise_clamp(const float* in, int64_t count, const TkNoiseLambda* lambda, const float* corr, const float* unit,
                        uint64_t seed, uint64_t step, int64_t elem_offset, float clip, float* noisy_out, float* clamped_out,
                        void* stream);
/* The unit draws tk_step_noise_clamp makes in Philox mode (parity checks). */
int tk_step_noise_units(int64_t count, int32_t distribution, uint64_t seed, uint64_t step, int64_t elem_offset, float* unit_out,
                        void* stream);

/* env_ids = reset_buf.nonzero().squeeze(-1) cast to int32 -- the id list every reset_idx hands to
 * gym.set_actor_root_state_tensor_indexed / gym.set_dof_state_tensor_indexed (humanoid.py:290,264-271;
 * gogoro_new.py:377,506,547,552; anymal.py:277-283).  The fused tk_*_post_physics kernels consume AND overwrite reset_buf,
 * so an adapter in front of a simulator that keeps its own copy of the state calls this FIRST, on the same stream: ids_out
 * [num_envs] int32 receives the ids of the non-zero entries in ascending order (what nonzero() returns), count_out [1]
 * int32 their number (both device memory); after the fused step it pushes exactly those rows.
 * scratch: tk_reset_ids_scratch_bytes(num_envs) bytes of device memory owned by the caller (4 bytes per 1024 envs), reusable
 * across calls on one stream.  No allocation, no synchronisation. */
int64_t tk_reset_ids_scratch_bytes(int64_t num_envs);
int tk_reset_ids(const int64_t* reset_buf, int64_t num_envs, int32_t* ids_out, int32_t* count_out, int32_t* scratch, void* stream);

/* Second destination of the Humanoid / Ant obs tile: every tk_humanoid_post_physics / _compute_observations after this call
 * also stores the [N,obs] rows at obs_copy (16-byte aligned; NULL switches it off) straight from the kernel's shared-memory
 * tile -- slot t of an rl_games [horizon, N, obs] rollout buffer when clipObservations is inf and no observation noise is
 * configured (Humanoid.yaml), which removes the clamp copy of vec_task.py:353 and the rollout copy after it. */
int tk_humanoid_set_obs_copy(TkHumanoid* h, float* obs_copy);

#ifdef __cplusplus
}
#endif
#endif /* THORMANG_B200_H */

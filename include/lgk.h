/*
 * lgk.h -- C ABI of the B200-native legged_gym step kernels (liblgk.so).
 *
 * Drop-in boundary for the per-env data-parallel path of the reference's
 * LeggedRobot.step() / post_physics_step()
 * (/root/reference/legged_gym/envs/base/legged_robot.py:79-231,337-474,818-966,
 *  envs/go1/go1.py:68-107, envs/cassie/cassie.py:43-46).
 *
 * Conventions
 *  - plain C, no torch / C++ types; every pointer is a raw CUDA device pointer unless
 *    the field name ends in _host.
 *  - the library never allocates, frees or retains caller memory; LgkConsts is read
 *    during the call only (copied by value into kernel parameters).
 *  - every entry point is asynchronous on the given stream (a cudaStream_t passed as
 *    void*), performs no host synchronisation and may be captured in a CUDA graph.
 *  - return value: 0 = success; negative = LgkStatus argument error; positive =
 *    cudaError_t raised by the launch.
 *  - all floating point data is fp32; index / counter types are called out per field.
 */
#ifndef LGK_H_
#define LGK_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LGK_ABI_VERSION 6

#define LGK_MAX_DOF     16   /* num_dof = num_actions (12 for every shipped robot)        */
#define LGK_MAX_FEET     8
#define LGK_MAX_PEN     16   /* penalised-contact bodies                                  */
#define LGK_MAX_TERM     8   /* termination-contact bodies                                */
#define LGK_MAX_REWARDS 32   /* active reward terms incl. termination                     */
#define LGK_MAX_HIP      4

typedef enum LgkStatus {
  LGK_OK = 0,
  LGK_ERR_NULL = -1,          /* required pointer is NULL                                  */
  LGK_ERR_DIM = -2,           /* dimension outside the supported range                     */
  LGK_ERR_CONTROL_TYPE = -3,  /* unknown control type (reference: NameError, :391)         */
  LGK_ERR_REWARD = -4,        /* unknown reward term id (reference: AttributeError, :593)  */
  LGK_ERR_ALIGN = -5,         /* pointer not aligned as required                           */
  LGK_ERR_VERSION = -6,       /* struct_size / abi_version mismatch                        */
  LGK_ERR_TERRAIN = -7,       /* heights requested but no height field given (:833-834)    */
  LGK_ERR_CAPACITY = -8       /* size beyond the fixed capacity of an allocation-free entry point */
} LgkStatus;

/* control_type of LeggedRobotCfg.control (legged_robot.py:383-391) */
typedef enum LgkControlType { LGK_CTRL_P = 0, LGK_CTRL_V = 1, LGK_CTRL_T = 2 } LgkControlType;

/* reward term ids: one per _reward_<name> of the reference (legged_robot.py:857-966,
 * cassie.py:43-46).  The host passes the ACTIVE terms in evaluation order (alphabetical,
 * the order class_to_dict yields -- utils/helpers.py:41-56). */
typedef enum LgkRewardTerm {
  LGK_REW_ACTION_RATE = 0,
  LGK_REW_ANG_VEL_XY,
  LGK_REW_BASE_HEIGHT,
  LGK_REW_COLLISION,
  LGK_REW_DOF_ACC,
  LGK_REW_DOF_POS_LIMITS,
  LGK_REW_DOF_VEL,
  LGK_REW_DOF_VEL_LIMITS,
  LGK_REW_ENERGY,
  LGK_REW_FEET_AIR_TIME,
  LGK_REW_FEET_CONTACT_FORCES,
  LGK_REW_HIP_MOTION,
  LGK_REW_LIN_VEL_Z,
  LGK_REW_NO_FLY,
  LGK_REW_ORIENTATION,
  LGK_REW_STAND_STILL,
  LGK_REW_STUMBLE,
  LGK_REW_TORQUE_LIMITS,
  LGK_REW_TORQUES,
  LGK_REW_TRACKING_ANG_VEL,
  LGK_REW_TRACKING_LIN_VEL,
  LGK_REW_EXTERNAL,          /* value supplied per env by the host in ext_rewards        */
  LGK_REW_COUNT
} LgkRewardTerm;

/* LgkConsts.index_math -- the one device-dependent rounding on the height-index path
 * (legged_robot.py:842): (p / horizontal_scale).long().  torch-CPU performs an IEEE division,
 * torch-CUDA multiplies by the fp32 reciprocal of the python scalar; the truncated indices differ
 * in ~2e-6 of samples (measured on the B200 box), so the caller names the torch device it must
 * agree with bit-for-bit.  Everything else on that path -- normalize([0,0,qz,qw]) as
 * sqrtf(z*z + w*w) with separately rounded squares, quat_apply without contraction -- was measured
 * to be identical on both devices (tools/calibrate_norm.py). */
typedef enum LgkIndexMath {
  LGK_INDEX_DIV = 0,      /* torch CPU  */
  LGK_INDEX_MUL_RCP = 1   /* torch CUDA */
} LgkIndexMath;

/* random-number source */
typedef enum LgkRngMode {
  LGK_RNG_PHILOX = 0,  /* in-kernel Philox4x32-7 keyed by (seed, step, env, site)          */
  LGK_RNG_TAPS = 1     /* caller-provided uniforms, one row per env (parity mode)          */
} LgkRngMode;

enum {                                   /* LgkConsts.flags                                */
  LGK_F_HEADING_COMMAND   = 1u << 0,     /* cfg.commands.heading_command                   */
  LGK_F_MEASURE_HEIGHTS   = 1u << 1,     /* cfg.terrain.measure_heights                    */
  LGK_F_TERRAIN_PLANE     = 1u << 2,     /* mesh_type == 'plane' -> heights are zeros      */
  LGK_F_ADD_NOISE         = 1u << 3,     /* cfg.noise.add_noise                            */
  LGK_F_PUSH_ROBOTS       = 1u << 4,     /* cfg.domain_rand.push_robots                    */
  LGK_F_TERRAIN_CURRICULUM= 1u << 5,     /* cfg.terrain.curriculum (rough terrain only)    */
  LGK_F_CUSTOM_ORIGINS    = 1u << 6,     /* rough terrain: xy jitter on reset (:422-425)   */
  LGK_F_ONLY_POSITIVE     = 1u << 7,     /* cfg.rewards.only_positive_rewards              */
  LGK_F_HAS_TERMINATION   = 1u << 8,     /* 'termination' scale is non-zero (:209-212)     */
  LGK_F_INIT_DONE         = 1u << 9,     /* self.init_done (curriculum skipped before)     */
  LGK_F_EP_INCREMENTED    = 1u << 10,    /* episode_length_buf already counts this step
                                            (lgk_pre_reward_state ran): do not add 1 again */
  LGK_F_CALLBACK_GIVEN    = 1u << 12,    /* commands are final: a _post_physics_step_callback() override
                                            (legged_robot.py:337-352) ran after lgk_pre_reward_state did the
                                            base resample / heading update; the step kernel does not redo them */
  LGK_F_FILTERED_CONTACTS = 1u << 13,    /* cfg.rewards.filtered_feet_contact: _reward_feet_air_time in the variant the
                                            reference keeps commented out (legged_robot.py:928-939) -- the foot counts
                                            as in contact if it is now OR was at the previous step (PhysX contact
                                            reporting is unreliable on meshes); needs LgkBuffers.last_contacts */
  LGK_F_RESET_GIVEN       = 1u << 11     /* reset_buf / time_out_buf are INPUTS: a check_termination()
                                            override (legged_robot.py:143-148) decided after
                                            lgk_pre_reward_state; the step kernel acts on them as they are */
};

typedef struct LgkRewardEntry {
  int32_t term;    /* LgkRewardTerm                                                        */
  float   scale;   /* cfg scale * dt, fp32 (legged_robot.py:584)                           */
} LgkRewardEntry;

/* Constant block: everything _parse_cfg/_init_buffers/_prepare_reward_function derive from
 * LeggedRobotCfg (legged_robot.py:503-598,769-779).  Host memory, passed by pointer. */
typedef struct LgkConsts {
  uint32_t struct_size;       /* sizeof(LgkConsts), checked                                */
  uint32_t abi_version;       /* LGK_ABI_VERSION                                           */
  /* dimensions */
  int32_t  num_envs;          /* N (this rank)                                             */
  int32_t  num_dof;           /* D                                                         */
  int32_t  num_bodies;        /* B                                                         */
  int32_t  num_feet;          /* F                                                         */
  int32_t  num_pen;           /* P                                                         */
  int32_t  num_term;          /* T                                                         */
  int32_t  num_height_points; /* H (0 when measure_heights is off)                         */
  int32_t  num_obs;           /* O                                                         */
  int32_t  num_rewards;       /* R: entries of rewards[] (termination excluded)            */
  int32_t  num_commands;      /* 4                                                         */
  int32_t  env_id_offset;     /* global id of local env 0 (multi-GPU sharding)             */
  uint32_t flags;             /* LGK_F_*                                                   */
  int32_t  control_type;      /* LgkControlType                                            */
  int32_t  index_math;        /* LgkIndexMath                                              */
  int32_t  rng_mode;          /* LgkRngMode                                                */
  int32_t  envs_per_cta;      /* envs per persistent CTA; 0 = library balances over the SMs */
  /* timing */
  float    dt;                /* decimation * sim.dt                                       */
  float    sim_dt;
  int64_t  max_episode_length;/* ceil(episode_length_s / dt)                               */
  int64_t  resample_interval; /* int(resampling_time / dt)                                 */
  int64_t  push_interval;     /* ceil(push_interval_s / dt)                                */
  float    max_episode_length_s;
  float    max_push_vel_xy;
  /* normalisation */
  float    clip_actions, clip_obs;
  float    obs_scale_lin_vel, obs_scale_ang_vel, obs_scale_dof_pos, obs_scale_dof_vel, obs_scale_height;
  float    commands_scale[3];
  /* control */
  float    action_scale;
  float    p_gains[LGK_MAX_DOF], d_gains[LGK_MAX_DOF], torque_limits[LGK_MAX_DOF];
  float    default_dof_pos[LGK_MAX_DOF];
  float    dof_pos_lo[LGK_MAX_DOF], dof_pos_hi[LGK_MAX_DOF];   /* soft limits (:305-309)   */
  float    dof_vel_limits[LGK_MAX_DOF];
  /* rewards */
  float    soft_dof_vel_limit, soft_torque_limit, base_height_target, tracking_sigma, max_contact_force;
  float    termination_scale;                                   /* scale * dt              */
  LgkRewardEntry rewards[LGK_MAX_REWARDS];
  int32_t  feet_idx[LGK_MAX_FEET], pen_idx[LGK_MAX_PEN], term_idx[LGK_MAX_TERM];
  int32_t  hip_idx[LGK_MAX_HIP];                                /* [0,3,6,9] (:966)        */
  int32_t  num_hip;
  /* reset */
  float    base_init_state[13];
  /* terrain */
  float    border_size, horizontal_scale, vertical_scale;
  float    half_env_length;                                     /* terrain.env_length / 2  */
  int32_t  terrain_rows, terrain_cols;                          /* height_samples.shape    */
  int32_t  terrain_pitch;                                       /* elements per row of height_min3 */
  int32_t  max_terrain_level;                                   /* cfg.terrain.num_rows    */
  int32_t  num_terrain_types;                                   /* cfg.terrain.num_cols    */
  /* rng */
  uint64_t seed;
} LgkConsts;

/* Device buffers of one env shard.  Names follow the reference's attributes
 * (legged_robot.py:503-598, base_task.py:70-78). */
typedef struct LgkBuffers {
  /* simulator-owned (gymtorch views): read every step, written on reset / push */
  float*       root_states;        /* [N,13] pos3 quat4(xyzw) linvel3 angvel3              */
  float*       dof_state;          /* [N,D,2] (pos, vel) interleaved                       */
  const float* contact_forces;     /* [N,B,3]                                              */
  const float* torques;            /* [N,D] simulator-reported dof forces                  */
  /* env state */
  const float* actions;            /* [N,D] already clipped (lgk_clip_actions)             */
  float*       last_actions;       /* [N,D]                                                */
  float*       last_dof_vel;       /* [N,D]                                                */
  float*       last_root_vel;      /* [N,6]                                                */
  float*       commands;           /* [N,4]                                                */
  float*       feet_air_time;      /* [N,F]                                                */
  float*       episode_sums;       /* [R(+1 if termination), N] term-major                 */
  int64_t*     episode_length_buf; /* [N]                                                  */
  int64_t*     terrain_levels;     /* [N]  (NULL without custom origins)                   */
  const int64_t* terrain_types;    /* [N]                                                  */
  float*       env_origins;        /* [N,3]                                                */
  /* outputs */
  float*       obs_buf;            /* [N,O]                                                */
  float*       rew_buf;            /* [N]                                                  */
  uint8_t*     reset_buf;          /* [N] torch.bool                                       */
  uint8_t*     time_out_buf;       /* [N] torch.bool                                       */
  float*       base_lin_vel;       /* [N,3]                                                */
  float*       base_ang_vel;       /* [N,3]                                                */
  float*       projected_gravity;  /* [N,3]                                                */
  float*       measured_heights;   /* [N,H] or NULL (not materialised)                     */
  int32_t*     reset_ids;          /* [N] ascending env ids reset this step                */
  int32_t*     resample_ids;       /* [N] ascending env ids whose command was resampled    */
  int32_t*     counts;             /* [2] {n_reset, n_resample}; counts_host mirrors it    */
  float*       episode_stats;      /* [2*(R'+4)], R' = R(+1): first half = this step: sum over reset
                                      ids of episode_sums[k], then n_reset, sum(terrain_levels), N;
                                      second half = the same sums accumulated over steps until
                                      the caller zeroes them (multi-GPU all-reduce payload)      */
  float*       extras;             /* [R(+1)+2]: rew_<k> means / max_episode_length_s,
                                      terrain_level mean; only rewritten when n_reset > 0  */
  const float* ext_rewards;        /* [n_ext, N] values of LGK_REW_EXTERNAL terms or NULL  */
  /* constant tables */
  const int16_t* height_min3;      /* [rows, pitch] min(h[x,y],h[x+1,y],h[x,y+1])          */
  const float* height_points_xy;   /* [H,2] body-frame scan offsets (:802-816)             */
  const float* noise_scale_vec;    /* [O]                                                  */
  const float* terrain_origins;    /* [num_rows, num_cols, 3]                              */
  float*       command_ranges;     /* [16] two sets of (lower, upper - lower) for lin_vel_x, lin_vel_y,
                                      ang_vel_yaw, heading: [0,8) for the callback's resample (:342), [8,16)
                                      for reset_idx's (:173).  They differ only on the step on which the
                                      command curriculum widens the range inside reset_idx (:166-173) */
  /* random taps (rng_mode == LGK_RNG_TAPS), uniform [0,1), one row per env */
  const float* tap_cmd_callback;   /* [N,3]                                                */
  const float* tap_cmd_reset;      /* [N,3]                                                */
  const float* tap_push;           /* [N,2]                                                */
  const float* tap_dof;            /* [N,D]                                                */
  const float* tap_root_xy;        /* [N,2]                                                */
  const float* tap_root_vel;       /* [N,6]                                                */
  const float* tap_noise;          /* [N,O]                                                */
  const int64_t* tap_level;        /* [N] ints in [0, max_terrain_level)                   */
  uint8_t*     last_contacts;      /* [N,F] torch.bool (legged_robot.py:547): foot contact flags of the previous step; read
                                      and rewritten only with LGK_F_FILTERED_CONTACTS (else may be NULL)          */
  /* scratch owned by the caller, sized by lgk_workspace_bytes() (about 330 B per env: header, per-tile bitmasks,
     height sums, hand-over flags, the [term][N] table of reward-term values, the side rows / list / episode sums of the
     envs reset in a step), zero-initialised once, never shared between two env instances */
  void*        workspace;
} LgkBuffers;

/* bytes of LgkBuffers.workspace for N envs and R episode_sums rows */
int64_t lgk_workspace_bytes(int32_t num_envs, int32_t num_sum_rows);

/* ABI version the library was built with */
int32_t lgk_abi_version(void);

/* Derived terrain table: out[x*pitch+y] = min(h[x,y], h[x+1,y], h[x,y+1]) for x<=rows-2,
 * y<=cols-2 (legged_robot.py:845-852); pitch >= cols, multiple of 8.  Run once per terrain. */
int lgk_build_height_min3(const int16_t* height_samples, int32_t rows, int32_t cols,
                          int16_t* out, int32_t pitch, void* stream);

/* Height-field generation on the device: the sub-terrain generators Terrain calls
 * (utils/terrain.py:109-145 make_terrain, :166-187 gap / pit, and the isaacgym.terrain_utils generators behind them)
 * plus add_terrain_to_map (:147-164).  The host draws the np.random stream in the reference's order and hands the
 * resolved draws over; the kernel fills the int16 field and the env-origin heights.  p[] per type:
 *   SLOPE        peak, clip lo, clip hi                       (pyramid_sloped_terrain)
 *   SLOPE_NOISE  the same, then coarse nx, ny; draws: nx*ny levels (random_uniform_terrain)
 *   STAIRS       step cells, step height, number of rings     (pyramid_stairs_terrain)
 *   OBSTACLES    rectangles, platform x1, x2, y1, y2; draws: (i, j, w, l, height) per rectangle
 *   STONES       stone cells, gap cells, depth, stones per row, platform x1, x2, y1, y2; draws: stone heights
 *   GAP          outer half x, y, inner half x, y             PIT  depth, half size */
typedef enum LgkTerrainType {
  LGK_TER_FLAT = 0, LGK_TER_SLOPE, LGK_TER_SLOPE_NOISE, LGK_TER_STAIRS, LGK_TER_OBSTACLES, LGK_TER_STONES, LGK_TER_GAP,
  LGK_TER_PIT
} LgkTerrainType;
typedef struct LgkSubTerrain {
  int32_t type;        /* LgkTerrainType                                              */
  int32_t row, col;    /* place in the grid (level, type column)                      */
  int32_t p[10];
  int32_t draws_off;   /* first entry of this sub-terrain in draws[]                  */
} LgkSubTerrain;
typedef struct LgkTerrainGrid {
  int32_t num_rows, num_cols;   /* sub-terrains                                       */
  int32_t sub_len, sub_wid;     /* cells per sub-terrain along x / y (length_per_env_pixels, width_per_env_pixels) */
  int32_t gen_w, gen_l;         /* the generators' own frame (the reference builds width x width squares, :111-115) */
  int32_t border;               /* cells                                              */
  int32_t tot_rows, tot_cols;   /* height_samples.shape                               */
  int32_t patch_x1, patch_x2, patch_y1, patch_y2;   /* env-origin patch (:158-161)    */
  double  vertical_scale;
} LgkTerrainGrid;
/* g: HOST memory (read during the call); subs [n_sub], draws, lin (gen_w + gen_l float64: np.linspace of the noise
 * up-sampling), field [tot_rows, tot_cols] int16 (fully written, border zero), origin_z [num_rows * num_cols] float64:
 * device memory. */
int lgk_generate_terrain(const LgkTerrainGrid* g, const LgkSubTerrain* subs, int32_t n_sub, const int32_t* draws,
                         const double* lin, int16_t* field, double* origin_z, void* stream);

/* actions = clip(actions_in, +-clip_actions)            (legged_robot.py:85-86) */
int lgk_clip_actions(const float* actions_in, float* actions_out, int64_t count,
                     float clip_actions, void* stream);

/* target_poses = clip(a*action_scale + default_dof_pos, lo, hi)  (legged_robot.py:394-397) */
int lgk_compute_poses(const LgkConsts* c, const float* actions, float* target_poses, void* stream);

/* The two above in one pass (step() with position control, legged_robot.py:85-86 then :94 -> :394-397):
 * actions_out = clip(actions_in, +-c->clip_actions); target_poses = poses(actions_out). */
int lgk_clip_actions_poses(const LgkConsts* c, const float* actions_in, float* actions_out, float* target_poses,
                           void* stream);

/* torques = clip(PD(...), +-torque_limits) for control_type P / V / T (legged_robot.py:370-392) */
int lgk_compute_torques(const LgkConsts* c, const float* actions, const float* dof_state,
                        const float* last_dof_vel, float* torques_out, void* stream);

/* Go1 / Aliengo actuator net (go1.py:79-107): shifts the 5-deep joint-error / velocity
 * histories, packs model_ins [N,120] and evaluates the 30-128-128-128-3 tanh MLP per leg.
 * weights: fp32 blob W0[128,30] b0[128] W1[128,128] b1 W2[128,128] b2 W3[3,128] b3[3].
 * norm: pos_err_mean[12] pos_err_std[12] vel_mean[12] vel_std[12].
 * workspace: lgk_actuator_workspace_bytes() of caller-owned device memory, 128-byte aligned, one per weight set in
 * use at a time (the tcgen05 kernel's weight images are rebuilt into it on every call); may be NULL when
 * LGK_MLP_KERNEL selects the mma / ffma kernel.
 * num_envs == 0 is the set-up call: it launches nothing and reads the biases and the output layer of `weights`
 * back to the host (a synchronous copy -- they travel in the MLP kernel's parameter block), so that later calls may be
 * made inside a CUDA-graph capture.  The first call with envs does the same if no set-up call was made for this
 * weights pointer; after changing the weights IN PLACE make the set-up call again.  Apart from that table of host
 * copies (read-only between set-up calls, keyed by device and weights pointer) the library keeps no state: envs with
 * different nets may run on different streams of one device. */
int64_t lgk_actuator_workspace_bytes(void);
int lgk_actuator_advance(int32_t num_envs, const float* actions, const float* dof_state,
                         float* pos_err_hist, float* vel_hist, float* model_ins,
                         const float* weights, const float* norm, float* dvel_out, void* workspace, void* stream);

/* Allocation-free variant of lgk_actuator_advance.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA-graph capture, tight serving loops, harnesses
 * that forbid it).  Same inputs, outputs and arithmetic as lgk_actuator_advance with the fp32 FFMA MLP kernel
 * (LGK_MLP_KERNEL=ffma), but the [N,120] model_ins scratch is a file-scope device array inside the library
 * (LGK_STATIC_MAX_ENVS * 480 B = 1 966 080 B held by every GPU that loads the library, until exit) and nothing is
 * allocated or freed: no cudaMalloc / cudaFree, no weight-image workspace.  num_envs > LGK_STATIC_MAX_ENVS returns
 * LGK_ERR_CAPACITY; num_envs == 0 is the set-up call (resolves the array's device address, launches nothing).
 * Every call on a device shares the array: calls that overlap on the GPU (different streams) race; calls on one
 * stream are ordered and safe.  The array keeps the last call's rows; nothing in it needs clearing (rows [0, 4N)
 * are rewritten before they are read). */
#define LGK_STATIC_MAX_ENVS 4096
int lgk_actuator_advance_static(int32_t num_envs, const float* actions, const float* dof_state,
                                float* pos_err_hist, float* vel_hist, const float* weights, const float* norm,
                                float* dvel_out, void* stream);

/* ANYmal series-elastic actuator net (envs/anymal_c/anymal.py:46-81, resources/actuator_nets/
 * anydrive_v3_lstm.pt): per joint a 2-layer torch.nn.LSTM(input 2, hidden 8, batch_first) followed by
 * Linear(8 -> 1); forward(x, (h, c)) = out_scale * linear(lstm(x * in_scale, (h, c))).  Parameter layout
 * is torch's: gate rows ordered i, f, g, o.  HOST memory, read during the call only. */
#define LGK_LSTM_HID 8
typedef struct LgkLstmWeights {
  float w_ih0[4 * LGK_LSTM_HID][2],            w_hh0[4 * LGK_LSTM_HID][LGK_LSTM_HID];
  float b_ih0[4 * LGK_LSTM_HID],               b_hh0[4 * LGK_LSTM_HID];
  float w_ih1[4 * LGK_LSTM_HID][LGK_LSTM_HID], w_hh1[4 * LGK_LSTM_HID][LGK_LSTM_HID];
  float b_ih1[4 * LGK_LSTM_HID],               b_hh1[4 * LGK_LSTM_HID];
  float lin_w[LGK_LSTM_HID], lin_b;
  float in_scale[2], out_scale;
} LgkLstmWeights;

/* Anymal._compute_torques with use_actuator_network (anymal.py:71-81), one call per simulator substep:
 *   sea_input[:,0,0] = actions*action_scale + default_dof_pos - dof_pos;  sea_input[:,0,1] = dof_vel;
 *   torques, (hidden, cell) = net(sea_input, (hidden, cell))
 * hidden / cell: sea_hidden_state / sea_cell_state [2, N*D, 8], updated in place; sea_input [N*D,1,2] may be
 * NULL (not materialised); torques_out [N*D] (the reference returns them unclipped).  Uses num_envs,
 * num_dof, action_scale and default_dof_pos of LgkConsts. */
int lgk_lstm_actuator_advance(const LgkConsts* c, const LgkLstmWeights* weights_host, const float* actions,
                              const float* dof_state, float* sea_input, float* hidden, float* cell,
                              float* torques_out, void* stream);

/* Anymal.reset_idx's extra lines (anymal.py:56-60): zero the hidden and cell state of the listed envs.
 * env_ids: int32 ids on the device (NULL = 0..n-1).  The count is read on the device from count_dev when
 * that is non-NULL (the step kernel's counts[0]: no host sync), else count_host is used. */
int lgk_lstm_reset_hidden(int32_t num_envs, int32_t num_dof, const int32_t* env_ids, const int32_t* count_dev,
                          int32_t count_host, float* hidden, float* cell, void* stream);

/* Everything post_physics_step() does for every env (legged_robot.py:109-141) in one launch. */
int lgk_post_physics_step(const LgkConsts* c, const LgkBuffers* b, void* stream);

/* For reward terms evaluated OUTSIDE the library (a subclass's python _reward_<name>, bound as
 * LGK_REW_EXTERNAL): brings the buffers such a term may read to the state the reference has when
 * compute_reward() runs (legged_robot.py:118-131) -- episode_length_buf += 1, base_lin_vel, base_ang_vel,
 * projected_gravity, commands after the callback's resample / heading update, measured_heights when
 * materialised, reset_buf and time_out_buf.  The caller then fills ext_rewards and calls
 * lgk_post_physics_step with LGK_F_EP_INCREMENTED set; that call recomputes all of the above identically
 * (the random draws are keyed by step, env and site).  Not needed when every term is built in. */
int lgk_pre_reward_state(const LgkConsts* c, const LgkBuffers* b, void* stream);

/* reset_idx(env_ids) (legged_robot.py:150-193) for an ascending int64 id list on the device
 * (env_ids == NULL means arange(n)); n == 0 returns at once like the reference (:160).
 * Writes reset_ids[0..n), counts[0] = n and the episode statistics / extras. */
int lgk_reset_idx(const LgkConsts* c, const LgkBuffers* b, const int64_t* env_ids, int32_t n, void* stream);

/* reset_idx(arange(N)) as issued by BaseTask.reset() (base_task.py:111-115). */
int lgk_reset_all(const LgkConsts* c, const LgkBuffers* b, void* stream);

/* Sets the device copy of common_step_counter (legged_robot.py:119,527) held in the workspace; the step
 * epilogue increments it itself.  One tiny launch on the stream: the rest of the workspace header (both parities'
 * statistics accumulators) starts over with it, the launch sequence number behind the scan hand-over flags stays. */
int lgk_set_step_counter(void* workspace, int64_t value, void* stream);

/* Calibration probe: out[i] = (qz, qw) / norm([0,0,qz,qw]) for quat [n,4], computed exactly as the
 * step kernel does (norm_mode 1); modes 0, 2, 3 are the alternatives that were ruled out
 * (fma-contracted sums, double-precision sqrt).  Lets a test pin the formula against torch. */
int lgk_debug_yaw_quat(const float* quat, int32_t n, int32_t norm_mode, float* out, void* stream);

/* sizeof(LgkConsts) / sizeof(LgkBuffers) as compiled, for binding self-checks */
int64_t lgk_sizeof_consts(void);
int64_t lgk_sizeof_buffers(void);

/* Number of kernel launches issued by this library since load (bench bookkeeping). */
int64_t lgk_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif  /* LGK_H_ */

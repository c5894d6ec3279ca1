// extern "C" entry points of liblgk.so (declared in include/lgk.h): argument validation,
// then one or two kernel launches on the caller's stream.  No allocation, no host sync.
#include <atomic>
#include "lgk_step.cuh"

extern "C" {
int lgk_post_physics_step_impl(const LgkConsts*, const LgkBuffers*, cudaStream_t);
int lgk_chain_mode(int);
int lgk_reset_idx_impl(const LgkConsts*, const LgkBuffers*, const long long*, int, cudaStream_t);
int lgk_pre_reward_state_impl(const LgkConsts*, const LgkBuffers*, cudaStream_t);
int lgk_build_height_min3_impl(const int16_t*, int, int, int16_t*, int, cudaStream_t);
int lgk_clip_actions_impl(const float*, float*, long long, float, cudaStream_t);
int lgk_compute_poses_impl(const LgkConsts*, const float*, float*, cudaStream_t);
int lgk_clip_actions_poses_impl(const LgkConsts*, const float*, float*, float*, cudaStream_t);
int lgk_compute_torques_impl(const LgkConsts*, const float*, const float*, const float*, float*, cudaStream_t);
int lgk_debug_yaw_quat_impl(const float*, int, int, float*, cudaStream_t);
int lgk_actuator_advance_impl(int, const float*, const float*, float*, float*, float*, const float*, const float*,
                              float*, void*, cudaStream_t);
long long lgk_actuator_workspace_bytes_impl();
int lgk_actuator_advance_static_impl(int, const float*, const float*, float*, float*, const float*, const float*, float*,
                                     cudaStream_t);
int lgk_lstm_actuator_advance_impl(const LgkConsts*, const LgkLstmWeights*, const float*, const float*, float*, float*,
                                   float*, float*, cudaStream_t);
int lgk_lstm_reset_hidden_impl(int, int, const int*, const int*, int, float*, float*, cudaStream_t);
int lgk_set_step_counter_impl(void*, long long, cudaStream_t);
int lgk_generate_terrain_impl(const LgkTerrainGrid*, const LgkSubTerrain*, int, const int32_t*, const double*, int16_t*,
                              double*, cudaStream_t);
}

static std::atomic<long long> g_launches{0};

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

static int check_consts(const LgkConsts* c) {
  if (!c) return LGK_ERR_NULL;
  if (c->struct_size != sizeof(LgkConsts) || c->abi_version != LGK_ABI_VERSION) return LGK_ERR_VERSION;
  if (c->num_envs <= 0 || c->num_dof <= 0 || c->num_dof > LGK_MAX_DOF) return LGK_ERR_DIM;
  if (c->control_type < LGK_CTRL_P || c->control_type > LGK_CTRL_T) return LGK_ERR_CONTROL_TYPE;
  return LGK_OK;
}

static int check_step(const LgkConsts* c, const LgkBuffers* b) {
  int rc = check_consts(c);
  if (rc) return rc;
  if (!b) return LGK_ERR_NULL;
  if (c->num_bodies <= 0 || c->num_feet < 0 || c->num_feet > LGK_MAX_FEET || c->num_pen < 0 ||
      c->num_pen > LGK_MAX_PEN || c->num_term < 0 || c->num_term > LGK_MAX_TERM || c->num_hip < 0 ||
      c->num_hip > LGK_MAX_HIP || c->num_commands != 4 || c->num_height_points < 0)
    return LGK_ERR_DIM;
  const int has_term = (c->flags & LGK_F_HAS_TERMINATION) ? 1 : 0;
  if (c->num_rewards < 0 || c->num_rewards + has_term > LGK_MAX_REWARDS) return LGK_ERR_DIM;
  if (c->num_obs != 12 + 3 * c->num_dof + c->num_height_points) return LGK_ERR_DIM;
  for (int i = 0; i < c->num_feet; ++i) if (c->feet_idx[i] < 0 || c->feet_idx[i] >= c->num_bodies) return LGK_ERR_DIM;
  for (int i = 0; i < c->num_pen; ++i) if (c->pen_idx[i] < 0 || c->pen_idx[i] >= c->num_bodies) return LGK_ERR_DIM;
  for (int i = 0; i < c->num_term; ++i) if (c->term_idx[i] < 0 || c->term_idx[i] >= c->num_bodies) return LGK_ERR_DIM;
  for (int i = 0; i < c->num_hip; ++i) if (c->hip_idx[i] < 0 || c->hip_idx[i] >= c->num_dof) return LGK_ERR_DIM;
  int n_ext = 0;
  for (int i = 0; i < c->num_rewards; ++i) {
    if (c->rewards[i].term < 0 || c->rewards[i].term >= LGK_REW_COUNT) return LGK_ERR_REWARD;
    n_ext += c->rewards[i].term == LGK_REW_EXTERNAL;
  }
  if (n_ext > 0 && !b->ext_rewards) return LGK_ERR_NULL;
  if (c->resample_interval <= 0 || c->push_interval <= 0 || c->max_episode_length <= 0) return LGK_ERR_DIM;
  const void* req[] = {b->root_states, b->dof_state, b->contact_forces, b->torques, b->actions, b->last_actions,
                       b->last_dof_vel, b->last_root_vel, b->commands, b->feet_air_time, b->episode_length_buf,
                       b->env_origins, b->obs_buf, b->rew_buf, b->reset_buf, b->time_out_buf, b->base_lin_vel,
                       b->base_ang_vel, b->projected_gravity, b->reset_ids, b->resample_ids, b->counts,
                       b->episode_stats, b->extras, b->noise_scale_vec, b->command_ranges, b->workspace};
  for (const void* p : req) if (!p) return LGK_ERR_NULL;
  if ((c->num_rewards + has_term) > 0 && !b->episode_sums) return LGK_ERR_NULL;
  const bool scan = c->num_height_points > 0 && !(c->flags & LGK_F_TERRAIN_PLANE);
  if (c->num_height_points > 0 && !b->height_points_xy) return LGK_ERR_NULL;
  if (scan) {
    if (!b->height_min3) return LGK_ERR_TERRAIN;
    if (c->terrain_rows < 2 || c->terrain_cols < 2 || c->terrain_pitch < c->terrain_cols || !(c->horizontal_scale > 0.f))
      return LGK_ERR_TERRAIN;
  }
  if ((c->flags & LGK_F_FILTERED_CONTACTS) && !b->last_contacts) return LGK_ERR_NULL;
  if ((c->flags & LGK_F_TERRAIN_CURRICULUM) && (!b->terrain_levels || !b->terrain_types || !b->terrain_origins))
    return LGK_ERR_NULL;
  if (c->rng_mode == LGK_RNG_TAPS) {
    const void* taps[] = {b->tap_cmd_callback, b->tap_cmd_reset, b->tap_dof, b->tap_root_xy, b->tap_root_vel};
    for (const void* p : taps) if (!p) return LGK_ERR_NULL;
    if ((c->flags & LGK_F_ADD_NOISE) && !b->tap_noise) return LGK_ERR_NULL;
    if ((c->flags & LGK_F_PUSH_ROBOTS) && !b->tap_push) return LGK_ERR_NULL;
    if ((c->flags & LGK_F_TERRAIN_CURRICULUM) && !b->tap_level) return LGK_ERR_NULL;
  } else if (c->rng_mode != LGK_RNG_PHILOX) {
    return LGK_ERR_DIM;
  }
  // vectorised rows: [N,4] commands and, for D % 4 == 0, the [N,D] / [N,D,2] rows
  if (!aligned16(b->commands) || !aligned16(b->contact_forces) || !aligned16(b->workspace)) return LGK_ERR_ALIGN;
  if (c->num_dof % 4 == 0 &&
      (!aligned16(b->actions) || !aligned16(b->last_actions) || !aligned16(b->torques) ||
       !aligned16(b->last_dof_vel) || !aligned16(b->dof_state)))
    return LGK_ERR_ALIGN;
  if ((reinterpret_cast<uintptr_t>(b->last_root_vel) & 7u) != 0) return LGK_ERR_ALIGN;
  return LGK_OK;
}

extern "C" {

int32_t lgk_abi_version(void) { return LGK_ABI_VERSION; }
int64_t lgk_sizeof_consts(void) { return (int64_t)sizeof(LgkConsts); }
int64_t lgk_sizeof_buffers(void) { return (int64_t)sizeof(LgkBuffers); }
int64_t lgk_launch_count(void) { return g_launches.load(); }

int64_t lgk_workspace_bytes(int32_t num_envs, int32_t num_sum_rows) {
  (void)num_sum_rows;
  // header, the two per-tile bitmask arrays, one float per env (the scan kernel's height sums), one flag per scan tile
  return (int64_t)lgk_ws_total_bytes(num_envs);
}

int lgk_build_height_min3(const int16_t* height_samples, int32_t rows, int32_t cols, int16_t* out, int32_t pitch,
                          void* stream) {
  if (!height_samples || !out) return LGK_ERR_NULL;
  if (rows < 2 || cols < 2 || pitch < cols) return LGK_ERR_DIM;
  g_launches += 1;
  return lgk_build_height_min3_impl(height_samples, rows, cols, out, pitch, (cudaStream_t)stream);
}

int lgk_generate_terrain(const LgkTerrainGrid* g, const LgkSubTerrain* subs, int32_t n_sub, const int32_t* draws,
                         const double* lin, int16_t* field, double* origin_z, void* stream) {
  if (!g || !subs || !draws || !lin || !field || !origin_z) return LGK_ERR_NULL;
  if (n_sub <= 0 || n_sub > g->num_rows * g->num_cols || g->sub_len <= 0 || g->sub_wid <= 0 || g->gen_w != g->sub_len ||
      g->gen_l != g->sub_wid || g->border < 0 || g->tot_rows != g->num_rows * g->sub_len + 2 * g->border ||
      g->tot_cols != g->num_cols * g->sub_wid + 2 * g->border || g->patch_x1 < 0 || g->patch_x2 > g->sub_len ||
      g->patch_x1 >= g->patch_x2 || g->patch_y1 < 0 || g->patch_y2 > g->sub_wid || g->patch_y1 >= g->patch_y2)
    return LGK_ERR_DIM;
  g_launches += 2;
  return lgk_generate_terrain_impl(g, subs, n_sub, draws, lin, field, origin_z, (cudaStream_t)stream);
}

int lgk_clip_actions(const float* in, float* out, int64_t count, float clip_actions, void* stream) {
  if (!in || !out) return LGK_ERR_NULL;
  if (count < 0) return LGK_ERR_DIM;
  g_launches += 1;
  return lgk_clip_actions_impl(in, out, count, clip_actions, (cudaStream_t)stream);
}

int lgk_compute_poses(const LgkConsts* c, const float* actions, float* target_poses, void* stream) {
  int rc = check_consts(c);
  if (rc) return rc;
  if (!actions || !target_poses) return LGK_ERR_NULL;
  g_launches += 1;
  return lgk_compute_poses_impl(c, actions, target_poses, (cudaStream_t)stream);
}

int lgk_clip_actions_poses(const LgkConsts* c, const float* actions_in, float* actions_out, float* target_poses,
                           void* stream) {
  int rc = check_consts(c);
  if (rc) return rc;
  if (!actions_in || !actions_out || !target_poses) return LGK_ERR_NULL;
  g_launches += 1;
  return lgk_clip_actions_poses_impl(c, actions_in, actions_out, target_poses, (cudaStream_t)stream);
}

int lgk_compute_torques(const LgkConsts* c, const float* actions, const float* dof_state, const float* last_dof_vel,
                        float* torques_out, void* stream) {
  int rc = check_consts(c);
  if (rc) return rc;
  if (!actions || !dof_state || !torques_out) return LGK_ERR_NULL;
  if (c->control_type == LGK_CTRL_V && !last_dof_vel) return LGK_ERR_NULL;
  if ((reinterpret_cast<uintptr_t>(dof_state) & 7u) != 0) return LGK_ERR_ALIGN;
  g_launches += 1;
  return lgk_compute_torques_impl(c, actions, dof_state, last_dof_vel, torques_out, (cudaStream_t)stream);
}

int64_t lgk_actuator_workspace_bytes(void) { return (int64_t)lgk_actuator_workspace_bytes_impl(); }

int lgk_actuator_advance(int32_t num_envs, const float* actions, const float* dof_state, float* pos_err_hist,
                         float* vel_hist, float* model_ins, const float* weights, const float* norm, float* dvel_out,
                         void* workspace, void* stream) {
  if (num_envs < 0) return LGK_ERR_DIM;
  if (!actions || !dof_state || !pos_err_hist || !vel_hist || !model_ins || !weights || !norm || !dvel_out)
    return LGK_ERR_NULL;
  if ((reinterpret_cast<uintptr_t>(dof_state) & 7u) != 0) return LGK_ERR_ALIGN;
  if (workspace && (reinterpret_cast<uintptr_t>(workspace) & 127u) != 0) return LGK_ERR_ALIGN;   // TMA source
  if (num_envs > 0) g_launches += 2;   /* zero envs: set-up only, nothing is launched */
  return lgk_actuator_advance_impl(num_envs, actions, dof_state, pos_err_hist, vel_hist, model_ins, weights, norm,
                                   dvel_out, workspace, (cudaStream_t)stream);
}

int lgk_actuator_advance_static(int32_t num_envs, const float* actions, const float* dof_state, float* pos_err_hist,
                                float* vel_hist, const float* weights, const float* norm, float* dvel_out,
                                void* stream) {
  if (num_envs < 0) return LGK_ERR_DIM;
  if (num_envs > LGK_STATIC_MAX_ENVS) return LGK_ERR_CAPACITY;   // the scratch array is fixed-size: no fallback
  if (!actions || !dof_state || !pos_err_hist || !vel_hist || !weights || !norm || !dvel_out) return LGK_ERR_NULL;
  if ((reinterpret_cast<uintptr_t>(dof_state) & 7u) != 0) return LGK_ERR_ALIGN;
  if (num_envs > 0) g_launches += 2;
  return lgk_actuator_advance_static_impl(num_envs, actions, dof_state, pos_err_hist, vel_hist, weights, norm, dvel_out,
                                          (cudaStream_t)stream);
}

int lgk_lstm_actuator_advance(const LgkConsts* c, const LgkLstmWeights* weights_host, const float* actions,
                              const float* dof_state, float* sea_input, float* hidden, float* cell,
                              float* torques_out, void* stream) {
  int rc = check_consts(c);
  if (rc) return rc;
  if (!weights_host || !actions || !dof_state || !hidden || !cell || !torques_out) return LGK_ERR_NULL;
  if (!aligned16(hidden) || !aligned16(cell) || (reinterpret_cast<uintptr_t>(dof_state) & 7u) != 0 ||
      (sea_input && (reinterpret_cast<uintptr_t>(sea_input) & 7u) != 0))
    return LGK_ERR_ALIGN;
  g_launches += 1;
  return lgk_lstm_actuator_advance_impl(c, weights_host, actions, dof_state, sea_input, hidden, cell, torques_out,
                                        (cudaStream_t)stream);
}

int lgk_lstm_reset_hidden(int32_t num_envs, int32_t num_dof, const int32_t* env_ids, const int32_t* count_dev,
                          int32_t count_host, float* hidden, float* cell, void* stream) {
  if (!hidden || !cell) return LGK_ERR_NULL;
  if (num_envs <= 0 || num_dof <= 0 || num_dof > LGK_MAX_DOF) return LGK_ERR_DIM;
  if (!count_dev && (count_host < 0 || count_host > num_envs)) return LGK_ERR_DIM;
  if (!aligned16(hidden) || !aligned16(cell)) return LGK_ERR_ALIGN;
  if (!count_dev && count_host == 0) return LGK_OK;
  g_launches += 1;
  return lgk_lstm_reset_hidden_impl(num_envs, num_dof, env_ids, count_dev, count_host, hidden, cell,
                                    (cudaStream_t)stream);
}

int lgk_post_physics_step(const LgkConsts* c, const LgkBuffers* b, void* stream) {
  int rc = check_step(c, b);
  if (rc) return rc;
  // default: step kernel, tail kernel (scan, rewards, statistics); LGK_CHAIN=3: (scan kernel,) step kernel, finish kernel
  g_launches += lgk_chain_mode(c->num_envs) ? 2 : 2 + (c->num_height_points > 0 ? 1 : 0);
  return lgk_post_physics_step_impl(c, b, (cudaStream_t)stream);
}

int lgk_pre_reward_state(const LgkConsts* c, const LgkBuffers* b, void* stream) {
  int rc = check_step(c, b);
  if (rc) return rc;
  if (c->flags & LGK_F_EP_INCREMENTED) return LGK_ERR_DIM;   // would count the step twice
  g_launches += 1;
  return lgk_pre_reward_state_impl(c, b, (cudaStream_t)stream);
}

int lgk_reset_idx(const LgkConsts* c, const LgkBuffers* b, const int64_t* env_ids, int32_t n, void* stream) {
  int rc = check_step(c, b);
  if (rc) return rc;
  if (n < 0 || n > c->num_envs) return LGK_ERR_DIM;
  if (n == 0) return LGK_OK;
  g_launches += 1 + (((c->flags & LGK_F_TERRAIN_CURRICULUM) && b->terrain_levels) ? 1 : 0);
  return lgk_reset_idx_impl(c, b, (const long long*)env_ids, n, (cudaStream_t)stream);
}

int lgk_reset_all(const LgkConsts* c, const LgkBuffers* b, void* stream) {
  if (!c) return LGK_ERR_NULL;
  return lgk_reset_idx(c, b, nullptr, c->num_envs, stream);
}

int lgk_debug_yaw_quat(const float* quat, int32_t n, int32_t norm_mode, float* out, void* stream) {
  if (!quat || !out) return LGK_ERR_NULL;
  if (n <= 0 || norm_mode < 0 || norm_mode > 3) return LGK_ERR_DIM;
  g_launches += 1;
  return lgk_debug_yaw_quat_impl(quat, n, norm_mode, out, (cudaStream_t)stream);
}

/* device-side common_step_counter (legged_robot.py:119,527) */
int lgk_set_step_counter(void* workspace, int64_t value, void* stream) {
  if (!workspace) return LGK_ERR_NULL;
  // The whole header goes with it, except the launch sequence number the scan flags are stamped with: the step
  // epilogue zeroes only the accumulators of the OTHER step parity, so a counter set by hand (load_state_dict into an
  // env that has stepped) must not leave this parity's sums behind.  One tiny launch; capturable.
  g_launches += 1;
  return lgk_set_step_counter_impl(workspace, value, (cudaStream_t)stream);
}

}  // extern "C"

// Fused post_physics_step kernel (sm_100a).
//
// One launch replaces everything LeggedRobot.post_physics_step() does per env
// (reference legged_robot.py:109-141): episode counters, body-frame velocities / projected
// gravity, command resampling + heading command, the yaw-rotated terrain height scan, pushes,
// terminations, every active _reward_* term with in-place episode sums, resets (terrain
// curriculum, dof / root state, commands), the noisy scaled observation row, last_* copies,
// ascending compacted reset / resample id lists and the episode statistics behind extras.
//
// Work decomposition: a warp owns `epw` consecutive envs.
//   S0  lane e does the per-env scalar work of env e (contacts staged through shared memory so
//       global traffic stays coalesced; 12-wide dof rows read as float4).
//   A   the warp walks its envs one at a time; lanes cover the 48-wide observation head and the
//       H scan points of that env (coalesced obs stores, one int16 gather per point from the
//       precomputed min-of-three-neighbours table).
//   S1  lane e sums the reward terms of env e in the reference's order and updates episode sums.
//   The last CTA to finish scans the per-warp reset / resample bitmasks into ascending id lists
//   and finalises the statistics.
#include "lgk_common.cuh"

struct StepArgs {
  LgkConsts c;
  LgkBuffers b;
  int epw;         // envs per warp (4, 8, 16 or 32)
  int head_stride; // shared-memory row stride of the staged observation head (odd)
  int row_floats;  // floats of shared memory per env  = max(3B, head_stride)
  int num_warps;   // warps with work
};

// ---------------------------------------------------------------------------------------------
struct ContactSummary {
  bool term;          // any termination body with |f| > 1           (:146)
  float collisions;   // # penalised bodies with |f| > 0.1            (:897)
  unsigned contact_bits;  // feet with f_z > 1                        (:943)
  int nofly_contacts; // feet with f_z > 0.1                          (cassie.py:44)
  bool stumble;       // any foot with |f_xy| > 5 |f_z|               (:953)
  float feet_force_pen;   // sum max(|f| - max_contact_force, 0)      (:962)
};

__device__ __forceinline__ float norm3(float x, float y, float z) {
  return sqrtf(fmaf(z, z, fmaf(y, y, x * x)));
}

__device__ __forceinline__ ContactSummary summarise_contacts(const LgkConsts& c, const float* __restrict__ cf) {
  ContactSummary s;
  s.term = false;
  for (int t = 0; t < c.num_term; ++t) {
    const float* f = cf + 3 * c.term_idx[t];
    s.term |= norm3(f[0], f[1], f[2]) > 1.0f;
  }
  s.collisions = 0.f;
  for (int p = 0; p < c.num_pen; ++p) {
    const float* f = cf + 3 * c.pen_idx[p];
    s.collisions += (norm3(f[0], f[1], f[2]) > 0.1f) ? 1.f : 0.f;
  }
  s.contact_bits = 0;
  s.nofly_contacts = 0;
  s.stumble = false;
  s.feet_force_pen = 0.f;
  for (int k = 0; k < c.num_feet; ++k) {
    const float* f = cf + 3 * c.feet_idx[k];
    const float fx = f[0], fy = f[1], fz = f[2];
    if (fz > 1.0f) s.contact_bits |= 1u << k;
    s.nofly_contacts += (fz > 0.1f) ? 1 : 0;
    s.stumble |= sqrtf(fmaf(fy, fy, fx * fx)) > 5.0f * fabsf(fz);
    s.feet_force_pen += fmaxf(norm3(fx, fy, fz) - c.max_contact_force, 0.f);
  }
  return s;
}

// _resample_commands for one env (legged_robot.py:354-368); cr = (lo, hi-lo) pairs
__device__ __forceinline__ void resample_command(float4& cmd, const float* __restrict__ cr, bool heading_mode,
                                                 float u0, float u1, float u2) {
  cmd.x = rand_affine(cr[1], cr[0], u0);
  cmd.y = rand_affine(cr[3], cr[2], u1);
  if (heading_mode) cmd.w = rand_affine(cr[7], cr[6], u2);
  else cmd.z = rand_affine(cr[5], cr[4], u2);
  const float keep = (sqrtf(fmaf(cmd.y, cmd.y, cmd.x * cmd.x)) > 0.2f) ? 1.f : 0.f;
  cmd.x *= keep;
  cmd.y *= keep;
}

struct DofAcc {   // per-env sums over joints feeding the reward terms (:874-916, :956-966)
  float tq2, energy, qd2, acc2, arate, poslim, vellim, tqlim, stand, hip;
};

template <int DT>
__global__ void __launch_bounds__(LGK_CTA_THREADS)
lgk_step_kernel(const __grid_constant__ StepArgs a) {
  const LgkConsts& c = a.c;
  const LgkBuffers& b = a.b;
  constexpr bool VEC = (DT > 0) && (DT % 4 == 0);
  const int D = DT > 0 ? DT : c.num_dof;
  const int N = c.num_envs;
  const int B3 = c.num_bodies * 3;
  const int H = c.num_height_points;
  const int O = c.num_obs;
  const int OH = 12 + 3 * D;               // width of the observation head (:217-224)
  const unsigned flags = c.flags;
  const bool taps = c.rng_mode == LGK_RNG_TAPS;
  const bool heading_mode = flags & LGK_F_HEADING_COMMAND;

  extern __shared__ float smem[];
  const int lane = threadIdx.x & 31;
  const int wic = threadIdx.x >> 5;
  const int gwarp = blockIdx.x * LGK_WARPS_PER_CTA + wic;
  const int epw = a.epw;
  const int env0 = gwarp * epw;
  const int tile_envs = max(0, min(epw, N - env0));
  float* wsm = smem + (size_t)wic * epw * a.row_floats;

  LgkWorkspace* ws = reinterpret_cast<LgkWorkspace*>(b.workspace);
  const unsigned long long csc = ws->step_counter + 1ull;   // common_step_counter after += 1 (:119)
  const int parity = (int)(csc & 1ull);
  const bool push_step = (flags & LGK_F_PUSH_ROBOTS) && (csc % (unsigned long long)c.push_interval == 0ull);

  const bool active = lane < tile_envs;
  const int env = env0 + (active ? lane : 0);
  const unsigned env_global = (unsigned)(c.env_id_offset + env);

  // ---- stage this warp's contact rows (contiguous, coalesced) --------------------------
  {
    const int total = tile_envs * B3;
    const float* src = b.contact_forces + (size_t)env0 * B3;
    if (((B3 * epw) & 3) == 0 && tile_envs == epw) {
      const float4* s4 = reinterpret_cast<const float4*>(src);
      float4* d4 = reinterpret_cast<float4*>(wsm);
      for (int i = lane; i < (total >> 2); i += 32) d4[i] = __ldg(s4 + i);
    } else {
      for (int i = lane; i < total; i += 32) wsm[i] = __ldg(src + i);
    }
  }
  __syncwarp();

  // per-env values that outlive S0
  float root_x = 0.f, root_y = 0.f, z_pre = 0.f, zobs = 0.f, yaw_z = 0.f, yaw_w = 1.f;
  float blv[3] = {0, 0, 0}, bav[3] = {0, 0, 0}, grav[3] = {0, 0, 0};
  float4 cmd_rew = make_float4(0, 0, 0, 0);
  ContactSummary cs = {};
  DofAcc acc = {};
  float r_air = 0.f;
  bool reset = false, time_out = false, resample_cb = false;
  float head[12];   // first 12 obs columns; the 3D joint columns go straight to shared memory
  int level_contrib = 0;

  if (active) cs = summarise_contacts(c, wsm + lane * B3);
  __syncwarp();   // every lane is done with the contact rows: the region is reused for obs heads
  float* head_row = wsm + lane * a.head_stride;

  if (active) {
    // ---- prologue (:118-125) ------------------------------------------------------------
    float root[13];
    const float* rp = b.root_states + (size_t)env * 13;
#pragma unroll
    for (int i = 0; i < 13; ++i) root[i] = rp[i];
    long long ep = b.episode_length_buf[env] + 1;
    const float q[4] = {root[3], root[4], root[5], root[6]};
    quat_rotate_inverse(q, root[7], root[8], root[9], blv);
    quat_rotate_inverse(q, root[10], root[11], root[12], bav);
    quat_rotate_inverse(q, 0.f, 0.f, -1.f, grav);
    float4 cmd = *reinterpret_cast<const float4*>(b.commands + (size_t)env * 4);
    const float* cr = b.command_ranges;

    // ---- _post_physics_step_callback (:337-352) -----------------------------------------
    resample_cb = (ep % c.resample_interval) == 0;
    if (resample_cb) {
      float u0, u1, u2;
      if (taps) {
        const float* t = b.tap_cmd_callback + (size_t)env * 3;
        u0 = t[0]; u1 = t[1]; u2 = t[2];
      } else {
        const uint4 r = lgk_philox(c.seed, csc, env_global, LGK_SITE_CMD_CB);
        u0 = u01(r.x); u1 = u01(r.y); u2 = u01(r.z);
      }
      resample_command(cmd, cr, heading_mode, u0, u1, u2);
    }
    if (heading_mode) {
      float fwd[3];
      quat_apply(q, 1.f, 0.f, 0.f, fwd);
      const float heading = atan2f(fwd[1], fwd[0]);
      cmd.z = clampf(0.5f * wrap_to_pi(cmd.w - heading), -1.f, 1.f);
    }
    cmd_rew = cmd;
    // yaw-only quaternion of quat_apply_yaw (utils/math.py:38-42): normalize([0,0,qz,qw])
    {
      // x.norm(p=2, dim=-1): separately rounded squares, correctly rounded sqrt (same on torch CPU/CUDA)
      const float n = fmaxf(sqrtf(add_rn(mul_rn(q[2], q[2]), mul_rn(q[3], q[3]))), 1e-9f);
      yaw_z = __fdiv_rn(q[2], n);
      yaw_w = __fdiv_rn(q[3], n);
    }
    root_x = root[0]; root_y = root[1]; z_pre = root[2];
    float rvel[6] = {root[7], root[8], root[9], root[10], root[11], root[12]};
    if (push_step) {   // _push_robots (:436-441): every env, after the body-frame velocities
      float u0, u1;
      if (taps) { u0 = b.tap_push[(size_t)env * 2]; u1 = b.tap_push[(size_t)env * 2 + 1]; }
      else { const uint4 r = lgk_philox(c.seed, csc, env_global, LGK_SITE_PUSH); u0 = u01(r.x); u1 = u01(r.y); }
      const float range = 2.0f * c.max_push_vel_xy;
      rvel[0] = rand_affine(range, -c.max_push_vel_xy, u0);
      rvel[1] = rand_affine(range, -c.max_push_vel_xy, u1);
    }

    // ---- check_termination (:143-148) ---------------------------------------------------
    time_out = ep > c.max_episode_length;
    reset = cs.term || time_out;

    // ---- terrain curriculum + new origin for reset envs (:443-463) --------------------
    float origin[3] = {0, 0, 0};
    const bool has_levels = b.terrain_levels != nullptr;
    long long level = has_levels ? b.terrain_levels[env] : 0;
    if (reset) {
      const float* op = b.env_origins + (size_t)env * 3;
      origin[0] = op[0]; origin[1] = op[1]; origin[2] = op[2];
      if ((flags & LGK_F_TERRAIN_CURRICULUM) && (flags & LGK_F_INIT_DONE) && has_levels) {
        const float dx = root[0] - origin[0], dy = root[1] - origin[1];
        const float dist = sqrtf(fmaf(dy, dy, dx * dx));
        const bool up = dist > c.half_env_length;
        const float cn = sqrtf(fmaf(cmd.y, cmd.y, cmd.x * cmd.x));
        const bool down = (dist < mul_rn(mul_rn(cn, c.max_episode_length_s), 0.5f)) && !up;
        level += (up ? 1 : 0) - (down ? 1 : 0);
        if (level >= c.max_terrain_level) {
          if (taps) level = b.tap_level[env];
          else level = (long long)(lgk_philox(c.seed, csc, env_global, LGK_SITE_LEVEL).x % (unsigned)c.max_terrain_level);
        } else if (level < 0) {
          level = 0;
        }
        b.terrain_levels[env] = level;
        const float* to = b.terrain_origins + ((size_t)level * c.num_terrain_types + (size_t)b.terrain_types[env]) * 3;
        origin[0] = to[0]; origin[1] = to[1]; origin[2] = to[2];
        float* ow = b.env_origins + (size_t)env * 3;
        ow[0] = origin[0]; ow[1] = origin[1]; ow[2] = origin[2];
      }
    }
    level_contrib = (int)level;

    // ---- joints: reward sums, reset values, obs head, last_* (:176-177,:399-408,:136-137) --
    float feet_air[LGK_MAX_FEET];
    const float inv_dt = 1.0f / c.dt;
    const float* act_p = b.actions + (size_t)env * D;
    const float* lact_p = b.last_actions + (size_t)env * D;
    const float* tq_p = b.torques + (size_t)env * D;
    const float* ldv_p = b.last_dof_vel + (size_t)env * D;
    float* dof_p = b.dof_state + (size_t)env * D * 2;
    float* lact_w = b.last_actions + (size_t)env * D;
    float* ldv_w = b.last_dof_vel + (size_t)env * D;

    unsigned hip_mask = 0u;
    for (int h = 0; h < c.num_hip; ++h) hip_mask |= 1u << c.hip_idx[h];
    auto joint = [&](int d, float av, float lav, float tqv, float ldvv, float& p, float& v, float u) {
      // rewards see the pre-reset joint state
      acc.tq2 = fmaf(tqv, tqv, acc.tq2);
      const float pw = tqv * v;
      acc.energy = fmaf(pw, pw, acc.energy);
      acc.qd2 = fmaf(v, v, acc.qd2);
      const float ja = (ldvv - v) * inv_dt;
      acc.acc2 = fmaf(ja, ja, acc.acc2);
      const float da = lav - av;
      acc.arate = fmaf(da, da, acc.arate);
      acc.poslim += -fminf(p - c.dof_pos_lo[d], 0.f) + fmaxf(p - c.dof_pos_hi[d], 0.f);
      acc.vellim += clampf(fabsf(v) - c.dof_vel_limits[d] * c.soft_dof_vel_limit, 0.f, 1.f);
      acc.tqlim += fmaxf(fabsf(tqv) - c.torque_limits[d] * c.soft_torque_limit, 0.f);
      const float dev = fabsf(p - c.default_dof_pos[d]);
      acc.stand += dev;
      if ((hip_mask >> d) & 1u) acc.hip += dev;
      if (reset) {   // _reset_dofs: default * U(0.5, 1.5), zero velocity
        p = mul_rn(c.default_dof_pos[d], rand_affine(1.0f, 0.5f, u));
        v = 0.f;
      }
      head_row[12 + d] = (p - c.default_dof_pos[d]) * c.obs_scale_dof_pos;
      head_row[12 + D + d] = v * c.obs_scale_dof_vel;
      head_row[12 + 2 * D + d] = av;
    };

    if constexpr (VEC) {
#pragma unroll
      for (int d0 = 0; d0 < DT; d0 += 4) {
        const float4 a4 = __ldg(reinterpret_cast<const float4*>(act_p + d0));
        const float4 la4 = *reinterpret_cast<const float4*>(lact_p + d0);
        const float4 tq4 = __ldg(reinterpret_cast<const float4*>(tq_p + d0));
        const float4 ldv4 = *reinterpret_cast<const float4*>(ldv_p + d0);
        float4 s0 = *reinterpret_cast<const float4*>(dof_p + 2 * d0);       // p0 v0 p1 v1
        float4 s1 = *reinterpret_cast<const float4*>(dof_p + 2 * d0 + 4);   // p2 v2 p3 v3
        float u[4] = {0, 0, 0, 0};
        if (reset) {
          if (taps) {
            const float* t = b.tap_dof + (size_t)env * D + d0;
            u[0] = t[0]; u[1] = t[1]; u[2] = t[2]; u[3] = t[3];
          } else {
            const uint4 r = lgk_philox(c.seed, csc, env_global, LGK_SITE_DOF + (d0 >> 2));
            u[0] = u01(r.x); u[1] = u01(r.y); u[2] = u01(r.z); u[3] = u01(r.w);
          }
        }
        joint(d0 + 0, a4.x, la4.x, tq4.x, ldv4.x, s0.x, s0.y, u[0]);
        joint(d0 + 1, a4.y, la4.y, tq4.y, ldv4.y, s0.z, s0.w, u[1]);
        joint(d0 + 2, a4.z, la4.z, tq4.z, ldv4.z, s1.x, s1.y, u[2]);
        joint(d0 + 3, a4.w, la4.w, tq4.w, ldv4.w, s1.z, s1.w, u[3]);
        *reinterpret_cast<float4*>(lact_w + d0) = a4;                              // last_actions = actions
        *reinterpret_cast<float4*>(ldv_w + d0) = make_float4(s0.y, s0.w, s1.y, s1.w);  // last_dof_vel = dof_vel
        if (reset) {
          *reinterpret_cast<float4*>(dof_p + 2 * d0) = s0;
          *reinterpret_cast<float4*>(dof_p + 2 * d0 + 4) = s1;
        }
      }
    } else {
      for (int d = 0; d < D; ++d) {
        const float av = act_p[d], lav = lact_p[d], tqv = tq_p[d], ldvv = ldv_p[d];
        float p = dof_p[2 * d], v = dof_p[2 * d + 1];
        float u = 0.f;
        if (reset) {
          if (taps) u = b.tap_dof[(size_t)env * D + d];
          else {
            const uint4 r = lgk_philox(c.seed, csc, env_global, LGK_SITE_DOF + (d >> 2));
            const unsigned rr[4] = {r.x, r.y, r.z, r.w};
            u = u01(rr[d & 3]);
          }
        }
        joint(d, av, lav, tqv, ldvv, p, v, u);
        lact_w[d] = av;
        ldv_w[d] = v;
        if (reset) { dof_p[2 * d] = p; dof_p[2 * d + 1] = v; }
      }
    }

    // ---- feet_air_time reward state (:941-949); only advanced when the term is active -------
    bool air_active = false;
    for (int i = 0; i < c.num_rewards; ++i) air_active |= (c.rewards[i].term == LGK_REW_FEET_AIR_TIME);
    {
      const int F = c.num_feet;
      float* fa = b.feet_air_time + (size_t)env * F;
      const float cn = sqrtf(fmaf(cmd_rew.y, cmd_rew.y, cmd_rew.x * cmd_rew.x));
      for (int k = 0; k < F; ++k) {
        float t = fa[k];
        if (air_active) {
          const bool contact = (cs.contact_bits >> k) & 1u;
          const bool first = (t > 0.f) && contact;
          t = add_rn(t, c.dt);
          r_air += first ? (t - 0.5f) : 0.f;
          if (contact) t = 0.f;
        }
        feet_air[k] = reset ? 0.f : t;
      }
      r_air *= (cn > 0.1f) ? 1.f : 0.f;
      if (air_active || reset)
        for (int k = 0; k < F; ++k) fa[k] = feet_air[k];
    }

    // ---- reset_idx for this env (:150-193) --------------------------------------------------
    if (reset) {
      float u[8];
      if (taps) {
        u[0] = b.tap_root_xy[(size_t)env * 2]; u[1] = b.tap_root_xy[(size_t)env * 2 + 1];
        for (int i = 0; i < 6; ++i) u[2 + i] = b.tap_root_vel[(size_t)env * 6 + i];
      } else {
        const uint4 r0 = lgk_philox(c.seed, csc, env_global, LGK_SITE_ROOT);
        const uint4 r1 = lgk_philox(c.seed, csc, env_global, LGK_SITE_ROOT + 1);
        u[0] = u01(r0.x); u[1] = u01(r0.y); u[2] = u01(r0.z); u[3] = u01(r0.w);
        u[4] = u01(r1.x); u[5] = u01(r1.y); u[6] = u01(r1.z); u[7] = u01(r1.w);
      }
#pragma unroll
      for (int i = 0; i < 13; ++i) root[i] = c.base_init_state[i];
      root[0] = add_rn(root[0], origin[0]);
      root[1] = add_rn(root[1], origin[1]);
      root[2] = add_rn(root[2], origin[2]);
      if (flags & LGK_F_CUSTOM_ORIGINS) {   // xy within 1 m of the platform centre (:425)
        root[0] = add_rn(root[0], rand_affine(2.0f, -1.0f, u[0]));
        root[1] = add_rn(root[1], rand_affine(2.0f, -1.0f, u[1]));
      }
#pragma unroll
      for (int i = 0; i < 6; ++i) { rvel[i] = rand_affine(1.0f, -0.5f, u[2 + i]); root[7 + i] = rvel[i]; }
      float* rw = b.root_states + (size_t)env * 13;
#pragma unroll
      for (int i = 0; i < 13; ++i) rw[i] = root[i];
      // commands for the new episode (:173)
      float c0, c1, c2;
      if (taps) { const float* t = b.tap_cmd_reset + (size_t)env * 3; c0 = t[0]; c1 = t[1]; c2 = t[2]; }
      else { const uint4 r = lgk_philox(c.seed, csc, env_global, LGK_SITE_CMD_RESET); c0 = u01(r.x); c1 = u01(r.y); c2 = u01(r.z); }
      resample_command(cmd, cr, heading_mode, c0, c1, c2);
      ep = 0;
    } else if (push_step) {
      float* rw = b.root_states + (size_t)env * 13;
      rw[7] = rvel[0];
      rw[8] = rvel[1];
    }
    zobs = sub_rn(root[2], 0.5f);   // root z after a possible reset, minus 0.5 (:227)

    // ---- per-env stores -------------------------------------------------------------------------
    b.episode_length_buf[env] = ep;
    b.reset_buf[env] = reset ? 1 : 0;
    b.time_out_buf[env] = time_out ? 1 : 0;
    *reinterpret_cast<float4*>(b.commands + (size_t)env * 4) = cmd;
    {
      float* p = b.base_lin_vel + (size_t)env * 3; p[0] = blv[0]; p[1] = blv[1]; p[2] = blv[2];
      p = b.base_ang_vel + (size_t)env * 3;        p[0] = bav[0]; p[1] = bav[1]; p[2] = bav[2];
      p = b.projected_gravity + (size_t)env * 3;   p[0] = grav[0]; p[1] = grav[1]; p[2] = grav[2];
      float2* lr = reinterpret_cast<float2*>(b.last_root_vel + (size_t)env * 6);   // (:138)
      lr[0] = make_float2(rvel[0], rvel[1]); lr[1] = make_float2(rvel[2], rvel[3]); lr[2] = make_float2(rvel[4], rvel[5]);
    }
    // obs head columns 0..11 (:217-220), post-reset commands
    head[0] = blv[0] * c.obs_scale_lin_vel; head[1] = blv[1] * c.obs_scale_lin_vel; head[2] = blv[2] * c.obs_scale_lin_vel;
    head[3] = bav[0] * c.obs_scale_ang_vel; head[4] = bav[1] * c.obs_scale_ang_vel; head[5] = bav[2] * c.obs_scale_ang_vel;
    head[6] = grav[0]; head[7] = grav[1]; head[8] = grav[2];
    head[9] = cmd.x * c.commands_scale[0]; head[10] = cmd.y * c.commands_scale[1]; head[11] = cmd.z * c.commands_scale[2];
#pragma unroll
    for (int i = 0; i < 12; ++i) head_row[i] = head[i];
  }
  __syncwarp();

  // ---- warp-wide bookkeeping: bitmasks for the ordered compaction, level sum -------------------
  const unsigned reset_mask = __ballot_sync(LGK_FULL_MASK, active && reset);
  const unsigned resample_mask = __ballot_sync(LGK_FULL_MASK, active && resample_cb);
  unsigned* mask_base = reinterpret_cast<unsigned*>(reinterpret_cast<char*>(ws) + LGK_WS_HEADER_BYTES);
  const int max_warps = lgk_max_warps(N);
  if (lane == 0 && gwarp < max_warps) {
    mask_base[gwarp] = reset_mask;
    mask_base[max_warps + gwarp] = resample_mask;
  }
  if ((flags & LGK_F_TERRAIN_CURRICULUM) && b.terrain_levels != nullptr) {
    const int s = warp_sum_int(active ? level_contrib : 0);
    if (lane == 0 && s != 0) atomicAdd(&ws->level_sum[parity], s);
  }

  // ---- phase A: observation row + height scan, one env at a time, lanes over columns ----------
  float bh_sum = 0.f;   // sum_h (root_z - measured_height) of this lane's env (:871)
  {
    const bool add_noise = flags & LGK_F_ADD_NOISE;
    const bool plane = flags & LGK_F_TERRAIN_PLANE;
    const bool mul_rcp = c.index_math == LGK_INDEX_MUL_RCP;
    const float inv_hs = 1.0f / c.horizontal_scale;
    const int max_px = c.terrain_rows - 2, max_py = c.terrain_cols - 2;
    // this lane's scan offsets (same for every env)
    float px_l[8], py_l[8];
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const int h = it * 32 + lane;
      const bool ok = h < H;
      px_l[it] = ok ? __ldg(b.height_points_xy + 2 * h) : 0.f;
      py_l[it] = ok ? __ldg(b.height_points_xy + 2 * h + 1) : 0.f;
    }
    for (int e = 0; e < tile_envs; ++e) {
      const int env_e = env0 + e;
      const float rx = __shfl_sync(LGK_FULL_MASK, root_x, e);
      const float ry = __shfl_sync(LGK_FULL_MASK, root_y, e);
      const float rz = __shfl_sync(LGK_FULL_MASK, z_pre, e);
      const float zo = __shfl_sync(LGK_FULL_MASK, zobs, e);
      const float qz = __shfl_sync(LGK_FULL_MASK, yaw_z, e);
      const float qw = __shfl_sync(LGK_FULL_MASK, yaw_w, e);
      float* orow = b.obs_buf + (size_t)env_e * O;
      const float* trow = taps ? b.tap_noise + (size_t)env_e * O : nullptr;
      uint4 rnd = make_uint4(0, 0, 0, 0);
      const unsigned eg = (unsigned)(c.env_id_offset + env_e);
      auto noisy = [&](float v, int col, unsigned rbits) {
        if (add_noise) {
          const float u = taps ? __ldg(trow + col) : u01(rbits);
          // obs += (2*rand - 1) * noise_scale_vec   (:231)
          v = add_rn(v, mul_rn(sub_rn(mul_rn(2.0f, u), 1.0f), __ldg(b.noise_scale_vec + col)));
        }
        return clampf(v, -c.clip_obs, c.clip_obs);   // step(): clip(obs, +-clip_observations) (:104)
      };
      if (add_noise && !taps) rnd = lgk_philox(c.seed, csc, eg, LGK_SITE_NOISE + lane);
      // head: columns 0..OH-1 (OH <= 64)
      {
        const float* hrow = wsm + e * a.head_stride;
        if (lane < OH) orow[lane] = noisy(hrow[lane], lane, rnd.x);
        if (lane + 32 < OH) orow[lane + 32] = noisy(hrow[lane + 32], lane + 32, rnd.y);
      }
      // scan points: column OH + h
      float part = 0.f;
      if (H > 0) {
        const float qz2 = mul_rn(qz, 2.0f);
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          if (it * 32 >= H) break;
          if (it == 2 && add_noise && !taps) rnd = lgk_philox(c.seed, csc, eg, LGK_SITE_NOISE + 32 + lane);
          if (it == 6 && add_noise && !taps) rnd = lgk_philox(c.seed, csc, eg, LGK_SITE_NOISE + 64 + lane);
          const int h = it * 32 + lane;
          const unsigned rbits = (it & 3) == 0 ? rnd.z : (it & 3) == 1 ? rnd.w : (it & 3) == 2 ? rnd.x : rnd.y;
          if (h < H) {
            float hgt = 0.f;
            if (!plane) {
              // quat_apply(normalize([0,0,qz,qw]), (px,py,0)) + root_pos, op-for-op (no contraction):
              //   t = 2 * cross(xyz, p); r = p + w*t + cross(xyz, t)
              const float bx = px_l[it], by = py_l[it];
              const float tx = -mul_rn(qz2, by), ty = mul_rn(qz2, bx);
              float wx = add_rn(add_rn(bx, mul_rn(qw, tx)), -mul_rn(qz, ty));
              float wy = add_rn(add_rn(by, mul_rn(qw, ty)), mul_rn(qz, tx));
              wx = add_rn(add_rn(wx, rx), c.border_size);
              wy = add_rn(add_rn(wy, ry), c.border_size);
              // (points / horizontal_scale).long(): truncation toward zero (:842)
              const float sx = mul_rcp ? mul_rn(wx, inv_hs) : __fdiv_rn(wx, c.horizontal_scale);
              const float sy = mul_rcp ? mul_rn(wy, inv_hs) : __fdiv_rn(wy, c.horizontal_scale);
              int ix = __float2int_rz(sx), iy = __float2int_rz(sy);
              ix = min(max(ix, 0), max_px);
              iy = min(max(iy, 0), max_py);
              const int raw = __ldg(b.height_min3 + (size_t)ix * c.terrain_pitch + iy);
              hgt = mul_rn((float)raw, c.vertical_scale);
            }
            if (b.measured_heights != nullptr) b.measured_heights[(size_t)env_e * H + h] = hgt;
            part += sub_rn(rz, hgt);
            const float v = clampf(sub_rn(zo, hgt), -1.f, 1.f) * c.obs_scale_height;
            orow[OH + h] = noisy(v, OH + h, rbits);
          }
        }
      }
      part = warp_sum(part);
      if (lane == e) bh_sum = part;
    }
  }

  // ---- S1: reward terms in the reference's order (:195-212), episode sums, statistics -----------
  const bool any_reset = reset_mask != 0u;
  const int num_rows = c.num_rewards + ((flags & LGK_F_HAS_TERMINATION) ? 1 : 0);
  float rew = 0.f;
  int ext_k = 0;
  for (int i = 0; i < num_rows; ++i) {
    float r = 0.f;
    if (i < c.num_rewards) {
      float v = 0.f;
      const float cn = sqrtf(fmaf(cmd_rew.y, cmd_rew.y, cmd_rew.x * cmd_rew.x));
      switch (c.rewards[i].term) {
        case LGK_REW_ACTION_RATE: v = acc.arate; break;
        case LGK_REW_ANG_VEL_XY: v = fmaf(bav[1], bav[1], bav[0] * bav[0]); break;
        case LGK_REW_BASE_HEIGHT: {
          const float bh = H > 0 ? bh_sum / (float)H : z_pre;
          const float d = bh - c.base_height_target;
          v = d * d;
        } break;
        case LGK_REW_COLLISION: v = cs.collisions; break;
        case LGK_REW_DOF_ACC: v = acc.acc2; break;
        case LGK_REW_DOF_POS_LIMITS: v = acc.poslim; break;
        case LGK_REW_DOF_VEL: v = acc.qd2; break;
        case LGK_REW_DOF_VEL_LIMITS: v = acc.vellim; break;
        case LGK_REW_ENERGY: v = acc.energy; break;
        case LGK_REW_FEET_AIR_TIME: v = r_air; break;
        case LGK_REW_FEET_CONTACT_FORCES: v = cs.feet_force_pen; break;
        case LGK_REW_HIP_MOTION: v = acc.hip; break;
        case LGK_REW_LIN_VEL_Z: v = blv[2] * blv[2]; break;
        case LGK_REW_NO_FLY: v = (cs.nofly_contacts == 1) ? 1.f : 0.f; break;
        case LGK_REW_ORIENTATION: v = fmaf(grav[1], grav[1], grav[0] * grav[0]); break;
        case LGK_REW_STAND_STILL: v = acc.stand * ((cn < 0.1f) ? 1.f : 0.f); break;
        case LGK_REW_STUMBLE: v = cs.stumble ? 1.f : 0.f; break;
        case LGK_REW_TORQUE_LIMITS: v = acc.tqlim; break;
        case LGK_REW_TORQUES: v = acc.tq2; break;
        case LGK_REW_TRACKING_ANG_VEL: {
          const float e = cmd_rew.z - bav[2];
          v = expf(-(e * e) / c.tracking_sigma);
        } break;
        case LGK_REW_TRACKING_LIN_VEL: {
          const float ex = cmd_rew.x - blv[0], ey = cmd_rew.y - blv[1];
          v = expf(-fmaf(ey, ey, ex * ex) / c.tracking_sigma);
        } break;
        case LGK_REW_EXTERNAL:
          v = (active && b.ext_rewards != nullptr) ? b.ext_rewards[(size_t)ext_k * N + env] : 0.f;
          ++ext_k;
          break;
        default: break;
      }
      r = mul_rn(v, c.rewards[i].scale);
      rew = add_rn(rew, r);
    } else {
      // termination term is added after the positive clip (:206-212)
      if (flags & LGK_F_ONLY_POSITIVE) rew = fmaxf(rew, 0.f);
      r = mul_rn((reset && !time_out) ? 1.f : 0.f, c.termination_scale);
      rew = add_rn(rew, r);
    }
    float total = 0.f;
    if (active) {
      float* sp = b.episode_sums + (size_t)i * N + env;
      total = add_rn(*sp, r);
      *sp = reset ? 0.f : total;    // zeroed for reset envs after the statistics read it (:184-185)
    }
    if (any_reset) {
      const float s = warp_sum((active && reset) ? total : 0.f);
      if (lane == 0) atomicAdd(&ws->stats[parity][i], s);
    }
  }
  if (!(flags & LGK_F_HAS_TERMINATION) && (flags & LGK_F_ONLY_POSITIVE)) rew = fmaxf(rew, 0.f);
  if (active) b.rew_buf[env] = rew;

  // ---- last CTA: ordered compaction of the id lists, statistics, counters ---------------------
  __shared__ unsigned s_is_last;
  __shared__ int s_scan[2][LGK_CTA_THREADS + 1];
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_is_last = (atomicAdd(&ws->done_ticket, 1u) == gridDim.x - 1) ? 1u : 0u;
  __syncthreads();
  if (!s_is_last) return;
  __threadfence();
  {
    const int nw = a.num_warps;
    const int per = (nw + LGK_CTA_THREADS - 1) / LGK_CTA_THREADS;
    const int w0 = min(nw, (int)threadIdx.x * per), w1 = min(nw, w0 + per);
    const volatile unsigned* rm = mask_base;
    const volatile unsigned* sm = mask_base + max_warps;
    int c0 = 0, c1 = 0;
    for (int w = w0; w < w1; ++w) { c0 += __popc(rm[w]); c1 += __popc(sm[w]); }
    s_scan[0][threadIdx.x + 1] = c0;
    s_scan[1][threadIdx.x + 1] = c1;
    if (threadIdx.x == 0) { s_scan[0][0] = 0; s_scan[1][0] = 0; }
    __syncthreads();
    if (threadIdx.x < 2) {   // serial scan of 128 partials per list: negligible next to the loads
      int* s = s_scan[threadIdx.x];
      for (int i = 1; i <= LGK_CTA_THREADS; ++i) s[i] += s[i - 1];
    }
    __syncthreads();
    int o0 = s_scan[0][threadIdx.x], o1 = s_scan[1][threadIdx.x];
    for (int w = w0; w < w1; ++w) {
      unsigned m = rm[w];
      while (m) { const int bit = __ffs(m) - 1; m &= m - 1; b.reset_ids[o0++] = w * epw + bit; }
      m = sm[w];
      while (m) { const int bit = __ffs(m) - 1; m &= m - 1; b.resample_ids[o1++] = w * epw + bit; }
    }
    const int n_reset = s_scan[0][LGK_CTA_THREADS], n_resample = s_scan[1][LGK_CTA_THREADS];
    const int level_sum = ((volatile int*)ws->level_sum)[parity];
    volatile float* st = ws->stats[parity];
    if ((int)threadIdx.x < num_rows) {
      const float s = st[threadIdx.x];
      b.episode_stats[threadIdx.x] = s;
      b.episode_stats[num_rows + 4 + threadIdx.x] += s;   // running sums since the host last cleared them
      if (n_reset > 0)   // extras["episode"]["rew_<k>"] = mean(sum[ids]) / max_episode_length_s (:184)
        b.extras[threadIdx.x] = (s / (float)n_reset) / c.max_episode_length_s;
      ws->stats[parity ^ 1][threadIdx.x] = 0.f;   // next step's accumulators
    }
    if (threadIdx.x == 0) {
      b.counts[0] = n_reset;
      b.counts[1] = n_resample;
      b.episode_stats[num_rows] = (float)n_reset;
      b.episode_stats[num_rows + 1] = (float)level_sum;
      b.episode_stats[num_rows + 2] = (float)N;
      b.episode_stats[2 * num_rows + 4] += (float)n_reset;
      b.episode_stats[2 * num_rows + 5] = (float)level_sum;
      b.episode_stats[2 * num_rows + 6] = (float)N;
      if (n_reset > 0 && (flags & LGK_F_TERRAIN_CURRICULUM))
        b.extras[num_rows] = (float)level_sum / (float)N;   // mean(terrain_levels.float()) (:188)
      ws->level_sum[parity ^ 1] = 0;
      ws->step_counter = csc;
      ws->done_ticket = 0u;
    }
  }
}

// ---------------------------------------------------------------------------------------------
static int pick_envs_per_warp(int num_envs, int requested) {
  if (requested == 4 || requested == 8 || requested == 16 || requested == 32) return requested;
  // enough warps to populate 148 SMs a few times over, else trade scalar-phase efficiency for spread
  if (num_envs >= 32 * 148 * 8) return 32;
  if (num_envs >= 16 * 148 * 8) return 16;
  if (num_envs >= 8 * 148 * 4) return 8;
  return 4;
}

extern "C" int lgk_post_physics_step_impl(const LgkConsts* c, const LgkBuffers* b, cudaStream_t stream) {
  StepArgs args;
  args.c = *c;
  args.b = *b;
  args.epw = pick_envs_per_warp(c->num_envs, c->envs_per_warp);
  const int OH = 12 + 3 * c->num_dof;
  args.head_stride = OH | 1;
  int row = c->num_bodies * 3;
  if (row < args.head_stride) row = args.head_stride;
  row = (row + 3) & ~3;   // keep every warp's region 16-byte aligned
  args.row_floats = row;
  args.num_warps = (c->num_envs + args.epw - 1) / args.epw;
  const int grid = (args.num_warps + LGK_WARPS_PER_CTA - 1) / LGK_WARPS_PER_CTA;
  const size_t smem = (size_t)LGK_WARPS_PER_CTA * args.epw * row * sizeof(float);
  if (c->num_dof == 12) {
    if (smem > 48 * 1024)
      LGK_CUDA_RETURN_IF(cudaFuncSetAttribute(lgk_step_kernel<12>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    lgk_step_kernel<12><<<grid, LGK_CTA_THREADS, smem, stream>>>(args);
  } else {
    if (smem > 48 * 1024)
      LGK_CUDA_RETURN_IF(cudaFuncSetAttribute(lgk_step_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    lgk_step_kernel<0><<<grid, LGK_CTA_THREADS, smem, stream>>>(args);
  }
  return (int)cudaGetLastError();
}

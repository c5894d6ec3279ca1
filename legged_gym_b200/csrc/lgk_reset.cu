// reset_idx(env_ids) for an explicit id list or for every env (legged_robot.py:150-193).
// Not on the per-step hot path (the fused step kernel resets terminated envs itself); used by
// BaseTask.reset() and by callers that reset a subset by hand.
#include "lgk_common.cuh"

struct ResetArgs {
  LgkConsts c;
  LgkBuffers b;
  const long long* ids;   // ascending int64 env ids, or nullptr for arange(N)
  int n;
};

__global__ void __launch_bounds__(128)
lgk_reset_kernel(const __grid_constant__ ResetArgs a) {
  const LgkConsts& c = a.c;
  const LgkBuffers& b = a.b;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const bool active = i < a.n;
  const int env = active ? (a.ids ? (int)a.ids[i] : i) : 0;
  const int D = c.num_dof, N = c.num_envs, F = c.num_feet;
  const unsigned flags = c.flags;
  const bool taps = c.rng_mode == LGK_RNG_TAPS;
  LgkWorkspace* ws = reinterpret_cast<LgkWorkspace*>(b.workspace);
  const unsigned long long csc = ws->step_counter;
  const int parity = (int)((csc + 1ull) & 1ull);   // accumulators the next step kernel will NOT use... see finalize
  const unsigned eg = (unsigned)(c.env_id_offset + env);
  const unsigned SITE = LGK_SITE_RESET_IDX;   // philox blocks disjoint from the step kernel's

  int level_contrib = 0;
  if (active) {
    float4 cmd = *reinterpret_cast<const float4*>(b.commands + (size_t)env * 4);
    float* rp = b.root_states + (size_t)env * 13;
    float origin[3];
    { const float* op = b.env_origins + (size_t)env * 3; origin[0] = op[0]; origin[1] = op[1]; origin[2] = op[2]; }
    long long level = b.terrain_levels ? b.terrain_levels[env] : 0;
    if ((flags & LGK_F_TERRAIN_CURRICULUM) && (flags & LGK_F_INIT_DONE) && b.terrain_levels) {   // (:443-463)
      const float dx = rp[0] - origin[0], dy = rp[1] - origin[1];
      const float dist = sqrtf(fmaf(dy, dy, dx * dx));
      const bool up = dist > c.half_env_length;
      const float cn = sqrtf(fmaf(cmd.y, cmd.y, cmd.x * cmd.x));
      const bool down = (dist < mul_rn(mul_rn(cn, c.max_episode_length_s), 0.5f)) && !up;
      level += (up ? 1 : 0) - (down ? 1 : 0);
      if (level >= c.max_terrain_level) {
        level = taps ? b.tap_level[env]
                     : (long long)(lgk_philox(c.seed, csc, eg, SITE + LGK_SITE_LEVEL).x % (unsigned)c.max_terrain_level);
      } else if (level < 0) {
        level = 0;
      }
      b.terrain_levels[env] = level;
      const float* to = b.terrain_origins + ((size_t)level * c.num_terrain_types + (size_t)b.terrain_types[env]) * 3;
      origin[0] = to[0]; origin[1] = to[1]; origin[2] = to[2];
      float* ow = b.env_origins + (size_t)env * 3;
      ow[0] = origin[0]; ow[1] = origin[1]; ow[2] = origin[2];
    }
    level_contrib = (int)level;
    // _reset_dofs (:399-408)
    float* dp = b.dof_state + (size_t)env * D * 2;
    for (int d = 0; d < D; ++d) {
      float u;
      if (taps) u = b.tap_dof[(size_t)env * D + d];
      else {
        const uint4 r = lgk_philox(c.seed, csc, eg, SITE + LGK_SITE_DOF + (d >> 2));
        const unsigned rr[4] = {r.x, r.y, r.z, r.w};
        u = u01(rr[d & 3]);
      }
      dp[2 * d] = mul_rn(c.default_dof_pos[d], rand_affine(1.0f, 0.5f, u));
      dp[2 * d + 1] = 0.f;
      b.last_actions[(size_t)env * D + d] = 0.f;
      b.last_dof_vel[(size_t)env * D + d] = 0.f;
    }
    // _reset_root_states (:414-430)
    float u[8];
    if (taps) {
      u[0] = b.tap_root_xy[(size_t)env * 2]; u[1] = b.tap_root_xy[(size_t)env * 2 + 1];
      for (int k = 0; k < 6; ++k) u[2 + k] = b.tap_root_vel[(size_t)env * 6 + k];
    } else {
      const uint4 r0 = lgk_philox(c.seed, csc, eg, SITE + LGK_SITE_ROOT);
      const uint4 r1 = lgk_philox(c.seed, csc, eg, SITE + LGK_SITE_ROOT + 1);
      u[0] = u01(r0.x); u[1] = u01(r0.y); u[2] = u01(r0.z); u[3] = u01(r0.w);
      u[4] = u01(r1.x); u[5] = u01(r1.y); u[6] = u01(r1.z); u[7] = u01(r1.w);
    }
    float root[13];
    for (int k = 0; k < 13; ++k) root[k] = c.base_init_state[k];
    root[0] = add_rn(root[0], origin[0]);
    root[1] = add_rn(root[1], origin[1]);
    root[2] = add_rn(root[2], origin[2]);
    if (flags & LGK_F_CUSTOM_ORIGINS) {
      root[0] = add_rn(root[0], rand_affine(2.0f, -1.0f, u[0]));
      root[1] = add_rn(root[1], rand_affine(2.0f, -1.0f, u[1]));
    }
    for (int k = 0; k < 6; ++k) root[7 + k] = rand_affine(1.0f, -0.5f, u[2 + k]);
    for (int k = 0; k < 13; ++k) rp[k] = root[k];
    // _resample_commands (:354-368)
    float c0, c1, c2;
    if (taps) { const float* t = b.tap_cmd_reset + (size_t)env * 3; c0 = t[0]; c1 = t[1]; c2 = t[2]; }
    else { const uint4 r = lgk_philox(c.seed, csc, eg, SITE + LGK_SITE_CMD_RESET); c0 = u01(r.x); c1 = u01(r.y); c2 = u01(r.z); }
    const float* cr = b.command_ranges + 8;   // reset_idx's ranges
    cmd.x = rand_affine(cr[1], cr[0], c0);
    cmd.y = rand_affine(cr[3], cr[2], c1);
    if (flags & LGK_F_HEADING_COMMAND) cmd.w = rand_affine(cr[7], cr[6], c2);
    else cmd.z = rand_affine(cr[5], cr[4], c2);
    const float keep = (sqrtf(fmaf(cmd.y, cmd.y, cmd.x * cmd.x)) > 0.2f) ? 1.f : 0.f;
    cmd.x *= keep;
    cmd.y *= keep;
    *reinterpret_cast<float4*>(b.commands + (size_t)env * 4) = cmd;
    for (int k = 0; k < F; ++k) b.feet_air_time[(size_t)env * F + k] = 0.f;
    b.episode_length_buf[env] = 0;
    b.reset_buf[env] = 1;
    b.reset_ids[i] = env;
  }
  // episode statistics over the reset ids (:182-185)
  const int num_rows = c.num_rewards + ((flags & LGK_F_HAS_TERMINATION) ? 1 : 0);
  for (int k = 0; k < num_rows; ++k) {
    float v = 0.f;
    if (active) {
      float* sp = b.episode_sums + (size_t)k * N + env;
      v = *sp;
      *sp = 0.f;
    }
    const float s = warp_sum(v);
    if ((threadIdx.x & 31) == 0 && s != 0.f) atomicAdd(&ws->stats[parity][k], s);
  }

  __shared__ unsigned s_is_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_is_last = (atomicAdd(&ws->done_ticket, 1u) == gridDim.x - 1) ? 1u : 0u;
  __syncthreads();
  if (!s_is_last) return;
  __threadfence();
  volatile float* st = ws->stats[parity];
  if ((int)threadIdx.x < num_rows) {
    const float s = st[threadIdx.x];
    b.episode_stats[threadIdx.x] = s;
    if (a.n > 0) b.extras[threadIdx.x] = (s / (float)a.n) / c.max_episode_length_s;
    st[threadIdx.x] = 0.f;
  }
  if (threadIdx.x == 0) {
    b.counts[0] = a.n;
    b.episode_stats[num_rows] = (float)a.n;
    ws->done_ticket = 0u;
  }
  (void)level_contrib;
}

// mean(terrain_levels.float()) over ALL envs (:188) after a reset_idx call
__global__ void lgk_level_mean_kernel(const int64_t* __restrict__ levels, int N, float* __restrict__ out) {
  __shared__ int s_part[32];
  int v = 0;
  for (int i = threadIdx.x; i < N; i += blockDim.x) v += (int)levels[i];
  v = warp_sum_int(v);
  if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += s_part[w];
    out[0] = (float)t / (float)N;
    out[1] = (float)t;
  }
}

extern "C" int lgk_reset_idx_impl(const LgkConsts* c, const LgkBuffers* b, const long long* ids, int n, cudaStream_t st) {
  if (n <= 0) return 0;   // early-out on an empty id list (:160)
  ResetArgs args;
  args.c = *c;
  args.b = *b;
  args.ids = ids;
  args.n = n;
  lgk_reset_kernel<<<(n + 127) / 128, 128, 0, st>>>(args);
  LGK_CUDA_RETURN_IF(cudaGetLastError());
  if ((c->flags & LGK_F_TERRAIN_CURRICULUM) && b->terrain_levels) {
    const int num_rows = c->num_rewards + ((c->flags & LGK_F_HAS_TERMINATION) ? 1 : 0);
    // extras[num_rows] = terrain level mean; episode_stats[num_rows+1] = level sum
    lgk_level_mean_kernel<<<1, 1024, 0, st>>>(b->terrain_levels, c->num_envs, b->extras + num_rows);
    LGK_CUDA_RETURN_IF(cudaGetLastError());
  }
  return 0;
}

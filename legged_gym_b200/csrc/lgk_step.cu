// Step kernel of post_physics_step (sm_100a): one CTA of four warps per tile of 32 envs, lanes = envs.
//
// post_physics_step() (reference legged_robot.py:109-141) is two launches chained with programmatic dependent launch:
//   lgk_step_kernel    (this file)    everything per env that needs no height: episode counters, body-frame
//                                     velocities / projected gravity, command resampling + heading command, pushes,
//                                     terminations, the VALUE of every _reward_* term, resets (terrain curriculum, dof
//                                     state, commands), the noisy scaled head of the observation row, last_* copies,
//                                     the per-tile reset / resample bitmasks behind the ascending id lists
//   lgk_tail_kernel    (lgk_scan.cu)  _get_heights + the scan columns of the observation row (lanes = points), then
//                                     -- by the CTAs that did the scan -- compute_reward's scaled in-place sums over
//                                     this kernel's term values (base_height from the scan's height sums), the held-back
//                                     root rows of the envs reset this step, the id lists (one CTA), extras and the
//                                     step counter (the same CTA, once everybody else has checked in)
// This kernel neither reads heights nor writes the root state: the new root row of a reset env goes to a side buffer
// in the workspace together with the scan operands of the old pose and the new root z (:227), so the tail kernel
// computes that env's scan columns once and stores the row when the pose is no longer needed.  From 12 288 envs on
// every CTA here raises a hand-over flag when its envs are through and the tail kernel's warps start on those envs
// under this kernel's last wave; below that the tail kernel waits for the whole grid (lgk_chain_mode).
// LGK_CHAIN=3 in the environment runs the three-launch chain of earlier rounds instead (lgk_scan_kernel, this kernel,
// lgk_finish_kernel below), kept for A/B measurements.
//
// Work decomposition of this kernel
//   grid    CTA c owns the contiguous env range [c*per, (c+1)*per) and walks it in sub-tiles of <= 32 envs.
//           per = 32 by default (the hardware scheduler backfills); LgkConsts.envs_per_cta selects longer ranges.
//   load    every per-env input row of a sub-tile is contiguous in HBM ([N,13], [N,D,2], [N,B,3], ...): the lanes
//           of W0 issue 1-D bulk async copies (cp.async.bulk -> SASS UBLKCP, the TMA engine) into shared memory,
//           completion counted on one mbarrier.
//   stage 1 four roles, lane = env (no divergence inside a warp):
//             W0  episode counter, commands (resample, heading command), push, root reset with the terrain
//                 curriculum, the per-tile bitmasks;  W1, W2  half of the joints each;  W3  body-frame
//                 velocities / gravity, contacts, feet air time.
//           Each role writes its columns of the observation head -- scaled, with noise, clipped -- into a staging
//           row per env; the noise fields come from Philox blocks dealt to the roles (lane = env: every lane of a
//           block computation is used).  The reset flag is cheap and is recomputed by each warp.
//   -- the only CTA barrier of a sub-tile --
//   store   W0..W2 write the staged heads to the observation rows, 128 B per warp store; last_actions and
//           last_dof_vel leave through bulk async stores;  W3 writes the term values of its 32 envs into the
//           [term][N] table of the workspace (one coalesced 128 B store per term) for the tail kernel.
//
// Code size matters (four roles live in different regions of the code at once): paths that run for ~1 % of the
// envs, once per launch, or only for ragged / unaligned tiles are __noinline__.
#include <cmath>
#include <cstdio>
#include <limits>
#include <type_traits>
#include "lgk_step.cuh"

#ifndef LGK_MIN_CTAS
#define LGK_MIN_CTAS 7   // register cap 72: 2 048 tiles are two waves of 7 x 148 slots.  (Six CTAs at 80 registers were as fast
                         // while a scan kernel ran ahead of this one; as the first launch of the two-launch chain, with the
                         // tail kernel moving in behind its last wave, seven are 1 us per step faster: 65.4 against 66.5 us)
#endif

#ifdef LGK_PROFILE
// development build only: per-CTA phase timestamps (clock64) of the first sub-tile, [cta][warp][16]
__device__ long long lgk_prof[4096 * 4 * 16];
#define LGK_STAMP(k) do { if (sub == 0 && lane == 0 && blockIdx.x < 4096) lgk_prof[(blockIdx.x * 4 + wic) * 16 + (k)] = clock64(); } while (0)
extern "C" int lgk_debug_read_prof(long long* out, int n) {
  return (int)cudaMemcpyFromSymbol(out, lgk_prof, sizeof(long long) * n);
}
// CTA timeline: [cta][globaltimer at start, globaltimer when W0 leaves, SM id, globaltimer when W3 leaves]
__device__ long long lgk_prof_cta[4096 * 4];
static __device__ __forceinline__ long long lgk_gtime() { long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
static __device__ __forceinline__ int lgk_smid() { int r; asm volatile("mov.u32 %0, %%smid;" : "=r"(r)); return r; }
#define LGK_CTA_STAMP(k) do { if (blockIdx.x < 4096) { lgk_prof_cta[blockIdx.x * 4 + (k)] = lgk_gtime(); lgk_prof_cta[blockIdx.x * 4 + 2] = lgk_smid(); } } while (0)
extern "C" int lgk_debug_read_prof_cta(long long* out, int n) {
  return (int)cudaMemcpyFromSymbol(out, lgk_prof_cta, sizeof(long long) * n);
}
#else
#define LGK_STAMP(k) do {} while (0)
#define LGK_CTA_STAMP(k) do {} while (0)
#endif

// scratch fields (SoA, [field][TILE]) exchanged between the roles
enum { SC_J1 = 0, SC_J2 = 10,    // joint partial sums of W1 / W2 (10 each)
       SC_CMDX = 20, SC_CMDY, SC_CMDZ,   // command after the callback, before any reset (what the rewards see)
       SC_ZPRE,                  // root z before any reset (_reward_base_height)
       SC_FIELDS };
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float norm3(float x, float y, float z) {
  return sqrtf(fmaf(z, z, fmaf(y, y, x * x)));
}

// any termination body with |f| > 1 (:146)
__device__ __forceinline__ float sumsq3(float x, float y, float z) { return fmaf(z, z, fmaf(y, y, x * x)); }
// offsets (in floats) of the contact rows the kernel looks at, copied from the parameter block into shared
// memory once per CTA: [0,8) feet, [8,24) penalised bodies, [24,32) termination bodies
#define IDX_FEET 0
#define IDX_PEN LGK_MAX_FEET
#define IDX_TERM (LGK_MAX_FEET + LGK_MAX_PEN)
static_assert(LGK_MAX_FEET + LGK_MAX_PEN + LGK_MAX_TERM == 32, "one lane per contact index");
__device__ __forceinline__ bool term_contact(const LgkConsts& c, const int* __restrict__ s_idx,
                                             const float* __restrict__ cf, float thr_term) {
  bool t = false;
#pragma unroll 1
  for (int k = 0; k < c.num_term; ++k) {
    const float* f = cf + s_idx[IDX_TERM + k];
    t |= sumsq3(f[0], f[1], f[2]) > thr_term;   // norm > 1
  }
  return t;
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
  float v[10];    // tq2, energy, qd2, acc2, arate, poslim, vellim, tqlim, stand, hip
};
enum { A_TQ2 = 0, A_ENERGY, A_QD2, A_ACC2, A_ARATE, A_POSLIM, A_VELLIM, A_TQLIM, A_STAND, A_HIP };

// one joint: reward partial sums on the pre-reset state, reset values; the three observation-head values of the
// joint (:221-223) come back unscaled-by-noise in hv[]
__device__ __forceinline__ void joint_update(const LgkConsts& c, int d, unsigned hip_mask, bool reset, float inv_dt,
                                             float av, float lav, float tqv, float ldvv, float& p, float& v, float u,
                                             DofAcc& acc, float (&hv)[3]) {
  acc.v[A_TQ2] = fmaf(tqv, tqv, acc.v[A_TQ2]);
  const float pw = tqv * v;
  acc.v[A_ENERGY] = fmaf(pw, pw, acc.v[A_ENERGY]);
  acc.v[A_QD2] = fmaf(v, v, acc.v[A_QD2]);
  const float ja = (ldvv - v) * inv_dt;
  acc.v[A_ACC2] = fmaf(ja, ja, acc.v[A_ACC2]);
  const float da = lav - av;
  acc.v[A_ARATE] = fmaf(da, da, acc.v[A_ARATE]);
  acc.v[A_POSLIM] += -fminf(p - c.dof_pos_lo[d], 0.f) + fmaxf(p - c.dof_pos_hi[d], 0.f);
  acc.v[A_VELLIM] += clampf(fabsf(v) - c.dof_vel_limits[d] * c.soft_dof_vel_limit, 0.f, 1.f);
  acc.v[A_TQLIM] += fmaxf(fabsf(tqv) - c.torque_limits[d] * c.soft_torque_limit, 0.f);
  const float dev = fabsf(p - c.default_dof_pos[d]);
  acc.v[A_STAND] += dev;
  if ((hip_mask >> d) & 1u) acc.v[A_HIP] += dev;
  if (reset) {   // _reset_dofs (:399-408): default * U(0.5, 1.5), zero velocity
    p = mul_rn(c.default_dof_pos[d], rand_affine(1.0f, 0.5f, u));
    v = 0.f;
  }
  hv[0] = (p - c.default_dof_pos[d]) * c.obs_scale_dof_pos;
  hv[1] = v * c.obs_scale_dof_vel;
  hv[2] = av;
}

// 16-bit field k (0..7) of a Philox block
__device__ __forceinline__ unsigned philox_word(const uint4& r, int w) {
  return w == 0 ? r.x : w == 1 ? r.y : w == 2 ? r.z : r.w;
}
// one column of the observation head: obs = clip(v + (2*rand - 1) * noise_scale_vec[col], +-clip_observations)
// (:231, :104).  Philox mode: field k of `blk` as the float 2^23 + r16 (one PRMT, no I2F on the quarter-rate
// conversion unit), (2^23 + r16) * 2^-15 - 257 = r16 / 2^15 - 1 exactly: uniform on [-1, 1) in steps of 2^-15.
// s_ns holds the head's noise scales (zeros without noise); trow: the env's tap row in tap mode with noise.
template <bool TAPS>
__device__ __forceinline__ float head_obs(float v, int col, const uint4& blk, int k, const float* __restrict__ s_ns,
                                          const float* __restrict__ trow, float clip_obs) {
  const float ns = s_ns[col];
  if (TAPS) {
    if (trow != nullptr) v = add_rn(v, mul_rn(sub_rn(mul_rn(2.0f, __ldg(trow + col)), 1.0f), ns));
  } else {
    const unsigned rw = philox_word(blk, k >> 1);
    const unsigned m = (k & 1) ? __byte_perm(rw, 0x4b000000u, 0x7632) : __byte_perm(rw, 0x4b000000u, 0x7610);
    v = fmaf(fmaf(__uint_as_float(m), 3.0517578125e-05f, -257.0f), ns, v);
  }
  return clampf(v, -clip_obs, clip_obs);
}

// ---- cold paths (kept out of line: they run for ~1 % of the envs, or once per launch) ---------------
static __device__ __noinline__ uint4 philox_cold(const StepArgs& a, unsigned long long step, unsigned env_global,
                                                 unsigned block) {
  return philox_keyed(a.pk, step, env_global, block);
}

static __device__ __noinline__ float wrap_to_pi_slow(float x) { return wrap_to_pi(x); }

// wrap_to_pi (utils/math.py:45-48); the common case |x| < 4*pi needs no fmodf
__device__ __forceinline__ float wrap_to_pi_fast(float x) {
  const float TWO_PI = 6.283185307179586f, PI_F = 3.141592653589793f;
  const float ax = fabsf(x);
  if (ax < 2.0f * TWO_PI) {
    float m = ax >= TWO_PI ? ax - TWO_PI : ax;   // exact (Sterbenz): fmodf(|x|, 2pi)
    m = copysignf(m, x);
    if (m < 0.0f) m += TWO_PI;                   // python-style modulo: result in [0, 2pi)
    if (m > PI_F) m -= TWO_PI;
    return m;
  }
  return wrap_to_pi_slow(x);
}

// reset of one env's root state, terrain level / origin and command (:443-463, :414-434, :173)
struct ResetOut {
  float4 cmd;      // in: command after the callback; out: command of the new episode
  float rvel[6];   // new root velocities (-> last_root_vel)
  float z;         // new root z (-> height observation offset, :227)
  int level;       // terrain level after the curriculum update
};
template <bool TAPS>
static __device__ __noinline__ void reset_root(const StepArgs& a, unsigned long long csc, int env, unsigned env_global,
                                               float rx, float ry, long long level, ResetOut& o, float* root_stage) {
  const LgkConsts& c = a.c;
  const LgkBuffers& b = a.b;
  const unsigned flags = c.flags;
  float origin[3];
  const float* op = b.env_origins + (size_t)env * 3;
  origin[0] = op[0]; origin[1] = op[1]; origin[2] = op[2];
  if ((flags & LGK_F_TERRAIN_CURRICULUM) && (flags & LGK_F_INIT_DONE) && b.terrain_levels != nullptr) {
    const float dx = rx - origin[0], dy = ry - origin[1];
    const float dist = sqrtf(fmaf(dy, dy, dx * dx));
    const bool up = dist > c.half_env_length;
    const float cn = sqrtf(fmaf(o.cmd.y, o.cmd.y, o.cmd.x * o.cmd.x));
    const bool down = (dist < mul_rn(mul_rn(cn, c.max_episode_length_s), 0.5f)) && !up;
    level += (up ? 1 : 0) - (down ? 1 : 0);
    if (level >= c.max_terrain_level) {
      if (TAPS) level = b.tap_level[env];
      else level = (long long)(philox_cold(a, csc, env_global, LGK_SITE_LEVEL).x % (unsigned)c.max_terrain_level);
    } else if (level < 0) {
      level = 0;
    }
    b.terrain_levels[env] = level;
    const float* to = b.terrain_origins + ((size_t)level * c.num_terrain_types + (size_t)b.terrain_types[env]) * 3;
    origin[0] = to[0]; origin[1] = to[1]; origin[2] = to[2];
    float* ow = b.env_origins + (size_t)env * 3;
    ow[0] = origin[0]; ow[1] = origin[1]; ow[2] = origin[2];
  }
  o.level = (int)level;
  float u[8];
  if (TAPS) {
    u[0] = b.tap_root_xy[(size_t)env * 2]; u[1] = b.tap_root_xy[(size_t)env * 2 + 1];
    for (int i = 0; i < 6; ++i) u[2 + i] = b.tap_root_vel[(size_t)env * 6 + i];
  } else {
    const uint4 r0 = philox_cold(a, csc, env_global, LGK_SITE_ROOT);
    const uint4 r1 = philox_cold(a, csc, env_global, LGK_SITE_ROOT + 1);
    u[0] = u01(r0.x); u[1] = u01(r0.y); u[2] = u01(r0.z); u[3] = u01(r0.w);
    u[4] = u01(r1.x); u[5] = u01(r1.y); u[6] = u01(r1.z); u[7] = u01(r1.w);
  }
  float root[13];
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
  for (int i = 0; i < 6; ++i) { o.rvel[i] = rand_affine(1.0f, -0.5f, u[2 + i]); root[7 + i] = o.rvel[i]; }
  // the new root state waits in the workspace's side buffer: the finish kernel stores it once the scan kernel, which
  // runs side by side with this one and needs the pose from before the reset (:818-854 precede reset_idx), is done
#pragma unroll
  for (int i = 0; i < 13; ++i) root_stage[i] = root[i];
  o.z = root[2];
  float c0, c1, c2;
  if (TAPS) { const float* t = b.tap_cmd_reset + (size_t)env * 3; c0 = t[0]; c1 = t[1]; c2 = t[2]; }
  else { const uint4 r = philox_cold(a, csc, env_global, LGK_SITE_CMD_RESET); c0 = u01(r.x); c1 = u01(r.y); c2 = u01(r.z); }
  resample_command(o.cmd, b.command_ranges + 8, flags & LGK_F_HEADING_COMMAND, c0, c1, c2);   // reset_idx's ranges
}

// generic (non-bulk) tile load / store for ragged or unaligned tiles
static __device__ __noinline__ void tile_load_fallback(const StepArgs& a, float* sm, int env0, int ne) {
  const LgkConsts& c = a.c;
  const LgkBuffers& b = a.b;
  const TileLayout& L = a.L;
  const int D = c.num_dof, B3 = c.num_bodies * 3, F = c.num_feet;
  const int tid = threadIdx.x;   // (physical thread index: every thread takes part, whatever its role)
  auto ld = [&](int off, const void* g, int words_per_env) {
    const unsigned* src = reinterpret_cast<const unsigned*>(g);
    unsigned* dst = reinterpret_cast<unsigned*>(sm + off);
#pragma unroll 1
    for (int i = tid; i < ne * words_per_env; i += LGK_CTA_THREADS) dst[i] = src[i];
  };
  ld(L.root, b.root_states + (size_t)env0 * 13, 13);
  ld(L.ep, b.episode_length_buf + env0, 2);
  ld(L.cf, b.contact_forces + (size_t)env0 * B3, B3);
  ld(L.cmd, b.commands + (size_t)env0 * 4, 4);
  ld(L.dof, b.dof_state + (size_t)env0 * 2 * D, 2 * D);
  ld(L.act, b.actions + (size_t)env0 * D, D);
  ld(L.lact, b.last_actions + (size_t)env0 * D, D);
  ld(L.tq, b.torques + (size_t)env0 * D, D);
  ld(L.ldv, b.last_dof_vel + (size_t)env0 * D, D);
  ld(L.fat, b.feet_air_time + (size_t)env0 * F, F);
  if (b.terrain_levels != nullptr) ld(L.lvl, b.terrain_levels + env0, 2);
}
static __device__ __noinline__ void tile_store_fallback(const StepArgs& a, const float* sm, int env0, int ne, int tid) {
  const int D = a.c.num_dof;
  const int t = tid - 32, nt = LGK_CTA_THREADS - 32;   // warps 1..3
  float* la = a.b.last_actions + (size_t)env0 * D;
  float* lv = a.b.last_dof_vel + (size_t)env0 * D;
#pragma unroll 1
  for (int i = t; i < ne * D; i += nt) { la[i] = sm[a.L.act + i]; lv[i] = sm[a.L.ldv + i]; }
}

template <int DT, bool TAPS, bool MUL_RCP>
__global__ void __launch_bounds__(LGK_CTA_THREADS, LGK_MIN_CTAS)
lgk_step_kernel(const __grid_constant__ StepArgs a) {
  // Fused chain: this is the first launch.  It waits for whatever ran before, and only then lets the scan kernel be
  // scheduled -- that kernel's CTAs take SM slots as the last wave here retires, and with hand-over flags they start
  // on the envs whose step CTAs are through; releasing them before the wait would let them run next to the PREVIOUS
  // step's last kernel when this grid fits the GPU at once.
  if (a.fused) pdl_wait();
  pdl_launch_dependents();   // the next kernel of the chain may be scheduled as soon as the last tile CTA has started
  // Behind the scan kernel this launch needs no wait on entry: that kernel waits for ITS predecessor before it lets
  // its dependents go, so every input here (simulator tensors, clipped actions, the previous step's state) is
  // complete when the first CTA starts, and nothing here depends on the scan itself.  Without a scan kernel: wait now.
  if (!a.scan_first && !a.fused) pdl_wait();
  const LgkConsts& c = a.c;
  const LgkBuffers& b = a.b;
  const TileLayout& L = a.L;
  constexpr bool VEC = (DT > 0) && (DT % 4 == 0);
  const int D = DT > 0 ? DT : c.num_dof;
  const int N = c.num_envs;
  const int B3 = c.num_bodies * 3;
  const int F = c.num_feet;
  const int O = c.num_obs;
  const int OH = 12 + 3 * D;               // width of the observation head (:217-224)
  const int HS = a.head_stride;
  const unsigned flags = c.flags;
  const bool heading_mode = flags & LGK_F_HEADING_COMMAND;
  const bool add_noise = flags & LGK_F_ADD_NOISE;

  extern __shared__ __align__(16) float sm[];
  __shared__ __align__(8) uint64_t s_bar;
  __shared__ int s_idx[32];
  __shared__ float s_ns[64];   // noise scales of the head columns (zeros without noise)

  // roles go by the logical warp index wic: rotated per CTA so that the same role of the CTAs resident on an SM
  // does not always sit on the same scheduler
  const int lane = threadIdx.x & 31;
  const int wic = ((int)(threadIdx.x >> 5) + (int)((blockIdx.x * 2654435761u) >> 30)) & (LGK_WARPS_PER_CTA - 1);
  const int tid = wic * 32 + lane;
  const bool big = (int)blockIdx.x < a.n_big;
  const int cta_env0 = big ? blockIdx.x * a.per : a.n_big * a.per + ((int)blockIdx.x - a.n_big) * a.per_small;
  const int cta_ne = min(big ? a.per : a.per_small, N - cta_env0);
  const int nsub = (cta_ne + TILE - 1) / TILE;
  const bool has_levels = b.terrain_levels != nullptr;

  // ---- first thing: W0 sets the mbarrier up and gets the bulk loads of the first sub-tile going; the rest of the
  // prologue runs under their latency.  Lane i copies input array i, then the lanes share the episode-sum rows.
  const char* ld_src = nullptr;
  unsigned ld_off = 0, ld_bpe = 0;
  if (wic == 0 && lane < LGK_LD_ARRAYS) { ld_src = a.ld_src[lane]; ld_off = a.ld_off[lane]; ld_bpe = a.ld_bpe[lane]; }
  auto issue_loads = [&](int env0, int ne) {   // W0, converged
    if (lane == 0) mbar_expect_tx(&s_bar, (unsigned)ne * a.ld_bytes_per_env);
    __syncwarp();
    if (ld_src != nullptr) bulk_g2s(sm + ld_off, ld_src + (size_t)env0 * ld_bpe, (unsigned)ne * ld_bpe, &s_bar);
  };
  if (wic == 0) {
    if (lane == 0) mbar_init(&s_bar, 1);
    __syncwarp();
    const int ne0 = min(TILE, cta_ne);
    if (a.tma_ok && (ne0 & 3) == 0) issue_loads(cta_env0, ne0);
  } else if (wic == 1) {
    s_idx[lane] = 3 * (lane < IDX_PEN ? c.feet_idx[lane] : lane < IDX_TERM ? c.pen_idx[lane - IDX_PEN] : c.term_idx[lane - IDX_TERM]);
  } else if (wic == 2) {
    s_ns[lane] = (add_noise && lane < OH) ? __ldg(b.noise_scale_vec + lane) : 0.f;
    s_ns[lane + 32] = (add_noise && lane + 32 < OH) ? __ldg(b.noise_scale_vec + lane + 32) : 0.f;
  }
  if (tid == 0) LGK_CTA_STAMP(0);
  LgkWorkspace* ws = reinterpret_cast<LgkWorkspace*>(b.workspace);
  const unsigned long long csc = ws->step_counter + 1ull;   // common_step_counter after += 1 (:119)
  const int parity = (int)(csc & 1ull);
  bool push_step = false;   // W0 only; 32-bit remainder unless the counter has outgrown it
  if (wic == 0 && (flags & LGK_F_PUSH_ROBOTS))
    push_step = (csc >> 32) == 0ull ? ((unsigned)csc % (unsigned)c.push_interval == 0u)
                                    : (csc % (unsigned long long)c.push_interval == 0ull);
  const float clip_obs = c.clip_obs;

  __syncthreads();          // barrier initialised before anyone polls it, tables written

  for (int sub = 0; sub < nsub; ++sub) {
    const int env0 = cta_env0 + sub * TILE;
    const int ne = min(TILE, cta_ne - sub * TILE);
    const bool full = a.tma_ok && ((ne & 3) == 0);   // bulk copies need 16-byte multiples: 4 envs x 4 B
    const int slot = blockIdx.x * a.max_sub + sub;   // where this sub-tile's reset / resample masks go

    LGK_STAMP(0);
    // ---- load the sub-tile --------------------------------------------------------------------
    if (full) {
      if (sub > 0) {            // (the first sub-tile's loads were issued at the top of the kernel)
        __syncthreads();        // every warp is through the previous sub-tile
        if (wic == 0) {
          if (lane == 0) fence_proxy_async_smem();   // generic reads of the previous sub-tile before the async writes
          issue_loads(env0, ne);
        }
      }
      mbar_wait(&s_bar, sub & 1);
    } else {
      __syncthreads();          // every warp is through the previous sub-tile
      tile_load_fallback(a, sm, env0, ne);
      __syncthreads();
    }

    LGK_STAMP(1);
    const bool active = lane < ne;
    const int env = env0 + (active ? lane : 0);
    const unsigned env_global = (unsigned)(c.env_id_offset + env);
    float* scr = sm + L.scr;
    float* head_row = sm + L.head + lane * HS;
    const float* cf_row = sm + L.cf + lane * B3;
    const float* trow = (TAPS && add_noise) ? b.tap_noise + (size_t)env * O : nullptr;
    const long long ep_in = reinterpret_cast<const long long*>(sm + L.ep)[lane];
    const long long ep_inc = (flags & LGK_F_EP_INCREMENTED) ? 0 : 1;   // lgk_pre_reward_state already counted this step

    // the reset decision (:143-148) is cheap: every warp forms it for its own lane
    bool time_out = false, reset = false;
    if (active) {
      if (flags & LGK_F_RESET_GIVEN) {   // a python check_termination() override has already decided (cold path)
        time_out = b.time_out_buf[env] != 0;
        reset = b.reset_buf[env] != 0;
      } else {
        time_out = (ep_in + ep_inc) > c.max_episode_length;
        reset = term_contact(c, s_idx, cf_row, a.thr_term) || time_out;
      }
    }
    const unsigned reset_mask = __ballot_sync(LGK_FULL_MASK, active && reset);

    // values W3 keeps in registers from stage 1 to stage 3
    float blv[3] = {0, 0, 0}, bav[3] = {0, 0, 0}, grav[3] = {0, 0, 0};
    float collisions = 0.f, ffp = 0.f, r_air = 0.f, stumble_f = 0.f, nofly_f = 0.f;

    // =========================================== stage 1 =========================================
    if (wic == 0) {
      // ---- W0: episode counter, commands, push, root reset, bitmasks ------------------------------
      bool resample_cb = false;
      int level_contrib = 0;
      if (active) {
        const float* rp = sm + L.root + lane * 13;
        long long ep = ep_in + ep_inc;                              // (:118)
        const float q[4] = {rp[3], rp[4], rp[5], rp[6]};
        float4 cmd = *reinterpret_cast<const float4*>(sm + L.cmd + lane * 4);

        // ---- _post_physics_step_callback (:337-352) -------------------------------------------
        resample_cb = (unsigned long long)ep < 0x80000000ull ? ((int)ep % (int)c.resample_interval) == 0
                                                             : (ep % c.resample_interval) == 0;
        // LGK_F_CALLBACK_GIVEN: a python _post_physics_step_callback() override ran on the commands that
        // lgk_pre_reward_state resampled; they are final (the id list still names the resampled envs)
        const bool cb_given = flags & LGK_F_CALLBACK_GIVEN;
        if (resample_cb && !cb_given) {
          float u0, u1, u2;
          if (TAPS) {
            const float* t = b.tap_cmd_callback + (size_t)env * 3;
            u0 = t[0]; u1 = t[1]; u2 = t[2];
          } else {
            const uint4 r = philox_cold(a, csc, env_global, LGK_SITE_CMD_CB);
            u0 = u01(r.x); u1 = u01(r.y); u2 = u01(r.z);
          }
          resample_command(cmd, b.command_ranges, heading_mode, u0, u1, u2);
        }
        if (heading_mode && !cb_given) {
          float fwd[3];
          quat_apply(q, 1.f, 0.f, 0.f, fwd);
          const float heading = atan2f(fwd[1], fwd[0]);
          cmd.z = clampf(0.5f * wrap_to_pi_fast(cmd.w - heading), -1.f, 1.f);
        }
        // what the rewards see: the command after the callback, the root height before any reset
        scr[SC_CMDX * TILE + lane] = cmd.x;
        scr[SC_CMDY * TILE + lane] = cmd.y;
        scr[SC_CMDZ * TILE + lane] = cmd.z;
        scr[SC_ZPRE * TILE + lane] = rp[2];
        float rvel[6] = {rp[7], rp[8], rp[9], rp[10], rp[11], rp[12]};
        if (push_step) {   // _push_robots (:436-441): every env, after the body-frame velocities
          float u0, u1;
          if (TAPS) { u0 = b.tap_push[(size_t)env * 2]; u1 = b.tap_push[(size_t)env * 2 + 1]; }
          else { const uint4 r = philox_cold(a, csc, env_global, LGK_SITE_PUSH); u0 = u01(r.x); u1 = u01(r.y); }
          const float range = 2.0f * c.max_push_vel_xy;
          rvel[0] = rand_affine(range, -c.max_push_vel_xy, u0);
          rvel[1] = rand_affine(range, -c.max_push_vel_xy, u1);
        }
        long long level = has_levels ? reinterpret_cast<const long long*>(sm + L.lvl)[lane] : 0;
        if (reset) {   // reset_idx for this env (:150-193): curriculum, root state, command
          ResetOut ro;
          ro.cmd = cmd;
          float* side = a.side + (size_t)env * LGK_SIDE_FLOATS;
          reset_root<TAPS>(a, csc, env, env_global, rp[0], rp[1], level, ro, side + 8);
          cmd = ro.cmd;
#pragma unroll
          for (int i = 0; i < 6; ++i) rvel[i] = ro.rvel[i];
          level = ro.level;
          ep = 0;
          if (a.scan_first) {   // the finish kernel redoes this env's scan columns with the new root z (:227)
            const ScanEnv e = scan_env_operands(rp[0], rp[1], ro.z, q[2], q[3]);
            float4* d = reinterpret_cast<float4*>(side);
            d[0] = make_float4(e.k_t.x, e.k_t.y, e.k_c.x, e.k_c.y);
            d[1] = make_float4(e.r2.x, e.r2.y, e.yw, e.zo);
          }
        } else if (push_step) {
          float* rw = b.root_states + (size_t)env * 13;
          rw[7] = rvel[0];
          rw[8] = rvel[1];
        }
        level_contrib = (int)level;

        // ---- per-env stores -------------------------------------------------------------------------
        b.episode_length_buf[env] = ep;
        b.reset_buf[env] = reset ? 1 : 0;
        b.time_out_buf[env] = time_out ? 1 : 0;
        *reinterpret_cast<float4*>(b.commands + (size_t)env * 4) = cmd;
        {
          float2* lr = reinterpret_cast<float2*>(b.last_root_vel + (size_t)env * 6);   // (:138)
          lr[0] = make_float2(rvel[0], rvel[1]); lr[1] = make_float2(rvel[2], rvel[3]); lr[2] = make_float2(rvel[4], rvel[5]);
        }
        // obs head columns 9..11 (:220): the commands after any reset
        uint4 nb = make_uint4(0, 0, 0, 0);
        if (!TAPS && add_noise) nb = philox_keyed(a.pk, csc, env_global, LGK_SITE_NOISE_HEAD + 2);
        head_row[9] = head_obs<TAPS>(cmd.x * c.commands_scale[0], 9, nb, 0, s_ns, trow, clip_obs);
        head_row[10] = head_obs<TAPS>(cmd.y * c.commands_scale[1], 10, nb, 1, s_ns, trow, clip_obs);
        head_row[11] = head_obs<TAPS>(cmd.z * c.commands_scale[2], 11, nb, 2, s_ns, trow, clip_obs);
      }
      // ---- sub-tile bookkeeping: bitmasks for the ordered compaction, level sum --------------------
      const unsigned resample_mask = __ballot_sync(LGK_FULL_MASK, active && resample_cb);
      if ((flags & LGK_F_TERRAIN_CURRICULUM) && has_levels) {
        const int s = warp_sum_int(active ? level_contrib : 0);
        if (lane == 0 && s != 0) atomicAdd(&ws->level_sum[parity], s);
      }
      if (lane == 0) {
        unsigned* mask_base = reinterpret_cast<unsigned*>(reinterpret_cast<char*>(ws) + LGK_WS_HEADER_BYTES);
        mask_base[slot] = reset_mask;
        mask_base[lgk_max_warps(N) + slot] = resample_mask;
      }
      {   // the envs reset in this sub-tile join the finish kernel's list (any order: the entries are independent)
        const int nr = __popc(reset_mask);
        if (nr != 0) {   // warp-uniform
          int base = 0;
          if (lane == 0) base = atomicAdd(&ws->reset_count[parity], nr);
          base = __shfl_sync(LGK_FULL_MASK, base, 0);
          if (active && reset) {
            const int idx = base + __popc(reset_mask & ((1u << lane) - 1u));
            a.fix_list[idx] = env;
          }
        }
      }
    }
    else if (wic == 1 || wic == 2) {
      // ---- W1 / W2: joints (:874-916, :956-966 sums; :399-408 reset; obs head 12..; last_dof_vel) --
      DofAcc acc;
#pragma unroll
      for (int i = 0; i < 10; ++i) acc.v[i] = 0.f;
      if (active) {
        const float inv_dt = 1.0f / c.dt;
        const unsigned hip_mask = a.hip_mask;
        const float* act_p = sm + L.act + lane * D;
        const float* lact_p = sm + L.lact + lane * D;
        const float* tq_p = sm + L.tq + lane * D;
        float* ldv_p = sm + L.ldv + lane * D;            // in: last_dof_vel, out: dof_vel (staged for the store)
        const float* dof_p = sm + L.dof + lane * 2 * D;
        float* dof_g = b.dof_state + (size_t)env * D * 2;
        if constexpr (VEC) {
          // two joints per trip (one 16-byte dof_state quad); W1 takes the first half of the pairs.  The six head
          // columns of a pair take fields 0..5 of Philox block LGK_SITE_NOISE_HEAD + 3 + pair
          constexpr int NPAIR = DT / 2, NP1 = (NPAIR + 1) / 2;
          const int pb = (wic == 1) ? 0 : NP1, pe = (wic == 1) ? NP1 : NPAIR;
#pragma unroll 1
          for (int pc = pb; pc < pe; ++pc) {
            const int d0 = pc * 2;
            const float2 a2 = *reinterpret_cast<const float2*>(act_p + d0);
            const float2 la2 = *reinterpret_cast<const float2*>(lact_p + d0);
            const float2 tq2 = *reinterpret_cast<const float2*>(tq_p + d0);
            const float2 ldv2 = *reinterpret_cast<const float2*>(ldv_p + d0);
            float4 s0 = *reinterpret_cast<const float4*>(dof_p + 2 * d0);       // p0 v0 p1 v1
            float u[2] = {0, 0};
            if (reset) {
              if (TAPS) {
                const float* t = b.tap_dof + (size_t)env * D + d0;
                u[0] = t[0]; u[1] = t[1];
              } else {
                const uint4 r = philox_cold(a, csc, env_global, LGK_SITE_DOF + (d0 >> 2));
                u[0] = u01((d0 & 2) ? r.z : r.x); u[1] = u01((d0 & 2) ? r.w : r.y);
              }
            }
            uint4 nb = make_uint4(0, 0, 0, 0);
            if (!TAPS && add_noise) nb = philox_keyed(a.pk, csc, env_global, LGK_SITE_NOISE_HEAD + 3 + pc);
            float hv[3];
            joint_update(c, d0 + 0, hip_mask, reset, inv_dt, a2.x, la2.x, tq2.x, ldv2.x, s0.x, s0.y, u[0], acc, hv);
            head_row[12 + d0] = head_obs<TAPS>(hv[0], 12 + d0, nb, 0, s_ns, trow, clip_obs);
            head_row[12 + D + d0] = head_obs<TAPS>(hv[1], 12 + D + d0, nb, 1, s_ns, trow, clip_obs);
            head_row[12 + 2 * D + d0] = head_obs<TAPS>(hv[2], 12 + 2 * D + d0, nb, 2, s_ns, trow, clip_obs);
            joint_update(c, d0 + 1, hip_mask, reset, inv_dt, a2.y, la2.y, tq2.y, ldv2.y, s0.z, s0.w, u[1], acc, hv);
            head_row[13 + d0] = head_obs<TAPS>(hv[0], 13 + d0, nb, 3, s_ns, trow, clip_obs);
            head_row[13 + D + d0] = head_obs<TAPS>(hv[1], 13 + D + d0, nb, 4, s_ns, trow, clip_obs);
            head_row[13 + 2 * D + d0] = head_obs<TAPS>(hv[2], 13 + 2 * D + d0, nb, 5, s_ns, trow, clip_obs);
            *reinterpret_cast<float2*>(ldv_p + d0) = make_float2(s0.y, s0.w);   // last_dof_vel = dof_vel (:137)
            if (reset) *reinterpret_cast<float4*>(dof_g + 2 * d0) = s0;
          }
        } else {
          const int half = (D + 1) / 2;
          const int db = (wic == 1) ? 0 : half, de = (wic == 1) ? half : D;
#pragma unroll 1
          for (int d = db; d < de; ++d) {
            const float av = act_p[d], lav = lact_p[d], tqv = tq_p[d], ldvv = ldv_p[d];
            float p = dof_p[2 * d], v = dof_p[2 * d + 1];
            float u = 0.f;
            if (reset) {
              if (TAPS) u = b.tap_dof[(size_t)env * D + d];
              else {
                const uint4 r = philox_cold(a, csc, env_global, LGK_SITE_DOF + (d >> 2));
                u = u01(philox_word(r, d & 3));
              }
            }
            uint4 nb = make_uint4(0, 0, 0, 0);
            if (!TAPS && add_noise) nb = philox_cold(a, csc, env_global, LGK_SITE_NOISE_HEAD + 3 + (d >> 1));
            // fields 3 * (d & 1) .. + 2 of the pair's block, rotated down to fields 0..2
            if (d & 1) nb = make_uint4((nb.y >> 16) | (nb.z << 16), nb.z >> 16, 0, 0);
            float hv[3];
            joint_update(c, d, hip_mask, reset, inv_dt, av, lav, tqv, ldvv, p, v, u, acc, hv);
            head_row[12 + d] = head_obs<TAPS>(hv[0], 12 + d, nb, 0, s_ns, trow, clip_obs);
            head_row[12 + D + d] = head_obs<TAPS>(hv[1], 12 + D + d, nb, 1, s_ns, trow, clip_obs);
            head_row[12 + 2 * D + d] = head_obs<TAPS>(hv[2], 12 + 2 * D + d, nb, 2, s_ns, trow, clip_obs);
            ldv_p[d] = v;
            if (reset) { dof_g[2 * d] = p; dof_g[2 * d + 1] = v; }
          }
        }
      }
      float* jp = scr + ((wic == 1) ? SC_J1 : SC_J2) * TILE + lane;
#pragma unroll
      for (int i = 0; i < 10; ++i) jp[i * TILE] = acc.v[i];
    } else {
      // ---- W3: body-frame velocities and gravity (:123-125), contacts (:895-897, :951-962, cassie.py:43-46),
      //      feet air time (:941-949) ---------------------------------------------------------------------
      if (active) {
        {
          const float* rp = sm + L.root + lane * 13;
          const float q[4] = {rp[3], rp[4], rp[5], rp[6]};
          quat_rotate_inverse(q, rp[7], rp[8], rp[9], blv);
          quat_rotate_inverse(q, rp[10], rp[11], rp[12], bav);
          quat_rotate_inverse(q, 0.f, 0.f, -1.f, grav);
          float* p = b.base_lin_vel + (size_t)env * 3; p[0] = blv[0]; p[1] = blv[1]; p[2] = blv[2];
          p = b.base_ang_vel + (size_t)env * 3;        p[0] = bav[0]; p[1] = bav[1]; p[2] = bav[2];
          p = b.projected_gravity + (size_t)env * 3;   p[0] = grav[0]; p[1] = grav[1]; p[2] = grav[2];
          // obs head columns 0..8 (:217-219): fields 0..7 of one block, field 0 of the next
          uint4 n0 = make_uint4(0, 0, 0, 0), n1 = n0;
          if (!TAPS && add_noise) {
            n0 = philox_keyed(a.pk, csc, env_global, LGK_SITE_NOISE_HEAD + 0);
            n1 = philox_keyed(a.pk, csc, env_global, LGK_SITE_NOISE_HEAD + 1);
          }
          head_row[0] = head_obs<TAPS>(blv[0] * c.obs_scale_lin_vel, 0, n0, 0, s_ns, trow, clip_obs);
          head_row[1] = head_obs<TAPS>(blv[1] * c.obs_scale_lin_vel, 1, n0, 1, s_ns, trow, clip_obs);
          head_row[2] = head_obs<TAPS>(blv[2] * c.obs_scale_lin_vel, 2, n0, 2, s_ns, trow, clip_obs);
          head_row[3] = head_obs<TAPS>(bav[0] * c.obs_scale_ang_vel, 3, n0, 3, s_ns, trow, clip_obs);
          head_row[4] = head_obs<TAPS>(bav[1] * c.obs_scale_ang_vel, 4, n0, 4, s_ns, trow, clip_obs);
          head_row[5] = head_obs<TAPS>(bav[2] * c.obs_scale_ang_vel, 5, n0, 5, s_ns, trow, clip_obs);
          head_row[6] = head_obs<TAPS>(grav[0], 6, n0, 6, s_ns, trow, clip_obs);
          head_row[7] = head_obs<TAPS>(grav[1], 7, n0, 7, s_ns, trow, clip_obs);
          head_row[8] = head_obs<TAPS>(grav[2], 8, n1, 0, s_ns, trow, clip_obs);
        }
        // four bodies per trip: the row reads of a trip are independent, so their latencies overlap
#pragma unroll 1
        for (int p0 = 0; p0 < c.num_pen; p0 += 4) {
          float n[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float* f = cf_row + s_idx[IDX_PEN + min(p0 + j, c.num_pen - 1)];
            n[j] = sumsq3(f[0], f[1], f[2]);
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) collisions += (p0 + j < c.num_pen && n[j] > a.thr_pen) ? 1.f : 0.f;   // norm > 0.1
        }
        const float* fa_in = sm + L.fat + lane * F;
        float* fa_out = b.feet_air_time + (size_t)env * F;
        bool stumble = false;
        int nofly = 0;
#pragma unroll 2
        for (int k = 0; k < F; ++k) {
          const float* f = cf_row + s_idx[IDX_FEET + k];
          const float fx = f[0], fy = f[1], fz = f[2];
          const bool contact = fz > 1.0f;
          nofly += (fz > 0.1f) ? 1 : 0;
          {   // |f_xy| > 5 |f_z| (:957): decided on the squares when they are not within rounding of each other
            const float sxy = fmaf(fy, fy, fx * fx), r = 5.0f * fabsf(fz), r2 = r * r;
            if (sxy > r2 * 1.00001f) stumble = true;
            else if (sxy >= r2 * 0.99999f) stumble |= sqrtf(sxy) > r;
          }
          {   // clip(norm - max_contact_force, min=0) (:962): zero unless norm > max_contact_force
            const float s3 = sumsq3(fx, fy, fz);
            if (s3 > a.thr_ffp) ffp += fmaxf(sqrtf(s3) - c.max_contact_force, 0.f);
          }
          float t = fa_in[k];
          if (a.air_active) {   // the state only advances inside _reward_feet_air_time
            bool on = contact;
            if (flags & LGK_F_FILTERED_CONTACTS) {   // contact now or at the previous step (:931-933, commented upstream)
              uint8_t* lc = b.last_contacts + (size_t)env * F + k;
              on = contact || (*lc != 0);
              *lc = contact ? 1 : 0;
            }
            const bool first = (t > 0.f) && on;
            t = add_rn(t, c.dt);
            r_air += first ? (t - 0.5f) : 0.f;
            if (on) t = 0.f;
          }
          if (a.air_active || reset) fa_out[k] = reset ? 0.f : t;   // zeroed on reset (:178)
        }
        stumble_f = stumble ? 1.f : 0.f;
        nofly_f = (nofly == 1) ? 1.f : 0.f;
      }
    }
    LGK_STAMP(2);
    fence_proxy_async_smem();   // last_dof_vel rows (W1 / W2) -> visible to the bulk store
    __syncthreads();
    LGK_STAMP(3);

    // ---- outputs that are complete after stage 1 ------------------------------------------------------
    if (full) {
      if (tid == 32) {
        bulk_s2g(b.last_actions + (size_t)env0 * D, sm + L.act, ne * 4 * D);   // last_actions = actions (:136)
        bulk_s2g(b.last_dof_vel + (size_t)env0 * D, sm + L.ldv, ne * 4 * D);
        bulk_commit();
      }
    } else if (wic > 0) {
      tile_store_fallback(a, sm, env0, ne, tid);
    }
    if (wic < 3) {
      // ---- W0..W2: the staged heads -> observation rows, lanes = columns, 128 B per warp store -------------
#pragma unroll 2
      for (int e = wic; e < ne; e += 3) {
        const float* hrow = sm + L.head + e * HS;
        float* orow = b.obs_buf + (size_t)(env0 + e) * O;
        const float o0 = hrow[lane];
        if (lane + 32 < OH) orow[lane + 32] = hrow[lane + 32];
        if (lane < OH) orow[lane] = o0;
      }
      if (full && tid == 32) bulk_wait_read_all();   // the bulk stores have read their rows before the next loads
      LGK_STAMP(4);
    } else {
      // ---- W3: every term's value for its env into the [term][N] table of the workspace (one coalesced store
      //      per term); the finish kernel scales, sums and accumulates them (:195-212).  Rows of terms that are not
      //      active are never read.
      if (active) {
        float* rvg = a.rv + env;
        const size_t NP = (size_t)a.rv_pitch;
        auto put = [&](int term, float v) { rvg[(size_t)term * NP] = v; };
        auto J = [&](int k) { return scr[(SC_J1 + k) * TILE + lane] + scr[(SC_J2 + k) * TILE + lane]; };
        const float cx = scr[SC_CMDX * TILE + lane], cy = scr[SC_CMDY * TILE + lane], cz = scr[SC_CMDZ * TILE + lane];
        const float cn = sqrtf(fmaf(cy, cy, cx * cx));
        put(LGK_REW_ANG_VEL_XY, fmaf(bav[1], bav[1], bav[0] * bav[0]));
        put(LGK_REW_LIN_VEL_Z, blv[2] * blv[2]);
        put(LGK_REW_ORIENTATION, fmaf(grav[1], grav[1], grav[0] * grav[0]));
        put(LGK_REW_FEET_AIR_TIME, r_air * ((cn > 0.1f) ? 1.f : 0.f));
        put(LGK_REW_STAND_STILL, J(A_STAND) * ((cn < 0.1f) ? 1.f : 0.f));
        {
          const float e = cz - bav[2];
          put(LGK_REW_TRACKING_ANG_VEL, expf(-(e * e) / c.tracking_sigma));
          const float ex = cx - blv[0], ey = cy - blv[1];
          put(LGK_REW_TRACKING_LIN_VEL, expf(-fmaf(ey, ey, ex * ex) / c.tracking_sigma));
        }
        put(LGK_REW_ACTION_RATE, J(A_ARATE));
        put(LGK_REW_DOF_ACC, J(A_ACC2));
        put(LGK_REW_DOF_POS_LIMITS, J(A_POSLIM));
        put(LGK_REW_DOF_VEL, J(A_QD2));
        put(LGK_REW_DOF_VEL_LIMITS, J(A_VELLIM));
        put(LGK_REW_ENERGY, J(A_ENERGY));
        put(LGK_REW_HIP_MOTION, J(A_HIP));
        put(LGK_REW_TORQUE_LIMITS, J(A_TQLIM));
        put(LGK_REW_TORQUES, J(A_TQ2));
        put(LGK_REW_COLLISION, collisions);
        put(LGK_REW_FEET_CONTACT_FORCES, ffp);
        put(LGK_REW_NO_FLY, nofly_f);
        put(LGK_REW_STUMBLE, stumble_f);
        // _reward_base_height (:871) needs the measured heights: the root height before any reset goes into the
        // term's row, the finish kernel subtracts the scan's mean height and squares the distance to the target
        put(LGK_REW_BASE_HEIGHT, scr[SC_ZPRE * TILE + lane]);
      }
      LGK_STAMP(6);
    }
  }   // sub-tiles
  if (a.fused == 2) {
    // hand this CTA's envs over to the scan kernel: everything the four warps wrote (and the bulk stores, complete --
    // not only read -- for the thread that committed them) is ordered before the flag
    if (tid == 32) bulk_wait_all();
    __syncthreads();
    if (tid == 32) flag_release(a.scan_flags + blockIdx.x, ws->launch_seq + 1u);
  }
  if (wic == 0 && lane == 0) LGK_CTA_STAMP(1);
#ifdef LGK_PROFILE
  if (wic == 3 && lane == 0 && blockIdx.x < 4096) lgk_prof_cta[blockIdx.x * 4 + 3] = lgk_gtime();   // W3 leaves
#endif
}
// ---------------------------------------------------------------------------------------------
// Finish kernel: compute_reward's sums (:195-212), the tail of reset_idx, the id lists and the step's statistics.
// Launched behind the step kernel (griddepcontrol.wait: its value table, flags, masks, reset list and side rows are
// complete); the scan kernel's part -- height sums, scan columns, its last read of the root states -- is taken over
// group by group through the flags it stamps with this launch's sequence number.  Three kinds of CTA, by blockIdx:
//   CTA 0          the ascending reset / resample id lists from the step kernel's per-tile bitmasks
//   fix CTAs       the envs reset this step (lanes = scan points; warp w takes list entries w, w + W, ...): the scan
//                  columns are rewritten with the root z after the reset (:227; indices, heights and noise are what the
//                  scan kernel computed) and the new root state (_reset_root_states, :414-434), held back in the side
//                  buffer while the scan kernel was reading poses, is stored.  CTAs of their own, scheduled first: the
//                  chain list -> side row -> flag -> gathers -> stores is as long as the reward chain and runs next
//                  to it instead of behind it
//   reward CTAs    lanes = envs, one warp per 32 consecutive envs; every access to the [R,N] episode-sum rows and the
//                  [term][N] value table is a coalesced 128 B row segment: r = value * scale, sum += r, reward += r in
//                  the reference's order; only_positive clip; termination term.  Sums of reset envs go by red.global.add
//                  into one of STAT_COPIES replicated accumulator rows picked by the env's warp (all into ONE row cost
//                  15 us: ~12 000 red.add on a single cache line at 65 536 envs)
// and the CTA that finishes last adds the replicated rows up into extras["episode"] and advances the step counter.
#define LGK_FIX_CTAS_MAX 296   // at most 2 x 148 CTAs of the finish kernel work on the reset envs
template <bool TAPS, bool MUL_RCP, int NP>
__device__ __forceinline__ void finish_fix_cta(const StepArgs& a, int fix_ctas) {
  const LgkBuffers& b = a.b;
  const int lane = threadIdx.x & 31;
  LgkWorkspace* ws = reinterpret_cast<LgkWorkspace*>(b.workspace);
  const unsigned long long csc = ws->step_counter + 1ull;
  const unsigned stamp = ws->launch_seq + 1u;
  const int n_reset = ws->reset_count[(int)(csc & 1ull)];
  const int gw = ((int)blockIdx.x - 1) * (FIN_THREADS / 32) + (int)(threadIdx.x >> 5), nw = fix_ctas * (FIN_THREADS / 32);
#pragma unroll 1
  for (int i = gw; i < n_reset; i += nw) {
    const int e = a.fix_list[i];
    const float* side = a.side + (size_t)e * LGK_SIDE_FLOATS;
    if (a.scan_first) {
      if (lane == 0) wait_scan_flag(a, ws, a.scan_flags + e / a.scan_epc, stamp);
      __syncwarp();
      lgk_scan_fixup_env<TAPS, MUL_RCP, NP>(a, side, e, csc);
    }
    if (lane < 13) b.root_states[(size_t)e * 13 + lane] = side[8 + lane];   // the row the step kernel held back
  }
}
__device__ __forceinline__ void finish_reward_cta(const StepArgs& a, int first_cta) {
  // the loads of a batch of FB rows are issued together, and the next batch's loads before this batch's stores (same
  // array: the compiler may not move them there by itself).  Nothing is waited for before the first batch is out.
  const LgkConsts& c = a.c;
  const LgkBuffers& b = a.b;
  const int N = c.num_envs;
  const int lane = threadIdx.x & 31;
  const unsigned flags = c.flags;
  LgkWorkspace* ws = reinterpret_cast<LgkWorkspace*>(b.workspace);
  const int env = ((int)blockIdx.x - first_cta) * FIN_THREADS + threadIdx.x;
  const int wenv0 = env - lane;
  if (wenv0 >= N) return;
  const int ne = min(32, N - wenv0);
  const bool active = lane < ne;
  const int nrew = c.num_rewards;
  const bool has_term = flags & LGK_F_HAS_TERMINATION;
  constexpr int FB = 8;
  const size_t NP_ = (size_t)a.rv_pitch;
  const int envc = active ? env : wenv0;        // inactive lanes shadow lane 0 (loads only)
  const float* rvg = a.rv + envc;
  float* gsum = b.episode_sums + envc;
  const int nrows = nrew + (has_term ? 1 : 0);
  float sv[2][FB], vv[2][FB];
  int xi = 0;   // index among the external rows
  auto load_batch = [&](int r0, float (&s8)[FB], float (&v8)[FB]) {
#pragma unroll
    for (int j = 0; j < FB; ++j) {
      const int r = r0 + j;
      s8[j] = 0.f; v8[j] = 0.f;
      if (r < nrows) s8[j] = gsum[(size_t)r * N];
      if (r < nrew) {
        const int term = c.rewards[r].term;
        if (term == LGK_REW_EXTERNAL) { v8[j] = b.ext_rewards[(size_t)xi * N + envc]; ++xi; }   // a python _reward_<name>
        else v8[j] = rvg[(size_t)term * NP_];      // (base_height: the root z before any reset)
      }
    }
  };
  const bool reset = b.reset_buf[envc] != 0, time_out = b.time_out_buf[envc] != 0;
  load_batch(0, sv[0], vv[0]);
  float bh = 0.f;
  if (a.scan_first && a.bh_active) {
    // _reward_base_height (:871): mean_h(root_z - measured_height) = root_z - mean_h(height); the height sums of this
    // warp's envs are out once every scan group overlapping them has been handed over (one lane per flag)
    const unsigned stamp = ws->launch_seq + 1u;
    const int f0 = wenv0 / a.scan_epc, f1 = (wenv0 + ne - 1) / a.scan_epc;
    if (f0 + lane <= f1) wait_scan_flag(a, ws, a.scan_flags + f0 + lane, stamp);
    __syncwarp();
    bh = ((float)a.hsum[envc] * c.vertical_scale) / (float)c.num_height_points;
  }
  if (!active) return;
  float* stat = a.stat_copies + (size_t)((wenv0 >> 5) & (STAT_COPIES - 1)) * LGK_MAX_REWARDS;
  float rew = 0.f;
  auto finish_batch = [&](int r0, const float (&s8)[FB], const float (&v8)[FB]) {
#pragma unroll
    for (int j = 0; j < FB; ++j) {
      const int r = r0 + j;
      if (r < nrew) {
        float v = v8[j];
        if (c.rewards[r].term == LGK_REW_BASE_HEIGHT) { const float d = (v - bh) - c.base_height_target; v = d * d; }
        const float rr = mul_rn(v, c.rewards[r].scale);
        const float total = add_rn(s8[j], rr);
        rew = add_rn(rew, rr);
        gsum[(size_t)r * N] = reset ? 0.f : total;   // zeroed for reset envs after the statistics took it (:184-185)
        if (reset) atomicAdd(stat + r, total);
      } else if (r == nrew) {
        if (flags & LGK_F_ONLY_POSITIVE) rew = fmaxf(rew, 0.f);
        if (has_term) {   // added after the positive clip (:206-212)
          const float rr = mul_rn((reset && !time_out) ? 1.f : 0.f, c.termination_scale);
          const float total = add_rn(s8[j], rr);
          rew = add_rn(rew, rr);
          gsum[(size_t)r * N] = reset ? 0.f : total;
          if (reset) atomicAdd(stat + r, total);
        }
      }
    }
  };
#pragma unroll 1
  for (int r0 = 0; r0 <= nrew; r0 += 2 * FB) {   // (<=: the batch holding row nrew does the clip / termination)
    if (r0 + FB <= nrew) load_batch(r0 + FB, sv[1], vv[1]);
    finish_batch(r0, sv[0], vv[0]);
    if (r0 + FB <= nrew) {
      if (r0 + 2 * FB <= nrew) load_batch(r0 + 2 * FB, sv[0], vv[0]);
      finish_batch(r0 + FB, sv[1], vv[1]);
    }
  }
  b.rew_buf[env] = rew;
}
template <bool TAPS, bool MUL_RCP, int NP>
__global__ void __launch_bounds__(FIN_THREADS) lgk_finish_kernel(const __grid_constant__ StepArgs a, int fix_ctas,
                                                                 int step_ctas) {
  pdl_launch_dependents();   // whatever follows on the stream may take its slots; it waits for this grid to complete
  __shared__ unsigned s_last;
  LgkWorkspace* ws = reinterpret_cast<LgkWorkspace*>(a.b.workspace);
  pdl_wait();
  const unsigned long long csc = ws->step_counter + 1ull;
  if (blockIdx.x == 0) compact_id_lists<FIN_THREADS>(a, step_ctas);
  else if ((int)blockIdx.x <= fix_ctas) finish_fix_cta<TAPS, MUL_RCP, NP>(a, fix_ctas);
  else finish_reward_cta(a, fix_ctas + 1);
  // ---- the last CTA to get here publishes the step's statistics and advances the counters ------------------------
  __syncthreads();   // (every thread of the CTA, whatever it returned from above)
  if (threadIdx.x == 0) {
    __threadfence();
    s_last = (atomicAdd(&ws->fin_ticket, 1u) == gridDim.x - 1) ? 1u : 0u;
  }
  __syncthreads();
  if (!s_last) return;   // (CTA-uniform)
  __threadfence();
  publish_step_statistics(a, csc);
}
// ---------------------------------------------------------------------------------------------
static inline int up4(int x) { return (x + 3) & ~3; }
// the largest float x with sqrtf(x) <= c  (c >= 0; host sqrtf is correctly rounded like the device's sqrt.rn)
static float sqrt_le_threshold(float c) {
  if (!(c >= 0.f)) return -1.f;   // sqrtf(s) > negative c: always (s >= 0)
  if (std::isinf(c)) return c;
  float x = c * c;
  if (std::isinf(x)) x = std::numeric_limits<float>::max();
  while (sqrtf(x) <= c && x < std::numeric_limits<float>::max()) x = std::nextafterf(x, INFINITY);
  while (sqrtf(x) > c) x = std::nextafterf(x, -INFINITY);
  return x;
}
static inline bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// ---------------------------------------------------------------------------------------------
// State the reference has when compute_reward() runs (:118-131), for reward terms that live in Python
// (a subclass's own _reward_<name>, LGK_REW_EXTERNAL): episode counter, body-frame velocities and gravity,
// commands after the callback's resample + heading update, measured heights (when materialised) and the
// termination flags.  One thread per env, plain global loads: this is the cold path of user-defined terms,
// not the tuned one.  Everything it writes is either a pure output of the step kernel or idempotent -- the
// step kernel that follows redraws the same random numbers (counter-based: same step, env, site) and
// recomputes the same values -- except episode_length_buf, which the step kernel is told not to count
// again (LGK_F_EP_INCREMENTED).  The push (:351) is left to the step kernel: body-frame velocities must be
// formed from the un-pushed state (:123-124).
template <bool TAPS>
__global__ void lgk_pre_reward_kernel(const __grid_constant__ StepArgs a, int mul_rcp) {
  const LgkConsts& c = a.c;
  const LgkBuffers& b = a.b;
  const int env = blockIdx.x * blockDim.x + threadIdx.x;
  if (env >= c.num_envs) return;
  const unsigned flags = c.flags;
  const bool heading_mode = flags & LGK_F_HEADING_COMMAND;
  const LgkWorkspace* ws = reinterpret_cast<const LgkWorkspace*>(b.workspace);
  const unsigned long long csc = ws->step_counter + 1ull;
  const unsigned env_global = (unsigned)(c.env_id_offset + env);
  float root[13];
#pragma unroll
  for (int i = 0; i < 13; ++i) root[i] = b.root_states[(size_t)env * 13 + i];
  const long long ep = b.episode_length_buf[env] + 1;            // (:118)
  b.episode_length_buf[env] = ep;
  const float q[4] = {root[3], root[4], root[5], root[6]};
  float v[3];
  quat_rotate_inverse(q, root[7], root[8], root[9], v);
  b.base_lin_vel[(size_t)env * 3] = v[0]; b.base_lin_vel[(size_t)env * 3 + 1] = v[1]; b.base_lin_vel[(size_t)env * 3 + 2] = v[2];
  quat_rotate_inverse(q, root[10], root[11], root[12], v);
  b.base_ang_vel[(size_t)env * 3] = v[0]; b.base_ang_vel[(size_t)env * 3 + 1] = v[1]; b.base_ang_vel[(size_t)env * 3 + 2] = v[2];
  quat_rotate_inverse(q, 0.f, 0.f, -1.f, v);
  b.projected_gravity[(size_t)env * 3] = v[0]; b.projected_gravity[(size_t)env * 3 + 1] = v[1]; b.projected_gravity[(size_t)env * 3 + 2] = v[2];
  // _post_physics_step_callback (:337-349)
  float4 cmd = *reinterpret_cast<const float4*>(b.commands + (size_t)env * 4);
  if ((ep % c.resample_interval) == 0) {
    float u0, u1, u2;
    if (TAPS) { const float* t = b.tap_cmd_callback + (size_t)env * 3; u0 = t[0]; u1 = t[1]; u2 = t[2]; }
    else { const uint4 r = philox_cold(a, csc, env_global, LGK_SITE_CMD_CB); u0 = u01(r.x); u1 = u01(r.y); u2 = u01(r.z); }
    resample_command(cmd, b.command_ranges, heading_mode, u0, u1, u2);
  }
  if (heading_mode) {
    float fwd[3];
    quat_apply(q, 1.f, 0.f, 0.f, fwd);
    cmd.z = clampf(0.5f * wrap_to_pi_fast(cmd.w - atan2f(fwd[1], fwd[0])), -1.f, 1.f);
  }
  *reinterpret_cast<float4*>(b.commands + (size_t)env * 4) = cmd;
  // _get_heights (:818-854), scalar op-for-op chain (the step kernel's packed chain rounds identically)
  const int H = c.num_height_points;
  if (b.measured_heights != nullptr && H > 0) {
    float* mrow = b.measured_heights + (size_t)env * H;
    if (flags & LGK_F_TERRAIN_PLANE) {
      for (int h = 0; h < H; ++h) mrow[h] = 0.f;
    } else {
      const float n = fmaxf(sqrtf(add_rn(mul_rn(q[2], q[2]), mul_rn(q[3], q[3]))), 1e-9f);
      const float yz = __fdiv_rn(q[2], n), yw = __fdiv_rn(q[3], n);
      const float yz2 = mul_rn(yz, 2.0f), inv_hs = 1.0f / c.horizontal_scale;
      for (int h = 0; h < H; ++h) {
        const float px = b.height_points_xy[2 * h], py = b.height_points_xy[2 * h + 1];
        const float tx = mul_rn(-yz2, py), ty = mul_rn(yz2, px);
        float wx = add_rn(add_rn(px, mul_rn(yw, tx)), mul_rn(-yz, ty));
        float wy = add_rn(add_rn(py, mul_rn(yw, ty)), mul_rn(yz, tx));
        wx = add_rn(add_rn(wx, root[0]), c.border_size);
        wy = add_rn(add_rn(wy, root[1]), c.border_size);
        const float sx = mul_rcp ? mul_rn(wx, inv_hs) : __fdiv_rn(wx, c.horizontal_scale);
        const float sy = mul_rcp ? mul_rn(wy, inv_hs) : __fdiv_rn(wy, c.horizontal_scale);
        // .long() truncates toward zero; clamp in floating point first so that the conversion cannot overflow
        int ix = (int)fminf(fmaxf(sx, -1.0f), 1.0e7f), iy = (int)fminf(fmaxf(sy, -1.0f), 1.0e7f);
        ix = min(max(ix, 0), c.terrain_rows - 2);
        iy = min(max(iy, 0), c.terrain_cols - 2);
        mrow[h] = mul_rn(s16_to_float(b.height_min3[(size_t)ix * c.terrain_pitch + iy]), c.vertical_scale);
      }
    }
  }
  // check_termination (:143-148)
  bool term = false;
  const float* cf = b.contact_forces + (size_t)env * c.num_bodies * 3;
  for (int k = 0; k < c.num_term; ++k) {
    const float* f = cf + 3 * c.term_idx[k];
    term |= sumsq3(f[0], f[1], f[2]) > a.thr_term;
  }
  const bool time_out = ep > c.max_episode_length;
  b.time_out_buf[env] = time_out ? 1 : 0;
  b.reset_buf[env] = (term || time_out) ? 1 : 0;
}

// lgk_set_step_counter: the header starts over (both parities' accumulators, tickets) with the given counter; the
// launch sequence number stays -- scan flags stamped by earlier launches must never match a later one
__global__ void lgk_set_counter_kernel(LgkWorkspace* ws, unsigned long long value) {
  const unsigned seq = ws->launch_seq;
  unsigned* w = reinterpret_cast<unsigned*>(ws);
  __syncthreads();
  for (int i = threadIdx.x; i < (int)(sizeof(LgkWorkspace) / 4); i += blockDim.x) w[i] = 0u;
  __syncthreads();
  if (threadIdx.x == 0) { ws->step_counter = value; ws->launch_seq = seq; }
}
extern "C" int lgk_set_step_counter_impl(void* workspace, long long value, cudaStream_t stream) {
  lgk_set_counter_kernel<<<1, 128, 0, stream>>>(reinterpret_cast<LgkWorkspace*>(workspace), (unsigned long long)value);
  return (int)cudaGetLastError();
}

static void fill_rng_and_thresholds(StepArgs& args, const LgkConsts* c) {
  for (int r = 0; r < 10; ++r) {
    args.pk[2 * r] = (unsigned)c->seed + (unsigned)r * 0x9E3779B9u;
    args.pk[2 * r + 1] = (unsigned)(c->seed >> 32) + (unsigned)r * 0xBB67AE85u;
  }
  args.thr_term = sqrt_le_threshold(1.0f);
  args.thr_pen = sqrt_le_threshold(0.1f);
  args.thr_ffp = sqrt_le_threshold(c->max_contact_force);
}

extern "C" int lgk_pre_reward_state_impl(const LgkConsts* c, const LgkBuffers* b, cudaStream_t stream) {
  StepArgs args = {};
  args.c = *c;
  args.b = *b;
  fill_rng_and_thresholds(args, c);
  const int grid = (c->num_envs + 127) / 128;
  const int mul = c->index_math == LGK_INDEX_MUL_RCP;
  if (c->rng_mode == LGK_RNG_TAPS) lgk_pre_reward_kernel<true><<<grid, 128, 0, stream>>>(args, mul);
  else lgk_pre_reward_kernel<false><<<grid, 128, 0, stream>>>(args, mul);
  return (int)cudaGetLastError();
}

// per-device facts and per-(device, kernel) set-up, looked up from the stream's device on every call: the library
// keeps no process-wide "current" state (two envs on two devices, or on two streams of one, never share any)
#define LGK_MAX_DEVICES 64
struct DeviceInfo { int num_sms; };
// The device the launch goes to is the calling thread's current device: CUDA rejects a launch on a stream that
// belongs to another device, so "current device" and "the stream's device" are the same thing for every call that
// can succeed (cudaStreamGetDevice would say the same but is not allowed under stream capture).
static int device_of_stream(cudaStream_t, int* dev) { return (int)cudaGetDevice(dev); }
static int device_info(int dev, DeviceInfo* out) {
  static DeviceInfo info[LGK_MAX_DEVICES];
  static bool known[LGK_MAX_DEVICES];   // idempotent: a race writes the same values twice
  if (dev < 0 || dev >= LGK_MAX_DEVICES) return LGK_ERR_DIM;
  if (!known[dev]) {
    LGK_CUDA_RETURN_IF(cudaDeviceGetAttribute(&info[dev].num_sms, cudaDevAttrMultiProcessorCount, dev));
    known[dev] = true;
  }
  *out = info[dev];
  return 0;
}

// The chain.  Default: two launches -- step kernel, then the tail kernel (scan + rewards + reset tail + statistics,
// lgk_scan.cu); from 12 288 envs on the tail kernel takes the step CTAs' results over through hand-over flags (it starts
// under the step kernel's last wave), below that it waits for the whole step grid (measured: 4 096 envs 25.3 us against
// 26.7 us with flags, 16 384 envs 31.8 / 30.9, 65 536 envs 68.9 / 67.5).  LGK_CHAIN in the environment overrides:
// "3": the three-launch chain of earlier rounds (scan, step, finish kernel), kept for A/B measurements; "2w": tail kernel
// with the grid-wide wait; "2f": tail kernel with hand-over flags.  Returns 0 (three launches), 1 (wait) or 2 (flags).
extern "C" int lgk_chain_mode(int num_envs) {
  static const int forced = [] {
    const char* e = getenv("LGK_CHAIN");
    if (e && e[0] == '3') return 0;
    if (e && e[0] == '2' && e[1] == 'w') return 1;
    if (e && e[0] == '2' && e[1] == 'f') return 2;
    return -1;
  }();
  if (forced >= 0) return forced;
  return num_envs >= 12288 ? 2 : 1;
}

template <int DT, bool TAPS, bool MUL_RCP>
static int launch_step(const StepArgs& args, int dev, int grid, size_t smem, cudaStream_t stream) {
  auto kern = lgk_step_kernel<DT, TAPS, MUL_RCP>;
  static size_t configured[LGK_MAX_DEVICES];   // per instantiation and device; the attribute is per device
  if (smem > 48 * 1024 && smem > configured[dev]) {
    LGK_CUDA_RETURN_IF(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured[dev] = smem;
  }
  static const bool dbg = [] { const char* e = getenv("LGK_DEBUG_SYNC"); return e && e[0] == '1'; }();
  auto dbg_sync = [&](const char* what) {   // development: name the launch a fault belongs to
    if (!dbg) return;
    const cudaError_t e = cudaStreamSynchronize(stream);
    fprintf(stderr, "[lgk] %s: %s\n", what, cudaGetErrorString(e));
  };
  dbg_sync("before post_physics_step");
  if (args.fused) {   // two launches: step kernel, then scan + rewards + statistics
    LGK_CUDA_RETURN_IF(lgk_launch_pdl(kern, dim3(grid), dim3(LGK_CTA_THREADS), smem, stream, args));
    dbg_sync("step kernel");
    const int rc = lgk_scan_launch(args, stream);
    if (rc) return rc;
    dbg_sync("scan kernel (fused)");
    return (int)cudaGetLastError();
  }
  if (args.scan_first) { const int rc = lgk_scan_launch(args, stream); if (rc) return rc; }
  dbg_sync("scan kernel");
  LGK_CUDA_RETURN_IF(lgk_launch_pdl(kern, dim3(grid), dim3(LGK_CTA_THREADS), smem, stream, args));
  dbg_sync("step kernel");
  {
    // CTA 0: id lists; then fix_ctas CTAs for the envs reset this step (about 1 % of them: one or two per warp); then
    // one CTA per 128 envs
    int fix_ctas = args.c.num_envs / 256;
    fix_ctas = fix_ctas < 1 ? 1 : fix_ctas > LGK_FIX_CTAS_MAX ? LGK_FIX_CTAS_MAX : fix_ctas;
    const dim3 fgrid(1 + fix_ctas + (args.c.num_envs + FIN_THREADS - 1) / FIN_THREADS);
    const int np = (args.c.num_height_points + 31) / 32;
    cudaError_t e;
    if (np == 6) e = lgk_launch_pdl(lgk_finish_kernel<TAPS, MUL_RCP, 6>, fgrid, dim3(FIN_THREADS), 0, stream, args, fix_ctas, grid);
    else if (np == 4) e = lgk_launch_pdl(lgk_finish_kernel<TAPS, MUL_RCP, 4>, fgrid, dim3(FIN_THREADS), 0, stream, args, fix_ctas, grid);
    else e = lgk_launch_pdl(lgk_finish_kernel<TAPS, MUL_RCP, 0>, fgrid, dim3(FIN_THREADS), 0, stream, args, fix_ctas, grid);
    LGK_CUDA_RETURN_IF(e);
  }
  dbg_sync("finish kernel");
  return (int)cudaGetLastError();
}

extern "C" int lgk_post_physics_step_impl(const LgkConsts* c, const LgkBuffers* b, cudaStream_t stream) {
  StepArgs args;
  args.c = *c;
  args.b = *b;
  const int D = c->num_dof, B3 = c->num_bodies * 3, F = c->num_feet;
  const int OH = 12 + 3 * D;
  args.num_rows = c->num_rewards + ((c->flags & LGK_F_HAS_TERMINATION) ? 1 : 0);
  args.head_stride = OH | 1;
  args.one = 1.0f;
  args.scan_first = c->num_height_points > 0 ? 1 : 0;
  args.hsum = reinterpret_cast<int*>(static_cast<char*>(b->workspace) + lgk_ws_hsum_offset(c->num_envs));
  args.expo23 = 0x4b000000u;
  args.scan_flags = reinterpret_cast<unsigned*>(static_cast<char*>(b->workspace) + lgk_ws_flags_offset(c->num_envs));
  args.rv = reinterpret_cast<float*>(static_cast<char*>(b->workspace) + lgk_ws_rv_offset(c->num_envs));
  args.rv_pitch = lgk_ws_rv_pitch(c->num_envs);
  args.side = reinterpret_cast<float*>(static_cast<char*>(b->workspace) + lgk_ws_side_offset(c->num_envs));
  args.fix_list = reinterpret_cast<int*>(static_cast<char*>(b->workspace) + lgk_ws_fix_offset(c->num_envs));
  args.stat_copies = reinterpret_cast<float*>(static_cast<char*>(b->workspace) + lgk_ws_tot_offset(c->num_envs));
  fill_rng_and_thresholds(args, c);
  args.hip_mask = 0u;
  for (int h = 0; h < c->num_hip; ++h) args.hip_mask |= 1u << c->hip_idx[h];
  args.air_active = 0;
  args.bh_active = 0;
  for (int i = 0; i < c->num_rewards; ++i) {
    args.air_active |= (c->rewards[i].term == LGK_REW_FEET_AIR_TIME);
    args.bh_active |= (c->rewards[i].term == LGK_REW_BASE_HEIGHT);
  }
  // shared-memory layout (floats); every region starts on a 16-byte boundary
  TileLayout& L = args.L;
  int o = 0;
  L.root = o; o += up4(TILE * 13);
  L.dof = o;  o += up4(TILE * 2 * D);
  L.act = o;  o += up4(TILE * D);
  L.ldv = o;  o += up4(TILE * D);
  L.cmd = o;  o += TILE * 4;
  L.fat = o;  o += up4(TILE * (F > 0 ? F : 1));
  L.ep = o;   o += TILE * 2;
  L.lvl = o;  o += TILE * 2;
  L.cf = o;   o += up4(TILE * B3);
  L.tq = o;   o += up4(TILE * D);
  L.lact = o; o += up4(TILE * D);
  L.head = o; o += up4(TILE * args.head_stride);
  L.scr = o;  o += TILE * SC_FIELDS;
  L.total = o;
  const size_t smem = (size_t)L.total * sizeof(float);
  if (smem > 200 * 1024) return LGK_ERR_DIM;
  {
    const void* src[LGK_LD_ARRAYS] = {b->root_states, b->episode_length_buf, b->contact_forces, b->commands, b->dof_state,
                                      b->actions, b->last_actions, b->torques, b->last_dof_vel, b->feet_air_time,
                                      b->terrain_levels};
    const int off[LGK_LD_ARRAYS] = {L.root, L.ep, L.cf, L.cmd, L.dof, L.act, L.lact, L.tq, L.ldv, L.fat, L.lvl};
    const int bpe[LGK_LD_ARRAYS] = {13 * 4, 8, B3 * 4, 16, D * 8, D * 4, D * 4, D * 4, D * 4, F * 4, 8};
    args.ld_bytes_per_env = 0;
    for (int i = 0; i < LGK_LD_ARRAYS; ++i) {
      const bool on = src[i] != nullptr && bpe[i] > 0;
      args.ld_src[i] = on ? static_cast<const char*>(src[i]) : nullptr;
      args.ld_off[i] = (unsigned)off[i];
      args.ld_bpe[i] = on ? (unsigned)bpe[i] : 0u;
      args.ld_bytes_per_env += args.ld_bpe[i];
    }
  }
  // bulk (TMA) path: all tile base addresses and the episode-sum rows must be 16-byte aligned
  args.tma_ok = (c->num_envs % 4 == 0) && al16(b->root_states) && al16(b->dof_state) && al16(b->contact_forces) &&
                al16(b->torques) && al16(b->actions) && al16(b->last_actions) && al16(b->last_dof_vel) &&
                al16(b->commands) && al16(b->feet_air_time) && al16(b->episode_length_buf) &&
                (args.num_rows == 0 || al16(b->episode_sums)) && (b->terrain_levels == nullptr || al16(b->terrain_levels));
  // Grid.  More envs than one wave of 32-env tiles: one tile per CTA, the hardware scheduler backfills the slots.
  // Fewer: the envs are dealt to all the resident slots in equal smaller tiles.  envs_per_cta > 0 overrides both.
  int dev = 0;
  { const int rc = device_of_stream(stream, &dev); if (rc) return rc; }
  DeviceInfo di;
  { const int rc = device_info(dev, &di); if (rc) return rc; }
  const int num_sms = di.num_sms;
  int grid;
  if (c->envs_per_cta > 0) {
    args.per = (c->envs_per_cta + 3) & ~3;
    grid = (c->num_envs + args.per - 1) / args.per;
    args.n_big = grid;
    args.per_small = args.per;
  } else {
    int per_sm = (int)((227u * 1024u) / (smem + 1024u));
    if (per_sm > LGK_MIN_CTAS) per_sm = LGK_MIN_CTAS;
    if (per_sm < 1) per_sm = 1;
    const int slots = num_sms * per_sm;
    args.per = TILE;
    args.n_big = c->num_envs > TILE * slots ? (c->num_envs + TILE - 1) / TILE : 0;
    const int rest = c->num_envs - args.n_big * TILE;   // > 0 only in the single-wave case
    args.per_small = (((rest + slots - 1) / slots) + 3) & ~3;
    if (args.per_small < 4) args.per_small = 4;
    grid = args.n_big + (rest > 0 ? (rest + args.per_small - 1) / args.per_small : 0);
  }
  args.max_sub = (args.per + TILE - 1) / TILE;
  args.step_grid = grid;
  args.fused = lgk_chain_mode(c->num_envs);
  {
    int xi = 0;
    args.bh_row = -2;
    for (int r = 0; r < LGK_MAX_REWARDS; ++r) {
      args.row_src[r] = args.rv;   // (rows past the terms: never used as a value, but always a valid address)
      args.row_scale[r] = r == c->num_rewards ? c->termination_scale : 0.f;
      if (r < c->num_rewards) {
        const int term = c->rewards[r].term;
        args.row_src[r] = term == LGK_REW_EXTERNAL ? b->ext_rewards + (size_t)(xi++) * c->num_envs
                                                   : args.rv + (size_t)term * args.rv_pitch;
        args.row_scale[r] = c->rewards[r].scale;
        if (term == LGK_REW_BASE_HEIGHT) args.bh_row = r;
      }
    }
  }
  // scan kernel: a persistent grid, resident at once; its warps walk groups of 8 envs (fewer when that would not
  // give every warp a group)
  {
    static const int scan_per_sm = [] {
      const char* e = getenv("LGK_SCAN_CTAS");
      const int v = e ? atoi(e) : 0;
      return (v >= 1 && v <= LGK_SCAN_REG_CTAS) ? v : LGK_SCAN_CTAS_PER_SM;
    }();
    const int cap = num_sms * (args.fused ? LGK_TAIL_CTAS_PER_SM : scan_per_sm);
    // fused chain: groups of 16 as long as they keep 4/5 of the warps busy (65 536 envs: 4 096 groups, one each)
    static const int tail_epg = [] { const char* e = getenv("LGK_TAIL_EPG"); const int v = e ? atoi(e) : 0; return v; }();
    int epg = args.fused ? SCAN_GROUP_FUSED : SCAN_GROUP;
    if (args.fused && (tail_epg == 8 || tail_epg == 4)) epg = tail_epg;   // (A/B: smaller units for the tail kernel's counter)
    const int want = args.fused ? cap * LGK_WARPS_PER_CTA * 4 / 5 : cap * LGK_WARPS_PER_CTA;
    while (epg > 1 && (c->num_envs + epg - 1) / epg < want) epg >>= 1;
    args.scan_epc = epg;
    args.scan_tiles = (c->num_envs + epg - 1) / epg;
    // (fused: a CTA takes the four groups of a unit, + the id-list CTA)
    const int ctas = (args.scan_tiles + LGK_WARPS_PER_CTA - 1) / LGK_WARPS_PER_CTA + (args.fused ? 1 : 0);
    args.scan_grid = ctas < cap ? ctas : cap;
  }
  {
    // staged reward rows: every group full (16 envs) and every row segment 16-byte aligned.  LGK_TAIL_STAGE=0: off
    static const bool stage_on = [] { const char* e = getenv("LGK_TAIL_STAGE"); return !(e && e[0] == '0'); }();
    bool ok = stage_on && args.fused && (args.scan_epc == 16 || args.scan_epc == 8 || args.scan_epc == 4) && c->num_envs % args.scan_epc == 0 &&
              args.num_rows > 0 && al16(b->episode_sums) && al16(args.rv);
    for (int r = 0; r < args.num_rows && ok; ++r) ok = al16(args.row_src[r]);
    args.tail_staged = ok ? 1 : 0;
  }
  const bool taps = c->rng_mode == LGK_RNG_TAPS;
  const bool mul = c->index_math == LGK_INDEX_MUL_RCP;
#define LGK_DISPATCH(DT)                                                             \
  (taps ? (mul ? launch_step<DT, true, true>(args, dev, grid, smem, stream)          \
               : launch_step<DT, true, false>(args, dev, grid, smem, stream))        \
        : (mul ? launch_step<DT, false, true>(args, dev, grid, smem, stream)         \
               : launch_step<DT, false, false>(args, dev, grid, smem, stream)))
  if (D == 12) return LGK_DISPATCH(12);
  return LGK_DISPATCH(0);
#undef LGK_DISPATCH
}

// Fused post_physics_step kernel (sm_100a): one CTA of four warps per tile of 32 envs.
//
// One launch replaces everything LeggedRobot.post_physics_step() does per env
// (reference legged_robot.py:109-141): episode counters, body-frame velocities / projected
// gravity, command resampling + heading command, the yaw-rotated terrain height scan, pushes,
// terminations, every active _reward_* term with in-place episode sums, resets (terrain
// curriculum, dof / root state, commands), the noisy scaled observation row, last_* copies,
// ascending compacted reset / resample id lists and the episode statistics behind extras.
//
// Work decomposition
//   grid    CTA c owns the contiguous env range [c*per, (c+1)*per) and walks it in sub-tiles of
//           <= 32 envs.  per = 32 by default (six CTAs resident per SM, the hardware scheduler
//           backfills); LgkConsts.envs_per_cta selects longer, persistent ranges.
//   load    every per-env input row of a sub-tile is contiguous in HBM ([N,13], [N,D,2], [N,B,3], ...):
//           the lanes of warp 0 issue 1-D bulk async copies (cp.async.bulk -> SASS UBLKCP, the TMA
//           engine) into shared memory, completion counted on an mbarrier.  No register staging,
//           full coalescing, ~22 KB in flight per CTA.  The [R,N] episode-sum rows (128 B per row
//           and tile) come in through 4-byte cp.async (LDGSTS) issued by W3.
//   stage 1 the four warps run four different per-env jobs with lane = env (no divergence inside a
//           warp): W0 base state / commands / termination / reset of the root, W1 + W2 the joints,
//           W3 contacts + feet air time.  The reset flag is cheap and is recomputed by each warp.
//   stage 2 the envs of the sub-tile are dealt to the warps; lanes cover the observation columns of
//           one env at a time: noise on the head, then the H scan points in passes of 32 (index
//           chain on packed f32x2 ops, one int16 gather per point from the precomputed
//           min-of-three-neighbours table, all gathers of an env in flight together).
//   stage 3 W0 alone, while W1..W3 already wait for the next sub-tile's loads: term values by env,
//           then lane = episode-sum row (scale, accumulate, statistics: independent lanes), then
//           lane = env again for the ordered reward sum and the coalesced stores.
//   store   observation rows and episode sums are written straight from registers (128-byte
//           coalesced); last_actions and last_dof_vel leave through bulk async stores.
//   The CTA that publishes the last reset / resample masks compacts them into ascending id lists;
//   the last W0 to finish finalises the episode statistics.
//
// Code size matters here (the four roles + two stages exceed the instruction cache when everything
// is inlined): paths that run once per launch or only for ragged / unaligned tiles are __noinline__.
#include <cmath>
#include <limits>
#include <type_traits>
#include "lgk_common.cuh"

#define TILE 32
#define HP_OFF (((RS * (LGK_REW_COUNT + 1)) + 3) & ~3)   // base-height partial sums, after the reward table
#define RS (TILE + 1)   // row stride of the [row][env] reward tables: conflict-free by row and by env
#ifndef LGK_STAGE2_ENVS
#define LGK_STAGE2_ENVS 1   // envs a warp keeps in flight in stage 2 (2 measured slower: stage 2 is issue-bound at full residency)
#endif
#ifndef LGK_MIN_CTAS
#define LGK_MIN_CTAS 6   // register cap 85: six 128-thread CTAs per SM
#endif

#ifdef LGK_PROFILE
// development build only: per-CTA phase timestamps (clock64) of the first sub-tile, [cta][warp][8]
__device__ long long lgk_prof[4096 * 4 * 16];
#ifndef LGK_PROF_SUB
#define LGK_PROF_SUB 0
#endif
#define LGK_STAMP(k) do { if (sub == LGK_PROF_SUB && lane == 0 && blockIdx.x < 4096) lgk_prof[(blockIdx.x * 4 + wic) * 16 + (k)] = clock64(); } while (0)
extern "C" int lgk_debug_read_prof(long long* out, int n) {
  return (int)cudaMemcpyFromSymbol(out, lgk_prof, sizeof(long long) * n);
}
// CTA timeline: [cta][globaltimer at start, globaltimer when W0 leaves, SM id, 0]
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

struct TileLayout {   // offsets (in floats, multiples of 4) into dynamic shared memory
  int root, dof, act, ldv, cmd, fat, ep, lvl, sums, head, scr, cf, tq, lact, total;
};

#define LGK_LD_ARRAYS 11
struct StepArgs {
  LgkConsts c;
  LgkBuffers b;
  TileLayout L;
  int per;           // envs per CTA (multiple of 4) of the first n_big CTAs, walked in sub-tiles of <= 32
  int n_big;         // CTAs past n_big own per_small envs each: the launch's last, partial wave runs on
  int per_small;     //   smaller tiles so that it is shorter than a full wave (per_small <= 32)
  int max_sub;       // sub-tiles per CTA = ceil(per / 32)
  int head_stride;   // odd row stride of the staged observation head
  int air_active;    // feet_air_time term active (:945-948 only run then)
  int bh_active;     // base_height term active
  int tma_ok;        // every tiled array 16-byte aligned and N % 4 == 0
  int num_rows;      // episode_sums rows: num_rewards (+1 with termination)
  unsigned pk[20];   // Philox4x32-10 round keys of c.seed
  // bulk-load table: lane i of W0 copies input array i of a sub-tile (source base, offset into shared
  // memory in floats, bytes per env); src == nullptr: nothing to copy
  const char* ld_src[LGK_LD_ARRAYS];
  unsigned ld_off[LGK_LD_ARRAYS], ld_bpe[LGK_LD_ARRAYS];
  unsigned ld_bytes_per_env;   // sum of ld_bpe
  // norm > c without the square root: sqrt is monotone and correctly rounded, so sqrtf(s) > c is the same
  // predicate as s > thr(c), thr(c) = the largest float whose square root rounds to <= c (found on the host)
  float thr_term, thr_pen, thr_ffp;   // c = 1 (:146), 0.1 (:897), max_contact_force (:962)
  unsigned hip_mask;                  // bit d set: joint d is a hip joint (_reward_hip_motion)
};

// scratch fields (SoA, [field][TILE]) exchanged between the stages
enum { SC_ES = 0,            // 12 floats per env (AoS), the scan's per-env operands as ready-made f32x2 pairs:
                             //   (-2qz, 2qz) (2qz, -2qz) | (qw, qw) (-qz, qz) | (rx, ry) rz zo      (yaw-only unit quaternion)
       ES_W = 12,
       SC_J1 = 12, SC_J2 = 22,  // joint partial sums of W1 / W2 (10 each)
       SC_COLL = 32, SC_FFP, SC_RAIR, SC_STUMBLE, SC_NOFLY, SC_FIELDS };

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

// one joint: reward partial sums on the pre-reset state, reset values, observation head columns
__device__ __forceinline__ void joint_update(const LgkConsts& c, int d, int D, unsigned hip_mask, bool reset,
                                             float inv_dt, float av, float lav, float tqv, float ldvv, float& p,
                                             float& v, float u, DofAcc& acc, float* __restrict__ head_row) {
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
  head_row[12 + d] = (p - c.default_dof_pos[d]) * c.obs_scale_dof_pos;
  head_row[12 + D + d] = v * c.obs_scale_dof_vel;
  head_row[12 + 2 * D + d] = av;
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
  float z;         // new root z (-> height observation offset)
  int level;       // terrain level after the curriculum update
};
template <bool TAPS>
static __device__ __noinline__ void reset_root(const StepArgs& a, unsigned long long csc, int env, unsigned env_global,
                                               float rx, float ry, long long level, ResetOut& o) {
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
  float* rw = b.root_states + (size_t)env * 13;
#pragma unroll
  for (int i = 0; i < 13; ++i) rw[i] = root[i];
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
  const int D = c.num_dof, B3 = c.num_bodies * 3, F = c.num_feet, N = c.num_envs;
  const int tid = threadIdx.x;
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

// ---------------------------------------------------------------------------------------------
// Step epilogue: the two once-per-step serial jobs, as a second small launch behind the tile kernel
// (one CTA of 1024 threads).  Being a separate launch, the kernel boundary orders it after every tile's
// masks and statistics: the tile CTAs need no tickets, fences or polling, and the job runs on an idle GPU
// instead of on one SM that is busy with tiles (there it took 16-20 us, most of it instruction fetch).
//   1. ordered compaction of the per-tile reset / resample bitmasks into ascending id lists (:150, :342)
//   2. episode statistics behind extras["episode"] (:182-190), step counter += 1 (:119)
#define EPI_THREADS 1024
__global__ void __launch_bounds__(EPI_THREADS, 1) lgk_step_epilogue_kernel(const __grid_constant__ StepArgs a, int n_ctas) {
  __shared__ int s_scan[2][EPI_THREADS / 32 + 1];
  const LgkConsts& c = a.c;
  const LgkBuffers& b = a.b;
  const int N = c.num_envs;
  const int tid = threadIdx.x, lane = tid & 31, wic = tid >> 5;
  LgkWorkspace* ws = reinterpret_cast<LgkWorkspace*>(b.workspace);
  pdl_wait();   // scheduled while the tile kernel drains; from here on every tile's writes are visible
  const unsigned long long csc = ws->step_counter + 1ull;
  const int parity = (int)(csc & 1ull);
  // ---- id lists --------------------------------------------------------------------------------
  const int nt = n_ctas * a.max_sub;      // mask slots, (cta, sub) order = ascending env order
  const unsigned* mask_base = reinterpret_cast<const unsigned*>(reinterpret_cast<const char*>(ws) + LGK_WS_HEADER_BYTES);
  const unsigned* rm = mask_base;
  const unsigned* sv = mask_base + lgk_max_warps(N);
  // slots of sub-tiles past a CTA's range were never written this launch: skip them
  const bool one_sub = a.max_sub == 1;
  auto slot_env0 = [&](int t) {
    if (one_sub) return t < a.n_big ? t * a.per : a.n_big * a.per + (t - a.n_big) * a.per_small;
    const int cta = t / a.max_sub, k = t - cta * a.max_sub;
    return cta * a.per + k * TILE;
  };
  auto slot_ok = [&](int t) {
    if (one_sub) return true;
    const int cta = t / a.max_sub, k = t - cta * a.max_sub;
    return k * TILE < min(a.per, N - cta * a.per);
  };
  const int per = (nt + EPI_THREADS - 1) / EPI_THREADS;   // 2 at 65536 envs
  const int t0 = min(nt, tid * per), t1 = min(nt, t0 + per);
  int c0 = 0, c1 = 0;
#pragma unroll 4
  for (int t = t0; t < t1; ++t)
    if (slot_ok(t)) { c0 += __popc(rm[t]); c1 += __popc(sv[t]); }
  // block-wide exclusive scan of (c0, c1): shuffles inside a warp, warp totals through shared memory
  int i0 = c0, i1 = c1;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int y0 = __shfl_up_sync(LGK_FULL_MASK, i0, o), y1 = __shfl_up_sync(LGK_FULL_MASK, i1, o);
    if (lane >= o) { i0 += y0; i1 += y1; }
  }
  if (lane == 31) { s_scan[0][wic] = i0; s_scan[1][wic] = i1; }
  __syncthreads();
  if (wic == 0) {   // exclusive scan of the 32 warp totals
    int w0 = s_scan[0][lane], w1 = s_scan[1][lane];
    const int t0w = w0, t1w = w1;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y0 = __shfl_up_sync(LGK_FULL_MASK, w0, o), y1 = __shfl_up_sync(LGK_FULL_MASK, w1, o);
      if (lane >= o) { w0 += y0; w1 += y1; }
    }
    s_scan[0][lane] = w0 - t0w;
    s_scan[1][lane] = w1 - t1w;
    if (lane == 31) { s_scan[0][32] = w0; s_scan[1][32] = w1; }
  }
  __syncthreads();
  int o0 = s_scan[0][wic] + i0 - c0, o1 = s_scan[1][wic] + i1 - c1;
  if (c0 + c1 != 0) {
#pragma unroll 1
    for (int t = t0; t < t1; ++t) {
      if (!slot_ok(t)) continue;
      unsigned m = rm[t], m2 = sv[t];
      const int e0 = slot_env0(t);
      while (m) { const int bit = __ffs(m) - 1; m &= m - 1; b.reset_ids[o0++] = e0 + bit; }
      while (m2) { const int bit = __ffs(m2) - 1; m2 &= m2 - 1; b.resample_ids[o1++] = e0 + bit; }
    }
  }
  if (tid == 0) {
    b.counts[0] = s_scan[0][32];
    b.counts[1] = s_scan[1][32];
  }
  if (wic != 1) return;
  // ---- episode statistics (one warp) ------------------------------------------------------------
  const int num_rows = a.num_rows;
  const int n_reset = ws->reset_count[parity];
  const int level_sum = ws->level_sum[parity];
  if (lane < num_rows) {
    const float s = ws->stats[parity][lane];
    b.episode_stats[lane] = s;
    b.episode_stats[num_rows + 4 + lane] += s;   // running sums since the host last cleared them
    if (n_reset > 0)   // extras["episode"]["rew_<k>"] = mean(sum[ids]) / max_episode_length_s (:184)
      b.extras[lane] = (s / (float)n_reset) / c.max_episode_length_s;
    ws->stats[parity ^ 1][lane] = 0.f;   // next step's accumulators
  }
  if (lane == 0) {
    b.episode_stats[num_rows] = (float)n_reset;
    b.episode_stats[num_rows + 1] = (float)level_sum;
    b.episode_stats[num_rows + 2] = (float)N;
    b.episode_stats[2 * num_rows + 4] += (float)n_reset;
    b.episode_stats[2 * num_rows + 5] = (float)level_sum;
    b.episode_stats[2 * num_rows + 6] = (float)N;
    if (n_reset > 0 && (c.flags & LGK_F_TERRAIN_CURRICULUM))
      b.extras[num_rows] = (float)level_sum / (float)N;   // mean(terrain_levels.float()) (:188)
    ws->level_sum[parity ^ 1] = 0;
    ws->reset_count[parity ^ 1] = 0;
    ws->step_counter = csc;
  }
}

template <int DT, bool TAPS, bool MUL_RCP, int NP>
__global__ void __launch_bounds__(LGK_CTA_THREADS, LGK_MIN_CTAS)
lgk_step_kernel(const __grid_constant__ StepArgs a) {
  pdl_launch_dependents();   // the epilogue may be scheduled as soon as the last tile CTA has started
  pdl_wait();                // launched behind clip_actions_poses: its writes (actions) are visible from here
  const LgkConsts& c = a.c;
  const LgkBuffers& b = a.b;
  const TileLayout& L = a.L;
  constexpr bool VEC = (DT > 0) && (DT % 4 == 0);
  const int D = DT > 0 ? DT : c.num_dof;
  const int N = c.num_envs;
  const int B3 = c.num_bodies * 3;
  const int F = c.num_feet;
  const int H = c.num_height_points;
  const int O = c.num_obs;
  const int OH = 12 + 3 * D;               // width of the observation head (:217-224)
  const int HS = a.head_stride;
  const unsigned flags = c.flags;
  const bool heading_mode = flags & LGK_F_HEADING_COMMAND;
  const int num_rows = a.num_rows;

  extern __shared__ __align__(16) float sm[];
  __shared__ __align__(8) uint64_t s_bar;

  // roles go by the logical warp index wic: rotated per CTA so that the W0s of the CTAs resident on an
  // SM (the longest role, and the only one running in stage 3) do not all sit on the same scheduler
  const int lane = threadIdx.x & 31;
  const int wic = ((int)(threadIdx.x >> 5) + (int)((blockIdx.x * 2654435761u) >> 30)) & (LGK_WARPS_PER_CTA - 1);
  const int tid = wic * 32 + lane;
  const bool big = (int)blockIdx.x < a.n_big;
  const int cta_env0 = big ? blockIdx.x * a.per : a.n_big * a.per + ((int)blockIdx.x - a.n_big) * a.per_small;
  const int cta_ne = min(big ? a.per : a.per_small, N - cta_env0);
  const int nsub = (cta_ne + TILE - 1) / TILE;
  const bool has_levels = b.terrain_levels != nullptr;

  // ---- first thing: W0 sets the mbarrier up and gets the bulk loads of the first sub-tile going; the
  // rest of the prologue runs under their latency.  Lane i copies input array i.
  const char* ld_src = nullptr;
  unsigned ld_off = 0, ld_bpe = 0;
  if (wic == 0) {
    if (lane < LGK_LD_ARRAYS) { ld_src = a.ld_src[lane]; ld_off = a.ld_off[lane]; ld_bpe = a.ld_bpe[lane]; }
    if (lane == 0) mbar_init(&s_bar, 1);
    __syncwarp();
    const int ne0 = min(TILE, cta_ne);
    if (a.tma_ok && (ne0 & 3) == 0) {
      if (lane == 0) mbar_expect_tx(&s_bar, (unsigned)ne0 * a.ld_bytes_per_env);
      __syncwarp();
      if (ld_src != nullptr) bulk_g2s(sm + ld_off, ld_src + (size_t)cta_env0 * ld_bpe, (unsigned)ne0 * ld_bpe, &s_bar);
    }
  }
  // stage 3, pass B: lane l of W0 owns episode-sum row l (term, scale, index among the external terms)
  __shared__ int s_row_term[32], s_row_ext[32], s_idx[32];
  __shared__ float s_row_scale[32];
  if (wic == 0) {
    const bool in = lane < c.num_rewards;
    const int term = in ? c.rewards[lane].term : LGK_REW_COUNT;   // LGK_REW_COUNT: the termination row
    const unsigned xm = __ballot_sync(LGK_FULL_MASK, in && term == LGK_REW_EXTERNAL);
    s_idx[lane] = 3 * (lane < IDX_PEN ? c.feet_idx[lane] : lane < IDX_TERM ? c.pen_idx[lane - IDX_PEN] : c.term_idx[lane - IDX_TERM]);
    s_row_term[lane] = term;
    s_row_ext[lane] = __popc(xm & ((1u << lane) - 1u));
    s_row_scale[lane] = in ? c.rewards[lane].scale : c.termination_scale;
  }
  if (tid == 0) LGK_CTA_STAMP(0);
#ifdef LGK_STAGGER_NS   // experiment: de-phase the CTAs that start together on an SM
  if (blockIdx.x < 888) {
    const unsigned k = min(blockIdx.x / 148u, 5u);
    for (unsigned i = 0; i < k; ++i) __nanosleep(LGK_STAGGER_NS);
  }
#endif
  LgkWorkspace* ws = reinterpret_cast<LgkWorkspace*>(b.workspace);
  const unsigned long long csc = ws->step_counter + 1ull;   // common_step_counter after += 1 (:119)
  const int parity = (int)(csc & 1ull);
  bool push_step = false;   // W0 only; 32-bit remainder unless the counter has outgrown it
  if (wic == 0 && (flags & LGK_F_PUSH_ROBOTS))
    push_step = (csc >> 32) == 0ull ? ((unsigned)csc % (unsigned)c.push_interval == 0u)
                                    : (csc % (unsigned long long)c.push_interval == 0ull);

  __syncthreads();          // barrier initialised before anyone polls it, row table written

  for (int sub = 0; sub < nsub; ++sub) {
    const int env0 = cta_env0 + sub * TILE;
    const int ne = min(TILE, cta_ne - sub * TILE);
    const bool full = a.tma_ok && ((ne & 3) == 0);   // bulk copies need 16-byte multiples: 4 envs x 4 B
    const int slot = blockIdx.x * a.max_sub + sub;   // where this sub-tile's reset / resample masks go
    const bool last_sub = sub == nsub - 1;

    LGK_STAMP(0);
    // ---- load the sub-tile --------------------------------------------------------------------
    if (full) {
      if (sub > 0) {            // (the first sub-tile's loads were issued at the top of the kernel)
        __syncthreads();        // every warp is through stage 3 of the previous sub-tile
        if (wic == 0) {
          if (lane == 0) {
            fence_proxy_async_smem();   // generic reads of the previous sub-tile before the async writes
            mbar_expect_tx(&s_bar, (unsigned)ne * a.ld_bytes_per_env);
          }
          __syncwarp();
          if (ld_src != nullptr) bulk_g2s(sm + ld_off, ld_src + (size_t)env0 * ld_bpe, (unsigned)ne * ld_bpe, &s_bar);
        }
      }
      mbar_wait(&s_bar, sub & 1);
    } else {
      __syncthreads();          // every warp is through stage 3 of the previous sub-tile
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
    const long long ep_in = reinterpret_cast<const long long*>(sm + L.ep)[lane];
    const long long ep_inc = (flags & LGK_F_EP_INCREMENTED) ? 0 : 1;   // lgk_pre_reward_state already counted this step

    // the reset decision (:143-148) is cheap: every stage-1 warp forms it for its own lane
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

    // values W0 keeps in registers from stage 1 to stage 3
    float blv[3] = {0, 0, 0}, bav[3] = {0, 0, 0}, grav[3] = {0, 0, 0};
    float4 cmd_rew = make_float4(0, 0, 0, 0);
    const unsigned reset_mask = __ballot_sync(LGK_FULL_MASK, active && reset);

    LGK_STAMP(12);
    // =========================================== stage 1 =========================================
    if (wic == 0) {
      // ---- W0: base state, commands, termination, root reset, per-env outputs --------------------
      bool resample_cb = false;
      int level_contrib = 0;
      if (active) {
        float root[13];
        const float* rp = sm + L.root + lane * 13;
#pragma unroll
        for (int i = 0; i < 13; ++i) root[i] = rp[i];
        long long ep = ep_in + ep_inc;                              // (:118)
        const float q[4] = {root[3], root[4], root[5], root[6]};
        quat_rotate_inverse(q, root[7], root[8], root[9], blv);     // (:123-125)
        quat_rotate_inverse(q, root[10], root[11], root[12], bav);
        quat_rotate_inverse(q, 0.f, 0.f, -1.f, grav);
        float4 cmd = *reinterpret_cast<const float4*>(sm + L.cmd + lane * 4);

        // ---- _post_physics_step_callback (:337-352) -------------------------------------------
        resample_cb = (unsigned long long)ep < 0x80000000ull ? ((int)ep % c.resample_interval) == 0
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
        cmd_rew = cmd;
        // yaw-only quaternion of quat_apply_yaw (utils/math.py:38-42): normalize([0,0,qz,qw]);
        // x.norm(p=2): separately rounded squares, correctly rounded sqrt (same on torch CPU / CUDA)
        float yaw_z, yaw_w;
        {
          const float n = fmaxf(sqrtf(add_rn(mul_rn(q[2], q[2]), mul_rn(q[3], q[3]))), 1e-9f);
          yaw_z = __fdiv_rn(q[2], n);
          yaw_w = __fdiv_rn(q[3], n);
        }
        float z_obs = root[2];
        float rvel[6] = {root[7], root[8], root[9], root[10], root[11], root[12]};
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
          reset_root<TAPS>(a, csc, env, env_global, root[0], root[1], level, ro);
          cmd = ro.cmd;
#pragma unroll
          for (int i = 0; i < 6; ++i) rvel[i] = ro.rvel[i];
          z_obs = ro.z;
          level = ro.level;
          ep = 0;
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
          float* p = b.base_lin_vel + (size_t)env * 3; p[0] = blv[0]; p[1] = blv[1]; p[2] = blv[2];
          p = b.base_ang_vel + (size_t)env * 3;        p[0] = bav[0]; p[1] = bav[1]; p[2] = bav[2];
          p = b.projected_gravity + (size_t)env * 3;   p[0] = grav[0]; p[1] = grav[1]; p[2] = grav[2];
          float2* lr = reinterpret_cast<float2*>(b.last_root_vel + (size_t)env * 6);   // (:138)
          lr[0] = make_float2(rvel[0], rvel[1]); lr[1] = make_float2(rvel[2], rvel[3]); lr[2] = make_float2(rvel[4], rvel[5]);
        }
        // obs head columns 0..11 (:217-220), post-reset commands
        head_row[0] = blv[0] * c.obs_scale_lin_vel; head_row[1] = blv[1] * c.obs_scale_lin_vel; head_row[2] = blv[2] * c.obs_scale_lin_vel;
        head_row[3] = bav[0] * c.obs_scale_ang_vel; head_row[4] = bav[1] * c.obs_scale_ang_vel; head_row[5] = bav[2] * c.obs_scale_ang_vel;
        head_row[6] = grav[0]; head_row[7] = grav[1]; head_row[8] = grav[2];
        head_row[9] = cmd.x * c.commands_scale[0]; head_row[10] = cmd.y * c.commands_scale[1]; head_row[11] = cmd.z * c.commands_scale[2];
        // scalars of the height scan: pre-reset position / yaw, post-reset z for the observation (:227)
        float4* es = reinterpret_cast<float4*>(scr + SC_ES * TILE + lane * ES_W);
        const float yz2 = mul_rn(yaw_z, 2.0f);
        es[0] = make_float4(-yz2, yz2, yz2, -yz2);
        es[1] = make_float4(yaw_w, yaw_w, -yaw_z, yaw_z);
        es[2] = make_float4(root[0], root[1], root[2], sub_rn(z_obs, 0.5f));
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
        const int nr = __popc(reset_mask);
        if (nr != 0) atomicAdd(&ws->reset_count[parity], nr);
      }
    }
    else if (wic == 1 || wic == 2) {
      // ---- W1 / W2: joints (:874-916, :956-966 sums; :399-408 reset; obs head; last_dof_vel) -----
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
          constexpr int NC = DT / 4, NC1 = (NC + 1) / 2;
          const int cb = (wic == 1) ? 0 : NC1, ce = (wic == 1) ? NC1 : NC;
#pragma unroll 1
          for (int ch = cb; ch < ce; ++ch) {
            const int d0 = ch * 4;
            const float4 a4 = *reinterpret_cast<const float4*>(act_p + d0);
            const float4 la4 = *reinterpret_cast<const float4*>(lact_p + d0);
            const float4 tq4 = *reinterpret_cast<const float4*>(tq_p + d0);
            const float4 ldv4 = *reinterpret_cast<const float4*>(ldv_p + d0);
            float4 s0 = *reinterpret_cast<const float4*>(dof_p + 2 * d0);       // p0 v0 p1 v1
            float4 s1 = *reinterpret_cast<const float4*>(dof_p + 2 * d0 + 4);   // p2 v2 p3 v3
            float u[4] = {0, 0, 0, 0};
            if (reset) {
              if (TAPS) {
                const float* t = b.tap_dof + (size_t)env * D + d0;
                u[0] = t[0]; u[1] = t[1]; u[2] = t[2]; u[3] = t[3];
              } else {
                const uint4 r = philox_cold(a, csc, env_global, LGK_SITE_DOF + ch);
                u[0] = u01(r.x); u[1] = u01(r.y); u[2] = u01(r.z); u[3] = u01(r.w);
              }
            }
            joint_update(c, d0 + 0, D, hip_mask, reset, inv_dt, a4.x, la4.x, tq4.x, ldv4.x, s0.x, s0.y, u[0], acc, head_row);
            joint_update(c, d0 + 1, D, hip_mask, reset, inv_dt, a4.y, la4.y, tq4.y, ldv4.y, s0.z, s0.w, u[1], acc, head_row);
            joint_update(c, d0 + 2, D, hip_mask, reset, inv_dt, a4.z, la4.z, tq4.z, ldv4.z, s1.x, s1.y, u[2], acc, head_row);
            joint_update(c, d0 + 3, D, hip_mask, reset, inv_dt, a4.w, la4.w, tq4.w, ldv4.w, s1.z, s1.w, u[3], acc, head_row);
            *reinterpret_cast<float4*>(ldv_p + d0) = make_float4(s0.y, s0.w, s1.y, s1.w);   // last_dof_vel = dof_vel (:137)
            if (reset) {
              *reinterpret_cast<float4*>(dof_g + 2 * d0) = s0;
              *reinterpret_cast<float4*>(dof_g + 2 * d0 + 4) = s1;
            }
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
                const unsigned rr[4] = {r.x, r.y, r.z, r.w};
                u = u01(rr[d & 3]);
              }
            }
            joint_update(c, d, D, hip_mask, reset, inv_dt, av, lav, tqv, ldvv, p, v, u, acc, head_row);
            ldv_p[d] = v;
            if (reset) { dof_g[2 * d] = p; dof_g[2 * d + 1] = v; }
          }
        }
      }
      float* jp = scr + ((wic == 1) ? SC_J1 : SC_J2) * TILE + lane;
#pragma unroll
      for (int i = 0; i < 10; ++i) jp[i * TILE] = acc.v[i];
    } else {
      // ---- W3: contacts (:895-897, :951-962, cassie.py:43-46) and feet air time (:941-949) -------
      float collisions = 0.f, ffp = 0.f, r_air = 0.f;
      bool stumble = false;
      int nofly = 0;
      if (active) {
        // four bodies per trip: the index loads (constant bank, dynamic offset), the row reads and the
        // square roots of a trip are independent, so their latencies overlap instead of adding up
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
        LGK_STAMP(10);
        const float* fa_in = sm + L.fat + lane * F;
        float* fa_out = b.feet_air_time + (size_t)env * F;
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
            const bool first = (t > 0.f) && contact;
            t = add_rn(t, c.dt);
            r_air += first ? (t - 0.5f) : 0.f;
            if (contact) t = 0.f;
          }
          if (a.air_active || reset) fa_out[k] = reset ? 0.f : t;   // zeroed on reset (:178)
        }
      }
      LGK_STAMP(11);
      scr[SC_COLL * TILE + lane] = collisions;
      scr[SC_FFP * TILE + lane] = ffp;
      scr[SC_RAIR * TILE + lane] = r_air;       // masked by |cmd_xy| > 0.1 in stage 3
      scr[SC_STUMBLE * TILE + lane] = stumble ? 1.f : 0.f;
      scr[SC_NOFLY * TILE + lane] = (nofly == 1) ? 1.f : 0.f;
    }
    LGK_STAMP(2);
    fence_proxy_async_smem();   // last_dof_vel rows (W1 / W2) -> visible to the bulk store
    __syncthreads();
    LGK_STAMP(3);

    // ---- outputs that are complete after stage 1; id-list compaction by the last mask publisher ---
    if (full) {
      if (tid == 32) {
        bulk_s2g(b.last_actions + (size_t)env0 * D, sm + L.act, ne * 4 * D);   // last_actions = actions (:136)
        bulk_s2g(b.last_dof_vel + (size_t)env0 * D, sm + L.ldv, ne * 4 * D);
        bulk_commit();
      }
    } else if (wic > 0) {
      tile_store_fallback(a, sm, env0, ne, tid);
    }
    if (wic == 3 && lane < ne) {
      // episode-sum rows ([R, N], 128 B per row and tile): 4-byte cp.async (LDGSTS) per lane and row into
      // rows the joints no longer need; in flight during stage 2, awaited before the stage-3 barrier
      const float* src = b.episode_sums + env0 + lane;
      const unsigned dst = smem_u32(sm + L.sums + lane);
#pragma unroll 1
      for (int r = 0; r < num_rows; ++r)
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst + (unsigned)r * (RS * 4)),
                     "l"(src + (size_t)r * N) : "memory");
      asm volatile("cp.async.commit_group;" ::: "memory");
    }

  // =========================================== stage 2 ===========================================
  // one env at a time per warp, lanes over observation columns: noise on the head, then the H scan
  // points in NP passes of 32.  All gathers of an env are issued before any is consumed.  The index
  // chain runs on packed f32x2 instructions (x and y of a point share every operation; each packed
  // op rounds its halves like the scalar op, so the result is bit-identical to the op-by-op chain).
  {
    constexpr int NPA = NP > 0 ? NP : 1;
    const bool add_noise = flags & LGK_F_ADD_NOISE;
    const bool plane = flags & LGK_F_TERRAIN_PLANE;
    const float inv_hs = 1.0f / c.horizontal_scale;
    const int max_px = c.terrain_rows - 2, max_py = c.terrain_cols - 2;
    const float clip_obs = c.clip_obs;
    // this lane's scan offsets and noise amplitudes (same for every env); lanes past H repeat point 0
    float2 pxy_l[NPA];
    float ns_h[NPA];
#pragma unroll
    for (int it = 0; it < NP; ++it) {
      const int h = it * 32 + lane;
      const bool ok = h < H;
      pxy_l[it].x = ok ? __ldg(b.height_points_xy + 2 * h) : 0.f;
      pxy_l[it].y = ok ? __ldg(b.height_points_xy + 2 * h + 1) : 0.f;
      ns_h[it] = (ok && add_noise) ? __ldg(b.noise_scale_vec + OH + h) : 0.f;
    }
    const float ns_a = (add_noise && lane < OH) ? __ldg(b.noise_scale_vec + lane) : 0.f;
    const float ns_b = (add_noise && lane + 32 < OH) ? __ldg(b.noise_scale_vec + lane + 32) : 0.f;
    // obs = clip(v + (2*rand - 1) * noise_scale_vec, +-clip_observations)   (:231, :104)
    auto noisy = [&](float v, const float* __restrict__ trow, int col, unsigned rw, int half, float ns, bool clip) {
      if (TAPS) {
        if (add_noise) v = add_rn(v, mul_rn(sub_rn(mul_rn(2.0f, __ldg(trow + col)), 1.0f), ns));
      } else {
        // 2^23 + r16 as a float (one PRMT, no I2F on the quarter-rate conversion unit), then
        // (2^23 + r16) * 2^-15 - 257 = r16 / 2^15 - 1, exact: uniform on [-1, 1) in steps of 2^-15; ns = 0 without noise
        const unsigned m = half ? __byte_perm(rw, 0x4b000000u, 0x7632) : __byte_perm(rw, 0x4b000000u, 0x7610);
        v = fmaf(fmaf(__uint_as_float(m), 3.0517578125e-05f, -257.0f), ns, v);
      }
      return clip ? clampf(v, -clip_obs, clip_obs) : v;
    };
    // the scan columns are bounded by obs_scale_height * 1 + noise: when that is inside clip_observations
    // for every column (warp-uniform), their clip is the identity and is skipped
    bool clip_h = false;
#pragma unroll
    for (int it = 0; it < NP; ++it) clip_h |= !(fabsf(c.obs_scale_height) + fabsf(ns_h[it]) <= clip_obs);
    clip_h = __any_sync(LGK_FULL_MASK, clip_h);
    float* hpart = sm + L.cf + HP_OFF;   // [env][RS]; torques / last_actions rows are dead after stage 1
    const float2 border2 = make_float2(c.border_size, c.border_size);
    const float2 inv2 = make_float2(inv_hs, inv_hs);
    // U envs in flight per warp (e, e + 4, ...): a single env's instruction stream is a chain of
    // dependent latencies (index chain -> gather -> scale / noise -> store), so the warp interleaves U of them
    constexpr int U = LGK_STAGE2_ENVS;
    for (int e0 = wic; e0 < ne; e0 += LGK_WARPS_PER_CTA * U) {
      int eu[U];
      bool live[U];
      float zo[U];
      int raw[U][NPA];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        eu[u] = e0 + u * LGK_WARPS_PER_CTA;
        live[u] = eu[u] < ne;
        if (!live[u]) eu[u] = e0;   // a dead slot recomputes the first env; its stores are predicated off
      }
      // ---- phase A: height-field indices of every pass, gathers in flight --------------------------
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const float4* es = reinterpret_cast<const float4*>(scr + SC_ES * TILE + eu[u] * ES_W);
        const float4 esa = es[0], esb = es[1], esc = es[2];
        zo[u] = esc.w;
        if (NP > 0 && !plane) {
          // quat_apply(normalize([0,0,qz,qw]), (px,py,0)) + root_pos, op for op (no contraction):
          //   t = 2 * cross(xyz, p) = (-2qz*py, 2qz*px); r = p + w*t + cross(xyz, t) = p + qw*t + (-qz*ty, qz*tx)
          const float2 k_t = make_float2(esa.x, esa.y), k_ts = make_float2(esa.z, esa.w);
          const float2 qw2 = make_float2(esb.x, esb.y), k_c = make_float2(esb.z, esb.w), r2 = make_float2(esc.x, esc.y);
#pragma unroll
          for (int it = 0; it < NP; ++it) {
            const float2 p = pxy_l[it], pyx = make_float2(p.y, p.x);
            const float2 t = __fmul2_rn(k_t, pyx);     // (tx, ty)
            const float2 ts = __fmul2_rn(k_ts, p);     // (ty, tx)
            float2 w = __fadd2_rn(p, __fmul2_rn(qw2, t));
            w = __fadd2_rn(w, __fmul2_rn(k_c, ts));
            w = __fadd2_rn(__fadd2_rn(w, r2), border2);
            // (points / horizontal_scale).long(): truncation toward zero (:842)
            float2 sxy;
            if (MUL_RCP) sxy = __fmul2_rn(w, inv2);
            else sxy = make_float2(__fdiv_rn(w.x, c.horizontal_scale), __fdiv_rn(w.y, c.horizontal_scale));
#ifdef LGK_F2I_ON_XU
            int ix = __float2int_rz(sxy.x), iy = __float2int_rz(sxy.y);
#else
            // trunc toward zero without the quarter-rate conversion unit: 2^23 + s, rounded toward zero, holds
            // trunc(s) in its mantissa for 0 <= s < 2^23; s < 0 lands below 2^23 (negative after the subtraction,
            // clipped to 0 like trunc(s) would be), s >= 2^23 / NaN above the clip's upper bound
            const float2 mg = __fadd2_rz(sxy, make_float2(8388608.0f, 8388608.0f));
            int ix = __float_as_int(mg.x) - 0x4b000000, iy = __float_as_int(mg.y) - 0x4b000000;
#endif
            ix = __vimin_s32_relu(ix, max_px);   // clip(0, max): one VIMNMX.RELU
            iy = __vimin_s32_relu(iy, max_py);
            raw[u][it] = __ldg(b.height_min3 + (unsigned)(ix * c.terrain_pitch + iy));
          }
        } else {
#pragma unroll
          for (int it = 0; it < NPA; ++it) raw[u][it] = 0;
        }
      }
      // ---- observation head (covers the gather latency) --------------------------------------------
      // noise field j of this lane (head: j = 0, 1; scan pass it: j = 2 + it) = 16-bit half (j & 1) of
      // word (j % 8) / 2 of Philox block LGK_SITE_NOISE + 3 * lane + j / 8: one block covers 6 passes
      uint4 rnd[U];
      float* orow[U];
      const float* trow[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int env_e = env0 + eu[u];
        orow[u] = b.obs_buf + (size_t)env_e * O;
        trow[u] = TAPS ? b.tap_noise + (size_t)env_e * O : nullptr;
        rnd[u] = make_uint4(0, 0, 0, 0);
        if (!TAPS && add_noise) rnd[u] = philox_keyed(a.pk, csc, (unsigned)(c.env_id_offset + env_e), LGK_SITE_NOISE + 3 * lane);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const float* hrow = sm + L.head + eu[u] * HS;
        const float o0 = noisy(hrow[lane], trow[u], lane, rnd[u].x, 0, ns_a, true);
        if (live[u] && lane < OH) orow[u][lane] = o0;
        if (lane + 32 < OH) {
          const float o1 = noisy(hrow[lane + 32], trow[u], lane + 32, rnd[u].x, 1, ns_b, true);
          if (live[u]) orow[u][lane + 32] = o1;
        }
      }
      // ---- phase B: heights -> observation columns (two copies: with and without the final clip) ----
      auto phase_b = [&](auto clip_tag) {
        constexpr bool CLIP = decltype(clip_tag)::value;
#pragma unroll
        for (int u = 0; u < U; ++u) {
          float hsum = 0.f;
#pragma unroll
          for (int it = 0; it < NP; ++it) {
            if (!TAPS && add_noise && it == 6)
              rnd[u] = philox_keyed(a.pk, csc, (unsigned)(c.env_id_offset + env0 + eu[u]), LGK_SITE_NOISE + 3 * lane + 1);
            const int w = ((it + 2) >> 1) & 3;
            const unsigned rw = w == 0 ? rnd[u].x : w == 1 ? rnd[u].y : w == 2 ? rnd[u].z : rnd[u].w;
            const int h = it * 32 + lane;
            // the specialised instantiations run with ceil(H / 32) == NP: only the last pass holds lanes past H
            const bool ok = (DT > 0 && it < NP - 1) ? true : h < H;
            const float hgt = mul_rn(s16_to_float(raw[u][it]), c.vertical_scale);
            hsum += ok ? hgt : 0.f;
            const float v = clampf(sub_rn(zo[u], hgt), -1.f, 1.f) * c.obs_scale_height;
            const float o = noisy(v, trow[u], ok ? OH + h : 0, rw, it & 1, ns_h[it], CLIP);
            if (ok && live[u]) orow[u][OH + h] = o;
          }
          if (b.measured_heights != nullptr && live[u]) {   // optional materialisation of measured_heights [N,H]
            float* mrow = b.measured_heights + (size_t)(env0 + eu[u]) * H;
#pragma unroll
            for (int it = 0; it < NP; ++it)
              if (it * 32 + lane < H) mrow[it * 32 + lane] = mul_rn(s16_to_float(raw[u][it]), c.vertical_scale);
          }
          // _reward_base_height (:871) needs mean_h(height): this lane's partial sum; added up in stage 3
          if (a.bh_active && live[u]) hpart[eu[u] * RS + lane] = hsum;
        }
      };
      if (clip_h) phase_b(std::true_type{});
      else phase_b(std::false_type{});
    }
  }
  LGK_STAMP(4);
  // =========================================== stage 3 ===========================================
  // reward terms in the reference's order (:195-212), episode sums, statistics; all four warps:
  //   A (lane = env)  each warp puts its share of the term values into the table rv[term][env]
  //   -- barrier --
  //   then warp w alone for the envs [w * epw, (w + 1) * epw) of the sub-tile:
  //   B (lane = row)  r = value * scale, episode sum += r, statistics over the reset envs; the lanes
  //                   work on independent rows, so nothing here waits on a neighbour or on a constant
  //   C (lane = (row group, env))  reward = ordered sum of the rows' r; the sums leave in 32-byte sectors
  {
    if (full && tid == 32) bulk_wait_read_all();   // the bulk stores have read their rows before anything refills them
    float* rv = sm + L.cf;     // [LGK_REW_COUNT + 1][RS]   (the contact rows are dead by now)
    float* rr = sm + L.head;   // [num_rows][RS]            (written after the barrier: the heads have left)
    float* sums = sm + L.sums; // [num_rows][RS]
    {
      auto J = [&](int k) { return scr[(SC_J1 + k) * TILE + lane] + scr[(SC_J2 + k) * TILE + lane]; };
      auto put = [&](int term, float v) { rv[term * RS + lane] = v; };
      if (wic == 0) {          // the terms on W0's registers
        const float cn = sqrtf(fmaf(cmd_rew.y, cmd_rew.y, cmd_rew.x * cmd_rew.x));
        put(LGK_REW_ANG_VEL_XY, fmaf(bav[1], bav[1], bav[0] * bav[0]));
        put(LGK_REW_LIN_VEL_Z, blv[2] * blv[2]);
        put(LGK_REW_ORIENTATION, fmaf(grav[1], grav[1], grav[0] * grav[0]));
        put(LGK_REW_FEET_AIR_TIME, scr[SC_RAIR * TILE + lane] * ((cn > 0.1f) ? 1.f : 0.f));
        put(LGK_REW_STAND_STILL, J(A_STAND) * ((cn < 0.1f) ? 1.f : 0.f));
        const float e = cmd_rew.z - bav[2];
        put(LGK_REW_TRACKING_ANG_VEL, expf(-(e * e) / c.tracking_sigma));
        const float ex = cmd_rew.x - blv[0], ey = cmd_rew.y - blv[1];
        put(LGK_REW_TRACKING_LIN_VEL, expf(-fmaf(ey, ey, ex * ex) / c.tracking_sigma));
      } else if (wic == 1) {
        put(LGK_REW_ACTION_RATE, J(A_ARATE));
        put(LGK_REW_DOF_ACC, J(A_ACC2));
        put(LGK_REW_DOF_POS_LIMITS, J(A_POSLIM));
        put(LGK_REW_DOF_VEL, J(A_QD2));
        put(LGK_REW_DOF_VEL_LIMITS, J(A_VELLIM));
      } else if (wic == 2) {
        put(LGK_REW_ENERGY, J(A_ENERGY));
        put(LGK_REW_HIP_MOTION, J(A_HIP));
        put(LGK_REW_TORQUE_LIMITS, J(A_TQLIM));
        put(LGK_REW_TORQUES, J(A_TQ2));
      } else {
        put(LGK_REW_COLLISION, scr[SC_COLL * TILE + lane]);
        put(LGK_REW_FEET_CONTACT_FORCES, scr[SC_FFP * TILE + lane]);
        put(LGK_REW_NO_FLY, scr[SC_NOFLY * TILE + lane]);
        put(LGK_REW_STUMBLE, scr[SC_STUMBLE * TILE + lane]);
        put(LGK_REW_COUNT, (reset && !time_out) ? 1.f : 0.f);   // the termination row's value (:210-212)
      }
    }
    if (wic == 3) asm volatile("cp.async.wait_group 0;" ::: "memory");   // this lane's episode-sum words have landed
    __syncthreads();   // table and base-height partial sums complete, observation heads consumed
    LGK_STAMP(5);
    const int epw = (ne + LGK_WARPS_PER_CTA - 1) / LGK_WARPS_PER_CTA;   // envs per warp (<= 8)
    const int e_lo = wic * epw, e_hi = min(e_lo + epw, ne);
    const int el = lane & 7, q = lane >> 3;
    {   // _reward_base_height (:871): mean_h(root_z - measured_height) = root_z - mean_h(height);
        // lane (q, el) adds a quarter of the partial sums of env e_lo + el
      const int e = min(e_lo + el, TILE - 1);
      float bh = scr[SC_ES * TILE + e * ES_W + 10];   // root z before the reset
      if (a.bh_active && H > 0) {
        const float* hp = sm + L.cf + HP_OFF + e * RS + q * 8;
        float hs = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) hs += hp[k];
        hs += __shfl_xor_sync(LGK_FULL_MASK, hs, 8);
        hs += __shfl_xor_sync(LGK_FULL_MASK, hs, 16);
        bh -= hs / (float)H;
      }
      const float d = bh - c.base_height_target;
      if (q == 0 && e_lo + el < e_hi) rv[LGK_REW_BASE_HEIGHT * RS + e] = d * d;
    }
    __syncwarp();
    LGK_STAMP(8);
    const unsigned my_resets = reset_mask & (e_hi > e_lo ? ((0xffffffffu >> (32 - (e_hi - e_lo))) << e_lo) : 0u);
    if (lane < num_rows) {
      const int row_term = s_row_term[lane];
      const float row_scale = s_row_scale[lane];
      float* srow = sums + lane * RS;
      float* rrow = rr + lane * RS;
      float stat = 0.f;
      if (row_term != LGK_REW_EXTERNAL) {
        const float* vrow = rv + row_term * RS;
#pragma unroll 4
        for (int e = e_lo; e < e_hi; ++e) {
          const float r = mul_rn(vrow[e], row_scale);
          const float total = add_rn(srow[e], r);
          const bool rs = (reset_mask >> e) & 1u;
          rrow[e] = r;
          srow[e] = rs ? 0.f : total;   // zeroed for reset envs after the statistics read it (:184-185)
          stat += rs ? total : 0.f;
        }
      } else {
        const float* xrow = b.ext_rewards != nullptr ? b.ext_rewards + (size_t)s_row_ext[lane] * N + env0 : nullptr;
#pragma unroll 1
        for (int e = e_lo; e < e_hi; ++e) {
          const float r = mul_rn(xrow != nullptr ? xrow[e] : 0.f, row_scale);
          const float total = add_rn(srow[e], r);
          const bool rs = (reset_mask >> e) & 1u;
          rrow[e] = r;
          srow[e] = rs ? 0.f : total;
          stat += rs ? total : 0.f;
        }
      }
      if (my_resets != 0u) atomicAdd(&ws->stats[parity][lane], stat);
    }
    LGK_STAMP(9);
    __syncwarp();
    LGK_STAMP(7);
    const int e_mine = e_lo + el;
    if (e_mine < e_hi) {
      if (q == 0) {
        float rew = 0.f;
        const int nrew = c.num_rewards;
#pragma unroll 4
        for (int i = 0; i < nrew; ++i) rew = add_rn(rew, rr[i * RS + e_mine]);
        if (flags & LGK_F_ONLY_POSITIVE) rew = fmaxf(rew, 0.f);
        // termination term is added after the positive clip (:206-212)
        if (flags & LGK_F_HAS_TERMINATION) rew = add_rn(rew, rr[nrew * RS + e_mine]);
        b.rew_buf[env0 + e_mine] = rew;
      }
      float* grow = b.episode_sums + env0 + e_mine;
#pragma unroll 2
      for (int i = q; i < num_rows; i += 4) grow[(size_t)i * N] = sums[i * RS + e_mine];
    }
  }
  LGK_STAMP(6);
  }   // sub-tiles
  if (wic == 0 && lane == 0) LGK_CTA_STAMP(1);
#ifdef LGK_PROFILE
  if (tid == 96 && blockIdx.x < 4096) lgk_prof_cta[blockIdx.x * 4 + 3] = lgk_gtime();   // W3 leaves
#endif
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

template <int DT, bool TAPS, bool MUL_RCP, int NP>
static int launch_step(const StepArgs& args, int grid, size_t smem, cudaStream_t stream) {
  auto kern = lgk_step_kernel<DT, TAPS, MUL_RCP, NP>;
  static size_t configured = 0;   // per instantiation
  if (smem > 48 * 1024 && smem > configured) {
    LGK_CUDA_RETURN_IF(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured = smem;
  }
  LGK_CUDA_RETURN_IF(lgk_launch_pdl(kern, dim3(grid), dim3(LGK_CTA_THREADS), smem, stream, args));
  LGK_CUDA_RETURN_IF(lgk_launch_pdl(lgk_step_epilogue_kernel, dim3(1), dim3(EPI_THREADS), 0, stream, args, grid));
  return (int)cudaGetLastError();
}

extern "C" int lgk_post_physics_step_impl(const LgkConsts* c, const LgkBuffers* b, cudaStream_t stream) {
  StepArgs args;
  args.c = *c;
  args.b = *b;
  const int D = c->num_dof, B3 = c->num_bodies * 3, F = c->num_feet, O = c->num_obs;
  const int OH = 12 + 3 * D;
  args.num_rows = c->num_rewards + ((c->flags & LGK_F_HAS_TERMINATION) ? 1 : 0);
  args.head_stride = OH | 1;
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
  L.head = o; o += (TILE * args.head_stride > RS * args.num_rows) ? up4(TILE * args.head_stride) : up4(RS * args.num_rows);   // later: reward rows
  L.scr = o;  o += TILE * SC_FIELDS;
  // contact rows; later the reward table [LGK_REW_COUNT + 1][RS] and, running on into the torque /
  // last_actions rows, the base-height partial sums [TILE][RS]
  {
    const int need = HP_OFF + TILE * RS - 2 * up4(TILE * D);
    L.cf = o; o += up4(TILE * B3 > need ? TILE * B3 : need);
  }
  L.tq = o;   o += up4(TILE * D);
  L.lact = o; o += up4(TILE * D);
  // episode-sum rows [num_rows][RS]: loaded after stage 1, over the dof_state rows when they fit
  if (RS * args.num_rows <= up4(TILE * 2 * D)) L.sums = L.dof;
  else { L.sums = o; o += up4(RS * args.num_rows); }
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
                al16(b->episode_sums) && al16(b->obs_buf) && (b->terrain_levels == nullptr || al16(b->terrain_levels));
  // Grid: the tile is latency-bound, so what counts is how many tiles are resident.  More
  // envs than one wave of 32-env tiles: one tile per CTA, the hardware scheduler backfills the slots
  // (measured at 65536 envs: 0.087 ms; equal persistent ranges 0.105 ms; smaller tiles for the last
  // wave 0.094 ms).  Fewer: the envs are dealt to all the resident slots in equal smaller tiles (4096
  // envs: 8-env tiles, 0.024 ms instead of 0.040 ms).  envs_per_cta > 0 overrides both.
  static int num_sms = 0;
  if (num_sms == 0) {
    int dev = 0;
    LGK_CUDA_RETURN_IF(cudaGetDevice(&dev));
    LGK_CUDA_RETURN_IF(cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev));
  }
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
  const bool taps = c->rng_mode == LGK_RNG_TAPS;
  const bool mul = c->index_math == LGK_INDEX_MUL_RCP;
  const int H = c->num_height_points;
  const int np = (H + 31) / 32;   // scan passes of 32 points
#define LGK_DISPATCH(DT, NP)                                                            \
  (taps ? (mul ? launch_step<DT, true, true, NP>(args, grid, smem, stream)              \
               : launch_step<DT, true, false, NP>(args, grid, smem, stream))            \
        : (mul ? launch_step<DT, false, true, NP>(args, grid, smem, stream)             \
               : launch_step<DT, false, false, NP>(args, grid, smem, stream)))
  if (D == 12 && np == 6) return LGK_DISPATCH(12, 6);   // 187-point scan
  if (D == 12 && np == 0) return LGK_DISPATCH(12, 0);   // blind
  if (D == 12 && np == 4) return LGK_DISPATCH(12, 4);   // Cassie's 121 points
  return LGK_DISPATCH(0, 8);
#undef LGK_DISPATCH
}

// Shared between the two kernels of post_physics_step (sm_100a):
//   lgk_scan_kernel  (lgk_scan.cu)  _get_heights + the scan columns of the observation row, lanes = points
//   lgk_step_kernel  (lgk_step.cu)  everything else, lanes = envs (four roles per 32-env tile)
// The scan depends on nothing but the simulator's root states, so it runs first; the step kernel, launched
// behind it with programmatic dependent launch, needs its result (the mean measured height of
// _reward_base_height, legged_robot.py:871) only in its last phase.
#pragma once
#include "lgk_common.cuh"

#define TILE 32
#ifndef LGK_SCAN_REG_CTAS
#define LGK_SCAN_REG_CTAS 8      // scan kernel: register cap 64 (eight 128-thread CTAs per SM would fit)
#endif
#define LGK_SCAN_CTAS_PER_SM LGK_SCAN_REG_CTAS   // ... and that many are launched per SM (LGK_SCAN_CTAS in the environment: fewer, to
                                 // leave room for step CTAs from the start -- measured slower)
#define RS (TILE + 1)   // row stride of [row][env] tables: conflict-free by row and by env
#define SCAN_GROUP 8    // envs a scan warp takes per group (one hand-over flag per group)
#ifndef LGK_TAIL_CTAS_PER_SM
#define LGK_TAIL_CTAS_PER_SM 7   // tail kernel: register cap 72; 65 536 envs are 1 024 units + the id-list CTA <= 7 x 148
#endif
#ifndef SCAN_GROUP_FUSED
#define SCAN_GROUP_FUSED 16   // ... in the fused chain (one trip of the persistent loop for most warps at 65 536 envs)
#endif

struct TileLayout {   // offsets (in floats, multiples of 4) into dynamic shared memory
  int root, dof, act, ldv, cmd, fat, ep, lvl, cf, tq, lact, head, scr, total;
};

#define LGK_LD_ARRAYS 11
struct StepArgs {
  LgkConsts c;
  LgkBuffers b;
  TileLayout L;
  int per;           // envs per CTA (multiple of 4) of the first n_big CTAs, walked in sub-tiles of <= 32
  int n_big;         // CTAs past n_big own per_small envs each (single-wave launches deal smaller tiles)
  int per_small;
  int max_sub;       // sub-tiles per CTA = ceil(per / 32)
  int head_stride;   // odd row stride of the staged observation head
  int air_active;    // feet_air_time term active (:945-948 only run then)
  int bh_active;     // base_height term active
  int tma_ok;        // every tiled array 16-byte aligned and N % 4 == 0
  int num_rows;      // episode_sums rows: num_rewards (+1 with termination)
  int scan_first;    // lgk_scan_kernel was launched ahead of the step kernel (num_height_points > 0)
  int scan_epc;      // envs per group of the scan kernel (1, 2, 4 or 8)
  float one;         // 1.0f the compiler cannot see: blocks ptxas' contraction of mul.rn.f32x2 + add.rn.f32x2
  unsigned expo23;   // 0x4b000000 (the float 2^23) the compiler cannot see: PRMT's register operand (field_as_float_bits)
  int* hsum;         // [N] sum over the scan points of the raw int16 heights (workspace), scan -> finish kernel (three-launch chain only)
  unsigned* scan_flags;   // [ceil(N / scan_epc)] (workspace): launch_seq + 1 once the scan group's columns and sums are out
  int scan_tiles;    // scan groups of this launch = ceil(N / scan_epc); the scan grid's warps walk them persistently
  int scan_grid;     // CTAs of the scan kernel: all resident at once
  int fused;         // two-launch chain: step kernel, then the scan kernel in its FUSED form (rewards, statistics)
  int step_grid;     // CTAs of the step kernel (the id-list compaction walks their mask slots)
  // fused chain, per episode-sum row (reward terms in order, then termination): where the term's values are ([N]
  // floats: a row of the step kernel's value table, or of ext_rewards), the scale, and which row is base_height (-2: none)
  const float* row_src[LGK_MAX_REWARDS];
  float row_scale[LGK_MAX_REWARDS];
  int bh_row;
  int tail_staged;   // the tail kernel copies a unit's reward rows into shared memory (cp.async) before its scan:
                     // full groups of 16 envs, every row 16-byte aligned
  float* rv;         // [LGK_REW_COUNT][rv_pitch] (workspace): value of every reward term by env, step kernel -> tail kernel
  int rv_pitch;      // floats per row of rv (N rounded up to 32)
  float* stat_copies;   // [64][LGK_MAX_REWARDS] (workspace): replicated accumulators of the reset envs' episode sums
  int* fix_list;     // [N] (workspace): the envs reset this step, in any order (reset_count of the step parity entries)
  float* side;       // [N][LGK_SIDE_FLOATS] (workspace): rows of the envs reset this step -- 8 scan operands with the new
                     // root z, the 13 floats of the new root state -- step kernel -> tail kernel
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

// where the scan sums live in the caller's workspace: behind the header and the two bitmask arrays
__host__ __device__ inline size_t lgk_ws_hsum_offset(int num_envs) {
  return ((size_t)LGK_WS_HEADER_BYTES + 2u * (size_t)lgk_max_warps(num_envs) * sizeof(unsigned) + 15u) & ~(size_t)15u;
}
// ... and behind them one flag per scan group (groups of >= 1 env)
__host__ __device__ inline size_t lgk_ws_flags_offset(int num_envs) {
  return (lgk_ws_hsum_offset(num_envs) + 4u * (size_t)num_envs + 15u) & ~(size_t)15u;
}
// ... the [term][N] table of reward-term values and the side rows of reset envs (step kernel -> tail kernel)
#define LGK_SIDE_FLOATS 24
__host__ __device__ inline int lgk_ws_rv_pitch(int num_envs) { return (num_envs + 31) & ~31; }
__host__ __device__ inline size_t lgk_ws_rv_offset(int num_envs) {
  return (lgk_ws_flags_offset(num_envs) + 4u * (size_t)(num_envs + 4) + 127u) & ~(size_t)127u;
}
__host__ __device__ inline size_t lgk_ws_side_offset(int num_envs) {
  return lgk_ws_rv_offset(num_envs) + 4u * (size_t)LGK_REW_COUNT * (size_t)lgk_ws_rv_pitch(num_envs);
}
// ... and the list of the envs reset this step (any order), filled by the step kernel for the finish kernel (three-launch chain only)
__host__ __device__ inline size_t lgk_ws_fix_offset(int num_envs) {
  return lgk_ws_side_offset(num_envs) + 4u * (size_t)LGK_SIDE_FLOATS * (size_t)num_envs;
}
// ... and 64 replicated accumulator rows of LGK_MAX_REWARDS floats for the episode sums of the reset envs
// (extras["episode"], legged_robot.py:182-185); all zero between launches
__host__ __device__ inline size_t lgk_ws_tot_offset(int num_envs) {
  return (lgk_ws_fix_offset(num_envs) + 4u * (size_t)num_envs + 127u) & ~(size_t)127u;
}
__host__ __device__ inline size_t lgk_ws_total_bytes(int num_envs) {
  return lgk_ws_tot_offset(num_envs) + 4u * (size_t)LGK_MAX_REWARDS * 64u;
}

// ---- the height scan of one env, lanes = scan points (legged_robot.py:818-854) ----------------------------
// Per-env operands of quat_apply_yaw (utils/math.py:38-42) for the unit yaw quaternion (0, 0, yz, yw):
//   t = 2 * cross(xyz, p) = (-2yz*py, 2yz*px);  r = p + yw*t + cross(xyz, t) = p + yw*t + (-yz*ty, yz*tx)
struct ScanEnv {
  float2 k_t;    // (-2yz, 2yz)
  float2 k_c;    // (-yz, yz)
  float2 r2;     // root x, y (before any reset)
  float yw;
  float zo;      // root z - 0.5 (after a reset, :227)
};
// normalize([0, 0, qz, qw]) as torch computes it: x.norm(p=2) with separately rounded squares and a correctly
// rounded square root (same on torch CPU / CUDA, tools/calibrate_norm.py), clamp(min=1e-9), true division
__device__ __forceinline__ ScanEnv scan_env_operands(float x, float y, float z_obs, float qz, float qw) {
  const float n = fmaxf(sqrtf(add_rn(mul_rn(qz, qz), mul_rn(qw, qw))), 1e-9f);
  const float yz = __fdiv_rn(qz, n), yw = __fdiv_rn(qw, n);
  const float yz2 = mul_rn(yz, 2.0f);
  ScanEnv e;
  e.k_t = make_float2(-yz2, yz2);
  e.k_c = make_float2(-yz, yz);
  e.r2 = make_float2(x, y);
  e.yw = yw;
  e.zo = sub_rn(z_obs, 0.5f);
  return e;
}

// Height-field cell of scan point p for env e, op for op like torch (no contraction): the chain runs on packed
// f32x2 instructions, x and y of the point sharing every operation; each half rounds like the scalar op.
// ptxas 12.9 contracts mul.rn.f32x2 followed by add.rn.f32x2 into one FFMA2 (a single rounding -- it does so
// even with --fmad=false; the scalar forms are left alone), which changes the truncated index in ~1e-7 of the
// points.  The two additions that follow a multiplication are therefore written as fma(p, one, q) with `one`
// a 1.0f from the kernel parameters: p * 1 + q rounds once, like p + q, and there is nothing left to contract.
template <bool MUL_RCP>
__device__ __forceinline__ int scan_cell(const ScanEnv& e, float2 p, float one, float border, float inv_hs,
                                         float hscale, int max_px, int max_py, int pitch) {
  const float2 one2 = make_float2(one, one);
  const float2 t = __fmul2_rn(e.k_t, make_float2(p.y, p.x));              // (tx, ty)
  float2 w = __ffma2_rn(p, one2, __fmul2_rn(make_float2(e.yw, e.yw), t));   // p + yw * t
  w = __ffma2_rn(w, one2, __fmul2_rn(e.k_c, make_float2(t.y, t.x)));        //   + cross(xyz, t)
  w = __fadd2_rn(__fadd2_rn(w, e.r2), make_float2(border, border));         // + root_pos, + border_size (:840-841)
  float2 s;   // (points / horizontal_scale).long(): truncation toward zero (:842)
  if (MUL_RCP) s = __fmul2_rn(w, make_float2(inv_hs, inv_hs));
  else s = make_float2(__fdiv_rn(w.x, hscale), __fdiv_rn(w.y, hscale));
  // trunc toward zero without the quarter-rate conversion unit: 2^23 + s, rounded toward zero, holds trunc(s) in
  // its mantissa for 0 <= s < 2^23; s < 0 lands below 2^23 (negative after the subtraction, clipped to 0 like
  // trunc(s) would be), s >= 2^23 / NaN above the clip's upper bound.  Subtract + clip(0, max): one VIADDMNMX.RELU
  const float2 mg = __fadd2_rz(s, make_float2(8388608.0f, 8388608.0f));
  const int ix = __vimin_s32_relu(__float_as_int(mg.x) - 0x4b000000, max_px);
  const int iy = __vimin_s32_relu(__float_as_int(mg.y) - 0x4b000000, max_py);
  return ix * pitch + iy;
}

// per-lane constants of one scan pass (the same for every env): body-frame offset of the lane's point and the
// noise factors of its observation column.  Philox mode: ka = ns * 2^-15, kb = -257 * ns (the 16-bit field r16
// enters as the float 2^23 + r16, so ns * ((2^23 + r16) * 2^-15 - 257) = ns * (r16 / 2^15 - 1)).  Tap mode: ka = ns.
struct ScanLane {
  float2 p;
  float ka, kb;
};
template <bool TAPS>
__device__ __forceinline__ ScanLane scan_lane_consts(const StepArgs& a, int h, int OH) {
  const bool ok = h < a.c.num_height_points;
  const bool add_noise = a.c.flags & LGK_F_ADD_NOISE;
  ScanLane s;
  s.p.x = ok ? __ldg(a.b.height_points_xy + 2 * h) : 0.f;   // lanes past H repeat point 0
  s.p.y = ok ? __ldg(a.b.height_points_xy + 2 * h + 1) : 0.f;
  const float ns = (ok && add_noise) ? __ldg(a.b.noise_scale_vec + OH + h) : 0.f;
  s.ka = TAPS ? ns : ns * 3.0517578125e-05f;
  s.kb = TAPS ? 0.f : ns * -257.0f;
  return s;
}

// Philox blocks of the observation noise: scan pass `it` of lane l takes 16-bit field (it & 7) of block
// LGK_SITE_NOISE_SCAN + 32 * (it / 8) + l; the head's blocks are dealt by the step kernel's roles
#define LGK_SITE_NOISE_SCAN 1024u
#define LGK_SITE_NOISE_HEAD 112u   // up to 16 blocks

// observation column of a scan point: clip(root_z - 0.5 - height, -1, 1) * scale (:227) + noise (:231), then the
// optional final clip (:104).  Philox mode folds scale and noise into two FFMAs; without noise (ka = kb = 0) they
// round like the plain product.
// 16-bit field `half` of rw under the exponent of 2^23: the float 2^23 + r16.  The exponent word travels in a register
// (from the kernel parameters, so that the compiler cannot fold it into the instruction): PRMT then takes the
// byte selector as its immediate, one instruction per field instead of a selector move + PRMT
__device__ __forceinline__ unsigned field_as_float_bits(unsigned rw, unsigned expo, int half) {
  unsigned m;
  if (half) asm("prmt.b32 %0, %1, %2, 0x7632;" : "=r"(m) : "r"(rw), "r"(expo));
  else asm("prmt.b32 %0, %1, %2, 0x7610;" : "=r"(m) : "r"(rw), "r"(expo));
  return m;
}
template <bool TAPS, bool CLIP>
__device__ __forceinline__ float scan_obs(float zo, float hgt, float scale, float ka, float kb, unsigned rw, int half,
                                          const float* __restrict__ tap, float clip_obs, unsigned expo) {
  const float d = clampf(sub_rn(zo, hgt), -1.f, 1.f);
  float v;
  if (TAPS) {
    v = mul_rn(d, scale);
    if (tap != nullptr) v = add_rn(v, mul_rn(sub_rn(mul_rn(2.0f, __ldg(tap)), 1.0f), ka));
  } else {
    v = fmaf(__uint_as_float(field_as_float_bits(rw, expo, half)), ka, fmaf(d, scale, kb));
  }
  return CLIP ? clampf(v, -clip_obs, clip_obs) : v;
}

// NP passes of 32 points (pass it0 .. it0 + NP - 1) of one env: indices, gathers (all in flight before the first is
// consumed), observation columns, measured_heights row when MAT.  Returns this lane's sum of the raw int16 heights
// (exact; times vertical_scale it is the sum of the measured heights up to the rounding of each product).
// rnd: the Philox block whose 16-bit fields 0..7 belong to passes it0..it0+7 of this lane.
// SLN: the per-lane constants are read from shared memory (sln[pass * 32 + lane] = {p.x, p.y, ka, kb}) where they are
// used instead of living in registers across envs (24 registers the scan kernel trades for resident warps).
// first half: the height-field cells of the NP points of this lane and their gathers (FMA-pipe work + loads)
template <bool MUL_RCP, int NP, bool SLN = false>
__device__ __forceinline__ void scan_gather(const StepArgs& a, const ScanEnv& e, const ScanLane (&ln)[NP], int lane,
                                            int (&raw)[NP], const float4* sln = nullptr) {
  const LgkConsts& c = a.c;
  if (!(c.flags & LGK_F_TERRAIN_PLANE)) {
    const float inv_hs = 1.0f / c.horizontal_scale;
    const int max_px = c.terrain_rows - 2, max_py = c.terrain_cols - 2;
#pragma unroll
    for (int it = 0; it < NP; ++it) {
      const float2 p = SLN ? reinterpret_cast<const float2*>(sln + it * 32 + lane)[0] : ln[it].p;
      const int cell = scan_cell<MUL_RCP>(e, p, a.one, c.border_size, inv_hs, c.horizontal_scale, max_px, max_py,
                                          c.terrain_pitch);
      raw[it] = __ldg(a.b.height_min3 + (unsigned)cell);
    }
  } else {
#pragma unroll
    for (int it = 0; it < NP; ++it) raw[it] = 0;
  }
}
// second half: heights -> observation columns (ALU-pipe work + stores); zo: root z - 0.5 of the env
template <bool TAPS, int NP, bool CLIP, bool MAT = true, bool SLN = false>
__device__ __forceinline__ int scan_emit(const StepArgs& a, float zo, const ScanLane (&ln)[NP], const int (&raw)[NP],
                                         int it0, int lane, uint4 rnd, float* __restrict__ orow,
                                         const float* __restrict__ trow, float* __restrict__ mrow, bool check_all,
                                         const float4* sln = nullptr) {
  const LgkConsts& c = a.c;
  const int H = c.num_height_points;
  int hsum = 0;
#pragma unroll
  for (int it = 0; it < NP; ++it) {
    const int h = (it0 + it) * 32 + lane;
    // the specialised instantiations run with ceil(H / 32) == NP: only the last pass holds lanes past H
    const bool ok = (!check_all && it < NP - 1) ? true : h < H;
    const unsigned rw = (it >> 1) == 0 ? rnd.x : (it >> 1) == 1 ? rnd.y : (it >> 1) == 2 ? rnd.z : rnd.w;
    const float hgt = mul_rn(s16_to_float(raw[it]), c.vertical_scale);
    hsum += ok ? raw[it] : 0;
    const float2 kk = SLN ? reinterpret_cast<const float2*>(sln + it * 32 + lane)[1] : make_float2(ln[it].ka, ln[it].kb);
    const float o = scan_obs<TAPS, CLIP>(zo, hgt, c.obs_scale_height, kk.x, kk.y, rw, it & 1,
                                         (TAPS && trow != nullptr) ? trow + (ok ? h : 0) : nullptr, c.clip_obs, a.expo23);
    if (ok) orow[h] = o;
    if (MAT && mrow != nullptr && ok) mrow[h] = hgt;   // optional materialisation of measured_heights [N,H]
  }
  return hsum;
}
template <bool TAPS, bool MUL_RCP, int NP, bool CLIP, bool MAT = true, bool SLN = false>
__device__ __forceinline__ int scan_passes(const StepArgs& a, const ScanEnv& e, const ScanLane (&ln)[NP], int it0,
                                             int lane, uint4 rnd, float* __restrict__ orow,
                                             const float* __restrict__ trow, float* __restrict__ mrow, bool check_all,
                                             const float4* sln = nullptr) {
  int raw[NP];
  scan_gather<MUL_RCP, NP, SLN>(a, e, ln, lane, raw, sln);
  return scan_emit<TAPS, NP, CLIP, MAT, SLN>(a, e.zo, ln, raw, it0, lane, rnd, orow, trow, mrow, check_all, sln);
}

// ---------------------------------------------------------------------------------------------
// The two once-per-step serial jobs of post_physics_step.  Both run on one CTA of the chain's last kernel (CTA 0 of the
// tail kernel does both; in the three-launch chain CTA 0 of the finish kernel does the first, its last CTA to finish the second):
//   1. ordered compaction of the per-tile reset / resample bitmasks into ascending id lists (:150, :342) -- needs
//      the step kernel's masks only, so a dedicated CTA does it next to the CTAs that scan / sum
//   2. episode statistics behind extras["episode"] (:182-190), step counter += 1 (:119) -- needs every CTA's sums
//      of the reset envs, so it waits for all of them (check-in counter in the workspace header)
// (earlier versions: a finisher CTA inside the fused tile kernel, 16-20 us on one busy SM; then an epilogue launch of
// its own, 7-11 us at the end of the chain with the GPU otherwise idle)
#define FIN_THREADS 128
#define STAT_COPIES 64   // replicated accumulator rows of the reset statistics (see lgk_tail_kernel / lgk_finish_kernel)
template <int THREADS>
__device__ __forceinline__ void compact_id_lists(const StepArgs& a, int n_ctas) {
  constexpr int NW = THREADS / 32;
  __shared__ int s_scan[2][NW + 1];
  const LgkBuffers& b = a.b;
  const int N = a.c.num_envs;
  const int tid = threadIdx.x, lane = tid & 31, wic = tid >> 5;
  const LgkWorkspace* ws = reinterpret_cast<const LgkWorkspace*>(b.workspace);
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
  const int per = (nt + THREADS - 1) / THREADS;   // 16 at 65536 envs
  const int t0 = min(nt, tid * per), t1 = min(nt, t0 + per);
  int c0 = 0, c1 = 0;
#pragma unroll 8
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
  if (wic == 0) {   // exclusive scan of the warp totals
    int w0 = lane < NW ? s_scan[0][lane] : 0, w1 = lane < NW ? s_scan[1][lane] : 0;
    const int t0w = w0, t1w = w1;
#pragma unroll
    for (int o = 1; o < NW; o <<= 1) {
      const int y0 = __shfl_up_sync(LGK_FULL_MASK, w0, o), y1 = __shfl_up_sync(LGK_FULL_MASK, w1, o);
      if (lane >= o) { w0 += y0; w1 += y1; }
    }
    if (lane < NW) { s_scan[0][lane] = w0 - t0w; s_scan[1][lane] = w1 - t1w; }
    if (lane == NW - 1) { s_scan[0][NW] = w0; s_scan[1][NW] = w1; }
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
    b.counts[0] = s_scan[0][NW];
    b.counts[1] = s_scan[1][NW];
  }
}

// episode statistics of the step (one CTA of the chain's last kernel, all of its threads, once every other CTA is done) and the
// counters of the next step
__device__ __forceinline__ void publish_step_statistics(const StepArgs& a, unsigned long long csc) {
  const LgkConsts& c = a.c;
  const LgkBuffers& b = a.b;
  const int N = c.num_envs;
  const int lane = threadIdx.x & 31, wic = threadIdx.x >> 5;
  constexpr int NW = FIN_THREADS / 32;
  __shared__ float s_part[NW][33];
  LgkWorkspace* ws = reinterpret_cast<LgkWorkspace*>(b.workspace);
  const int parity = (int)(csc & 1ull);
  const int num_rows = a.num_rows;
  // the reset envs' episode sums were added into STAT_COPIES replicated rows (spread by env, see the tail / finish kernel):
  // thread (row = lane, quarter = warp) adds a quarter of the copies -- independent L2 loads, all in flight -- and
  // zeroes them for the next step
  {
    float* cp = a.stat_copies + (size_t)wic * (STAT_COPIES / NW) * LGK_MAX_REWARDS + lane;
    float v[STAT_COPIES / NW];
#pragma unroll
    for (int k = 0; k < STAT_COPIES / NW; ++k) v[k] = __ldcg(cp + k * LGK_MAX_REWARDS);
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < STAT_COPIES / NW; ++k) { acc += v[k]; cp[k * LGK_MAX_REWARDS] = 0.f; }
    s_part[wic][lane] = acc;
  }
  __syncthreads();
  if (wic != 0) return;
  const int n_reset = ws->reset_count[parity];
  const int level_sum = ws->level_sum[parity];
  if (lane < num_rows) {
    float s = ws->stats[parity][lane];   // (+ what lgk_reset_idx calls may have left: nothing today)
#pragma unroll
    for (int k = 0; k < NW; ++k) s += s_part[k][lane];
    b.episode_stats[lane] = s;
    b.episode_stats[num_rows + 4 + lane] += s;   // running sums since the host last cleared them
    if (n_reset > 0)   // extras["episode"]["rew_<k>"] = mean(sum[ids]) / max_episode_length_s (:184)
      b.extras[lane] = (s / (float)n_reset) / c.max_episode_length_s;
    ws->stats[parity][lane] = 0.f;
    ws->stats[parity ^ 1][lane] = 0.f;
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
    ws->fin_ticket = 0u;
    ws->unit_next = 0u;
    ws->step_counter = csc;
    ws->launch_seq += 1u;
  }
}

// hand-over flags between kernels that run side by side: wait until *f carries this launch's stamp
__device__ __forceinline__ void wait_scan_flag(const StepArgs& a, LgkWorkspace* ws, const unsigned* f, unsigned stamp) {
  // Bounded (~0.3 s): a hand-over that never comes -- it cannot, the scan grid was resident before the step kernel
  // even started -- must fail the step's results, not hang the GPU; the header counts such give-ups
  unsigned polls = 0;
  while (flag_acquire(f) != stamp) {
    if (++polls > (1u << 22)) { atomicAdd(&ws->scan_timeouts, 1u); break; }
    __nanosleep(40);
  }
}

// launches lgk_scan_kernel for args.c.num_envs envs (lgk_scan.cu)
int lgk_scan_launch(const StepArgs& args, cudaStream_t stream);

// The scan columns of one env by one warp (three-launch chain only): the finish kernel's fix-up for the envs reset this step (their
// observation uses the new root z, :227; everything else -- indices, heights, noise -- is what the scan kernel
// computed).  sc: the env's 8 operands (k_t, k_c, r2, yw, zo).  NP = ceil(H / 32) for the specialised widths (every
// gather of the env in flight together, like the scan kernel), 0 = any H, pass by pass.
template <bool TAPS, bool MUL_RCP, int NP>
static __device__ __forceinline__ void lgk_scan_fixup_env(const StepArgs& a, const float* sc, int env,
                                                       unsigned long long csc) {
  const LgkConsts& c = a.c;
  const int lane = threadIdx.x & 31;
  const int H = c.num_height_points, OH = 12 + 3 * c.num_dof;
  ScanEnv e;
  {
    const float4 ea = reinterpret_cast<const float4*>(sc)[0], eb = reinterpret_cast<const float4*>(sc)[1];
    e.k_t = make_float2(ea.x, ea.y);
    e.k_c = make_float2(ea.z, ea.w);
    e.r2 = make_float2(eb.x, eb.y);
    e.yw = eb.z;
    e.zo = eb.w;
  }
  float* orow = a.b.obs_buf + (size_t)env * c.num_obs + OH;
  const bool noise = c.flags & LGK_F_ADD_NOISE;
  const float* trow = (TAPS && noise) ? a.b.tap_noise + (size_t)env * c.num_obs + OH : nullptr;
  const unsigned eg = (unsigned)(c.env_id_offset + env);
  if constexpr (NP > 0) {
    ScanLane ln[NP];
#pragma unroll
    for (int it = 0; it < NP; ++it) ln[it] = scan_lane_consts<TAPS>(a, it * 32 + lane, OH);
    uint4 rnd = make_uint4(0, 0, 0, 0);
    if (!TAPS && noise) rnd = philox_keyed(a.pk, csc, eg, LGK_SITE_NOISE_SCAN + lane);
    (void)scan_passes<TAPS, MUL_RCP, NP, true, false>(a, e, ln, 0, lane, rnd, orow, trow, nullptr, false);
  } else {
    const int np = (H + 31) / 32;
    uint4 rnd = make_uint4(0, 0, 0, 0);
#pragma unroll 1
    for (int it = 0; it < np; ++it) {
      if (!TAPS && noise && (it & 7) == 0)
        rnd = philox_keyed(a.pk, csc, eg, LGK_SITE_NOISE_SCAN + 32u * (unsigned)(it >> 3) + lane);
      ScanLane ln[1];
      ln[0] = scan_lane_consts<TAPS>(a, it * 32 + lane, OH);
      // field (it & 7) of the block: rotate it into the position pass 0 reads
      const int f = it & 7;
      const unsigned w = (f >> 1) == 0 ? rnd.x : (f >> 1) == 1 ? rnd.y : (f >> 1) == 2 ? rnd.z : rnd.w;
      const uint4 r1 = make_uint4((f & 1) ? (w >> 16) : w, 0, 0, 0);
      (void)scan_passes<TAPS, MUL_RCP, 1, true, false>(a, e, ln, it, lane, r1, orow, trow, nullptr, true);
    }
  }
}

// Height scan (sm_100a): _get_heights (legged_robot.py:818-854) and the scan columns of the observation row
// (:225-231) for every env, lanes = scan points -- in two kernels that share the per-env code (lgk_step.cuh):
//   lgk_tail_kernel  (second half of this file) the default: the last launch of the two-launch chain, behind the step
//                    kernel; scan, then reward sums, reset tail, id lists and statistics by the same CTAs
//   lgk_scan_kernel  (below) the scan alone, first launch of the three-launch chain of earlier rounds (LGK_CHAIN=3):
//                    it hands the sum of the measured heights (for _reward_base_height, :871) and the finished columns
//                    over to lgk_finish_kernel group by group through flags in the workspace; envs the step kernel
//                    resets get their scan columns rewritten there (the observation uses the root z after the reset,
//                    :227).  The rest of this comment describes that kernel.
//
// The scan and the step kernel do not depend on each other (the step kernel neither reads heights nor writes the
// root state: the new root row of a reset env waits in a side buffer until the finish kernel stores it).  This grid
// is persistent and resident at once; every CTA waits for the launch's predecessor, then lets the step kernel go,
// whose CTAs take SM slots as the warps here retire.  Hand-over to the finish kernel is per group of envs: when a
// group's columns and sums are out, lane 0 stamps the group's flag in the workspace with this launch's sequence
// number (st.release.gpu); the finish kernel polls the flags of its envs (ld.acquire.gpu).
//
// Work decomposition
//   grid     min(groups / 4, LGK_SCAN_CTAS_PER_SM x SMs) CTAs of four warps.  Warps are independent -- no CTA
//            barrier anywhere: warp w of the grid walks groups w, w + W, ... of `scan_epc` envs (8, fewer when there
//            are not enough envs to give every warp a group).
//   prologue lane = env of the group: root position and quaternion -> the env's eight scan operands (normalised
//            yaw quaternion as ready-made f32x2 pairs) into the warp's slice of shared memory.
//   scan     the warp takes the group's envs one at a time, lanes over the H points in passes of 32: index chain
//            on packed f32x2 instructions (bit-identical to torch's op-by-op chain), one int16 gather per point from
//            the L2-resident min-of-three-neighbours table with all gathers of an env in flight together, then
//            height -> clipped, scaled, noisy observation column, stored straight from registers (128 B per
//            warp store).  The per-lane constants (point offsets, noise factors) stay in registers across groups.
//   sums     per-lane partial sums of the heights go through shared memory; four lanes per env add them up.
#include "lgk_step.cuh"

#define LGK_SCAN_MIN_CTAS LGK_SCAN_REG_CTAS

#ifdef LGK_SCAN_STAGED
// A/B variant (north star: "shared-memory or TMA-staged terrain tiles"), built with -DLGK_SCAN_STAGED: the cells an
// env's scan can touch -- a window of ST_ROWS x ST_COLS int16 of the min-of-three table around the root -- are staged
// into the warp's shared memory by ST_ROWS 64-byte bulk async copies (double-buffered: env k + 1 travels while env k
// is scanned), and the per-point gathers read shared memory instead of L2.  Envs whose window leaves the table fall
// back to the gathers.  Measured against the default: DESIGN.md section 3.
constexpr int ST_ROWS = 24, ST_COLS = 32;
template <bool MUL_RCP, int NP>
__device__ __forceinline__ void scan_gather_tile(const StepArgs& a, const ScanEnv& e, const ScanLane (&ln)[NP],
                                                 int (&raw)[NP], const short* tile, int base32) {
  const LgkConsts& c = a.c;
  const float inv_hs = 1.0f / c.horizontal_scale;
  const int max_px = c.terrain_rows - 2, max_py = c.terrain_cols - 2;
#pragma unroll
  for (int it = 0; it < NP; ++it) {
    const int cell = scan_cell<MUL_RCP>(e, ln[it].p, a.one, c.border_size, inv_hs, c.horizontal_scale, max_px, max_py,
                                        ST_COLS);
    raw[it] = tile[cell - base32];
  }
}
#endif

template <bool TAPS, bool MUL_RCP, int NP>   // NP = 0: any H, passes in chunks of 8, constants re-read per env
__global__ void __launch_bounds__(LGK_CTA_THREADS, LGK_SCAN_MIN_CTAS)
lgk_scan_kernel(const __grid_constant__ StepArgs a) {
  const LgkConsts& c = a.c;
  const LgkBuffers& b = a.b;
  constexpr bool GENERIC = NP == 0;
  constexpr int CH = GENERIC ? 8 : NP;   // passes per chunk
  const int N = c.num_envs, H = c.num_height_points, O = c.num_obs;
  const int OH = 12 + 3 * c.num_dof;
  const int lane = threadIdx.x & 31, wic = threadIdx.x >> 5;
  const bool add_noise = c.flags & LGK_F_ADD_NOISE;

  __shared__ __align__(16) float s_es_all[LGK_WARPS_PER_CTA][SCAN_GROUP * 8];
  __shared__ int s_hp_all[LGK_WARPS_PER_CTA][SCAN_GROUP * RS];
  float* s_es = s_es_all[wic];
  int* s_hp = s_hp_all[wic];
#ifdef LGK_SCAN_STAGED
  __shared__ __align__(128) short s_tile_all[LGK_WARPS_PER_CTA][2][ST_ROWS * ST_COLS];
  __shared__ __align__(8) uint64_t s_bar_all[LGK_WARPS_PER_CTA][2];
  short (*s_tile)[ST_ROWS * ST_COLS] = s_tile_all[wic];
  uint64_t* s_bar = s_bar_all[wic];
  if (lane == 0) { mbar_init(s_bar, 1); mbar_init(s_bar + 1, 1); }
  __syncwarp();
  unsigned st_phase = 0;   // bit b: parity of buffer b's barrier
#endif

  const int epg = a.scan_epc;   // envs per group (= per flag): 8, fewer for small N

  // this lane's scan offsets and noise factors (same for every env), before the wait: they do not depend on
  // anything an earlier kernel wrote
  ScanLane ln[CH];
  bool clip_h = false;
  if (!GENERIC) {
#pragma unroll
    for (int it = 0; it < CH; ++it) {
      ln[it] = scan_lane_consts<TAPS>(a, it * 32 + lane, OH);
      // the scan columns are bounded by obs_scale_height * 1 + noise: when that is inside clip_observations for
      // every column (warp-uniform), the final clip is the identity and is skipped
      const float ns = TAPS ? ln[it].ka : ln[it].ka * 32768.0f;
      clip_h |= !(fabsf(c.obs_scale_height) + fabsf(ns) <= c.clip_obs);
    }
    clip_h = __any_sync(LGK_FULL_MASK, clip_h);
  }

#ifdef LGK_SCAN_STAGED
  float st_rc = 1.0e9f;   // window half-width in cells (+1.5 cells of margin); staging is off when it does not fit
  if (!GENERIC && !(c.flags & LGK_F_TERRAIN_PLANE)) {
    float r2m = 0.f;
#pragma unroll
    for (int it = 0; it < CH; ++it) r2m = fmaxf(r2m, fmaf(ln[it].p.x, ln[it].p.x, ln[it].p.y * ln[it].p.y));
#pragma unroll
    for (int o = 16; o; o >>= 1) r2m = fmaxf(r2m, __shfl_xor_sync(LGK_FULL_MASK, r2m, o));
    st_rc = sqrtf(r2m) / c.horizontal_scale + 1.5f;
  }
  const bool st_on = !clip_h && a.b.measured_heights == nullptr && 2.f * st_rc + 1.f <= (float)ST_ROWS &&
                     2.f * st_rc + 8.f <= (float)ST_COLS && (c.terrain_pitch & 7) == 0;
  const float st_inv_hs = 1.0f / c.horizontal_scale;
  // window of env `el` of the group (operands in s_es) -> buffer `buf`; returns ix0 * ST_COLS + iy0, or -1 (not staged)
  auto stage = [&](int el, int buf) -> int {
    if (!st_on) return -1;
    const float cx = (s_es[el * 8 + 4] + c.border_size) * st_inv_hs, cy = (s_es[el * 8 + 5] + c.border_size) * st_inv_hs;
    if (!(fabsf(cx) < 1.0e6f && fabsf(cy) < 1.0e6f)) return -1;
    const int ix0 = (int)floorf(cx - st_rc), iy0 = ((int)floorf(cy - st_rc)) & ~7;
    if (ix0 < 0 || iy0 < 0 || ix0 + ST_ROWS > c.terrain_rows - 1 || iy0 + ST_COLS > c.terrain_pitch) return -1;
    if (lane == 0) mbar_expect_tx(s_bar + buf, ST_ROWS * ST_COLS * 2);
    __syncwarp();
    if (lane < ST_ROWS)
      bulk_g2s(s_tile[buf] + lane * ST_COLS, a.b.height_min3 + (size_t)(ix0 + lane) * c.terrain_pitch + iy0, ST_COLS * 2,
               s_bar + buf);
    return ix0 * ST_COLS + iy0;
  };
#endif
  pdl_wait();   // root_states belong to the simulator: written by whatever ran before this launch
  // The whole grid is resident: the step kernel behind this launch may take SM slots as soon as warps here retire.
  pdl_launch_dependents();

  const LgkWorkspace* ws = reinterpret_cast<const LgkWorkspace*>(b.workspace);
  const unsigned long long csc = ws->step_counter + 1ull;   // common_step_counter after += 1 (:119)
  const unsigned stamp = ws->launch_seq + 1u;
  const int nchunks = GENERIC ? (H + 255) / 256 : 1;
  const int gw = (int)blockIdx.x * LGK_WARPS_PER_CTA + wic, nw = (int)gridDim.x * LGK_WARPS_PER_CTA;

#pragma unroll 1
  for (int grp = gw; grp < a.scan_tiles; grp += nw) {
    const int env0 = grp * epg;
    const int ne = min(epg, N - env0);
    if (lane < ne) {
      const float* r = b.root_states + (size_t)(env0 + lane) * 13;
      const ScanEnv e = scan_env_operands(r[0], r[1], r[2], r[5], r[6]);
      float4* d = reinterpret_cast<float4*>(s_es + lane * 8);
      d[0] = make_float4(e.k_t.x, e.k_t.y, e.k_c.x, e.k_c.y);
      d[1] = make_float4(e.r2.x, e.r2.y, e.yw, e.zo);
    }
    __syncwarp();
#ifdef LGK_SCAN_STAGED
    int st_base = stage(0, 0), st_base_next = -1;
#endif

#pragma unroll 1
    for (int el = 0; el < ne; ++el) {
      const int env = env0 + el;
#ifdef LGK_SCAN_STAGED
      if (el + 1 < ne) st_base_next = stage(el + 1, (el + 1) & 1);
#endif
      const float4 ea = reinterpret_cast<const float4*>(s_es + el * 8)[0];
      const float4 eb = reinterpret_cast<const float4*>(s_es + el * 8)[1];
      ScanEnv e;
      e.k_t = make_float2(ea.x, ea.y);
      e.k_c = make_float2(ea.z, ea.w);
      e.r2 = make_float2(eb.x, eb.y);
      e.yw = eb.z;
      e.zo = eb.w;
      float* orow = b.obs_buf + (size_t)env * O + OH;
      const float* trow = (TAPS && add_noise) ? b.tap_noise + (size_t)env * O + OH : nullptr;
      float* mrow = b.measured_heights != nullptr ? b.measured_heights + (size_t)env * H : nullptr;
      const unsigned eg = (unsigned)(c.env_id_offset + env);
      int hsum = 0;
#pragma unroll 1
      for (int ch = 0; ch < nchunks; ++ch) {
        uint4 rnd = make_uint4(0, 0, 0, 0);
        if (!TAPS && add_noise) rnd = philox_keyed(a.pk, csc, eg, LGK_SITE_NOISE_SCAN + 32u * (unsigned)ch + lane);
        if (GENERIC) {
#pragma unroll
          for (int it = 0; it < CH; ++it) ln[it] = scan_lane_consts<TAPS>(a, (ch * CH + it) * 32 + lane, OH);
          hsum += scan_passes<TAPS, MUL_RCP, CH, true>(a, e, ln, ch * CH, lane, rnd, orow, trow, mrow, true);
        } else if (clip_h || mrow != nullptr) {   // (warp-uniform)
          hsum += scan_passes<TAPS, MUL_RCP, CH, true>(a, e, ln, 0, lane, rnd, orow, trow, mrow, false);
        } else {          // the common case: no final clip needed, heights not materialised
#ifdef LGK_SCAN_STAGED
          if (st_base >= 0) {   // (warp-uniform)
            const int buf = el & 1;
            mbar_wait(s_bar + buf, (st_phase >> buf) & 1u);
            st_phase ^= 1u << buf;
            int raw[CH];
            scan_gather_tile<MUL_RCP, CH>(a, e, ln, raw, s_tile[buf], st_base);
            hsum += scan_emit<TAPS, CH, false, false, false>(a, e.zo, ln, raw, 0, lane, rnd, orow, trow, nullptr, false);
          } else
#endif
          hsum += scan_passes<TAPS, MUL_RCP, CH, false, false>(a, e, ln, 0, lane, rnd, orow, trow, nullptr, false);
        }
      }
#ifdef LGK_SCAN_STAGED
      st_base = st_base_next;
      __syncwarp();   // every lane has read the tile before the copy after next overwrites it
#endif
      s_hp[el * RS + lane] = hsum;
    }
    __syncwarp();
    // sum over the points of each env's heights: lane (q, k) adds a quarter of the partial sums of env k
    {
      const int k = lane & 7, q = lane >> 3;
      const int* hp = s_hp + k * RS + q * 8;
      int hs = 0;
#pragma unroll
      for (int i = 0; i < 8; ++i) hs += hp[i];
      hs += __shfl_xor_sync(LGK_FULL_MASK, hs, 8);
      hs += __shfl_xor_sync(LGK_FULL_MASK, hs, 16);
      if (q == 0 && k < ne) a.hsum[env0 + k] = hs;
    }
    // the group's columns and sums are out (the warp barrier also frees s_es / s_hp for the next group): hand it over
    __syncwarp();
    if (lane == 0) flag_release(a.scan_flags + grp, stamp);
  }
}

// ---------------------------------------------------------------------------------------------
// Tail kernel: the second (and last) launch of the two-launch chain, behind the step kernel -- the height scan of
// lgk_scan_kernel plus everything that is left of post_physics_step once the scan is known, by the CTAs that did the
// scan.  A CTA takes units of 4 groups (<= 64 consecutive envs) from a counter in the workspace header, warp w the
// unit's group w:
//   stage     the unit's rows of the [R,N] episode sums and of the step kernel's [term][N] value table -> shared
//             memory by 16-byte cp.async, issued before the scan: their latency passes under the CTA's own arithmetic
//             (every CTA of the grid does the same amount of work at the same time, so nobody else's would hide it)
//   scan      as in lgk_scan_kernel.  Envs the step kernel reset take their operands from the side row (old pose,
//             root z after the reset, :227): their columns are computed once, there is no fix-up pass.  The sums of the
//             heights never leave the CTA
//   rewards   compute_reward (:195-212) after a CTA barrier, from the staged copies, two threads per env (first / second
//             half of the rows): r = value * scale, sum += r in place (zeroed for reset envs after the statistics took
//             it, :184-185); base_height from the scan's height sum; then one thread per env adds the scaled values
//             up in the reference's order, clips (only_positive) and adds the termination term.  Row stores are
//             coalesced 128-byte segments.  Out-of-line functions with the row constants in shared memory (see
//             tail_rewards_env; that one -- one env per thread, rows read in place -- serves ragged env counts)
//   resets    warps 2, 3 store the root rows the step kernel held back while poses were still being read (:414-434)
//   lists     CTA 0 compacts the step kernel's per-tile bitmasks into the ascending id lists (:150, :342) ...
//   counters  ... and, once every other CTA has left its release-ordered increment, publishes extras["episode"]
//             (:182-190) and advances the step counter (:119)
// a.fused == 2: no grid-wide wait for the step kernel -- a warp waits for the flags of the step CTAs that own its
// group's envs, so the scan starts under the step kernel's last wave.
#define TAIL_UNIT (4 * SCAN_GROUP_FUSED)
// compute_reward (:195-212) for one env per thread: every row in the reference's order -- r = value * scale, sum += r
// in place (zeroed for reset envs after the statistics took it, :184-185), reward += r; only_positive clip; the
// termination row after the clip.  The loads of up to TAIL_BATCH rows (sum and value each) are in flight together:
// one or two L2 round trips per env in all.  Out of line on purpose: inlined into the scan loop -- two dozen per-lane
// constants live across it under a tight register cap -- the compiler rematerialised every address in every row;
// as a function of its own it has the registers to keep a batch of rows.  Everything comes by value or through
// shared memory (rows: per-row {source row, scale, kind}, filled from the parameter block once per CTA).
struct TailRow { float scale; int kind; };   // kind 1: base_height, 2: termination
#define TAIL_BATCH 8
static __device__ __noinline__ void tail_rewards_env(float* gp, unsigned n_envs, int nrows, int env, bool reset, float bh,
                                                     float target, float term_v, bool only_positive,
                                                     float* __restrict__ stat, const float* const* __restrict__ srcs,
                                                     const TailRow* __restrict__ rows, float* __restrict__ rew_out) {
  float rew = 0.f;
  bool clip_due = only_positive;
  const int last = nrows - 1;
#pragma unroll 1
  for (int r0 = 0; r0 < nrows; r0 += TAIL_BATCH) {
    // (slots past the last row shadow it: every load is unconditional, only the stores look at the row index)
    float s[TAIL_BATCH], v[TAIL_BATCH];
#pragma unroll
    for (int j = 0; j < TAIL_BATCH; ++j) {
      const int r = min(r0 + j, last);
      s[j] = gp[(unsigned)r * n_envs];   // (element offsets fit 32 bits: rows x envs < 2^32)
      v[j] = srcs[r][env];
    }
#pragma unroll
    for (int j = 0; j < TAIL_BATCH; ++j) {
      const int r = min(r0 + j, last);
      const bool live = r0 + j <= last;
      const TailRow t = rows[r];
      float val = v[j];
      if (t.kind != 0) {
        if (t.kind == 1) { const float d = (val - bh) - target; val = d * d; }
        else if (live) {   // the termination row: added after the positive clip (:206-212)
          val = term_v;
          if (clip_due) { rew = fmaxf(rew, 0.f); clip_due = false; }
        }
      }
      const float rr = mul_rn(val, t.scale);
      const float total = add_rn(s[j], rr);
      rew = add_rn(rew, live ? rr : 0.f);
      if (live) {
        gp[(unsigned)r * n_envs] = reset ? 0.f : total;
        if (reset) atomicAdd(stat + r, total);
      }
    }
  }
  if (clip_due) rew = fmaxf(rew, 0.f);
  *rew_out = rew;
}
// The same from staged rows, by two threads per env: the unit's episode-sum rows and term-value rows were copied into
// shared memory ([2][nrows][TAIL_UNIT] floats) by cp.async before the scan began, so nothing here waits for global
// memory -- the latency of the rows passed under the CTA's own scan.  Thread (half, env) scales and sums the first /
// second half of the rows and leaves the scaled values in the value slots; after a CTA barrier (every thread of the
// CTA calls this, active or not) thread (0, env) adds them up in the reference's order.  st: this env's column.
static __device__ __noinline__ void tail_rewards_env_staged(float* __restrict__ st, float* gp, unsigned n_envs, int nrows,
                                                            int nrew, int half, bool act, bool reset, float bh,
                                                            float target, float term_v, bool only_positive,
                                                            float* __restrict__ stat, const TailRow* __restrict__ rows,
                                                            float* __restrict__ rew_out) {
  float* sv = st + nrows * TAIL_UNIT;
  const int nh = (nrows + 1) >> 1;
  if (act) {
    const int r_end = half ? nrows : nh;
#pragma unroll 4
    for (int r = half ? nh : 0; r < r_end; ++r) {
      const TailRow t = rows[r];
      float val = sv[r * TAIL_UNIT];
      if (t.kind != 0) {
        if (t.kind == 1) { const float d = (val - bh) - target; val = d * d; }
        else val = term_v;   // the termination row (:209-212)
      }
      const float rr = mul_rn(val, t.scale);
      const float total = add_rn(st[r * TAIL_UNIT], rr);
      sv[r * TAIL_UNIT] = rr;
      gp[(unsigned)r * n_envs] = reset ? 0.f : total;   // zeroed for reset envs after the statistics took it (:184-185)
      if (reset) atomicAdd(stat + r, total);
    }
  }
  __syncthreads();
  if (act && half == 0) {
    float rew = 0.f;
#pragma unroll 4
    for (int r = 0; r < nrew; ++r) rew = add_rn(rew, sv[r * TAIL_UNIT]);
    if (only_positive) rew = fmaxf(rew, 0.f);
    if (nrows > nrew) rew = add_rn(rew, sv[nrew * TAIL_UNIT]);   // added after the positive clip (:206-212)
    *rew_out = rew;
  }
}
__device__ __forceinline__ void cp_async16(void* dst_smem, const void* src_gmem) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst_smem)), "l"(src_gmem) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
  asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
}
template <bool TAPS, bool MUL_RCP, int NP>   // NP = 0: any H (0 included), passes in chunks of 8, constants re-read per env
__global__ void __launch_bounds__(LGK_CTA_THREADS, LGK_TAIL_CTAS_PER_SM)
lgk_tail_kernel(const __grid_constant__ StepArgs a) {
  const LgkConsts& c = a.c;
  const LgkBuffers& b = a.b;
  constexpr bool GENERIC = NP == 0;
  constexpr int CH = GENERIC ? 8 : NP;   // passes per chunk
  constexpr int GS = SCAN_GROUP_FUSED;
  const int N = c.num_envs, H = c.num_height_points, O = c.num_obs;
  const int OH = 12 + 3 * c.num_dof;
  const int lane = threadIdx.x & 31, wic = threadIdx.x >> 5;
  const bool add_noise = c.flags & LGK_F_ADD_NOISE;

  __shared__ __align__(16) float s_es_all[LGK_WARPS_PER_CTA][GS * 8];
  __shared__ int s_hp_all[LGK_WARPS_PER_CTA][GS * RS];   // per-lane partial height sums of the group's envs
  __shared__ int s_hs[TAIL_UNIT];             // height sum of each env of the unit
  __shared__ unsigned char s_fl[TAIL_UNIT];   // bit 0: reset, bit 1: time-out (the step kernel's decision)
  extern __shared__ __align__(16) float s_stage[];   // a.tail_staged: [2][nrows][TAIL_UNIT] rows of the unit (sums, values)
  __shared__ __align__(8) TailRow s_rows[LGK_MAX_REWARDS];
  __shared__ const float* s_srcs[LGK_MAX_REWARDS];
  if (threadIdx.x < LGK_MAX_REWARDS) {   // (kernel parameters: nothing an earlier kernel wrote)
    const int r = threadIdx.x;
    TailRow t;
    s_srcs[r] = a.row_src[r];
    t.scale = a.row_scale[r];
    t.kind = r == a.bh_row ? 1 : (r == c.num_rewards ? 2 : 0);
    s_rows[r] = t;
  }
  __syncthreads();   // the row tables are read by every warp's staging loop
  float* s_es = s_es_all[wic];
  int* s_hp = s_hp_all[wic];

  const int epg = a.scan_epc;   // envs per group: 16, fewer for small N

  // (The step kernel lets this grid go only after its own wait, so the header read below is never ahead of the
  // previous step's last kernel.)
  const bool by_flags = a.fused == 2;
  if (!by_flags) pdl_wait();
  pdl_launch_dependents();

  LgkWorkspace* ws = reinterpret_cast<LgkWorkspace*>(b.workspace);
  const unsigned long long csc = ws->step_counter + 1ull;   // common_step_counter after += 1 (:119)
  const unsigned stamp = ws->launch_seq + 1u;
  const int nchunks = GENERIC ? (H + 255) / 256 : 1;
  // CTA 0 compacts the id lists instead of walking units (the first CTA, so that it is resident from the start)
  const bool list_cta = blockIdx.x == 0;
  if (list_cta) {
    if (by_flags) pdl_wait();   // needs every step CTA's masks
    compact_id_lists<LGK_CTA_THREADS>(a, a.step_grid);
  }
  const int nb_envs = a.n_big * a.per;   // envs of the step kernel's full-size CTAs (the rest: per_small each)
  const int unit_envs = LGK_WARPS_PER_CTA * epg;
  const int n_units = (N + unit_envs - 1) / unit_envs;
  const int nrows = a.num_rows;

  // Units are drawn from a counter in the workspace header, not dealt by CTA index: the CTAs of this grid become
  // resident one by one as the step kernel's last wave retires, and the ones that start early must take more of the
  // work -- with one fixed unit per CTA the kernel ended one full unit after the LAST CTA had started, whatever the
  // early ones had done by then.  It also takes the CTAs of an SM out of step with each other: the reward pass of
  // one (latency) meets the scan of another (arithmetic).
  __shared__ int s_unit;
  int unit = n_units;
  if (!list_cta) {
    if (threadIdx.x == 0) s_unit = (int)atomicAdd(&ws->unit_next, 1u);
    __syncthreads();
    unit = s_unit;
  }
#pragma unroll 1
  while (unit < n_units) {   // (CTA-uniform)
    const int uenv0 = unit * unit_envs;
    const int une = min(unit_envs, N - uenv0);
    // this lane's scan offsets and noise factors (same for every env).  Read per unit, not once per CTA: they must not
    // be live across the out-of-line reward pass below (the compiler kept them in local memory for the whole kernel
    // and re-read them for every env); most CTAs take one unit anyway
    ScanLane ln[CH];
    bool clip_h = false;
    if (!GENERIC) {
#pragma unroll
      for (int it = 0; it < CH; ++it) {
        ln[it] = scan_lane_consts<TAPS>(a, it * 32 + lane, OH);
        // the scan columns are bounded by obs_scale_height * 1 + noise: when that is inside clip_observations for
        // every column (warp-uniform), the final clip is the identity and is skipped
        const float ns = TAPS ? ln[it].ka : ln[it].ka * 32768.0f;
        clip_h |= !(fabsf(c.obs_scale_height) + fabsf(ns) <= c.clip_obs);
      }
      clip_h = __any_sync(LGK_FULL_MASK, clip_h);
    }
    // ---- this warp's group: scan ------------------------------------------------------------------------------
    const int env0 = uenv0 + wic * epg;
    const int ne = max(0, min(epg, N - env0));
    if (by_flags && ne > 0) {   // the step CTAs that own envs env0 .. env0 + ne - 1 are through (one lane per flag)
      const int el = env0 + ne - 1;
      const int f0 = env0 < nb_envs ? env0 / a.per : a.n_big + (env0 - nb_envs) / a.per_small;
      const int f1 = el < nb_envs ? el / a.per : a.n_big + (el - nb_envs) / a.per_small;
      if (f0 + lane <= f1) wait_scan_flag(a, ws, a.scan_flags + f0 + lane, stamp);
      __syncwarp();
    }
    // the pose and the step kernel's decision for env0 + lane: read next to each other (one round trip, not two), and
    // before the staging copies are issued, so that their latency passes under that loop
    float rx = 0.f, ry = 0.f, rz = 0.f, qz = 0.f, qw = 1.f;
    bool rst = false, tmo = false;
    if (lane < ne) {
      const float* r = b.root_states + (size_t)(env0 + lane) * 13;
      if (a.scan_first) { rx = r[0]; ry = r[1]; rz = r[2]; qz = r[5]; qw = r[6]; }
      rst = b.reset_buf[env0 + lane] != 0;
      tmo = b.time_out_buf[env0 + lane] != 0;
    }
    if (a.tail_staged && ne > 0) {   // (a unit's last groups may lie past the last env: N % 64 != 0)
      // the reward pass's rows of this group's 16 envs -> shared memory, 16 bytes per copy, asynchronously: they arrive
      // under the scan below (chunk i: row i >> 2 of the sums, then of the values; quarter i & 3 of the 64 bytes)
      const int qsh = epg == 16 ? 2 : epg == 8 ? 1 : 0;   // 4, 2 or 1 copies per row and group
      for (int i = lane; i < (nrows * 2) << qsh; i += 32) {
        const int row2 = i >> qsh, qd = i & ((1 << qsh) - 1);
        const float* src = (row2 < nrows ? b.episode_sums + (size_t)row2 * N : a.row_src[row2 - nrows]) + env0 + 4 * qd;
        cp_async16(s_stage + row2 * TAIL_UNIT + wic * epg + 4 * qd, src);
      }
    } else if (ne > 0 && lane < nrows) {   // ... or at least on their way into L2
      prefetch_l2(b.episode_sums + (size_t)lane * N + env0);
      prefetch_l2(a.row_src[lane] + env0);
    }
    if (lane < ne) {
      s_fl[wic * epg + lane] = (unsigned char)((rst ? 1 : 0) | (tmo ? 2 : 0));
      if (a.scan_first) {
        float4* d = reinterpret_cast<float4*>(s_es + lane * 8);
        if (rst) {   // reset this step: the old pose with the new root z, as the step kernel left it
          const float4* sd = reinterpret_cast<const float4*>(a.side + (size_t)(env0 + lane) * LGK_SIDE_FLOATS);
          d[0] = sd[0];
          d[1] = sd[1];
        } else {
          const ScanEnv e = scan_env_operands(rx, ry, rz, qz, qw);
          d[0] = make_float4(e.k_t.x, e.k_t.y, e.k_c.x, e.k_c.y);
          d[1] = make_float4(e.r2.x, e.r2.y, e.yw, e.zo);
        }
      }
    }
    __syncwarp();
#pragma unroll 1
    for (int el = 0; el < ne; ++el) {
      const int env = env0 + el;
      const float4 ea = reinterpret_cast<const float4*>(s_es + el * 8)[0];
      const float4 eb = reinterpret_cast<const float4*>(s_es + el * 8)[1];
      ScanEnv e;
      e.k_t = make_float2(ea.x, ea.y);
      e.k_c = make_float2(ea.z, ea.w);
      e.r2 = make_float2(eb.x, eb.y);
      e.yw = eb.z;
      e.zo = eb.w;
      float* orow = b.obs_buf + (size_t)env * O + OH;
      const float* trow = (TAPS && add_noise) ? b.tap_noise + (size_t)env * O + OH : nullptr;
      float* mrow = b.measured_heights != nullptr ? b.measured_heights + (size_t)env * H : nullptr;
      const unsigned eg = (unsigned)(c.env_id_offset + env);
      int hsum = 0;
#pragma unroll 1
      for (int ch = 0; ch < nchunks; ++ch) {
        uint4 rnd = make_uint4(0, 0, 0, 0);
        if (!TAPS && add_noise) rnd = philox_keyed(a.pk, csc, eg, LGK_SITE_NOISE_SCAN + 32u * (unsigned)ch + lane);
        if (GENERIC) {
#pragma unroll
          for (int it = 0; it < CH; ++it) ln[it] = scan_lane_consts<TAPS>(a, (ch * CH + it) * 32 + lane, OH);
          hsum += scan_passes<TAPS, MUL_RCP, CH, true>(a, e, ln, ch * CH, lane, rnd, orow, trow, mrow, true);
        } else if (clip_h || mrow != nullptr) {   // (warp-uniform)
          hsum += scan_passes<TAPS, MUL_RCP, CH, true>(a, e, ln, 0, lane, rnd, orow, trow, mrow, false);
        } else {          // the common case: no final clip needed, heights not materialised
          hsum += scan_passes<TAPS, MUL_RCP, CH, false, false>(a, e, ln, 0, lane, rnd, orow, trow, nullptr, false);
        }
      }
      s_hp[el * RS + lane] = hsum;
    }
    __syncwarp();
    {   // sum over the points of each env's heights: lane (q, k) adds half of the 32 partial sums of env k
      const int k = lane & (GS - 1), q = lane / GS;
      const int* hp = s_hp + k * RS + q * GS;
      int hs = 0;
#pragma unroll
      for (int i = 0; i < GS; ++i) hs += hp[i];
#pragma unroll
      for (int o = GS; o < 32; o <<= 1) hs += __shfl_xor_sync(LGK_FULL_MASK, hs, o);
      if (q == 0 && k < ne) s_hs[wic * epg + k] = hs;
    }
    if (a.tail_staged) cp_async_wait_all();   // (this thread's copies; the barrier makes everybody's visible)
    __syncthreads();   // every group of the unit is scanned: height sums and flags are in shared memory
    // ---- rewards of the unit: rows in the reference's order; staged rows: two threads per env ------------------------
    if (a.tail_staged) {   // (CTA-uniform: there is a barrier inside)
      const int u = threadIdx.x & (TAIL_UNIT - 1), half = threadIdx.x / TAIL_UNIT;
      const bool act = u < une;
      const int uc = act ? u : 0, env = uenv0 + uc;
      const unsigned fl = s_fl[uc];
      const bool reset = fl & 1u, time_out = fl & 2u;
      float bh = 0.f;   // _reward_base_height (:871): mean_h(root_z - measured_height) = root_z - mean_h(height)
      if (a.scan_first && a.bh_active) bh = ((float)s_hs[uc] * c.vertical_scale) / (float)H;
      tail_rewards_env_staged(s_stage + uc, b.episode_sums + env, (unsigned)N, nrows, c.num_rewards, half, act, reset, bh,
                              c.base_height_target, (reset && !time_out) ? 1.f : 0.f, (c.flags & LGK_F_ONLY_POSITIVE) != 0,
                              a.stat_copies + (size_t)((env >> 5) & (STAT_COPIES - 1)) * LGK_MAX_REWARDS, s_rows,
                              b.rew_buf + env);
    } else if ((int)threadIdx.x < une) {   // one env per thread (warps 0, 1), rows read in place
      const int u = threadIdx.x, env = uenv0 + u;
      const unsigned fl = s_fl[u];
      const bool reset = fl & 1u, time_out = fl & 2u;
      float bh = 0.f;
      if (a.scan_first && a.bh_active) bh = ((float)s_hs[u] * c.vertical_scale) / (float)H;
      tail_rewards_env(b.episode_sums + env, (unsigned)N, nrows, env, reset, bh, c.base_height_target,
                       (reset && !time_out) ? 1.f : 0.f, (c.flags & LGK_F_ONLY_POSITIVE) != 0,
                       a.stat_copies + (size_t)((env >> 5) & (STAT_COPIES - 1)) * LGK_MAX_REWARDS, s_srcs, s_rows,
                       b.rew_buf + env);
    }
    if (wic >= 2) {   // the root rows the step kernel held back while poses were still being read (:414-434)
      const int u = (wic & 1) * 32 + lane;
      unsigned m = __ballot_sync(LGK_FULL_MASK, u < une && (s_fl[u < une ? u : 0] & 1u));
      for (; m; m &= m - 1u) {   // (warp-uniform)
        const int e = uenv0 + (wic & 1) * 32 + __ffs(m) - 1;
        if (lane < 13) b.root_states[(size_t)e * 13 + lane] = a.side[(size_t)e * LGK_SIDE_FLOATS + 8 + lane];
      }
    }
    // (s_unit was read by everybody behind an earlier barrier of this trip)
    if (threadIdx.x == 0) s_unit = (int)atomicAdd(&ws->unit_next, 1u);
    __syncthreads();   // shared memory is free for the next unit, and the next unit is known
    unit = s_unit;
  }
  // ---- statistics and counters: CTA 0, once every other CTA has checked in ---------------------------------------
  // The other CTAs leave a release-ordered increment behind and go (no fence + atomic round trip + second barrier at
  // the end of every CTA: with all of them finishing together those 1 024 same-address atomics were ~4 us of every
  // CTA's life); CTA 0, idle since it compacted the id lists, polls the count and publishes.
  if (!list_cta) {
    __syncthreads();   // every warp of this CTA is through: its statistics atomics are ordered before the increment
    if (threadIdx.x == 0) ticket_release(&ws->fin_ticket);
    return;
  }
  if (threadIdx.x == 0) {   // bounded like every hand-over wait here: a count that never comes is reported, not hung on
    unsigned polls = 0;
    while (flag_acquire(&ws->fin_ticket) != gridDim.x - 1) {
      if (++polls > (1u << 22)) { atomicAdd(&ws->scan_timeouts, 1u); break; }
      __nanosleep(64);
    }
  }
  __syncthreads();
  __threadfence();   // (also drops this SM's L1 copy of the header line, read at the start of the kernel)
  if (by_flags) pdl_wait();   // (every step CTA has handed over: returns at once; the dependency is now formal)
  publish_step_statistics(a, csc);
}

template <bool TAPS, bool MUL_RCP, int NP>
static int launch_scan(const StepArgs& args, int grid, cudaStream_t stream) {
  if (args.fused)   // (the id-list CTA is part of scan_grid)
    return (int)lgk_launch_pdl(lgk_tail_kernel<TAPS, MUL_RCP, NP>, dim3(grid), dim3(LGK_CTA_THREADS),
                               args.tail_staged ? (size_t)2 * args.num_rows * TAIL_UNIT * sizeof(float) : 0, stream, args);
  return (int)lgk_launch_pdl(lgk_scan_kernel<TAPS, MUL_RCP, NP>, dim3(grid), dim3(LGK_CTA_THREADS), 0, stream, args);
}

int lgk_scan_launch(const StepArgs& args, cudaStream_t stream) {
  const LgkConsts& c = args.c;
  const int grid = args.scan_grid;
  const bool taps = c.rng_mode == LGK_RNG_TAPS;
  const bool mul = c.index_math == LGK_INDEX_MUL_RCP;
  const int np = (c.num_height_points + 31) / 32;
#define LGK_SCAN_DISPATCH(NP)                                                  \
  (taps ? (mul ? launch_scan<true, true, NP>(args, grid, stream)               \
               : launch_scan<true, false, NP>(args, grid, stream))             \
        : (mul ? launch_scan<false, true, NP>(args, grid, stream)              \
               : launch_scan<false, false, NP>(args, grid, stream)))
  if (np == 6) return LGK_SCAN_DISPATCH(6);   // 187-point scan
  if (np == 4) return LGK_SCAN_DISPATCH(4);   // Cassie's 121 points
  return LGK_SCAN_DISPATCH(0);
#undef LGK_SCAN_DISPATCH
}

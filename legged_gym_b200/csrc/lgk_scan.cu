// Height-scan kernel (sm_100a): _get_heights (legged_robot.py:818-854) and the scan columns of the observation
// row (:225-231) for every env, lanes = scan points.
//
// The scan reads nothing but the simulator's root states (position, yaw) and the terrain table, so it is the first
// launch of post_physics_step; the step kernel (lgk_step.cu) follows under programmatic dependent launch and
// needs the one number this kernel hands over per env -- the sum of the measured heights, for _reward_base_height
// (:871) -- only in its last phase.  Envs the step kernel resets get their scan columns rewritten there (the
// observation uses the root z after the reset, :227); for everyone else what this kernel writes is final.
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

template <bool TAPS, bool MUL_RCP, int NP>
static int launch_scan(const StepArgs& args, int grid, cudaStream_t stream) {
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

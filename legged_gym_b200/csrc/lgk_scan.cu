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

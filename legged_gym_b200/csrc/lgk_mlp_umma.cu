// Go1 / Aliengo actuator net on the 5th-generation tensor cores (reference envs/go1/go1.py:22-35,102-105):
// the per-leg MLP 30 -> 128 -> 128 -> 128 -> 3 (tanh) as 3xTF32 tcgen05.mma with the accumulator AND the
// activations in tensor memory (sm_100a).
//
// fp32 parity (1e-5 relative) rules out single-pass TF32, so every product is split as in lgk_mlp_tc.cu:
//   a*w ~= a_lo*w_hi + a_hi*w_hi + a_hi*w_lo   (hi = fp32 rounded to the 10-bit TF32 mantissa, lo = the
//                                               remainder, itself rounded; the dropped lo*lo is 2^-22 relative)
// Decomposition: one CTA per SM, 128 leg-rows per tile = the 128 lanes of tensor memory.
//   * A operand (activations) lives in TMEM: columns [256,384) hold a_hi, [384,512) a_lo, one row per lane.
//     The MMA reads them from there (tcgen05.mma [d], [a], b-desc) -- activations never touch shared memory.
//   * D (fp32 accumulators, 128 x 128 each) live in TMEM columns [0,128) and [128,256).
//   * B operand (weights, [out][in] = K-major as torch stores them) sits in shared memory as ready-made
//     no-swizzle UMMA images: 8x(16 B) core matrices, LBO 128 B along K, SBO (K/4)*128 B along N.  hi and lo
//     images of the two 128x128 layers are 256 KB, more than an SM has, so W0 hi/lo, W1 hi and W2 hi are
//     resident (224 KB) and one 64 KB slot is re-filled by the TMA engine with W1 lo, then W2 lo, every tile
//     (cp.async.bulk from L2, hidden behind the epilogues; the lo pass of a layer is issued last).
//   * epilogue: eight warps, thread = row (warps e and e+4 share a lane quadrant and split every 32-column
//     group).  tcgen05.ld 16 columns -> + bias, tanh, split hi/lo -> tcgen05.st into the A columns of the next
//     layer; after every group an mbarrier tells the control warp, which issues that group's k-steps of the next
//     layer into the OTHER accumulator (D0 / D1 alternate), so the tensor core runs under the epilogue.  The
//     128 -> 3 output layer is folded into the last epilogue on the FMA pipe.  Biases and W3 are immediate
//     constant-bank operands (the column index is uniform across the CTA and known at compile time).
//   * control warp: one thread issues the bulk copies and every tcgen05.mma / tcgen05.commit.
// The images and the constant block are built from the fp32 weight blob on every call by 64 extra CTAs of the
// history kernel (lgk_umma_prep.cuh, lgk_control.cu), which this kernel is launched behind with programmatic
// serialisation; the image buffer is a per-device workspace, so calls for different weight blobs must not run
// concurrently on two streams of one device.
#include "lgk_common.cuh"
#include <cstdio>
#include <cstdlib>

#include "lgk_umma_prep.cuh"

namespace {
using namespace um;
constexpr int UM_ROWS = 128;
constexpr int UM_EPI_WARPS = 8;                        // two per lane quadrant, each takes half of the columns
constexpr int UM_THREADS = (UM_EPI_WARPS + 1) * 32;      // + the control warp (TMA and MMA issue)
constexpr int UM_K0 = K0;
// shared memory: W0hi W0lo W1hi W2hi slot
constexpr uint32_t S_W0HI = 0, S_W0LO = IMG0, S_W1HI = 2 * IMG0, S_W2HI = S_W1HI + IMG, S_SLOT = S_W2HI + IMG,
                   S_BARS = S_SLOT + IMG, S_TOTAL = S_BARS + 128 + UM_ROWS * 3 * 4;
// tensor-memory columns (the two accumulators D0, D1 are [0,128) and [128,256))
constexpr uint32_t T_AHI = 256, T_ALO = 384, T_COLS = 512;
// instruction descriptor: D fp32 (1<<4), A, B tf32 (2<<7, 2<<10), both K-major, N=128 (>>3 at 17), M=128 (>>4 at 24)
constexpr uint32_t UM_IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((ACT_HID >> 3) << 17) | ((UM_ROWS >> 4) << 24);
}  // namespace

// biases of the three hidden layers times 2*log2(e) (the epilogue's tanh takes its argument in that scale),
// W3 [3][128], b3 [3] (+1 pad).  They travel BY VALUE in the kernel's parameter block (3 KB of the 32 KB sm_100
// allows): parameters live in the constant bank, so the epilogue's FFMAs still take them as immediate constant-bank
// operands, but every launch has its own copy -- no __constant__ symbol shared by the launches of a device, no
// device-side write into constant memory.  `P` is the kernel's parameter in every user below.
struct UmConsts { float v[um::CBLOCK_FLOATS]; };
#define C_BS(l, col) P.v[(l) * ACT_HID + (col)]
#define C_W3(o, col) P.v[3 * ACT_HID + (o) * ACT_HID + (col)]
#define C_B3(o) P.v[3 * ACT_HID + ACT_LEG_DOF * ACT_HID + (o)]

// ---- tcgen05 / tensor-memory helpers ---------------------------------------------------------------------
__device__ __forceinline__ void um_tmem_alloc(uint32_t* slot_smem, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot_smem)), "r"(cols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void um_tmem_free(uint32_t addr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
__device__ __forceinline__ void um_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void um_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void um_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void um_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// all MMAs issued so far by this thread -> one arrival on `bar` when they have completed
__device__ __forceinline__ void um_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
// D[128x128] (+)= A[128x8] (TMEM, one row per lane, k along the columns) * B[128x8]^T (shared-memory image)
__device__ __forceinline__ void um_mma(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(UM_IDESC), "r"(accumulate)
      : "memory");
}
// K-major, no swizzle: start address, LBO = 128 B (next core matrix along K), SBO = (K/4)*128 B (next 8 rows
// of N), descriptor version 1 (bits 46-47), layout type 0
__device__ __forceinline__ uint64_t um_desc(uint32_t smem_addr, uint32_t sbo_bytes) {
  return (uint64_t)((smem_addr & 0x3ffffu) >> 4) | ((uint64_t)(128u >> 4) << 16) | ((uint64_t)(sbo_bytes >> 4) << 32) |
         (1ull << 46);
}

#define UM_R8(v, o) "=r"(v[o]), "=r"(v[o + 1]), "=r"(v[o + 2]), "=r"(v[o + 3]), "=r"(v[o + 4]), "=r"(v[o + 5]), "=r"(v[o + 6]), "=r"(v[o + 7])
#define UM_W8(v, o) "r"(v[o]), "r"(v[o + 1]), "r"(v[o + 2]), "r"(v[o + 3]), "r"(v[o + 4]), "r"(v[o + 5]), "r"(v[o + 6]), "r"(v[o + 7])
// this warp's 32 lanes x 32 consecutive columns: thread i <-> lane (addr.lane + i), register j <-> column + j
__device__ __forceinline__ void um_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : UM_R8(v, 0), UM_R8(v, 8), UM_R8(v, 16), UM_R8(v, 24)
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void um_st32(uint32_t taddr, const uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
      ::"r"(taddr), UM_W8(v, 0), UM_W8(v, 8), UM_W8(v, 16), UM_W8(v, 24)
      : "memory");
}
__device__ __forceinline__ void um_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : UM_R8(v, 0), UM_R8(v, 8)
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void um_st16(uint32_t taddr, const uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
      ::"r"(taddr), UM_W8(v, 0), UM_W8(v, 8)
      : "memory");
}
// mbarrier wait that traps instead of hanging the device if the arrival never comes
__device__ __forceinline__ void um_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok = 0;
  const uint32_t addr = smem_u32(bar);
  uint32_t spins = 0;
  while (true) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(addr), "r"(parity)
        : "memory");
    if (ok) break;
    if (++spins > (1u << 28)) __trap();
  }
}

#ifdef UM_PROFILE
__device__ long long um_prof[64];
__device__ long long um_prof_cta[160][4];   // per CTA: start, resident images landed, end (ctl thread), tiles
#define UM_STAMP(cond, i) do { if (cond) um_prof[i] = clock64(); } while (0)
#else
#define UM_STAMP(cond, i) do { } while (0)
#endif
__device__ __forceinline__ void um_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// Activation split on three instructions: hi = v truncated to the TF32 mantissa (one LOP3; the remainder v - hi is
// exact either way), lo = the remainder plus half an ulp of ITS TF32 mantissa (the tensor core truncates, so this
// rounds it to nearest and keeps the error zero-mean).  The weights are split with hi rounded (um_split).
__device__ __forceinline__ void um_split_act(float v, uint32_t& hi, uint32_t& lo) {
  hi = __float_as_uint(v) & 0xffffe000u;
  lo = __float_as_uint(v - __uint_as_float(hi)) + 0x1000u;
}
// tanh(x) for x*2*log2(e) given: 1 - 2 / (1 + 2^t) on two bare MUFU ops; absolute error ~1e-7 (it cancels for
// small |x|), which is the rounding of the O(1) sums the activations feed; saturates exactly (ex2 -> inf / 0)
__device__ __forceinline__ float um_tanh_scaled(float t) {
  float e, r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(t));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.0f + e));
  return fmaf(-2.0f, r, 1.0f);
}
#define UM_2LOG2E 2.885390081777927f

// k-steps [ks0, ks0+n) of the two w_hi passes of a layer: D (+)= a_lo*w_hi + a_hi*w_hi.  One k-step = 8 tf32 =
// two 128 B core matrices = 256 B of the image = 8 TMEM columns.
template <int K>
__device__ __forceinline__ void um_issue_hi(uint32_t d, uint32_t tmem, uint32_t w_hi_smem, int ks0, int n, bool& first) {
  constexpr uint32_t SBO = (K / 4) * 128;
#pragma unroll
  for (int i = 0; i < n; ++i) {
    const int ks = ks0 + i;
    const uint64_t bh = um_desc(w_hi_smem + ks * 256, SBO);
    um_mma(d, tmem + T_ALO + ks * 8, bh, first ? 0u : 1u);
    first = false;
    um_mma(d, tmem + T_AHI + ks * 8, bh, 1u);
  }
}
// the w_lo pass: D += a_hi*w_lo over k-steps [ks0, ks0+n)
template <int K>
__device__ __forceinline__ void um_issue_lo(uint32_t d, uint32_t tmem, uint32_t w_lo_smem, int ks0, int n) {
  constexpr uint32_t SBO = (K / 4) * 128;
#pragma unroll
  for (int i = 0; i < n; ++i) {
    const int ks = ks0 + i;
    um_mma(d, tmem + T_AHI + ks * 8, um_desc(w_lo_smem + ks * 256, SBO), 1u);
  }
}

// hidden-layer epilogue for this thread's row, 16 columns at col0 = 32G + 16*HALF: D -> tanh(D + b) -> hi / lo
// into the A columns.  L, HALF and G are compile-time so that the biases are immediate constant-bank operands.
template <int L, int HALF, int G>
__device__ __forceinline__ void um_epilogue_group(const UmConsts& P, uint32_t trow, uint32_t d_col) {
  constexpr int col0 = G * 32 + HALF * 16;
  uint32_t v[16], hi[16], lo[16];
  um_ld16(trow + d_col + col0, v);
  um_wait_ld();
#pragma unroll
  for (int j = 0; j < 16; ++j)
    um_split_act(um_tanh_scaled(fmaf(__uint_as_float(v[j]), UM_2LOG2E, C_BS(L, col0 + j))), hi[j], lo[j]);
  um_st16(trow + T_AHI + col0, hi);
  um_st16(trow + T_ALO + col0, lo);
  um_wait_st();
}
// last hidden layer: tanh, then this thread's share of the 128 -> 3 output layer (FMA pipe, W3 from the constant bank)
template <int HALF, int G>
__device__ __forceinline__ void um_final_group(const UmConsts& P, uint32_t trow, uint32_t d_col, float (&p)[ACT_LEG_DOF]) {
  constexpr int col0 = G * 32 + HALF * 16;
  uint32_t v[16];
  um_ld16(trow + d_col + col0, v);
  um_wait_ld();
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    const float h = um_tanh_scaled(fmaf(__uint_as_float(v[j]), UM_2LOG2E, C_BS(2, col0 + j)));
#pragma unroll
    for (int o = 0; o < ACT_LEG_DOF; ++o) p[o] = fmaf(h, C_W3(o, col0 + j), p[o]);
  }
}

// a row's 30 inputs: fetched early (registers), split and stored into A columns [0,32) (+2 zero columns) later
__device__ __forceinline__ void um_fetch_inputs(float2 (&x)[ACT_IN / 2], const float* __restrict__ model_ins, int row, int rows_total) {
  const float2* src = reinterpret_cast<const float2*>(model_ins + (size_t)row * ACT_IN);
#pragma unroll
  for (int j = 0; j < ACT_IN / 2; ++j) x[j] = row < rows_total ? __ldg(src + j) : make_float2(0.f, 0.f);
}
__device__ __forceinline__ void um_stage_inputs(uint32_t trow, const float2 (&x)[ACT_IN / 2]) {
  uint32_t hi[32], lo[32];
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    const float2 v = j < ACT_IN / 2 ? x[j < ACT_IN / 2 ? j : 0] : make_float2(0.f, 0.f);
    um_split(v.x, hi[2 * j], lo[2 * j]);
    um_split(v.y, hi[2 * j + 1], lo[2 * j + 1]);
  }
  um_st32(trow + T_AHI, hi);
  um_st32(trow + T_ALO, lo);
  um_wait_st();
}

__global__ void __launch_bounds__(UM_THREADS, 1)
lgk_act_mlp_umma_kernel(int rows_total, const float* __restrict__ model_ins, const unsigned char* __restrict__ wimg,
                        const float* __restrict__ norm, float* __restrict__ dvel, const __grid_constant__ UmConsts P) {
  extern __shared__ __align__(1024) unsigned char um_smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(um_smem + S_BARS);
  uint64_t* bar_res = bars + 0;    // resident weight images landed (TMA)
  uint64_t* bar_lo = bars + 1;     // streamed lo image landed (TMA)
  uint64_t* bar_mma = bars + 2;    // a layer's MMAs completed: D readable, A writable (tcgen05.commit)
  uint64_t* bar_free = bars + 3;   // same event, for the control thread: the slot may be re-filled
  uint64_t* bar_in = bars + 4;     // next tile's inputs staged in A columns [0,32)      (4 warps)
  uint64_t* bar_fin = bars + 5;    // last epilogue of a tile has read D                  (8 warps)
  uint64_t* bar_a = bars + 6;      // [4] A columns [32g, 32g+32) of the next layer written (8 warps each)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 10);
  float* scratch = reinterpret_cast<float*>(um_smem + S_BARS + 128);   // [128][3] partial outputs of the upper half
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t sbase = smem_u32(um_smem);

  if (tid == 0) {
    mbar_init(bar_res, 1);
    mbar_init(bar_lo, 1);
    mbar_init(bar_mma, 1);
    mbar_init(bar_free, 1);
    mbar_init(bar_in, 4);
    mbar_init(bar_fin, UM_EPI_WARPS);
    for (int c = 0; c < 4; ++c) mbar_init(bar_a + c, UM_EPI_WARPS);
  }
  if (warp == UM_EPI_WARPS) um_tmem_alloc(tmem_slot, T_COLS);
  um_fence_before();
  __syncthreads();
  um_fence_after();
  const uint32_t tmem = *tmem_slot;
  const int num_tiles = (rows_total + UM_ROWS - 1) / UM_ROWS;
  // launched behind the history kernel with programmatic serialisation: the set-up above ran under its tail; its
  // writes (model_ins, the weight images, the constant block) are visible from here
  pdl_wait();

  if (warp == UM_EPI_WARPS) {
    // ---------------- control warp: one thread feeds the TMA engine and the tensor core ----------------
    if (lane == 0) {
      mbar_expect_tx(bar_res, 2 * IMG0 + 2 * IMG);
      bulk_g2s(um_smem + S_W0HI, wimg + G_W0HI, 2 * IMG0, bar_res);    // W0 hi and lo are adjacent in both
      bulk_g2s(um_smem + S_W1HI, wimg + G_W1HI, IMG, bar_res);
      bulk_g2s(um_smem + S_W2HI, wimg + G_W2HI, IMG, bar_res);
      uint32_t ph_in = 0, ph_lo = 0, ph_free = 0, ph_fin = 0, ph_a = 0, dsel = 0;
#ifdef UM_PROFILE
      um_prof_cta[blockIdx.x][0] = clock64();
#endif
      um_wait(bar_res, 0);
#ifdef UM_PROFILE
      um_prof_cta[blockIdx.x][1] = clock64();
#endif
      for (int tile = blockIdx.x, it = 0; tile < num_tiles; tile += gridDim.x, ++it) {
        const bool pr = blockIdx.x == 0 && it == 3;
        (void)pr;
        UM_STAMP(pr, 0);
        mbar_expect_tx(bar_lo, IMG);                                   // the slot is free (bar_free below)
        bulk_g2s(um_smem + S_SLOT, wimg + G_W1LO, IMG, bar_lo);
        // layer 0: all three passes from resident images
        um_wait(bar_in, ph_in); ph_in ^= 1;
        um_fence_after();
        UM_STAMP(pr, 1);
        bool first = true;
        um_issue_hi<UM_K0>(tmem + dsel * ACT_HID, tmem, sbase + S_W0HI, 0, UM_K0 / 8, first);
        um_issue_lo<UM_K0>(tmem + dsel * ACT_HID, tmem, sbase + S_W0LO, 0, UM_K0 / 8);
        um_commit(bar_mma);
        dsel ^= 1;
#pragma unroll 1
        for (int l = 1; l <= 2; ++l) {
          // this layer's accumulator is the buffer the previous tile's last epilogue read: let it finish
          if (l == 1 && it > 0) { um_wait(bar_fin, ph_fin); ph_fin ^= 1; }
          const uint32_t d = tmem + dsel * ACT_HID;
          const uint32_t w_hi = sbase + (l == 1 ? S_W1HI : S_W2HI);
          first = true;
          // all three passes follow the epilogue of the layer before, 32 columns (4 k-steps) at a time; the w_lo
          // pass joins from the second group on (the streamed image has had a group's time more to land), so
          // only 12 MMAs remain to be run after the last group arrives
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            um_wait(bar_a + g, ph_a);
            um_fence_after();
            UM_STAMP(pr, l * 8 + g);
            um_issue_hi<ACT_HID>(d, tmem, w_hi, g * 4, 4, first);
            if (g == 1) {
              um_wait(bar_lo, ph_lo); ph_lo ^= 1;
              UM_STAMP(pr, l * 8 + 5);
              um_issue_lo<ACT_HID>(d, tmem, sbase + S_SLOT, 0, 8);
            } else if (g > 1) {
              um_issue_lo<ACT_HID>(d, tmem, sbase + S_SLOT, g * 4, 4);
            }
          }
          ph_a ^= 1;
          UM_STAMP(pr, l * 8 + 4);
          um_commit(bar_mma);
          um_commit(bar_free);
          dsel ^= 1;
          UM_STAMP(pr, l * 8 + 6);
          um_wait(bar_free, ph_free); ph_free ^= 1;                    // the MMAs have read the slot
          UM_STAMP(pr, l * 8 + 7);
          if (l == 1) {
            mbar_expect_tx(bar_lo, IMG);
            bulk_g2s(um_smem + S_SLOT, wimg + G_W2LO, IMG, bar_lo);
          }
        }
#ifdef UM_PROFILE
        um_prof_cta[blockIdx.x][2] = clock64();
        um_prof_cta[blockIdx.x][3] = it + 1;
#endif
      }
    }
    __syncwarp();
  } else {
    // ---------------- epilogue warps: thread = row; warps e and e+4 share a lane quadrant ----------------
    // and split every 32-column group between them (16 columns each), so that a group is ready for the
    // tensor core after one eighth of the layer's epilogue work
    const int q = warp & 3, half = warp >> 2;
    const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16);
    uint32_t ph_mma = 0, dsel = 0;
    float2 xin[ACT_IN / 2];
    if (half == 0 && (int)blockIdx.x < num_tiles) {
      um_fetch_inputs(xin, model_ins, blockIdx.x * UM_ROWS + q * 32 + lane, rows_total);
      um_stage_inputs(trow, xin);
      um_fence_before();
      __syncwarp();
      if (lane == 0) um_arrive(bar_in);
    }
#define UM_GROUP_DONE(g)                   \
  do {                                     \
    um_fence_before();                     \
    __syncwarp();                          \
    if (lane == 0) um_arrive(bar_a + (g)); \
    UM_STAMP(pr, pb + l * 5 + 1 + (g));    \
  } while (0)
#define UM_HIDDEN_LAYER(L, H)                                           \
  do {                                                                  \
    um_epilogue_group<L, H, 0>(P, trow, dsel * ACT_HID); UM_GROUP_DONE(0); \
    um_epilogue_group<L, H, 1>(P, trow, dsel * ACT_HID); UM_GROUP_DONE(1); \
    um_epilogue_group<L, H, 2>(P, trow, dsel * ACT_HID); UM_GROUP_DONE(2); \
    um_epilogue_group<L, H, 3>(P, trow, dsel * ACT_HID); UM_GROUP_DONE(3); \
  } while (0)
    for (int tile = blockIdx.x, it = 0; tile < num_tiles; tile += gridDim.x, ++it) {
      const int row = tile * UM_ROWS + q * 32 + lane;
      const int next = tile + gridDim.x;
      const bool pr = blockIdx.x == 0 && it == 3 && lane == 0 && q == 0;
      const int pb = 32 + half * 16;
      (void)pr; (void)pb;
      // the next tile's inputs travel while this tile's epilogues run
      if (half == 0 && next < num_tiles) um_fetch_inputs(xin, model_ins, next * UM_ROWS + q * 32 + lane, rows_total);
#pragma unroll 1
      for (int l = 0; l < 2; ++l) {
        um_wait(bar_mma, ph_mma); ph_mma ^= 1;
        __syncwarp();
        um_fence_after();
        UM_STAMP(pr, pb + l * 5);
        if (half == 0) {
          if (l == 0) UM_HIDDEN_LAYER(0, 0); else UM_HIDDEN_LAYER(1, 0);
        } else {
          if (l == 0) UM_HIDDEN_LAYER(0, 1); else UM_HIDDEN_LAYER(1, 1);
        }
        dsel ^= 1;
      }
      um_wait(bar_mma, ph_mma); ph_mma ^= 1;
      __syncwarp();
      um_fence_after();
      UM_STAMP(pr, pb + 10);
      // A is free again: the lower half stages the next tile's inputs so that its layer 0 runs under this epilogue
      if (half == 0 && next < num_tiles) {
        um_stage_inputs(trow, xin);
        um_fence_before();
        __syncwarp();
        if (lane == 0) um_arrive(bar_in);
      }
      UM_STAMP(pr, pb + 11);
      float p[ACT_LEG_DOF] = {0.f, 0.f, 0.f};
      if (half == 0) {
        um_final_group<0, 0>(P, trow, dsel * ACT_HID, p); um_final_group<0, 1>(P, trow, dsel * ACT_HID, p);
        um_final_group<0, 2>(P, trow, dsel * ACT_HID, p); um_final_group<0, 3>(P, trow, dsel * ACT_HID, p);
      } else {
        um_final_group<1, 0>(P, trow, dsel * ACT_HID, p); um_final_group<1, 1>(P, trow, dsel * ACT_HID, p);
        um_final_group<1, 2>(P, trow, dsel * ACT_HID, p); um_final_group<1, 3>(P, trow, dsel * ACT_HID, p);
      }
      UM_STAMP(pr, pb + 12);
      um_fence_before();
      __syncwarp();
      if (lane == 0) um_arrive(bar_fin);
      dsel ^= 1;
      // the two column halves of a row meet through shared memory
      if (half == 1) {
#pragma unroll
        for (int o = 0; o < ACT_LEG_DOF; ++o) scratch[(q * 32 + lane) * ACT_LEG_DOF + o] = p[o];
      }
      asm volatile("bar.sync 1, %0;" ::"n"(UM_EPI_WARPS * 32) : "memory");
      if (half == 0 && row < rows_total) {
        // dVel = net(x) * vel_std   (go1.py:102-105); leg = row % 4
#pragma unroll
        for (int o = 0; o < ACT_LEG_DOF; ++o)
          dvel[(size_t)row * ACT_LEG_DOF + o] = (p[o] + scratch[(q * 32 + lane) * ACT_LEG_DOF + o] + C_B3(o)) *
                                                norm[3 * ACT_DOF + (row % ACT_LEGS) * ACT_LEG_DOF + o];
      }
    }
#undef UM_HIDDEN_LAYER
#undef UM_GROUP_DONE
  }
  um_fence_before();
  __syncthreads();
  if (warp == UM_EPI_WARPS) um_tmem_free(tmem, T_COLS);
}

// Host side.  The weight images (hi / lo UMMA images of W0, W1, W2: um::G_TOTAL bytes) live in memory the CALLER owns
// (lgk_actuator_workspace_bytes); the history kernel's extra CTAs rebuild them from the fp32 blob on every call.  The
// 772 floats that travel in the parameter block (scaled biases, W3, b3) must be on the HOST at launch time: the
// set-up call (lgk_actuator_advance with zero envs; not inside a stream capture) copies them from the device blob once
// per (device, weights pointer) into a small table here.  That table is the library's only state on this path: it is
// read-only between set-up calls and keyed by the caller's pointer, so two envs with different nets, on one stream or
// two, never share anything.  Changing the weights in place requires another set-up call.
#include <mutex>
#include <vector>
namespace {
struct UmHostEntry { int dev; const float* weights; UmConsts c; };
std::mutex g_um_mutex;
std::vector<UmHostEntry> g_um_table;
int g_um_configured[64] = {0};
}  // namespace

extern "C" int lgk_umma_setup(const float* weights, cudaStream_t st, bool refresh) {
  int dev = 0;
  LGK_CUDA_RETURN_IF(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return (int)cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lock(g_um_mutex);
  UmHostEntry* e = nullptr;
  for (auto& x : g_um_table) if (x.dev == dev && x.weights == weights) e = &x;
  if (e && !refresh) return 0;
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  LGK_CUDA_RETURN_IF(cudaStreamIsCapturing(st, &cap));
  if (cap != cudaStreamCaptureStatusNone) return (int)cudaErrorStreamCaptureUnsupported;   // make the set-up call first
  if (!g_um_configured[dev]) {
    LGK_CUDA_RETURN_IF(cudaFuncSetAttribute(lgk_act_mlp_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S_TOTAL));
    g_um_configured[dev] = 1;
  }
  // blob: W0[128,30] b0 W1[128,128] b1 W2[128,128] b2 W3[3,128] b3
  const size_t o_b0 = (size_t)ACT_HID * ACT_IN, o_b1 = o_b0 + ACT_HID + (size_t)ACT_HID * ACT_HID,
               o_b2 = o_b1 + ACT_HID + (size_t)ACT_HID * ACT_HID, o_w3 = o_b2 + ACT_HID;
  UmConsts c = {};
  LGK_CUDA_RETURN_IF(cudaStreamSynchronize(st));   // the blob may have been written on this stream just now
  LGK_CUDA_RETURN_IF(cudaMemcpy(c.v, weights + o_b0, sizeof(float) * ACT_HID, cudaMemcpyDeviceToHost));
  LGK_CUDA_RETURN_IF(cudaMemcpy(c.v + ACT_HID, weights + o_b1, sizeof(float) * ACT_HID, cudaMemcpyDeviceToHost));
  LGK_CUDA_RETURN_IF(cudaMemcpy(c.v + 2 * ACT_HID, weights + o_b2, sizeof(float) * ACT_HID, cudaMemcpyDeviceToHost));
  LGK_CUDA_RETURN_IF(cudaMemcpy(c.v + 3 * ACT_HID, weights + o_w3, sizeof(float) * (ACT_LEG_DOF * ACT_HID + ACT_LEG_DOF),
                                cudaMemcpyDeviceToHost));
  for (int i = 0; i < 3 * ACT_HID; ++i) c.v[i] *= 2.885390081777927f;   // (fp32 product, as the device prep did)
  if (e) e->c = c; else g_um_table.push_back(UmHostEntry{dev, weights, c});
  return 0;
}

// the MLP on images that are already built for these weights (the history kernel's extra CTAs build them)
extern "C" int lgk_act_mlp_umma_impl(int rows, const float* model_ins, const float* weights, const unsigned char* images,
                                     const float* norm, float* dvel, cudaStream_t st) {
  int dev = 0, sms = 148;
  LGK_CUDA_RETURN_IF(cudaGetDevice(&dev));
  if (rows <= 0) return 0;
  { const int rc = lgk_umma_setup(weights, st, false); if (rc) return rc; }   // (a table look-up after the set-up call)
  UmConsts c;
  {
    std::lock_guard<std::mutex> lock(g_um_mutex);
    const UmHostEntry* e = nullptr;
    for (const auto& x : g_um_table) if (x.dev == dev && x.weights == weights) e = &x;
    if (!e) return (int)cudaErrorInvalidValue;
    c = e->c;
  }
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int tiles = (rows + UM_ROWS - 1) / UM_ROWS;
  const int grid = tiles < sms ? tiles : sms;
  LGK_CUDA_RETURN_IF(lgk_launch_pdl(lgk_act_mlp_umma_kernel, dim3(grid), dim3(UM_THREADS), (size_t)S_TOTAL, st, rows, model_ins,
                                    images, norm, dvel, c));
#ifdef UM_PROFILE
  if (getenv("LGK_UMMA_PROFILE")) {
    long long h[64];
    cudaStreamSynchronize(st);
    cudaMemcpyFromSymbol(h, um_prof, sizeof(h));
    fprintf(stderr, "umma profile, cycles from tile start\n ctl  (in | L1: a0 a1 a2 a3 hi-issued lo-landed committed done | L2 ...):");
    for (int i = 1; i < 24; ++i) fprintf(stderr, " %lld", h[i] ? h[i] - h[0] : -1);
    fprintf(stderr, "\n epi half0 (mma0 g0 g1 g2 g3 | mma1 g0 g1 g2 g3 | mma2 staged fin):");
    for (int i = 32; i < 45; ++i) fprintf(stderr, " %lld", h[i] ? h[i] - h[0] : -1);
    fprintf(stderr, "\n epi half1:");
    for (int i = 48; i < 61; ++i) fprintf(stderr, " %lld", h[i] ? h[i] - h[0] : -1);
    long long c[160][4];
    cudaMemcpyFromSymbol(c, um_prof_cta, sizeof(c));
    long long tmax = 0, tsum = 0, rmax = 0, per_tile_max = 0, per_tile_min = 1ll << 60;
    for (int b = 0; b < grid; ++b) {
      const long long t = c[b][2] - c[b][0], r = c[b][1] - c[b][0], pt = (c[b][2] - c[b][1]) / (c[b][3] ? c[b][3] : 1);
      tmax = t > tmax ? t : tmax; tsum += t; rmax = r > rmax ? r : rmax;
      per_tile_max = pt > per_tile_max ? pt : per_tile_max; per_tile_min = pt < per_tile_min ? pt : per_tile_min;
    }
    fprintf(stderr, "\n per CTA: cycles start->end max %lld avg %lld; resident images landed after max %lld; cycles per tile min %lld max %lld\n",
            tmax, tsum / grid, rmax, per_tile_min, per_tile_max);
  }
#endif
  return (int)cudaGetLastError();
}

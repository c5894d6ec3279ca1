// Device helpers shared by the legged_gym step kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/lgk.h"

#define LGK_WARPS_PER_CTA 4
#define LGK_CTA_THREADS (LGK_WARPS_PER_CTA * 32)
#define LGK_FULL_MASK 0xffffffffu

// ---- workspace (device memory owned by the caller, zeroed once) ---------------------
struct LgkWorkspace {
  unsigned long long step_counter;   // == reference common_step_counter (legged_robot.py:119)
  unsigned int done_ticket;          // CTAs finished in the current launch
  unsigned int launch_seq;           // post_physics_step launches so far (the epilogue adds 1); value of the per-tile
                                     // scan flags of the NEXT launch is launch_seq + 1.  lgk_set_step_counter keeps it
  unsigned int fin_ticket;           // CTAs of the chain's last kernel that are done (statistics are published when all are)
  unsigned int scan_timeouts;        // step CTAs that gave up waiting for a scan tile (never, unless the launch order broke)
  unsigned int unit_next;            // tail kernel: next unit of envs to hand out (its CTAs draw units from this counter;
                                     // zero between launches)
  int level_sum[2];                  // sum(terrain_levels) accumulators, indexed by step parity
  int reset_count[2];                // number of reset envs, indexed by step parity
  float stats[2][LGK_MAX_REWARDS + 4];   // per-parity: sum over reset ids of episode_sums[k]
  // followed by: uint32 reset_mask[max_warps]; uint32 resample_mask[max_warps];
};
#define LGK_WS_HEADER_BYTES 512
#ifdef __CUDACC__
// Programmatic dependent launch (sm_90+): a kernel launched with the programmatic-serialisation attribute
// may be scheduled once every CTA of the kernel before it has passed pdl_launch_dependents(); it must call
// pdl_wait() before it touches global memory -- that returns when the earlier grid has completed and its
// writes are visible.  Both are no-ops in a launch without the attribute.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
// ticket counters: release-ordered increment (MEMBAR.ALL.GPU + RED; no L1 invalidation, unlike
// __threadfence() + atomicAdd), so everything the CTA wrote before it is visible to whoever sees the count
__device__ __forceinline__ void ticket_relaxed(unsigned int* p) {
  asm volatile("red.relaxed.gpu.global.add.u32 [%0], 1;" ::"l"(p) : "memory");
}
__device__ __forceinline__ void ticket_release(unsigned int* p) {
  asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(p) : "memory");
}
// per-tile hand-over flags between two kernels that run side by side (scan kernel -> step kernel)
__device__ __forceinline__ void flag_release(unsigned int* p, unsigned int v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int flag_acquire(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
#endif
#ifdef __CUDACC__
#include <cstdlib>
// launch with the programmatic-serialisation attribute (LGK_NO_PDL=1 in the environment: plain launch)
template <typename... KArgs, typename... Args>
static inline cudaError_t lgk_launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                         Args... args) {
  static const bool off = [] { const char* e = getenv("LGK_NO_PDL"); return e && e[0] == '1'; }();
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = off ? 0 : 1;
  return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}
#endif
static_assert(sizeof(LgkWorkspace) <= LGK_WS_HEADER_BYTES, "workspace header too large");

__host__ __device__ inline int lgk_max_warps(int num_envs) { return (num_envs + 3) / 4 + LGK_WARPS_PER_CTA; }

// ---- Philox4x32-R (counter-based; same generator family torch/curand use) ----------
// R = LGK_PHILOX_ROUNDS = 7: the fewest rounds with which Philox4x32 passes BigCrush (Salmon et al., SC'11, table 2;
// Random123 ships it as philox4x32_R(7)); the noise fields of an observation row do not need the safety margin of the
// 10-round default of curand / torch, and every round is 4 instructions per block (-DLGK_PHILOX_ROUNDS=10 builds that)
#ifndef LGK_PHILOX_ROUNDS
#define LGK_PHILOX_ROUNDS 7
#endif
struct Philox {
  static __device__ __forceinline__ uint4 round(uint4 ctr, uint2 key) {
    const unsigned int M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
    unsigned int hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
    unsigned int hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
    return make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
  }
  template <int ROUNDS = 10>
  static __device__ __forceinline__ uint4 gen(uint4 ctr, uint2 key) {
#pragma unroll
    for (int i = 0; i < ROUNDS; ++i) {
      ctr = round(ctr, key);
      key.x += 0x9E3779B9u;
      key.y += 0xBB67AE85u;
    }
    return ctr;
  }
};
// uniform in [0,1) with 24 random bits (torch's float convention)
__device__ __forceinline__ float u01(unsigned int x) { return (float)(x >> 8) * 5.9604644775390625e-08f; }

// RNG "sites" of one env-step; each owns a range of Philox blocks
enum : unsigned int {
  LGK_SITE_NOISE = 0,        // blocks 0..95: 3 * lane + k (head A, head B, then one word per scan pass)
  LGK_SITE_CMD_CB = 96,
  LGK_SITE_CMD_RESET = 97,
  LGK_SITE_PUSH = 98,
  LGK_SITE_LEVEL = 99,
  LGK_SITE_ROOT = 100,       // 2 blocks: xy(2) + vel(6)
  LGK_SITE_DOF = 102,        // ceil(D/4) blocks
  LGK_SITE_RESET_IDX = 256   // offset added by lgk_reset_idx (draws disjoint from the step's)
};
__device__ __forceinline__ uint4 lgk_philox(const uint64_t seed, unsigned long long step, unsigned int env_global,
                                            unsigned int block) {
  return Philox::gen<LGK_PHILOX_ROUNDS>(make_uint4(env_global, block, (unsigned int)step, (unsigned int)(step >> 32)),
                         make_uint2((unsigned int)seed, (unsigned int)(seed >> 32)));
}

// same generator with the 10 round keys precomputed on the host (they depend on the seed only)
__device__ __forceinline__ uint4 philox_keyed(const unsigned int (&pk)[20], unsigned long long step,
                                              unsigned int env_global, unsigned int block) {
  uint4 ctr = make_uint4(env_global, block, (unsigned int)step, (unsigned int)(step >> 32));
#pragma unroll
  for (int r = 0; r < LGK_PHILOX_ROUNDS; ++r) ctr = Philox::round(ctr, make_uint2(pk[2 * r], pk[2 * r + 1]));
  return ctr;
}

// ---- exact-order fp32 helpers (no contraction) ----------------------------------------
// int16 sample (sign-extended in a 32-bit register) -> float on the ALU's I2FP instead of the
// quarter-rate conversion unit's I2F.S16 the compiler picks when it knows the range
__device__ __forceinline__ float s16_to_float(int v) {
  float f;
  asm("cvt.rn.f32.s32 %0, %1;" : "=f"(f) : "r"(v));
  return f;
}
__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub_rn(float a, float b) { return __fsub_rn(a, b); }
// lower + (upper-lower)*u exactly as torch_rand_float evaluates it (two roundings)
__device__ __forceinline__ float rand_affine(float range, float lo, float u) { return add_rn(mul_rn(range, u), lo); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(LGK_FULL_MASK, v, o);
  return v;
}
__device__ __forceinline__ int warp_sum_int(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(LGK_FULL_MASK, v, o);
  return v;
}

// quat_rotate_inverse(q, v), q = xyzw  (isaacgym.torch_utils; SURVEY A-1)
__device__ __forceinline__ void quat_rotate_inverse(const float q[4], float vx, float vy, float vz, float out[3]) {
  const float w = q[3];
  const float s = 2.0f * w * w - 1.0f;
  const float cx = q[1] * vz - q[2] * vy, cy = q[2] * vx - q[0] * vz, cz = q[0] * vy - q[1] * vx;
  const float d = q[0] * vx + q[1] * vy + q[2] * vz;
  const float w2 = w * 2.0f, d2 = d * 2.0f;
  out[0] = vx * s - cx * w2 + q[0] * d2;
  out[1] = vy * s - cy * w2 + q[1] * d2;
  out[2] = vz * s - cz * w2 + q[2] * d2;
}
// quat_apply(q, v)
__device__ __forceinline__ void quat_apply(const float q[4], float vx, float vy, float vz, float out[3]) {
  const float tx = 2.0f * (q[1] * vz - q[2] * vy), ty = 2.0f * (q[2] * vx - q[0] * vz),
              tz = 2.0f * (q[0] * vy - q[1] * vx);
  out[0] = vx + q[3] * tx + (q[1] * tz - q[2] * ty);
  out[1] = vy + q[3] * ty + (q[2] * tx - q[0] * tz);
  out[2] = vz + q[3] * tz + (q[0] * ty - q[1] * tx);
}

// wrap_to_pi (utils/math.py:45-48): python-style modulo by fp32(2*pi), then fold (pi, 2pi) down
__device__ __forceinline__ float wrap_to_pi(float a) {
  const float TWO_PI = 6.283185307179586f, PI_F = 3.141592653589793f;
  float m = fmodf(a, TWO_PI);
  if (m != 0.0f && m < 0.0f) m += TWO_PI;
  if (m > PI_F) m -= TWO_PI;
  return m;
}

__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
__device__ __forceinline__ float clampf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }

// tanh to ~2 ulp with two MUFU ops instead of the ~25-instruction library routine: 1 - 2 / (exp(2x) + 1) away
// from zero, the odd Taylor polynomial where that form cancels (|x| < 0.15: truncation error < 2e-9 relative)
__device__ __forceinline__ float tanh_fast(float x) {
  const float ax = fabsf(x);
  const float e = exp2f(ax * 2.885390081777927f);          // exp(2|x|); inf for large |x| -> 1
  const float big = 1.0f - __fdividef(2.0f, e + 1.0f);
  const float x2 = x * x;
  const float small = ax * fmaf(x2, fmaf(x2, fmaf(x2, -0.05396825396825397f, 0.13333333333333333f), -0.3333333333333333f), 1.0f);
  return copysignf(ax < 0.15f ? small : big, x);
}

// logistic function with two MUFU ops (ex2, rcp): 1 / (1 + exp(-x)); exp(-x) = inf for very negative x -> 0
__device__ __forceinline__ float sigmoid_fast(float x) {
  return __fdividef(1.0f, 1.0f + exp2f(x * -1.4426950408889634f));
}

#define LGK_CUDA_RETURN_IF(err)            \
  do {                                     \
    cudaError_t e__ = (err);               \
    if (e__ != cudaSuccess) return (int)e__; \
  } while (0)

// ---- sm_100a async-proxy helpers: 1-D bulk copies (TMA engine, SASS UBLKCP) + mbarrier ----------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok = 0;
  const uint32_t addr = smem_u32(bar);
  while (!ok) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(addr), "r"(parity)
        : "memory");
  }
}
// global -> shared, completion counted in bytes on `bar`; dst/src 16-byte aligned, bytes % 16 == 0
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst_smem)),
               "l"(__cvta_generic_to_global(src_gmem)), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// shared -> global (bulk async-group)
__device__ __forceinline__ void bulk_s2g(void* dst_gmem, const void* src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(__cvta_generic_to_global(dst_gmem)),
               "r"(smem_u32(src_smem)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read_all() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
// generic-proxy shared-memory writes -> visible to the async proxy (before a bulk store reads them)
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// Go1 / Aliengo actuator net on the tensor pipe (reference envs/go1/go1.py:22-35,102-105): the per-leg MLP
// 30 -> 128 -> 128 -> 128 -> 3 (tanh) as 3xTF32 warp-level MMAs (sm_100a).
//
// fp32 parity (1e-5 relative) rules out single-pass TF32 (10-bit mantissa), so every product is split:
//   a = a_hi + a_lo,  w = w_hi + w_lo   (hi = the top 19 bits of the fp32 word, lo = the exact remainder)
//   a*w ~= a_lo*w_hi + a_hi*w_lo + a_hi*w_hi      (three m16n8k8 TF32 MMAs, fp32 accumulate; the dropped
//                                                  lo*lo term is 2^-22 relative)
// Decomposition: one warp carries 16 leg-rows through ALL four layers without leaving its registers.  The
// accumulator fragment of layer L (lane (g,t) holds rows g, g+8 and columns 8nb+2t, 8nb+2t+1) is used as the
// A fragment of layer L+1 as it is: inside each block of 8 the reduction index k is relabelled (fragment
// slot t <- column 2t, slot t+4 <- column 2t+1), and the B fragments are read from the weight matrix with the
// same relabelling -- one 8-byte shared-memory load W[n][8kb+2t .. +1] per fragment.  A sum does not care in
// which order its terms are labelled, so no shuffle or shared-memory round trip is needed between layers.
// Weights sit in shared memory as plain fp32 (hi + lo copies of both 128x128 layers would need 256 KB) and
// are split when a fragment is loaded: four ALU operations per weight pair-half against three MMAs per fragment.
#include "lgk_common.cuh"

#define ACT_LEGS 4
#define ACT_LEG_DOF 3
#define ACT_DOF 12
#define ACT_IN 30
#define ACT_HID 128
#define TC_THREADS 256
#define TC_ROWS (16 * TC_THREADS / 32)   // 128 leg-rows per CTA pass
#define TC_WS (ACT_HID + 8)              // row stride of W1 / W2: lanes g*TC_WS + 2t hit 32 distinct bank pairs
#define TC_W0S 40                        // W0 rows: 30 inputs padded to 32 (+8)

struct TcSmem {
  float w1[ACT_HID][TC_WS];
  float w2[ACT_HID][TC_WS];
  float w0[ACT_HID][TC_W0S];
  float w3[4][ACT_HID];
  float b0[ACT_HID], b1[ACT_HID], b2[ACT_HID], b3[4];
};

// The tensor core reads the top 19 bits of an fp32 word (it truncates).  Truncated splits would bias every
// product the same way and the bias would add up over the 128-term sums; rounding to nearest (add half an
// ulp of the 10-bit mantissa, then truncate) keeps the errors zero-mean.  hi is rounded and masked because
// the remainder is formed from it (v - hi is exact); lo only gets the half-ulp added, the hardware drops the rest.
__device__ __forceinline__ void split_tf32(float v, unsigned& hi, unsigned& lo) {
  hi = (__float_as_uint(v) + 0x1000u) & 0xffffe000u;
  lo = __float_as_uint(v - __uint_as_float(hi)) + 0x1000u;
}
__device__ __forceinline__ void mma_tf32(float (&c)[4], const unsigned (&a)[4], unsigned b0, unsigned b1) {
  asm(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// out = tanh(act . W^T + bias) for this warp's 16 rows.  Fragment convention for act / out (index 4*blk + i):
//   i = 0: (row g, col 8blk+2t)  1: (row g, col 8blk+2t+1)  2: (row g+8, col 8blk+2t)  3: (row g+8, col 8blk+2t+1)
template <int KB, int WS>
__device__ __forceinline__ void tc_layer(const float (*__restrict__ W)[WS], const float* __restrict__ bias,
                                         const float (&act)[KB * 4], float (&out)[64], int g, int t) {
  float acc[16][4];
#pragma unroll
  for (int nb = 0; nb < 16; ++nb) { acc[nb][0] = 0.f; acc[nb][1] = 0.f; acc[nb][2] = 0.f; acc[nb][3] = 0.f; }
#pragma unroll
  for (int kb = 0; kb < KB; ++kb) {
    unsigned ah[4], al[4];
    split_tf32(act[kb * 4 + 0], ah[0], al[0]);   // a0: row g,   slot t   <- column 2t
    split_tf32(act[kb * 4 + 2], ah[1], al[1]);   // a1: row g+8, slot t
    split_tf32(act[kb * 4 + 1], ah[2], al[2]);   // a2: row g,   slot t+4 <- column 2t+1
    split_tf32(act[kb * 4 + 3], ah[3], al[3]);   // a3: row g+8, slot t+4
    // four column blocks at a time: the three MMAs on one accumulator depend on each other, so the passes are
    // issued across the group (a dependent MMA is four instructions away, eight with the SM's other warp)
#pragma unroll
    for (int nb0 = 0; nb0 < 16; nb0 += 4) {
      unsigned bh[4][2], bl[4][2];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 w = *reinterpret_cast<const float2*>(&W[(nb0 + j) * 8 + g][kb * 8 + 2 * t]);
        split_tf32(w.x, bh[j][0], bl[j][0]);
        split_tf32(w.y, bh[j][1], bl[j][1]);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) mma_tf32(acc[nb0 + j], al, bh[j][0], bh[j][1]);   // small terms first
#pragma unroll
      for (int j = 0; j < 4; ++j) mma_tf32(acc[nb0 + j], ah, bl[j][0], bl[j][1]);
#pragma unroll
      for (int j = 0; j < 4; ++j) mma_tf32(acc[nb0 + j], ah, bh[j][0], bh[j][1]);
    }
  }
#pragma unroll
  for (int nb = 0; nb < 16; ++nb) {
    const float2 bb = *reinterpret_cast<const float2*>(bias + nb * 8 + 2 * t);
    out[nb * 4 + 0] = tanh_fast(acc[nb][0] + bb.x);
    out[nb * 4 + 1] = tanh_fast(acc[nb][1] + bb.y);
    out[nb * 4 + 2] = tanh_fast(acc[nb][2] + bb.x);
    out[nb * 4 + 3] = tanh_fast(acc[nb][3] + bb.y);
  }
}

__global__ void __launch_bounds__(TC_THREADS, 1)
lgk_act_mlp_tc_kernel(int rows_total, const float* __restrict__ model_ins, const float* __restrict__ weights,
                      const float* __restrict__ norm, float* __restrict__ dvel) {
  extern __shared__ __align__(16) unsigned char tc_smem_raw[];
  TcSmem& s = *reinterpret_cast<TcSmem*>(tc_smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, t = lane & 3;
  // blob: W0[128,30] b0[128] W1[128,128] b1 W2[128,128] b2 W3[3,128] b3[3]
  const float* W0 = weights;
  const float* B0 = W0 + ACT_HID * ACT_IN;
  const float* W1 = B0 + ACT_HID;
  const float* B1 = W1 + ACT_HID * ACT_HID;
  const float* W2 = B1 + ACT_HID;
  const float* B2 = W2 + ACT_HID * ACT_HID;
  const float* W3 = B2 + ACT_HID;
  const float* B3 = W3 + ACT_LEG_DOF * ACT_HID;
  for (int i = tid; i < ACT_HID * 32; i += TC_THREADS) {
    const int n = i >> 5, k = i & 31;
    s.w0[n][k] = k < ACT_IN ? W0[n * ACT_IN + k] : 0.f;
  }
  for (int i = tid; i < ACT_HID * ACT_HID; i += TC_THREADS) {
    const int n = i >> 7, k = i & 127;
    s.w1[n][k] = W1[i];
    s.w2[n][k] = W2[i];
  }
  for (int i = tid; i < ACT_LEG_DOF * ACT_HID; i += TC_THREADS) s.w3[i >> 7][i & 127] = W3[i];
  if (tid < ACT_HID) { s.b0[tid] = B0[tid]; s.b1[tid] = B1[tid]; s.b2[tid] = B2[tid]; }
  if (tid < ACT_LEG_DOF) s.b3[tid] = B3[tid];
  __syncthreads();

  const int num_tiles = (rows_total + TC_ROWS - 1) / TC_ROWS;
  for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
    const int r0 = tile * TC_ROWS + warp * 16 + g, r1 = r0 + 8;      // this lane's two rows
    // layer-0 input fragments straight from model_ins [rows, 30] (row stride 120 B: 8-byte aligned pairs)
    float x[16];
#pragma unroll
    for (int kb = 0; kb < 4; ++kb) {
      const int col = kb * 8 + 2 * t;
      float2 v0 = make_float2(0.f, 0.f), v1 = make_float2(0.f, 0.f);
      if (col < ACT_IN) {
        if (r0 < rows_total) v0 = __ldg(reinterpret_cast<const float2*>(model_ins + (size_t)r0 * ACT_IN + col));
        if (r1 < rows_total) v1 = __ldg(reinterpret_cast<const float2*>(model_ins + (size_t)r1 * ACT_IN + col));
      }
      x[kb * 4 + 0] = v0.x; x[kb * 4 + 1] = v0.y; x[kb * 4 + 2] = v1.x; x[kb * 4 + 3] = v1.y;
    }
    float a[64], b[64];
    tc_layer<4, TC_W0S>(s.w0, s.b0, x, a, g, t);
#pragma unroll 1
    for (int l = 0; l < 2; ++l) {     // one copy of the 128x128 layer's code (768 MMAs) serves both hidden layers
      tc_layer<16, TC_WS>(l == 0 ? s.w1 : s.w2, l == 0 ? s.b1 : s.b2, a, b, g, t);
#pragma unroll
      for (int i = 0; i < 64; ++i) a[i] = b[i];
    }
    // output layer 128 -> 3 on the FMA pipe: each lane owns 32 of the 128 columns of its two rows
    float p0[ACT_LEG_DOF] = {0.f, 0.f, 0.f}, p1[ACT_LEG_DOF] = {0.f, 0.f, 0.f};
#pragma unroll
    for (int nb = 0; nb < 16; ++nb) {
#pragma unroll
      for (int o = 0; o < ACT_LEG_DOF; ++o) {
        const float2 w = *reinterpret_cast<const float2*>(&s.w3[o][nb * 8 + 2 * t]);
        p0[o] = fmaf(a[nb * 4 + 1], w.y, fmaf(a[nb * 4 + 0], w.x, p0[o]));
        p1[o] = fmaf(a[nb * 4 + 3], w.y, fmaf(a[nb * 4 + 2], w.x, p1[o]));
      }
    }
#pragma unroll
    for (int o = 0; o < ACT_LEG_DOF; ++o) {
      p0[o] += __shfl_xor_sync(LGK_FULL_MASK, p0[o], 1); p0[o] += __shfl_xor_sync(LGK_FULL_MASK, p0[o], 2);
      p1[o] += __shfl_xor_sync(LGK_FULL_MASK, p1[o], 1); p1[o] += __shfl_xor_sync(LGK_FULL_MASK, p1[o], 2);
    }
    if (t < ACT_LEG_DOF) {
      const float v0 = t == 0 ? p0[0] : t == 1 ? p0[1] : p0[2];
      const float v1 = t == 0 ? p1[0] : t == 1 ? p1[1] : p1[2];
      // dVel = net(x) * vel_std   (go1.py:102-105); leg = row % 4
      if (r0 < rows_total) dvel[(size_t)r0 * ACT_LEG_DOF + t] = (v0 + s.b3[t]) * norm[3 * ACT_DOF + (r0 % ACT_LEGS) * ACT_LEG_DOF + t];
      if (r1 < rows_total) dvel[(size_t)r1 * ACT_LEG_DOF + t] = (v1 + s.b3[t]) * norm[3 * ACT_DOF + (r1 % ACT_LEGS) * ACT_LEG_DOF + t];
    }
  }
}

extern "C" int lgk_act_mlp_tc_impl(int rows, const float* model_ins, const float* weights, const float* norm,
                                   float* dvel, cudaStream_t st) {
  const int tiles = (rows + TC_ROWS - 1) / TC_ROWS;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int grid = tiles < sms ? tiles : sms;
  const int smem = (int)sizeof(TcSmem);
  LGK_CUDA_RETURN_IF(cudaFuncSetAttribute(lgk_act_mlp_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  lgk_act_mlp_tc_kernel<<<grid, TC_THREADS, smem, st>>>(rows, model_ins, weights, norm, dvel);
  return (int)cudaGetLastError();
}

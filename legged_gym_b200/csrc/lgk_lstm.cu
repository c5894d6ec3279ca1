// ANYmal series-elastic actuator net (reference envs/anymal_c/anymal.py:46-81): per joint a 2-layer
// LSTM (2 -> 8 -> 8, one time step per simulator substep) and a Linear(8 -> 1), state carried between
// substeps in sea_hidden_state / sea_cell_state [2, N*D, 8] (sm_100a).
//
// One thread per joint ("actuator").  The 972 weights travel by value in the kernel parameters, so every
// FFMA takes its weight straight from the constant bank (all lanes of a warp read the same word: no
// shared-memory staging, no load instructions for weights).  Per actuator and substep the kernel moves
// 12 B of inputs + 128 B of state in and 128 B of state + 4 B of torque out (272 B) against 1.7 kFLOP
// and 80 activations: at 49 152 actuators (4096 envs) both the byte time and the issue time are ~2 us.
#include "lgk_common.cuh"

#define LSTM_H LGK_LSTM_HID
#define LSTM_G (4 * LSTM_H)

struct LstmArgs {
  LgkLstmWeights w;
  float default_dof_pos[LGK_MAX_DOF];
  float action_scale;
  int D, total;
};
static_assert(sizeof(LstmArgs) + 64 < 32000, "kernel parameter block too large");

__device__ __forceinline__ void ld8(const float* __restrict__ p, float (&v)[LSTM_H]) {
  const float4 a = *reinterpret_cast<const float4*>(p);
  const float4 b = *reinterpret_cast<const float4*>(p + 4);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void st8(float* __restrict__ p, const float (&v)[LSTM_H]) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
}

// Gate activations on two MUFU ops each (ex2.approx + rcp.approx, ~2 ulp apiece) without the range fix-ups
// of exp2f() / division: saturation is exact (ex2 -> +inf gives rcp -> 0, ex2 -> 0 gives 1).  tanh as
// 1 - 2/(exp(2x)+1) has an ABSOLUTE error of ~1e-7 for small |x| (it cancels there); the cell state and the
// outputs it feeds are O(1) sums, so that is below the 1e-5 relative bar of the parity tests.
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float gate_sigmoid(float x) { return rcp_approx(1.0f + ex2_approx(x * -1.4426950408889634f)); }
__device__ __forceinline__ float gate_tanh(float x) {
  return fmaf(-2.0f, rcp_approx(1.0f + ex2_approx(x * 2.885390081777927f)), 1.0f);
}

// one LSTM cell step (torch.nn.LSTM, gate rows ordered i, f, g, o):
//   gates = (W_ih x + b_ih) + (W_hh h + b_hh);  c' = sig(f) c + sig(i) tanh(g);  h' = sig(o) tanh(c')
template <int IN>
__device__ __forceinline__ void lstm_cell(const float (&w_ih)[LSTM_G][IN], const float (&w_hh)[LSTM_G][LSTM_H],
                                          const float (&b_ih)[LSTM_G], const float (&b_hh)[LSTM_G],
                                          const float (&x)[IN], const float (&h)[LSTM_H], float (&c)[LSTM_H],
                                          float (&h_out)[LSTM_H]) {
#pragma unroll
  for (int j = 0; j < LSTM_H; ++j) {
    float g4[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int r = q * LSTM_H + j;
      float a = b_ih[r], b = b_hh[r];
#pragma unroll
      for (int k = 0; k < IN; ++k) a = fmaf(w_ih[r][k], x[k], a);
#pragma unroll
      for (int k = 0; k < LSTM_H; ++k) b = fmaf(w_hh[r][k], h[k], b);
      g4[q] = a + b;
    }
    const float cn = fmaf(gate_sigmoid(g4[1]), c[j], gate_sigmoid(g4[0]) * gate_tanh(g4[2]));
    c[j] = cn;
    h_out[j] = gate_sigmoid(g4[3]) * gate_tanh(cn);
  }
}

__global__ void __launch_bounds__(128)
lgk_lstm_kernel(const __grid_constant__ LstmArgs a, const float* __restrict__ actions,
                const float* __restrict__ dof_state, float* __restrict__ sea_input, float* __restrict__ hidden,
                float* __restrict__ cell, float* __restrict__ torques) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;   // actuator = env * D + joint
  if (g >= a.total) return;
  const int d = g % a.D;
  const size_t layer = (size_t)a.total * LSTM_H;
  float h0[LSTM_H], c0[LSTM_H], h1[LSTM_H], c1[LSTM_H];
  ld8(hidden + (size_t)g * LSTM_H, h0);
  ld8(cell + (size_t)g * LSTM_H, c0);
  ld8(hidden + layer + (size_t)g * LSTM_H, h1);
  ld8(cell + layer + (size_t)g * LSTM_H, c1);
  const float2 qv = __ldg(reinterpret_cast<const float2*>(dof_state) + g);
  // sea_input[:,0,0] = actions * action_scale + default_dof_pos - dof_pos; [:,0,1] = dof_vel  (anymal.py:75-76)
  const float e = add_rn(add_rn(mul_rn(__ldg(actions + g), a.action_scale), a.default_dof_pos[d]), -qv.x);
  if (sea_input) reinterpret_cast<float2*>(sea_input)[g] = make_float2(e, qv.y);
  const float x[2] = {e * a.w.in_scale[0], qv.y * a.w.in_scale[1]};
  float n0[LSTM_H], n1[LSTM_H];
  lstm_cell<2>(a.w.w_ih0, a.w.w_hh0, a.w.b_ih0, a.w.b_hh0, x, h0, c0, n0);
  lstm_cell<LSTM_H>(a.w.w_ih1, a.w.w_hh1, a.w.b_ih1, a.w.b_hh1, n0, h1, c1, n1);
  float out = a.w.lin_b;
#pragma unroll
  for (int k = 0; k < LSTM_H; ++k) out = fmaf(a.w.lin_w[k], n1[k], out);
  torques[g] = a.w.out_scale * out;
  st8(hidden + (size_t)g * LSTM_H, n0);
  st8(cell + (size_t)g * LSTM_H, c0);
  st8(hidden + layer + (size_t)g * LSTM_H, n1);
  st8(cell + layer + (size_t)g * LSTM_H, c1);
}

extern "C" int lgk_lstm_actuator_advance_impl(const LgkConsts* c, const LgkLstmWeights* w, const float* actions,
                                              const float* dof_state, float* sea_input, float* hidden, float* cell,
                                              float* torques, cudaStream_t st) {
  LstmArgs a;
  a.w = *w;
  for (int i = 0; i < LGK_MAX_DOF; ++i) a.default_dof_pos[i] = c->default_dof_pos[i];
  a.action_scale = c->action_scale;
  a.D = c->num_dof;
  a.total = c->num_envs * c->num_dof;
  lgk_lstm_kernel<<<(a.total + 127) / 128, 128, 0, st>>>(a, actions, dof_state, sea_input, hidden, cell, torques);
  return (int)cudaGetLastError();
}

// sea_hidden_state_per_env[:, env_ids] = 0; sea_cell_state_per_env[:, env_ids] = 0  (anymal.py:56-60) for a
// device-side id list: one thread per (list slot, joint, layer); slots past the count do nothing.
__global__ void lgk_lstm_reset_kernel(const int* __restrict__ ids, const int* __restrict__ count_dev, int count_host,
                                      int max_ids, int D, int total, float* __restrict__ hidden,
                                      float* __restrict__ cell) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int n = count_dev ? min(*count_dev, max_ids) : count_host;
  const int slot = t / (2 * D), rem = t % (2 * D);
  if (slot >= n) return;
  const int env = ids ? ids[slot] : slot;
  const size_t off = ((size_t)(rem / D) * total + (size_t)env * D + (rem % D)) * LSTM_H;
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
  *reinterpret_cast<float4*>(hidden + off) = z;
  *reinterpret_cast<float4*>(hidden + off + 4) = z;
  *reinterpret_cast<float4*>(cell + off) = z;
  *reinterpret_cast<float4*>(cell + off + 4) = z;
}

extern "C" int lgk_lstm_reset_hidden_impl(int num_envs, int D, const int* ids, const int* count_dev, int count_host,
                                          float* hidden, float* cell, cudaStream_t st) {
  const int max_ids = count_dev ? num_envs : count_host;
  if (max_ids <= 0) return 0;
  const long long threads = (long long)max_ids * 2 * D;
  lgk_lstm_reset_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(ids, count_dev, count_host, max_ids, D,
                                                                            num_envs * D, hidden, cell);
  return (int)cudaGetLastError();
}

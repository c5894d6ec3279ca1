// Control-side kernels: action clip, position targets, PD torques, terrain min-of-3 table,
// Go1 actuator-net history + MLP (sm_100a).
#include "lgk_common.cuh"

// ---------------------------------------------------------------------------------------------
// height_min3[x,y] = min(h[x,y], h[x+1,y], h[x,y+1])   (legged_robot.py:848-852)
__global__ void lgk_min3_kernel(const int16_t* __restrict__ hs, int rows, int cols, int16_t* __restrict__ out, int pitch) {
  const int y = blockIdx.x * blockDim.x + threadIdx.x;
  const int x = blockIdx.y;
  if (y >= pitch) return;
  int16_t v = 0;
  if (y < cols) {
    v = hs[(size_t)x * cols + y];
    if (x + 1 < rows) v = min(v, hs[(size_t)(x + 1) * cols + y]);
    if (y + 1 < cols) v = min(v, hs[(size_t)x * cols + y + 1]);
  }
  out[(size_t)x * pitch + y] = v;
}

extern "C" int lgk_build_height_min3_impl(const int16_t* hs, int rows, int cols, int16_t* out, int pitch, cudaStream_t st) {
  dim3 grid((pitch + 255) / 256, rows);
  lgk_min3_kernel<<<grid, 256, 0, st>>>(hs, rows, cols, out, pitch);
  return (int)cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// Calibration probe: normalize([0,0,qz,qw]) in each LgkIndexMath norm mode, so a test can find the
// mode that reproduces torch's `x / x.norm(p=2, dim=-1).clamp(min=eps)` bit-for-bit on a device.
__global__ void lgk_yaw_quat_kernel(const float* __restrict__ quat, int n, int mode, float* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float z = quat[4 * i + 2], w = quat[4 * i + 3];
  const float zz = mul_rn(z, z), ww = mul_rn(w, w);
  float nn;
  if (mode == 0) nn = sqrtf(fmaf(w, w, zz));
  else if (mode == 1) nn = sqrtf(add_rn(zz, ww));
  else if (mode == 2) nn = sqrtf(fmaf(z, z, ww));
  else nn = (float)sqrt((double)add_rn(zz, ww));
  nn = fmaxf(nn, 1e-9f);
  out[2 * i] = __fdiv_rn(z, nn);
  out[2 * i + 1] = __fdiv_rn(w, nn);
}
extern "C" int lgk_debug_yaw_quat_impl(const float* quat, int n, int mode, float* out, cudaStream_t st) {
  lgk_yaw_quat_kernel<<<(n + 255) / 256, 256, 0, st>>>(quat, n, mode, out);
  return (int)cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
__global__ void lgk_clip_kernel(const float* __restrict__ in, float* __restrict__ out, long long n, float lim) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long n4 = n >> 2;
  if (i < n4) {
    float4 v = __ldg(reinterpret_cast<const float4*>(in) + i);
    v.x = clampf(v.x, -lim, lim); v.y = clampf(v.y, -lim, lim);
    v.z = clampf(v.z, -lim, lim); v.w = clampf(v.w, -lim, lim);
    reinterpret_cast<float4*>(out)[i] = v;
  }
  if (i < (n & 3)) out[n4 * 4 + i] = clampf(in[n4 * 4 + i], -lim, lim);
}
__global__ void lgk_clip_kernel_scalar(const float* __restrict__ in, float* __restrict__ out, long long n, float lim) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = clampf(in[i], -lim, lim);
}

extern "C" int lgk_clip_actions_impl(const float* in, float* out, long long n, float lim, cudaStream_t st) {
  if (n <= 0) return 0;
  const bool aligned = (((uintptr_t)in | (uintptr_t)out) & 15) == 0;
  if (aligned) {
    const long long work = (n >> 2) > 3 ? (n >> 2) : 4;
    lgk_clip_kernel<<<(unsigned)((work + 255) / 256), 256, 0, st>>>(in, out, n, lim);
  } else {
    lgk_clip_kernel_scalar<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(in, out, n, lim);
  }
  return (int)cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
struct CtrlArgs {
  float action_scale, sim_dt;
  int D, control_type;
  float def[LGK_MAX_DOF], lo[LGK_MAX_DOF], hi[LGK_MAX_DOF];
  float kp[LGK_MAX_DOF], kd[LGK_MAX_DOF], tl[LGK_MAX_DOF];
};

// _compute_poses (:394-397): clip(a * action_scale + default, lo, hi)
__global__ void lgk_poses_kernel(const __grid_constant__ CtrlArgs a, const float* __restrict__ actions,
                                 float* __restrict__ out, long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int d = (int)(i % a.D);
  const float t = add_rn(mul_rn(__ldg(actions + i), a.action_scale), a.def[d]);
  out[i] = clampf(t, a.lo[d], a.hi[d]);
}

// _compute_torques (:370-392)
__global__ void lgk_torques_kernel(const __grid_constant__ CtrlArgs a, const float* __restrict__ actions,
                                   const float* __restrict__ dof_state, const float* __restrict__ last_dof_vel,
                                   float* __restrict__ out, long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int d = (int)(i % a.D);
  const float as = mul_rn(__ldg(actions + i), a.action_scale);
  const float2 s = __ldg(reinterpret_cast<const float2*>(dof_state) + i);   // (pos, vel)
  float t;
  if (a.control_type == LGK_CTRL_P) {
    t = sub_rn(mul_rn(a.kp[d], sub_rn(add_rn(as, a.def[d]), s.x)), mul_rn(a.kd[d], s.y));
  } else if (a.control_type == LGK_CTRL_V) {
    const float dv = sub_rn(s.y, __ldg(last_dof_vel + i));
    t = sub_rn(mul_rn(a.kp[d], sub_rn(as, s.y)), __fdiv_rn(mul_rn(a.kd[d], dv), a.sim_dt));
  } else {
    t = as;
  }
  out[i] = clampf(t, -a.tl[d], a.tl[d]);
}

static void fill_ctrl(const LgkConsts* c, CtrlArgs& a) {
  a.action_scale = c->action_scale;
  a.sim_dt = c->sim_dt;
  a.D = c->num_dof;
  a.control_type = c->control_type;
  for (int d = 0; d < LGK_MAX_DOF; ++d) {
    a.def[d] = c->default_dof_pos[d]; a.lo[d] = c->dof_pos_lo[d]; a.hi[d] = c->dof_pos_hi[d];
    a.kp[d] = c->p_gains[d]; a.kd[d] = c->d_gains[d]; a.tl[d] = c->torque_limits[d];
  }
}

extern "C" int lgk_compute_poses_impl(const LgkConsts* c, const float* actions, float* out, cudaStream_t st) {
  CtrlArgs a;
  fill_ctrl(c, a);
  const long long n = (long long)c->num_envs * c->num_dof;
  lgk_poses_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(a, actions, out, n);
  return (int)cudaGetLastError();
}

// clip_actions (:85-86) and _compute_poses (:394-397) in one pass over the actions: 4 joints per thread
__global__ void lgk_clip_poses_kernel(const __grid_constant__ CtrlArgs a, const float* __restrict__ in,
                                      float* __restrict__ act_out, float* __restrict__ poses_out, long long n4,
                                      float lim) {
  pdl_launch_dependents();   // the step kernel behind this one may take its SM slots now (it waits on entry)
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n4) return;
  const int d = (int)((unsigned)(i % (unsigned)(a.D >> 2)) << 2);   // D % 4 == 0: a quad never straddles two envs
  float4 v = __ldg(reinterpret_cast<const float4*>(in) + i);
  v.x = clampf(v.x, -lim, lim); v.y = clampf(v.y, -lim, lim);
  v.z = clampf(v.z, -lim, lim); v.w = clampf(v.w, -lim, lim);
  reinterpret_cast<float4*>(act_out)[i] = v;
  float4 p;
  p.x = clampf(add_rn(mul_rn(v.x, a.action_scale), a.def[d]), a.lo[d], a.hi[d]);
  p.y = clampf(add_rn(mul_rn(v.y, a.action_scale), a.def[d + 1]), a.lo[d + 1], a.hi[d + 1]);
  p.z = clampf(add_rn(mul_rn(v.z, a.action_scale), a.def[d + 2]), a.lo[d + 2], a.hi[d + 2]);
  p.w = clampf(add_rn(mul_rn(v.w, a.action_scale), a.def[d + 3]), a.lo[d + 3], a.hi[d + 3]);
  reinterpret_cast<float4*>(poses_out)[i] = p;
}
extern "C" int lgk_clip_actions_poses_impl(const LgkConsts* c, const float* in, float* act_out, float* poses_out,
                                           cudaStream_t st) {
  const long long n = (long long)c->num_envs * c->num_dof;
  const bool vec = (c->num_dof % 4 == 0) && ((((uintptr_t)in | (uintptr_t)act_out | (uintptr_t)poses_out) & 15) == 0);
  if (!vec) {   // odd joint counts / unaligned views: the two separate kernels
    const int rc = lgk_clip_actions_impl(in, act_out, n, c->clip_actions, st);
    return rc != 0 ? rc : lgk_compute_poses_impl(c, act_out, poses_out, st);
  }
  CtrlArgs a;
  fill_ctrl(c, a);
  const long long n4 = n >> 2;
  lgk_clip_poses_kernel<<<(unsigned)((n4 + 255) / 256), 256, 0, st>>>(a, in, act_out, poses_out, n4, c->clip_actions);
  return (int)cudaGetLastError();
}

extern "C" int lgk_compute_torques_impl(const LgkConsts* c, const float* actions, const float* dof_state,
                                        const float* last_dof_vel, float* out, cudaStream_t st) {
  CtrlArgs a;
  fill_ctrl(c, a);
  const long long n = (long long)c->num_envs * c->num_dof;
  lgk_torques_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(a, actions, dof_state, last_dof_vel, out, n);
  return (int)cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// Go1 / Aliengo actuator net (go1.py:79-107)
#define ACT_LEGS 4
#define ACT_LEG_DOF 3
#define ACT_HIST 5
#define ACT_DOF (ACT_LEGS * ACT_LEG_DOF)
#define ACT_IN (2 * ACT_LEG_DOF * ACT_HIST)   // 30
#define ACT_HID 128
#include "lgk_umma_prep.cuh"   // after the ACT_* constants above (it only defines what is missing)

// history shift + pack: one thread per (env, joint).  A thread owns 5 + 5 history floats and 10 floats of
// model_ins, 20 / 40 B apart from its neighbour's: the CTA's slices of the three arrays are contiguous, so they
// move between HBM and shared memory as float4 and the threads do their stride-5 / stride-10 work there.
#define HIST_THREADS 256
__global__ void __launch_bounds__(HIST_THREADS)
lgk_act_hist_kernel(int N, const float* __restrict__ actions, const float* __restrict__ dof_state,
                    float* __restrict__ pe_hist, float* __restrict__ v_hist, float* __restrict__ model_ins,
                    const float* __restrict__ norm, int hist_ctas, const float* __restrict__ weights,
                    float* __restrict__ img) {
  pdl_launch_dependents();   // the MLP kernel behind this one may set itself up on free SMs (it waits before reading)
  if ((int)blockIdx.x >= hist_ctas) {
    // CTAs past the history work build the tcgen05 MLP's weight images (lgk_umma_prep.cuh) under it
    lgk_umma_prep_element(((int)blockIdx.x - hist_ctas) * HIST_THREADS + (int)threadIdx.x, weights, img);
    return;
  }
  __shared__ __align__(16) float s_pe[HIST_THREADS * ACT_HIST];
  __shared__ __align__(16) float s_v[HIST_THREADS * ACT_HIST];
  __shared__ __align__(16) float s_mi[HIST_THREADS * 2 * ACT_HIST];
  const int total = N * ACT_DOF;                       // a multiple of 4, and so is every CTA's first index
  const int i0 = blockIdx.x * HIST_THREADS, tid = threadIdx.x;
  const int cnt = min(HIST_THREADS, total - i0);
  const int nv4 = cnt * ACT_HIST / 4;                  // float4s of this CTA's history slices
  const float4* pe4 = reinterpret_cast<const float4*>(pe_hist + (size_t)i0 * ACT_HIST);
  const float4* v4 = reinterpret_cast<const float4*>(v_hist + (size_t)i0 * ACT_HIST);
  for (int t = tid; t < nv4; t += HIST_THREADS) {
    reinterpret_cast<float4*>(s_pe)[t] = pe4[t];
    reinterpret_cast<float4*>(s_v)[t] = v4[t];
  }
  __syncthreads();
  const int i = i0 + tid;
  if (tid < cnt) {
    const int j = i % ACT_DOF;
    const float2 s = __ldg(reinterpret_cast<const float2*>(dof_state) + i);
    // pos_err uses the RAW action, not the scaled one (go1.py:81)
    const float pe = __fdiv_rn(sub_rn(sub_rn(__ldg(actions + i), s.x), norm[j]), norm[ACT_DOF + j]);
    const float vs = __fdiv_rn(sub_rn(s.y, norm[2 * ACT_DOF + j]), norm[3 * ACT_DOF + j]);
    float* ph = s_pe + tid * ACT_HIST;
    float* vh = s_v + tid * ACT_HIST;
    float* mi = s_mi + tid * 2 * ACT_HIST;             // model_ins[env][j][0..4] = pos_err window, [5..9] = velocity window
#pragma unroll
    for (int k = 0; k < ACT_HIST; ++k) {
      const float a = (k < ACT_HIST - 1) ? ph[k + 1] : pe;   // drop the oldest, append the newest
      const float b = (k < ACT_HIST - 1) ? vh[k + 1] : vs;
      ph[k] = a;
      vh[k] = b;
      mi[k] = a;
      mi[ACT_HIST + k] = b;
    }
  }
  __syncthreads();
  float4* pe4o = reinterpret_cast<float4*>(pe_hist + (size_t)i0 * ACT_HIST);
  float4* v4o = reinterpret_cast<float4*>(v_hist + (size_t)i0 * ACT_HIST);
  float4* mi4o = reinterpret_cast<float4*>(model_ins + (size_t)i0 * 2 * ACT_HIST);
  for (int t = tid; t < nv4; t += HIST_THREADS) {
    pe4o[t] = reinterpret_cast<const float4*>(s_pe)[t];
    v4o[t] = reinterpret_cast<const float4*>(s_v)[t];
  }
  for (int t = tid; t < 2 * nv4; t += HIST_THREADS) mi4o[t] = reinterpret_cast<const float4*>(s_mi)[t];
}

// MLP 30 -> 128 -> 128 -> 128 -> 3 (tanh), fp32 SIMT; weights resident in shared memory,
// persistent CTAs over tiles of 64 leg-rows.
#define MLP_ROWS 64
#define MLP_THREADS 256
// row strides padded by 4 floats: the transposing stores (lanes run over k) then hit 8 banks instead of 1
#define MLP_WS (ACT_HID + 4)
#define MLP_AS (MLP_ROWS + 4)
struct MlpSmem {
  float w0[32][MLP_WS];         // [k][n], k padded 30 -> 32
  float w1[ACT_HID][MLP_WS];    // [k][n]
  float w2[ACT_HID][MLP_WS];
  float w3[ACT_LEG_DOF][ACT_HID];
  float b0[ACT_HID], b1[ACT_HID], b2[ACT_HID], b3[4];
  float act[2][ACT_HID][MLP_AS];   // [k][row]
};

template <int K>
__device__ __forceinline__ void mlp_layer(const float (*__restrict__ wt)[MLP_WS], const float* __restrict__ bias,
                                          const float (*__restrict__ in)[MLP_AS], float (*__restrict__ out)[MLP_AS],
                                          int tr, int tn) {
  float acc[4][8];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
#pragma unroll 4
  for (int k = 0; k < K; ++k) {
    const float4 a4 = *reinterpret_cast<const float4*>(&in[k][tr * 4]);
    const float4 wa = *reinterpret_cast<const float4*>(&wt[k][tn * 4]);
    const float4 wb = *reinterpret_cast<const float4*>(&wt[k][64 + tn * 4]);
    const float av[4] = {a4.x, a4.y, a4.z, a4.w};
    const float wv[8] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w};
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(av[i], wv[j], acc[i][j]);
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int n = (j < 4) ? tn * 4 + j : 64 + tn * 4 + (j - 4);
    const float bb = bias[n];
    float4 o;
    o.x = tanh_fast(acc[0][j] + bb); o.y = tanh_fast(acc[1][j] + bb);
    o.z = tanh_fast(acc[2][j] + bb); o.w = tanh_fast(acc[3][j] + bb);
    *reinterpret_cast<float4*>(&out[n][tr * 4]) = o;
  }
}

__global__ void __launch_bounds__(MLP_THREADS, 1)
lgk_act_mlp_kernel(int rows_total, const float* __restrict__ model_ins, const float* __restrict__ weights,
                   const float* __restrict__ norm, float* __restrict__ dvel) {
  extern __shared__ __align__(16) unsigned char raw[];
  MlpSmem& s = *reinterpret_cast<MlpSmem*>(raw);
  const int tid = threadIdx.x;
  // weights blob: W0[128,30] b0 W1[128,128] b1 W2[128,128] b2 W3[3,128] b3   (torch Linear: [out,in])
  const float* W0 = weights;
  const float* B0 = W0 + ACT_HID * ACT_IN;
  const float* W1 = B0 + ACT_HID;
  const float* B1 = W1 + ACT_HID * ACT_HID;
  const float* W2 = B1 + ACT_HID;
  const float* B2 = W2 + ACT_HID * ACT_HID;
  const float* W3 = B2 + ACT_HID;
  const float* B3 = W3 + ACT_LEG_DOF * ACT_HID;
  for (int i = tid; i < 32 * ACT_HID; i += MLP_THREADS) {
    const int k = i / ACT_HID, n = i % ACT_HID;
    s.w0[k][n] = (k < ACT_IN) ? W0[n * ACT_IN + k] : 0.f;
  }
  for (int i = tid; i < ACT_HID * ACT_HID; i += MLP_THREADS) {
    const int n = i / ACT_HID, k = i % ACT_HID;   // coalesced global read, transposed store
    s.w1[k][n] = W1[i];
    s.w2[k][n] = W2[i];
  }
  for (int i = tid; i < ACT_LEG_DOF * ACT_HID; i += MLP_THREADS) s.w3[i / ACT_HID][i % ACT_HID] = W3[i];
  if (tid < ACT_HID) { s.b0[tid] = B0[tid]; s.b1[tid] = B1[tid]; s.b2[tid] = B2[tid]; }
  if (tid < ACT_LEG_DOF) s.b3[tid] = B3[tid];
  __syncthreads();

  const int tr = tid % 16, tn = tid / 16;
  const int num_tiles = (rows_total + MLP_ROWS - 1) / MLP_ROWS;
  for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
    const int row0 = tile * MLP_ROWS;
    // stage inputs: act[0][k][row] = model_ins[row0+row][k]
    for (int i = tid; i < MLP_ROWS * 32; i += MLP_THREADS) {
      const int r = i / 32, k = i % 32;
      float v = 0.f;
      if (k < ACT_IN && row0 + r < rows_total) v = __ldg(model_ins + (size_t)(row0 + r) * ACT_IN + k);
      s.act[0][k][r] = v;
    }
    __syncthreads();
    mlp_layer<32>(s.w0, s.b0, s.act[0], s.act[1], tr, tn);
    __syncthreads();
    mlp_layer<ACT_HID>(s.w1, s.b1, s.act[1], s.act[0], tr, tn);
    __syncthreads();
    mlp_layer<ACT_HID>(s.w2, s.b2, s.act[0], s.act[1], tr, tn);
    __syncthreads();
    if (tid < MLP_ROWS * ACT_LEG_DOF) {
      const int r = tid % MLP_ROWS, o = tid / MLP_ROWS;
      float acc = 0.f;
#pragma unroll 8
      for (int k = 0; k < ACT_HID; ++k) acc = fmaf(s.act[1][k][r], s.w3[o][k], acc);
      const int g = row0 + r;
      if (g < rows_total) {
        const int leg = g % ACT_LEGS;
        // dVel = net(x) * vel_std   (go1.py:102-105)
        dvel[(size_t)g * ACT_LEG_DOF + o] = (acc + s.b3[o]) * norm[3 * ACT_DOF + leg * ACT_LEG_DOF + o];
      }
    }
    __syncthreads();
  }
}

extern "C" int lgk_act_mlp_tc_impl(int rows, const float* model_ins, const float* weights, const float* norm,
                                   float* dvel, cudaStream_t st);

extern "C" int lgk_act_mlp_umma_impl(int rows, const float* model_ins, const float* weights, const unsigned char* images,
                                     const float* norm, float* dvel, cudaStream_t st);
extern "C" int lgk_umma_setup(const float* weights, cudaStream_t st, bool refresh);

extern "C" long long lgk_actuator_workspace_bytes_impl() { return (long long)um::G_TOTAL; }

extern "C" int lgk_actuator_advance_impl(int N, const float* actions, const float* dof_state, float* pe_hist,
                                         float* v_hist, float* model_ins, const float* weights, const float* norm,
                                         float* dvel, void* workspace, cudaStream_t st) {
  // LGK_MLP_KERNEL picks the MLP kernel (read per call, cheap, so that one process can time / test them all):
  //   "umma" tcgen05 3xTF32, activations + accumulator in tensor memory (lgk_mlp_umma.cu; the default)
  //   "mma"  warp-level 3xTF32 mma.sync (lgk_mlp_tc.cu)      "ffma" the fp32 FFMA kernel below
  // LGK_MLP_FFMA=1 is the older spelling of "ffma".
  const char* e = getenv("LGK_MLP_FFMA");
  const char* k = getenv("LGK_MLP_KERNEL");
  const bool ffma = (e && e[0] == '1') || (k && k[0] == 'f');
  const bool mma = !ffma && k && k[0] == 'm';
  const bool umma = !ffma && !mma;
  float* img = static_cast<float*>(workspace);
  if (umma) {
    if (img == nullptr) return LGK_ERR_NULL;   // the tcgen05 kernel's weight images live in the caller's workspace
    // zero envs = the set-up call: the host copy of the biases / output layer this launch carries in its parameters
    // is (re)read from the device blob (a synchronous copy: not inside a stream capture)
    const int rc = lgk_umma_setup(weights, st, /*refresh=*/N <= 0);
    if (rc) return rc;
  }
  if (N <= 0) return 0;
  const int n = N * ACT_DOF;
  const int hist_ctas = (n + HIST_THREADS - 1) / HIST_THREADS;
  const int prep_ctas = umma ? um::PREP_ELEMENTS / HIST_THREADS : 0;
  lgk_act_hist_kernel<<<hist_ctas + prep_ctas, HIST_THREADS, 0, st>>>(N, actions, dof_state, pe_hist, v_hist, model_ins, norm,
                                                                     hist_ctas, weights, img);
  LGK_CUDA_RETURN_IF(cudaGetLastError());
  const int rows = N * ACT_LEGS;
  if (mma) return lgk_act_mlp_tc_impl(rows, model_ins, weights, norm, dvel, st);
  if (umma) return lgk_act_mlp_umma_impl(rows, model_ins, weights, static_cast<const unsigned char*>(workspace), norm, dvel, st);
  const int tiles = (rows + MLP_ROWS - 1) / MLP_ROWS;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int grid = tiles < sms ? tiles : sms;
  const int smem = (int)sizeof(MlpSmem);
  LGK_CUDA_RETURN_IF(cudaFuncSetAttribute(lgk_act_mlp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  lgk_act_mlp_kernel<<<grid, MLP_THREADS, smem, st>>>(rows, model_ins, weights, norm, dvel);
  return (int)cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// Allocation-free actuator substep (lgk_actuator_advance_static, include/lgk.h).  The [N,120] model_ins scratch
// of the history kernel / MLP pair lives in this file-scope device array instead of a caller buffer, and the MLP
// is the fp32 FFMA kernel above, which needs no weight images: the call makes no cudaMalloc / cudaFree and uses no
// allocating routine.  Every GPU that loads the library holds the array (1 966 080 B) until the process exits.
// All calls on one device share it: calls that overlap on the GPU (two streams) race.  Nothing in it has to start
// at zero -- the history kernel rewrites rows [0, 4N) before the MLP reads exactly those rows.
static __device__ __align__(256) float g_static_model_ins[(size_t)LGK_STATIC_MAX_ENVS * ACT_DOF * 2 * ACT_HIST];

// In host code the array's name is only a handle: the device address comes from cudaGetSymbolAddress, asked once
// per device with that device current.
#define LGK_STATIC_MAX_DEVICES 64
static int static_model_ins_ptr(float** out) {
  static float* ptr[LGK_STATIC_MAX_DEVICES];   // idempotent: a race stores the same address twice
  int dev = 0;
  LGK_CUDA_RETURN_IF(cudaGetDevice(&dev));
  if (dev < 0 || dev >= LGK_STATIC_MAX_DEVICES) return (int)cudaErrorInvalidDevice;
  if (!ptr[dev]) {
    float* p = nullptr;
    LGK_CUDA_RETURN_IF(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_static_model_ins));
    if (!p) return (int)cudaErrorInvalidSymbol;
    ptr[dev] = p;
  }
  *out = ptr[dev];
  return 0;
}

extern "C" int lgk_actuator_advance_static_impl(int N, const float* actions, const float* dof_state, float* pe_hist,
                                                float* v_hist, const float* weights, const float* norm, float* dvel,
                                                cudaStream_t st) {
  float* model_ins = nullptr;
  const int rc = static_model_ins_ptr(&model_ins);
  if (rc) return rc;
  if (N <= 0) return 0;   // set-up call: the address is known from here on
  const int n = N * ACT_DOF;
  const int hist_ctas = (n + HIST_THREADS - 1) / HIST_THREADS;
  lgk_act_hist_kernel<<<hist_ctas, HIST_THREADS, 0, st>>>(N, actions, dof_state, pe_hist, v_hist, model_ins, norm,
                                                          hist_ctas, weights, nullptr);
  LGK_CUDA_RETURN_IF(cudaGetLastError());
  const int rows = N * ACT_LEGS;
  const int tiles = (rows + MLP_ROWS - 1) / MLP_ROWS;
  int dev = 0, sms = 148;
  LGK_CUDA_RETURN_IF(cudaGetDevice(&dev));
  LGK_CUDA_RETURN_IF(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int grid = tiles < sms ? tiles : sms;
  const int smem = (int)sizeof(MlpSmem);
  LGK_CUDA_RETURN_IF(cudaFuncSetAttribute(lgk_act_mlp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  lgk_act_mlp_kernel<<<grid, MLP_THREADS, smem, st>>>(rows, model_ins, weights, norm, dvel);
  return (int)cudaGetLastError();
}

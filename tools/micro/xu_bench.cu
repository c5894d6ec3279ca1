// Throughput of the conversion / special-function instructions the step kernel leans on (per SM per clock).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/xu_bench tools/micro/xu_bench.cu
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 2048
#define CHAINS 8
template <int OP>
__global__ void k(float* out, float seed) {
  float x[CHAINS]; int y[CHAINS];
#pragma unroll
  for (int c = 0; c < CHAINS; ++c) { x[c] = seed + threadIdx.x * 0.37f + c; y[c] = threadIdx.x * 7 + c; }
  for (int i = 0; i < ITERS; ++i) {
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) {
      if (OP == 0) { asm volatile("cvt.rzi.s32.f32 %0, %1;" : "=r"(y[c]) : "f"(x[c])); x[c] = __int_as_float(y[c] | 0x3f800000); }   // F2I + LOP
      if (OP == 1) { asm volatile("cvt.rn.f32.s32 %0, %1;" : "=f"(x[c]) : "r"(y[c])); y[c] = __float_as_int(x[c]) ^ i; }             // I2F + LOP
      if (OP == 2) { unsigned short h = (unsigned short)y[c]; asm volatile("cvt.rn.f32.u16 %0, %1;" : "=f"(x[c]) : "h"(h)); y[c] = __float_as_int(x[c]) ^ i; }
      if (OP == 3) { asm volatile("rcp.approx.ftz.f32 %0, %1;" : "=f"(x[c]) : "f"(x[c])); }
      if (OP == 4) { asm volatile("fma.rn.f32 %0, %0, %1, %1;" : "+f"(x[c]) : "f"(seed)); }
      if (OP == 5) { y[c] = __float_as_int(x[c]) ^ (y[c] + i); x[c] = __int_as_float(y[c] | 0x3f800000); }                        // the LOP overhead alone
      if (OP == 6) { short h = (short)y[c]; asm volatile("cvt.rn.f32.s16 %0, %1;" : "=f"(x[c]) : "h"(h)); y[c] = __float_as_int(x[c]) ^ i; }
    }
  }
  float s = 0; for (int c = 0; c < CHAINS; ++c) s += x[c] + y[c];
  if (s == 12345.678f) out[0] = s;
}
template <int OP> void run(const char* name, int sms, float mhz) {
  float* d; cudaMalloc(&d, 4);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  const int blocks = sms * 8, threads = 256;
  k<OP><<<blocks, threads>>>(d, 1.5f); cudaDeviceSynchronize();
  cudaEventRecord(a); k<OP><<<blocks, threads>>>(d, 1.5f); cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  const double ops = (double)blocks * threads * ITERS * CHAINS;
  printf("%-28s %8.3f ms  %7.1f thread-ops/clk/SM (at %.0f MHz)\n", name, ms, ops / (ms * 1e-3) / (mhz * 1e6) / sms, mhz);
  cudaFree(d);
}
int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  int khz; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  const float mhz = khz / 1000.f; const int sms = p.multiProcessorCount;
  run<5>("LOP pair alone", sms, mhz);
  run<4>("FFMA", sms, mhz);
  run<0>("F2I.TRUNC (+LOP)", sms, mhz);
  run<1>("I2F.S32 (+LOP)", sms, mhz);
  run<2>("I2F.U16 (+LOP)", sms, mhz);
  run<6>("I2F.S16 (+LOP)", sms, mhz);
  run<3>("MUFU.RCP", sms, mhz);
  return 0;
}

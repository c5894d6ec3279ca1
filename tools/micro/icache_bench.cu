// Does a loop body larger than the L0 instruction cache cap the issue rate when many warps run it?
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/icache_bench tools/micro/icache_bench.cu
#include <cstdio>
#include <cuda_runtime.h>
template <int BODY>   // BODY = FFMAs per loop iteration (straight-line, 8 independent chains)
__global__ void __launch_bounds__(128) k(float* out, float s, int iters) {
  float x[8];
#pragma unroll
  for (int c = 0; c < 8; ++c) x[c] = s + threadIdx.x + c;
#pragma unroll 1
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int j = 0; j < BODY; ++j) x[j & 7] = fmaf(x[j & 7], s, x[(j + 3) & 7]);
  }
  float t = 0; for (int c = 0; c < 8; ++c) t += x[c];
  if (t == 1234.5f) out[0] = t;
}
template <int BODY> void run(int sms, float mhz, int ctas_per_sm) {
  float* d; cudaMalloc(&d, 4);
  const long total = 1 << 20;           // FFMAs per thread
  const int iters = (int)(total / BODY);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  k<BODY><<<sms * ctas_per_sm, 128>>>(d, 1.0001f, iters); cudaDeviceSynchronize();
  cudaEventRecord(a); k<BODY><<<sms * ctas_per_sm, 128>>>(d, 1.0001f, iters); cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  const double winst = (double)sms * ctas_per_sm * 4 * iters * BODY;   // warp instructions
  printf("body %5d instrs (%6.1f KB)  %d CTAs/SM: %7.3f ms  IPC/SMSP %.3f\n", BODY, BODY * 16 / 1024.0, ctas_per_sm, ms,
         winst / (ms * 1e-3 * mhz * 1e6) / (sms * 4));
  cudaFree(d);
}
int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  int khz; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  const float mhz = khz / 1000.f; const int sms = p.multiProcessorCount;
  for (int c : {1, 6}) {
    run<128>(sms, mhz, c); run<256>(sms, mhz, c); run<384>(sms, mhz, c); run<512>(sms, mhz, c);
    run<1024>(sms, mhz, c); run<2048>(sms, mhz, c); run<4096>(sms, mhz, c);
  }
  return 0;
}

// Throughput of the special-function unit per SM on this GPU: independent chains of rcp / ex2 / tanh / rsqrt approximations
// (dev tool; build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_bin/mufu_rate tools/mufu_rate.cu).
#include <cstdio>
#include <cuda_runtime.h>

template <int OP>
__device__ __forceinline__ float op(float x) {
  float r;
  if (OP == 0) asm volatile("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  else if (OP == 1) asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  else if (OP == 2) asm volatile("tanh.approx.f32 %0, %1;" : "=f"(r) : "f"(x));
  else if (OP == 3) asm volatile("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  else if (OP == 4) asm volatile("fma.rn.f32 %0, %1, %1, %1;" : "=f"(r) : "f"(x));
  else asm volatile("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

template <int OP>
__global__ void k(float* out, int iters) {
  float v[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = 1.0f + 0.001f * (threadIdx.x + i);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = fmaf(op<OP>(v[i]), 0.999f, 0.5f);     // the fma keeps ptxas from folding rcp(rcp(x))
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += v[i];
  if (s == 123.456f) out[0] = s;
}

template <int OP>
void run(const char* name, int sms, double ghz) {
  float* d;
  cudaMalloc(&d, 4);
  const int iters = 4096, threads = 1024, blocks = sms * 2;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<OP><<<blocks, threads>>>(d, iters);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  k<OP><<<blocks, threads>>>(d, iters);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  const double ops = (double)blocks * threads * iters * 8;
  printf("%-6s %8.1f Gop/s  = %5.1f ops/clk/SM at %.3f GHz\n", name, ops / ms * 1e-6, ops / (ms * 1e-3) / (sms * ghz * 1e9), ghz);
  cudaFree(d);
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  int khz = 0;
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  const double ghz = khz * 1e-6;
  printf("%s, %d SMs, %.3f GHz nominal\n", p.name, p.multiProcessorCount, ghz);
  run<4>("fma", p.multiProcessorCount, ghz);
  run<0>("rcp", p.multiProcessorCount, ghz);
  run<1>("ex2", p.multiProcessorCount, ghz);
  run<2>("tanh", p.multiProcessorCount, ghz);
  run<3>("rsqrt", p.multiProcessorCount, ghz);
  run<5>("lg2", p.multiProcessorCount, ghz);
  return 0;
}

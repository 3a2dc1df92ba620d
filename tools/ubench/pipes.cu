// Micro-benchmark (scratch evidence for DESIGN.md): issue rate of scalar FFMA vs packed fma.rn.f32x2 (FFMA2),
// MUFU.EX2, and a 50/50 FFMA2 + MUFU mix, on one B200.  nvcc -gencode arch=compute_100a,code=sm_100a -O3
#include <cstdio>
#include <cuda_runtime.h>
constexpr int ITER = 4096, UNROLL = 8;

__global__ void k_ffma(float* out, float a, float b) {
  float x[UNROLL];
#pragma unroll
  for (int i = 0; i < UNROLL; ++i) x[i] = threadIdx.x + i;
  for (int it = 0; it < ITER; ++it) {
#pragma unroll
    for (int i = 0; i < UNROLL; ++i) x[i] = fmaf(x[i], a, b);
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < UNROLL; ++i) s += x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__device__ __forceinline__ unsigned long long fma2(unsigned long long x, unsigned long long a, unsigned long long b) {
  unsigned long long r;
  asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(x), "l"(a), "l"(b));
  return r;
}
__global__ void k_ffma2(float* out, float a, float b) {
  unsigned long long x[UNROLL], aa, bb;
  asm("mov.b64 %0, {%1, %1};" : "=l"(aa) : "f"(a));
  asm("mov.b64 %0, {%1, %1};" : "=l"(bb) : "f"(b));
#pragma unroll
  for (int i = 0; i < UNROLL; ++i) { float v = threadIdx.x + i; asm("mov.b64 %0, {%1, %1};" : "=l"(x[i]) : "f"(v)); }
  for (int it = 0; it < ITER; ++it) {
#pragma unroll
    for (int i = 0; i < UNROLL; ++i) x[i] = fma2(x[i], aa, bb);
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < UNROLL; ++i) { float lo, hi; asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(x[i])); s += lo + hi; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_mufu(float* out, float a) {
  float x[UNROLL];
#pragma unroll
  for (int i = 0; i < UNROLL; ++i) x[i] = (threadIdx.x + i) * 1e-3f;
  for (int it = 0; it < ITER; ++it) {
#pragma unroll
    for (int i = 0; i < UNROLL; ++i) asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(x[i]) : "f"(x[i]));
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < UNROLL; ++i) s += x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s + a;
}
// 6 FFMA (or 3 FFMA2) per MUFU: close to the gradient kernel's mix
template <bool PACKED> __global__ void k_mix(float* out, float a, float b) {
  float m[2] = {threadIdx.x * 1e-3f, threadIdx.x * 2e-3f};
  float x[12];
  unsigned long long y[6], aa, bb;
  asm("mov.b64 %0, {%1, %1};" : "=l"(aa) : "f"(a));
  asm("mov.b64 %0, {%1, %1};" : "=l"(bb) : "f"(b));
#pragma unroll
  for (int i = 0; i < 12; ++i) x[i] = threadIdx.x + i;
#pragma unroll
  for (int i = 0; i < 6; ++i) asm("mov.b64 %0, {%1, %2};" : "=l"(y[i]) : "f"(x[2 * i]), "f"(x[2 * i + 1]));
  for (int it = 0; it < ITER; ++it) {
#pragma unroll
    for (int j = 0; j < 2; ++j) asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(m[j]) : "f"(m[j]));
    if (PACKED) {
#pragma unroll
      for (int i = 0; i < 6; ++i) y[i] = fma2(y[i], aa, bb);
    } else {
#pragma unroll
      for (int i = 0; i < 12; ++i) x[i] = fmaf(x[i], a, b);
    }
  }
  float s = m[0] + m[1];
#pragma unroll
  for (int i = 0; i < 12; ++i) s += x[i];
#pragma unroll
  for (int i = 0; i < 6; ++i) { float lo, hi; asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(y[i])); s += lo + hi; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <class F> float timeit(F f) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  f(); cudaDeviceSynchronize();
  cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b); return ms;
}
int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  const int sms = p.multiProcessorCount, blocks = sms * 8, threads = 256;
  float* out; cudaMalloc(&out, sizeof(float) * blocks * threads);
  int khz; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  const double clk = khz * 1e3, warps = double(blocks) * threads / 32;
  auto rep = [&](const char* name, float ms, double warp_instr_per_iter, double flop_per_thread_iter) {
    const double wi = warps * ITER * warp_instr_per_iter, cyc = ms * 1e-3 * clk;
    printf("%-28s %8.3f ms  %6.2f warp-instr/clk/SM  %7.2f Tlane-op/s\n", name, ms, wi / cyc / sms, warps * 32 * ITER * flop_per_thread_iter / (ms * 1e-3) / 1e12);
  };
  printf("device %s, %d SMs, clock %.0f MHz (nominal)\n", p.name, sms, clk / 1e6);
  rep("FFMA x8", timeit([&] { k_ffma<<<blocks, threads>>>(out, 1.0001f, 0.5f); }), UNROLL, UNROLL);
  rep("FFMA2 x8 (fma.rn.f32x2)", timeit([&] { k_ffma2<<<blocks, threads>>>(out, 1.0001f, 0.5f); }), UNROLL, 2 * UNROLL);
  rep("MUFU.EX2 x8", timeit([&] { k_mufu<<<blocks, threads>>>(out, 1.f); }), UNROLL, UNROLL);
  rep("2 MUFU + 12 FFMA", timeit([&] { k_mix<false><<<blocks, threads>>>(out, 1.0001f, 0.5f); }), 14, 14);
  rep("2 MUFU + 6 FFMA2", timeit([&] { k_mix<true><<<blocks, threads>>>(out, 1.0001f, 0.5f); }), 8, 14);
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}

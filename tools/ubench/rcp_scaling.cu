// Does MUFU.RCP commute with a power of two?  rcp.approx.ftz.f32(2 x) == 0.5f * rcp.approx.ftz.f32(x), bit for bit, for
// every mantissa and a spread of exponents -- what hgf_math.cuh::recip_twice relies on to share one reciprocal between a
// quotient by x and a quotient by 2 x.    nvcc -arch=sm_100a -o rcp_scaling tools/ubench/rcp_scaling.cu && ./rcp_scaling
#include <cstdio>
#include <cuda_runtime.h>
__device__ float rcp(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
// every mantissa, both signs, every normal exponent e: x = +-1.m * 2^(e-127)
__global__ void check(unsigned long long* bad) {  // bad[e]
  for (unsigned m = blockIdx.x * blockDim.x + threadIdx.x; m < (1u << 23); m += gridDim.x * blockDim.x)
    for (int e = 1; e < 254; ++e)
      for (unsigned s = 0; s < 2; ++s) {
        const float x = __uint_as_float((s << 31) | (static_cast<unsigned>(e) << 23) | m);
        const float a = rcp(2.0f * x), h = 0.5f * rcp(x);
        if (__float_as_uint(a) != __float_as_uint(h) && !(a != a && h != h)) atomicAdd(bad + e, 1ull);
      }
}
int main() {
  unsigned long long *d, h[256] = {};
  cudaMalloc(&d, sizeof h); cudaMemset(d, 0, sizeof h);
  check<<<592, 256>>>(d);
  cudaMemcpy(h, d, sizeof h, cudaMemcpyDeviceToHost);
  // |x| >= 2^125: rcp(2 x) is subnormal and flushed (ftz) while rcp(x) / 2 keeps it -- far beyond max_posterior_precision (1e10)
  unsigned long long in_range = 0, edge = 0;
  for (int e = 1; e < 254; ++e) {
    if (h[e]) printf("  exponent %d (|x| in [2^%d, 2^%d)): %llu mismatches\n", e, e - 127, e - 126, h[e]);
    (e < 252 ? in_range : edge) += h[e];
  }
  printf("rcp(2x) vs 0.5*rcp(x), 2 x 2^23 x 253 values: %llu mismatches for |x| < 2^125, %llu beyond (%s)\n", in_range, edge,
         cudaGetErrorString(cudaGetLastError()));
  return in_range != 0;
}

// Exponential-family state node with a fixed learning rate (include/pyhgf_cuda.h: hgf_ef_filter).
// One thread per instance, the K sufficient statistics in registers (D is a template parameter), trajectory
// stores coalesced over instances.  The node has no parents, so it is its own (tiny) time loop.
#include <cstdio>

#include "../../include/pyhgf_cuda.h"
#include "hgf_math.cuh"

namespace hgf {

void count_launches(int n);       // hgf_capi.cu
int report_error(const char* msg);  // hgf_capi.cu: sets hgf_last_error, returns 1

template <class B, int D>
__global__ void __launch_bounds__(128) ef_fixed_kernel(hgf_ef_call_t c) {
  constexpr int K = D + D * (D + 1) / 2;
  const long long n = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (n >= c.n_instances) return;
  const long long N = c.n_instances;
  const int T = c.n_steps;
  const long long series = c.input_group_size > 0 ? n / c.input_group_size : 0;
  const B* x = static_cast<const B*>(c.input_data) + series * T * D;
  const uint8_t* obs = c.observed ? c.observed + series * T : nullptr;
  // exponential.py:50-52: xis + (1 / (1 + nus)) * (observation_ss - xis)
  const B* implied = static_cast<const B*>(c.implied);
  B w = implied ? B(0) : b_div(B(1), B(1) + HGF_LDG(static_cast<const B*>(c.nus) + n));
  B xis[K];
#pragma unroll
  for (int k = 0; k < K; ++k) xis[k] = HGF_LDG(static_cast<const B*>(c.xis0) + k);
  B* traj = static_cast<B*>(c.xis_trajectory);
  for (int t = 0; t < T; ++t) {
    B v[D], ss[K];
#pragma unroll
    for (int i = 0; i < D; ++i) v[i] = HGF_LDG(x + static_cast<long long>(t) * D + i);
#pragma unroll
    for (int i = 0; i < D; ++i) ss[i] = v[i];
    int k = D;
#pragma unroll
    for (int r = 0; r < D; ++r)  // jnp.tril_indices order: (0,0), (1,0), (1,1), (2,0), ...
#pragma unroll
      for (int q = 0; q <= r; ++q) ss[k++] = v[r] * v[q];
    const bool seen = obs ? HGF_LDG(obs + t) != 0 : true;
    if (implied) {  // posterior/exponential.py: pe / expectation differential, averaged over the statistics
      const long long plane = static_cast<long long>(K) * T * N, at = static_cast<long long>(t) * N + n;
      B nu = B(0);
#pragma unroll
      for (int j = 0; j < K; ++j) {
        const long long o = static_cast<long long>(j) * T * N + at;
        nu += b_div(HGF_LDG(implied + o) - HGF_LDG(implied + plane + o), HGF_LDG(implied + 2 * plane + o) - HGF_LDG(implied + 3 * plane + o));
      }
      nu = nu / B(K);
      w = b_div(B(1), B(1) + nu);
      if (c.nus_trajectory) static_cast<B*>(c.nus_trajectory)[at] = nu;
    }
#pragma unroll
    for (int j = 0; j < K; ++j) {
      const B next = xis[j] + w * (ss[j] - xis[j]);
      xis[j] = seen ? next : xis[j];  // exponential.py:55-58
      if (traj) traj[(static_cast<long long>(j) * T + t) * N + n] = xis[j];
    }
  }
  if (c.xis_last) {
#pragma unroll
    for (int j = 0; j < K; ++j) static_cast<B*>(c.xis_last)[static_cast<long long>(j) * N + n] = xis[j];
  }
}

template <class B>
static cudaError_t launch_ef(const hgf_ef_call_t& c, cudaStream_t s) {
  const unsigned blocks = static_cast<unsigned>((c.n_instances + 127) / 128);
  switch (c.dimension) {
#define HGF_EF_CASE(D) case D: ef_fixed_kernel<B, D><<<blocks, 128, 0, s>>>(c); break;
    HGF_EF_CASE(1) HGF_EF_CASE(2) HGF_EF_CASE(3) HGF_EF_CASE(4) HGF_EF_CASE(5) HGF_EF_CASE(6) HGF_EF_CASE(7) HGF_EF_CASE(8)
#undef HGF_EF_CASE
    default: return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

}  // namespace hgf

extern "C" int hgf_ef_filter(const hgf_ef_call_t* c, void* stream) {
  using namespace hgf;
  if (!c) return report_error("hgf_ef_filter: null call");
  if (c->dtype != 32 && c->dtype != 64) return report_error("hgf_ef_filter: dtype must be 32 or 64");
  if (c->dimension < 1 || c->dimension > 8) return report_error("hgf_ef_filter: dimension outside [1, 8]");
  if (c->n_instances < 0 || c->n_steps < 0 || c->input_group_size < 0) return report_error("hgf_ef_filter: negative size");
  if (c->n_instances == 0) return 0;
  if ((!c->nus && !c->implied) || !c->xis0 || (c->n_steps > 0 && !c->input_data)) return report_error("hgf_ef_filter: null input");
  const cudaError_t e = c->dtype == 32 ? launch_ef<float>(*c, static_cast<cudaStream_t>(stream)) : launch_ef<double>(*c, static_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) {
    char buf[256];
    snprintf(buf, sizeof buf, "hgf_ef_filter: launch failed: %s", cudaGetErrorString(e));
    return report_error(buf);
  }
  count_launches(1);
  return 0;
}

// Categorical state node (include/pyhgf_cuda.h: hgf_categorical_update).  The implied binary HGFs are ordinary nodes of
// the scan; what is left of the node is a per-step function of their trajectories (posterior/categorical.py:44-89):
// the Dirichlet concentration implied by each step, its KL divergence from the previous step's, and the summed binary
// surprise.  One thread per (step, instance); stores coalesced over instances.  alpha is computed in the call's dtype
// (as the reference does); the Gamma / digamma terms of the divergence are evaluated in double and rounded once.
#include <cmath>
#include <cstdio>

#include "../../include/pyhgf_cuda.h"
#include "hgf_math.cuh"

namespace hgf {

void count_launches(int n);         // hgf_capi.cu
int report_error(const char* msg);  // hgf_capi.cu

// psi(x): reflection for x <= 0, upward recurrence to x >= 10, then the asymptotic series (error < 1e-13 there)
__device__ inline double digamma(double x) {
  double r = 0.0;
  if (x <= 0.0) {
    if (x == floor(x)) return nan("");
    r = -3.14159265358979323846 / tan(3.14159265358979323846 * x);
    x = 1.0 - x;
  }
  while (x < 10.0) {
    r -= 1.0 / x;
    x += 1.0;
  }
  const double f = 1.0 / (x * x);
  return r + log(x) - 0.5 / x - f * (1.0 / 12 - f * (1.0 / 120 - f * (1.0 / 252 - f * (1.0 / 240 - f * (1.0 / 132)))));
}

template <class B>
struct CategoricalStep {
  const B* x;          // this step's observation row [n]
  const B* em;         // expected_mean [n][T][N] at (t, instance): stride T*N between categories
  const B* pm;         // parent mean, same layout
  long long stride;
  // posterior/categorical.py:56-67
  __device__ B updated_mean(int k) const { return B(1) / (B(1) + exp(-HGF_LDG(pm + k * stride))); }
  __device__ B alpha(int k) const {
    const B e = HGF_LDG(em + k * stride);
    const B pe = HGF_LDG(x + k) - e;
    const B delta_xi = updated_mean(k) - e;
    const B nu = pe / delta_xi - B(1);
    return nu * e + B(1);
  }
};

template <class B>
__global__ void __launch_bounds__(128) categorical_kernel(hgf_categorical_call_t c) {
  const long long N = c.n_instances, T = c.n_steps;
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= N * T) return;
  const long long t = i / N, inst = i - t * N;
  const int n = c.n_categories;
  const long long series = c.input_group_size > 0 ? inst / c.input_group_size : 0;
  const B* obs = static_cast<const B*>(c.observations) + (series * T + t) * n;
  CategoricalStep<B> now{obs, static_cast<const B*>(c.expected_mean) + i, static_cast<const B*>(c.parent_mean) + i, T * N};
  CategoricalStep<B> before{obs - n, now.em - N, now.pm - N, T * N};
  const B* alpha0 = static_cast<const B*>(c.alpha0);
  auto previous = [&](int k) { return t > 0 ? before.alpha(k) : HGF_LDG(alpha0 + k); };

  // math.py:337-341 -- KL(Dir(alpha_1) || Dir(alpha_2)), alpha_1 the previous concentration
  B sum1 = B(0), sum2 = B(0), surprise = B(0);
  for (int k = 0; k < n; ++k) {
    const B a2 = now.alpha(k);
    sum1 += previous(k);
    sum2 += a2;
    const B e = HGF_LDG(now.em + k * now.stride);
    surprise += HGF_LDG(obs + k) != B(0) ? -log(e) : -log(B(1) - e);  // binary_surprise, summed over the categories
    if (c.alpha) static_cast<B*>(c.alpha)[k * now.stride + i] = a2;
    if (c.mean) static_cast<B*>(c.mean)[k * now.stride + i] = now.updated_mean(k);
  }
  const double psi_sum = digamma(static_cast<double>(sum1));
  double kl = log(tgamma(static_cast<double>(sum1)) / tgamma(static_cast<double>(sum2)));
  for (int k = 0; k < n; ++k) {
    const double a1 = static_cast<double>(previous(k)), a2 = static_cast<double>(now.alpha(k));
    kl += log(tgamma(a2) / tgamma(a1)) + (a1 - a2) * (digamma(a1) - psi_sum);
  }
  static_cast<B*>(c.kl_divergence)[i] = static_cast<B>(kl);
  static_cast<B*>(c.surprise)[i] = surprise;
}

}  // namespace hgf

extern "C" int hgf_categorical_update(const hgf_categorical_call_t* c, void* stream) {
  using namespace hgf;
  if (!c) return report_error("hgf_categorical_update: null call");
  if (c->dtype != 32 && c->dtype != 64) return report_error("hgf_categorical_update: dtype must be 32 or 64");
  if (c->n_categories < 2 || c->n_categories > HGF_MAX_CATEGORIES) return report_error("hgf_categorical_update: n_categories outside [2, 16]");
  if (c->n_instances < 0 || c->n_steps < 0 || c->input_group_size < 0) return report_error("hgf_categorical_update: negative size");
  const long long work = static_cast<long long>(c->n_instances) * c->n_steps;
  if (work == 0) return 0;
  if (!c->observations || !c->expected_mean || !c->parent_mean || !c->alpha0 || !c->kl_divergence || !c->surprise)
    return report_error("hgf_categorical_update: null input or output");
  const unsigned blocks = static_cast<unsigned>((work + 127) / 128);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (c->dtype == 32) categorical_kernel<float><<<blocks, 128, 0, s>>>(*c);
  else categorical_kernel<double><<<blocks, 128, 0, s>>>(*c);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    char buf[256];
    snprintf(buf, sizeof buf, "hgf_categorical_update: launch failed: %s", cudaGetErrorString(e));
    return report_error(buf);
  }
  count_launches(1);
  return 0;
}

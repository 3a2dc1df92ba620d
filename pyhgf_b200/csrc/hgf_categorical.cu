// Categorical state node (include/pyhgf_cuda.h: hgf_categorical_update).  The implied binary HGFs are ordinary nodes of
// the scan; what is left of the node is a per-step function of their trajectories (posterior/categorical.py:44-89):
// the Dirichlet concentration implied by each step, its KL divergence from the previous step's, and the summed binary
// surprise.  One thread per (chunk of steps, instance); stores coalesced over instances.  alpha is computed in the call's dtype
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

// log Gamma(x) and psi(x) for x > 0 with the work shared: shift x up to >= 10 with ONE product (and its derivative, so
// the recurrence sum 1/x + 1/(x+1) + ... costs one division), then the Stirling / asymptotic series, which share log x
// and 1/x.  Absolute error < 1e-13, a third of the instructions of tgamma + digamma above.
template <bool WITH_PSI>
__device__ inline void lgamma_psi(double a, double& lg, double& psi) {
  double p = 1.0, dp = 0.0, x = a;
  while (x < 10.0) {
    if (WITH_PSI) dp = dp * x + p;
    p *= x;
    x += 1.0;
  }
  const double lx = log(x), r = 1.0 / x, f = r * r;
  lg = (x - 0.5) * lx - x + 0.91893853320467274178 - log(p) +
       r * (1.0 / 12 - f * (1.0 / 360 - f * (1.0 / 1260 - f * (1.0 / 1680 - f * (1.0 / 1188)))));
  if (WITH_PSI) psi = lx - 0.5 * r - f * (1.0 / 12 - f * (1.0 / 120 - f * (1.0 / 252 - f * (1.0 / 240 - f * (1.0 / 132))))) - dp / p;
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

// One thread walks CHUNK consecutive steps of one instance, so that what step t computed for its own concentration
// (alpha_t, log Gamma and psi of every entry and of the sum) is step t+1's "previous" side for free; only the first step
// of a chunk recomputes its predecessor.  Threads of a warp are consecutive instances: loads and stores stay coalesced.
constexpr int CATEGORICAL_CHUNK = 8;

template <class B>
__global__ void __launch_bounds__(128) categorical_kernel(hgf_categorical_call_t c) {
  constexpr int MAXN = HGF_MAX_CATEGORIES;
  const long long N = c.n_instances, T = c.n_steps;
  const long long chunks = (T + CATEGORICAL_CHUNK - 1) / CATEGORICAL_CHUNK;
  const long long id = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (id >= N * chunks) return;
  const long long chunk = id / N, inst = id - chunk * N;
  const int n = c.n_categories;
  const long long series = c.input_group_size > 0 ? inst / c.input_group_size : 0;
  const B* obs0 = static_cast<const B*>(c.observations) + series * T * n;
  const B* em0 = static_cast<const B*>(c.expected_mean) + inst;
  const B* pm0 = static_cast<const B*>(c.parent_mean) + inst;
  const long long stride = T * N;
  const long long t_begin = chunk * CATEGORICAL_CHUNK, t_end = t_begin + CATEGORICAL_CHUNK < T ? t_begin + CATEGORICAL_CHUNK : T;

  // the "previous" side: alpha_{t-1}, log Gamma / psi of its entries and of its sum (valid while prev_positive)
  B a1[MAXN], sum1 = B(0);
  double lg1[MAXN], psi1[MAXN], lg_sum1 = 0.0, psi_sum1 = 0.0;
  bool prev_positive = true;
  {
    CategoricalStep<B> before{obs0 + (t_begin - 1) * n, em0 + (t_begin - 1) * N, pm0 + (t_begin - 1) * N, stride};
    for (int k = 0; k < n; ++k) {
      a1[k] = t_begin > 0 ? before.alpha(k) : HGF_LDG(static_cast<const B*>(c.alpha0) + k);
      sum1 += a1[k];
      prev_positive = prev_positive && a1[k] > B(0);
    }
    prev_positive = prev_positive && sum1 > B(0);
    if (prev_positive) {
      for (int k = 0; k < n; ++k) lgamma_psi<true>(static_cast<double>(a1[k]), lg1[k], psi1[k]);
      lgamma_psi<true>(static_cast<double>(sum1), lg_sum1, psi_sum1);
    }
  }

  for (long long t = t_begin; t < t_end; ++t) {
    const long long i = t * N + inst;
    const B* obs = obs0 + t * n;
    CategoricalStep<B> now{obs, em0 + t * N, pm0 + t * N, stride};
    B a2[MAXN], sum2 = B(0), surprise = B(0);
    bool now_positive = true;
    for (int k = 0; k < n; ++k) {
      a2[k] = now.alpha(k);
      sum2 += a2[k];
      now_positive = now_positive && a2[k] > B(0);
      // binary_surprise(..., clipping=True), summed over the categories (categorical.py:77-82; math.py:248-253 clips the
      // expected mean to [1e-6, 1 - 1e-6] itself, whatever the network's precision_clipping_value is)
      B e = HGF_LDG(now.em + k * stride);
      e = e < B(1e-6) ? B(1e-6) : (e > B(1.0 - 1e-6) ? B(1.0 - 1e-6) : e);
      surprise += HGF_LDG(obs + k) != B(0) ? -log(e) : -log(B(1) - e);
      if (c.alpha) static_cast<B*>(c.alpha)[k * stride + i] = a2[k];
      if (c.mean) static_cast<B*>(c.mean)[k * stride + i] = now.updated_mean(k);
    }
    now_positive = now_positive && sum2 > B(0);
    double lg2[MAXN], psi2[MAXN], lg_sum2 = 0.0, psi_sum2 = 0.0;
    if (now_positive) {
      for (int k = 0; k < n; ++k) lgamma_psi<true>(static_cast<double>(a2[k]), lg2[k], psi2[k]);
      lgamma_psi<true>(static_cast<double>(sum2), lg_sum2, psi_sum2);
    }
    // math.py:337-341 -- KL(Dir(alpha_1) || Dir(alpha_2)), alpha_1 the previous concentration.  Every concentration
    // positive (the usual case): log-Gamma differences; otherwise the reference's own expression, whose Gamma ratios
    // carry the signs (and give NaN where the reference does)
    double kl;
    if (prev_positive && now_positive) {
      kl = lg_sum1 - lg_sum2;
      for (int k = 0; k < n; ++k)
        kl += (lg2[k] - lg1[k]) + (static_cast<double>(a1[k]) - static_cast<double>(a2[k])) * (psi1[k] - psi_sum1);
    } else {
      const double psi_s = digamma(static_cast<double>(sum1));
      kl = log(tgamma(static_cast<double>(sum1)) / tgamma(static_cast<double>(sum2)));
      for (int k = 0; k < n; ++k) {
        const double x1 = static_cast<double>(a1[k]), x2 = static_cast<double>(a2[k]);
        kl += log(tgamma(x2) / tgamma(x1)) + (x1 - x2) * (digamma(x1) - psi_s);
      }
    }
    static_cast<B*>(c.kl_divergence)[i] = static_cast<B>(kl);
    static_cast<B*>(c.surprise)[i] = surprise;
    for (int k = 0; k < n; ++k) {
      a1[k] = a2[k];
      lg1[k] = lg2[k];
      psi1[k] = psi2[k];
    }
    sum1 = sum2, lg_sum1 = lg_sum2, psi_sum1 = psi_sum2, prev_positive = now_positive;
  }
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
  const long long threads = static_cast<long long>(c->n_instances) * ((c->n_steps + hgf::CATEGORICAL_CHUNK - 1) / hgf::CATEGORICAL_CHUNK);
  const unsigned blocks = static_cast<unsigned>((threads + 127) / 128);
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

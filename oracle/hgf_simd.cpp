// TEST / BASELINE INFRASTRUCTURE ONLY -- the instance-vectorised float32 CPU arm of bench.py.
//
// What the reference does on a CPU is `jit(vmap(...))` / `jit(vmap(grad(...)))` compiled by XLA: float32, every primitive
// vectorised ACROSS the instances of the batch, reverse-mode gradient (pyhgf/model/network.py:612-624 under jax.grad).
// hgf_oracle.cpp (the parity checker) is scalar per instance, which flatters a GPU/CPU ratio.  This file is the fair
// denominator: W = 16 (AVX-512) or 8 (AVX2) instances per SIMD vector, glibc's vector math (libmvec: expf, logf, log1pf,
// powf), OpenMP over vectors of instances, and for logp + gradient the checkpointed reverse sweep -- the generated
// straight-line step of pyhgf_b200/revgen.py (csrc/generated/reverse_<topology>.h; it restates the reference's update
// equations operation by operation and cites them) instantiated with F = a GCC vector type.  IEEE division, no FMA
// contraction games: arithmetic is what gcc -O3 makes of it.  It is checked against the scalar oracle in
// tests/test_simd_cpu_arm.py and is never the thing a parity test trusts.
//
// Only tests/, bench.py's cpu_baseline / --impl reference legs may load this library.
#include <immintrin.h>
#include <omp.h>

#include <cstdint>
#include <cstring>
#include <vector>

#define HGF_LIT(c) ::hgf::simd_lit<F>(c)
#include "hgf_math.cuh"

namespace hgf {

#if defined(__AVX512F__)
constexpr int W = 16;
typedef float vf __attribute__((vector_size(64)));
typedef int32_t vi __attribute__((vector_size(64)));
extern "C" __m512 _ZGVeN16v_expf(__m512);
extern "C" __m512 _ZGVeN16v_logf(__m512);
extern "C" __m512 _ZGVeN16v_log1pf(__m512);
extern "C" __m512 _ZGVeN16vv_powf(__m512, __m512);
inline vf v_exp(vf x) { return (vf)_ZGVeN16v_expf((__m512)x); }
inline vf v_log(vf x) { return (vf)_ZGVeN16v_logf((__m512)x); }
inline vf v_log1p(vf x) { return (vf)_ZGVeN16v_log1pf((__m512)x); }
inline vf v_pow(vf x, vf y) { return (vf)_ZGVeN16vv_powf((__m512)x, (__m512)y); }
#else
constexpr int W = 8;
typedef float vf __attribute__((vector_size(32)));
typedef int32_t vi __attribute__((vector_size(32)));
extern "C" __m256 _ZGVdN8v_expf(__m256);
extern "C" __m256 _ZGVdN8v_logf(__m256);
extern "C" __m256 _ZGVdN8v_log1pf(__m256);
extern "C" __m256 _ZGVdN8vv_powf(__m256, __m256);
inline vf v_exp(vf x) { return (vf)_ZGVdN8v_expf((__m256)x); }
inline vf v_log(vf x) { return (vf)_ZGVdN8v_logf((__m256)x); }
inline vf v_log1p(vf x) { return (vf)_ZGVdN8v_log1pf((__m256)x); }
inline vf v_pow(vf x, vf y) { return (vf)_ZGVdN8vv_powf((__m256)x, (__m256)y); }
#endif

inline vf splat(float c) { vf r = {}; return r + c; }
template <class F> inline F simd_lit(double c) { return static_cast<F>(c); }  // a literal adopts the array dtype (1e-128 -> 0.0f)
template <> inline vf simd_lit<vf>(double c) { return splat(static_cast<float>(c)); }

// the primitives the generated step calls (hgf_math.cuh has them for float / double), on vectors of instances
inline vf b_div(vf a, vf b) { return a / b; }
inline vf b_rcp(vf x) { return splat(1.0f) / x; }
inline vf t_rcp(vf x) { return splat(1.0f) / x; }
inline vf exact_rcp(vf x) { return splat(1.0f) / x; }
inline vf exact_quot(vf a, vf b) { return a / b; }
inline vf b_exp(vf x) { return v_exp(x); }
inline vf b_log(vf x) { return v_log(x); }
inline vf exact_pow(vf x, vf y) { return v_pow(x, y); }
inline vf b_pow(vf x, vf y) { return v_pow(x, y); }
inline vf b_abs(vf x) { return (vf)((vi)x & 0x7fffffff); }
inline vi b_isnan(vf x) { return x != x; }
inline vi b_isfinite(vf x) { return b_abs(x) <= splat(3.402823466e+38f); }
template <> inline vf b_nan<vf>() { return splat(__builtin_nanf("")); }
template <> inline vf b_inf<vf>() { return splat(__builtin_huge_valf()); }
template <> inline vf b_log_max<vf>() { return splat(88.72283905206835f); }
// jnp.logaddexp: max + log1p(exp(-|a - b|)); NaN delta -> a + b
template <> inline vf b_logaddexp<vf>(vf a, vf b) {
  vf amax = ((a >= b) || b_isnan(a)) ? a : b;
  amax = b_isnan(b) ? b : amax;
  const vf delta = a - b;
  return b_isnan(delta) ? (a + b) : (amax + v_log1p(v_exp(-b_abs(delta))));
}
// pyhgf/math.py:450-488: w0 = log(z + 1), six Halley steps
template <> inline vf b_lambert_w0<vf>(vf z) {
  vf w = v_log(z + splat(1.0f));
  for (int it = 0; it < 6; ++it) {
    const vf ew = v_exp(w);
    const vf f = w * ew - z;
    const vf f1 = (w + splat(1.0f)) * ew;
    const vf f2 = (w + splat(2.0f)) * ew;
    w = w - (splat(2.0f) * f * f1) / (splat(2.0f) * (f1 * f1) - f * f2);
  }
  return w;
}

}  // namespace hgf

#include <generated/reverse_binary3.h>
#include <generated/reverse_continuous3.h>
#include <generated/reverse_ghgf16.h>

namespace hgf {

struct SimdCall {
  long long n_instances;
  int n_steps;
  long long input_group_size, response_group_size;
  const float* input_data;       // [G][T][kInputs]
  const float* time_steps;       // [T] or null
  const float* response_inputs;  // [G][T] or null
  const float* attributes;       // [kSlots]
  int n_params;
  const int* param_slots;        // [P]
  const float* params;           // [P][N]
  int response_fn;
  float max_posterior_precision;
  float* surprise;               // [N]
  float* gradient;               // [P][N] or null: forward filter only
};

// One vector of W consecutive instances through the whole series (the time loop of hgf_reverse.cuh::reverse_instance,
// on vectors): forward with checkpoints, then -- if a gradient is wanted -- the backward sweep.
template <class Rev>
static void run_pack(const SimdCall& c, long long first, std::vector<vf>& state) {
  const long long N = c.n_instances;
  const int T = c.n_steps;
  long long idx[W];
  for (int l = 0; l < W; ++l) idx[l] = first + l < N ? first + l : N - 1;
  vf s[Rev::kSlots];
  for (int k = 0; k < Rev::kSlots; ++k) s[k] = splat(c.attributes[k]);
  for (int q = 0; q < c.n_params; ++q)
    for (int l = 0; l < W; ++l) s[c.param_slots[q]][l] = c.params[static_cast<long long>(q) * N + idx[l]];
  // observations / response inputs: one series per group of instances
  const float* up[W];
  const float* rp[W];
  bool u_uniform = true, r_uniform = true;
  for (int l = 0; l < W; ++l) {
    const long long g = c.input_group_size > 0 ? idx[l] / c.input_group_size : 0;
    up[l] = c.input_data + g * T * Rev::kInputs;
    const long long gr = c.response_group_size > 0 ? idx[l] / c.response_group_size : 0;
    rp[l] = c.response_inputs ? c.response_inputs + gr * T : nullptr;
    u_uniform = u_uniform && up[l] == up[0];
    r_uniform = r_uniform && rp[l] == rp[0];
  }
  const vf cap = splat(c.max_posterior_precision);
  auto gather = [&](const float* const* p, bool uniform, long long offset) {
    if (uniform) return splat(p[0][offset]);
    vf r;
    for (int l = 0; l < W; ++l) r[l] = p[l][offset];
    return r;
  };
  auto observe = [&](int t) {
    for (int k = 0; k < Rev::kInputs; ++k) {
      const int node = Rev::input_node(k);
      s[node * NF + F_MEAN] = gather(up, u_uniform, static_cast<long long>(t) * Rev::kInputs + k);
      s[node * NF + F_OBSERVED] = splat(1.0f);
    }
  };
  auto step_size = [&](int t, vf& dt, vf& log_dt) {
    const float d = c.time_steps ? c.time_steps[t] : 1.0f;
    dt = splat(d);
    log_dt = splat(c.time_steps ? ::logf(d) : 0.0f);
  };
  const bool want_grad = c.gradient != nullptr;
  vf total = splat(0.0f);
  vi bad = {};  // total_gaussian_surprise's any-NaN rule (response.py:86-89): +inf, zero gradient
  for (int t = 0; t < T; ++t) {
    if (want_grad)
      for (int k = 0; k < Rev::kCarried; ++k) state[static_cast<size_t>(t) * Rev::kCarried + k] = s[Rev::carried_slot(k)];
    vf dt, log_dt;
    step_size(t, dt, log_dt);
    Rev::predict(s, dt, log_dt, cap);
    observe(t);
    Rev::update(s, dt, log_dt, cap);
    total = total + Rev::response(c.response_fn, s, rp[0] ? gather(rp, r_uniform, t) : splat(0.0f));
    if (c.response_fn == RESP_TOTAL_GAUSSIAN && Rev::kNodes > 1) bad = bad | b_isnan(s[1 * NF + F_MEAN]);
  }
  total = bad ? splat(__builtin_huge_valf()) : total;
  for (int l = 0; l < W; ++l)
    if (first + l < N) c.surprise[first + l] = total[l];
  if (!want_grad) return;
  vf sb[Rev::kSlots];
  for (int k = 0; k < Rev::kSlots; ++k) sb[k] = splat(0.0f);
  vf log[Rev::kLog > 0 ? Rev::kLog : 1];
  for (int t = T - 1; t >= 0; --t) {
    for (int k = 0; k < Rev::kCarried; ++k) s[Rev::carried_slot(k)] = state[static_cast<size_t>(t) * Rev::kCarried + k];
    vf dt, log_dt;
    step_size(t, dt, log_dt);
    Rev::predict_log(s, dt, log_dt, cap, log);
    observe(t);
    Rev::update_log(s, dt, log_dt, cap, log);
    Rev::response_reverse(c.response_fn, s, rp[0] ? gather(rp, r_uniform, t) : splat(0.0f), splat(1.0f), sb);
    Rev::update_reverse(s, sb, log, dt, log_dt, cap);
    for (int k = 0; k < Rev::kInputs; ++k) {
      sb[Rev::input_node(k) * NF + F_MEAN] = splat(0.0f);
      sb[Rev::input_node(k) * NF + F_OBSERVED] = splat(0.0f);
    }
    Rev::predict_reverse(s, sb, log, dt, log_dt, cap);
  }
  for (int q = 0; q < c.n_params; ++q)
    for (int l = 0; l < W; ++l)
      if (first + l < N) c.gradient[static_cast<long long>(q) * N + first + l] = bad[l] ? 0.0f : sb[c.param_slots[q]][l];
}

template <class Rev>
static int run(const SimdCall& c, int n_threads) {
  for (int q = 0; q < c.n_params; ++q)
    if (c.param_slots[q] < 0 || c.param_slots[q] >= Rev::kSlots) return 2;
  const long long packs = (c.n_instances + W - 1) / W;
  if (n_threads <= 0) n_threads = omp_get_max_threads();
#pragma omp parallel num_threads(n_threads)
  {
    std::vector<vf> state(c.gradient ? static_cast<size_t>(c.n_steps > 0 ? c.n_steps : 1) * Rev::kCarried : 1);
#pragma omp for schedule(dynamic, 4)
    for (long long p = 0; p < packs; ++p) run_pack<Rev>(c, p * W, state);
  }
  return 0;
}

}  // namespace hgf

extern "C" {
int hgf_simd_width(void) { return hgf::W; }
int hgf_simd_n_slots(int topology) {
  return topology == 0 ? hgf::Rev_binary3<hgf::vf>::kSlots : topology == 1 ? hgf::Rev_continuous3<hgf::vf>::kSlots
       : topology == 2 ? hgf::Rev_ghgf16<hgf::vf>::kSlots : -1;
}
// topology 0: the 3-level binary HGF, 1: the 4-node continuous HGF, 2: the 16-node generalised HGF (the networks of the
// BASELINE configurations)
int hgf_simd_batch(int topology, const hgf::SimdCall* call, int n_threads) {
  if (!call || call->n_instances < 0 || call->n_steps < 0 || !call->attributes || !call->surprise) return 1;
  if (call->n_instances == 0) return 0;
  if (topology == 0) return hgf::run<hgf::Rev_binary3<hgf::vf>>(*call, n_threads);
  if (topology == 1) return hgf::run<hgf::Rev_continuous3<hgf::vf>>(*call, n_threads);
  if (topology == 2) return hgf::run<hgf::Rev_ghgf16<hgf::vf>>(*call, n_threads);
  return 1;
}
}

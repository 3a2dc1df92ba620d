// Node update equations of the belief-propagation time slice, generic in
//   T -- the scalar (float, double or Dual<.,P>), and
//   A -- the topology accessor: RuntimeTopo reads the lowered Program from the
//        kernel's constant parameter bank (opcode interpreter, attributes in
//        per-thread local memory); a generated StaticTopo answers every query
//        with a compile-time constant, so that after unrolling every attribute
//        index is static and the whole instance lives in registers.
//
// `s` is one instance's attribute vector (layout: hgf_program.h).  Each
// function names the reference code whose result it must reproduce.
#pragma once
#include "hgf_math.cuh"
#include "hgf_program.h"

namespace hgf {

struct RuntimeTopo {
  static constexpr bool is_static = false;
  static constexpr bool kConstAttributes = false;
  static constexpr int kUnroll = 1;  // interpreter: keep the schedule and edge loops rolled
  const Program& p;
  HGF_D explicit RuntimeTopo(const Program& prog) : p(prog) {}
  HGF_D int n_nodes() const { return p.n_nodes; }
  HGF_D int n_slots() const { return p.n_slots; }
  HGF_D int n_pred() const { return p.n_pred; }
  HGF_D int pred_node(int i) const { return p.pred_node[i]; }
  HGF_D int pred_op(int i) const { return p.pred_op[i]; }
  HGF_D int n_upd() const { return p.n_upd; }
  HGF_D int upd_node(int i) const { return p.upd_node[i]; }
  HGF_D int upd_op(int i) const { return p.upd_op[i]; }
  HGF_D int n_branches() const { return p.n_branches; }
  HGF_D int br_pred(int b) const { return p.br_pred[b]; }
  HGF_D int br_upd(int b) const { return p.br_upd[b]; }
  HGF_D int n_inputs() const { return p.n_inputs; }
  HGF_D int n_fed() const { return p.n_fed; }
  HGF_D int input_idx(int k) const { return p.input_idx[k]; }
  HGF_D int vol_updates() const { return p.vol_updates; }
  HGF_D double max_posterior_precision() const { return p.max_posterior_precision; }
  HGF_D int node_type(int n) const { return p.node_type[n]; }
  HGF_D int n_vp(int n) const { return p.n_vp[n]; }
  HGF_D int n_volp(int n) const { return p.n_volp[n]; }
  HGF_D int n_vc(int n) const { return p.n_vc[n]; }
  HGF_D int n_volc(int n) const { return p.n_volc[n]; }
  HGF_D int vp(int n, int k) const { return p.e_vp[p.off_vp[n] + k]; }
  HGF_D int vp_fn(int n, int k) const { return p.e_vp_fn[p.off_vp[n] + k]; }
  HGF_D int volp(int n, int k) const { return p.e_volp[p.off_volp[n] + k]; }
  HGF_D int vc(int n, int k) const { return p.e_vc[p.off_vc[n] + k]; }
  HGF_D int vc_fn(int n, int k) const { return p.e_vc_fn[p.off_vc[n] + k]; }
  HGF_D int volc(int n, int k) const { return p.e_volc[p.off_volc[n] + k]; }
  HGF_D int slot_cvp(int n, int k) const { return p.base_cvp + p.off_vp[n] + k; }
  HGF_D int slot_cvolp(int n, int k) const { return p.base_cvolp + p.off_volp[n] + k; }
  HGF_D int slot_cvc(int n, int k) const { return p.base_cvc + p.off_vc[n] + k; }
  HGF_D int slot_cvolc(int n, int k) const { return p.base_cvolc + p.off_volc[n] + k; }
  HGF_D int slot_clip() const { return p.slot_clip; }
  HGF_D int slot_resp() const { return p.slot_resp; }
};

#define HGF_S(node, field) s[(node) * NF + (field)]

// The time step of the current slice and its logarithm (the "unbounded" update works in log
// space); computed once per slice by the scan -- and folded to (1, 0) when the call has no
// time_steps vector.
// The time step is data, never a differentiated parameter: it is held in the base type, so that a Dual filter spends no
// tangent arithmetic on it.
template <class T> struct Step {
  base_t<T> dt, log_dt;
};

// Attributes that are PARAMETERS OF THE MODEL'S STRUCTURE -- autoconnection strength, tonic drift, coupling strengths, the
// clipping value -- are read through attr().  The general kernels return the full scalar (such a slot may be a
// differentiated per-instance parameter); a topology accessor wrapped in ConstAttributes<> promises that the call
// differentiates none of them (the usual case: gradients wrt tonic volatilities / initial beliefs / response parameters)
// and gets the bare primal, so that every product with them costs 1 + P instructions instead of 1 + 2 P and their
// tangent registers disappear.  Primal arithmetic is identical in both flavours.
template <class T, class A> HGF_D auto attr(const A&, const T* s, int idx) {
  if constexpr (A::kConstAttributes) return pv(s[idx]); else return s[idx];
}
template <class Topo> struct ConstAttributes : Topo {
  static constexpr bool kConstAttributes = true;
  HGF_D explicit ConstAttributes(const Program& prog) : Topo(prog) {}
};
// the `observed` flag of a node is data
#define HGF_OBSERVED(node) pv(HGF_S(node, F_OBSERVED))

// ---------------------------------------------------------------------------
// coupling functions g, g', g''.  The reference differentiates the user's
// callable with jax.grad (prediction/continuous.py:252,
// posterior_update_precision_continuous_node.py:116-126); across a C-ABI the
// callable becomes an enum with analytic derivatives.
// ---------------------------------------------------------------------------
template <class T> HGF_D T cf_value(int fn, const T& x) {
  using B = base_t<T>;
  switch (fn) {
    case CF_RELU: return x_max(x, B(0));
    case CF_SIGMOID: return x_sigmoid(x);
    case CF_TANH: return x_tanh(x);
    case CF_LEAKY_RELU: return (pv(x) >= B(0)) ? x : x * B(0.01);
    case CF_GELU: return x * B(0.5) * (B(1) + x_erf(x / B(1.4142135623730951)));
    case CF_SIN: return x_sin(x);
    default: return x;
  }
}
template <class T> HGF_D T cf_prime(int fn, const T& x) {
  using B = base_t<T>;
  switch (fn) {
    case CF_RELU: return T((pv(x) > B(0)) ? B(1) : B(0));
    case CF_SIGMOID: { const T g = x_sigmoid(x); return g * (B(1) - g); }
    case CF_TANH: { const T g = x_tanh(x); return B(1) - g * g; }
    case CF_LEAKY_RELU: return T((pv(x) >= B(0)) ? B(1) : B(0.01));
    case CF_GELU: {
      const T cdf = B(0.5) * (B(1) + x_erf(x / B(1.4142135623730951)));
      const T pdf = x_exp(-(x * x) / B(2)) / B(2.5066282746310002);
      return cdf + x * pdf;
    }
    case CF_SIN: return x_cos(x);
    default: return T(B(1));
  }
}
template <class T> HGF_D T cf_second(int fn, const T& x) {
  using B = base_t<T>;
  switch (fn) {
    case CF_SIGMOID: { const T g = x_sigmoid(x); return g * (B(1) - g) * (B(1) - B(2) * g); }
    case CF_TANH: { const T g = x_tanh(x); return B(-2) * g * (B(1) - g * g); }
    case CF_GELU: {
      const T pdf = x_exp(-(x * x) / B(2)) / B(2.5066282746310002);
      return pdf * (B(2) - x * x);
    }
    case CF_SIN: return -x_sin(x);
    default: return T(B(0));
  }
}

// ===========================================================================
// Prediction
// ===========================================================================

// binary_state_node_prediction, pyhgf/updates/prediction/binary.py:12-93:
// sigmoid of the summed parent predictions (no coupling strength), clipped.
template <class T, class A> HGF_D void predict_binary(const A& a, T* s, int node) {
  using B = base_t<T>;
  T mu = T(B(0));
#pragma unroll (A::kUnroll)
  for (int k = 0; k < a.n_vp(node); ++k) mu = mu + HGF_S(a.vp(node, k), F_EXPECTED_MEAN);
#if HGF_RESP_MODE != 0
  mu = x_sigmoid_exact(mu);
#else
  mu = x_sigmoid(mu);
#endif
  const auto c = attr(a, s, a.slot_clip());
  mu = x_clip(mu, c, B(1) - c);
  HGF_S(node, F_EXPECTED_MEAN) = mu;
  HGF_S(node, F_EXPECTED_PRECISION) = mu * (B(1) - mu);
}

// predict_mean, prediction/continuous.py:11-103 (volatile twin: volatile.py:51-108,
// which has no tonic drift).
template <class T, class A> HGF_D T predicted_mean(const A& a, const T* s, const Step<T>& st, int node, bool is_volatile) {
  using B = base_t<T>;
  const auto lambda = attr(a, s, node * NF + F_AUTOCONNECTION_STRENGTH);
  if (a.n_vp(node) == 0) {  // no value parents: the drift is the tonic drift alone (same operations, fewer scalar types)
    if (is_volatile) return lambda * HGF_S(node, F_MEAN) + st.dt * T(B(0));
    return lambda * HGF_S(node, F_MEAN) + st.dt * attr(a, s, node * NF + F_TONIC_DRIFT);
  }
  T drift = is_volatile ? T(B(0)) : T(attr(a, s, node * NF + F_TONIC_DRIFT));
#pragma unroll (A::kUnroll)
  for (int k = 0; k < a.n_vp(node); ++k) {
    const int par = a.vp(node, k);
    const int fn = a.vp_fn(node, k);
    const T mu_par = HGF_S(par, F_EXPECTED_MEAN);
    drift = drift + attr(a, s, a.slot_cvp(node, k)) * (fn == CF_LINEAR ? mu_par : cf_value(fn, mu_par));
  }
  return lambda * HGF_S(node, F_MEAN) + st.dt * drift;
}

// Laplace term of the value parents, prediction/continuous.py:240-257.
template <class T, class A> HGF_D T value_parent_variance(const A& a, const T* s, const Step<T>& st, int node) {
  using B = base_t<T>;
  T var = T(B(0));
#pragma unroll (A::kUnroll)
  for (int k = 0; k < a.n_vp(node); ++k) {
    const int par = a.vp(node, k);
    const int fn = a.vp_fn(node, k);
    const T slope = (fn == CF_LINEAR) ? T(B(1)) : cf_prime(fn, HGF_S(par, F_EXPECTED_MEAN));
    const T c = st.dt * attr(a, s, a.slot_cvp(node, k)) * slope;
    var = var + (c * c) / HGF_S(par, F_EXPECTED_PRECISION);
  }
  return var;
}

template <class T> HGF_D T nan_unless_positive(const T& x) {
  // `jnp.where(x > 1e-128, x, nan)`: the literal is 0 in float32
  using B = base_t<T>;
  return x_where(pv(x) > static_cast<B>(1e-128), x, T(b_nan<B>()));
}

// continuous_node_prediction, prediction/continuous.py:282-366 (relaxed scheme
// :106-279) and its mean-field variant :369-512.
template <class T, class A> HGF_D void predict_continuous(const A& a, T* s, const Step<T>& st, int node, bool mean_field) {
  using B = base_t<T>;
  const T variance = B(1) / HGF_S(node, F_PRECISION);
  HGF_S(node, F_CURRENT_VARIANCE) = variance;
  const T mu_hat = predicted_mean(a, s, st, node, false);
  T log_vol = HGF_S(node, F_TONIC_VOLATILITY);
#pragma unroll (A::kUnroll)
  for (int k = 0; k < a.n_volp(node); ++k) {
    const int par = a.volp(node, k);
    const auto kappa = attr(a, s, a.slot_cvolp(node, k));
    log_vol = log_vol + kappa * HGF_S(par, F_EXPECTED_MEAN);
    if (!mean_field) log_vol = log_vol + (kappa * kappa) / (B(2) * HGF_S(par, F_EXPECTED_PRECISION));  // MGF term
  }
  const T omega = nan_unless_positive(st.dt * x_exp(log_vol));
  T pi_cond, pi_marg;
  if (mean_field) {
    pi_cond = B(1) / (variance + omega);
    pi_marg = pi_cond;
  } else {
    pi_cond = B(1) / (variance + omega);
    // without value parents the Laplace term is exactly 0 and the marginal precision IS the conditional one
    // (same bits: x + 0 = x); skipping it saves a reciprocal per node per step
    pi_marg = a.n_vp(node) == 0 ? pi_cond : B(1) / (variance + omega + value_parent_variance(a, s, st, node));
  }
  HGF_S(node, F_EFFECTIVE_PRECISION) = omega * pi_marg;
  if (a.n_vc(node) == 0 && a.n_volc(node) == 0 && a.n_volp(node) == 0) {  // :345-354
    HGF_S(node, F_EXPECTED_PRECISION) = HGF_S(node, F_PRECISION);
    HGF_S(node, F_CONDITIONAL_EXPECTED_PRECISION) = HGF_S(node, F_PRECISION);
  } else {
    HGF_S(node, F_EXPECTED_PRECISION) = pi_marg;
    HGF_S(node, F_CONDITIONAL_EXPECTED_PRECISION) = pi_cond;
  }
  HGF_S(node, F_EXPECTED_MEAN) = mu_hat;
}

// volatile_node_prediction, prediction/volatile.py:230-309 (helpers :9-227),
// mean-field :312-429: implicit volatility level first, then the value level.
template <class T, class A> HGF_D void predict_volatile(const A& a, T* s, const Step<T>& st, int node, bool mean_field) {
  using B = base_t<T>;
  const T variance = B(1) / HGF_S(node, F_PRECISION);
  HGF_S(node, F_CURRENT_VARIANCE) = variance;
  const T omega_vol = nan_unless_positive(st.dt * x_exp(HGF_S(node, F_TONIC_VOLATILITY_VOL)));
  const T pi_hat_vol = B(1) / (B(1) / HGF_S(node, F_PRECISION_VOL) + omega_vol);
  const T mu_hat_vol = HGF_S(node, F_MEAN_VOL);
  HGF_S(node, F_EXPECTED_MEAN_VOL) = mu_hat_vol;
  HGF_S(node, F_EXPECTED_PRECISION_VOL) = pi_hat_vol;
  HGF_S(node, F_EFFECTIVE_PRECISION_VOL) = omega_vol * pi_hat_vol;
  const T log_vol = mean_field ? mu_hat_vol : mu_hat_vol + B(1) / (B(2) * pi_hat_vol);
  const T omega = nan_unless_positive(st.dt * x_exp(log_vol));
  T pi_cond, pi_marg;
  if (mean_field) {
    pi_cond = B(1) / (variance + omega);
    pi_marg = pi_cond;
  } else {
    pi_cond = B(1) / (variance + omega);
    // without value parents the Laplace term is exactly 0 and the marginal precision IS the conditional one
    // (same bits: x + 0 = x); skipping it saves a reciprocal per node per step
    pi_marg = a.n_vp(node) == 0 ? pi_cond : B(1) / (variance + omega + value_parent_variance(a, s, st, node));
  }
  HGF_S(node, F_EXPECTED_MEAN) = predicted_mean(a, s, st, node, true);
  if (a.n_vc(node) == 0 && a.n_volc(node) == 0 && a.n_volp(node) == 0) {  // :288-301
    HGF_S(node, F_EXPECTED_PRECISION) = HGF_S(node, F_PRECISION);
    HGF_S(node, F_CONDITIONAL_EXPECTED_PRECISION) = HGF_S(node, F_PRECISION);
    HGF_S(node, F_EFFECTIVE_PRECISION) = T(B(0));
  } else {
    HGF_S(node, F_EXPECTED_PRECISION) = pi_marg;
    HGF_S(node, F_CONDITIONAL_EXPECTED_PRECISION) = pi_cond;
    HGF_S(node, F_EFFECTIVE_PRECISION) = omega * pi_marg;
  }
}

// ===========================================================================
// Prediction errors
// ===========================================================================

// binary_state_node_prediction_error, prediction_error/binary.py:11-51
template <class T> HGF_D void pe_binary(T* s, int node) {
  const T pi_hat = HGF_S(node, F_EXPECTED_PRECISION);
  HGF_S(node, F_VALUE_PREDICTION_ERROR) =
      ((HGF_S(node, F_MEAN) - HGF_S(node, F_EXPECTED_MEAN)) * HGF_OBSERVED(node)) / pi_hat;
  HGF_S(node, F_PRECISION) = pi_hat;
}

// continuous_node_prediction_error, prediction_error/continuous.py:10-176
template <class T, class A> HGF_D void pe_continuous(const A& a, T* s, int node) {
  using B = base_t<T>;
  const T delta = HGF_S(node, F_MEAN) - HGF_S(node, F_EXPECTED_MEAN);
  const T pi_hat = HGF_S(node, F_EXPECTED_PRECISION);
  HGF_S(node, F_VALUE_PREDICTION_ERROR) = delta;
  T vope = pi_hat / HGF_S(node, F_PRECISION) + pi_hat * x_square(delta) - B(1);
  if (a.n_volp(node) > 0) vope = vope / static_cast<B>(a.n_volp(node));
  HGF_S(node, F_VOLATILITY_PREDICTION_ERROR) = vope;
}

// ===========================================================================
// Posterior updates, continuous nodes
// ===========================================================================

template <class A> HGF_D bool is_gaussian_interior(const A& a, int child) {
  // posterior_update_precision_continuous_node.py:56-60
  const int t = a.node_type(child);
  return (t == NODE_CONTINUOUS || t == NODE_VOLATILE) && (a.n_vc(child) > 0 || a.n_volc(child) > 0);
}

// Effective precision a value child contributes to its parent's posterior
// precision (Schur complement for interior Gaussian children),
// posterior_update_precision_continuous_node.py:41-80.
template <class T, class A> HGF_D T child_precision_weight(const A& a, const T* s, int child, bool mean_field) {
  const T pi_hat = HGF_S(child, F_EXPECTED_PRECISION);
  if (!mean_field && is_gaussian_interior(a, child)) {
    const T pi_y = HGF_S(child, F_PRECISION) - pi_hat;
    const T pi_c = HGF_S(child, F_CONDITIONAL_EXPECTED_PRECISION);
    return pi_c * pi_y / (pi_c + pi_y);
  }
  return pi_hat;
}

// RTS-smoother gain precision used by the mean update,
// posterior_update_mean_continuous_node.py:130-144.
template <class T, class A> HGF_D T child_gain_precision(const A& a, const T* s, int child, bool mean_field) {
  const T pi_hat = HGF_S(child, F_EXPECTED_PRECISION);
  if (!mean_field && is_gaussian_interior(a, child)) {
    const T pi_child = HGF_S(child, F_PRECISION);
    const T pi_y = pi_child - pi_hat;
    const T pi_c = HGF_S(child, F_CONDITIONAL_EXPECTED_PRECISION);
    return pi_c * pi_child / (pi_c + pi_y);
  }
  return pi_hat;
}

// eHGF re-predicted volatility increment,
// posterior_update_precision_continuous_node.py:564-624.
template <class T, class K> HGF_D T ehgf_increment(const T* s, const Step<T>& st, int node, int child, const K& kappa) {
  using B = base_t<T>;
  const T prev_var = HGF_S(child, F_CURRENT_VARIANCE);
  const T omega = st.dt * x_exp(kappa * HGF_S(node, F_MEAN) + HGF_S(child, F_TONIC_VOLATILITY));
  const T pi_hat = B(1) / (prev_var + omega);
  const T gamma = omega * pi_hat;
  const T weight = (omega - prev_var) * pi_hat;
  const T delta = HGF_S(child, F_MEAN) - HGF_S(child, F_EXPECTED_MEAN);
  const T vope = (B(1) / HGF_S(child, F_PRECISION) + delta * delta) * pi_hat - B(1);
  const T inc = x_max(B(0), B(0.5) * (kappa * kappa) * gamma * (gamma + weight * vope));
  return inc * HGF_OBSERVED(child);
}

// posterior_update_precision_continuous_node (+ mean-field, + eHGF forms),
// posterior_update_precision_continuous_node.py:208-365, 425-561, 627-750;
// the unobserved branch of its lax.cond is precision_update_missing_values :368-422.
template <class T, class A>
HGF_D T posterior_precision(const A& a, const T* s, const Step<T>& st, int node, bool mean_field, bool ehgf) {
  using B = base_t<T>;
  bool any_observed = false;
#pragma unroll (A::kUnroll)
  for (int k = 0; k < a.n_vc(node); ++k) any_observed = any_observed || (HGF_OBSERVED(a.vc(node, k)) != B(0));
#pragma unroll (A::kUnroll)
  for (int k = 0; k < a.n_volc(node); ++k) any_observed = any_observed || (HGF_OBSERVED(a.volc(node, k)) != B(0));
  if (!any_observed) {
    T log_vol = HGF_S(node, F_TONIC_VOLATILITY);
#pragma unroll (A::kUnroll)
    for (int k = 0; k < a.n_volp(node); ++k) log_vol = log_vol + attr(a, s, a.slot_cvolp(node, k)) * HGF_S(a.volp(node, k), F_MEAN);
    return B(1) / (B(1) / HGF_S(node, F_PRECISION) + st.dt * x_exp(log_vol));
  }
  T from_values = T(B(0));
#pragma unroll (A::kUnroll)
  for (int k = 0; k < a.n_vc(node); ++k) {
    const int child = a.vc(node, k);
    const int fn = a.vc_fn(node, k);
    const auto psi = attr(a, s, a.slot_cvc(node, k));
    T curvature;
    if (fn == CF_LINEAR) {
      curvature = T(psi * psi);
    } else {
      const T g1 = cf_prime(fn, HGF_S(node, F_MEAN));
      curvature = (psi * psi) * (g1 * g1) - psi * cf_second(fn, HGF_S(node, F_MEAN)) * HGF_S(child, F_VALUE_PREDICTION_ERROR);
    }
    from_values = from_values + child_precision_weight(a, s, child, mean_field) * curvature * HGF_OBSERVED(child);
  }
  T from_volatility = T(B(0));
#pragma unroll (A::kUnroll)
  for (int k = 0; k < a.n_volc(node); ++k) {
    const int child = a.volc(node, k);
    const auto kappa = attr(a, s, a.slot_cvolc(node, k));
    if (ehgf) {
      from_volatility = from_volatility + ehgf_increment(s, st, node, child, kappa);
    } else {  // :142-173
      const T gamma = HGF_S(child, F_EFFECTIVE_PRECISION);
      const T vope = HGF_S(child, F_VOLATILITY_PREDICTION_ERROR);
      const T kg = kappa * gamma;
      from_volatility = from_volatility +
                        (B(0.5) * (kg * kg) + (kg * kg) * vope - B(0.5) * (kappa * kappa) * gamma * vope) * HGF_OBSERVED(child);
    }
  }
  return nan_unless_positive(HGF_S(node, F_EXPECTED_PRECISION) + (from_values + from_volatility));
}

// posterior_update_mean_continuous_node.py:10-195 (mean-field :198-361)
template <class T, class A>
HGF_D T posterior_mean(const A& a, const T* s, int node, const T& node_precision, bool mean_field) {
  using B = base_t<T>;
  T from_values = T(B(0)), from_volatility = T(B(0));
#pragma unroll (A::kUnroll)
  for (int k = 0; k < a.n_vc(node); ++k) {
    const int child = a.vc(node, k);
    const int fn = a.vc_fn(node, k);
    const auto psi = attr(a, s, a.slot_cvc(node, k));
    const T vape = HGF_S(child, F_VALUE_PREDICTION_ERROR) * HGF_OBSERVED(child);
    if (fn == CF_LINEAR) {  // slope 1: x * 1 = x bit for bit, and no tangent arithmetic is spent on the constant
      from_values = from_values + ((psi * child_gain_precision(a, s, child, mean_field)) / node_precision) * vape;
    } else {
      const T slope = cf_prime(fn, HGF_S(node, F_MEAN));
      from_values = from_values + ((psi * child_gain_precision(a, s, child, mean_field) * slope) / node_precision) * vape;
    }
  }
#pragma unroll (A::kUnroll)
  for (int k = 0; k < a.n_volc(node); ++k) {
    const int child = a.volc(node, k);
    T w = attr(a, s, a.slot_cvolc(node, k)) * HGF_S(child, F_EFFECTIVE_PRECISION) * HGF_S(child, F_VOLATILITY_PREDICTION_ERROR);
    w = w * (B(1) / (B(2) * node_precision));
    from_volatility = from_volatility + w * HGF_OBSERVED(child);
  }
  return HGF_S(node, F_EXPECTED_MEAN) + from_values + from_volatility;
}

// continuous_node_posterior_update.py:20-128: precision, then mean.
template <class T, class A> HGF_D void posterior_continuous(const A& a, T* s, const Step<T>& st, int node, bool mean_field) {
  using B = base_t<T>;
  const T pi = x_min(posterior_precision(a, s, st, node, mean_field, false), static_cast<B>(a.max_posterior_precision()));
  HGF_S(node, F_PRECISION) = pi;
  HGF_S(node, F_MEAN) = posterior_mean(a, s, node, pi, mean_field);
}

// continuous_node_posterior_update_ehgf.py:20-155: mean (with the expected
// precision), then precision re-predicted from that mean.
template <class T, class A> HGF_D void posterior_continuous_ehgf(const A& a, T* s, const Step<T>& st, int node, bool mean_field) {
  using B = base_t<T>;
  HGF_S(node, F_MEAN) = posterior_mean(a, s, node, HGF_S(node, F_EXPECTED_PRECISION), mean_field);
  HGF_S(node, F_PRECISION) =
      x_min(posterior_precision(a, s, st, node, mean_field, true), static_cast<B>(a.max_posterior_precision()));
}

// Quantities of the "unbounded" volatility-parent update that the child node
// supplies: previous variance 1/pi(k-1), posterior (mean, precision),
// prediction (mean, precision).
template <class T> struct VolatilityChild {
  T prev_variance, precision, mean, expected_mean, expected_precision;
};

// The two-expansion Lambert-W mixture,
// continuous_node_posterior_update_unbounded.py:51-220 (kappa, omega of the
// child) and volatile_node_posterior_update_unbounded.py:12-178 (kappa = 1, no
// omega: EXPLICIT = false).  mu_hat / pi_hat are the volatility parent's (or
// implicit volatility level's) prediction.
template <bool EXPLICIT, class T, class K>
HGF_D void unbounded_mixture(const Step<T>& st, const VolatilityChild<T>& c, const T& mu_hat, const T& pi_hat, const K& kappa,
                             const T& omega, T& post_precision, T& post_mean) {
  using B = base_t<T>;
  const T prev_var = x_max(c.prev_variance, static_cast<B>(1e-128));
  const T delta = c.mean - c.expected_mean;
  const T be = B(1) / c.precision + x_square(delta);  // E[(x - mu_hat)^2] under the child's posterior
  const B log_dt = st.log_dt;
  const T log_prev_var = x_log(prev_var);
  T gamma;
  if constexpr (EXPLICIT) gamma = log_dt + kappa * mu_hat + omega; else gamma = log_dt + mu_hat;
  // expansion 1: around the prediction
  const T w1 = x_sigmoid(gamma - log_prev_var);
  const T da1 = c.expected_precision * be - B(1);
  // (quotients by pi1, 2 pi1 and by pi2 share one reciprocal per denominator: hgf_math.cuh::Recip -- the same seeds, fewer SFU ops)
  T pi1, mu1, pi_y;
  Recip<T> over_pi1;
  if constexpr (EXPLICIT) {
    pi1 = pi_hat + B(0.5) * (kappa * kappa) * w1 * (B(1) - w1);
    over_pi1 = x_recip(pi1);
    mu1 = mu_hat + (kappa * w1 / recip_twice(over_pi1)) * da1;
    pi_y = pi_hat / (kappa * kappa);
  } else {
    pi1 = pi_hat + B(0.5) * w1 * (B(1) - w1);
    over_pi1 = x_recip(pi1);
    mu1 = mu_hat + (w1 / recip_twice(over_pi1)) * da1;
    pi_y = pi_hat;
  }
  // expansion 2: around the Lambert-W mode
  const T log_w_arg = x_log(be) - x_log(B(2) * pi_y) + B(0.5) / pi_y - gamma;
  const T w_arg = x_exp(x_min(log_w_arg, b_log_max<B>()));
  const T y_star = gamma + x_lambert_w0(w_arg) - B(0.5) / pi_y;
  T x_star, log_s2;
  if constexpr (EXPLICIT) {
    x_star = (y_star - log_dt - omega) / kappa;
    log_s2 = log_dt + kappa * x_star + omega;
  } else {
    x_star = y_star - log_dt;
    log_s2 = log_dt + x_star;
  }
  const T log_den = x_logaddexp(log_prev_var, log_s2);
  const T w2 = x_sigmoid(log_s2 - log_prev_var);
  const T da2 = be * x_exp_neg(log_den) - B(1);
  T pi2_full, pi2_alt;
  if constexpr (EXPLICIT) {
    const auto k2 = kappa * kappa;
    pi2_full = pi_hat + B(0.5) * k2 * w2 * (w2 + (B(2) * w2 - B(1)) * da2);
    pi2_alt = pi_hat + B(0.5) * k2 * w2 * (B(1) - w2);
  } else {
    pi2_full = pi_hat + B(0.5) * w2 * (w2 + (B(2) * w2 - B(1)) * da2);
    pi2_alt = pi_hat + B(0.5) * w2 * (B(1) - w2);
  }
  const T pi2_safe = x_where(pv(pi2_full) <= B(0), pi2_alt, pi2_full);
  const Recip<T> over_pi2_safe = x_recip(pi2_safe);
  T mu2_safe;
  if constexpr (EXPLICIT) mu2_safe = x_star + (B(0.5) * kappa * w2 * da2 - pi_hat * (x_star - mu_hat)) / over_pi2_safe;
  else mu2_safe = x_star + (B(0.5) * w2 * da2 - pi_hat * (x_star - mu_hat)) / over_pi2_safe;
  // fall back to expansion 1 wherever expansion 2 is not finite ("double where",
  // continuous_node_posterior_update_unbounded.py:173-177)
  const bool ok = x_isfinite(pi2_safe) && x_isfinite(mu2_safe);
  const Recip<T> over_pi2 = x_where(ok, over_pi2_safe, over_pi1);  // pi2 = where(ok, pi2_safe, pi1), with its reciprocal
  const T mu2 = x_where(ok, mu2_safe, mu1);
  // variational energies at both candidate means
  T arg1, arg2;
  if constexpr (EXPLICIT) { arg1 = log_dt + kappa * mu1 + omega; arg2 = log_dt + kappa * mu2 + omega; }
  else { arg1 = log_dt + mu1; arg2 = log_dt + mu2; }
  const T l1 = x_logaddexp(log_prev_var, arg1);
  const T e1 = mu1 - mu_hat;
  const T i1 = B(-0.5) * l1 - B(0.5) * be * x_exp_neg(l1) - B(0.5) * pi_hat * x_square(e1);
  const T l2 = x_logaddexp(log_prev_var, arg2);
  const T e2 = mu2 - mu_hat;
  const T i2 = B(-0.5) * l2 - B(0.5) * be * x_exp_neg(l2) - B(0.5) * pi_hat * x_square(e2);
  const T b = x_sigmoid(i2 - i1);
  // moment-matched mixture
  post_mean = (B(1) - b) * mu1 + b * mu2;
  const T dm = mu1 - mu2;
  post_precision = B(1) / ((B(1) - b) / over_pi1 + b / over_pi2 + b * (B(1) - b) * x_square(dm));
}

// continuous_node_posterior_update_unbounded.py:13-48; reads only the first
// volatility child (:71) and never its `observed` flag.
template <class T, class A> HGF_D void posterior_continuous_unbounded(const A& a, T* s, const Step<T>& st, int node) {
  using B = base_t<T>;
  const int child = a.volc(node, 0);
  VolatilityChild<T> c;
  c.prev_variance = HGF_S(child, F_CURRENT_VARIANCE);
  c.precision = HGF_S(child, F_PRECISION);
  c.mean = HGF_S(child, F_MEAN);
  c.expected_mean = HGF_S(child, F_EXPECTED_MEAN);
  c.expected_precision = HGF_S(child, F_EXPECTED_PRECISION);
  T pi, mu;
  unbounded_mixture<true>(st, c, HGF_S(node, F_EXPECTED_MEAN), HGF_S(node, F_EXPECTED_PRECISION), attr(a, s, a.slot_cvolc(node, 0)),
                          HGF_S(child, F_TONIC_VOLATILITY), pi, mu);
  HGF_S(node, F_PRECISION) = x_min(pi, static_cast<B>(a.max_posterior_precision()));
  HGF_S(node, F_MEAN) = mu;
}

// ===========================================================================
// Volatile-state nodes
// ===========================================================================

// Value level: posterior/volatile/posterior_update_value_level.py:11-245 (mean-
// field :248-376), clipped per volatile_node_posterior_update.py:56-63.
// Coupling derivatives are taken at the EXPECTED mean (:120-125); no `observed` gating.
template <class T, class A> HGF_D void posterior_volatile(const A& a, T* s, int node, bool mean_field) {
  using B = base_t<T>;
  const T pi_hat = HGF_S(node, F_EXPECTED_PRECISION);
  T pi = pi_hat;
#pragma unroll (A::kUnroll)
  for (int k = 0; k < a.n_vc(node); ++k) {
    const int child = a.vc(node, k);
    const int fn = a.vc_fn(node, k);
    const auto psi = attr(a, s, a.slot_cvc(node, k));
    const T w = child_precision_weight(a, s, child, mean_field);
    if (fn == CF_LINEAR) {
      pi = pi + (psi * psi) * w;
    } else {
      const T g1 = cf_prime(fn, HGF_S(node, F_EXPECTED_MEAN));
      const T g2 = cf_second(fn, HGF_S(node, F_EXPECTED_MEAN));
      pi = pi + w * ((psi * psi) * (g1 * g1) - psi * g2 * HGF_S(child, F_VALUE_PREDICTION_ERROR));
    }
  }
  pi = x_min(x_max(pi, pi_hat), static_cast<B>(a.max_posterior_precision()));
  HGF_S(node, F_PRECISION) = pi;
  T gain = T(B(0));
#pragma unroll (A::kUnroll)
  for (int k = 0; k < a.n_vc(node); ++k) {
    const int child = a.vc(node, k);
    const int fn = a.vc_fn(node, k);
    if (fn == CF_LINEAR) {
      gain = gain + ((attr(a, s, a.slot_cvc(node, k)) * child_gain_precision(a, s, child, mean_field)) / pi) * HGF_S(child, F_VALUE_PREDICTION_ERROR);
    } else {
      const T slope = cf_prime(fn, HGF_S(node, F_EXPECTED_MEAN));
      gain = gain + ((attr(a, s, a.slot_cvc(node, k)) * child_gain_precision(a, s, child, mean_field) * slope) / pi) *
                        HGF_S(child, F_VALUE_PREDICTION_ERROR);
    }
  }
  HGF_S(node, F_MEAN) = HGF_S(node, F_EXPECTED_MEAN) + gain;
}

// volatile_node_prediction_error, prediction_error/volatile.py:13-154: the two
// prediction errors, then the implicit volatility level's posterior
// (unbounded: volatile_node_posterior_update_unbounded.py:12-178; eHGF:
// volatile_node_posterior_update_ehgf.py:16-75 + posterior_update_volatility_level.py;
// standard: volatile_node_posterior_update.py:75-107).
template <class T, class A> HGF_D void pe_volatile(const A& a, T* s, const Step<T>& st, int node) {
  using B = base_t<T>;
  const T delta = HGF_S(node, F_MEAN) - HGF_S(node, F_EXPECTED_MEAN);
  const T pi_hat = HGF_S(node, F_EXPECTED_PRECISION);
  const T vope = pi_hat / HGF_S(node, F_PRECISION) + pi_hat * x_square(delta) - B(1);
  HGF_S(node, F_VALUE_PREDICTION_ERROR) = delta;
  HGF_S(node, F_VOLATILITY_PREDICTION_ERROR) = vope;
  const B cap = static_cast<B>(a.max_posterior_precision());
  const int vu = a.vol_updates();
  if (vu == VU_UNBOUNDED) {
    VolatilityChild<T> c;
    c.prev_variance = HGF_S(node, F_CURRENT_VARIANCE);
    c.precision = HGF_S(node, F_PRECISION);
    c.mean = HGF_S(node, F_MEAN);
    c.expected_mean = HGF_S(node, F_EXPECTED_MEAN);
    c.expected_precision = pi_hat;
    T pi, mu;
    unbounded_mixture<false>(st, c, HGF_S(node, F_EXPECTED_MEAN_VOL), HGF_S(node, F_EXPECTED_PRECISION_VOL), B(1), T(B(0)),
                             pi, mu);
    HGF_S(node, F_PRECISION_VOL) = x_min(pi, cap);
    HGF_S(node, F_MEAN_VOL) = mu;
  } else if (vu == VU_EHGF) {
    const T gamma = HGF_S(node, F_EFFECTIVE_PRECISION);
    const T mean_vol = HGF_S(node, F_EXPECTED_MEAN_VOL) + (gamma * vope) / (B(2) * HGF_S(node, F_EXPECTED_PRECISION_VOL));
    HGF_S(node, F_MEAN_VOL) = mean_vol;
    const T prev_var = HGF_S(node, F_CURRENT_VARIANCE);
    const T omega = st.dt * x_exp(mean_vol);
    const T pi_rep = B(1) / (prev_var + omega);
    const T gamma_rep = omega * pi_rep;
    const T weight = (omega - prev_var) * pi_rep;
    const T vope_rep = (B(1) / HGF_S(node, F_PRECISION) + delta * delta) * pi_rep - B(1);
    const T inc = x_max(B(0), B(0.5) * gamma_rep * (gamma_rep + weight * vope_rep));
    HGF_S(node, F_PRECISION_VOL) = x_min(HGF_S(node, F_EXPECTED_PRECISION_VOL) + inc, cap);
  } else {
    const T gamma = HGF_S(node, F_EFFECTIVE_PRECISION);
    const T pi_vol = x_min(
        HGF_S(node, F_EXPECTED_PRECISION_VOL) + (B(0.5) * (gamma * gamma) + (gamma * gamma) * vope - B(0.5) * gamma * vope), cap);
    HGF_S(node, F_PRECISION_VOL) = pi_vol;
    HGF_S(node, F_MEAN_VOL) = HGF_S(node, F_EXPECTED_MEAN_VOL) + (gamma * vope) / (B(2) * pi_vol);
  }
}

// ===========================================================================
// One time slice -- pyhgf/utils/beliefs_propagation.py:24-159
// ===========================================================================
template <class T, class A> HGF_D void apply_op(const A& a, T* s, const Step<T>& st, int op, int node) {
  switch (op) {
    case OP_PRED_BINARY: predict_binary(a, s, node); break;
    case OP_PRED_CONT: predict_continuous(a, s, st, node, false); break;
    case OP_PRED_CONT_MF: predict_continuous(a, s, st, node, true); break;
    case OP_PRED_VOLATILE: predict_volatile(a, s, st, node, false); break;
    case OP_PRED_VOLATILE_MF: predict_volatile(a, s, st, node, true); break;
    case OP_PE_BINARY: pe_binary(s, node); break;
    case OP_PE_CONT: pe_continuous(a, s, node); break;
    case OP_PE_VOLATILE: pe_volatile(a, s, st, node); break;
    case OP_POST_CONT: posterior_continuous(a, s, st, node, false); break;
    case OP_POST_CONT_MF: posterior_continuous(a, s, st, node, true); break;
    case OP_POST_CONT_EHGF: posterior_continuous_ehgf(a, s, st, node, false); break;
    case OP_POST_CONT_EHGF_MF: posterior_continuous_ehgf(a, s, st, node, true); break;
    case OP_POST_CONT_UNBOUNDED: posterior_continuous_unbounded(a, s, st, node); break;
    case OP_POST_VOLATILE: posterior_volatile(a, s, node, false); break;
    case OP_POST_VOLATILE_MF: posterior_volatile(a, s, node, true); break;
    default: break;
  }
}

// ===========================================================================
// Response functions, evaluated on the post-step attributes (what the
// reference's recorded trajectory holds at that index) -- pyhgf/response.py,
// pyhgf/math.py:209-297
// ===========================================================================
template <class T> HGF_D T binary_surprise(bool x_is_one, const T& mu_hat) {
  using B = base_t<T>;
  const T m = x_clip(mu_hat, static_cast<B>(1e-6), static_cast<B>(1.0 - 1e-6));
  return x_is_one ? -x_log(m) : -x_log(B(1) - m);
}

template <class T> HGF_D T gaussian_surprise(const T& x, const T& mu_hat, const T& pi_hat) {
  using B = base_t<T>;
  const T d = x - mu_hat;
  const B log_2pi = b_log(static_cast<B>(2.0) * static_cast<B>(3.141592653589793));
  return B(0.5) * (log_2pi - x_log(pi_hat) + pi_hat * (d * d));
}

template <class T> HGF_D T nan_to_inf(const T& x) { return x_isnan(x) ? T(b_inf<base_t<T>>()) : x; }

// Widen a filter scalar to the response function's scalar: same value and leading tangents, the extra
// (response-parameter) tangent directions start at zero.
template <class TR, class T> HGF_D TR widen(const T& x) {
  if constexpr (scalar_traits<TR>::n_tangents == scalar_traits<T>::n_tangents) {
    return x;
  } else {
    TR r(pv(x));
    r.d.widen_from(x.d, scalar_traits<T>::n_tangents);
    return r;
  }
}

// `s` holds the beliefs in the filter's scalar T; `rp` the response-function parameters in the
// response scalar TR (which may carry more tangent directions than T).
template <class TR, class T, class A>
HGF_D TR response_step(const A& a, const T* s, const TR* rp, int response_fn, base_t<T> resp_input) {
  using B = base_t<T>;
#define HGF_R(node, field) widen<TR>(HGF_S(node, field))
  switch (response_fn) {
    case RESP_FIRST_LEVEL_GAUSSIAN:  // node 0's observation against node 1's prediction, response.py:41-48
      return nan_to_inf(gaussian_surprise(HGF_R(0, F_MEAN), HGF_R(1, F_EXPECTED_MEAN), HGF_R(1, F_EXPECTED_PRECISION)));
    case RESP_TOTAL_GAUSSIAN: {  // response.py:51-89 (its any-NaN rule is applied by the caller)
      TR total = TR(B(0));
#pragma unroll (A::kUnroll)
      for (int n = 0; n < a.n_nodes(); ++n)
        if (a.node_type(n) == NODE_CONTINUOUS)
          total = total + gaussian_surprise(HGF_R(n, F_MEAN), HGF_R(n, F_EXPECTED_MEAN), HGF_R(n, F_EXPECTED_PRECISION));
      return total;
    }
    case RESP_FIRST_LEVEL_BINARY: {  // response.py:92-130
      TR total = TR(B(0));
#pragma unroll (A::kUnroll)
      for (int k = 0; k < a.n_inputs(); ++k) {
        const int n = a.input_idx(k);
        total = total + binary_surprise(pv(HGF_S(n, F_MEAN)) != B(0), HGF_R(n, F_EXPECTED_MEAN));
      }
      return nan_to_inf(total);
    }
    case RESP_BINARY_SOFTMAX:  // response.py:133-164
      return nan_to_inf(binary_surprise(resp_input != B(0), HGF_R(0, F_EXPECTED_MEAN)));
    case RESP_BINARY_SOFTMAX_INVERSE_TEMPERATURE: {  // response.py:167-215
      const TR mu = HGF_R(0, F_EXPECTED_MEAN);
      const TR temp = rp[0];
      const bool decided_one = resp_input != B(0);
#if HGF_RESP_MODE == 0
      const TR p1 = x_pow(mu, temp);
      const TR p0 = x_pow(B(1) - mu, temp);
      return nan_to_inf(binary_surprise(decided_one, p1 / (p1 + p0)));
#else
      // The surprise of a decision 0 is -log(1 - p): next to p = 1 (a saturated belief contradicted by the decision)
      // ONE ulp of p is per cents of 1 - p, and a belief pinned at the clipping value repeats the same rounding at
      // every such trial.  Matching the reference there takes its arithmetic to the last bit: libm-grade pow and an
      // IEEE quotient (HGF_RESP_MODE 1: always; 2: only when a lane of the warp is in that regime -- the branch is
      // warp-uniform), while everywhere else the SFU pow's 1e-6 relative error moves the surprise by 1e-6.
      const TR one_minus_mu = B(1) - mu;
      B v1, v0;
#if HGF_RESP_MODE == 2 && defined(__CUDA_ARCH__)
      bool exact = true;
      if constexpr (sizeof(B) == 4) {
        v1 = b_pow(pv(mu), pv(temp));
        v0 = b_pow(pv(one_minus_mu), pv(temp));
        const B p_fast = v1 * sfu_rcp(v1 + v0);
        exact = __any_sync(0xffffffffu, !decided_one && !(p_fast < B(0.96875)));
      }
      if (exact)
#endif
      {
        v1 = exact_pow(pv(mu), pv(temp));
        v0 = exact_pow(pv(one_minus_mu), pv(temp));
      }
      const TR p1 = pow_with_primal(v1, mu, temp);
      const TR p0 = pow_with_primal(v0, one_minus_mu, temp);
      return nan_to_inf(binary_surprise(decided_one, exact_div(p1, p1 + p0)));
#endif
    }
    default: return TR(B(0));
  }
#undef HGF_R
}

}  // namespace hgf

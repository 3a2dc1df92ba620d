// TEST INFRASTRUCTURE ONLY -- CPU oracle for the pyhgf belief-propagation path.
//
// A restatement, in plain C++17, of the reference algorithm of
// ComputationalPsychiatry/pyhgf for the hot path named in BASELINE.json:
// the per-observation time loop over `beliefs_propagation` (prediction ->
// observation -> prediction-error / posterior updates) for binary,
// continuous and volatile state nodes, plus the response (surprise)
// functions.  Every function cites the reference file:line it follows; the
// operation order of each floating-point expression is the reference's, so
// that the float instantiation reproduces the float32 JAX goldens.
//
// Parity status: PINNED against the reference's own test vectors
// (tests/test_oracle_golden.py checks every golden listed in SURVEY.md §4).
// Gradients are NOT pinned by the reference (it has no gradient test on this
// path, and JAX cannot run here): the Dual<double,P> instantiation below is a
// forward-mode restatement of JAX's differentiation rules, cross-checked by
// central finite differences.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may load this library.  The product (pyhgf_b200/)
// never does.
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <vector>

#include "hgf_scalar.hpp"

#ifdef _OPENMP
#include <omp.h>
#endif

namespace oracle {

// node_type codes: pyhgf/typing/typing.py:12-52
enum NodeType { CONSTANT = 0, BINARY = 1, CONTINUOUS = 2, VOLATILE = 6 };

// Coupling functions.  LINEAR is the reference's `coupling_fn is None` branch;
// the others are the enumerated non-linear set (src/math.rs:41-55) plus
// IDENTITY / SIN which the reference tests use (tests/test_continuous.py:59-98).
enum CouplingFn { LINEAR = 0, IDENTITY = 1, RELU = 2, SIGMOID = 3, TANH = 4, LEAKY_RELU = 5, GELU = 6, SIN = 7 };

// update-step opcodes (the oracle's own numbering; produced by oracle/network.py)
enum Op {
  PRED_BINARY = 1, PRED_CONT = 2, PRED_CONT_MF = 3, PRED_VOLATILE = 4, PRED_VOLATILE_MF = 5,
  PE_BINARY = 10, PE_CONT = 11, PE_VOLATILE = 12,
  POST_CONT = 20, POST_CONT_MF = 21, POST_CONT_EHGF = 22, POST_CONT_EHGF_MF = 23,
  POST_CONT_UNBOUNDED = 24, POST_VOLATILE = 25, POST_VOLATILE_MF = 26,
};

enum VolUpdates { UNBOUNDED = 0, EHGF = 1, STANDARD = 2 };

// response functions: pyhgf/response.py
enum Response {
  RESP_NONE = 0, RESP_FIRST_LEVEL_GAUSSIAN = 1, RESP_TOTAL_GAUSSIAN = 2, RESP_FIRST_LEVEL_BINARY = 3,
  RESP_BINARY_SOFTMAX = 4, RESP_BINARY_SOFTMAX_INVERSE_TEMPERATURE = 5,
};

// Per-node scalar attribute fields (flat I/O layout: slot = node * NF + field).
// Names are the reference's attribute keys (pyhgf/model/add_nodes.py:65-85,
// 159-184, 330-340); `temp.*` keys are flattened.
enum Field {
  F_MEAN = 0, F_EXPECTED_MEAN, F_PRECISION, F_EXPECTED_PRECISION,
  F_TONIC_VOLATILITY, F_TONIC_DRIFT, F_AUTOCONNECTION_STRENGTH, F_OBSERVED,
  F_MEAN_VOL, F_EXPECTED_MEAN_VOL, F_PRECISION_VOL, F_EXPECTED_PRECISION_VOL, F_TONIC_VOLATILITY_VOL,
  F_EFFECTIVE_PRECISION, F_CONDITIONAL_EXPECTED_PRECISION, F_VALUE_PREDICTION_ERROR,
  F_VOLATILITY_PREDICTION_ERROR, F_CURRENT_VARIANCE, F_EFFECTIVE_PRECISION_VOL,
  NF
};
constexpr int N_RESP_PARAMS = 4;

struct Edges {  // pyhgf/typing/typing.py:12-52 (AdjacencyLists)
  int node_type = CONTINUOUS;
  bool has_value_parents = false, has_volatility_parents = false;
  bool has_value_children = false, has_volatility_children = false;
  std::vector<int> value_parents, volatility_parents, value_children, volatility_children;
  std::vector<int> coupling_fn;  // one per value child
};

struct Net {
  int n_nodes = 0;
  std::vector<Edges> edges;
  std::vector<int> input_idxs;
  int n_fed = 0;  // leading input nodes that actually receive a data column (zip truncation, beliefs_propagation.py:128)
  std::vector<std::pair<int, int>> prediction_steps, update_steps;  // (node, op)
  int volatility_updates = UNBOUNDED;
  bool mean_field = false;
  double max_posterior_precision = 1e10;
  // flat slot layout
  std::vector<int> off_vcp, off_volcp, off_vcc, off_volcc;  // per node offsets into the slot vector
  int slot_clip = 0, slot_resp = 0, n_slots = 0;
};

template <class T>
struct Attr {
  T f[NF];
  std::vector<T> value_coupling_parents, volatility_coupling_parents, value_coupling_children,
      volatility_coupling_children;
};

template <class T>
struct State {
  std::vector<Attr<T>> a;
  T time_step;
  T precision_clipping_value;
  T resp_params[N_RESP_PARAMS];
};

template <class T>
static void unpack(const Net& net, const T* slots, State<T>& s) {
  s.a.resize(net.n_nodes);
  for (int i = 0; i < net.n_nodes; ++i) {
    Attr<T>& a = s.a[i];
    for (int k = 0; k < NF; ++k) a.f[k] = slots[i * NF + k];
    const Edges& e = net.edges[i];
    a.value_coupling_parents.assign(slots + net.off_vcp[i], slots + net.off_vcp[i] + e.value_parents.size());
    a.volatility_coupling_parents.assign(slots + net.off_volcp[i], slots + net.off_volcp[i] + e.volatility_parents.size());
    a.value_coupling_children.assign(slots + net.off_vcc[i], slots + net.off_vcc[i] + e.value_children.size());
    a.volatility_coupling_children.assign(slots + net.off_volcc[i], slots + net.off_volcc[i] + e.volatility_children.size());
  }
  s.precision_clipping_value = slots[net.slot_clip];
  for (int k = 0; k < N_RESP_PARAMS; ++k) s.resp_params[k] = slots[net.slot_resp + k];
  s.time_step = lit<T>(0.0);
}

template <class T>
static void pack_fields(const Net& net, const State<T>& s, T* slots) {
  for (int i = 0; i < net.n_nodes; ++i)
    for (int k = 0; k < NF; ++k) slots[i * NF + k] = s.a[i].f[k];
}

// ---------------------------------------------------------------------------
// coupling functions g, g', g''  (jax.grad of the user's callable in the
// reference; analytic forms here, src/math.rs:159-381 for the enumerated set)
// ---------------------------------------------------------------------------
template <class T> static T cf_value(int fn, const T& x) {
  switch (fn) {
    case LINEAR: case IDENTITY: return x;
    case RELU: return x_max(x, lit<T>(0.0));
    case SIGMOID: return x_sigmoid(x);
    case TANH: return x_tanh(x);
    case LEAKY_RELU: return (pv(x) >= 0) ? x : lit<T>(0.01) * x;
    case GELU: return x * lit<T>(0.5) * (lit<T>(1.0) + x_erf(x / lit<T>(1.4142135623730951)));
    case SIN: return x_sin(x);
  }
  return x;
}
template <class T> static T cf_prime(int fn, const T& x) {
  switch (fn) {
    case LINEAR: case IDENTITY: return lit<T>(1.0);
    case RELU: return (pv(x) > 0) ? lit<T>(1.0) : lit<T>(0.0);
    case SIGMOID: { T s = x_sigmoid(x); return s * (lit<T>(1.0) - s); }
    case TANH: { T t = x_tanh(x); return lit<T>(1.0) - t * t; }
    case LEAKY_RELU: return (pv(x) >= 0) ? lit<T>(1.0) : lit<T>(0.01);
    case GELU: {
      T phi = lit<T>(0.5) * (lit<T>(1.0) + x_erf(x / lit<T>(1.4142135623730951)));
      T pdf = x_exp(-(x * x) / lit<T>(2.0)) / lit<T>(2.5066282746310002);
      return phi + x * pdf;
    }
    case SIN: return x_cos(x);
  }
  return lit<T>(1.0);
}
template <class T> static T cf_second(int fn, const T& x) {
  switch (fn) {
    case LINEAR: case IDENTITY: case RELU: case LEAKY_RELU: return lit<T>(0.0);
    case SIGMOID: { T s = x_sigmoid(x); return s * (lit<T>(1.0) - s) * (lit<T>(1.0) - lit<T>(2.0) * s); }
    case TANH: { T t = x_tanh(x); return lit<T>(-2.0) * t * (lit<T>(1.0) - t * t); }
    case GELU: {
      T pdf = x_exp(-(x * x) / lit<T>(2.0)) / lit<T>(2.5066282746310002);
      return pdf * (lit<T>(2.0) - x * x);
    }
    case SIN: return -x_sin(x);
  }
  return lit<T>(0.0);
}

static int child_position(const Net& net, int parent, int child) {
  const std::vector<int>& vc = net.edges[parent].value_children;
  return static_cast<int>(std::find(vc.begin(), vc.end(), child) - vc.begin());
}

// ===========================================================================
// PREDICTION
// ===========================================================================

// pyhgf/updates/prediction/binary.py:12-93
template <class T>
static void binary_state_node_prediction(const Net& net, State<T>& s, int node_idx) {
  T expected_mean = lit<T>(0.0);
  for (int value_parent_idx : net.edges[node_idx].value_parents)  // :67-69 (no coupling strength)
    expected_mean = expected_mean + s.a[value_parent_idx].f[F_EXPECTED_MEAN];
  expected_mean = x_sigmoid(expected_mean);  // :73
  T clipping_value = s.precision_clipping_value;  // :81
  expected_mean = x_clip(expected_mean, clipping_value, lit<T>(1.0) - clipping_value);  // :82
  s.a[node_idx].f[F_EXPECTED_MEAN] = expected_mean;
  s.a[node_idx].f[F_EXPECTED_PRECISION] = expected_mean * (lit<T>(1.0) - expected_mean);  // :91
}

// pyhgf/updates/prediction/continuous.py:11-103 and prediction/volatile.py:51-108
template <class T>
static T predict_mean(const Net& net, const State<T>& s, int node_idx, bool volatile_node) {
  const Edges& e = net.edges[node_idx];
  const Attr<T>& a = s.a[node_idx];
  // volatile nodes carry no tonic drift (volatile.py:83)
  T driftrate = volatile_node ? lit<T>(0.0) : a.f[F_TONIC_DRIFT];
  for (size_t k = 0; k < e.value_parents.size(); ++k) {
    int value_parent_idx = e.value_parents[k];
    T psi = a.value_coupling_parents[k];
    int fn = net.edges[value_parent_idx].coupling_fn[child_position(net, value_parent_idx, node_idx)];
    T parent_value = (fn == LINEAR) ? s.a[value_parent_idx].f[F_EXPECTED_MEAN]
                                    : cf_value(fn, s.a[value_parent_idx].f[F_EXPECTED_MEAN]);
    driftrate = driftrate + psi * parent_value;  // :95
  }
  return (a.f[F_AUTOCONNECTION_STRENGTH] * a.f[F_MEAN]) + (s.time_step * driftrate);  // :98-101
}

// Laplace value-coupling variance, continuous.py:240-257 / volatile.py:192-209
template <class T>
static T value_coupling_variance(const Net& net, const State<T>& s, int node_idx) {
  const Edges& e = net.edges[node_idx];
  const Attr<T>& a = s.a[node_idx];
  T vcv = lit<T>(0.0);
  for (size_t k = 0; k < e.value_parents.size(); ++k) {
    int value_parent_idx = e.value_parents[k];
    T psi = a.value_coupling_parents[k];
    int fn = net.edges[value_parent_idx].coupling_fn[child_position(net, value_parent_idx, node_idx)];
    T mu_b = s.a[value_parent_idx].f[F_EXPECTED_MEAN];
    T g_prime = (fn == LINEAR) ? lit<T>(1.0) : cf_prime(fn, mu_b);
    T c = s.time_step * psi * g_prime;
    vcv = vcv + (c * c) / s.a[value_parent_idx].f[F_EXPECTED_PRECISION];
  }
  return vcv;
}

// pyhgf/updates/prediction/continuous.py:106-279 (relaxed) and :369-452 (mean-field)
template <class T>
static void predict_precision(const Net& net, const State<T>& s, int node_idx, bool mean_field,
                              T& expected_precision, T& conditional_expected_precision, T& effective_precision) {
  const Edges& e = net.edges[node_idx];
  const Attr<T>& a = s.a[node_idx];
  T total_volatility = a.f[F_TONIC_VOLATILITY];
  for (size_t k = 0; k < e.volatility_parents.size(); ++k) {
    int p = e.volatility_parents[k];
    T volatility_coupling = a.volatility_coupling_parents[k];
    total_volatility = total_volatility + volatility_coupling * s.a[p].f[F_EXPECTED_MEAN];  // :219-222
    if (!mean_field)  // MGF correction :223-225
      total_volatility = total_volatility +
                         (volatility_coupling * volatility_coupling) / (lit<T>(2.0) * s.a[p].f[F_EXPECTED_PRECISION]);
  }
  T predicted_volatility = s.time_step * x_exp(total_volatility);  // :228
  predicted_volatility = x_where(pv(predicted_volatility) > pv(lit<T>(1e-128)), predicted_volatility,
                                 lit<T>(std::numeric_limits<double>::quiet_NaN()));  // :229-231
  if (mean_field) {  // :446-452
    expected_precision = lit<T>(1.0) / ((lit<T>(1.0) / a.f[F_PRECISION]) + predicted_volatility);
    conditional_expected_precision = expected_precision;
    effective_precision = predicted_volatility * expected_precision;
    return;
  }
  T vcv = value_coupling_variance(net, s, node_idx);
  conditional_expected_precision = lit<T>(1.0) / ((lit<T>(1.0) / a.f[F_PRECISION]) + predicted_volatility);  // :262-264
  expected_precision = lit<T>(1.0) / ((lit<T>(1.0) / a.f[F_PRECISION]) + predicted_volatility + vcv);  // :269-273
  effective_precision = predicted_volatility * expected_precision;  // :277
}

// pyhgf/updates/prediction/continuous.py:282-366 and :455-512
template <class T>
static void continuous_node_prediction(const Net& net, State<T>& s, int node_idx, bool mean_field) {
  const Edges& e = net.edges[node_idx];
  Attr<T>& a = s.a[node_idx];
  a.f[F_CURRENT_VARIANCE] = lit<T>(1.0) / a.f[F_PRECISION];  // :329-331
  T expected_mean = predict_mean(net, s, node_idx, false);
  T ep, cep, gamma;
  predict_precision(net, s, node_idx, mean_field, ep, cep, gamma);
  if (!e.has_value_children && !e.has_volatility_children && !e.has_volatility_parents) {  // :345-354
    a.f[F_EXPECTED_PRECISION] = a.f[F_PRECISION];
    a.f[F_CONDITIONAL_EXPECTED_PRECISION] = a.f[F_PRECISION];
  } else {
    a.f[F_EXPECTED_PRECISION] = ep;
    a.f[F_CONDITIONAL_EXPECTED_PRECISION] = cep;
  }
  a.f[F_EFFECTIVE_PRECISION] = gamma;
  a.f[F_EXPECTED_MEAN] = expected_mean;
}

// pyhgf/updates/prediction/volatile.py:9-309 (relaxed) and :312-429 (mean-field)
template <class T>
static void volatile_node_prediction(const Net& net, State<T>& s, int node_idx, bool mean_field) {
  const Edges& e = net.edges[node_idx];
  Attr<T>& a = s.a[node_idx];
  const T one = lit<T>(1.0);
  const T nan = lit<T>(std::numeric_limits<double>::quiet_NaN());
  a.f[F_CURRENT_VARIANCE] = one / a.f[F_PRECISION];  // :259-261
  // volatility level :30-48
  T predicted_volatility_vol = s.time_step * x_exp(a.f[F_TONIC_VOLATILITY_VOL]);
  predicted_volatility_vol = x_where(pv(predicted_volatility_vol) > pv(lit<T>(1e-128)), predicted_volatility_vol, nan);
  T expected_precision_vol = one / ((one / a.f[F_PRECISION_VOL]) + predicted_volatility_vol);
  T effective_precision_vol = predicted_volatility_vol * expected_precision_vol;
  a.f[F_EXPECTED_MEAN_VOL] = a.f[F_MEAN_VOL];  // :268
  a.f[F_EXPECTED_PRECISION_VOL] = expected_precision_vol;
  a.f[F_EFFECTIVE_PRECISION_VOL] = effective_precision_vol;
  // value level :163-227 / :342-358
  T total_volatility = mean_field ? a.f[F_EXPECTED_MEAN_VOL]
                                  : a.f[F_EXPECTED_MEAN_VOL] + one / (lit<T>(2.0) * a.f[F_EXPECTED_PRECISION_VOL]);
  T predicted_volatility = s.time_step * x_exp(total_volatility);
  predicted_volatility = x_where(pv(predicted_volatility) > pv(lit<T>(1e-128)), predicted_volatility, nan);
  T ep, cep, gamma;
  if (mean_field) {
    ep = one / ((one / a.f[F_PRECISION]) + predicted_volatility);
    cep = ep;
    gamma = predicted_volatility * ep;
  } else {
    T vcv = value_coupling_variance(net, s, node_idx);
    cep = one / ((one / a.f[F_PRECISION]) + predicted_volatility);
    ep = one / ((one / a.f[F_PRECISION]) + predicted_volatility + vcv);
    gamma = predicted_volatility * ep;
  }
  a.f[F_EXPECTED_MEAN] = predict_mean(net, s, node_idx, true);  // :279-281
  if (!e.has_value_children && !e.has_volatility_children && !e.has_volatility_parents) {  // :288-301
    a.f[F_EXPECTED_PRECISION] = a.f[F_PRECISION];
    a.f[F_CONDITIONAL_EXPECTED_PRECISION] = a.f[F_PRECISION];
    a.f[F_EFFECTIVE_PRECISION] = lit<T>(0.0);
  } else {
    a.f[F_EXPECTED_PRECISION] = ep;
    a.f[F_CONDITIONAL_EXPECTED_PRECISION] = cep;
    a.f[F_EFFECTIVE_PRECISION] = gamma;
  }
}

// ===========================================================================
// OBSERVATION + PREDICTION ERRORS
// ===========================================================================

// pyhgf/updates/observation.py:10-38
template <class T>
static void set_observation(State<T>& s, int node_idx, const T& value, const T& observed) {
  s.a[node_idx].f[F_MEAN] = value;
  s.a[node_idx].f[F_OBSERVED] = observed;
}

// pyhgf/updates/prediction_error/binary.py:11-51
template <class T>
static void binary_state_node_prediction_error(State<T>& s, int node_idx) {
  Attr<T>& a = s.a[node_idx];
  T value_prediction_error = a.f[F_MEAN] - a.f[F_EXPECTED_MEAN];           // :34-36
  value_prediction_error = value_prediction_error * a.f[F_OBSERVED];       // :39
  value_prediction_error = value_prediction_error / a.f[F_EXPECTED_PRECISION];  // :43
  a.f[F_VALUE_PREDICTION_ERROR] = value_prediction_error;
  a.f[F_PRECISION] = a.f[F_EXPECTED_PRECISION];  // :49
}

// pyhgf/updates/prediction_error/continuous.py:10-176
template <class T>
static void continuous_node_prediction_error(const Net& net, State<T>& s, int node_idx) {
  Attr<T>& a = s.a[node_idx];
  a.f[F_VALUE_PREDICTION_ERROR] = a.f[F_MEAN] - a.f[F_EXPECTED_MEAN];  // :53-58
  T d = a.f[F_MEAN] - a.f[F_EXPECTED_MEAN];
  T vope = (a.f[F_EXPECTED_PRECISION] / a.f[F_PRECISION]) + a.f[F_EXPECTED_PRECISION] * (d * d) - lit<T>(1.0);  // :103-108
  if (net.edges[node_idx].has_volatility_parents)  // :110-114
    vope = vope / lit<T>(static_cast<double>(net.edges[node_idx].volatility_parents.size()));
  a.f[F_VOLATILITY_PREDICTION_ERROR] = vope;
}

// ===========================================================================
// POSTERIOR UPDATES -- continuous nodes
// ===========================================================================

// posterior_update_precision_continuous_node.py:24-38
template <class T>
static bool has_observations(const Net& net, const State<T>& s, int node_idx) {
  bool any = false;
  for (int c : net.edges[node_idx].value_children) any = any || (pv(s.a[c].f[F_OBSERVED]) != 0);
  for (int c : net.edges[node_idx].volatility_children) any = any || (pv(s.a[c].f[F_OBSERVED]) != 0);
  return any;
}

static bool child_is_gaussian_interior(const Net& net, int child) {  // :56-60
  const Edges& e = net.edges[child];
  return (e.node_type == CONTINUOUS || e.node_type == VOLATILE) && (e.has_value_children || e.has_volatility_children);
}

// posterior_update_precision_continuous_node.py:41-80
template <class T>
static T effective_child_precision(const Net& net, const State<T>& s, int value_child_idx, bool mean_field) {
  const Attr<T>& c = s.a[value_child_idx];
  T child_expected_precision = c.f[F_EXPECTED_PRECISION];
  if (!mean_field && child_is_gaussian_interior(net, value_child_idx)) {
    T pi_y = c.f[F_PRECISION] - child_expected_precision;
    T child_cond = c.f[F_CONDITIONAL_EXPECTED_PRECISION];
    return child_cond * pi_y / (child_cond + pi_y);
  }
  return child_expected_precision;
}

// posterior_update_precision_continuous_node.py:83-139
template <class T>
static T value_coupling_pwpe(const Net& net, const State<T>& s, int node_idx, bool mean_field) {
  const Edges& e = net.edges[node_idx];
  const Attr<T>& a = s.a[node_idx];
  T pwpe = lit<T>(0.0);
  for (size_t k = 0; k < e.value_children.size(); ++k) {
    int value_child_idx = e.value_children[k];
    T value_coupling = a.value_coupling_children[k];
    int fn = e.coupling_fn[k];
    T coupling_fn_prime, coupling_fn_second;
    if (fn == LINEAR) {
      coupling_fn_prime = value_coupling * value_coupling;
      coupling_fn_second = lit<T>(0.0);
    } else {
      T g1 = cf_prime(fn, a.f[F_MEAN]);
      coupling_fn_prime = (value_coupling * value_coupling) * (g1 * g1);  // :116-118
      coupling_fn_second = value_coupling * cf_second(fn, a.f[F_MEAN]) * s.a[value_child_idx].f[F_VALUE_PREDICTION_ERROR];
    }
    T ecp = effective_child_precision(net, s, value_child_idx, mean_field);
    pwpe = pwpe + ecp * (coupling_fn_prime - coupling_fn_second) * s.a[value_child_idx].f[F_OBSERVED];  // :133-137
  }
  return pwpe;
}

// posterior_update_precision_continuous_node.py:142-173
template <class T>
static T standard_volatility_increment(const State<T>& s, int volatility_child_idx, const T& kappa) {
  const Attr<T>& c = s.a[volatility_child_idx];
  T gamma = c.f[F_EFFECTIVE_PRECISION];
  T vope = c.f[F_VOLATILITY_PREDICTION_ERROR];
  T kg = kappa * gamma;
  return (lit<T>(0.5) * (kg * kg) + (kg * kg) * vope - lit<T>(0.5) * (kappa * kappa) * gamma * vope) * c.f[F_OBSERVED];
}

// posterior_update_precision_continuous_node.py:564-624
template <class T>
static T ehgf_volatility_precision_increment(const State<T>& s, int node_idx, int volatility_child_idx, const T& kappa) {
  const Attr<T>& c = s.a[volatility_child_idx];
  T mean = s.a[node_idx].f[F_MEAN];
  T previous_variance = c.f[F_CURRENT_VARIANCE];
  T predicted_volatility = s.time_step * x_exp(kappa * mean + c.f[F_TONIC_VOLATILITY]);
  T expected_precision = lit<T>(1.0) / (previous_variance + predicted_volatility);
  T effective_precision = predicted_volatility * expected_precision;
  T volatility_error_weight = (predicted_volatility - previous_variance) * expected_precision;
  T d = c.f[F_MEAN] - c.f[F_EXPECTED_MEAN];
  T vope = (lit<T>(1.0) / c.f[F_PRECISION] + d * d) * expected_precision - lit<T>(1.0);
  T increment = x_max(lit<T>(0.0), lit<T>(0.5) * (kappa * kappa) * effective_precision *
                                       (effective_precision + volatility_error_weight * vope));
  return increment * c.f[F_OBSERVED];
}

// posterior_update_precision_continuous_node.py:368-422
template <class T>
static T precision_update_missing_values(const Net& net, const State<T>& s, int node_idx) {
  const Edges& e = net.edges[node_idx];
  const Attr<T>& a = s.a[node_idx];
  T total_volatility = a.f[F_TONIC_VOLATILITY];
  for (size_t k = 0; k < e.volatility_parents.size(); ++k)
    total_volatility = total_volatility + a.volatility_coupling_parents[k] * s.a[e.volatility_parents[k]].f[F_MEAN];  // :409-411
  T predicted_volatility = s.time_step * x_exp(total_volatility);
  return lit<T>(1.0) / ((lit<T>(1.0) / a.f[F_PRECISION]) + predicted_volatility);
}

// posterior_update_precision_continuous_node.py:208-365, 425-561, 627-750
template <class T>
static T posterior_update_precision(const Net& net, const State<T>& s, int node_idx, bool mean_field, bool ehgf) {
  if (!has_observations(net, s, node_idx)) return precision_update_missing_values(net, s, node_idx);  // lax.cond :308-313
  const Edges& e = net.edges[node_idx];
  const Attr<T>& a = s.a[node_idx];
  T pwpe = value_coupling_pwpe(net, s, node_idx, mean_field);
  T vol_pwpe = lit<T>(0.0);  // :176-198
  for (size_t k = 0; k < e.volatility_children.size(); ++k) {
    T kappa = a.volatility_coupling_children[k];
    vol_pwpe = vol_pwpe + (ehgf ? ehgf_volatility_precision_increment(s, node_idx, e.volatility_children[k], kappa)
                                : standard_volatility_increment(s, e.volatility_children[k], kappa));
  }
  pwpe = pwpe + vol_pwpe;
  T posterior_precision = a.f[F_EXPECTED_PRECISION] + pwpe;  // :201-205
  return x_where(pv(posterior_precision) > pv(lit<T>(1e-128)), posterior_precision,
                 lit<T>(std::numeric_limits<double>::quiet_NaN()));
}

// posterior_update_mean_continuous_node.py:10-195 (smoothing) and :198-361 (mean-field)
template <class T>
static T posterior_update_mean(const Net& net, const State<T>& s, int node_idx, const T& node_precision, bool mean_field) {
  const Edges& e = net.edges[node_idx];
  const Attr<T>& a = s.a[node_idx];
  T value_pwpe = lit<T>(0.0), volatility_pwpe = lit<T>(0.0);
  for (size_t k = 0; k < e.value_children.size(); ++k) {
    int c = e.value_children[k];
    T value_coupling = a.value_coupling_children[k];
    int fn = e.coupling_fn[k];
    T vape = s.a[c].f[F_VALUE_PREDICTION_ERROR] * s.a[c].f[F_OBSERVED];  // :108-113
    T coupling_fn_prime = (fn == LINEAR) ? lit<T>(1.0) : cf_prime(fn, a.f[F_MEAN]);  // :116-120
    T gain_precision;
    T child_expected_precision = s.a[c].f[F_EXPECTED_PRECISION];
    if (!mean_field && child_is_gaussian_interior(net, c)) {  // :130-144
      T child_precision = s.a[c].f[F_PRECISION];
      T pi_y = child_precision - child_expected_precision;
      T child_cond = s.a[c].f[F_CONDITIONAL_EXPECTED_PRECISION];
      gain_precision = child_cond * child_precision / (child_cond + pi_y);
    } else {
      gain_precision = child_expected_precision;
    }
    value_pwpe = value_pwpe + ((value_coupling * gain_precision * coupling_fn_prime) / node_precision) * vape;  // :147-149
  }
  for (size_t k = 0; k < e.volatility_children.size(); ++k) {
    int c = e.volatility_children[k];
    T kappa = a.volatility_coupling_children[k];
    T pwpe = kappa * s.a[c].f[F_EFFECTIVE_PRECISION] * s.a[c].f[F_VOLATILITY_PREDICTION_ERROR];  // :170-172
    pwpe = pwpe * (lit<T>(1.0) / (lit<T>(2.0) * node_precision));  // :175
    pwpe = pwpe * s.a[c].f[F_OBSERVED];  // :178-180
    volatility_pwpe = volatility_pwpe + pwpe;
  }
  return a.f[F_EXPECTED_MEAN] + value_pwpe + volatility_pwpe;  // :189-193
}

// continuous_node_posterior_update.py:20-128
template <class T>
static void continuous_node_posterior_update(const Net& net, State<T>& s, int node_idx, bool mean_field) {
  T posterior_precision = posterior_update_precision(net, s, node_idx, mean_field, false);
  s.a[node_idx].f[F_PRECISION] = x_min(posterior_precision, lit<T>(net.max_posterior_precision));  // :62-64
  T m = posterior_update_mean(net, s, node_idx, s.a[node_idx].f[F_PRECISION], mean_field);
  s.a[node_idx].f[F_MEAN] = m;
}

// continuous_node_posterior_update_ehgf.py:20-155
template <class T>
static void continuous_node_posterior_update_ehgf(const Net& net, State<T>& s, int node_idx, bool mean_field) {
  T m = posterior_update_mean(net, s, node_idx, s.a[node_idx].f[F_EXPECTED_PRECISION], mean_field);  // :66-72
  s.a[node_idx].f[F_MEAN] = m;
  T posterior_precision = posterior_update_precision(net, s, node_idx, mean_field, true);  // :74-78
  s.a[node_idx].f[F_PRECISION] = x_min(posterior_precision, lit<T>(net.max_posterior_precision));
}

// The two-expansion Lambert-W mixture shared by
// continuous_node_posterior_update_unbounded.py:51-220 (kappa, omega of the child)
// and volatile_node_posterior_update_unbounded.py:12-178 (kappa = 1, no omega).
template <class T>
static void unbounded_core(const T& time_step, const T& current_variance, const T& child_precision, const T& child_mean,
                           const T& child_expected_mean, const T& child_expected_precision, const T& expected_mean,
                           const T& expected_precision, bool explicit_form, const T& volatility_coupling,
                           const T& tonic_volatility, T& posterior_precision, T& posterior_mean) {
  const T one = lit<T>(1.0), half = lit<T>(0.5), two = lit<T>(2.0);
  T previous_variance = x_max(current_variance, lit<T>(1e-128));  // :77-79
  T d = child_mean - child_expected_mean;
  T be_aux = (one / child_precision) + d * d;  // :80-83
  T log_time_step = x_log(time_step);          // :96
  T log_previous_variance = x_log(previous_variance);
  // explicit: gamma = log t + kappa*mu_hat + omega (:100); volatile: gamma = log t + mu_hat_vol (:71)
  T gamma_c = explicit_form ? (log_time_step + volatility_coupling * expected_mean + tonic_volatility)
                            : (log_time_step + expected_mean);
  T w_jm1 = x_sigmoid(gamma_c - log_previous_variance);  // :105
  T da_jm1 = child_expected_precision * be_aux - one;      // :110
  T pi1, mu1, pihat_y;
  if (explicit_form) {
    pi1 = expected_precision + half * (volatility_coupling * volatility_coupling) * w_jm1 * (one - w_jm1);  // :115
    mu1 = expected_mean + (volatility_coupling * w_jm1 / (two * pi1)) * da_jm1;                              // :116
    pihat_y = expected_precision / (volatility_coupling * volatility_coupling);                              // :121
  } else {
    pi1 = expected_precision + half * w_jm1 * (one - w_jm1);  // volatile :84
    mu1 = expected_mean + (w_jm1 / (two * pi1)) * da_jm1;     // :85
    pihat_y = expected_precision;                              // :90
  }
  T log_W_arg = x_log(be_aux) - x_log(two * pihat_y) + half / pihat_y - gamma_c;  // :125
  base_t<T> log_float_max = std::log(std::numeric_limits<base_t<T>>::max());     // :126
  T W_arg = x_exp(x_min(log_W_arg, T(log_float_max)));                            // :127
  T v_W = x_lambert_w0(W_arg);
  T y_star = gamma_c + v_W - half / pihat_y;  // :129
  T x_star, log_s2;
  if (explicit_form) {
    x_star = (y_star - log_time_step - tonic_volatility) / volatility_coupling;  // :130
    log_s2 = log_time_step + volatility_coupling * x_star + tonic_volatility;    // :138
  } else {
    x_star = y_star - log_time_step;  // volatile :99
    log_s2 = log_time_step + x_star;  // :105
  }
  T log_denom_s = x_logaddexp(log_previous_variance, log_s2);  // :139-141
  T w2 = x_sigmoid(log_s2 - log_previous_variance);
  T da2 = be_aux * x_exp(-log_denom_s) - one;  // :143
  T pi2_full, pi2_alt, mu2_safe;
  if (explicit_form) {
    T k2 = volatility_coupling * volatility_coupling;
    pi2_full = expected_precision + half * k2 * w2 * (w2 + (two * w2 - one) * da2);  // :145-147
    pi2_alt = expected_precision + half * k2 * w2 * (one - w2);
  } else {
    pi2_full = expected_precision + half * w2 * (w2 + (two * w2 - one) * da2);  // volatile :112
    pi2_alt = expected_precision + half * w2 * (one - w2);
  }
  T pi2_safe = x_where(pv(pi2_full) <= 0, pi2_alt, pi2_full);  // :148-152
  if (explicit_form)
    mu2_safe = x_star + (half * volatility_coupling * w2 * da2 - expected_precision * (x_star - expected_mean)) / pi2_safe;
  else
    mu2_safe = x_star + (half * w2 * da2 - expected_precision * (x_star - expected_mean)) / pi2_safe;
  bool exp2_finite = x_isfinite(pi2_safe) && x_isfinite(mu2_safe);  // :173
  T pi2 = x_where(exp2_finite, x_where(exp2_finite, pi2_safe, one), pi1);          // :174-176
  T mu2 = x_where(exp2_finite, x_where(exp2_finite, mu2_safe, lit<T>(0.0)), mu1);  // :175-177
  T log_ey1 = explicit_form ? (log_time_step + volatility_coupling * mu1 + tonic_volatility) : (log_time_step + mu1);
  T log_denom_1 = x_logaddexp(log_previous_variance, log_ey1);
  T e1 = mu1 - expected_mean;
  T I1 = lit<T>(-0.5) * log_denom_1 - half * be_aux * x_exp(-log_denom_1) - half * expected_precision * (e1 * e1);  // :196-200
  T log_ey2 = explicit_form ? (log_time_step + volatility_coupling * mu2 + tonic_volatility) : (log_time_step + mu2);
  T log_denom_2 = x_logaddexp(log_previous_variance, log_ey2);
  T e2 = mu2 - expected_mean;
  T I2 = lit<T>(-0.5) * log_denom_2 - half * be_aux * x_exp(-log_denom_2) - half * expected_precision * (e2 * e2);  // :204-208
  T b = x_sigmoid(I2 - I1);  // :211
  posterior_mean = (one - b) * mu1 + b * mu2;  // :216
  T dm = mu1 - mu2;
  T sig2 = (one - b) / pi1 + b / pi2 + b * (one - b) * (dm * dm);  // :217
  posterior_precision = one / sig2;
}

// continuous_node_posterior_update_unbounded.py:13-48
template <class T>
static void continuous_node_posterior_update_unbounded(const Net& net, State<T>& s, int node_idx) {
  int c = net.edges[node_idx].volatility_children[0];  // only the first volatility child (:71)
  Attr<T>& a = s.a[node_idx];
  const Attr<T>& ch = s.a[c];
  T pp, pm;
  unbounded_core<T>(s.time_step, ch.f[F_CURRENT_VARIANCE], ch.f[F_PRECISION], ch.f[F_MEAN], ch.f[F_EXPECTED_MEAN],
                    ch.f[F_EXPECTED_PRECISION], a.f[F_EXPECTED_MEAN], a.f[F_EXPECTED_PRECISION], true,
                    a.volatility_coupling_children[0], ch.f[F_TONIC_VOLATILITY], pp, pm);
  a.f[F_PRECISION] = x_min(pp, lit<T>(net.max_posterior_precision));  // :43-45
  a.f[F_MEAN] = pm;
}

// ===========================================================================
// VOLATILE-STATE NODES
// ===========================================================================

// posterior/volatile/posterior_update_value_level.py:11-133 and :248-308,
// clipped as in volatile_node_posterior_update.py:56-63
template <class T>
static void volatile_node_posterior_update(const Net& net, State<T>& s, int node_idx, bool mean_field) {
  const Edges& e = net.edges[node_idx];
  Attr<T>& a = s.a[node_idx];
  T posterior_precision = a.f[F_EXPECTED_PRECISION];
  for (size_t k = 0; k < e.value_children.size(); ++k) {
    int c = e.value_children[k];
    T value_coupling = a.value_coupling_children[k];
    int fn = e.coupling_fn[k];
    T ecp = effective_child_precision(net, s, c, mean_field);
    if (fn == LINEAR) {
      posterior_precision = posterior_precision + (value_coupling * value_coupling) * ecp;  // :117
    } else {
      T g1 = cf_prime(fn, a.f[F_EXPECTED_MEAN]);  // derivatives at the EXPECTED mean (:120-125)
      T g2 = cf_second(fn, a.f[F_EXPECTED_MEAN]);
      T value_pe = s.a[c].f[F_VALUE_PREDICTION_ERROR];
      posterior_precision = posterior_precision +
                            ecp * ((value_coupling * value_coupling) * (g1 * g1) - value_coupling * g2 * value_pe);  // :128-131
    }
  }
  T precision_value = x_clip(posterior_precision, a.f[F_EXPECTED_PRECISION], lit<T>(net.max_posterior_precision));
  a.f[F_PRECISION] = precision_value;
  // mean: posterior_update_value_level.py:137-245 / :312-376 (no `observed` gating)
  T value_pwpe = lit<T>(0.0);
  for (size_t k = 0; k < e.value_children.size(); ++k) {
    int c = e.value_children[k];
    T value_coupling = a.value_coupling_children[k];
    int fn = e.coupling_fn[k];
    T value_pe = s.a[c].f[F_VALUE_PREDICTION_ERROR];
    T coupling_fn_prime = (fn == LINEAR) ? lit<T>(1.0) : cf_prime(fn, a.f[F_EXPECTED_MEAN]);
    T child_expected_precision = s.a[c].f[F_EXPECTED_PRECISION];
    T gain_precision;
    if (!mean_field && child_is_gaussian_interior(net, c)) {
      T child_precision = s.a[c].f[F_PRECISION];
      T pi_y = child_precision - child_expected_precision;
      T child_cond = s.a[c].f[F_CONDITIONAL_EXPECTED_PRECISION];
      gain_precision = child_cond * child_precision / (child_cond + pi_y);
    } else {
      gain_precision = child_expected_precision;
    }
    value_pwpe = value_pwpe + ((value_coupling * gain_precision * coupling_fn_prime) / precision_value) * value_pe;
  }
  a.f[F_MEAN] = a.f[F_EXPECTED_MEAN] + value_pwpe;
}

// prediction_error/volatile.py:13-154 (PEs, then the volatility-level posterior)
template <class T>
static void volatile_node_prediction_error(const Net& net, State<T>& s, int node_idx) {
  Attr<T>& a = s.a[node_idx];
  const T one = lit<T>(1.0);
  a.f[F_VALUE_PREDICTION_ERROR] = a.f[F_MEAN] - a.f[F_EXPECTED_MEAN];  // :38-42
  T d = a.f[F_MEAN] - a.f[F_EXPECTED_MEAN];
  a.f[F_VOLATILITY_PREDICTION_ERROR] =
      (a.f[F_EXPECTED_PRECISION] / a.f[F_PRECISION]) + a.f[F_EXPECTED_PRECISION] * (d * d) - one;  // :71-76
  T maxp = lit<T>(net.max_posterior_precision);
  if (net.volatility_updates == UNBOUNDED) {  // volatile_node_posterior_update_unbounded.py:12-178
    T pp, pm;
    unbounded_core<T>(s.time_step, a.f[F_CURRENT_VARIANCE], a.f[F_PRECISION], a.f[F_MEAN], a.f[F_EXPECTED_MEAN],
                      a.f[F_EXPECTED_PRECISION], a.f[F_EXPECTED_MEAN_VOL], a.f[F_EXPECTED_PRECISION_VOL], false, one,
                      lit<T>(0.0), pp, pm);
    a.f[F_PRECISION_VOL] = x_min(pp, maxp);
    a.f[F_MEAN_VOL] = pm;
  } else if (net.volatility_updates == EHGF) {  // volatile_node_posterior_update_ehgf.py:16-75
    T vope = a.f[F_VOLATILITY_PREDICTION_ERROR];
    T gamma = a.f[F_EFFECTIVE_PRECISION];
    // mean first, weighted by the expected precision (posterior_update_volatility_level.py:101-139)
    T mean_vol = a.f[F_EXPECTED_MEAN_VOL] + (gamma * vope) / (lit<T>(2.0) * a.f[F_EXPECTED_PRECISION_VOL]);
    a.f[F_MEAN_VOL] = mean_vol;
    // posterior_update_volatility_level.py:50-97
    T previous_variance = a.f[F_CURRENT_VARIANCE];
    T predicted_volatility = s.time_step * x_exp(mean_vol);
    T expected_precision = one / (previous_variance + predicted_volatility);
    T effective_precision = predicted_volatility * expected_precision;
    T volatility_error_weight = (predicted_volatility - previous_variance) * expected_precision;
    T vpe = (one / a.f[F_PRECISION] + d * d) * expected_precision - one;
    T increment = x_max(lit<T>(0.0), lit<T>(0.5) * effective_precision * (effective_precision + volatility_error_weight * vpe));
    a.f[F_PRECISION_VOL] = x_min(a.f[F_EXPECTED_PRECISION_VOL] + increment, maxp);
  } else {  // standard: volatile_node_posterior_update.py:75-107
    T vope = a.f[F_VOLATILITY_PREDICTION_ERROR];
    T gamma = a.f[F_EFFECTIVE_PRECISION];
    T precision_vol = a.f[F_EXPECTED_PRECISION_VOL] +
                      (lit<T>(0.5) * (gamma * gamma) + (gamma * gamma) * vope - lit<T>(0.5) * gamma * vope);  // :40-44
    precision_vol = x_min(precision_vol, maxp);
    a.f[F_PRECISION_VOL] = precision_vol;
    a.f[F_MEAN_VOL] = a.f[F_EXPECTED_MEAN_VOL] + (gamma * vope) / (lit<T>(2.0) * precision_vol);  // :133-137
  }
}

// ===========================================================================
// ONE TIME SLICE -- pyhgf/utils/beliefs_propagation.py:24-159
// ===========================================================================
template <class T>
static void apply_op(const Net& net, State<T>& s, int op, int node) {
  switch (op) {
    case PRED_BINARY: binary_state_node_prediction(net, s, node); break;
    case PRED_CONT: continuous_node_prediction(net, s, node, false); break;
    case PRED_CONT_MF: continuous_node_prediction(net, s, node, true); break;
    case PRED_VOLATILE: volatile_node_prediction(net, s, node, false); break;
    case PRED_VOLATILE_MF: volatile_node_prediction(net, s, node, true); break;
    case PE_BINARY: binary_state_node_prediction_error(s, node); break;
    case PE_CONT: continuous_node_prediction_error(net, s, node); break;
    case PE_VOLATILE: volatile_node_prediction_error(net, s, node); break;
    case POST_CONT: continuous_node_posterior_update(net, s, node, false); break;
    case POST_CONT_MF: continuous_node_posterior_update(net, s, node, true); break;
    case POST_CONT_EHGF: continuous_node_posterior_update_ehgf(net, s, node, false); break;
    case POST_CONT_EHGF_MF: continuous_node_posterior_update_ehgf(net, s, node, true); break;
    case POST_CONT_UNBOUNDED: continuous_node_posterior_update_unbounded(net, s, node); break;
    case POST_VOLATILE: volatile_node_posterior_update(net, s, node, false); break;
    case POST_VOLATILE_MF: volatile_node_posterior_update(net, s, node, true); break;
    default: break;
  }
}

template <class T>
static void beliefs_propagation(const Net& net, State<T>& s, const T* values, const T* observed, const T& time_step) {
  s.time_step = time_step;                                              // :76
  for (auto& st : net.prediction_steps) apply_op(net, s, st.second, st.first);  // :91-92
  for (int k = 0; k < net.n_fed; ++k)                                   // :126-134
    set_observation(s, net.input_idxs[k], values[k], observed ? observed[k] : lit<T>(1.0));
  for (auto& st : net.update_steps) apply_op(net, s, st.second, st.first);      // :145-146
}

// ===========================================================================
// RESPONSE FUNCTIONS (per time step) -- pyhgf/response.py, pyhgf/math.py
// ===========================================================================

// pyhgf/math.py:209-253
template <class T>
static T binary_surprise(const T& x, T expected_mean) {
  expected_mean = x_clip(expected_mean, lit<T>(1e-6), lit<T>(1.0 - 1e-6));
  return (pv(x) != 0) ? -x_log(expected_mean) : -x_log(lit<T>(1.0) - expected_mean);
}

// pyhgf/math.py:256-297
template <class T>
static T gaussian_surprise(const T& x, const T& expected_mean, const T& expected_precision) {
  T d = x - expected_mean;
  base_t<T> two_pi = static_cast<base_t<T>>(2.0) * static_cast<base_t<T>>(M_PI);
  return lit<T>(0.5) * (T(std::log(two_pi)) - x_log(expected_precision) + expected_precision * (d * d));
}

template <class T>
static T nan_to_inf(const T& x) { return x_isnan(x) ? lit<T>(std::numeric_limits<double>::infinity()) : x; }

// Surprise of one time step, evaluated on the post-step attributes (which is
// what the recorded trajectory holds at that index).  `total_gaussian`'s
// any-NaN-anywhere rule (response.py:87-89) is applied by the caller.
template <class T>
static T response_step(const Net& net, const State<T>& s, int response_fn, const T& resp_input) {
  switch (response_fn) {
    case RESP_FIRST_LEVEL_GAUSSIAN:  // response.py:15-48 (node 0's observation vs node 1's prediction)
      return nan_to_inf(gaussian_surprise(s.a[0].f[F_MEAN], s.a[1].f[F_EXPECTED_MEAN], s.a[1].f[F_EXPECTED_PRECISION]));
    case RESP_TOTAL_GAUSSIAN: {  // response.py:51-89
      T surprise = lit<T>(0.0);
      for (int i = 0; i < net.n_nodes; ++i)
        if (net.edges[i].node_type == CONTINUOUS)
          surprise = surprise + gaussian_surprise(s.a[i].f[F_MEAN], s.a[i].f[F_EXPECTED_MEAN], s.a[i].f[F_EXPECTED_PRECISION]);
      return surprise;
    }
    case RESP_FIRST_LEVEL_BINARY: {  // response.py:92-130
      T surprise = lit<T>(0.0);
      for (int idx : net.input_idxs)
        surprise = surprise + binary_surprise(s.a[idx].f[F_MEAN], s.a[idx].f[F_EXPECTED_MEAN]);
      return nan_to_inf(surprise);
    }
    case RESP_BINARY_SOFTMAX:  // response.py:133-164
      return nan_to_inf(binary_surprise(resp_input, s.a[0].f[F_EXPECTED_MEAN]));
    case RESP_BINARY_SOFTMAX_INVERSE_TEMPERATURE: {  // response.py:167-215
      T mu = s.a[0].f[F_EXPECTED_MEAN];
      T t = s.resp_params[0];
      T a = x_pow(mu, t);
      T b = x_pow(lit<T>(1.0) - mu, t);
      T beliefs = a / (a + b);
      return nan_to_inf(binary_surprise(resp_input, beliefs));
    }
  }
  return lit<T>(0.0);
}

// ===========================================================================
// DRIVERS
// ===========================================================================

// Network.input_data (pyhgf/model/network.py:528-628): scan over T slices.
// values [T][n_inputs], observed [T][n_inputs] or null, time_steps [T] or null.
template <class T>
static void run_scan(const Net& net, State<T>& s, int n_steps, const double* values, const double* observed,
                     const double* time_steps, double* traj /* [T][n_nodes*NF] or null */) {
  const int n_in = net.n_fed;
  std::vector<T> v(n_in), o(n_in);
  std::vector<T> flat(static_cast<size_t>(net.n_nodes) * NF);
  for (int t = 0; t < n_steps; ++t) {
    for (int k = 0; k < n_in; ++k) {
      v[k] = lit<T>(values[static_cast<size_t>(t) * n_in + k]);
      o[k] = lit<T>(observed ? observed[static_cast<size_t>(t) * n_in + k] : 1.0);
    }
    beliefs_propagation<T>(net, s, v.data(), o.data(), lit<T>(time_steps ? time_steps[t] : 1.0));
    if (traj) {
      pack_fields(net, s, flat.data());
      double* row = traj + static_cast<size_t>(t) * net.n_nodes * NF;
      for (size_t k = 0; k < flat.size(); ++k) row[k] = static_cast<double>(pv(flat[k]));
    }
  }
}

// Summed surprise of one instance with the response fused into the scan
// (the user-level pattern `net.input_data(u).surprise(fn).sum()`).
template <class T>
static T run_surprise_sum(const Net& net, State<T>& s, int n_steps, const double* values, const double* observed,
                          const double* time_steps, int response_fn, const double* resp_inputs) {
  const int n_in = net.n_fed;
  std::vector<T> v(n_in), o(n_in);
  T total = lit<T>(0.0);
  bool node1_mean_nan = false;
  for (int t = 0; t < n_steps; ++t) {
    for (int k = 0; k < n_in; ++k) {
      v[k] = lit<T>(values[static_cast<size_t>(t) * n_in + k]);
      o[k] = lit<T>(observed ? observed[static_cast<size_t>(t) * n_in + k] : 1.0);
    }
    beliefs_propagation<T>(net, s, v.data(), o.data(), lit<T>(time_steps ? time_steps[t] : 1.0));
    total = total + response_step<T>(net, s, response_fn, lit<T>(resp_inputs ? resp_inputs[t] : 0.0));
    if (net.n_nodes > 1 && x_isnan(s.a[1].f[F_MEAN])) node1_mean_nan = true;
  }
  if (response_fn == RESP_TOTAL_GAUSSIAN && node1_mean_nan) total = lit<T>(std::numeric_limits<double>::infinity());
  return total;
}

constexpr int MAXP = 8;  // max simultaneous tangent directions of the gradient oracle
using DualD = Dual<double, MAXP>;

}  // namespace oracle

// ===========================================================================
// C API (ctypes)
// ===========================================================================
using namespace oracle;

template <class T>
static void load_state(const Net& net, const double* slots, State<T>& s) {
  std::vector<T> tmp(net.n_slots);
  for (int k = 0; k < net.n_slots; ++k) tmp[k] = lit<T>(slots[k]);
  unpack<T>(net, tmp.data(), s);
}
template <class T>
static void store_fields(const Net& net, const State<T>& s, double* slots) {
  for (int i = 0; i < net.n_nodes; ++i)
    for (int k = 0; k < NF; ++k) slots[i * NF + k] = static_cast<double>(pv(s.a[i].f[k]));
}

template <class B>
static void batch_grad_impl(const Net& net, long long n_instances, int n_steps, long long group_size, const double* values,
                            const double* observed, const double* time_steps, const double* slots0, int n_params,
                            const int* param_slots, const double* params, int response_fn, const double* resp_inputs,
                            double* surprise_sum, double* grad) {
  using D = Dual<B, MAXP>;
  const int n_in = net.n_fed;
#pragma omp parallel for schedule(static)
  for (long long n = 0; n < n_instances; ++n) {
    std::vector<D> slots(net.n_slots);
    for (int k = 0; k < net.n_slots; ++k) slots[k] = D(static_cast<B>(slots0[k]));
    for (int q = 0; q < n_params; ++q) {
      D x(static_cast<B>(params[static_cast<size_t>(q) * n_instances + n]));
      x.d[q] = B(1);
      slots[param_slots[q]] = x;
    }
    State<D> s; unpack<D>(net, slots.data(), s);
    long long g = group_size > 0 ? n / group_size : 0;
    const double* v = values + static_cast<size_t>(g) * n_steps * n_in;
    const double* o = observed ? observed + static_cast<size_t>(g) * n_steps * n_in : nullptr;
    const double* r = resp_inputs ? resp_inputs + static_cast<size_t>(g) * n_steps : nullptr;
    D total = run_surprise_sum<D>(net, s, n_steps, v, o, time_steps, response_fn, r);
    surprise_sum[n] = static_cast<double>(total.v);
    for (int q = 0; q < n_params; ++q) grad[static_cast<size_t>(q) * n_instances + n] = static_cast<double>(total.d[q]);
  }
}

extern "C" {

void* oracle_net_create(int n_nodes, const int* node_type, const int* vp_ptr, const int* vp_idx, const int* volp_ptr,
                        const int* volp_idx, const int* vc_ptr, const int* vc_idx, const int* vc_fn, const int* volc_ptr,
                        const int* volc_idx, int volatility_updates, int mean_field, double max_posterior_precision) {
  Net* net = new Net();
  net->n_nodes = n_nodes;
  net->edges.resize(n_nodes);
  net->volatility_updates = volatility_updates;
  net->mean_field = mean_field != 0;
  net->max_posterior_precision = max_posterior_precision;
  for (int i = 0; i < n_nodes; ++i) {
    Edges& e = net->edges[i];
    e.node_type = node_type[i];
    e.value_parents.assign(vp_idx + vp_ptr[i], vp_idx + vp_ptr[i + 1]);
    e.volatility_parents.assign(volp_idx + volp_ptr[i], volp_idx + volp_ptr[i + 1]);
    e.value_children.assign(vc_idx + vc_ptr[i], vc_idx + vc_ptr[i + 1]);
    e.coupling_fn.assign(vc_fn + vc_ptr[i], vc_fn + vc_ptr[i + 1]);
    e.volatility_children.assign(volc_idx + volc_ptr[i], volc_idx + volc_ptr[i + 1]);
    e.has_value_parents = !e.value_parents.empty();
    e.has_volatility_parents = !e.volatility_parents.empty();
    e.has_value_children = !e.value_children.empty();
    e.has_volatility_children = !e.volatility_children.empty();
  }
  int off = n_nodes * NF;
  net->off_vcp.resize(n_nodes); net->off_volcp.resize(n_nodes); net->off_vcc.resize(n_nodes); net->off_volcc.resize(n_nodes);
  for (int i = 0; i < n_nodes; ++i) { net->off_vcp[i] = off; off += static_cast<int>(net->edges[i].value_parents.size()); }
  for (int i = 0; i < n_nodes; ++i) { net->off_volcp[i] = off; off += static_cast<int>(net->edges[i].volatility_parents.size()); }
  for (int i = 0; i < n_nodes; ++i) { net->off_vcc[i] = off; off += static_cast<int>(net->edges[i].value_children.size()); }
  for (int i = 0; i < n_nodes; ++i) { net->off_volcc[i] = off; off += static_cast<int>(net->edges[i].volatility_children.size()); }
  net->slot_clip = off++;
  net->slot_resp = off; off += N_RESP_PARAMS;
  net->n_slots = off;
  return net;
}

void oracle_net_destroy(void* p) { delete static_cast<Net*>(p); }

void oracle_net_set_schedule(void* p, int n_pred, const int* pred_node, const int* pred_op, int n_upd, const int* upd_node,
                             const int* upd_op) {
  Net* net = static_cast<Net*>(p);
  net->prediction_steps.clear(); net->update_steps.clear();
  for (int i = 0; i < n_pred; ++i) net->prediction_steps.push_back({pred_node[i], pred_op[i]});
  for (int i = 0; i < n_upd; ++i) net->update_steps.push_back({upd_node[i], upd_op[i]});
}

void oracle_net_set_inputs(void* p, int n_inputs, const int* input_idxs, int n_fed) {
  static_cast<Net*>(p)->input_idxs.assign(input_idxs, input_idxs + n_inputs);
  static_cast<Net*>(p)->n_fed = n_fed;
}

int oracle_n_fields(void) { return NF; }
int oracle_n_slots(void* p) { return static_cast<Net*>(p)->n_slots; }
int oracle_slot_coupling(void* p, int kind, int node, int pos) {
  Net* net = static_cast<Net*>(p);
  switch (kind) {
    case 0: return net->off_vcp[node] + pos;
    case 1: return net->off_volcp[node] + pos;
    case 2: return net->off_vcc[node] + pos;
    case 3: return net->off_volcc[node] + pos;
  }
  return -1;
}
int oracle_slot_clip(void* p) { return static_cast<Net*>(p)->slot_clip; }
int oracle_slot_resp(void* p) { return static_cast<Net*>(p)->slot_resp; }

// One instance.  slots: in = initial attributes, out = last attributes.
int oracle_run(void* p, int dtype, int n_steps, const double* values, const double* observed, const double* time_steps,
               double* slots, double* traj) {
  const Net& net = *static_cast<Net*>(p);
  if (dtype == 32) {
    State<float> s; load_state<float>(net, slots, s);
    run_scan<float>(net, s, n_steps, values, observed, time_steps, traj);
    store_fields(net, s, slots);
  } else {
    State<double> s; load_state<double>(net, slots, s);
    run_scan<double>(net, s, n_steps, values, observed, time_steps, traj);
    store_fields(net, s, slots);
  }
  return 0;
}

// One update step applied in place (for the isolated-kernel goldens,
// tests/test_updates/posterior/test_posterior_update_continuous.py).
int oracle_apply_op(void* p, int dtype, int op, int node, double time_step, double* slots) {
  const Net& net = *static_cast<Net*>(p);
  if (dtype == 32) {
    State<float> s; load_state<float>(net, slots, s); s.time_step = static_cast<float>(time_step);
    apply_op<float>(net, s, op, node); store_fields(net, s, slots);
  } else {
    State<double> s; load_state<double>(net, slots, s); s.time_step = time_step;
    apply_op<double>(net, s, op, node); store_fields(net, s, slots);
  }
  return 0;
}

// Per-step surprise of a response function from a recorded trajectory
// (Network.surprise, pyhgf/model/network.py:986-1022).
int oracle_response(void* p, int dtype, int response_fn, int n_steps, const double* traj, const double* slots0,
                    const double* resp_inputs, double* out) {
  const Net& net = *static_cast<Net*>(p);
  std::vector<double> slots(slots0, slots0 + net.n_slots);
  bool node1_nan = false;
  for (int t = 0; t < n_steps; ++t) {
    std::memcpy(slots.data(), traj + static_cast<size_t>(t) * net.n_nodes * NF, sizeof(double) * net.n_nodes * NF);
    double ri = resp_inputs ? resp_inputs[t] : 0.0;
    if (dtype == 32) {
      State<float> s; load_state<float>(net, slots.data(), s);
      out[t] = static_cast<double>(response_step<float>(net, s, response_fn, static_cast<float>(ri)));
    } else {
      State<double> s; load_state<double>(net, slots.data(), s);
      out[t] = response_step<double>(net, s, response_fn, ri);
    }
    if (net.n_nodes > 1 && std::isnan(slots[1 * NF + F_MEAN])) node1_nan = true;
  }
  if (response_fn == RESP_TOTAL_GAUSSIAN && node1_nan)
    for (int t = 0; t < n_steps; ++t) out[t] = std::numeric_limits<double>::infinity();
  return 0;
}

// Batched summed surprise.  Instance n starts from slots0 with
// slots0[param_slots[p]] replaced by params[p*N + n]; it reads observation
// group n / group_size (values [G][T][n_inputs], resp_inputs [G][T]).
int oracle_batch_surprise(void* p, int dtype, long long n_instances, int n_steps, long long group_size,
                          const double* values, const double* observed, const double* time_steps, const double* slots0,
                          int n_params, const int* param_slots, const double* params, int response_fn,
                          const double* resp_inputs, double* surprise_sum, double* last_fields, int n_threads) {
  const Net& net = *static_cast<Net*>(p);
  const int n_in = net.n_fed;
#ifdef _OPENMP
  if (n_threads > 0) omp_set_num_threads(n_threads);
#endif
#pragma omp parallel for schedule(static)
  for (long long n = 0; n < n_instances; ++n) {
    std::vector<double> slots(slots0, slots0 + net.n_slots);
    for (int q = 0; q < n_params; ++q) slots[param_slots[q]] = params[static_cast<size_t>(q) * n_instances + n];
    long long g = group_size > 0 ? n / group_size : 0;
    const double* v = values + static_cast<size_t>(g) * n_steps * n_in;
    const double* o = observed ? observed + static_cast<size_t>(g) * n_steps * n_in : nullptr;
    const double* r = resp_inputs ? resp_inputs + static_cast<size_t>(g) * n_steps : nullptr;
    if (dtype == 32) {
      State<float> s; load_state<float>(net, slots.data(), s);
      surprise_sum[n] = static_cast<double>(run_surprise_sum<float>(net, s, n_steps, v, o, time_steps, response_fn, r));
      if (last_fields) store_fields(net, s, last_fields + static_cast<size_t>(n) * net.n_nodes * NF);
    } else {
      State<double> s; load_state<double>(net, slots.data(), s);
      surprise_sum[n] = run_surprise_sum<double>(net, s, n_steps, v, o, time_steps, response_fn, r);
      if (last_fields) store_fields(net, s, last_fields + static_cast<size_t>(n) * net.n_nodes * NF);
    }
  }
  return 0;
}

// Batched summed surprise and its gradient wrt the n_params (<= MAXP) batched
// slots, by forward-mode Dual propagation (dtype 64: Dual<double>, the reference
// gradient; dtype 32: Dual<float>, the float32 twin used to check float32
// kernels like-for-like in rounding-sensitive regimes).  grad [n_params][N].
int oracle_batch_grad(void* p, int dtype, long long n_instances, int n_steps, long long group_size, const double* values,
                      const double* observed, const double* time_steps, const double* slots0, int n_params,
                      const int* param_slots, const double* params, int response_fn, const double* resp_inputs,
                      double* surprise_sum, double* grad, int n_threads) {
  const Net& net = *static_cast<Net*>(p);
  if (n_params > MAXP) return 1;
#ifdef _OPENMP
  if (n_threads > 0) omp_set_num_threads(n_threads);
#endif
  if (dtype == 32)
    batch_grad_impl<float>(net, n_instances, n_steps, group_size, values, observed, time_steps, slots0, n_params, param_slots,
                           params, response_fn, resp_inputs, surprise_sum, grad);
  else
    batch_grad_impl<double>(net, n_instances, n_steps, group_size, values, observed, time_steps, slots0, n_params, param_slots,
                            params, response_fn, resp_inputs, surprise_sum, grad);
  return 0;
}

// scalar surprise functions (pyhgf/math.py:209-297) for the KATs of tests/test_math.py
double oracle_binary_surprise(int dtype, double x, double expected_mean) {
  return dtype == 32 ? static_cast<double>(binary_surprise<float>(static_cast<float>(x), static_cast<float>(expected_mean)))
                     : binary_surprise<double>(x, expected_mean);
}
double oracle_gaussian_surprise(int dtype, double x, double expected_mean, double expected_precision) {
  return dtype == 32 ? static_cast<double>(gaussian_surprise<float>(static_cast<float>(x), static_cast<float>(expected_mean),
                                                                     static_cast<float>(expected_precision)))
                     : gaussian_surprise<double>(x, expected_mean, expected_precision);
}
double oracle_lambert_w0(int dtype, double z) {
  return dtype == 32 ? static_cast<double>(x_lambert_w0(static_cast<float>(z))) : x_lambert_w0(z);
}

}  // extern "C"

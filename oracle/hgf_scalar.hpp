// TEST INFRASTRUCTURE ONLY -- CPU oracle for the pyhgf belief-propagation path.
//
// Scalar abstraction used by hgf_oracle.cpp: the update equations are written
// once, templated on T in {float, double, Dual<double,P>}.
//
//  * float  : reproduces the reference's default float32 JAX arithmetic
//             (one rounding per primitive, no FMA contraction -- build with
//             -ffp-contract=off).
//  * double : the jax_enable_x64 arithmetic / Rust-backend cross-check.
//  * Dual   : forward-mode derivative carrier.  It is the gradient oracle that
//             stands in for `jax.grad` (which cannot run in this container):
//             it follows JAX's differentiation rules for every primitive on
//             the path, including the custom JVPs of `lambert_w0`
//             (pyhgf/math.py:491-498) and `jnp.logaddexp`.
//
// Nothing outside tests/, __graft_entry__.smoke() and bench.py's cpu_baseline
// leg may include or link this file.
#pragma once
#include <cmath>
#include <limits>

namespace oracle {

// ---------------------------------------------------------------------------
// Dual numbers (value + P tangents)
// ---------------------------------------------------------------------------
template <class B, int P>
struct Dual {
  B v;
  B d[P];
  Dual() : v(0) { for (int i = 0; i < P; ++i) d[i] = 0; }
  Dual(B x) : v(x) { for (int i = 0; i < P; ++i) d[i] = 0; }  // NOLINT implicit
};

template <class B, int P> inline Dual<B, P> operator+(const Dual<B, P>& a, const Dual<B, P>& b) {
  Dual<B, P> r; r.v = a.v + b.v; for (int i = 0; i < P; ++i) r.d[i] = a.d[i] + b.d[i]; return r;
}
template <class B, int P> inline Dual<B, P> operator-(const Dual<B, P>& a, const Dual<B, P>& b) {
  Dual<B, P> r; r.v = a.v - b.v; for (int i = 0; i < P; ++i) r.d[i] = a.d[i] - b.d[i]; return r;
}
template <class B, int P> inline Dual<B, P> operator-(const Dual<B, P>& a) {
  Dual<B, P> r; r.v = -a.v; for (int i = 0; i < P; ++i) r.d[i] = -a.d[i]; return r;
}
template <class B, int P> inline Dual<B, P> operator*(const Dual<B, P>& a, const Dual<B, P>& b) {
  Dual<B, P> r; r.v = a.v * b.v; for (int i = 0; i < P; ++i) r.d[i] = a.d[i] * b.v + a.v * b.d[i]; return r;
}
template <class B, int P> inline Dual<B, P> operator/(const Dual<B, P>& a, const Dual<B, P>& b) {
  Dual<B, P> r; r.v = a.v / b.v;
  for (int i = 0; i < P; ++i) r.d[i] = (a.d[i] - r.v * b.d[i]) / b.v;
  return r;
}
#define ORACLE_DUAL_MIXED(op)                                                                   \
  template <class B, int P> inline Dual<B, P> operator op(const Dual<B, P>& a, double b) {      \
    return a op Dual<B, P>(static_cast<B>(b));                                                  \
  }                                                                                             \
  template <class B, int P> inline Dual<B, P> operator op(double a, const Dual<B, P>& b) {      \
    return Dual<B, P>(static_cast<B>(a)) op b;                                                  \
  }
ORACLE_DUAL_MIXED(+)
ORACLE_DUAL_MIXED(-)
ORACLE_DUAL_MIXED(*)
ORACLE_DUAL_MIXED(/)
#undef ORACLE_DUAL_MIXED
template <class B, int P> inline Dual<B, P>& operator+=(Dual<B, P>& a, const Dual<B, P>& b) { a = a + b; return a; }
template <class B, int P> inline Dual<B, P>& operator*=(Dual<B, P>& a, const Dual<B, P>& b) { a = a * b; return a; }
template <class B, int P> inline Dual<B, P>& operator/=(Dual<B, P>& a, const Dual<B, P>& b) { a = a / b; return a; }

// ---------------------------------------------------------------------------
// Traits: base (storage) type and primal value
// ---------------------------------------------------------------------------
template <class T> struct base_of { using type = T; };
template <class B, int P> struct base_of<Dual<B, P>> { using type = B; };
template <class T> using base_t = typename base_of<T>::type;

inline float pv(float x) { return x; }
inline double pv(double x) { return x; }
template <class B, int P> inline B pv(const Dual<B, P>& x) { return x.v; }

// A literal as the reference would see it: a weak-typed Python scalar that
// adopts the array dtype.  In float32 `1e-128` therefore becomes 0.0f, which
// is what the JAX reference compares against (SURVEY.md "f32 literal
// semantics").
template <class T> inline T lit(double c) { return T(static_cast<base_t<T>>(c)); }

// ---------------------------------------------------------------------------
// Elementary functions
// ---------------------------------------------------------------------------
inline float x_exp(float x) { return std::exp(x); }
inline double x_exp(double x) { return std::exp(x); }
inline float x_log(float x) { return std::log(x); }
inline double x_log(double x) { return std::log(x); }
inline float x_log1p(float x) { return std::log1p(x); }
inline double x_log1p(double x) { return std::log1p(x); }
inline float x_sin(float x) { return std::sin(x); }
inline double x_sin(double x) { return std::sin(x); }
inline float x_cos(float x) { return std::cos(x); }
inline double x_cos(double x) { return std::cos(x); }
inline float x_tanh(float x) { return std::tanh(x); }
inline double x_tanh(double x) { return std::tanh(x); }
inline float x_erf(float x) { return std::erf(x); }
inline double x_erf(double x) { return std::erf(x); }
inline float x_abs(float x) { return std::fabs(x); }
inline double x_abs(double x) { return std::fabs(x); }
inline float x_pow(float x, float y) { return std::pow(x, y); }
inline double x_pow(double x, double y) { return std::pow(x, y); }

template <class B, int P> inline Dual<B, P> x_exp(const Dual<B, P>& a) {
  Dual<B, P> r; r.v = std::exp(a.v); for (int i = 0; i < P; ++i) r.d[i] = r.v * a.d[i]; return r;
}
template <class B, int P> inline Dual<B, P> x_log(const Dual<B, P>& a) {
  Dual<B, P> r; r.v = std::log(a.v); for (int i = 0; i < P; ++i) r.d[i] = a.d[i] / a.v; return r;
}
template <class B, int P> inline Dual<B, P> x_log1p(const Dual<B, P>& a) {
  Dual<B, P> r; r.v = std::log1p(a.v); for (int i = 0; i < P; ++i) r.d[i] = a.d[i] / (B(1) + a.v); return r;
}
template <class B, int P> inline Dual<B, P> x_sin(const Dual<B, P>& a) {
  Dual<B, P> r; r.v = std::sin(a.v); B c = std::cos(a.v); for (int i = 0; i < P; ++i) r.d[i] = c * a.d[i]; return r;
}
template <class B, int P> inline Dual<B, P> x_cos(const Dual<B, P>& a) {
  Dual<B, P> r; r.v = std::cos(a.v); B s = -std::sin(a.v); for (int i = 0; i < P; ++i) r.d[i] = s * a.d[i]; return r;
}
template <class B, int P> inline Dual<B, P> x_tanh(const Dual<B, P>& a) {
  Dual<B, P> r; r.v = std::tanh(a.v); B c = B(1) - r.v * r.v; for (int i = 0; i < P; ++i) r.d[i] = c * a.d[i]; return r;
}
template <class B, int P> inline Dual<B, P> x_erf(const Dual<B, P>& a) {
  Dual<B, P> r; r.v = std::erf(a.v);
  B c = B(1.1283791670955126) * std::exp(-a.v * a.v);
  for (int i = 0; i < P; ++i) r.d[i] = c * a.d[i];
  return r;
}
template <class B, int P> inline Dual<B, P> x_abs(const Dual<B, P>& a) {
  // jnp.abs JVP: sign(x) * t
  Dual<B, P> r; r.v = std::fabs(a.v);
  B s = (a.v > 0) ? B(1) : ((a.v < 0) ? B(-1) : B(0));
  for (int i = 0; i < P; ++i) r.d[i] = s * a.d[i];
  return r;
}
template <class B, int P> inline Dual<B, P> x_pow(const Dual<B, P>& a, const Dual<B, P>& b) {
  // jnp.power JVP: y x^(y-1) dx + log(x) x^y dy   (log term only where dy != 0)
  Dual<B, P> r; r.v = std::pow(a.v, b.v);
  B da = b.v * std::pow(a.v, b.v - B(1));
  for (int i = 0; i < P; ++i) {
    B t = da * a.d[i];
    if (b.d[i] != B(0)) t += std::log(a.v) * r.v * b.d[i];
    r.d[i] = t;
  }
  return r;
}

// jnp.where(cond, a, b): the derivative follows the selected branch.
template <class T> inline T x_where(bool c, const T& a, const T& b) { return c ? a : b; }

// jnp.minimum / jnp.maximum: NaN-propagating; derivative follows the winner.
template <class T> inline T x_min(const T& a, const T& b) {
  if (std::isnan(pv(a))) return a;
  if (std::isnan(pv(b))) return b;
  return (pv(a) <= pv(b)) ? a : b;
}
template <class T> inline T x_max(const T& a, const T& b) {
  if (std::isnan(pv(a))) return a;
  if (std::isnan(pv(b))) return b;
  return (pv(a) >= pv(b)) ? a : b;
}
// jnp.clip(x, lo, hi) == minimum(hi, maximum(lo, x))
template <class T> inline T x_clip(const T& x, const T& lo, const T& hi) { return x_min(x_max(x, lo), hi); }

template <class T> inline bool x_isfinite(const T& a) { return std::isfinite(pv(a)); }
template <class T> inline bool x_isnan(const T& a) { return std::isnan(pv(a)); }

// jax.nn.sigmoid == lax.logistic, which XLA expands as 1 / (1 + exp(-x)).
template <class T> inline T x_sigmoid(const T& x) { return lit<T>(1.0) / (lit<T>(1.0) + x_exp(-x)); }

// jnp.logaddexp(x1, x2) = max + log1p(exp(-|x1 - x2|)), NaN delta -> x1 + x2.
inline float x_logaddexp(float a, float b) {
  float amax = (a >= b || std::isnan(a)) ? a : b;
  if (std::isnan(b)) amax = b;
  float delta = a - b;
  if (std::isnan(delta)) return a + b;
  return amax + std::log1p(std::exp(-std::fabs(delta)));
}
inline double x_logaddexp(double a, double b) {
  double amax = (a >= b || std::isnan(a)) ? a : b;
  if (std::isnan(b)) amax = b;
  double delta = a - b;
  if (std::isnan(delta)) return a + b;
  return amax + std::log1p(std::exp(-std::fabs(delta)));
}
template <class B, int P> inline Dual<B, P> x_logaddexp(const Dual<B, P>& a, const Dual<B, P>& b) {
  // custom JVP of jnp.logaddexp: t1 * exp(x1 - out) + t2 * exp(x2 - out)
  Dual<B, P> r; r.v = x_logaddexp(a.v, b.v);
  B wa = std::exp(a.v - r.v), wb = std::exp(b.v - r.v);
  for (int i = 0; i < P; ++i) r.d[i] = a.d[i] * wa + b.d[i] * wb;
  return r;
}

// pyhgf/math.py:450-488 -- principal Lambert W by 6 Halley steps from log(z+1).
template <class B> inline B lambert_w0_primal(B z) {
  B w = std::log(z + B(1));
  for (int it = 0; it < 6; ++it) {
    B ew = std::exp(w);
    B f = w * ew - z;
    B f1 = (w + B(1)) * ew;
    B f2 = (w + B(2)) * ew;
    w = w - (B(2) * f * f1) / (B(2) * (f1 * f1) - f * f2);
  }
  return w;
}
inline float x_lambert_w0(float z) { return lambert_w0_primal<float>(z); }
inline double x_lambert_w0(double z) { return lambert_w0_primal<double>(z); }
template <class B, int P> inline Dual<B, P> x_lambert_w0(const Dual<B, P>& z) {
  // pyhgf/math.py:491-498 -- custom JVP: z_dot / (exp(w) * (1 + w))
  Dual<B, P> r; r.v = lambert_w0_primal<B>(z.v);
  B den = std::exp(r.v) * (B(1) + r.v);
  for (int i = 0; i < P; ++i) r.d[i] = z.d[i] / den;
  return r;
}

}  // namespace oracle

// Scalar arithmetic of the belief-propagation kernels (sm_100a).
//
// Every update equation is written once, generic in the scalar type T:
//   * float / double            -- the filter itself;
//   * Dual<float|double, P>     -- the filter carrying P tangent directions, i.e.
//                                  d(surprise)/d(params) by forward-mode propagation
//                                  inside the same time loop (no tape, no checkpoints).
// The differentiation rules are JAX's, including the custom JVPs the reference
// registers for `lambert_w0` (pyhgf/math.py:491-498) and `jnp.logaddexp`, and
// the branch-following rules of where / minimum / maximum / clip.
//
// Literal semantics follow the reference's weak-typed Python scalars: a literal
// adopts the array dtype, so `1e-128` is 0.0f in float32.
//
// Math policy (float32 only; float64 always uses IEEE division and libdevice):
//   default            -- SFU primitives: division = MUFU.RCP x multiply, exp = MUFU.EX2,
//                         log = MUFU.LG2 (<= 2 ulp each), mul+add contraction allowed.  This is
//                         what makes the time loop SFU/FP32-pipe bound instead of instruction bound;
//                         parity with the reference stays inside the stated tolerances
//                         (tests/test_gpu_parity.py).
//   -DHGF_MATH_ACCURATE -- IEEE division, libdevice expf/logf/log1pf/powf and -fmad=false
//                         (one rounding per primitive, the oracle's arithmetic): the validation
//                         build, `HGF_MATH=accurate python -m pyhgf_b200.build`.
// Kernels never instantiate a bare float: the plain filter runs on Dual<B, 0>, so every
// operator below -- in particular operator/ -- is ours.
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>

#define HGF_HD __host__ __device__ __forceinline__
// The update equations are also compilable for the host: tests/ builds them into a
// debugging harness (never part of the shipped library) to check the device source
// against the oracle on machines without a GPU.
#define HGF_D __host__ __device__ __forceinline__
#ifdef __CUDA_ARCH__
#define HGF_LDG(p) __ldg(p)
#else
#define HGF_LDG(p) (*(p))
#endif

namespace hgf {

// ----------------------------------------------------------------------------
// Dual numbers
// ----------------------------------------------------------------------------
template <class B, int P>
struct Dual {
  B v;
  B d[P > 0 ? P : 1];  // P = 0: the plain filter (the unused element is never touched and is optimised away)
  HGF_HD Dual() {}
  HGF_HD Dual(B x) : v(x) {  // NOLINT(runtime/explicit)
#pragma unroll
    for (int i = 0; i < P; ++i) d[i] = B(0);
  }
};

template <class T> struct scalar_traits { using base = T; static constexpr int n_tangents = 0; };
template <class B, int P> struct scalar_traits<Dual<B, P>> { using base = B; static constexpr int n_tangents = P; };
template <class T> using base_t = typename scalar_traits<T>::base;

// The library is compiled with -fmad=false so that primal arithmetic rounds once per operation like the
// reference's; derivative (tangent) arithmetic has no such contract and uses fused multiply-adds explicitly.
HGF_HD float b_fma(float a, float b, float c) { return ::fmaf(a, b, c); }
HGF_HD double b_fma(double a, double b, double c) { return ::fma(a, b, c); }

#if defined(__CUDA_ARCH__) && !defined(HGF_MATH_ACCURATE)
#define HGF_SFU_MATH 1
#endif
// reciprocal and division
HGF_HD float b_rcp(float x) {
#ifdef HGF_SFU_MATH
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
#else
  return 1.0f / x;
#endif
}
HGF_HD double b_rcp(double x) { return 1.0 / x; }
HGF_HD float b_div(float a, float b) {
#ifdef HGF_SFU_MATH
  return a * b_rcp(b);
#else
  return a / b;
#endif
}
HGF_HD double b_div(double a, double b) { return a / b; }

HGF_HD float pv(float x) { return x; }
HGF_HD double pv(double x) { return x; }
template <class B, int P> HGF_HD B pv(const Dual<B, P>& x) { return x.v; }

template <class T> HGF_HD T lit(double c) { return T(static_cast<base_t<T>>(c)); }

template <class B, int P> HGF_HD Dual<B, P> operator+(const Dual<B, P>& a, const Dual<B, P>& b) {
  Dual<B, P> r; r.v = a.v + b.v;
#pragma unroll
  for (int i = 0; i < P; ++i) r.d[i] = a.d[i] + b.d[i];
  return r;
}
template <class B, int P> HGF_HD Dual<B, P> operator-(const Dual<B, P>& a, const Dual<B, P>& b) {
  Dual<B, P> r; r.v = a.v - b.v;
#pragma unroll
  for (int i = 0; i < P; ++i) r.d[i] = a.d[i] - b.d[i];
  return r;
}
template <class B, int P> HGF_HD Dual<B, P> operator-(const Dual<B, P>& a) {
  Dual<B, P> r; r.v = -a.v;
#pragma unroll
  for (int i = 0; i < P; ++i) r.d[i] = -a.d[i];
  return r;
}
template <class B, int P> HGF_HD Dual<B, P> operator*(const Dual<B, P>& a, const Dual<B, P>& b) {
  Dual<B, P> r; r.v = a.v * b.v;
#pragma unroll
  for (int i = 0; i < P; ++i) r.d[i] = b_fma(a.d[i], b.v, a.v * b.d[i]);
  return r;
}
template <class B, int P> HGF_HD Dual<B, P> operator/(const Dual<B, P>& a, const Dual<B, P>& b) {
  Dual<B, P> r;
  const B inv = b_rcp(b.v);
#ifdef HGF_SFU_MATH
  r.v = a.v * inv;
#else
  r.v = b_div(a.v, b.v);
#endif
#pragma unroll
  for (int i = 0; i < P; ++i) r.d[i] = b_fma(-r.v, b.d[i], a.d[i]) * inv;
  return r;
}
// scalar (tangent-free) operands
template <class B, int P> HGF_HD Dual<B, P> operator+(const Dual<B, P>& a, B b) { Dual<B, P> r = a; r.v = a.v + b; return r; }
template <class B, int P> HGF_HD Dual<B, P> operator+(B a, const Dual<B, P>& b) { Dual<B, P> r = b; r.v = a + b.v; return r; }
template <class B, int P> HGF_HD Dual<B, P> operator-(const Dual<B, P>& a, B b) { Dual<B, P> r = a; r.v = a.v - b; return r; }
template <class B, int P> HGF_HD Dual<B, P> operator-(B a, const Dual<B, P>& b) { Dual<B, P> r = -b; r.v = a - b.v; return r; }
template <class B, int P> HGF_HD Dual<B, P> operator*(const Dual<B, P>& a, B b) {
  Dual<B, P> r; r.v = a.v * b;
#pragma unroll
  for (int i = 0; i < P; ++i) r.d[i] = a.d[i] * b;
  return r;
}
template <class B, int P> HGF_HD Dual<B, P> operator*(B a, const Dual<B, P>& b) { return b * a; }
template <class B, int P> HGF_HD Dual<B, P> operator/(const Dual<B, P>& a, B b) {
  Dual<B, P> r; r.v = b_div(a.v, b);
  if constexpr (P > 0) {
    const B inv = b_rcp(b);
#pragma unroll
    for (int i = 0; i < P; ++i) r.d[i] = a.d[i] * inv;
  }
  return r;
}
template <class B, int P> HGF_HD Dual<B, P> operator/(B a, const Dual<B, P>& b) {
  Dual<B, P> r;
  const B inv = b_rcp(b.v);
#ifdef HGF_SFU_MATH
  r.v = a * inv;
#else
  r.v = b_div(a, b.v);
#endif
  if constexpr (P > 0) {
    const B c = -r.v * inv;
#pragma unroll
    for (int i = 0; i < P; ++i) r.d[i] = c * b.d[i];
  }
  return r;
}

// ----------------------------------------------------------------------------
// Elementary functions (float32: SFU primitives by default, libdevice with
// -DHGF_MATH_ACCURATE; see the math policy at the top of this file)
// ----------------------------------------------------------------------------
#ifdef HGF_SFU_MATH
HGF_D float sfu_ex2(float x) { float r; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
HGF_D float sfu_lg2(float x) { float r; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
HGF_D float b_exp(float x) { return sfu_ex2(x * 1.4426950408889634f); }
HGF_D float b_log(float x) { return sfu_lg2(x) * 0.6931471805599453f; }
HGF_D float b_log1p(float x) { return b_log(1.0f + x); }
HGF_D float b_pow(float x, float y) { return sfu_ex2(y * sfu_lg2(x)); }
#else
HGF_D float b_exp(float x) { return ::expf(x); }
HGF_D float b_log(float x) { return ::logf(x); }
HGF_D float b_log1p(float x) { return ::log1pf(x); }
HGF_D float b_pow(float x, float y) { return ::powf(x, y); }
#endif
HGF_D double b_exp(double x) { return ::exp(x); }
HGF_D double b_log(double x) { return ::log(x); }
HGF_D double b_log1p(double x) { return ::log1p(x); }
HGF_D float b_sin(float x) { return ::sinf(x); }
HGF_D double b_sin(double x) { return ::sin(x); }
HGF_D float b_cos(float x) { return ::cosf(x); }
HGF_D double b_cos(double x) { return ::cos(x); }
HGF_D float b_tanh(float x) { return ::tanhf(x); }
HGF_D double b_tanh(double x) { return ::tanh(x); }
HGF_D float b_erf(float x) { return ::erff(x); }
HGF_D double b_erf(double x) { return ::erf(x); }
HGF_D float b_abs(float x) { return ::fabsf(x); }
HGF_D double b_abs(double x) { return ::fabs(x); }
HGF_D double b_pow(double x, double y) { return ::pow(x, y); }
HGF_D bool b_isnan(float x) { return x != x; }
HGF_D bool b_isnan(double x) { return x != x; }
HGF_D bool b_isfinite(float x) { return ::fabsf(x) <= 3.402823466e+38f; }
HGF_D bool b_isfinite(double x) { return ::fabs(x) <= 1.7976931348623157e+308; }
template <class B> HGF_D B b_nan();
template <> HGF_D float b_nan<float>() { return __builtin_nanf(""); }
template <> HGF_D double b_nan<double>() { return __builtin_nan(""); }
template <class B> HGF_D B b_inf();
template <> HGF_D float b_inf<float>() { return __builtin_huge_valf(); }
template <> HGF_D double b_inf<double>() { return __builtin_huge_val(); }
// log(finfo(dtype).max): continuous_node_posterior_update_unbounded.py:126
template <class B> HGF_D B b_log_max();
template <> HGF_D float b_log_max<float>() { return 88.72283905206835f; }
template <> HGF_D double b_log_max<double>() { return 709.782712893384; }

HGF_D float x_exp(float x) { return b_exp(x); }
HGF_D double x_exp(double x) { return b_exp(x); }
HGF_D float x_log(float x) { return b_log(x); }
HGF_D double x_log(double x) { return b_log(x); }
HGF_D float x_sin(float x) { return b_sin(x); }
HGF_D double x_sin(double x) { return b_sin(x); }
HGF_D float x_cos(float x) { return b_cos(x); }
HGF_D double x_cos(double x) { return b_cos(x); }
HGF_D float x_tanh(float x) { return b_tanh(x); }
HGF_D double x_tanh(double x) { return b_tanh(x); }
HGF_D float x_erf(float x) { return b_erf(x); }
HGF_D double x_erf(double x) { return b_erf(x); }
HGF_D float x_pow(float x, float y) { return b_pow(x, y); }
HGF_D double x_pow(double x, double y) { return b_pow(x, y); }

template <class B, int P> HGF_D Dual<B, P> x_exp(const Dual<B, P>& a) {
  Dual<B, P> r; r.v = b_exp(a.v);
#pragma unroll
  for (int i = 0; i < P; ++i) r.d[i] = r.v * a.d[i];
  return r;
}
template <class B, int P> HGF_D Dual<B, P> x_log(const Dual<B, P>& a) {
  Dual<B, P> r; r.v = b_log(a.v);
  if constexpr (P > 0) {
    const B inv = b_rcp(a.v);
#pragma unroll
    for (int i = 0; i < P; ++i) r.d[i] = a.d[i] * inv;
  }
  return r;
}
template <class B, int P> HGF_D Dual<B, P> x_sin(const Dual<B, P>& a) {
  Dual<B, P> r; r.v = b_sin(a.v); const B c = b_cos(a.v);
#pragma unroll
  for (int i = 0; i < P; ++i) r.d[i] = c * a.d[i];
  return r;
}
template <class B, int P> HGF_D Dual<B, P> x_cos(const Dual<B, P>& a) {
  Dual<B, P> r; r.v = b_cos(a.v); const B s = -b_sin(a.v);
#pragma unroll
  for (int i = 0; i < P; ++i) r.d[i] = s * a.d[i];
  return r;
}
template <class B, int P> HGF_D Dual<B, P> x_tanh(const Dual<B, P>& a) {
  Dual<B, P> r; r.v = b_tanh(a.v); const B c = B(1) - r.v * r.v;
#pragma unroll
  for (int i = 0; i < P; ++i) r.d[i] = c * a.d[i];
  return r;
}
template <class B, int P> HGF_D Dual<B, P> x_erf(const Dual<B, P>& a) {
  Dual<B, P> r; r.v = b_erf(a.v);
  const B c = B(1.1283791670955126) * b_exp(-a.v * a.v);
#pragma unroll
  for (int i = 0; i < P; ++i) r.d[i] = c * a.d[i];
  return r;
}
template <class B, int P> HGF_D Dual<B, P> x_pow(const Dual<B, P>& a, const Dual<B, P>& b) {
  // jnp.power JVP: y x^(y-1) dx + log(x) x^y dy
  Dual<B, P> r; r.v = b_pow(a.v, b.v);
  if constexpr (P > 0) {
#ifdef HGF_SFU_MATH
    const B da = b.v * r.v * b_rcp(a.v);
#else
    const B da = b.v * b_pow(a.v, b.v - B(1));
#endif
    const B db = b_log(a.v) * r.v;
#pragma unroll
    for (int i = 0; i < P; ++i) {
      B t = da * a.d[i];
      if (b.d[i] != B(0)) t = b_fma(db, b.d[i], t);
      r.d[i] = t;
    }
  }
  return r;
}

// jnp.where: the derivative follows the selected branch.
template <class T> HGF_D T x_where(bool c, const T& a, const T& b) { return c ? a : b; }

// jnp.minimum / jnp.maximum: NaN-propagating, derivative follows the winner.
template <class T> HGF_D T x_min(const T& a, const T& b) {
  if (b_isnan(pv(a))) return a;
  if (b_isnan(pv(b))) return b;
  return (pv(a) <= pv(b)) ? a : b;
}
template <class T> HGF_D T x_max(const T& a, const T& b) {
  if (b_isnan(pv(a))) return a;
  if (b_isnan(pv(b))) return b;
  return (pv(a) >= pv(b)) ? a : b;
}
template <class T> HGF_D T x_clip(const T& x, const T& lo, const T& hi) { return x_min(x_max(x, lo), hi); }
template <class T> HGF_D bool x_isfinite(const T& a) { return b_isfinite(pv(a)); }
template <class T> HGF_D bool x_isnan(const T& a) { return b_isnan(pv(a)); }

// jax.nn.sigmoid == lax.logistic == 1 / (1 + exp(-x))
HGF_D float x_sigmoid(float x) { return b_div(1.0f, 1.0f + b_exp(-x)); }
HGF_D double x_sigmoid(double x) { return 1.0 / (1.0 + b_exp(-x)); }
template <class B, int P> HGF_D Dual<B, P> x_sigmoid(const Dual<B, P>& x) {
  // primal exactly as above; tangent in closed form, s (1 - s) dx, instead of differentiating exp and the division
  Dual<B, P> r;
#ifdef HGF_SFU_MATH
  r.v = b_rcp(B(1) + b_exp(-x.v));
#else
  r.v = b_div(B(1), B(1) + b_exp(-x.v));
#endif
  if constexpr (P > 0) {
    const B c = r.v * (B(1) - r.v);
#pragma unroll
    for (int i = 0; i < P; ++i) r.d[i] = c * x.d[i];
  }
  return r;
}

// jnp.logaddexp(a, b) = max + log1p(exp(-|a - b|)); NaN delta -> a + b
template <class B> HGF_D B b_logaddexp(B a, B b) {
  B amax = (a >= b || b_isnan(a)) ? a : b;
  if (b_isnan(b)) amax = b;
  const B delta = a - b;
  if (b_isnan(delta)) return a + b;
  return amax + b_log1p(b_exp(-b_abs(delta)));
}
HGF_D float x_logaddexp(float a, float b) { return b_logaddexp<float>(a, b); }
HGF_D double x_logaddexp(double a, double b) { return b_logaddexp<double>(a, b); }
template <class B, int P> HGF_D Dual<B, P> x_logaddexp(const Dual<B, P>& a, const Dual<B, P>& b) {
  // JVP of jnp.logaddexp: t1 * exp(x1 - out) + t2 * exp(x2 - out)
  Dual<B, P> r; r.v = b_logaddexp<B>(a.v, b.v);
  if constexpr (P > 0) {
    // the weights exp(x1 - out), exp(x2 - out) sum to one: w(larger) = 1 / (1 + exp(-|x1 - x2|)), w(smaller) = 1 - it
    // (one reciprocal instead of two exponentials); a NaN anywhere gives NaN weights, as the JVP's would be
    const B e = b_exp(-b_abs(a.v - b.v));
    const B w_hi = b_rcp(B(1) + e), w_lo = B(1) - w_hi;
    const bool a_hi = a.v >= b.v;
    const B wa = a_hi ? w_hi : w_lo, wb = a_hi ? w_lo : w_hi;
#pragma unroll
    for (int i = 0; i < P; ++i) r.d[i] = b_fma(a.d[i], wa, b.d[i] * wb);
  }
  return r;
}

// pyhgf/math.py:450-488 -- principal Lambert W: w0 = log(z + 1), six Halley steps.
template <class B> HGF_D B b_lambert_w0(B z) {
  B w = b_log(z + B(1));
#pragma unroll
  for (int it = 0; it < 6; ++it) {
    const B ew = b_exp(w);
    const B f = w * ew - z;
    const B f1 = (w + B(1)) * ew;
    const B f2 = (w + B(2)) * ew;
    w = w - b_div(B(2) * f * f1, B(2) * (f1 * f1) - f * f2);
  }
  return w;
}
HGF_D float x_lambert_w0(float z) { return b_lambert_w0<float>(z); }
HGF_D double x_lambert_w0(double z) { return b_lambert_w0<double>(z); }
template <class B, int P> HGF_D Dual<B, P> x_lambert_w0(const Dual<B, P>& z) {
  // custom JVP, pyhgf/math.py:491-498: z_dot / (exp(w) (1 + w))
  Dual<B, P> r; r.v = b_lambert_w0<B>(z.v);
  if constexpr (P > 0) {
    const B inv = b_rcp(b_exp(r.v) * (B(1) + r.v));
#pragma unroll
    for (int i = 0; i < P; ++i) r.d[i] = z.d[i] * inv;
  }
  return r;
}

}  // namespace hgf

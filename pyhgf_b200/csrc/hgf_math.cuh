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
// Math policy (float32 device code only; float64 and host code always use IEEE division and libm/libdevice).
// Four independent knobs, one per primitive class, chosen at compile time (-DHGF_<X>_MODE=n); what the shipped
// library uses, and what each costs in parity and time, is measured by tools/parity_ablation.py
// (profiles/r2_parity_ablation.json):
//   HGF_DIV_MODE  0  a * MUFU.RCP(b)                      (2 roundings, <= 2 ulp)
//                 1  MUFU.RCP + one Newton step on the quotient (2 FMA + NaN guard; <= 1 ulp, special values kept)
//                 2  IEEE division (div.rn)
//   HGF_EXP_MODE  0  MUFU.EX2(x * log2 e)                 (argument rounding: relative error ~ |x| 2^-24, + 2 ulp)
//                 2  libdevice expf (Cody-Waite reduction around the same MUFU.EX2, <= 2 ulp over the whole range)
//   HGF_LOG_MODE  0  MUFU.LG2(x) * ln 2, log1p(x) = log(1 + x)   (absolute error 2^-22 near 1)
//                 1  libdevice logf; log1p(x) = logf(u) + (x - (u - 1)) / u with u = 1 + x (first-order repair of the
//                    rounding of 1 + x: one MUFU.RCP instead of log1pf's second polynomial)
//                 2  libdevice logf and log1pf
//   HGF_POW_MODE  0  MUFU.EX2(y * MUFU.LG2(x))            (only the inverse-temperature response uses pow; its
//                 2  libdevice powf                         result does not feed back into the beliefs)
//   HGF_FTZ       1  the MUFU wrappers flush subnormals (.ftz); 0: they do not
//   HGF_RESP_MODE 0  the response functions and the binary node's sigmoid follow the knobs above
//                 1  libdevice powf + an IEEE-accurate quotient in the inverse-temperature response, an IEEE-accurate
//                    reciprocal in the binary node's sigmoid: a probability next to 1 whose complement is taken
//                    afterwards (-log(1 - p)) needs its LAST BIT -- one ulp of p is per cents of 1 - p ~ 1e-6
//                 2  the same, but powf only when a lane of the warp is in that regime (warp-uniform branch)
// Shipped: DIV/EXP/LOG/POW = 0, FTZ = 1, RESP = 2.  Measured (profiles/r2_parity_ablation.json): RESP is what the
// round-1 build lacked (260 of 32 768 instances of the omega in [-6, -3] box off by more than 1e-4, 3-6 with it --
// the same count, and mostly the same instances, as the all-IEEE validation build, which is 2.7x slower); making
// division, exp or log more accurate on top of it changes nothing measurable and costs 13-29 % each.
// Tangent (derivative) arithmetic has no rounding contract with the reference: it always takes MUFU.RCP
// reciprocals and explicit FMAs, except under -DHGF_MATH_ACCURATE.
//   -DHGF_MATH_ACCURATE -- every knob at 2, IEEE reciprocals for the tangents too, and (build.py) -fmad=false:
//                         one rounding per primitive, the oracle's arithmetic.  The validation build,
//                         `HGF_MATH=accurate python -m pyhgf_b200.build`.
// Kernels never instantiate a bare float: the plain filter runs on Dual<B, 0>, so every
// operator below -- in particular operator/ -- is ours.
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#ifndef __CUDACC__
#include <math.h>  // a plain C++ compiler (tools/hostsim) does not get the math functions from cuda_runtime.h
#endif

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
// Tangent arithmetic uses fused multiply-adds explicitly (the primal's contraction is the compiler's: on in the
// shipped build, off with -fmad=false in the validation build).
HGF_HD float b_fma(float a, float b, float c) { return ::fmaf(a, b, c); }
HGF_HD double b_fma(double a, double b, double c) { return ::fma(a, b, c); }

// The tangent vector of a Dual.  Generic form: P scalars, every operation a loop over them.
// PACKED form (float, P >= 2; HGF_PACKED_TANGENTS): the tangents live in 64-bit register PAIRS and every operation is one
// Blackwell packed-FP32 instruction per pair -- fma / mul / add .rn.f32x2 (SASS FFMA2 / FMUL2 / FADD2), which take the scalar
// multiplier as a broadcast operand.  Lane throughput is that of two scalar instructions (tools/ubench/pipes.cu), but it
// is ONE issue slot, and the logp+grad kernels are issue-bound: the 3-level binary HGF with two filter tangents executes
// 797 warp-instructions per instance-step unpacked, most of them tangent arithmetic on the FP32 pipe.  Same IEEE
// operations in the same order as the generic form (an odd P keeps its last tangent as a scalar).
#ifndef HGF_PACKED_TANGENTS
#define HGF_PACKED_TANGENTS 1
#endif
struct F2 { unsigned long long u; };  // two float32 {lo, hi}
HGF_HD F2 f2_make(float lo, float hi) {
  F2 r;
#ifdef __CUDA_ARCH__
  asm("mov.b64 %0, {%1, %2};" : "=l"(r.u) : "f"(lo), "f"(hi));
#else
  unsigned int a, b;
  __builtin_memcpy(&a, &lo, 4); __builtin_memcpy(&b, &hi, 4);
  r.u = static_cast<unsigned long long>(a) | (static_cast<unsigned long long>(b) << 32);
#endif
  return r;
}
HGF_HD float f2_lo(F2 x) {
#ifdef __CUDA_ARCH__
  float lo; asm("mov.b64 {%0, _}, %1;" : "=f"(lo) : "l"(x.u)); return lo;
#else
  const unsigned int a = static_cast<unsigned int>(x.u); float f; __builtin_memcpy(&f, &a, 4); return f;
#endif
}
HGF_HD float f2_hi(F2 x) {
#ifdef __CUDA_ARCH__
  float hi; asm("mov.b64 {_, %0}, %1;" : "=f"(hi) : "l"(x.u)); return hi;
#else
  const unsigned int a = static_cast<unsigned int>(x.u >> 32); float f; __builtin_memcpy(&f, &a, 4); return f;
#endif
}
HGF_HD F2 f2_add(F2 a, F2 b) {
#ifdef __CUDA_ARCH__
  F2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r.u) : "l"(a.u), "l"(b.u)); return r;
#else
  return f2_make(f2_lo(a) + f2_lo(b), f2_hi(a) + f2_hi(b));
#endif
}
HGF_HD F2 f2_sub(F2 a, F2 b) {
#ifdef __CUDA_ARCH__
  F2 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r.u) : "l"(a.u), "l"(b.u)); return r;
#else
  return f2_make(f2_lo(a) - f2_lo(b), f2_hi(a) - f2_hi(b));
#endif
}
HGF_HD F2 f2_scale(F2 a, float s) {  // a * s
#ifdef __CUDA_ARCH__
  F2 r, b; asm("mov.b64 %0, {%1, %1};" : "=l"(b.u) : "f"(s)); asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r.u) : "l"(a.u), "l"(b.u)); return r;
#else
  return f2_make(f2_lo(a) * s, f2_hi(a) * s);
#endif
}
HGF_HD F2 f2_fma(F2 a, float s, F2 c) {  // fma(a, s, c)
#ifdef __CUDA_ARCH__
  F2 r, b; asm("mov.b64 %0, {%1, %1};" : "=l"(b.u) : "f"(s)); asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.u) : "l"(a.u), "l"(b.u), "l"(c.u)); return r;
#else
  return f2_make(::fmaf(f2_lo(a), s, f2_lo(c)), ::fmaf(f2_hi(a), s, f2_hi(c)));
#endif
}

template <class B, int P> struct use_packed { static constexpr bool value = false; };
#if HGF_PACKED_TANGENTS
template <int P> struct use_packed<float, P> { static constexpr bool value = P >= 2; };
#endif

template <class B, int P, bool PACKED = use_packed<B, P>::value> struct Tan;
template <class B, int P> struct Tan<B, P, false> {
  B e[P > 0 ? P : 1];  // P = 0: the plain filter (the unused element is never touched and is optimised away)
  HGF_HD B get(int i) const { return e[i]; }
  HGF_HD void set(int i, B x) { e[i] = x; }
  HGF_HD void zero() {
#pragma unroll
    for (int i = 0; i < P; ++i) e[i] = B(0);
  }
  HGF_HD void add(const Tan& a, const Tan& b) {  // a + b
#pragma unroll
    for (int i = 0; i < P; ++i) e[i] = a.e[i] + b.e[i];
  }
  HGF_HD void sub(const Tan& a, const Tan& b) {  // a - b
#pragma unroll
    for (int i = 0; i < P; ++i) e[i] = a.e[i] - b.e[i];
  }
  HGF_HD void neg(const Tan& a) {
#pragma unroll
    for (int i = 0; i < P; ++i) e[i] = -a.e[i];
  }
  HGF_HD void scale(const Tan& a, B s) {  // a * s
#pragma unroll
    for (int i = 0; i < P; ++i) e[i] = a.e[i] * s;
  }
  HGF_HD void product(const Tan& a, B bv, B av, const Tan& b) {  // fma(a, bv, av * b): the product rule
#pragma unroll
    for (int i = 0; i < P; ++i) e[i] = b_fma(a.e[i], bv, av * b.e[i]);
  }
  HGF_HD void quotient(B q, const Tan& b, const Tan& a, B inv) {  // fma(-q, b, a) * inv: the quotient rule
#pragma unroll
    for (int i = 0; i < P; ++i) e[i] = b_fma(-q, b.e[i], a.e[i]) * inv;
  }
  HGF_HD void blend(const Tan& a, B wa, const Tan& b, B wb) {  // fma(a, wa, b * wb)
#pragma unroll
    for (int i = 0; i < P; ++i) e[i] = b_fma(a.e[i], wa, b.e[i] * wb);
  }
  template <class T2> HGF_HD void widen_from(const T2& a, int n) {  // the first n tangents of a (the others keep their value)
#pragma unroll
    for (int i = 0; i < P; ++i)
      if (i < n) e[i] = a.get(i);
  }
};
template <int P> struct Tan<float, P, true> {
  static constexpr int NP = P / 2;        // whole pairs
  static constexpr bool ODD = (P & 1) != 0;  // an odd P keeps its last tangent as a scalar (a padded pair would spend a lane)
  F2 e[NP];
  float last;
  HGF_HD float get(int i) const {
    if (ODD && i == P - 1) return last;
    return (i & 1) ? f2_hi(e[i >> 1]) : f2_lo(e[i >> 1]);
  }
  HGF_HD void set(int i, float x) {
    if (ODD && i == P - 1) { last = x; return; }
    F2& p = e[i >> 1];
    p = (i & 1) ? f2_make(f2_lo(p), x) : f2_make(x, f2_hi(p));
  }
  HGF_HD void zero() {
#pragma unroll
    for (int k = 0; k < NP; ++k) e[k].u = 0ull;
    last = 0.0f;
  }
  HGF_HD void add(const Tan& a, const Tan& b) {
#pragma unroll
    for (int k = 0; k < NP; ++k) e[k] = f2_add(a.e[k], b.e[k]);
    if constexpr (ODD) last = a.last + b.last;
  }
  HGF_HD void sub(const Tan& a, const Tan& b) {
#pragma unroll
    for (int k = 0; k < NP; ++k) e[k] = f2_sub(a.e[k], b.e[k]);
    if constexpr (ODD) last = a.last - b.last;
  }
  HGF_HD void neg(const Tan& a) {  // (a * -1: an exact sign flip, zeros included)
#pragma unroll
    for (int k = 0; k < NP; ++k) e[k] = f2_scale(a.e[k], -1.0f);
    if constexpr (ODD) last = -a.last;
  }
  HGF_HD void scale(const Tan& a, float s) {
#pragma unroll
    for (int k = 0; k < NP; ++k) e[k] = f2_scale(a.e[k], s);
    if constexpr (ODD) last = a.last * s;
  }
  HGF_HD void product(const Tan& a, float bv, float av, const Tan& b) {
#pragma unroll
    for (int k = 0; k < NP; ++k) e[k] = f2_fma(a.e[k], bv, f2_scale(b.e[k], av));
    if constexpr (ODD) last = b_fma(a.last, bv, av * b.last);
  }
  HGF_HD void quotient(float q, const Tan& b, const Tan& a, float inv) {
#pragma unroll
    for (int k = 0; k < NP; ++k) e[k] = f2_scale(f2_fma(b.e[k], -q, a.e[k]), inv);
    if constexpr (ODD) last = b_fma(-q, b.last, a.last) * inv;
  }
  HGF_HD void blend(const Tan& a, float wa, const Tan& b, float wb) {
#pragma unroll
    for (int k = 0; k < NP; ++k) e[k] = f2_fma(a.e[k], wa, f2_scale(b.e[k], wb));
    if constexpr (ODD) last = b_fma(a.last, wa, b.last * wb);
  }
  template <class T2> HGF_HD void widen_from(const T2& a, int n) {  // the first n tangents of a (the others keep their value)
#pragma unroll
    for (int i = 0; i < P; ++i)
      if (i < n) set(i, a.get(i));
  }
  template <int P2> HGF_HD void widen_from(const Tan<float, P2, true>& a, int n) {  // whole pairs move as pairs
    constexpr int NP2 = P2 / 2;
#pragma unroll
    for (int k = 0; k < NP; ++k)
      if (k < NP2 && 2 * k + 1 < n) e[k] = a.e[k];
#pragma unroll
    for (int i = 0; i < P; ++i)
      if (i < n && i < P2 && !((i >> 1) < NP && (i >> 1) < NP2 && 2 * (i >> 1) + 1 < n)) set(i, a.get(i));
  }
};

template <class B, int P>
struct Dual {
  B v;
  Tan<B, P> d;
  HGF_HD Dual() {}
  HGF_HD Dual(B x) : v(x) { d.zero(); }  // NOLINT(runtime/explicit)
};

template <class T> struct scalar_traits { using base = T; static constexpr int n_tangents = 0; };
template <class B, int P> struct scalar_traits<Dual<B, P>> { using base = B; static constexpr int n_tangents = P; };
template <class T> using base_t = typename scalar_traits<T>::base;


#ifdef HGF_MATH_ACCURATE
#undef HGF_DIV_MODE
#undef HGF_EXP_MODE
#undef HGF_LOG_MODE
#undef HGF_POW_MODE
#define HGF_DIV_MODE 2
#define HGF_EXP_MODE 2
#define HGF_LOG_MODE 2
#define HGF_POW_MODE 2
#endif
#if !defined(__CUDA_ARCH__) && !defined(HGF_HOST_EMULATE_SFU)  // host compilation of the equations: IEEE / libm throughout
#undef HGF_DIV_MODE
#undef HGF_EXP_MODE
#undef HGF_LOG_MODE
#undef HGF_POW_MODE
#define HGF_DIV_MODE 2
#define HGF_EXP_MODE 2
#define HGF_LOG_MODE 2
#define HGF_POW_MODE 2
#endif
#ifndef HGF_DIV_MODE
#define HGF_DIV_MODE 0
#endif
#ifndef HGF_EXP_MODE
#define HGF_EXP_MODE 0
#endif
#ifndef HGF_LOG_MODE
#define HGF_LOG_MODE 0
#endif
#ifndef HGF_POW_MODE
#define HGF_POW_MODE 0
#endif
#ifndef HGF_FTZ
#define HGF_FTZ 1
#endif
#ifndef HGF_RESP_MODE  // last-bit arithmetic for probabilities whose complement is taken: hgf_ops.cuh, response_step
#define HGF_RESP_MODE 2
#endif
#if defined(__CUDA_ARCH__) && !defined(HGF_MATH_ACCURATE)
#define HGF_SFU_TANGENTS 1  // tangents take MUFU.RCP reciprocals
#endif

// raw SFU primitives (device only)
#ifdef __CUDA_ARCH__
#if HGF_FTZ
HGF_D float sfu_rcp(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
HGF_D float sfu_ex2(float x) { float r; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
HGF_D float sfu_lg2(float x) { float r; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
#else
HGF_D float sfu_rcp(float x) { float r; asm("rcp.approx.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
HGF_D float sfu_ex2(float x) { float r; asm("ex2.approx.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
HGF_D float sfu_lg2(float x) { float r; asm("lg2.approx.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
#endif
#elif defined(HGF_HOST_EMULATE_SFU)
// tools/hostsim: the SFU primitives EMULATED on the host for numerics experiments without a GPU -- the correctly
// rounded result moved by a deterministic pseudo-random amount inside the documented error bound of the PTX
// instruction (rcp.approx 1 ulp; ex2.approx 2 ulp; lg2.approx 2^-22 absolute in (0.5, 2), relative elsewhere).
// Like the hardware it is a pure function of the input bits; it is NOT the hardware's function.
static inline float sfu_emu_jitter(float x) {  // in [-1, 1)
  unsigned u; __builtin_memcpy(&u, &x, 4);
  u ^= u >> 16; u *= 0x7feb352du; u ^= u >> 15; u *= 0x846ca68bu; u ^= u >> 16;
  return static_cast<float>(static_cast<int>(u >> 8) - (1 << 23)) * (1.0f / (1 << 23));
}
static inline float sfu_emu_rel(double exact, float key, double ulps) {
  const float r = static_cast<float>(exact);
  if (!(__builtin_fabsf(r) < 3.0e38f) || r == 0.0f) return r;
  return static_cast<float>(exact * (1.0 + ulps * 5.9604645e-8 * sfu_emu_jitter(key)));
}
static inline float sfu_rcp(float x) { return sfu_emu_rel(1.0 / static_cast<double>(x), x, 1.0); }
static inline float sfu_ex2(float x) { return sfu_emu_rel(::exp2(static_cast<double>(x)), x, 2.0); }
static inline float sfu_lg2(float x) {
  const double e = ::log2(static_cast<double>(x));
  if (x > 0.5f && x < 2.0f) return static_cast<float>(e + 2.3841858e-7 * sfu_emu_jitter(x));
  return sfu_emu_rel(e, x, 4.0);
}
#define HGF_SFU_TANGENTS 1
#else
HGF_D float sfu_rcp(float x) { return 1.0f / x; }
#endif

// reciprocal of a tangent-grade quantity
HGF_HD float t_rcp(float x) {
#ifdef HGF_SFU_TANGENTS
  return sfu_rcp(x);
#else
  return 1.0f / x;
#endif
}
HGF_HD double t_rcp(double x) { return 1.0 / x; }

// a / b given r0 = sfu_rcp(b): the primal's division under the policy
HGF_HD float div_from_rcp(float a, float b, float r0) {
#if HGF_DIV_MODE == 0
  (void)b;
  return a * r0;
#elif HGF_DIV_MODE == 1
  // one Newton step on the quotient; the residual is NaN exactly where q is already the answer (b = 0, inf;
  // q overflowed), so a NaN refinement falls back to q and special values flow as they do in a true division
  const float q = a * r0;
  const float q1 = ::fmaf(::fmaf(-b, q, a), r0, q);
  return (q1 == q1) ? q1 : q;
#else
  (void)r0;
  return a / b;
#endif
}
HGF_HD float b_rcp(float x) {
#if HGF_DIV_MODE == 2
  return 1.0f / x;
#else
  return div_from_rcp(1.0f, x, sfu_rcp(x));
#endif
}
HGF_HD double b_rcp(double x) { return 1.0 / x; }
HGF_HD float b_div(float a, float b) {
#if HGF_DIV_MODE == 2
  return a / b;
#else
  return div_from_rcp(a, b, sfu_rcp(b));
#endif
}
HGF_HD double b_div(double a, double b) { return a / b; }
// a / b to IEEE accuracy whatever the policy (MUFU.RCP + Newton step, or div.rn): for the few quotients whose
// LAST BIT matters -- a probability next to 1 whose complement is taken afterwards (1 - p), see the response
// functions in hgf_ops.cuh.  On this path's data the Newton form returned the IEEE quotient everywhere
// (profiles/r2_parity_ablation.json: div1 == div2 to the last digit).
HGF_HD float exact_div_from_rcp(float a, float b, float r0) {
#if HGF_DIV_MODE == 2
  (void)r0;
  return a / b;
#else
  const float q = a * r0;
  const float q1 = ::fmaf(::fmaf(-b, q, a), r0, q);
  return (q1 == q1) ? q1 : q;
#endif
}
HGF_HD float exact_rcp(float x) {
#if HGF_DIV_MODE == 2
  return 1.0f / x;
#else
  return exact_div_from_rcp(1.0f, x, sfu_rcp(x));
#endif
}
HGF_HD double exact_rcp(double x) { return 1.0 / x; }
HGF_HD float exact_quot(float a, float b) {
#if HGF_DIV_MODE == 2
  return a / b;
#else
  return exact_div_from_rcp(a, b, sfu_rcp(b));
#endif
}
HGF_HD double exact_quot(double a, double b) { return a / b; }

HGF_HD float pv(float x) { return x; }
HGF_HD double pv(double x) { return x; }
template <class B, int P> HGF_HD B pv(const Dual<B, P>& x) { return x.v; }

template <class T> HGF_HD T lit(double c) { return T(static_cast<base_t<T>>(c)); }

template <class B, int P> HGF_HD Dual<B, P> operator+(const Dual<B, P>& a, const Dual<B, P>& b) {
  Dual<B, P> r; r.v = a.v + b.v;
  r.d.add(a.d, b.d);
  return r;
}
template <class B, int P> HGF_HD Dual<B, P> operator-(const Dual<B, P>& a, const Dual<B, P>& b) {
  Dual<B, P> r; r.v = a.v - b.v;
  r.d.sub(a.d, b.d);
  return r;
}
template <class B, int P> HGF_HD Dual<B, P> operator-(const Dual<B, P>& a) {
  Dual<B, P> r; r.v = -a.v;
  r.d.neg(a.d);
  return r;
}
template <class B, int P> HGF_HD Dual<B, P> operator*(const Dual<B, P>& a, const Dual<B, P>& b) {
  Dual<B, P> r; r.v = a.v * b.v;
  r.d.product(a.d, b.v, a.v, b.d);
  return r;
}
// a / b for the primal of a Dual, and the tangent-grade reciprocal of b for its derivative: with the SFU
// policies both come from ONE MUFU.RCP; `tangents(inv)` is only run when the Dual carries tangents.
template <class F> HGF_HD void dual_div_primal(float a, float b, float& q, bool has_tangents, F tangents) {
#if defined(HGF_SFU_TANGENTS) && HGF_DIV_MODE != 2
  const float r0 = sfu_rcp(b);
  q = div_from_rcp(a, b, r0);
  if (has_tangents) tangents(r0);
#else
  q = b_div(a, b);
  if (has_tangents) tangents(t_rcp(b));
#endif
}
template <class F> HGF_HD void dual_div_primal(double a, double b, double& q, bool has_tangents, F tangents) {
  q = a / b;
  if (has_tangents) tangents(1.0 / b);
}
template <class B, int P> HGF_HD Dual<B, P> operator/(const Dual<B, P>& a, const Dual<B, P>& b) {
  Dual<B, P> r;
  dual_div_primal(a.v, b.v, r.v, P > 0, [&](B inv) { r.d.quotient(r.v, b.d, a.d, inv); });
  return r;
}
// A denominator together with its MUFU.RCP seed, so that SEVERAL quotients by the same value -- or by twice it -- cost ONE SFU
// operation: a / x and c / x, and a / (2 x) through rcp(2 x) = rcp(x) / 2 (the approximation acts on the mantissa: a power of
// two changes the exponent only, so the shared seed gives the bits a seed of its own would: tools/ubench/rcp_scaling.cu finds
// 0 mismatches over every mantissa, sign and exponent with |x| < 2^125, profiles/r2_rcp_scaling.log; beyond, rcp(2 x) is
// subnormal and flushed -- the denominators here are precisions capped at max_posterior_precision = 1e10).  Under the IEEE policies (HGF_DIV_MODE 2, the validation build, double, the host build without
// SFU emulation) the seed is unused and `a / recip` is the true division a / x: nothing changes there.  (What the KERNELS
// return with and without the sharing -- identical gradient kernels, a plain filter that rounds one product elsewhere:
// tools/shared_rcp_check.py.)
#if defined(HGF_SFU_TANGENTS) && HGF_DIV_MODE != 2 && !defined(HGF_NO_SHARED_RCP)  // (-DHGF_NO_SHARED_RCP: the test's twin)
#define HGF_SHARED_RCP 1
#else
#define HGF_SHARED_RCP 0
#endif
template <class T> struct Recip {
  T x;
  base_t<T> r0;
};
template <class B, int P> HGF_HD Recip<Dual<B, P>> x_recip(const Dual<B, P>& x) {
  Recip<Dual<B, P>> r;
  r.x = x;
  r.r0 = B(0);
#if HGF_SHARED_RCP
  if constexpr (sizeof(B) == 4) r.r0 = sfu_rcp(x.v);
#endif
  return r;
}
template <class B, int P> HGF_HD Recip<Dual<B, P>> recip_twice(const Recip<Dual<B, P>>& a) {  // of 2 x
  Recip<Dual<B, P>> r;
  r.x = B(2) * a.x;
  r.r0 = B(0.5) * a.r0;
  return r;
}
template <class T> HGF_HD Recip<T> x_where(bool c, const Recip<T>& a, const Recip<T>& b) { return c ? a : b; }
template <class B, int P> HGF_HD Dual<B, P> operator/(const Dual<B, P>& a, const Recip<Dual<B, P>>& b) {
#if HGF_SHARED_RCP
  if constexpr (sizeof(B) == 4) {
    Dual<B, P> r;
    r.v = div_from_rcp(a.v, b.x.v, b.r0);
    if constexpr (P > 0) r.d.quotient(r.v, b.x.d, a.d, b.r0);
    return r;
  } else {
    return a / b.x;
  }
#else
  return a / b.x;
#endif
}

// a / b with the primal quotient at IEEE accuracy under every policy (see exact_div_from_rcp)
template <int P> HGF_HD Dual<float, P> exact_div(const Dual<float, P>& a, const Dual<float, P>& b) {
  Dual<float, P> r;
#if defined(HGF_SFU_TANGENTS) && HGF_DIV_MODE != 2
  const float inv = sfu_rcp(b.v);
  r.v = exact_div_from_rcp(a.v, b.v, inv);
#else
  const float inv = t_rcp(b.v);
  r.v = a.v / b.v;
#endif
  r.d.quotient(r.v, b.d, a.d, inv);
  return r;
}
template <int P> HGF_HD Dual<double, P> exact_div(const Dual<double, P>& a, const Dual<double, P>& b) { return a / b; }
// scalar (tangent-free) operands
template <class B, int P> HGF_HD Dual<B, P> operator+(const Dual<B, P>& a, B b) { Dual<B, P> r = a; r.v = a.v + b; return r; }
template <class B, int P> HGF_HD Dual<B, P> operator+(B a, const Dual<B, P>& b) { Dual<B, P> r = b; r.v = a + b.v; return r; }
template <class B, int P> HGF_HD Dual<B, P> operator-(const Dual<B, P>& a, B b) { Dual<B, P> r = a; r.v = a.v - b; return r; }
template <class B, int P> HGF_HD Dual<B, P> operator-(B a, const Dual<B, P>& b) { Dual<B, P> r = -b; r.v = a - b.v; return r; }
template <class B, int P> HGF_HD Dual<B, P> operator*(const Dual<B, P>& a, B b) {
  Dual<B, P> r; r.v = a.v * b;
  r.d.scale(a.d, b);
  return r;
}
template <class B, int P> HGF_HD Dual<B, P> operator*(B a, const Dual<B, P>& b) { return b * a; }
template <class B, int P> HGF_HD Dual<B, P> operator/(const Dual<B, P>& a, B b) {
  Dual<B, P> r;
  dual_div_primal(a.v, b, r.v, P > 0, [&](B inv) { r.d.scale(a.d, inv); });
  return r;
}
template <class B, int P> HGF_HD Dual<B, P> operator/(B a, const Dual<B, P>& b) {
  Dual<B, P> r;
  dual_div_primal(a, b.v, r.v, P > 0, [&](B inv) { r.d.scale(b.d, -r.v * inv); });
  return r;
}

// ----------------------------------------------------------------------------
// Elementary functions (float32: see the math policy at the top of this file)
// ----------------------------------------------------------------------------
HGF_D float b_exp(float x) {
#if HGF_EXP_MODE == 0
  return sfu_ex2(x * 1.4426950408889634f);
#else
  return ::expf(x);
#endif
}
HGF_D float b_log(float x) {
#if HGF_LOG_MODE == 0
  return sfu_lg2(x) * 0.6931471805599453f;
#else
  return ::logf(x);
#endif
}
HGF_D float b_log1p(float x) {
#if HGF_LOG_MODE == 0
  return b_log(1.0f + x);
#elif HGF_LOG_MODE == 1
  // log(u) + (x - (u - 1)) / u, u = fl(1 + x): repairs the rounding of 1 + x to first order.  u - 1 is exact for
  // x >= -1/2 (Sterbenz) -- the only caller, logaddexp, passes x in [0, 1].  An infinite or NaN x gives log(u).
  const float u = 1.0f + x;
  const float l = ::logf(u);
  const float c = (x - (u - 1.0f)) * sfu_rcp(u);
  return (c == c) ? l + c : l;
#else
  return ::log1pf(x);
#endif
}
HGF_D float b_pow(float x, float y) {
#if HGF_POW_MODE == 0
  return sfu_ex2(y * sfu_lg2(x));
#else
  return ::powf(x, y);
#endif
}
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
  r.d.scale(a.d, r.v);
  return r;
}
// exp(-x) with the negation folded into the tangent's scale factor (same values as x_exp(-x), one instruction less per
// tangent pair), and x * x with the product rule collapsed to 2 x dx
HGF_D float x_exp_neg(float x) { return b_exp(-x); }
HGF_D double x_exp_neg(double x) { return b_exp(-x); }
template <class B, int P> HGF_D Dual<B, P> x_exp_neg(const Dual<B, P>& a) {
  Dual<B, P> r; r.v = b_exp(-a.v);
  r.d.scale(a.d, -r.v);
  return r;
}
HGF_HD float x_square(float x) { return x * x; }
HGF_HD double x_square(double x) { return x * x; }
template <class B, int P> HGF_HD Dual<B, P> x_square(const Dual<B, P>& a) {
  Dual<B, P> r; r.v = a.v * a.v;
  r.d.scale(a.d, a.v + a.v);
  return r;
}
template <class B, int P> HGF_D Dual<B, P> x_log(const Dual<B, P>& a) {
  Dual<B, P> r; r.v = b_log(a.v);
  if constexpr (P > 0) {
    const B inv = t_rcp(a.v);
    r.d.scale(a.d, inv);
  }
  return r;
}
template <class B, int P> HGF_D Dual<B, P> x_sin(const Dual<B, P>& a) {
  Dual<B, P> r; r.v = b_sin(a.v); const B c = b_cos(a.v);
  r.d.scale(a.d, c);
  return r;
}
template <class B, int P> HGF_D Dual<B, P> x_cos(const Dual<B, P>& a) {
  Dual<B, P> r; r.v = b_cos(a.v); const B s = -b_sin(a.v);
  r.d.scale(a.d, s);
  return r;
}
template <class B, int P> HGF_D Dual<B, P> x_tanh(const Dual<B, P>& a) {
  Dual<B, P> r; r.v = b_tanh(a.v); const B c = B(1) - r.v * r.v;
  r.d.scale(a.d, c);
  return r;
}
template <class B, int P> HGF_D Dual<B, P> x_erf(const Dual<B, P>& a) {
  Dual<B, P> r; r.v = b_erf(a.v);
  const B c = B(1.1283791670955126) * b_exp(-a.v * a.v);
  r.d.scale(a.d, c);
  return r;
}
// jnp.power JVP given the primal r = x^y: y x^(y-1) dx + log(x) x^y dy
template <class B, int P> HGF_D Dual<B, P> pow_with_primal(B primal, const Dual<B, P>& a, const Dual<B, P>& b) {
  Dual<B, P> r; r.v = primal;
  r.d.zero();
  if constexpr (P > 0) {
#ifdef HGF_SFU_TANGENTS
    const B da = b.v * r.v * t_rcp(a.v);
#else
    const B da = b.v * b_pow(a.v, b.v - B(1));
#endif
    const B db = b_log(a.v) * r.v;
#pragma unroll
    for (int i = 0; i < P; ++i) {  // (element by element: the log term only where dy != 0)
      B t = da * a.d.get(i);
      if (b.d.get(i) != B(0)) t = b_fma(db, b.d.get(i), t);
      r.d.set(i, t);
    }
  }
  return r;
}
template <class B, int P> HGF_D Dual<B, P> x_pow(const Dual<B, P>& a, const Dual<B, P>& b) {
  return pow_with_primal(b_pow(a.v, b.v), a, b);
}
// libm / libdevice pow whatever the policy (<= 1 ulp in practice): see the inverse-temperature response
HGF_D float exact_pow(float x, float y) { return ::powf(x, y); }
HGF_D double exact_pow(double x, double y) { return ::pow(x, y); }

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
// The same against a bound held in the base type (a constant, never NaN): one comparison instead of three, same ties
// (the first argument wins) and the same NaN propagation from the variable argument.
template <class B, int P> HGF_D Dual<B, P> x_min(const Dual<B, P>& a, B b) { return (a.v > b) ? Dual<B, P>(b) : a; }
template <class B, int P> HGF_D Dual<B, P> x_max(const Dual<B, P>& a, B b) { return (a.v < b) ? Dual<B, P>(b) : a; }
template <class B, int P> HGF_D Dual<B, P> x_max(B a, const Dual<B, P>& b) { return (b.v <= a) ? Dual<B, P>(a) : b; }
template <class B, int P> HGF_D Dual<B, P> x_clip(const Dual<B, P>& x, B lo, B hi) { return x_min(x_max(x, lo), hi); }
template <class T> HGF_D bool x_isfinite(const T& a) { return b_isfinite(pv(a)); }
template <class T> HGF_D bool x_isnan(const T& a) { return b_isnan(pv(a)); }

// jax.nn.sigmoid == lax.logistic == 1 / (1 + exp(-x))
HGF_D float x_sigmoid(float x) { return b_rcp(1.0f + b_exp(-x)); }
HGF_D double x_sigmoid(double x) { return 1.0 / (1.0 + b_exp(-x)); }
template <class B, int P> HGF_D Dual<B, P> x_sigmoid(const Dual<B, P>& x) {
  // primal exactly as above; tangent in closed form, s (1 - s) dx, instead of differentiating exp and the division
  Dual<B, P> r;
  r.v = b_rcp(B(1) + b_exp(-x.v));
  if constexpr (P > 0) {
    const B c = r.v * (B(1) - r.v);
    r.d.scale(x.d, c);
  }
  return r;
}

// The same with the reciprocal at IEEE accuracy: for a sigmoid whose complement 1 - s is taken when s is next to 1
// (the binary node's prediction: 1 - mu_hat feeds its surprise), where one ulp of s is per cents of 1 - s.
template <class B, int P> HGF_D Dual<B, P> x_sigmoid_exact(const Dual<B, P>& x) {
  Dual<B, P> r;
  r.v = exact_rcp(B(1) + b_exp(-x.v));
  if constexpr (P > 0) {
    const B c = r.v * (B(1) - r.v);
    r.d.scale(x.d, c);
  }
  return r;
}

// jnp.logaddexp(a, b) = max + log1p(exp(-|a - b|)); NaN delta -> a + b
template <class B> HGF_D B b_logaddexp(B a, B b) {
  B amax = (a >= b || b_isnan(a)) ? a : b;
  if (b_isnan(b)) amax = b;
  const B delta = a - b;
  const B r = amax + b_log1p(b_exp(-b_abs(delta)));
  return b_isnan(delta) ? a + b : r;  // (a select, not an early return: no branch in the step's straight-line code)
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
    const B w_hi = t_rcp(B(1) + e), w_lo = B(1) - w_hi;
    const bool a_hi = a.v >= b.v;
    const B wa = a_hi ? w_hi : w_lo, wb = a_hi ? w_lo : w_hi;
    r.d.blend(a.d, wa, b.d, wb);
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
    const B inv = t_rcp(b_exp(r.v) * (B(1) + r.v));
    r.d.scale(z.d, inv);
  }
  return r;
}

}  // namespace hgf

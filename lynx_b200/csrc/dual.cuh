// Forward-mode dual numbers for the adjoint kernels.  The per-setting map algebra
// (maps.cuh, walk.cuh) is templated on its scalar type; instantiated with float it is the
// forward composer (K1 / K6), instantiated with Dual<float> the very same code yields
// d(map)/d(parameter) for one seeded parameter column -- no hand-derived per-element
// derivative code.  The reference has no custom_vjp: its gradients come
// from the array framework differentiating track_methods.py / cavity.py op by op, which
// is what the dual arithmetic reproduces.
#pragma once
#include <cuda_runtime.h>
#include <math.h>

namespace lynx {

template <typename F>
struct Dual {
  F v, d;
  Dual() = default;
  __host__ __device__ Dual(F v_) : v(v_), d(F(0)) {}
  __host__ __device__ Dual(F v_, F d_) : v(v_), d(d_) {}
};

// ---- underlying value / conversions --------------------------------------------------
__host__ __device__ __forceinline__ float val(float x) { return x; }
__host__ __device__ __forceinline__ double val(double x) { return x; }
template <typename F>
__host__ __device__ __forceinline__ F val(const Dual<F>& x) { return x.v; }

template <typename T> struct to_double_t { using type = double; };
template <typename F> struct to_double_t<Dual<F>> { using type = Dual<double>; };
template <typename T> struct to_float_t { using type = float; };
template <typename F> struct to_float_t<Dual<F>> { using type = Dual<float>; };

__host__ __device__ __forceinline__ double to_double(float x) { return (double)x; }
__host__ __device__ __forceinline__ double to_double(double x) { return x; }
template <typename F>
__host__ __device__ __forceinline__ Dual<double> to_double(const Dual<F>& x) {
  return Dual<double>((double)x.v, (double)x.d);
}
__host__ __device__ __forceinline__ float to_float(double x) { return (float)x; }
__host__ __device__ __forceinline__ float to_float(float x) { return x; }
template <typename F>
__host__ __device__ __forceinline__ Dual<float> to_float(const Dual<F>& x) {
  return Dual<float>((float)x.v, (float)x.d);
}

// "not identically zero": value or tangent non-zero.  Branches that skip work for a zero
// parameter (tilt, hx) must still be taken when that parameter carries the seeded tangent.
__host__ __device__ __forceinline__ bool nz(float x) { return x != 0.f; }
__host__ __device__ __forceinline__ bool nz(double x) { return x != 0.0; }
template <typename F>
__host__ __device__ __forceinline__ bool nz(const Dual<F>& x) { return x.v != F(0) || x.d != F(0); }

// ---- arithmetic ----------------------------------------------------------------------
#define LYNX_HD __host__ __device__ __forceinline__
// Scalar transcendentals are called once per setting and sit on rarely taken branches of
// the map builders (|k^2 L^2| >= 1, tilt != 0, dipoles, cavities): out of line, so their
// slow paths (Payne-Hanek, exp) do not inflate the composer kernels -- whose launch-to-first-
// result latency grows with code size when the L2 is cold.
#define LYNX_COLD static __device__ __noinline__
template <typename F> LYNX_HD Dual<F> operator+(const Dual<F>& a, const Dual<F>& b) { return {a.v + b.v, a.d + b.d}; }
template <typename F> LYNX_HD Dual<F> operator+(const Dual<F>& a, F b) { return {a.v + b, a.d}; }
template <typename F> LYNX_HD Dual<F> operator+(F a, const Dual<F>& b) { return {a + b.v, b.d}; }
template <typename F> LYNX_HD Dual<F> operator-(const Dual<F>& a, const Dual<F>& b) { return {a.v - b.v, a.d - b.d}; }
template <typename F> LYNX_HD Dual<F> operator-(const Dual<F>& a, F b) { return {a.v - b, a.d}; }
template <typename F> LYNX_HD Dual<F> operator-(F a, const Dual<F>& b) { return {a - b.v, -b.d}; }
template <typename F> LYNX_HD Dual<F> operator-(const Dual<F>& a) { return {-a.v, -a.d}; }
template <typename F> LYNX_HD Dual<F> operator*(const Dual<F>& a, const Dual<F>& b) { return {a.v * b.v, a.d * b.v + a.v * b.d}; }
template <typename F> LYNX_HD Dual<F> operator*(const Dual<F>& a, F b) { return {a.v * b, a.d * b}; }
template <typename F> LYNX_HD Dual<F> operator*(F a, const Dual<F>& b) { return {a * b.v, a * b.d}; }
template <typename F> LYNX_HD Dual<F> operator/(const Dual<F>& a, const Dual<F>& b) {
  const F q = a.v / b.v;
  return {q, (a.d - q * b.d) / b.v};
}
template <typename F> LYNX_HD Dual<F> operator/(const Dual<F>& a, F b) { return {a.v / b, a.d / b}; }
template <typename F> LYNX_HD Dual<F> operator/(F a, const Dual<F>& b) {
  const F q = a / b.v;
  return {q, -q * b.d / b.v};
}
template <typename F> LYNX_HD Dual<F>& operator+=(Dual<F>& a, const Dual<F>& b) { a.v += b.v; a.d += b.d; return a; }
template <typename F> LYNX_HD Dual<F>& operator-=(Dual<F>& a, const Dual<F>& b) { a.v -= b.v; a.d -= b.d; return a; }
template <typename F> LYNX_HD Dual<F>& operator+=(Dual<F>& a, F b) { a.v += b; return a; }
template <typename F> LYNX_HD Dual<F>& operator-=(Dual<F>& a, F b) { a.v -= b; return a; }
template <typename F> LYNX_HD Dual<F>& operator*=(Dual<F>& a, const Dual<F>& b) { a = a * b; return a; }
template <typename F> LYNX_HD Dual<F>& operator*=(Dual<F>& a, F b) { a.v *= b; a.d *= b; return a; }

// comparisons act on the value (branch selection follows the primal, as autograd does)
template <typename F> LYNX_HD bool operator==(const Dual<F>& a, F b) { return a.v == b; }
template <typename F> LYNX_HD bool operator!=(const Dual<F>& a, F b) { return a.v != b; }
template <typename F> LYNX_HD bool operator<(const Dual<F>& a, F b) { return a.v < b; }
template <typename F> LYNX_HD bool operator>(const Dual<F>& a, F b) { return a.v > b; }

// ---- math: one overload set for float, double and Dual -------------------------------
LYNX_HD float fma_(float a, float b, float c) { return fmaf(a, b, c); }
LYNX_HD double fma_(double a, double b, double c) { return fma(a, b, c); }
template <typename F> LYNX_HD Dual<F> fma_(const Dual<F>& a, const Dual<F>& b, const Dual<F>& c) {
  return {fma_(a.v, b.v, c.v), fma_(a.d, b.v, fma_(a.v, b.d, c.d))};
}
template <typename F> LYNX_HD Dual<F> fma_(F a, const Dual<F>& b, const Dual<F>& c) {
  return {fma_(a, b.v, c.v), fma_(a, b.d, c.d)};
}
template <typename F> LYNX_HD Dual<F> fma_(const Dual<F>& a, F b, const Dual<F>& c) {
  return {fma_(a.v, b, c.v), fma_(a.d, b, c.d)};
}
template <typename F> LYNX_HD Dual<F> fma_(const Dual<F>& a, const Dual<F>& b, F c) {
  return {fma_(a.v, b.v, c), fma_(a.d, b.v, a.v * b.d)};
}
template <typename F> LYNX_HD Dual<F> fma_(const Dual<F>& a, F b, F c) { return {fma_(a.v, b, c), a.d * b}; }
template <typename F> LYNX_HD Dual<F> fma_(F a, const Dual<F>& b, F c) { return {fma_(a, b.v, c), a * b.d}; }

LYNX_HD float sqrt_(float x) { return sqrtf(x); }
LYNX_HD double sqrt_(double x) { return sqrt(x); }
template <typename F> LYNX_HD Dual<F> sqrt_(const Dual<F>& x) {
  const F r = sqrt_(x.v);
  return {r, x.d / (F(2) * r)};
}
__device__ __forceinline__ float rsqrt_(float x) { return rsqrtf(x); }
__device__ __forceinline__ double rsqrt_(double x) { return rsqrt(x); }
template <typename F> __device__ __forceinline__ Dual<F> rsqrt_(const Dual<F>& x) {
  const F r = rsqrt_(x.v);
  return {r, F(-0.5) * r / x.v * x.d};
}
LYNX_COLD void sincos_(float x, float* s, float* c) { sincosf(x, s, c); }
LYNX_COLD void sincos_(double x, double* s, double* c) { sincos(x, s, c); }
template <typename F> __device__ __forceinline__ void sincos_(const Dual<F>& x, Dual<F>* s, Dual<F>* c) {
  F sv, cv;
  sincos_(x.v, &sv, &cv);
  *s = {sv, cv * x.d};
  *c = {cv, -sv * x.d};
}
LYNX_COLD float sin_(float x) { return sinf(x); }
LYNX_COLD double sin_(double x) { return sin(x); }
LYNX_COLD float cos_(float x) { return cosf(x); }
LYNX_COLD double cos_(double x) { return cos(x); }
template <typename F> __device__ __forceinline__ Dual<F> sin_(const Dual<F>& x) {
  F sv, cv;
  sincos_(x.v, &sv, &cv);
  return {sv, cv * x.d};
}
template <typename F> __device__ __forceinline__ Dual<F> cos_(const Dual<F>& x) {
  F sv, cv;
  sincos_(x.v, &sv, &cv);
  return {cv, -sv * x.d};
}
LYNX_COLD float sinh_(float x) { return sinhf(x); }
LYNX_COLD double sinh_(double x) { return sinh(x); }
LYNX_COLD float cosh_(float x) { return coshf(x); }
LYNX_COLD double cosh_(double x) { return cosh(x); }
template <typename F> LYNX_HD Dual<F> sinh_(const Dual<F>& x) { return {sinh_(x.v), cosh_(x.v) * x.d}; }
template <typename F> LYNX_HD Dual<F> cosh_(const Dual<F>& x) { return {cosh_(x.v), sinh_(x.v) * x.d}; }
LYNX_COLD float tan_(float x) { return tanf(x); }
LYNX_COLD double tan_(double x) { return tan(x); }
template <typename F> LYNX_HD Dual<F> tan_(const Dual<F>& x) {
  const F t = tan_(x.v);
  return {t, (F(1) + t * t) * x.d};
}
LYNX_COLD float log1p_(float x) { return log1pf(x); }
LYNX_COLD double log1p_(double x) { return log1p(x); }
template <typename F> LYNX_HD Dual<F> log1p_(const Dual<F>& x) { return {log1p_(x.v), x.d / (F(1) + x.v)}; }

}  // namespace lynx

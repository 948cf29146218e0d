// Per-setting transfer-map algebra, device side.  One thread owns one setting's 6x7
// affine map in registers and LEFT-multiplies it by each element's (sparse) map:
// no 7x7 temporaries, no scatter writes, no complex arithmetic.
//
// Arithmetic follows:
//   lynx/track_methods.py:14-122, drift.py:44-62, quadrupole.py:66-80,
//   horizontal_corrector.py:52-67, vertical_corrector.py:52-66, dipole.py:96-181,
//   rbend.py:79-80, cavity.py:248-325, solenoid.py:61-105, undulator.py:48-60,
//   custom_transfer_map.py:87-88
// with the complex sqrt/cos/sin of track_methods.py:72-79 replaced by the real
// cos/sin | cosh/sinh pair selected by the sign of k^2, and the subtractive forms
// (1-cx)/kx2, (L-sx)/kx2 replaced by cancellation-free series / half-angle forms, so
// that fp32 results sit within ~1e-6 of the fp64 value of the reference's formula.
//
// Everything is templated on the scalar type T: float for the forward composer (K1/K6),
// Dual<float> (dual.cuh) for the parameter tangents of the adjoint kernels (K3b, K6-bwd).
#pragma once
#include "common.cuh"
#include "dual.cuh"

namespace lynx {

// LT ("light"): the map of a run made only of kOpsLight elements (walk.cuh: drifts, quadrupoles,
// correctors, undulators, markers).  No such element reads rows 4 / 5 into rows 0..3 (no
// dispersion) or writes row 5, so columns 4 and 5 of rows 0..3 stay zero, row 4 stays
// e4 + r56 e5 and row 5 stays e5.  The row primitives skip those entries and every consumer
// skips the products with them: the results are the bits the dense arithmetic produces (an fma
// with a structural zero is exact), at 21 live entries instead of 42.
template <typename T, bool LT = false>
struct MapT {
  T m[6][7];
};
using Map = MapT<float>;

// is entry (i, j) of a map carried as a variable (true) or a structural 0 / 1 (false)
template <bool LT>
__host__ __device__ constexpr bool map_live(int i, int j) {
  return !LT || (i < 4 && j != 4 && j != 5) || (i == 4 && j == 5);
}

template <typename T, bool LT>
__device__ __forceinline__ void map_identity(MapT<T, LT>& M) {
#pragma unroll
  for (int i = 0; i < 6; ++i)
#pragma unroll
    for (int j = 0; j < 7; ++j) M.m[i][j] = T((i == j) ? 1.f : 0.f);
}

// row_a += f * row_b
template <typename T, bool LT>
__device__ __forceinline__ void row_axpy(MapT<T, LT>& M, int a, T f, int b) {
#pragma unroll
  for (int j = 0; j < 7; ++j)
    if (map_live<LT>(a, j)) M.m[a][j] = fma_(f, M.m[b][j], M.m[a][j]);
}

// [row_a; row_b] <- [[p, q], [r, s]] [row_a; row_b]
template <typename T, bool LT>
__device__ __forceinline__ void rows_2x2(MapT<T, LT>& M, int a, int b, T p, T q, T r, T s) {
#pragma unroll
  for (int j = 0; j < 7; ++j) {
    if (!map_live<LT>(a, j)) continue;
    T x = M.m[a][j], y = M.m[b][j];
    M.m[a][j] = fma_(p, x, q * y);
    M.m[b][j] = fma_(r, x, s * y);
  }
}

// Relativistic factors of the reference energy, computed once per energy value.
template <typename T>
struct RelT {
  T igamma2_drift;  // drift.py:53-54: 0 where gamma == 0
  T igamma2_base;   // track_methods.py:61-62: 1 where gamma == 0 (quirk B9)
  T inv_beta2_drift, inv_beta_base, inv_beta2_base;
};
using Rel = RelT<float>;

template <typename T>
__device__ __forceinline__ RelT<T> make_rel(T energy) {
  RelT<T> r;
  T gamma = energy / (float)kRestEnergyEV;
  T ig2 = (val(gamma) != 0.f) ? 1.f / (gamma * gamma) : T(0.f);
  r.igamma2_drift = ig2;
  r.igamma2_base = (val(gamma) != 0.f) ? ig2 : T(1.f);
  r.inv_beta2_drift = 1.f / (1.f - r.igamma2_drift);
  T b2 = 1.f - r.igamma2_base;
  r.inv_beta2_base = 1.f / b2;
  r.inv_beta_base = rsqrt_(b2);
  return r;
}

// c = cos(sqrt(t)) / cosh(sqrt(-t)),  s = sin(sqrt(t))/sqrt(t) / sinh(..)/..  (s for L=1)
// |t| < 1: Taylor series in t (entire functions of t; also keeps the dual tangent free of
// the (u cos u - sin u)/u^3 cancellation of the closed forms).
template <typename T>
__device__ __forceinline__ void cos_sinc(T t, T& c, T& s) {
  const float tv = val(t);
  if (fabsf(tv) < 1.f) {
    c = fma_(t, fma_(t, fma_(t, fma_(t, fma_(t, fma_(t, T(1.f / 479001600.f), -1.f / 3628800.f),
                                              1.f / 40320.f), -1.f / 720.f), 1.f / 24.f), -0.5f), 1.f);
    s = fma_(t, fma_(t, fma_(t, fma_(t, fma_(t, fma_(t, T(1.f / 6227020800.f), -1.f / 39916800.f),
                                              1.f / 362880.f), -1.f / 5040.f), 1.f / 120.f),
                     -1.f / 6.f), 1.f);
  } else if (tv > 0.f) {
    T u = sqrt_(t), sn, cs;
    sincos_(u, &sn, &cs);
    c = cs;
    s = sn / u;
  } else {
    T u = sqrt_(-t);
    c = cosh_(u);
    s = sinh_(u) / u;
  }
}

// g(t) = (1 - cos sqrt t)/t ;  h(t) = (1 - sinc sqrt t)/t   (analytic through t <= 0)
template <typename T>
__device__ __forceinline__ void disp_fns(T t, T& g, T& h) {
  const float tv = val(t);
  if (fabsf(tv) < 1.f) {
    g = fma_(t, fma_(t, fma_(t, fma_(t, fma_(t, fma_(t, T(1.f / 87178291200.f), -1.f / 479001600.f),
                                              1.f / 3628800.f), -1.f / 40320.f), 1.f / 720.f),
                     -1.f / 24.f), 0.5f);
    h = fma_(t, fma_(t, fma_(t, fma_(t, fma_(t, fma_(t, T(1.f / 1307674368000.f), -1.f / 6227020800.f),
                                              1.f / 39916800.f), -1.f / 362880.f), 1.f / 5040.f),
                     -1.f / 120.f), 1.f / 6.f);
  } else if (tv > 0.f) {
    T u = sqrt_(t);
    T sh = sin_(0.5f * u);
    g = 2.f * sh * sh / t;
    h = (1.f - sin_(u) / u) / t;
  } else {
    T u = sqrt_(-t);
    T sh = sinh_(0.5f * u);
    g = 2.f * sh * sh / (-t);
    h = (sinh_(u) / u - 1.f) / (-t);
  }
}

// rotation_matrix(a) applied from the left, track_methods.py:14-34
template <typename T, bool LT>
__device__ __forceinline__ void lm_rotation(MapT<T, LT>& M, T a) {
  T sn, cs;
  sincos_(a, &sn, &cs);
  rows_2x2(M, 0, 2, cs, sn, -sn, cs);
  rows_2x2(M, 1, 3, cs, sn, -sn, cs);
}

// drift.py:57-60 (also correctors, undulator with its own r56)
template <typename T, bool LT>
__device__ __forceinline__ void lm_drift(MapT<T, LT>& M, T L, T r56) {
  row_axpy(M, 0, L, 1);
  row_axpy(M, 2, L, 3);
  row_axpy(M, 4, r56, 5);
}

template <typename T>
__device__ __forceinline__ T drift_r56(T L, const RelT<T>& rel) {
  return -L * rel.inv_beta2_drift * rel.igamma2_drift;
}

// base_rmatrix(length, k1, hx, tilt=0, energy), track_methods.py:37-98 (tilt by caller)
template <typename T, bool LT>
__device__ __forceinline__ void lm_base(MapT<T, LT>& M, T L, T k1, T hx, const RelT<T>& rel) {
  // :67-68 (k1 == 0 -> 1e-12).  The reference does this with a masked write, which also
  // zeroes d/dk1 at exactly k1 == 0; the tangent is kept here.
  if (val(k1) == 0.f) k1 = k1 + 1e-12f;
  T kx2 = fma_(hx, hx, k1), ky2 = -k1;
  T L2 = L * L;
  T cx, sx, cy, sy;
  cos_sinc(kx2 * L2, cx, sx);
  cos_sinc(ky2 * L2, cy, sy);
  sx = sx * L;
  sy = sy * L;
  T r56 = -L * rel.inv_beta2_base * rel.igamma2_base;
  if (!LT && nz(hx)) {  // dispersion: dipoles only, never part of a light run
    T g, h;
    disp_fns(kx2 * L2, g, h);
    T dxb = hx * L2 * g * rel.inv_beta_base;   // dx / beta
    T sxb = sx * hx * rel.inv_beta_base;       // sx hx / beta
    r56 = fma_(hx * hx * L2 * L * h, rel.inv_beta2_base, r56);
    // rows 0,1,4 read old rows 0,1,5 -> compute row 4 first
#pragma unroll
    for (int j = 0; j < 7; ++j) {
      T r0 = M.m[0][j], r1 = M.m[1][j], r5 = M.m[5][j];
      M.m[4][j] = fma_(sxb, r0, fma_(dxb, r1, fma_(r56, r5, M.m[4][j])));
      M.m[0][j] = fma_(cx, r0, fma_(sx, r1, dxb * r5));
      M.m[1][j] = fma_(-kx2 * sx, r0, fma_(cx, r1, sxb * r5));
    }
  } else {
    rows_2x2(M, 0, 1, cx, sx, -kx2 * sx, cx);
    row_axpy(M, 4, r56, 5);
  }
  rows_2x2(M, 2, 3, cy, sy, -ky2 * sy, cy);
}

// quadrupole.py:66-80
template <typename T, bool LT>
__device__ __forceinline__ void lm_quad(MapT<T, LT>& M, T L, T k1, T mx, T my, T tilt,
                                         const RelT<T>& rel) {
  // R_exit . [rot(-tilt) . base . rot(tilt)] . R_entry   (right-most acts first)
  M.m[0][6] -= mx;  // R_entry, track_methods.py:118-120
  M.m[2][6] -= my;
  if (nz(tilt)) lm_rotation(M, tilt);
  lm_base(M, L, k1, T(0.f), rel);
  if (nz(tilt)) lm_rotation(M, -tilt);
  M.m[0][6] += mx;  // R_exit, :114-116
  M.m[2][6] += my;
}

// dipole.py:143-181 edge map
template <typename T>
__device__ __forceinline__ void lm_dipole_edge(MapT<T>& M, T hx, T e, T fint, T gap) {
  T se, ce;
  sincos_(e, &se, &ce);
  T phi = fint * hx * gap / ce * (1.f + se * se);
  row_axpy(M, 1, hx * tan_(e), 0);
  row_axpy(M, 3, -hx * tan_(e - phi), 2);
}

// dipole.py:112-141 (+ rbend.py:79-80 via `rbend`)
template <typename T>
__device__ __forceinline__ void lm_dipole(MapT<T>& M, T L, T angle, T e1, T e2, T tilt, T fint,
                                           T fint_exit, T gap, bool rbend, const RelT<T>& rel) {
  if (rbend) {
    e1 = e1 + 0.5f * angle;
    e2 = e2 + 0.5f * angle;
  }
  T hx = (val(L) != 0.f) ? angle / L : T(0.f);  // :96-102
  if (nz(tilt)) lm_rotation(M, tilt);
  lm_dipole_edge(M, hx, e1, fint, gap);
  if (val(L) != 0.f) {
    lm_base(M, L, T(0.f), hx, rel);
  } else {
    M.m[2][6] += angle;  // :131-133 as written: the kick lands in R[2,6] (quirk B3)
  }
  lm_dipole_edge(M, hx, e2, fint_exit, gap);
  if (nz(tilt)) lm_rotation(M, -tilt);
}

// solenoid.py:61-105
template <typename T>
__device__ __forceinline__ void lm_solenoid(MapT<T>& M, T L, T k, T mx, T my,
                                             const RelT<T>& rel) {
  M.m[0][6] -= mx;
  M.m[2][6] -= my;
  T s, c;
  sincos_(L * k, &s, &c);
  T t = L * k;
  T c1, sk;
  cos_sinc(t * t, c1, sk);  // sk = sin(Lk)/(Lk), exact limit at k = 0
  sk = sk * L;
  T c2 = c * c, sc = s * c, s2 = s * s;
#pragma unroll
  for (int j = 0; j < 7; ++j) {
    T r0 = M.m[0][j], r1 = M.m[1][j], r2 = M.m[2][j], r3 = M.m[3][j];
    M.m[0][j] = c2 * r0 + c * sk * r1 + sc * r2 + s * sk * r3;
    M.m[1][j] = -k * sc * r0 + c2 * r1 - k * s2 * r2 + sc * r3;
    M.m[2][j] = -sc * r0 - s * sk * r1 + c2 * r2 + c * sk * r3;
    M.m[3][j] = k * s2 * r0 - sc * r1 - k * sc * r2 + c2 * r3;
  }
  // :74-78 r56 = -L/(beta^2 gamma^2) where gamma != 0
  row_axpy(M, 4, -L * rel.inv_beta2_drift * rel.igamma2_drift, 5);
  M.m[0][6] += mx;
  M.m[2][6] += my;
}

// custom_transfer_map.py:87-88: dense 6x7 (rows 0..5 of the stored 7x7)
template <typename T, typename F>
__device__ __forceinline__ void lm_custom(MapT<T>& M, F&& coef /* (row, col) -> T */) {
  MapT<T> R;
#pragma unroll
  for (int i = 0; i < 6; ++i) {
    T c[7];
#pragma unroll
    for (int k = 0; k < 7; ++k) c[k] = coef(i, k);
#pragma unroll
    for (int j = 0; j < 7; ++j) {
      T acc = (j == 6) ? c[6] : T(0.f);
#pragma unroll
      for (int k = 0; k < 6; ++k) acc = fma_(c[k], M.m[k][j], acc);
      R.m[i][j] = acc;
    }
  }
  M = R;
}

// Standing-wave cavity, cavity.py:248-325 (linear map) and :101-200 (non-linear
// scalars).  Evaluated in double: the formulas subtract nearly equal quantities
// ((b0 b1 - 1) g0 g1 + 1, (b0 g0)^3 - (b1 g1)^3) and run once per setting, not per
// particle.  V == 0: exact Drift map (deviation from quirk B6, the 0*inf NaN).
template <typename T>
struct CavityOutT {
  T sc[kCavFloats];
  T dE;
};
using CavityOut = CavityOutT<float>;

template <typename T>
__device__ inline void cavity_setting(T Lf, T Vf, T phase_deg, T freq, T energy, MapT<T>& C,
                                       CavityOutT<T>& out) {
  using D = typename to_double_t<T>::type;
  map_identity(C);
  const D L = to_double(Lf), V = to_double(Vf), E = to_double(energy);
  const double me = kElectronMassEV;
  const D phi = to_double(phase_deg) * (kPi / 180.0);
  D sphi, cphi;
  sincos_(phi, &sphi, &cphi);
  const D dE = V * cphi;
  const D k = (2.0 * kPi / kSpeedOfLight) * to_double(freq);
#pragma unroll
  for (int i = 0; i < kCavFloats; ++i) out.sc[i] = T(0.f);
  out.dE = to_float(dE);
  out.sc[CV_DE] = to_float(dE);

  // cavity.py:101-112
  const bool e_nz = val(E) != 0.0;
  D g0 = e_nz ? E / me : D(1e10);
  D igamma2 = e_nz ? 1.0 / (g0 * g0) : D(0.0);
  D beta0 = e_nz ? sqrt_(1.0 - igamma2) : D(1.0);

  if (val(V) == 0.0) {
    // drift.py:57-60 with REST_ENERGY-based gamma (what a skippable cavity should be)
    D g = E / kRestEnergyEV;
    D ig2 = (val(g) != 0.0) ? 1.0 / (g * g) : D(0.0);
    C.m[0][1] = Lf;
    C.m[2][3] = Lf;
    C.m[4][5] = to_float(-L / (1.0 - ig2) * ig2);
  } else {
    const D Ei = E / me, Ef = (E + dE) / me;
    const D dG = dE / me;           // Ef - Ei without the subtraction
    const D Ep = dG / L;
    const D alpha = sqrt(0.125) / cphi * log1p_(dG / Ei);
    D sa, ca;
    sincos_(alpha, &sa, &ca);
    const D r11 = ca - sqrt(2.0) * cphi * sa;
    // r12 = sqrt(8) Ei/Ep cphi sin(alpha); finite limit L as dG -> 0 (phi = +-90 deg)
    const D r12 = (val(dG) != 0.0) ? sqrt(8.0) * Ei / Ep * cphi * sa : L;
    const D r21 = -Ep / Ef * (cphi / sqrt(2.0) + sqrt(0.125) / cphi) * sa;
    const D r22 = Ei / Ef * (ca + sqrt(2.0) * cphi * sa);
    D r56 = D(0.0), b0 = D(1.0), b1 = D(1.0), r55c = D(0.0);
    if (e_nz) {  // :290 branch, per setting
      b0 = sqrt_(1.0 - 1.0 / (Ei * Ei));
      b1 = sqrt_(1.0 - 1.0 / (Ef * Ef));
      r56 = -L / (Ef * Ef * Ei * b1) * (Ef + Ei) / (b1 + b0);
      r55c = k * L * b0 * V / me * sphi * (Ei * Ef * (b0 * b1 - 1.0) + 1.0) /
             (b1 * Ef * dG * dG);
    }
    const D r66 = Ei / Ef * b0 / b1;
    const D r65 = k * sphi * V / (Ef * b1 * me);
    C.m[0][0] = C.m[2][2] = to_float(r11);
    C.m[0][1] = C.m[2][3] = to_float(r12);
    C.m[1][0] = C.m[3][2] = to_float(r21);
    C.m[1][1] = C.m[3][3] = to_float(r22);
    C.m[4][4] = to_float(1.0 + r55c);
    C.m[4][5] = to_float(r56);
    C.m[5][4] = to_float(r65);
    C.m[5][5] = to_float(r66);
  }

  // non-linear scalars, cavity.py:121-200
  D T566 = 1.5 * L * igamma2 / (beta0 * beta0 * beta0), T556 = D(0.0), T555 = D(0.0);
  const D E1 = E + dE;
  const bool on = val(E1) > 0.0;  // :128, per setting
  if (on) {
    const D g1 = E1 / me;
    const D beta1 = sqrt_(1.0 - 1.0 / (g1 * g1));
    out.sc[CV_AP] = to_float(E * beta0 / (E1 * beta1));
    out.sc[CV_BP] = to_float(V * beta0 / (E1 * beta1));
    out.sc[CV_KB] = to_float(beta0 * k);
    out.sc[CV_PHI] = to_float(phi);
    out.sc[CV_SINPHI] = to_float(sphi);
    out.sc[CV_COSPHI] = to_float(cphi);
    if (val(dE) > 0.0) {  // :164, per setting
      const D dgamma = V / me;
      const D b03 = beta0 * beta0 * beta0, b13 = beta1 * beta1 * beta1;
      const D g03 = g0 * g0 * g0, g13 = g1 * g1 * g1;
      const D d01 = g0 - g1;
      T566 = L * (b03 * g03 - b13 * g13) / (2.0 * beta0 * b13 * g0 * d01 * g13);
      T556 = beta0 * k * L * dgamma * g0 * (b13 * g13 + beta0 * (g0 - g13)) * sphi /
             (b13 * g13 * d01 * d01);
      T555 = beta0 * beta0 * k * k * L * dgamma / 2.0 *
             (dgamma * (2.0 * g0 * g13 * (beta0 * b13 - 1.0) + g0 * g0 + 3.0 * g1 * g1 - 2.0) /
                  (b13 * g13 * d01 * d01 * d01) * sphi * sphi -
              (g1 * g0 * (beta1 * beta0 - 1.0) + 1.0) / (beta1 * g1 * d01 * d01) * cphi);
    }
  }
  out.sc[CV_T566] = to_float(T566);
  out.sc[CV_T556] = to_float(T556);
  out.sc[CV_T555] = to_float(T555);
  out.sc[CV_ON] = T(on ? 1.f : 0.f);
}

// Non-linear part of Cavity._track_beam for one coordinate vector, cavity.py:142-161,
// :220-226.  (s0, p0) incoming, lin4 = (R v0)[4].  cos(phi - d) - cos(phi) is taken as
// 2 sin(phi - d/2) sin(d/2): the subtraction form loses every digit in fp32 for
// d = s0 beta0 k ~ 1e-4.
// sin / cos of the half phase slip h = s0 beta0 k / 2.  For a bunch (|h| ~ 1e-5) short series
// are exact to fp32 and cost a tenth of the library sincosf; the branch is warp-uniform in
// practice.  Generic scalar types (double, dual numbers) take the library route.
template <typename T>
__device__ __forceinline__ void sincos_half_slip(T h, T& s, T& c) {
  sincos_(h, &s, &c);
}
__device__ __forceinline__ void sincos_half_slip(float h, float& s, float& c) {
  if (fabsf(h) < 0.03125f) {  // truncation: h^8/9! and h^8/8! relative, < 1e-15
    const float h2 = h * h;
    s = h * fmaf(h2, fmaf(h2, fmaf(h2, -1.9841270e-4f, 8.3333333e-3f), -0.16666667f), 1.0f);
    c = fmaf(h2, fmaf(h2, fmaf(h2, -1.3888889e-3f, 4.1666667e-2f), -0.5f), 1.0f);
  } else {
    sincosf(h, &s, &c);
  }
}

template <typename T>
__device__ __forceinline__ void cavity_nonlinear(const T* sc, T s0, T p0, T lin4, T lin5, T& s1,
                                                  T& p1) {
  if (val(sc[CV_ON]) != 0.f) {
    // cos(phi - d) - cos(phi) = 2 sin(phi - h) sin(h), h = d / 2, and
    // sin(phi - h) = sin(phi) cos(h) - cos(phi) sin(h) with sin / cos(phi) from K1 (fp64)
    T sh, ch;
    sincos_half_slip(T(0.5f) * s0 * sc[CV_KB], sh, ch);
    T dc = 2.f * (sc[CV_SINPHI] * ch - sc[CV_COSPHI] * sh) * sh;
    p1 = fma_(p0, sc[CV_AP], sc[CV_BP] * dc);
    s1 = lin4 + fma_(sc[CV_T566] * p0, p0, fma_(sc[CV_T556] * s0, p0, sc[CV_T555] * s0 * s0));
  } else {
    p1 = lin5;
    s1 = lin4;
  }
}

}  // namespace lynx

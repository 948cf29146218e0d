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
#pragma once
#include "common.cuh"

namespace lynx {

struct Map {
  float m[6][7];
};

__device__ __forceinline__ void map_identity(Map& M) {
#pragma unroll
  for (int i = 0; i < 6; ++i)
#pragma unroll
    for (int j = 0; j < 7; ++j) M.m[i][j] = (i == j) ? 1.f : 0.f;
}

// row_a += f * row_b
__device__ __forceinline__ void row_axpy(Map& M, int a, float f, int b) {
#pragma unroll
  for (int j = 0; j < 7; ++j) M.m[a][j] = fmaf(f, M.m[b][j], M.m[a][j]);
}

// [row_a; row_b] <- [[p, q], [r, s]] [row_a; row_b]
__device__ __forceinline__ void rows_2x2(Map& M, int a, int b, float p, float q, float r,
                                          float s) {
#pragma unroll
  for (int j = 0; j < 7; ++j) {
    float x = M.m[a][j], y = M.m[b][j];
    M.m[a][j] = fmaf(p, x, q * y);
    M.m[b][j] = fmaf(r, x, s * y);
  }
}

// Relativistic factors of the reference energy, computed once per energy value.
struct Rel {
  float igamma2_drift;  // drift.py:53-54: 0 where gamma == 0
  float igamma2_base;   // track_methods.py:61-62: 1 where gamma == 0 (quirk B9)
  float inv_beta2_drift, inv_beta_base, inv_beta2_base;
};

__device__ __forceinline__ Rel make_rel(float energy) {
  Rel r;
  float gamma = energy / (float)kRestEnergyEV;
  float ig2 = (gamma != 0.f) ? 1.f / (gamma * gamma) : 0.f;
  r.igamma2_drift = ig2;
  r.igamma2_base = (gamma != 0.f) ? ig2 : 1.f;
  r.inv_beta2_drift = 1.f / (1.f - r.igamma2_drift);
  float b2 = 1.f - r.igamma2_base;
  r.inv_beta2_base = 1.f / b2;
  r.inv_beta_base = rsqrtf(b2) ;
  return r;
}

// c = cos(sqrt(t)) / cosh(sqrt(-t)),  s = sin(sqrt(t))/sqrt(t) / sinh(..)/..  (s for L=1)
__device__ __forceinline__ void cos_sinc(float t, float& c, float& s) {
  if (fabsf(t) < 0.0625f) {
    c = fmaf(t, fmaf(t, fmaf(t, fmaf(t, 1.f / 40320.f, -1.f / 720.f), 1.f / 24.f), -0.5f), 1.f);
    s = fmaf(t, fmaf(t, fmaf(t, fmaf(t, 1.f / 362880.f, -1.f / 5040.f), 1.f / 120.f),
                     -1.f / 6.f), 1.f);
  } else if (t > 0.f) {
    float u = sqrtf(t), sn, cs;
    sincosf(u, &sn, &cs);
    c = cs;
    s = sn / u;
  } else {
    float u = sqrtf(-t);
    c = coshf(u);
    s = sinhf(u) / u;
  }
}

// g(t) = (1 - cos sqrt t)/t ;  h(t) = (1 - sinc sqrt t)/t   (analytic through t <= 0)
__device__ __forceinline__ void disp_fns(float t, float& g, float& h) {
  if (fabsf(t) < 1.f) {
    g = fmaf(t, fmaf(t, fmaf(t, fmaf(t, fmaf(t, -1.f / 479001600.f, 1.f / 3628800.f),
                                     -1.f / 40320.f), 1.f / 720.f), -1.f / 24.f), 0.5f);
    h = fmaf(t, fmaf(t, fmaf(t, fmaf(t, fmaf(t, -1.f / 6227020800.f, 1.f / 39916800.f),
                                     -1.f / 362880.f), 1.f / 5040.f), -1.f / 120.f),
             1.f / 6.f);
  } else if (t > 0.f) {
    float u = sqrtf(t);
    float sh = sinf(0.5f * u);
    g = 2.f * sh * sh / t;
    h = (1.f - sinf(u) / u) / t;
  } else {
    float u = sqrtf(-t);
    float sh = sinhf(0.5f * u);
    g = 2.f * sh * sh / (-t);
    h = (sinhf(u) / u - 1.f) / (-t);
  }
}

// rotation_matrix(a) applied from the left, track_methods.py:14-34
__device__ __forceinline__ void lm_rotation(Map& M, float a) {
  float sn, cs;
  sincosf(a, &sn, &cs);
  rows_2x2(M, 0, 2, cs, sn, -sn, cs);
  rows_2x2(M, 1, 3, cs, sn, -sn, cs);
}

// drift.py:57-60 (also correctors, undulator with its own r56)
__device__ __forceinline__ void lm_drift(Map& M, float L, float r56) {
  row_axpy(M, 0, L, 1);
  row_axpy(M, 2, L, 3);
  row_axpy(M, 4, r56, 5);
}

__device__ __forceinline__ float drift_r56(float L, const Rel& rel) {
  return -L * rel.inv_beta2_drift * rel.igamma2_drift;
}

// base_rmatrix(length, k1, hx, tilt=0, energy), track_methods.py:37-98 (tilt by caller)
__device__ __forceinline__ void lm_base(Map& M, float L, float k1, float hx, const Rel& rel) {
  if (k1 == 0.f) k1 = 1e-12f;  // :67-68
  float kx2 = fmaf(hx, hx, k1), ky2 = -k1;
  float L2 = L * L;
  float cx, sx, cy, sy;
  cos_sinc(kx2 * L2, cx, sx);
  cos_sinc(ky2 * L2, cy, sy);
  sx *= L;
  sy *= L;
  float r56 = -L * rel.inv_beta2_base * rel.igamma2_base;
  if (hx != 0.f) {
    float g, h;
    disp_fns(kx2 * L2, g, h);
    float dxb = hx * L2 * g * rel.inv_beta_base;   // dx / beta
    float sxb = sx * hx * rel.inv_beta_base;       // sx hx / beta
    r56 = fmaf(hx * hx * L2 * L * h, rel.inv_beta2_base, r56);
    // rows 0,1,4 read old rows 0,1,5 -> compute row 4 first
#pragma unroll
    for (int j = 0; j < 7; ++j) {
      float r0 = M.m[0][j], r1 = M.m[1][j], r5 = M.m[5][j];
      M.m[4][j] = fmaf(sxb, r0, fmaf(dxb, r1, fmaf(r56, r5, M.m[4][j])));
      M.m[0][j] = fmaf(cx, r0, fmaf(sx, r1, dxb * r5));
      M.m[1][j] = fmaf(-kx2 * sx, r0, fmaf(cx, r1, sxb * r5));
    }
  } else {
    rows_2x2(M, 0, 1, cx, sx, -kx2 * sx, cx);
    row_axpy(M, 4, r56, 5);
  }
  rows_2x2(M, 2, 3, cy, sy, -ky2 * sy, cy);
}

// quadrupole.py:66-80
__device__ __forceinline__ void lm_quad(Map& M, float L, float k1, float mx, float my,
                                         float tilt, const Rel& rel) {
  // R_exit . [rot(-tilt) . base . rot(tilt)] . R_entry   (right-most acts first)
  M.m[0][6] -= mx;  // R_entry, track_methods.py:118-120
  M.m[2][6] -= my;
  if (tilt != 0.f) lm_rotation(M, tilt);
  lm_base(M, L, k1, 0.f, rel);
  if (tilt != 0.f) lm_rotation(M, -tilt);
  M.m[0][6] += mx;  // R_exit, :114-116
  M.m[2][6] += my;
}

// dipole.py:143-181 edge map
__device__ __forceinline__ void lm_dipole_edge(Map& M, float hx, float e, float fint,
                                                float gap) {
  float se, ce;
  sincosf(e, &se, &ce);
  float phi = fint * hx * gap / ce * (1.f + se * se);
  row_axpy(M, 1, hx * tanf(e), 0);
  row_axpy(M, 3, -hx * tanf(e - phi), 2);
}

// dipole.py:112-141 (+ rbend.py:79-80 via `rbend`)
__device__ __forceinline__ void lm_dipole(Map& M, float L, float angle, float e1, float e2,
                                           float tilt, float fint, float fint_exit, float gap,
                                           bool rbend, const Rel& rel) {
  if (rbend) {
    e1 += 0.5f * angle;
    e2 += 0.5f * angle;
  }
  float hx = (L != 0.f) ? angle / L : 0.f;  // :96-102
  if (tilt != 0.f) lm_rotation(M, tilt);
  lm_dipole_edge(M, hx, e1, fint, gap);
  if (L != 0.f) {
    lm_base(M, L, 0.f, hx, rel);
  } else {
    M.m[2][6] += angle;  // :131-133 as written: the kick lands in R[2,6] (quirk B3)
  }
  lm_dipole_edge(M, hx, e2, fint_exit, gap);
  if (tilt != 0.f) lm_rotation(M, -tilt);
}

// solenoid.py:61-105
__device__ __forceinline__ void lm_solenoid(Map& M, float L, float k, float mx, float my,
                                             const Rel& rel) {
  M.m[0][6] -= mx;
  M.m[2][6] -= my;
  float s, c;
  sincosf(L * k, &s, &c);
  float t = L * k;
  float c1, sk;
  cos_sinc(t * t, c1, sk);  // sk = sin(Lk)/(Lk), exact limit at k = 0
  sk *= L;
  float c2 = c * c, sc = s * c, s2 = s * s;
#pragma unroll
  for (int j = 0; j < 7; ++j) {
    float r0 = M.m[0][j], r1 = M.m[1][j], r2 = M.m[2][j], r3 = M.m[3][j];
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
template <typename F>
__device__ __forceinline__ void lm_custom(Map& M, F&& coef /* (row, col) -> float */) {
  Map T;
#pragma unroll
  for (int i = 0; i < 6; ++i) {
    float c[7];
#pragma unroll
    for (int k = 0; k < 7; ++k) c[k] = coef(i, k);
#pragma unroll
    for (int j = 0; j < 7; ++j) {
      float acc = (j == 6) ? c[6] : 0.f;
#pragma unroll
      for (int k = 0; k < 6; ++k) acc = fmaf(c[k], M.m[k][j], acc);
      T.m[i][j] = acc;
    }
  }
  M = T;
}

// Standing-wave cavity, cavity.py:248-325 (linear map) and :101-200 (non-linear
// scalars).  Evaluated in double: the formulas subtract nearly equal quantities
// ((b0 b1 - 1) g0 g1 + 1, (b0 g0)^3 - (b1 g1)^3) and run once per setting, not per
// particle.  V == 0: exact Drift map (deviation from quirk B6, the 0*inf NaN).
struct CavityOut {
  float sc[kCavFloats];
  float dE;
};

__device__ inline void cavity_setting(float Lf, float Vf, float phase_deg, float freq,
                                       float energy, Map& C, CavityOut& out) {
  map_identity(C);
  const double L = Lf, V = Vf, E = energy, me = kElectronMassEV;
  const double phi = (double)phase_deg * (kPi / 180.0);
  const double cphi = cos(phi), sphi = sin(phi);
  const double dE = V * cphi;
  const double k = 2.0 * kPi * (double)freq / kSpeedOfLight;
#pragma unroll
  for (int i = 0; i < kCavFloats; ++i) out.sc[i] = 0.f;
  out.dE = (float)dE;
  out.sc[CV_DE] = (float)dE;

  // cavity.py:101-112
  double g0 = (E != 0.0) ? E / me : 1e10;
  double igamma2 = (E != 0.0) ? 1.0 / (g0 * g0) : 0.0;
  double beta0 = (E != 0.0) ? sqrt(1.0 - igamma2) : 1.0;

  if (V == 0.0) {
    // drift.py:57-60 with REST_ENERGY-based gamma (what a skippable cavity should be)
    double g = E / kRestEnergyEV;
    double ig2 = (g != 0.0) ? 1.0 / (g * g) : 0.0;
    C.m[0][1] = Lf;
    C.m[2][3] = Lf;
    C.m[4][5] = (float)(-L / (1.0 - ig2) * ig2);
  } else {
    const double Ei = E / me, Ef = (E + dE) / me;
    const double dG = dE / me;           // Ef - Ei without the subtraction
    const double Ep = dG / L;
    const double alpha = sqrt(0.125) / cphi * log1p(dG / Ei);
    const double sa = sin(alpha), ca = cos(alpha);
    const double r11 = ca - sqrt(2.0) * cphi * sa;
    // r12 = sqrt(8) Ei/Ep cphi sin(alpha); finite limit L as dG -> 0 (phi = +-90 deg)
    const double r12 = (dG != 0.0) ? sqrt(8.0) * Ei / Ep * cphi * sa : L;
    const double r21 = -Ep / Ef * (cphi / sqrt(2.0) + sqrt(0.125) / cphi) * sa;
    const double r22 = Ei / Ef * (ca + sqrt(2.0) * cphi * sa);
    double r56 = 0.0, b0 = 1.0, b1 = 1.0, r55c = 0.0;
    if (E != 0.0) {  // :290 branch, per setting
      b0 = sqrt(1.0 - 1.0 / (Ei * Ei));
      b1 = sqrt(1.0 - 1.0 / (Ef * Ef));
      r56 = -L / (Ef * Ef * Ei * b1) * (Ef + Ei) / (b1 + b0);
      r55c = k * L * b0 * V / me * sphi * (Ei * Ef * (b0 * b1 - 1.0) + 1.0) /
             (b1 * Ef * dG * dG);
    }
    const double r66 = Ei / Ef * b0 / b1;
    const double r65 = k * sphi * V / (Ef * b1 * me);
    C.m[0][0] = C.m[2][2] = (float)r11;
    C.m[0][1] = C.m[2][3] = (float)r12;
    C.m[1][0] = C.m[3][2] = (float)r21;
    C.m[1][1] = C.m[3][3] = (float)r22;
    C.m[4][4] = (float)(1.0 + r55c);
    C.m[4][5] = (float)r56;
    C.m[5][4] = (float)r65;
    C.m[5][5] = (float)r66;
  }

  // non-linear scalars, cavity.py:121-200
  double T566 = 1.5 * L * igamma2 / (beta0 * beta0 * beta0), T556 = 0.0, T555 = 0.0;
  const double E1 = E + dE;
  const bool on = E1 > 0.0;  // :128, per setting
  if (on) {
    const double g1 = E1 / me;
    const double beta1 = sqrt(1.0 - 1.0 / (g1 * g1));
    out.sc[CV_AP] = (float)(E * beta0 / (E1 * beta1));
    out.sc[CV_BP] = (float)(V * beta0 / (E1 * beta1));
    out.sc[CV_KB] = (float)(beta0 * k);
    out.sc[CV_PHI] = (float)phi;
    if (dE > 0.0) {  // :164, per setting
      const double dgamma = V / me;
      const double b03 = beta0 * beta0 * beta0, b13 = beta1 * beta1 * beta1;
      const double g03 = g0 * g0 * g0, g13 = g1 * g1 * g1;
      const double d01 = g0 - g1;
      T566 = L * (b03 * g03 - b13 * g13) / (2.0 * beta0 * b13 * g0 * d01 * g13);
      T556 = beta0 * k * L * dgamma * g0 * (b13 * g13 + beta0 * (g0 - g13)) * sphi /
             (b13 * g13 * d01 * d01);
      T555 = beta0 * beta0 * k * k * L * dgamma / 2.0 *
             (dgamma * (2.0 * g0 * g13 * (beta0 * b13 - 1.0) + g0 * g0 + 3.0 * g1 * g1 - 2.0) /
                  (b13 * g13 * d01 * d01 * d01) * sphi * sphi -
              (g1 * g0 * (beta1 * beta0 - 1.0) + 1.0) / (beta1 * g1 * d01 * d01) * cphi);
    }
  }
  out.sc[CV_T566] = (float)T566;
  out.sc[CV_T556] = (float)T556;
  out.sc[CV_T555] = (float)T555;
  out.sc[CV_ON] = on ? 1.f : 0.f;
}

// Non-linear part of Cavity._track_beam for one coordinate vector, cavity.py:142-161,
// :220-226.  (s0, p0) incoming, lin4 = (R v0)[4].  cos(phi - d) - cos(phi) is taken as
// 2 sin(phi - d/2) sin(d/2): the subtraction form loses every digit in fp32 for
// d = s0 beta0 k ~ 1e-4.
__device__ __forceinline__ void cavity_nonlinear(const float* sc, float s0, float p0,
                                                  float lin4, float lin5, float& s1,
                                                  float& p1) {
  if (sc[CV_ON] != 0.f) {
    float d = s0 * sc[CV_KB];
    float dc = 2.f * sinf(sc[CV_PHI] - 0.5f * d) * sinf(0.5f * d);
    p1 = fmaf(p0, sc[CV_AP], sc[CV_BP] * dc);
    s1 = lin4 + fmaf(sc[CV_T566] * p0, p0, fmaf(sc[CV_T556] * s0, p0, sc[CV_T555] * s0 * s0));
  } else {
    p1 = lin5;
    s1 = lin4;
  }
}

}  // namespace lynx

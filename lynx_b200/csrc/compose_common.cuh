// Argument blocks and sinks of K1 / K6 (compose.cu), also used by the along-s kernel and the
// adjoint kernels.
#pragma once
#include <type_traits>
#include "walk.cuh"

namespace lynx {

// ------------------------------------------------------------------ K1
// WITH_MAPS: also emit the 7x7 stage maps (lynx_compose_maps); the tracking launchers use
// the variant without that code (smaller K1).
template <bool WITH_MAPS>
struct RecordSinkT {
  float* rec;
  float* maps_out;  // nullable, this setting's [n_stages][7][7]
  int off_cav, off_centre, off_screen;
  float ctr[6];

  template <bool LT>
  __device__ __forceinline__ void write_map(const MapT<float, LT>& M, int idx) {
    float* d = rec + idx * kMapStride;
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
      for (int j = 0; j < 7; ++j) d[i * 7 + j] = M.m[i][j];
    // slot 42 of the 44-float record: 1 if the map is three 2x2 blocks on (x, xp), (y, yp),
    // (s, p) with a zero kick column -- drifts, untilted and aligned quadrupoles, a cavity's own
    // map -- which the particle kernels apply with 24 instead of 42 operations
    // (track_common.cuh apply_block2, the same bits).  -0 counts as zero: 0 * v + acc == acc.
    bool blocks = true;
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
      for (int j = 0; j < 7; ++j)
        if ((i >> 1) != (j >> 1) && map_live<LT>(i, j)) blocks = blocks && (M.m[i][j] == 0.f);
    d[kMapFloats] = blocks ? 1.f : 0.f;
    if (WITH_MAPS && maps_out) {
      float* o = maps_out + idx * 49;
#pragma unroll
      for (int i = 0; i < 6; ++i)
#pragma unroll
        for (int j = 0; j < 7; ++j) o[i * 7 + j] = M.m[i][j];
#pragma unroll
      for (int j = 0; j < 7; ++j) o[42 + j] = (j == 6) ? 1.f : 0.f;
    }
  }
  // structural entries of a light map are skipped where the dense sum is unchanged by them
  // (0 * ctr added to a non-zero partial sum); rows 4 / 5 keep the dense order of operations
  template <bool LT>
  __device__ __forceinline__ void push_centre(const MapT<float, LT>& M, float* lin) {
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      float a = M.m[i][6];
#pragma unroll
      for (int j = 0; j < 6; ++j)
        if (map_live<LT>(i, j) || i == j) a = fmaf(M.m[i][j], ctr[j], a);
      lin[i] = a;
    }
  }
  template <bool LT>
  __device__ __forceinline__ void stage(const MapT<float, LT>& M, int idx) {
    write_map(M, idx);
    float lin[6];
    push_centre(M, lin);
#pragma unroll
    for (int i = 0; i < 6; ++i) ctr[i] = lin[i];
  }
  __device__ __forceinline__ void cavity(const Map& C, const CavityOut& co, int idx,
                                         int cav_idx) {
    write_map(C, idx);
    float* d = rec + off_cav + cav_idx * kCavFloats;
#pragma unroll
    for (int i = 0; i < kCavFloats; ++i) d[i] = co.sc[i];
    float lin[6];
    push_centre(C, lin);
    float s1, p1;
    cavity_nonlinear(co.sc, ctr[4], ctr[5], lin[4], lin[5], s1, p1);
    lin[4] = s1;
    lin[5] = p1;
#pragma unroll
    for (int i = 0; i < 6; ++i) ctr[i] = lin[i];
  }
  __device__ __forceinline__ bool flush_before(const DevOp& op) const { return op.flush_stage >= 0; }
  template <bool LT>
  __device__ __forceinline__ void finish(const MapT<float, LT>& M, int final_stage) {
    if (final_stage >= 0) stage(M, final_stage);
  }
  template <bool LT>
  __device__ __forceinline__ void bpm(const MapT<float, LT>&, int, int snap) {
    float* d = rec + off_centre + snap * 6;
#pragma unroll
    for (int i = 0; i < 6; ++i) d[i] = ctr[i];
  }
  template <bool LT>
  __device__ __forceinline__ void screen(const MapT<float, LT>&, int, int sidx, float mx, float my, int) {
    rec[off_screen + sidx * 2 + 0] = mx;
    rec[off_screen + sidx * 2 + 1] = my;
  }
};

struct ComposeArgs {
  const DevOp* ops;
  int n_ops, final_stage;
  const int64_t* col_offset;
  const int32_t* col_stride;
  const float* params;
  int64_t B;
  const float* energy_in;
  const float* seed;
  int64_t seed_row_stride, seed_ld;
  float* records;
  int rec_floats, off_cav, off_centre, off_screen, final_centre, n_stages;
  float* maps_out;
  float* energy_out;
  float* zero_ptr;     // optional buffer zeroed by the extra blocks of the same launch
  int64_t zero_vec4;   // its size in float4
  // optional: screen edge tables checked / summarised by the LAST block of the launch
  const float* hist_ex;
  const float* hist_ey;
  int hist_W, hist_H;
  void* hist_out;      // HistAxis[2]
  // persistent screen image (lynx_track_particles_fwd_persistent): instead of the whole
  // buffer the extra blocks clear only the rectangle the PREVIOUS call lit (its bounding-box
  // cell, zero_rect[4]); the last block clears this call's cell
  const int* zero_rect;
  int* bbox_clear;
};

using RecordSink = RecordSinkT<false>;

// ------------------------------------------------------------------ K6
template <typename T>
struct MomentSinkT {
  T mu[7];
  T cov[7][7];
  T* bpm_out;
  int64_t B, b;

  // element.py:73-74 with R = [M; e6].  A light map (maps.cuh) skips the products with its
  // structural zeros and ones, the same bits as the dense arithmetic.  The second product
  // cov R^T is evaluated for the upper triangle only and mirrored: R cov R^T of a symmetric
  // cov is symmetric, the reference's full product differs from its own transpose by rounding
  // (an input that is NOT symmetric -- not a covariance -- comes out symmetrised here).
  // 686 fma dense / 315 light for the full product, 539 / 223 with the triangle.  The
  // dual-number instantiation (K6-bwd) keeps the full product: its tangent seeds ONE entry of
  // cov, which is not a symmetric direction, and d / d cov[i][j] must stay what autograd through
  // the reference's matmul gives.
  static constexpr bool kSymmetric = std::is_same<T, float>::value;
  template <bool LT>
  __device__ __forceinline__ void apply(const MapT<T, LT>& M) {
    T t[7];
    // t[i] = (R v)[i] for i in [lo, 6) and a 7-vector v (mu, a column or a row of cov), dense
    // order of summation
    auto rows_times = [&](auto&& v, int lo) {
#pragma unroll
      for (int i = 0; i < 6; ++i) {
        if (i < lo) continue;
        if (LT && i == 5) {
          t[5] = v(5);
        } else if (LT && i == 4) {
          t[4] = fma_(M.m[4][5], v(5), v(4));
        } else {
          T acc = T(0.f);
#pragma unroll
          for (int j = 0; j < 7; ++j)
            if (map_live<LT>(i, j)) acc = fma_(M.m[i][j], v(j), acc);
          t[i] = acc;
        }
      }
    };
    rows_times([&](int j) { return mu[j]; }, 0);
#pragma unroll
    for (int i = 0; i < 6; ++i) mu[i] = t[i];
    // cov <- R cov  (column by column, in place)
#pragma unroll
    for (int k = 0; k < 7; ++k) {
      rows_times([&](int j) { return cov[j][k]; }, 0);
#pragma unroll
      for (int i = 0; i < 6; ++i) cov[i][k] = t[i];
    }
    if constexpr (kSymmetric) {
      // cov <- cov R^T: entries (i, l >= i) row by row, in place (row i only reads row i) ...
#pragma unroll
      for (int i = 0; i < 6; ++i) {
        rows_times([&](int k) { return cov[i][k]; }, i);
#pragma unroll
        for (int l = 0; l < 6; ++l)
          if (l >= i) cov[i][l] = t[l];
      }
      // ... and the mirror image (column 6 of the product is column 6 of R cov)
#pragma unroll
      for (int i = 0; i < 7; ++i)
#pragma unroll
        for (int l = 0; l < 7; ++l)
          if (l > i) cov[l][i] = cov[i][l];
    } else {
      // cov <- cov R^T  (row by row, in place)
#pragma unroll
      for (int i = 0; i < 7; ++i) {
        rows_times([&](int k) { return cov[i][k]; }, 0);
#pragma unroll
        for (int l = 0; l < 6; ++l) cov[i][l] = t[l];
      }
    }
  }
  template <bool LT>
  __device__ __forceinline__ void stage(const MapT<T, LT>& M, int) { apply(M); }
  // cavity.py:113-140, 202-218 (ParameterBeam branch; overwrites as written, quirk B5)
  __device__ __forceinline__ void cavity(const MapT<T>& C, const CavityOutT<T>& co, int, int) {
    const T mu4 = mu[4], mu5 = mu[5];
    const T c44 = cov[4][4], c45 = cov[4][5], c55 = cov[5][5];
    apply(C);
    const T* sc = co.sc;
    if (val(sc[CV_ON]) != 0.f) {
      T s1, p1;
      cavity_nonlinear(sc, mu4, mu5, mu[4], mu[5], s1, p1);
      mu[4] = s1;
      mu[5] = p1;
      cov[5][5] = c55;
      T v = sc[CV_T566] * c55 * c55 + sc[CV_T556] * c45 * c55 + sc[CV_T555] * c44 * c44;
      cov[4][4] = v;
      cov[4][5] = v;
      cov[5][4] = v;
    }
  }
  __device__ __forceinline__ bool flush_before(const DevOp& op) const { return op.flush_stage >= 0; }
  template <bool LT>
  __device__ __forceinline__ void finish(const MapT<T, LT>& M, int final_stage) {
    if (final_stage >= 0) apply(M);
  }
  template <bool LT>
  __device__ __forceinline__ void bpm(const MapT<T, LT>&, int slot, int) {
    if (bpm_out) {
      bpm_out[(slot * B + b) * 2 + 0] = mu[0];  // bpm.py:52
      bpm_out[(slot * B + b) * 2 + 1] = mu[2];
    }
  }
  template <bool LT>
  __device__ __forceinline__ void screen(const MapT<T, LT>&, int, int, T mx, T my, int) {
    mu[0] -= mx;  // screen.py:131-132
    mu[2] -= my;
  }
};
using MomentSink = MomentSinkT<float>;

struct ParamBeamArgs {
  const DevOp* ops;
  int n_ops, final_stage;
  const int64_t* col_offset;
  const int32_t* col_stride;
  const float* params;
  int64_t B;
  const float* mu_in;
  const float* cov_in;
  int64_t beam_stride;  // 0 or 1 (in beams)
  const float* energy_in;
  float* mu_out;
  float* cov_out;
  float* bpm_out;
  float* energy_out;
};

}  // namespace lynx

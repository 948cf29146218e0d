// K1 / K6, warp-per-setting variant for small batches.
//
// With few settings the thread-per-setting kernels are pure latency: one thread walks
// ~2-3k dependent instructions (12 us for 11 elements).  Here one warp owns a setting:
//   phase B  lane e builds the dense 6x7 map of element e (all elements in parallel; the
//            energy seen by each element is a warp prefix sum over active cavities)
//   phase C  lanes 0..6 each own one COLUMN of the accumulated map -- left-multiplication
//            acts on columns independently -- and fold the element maps in beam order
//   stages   are handed to the same RecordSink / MomentSink logic, with mu / cov
//            products spread over the lanes (K6)
// One setting per warp, shuffle prefix sums and reductions, as the north star sketches
// for the ParameterBeam path.
#include <algorithm>

#include "compose_common.cuh"

namespace lynx {

constexpr int kCWWarps = 4;                 // settings per CTA
constexpr int kCWThreads = kCWWarps * 32;
constexpr int kCWMap = 44;                  // floats per element map slot
// per-warp scratch (floats)
constexpr int kCWScratch = 32 * kCWMap + 32 * kCavFloats + kCWMap + 8 + 56 + 56 + 8;

struct WarpScratch {
  float* maps;  // [32][44] element maps of the current chunk
  float* cav;   // [32][12] cavity scalars of the current chunk
  float* M;     // [44]     current stage map (rows 0..5, row-major 6x7)
  float* mu;    // [8]
  float* cov;   // [56]     7x7 row-major
  float* T;     // [56]
  float* ctr;   // [8]
};

__device__ __forceinline__ WarpScratch warp_scratch(float* base) {
  WarpScratch w;
  w.maps = base;
  w.cav = w.maps + 32 * kCWMap;
  w.M = w.cav + 32 * kCavFloats;
  w.mu = w.M + kCWMap;
  w.cov = w.mu + 8;
  w.T = w.cov + 56;
  w.ctr = w.T + 56;
  return w;
}

// mu <- R mu, cov <- R cov R^T with R = [M; e6], M (6x7) in shared memory; 32 lanes.
__device__ __forceinline__ void warp_apply_moments(const WarpScratch& w, int lane) {
  float nmu = 0.f;
  if (lane < 6) {
#pragma unroll
    for (int j = 0; j < 7; ++j) nmu = fmaf(w.M[lane * 7 + j], w.mu[j], nmu);
  }
  // T = R cov
  for (int idx = lane; idx < 49; idx += 32) {
    const int i = idx / 7, k = idx - i * 7;
    float acc;
    if (i < 6) {
      acc = 0.f;
#pragma unroll
      for (int j = 0; j < 7; ++j) acc = fmaf(w.M[i * 7 + j], w.cov[j * 7 + k], acc);
    } else {
      acc = w.cov[42 + k];
    }
    w.T[idx] = acc;
  }
  __syncwarp();
  if (lane < 6) w.mu[lane] = nmu;
  // cov = T R^T
  for (int idx = lane; idx < 49; idx += 32) {
    const int i = idx / 7, l = idx - i * 7;
    float acc;
    if (l < 6) {
      acc = 0.f;
#pragma unroll
      for (int k = 0; k < 7; ++k) acc = fmaf(w.T[i * 7 + k], w.M[l * 7 + k], acc);
    } else {
      acc = w.T[i * 7 + 6];
    }
    w.cov[idx] = acc;
  }
  __syncwarp();
}

struct CWArgs {
  const DevOp* ops;
  int n_ops, final_stage, n_cols;
  const int64_t* col_offset;
  const int32_t* col_stride;
  const float* params;
  int64_t B;
  const float* energy_in;
  float* energy_out;
  // K1
  const float* seed;
  int64_t seed_row_stride, seed_ld;
  float* records;
  int rec_floats, off_cav, off_centre, off_screen, final_centre, n_stages;
  float* maps_out;
  // K6
  const float* mu_in;
  const float* cov_in;
  int64_t beam_stride;
  float* mu_out;
  float* cov_out;
  float* bpm_out;
};

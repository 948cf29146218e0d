// Shared pieces of the K2 particle kernels (LDG variant in track.cu, TMA-bulk pipelined
// variant in track_pipe.cu).
#pragma once
#include "maps.cuh"

namespace lynx {

// Per-axis histogram geometry (hist_axis_setup): prepared by K1 for the pipelined kernel,
// per CTA by the LDG kernel.
struct HistAxis {
  float e0, en, inv, margin;  // margin >= 0.5: the table is always consulted
};

struct TrackArgs {
  const float* in;
  int64_t in_row_stride;  // floats between beam rows of `in` (0: one beam for all settings)
  int64_t ld_in, N;
  float* out;
  int64_t ld_out;
  const float* records;
  int rec_floats;
  const PlanItem* plan;
  int n_plan;
  int n_snap;
  int final_centre;
  double* bpm_sums;
  float* img;
  int W, H;
  const float* ex;
  const float* ey;
  const HistAxis* hist;  // [2] x, y geometry written by K1 (pipelined kernel)
  unsigned long long* stamp;  // [4] debug %globaltimer cell in a symmetric slot's box, or null
  int tile;     // particles per pipeline tile (K2 pipe variant; set by its launcher)
  float* rep;   // histogram replicas 1..n_rep-1: [n_rep - 1][B][H][W], all zero on entry
  int n_rep;    // CTA x histograms into replica x % n_rep (0 = img); folded by hist_fold_kernel
  int* bbox;             // nullable: [4] max(W - ix), max(ix + 1), max(H - iy), max(iy + 1)
  double* mom;
  int64_t B;
  int dbg;  // experiments only (LYNX_DBG): bit0 = skip the screen atomics
  // ---- fused composer (track_pipe_kernel<.., OPS != 0>): every CTA walks its own setting's
  //      lattice into the shared-memory record; no K1 launch, no record round trip
  const DevOp* ops;
  int n_ops, final_stage, n_cols;
  const int64_t* col_offset;
  const int32_t* col_stride;
  const float* params;
  const float* energy_in;
  float* energy_out;
  int off_cav, off_centre, off_screen;
  unsigned int* grid_bar;  // [2] arrival count, generation (library-owned, self-resetting)
  int zero_img;            // the kernel zeroes the image (+ bbox) itself, behind a grid barrier
};

__device__ __forceinline__ float4 ld_stream(const float* p) {
  return __ldg(reinterpret_cast<const float4*>(p));
}
__device__ __forceinline__ void st_stream(float* p, float4 v) {
  __stcs(reinterpret_cast<float4*>(p), v);
}

// v <- A v + c for kNP particles; A|c = 6x7 row-major in shared memory (16-B aligned)
template <int NP>
__device__ __forceinline__ void apply_affine(const float* __restrict__ m, float (&v)[6][NP]) {
  float r[kMapStride];
  const float4* m4 = reinterpret_cast<const float4*>(m);
#pragma unroll
  for (int i = 0; i < kMapStride / 4; ++i) {
    float4 t = m4[i];
    r[4 * i] = t.x;
    r[4 * i + 1] = t.y;
    r[4 * i + 2] = t.z;
    r[4 * i + 3] = t.w;
  }
#pragma unroll
  for (int j = 0; j < NP; ++j) {
    float n[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      float acc = r[i * 7 + 6];
#pragma unroll
      for (int k = 0; k < 6; ++k) acc = fmaf(r[i * 7 + k], v[k][j], acc);
      n[i] = acc;
    }
#pragma unroll
    for (int i = 0; i < 6; ++i) v[i][j] = n[i];
  }
}

// The linear map of an active cavity (cavity.py:248-325, maps.cuh cavity_setting) is three
// 2x2 blocks on (x, xp), (y, yp), (s, p) with a zero kick column: 12 of the 42 entries.  Same
// order of operations as apply_affine on the non-zero entries, so the result is bit-identical
// for finite coordinates (fma(0, v, acc) == acc).  A dense product also turns EVERY coordinate
// of a particle into NaN as soon as one of them is not finite (0 * inf, 0 * NaN); the term z
// (0 for a finite particle, NaN otherwise) reproduces that.  24 instead of 42 operations.
template <int NP>
__device__ __forceinline__ void apply_block2(const float* __restrict__ m, float (&v)[6][NP]) {
  float a[3][4];
#pragma unroll
  for (int b = 0; b < 3; ++b) {
    a[b][0] = m[(2 * b) * 7 + 2 * b];
    a[b][1] = m[(2 * b) * 7 + 2 * b + 1];
    a[b][2] = m[(2 * b + 1) * 7 + 2 * b];
    a[b][3] = m[(2 * b + 1) * 7 + 2 * b + 1];
  }
#pragma unroll
  for (int j = 0; j < NP; ++j) {
    float z = 0.f;
#pragma unroll
    for (int k = 0; k < 6; ++k) z = fmaf(v[k][j], 0.f, z);
#pragma unroll
    for (int b = 0; b < 3; ++b) {
      const float u = v[2 * b][j], w = v[2 * b + 1][j];
      v[2 * b][j] = fmaf(a[b][1], w, fmaf(a[b][0], u, 0.f)) + z;
      v[2 * b + 1][j] = fmaf(a[b][3], w, fmaf(a[b][2], u, 0.f)) + z;
    }
  }
}

// Bin lookup against the reference's fp32 edge table: left-closed bins, last bin
// right-closed, out-of-range dropped (torch.histogramdd semantics).
// Slow path: the table decides (exact).
static __device__ __noinline__ int find_bin_table(float x, const float* __restrict__ e, int n, int i) {
  i = min(max(i, 0), n - 1);
  if (x < e[i]) {
    do { --i; } while (i > 0 && x < e[i]);
  } else if (x >= e[i + 1] && i < n - 1) {
    do { ++i; } while (i < n - 1 && x >= e[i + 1]);
  }
  return i;
}


// Fast path: t = (x - e0) * n/(en - e0) names the bin outright unless its fractional
// part is within `margin` of an edge; margin covers the fp32 error of t (<= ~3 eps n)
// plus the verified deviation of the table from an ideal linspace (hist_axis_setup),
// so every particle lands in exactly the bin the table would give.
__device__ __forceinline__ int find_bin(float x, const float* __restrict__ e, int n,
                                        const HistAxis& h) {
  if (!(x >= h.e0 && x <= h.en)) return -1;
  const float t = (x - h.e0) * h.inv;
  const int i = (int)t;
  const float fr = t - (float)i;
  if (i < n && fr > h.margin && fr < 1.f - h.margin) return i;
  return find_bin_table(x, e, n, i);
}

// Cooperative (whole CTA) check that edge table `e` (n+1 values) deviates from the ideal
// linspace by < 2^-10 of a bin; fills `h`.  Contains __syncthreads.
__device__ __forceinline__ void hist_axis_setup(const float* __restrict__ e, int n, HistAxis* h,
                                                int tid, int nthreads) {
  const double e0 = (double)e[0], en = (double)e[n];
  const double step = (en - e0) / n;
  int bad = 0;
  for (int i = tid; i <= n; i += nthreads)
    bad |= fabs((double)e[i] - (e0 + step * i)) > step * (1.0 / 1024.0);
  bad = __syncthreads_or(bad);
  if (tid == 0) {
    h->e0 = (float)e0;
    h->en = (float)en;
    h->inv = (float)n / ((float)en - (float)e0);
    const float m = fmaxf(1.f / 256.f, (float)n * (1.f / 1048576.f));
    h->margin = bad ? 1.f : m;
  }
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

template <int MOM>
struct MomAcc {
  static constexpr int kN = (MOM == 2) ? 27 : (MOM == 1 ? 14 : 1);
  float a[kN];
};

// index of (i,j), i<=j in the 21-entry upper triangle, row-major
__host__ __device__ constexpr int tri(int i, int j) { return i * 6 - i * (i - 1) / 2 + (j - i); }

// TMA-bulk pipelined variant (track_pipe.cu); returns 1 if it cannot take the problem.
int launch_track_pipe(const TrackArgs& a, int moment_mode, bool has_out, int sm_count,
                      cudaStream_t st, int fused_ops = 0);

// Accumulate shifted first / second moments of one particle (d = v - centre).
template <int MOM>
__device__ __forceinline__ void mom_accumulate(MomAcc<MOM>& acc, const float (&d)[6]) {
#pragma unroll
  for (int k = 0; k < 6; ++k) acc.a[k] += d[k];
  if (MOM == 2) {
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
      for (int j = i; j < 6; ++j) acc.a[6 + tri(i, j)] = fmaf(d[i], d[j], acc.a[6 + tri(i, j)]);
  } else if (MOM == 1) {
#pragma unroll
    for (int k = 0; k < 6; ++k) acc.a[6 + k] = fmaf(d[k], d[k], acc.a[6 + k]);
    acc.a[12] = fmaf(d[0], d[1], acc.a[12]);
    acc.a[13] = fmaf(d[2], d[3], acc.a[13]);
  }
}

// CTA epilogue: per-thread fp32 shifted sums -> double warp sums -> raw (unshifted)
// double sums -> one atomicAdd per moment per CTA.  `w` = warp index in [0, NWARPS) for
// participating warps (others pass w < 0 and only take part in the barrier).
template <int MOM, int NWARPS>
__device__ __forceinline__ void mom_epilogue(const MomAcc<MOM>& acc, double (*s_red)[28], int w,
                                             int lane, bool leader, int64_t n_mine,
                                             const float (&ctr)[6], double* g) {
  constexpr int kN = MomAcc<MOM>::kN;
  if (w >= 0) {
#pragma unroll
    for (int i = 0; i < kN; ++i) {
      double s = warp_sum((double)acc.a[i]);
      if (lane == 0) s_red[w][i] = s;
    }
  }
  __syncthreads();
  if (!leader) return;
  double S1[6], S2[21];
  bool have[21];
  for (int i = 0; i < 21; ++i) {
    S2[i] = 0.0;
    have[i] = (MOM == 2);
  }
  auto col = [&](int i) {
    double s = 0.0;
    for (int ww = 0; ww < NWARPS; ++ww) s += s_red[ww][i];
    return s;
  };
  for (int i = 0; i < 6; ++i) S1[i] = col(i);
  if (MOM == 2) {
    for (int i = 0; i < 21; ++i) S2[i] = col(6 + i);
  } else {
    for (int k = 0; k < 6; ++k) {
      S2[tri(k, k)] = col(6 + k);
      have[tri(k, k)] = true;
    }
    S2[tri(0, 1)] = col(12);
    have[tri(0, 1)] = true;
    S2[tri(2, 3)] = col(13);
    have[tri(2, 3)] = true;
  }
  const double n = (double)n_mine;
  atomicAdd(g, n);
  double c[6];
  for (int i = 0; i < 6; ++i) c[i] = (double)ctr[i];
  for (int i = 0; i < 6; ++i) atomicAdd(g + 1 + i, S1[i] + n * c[i]);
  for (int i = 0; i < 6; ++i)
    for (int j = i; j < 6; ++j)
      if (have[tri(i, j)])
        atomicAdd(g + 7 + tri(i, j), S2[tri(i, j)] + c[i] * S1[j] + c[j] * S1[i] + n * c[i] * c[j]);
}

}  // namespace lynx

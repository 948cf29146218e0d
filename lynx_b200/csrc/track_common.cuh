// Shared pieces of the K2 particle kernels (LDG variant in track.cu, TMA-bulk pipelined
// variant in track_pipe.cu).
#pragma once
#include "maps.cuh"

namespace lynx {

// Per-axis histogram geometry (hist_axis_setup): prepared by K1 for the pipelined kernel,
// per CTA by the LDG kernel.
struct HistAxis {
  float e0, en, inv, margin;  // inv = n / (en - e0); margin >= 0.5: the table is always consulted
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
  PlanItem plan_inline[8];  // the first plan items by value: short plans need no global load in the prologue
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
  int tile;          // particles per pipeline tile (K2 pipe variant; set by its launcher)
  int stages;        // ring depth of the pipelined kernel (2..4; the histogram tile takes the rest)
  int ctas_per_row;  // CTAs that share one setting: b = blockIdx.x / ctas_per_row
  int xy_only;       // the plan is [AFFINE, SCREEN] and nothing else reads the beam: rows x, y only
  int xy_tail;       // longer plans ending [.., AFFINE, SCREEN] with nothing else reading the beam:
                     // plan index of that last stage (x, y rows only), else -1
  int* bbox;         // nullable: [4] max(W - ix), max(ix + 1), max(H - iy), max(iy + 1)
  double* mom;
  int64_t B;
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

// Rows 0 and 2 only: (x, y) of A v + c, in apply_affine's order of operations (bit-identical).
// What a terminal screen needs of the last stage when nothing else reads the beam.
template <int NP>
__device__ __forceinline__ void apply_affine_xy(const float* __restrict__ m, const float (&v)[6][NP],
                                                float (&x)[NP], float (&y)[NP]) {
  float rx[7], ry[7];
#pragma unroll
  for (int k = 0; k < 7; ++k) {
    rx[k] = m[k];
    ry[k] = m[14 + k];
  }
#pragma unroll
  for (int j = 0; j < NP; ++j) {
    float ax = rx[6], ay = ry[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      ax = fmaf(rx[k], v[k][j], ax);
      ay = fmaf(ry[k], v[k][j], ay);
    }
    x[j] = ax;
    y[j] = ay;
  }
}

// A map that is three 2x2 blocks on (x, xp), (y, yp), (s, p) with a zero kick column: 12 of the
// 42 entries -- the linear map of an active cavity (cavity.py:248-325, maps.cuh cavity_setting)
// and every stage K1 flags (compose_common.cuh write_map).  Same order of operations as
// apply_affine on the non-zero entries, so the result is bit-identical for finite coordinates
// (fma(0, v, acc) == acc).  A dense product also multiplies every coordinate OUTSIDE a row's
// block by zero: one that is not finite turns the row into NaN (0 * inf, 0 * NaN), while inside
// its own block an infinity stays an infinity.  The terms t (0 for a finite pair, NaN otherwise)
// reproduce exactly that: row block b gets the t of the other two blocks.  27 instead of 42
// operations.
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
    float t[3];
#pragma unroll
    for (int b = 0; b < 3; ++b) t[b] = fmaf(v[2 * b + 1][j], 0.f, fmaf(v[2 * b][j], 0.f, 0.f));
#pragma unroll
    for (int b = 0; b < 3; ++b) {
      const float z = t[(b + 1) % 3] + t[(b + 2) % 3];
      const float u = v[2 * b][j], w = v[2 * b + 1][j];
      v[2 * b][j] = fmaf(a[b][1], w, fmaf(a[b][0], u, 0.f)) + z;
      v[2 * b + 1][j] = fmaf(a[b][3], w, fmaf(a[b][2], u, 0.f)) + z;
    }
  }
}

// Bin lookup against the reference's fp32 edge table: left-closed bins, last bin
// right-closed, out-of-range dropped (torch.histogramdd semantics).
// Slow path: the range check and the table decide (exact).  Out of line: it holds dependent
// table loads and is taken by a few particles in a thousand.
static __device__ __noinline__ int find_bin_slow(float x, const float* __restrict__ e, int n, int i,
                                                 float e0, float en) {
  if (!(x >= e0 && x <= en)) return -1;
  i = min(max(i, 0), n - 1);
  if (x < e[i]) {
    do { --i; } while (i > 0 && x < e[i]);
  } else if (x >= e[i + 1] && i < n - 1) {
    do { ++i; } while (i < n - 1 && x >= e[i + 1]);
  }
  return i;
}

// Regular tables (every screen the reference builds: torch.linspace), exact.
// t = x * inv + off = (x - e0) * n / (en - e0).  margin = d + delta (hist_axis_setup): d < 0.2 is
// the MEASURED deviation of the table from the ideal linspace through its end points, in bins;
// delta >= |t - t*| bounds the fp32 error of t for x in [e0, en] and |e0|, |en| <= en - e0.  With
// inv_f = (n / D)(1 + a), |a| <= 2^-23 (the rounded difference D_f, then the rounded quotient),
// off_f = -e0 inv_f (1 + b), |b| <= 2^-24, and one rounding c of the fma:
//   t = ((x - e0) inv_f - e0 inv_f b)(1 + c) = u (1 + a) - (e0 n / D)(1 + a) b + t c,
// the error of inv is common to both terms and acts on u = (x - e0) n / D <= n, not on x inv and
// e0 inv separately: |t - u| <= n (2^-23 + 2^-24 + 2^-24) = n 2^-22 in the worst case.
// hist_axis_finish evaluates the first two terms for the table at hand instead of bounding them.
//   frac(t) in (margin, 1 - margin): the particle lies in bin floor(t) of the TABLE.
//   otherwise it is within margin + delta < 0.26 bins of the ideal edge j = round(t), hence more
//   than 0.49 bins from the ideal edges j - 1 and j + 1, whose table values are off by < 0.2:
//   x lies strictly between e[j - 1] and e[j + 1], and ONE comparison with e[j] (the 10 KB table
//   is L1-resident) decides between bins j - 1 and j exactly as a search of the table would.
// The near-edge block is a short inline branch, not a call: an out-of-line search cost a whole
// warp ~50 instructions each time ONE lane took it (0.3 % per lane = 9 % per warp instruction:
// 100 instructions per warp and tile, measured); doing the near-edge work branch-free for every
// lane cost 29 instructions per lookup (measured as well).  x in [e0, en] implies floor(t) in
// [-1, n] and j in [0, n], so nothing needs clamping.
// Two-step form, so that a thread can do the cheap part for all its lookups first and then take
// ONE branch for the rare near-edge cases (a branch per lookup cost ~9 instructions each in
// reconvergence bookkeeping, measured):
//   bin_guess: floor(t), or -1 outside [e0, en]; `near` = the table has to decide
//   bin_near_edge: the decision
__device__ __forceinline__ int bin_guess(float x, float inv, float off, float margin, float e0,
                                         float en, bool& near, float& fr_out) {
  const float t = fmaf(x, inv, off);
  const int i = __float2int_rd(t);
  const float fr = t - (float)i;
  const bool valid = x >= e0 && x <= en;  // NaN fails both
  near = valid && !(fabsf(fr - 0.5f) < 0.5f - margin);
  fr_out = fr;
  return valid ? i : -1;
}
__device__ __forceinline__ int bin_near_edge(float x, const float* __restrict__ e, int n, int i,
                                             float fr) {
  const int j = i + (fr > 0.5f ? 1 : 0);  // the nearest edge
  return min(x >= __ldg(e + j) ? j : j - 1, n - 1);  // x == en: the last bin is right-closed
}
__device__ __forceinline__ int find_bin_regular(float x, const float* __restrict__ e, int n,
                                                float inv, float off, float margin, float e0,
                                                float en) {
  bool near;
  float fr;
  int b = bin_guess(x, inv, off, margin, e0, en, near, fr);
  if (near) b = bin_near_edge(x, e, n, b, fr);
  return b;
}

// Any table: irregular ones (margin >= 0.5) are searched.
__device__ __forceinline__ int find_bin(float x, const float* __restrict__ e, int n,
                                        const HistAxis& h) {
  const float off = __fmul_rn(-h.e0, h.inv);  // the value whose error hist_axis_finish measured
  if (h.margin >= 0.5f) {
    const float t = fmaf(x, h.inv, off);
    return find_bin_slow(x, e, n, (t >= 0.f && t < (float)n) ? (int)t : 0, h.e0, h.en);
  }
  return find_bin_regular(x, e, n, h.inv, off, h.margin, h.e0, h.en);
}

// Geometry of one axis from its end points and the measured deviation d of the table (bins).
// The rounding error of t is bounded with THIS table's numbers instead of their worst case:
//   t = fl(x inv_f + off_f),  x inv_f + off_f = u (inv_f / inv) + (off_f + e0 inv_f),
// so |t - u| <= n a + b + (n + 1) 2^-24 with a = |inv_f / inv - 1| (<= 2^-23) and
// b = |off_f + e0 inv_f| (<= n 2^-24), both evaluated here in fp64 from the very fp32 values the
// particle kernel uses (it forms off_f = -e0 * inv_f with one fp32 multiplication, as below).
// Worst case n 2^-22; the reference's 2448 x 2040 screens come out at 0.35 - 0.6 of that, and
// with them the share of lookups the table has to decide.  2^-20 of slack covers the rounding of
// this evaluation, of frac(t) - 0.5 and of 0.5 - margin; 0.25 d that d is a rounded measurement.
__device__ __forceinline__ void hist_axis_finish(HistAxis* h, float fe0, float fen, int n, float d,
                                                 double step) {
  const float inv_f = (float)n / (fen - fe0);
  const float off_f = __fmul_rn(-fe0, inv_f);
  const double inv = (double)n / ((double)fen - (double)fe0);
  const double a = fabs((double)inv_f - inv) / inv;
  const double b = fabs((double)off_f + (double)fe0 * (double)inv_f);
  const float eps = (float)((double)n * a + b + (double)(n + 1) * 5.9604644775390625e-08);
  const float m = 1.25f * d + 1.001f * eps + (1.f / 1048576.f);
  const bool ok = d < 0.2f && m < 0.25f && step > 0.0 && fmaxf(fabsf(fe0), fabsf(fen)) <= fen - fe0;
  h->e0 = fe0;
  h->en = fen;
  h->inv = inv_f;
  h->margin = ok ? m : 1.f;  // margin >= 0.5 means: irregular table, always searched
}

// Cooperative (whole CTA) measurement of how far edge table `e` (n+1 values) deviates from the
// ideal linspace through its end points; fills `h`.  Contains __syncthreads.
__device__ __forceinline__ void hist_axis_setup(const float* __restrict__ e, int n, HistAxis* h,
                                                int tid, int nthreads) {
  __shared__ unsigned int s_dev;  // max deviation in bins (bits of a non-negative float)
  if (tid == 0) s_dev = 0u;
  __syncthreads();
  const double e0 = (double)e[0], en = (double)e[n];
  const double step = (en - e0) / n;
  const double inv_step = 1.0 / step;  // one division, not one per table entry: this runs in K1's
                                       // last block, on the critical path of the 1 M-particle step
  float dev = 0.f;
  for (int i = tid; i <= n; i += nthreads) {
    const float d = (float)(fabs((double)e[i] - (e0 + step * i)) * inv_step);
    dev = (d <= 1e30f) ? fmaxf(dev, d) : 2.f;  // NaN / inf in the table: never trust the fast path
  }
  atomicMax(&s_dev, __float_as_uint(dev));
  __syncthreads();
  if (tid == 0) {
    hist_axis_finish(h, (float)e0, (float)en, n, __uint_as_float(s_dev), step);
  }
  __syncthreads();
}

// One axis by one block of K1 (two histogram blocks per launch, x and y side by side).  The
// table is staged in shared memory by cp.async -- every element in flight at once, ONE DRAM
// round trip per 4096-entry chunk -- where hist_axis_setup's loop paid one cold miss per
// iteration (20 per axis on the 2448 x 2040 screen; the block was K1's critical path,
// tools/k2_timeline.py).  Same arithmetic, hence the same margin, as hist_axis_setup: the
// deviation is measured in fp64 -- an fp32 measurement needs ~0.002 bins of extra margin, which
// doubles the near-edge lookups of the particle kernel and cost it 5 % at 32 M particles.
// Contains __syncthreads.
constexpr int kHistChunk = 4096;
#ifdef LYNX_EXPERIMENTS
#define LYNX_TSTAMP(ptr, i)                                             \
  do {                                                                  \
    if (threadIdx.x == 0 && (ptr)) {                                    \
      unsigned long long t_;                                            \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));            \
      reinterpret_cast<unsigned long long*>(ptr)[i] = t_;               \
    }                                                                   \
  } while (0)
#else
#define LYNX_TSTAMP(ptr, i) do {} while (0)
#endif
__device__ __forceinline__ void hist_axis_setup_staged(const float* __restrict__ e, int n,
                                                       HistAxis* h, int tid, int nthreads,
                                                       float* s_e /* [kHistChunk] */, void* stamps) {
  __shared__ unsigned int s_dev;
  if (tid == 0) s_dev = 0u;
  const float f0 = __ldg(e), fn = __ldg(e + n);  // in flight together with the first chunk
  float dev = 0.f;
  double e0 = 0.0, step = 0.0, inv_step = 0.0;
  const int chunks = (n + kHistChunk) / kHistChunk;  // covers n + 1 entries
  for (int c = 0; c < chunks; ++c) {
    const int base = c * kHistChunk;
    for (int i = tid; i < kHistChunk && base + i <= n; i += nthreads)
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(
                       (uint32_t)__cvta_generic_to_shared(&s_e[i])),
                   "l"(e + base + i)
                   : "memory");
    LYNX_TSTAMP(stamps, 12);
    asm volatile("cp.async.wait_all;" ::: "memory");
    __syncthreads();
    LYNX_TSTAMP(stamps, 13);
    if (c == 0) {
      e0 = (double)f0;
      step = ((double)fn - e0) / n;
      inv_step = 1.0 / step;
    }
    for (int i = tid; i < kHistChunk && base + i <= n; i += nthreads) {
      const float d = (float)(fabs((double)s_e[i] - (e0 + step * (base + i))) * inv_step);
      dev = (d <= 1e30f) ? fmaxf(dev, d) : 2.f;  // NaN / inf: never trust the fast path
    }
    __syncthreads();  // the chunk buffer is refilled next
  }
  LYNX_TSTAMP(stamps, 14);
  atomicMax(&s_dev, __float_as_uint(dev));
  __syncthreads();
  if (tid == 0) {
    HistAxis r;
    hist_axis_finish(&r, f0, fn, n, __uint_as_float(s_dev), step);
    *h = r;
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
                      cudaStream_t st);

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

// K3 / K6-bwd: adjoint of the tracking path.
//
// The reference has no custom_vjp: gradients there come from the array framework
// differentiating segment.py:340-356 / element.py:61-94 / cavity.py:97-246 op by op
// (SURVEY.md 3.4).  Here the vjp is three device steps:
//
//   adj_compose_kernel  per setting: the total 6x7 map T_k of every cavity-free SEGMENT
//                       (BPMs do not split a segment), the constant cotangent injections
//                       of BPM readings, cavity maps + scalars, moment cotangents
//   adj_stream_kernel   (K3a) one pass over the particles: forward through the segments,
//                       then backward, accumulating per segment the single 6x7 sum
//                       H_k = sum_n gbar_n (x) [u_n; 1] (every parameter gradient of the
//                       segment is <H_k, dT_k/dtheta>), per cavity a 6x7 sum + 7 scalar
//                       sums; writes g_in if requested.       24 B (x) + 24 B (g_out)
//                       [+ 24 B g_in] per particle, independent of the element count
//   adj_param_kernel    (K3b) one thread per (setting, parameter column): walks the
//                       program in forward-mode dual arithmetic (dual.cuh) through the
//                       same templated map algebra K1 uses and contracts dT_k with H_k
//
// Moment collapse: without active cavities, per-particle cotangents and g_in, H follows
// from the 28 raw moment sums of the INCOMING beam (one pass per beam row, not per
// setting): adj_collapse_kernel.
#include <algorithm>
#include <cstdlib>

#include "pipe.cuh"
#include "compose_common.cuh"
#include "track_common.cuh"

namespace lynx {

constexpr int kAdjBlock = 56;   // floats per segment / cavity block of the adjoint record
constexpr int kAdjInj = 44;     // segment block: injection vector (6)
constexpr int kAdjCu = 50;      // segment block: centre of the segment's start state (6)
constexpr int kAdjSc = 44;      // cavity block: scalars (kCavFloats)
constexpr int kAdjBpm = 16;     // floats per BPM block: rows 0 and 2 of the prefix map
constexpr int kAdjFinal = 48;   // ctr_out[6], gm1c[6], G2sym[36]
constexpr int kAdjSums = 56;    // doubles per target in the sums buffer
constexpr int kMaxSeg = 8;      // segments the streaming kernel keeps per particle

struct AdjLayout {
  int n_seg, n_cav, n_snap;
  int off_bpm, off_final, rec_floats;
};

__host__ __device__ inline AdjLayout adj_layout(int n_cav, int n_snap) {
  AdjLayout L;
  L.n_cav = n_cav;
  L.n_seg = n_cav + 1;
  L.n_snap = n_snap;
  L.off_bpm = (L.n_seg + L.n_cav) * kAdjBlock;
  L.off_final = L.off_bpm + n_snap * kAdjBpm;
  L.rec_floats = L.off_final + kAdjFinal;
  return L;
}

__device__ __forceinline__ bool is_active_cavity(const DevOp& op) {
  return op.opcode == LYNX_OP_CAVITY && (op.flags & LYNX_F_ACTIVE);
}

// ------------------------------------------------------------------ adjoint compose
struct AdjRecordSink {
  float* rec;
  AdjLayout L;
  const double* g_bpm;  // [n_slots][B][2] cotangent of bpm_sums, nullable
  int64_t B, b;
  int k;         // current segment
  float inj[6];  // sum over this segment's BPMs of P_b^T g_b
  float ctr[6];  // centre (seed particle) at the current position

  __device__ __forceinline__ bool flush_before(const DevOp& op) const { return is_active_cavity(op); }

  __device__ __forceinline__ void push_centre(const Map& M, float* lin) const {
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      float a = M.m[i][6];
#pragma unroll
      for (int j = 0; j < 6; ++j) a = fmaf(M.m[i][j], ctr[j], a);
      lin[i] = a;
    }
  }
  __device__ __forceinline__ void write_block(float* d, const Map& M) const {
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
      for (int j = 0; j < 7; ++j) d[i * 7 + j] = M.m[i][j];
  }
  __device__ __forceinline__ void segment(const Map& M) {
    float* d = rec + k * kAdjBlock;
    write_block(d, M);
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      d[kAdjInj + i] = inj[i];
      d[kAdjCu + i] = ctr[i];
      inj[i] = 0.f;
    }
    float lin[6];
    push_centre(M, lin);
#pragma unroll
    for (int i = 0; i < 6; ++i) ctr[i] = lin[i];
    ++k;
  }
  __device__ __forceinline__ void stage(const Map& M, int) { segment(M); }
  __device__ __forceinline__ void cavity(const Map& C, const CavityOut& co, int, int cav_idx) {
    float* d = rec + (L.n_seg + cav_idx) * kAdjBlock;
    write_block(d, C);
#pragma unroll
    for (int i = 0; i < kCavFloats; ++i) d[kAdjSc + i] = co.sc[i];
    float lin[6];
    push_centre(C, lin);
    float s1, p1;
    cavity_nonlinear(co.sc, ctr[4], ctr[5], lin[4], lin[5], s1, p1);
    lin[4] = s1;
    lin[5] = p1;
#pragma unroll
    for (int i = 0; i < 6; ++i) ctr[i] = lin[i];
  }
  __device__ __forceinline__ void bpm(const Map& M, int slot, int snap) {
    float* d = rec + L.off_bpm + snap * kAdjBpm;
#pragma unroll
    for (int j = 0; j < 7; ++j) {
      d[j] = M.m[0][j];
      d[7 + j] = M.m[2][j];
    }
    if (g_bpm) {
      const float gx = (float)g_bpm[(slot * B + b) * 2 + 0];
      const float gy = (float)g_bpm[(slot * B + b) * 2 + 1];
#pragma unroll
      for (int j = 0; j < 6; ++j) inj[j] = fmaf(M.m[0][j], gx, fmaf(M.m[2][j], gy, inj[j]));
    }
  }
  __device__ __forceinline__ void screen(const Map&, int, int, float, float, int) {}
  __device__ __forceinline__ void finish(const Map& M, int) { segment(M); }
};

struct AdjComposeArgs {
  const DevOp* ops;
  int n_ops, final_stage;
  const int64_t* col_offset;
  const int32_t* col_stride;
  const float* params;
  int64_t B;
  const float* energy_in;
  const float* seed;
  int64_t seed_row_stride, seed_ld;
  float* arec;
  AdjLayout L;
  const double* g_bpm;
  const double* g_mom;  // [B][28] nullable
};

__global__ void __launch_bounds__(128) adj_compose_kernel(AdjComposeArgs a) {
  const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= a.B) return;
  AdjRecordSink sink;
  sink.rec = a.arec + b * a.L.rec_floats;
  sink.L = a.L;
  sink.g_bpm = a.g_bpm;
  sink.B = a.B;
  sink.b = b;
  sink.k = 0;
#pragma unroll
  for (int i = 0; i < 6; ++i) {
    sink.inj[i] = 0.f;
    sink.ctr[i] = a.seed ? __ldg(a.seed + b * a.seed_row_stride + i * a.seed_ld) : 0.f;
  }
  float energy = a.energy_in[b];
  ParamView P{a.params, a.col_offset, a.col_stride, b};
  walk_program<kOpsAll>(a.ops, a.n_ops, a.final_stage, P, energy, sink);
  // final block: outgoing centre, moment cotangents re-centred on it
  float* f = sink.rec + a.L.off_final;
  double G2[6][6], gm1[6];
#pragma unroll
  for (int i = 0; i < 6; ++i) {
    gm1[i] = a.g_mom ? a.g_mom[b * 28 + 1 + i] : 0.0;
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      const int t = 7 + tri(min(i, j), max(i, j));
      G2[i][j] = a.g_mom ? (i == j ? 2.0 : 1.0) * a.g_mom[b * 28 + t] : 0.0;
    }
  }
#pragma unroll
  for (int i = 0; i < 6; ++i) {
    double s = gm1[i];
#pragma unroll
    for (int j = 0; j < 6; ++j) s += G2[i][j] * (double)sink.ctr[j];
    f[i] = sink.ctr[i];
    f[6 + i] = (float)s;
#pragma unroll
    for (int j = 0; j < 6; ++j) f[12 + i * 6 + j] = (float)G2[i][j];
  }
}

// ------------------------------------------------------------------ K3a streaming adjoint
struct AdjStreamArgs {
  const float* in;
  int64_t in_row_stride, ld_in, N;
  const float* g_out;  // [B][6][ld_g] nullable
  int64_t ld_g;
  float* g_in;         // nullable; row stride gin_row_stride (0: reduced over settings)
  int64_t gin_row_stride, ld_gin;
  int atomic_gin;
  const float* arec;
  AdjLayout L;
  double* sums;        // [B][n_seg + n_cav][kAdjSums]
  int target;          // which segment / cavity this launch accumulates
  int has_mom, has_bpm;
  int64_t B;
};

constexpr int kAdjTPB = 256;

__device__ __forceinline__ void affine6(const float* __restrict__ m, const float (&x)[6],
                                        float (&y)[6]) {
#pragma unroll
  for (int i = 0; i < 6; ++i) {
    float a = m[i * 7 + 6];
#pragma unroll
    for (int j = 0; j < 6; ++j) a = fmaf(m[i * 7 + j], x[j], a);
    y[i] = a;
  }
}

// One particle through the adjoint.  NOCAV: a single segment, everything in registers.
template <bool NOCAV>
__device__ __forceinline__ void adj_particle(const AdjStreamArgs& a, const float* __restrict__ rec,
                                             const float (&x)[6], const float (&go)[6],
                                             bool has_go, float (&acc)[kAdjSums],
                                             float (&gin)[6], const float* __restrict__ seg0 = nullptr) {
  // seg0 (NOCAV only): the single segment block held by the caller, e.g. in registers
  if (!NOCAV || !seg0) seg0 = rec;
  const int n_seg = NOCAV ? 1 : a.L.n_seg;
  float u[NOCAV ? 1 : kMaxSeg][6];
  float v[6];
#pragma unroll
  for (int i = 0; i < 6; ++i) v[i] = x[i];
  // ---- forward
  for (int k = 0; k < n_seg; ++k) {
#pragma unroll
    for (int i = 0; i < 6; ++i) u[k][i] = v[i];
    float w[6];
    affine6(NOCAV ? seg0 : rec + k * kAdjBlock, v, w);
    if (!NOCAV && k < n_seg - 1) {
      const float* cb = rec + (n_seg + k) * kAdjBlock;
      float vl[6];
      affine6(cb, w, vl);
      float s1, p1;
      cavity_nonlinear(cb + kAdjSc, w[4], w[5], vl[4], vl[5], s1, p1);
      vl[4] = s1;
      vl[5] = p1;
#pragma unroll
      for (int i = 0; i < 6; ++i) v[i] = vl[i];
    } else {
#pragma unroll
      for (int i = 0; i < 6; ++i) v[i] = w[i];
    }
  }
  // ---- cotangent of the outgoing particle
  float g[6];
  const float* fin = rec + a.L.off_final;
#pragma unroll
  for (int i = 0; i < 6; ++i) g[i] = has_go ? go[i] : 0.f;
  if (a.has_mom) {
    float d[6];
#pragma unroll
    for (int j = 0; j < 6; ++j) d[j] = v[j] - fin[j];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      float s = fin[6 + i];
#pragma unroll
      for (int j = 0; j < 6; ++j) s = fmaf(fin[12 + i * 6 + j], d[j], s);
      g[i] += s;
    }
  }
  // ---- backward
  for (int k = n_seg - 1; k >= 0; --k) {
    const float* sb = NOCAV ? seg0 : rec + k * kAdjBlock;
    if (a.target == k) {
      float d[6];
#pragma unroll
      for (int j = 0; j < 6; ++j) d[j] = u[k][j] - sb[kAdjCu + j];
#pragma unroll
      for (int i = 0; i < 6; ++i) {
#pragma unroll
        for (int j = 0; j < 6; ++j) acc[i * 7 + j] = fmaf(g[i], d[j], acc[i * 7 + j]);
        acc[i * 7 + 6] += g[i];
      }
      if (a.has_bpm) {
#pragma unroll
        for (int j = 0; j < 6; ++j) acc[42 + j] += d[j];
      }
    }
    float gn[6];
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      float s = sb[kAdjInj + j];
#pragma unroll
      for (int i = 0; i < 6; ++i) s = fmaf(sb[i * 7 + j], g[i], s);
      gn[j] = s;
    }
    if (!NOCAV && k > 0) {
      // cavity k-1 sits between segment k-1 and segment k: v1 = cav(w), w = T_{k-1} u_{k-1}
      const float* cb = rec + (n_seg + k - 1) * kAdjBlock;
      const float* sc = cb + kAdjSc;
      float w[6];
      affine6(rec + (k - 1) * kAdjBlock, u[k - 1], w);
      const bool on = sc[CV_ON] != 0.f;
      const bool mine = a.target == n_seg + k - 1;
      const float s0 = w[4], p0 = w[5];
      const int rows = on ? 4 : 6;  // rows of C that reach the output linearly
      float gw[6];
#pragma unroll
      for (int j = 0; j < 6; ++j) gw[j] = 0.f;
#pragma unroll
      for (int i = 0; i < 6; ++i) {
        // row 4 feeds s1 also when the non-linear part is on; row 5 only when it is off
        const bool lin = i < rows || (on && i == 4);
        if (lin) {
#pragma unroll
          for (int j = 0; j < 6; ++j) gw[j] = fmaf(cb[i * 7 + j], gn[i], gw[j]);
          if (mine) {
#pragma unroll
            for (int j = 0; j < 6; ++j) acc[i * 7 + j] = fmaf(gn[i], w[j], acc[i * 7 + j]);
            acc[i * 7 + 6] += gn[i];
          }
        }
      }
      if (on) {
        const float kb = sc[CV_KB], phi = sc[CV_PHI], bp = sc[CV_BP];
        const float d = s0 * kb;
        float sn, sphi;
        sn = sinf(phi - d);
        sphi = sinf(phi);
        const float dc = 2.f * sinf(phi - 0.5f * d) * sinf(0.5f * d);
        gw[4] += gn[4] * fmaf(sc[CV_T556], p0, 2.f * sc[CV_T555] * s0) + gn[5] * bp * sn * kb;
        gw[5] += gn[4] * fmaf(2.f * sc[CV_T566], p0, sc[CV_T556] * s0) + gn[5] * sc[CV_AP];
        if (mine) {
          acc[42 + CV_T566] = fmaf(gn[4], p0 * p0, acc[42 + CV_T566]);
          acc[42 + CV_T556] = fmaf(gn[4], s0 * p0, acc[42 + CV_T556]);
          acc[42 + CV_T555] = fmaf(gn[4], s0 * s0, acc[42 + CV_T555]);
          acc[42 + CV_AP] = fmaf(gn[5], p0, acc[42 + CV_AP]);
          acc[42 + CV_BP] = fmaf(gn[5], dc, acc[42 + CV_BP]);
          acc[42 + CV_KB] = fmaf(gn[5], bp * sn * s0, acc[42 + CV_KB]);
          // d/dphi [cos(phi - d) - cos(phi)] = -sin(phi - d) + sin(phi)
          //   = 2 cos(phi - d/2) sin(d/2)  (cancellation-free)
          acc[42 + CV_PHI] = fmaf(gn[5], bp * 2.f * cosf(phi - 0.5f * d) * sinf(0.5f * d),
                                  acc[42 + CV_PHI]);
        }
        (void)sphi;
      }
#pragma unroll
      for (int j = 0; j < 6; ++j) g[j] = gw[j];
    } else {
#pragma unroll
      for (int j = 0; j < 6; ++j) g[j] = gn[j];
    }
  }
#pragma unroll
  for (int j = 0; j < 6; ++j) gin[j] = g[j];
}

template <bool NOCAV>
__global__ void __launch_bounds__(kAdjTPB) adj_stream_kernel(AdjStreamArgs a) {
  extern __shared__ __align__(16) float arec_s[];
  __shared__ double s_red[kAdjTPB / 32][kAdjSums];
  const int tid = threadIdx.x;
  const int64_t b = blockIdx.y;
  {
    const float* src = a.arec + b * a.L.rec_floats;
    for (int i = tid; i < a.L.rec_floats; i += kAdjTPB) arec_s[i] = src[i];
  }
  __syncthreads();
  const float* in_b = a.in + b * a.in_row_stride;
  const float* go_b = a.g_out ? a.g_out + b * 6 * a.ld_g : nullptr;
  float* gi_b = a.g_in ? a.g_in + b * a.gin_row_stride : nullptr;
  float acc[kAdjSums];
#pragma unroll
  for (int i = 0; i < kAdjSums; ++i) acc[i] = 0.f;

  for (int64_t idx = ((int64_t)blockIdx.x * kAdjTPB + tid) * 4; idx < a.N;
       idx += (int64_t)gridDim.x * kAdjTPB * 4) {
    float4 xv[6], gv[6], ov[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      xv[k] = ld_stream(in_b + k * a.ld_in + idx);
      gv[k] = go_b ? ld_stream(go_b + k * a.ld_g + idx) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int l = 0; l < 4; ++l) {
      float x[6], go[6], gin[6];
#pragma unroll
      for (int k = 0; k < 6; ++k) {
        x[k] = reinterpret_cast<const float*>(&xv[k])[l];
        go[k] = reinterpret_cast<const float*>(&gv[k])[l];
      }
      if (idx + l < a.N) {
        adj_particle<NOCAV>(a, arec_s, x, go, go_b != nullptr, acc, gin);
      } else {
#pragma unroll
        for (int k = 0; k < 6; ++k) gin[k] = 0.f;
      }
#pragma unroll
      for (int k = 0; k < 6; ++k) reinterpret_cast<float*>(&ov[k])[l] = gin[k];
    }
    if (gi_b) {
      if (a.atomic_gin) {
#pragma unroll
        for (int k = 0; k < 6; ++k)
#pragma unroll
          for (int l = 0; l < 4; ++l)
            if (idx + l < a.N)
              atomicAdd(gi_b + k * a.ld_gin + idx + l, reinterpret_cast<float*>(&ov[k])[l]);
      } else {
#pragma unroll
        for (int k = 0; k < 6; ++k) st_stream(gi_b + k * a.ld_gin + idx, ov[k]);
      }
    }
  }

  // ---- CTA reduction -> double atomics
  const int w = tid >> 5, lane = tid & 31;
#pragma unroll
  for (int i = 0; i < kAdjSums; ++i) {
    const double s = warp_sum((double)acc[i]);
    if (lane == 0) s_red[w][i] = s;
  }
  __syncthreads();
  if (tid < kAdjSums) {
    double s = 0.0;
#pragma unroll
    for (int ww = 0; ww < kAdjTPB / 32; ++ww) s += s_red[ww][tid];
    const int n_t = a.L.n_seg + a.L.n_cav;
    if (s != 0.0) atomicAdd(a.sums + ((int64_t)b * n_t + a.target) * kAdjSums + tid, s);
  }
}

// ------------------------------------------------------------------ K3a, pipelined
// Same arithmetic (adj_particle), data path of K2 (track_pipe.cu): one persistent CTA per SM,
// warp 0 streams the x (and g_out) planes of a tile into shared memory with cp.async.bulk +
// mbarrier, 8 consumer warps run the adjoint per particle out of shared memory and write
// g_in over the x planes in place, warp 1 drains finished tiles with bulk stores.  The
// grid-stride kernel above keeps 18 float4 of particle data in registers (182 registers, one
// 256-thread CTA per SM, no prefetch): 0.60 of the HBM peak at 32 M particles.  Here the bytes
// in flight are set by shared memory (4 stages x 48 KB) and the consumers hold the accumulators
// and -- for a cavity-free lattice -- the segment's map in registers (the first version read
// the 6x7 map from shared memory twice per particle: LSU pipe 58 % busy, 35 % short-scoreboard
// stalls, 0.69 of the HBM peak).  Particle q of a tile belongs to consumer q mod 256:
// LDS.32 / STS.32 are conflict-free without transposing.
constexpr int kAPConsumers = 256;
constexpr int kAPWarps = kAPConsumers / 32;
constexpr int kAPThreads = kAPConsumers + 64;
constexpr int kAPTile = 4 * kAPConsumers;  // particles per stage
constexpr int kAPStages = 4;

template <bool NOCAV, bool HAS_GO, bool HAS_GIN>
__global__ void __launch_bounds__(kAPThreads, 1) adj_pipe_kernel(AdjStreamArgs a, int tile_i) {
  extern __shared__ __align__(128) unsigned char apsm[];
  constexpr int kPlanes = HAS_GO ? 12 : 6;
  constexpr int kStage = kPlanes * kAPTile;  // floats
  float* stages = reinterpret_cast<float*>(apsm);
  float* arec_s = stages + kAPStages * kStage;
  uint64_t* full = reinterpret_cast<uint64_t*>(arec_s + ((a.L.rec_floats + 3) & ~3));
  uint64_t* empty = full + kAPStages;
  uint64_t* done = empty + kAPStages;
  __shared__ double s_red[kAPWarps][kAdjSums];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t b = blockIdx.y;
  const int64_t tile = tile_i;
  const int64_t n_tiles = (a.N + tile - 1) / tile;
  const float* in_b = a.in + b * a.in_row_stride;
  const float* go_b = HAS_GO ? a.g_out + b * 6 * a.ld_g : nullptr;
  float* gi_b = HAS_GIN ? a.g_in + b * a.gin_row_stride : nullptr;

  auto issue_tile = [&](int it, int64_t t) {
    const int s = it % kAPStages;
    const int64_t p0 = t * tile;
    const uint32_t cnt = (uint32_t)min(tile, a.N - p0);
    const uint32_t bytes = ((cnt + 3u) & ~3u) * 4u;
    mbar_arrive_expect_tx(full + s, (uint32_t)kPlanes * bytes);
    float* dst = stages + s * kStage;
#pragma unroll
    for (int k = 0; k < 6; ++k) bulk_g2s(dst + k * kAPTile, in_b + k * a.ld_in + p0, bytes, full + s);
    if (HAS_GO) {
#pragma unroll
      for (int k = 0; k < 6; ++k)
        bulk_g2s(dst + (6 + k) * kAPTile, go_b + k * a.ld_g + p0, bytes, full + s);
    }
  };
  if (tid == 0) {
    for (int s = 0; s < kAPStages; ++s) {
      mbar_init(full + s, 1);
      mbar_init(empty + s, HAS_GIN ? 1 : kAPWarps);
      mbar_init(done + s, kAPWarps);
    }
    fence_mbar_init();
    int it = 0;
    for (int64_t t = blockIdx.x; t < n_tiles && it < kAPStages; t += gridDim.x, ++it) issue_tile(it, t);
  }
  {
    const float* src = a.arec + b * a.L.rec_floats;
    for (int i = tid; i < a.L.rec_floats; i += kAPThreads) arec_s[i] = src[i];
  }
  __syncthreads();

  float acc[kAdjSums];
#pragma unroll
  for (int i = 0; i < kAdjSums; ++i) acc[i] = 0.f;

  if (warp == 0) {
    if (lane == 0) {
      int it = 0;
      for (int64_t t = blockIdx.x; t < n_tiles; t += gridDim.x, ++it) {
        if (it < kAPStages) continue;
        mbar_wait(empty + (it % kAPStages), ((uint32_t)(it / kAPStages) - 1) & 1);
        issue_tile(it, t);
      }
    }
  } else if (warp == 1) {
    if (HAS_GIN && lane == 0) {
      int it = 0;
      for (int64_t t = blockIdx.x; t < n_tiles; t += gridDim.x, ++it) {
        const int s = it % kAPStages;
        mbar_wait(done + s, (uint32_t)(it / kAPStages) & 1);
        fence_proxy_async();
        const int64_t p0 = t * tile;
        const uint32_t cnt = (uint32_t)min(tile, a.N - p0);
        const uint32_t bytes = ((cnt + 3u) & ~3u) * 4u;
        const float* src = stages + s * kStage;
#pragma unroll
        for (int k = 0; k < 6; ++k) bulk_s2g(gi_b + k * a.ld_gin + p0, src + k * kAPTile, bytes);
        bulk_commit();
        bulk_wait_read<0>();  // the stage has been read: hand it back to the loader
        mbar_arrive(empty + s);
      }
    }
  } else {
    const int cid = tid - 64;
    float seg[kAdjBlock];  // NOCAV: map, injection and centre of the one segment, in registers
    if (NOCAV) {
#pragma unroll
      for (int i = 0; i < kAdjBlock; ++i) seg[i] = arec_s[i];
    }
    int it = 0;
    for (int64_t t = blockIdx.x; t < n_tiles; t += gridDim.x, ++it) {
      const int s = it % kAPStages;
      const int64_t p0 = t * tile;
      const int cnt = (int)min(tile, a.N - p0);
      float* st = stages + s * kStage;
      mbar_wait(full + s, (uint32_t)(it / kAPStages) & 1);
#pragma unroll 1
      for (int q = cid; q < ((cnt + 3) & ~3); q += kAPConsumers) {
        float x[6], go[6], gin[6];
#pragma unroll
        for (int k = 0; k < 6; ++k) {
          x[k] = st[k * kAPTile + q];
          go[k] = HAS_GO ? st[(6 + k) * kAPTile + q] : 0.f;
        }
        if (q < cnt) {
          adj_particle<NOCAV>(a, arec_s, x, go, HAS_GO, acc, gin, NOCAV ? seg : nullptr);
        } else {  // padding lanes [N, ld) of the last tile
#pragma unroll
          for (int k = 0; k < 6; ++k) gin[k] = 0.f;
        }
        if (HAS_GIN) {
#pragma unroll
          for (int k = 0; k < 6; ++k) st[k * kAPTile + q] = gin[k];
        }
      }
      if (HAS_GIN) fence_proxy_async();  // generic-proxy writes before the bulk store reads
      __syncwarp();
      if (lane == 0) mbar_arrive(HAS_GIN ? done + s : empty + s);
    }
  }

  // ---- CTA reduction -> double atomics (consumer warps hold the sums)
  if (warp >= 2) {
#pragma unroll
    for (int i = 0; i < kAdjSums; ++i) {
      const double sdbl = warp_sum((double)acc[i]);
      if (lane == 0) s_red[warp - 2][i] = sdbl;
    }
  }
  __syncthreads();
  if (tid < kAdjSums) {
    double sdbl = 0.0;
#pragma unroll
    for (int ww = 0; ww < kAPWarps; ++ww) sdbl += s_red[ww][tid];
    const int n_t = a.L.n_seg + a.L.n_cav;
    if (sdbl != 0.0) atomicAdd(a.sums + ((int64_t)b * n_t + a.target) * kAdjSums + tid, sdbl);
  }
}

template <bool NOCAV, bool HAS_GO, bool HAS_GIN>
static cudaError_t launch_adj_pipe(const AdjStreamArgs& sa, dim3 grid, size_t smem, int tile,
                                   cudaStream_t st) {
  auto k = adj_pipe_kernel<NOCAV, HAS_GO, HAS_GIN>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 8192);
    if (e != cudaSuccess) return e;
    attr_set = true;
  }
  k<<<grid, kAPThreads, smem, st>>>(sa, tile);
  return cudaGetLastError();
}

// Returns cudaSuccess after enqueueing, cudaErrorNotSupported if the variant cannot take it.
static cudaError_t adj_pipe_dispatch(const AdjStreamArgs& sa, bool nocav, int sm_count, cudaStream_t st) {
  const bool has_go = sa.g_out != nullptr, has_gin = sa.g_in != nullptr;
  if (has_gin && sa.atomic_gin) return cudaErrorNotSupported;  // reduced over settings: atomics
  if ((sa.ld_in & 3) || (has_go && (sa.ld_g & 3)) || (has_gin && (sa.ld_gin & 3))) return cudaErrorNotSupported;
  const size_t planes = has_go ? 12 : 6;
  const size_t smem = sizeof(float) * (kAPStages * planes * kAPTile + ((sa.L.rec_floats + 3) & ~3)) +
                      8 * 3 * kAPStages;
  if (smem > 227 * 1024 - 8192) return cudaErrorNotSupported;
  int64_t per_b = std::max<int64_t>(1, sm_count / sa.B);
  const int64_t full_tiles = (sa.N + kAPTile - 1) / kAPTile;
  const int64_t gx = std::min<int64_t>(full_tiles, per_b);
  const int64_t rounds = (sa.N + gx * kAPTile - 1) / (gx * kAPTile);
  int64_t tile = (sa.N + gx * rounds - 1) / (gx * rounds);
  tile = std::min<int64_t>(kAPTile, (tile + 31) & ~(int64_t)31);
  const int64_t n_tiles = (sa.N + tile - 1) / tile;
  dim3 grid((unsigned)std::min<int64_t>(n_tiles, gx), (unsigned)sa.B);
  const int v = (nocav ? 4 : 0) + (has_go ? 2 : 0) + (has_gin ? 1 : 0);
  switch (v) {
    case 0: return launch_adj_pipe<false, false, false>(sa, grid, smem, (int)tile, st);
    case 1: return launch_adj_pipe<false, false, true>(sa, grid, smem, (int)tile, st);
    case 2: return launch_adj_pipe<false, true, false>(sa, grid, smem, (int)tile, st);
    case 3: return launch_adj_pipe<false, true, true>(sa, grid, smem, (int)tile, st);
    case 4: return launch_adj_pipe<true, false, false>(sa, grid, smem, (int)tile, st);
    case 5: return launch_adj_pipe<true, false, true>(sa, grid, smem, (int)tile, st);
    case 6: return launch_adj_pipe<true, true, false>(sa, grid, smem, (int)tile, st);
    default: return launch_adj_pipe<true, true, true>(sa, grid, smem, (int)tile, st);
  }
}

// ------------------------------------------------------------------ K3b parameter gradients
using DF = Dual<float>;

// Parameter view whose column `seed_col` carries the unit tangent.
struct DualParamView {
  const float* params;
  const int64_t* col_offset;
  const int32_t* col_stride;
  int64_t b;
  int seed_col;
  __device__ __forceinline__ DF operator()(int col) const {
    const float v = __ldg(params + __ldg(col_offset + col) + b * (int64_t)__ldg(col_stride + col));
    return DF(v, col == seed_col ? 1.f : 0.f);
  }
};

// Contracts the tangents of the segment maps / cavity maps / cavity scalars with the sums
// of the streaming pass.
struct AdjGradSink {
  const double* sums;   // this setting's [n_t][kAdjSums]
  const float* rec;     // this setting's adjoint record (centres)
  AdjLayout L;
  const double* g_bpm;
  int64_t B, b;
  double n_particles;
  int k;
  double g;

  __device__ __forceinline__ bool flush_before(const DevOp& op) const { return is_active_cavity(op); }

  __device__ __forceinline__ void segment(const MapT<DF>& M) {
    const double* H = sums + k * kAdjSums;
    const float* cu = rec + k * kAdjBlock + kAdjCu;
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      const double h6 = H[i * 7 + 6];
#pragma unroll
      for (int j = 0; j < 6; ++j) s += (H[i * 7 + j] + (double)cu[j] * h6) * (double)M.m[i][j].d;
      s += h6 * (double)M.m[i][6].d;
    }
    g += s;
    ++k;
  }
  __device__ __forceinline__ void stage(const MapT<DF>& M, int) { segment(M); }
  __device__ __forceinline__ void finish(const MapT<DF>& M, int) { segment(M); }
  __device__ __forceinline__ void cavity(const MapT<DF>& C, const CavityOutT<DF>& co, int, int cav) {
    const double* G = sums + (L.n_seg + cav) * kAdjSums;
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
      for (int j = 0; j < 7; ++j) s += G[i * 7 + j] * (double)C.m[i][j].d;
#pragma unroll
    for (int i = 0; i <= CV_T555; ++i) s += G[42 + i] * (double)co.sc[i].d;
    g += s;
  }
  __device__ __forceinline__ void bpm(const MapT<DF>& M, int slot, int) {
    if (!g_bpm) return;
    // reading sums = (P_b S1)_{0,2},  S1 = sum_n [u_n; 1] of this segment's start state
    const double gx = g_bpm[(slot * B + b) * 2 + 0], gy = g_bpm[(slot * B + b) * 2 + 1];
    const double* S = sums + k * kAdjSums + 42;
    const float* cu = rec + k * kAdjBlock + kAdjCu;
    double s = 0.0;
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      const double s1 = S[j] + n_particles * (double)cu[j];
      s += s1 * (gx * (double)M.m[0][j].d + gy * (double)M.m[2][j].d);
    }
    s += n_particles * (gx * (double)M.m[0][6].d + gy * (double)M.m[2][6].d);
    g += s;
  }
  __device__ __forceinline__ void screen(const MapT<DF>&, int, int, DF, DF, int) {}
};

struct AdjParamArgs {
  const DevOp* ops;
  int n_ops, final_stage, n_cols;
  const int64_t* col_offset;
  const int32_t* col_stride;
  const float* params;
  int64_t B;
  const float* energy_in;
  const float* arec;
  AdjLayout L;
  const double* sums;
  const double* g_bpm;
  double n_particles;
  float* g_params;
};

__global__ void __launch_bounds__(128) adj_param_kernel(AdjParamArgs a) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= a.B * a.n_cols) return;
  const int c = (int)(idx / a.B);
  const int64_t b = idx - (int64_t)c * a.B;
  AdjGradSink sink;
  const int n_t = a.L.n_seg + a.L.n_cav;
  sink.sums = a.sums + b * n_t * kAdjSums;
  sink.rec = a.arec + b * a.L.rec_floats;
  sink.L = a.L;
  sink.g_bpm = a.g_bpm;
  sink.B = a.B;
  sink.b = b;
  sink.n_particles = a.n_particles;
  sink.k = 0;
  sink.g = 0.0;
  DF energy(a.energy_in[b], 0.f);
  DualParamView P{a.params, a.col_offset, a.col_stride, b, c};
  walk_program<kOpsAll>(a.ops, a.n_ops, a.final_stage, P, energy, sink);
  float* dst = a.g_params + a.col_offset[c];
  if (a.col_stride[c])
    dst[b] = (float)sink.g;
  else
    atomicAdd(dst, (float)sink.g);
}

// ------------------------------------------------------------------ moment collapse
// Raw moment sums of beam rows: [R][28] double (N, sum v_i, sum v_i v_j i<=j).
constexpr int kMomTPB = 256;
__global__ void __launch_bounds__(kMomTPB) beam_moments_kernel(const float* __restrict__ in,
                                                               int64_t row_stride, int64_t ld,
                                                               int64_t N, double* __restrict__ out) {
  __shared__ double s_red[kMomTPB / 32][28];
  const int tid = threadIdx.x;
  const int64_t r = blockIdx.y;
  const float* in_r = in + r * row_stride;
  float ctr[6];
#pragma unroll
  for (int k = 0; k < 6; ++k) ctr[k] = __ldg(in_r + k * ld);
  MomAcc<2> acc;
#pragma unroll
  for (int i = 0; i < 27; ++i) acc.a[i] = 0.f;
  int64_t n_mine = 0;
  for (int64_t idx = ((int64_t)blockIdx.x * kMomTPB + tid) * 4; idx < N;
       idx += (int64_t)gridDim.x * kMomTPB * 4) {
    float4 xv[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) xv[k] = ld_stream(in_r + k * ld + idx);
#pragma unroll
    for (int l = 0; l < 4; ++l) {
      const bool live = idx + l < N;
      float d[6];
#pragma unroll
      for (int k = 0; k < 6; ++k)
        d[k] = live ? reinterpret_cast<const float*>(&xv[k])[l] - ctr[k] : 0.f;
      mom_accumulate(acc, d);
    }
  }
  // every CTA counts its own particles
  {
    const int64_t per = (int64_t)kMomTPB * 4;
    for (int64_t t = blockIdx.x; t * per < N; t += gridDim.x) n_mine += min(per, N - t * per);
  }
  mom_epilogue<2, kMomTPB / 32>(acc, s_red, tid >> 5, tid & 31, tid == 0, n_mine, ctr, out + r * 28);
}

// H'_0 (centred like the streaming pass) and S1' from the incoming beam's raw sums:
//   H = gm1 (x) S1 + G2 T S2,  S1 = [sum x; N],  S2 = sum [x;1][x;1]^T
struct AdjCollapseArgs {
  const float* arec;
  AdjLayout L;
  const double* in_sums;  // [Bb][28]
  int64_t in_sums_stride; // 0 or 28
  const double* g_mom;    // [B][28] nullable
  double* sums;           // [B][1][kAdjSums]
  int64_t B;
};

__global__ void __launch_bounds__(128) adj_collapse_kernel(AdjCollapseArgs a) {
  const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= a.B) return;
  const float* rec = a.arec + b * a.L.rec_floats;
  const double* S = a.in_sums + b * a.in_sums_stride;
  double* out = a.sums + b * kAdjSums;
  double S1[7], S2[7][7];
  const double n = S[0];
  for (int i = 0; i < 6; ++i) {
    S1[i] = S[1 + i];
    S2[i][6] = S2[6][i] = S[1 + i];
    for (int j = i; j < 6; ++j) S2[i][j] = S2[j][i] = S[7 + tri(i, j)];
  }
  S1[6] = n;
  S2[6][6] = n;
  double H[6][7];
  for (int i = 0; i < 6; ++i)
    for (int j = 0; j < 7; ++j) H[i][j] = 0.0;
  if (a.g_mom) {
    const double* gm = a.g_mom + b * 28;
    // TS = T S2  (6x7), T = [rec rows; affine column 6]
    double TS[6][7];
    for (int l = 0; l < 6; ++l)
      for (int j = 0; j < 7; ++j) {
        double s = 0.0;
        for (int m = 0; m < 7; ++m) s += (double)rec[l * 7 + m] * S2[m][j];
        TS[l][j] = s;
      }
    for (int i = 0; i < 6; ++i)
      for (int j = 0; j < 7; ++j) {
        double s = gm[1 + i] * S1[j];
        for (int l = 0; l < 6; ++l) {
          const double g2 = (i == l ? 2.0 : 1.0) * gm[7 + tri(min(i, l), max(i, l))];
          s += g2 * TS[l][j];
        }
        H[i][j] = s;
      }
  }
  // BPM injections are constant per particle: gbar at the segment start gains inj, so
  // H += inj-independent...  (injections act through AdjGradSink::bpm, not through H)
  const float* cu = rec + kAdjCu;
  for (int i = 0; i < 6; ++i) {
    for (int j = 0; j < 6; ++j) out[i * 7 + j] = H[i][j] - (double)cu[j] * H[i][6];
    out[i * 7 + 6] = H[i][6];
  }
  for (int j = 0; j < 6; ++j) out[42 + j] = S1[j] - n * (double)cu[j];
}

// ------------------------------------------------------------------ K6-bwd
// vjp of lynx_parambeam_fwd: one thread per (setting, direction); directions are the
// parameter columns, then (if requested) the 7 entries of mu_in and the 49 of cov_in.
// The forward walk is repeated in dual arithmetic with the direction seeded, and the
// output tangents are contracted with (g_mu, g_cov, g_bpm).
struct PBGradSink : MomentSinkT<DF> {
  const float* g_bpm;  // [n_slots][B][2] nullable
  float gb;
  __device__ __forceinline__ void bpm(const MapT<DF>&, int slot, int) {
    if (g_bpm) {
      gb = fmaf(g_bpm[(slot * B + b) * 2 + 0], mu[0].d, gb);
      gb = fmaf(g_bpm[(slot * B + b) * 2 + 1], mu[2].d, gb);
    }
  }
};

struct PBBwdArgs {
  const DevOp* ops;
  int n_ops, final_stage, n_cols, n_dirs;
  const int64_t* col_offset;
  const int32_t* col_stride;
  const float* params;
  int64_t B;
  const float* mu_in;
  const float* cov_in;
  int64_t beam_stride;
  const float* energy_in;
  const float* g_mu;
  const float* g_cov;
  const float* g_bpm;
  float* g_params;
  float* g_mu_in;
  float* g_cov_in;
};

__global__ void __launch_bounds__(64) parambeam_bwd_kernel(PBBwdArgs a) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= a.B * a.n_dirs) return;
  const int dir = (int)(idx / a.B);
  const int64_t b = idx - (int64_t)dir * a.B;
  const int beam_dir = dir - a.n_cols;  // >= 0: 0..6 mu, 7..55 cov
  PBGradSink sink;
  const float* mi = a.mu_in + b * a.beam_stride * 7;
  const float* ci = a.cov_in + b * a.beam_stride * 49;
#pragma unroll
  for (int i = 0; i < 7; ++i) sink.mu[i] = DF(__ldg(mi + i), beam_dir == i ? 1.f : 0.f);
#pragma unroll
  for (int i = 0; i < 7; ++i)
#pragma unroll
    for (int j = 0; j < 7; ++j)
      sink.cov[i][j] = DF(__ldg(ci + i * 7 + j), beam_dir == 7 + i * 7 + j ? 1.f : 0.f);
  sink.bpm_out = nullptr;
  sink.B = a.B;
  sink.b = b;
  sink.g_bpm = a.g_bpm;
  sink.gb = 0.f;
  DF energy(a.energy_in[b], 0.f);
  DualParamView P{a.params, a.col_offset, a.col_stride, b, dir < a.n_cols ? dir : -1};
  walk_program<kOpsAll>(a.ops, a.n_ops, a.final_stage, P, energy, sink);
  float g = sink.gb;
  if (a.g_mu) {
#pragma unroll
    for (int i = 0; i < 7; ++i) g = fmaf(__ldg(a.g_mu + b * 7 + i), sink.mu[i].d, g);
  }
  if (a.g_cov) {
#pragma unroll
    for (int i = 0; i < 7; ++i)
#pragma unroll
      for (int j = 0; j < 7; ++j) g = fmaf(__ldg(a.g_cov + b * 49 + i * 7 + j), sink.cov[i][j].d, g);
  }
  if (dir < a.n_cols) {
    float* dst = a.g_params + a.col_offset[dir];
    if (a.col_stride[dir]) dst[b] = g; else atomicAdd(dst, g);
  } else if (beam_dir < 7) {
    if (a.g_mu_in) {
      if (a.beam_stride) a.g_mu_in[b * 7 + beam_dir] = g; else atomicAdd(a.g_mu_in + beam_dir, g);
    }
  } else if (a.g_cov_in) {
    if (a.beam_stride) a.g_cov_in[b * 49 + beam_dir - 7] = g; else atomicAdd(a.g_cov_in + beam_dir - 7, g);
  }
}

// ------------------------------------------------------------------ forward moment collapse
// One beam shared by B settings, no outgoing particles, no image, no active cavity: the
// outgoing moment sums are the incoming ones pushed through each setting's total map,
//   S1' = T S1,  S2' = T S2 T^T  (homogeneous 7x7),  bpm sums = (P_b S1)_{x,y},
// so the beam is streamed ONCE (beam_moments_kernel), not once per setting.
struct FwdCollapseArgs {
  const float* arec;
  AdjLayout L;
  const double* in_sums;  // [28] of the shared beam
  const PlanItem* plan;
  int n_plan;
  double* mom;            // [B][28] nullable
  double* bpm_sums;       // [n_slots][B][2] nullable
  const float* energy_in;
  float* energy_out;
  int64_t B;
};

__global__ void __launch_bounds__(128) fwd_collapse_kernel(FwdCollapseArgs a) {
  const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= a.B) return;
  const float* rec = a.arec + b * a.L.rec_floats;
  const double* S = a.in_sums;
  const double n = S[0];
  double S1[7], S2[7][7];
  for (int i = 0; i < 6; ++i) {
    S1[i] = S[1 + i];
    S2[i][6] = S2[6][i] = S[1 + i];
    for (int j = i; j < 6; ++j) S2[i][j] = S2[j][i] = S[7 + tri(i, j)];
  }
  S1[6] = n;
  S2[6][6] = n;
  if (a.mom) {
    double* o = a.mom + b * 28;
    double TS[6][7];
    for (int l = 0; l < 6; ++l)
      for (int j = 0; j < 7; ++j) {
        double s = 0.0;
        for (int m = 0; m < 7; ++m) s += (double)rec[l * 7 + m] * S2[m][j];
        TS[l][j] = s;
      }
    o[0] = n;
    for (int i = 0; i < 6; ++i) o[1 + i] = TS[i][6];
    for (int i = 0; i < 6; ++i)
      for (int j = i; j < 6; ++j) {
        double s = 0.0;
        for (int m = 0; m < 7; ++m) s += TS[i][m] * (double)rec[j * 7 + m];
        o[7 + tri(i, j)] = s;
      }
  }
  if (a.bpm_sums) {
    for (int it = 0; it < a.n_plan; ++it) {
      const PlanItem p = a.plan[it];
      if (p.kind != P_BPM) continue;
      const float* r = rec + a.L.off_bpm + p.c * kAdjBpm;
      double sx = 0.0, sy = 0.0;
      for (int j = 0; j < 7; ++j) {
        sx += (double)r[j] * S1[j];
        sy += (double)r[7 + j] * S1[j];
      }
      a.bpm_sums[((int64_t)p.a * a.B + b) * 2 + 0] = sx;
      a.bpm_sums[((int64_t)p.a * a.B + b) * 2 + 1] = sy;
    }
  }
  if (a.energy_out) a.energy_out[b] = a.energy_in[b];
}

static size_t align256_(size_t x) { return (x + 255) & ~(size_t)255; }

size_t fwd_collapse_workspace_bytes(const lynx_program_t* p, int64_t B) {
  const AdjLayout L = adj_layout(p->n_cav, p->n_snap);
  return align256_(sizeof(float) * (size_t)B * L.rec_floats) + 256;
}

int fwd_collapse_launch(const lynx_program_t* p, const float* d_params, int64_t B,
                        const float* d_in_soa, int64_t ld_in, int64_t N, double* d_bpm_sums,
                        double* d_moment_sums, const float* d_energy_in, float* d_energy_out,
                        void* d_workspace, cudaStream_t st) {
  const AdjLayout L = adj_layout(p->n_cav, p->n_snap);
  float* arec = reinterpret_cast<float*>(d_workspace);
  double* in_sums = reinterpret_cast<double*>((char*)d_workspace +
                                              align256_(sizeof(float) * (size_t)B * L.rec_floats));
  AdjComposeArgs ca;
  ca.ops = p->d_ops;
  ca.n_ops = (int)p->dev_ops.size();
  ca.final_stage = p->final_stage;
  ca.col_offset = p->d_col_offset;
  ca.col_stride = p->d_col_stride;
  ca.params = d_params;
  ca.B = B;
  ca.energy_in = d_energy_in;
  ca.seed = nullptr;
  ca.seed_row_stride = 0;
  ca.seed_ld = 0;
  ca.arec = arec;
  ca.L = L;
  ca.g_bpm = nullptr;
  ca.g_mom = nullptr;
  adj_compose_kernel<<<(unsigned)((B + 127) / 128), 128, 0, st>>>(ca);
  count_launch();
  LYNX_CUDA_OK(cudaMemsetAsync(in_sums, 0, sizeof(double) * 28, st));
  const int64_t per = (int64_t)kMomTPB * 4;
  const int64_t gx = std::min<int64_t>((N + per - 1) / per, 4LL * p->sm_count);
  {
    ProfileScope ps(1, st);
    beam_moments_kernel<<<dim3((unsigned)gx, 1), kMomTPB, 0, st>>>(d_in_soa, 0, ld_in, N, in_sums);
  }
  count_launch();
  FwdCollapseArgs fa;
  fa.arec = arec;
  fa.L = L;
  fa.in_sums = in_sums;
  fa.plan = p->d_plan;
  fa.n_plan = (int)p->plan.size();
  fa.mom = d_moment_sums;
  fa.bpm_sums = d_bpm_sums;
  fa.energy_in = d_energy_in;
  fa.energy_out = d_energy_out;
  fa.B = B;
  fwd_collapse_kernel<<<(unsigned)((B + 127) / 128), 128, 0, st>>>(fa);
  count_launch();
  LYNX_CUDA_OK(cudaGetLastError());
  return LYNX_OK;
}

}  // namespace lynx

using namespace lynx;

// ---------------------------------------------------------------------------- C ABI
static int64_t param_extent(const lynx_program_t* p, int64_t B) {
  int64_t ext = 0;
  for (size_t c = 0; c < p->col_offset.size(); ++c)
    ext = std::max<int64_t>(ext, p->col_offset[c] + (p->col_stride[c] ? B : 1));
  return ext;
}

static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

struct BwdWorkspace {
  float* arec;
  double* sums;
  double* in_sums;
  size_t total;
};

static BwdWorkspace bwd_workspace(const lynx_program_t* p, int64_t B, void* base) {
  const AdjLayout L = adj_layout(p->n_cav, p->n_snap);
  BwdWorkspace w;
  size_t o = 0;
  w.arec = reinterpret_cast<float*>((char*)base + o);
  o += align256(sizeof(float) * (size_t)B * L.rec_floats);
  w.sums = reinterpret_cast<double*>((char*)base + o);
  o += align256(sizeof(double) * (size_t)B * (L.n_seg + L.n_cav) * kAdjSums);
  w.in_sums = reinterpret_cast<double*>((char*)base + o);
  o += align256(sizeof(double) * (size_t)B * 28);
  w.total = o;
  return w;
}

extern "C" size_t lynx_bwd_workspace_bytes(const lynx_program_t* p, int64_t B) {
  if (!p || B <= 0) return 0;
  return bwd_workspace(p, B, nullptr).total;
}

extern "C" int lynx_track_particles_bwd(
    const lynx_program_t* p, const float* d_params, int64_t B, const float* d_in_soa,
    int64_t in_row_stride, int64_t ld_in, int64_t N, const float* d_g_out_soa, int64_t ld_g,
    const double* d_g_moment_sums, const double* d_g_bpm_sums, const float* d_energy_in,
    float* d_g_params, float* d_g_in_soa, int64_t ld_gin, void* d_workspace,
    lynx_stream_t stream) {
  if (!p || !d_params || !d_in_soa || !d_energy_in || !d_g_params || !d_workspace || B <= 0 ||
      N <= 0)
    return LYNX_E_BADARG;
  if ((ld_in & 3) || ld_in < N || ((uintptr_t)d_in_soa & 15)) return LYNX_E_BADARG;
  if (in_row_stride != 0 && in_row_stride != 1) return LYNX_E_BADARG;
  if (d_g_out_soa && ((ld_g & 3) || ld_g < N || ((uintptr_t)d_g_out_soa & 15))) return LYNX_E_BADARG;
  if (d_g_in_soa && ((ld_gin & 3) || ld_gin < N || ((uintptr_t)d_g_in_soa & 15))) return LYNX_E_BADARG;
  if (d_g_out_soa && p->terminal) return LYNX_E_BADARG;
  if (B > 65535) return LYNX_E_BADARG;
  cudaStream_t st = (cudaStream_t)stream;
  const AdjLayout L = adj_layout(p->n_cav, p->n_snap);
  if (L.n_seg > kMaxSeg) return LYNX_E_BADPROGRAM;
  const BwdWorkspace w = bwd_workspace(p, B, d_workspace);
  const int n_t = L.n_seg + L.n_cav;
  const int n_cols = (int)p->col_offset.size();
  const bool shared_beam = in_row_stride == 0;
  const int64_t Bb = shared_beam ? 1 : B;

  const int64_t ext = param_extent(p, B);
  if (ext > 0) LYNX_CUDA_OK(cudaMemsetAsync(d_g_params, 0, sizeof(float) * ext, st));
  LYNX_CUDA_OK(cudaMemsetAsync(w.sums, 0, sizeof(double) * (size_t)B * n_t * kAdjSums, st));

  // ---- per-setting segment maps, injections, cotangent constants
  AdjComposeArgs ca;
  ca.ops = p->d_ops;
  ca.n_ops = (int)p->dev_ops.size();
  ca.final_stage = p->final_stage;
  ca.col_offset = p->d_col_offset;
  ca.col_stride = p->d_col_stride;
  ca.params = d_params;
  ca.B = B;
  ca.energy_in = d_energy_in;
  ca.seed = d_in_soa;
  ca.seed_row_stride = in_row_stride * 6 * ld_in;
  ca.seed_ld = ld_in;
  ca.arec = w.arec;
  ca.L = L;
  ca.g_bpm = d_g_bpm_sums;
  ca.g_mom = d_g_moment_sums;
  adj_compose_kernel<<<(unsigned)((B + 127) / 128), 128, 0, st>>>(ca);
  count_launch();
  LYNX_CUDA_OK(cudaGetLastError());

  const bool collapse = p->n_cav == 0 && !d_g_out_soa && !d_g_in_soa;
  bool no_collapse = false;
#ifdef LYNX_EXPERIMENTS
  static const bool env_nc = [] { const char* e = getenv("LYNX_NO_COLLAPSE"); return e && e[0] == '1'; }();
  no_collapse = env_nc;
#endif
  if (collapse && !no_collapse) {
    // one pass per beam row (not per setting): raw moment sums of the incoming beam
    LYNX_CUDA_OK(cudaMemsetAsync(w.in_sums, 0, sizeof(double) * 28 * Bb, st));
    const int64_t per = (int64_t)kMomTPB * 4;
    int64_t gx = std::min<int64_t>((N + per - 1) / per, std::max<int64_t>(1, 4LL * p->sm_count / Bb));
    beam_moments_kernel<<<dim3((unsigned)gx, (unsigned)Bb), kMomTPB, 0, st>>>(
        d_in_soa, in_row_stride * 6 * ld_in, ld_in, N, w.in_sums);
    count_launch();
    AdjCollapseArgs cc;
    cc.arec = w.arec;
    cc.L = L;
    cc.in_sums = w.in_sums;
    cc.in_sums_stride = shared_beam ? 0 : 28;
    cc.g_mom = d_g_moment_sums;
    cc.sums = w.sums;
    cc.B = B;
    adj_collapse_kernel<<<(unsigned)((B + 127) / 128), 128, 0, st>>>(cc);
    count_launch();
    LYNX_CUDA_OK(cudaGetLastError());
  } else {
    AdjStreamArgs sa;
    sa.in = d_in_soa;
    sa.in_row_stride = in_row_stride * 6 * ld_in;
    sa.ld_in = ld_in;
    sa.N = N;
    sa.g_out = d_g_out_soa;
    sa.ld_g = ld_g;
    sa.g_in = d_g_in_soa;
    sa.gin_row_stride = shared_beam ? 0 : 6 * ld_gin;
    sa.ld_gin = ld_gin;
    sa.atomic_gin = shared_beam && B > 1;
    sa.arec = w.arec;
    sa.L = L;
    sa.sums = w.sums;
    sa.has_mom = d_g_moment_sums != nullptr;
    sa.has_bpm = d_g_bpm_sums != nullptr && p->n_snap > 0;
    sa.B = B;
    if (d_g_in_soa && sa.atomic_gin)
      LYNX_CUDA_OK(cudaMemsetAsync(d_g_in_soa, 0, sizeof(float) * 6 * ld_gin, st));
    const int64_t per = (int64_t)kAdjTPB * 4;
    int64_t gx = std::min<int64_t>((N + per - 1) / per, std::max<int64_t>(1, (4LL * p->sm_count + B - 1) / B));
    const size_t smem = sizeof(float) * L.rec_floats;
    if (smem > 40 * 1024) return LYNX_E_BADPROGRAM;
    for (int t = 0; t < n_t; ++t) {
      sa.target = t;
      if (t > 0) sa.g_in = nullptr;  // written by the first pass only
      ProfileScope ps(3, st);
      bool no_pipe = false;
#ifdef LYNX_EXPERIMENTS
      static const bool env_np = [] { const char* e = getenv("LYNX_K3"); return e && e[0] == 'l'; }();
      no_pipe = env_np;
#endif
      cudaError_t pe = (no_pipe || B > 65535) ? cudaErrorNotSupported
                                              : adj_pipe_dispatch(sa, p->n_cav == 0, p->sm_count, st);
      if (pe == cudaErrorNotSupported) {  // grid-stride LDG variant
        if (p->n_cav == 0)
          adj_stream_kernel<true><<<dim3((unsigned)gx, (unsigned)B), kAdjTPB, smem, st>>>(sa);
        else
          adj_stream_kernel<false><<<dim3((unsigned)gx, (unsigned)B), kAdjTPB, smem, st>>>(sa);
      } else if (pe != cudaSuccess) {
        return LYNX_E_LAUNCH;
      }
      count_launch();
    }
    LYNX_CUDA_OK(cudaGetLastError());
  }

  if (n_cols > 0) {
    AdjParamArgs pa;
    pa.ops = p->d_ops;
    pa.n_ops = (int)p->dev_ops.size();
    pa.final_stage = p->final_stage;
    pa.n_cols = n_cols;
    pa.col_offset = p->d_col_offset;
    pa.col_stride = p->d_col_stride;
    pa.params = d_params;
    pa.B = B;
    pa.energy_in = d_energy_in;
    pa.arec = w.arec;
    pa.L = L;
    pa.sums = w.sums;
    pa.g_bpm = (d_g_bpm_sums && p->n_snap > 0) ? d_g_bpm_sums : nullptr;
    pa.n_particles = (double)N;
    pa.g_params = d_g_params;
    const int64_t total = B * n_cols;
    adj_param_kernel<<<(unsigned)((total + 127) / 128), 128, 0, st>>>(pa);
    count_launch();
    LYNX_CUDA_OK(cudaGetLastError());
  }
  return LYNX_OK;
}

extern "C" int lynx_parambeam_bwd(const lynx_program_t* p, const float* d_params, int64_t B,
                                  const float* d_mu_in, const float* d_cov_in,
                                  int64_t beam_stride, const float* d_energy_in,
                                  const float* d_g_mu, const float* d_g_cov,
                                  const float* d_g_bpm, float* d_g_params, float* d_g_mu_in,
                                  float* d_g_cov_in, lynx_stream_t stream) {
  if (!p || !d_params || !d_mu_in || !d_cov_in || !d_energy_in || !d_g_params || B <= 0 ||
      (beam_stride != 0 && beam_stride != 1))
    return LYNX_E_BADARG;
  cudaStream_t st = (cudaStream_t)stream;
  const int n_cols = (int)p->col_offset.size();
  const bool beam_grads = d_g_mu_in || d_g_cov_in;
  const int64_t ext = param_extent(p, B);
  if (ext > 0) LYNX_CUDA_OK(cudaMemsetAsync(d_g_params, 0, sizeof(float) * ext, st));
  if (beam_stride == 0) {
    if (d_g_mu_in) LYNX_CUDA_OK(cudaMemsetAsync(d_g_mu_in, 0, sizeof(float) * 7, st));
    if (d_g_cov_in) LYNX_CUDA_OK(cudaMemsetAsync(d_g_cov_in, 0, sizeof(float) * 49, st));
  }
  PBBwdArgs a;
  a.ops = p->d_ops;
  a.n_ops = (int)p->dev_ops.size();
  a.final_stage = p->final_stage;
  a.n_cols = n_cols;
  a.n_dirs = n_cols + (beam_grads ? 56 : 0);
  a.col_offset = p->d_col_offset;
  a.col_stride = p->d_col_stride;
  a.params = d_params;
  a.B = B;
  a.mu_in = d_mu_in;
  a.cov_in = d_cov_in;
  a.beam_stride = beam_stride;
  a.energy_in = d_energy_in;
  a.g_mu = d_g_mu;
  a.g_cov = d_g_cov;
  a.g_bpm = (p->n_bpm_slots > 0) ? d_g_bpm : nullptr;
  a.g_params = d_g_params;
  a.g_mu_in = d_g_mu_in;
  a.g_cov_in = d_g_cov_in;
  const int64_t total = B * a.n_dirs;
  if (total > 0) {
    ProfileScope ps(3, st);
    parambeam_bwd_kernel<<<(unsigned)((total + 63) / 64), 64, 0, st>>>(a);
    count_launch();
  }
  LYNX_CUDA_OK(cudaGetLastError());
  return LYNX_OK;
}

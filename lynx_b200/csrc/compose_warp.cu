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
#include <cstdio>
#include <cstdlib>

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
  // optional buffer zeroed by extra blocks of the same launch
  float* zero_ptr;
  int64_t zero_vec4;
  int compose_blocks;
  int dbg;
};

// Stage map finished: M (columns in lanes 0..6, c[6]) -> shared (row-major 6x7) and,
// depending on the mode, the record / the moments.
template <bool K6>
__device__ __forceinline__ void cw_stage(const CWArgs& a, const WarpScratch& w, const float (&c)[6],
                                         int lane, int stage, int64_t b) {
  if (lane < 7) {
#pragma unroll
    for (int i = 0; i < 6; ++i) w.M[i * 7 + lane] = c[i];
  }
  __syncwarp();
  if (K6) {
    warp_apply_moments(w, lane);
  } else {
    float* rec = a.records + b * a.rec_floats + stage * kMapStride;
    for (int i = lane; i < kMapFloats; i += 32) rec[i] = w.M[i];
    if (a.maps_out) {
      float* o = a.maps_out + (b * a.n_stages + stage) * 49;
      for (int i = lane; i < 49; i += 32) o[i] = (i < 42) ? w.M[i] : (i == 48 ? 1.f : 0.f);
    }
    // moment-shift centre: ctr <- M ctr
    float n = 0.f;
    if (lane < 6) {
      n = w.M[lane * 7 + 6];
#pragma unroll
      for (int j = 0; j < 6; ++j) n = fmaf(w.M[lane * 7 + j], w.ctr[j], n);
    }
    __syncwarp();
    if (lane < 6) w.ctr[lane] = n;
    __syncwarp();
  }
}

template <bool K6, int OPS>
__global__ void __launch_bounds__(kCWThreads) compose_warp_kernel(CWArgs a) {
  extern __shared__ __align__(16) float cw_smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  long long t0 = clock64(), t1 = 0, t2 = 0, t3 = 0;
  if ((int)blockIdx.x >= a.compose_blocks) {  // spare blocks: zero the caller's buffer
    const int zb = blockIdx.x - a.compose_blocks, nzb = gridDim.x - a.compose_blocks;
    float4* z = reinterpret_cast<float4*>(a.zero_ptr);
    const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int64_t i = (int64_t)zb * kCWThreads + threadIdx.x; i < a.zero_vec4;
         i += (int64_t)nzb * kCWThreads)
      z[i] = zero;
    return;
  }
  const int64_t b = (int64_t)blockIdx.x * kCWWarps + warp;
  DevOp* s_ops = reinterpret_cast<DevOp*>(cw_smem);
  float* base = cw_smem + a.n_ops * (sizeof(DevOp) / 4);
  float* s_vals = base + warp * (kCWScratch + a.n_cols);  // this warp's parameter columns
  const WarpScratch w = warp_scratch(s_vals + a.n_cols);

  // ---- stage ops (CTA) and this setting's parameter columns (warp): one round of loads
  for (int i = threadIdx.x; i < a.n_ops * (int)(sizeof(DevOp) / 4); i += kCWThreads)
    cw_smem[i] = reinterpret_cast<const float*>(a.ops)[i];
  const bool live = b < a.B;
  if (live) {
    for (int c = lane; c < a.n_cols; c += 32)
      s_vals[c] = __ldg(a.params + __ldg(a.col_offset + c) + b * (int64_t)__ldg(a.col_stride + c));
    if (K6) {
      const float* mi = a.mu_in + b * a.beam_stride * 7;
      const float* ci = a.cov_in + b * a.beam_stride * 49;
      if (lane < 7) w.mu[lane] = __ldg(mi + lane);
      for (int i = lane; i < 49; i += 32) w.cov[i] = __ldg(ci + i);
    } else if (lane < 6) {
      w.ctr[lane] = a.seed ? __ldg(a.seed + b * a.seed_row_stride + lane * a.seed_ld) : 0.f;
    }
  }
  __syncthreads();
  if (!live) return;
  t1 = clock64();

  SmemParamView P{s_vals, 1, 0};
  float energy = a.energy_in[b];  // energy at the start of the current chunk
  float c[6];                     // lane j < 7: column j of the accumulated map
#pragma unroll
  for (int i = 0; i < 6; ++i) c[i] = (i == lane) ? 1.f : 0.f;

  for (int e0 = 0; e0 < a.n_ops; e0 += 32) {
    const int ne = min(32, a.n_ops - e0);
    // ---- phase B: lane -> element map, in parallel
    DevOp op;
    op.opcode = LYNX_OP_MARKER;
    op.flags = 0;
    op.col = -1;
    if (lane < ne) op = s_ops[e0 + lane];
    const bool active_cav = op.opcode == LYNX_OP_CAVITY && (op.flags & LYNX_F_ACTIVE);
    float dE = 0.f;
    if (active_cav)
      dE = (float)((double)P(op.col + 1) * cos((double)P(op.col + 2) * (kPi / 180.0)));
    float incl = dE;  // inclusive prefix sum of energy gains over the chunk
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const float v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    const float e_before = energy + (incl - dE);
    if (lane < ne) {
      Map R;
      map_identity(R);
      const Rel rel = make_rel(e_before);
      if (active_cav) {
        if constexpr ((OPS & op_bit(LYNX_OP_CAVITY)) != 0) {
          CavityOut co;
          cavity_setting(P(op.col), P(op.col + 1), P(op.col + 2), P(op.col + 3), e_before, R, co);
#pragma unroll
          for (int i = 0; i < kCavFloats; ++i) w.cav[lane * kCavFloats + i] = co.sc[i];
        }
      } else {
        lm_linear_op<OPS>(R, op, P, rel, e_before);
      }
      float* d = w.maps + lane * kCWMap;
#pragma unroll
      for (int i = 0; i < 6; ++i)
#pragma unroll
        for (int j = 0; j < 7; ++j) d[i * 7 + j] = R.m[i][j];
    }
    energy += __shfl_sync(0xffffffffu, incl, 31);
    __syncwarp();
    t2 = clock64();

    // ---- phase C: fold the chunk in beam order; lanes 0..6 own one column each
    for (int el = 0; el < ne; ++el) {
      const DevOp o = s_ops[e0 + el];
      if (o.flush_stage >= 0) {
        cw_stage<K6>(a, w, c, lane, o.flush_stage, b);
#pragma unroll
        for (int i = 0; i < 6; ++i) c[i] = (i == lane) ? 1.f : 0.f;
      }
      const float* R = w.maps + el * kCWMap;
      const bool acav = o.opcode == LYNX_OP_CAVITY && (o.flags & LYNX_F_ACTIVE);
      if (acav) {
        // cavity = its own stage (linear map) + non-linear overrides (cavity.py:97-246)
        const float* sc = w.cav + el * kCavFloats;
        float s0 = 0.f, p0 = 0.f, c44 = 0.f, c45 = 0.f, c55 = 0.f;
        if (K6) {
          s0 = w.mu[4];
          p0 = w.mu[5];
          c44 = w.cov[4 * 7 + 4];
          c45 = w.cov[4 * 7 + 5];
          c55 = w.cov[5 * 7 + 5];
        } else {
          s0 = w.ctr[4];
          p0 = w.ctr[5];
          float* rec = a.records + b * a.rec_floats + a.off_cav + o.cav_idx * kCavFloats;
          if (lane < kCavFloats) rec[lane] = sc[lane];
        }
        __syncwarp();
        float cc[6];
#pragma unroll
        for (int i = 0; i < 6; ++i) cc[i] = (lane < 7) ? R[i * 7 + lane] : 0.f;
        cw_stage<K6>(a, w, cc, lane, o.self_stage, b);
        if (lane == 0) {
          float* v = K6 ? w.mu : w.ctr;
          float s1, p1;
          cavity_nonlinear(sc, s0, p0, v[4], v[5], s1, p1);
          v[4] = s1;
          v[5] = p1;
          if (K6 && sc[CV_ON] != 0.f) {  // overwrites as written (quirk B5)
            w.cov[5 * 7 + 5] = c55;
            const float t = sc[CV_T566] * c55 * c55 + sc[CV_T556] * c45 * c55 + sc[CV_T555] * c44 * c44;
            w.cov[4 * 7 + 4] = t;
            w.cov[4 * 7 + 5] = t;
            w.cov[5 * 7 + 4] = t;
          }
        }
        __syncwarp();
      } else if (o.opcode == LYNX_OP_BPM) {
        if (K6) {
          if (lane == 0 && a.bpm_out) {
            a.bpm_out[(o.aux * a.B + b) * 2 + 0] = w.mu[0];
            a.bpm_out[(o.aux * a.B + b) * 2 + 1] = w.mu[2];
          }
        } else if (lane < 6) {
          a.records[b * a.rec_floats + a.off_centre + o.snap_idx * 6 + lane] = w.ctr[lane];
        }
      } else if (o.opcode == LYNX_OP_SCREEN) {
        const float mx = P(o.col), my = P(o.col + 1);
        if (K6) {
          if (lane == 0) {
            w.mu[0] -= mx;
            w.mu[2] -= my;
          }
          __syncwarp();
        } else if (lane == 0) {
          a.records[b * a.rec_floats + a.off_screen + o.snap_idx * 2 + 0] = mx;
          a.records[b * a.rec_floats + a.off_screen + o.snap_idx * 2 + 1] = my;
        }
      } else if (o.opcode != LYNX_OP_MARKER) {
        // column <- R_e column  (+ constant column for lane 6)
        float n[6];
#pragma unroll
        for (int i = 0; i < 6; ++i) {
          float acc = (lane == 6) ? R[i * 7 + 6] : 0.f;
#pragma unroll
          for (int k = 0; k < 6; ++k) acc = fmaf(R[i * 7 + k], c[k], acc);
          n[i] = acc;
        }
#pragma unroll
        for (int i = 0; i < 6; ++i) c[i] = n[i];
      }
    }
    __syncwarp();
  }
  if (a.final_stage >= 0) cw_stage<K6>(a, w, c, lane, a.final_stage, b);

  if (K6) {
    if (lane < 7) a.mu_out[b * 7 + lane] = w.mu[lane];
    for (int i = lane; i < 49; i += 32) a.cov_out[b * 49 + i] = w.cov[i];
  } else if (lane < 6) {
    a.records[b * a.rec_floats + a.final_centre + lane] = w.ctr[lane];
  }
  if (lane == 0 && a.energy_out) a.energy_out[b] = energy;
  t3 = clock64();
  if ((a.dbg & 2) && b == 0 && lane == 0)
    printf("K1 warp clocks: setup %lld phaseB %lld phaseC+out %lld\n", t1 - t0, t2 - t1, t3 - t2);
}

constexpr int64_t kWarpKernelMaxB = 8192;

template <bool K6>
static int cw_launch(const lynx_program_t* p, CWArgs& a, int which, cudaStream_t st) {
  if (a.B > kWarpKernelMaxB) return 1;
  a.n_cols = (int)p->col_offset.size();
  const size_t smem = a.n_ops * sizeof(DevOp) + sizeof(float) * kCWWarps * (size_t)(kCWScratch + a.n_cols);
  if (smem > 48 * 1024) return 1;
  static const int dbg = [] { const char* e = getenv("LYNX_DBG"); return e ? atoi(e) : 0; }();
  a.dbg = dbg;
  a.compose_blocks = (int)((a.B + kCWWarps - 1) / kCWWarps);
  unsigned blocks = (unsigned)a.compose_blocks;
  if (a.zero_vec4 > 0)
    blocks += (unsigned)std::min<int64_t>(2 * p->sm_count, (a.zero_vec4 + kCWThreads - 1) / kCWThreads);
  {
    ProfileScope ps(which, st);
    if ((p->op_mask & ~kOpsLight) == 0)
      compose_warp_kernel<K6, kOpsLight><<<blocks, kCWThreads, smem, st>>>(a);
    else
      compose_warp_kernel<K6, kOpsAll><<<blocks, kCWThreads, smem, st>>>(a);
  }
  count_launch();
  return cudaGetLastError() == cudaSuccess ? LYNX_OK : LYNX_E_LAUNCH;
}

int compose_warp_launch(const lynx_program_t* p, const ComposeArgs& c, cudaStream_t st) {
  CWArgs a{};
  a.ops = c.ops;
  a.n_ops = c.n_ops;
  a.final_stage = c.final_stage;
  a.col_offset = c.col_offset;
  a.col_stride = c.col_stride;
  a.params = c.params;
  a.B = c.B;
  a.energy_in = c.energy_in;
  a.energy_out = c.energy_out;
  a.seed = c.seed;
  a.seed_row_stride = c.seed_row_stride;
  a.seed_ld = c.seed_ld;
  a.records = c.records;
  a.rec_floats = c.rec_floats;
  a.off_cav = c.off_cav;
  a.off_centre = c.off_centre;
  a.off_screen = c.off_screen;
  a.final_centre = c.final_centre;
  a.n_stages = c.n_stages;
  a.maps_out = c.maps_out;
  a.zero_ptr = c.zero_ptr;
  a.zero_vec4 = c.zero_vec4;
  return cw_launch<false>(p, a, 0, st);
}

int parambeam_warp_launch(const lynx_program_t* p, const ParamBeamArgs& c, cudaStream_t st) {
  CWArgs a{};
  a.ops = c.ops;
  a.n_ops = c.n_ops;
  a.final_stage = c.final_stage;
  a.col_offset = c.col_offset;
  a.col_stride = c.col_stride;
  a.params = c.params;
  a.B = c.B;
  a.energy_in = c.energy_in;
  a.energy_out = c.energy_out;
  a.mu_in = c.mu_in;
  a.cov_in = c.cov_in;
  a.beam_stride = c.beam_stride;
  a.mu_out = c.mu_out;
  a.cov_out = c.cov_out;
  a.bpm_out = c.bpm_out;
  return cw_launch<true>(p, a, 2, st);
}

}  // namespace lynx

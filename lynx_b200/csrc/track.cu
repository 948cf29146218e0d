// K2: fused particle tracker (forward).  Streams each particle's six coordinates from
// HBM once (SoA planes, 128-bit coalesced loads), applies every merged affine stage and
// non-linear cavity in fp32 registers, snapshots BPM sums, histograms at an active
// screen, accumulates beam moments, and writes once.
//
// Replaces segment.py:340-356 + element.py:83-92 + cavity.py:97-246 + bpm.py:48-58 +
// screen.py:126-141,195-211 + particle_beam.py:736-836 (one pass instead of one
// (B,N,7)x(B,7,7) matmul + full copy per run and ~15 elementwise passes per cavity).
//
// Algorithmic HBM bytes: 24 B/particle read (+24 B written when outgoing particles are
// requested); the per-setting record (<1 KB) is read once per CTA into shared memory.
#include <algorithm>
#include <cstdlib>

#include "walk.cuh"
#include "track_common.cuh"

namespace lynx {

constexpr int kTPB = 256;          // threads per CTA
constexpr int kVec = 4;            // particles per 128-bit access
constexpr int kUnroll = 2;         // 128-bit accesses per plane per thread per tile
constexpr int kNP = kVec * kUnroll;            // particles per thread per tile
constexpr int kTile = kTPB * kNP;              // particles per CTA per tile

template <int MOM, bool HAS_OUT>
__global__ void __launch_bounds__(kTPB, 2) track_fwd_kernel(TrackArgs a) {
  extern __shared__ __align__(16) float smem[];
  float* rec = smem;                                                  // rec_floats
  PlanItem* plan = reinterpret_cast<PlanItem*>(smem + a.rec_floats);  // n_plan
  double* s_bpm = reinterpret_cast<double*>(plan + a.n_plan);         // n_snap*2
  __shared__ double s_red[kTPB / 32][28];
  __shared__ HistAxis s_hx, s_hy;

  const int tid = threadIdx.x;
  const int64_t b = blockIdx.x / (unsigned)a.ctas_per_row;
  const int cta = (int)(blockIdx.x - b * a.ctas_per_row), n_cta = a.ctas_per_row;
  {
    const float* src = a.records + b * a.rec_floats;
    for (int i = tid; i < a.rec_floats; i += kTPB) rec[i] = src[i];
    for (int i = tid; i < a.n_plan; i += kTPB) plan[i] = a.plan[i];
    for (int i = tid; i < a.n_snap * 2; i += kTPB) s_bpm[i] = 0.0;
  }
  if (a.img) {
    hist_axis_setup(a.ex, a.W, &s_hx, tid, kTPB);
    hist_axis_setup(a.ey, a.H, &s_hy, tid, kTPB);
  }
  __syncthreads();

  const float* in_b = a.in + b * a.in_row_stride;
  float* out_b = HAS_OUT ? a.out + b * 6 * a.ld_out : nullptr;
  float* img_b = a.img ? a.img + b * (int64_t)a.W * a.H : nullptr;

  MomAcc<MOM> acc;
#pragma unroll
  for (int i = 0; i < MomAcc<MOM>::kN; ++i) acc.a[i] = 0.f;
  float ctr[6];
#pragma unroll
  for (int i = 0; i < 6; ++i) ctr[i] = rec[a.final_centre + i];

  const int64_t n_tiles = (a.N + kTile - 1) / kTile;
  int64_t n_mine = 0;
  for (int64_t tile = cta; tile < n_tiles; tile += n_cta) {
    const int64_t base = tile * kTile + (int64_t)tid * kVec;
    n_mine += min((int64_t)kTile, a.N - tile * kTile);
    float v[6][kNP];
    bool ok[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int64_t idx = base + (int64_t)u * kTPB * kVec;
      ok[u] = idx < a.N;
#pragma unroll
      for (int k = 0; k < 6; ++k) {
        float4 t = ok[u] ? ld_stream(in_b + k * a.ld_in + idx) : make_float4(0.f, 0.f, 0.f, 0.f);
        v[k][u * 4 + 0] = t.x;
        v[k][u * 4 + 1] = t.y;
        v[k][u * 4 + 2] = t.z;
        v[k][u * 4 + 3] = t.w;
      }
    }

    for (int it = 0; it < a.n_plan; ++it) {
      const PlanItem p = plan[it];
      if (p.kind == P_AFFINE) {
        apply_affine(rec + p.a, v);
      } else if (p.kind == P_CAVITY) {
        float s0[kNP], p0[kNP];
#pragma unroll
        for (int j = 0; j < kNP; ++j) {
          s0[j] = v[4][j];
          p0[j] = v[5][j];
        }
        apply_block2(rec + p.a, v);  // the cavity's own map is block-diagonal
        const float* sc = rec + p.b;
#pragma unroll
        for (int j = 0; j < kNP; ++j) {
          float s1, p1;
          cavity_nonlinear(sc, s0[j], p0[j], v[4][j], v[5][j], s1, p1);
          v[4][j] = s1;
          v[5][j] = p1;
        }
      } else if (p.kind == P_BPM) {
        // bpm.py:54: reading = (mean x, mean y); shifted sums, reduced per warp
        const float cx = rec[p.b], cy = rec[p.b + 2];
        float sx = 0.f, sy = 0.f;
#pragma unroll
        for (int u = 0; u < kUnroll; ++u)
#pragma unroll
          for (int l = 0; l < 4; ++l) {
            const bool live = ok[u] && (base + (int64_t)u * kTPB * kVec + l < a.N);
            sx += live ? v[0][u * 4 + l] - cx : 0.f;
            sy += live ? v[2][u * 4 + l] - cy : 0.f;
          }
        sx = warp_sum(sx);
        sy = warp_sum(sy);
        if ((tid & 31) == 0) {
          atomicAdd(&s_bpm[2 * p.c + 0], (double)sx);
          atomicAdd(&s_bpm[2 * p.c + 1], (double)sy);
        }
      } else {  // P_SCREEN, screen.py:126-141 + 195-211
        const float mx = rec[p.a], my = rec[p.a + 1];
        const bool yfix = (p.c & LYNX_F_SCREEN_Y_FIX) != 0;
#pragma unroll
        for (int u = 0; u < kUnroll; ++u)
#pragma unroll
          for (int l = 0; l < 4; ++l) {
            const bool live = ok[u] && (base + (int64_t)u * kTPB * kVec + l < a.N);
            if (live && img_b) {
              const float x = v[0][u * 4 + l] - mx;
              const float y = yfix ? v[2][u * 4 + l] - my : v[2][u * 4 + l];
              const int ix = find_bin(x, a.ex, a.W, s_hx);
              const int iy = find_bin(y, a.ey, a.H, s_hy);
              if (ix >= 0 && iy >= 0)
                atomicAdd(img_b + (int64_t)(a.H - 1 - iy) * a.W + ix, 1.0f);
            }
          }
      }
    }

    if (HAS_OUT) {
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const int64_t idx = base + (int64_t)u * kTPB * kVec;
        if (ok[u]) {
#pragma unroll
          for (int k = 0; k < 6; ++k)
            st_stream(out_b + k * a.ld_out + idx,
                      make_float4(v[k][u * 4], v[k][u * 4 + 1], v[k][u * 4 + 2], v[k][u * 4 + 3]));
        }
      }
    }

    if (MOM != 0) {
#pragma unroll
      for (int u = 0; u < kUnroll; ++u)
#pragma unroll
        for (int l = 0; l < 4; ++l) {
          const bool live = ok[u] && (base + (int64_t)u * kTPB * kVec + l < a.N);
          float d[6];
#pragma unroll
          for (int k = 0; k < 6; ++k) d[k] = live ? v[k][u * 4 + l] - ctr[k] : 0.f;
#pragma unroll
          for (int k = 0; k < 6; ++k) acc.a[k] += d[k];
          if (MOM == 2) {
#pragma unroll
            for (int i = 0; i < 6; ++i)
#pragma unroll
              for (int j = i; j < 6; ++j)
                acc.a[6 + tri(i, j)] = fmaf(d[i], d[j], acc.a[6 + tri(i, j)]);
          } else {
#pragma unroll
            for (int k = 0; k < 6; ++k) acc.a[6 + k] = fmaf(d[k], d[k], acc.a[6 + k]);
            acc.a[12] = fmaf(d[0], d[1], acc.a[12]);
            acc.a[13] = fmaf(d[2], d[3], acc.a[13]);
          }
        }
    }
  }

  // ---- CTA epilogue: moments -> raw double sums -> global atomics
  if (MOM != 0) {
    constexpr int kN = MomAcc<MOM>::kN;
    const int w = tid >> 5, lane = tid & 31;
#pragma unroll
    for (int i = 0; i < kN; ++i) {
      double s = warp_sum((double)acc.a[i]);
      if (lane == 0) s_red[w][i] = s;
    }
    __syncthreads();
    if (tid == 0) {
      double S1[6], S2[21];
      for (int i = 0; i < 21; ++i) S2[i] = 0.0;
      bool have[21];
      for (int i = 0; i < 21; ++i) have[i] = (MOM == 2);
      for (int i = 0; i < 6; ++i) {
        double s = 0.0;
        for (int ww = 0; ww < kTPB / 32; ++ww) s += s_red[ww][i];
        S1[i] = s;
      }
      if (MOM == 2) {
        for (int i = 0; i < 21; ++i) {
          double s = 0.0;
          for (int ww = 0; ww < kTPB / 32; ++ww) s += s_red[ww][6 + i];
          S2[i] = s;
        }
      } else {
        for (int k = 0; k < 6; ++k) {
          double s = 0.0;
          for (int ww = 0; ww < kTPB / 32; ++ww) s += s_red[ww][6 + k];
          S2[tri(k, k)] = s;
          have[tri(k, k)] = true;
        }
        double s01 = 0.0, s23 = 0.0;
        for (int ww = 0; ww < kTPB / 32; ++ww) {
          s01 += s_red[ww][12];
          s23 += s_red[ww][13];
        }
        S2[tri(0, 1)] = s01;
        have[tri(0, 1)] = true;
        S2[tri(2, 3)] = s23;
        have[tri(2, 3)] = true;
      }
      const double n = (double)n_mine;
      double* g = a.mom + b * 28;
      atomicAdd(g, n);
      double c[6];
      for (int i = 0; i < 6; ++i) c[i] = (double)ctr[i];
      for (int i = 0; i < 6; ++i) atomicAdd(g + 1 + i, S1[i] + n * c[i]);
      for (int i = 0; i < 6; ++i)
        for (int j = i; j < 6; ++j)
          if (have[tri(i, j)])
            atomicAdd(g + 7 + tri(i, j),
                      S2[tri(i, j)] + c[i] * S1[j] + c[j] * S1[i] + n * c[i] * c[j]);
    }
  }
  // ---- BPM sums
  if (a.n_snap > 0) {
    __syncthreads();
    for (int it = tid; it < a.n_plan; it += kTPB) {
      const PlanItem p = plan[it];
      if (p.kind == P_BPM) {
        const double n = (double)n_mine;
        double* g = a.bpm_sums + ((int64_t)p.a * a.B + b) * 2;
        atomicAdd(g, s_bpm[2 * p.c] + n * (double)rec[p.b]);
        atomicAdd(g + 1, s_bpm[2 * p.c + 1] + n * (double)rec[p.b + 2]);
      }
    }
  }
}

template <int MOM, bool HAS_OUT>
static int launch_track(const TrackArgs& a, unsigned grid, size_t smem, cudaStream_t st) {
  auto k = track_fwd_kernel<MOM, HAS_OUT>;
  if (smem > 48 * 1024) {
    if (cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) !=
        cudaSuccess)
      return LYNX_E_BADPROGRAM;
  }
  {
    ProfileScope ps(1, st);
    k<<<grid, kTPB, smem, st>>>(a);
  }
  count_launch();
  return cudaGetLastError() == cudaSuccess ? LYNX_OK : LYNX_E_LAUNCH;
}

}  // namespace lynx

using namespace lynx;

extern "C" size_t lynx_workspace_bytes(const lynx_program_t* p, int64_t B) {
  if (!p || B <= 0) return 0;
  size_t bytes = (size_t)B * p->rec_floats * sizeof(float);
  bytes = std::max(bytes, fwd_collapse_workspace_bytes(p, B));
  return ((bytes + 255) & ~(size_t)255) + 256;  // + HistAxis[2] written by K1
}

extern "C" int lynx_track_particles_fwd(
    const lynx_program_t* p, const float* d_params, int64_t B, const float* d_in_soa,
    int64_t in_row_stride, int64_t ld_in, int64_t N, float* d_out_soa, int64_t ld_out,
    double* d_bpm_sums, float* d_screen_img, int32_t W, int32_t H, const float* d_edges_x,
    const float* d_edges_y, double* d_moment_sums, int32_t moment_mode,
    const float* d_energy_in, float* d_energy_out, void* d_workspace, lynx_stream_t stream) {
  return lynx_track_particles_fwd_bbox(p, d_params, B, d_in_soa, in_row_stride, ld_in, N, d_out_soa,
                                       ld_out, d_bpm_sums, d_screen_img, W, H, d_edges_x, d_edges_y,
                                       d_moment_sums, moment_mode, d_energy_in, d_energy_out,
                                       d_workspace, nullptr, stream);
}

using lynx::track_fwd_impl;

// does this launch take the collapsed path (one beam, B > 1 settings, additive outputs only)?
bool lynx::fwd_collapses(const lynx_program_t* p, int64_t B, int64_t in_row_stride, bool has_out,
                         int moment_mode) {
  return in_row_stride == 0 && B > 1 && !has_out && p->n_screen == 0 && p->n_cav == 0 &&
         (moment_mode != 0 || p->n_bpm_slots > 0);
}

extern "C" int lynx_track_particles_fwd_bbox(
    const lynx_program_t* p, const float* d_params, int64_t B, const float* d_in_soa,
    int64_t in_row_stride, int64_t ld_in, int64_t N, float* d_out_soa, int64_t ld_out,
    double* d_bpm_sums, float* d_screen_img, int32_t W, int32_t H, const float* d_edges_x,
    const float* d_edges_y, double* d_moment_sums, int32_t moment_mode,
    const float* d_energy_in, float* d_energy_out, void* d_workspace, int32_t* d_screen_bbox,
    lynx_stream_t stream) {
  return track_fwd_impl(p, d_params, B, d_in_soa, in_row_stride, ld_in, N, d_out_soa, ld_out,
                        d_bpm_sums, d_screen_img, W, H, d_edges_x, d_edges_y, d_moment_sums,
                        moment_mode, d_energy_in, d_energy_out, d_workspace, d_screen_bbox, nullptr,
                        0u, stream, false);
}

extern "C" int lynx_track_particles_fwd_persistent(
    const lynx_program_t* p, const float* d_params, int64_t B, const float* d_in_soa,
    int64_t in_row_stride, int64_t ld_in, int64_t N, float* d_out_soa, int64_t ld_out,
    double* d_bpm_sums, float* d_screen_img, int32_t W, int32_t H, const float* d_edges_x,
    const float* d_edges_y, double* d_moment_sums, int32_t moment_mode,
    const float* d_energy_in, float* d_energy_out, void* d_workspace, int32_t* d_img_state,
    uint32_t epoch, lynx_stream_t stream) {
  if (!d_img_state || !d_screen_img || !p || p->n_screen == 0) return LYNX_E_BADARG;
  return track_fwd_impl(p, d_params, B, d_in_soa, in_row_stride, ld_in, N, d_out_soa, ld_out,
                        d_bpm_sums, d_screen_img, W, H, d_edges_x, d_edges_y, d_moment_sums,
                        moment_mode, d_energy_in, d_energy_out, d_workspace, nullptr, d_img_state,
                        epoch, stream, false);
}

// d_img_state != nullptr: persistent screen image (lynx_track_particles_fwd_persistent) -- the
// image is NOT zeroed as a whole; K1's extra blocks clear the rectangle the previous call lit
// (bounding-box cell (epoch + 1) & 1 of the state), K2 records this call's box in cell epoch & 1.
int lynx::track_fwd_impl(
    const lynx_program_t* p, const float* d_params, int64_t B, const float* d_in_soa,
    int64_t in_row_stride, int64_t ld_in, int64_t N, float* d_out_soa, int64_t ld_out,
    double* d_bpm_sums, float* d_screen_img, int32_t W, int32_t H, const float* d_edges_x,
    const float* d_edges_y, double* d_moment_sums, int32_t moment_mode,
    const float* d_energy_in, float* d_energy_out, void* d_workspace, int32_t* d_screen_bbox,
    int32_t* d_img_state, uint32_t epoch, lynx_stream_t stream, bool accumulate) {
  if (!p || !d_params || !d_in_soa || !d_energy_in || !d_workspace || B <= 0 || N <= 0)
    return LYNX_E_BADARG;
  if ((ld_in & 3) || ld_in < N || ((uintptr_t)d_in_soa & 15)) return LYNX_E_BADARG;
  if (d_out_soa && ((ld_out & 3) || ld_out < N || ((uintptr_t)d_out_soa & 15)))
    return LYNX_E_BADARG;
  if (in_row_stride != 0 && in_row_stride != 1) return LYNX_E_BADARG;
  if (p->n_bpm_slots > 0 && !d_bpm_sums) return LYNX_E_BADARG;
  if (d_screen_img && (!d_edges_x || !d_edges_y || W <= 0 || H <= 0)) return LYNX_E_BADARG;
  if (moment_mode < 0 || moment_mode > 2 || (moment_mode && !d_moment_sums))
    return LYNX_E_BADARG;
  if (p->terminal && d_out_soa) return LYNX_E_BADARG;  // Beam.empty after an active screen
  cudaStream_t st = (cudaStream_t)stream;

  // one beam, B > 1 settings, only additive moment / BPM outputs, no cavity: stream the
  // beam once and push its moment sums through every setting's map (adjoint.cu)
  bool no_collapse = false;
#ifdef LYNX_EXPERIMENTS
  static const bool env_nc = [] { const char* e = getenv("LYNX_NO_COLLAPSE"); return e && e[0] == '1'; }();
  no_collapse = env_nc;
#endif
  if (accumulate && fwd_collapses(p, B, in_row_stride, d_out_soa != nullptr, moment_mode))
    return LYNX_E_BADARG;  // the collapsed path has no continuation (callers ask fwd_collapses first)
  if (!no_collapse && fwd_collapses(p, B, in_row_stride, d_out_soa != nullptr, moment_mode)) {
    return fwd_collapse_launch(p, d_params, B, d_in_soa, ld_in, N, d_bpm_sums,
                               moment_mode ? d_moment_sums : nullptr, d_energy_in, d_energy_out,
                               d_workspace, st);
  }
  float* records = reinterpret_cast<float*>(d_workspace);
  void* hist_ws = (char*)d_workspace + lynx_workspace_bytes(p, B) - 256;
  // the screen image is zeroed by spare blocks of the K1 launch when its size allows
  const bool has_img = d_screen_img && p->n_screen > 0;
  const size_t img_bytes = has_img ? sizeof(float) * (size_t)W * H * B : 0;
  const bool img_aligned = has_img && !(img_bytes & 15) && !((uintptr_t)d_screen_img & 15);
  const bool persistent = d_img_state != nullptr && has_img;
  if (persistent && (!img_aligned || (W & 3) || d_screen_bbox)) return LYNX_E_BADARG;
  int32_t* cell_cur = persistent ? d_img_state + 4 * (epoch & 1u) : nullptr;
  const int32_t* cell_prev = persistent ? d_img_state + 4 * ((epoch + 1u) & 1u) : nullptr;
  const bool fuse_zero = img_aligned && !persistent;
  // the bounding box sits right in front of the image in the symmetric exchange buffer
  // (exchange.cu): one fused zeroing range covers both
  const bool bbox_adjacent = d_screen_bbox && has_img && fuse_zero &&
                             (char*)d_screen_bbox + 64 == (char*)d_screen_img;

  TrackArgs a{};
  a.in = d_in_soa;
  a.in_row_stride = in_row_stride * 6 * ld_in;
  a.ld_in = ld_in;
  a.N = N;
  a.out = d_out_soa;
  a.ld_out = ld_out;
  a.records = records;
  a.rec_floats = p->rec_floats;
  a.plan = p->d_plan;
  a.n_plan = (int)p->plan.size();
  for (size_t i = 0; i < p->plan.size() && i < 8; ++i) a.plan_inline[i] = p->plan[i];
  a.n_snap = p->n_snap;
  a.final_centre = p->final_centre;
  a.bpm_sums = d_bpm_sums;
  a.img = (p->n_screen > 0) ? d_screen_img : nullptr;
  a.W = W;
  a.H = H;
  a.ex = d_edges_x;
  a.ey = d_edges_y;
  a.hist = reinterpret_cast<const HistAxis*>(hist_ws);
  a.bbox = has_img ? (persistent ? cell_cur : d_screen_bbox) : nullptr;
  a.stamp = bbox_adjacent ? reinterpret_cast<unsigned long long*>(d_screen_bbox + 8) : nullptr;
  a.mom = d_moment_sums;
  a.B = B;
  // one merged stage, then the terminal screen, and nothing else reads the beam
  a.xy_only = (has_img && !d_out_soa && moment_mode == 0 && p->plan.size() == 2 &&
               p->plan[0].kind == P_AFFINE && p->plan[1].kind == P_SCREEN) ? 1 : 0;
  {
    const size_t np_ = p->plan.size();
    a.xy_tail = (has_img && !d_out_soa && moment_mode == 0 && np_ > 2 && img_bytes &&
                 p->plan[np_ - 2].kind == P_AFFINE && p->plan[np_ - 1].kind == P_SCREEN)
                    ? (int)np_ - 2 : -1;
  }

  // accumulators are cleared first so that the kernels stay adjacent in the stream
  // (accumulate: a further range of the SAME beam -- records, zeroed image and sums of the
  // first range's call are in place, only the particle kernel runs)
  if (!accumulate) {
    if (p->n_bpm_slots > 0)
      LYNX_CUDA_OK(cudaMemsetAsync(d_bpm_sums, 0, sizeof(double) * 2 * p->n_bpm_slots * B, st));
    if (has_img && !fuse_zero && !persistent)
      LYNX_CUDA_OK(cudaMemsetAsync(d_screen_img, 0, img_bytes, st));
    if (moment_mode) LYNX_CUDA_OK(cudaMemsetAsync(d_moment_sums, 0, sizeof(double) * 28 * B, st));
  }

  bool force_ldg = false;
#ifdef LYNX_EXPERIMENTS
  static const bool env_ldg = [] { const char* e = getenv("LYNX_K2"); return e && e[0] == 'l'; }();
  force_ldg = env_ldg;
#endif
  // ---- K1 (records, image zeroing, histogram geometry) then K2
  if (d_screen_bbox && !bbox_adjacent && !accumulate)
    LYNX_CUDA_OK(cudaMemsetAsync(d_screen_bbox, 0, 16, st));
  int rc = accumulate ? LYNX_OK : compose_launch(p, d_params, B, d_energy_in, d_in_soa, in_row_stride * 6 * ld_in, ld_in,
                          records, nullptr, d_energy_out,
                          fuse_zero ? (bbox_adjacent ? (void*)d_screen_bbox : (void*)d_screen_img)
                                    : (persistent ? (void*)d_screen_img : nullptr),
                          fuse_zero ? img_bytes + (bbox_adjacent ? 64 : 0) : 0, st, has_img ? d_edges_x : nullptr,
                          has_img ? d_edges_y : nullptr, W, H, has_img ? hist_ws : nullptr,
                          cell_prev, cell_cur);
  if (rc != LYNX_OK) return rc;
  if (!force_ldg) {
    const int r = launch_track_pipe(a, moment_mode, d_out_soa != nullptr, p->sm_count, st);
    if (r <= 0) return r;
  }
  // the bounding box (and with it the persistent image) is a feature of the pipelined kernel
  if (d_screen_bbox || persistent) return LYNX_E_BADPROGRAM;
  a.bbox = nullptr;
  const int64_t n_tiles = (N + kTile - 1) / kTile;
  int64_t per_b = (2LL * p->sm_count + B - 1) / B;
  if (per_b < 1) per_b = 1;
  a.ctas_per_row = (int)std::min<int64_t>(n_tiles, per_b);
  if ((int64_t)a.ctas_per_row * B > 0x7fffffffLL) return LYNX_E_BADARG;
  const unsigned grid = (unsigned)((int64_t)a.ctas_per_row * B);
  size_t smem = sizeof(float) * p->rec_floats + sizeof(PlanItem) * p->plan.size() +
                sizeof(double) * 2 * p->n_snap + 16;
  if (smem > 200 * 1024) return LYNX_E_BADPROGRAM;
  const bool out = d_out_soa != nullptr;
  switch (moment_mode * 2 + (out ? 1 : 0)) {
    case 0: return launch_track<0, false>(a, grid, smem, st);
    case 1: return launch_track<0, true>(a, grid, smem, st);
    case 2: return launch_track<1, false>(a, grid, smem, st);
    case 3: return launch_track<1, true>(a, grid, smem, st);
    case 4: return launch_track<2, false>(a, grid, smem, st);
    default: return launch_track<2, true>(a, grid, smem, st);
  }
}

// K1: per-setting map composer -> records for the particle kernels.
// K6: ParameterBeam propagation (mu' = R mu, cov' = R cov R^T) fused with K1.
// One thread owns one setting; parameters are read as columns (coalesced across the
// warp for per-setting columns, broadcast for shared ones).
#include "walk.cuh"

namespace lynx {

// ------------------------------------------------------------------ K1
struct RecordSink {
  float* rec;
  float* maps_out;  // nullable, this setting's [n_stages][7][7]
  int off_cav, off_centre, off_screen;
  float ctr[6];

  __device__ __forceinline__ void write_map(const Map& M, int idx) {
    float* d = rec + idx * kMapStride;
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
      for (int j = 0; j < 7; ++j) d[i * 7 + j] = M.m[i][j];
    if (maps_out) {
      float* o = maps_out + idx * 49;
#pragma unroll
      for (int i = 0; i < 6; ++i)
#pragma unroll
        for (int j = 0; j < 7; ++j) o[i * 7 + j] = M.m[i][j];
#pragma unroll
      for (int j = 0; j < 7; ++j) o[42 + j] = (j == 6) ? 1.f : 0.f;
    }
  }
  __device__ __forceinline__ void push_centre(const Map& M, float* lin) {
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      float a = M.m[i][6];
#pragma unroll
      for (int j = 0; j < 6; ++j) a = fmaf(M.m[i][j], ctr[j], a);
      lin[i] = a;
    }
  }
  __device__ __forceinline__ void stage(const Map& M, int idx) {
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
  __device__ __forceinline__ void bpm(int, int snap) {
    float* d = rec + off_centre + snap * 6;
#pragma unroll
    for (int i = 0; i < 6; ++i) d[i] = ctr[i];
  }
  __device__ __forceinline__ void screen(int, int sidx, float mx, float my, int) {
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
};

__global__ void __launch_bounds__(128) compose_kernel(ComposeArgs a) {
  int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= a.B) return;
  ParamView P{a.params, a.col_offset, a.col_stride, b};
  RecordSink sink;
  sink.rec = a.records + b * a.rec_floats;
  sink.maps_out = a.maps_out ? a.maps_out + b * (int64_t)a.n_stages * 49 : nullptr;
  sink.off_cav = a.off_cav;
  sink.off_centre = a.off_centre;
  sink.off_screen = a.off_screen;
#pragma unroll
  for (int i = 0; i < 6; ++i)
    sink.ctr[i] = a.seed ? __ldg(a.seed + b * a.seed_row_stride + i * a.seed_ld) : 0.f;
  float energy = a.energy_in[b];
  walk_program(a.ops, a.n_ops, a.final_stage, P, energy, sink);
  float* fc = sink.rec + a.final_centre;
#pragma unroll
  for (int i = 0; i < 6; ++i) fc[i] = sink.ctr[i];
  if (a.energy_out) a.energy_out[b] = energy;
}

// ------------------------------------------------------------------ K6
struct MomentSink {
  float mu[7];
  float cov[7][7];
  float* bpm_out;
  int64_t B, b;

  // element.py:73-74 with R = [M; e6]
  __device__ __forceinline__ void apply(const Map& M) {
    float t[7];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      float acc = 0.f;
#pragma unroll
      for (int j = 0; j < 7; ++j) acc = fmaf(M.m[i][j], mu[j], acc);
      t[i] = acc;
    }
#pragma unroll
    for (int i = 0; i < 6; ++i) mu[i] = t[i];
    // cov <- R cov  (column by column, in place)
#pragma unroll
    for (int k = 0; k < 7; ++k) {
#pragma unroll
      for (int i = 0; i < 6; ++i) {
        float acc = 0.f;
#pragma unroll
        for (int j = 0; j < 7; ++j) acc = fmaf(M.m[i][j], cov[j][k], acc);
        t[i] = acc;
      }
#pragma unroll
      for (int i = 0; i < 6; ++i) cov[i][k] = t[i];
    }
    // cov <- cov R^T  (row by row, in place)
#pragma unroll
    for (int i = 0; i < 7; ++i) {
#pragma unroll
      for (int l = 0; l < 6; ++l) {
        float acc = 0.f;
#pragma unroll
        for (int k = 0; k < 7; ++k) acc = fmaf(cov[i][k], M.m[l][k], acc);
        t[l] = acc;
      }
#pragma unroll
      for (int l = 0; l < 6; ++l) cov[i][l] = t[l];
    }
  }
  __device__ __forceinline__ void stage(const Map& M, int) { apply(M); }
  // cavity.py:113-140, 202-218 (ParameterBeam branch; overwrites as written, quirk B5)
  __device__ __forceinline__ void cavity(const Map& C, const CavityOut& co, int, int) {
    const float mu4 = mu[4], mu5 = mu[5];
    const float c44 = cov[4][4], c45 = cov[4][5], c55 = cov[5][5];
    apply(C);
    const float* sc = co.sc;
    if (sc[CV_ON] != 0.f) {
      float s1, p1;
      cavity_nonlinear(sc, mu4, mu5, mu[4], mu[5], s1, p1);
      mu[4] = s1;
      mu[5] = p1;
      cov[5][5] = c55;
      float v = sc[CV_T566] * c55 * c55 + sc[CV_T556] * c45 * c55 + sc[CV_T555] * c44 * c44;
      cov[4][4] = v;
      cov[4][5] = v;
      cov[5][4] = v;
    }
  }
  __device__ __forceinline__ void bpm(int slot, int) {
    if (bpm_out) {
      bpm_out[(slot * B + b) * 2 + 0] = mu[0];  // bpm.py:52
      bpm_out[(slot * B + b) * 2 + 1] = mu[2];
    }
  }
  __device__ __forceinline__ void screen(int, int, float mx, float my, int) {
    mu[0] -= mx;  // screen.py:131-132
    mu[2] -= my;
  }
};

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

constexpr int kPBThreads = 64;

__global__ void __launch_bounds__(kPBThreads) parambeam_fwd_kernel(ParamBeamArgs a) {
  // staging tile so the [B][7] / [B][7][7] outputs leave the SM as full coalesced rows
  __shared__ float tile[kPBThreads * 57];
  const int64_t b0 = (int64_t)blockIdx.x * kPBThreads;
  const int64_t b = b0 + threadIdx.x;
  const bool live = b < a.B;
  MomentSink sink;
  float energy = 0.f;
  if (live) {
    ParamView P{a.params, a.col_offset, a.col_stride, b};
    const float* mi = a.mu_in + b * a.beam_stride * 7;
    const float* ci = a.cov_in + b * a.beam_stride * 49;
#pragma unroll
    for (int i = 0; i < 7; ++i) sink.mu[i] = __ldg(mi + i);
#pragma unroll
    for (int i = 0; i < 7; ++i)
#pragma unroll
      for (int j = 0; j < 7; ++j) sink.cov[i][j] = __ldg(ci + i * 7 + j);
    sink.bpm_out = a.bpm_out;
    sink.B = a.B;
    sink.b = b;
    energy = a.energy_in[b];
    walk_program(a.ops, a.n_ops, a.final_stage, P, energy, sink);
    if (a.energy_out) a.energy_out[b] = energy;
  }
  const int nb = (int)min((int64_t)kPBThreads, a.B - b0);
  // mu: [nb][7]
  if (live) {
#pragma unroll
    for (int i = 0; i < 7; ++i) tile[threadIdx.x * 7 + i] = sink.mu[i];
  }
  __syncthreads();
  for (int i = threadIdx.x; i < nb * 7; i += kPBThreads) a.mu_out[b0 * 7 + i] = tile[i];
  __syncthreads();
  if (live) {
#pragma unroll
    for (int i = 0; i < 7; ++i)
#pragma unroll
      for (int j = 0; j < 7; ++j) tile[threadIdx.x * 49 + i * 7 + j] = sink.cov[i][j];
  }
  __syncthreads();
  for (int i = threadIdx.x; i < nb * 49; i += kPBThreads) a.cov_out[b0 * 49 + i] = tile[i];
}

}  // namespace lynx

using namespace lynx;

extern "C" int lynx_compose_maps(const lynx_program_t* p, const float* d_params, int64_t B,
                                 const float* d_energy_in, const float* d_seed_soa,
                                 int64_t seed_row_stride, int64_t seed_ld, float* d_records,
                                 float* d_maps_out, float* d_energy_out,
                                 lynx_stream_t stream) {
  if (!p || !d_params || !d_energy_in || !d_records || B <= 0) return LYNX_E_BADARG;
  ComposeArgs a;
  a.ops = p->d_ops;
  a.n_ops = (int)p->dev_ops.size();
  a.final_stage = p->final_stage;
  a.col_offset = p->d_col_offset;
  a.col_stride = p->d_col_stride;
  a.params = d_params;
  a.B = B;
  a.energy_in = d_energy_in;
  a.seed = d_seed_soa;
  a.seed_row_stride = seed_row_stride;
  a.seed_ld = seed_ld;
  a.records = d_records;
  a.rec_floats = p->rec_floats;
  a.off_cav = p->off_cav;
  a.off_centre = p->off_centre;
  a.off_screen = p->off_screen;
  a.final_centre = p->final_centre;
  a.n_stages = p->n_stages;
  a.maps_out = d_maps_out;
  a.energy_out = d_energy_out;
  const int threads = 128;
  const unsigned blocks = (unsigned)((B + threads - 1) / threads);
  {
    ProfileScope ps(0, (cudaStream_t)stream);
    compose_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(a);
  }
  count_launch();
  LYNX_CUDA_OK(cudaGetLastError());
  return LYNX_OK;
}

extern "C" int lynx_parambeam_fwd(const lynx_program_t* p, const float* d_params, int64_t B,
                                  const float* d_mu_in, const float* d_cov_in,
                                  int64_t beam_stride, const float* d_energy_in,
                                  float* d_mu_out, float* d_cov_out, float* d_bpm_out,
                                  float* d_energy_out, lynx_stream_t stream) {
  if (!p || !d_params || !d_mu_in || !d_cov_in || !d_energy_in || !d_mu_out || !d_cov_out ||
      B <= 0 || (beam_stride != 0 && beam_stride != 1))
    return LYNX_E_BADARG;
  if (p->n_bpm_slots > 0 && !d_bpm_out) return LYNX_E_BADARG;
  ParamBeamArgs a;
  a.ops = p->d_ops;
  a.n_ops = (int)p->dev_ops.size();
  a.final_stage = p->final_stage;
  a.col_offset = p->d_col_offset;
  a.col_stride = p->d_col_stride;
  a.params = d_params;
  a.B = B;
  a.mu_in = d_mu_in;
  a.cov_in = d_cov_in;
  a.beam_stride = beam_stride;
  a.energy_in = d_energy_in;
  a.mu_out = d_mu_out;
  a.cov_out = d_cov_out;
  a.bpm_out = d_bpm_out;
  a.energy_out = d_energy_out;
  const unsigned blocks = (unsigned)((B + kPBThreads - 1) / kPBThreads);
  {
    ProfileScope ps(2, (cudaStream_t)stream);
    parambeam_fwd_kernel<<<blocks, kPBThreads, 0, (cudaStream_t)stream>>>(a);
  }
  count_launch();
  LYNX_CUDA_OK(cudaGetLastError());
  return LYNX_OK;
}

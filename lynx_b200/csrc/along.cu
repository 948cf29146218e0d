// K6-along: ParameterBeam tracked element by element with a snapshot after EVERY element
// (the "snapshot every element" mode behind Segment.plot_twiss,
// lynx/accelerator/segment.py:490-518, which calls element.track(beam) once per element and
// keeps every intermediate beam).  Unlike K6 nothing is merged: each element's own map is
// applied on its own (element.py:71-82), exactly the reference's sequence of roundings.
// One thread per setting; snapshots leave as [n_ops][B][7] / [n_ops][B][7][7] / [n_ops][B].
#include "compose_common.cuh"

namespace lynx {

constexpr int kAlongTPB = 64;

__global__ void __launch_bounds__(kAlongTPB) parambeam_along_kernel(ParamBeamArgs a,
                                                                    float* __restrict__ snap_mu,
                                                                    float* __restrict__ snap_cov,
                                                                    float* __restrict__ snap_energy) {
  const int64_t b = (int64_t)blockIdx.x * kAlongTPB + threadIdx.x;
  if (b >= a.B) return;
  MomentSink sink;
  const float* mi = a.mu_in + b * a.beam_stride * 7;
  const float* ci = a.cov_in + b * a.beam_stride * 49;
#pragma unroll
  for (int i = 0; i < 7; ++i) sink.mu[i] = __ldg(mi + i);
#pragma unroll
  for (int i = 0; i < 7; ++i)
#pragma unroll
    for (int j = 0; j < 7; ++j) sink.cov[i][j] = __ldg(ci + i * 7 + j);
  sink.bpm_out = nullptr;
  sink.B = a.B;
  sink.b = b;
  float energy = a.energy_in[b];
  RelT<float> rel = make_rel(energy);
  ParamView P{a.params, a.col_offset, a.col_stride, b};
  for (int i = 0; i < a.n_ops; ++i) {
    const DevOp op = a.ops[i];
    Map M;
    map_identity(M);
    if (lm_linear_op<kOpsAll>(M, op, P, rel, energy)) {
      if (op.opcode != LYNX_OP_MARKER) sink.apply(M);
    } else if (op.opcode == LYNX_OP_CAVITY) {
      Map C;
      CavityOut co;
      cavity_setting<float>(P(op.col), P(op.col + 1), P(op.col + 2), P(op.col + 3), energy, C, co);
      sink.cavity(C, co, 0, 0);
      energy += co.dE;  // cavity.py:130
      rel = make_rel(energy);
    }
    // BPM / Screen: zero length, the beam passes through (plot_twiss skips them anyway)
    float* m = snap_mu + ((int64_t)i * a.B + b) * 7;
    float* c = snap_cov + ((int64_t)i * a.B + b) * 49;
#pragma unroll
    for (int k = 0; k < 7; ++k) m[k] = sink.mu[k];
#pragma unroll
    for (int k = 0; k < 7; ++k)
#pragma unroll
      for (int l = 0; l < 7; ++l) c[k * 7 + l] = sink.cov[k][l];
    snap_energy[(int64_t)i * a.B + b] = energy;
  }
}

}  // namespace lynx

using namespace lynx;

extern "C" int lynx_parambeam_along(const lynx_program_t* p, const float* d_params, int64_t B,
                                    const float* d_mu_in, const float* d_cov_in,
                                    int64_t beam_stride, const float* d_energy_in,
                                    float* d_snap_mu, float* d_snap_cov, float* d_snap_energy,
                                    lynx_stream_t stream) {
  if (!p || !d_params || !d_mu_in || !d_cov_in || !d_energy_in || !d_snap_mu || !d_snap_cov ||
      !d_snap_energy || B <= 0 || (beam_stride != 0 && beam_stride != 1))
    return LYNX_E_BADARG;
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
  a.mu_out = nullptr;
  a.cov_out = nullptr;
  a.bpm_out = nullptr;
  a.energy_out = nullptr;
  if (a.n_ops == 0) return LYNX_OK;
  const unsigned blocks = (unsigned)((B + kAlongTPB - 1) / kAlongTPB);
  parambeam_along_kernel<<<blocks, kAlongTPB, 0, (cudaStream_t)stream>>>(a, d_snap_mu, d_snap_cov,
                                                                         d_snap_energy);
  count_launch();
  LYNX_CUDA_OK(cudaGetLastError());
  return LYNX_OK;
}

// Walks one setting through the compiled program, composing merged affine stages
// (Segment.transfer_map, segment.py:329-338; run grouping of segment.py:344-351) and
// handing them to a Sink.  Shared by K1 (records for the particle kernels) and K6
// (ParameterBeam propagation).
#pragma once
#include "maps.cuh"

namespace lynx {

struct ParamView {
  const float* params;
  const int64_t* col_offset;
  const int32_t* col_stride;
  int64_t b;
  __device__ __forceinline__ float operator()(int col) const {
    return __ldg(params + __ldg(col_offset + col) + b * (int64_t)__ldg(col_stride + col));
  }
};

// Sink concept:
//   bool flush_before(const DevOp&);   forward sinks: op.flush_stage >= 0 (the run grouping
//                                      of segment.py:344-351); the adjoint sinks flush only
//                                      before active cavities (they merge whole cavity-free
//                                      segments, BPMs included)
//   void finish(const Map& M, int final_stage);   end of program; forward sinks emit M as
//                                      stage `final_stage` if it is >= 0
//   void stage(const Map& M, int stage_idx);                 merged run finished
//   void cavity(const Map& C, const CavityOut& co, int stage_idx, int cav_idx);
//   void bpm(const Map& M, int slot, int snap_idx);     M = map composed since the last flush
//   void screen(const Map& M, int slot, int screen_idx, T mis_x, T mis_y, int flags);
// Parameter columns of a tile of settings staged in shared memory: vals[col * S + s].
struct SmemParamView {
  const float* vals;
  int S, s;
  __device__ __forceinline__ float operator()(int col) const { return vals[col * S + s]; }
};

// Left-multiplies M by the map of one linear (skippable) op.  Returns false for ops that
// are not part of a merged run (active cavity, BPM, screen).
// Opcode sets a kernel is compiled for.  The walk is bound by instruction fetch at small B
// (one thread executes ~1.3k instructions once; the full switch is 64 KB of SASS), so
// lattices made only of the cheap elements get a kernel without the dipole / cavity /
// solenoid / dense-custom code.
__host__ __device__ constexpr int op_bit(int opcode) { return 1 << opcode; }
constexpr int kOpsAll = (1 << LYNX_OP__COUNT) - 1;
constexpr int kOpsLight = op_bit(LYNX_OP_MARKER) | op_bit(LYNX_OP_DRIFT) | op_bit(LYNX_OP_QUAD) |
                          op_bit(LYNX_OP_HCOR) | op_bit(LYNX_OP_VCOR) | op_bit(LYNX_OP_UNDULATOR) |
                          op_bit(LYNX_OP_BPM) | op_bit(LYNX_OP_SCREEN);

template <int OPS, typename T, bool LT, typename PV>
__device__ __forceinline__ bool lm_linear_op(MapT<T, LT>& M, const DevOp& op, const PV& P,
                                             const RelT<T>& rel, T energy) {
  const int c = op.col;
  switch (op.opcode) {
    case LYNX_OP_MARKER:
      return true;
    case LYNX_OP_DRIFT: {
      T L = P(c);
      lm_drift(M, L, drift_r56(L, rel));
    } return true;
    case LYNX_OP_QUAD:
      lm_quad(M, P(c), P(c + 1), P(c + 2), P(c + 3), P(c + 4), rel);
      return true;
    case LYNX_OP_HCOR: {
      T L = P(c);
      lm_drift(M, L, drift_r56(L, rel));
      M.m[1][6] += P(c + 1);  // horizontal_corrector.py:63
    } return true;
    case LYNX_OP_VCOR: {
      T L = P(c);
      lm_drift(M, L, drift_r56(L, rel));
      M.m[3][6] += P(c + 1);  // vertical_corrector.py:63
    } return true;
    case LYNX_OP_DIPOLE:
      if constexpr ((OPS & op_bit(LYNX_OP_DIPOLE)) != 0)
        lm_dipole(M, P(c), P(c + 1), P(c + 2), P(c + 3), P(c + 4), P(c + 5), P(c + 6), P(c + 7),
                  (op.flags & LYNX_F_RBEND) != 0, rel);
      return true;
    case LYNX_OP_CAVITY: {
      if (op.flags & LYNX_F_ACTIVE) return false;
      // skippable cavity (voltage == 0 everywhere): linear map merged into the run
      if constexpr ((OPS & op_bit(LYNX_OP_CAVITY)) != 0) {
        MapT<T> C;
        CavityOutT<T> co;
        cavity_setting<T>(P(c), P(c + 1), P(c + 2), P(c + 3), energy, C, co);
        lm_custom(M, [&](int r, int k) { return C.m[r][k]; });
      }
    } return true;
    case LYNX_OP_SOLENOID:
      if constexpr ((OPS & op_bit(LYNX_OP_SOLENOID)) != 0)
        lm_solenoid(M, P(c), P(c + 1), P(c + 2), P(c + 3), rel);
      return true;
    case LYNX_OP_UNDULATOR: {
      T L = P(c);
      lm_drift(M, L, L * rel.igamma2_drift);  // undulator.py:52-58 (quirk B4)
    } return true;
    case LYNX_OP_CUSTOM:
      if constexpr ((OPS & op_bit(LYNX_OP_CUSTOM)) != 0)
        lm_custom(M, [&](int r, int k) { return P(c + r * 7 + k); });
      return true;
    default:
      return false;
  }
}

// LT: carry the merged map as a light map (maps.cuh); only for OPS within kOpsLight, and
// only with sinks whose stage / bpm / screen / finish accept MapT<T, true>.
template <int OPS, bool LT = false, typename T, typename Sink, typename PV>
__device__ __forceinline__ void walk_program(const DevOp* __restrict__ ops, int n_ops,
                                             int final_stage, const PV& P,
                                             T& energy, Sink& sink) {
  static_assert(!LT || (OPS & ~kOpsLight) == 0, "light maps need a light opcode set");
  MapT<T, LT> M;
  map_identity(M);
  RelT<T> rel = make_rel(energy);
  for (int i = 0; i < n_ops; ++i) {
    const DevOp op = ops[i];
    if (sink.flush_before(op)) {
      sink.stage(M, op.flush_stage);
      map_identity(M);
    }
    if (lm_linear_op<OPS>(M, op, P, rel, energy)) continue;
    const int c = op.col;
    if (op.opcode == LYNX_OP_CAVITY) {
      if constexpr ((OPS & op_bit(LYNX_OP_CAVITY)) != 0) {
        MapT<T> C;
        CavityOutT<T> co;
        cavity_setting<T>(P(c), P(c + 1), P(c + 2), P(c + 3), energy, C, co);
        sink.cavity(C, co, op.self_stage, op.cav_idx);
        energy += co.dE;  // cavity.py:130
        rel = make_rel(energy);
      }
    } else if (op.opcode == LYNX_OP_BPM) {
      sink.bpm(M, op.aux, op.snap_idx);
    } else if (op.opcode == LYNX_OP_SCREEN) {
      sink.screen(M, op.aux, op.snap_idx, P(c), P(c + 1), op.flags);
    }
  }
  sink.finish(M, final_stage);
}

}  // namespace lynx

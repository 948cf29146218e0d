// Internal shared definitions for the lynx_b200 CUDA library (not part of the C ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <vector>

#include "lynx_b200.h"

namespace lynx {

// ---- physical constants (CODATA 2018, the scipy.constants values the reference uses)
// lynx/track_methods.py:9-11  REST_ENERGY = m_e c^2 / e
constexpr double kRestEnergyEV = 510998.9506917531;
// lynx/accelerator/cavity.py:20  physical_constants["electron mass energy equivalent in MeV"]*1e6
constexpr double kElectronMassEV = 510998.95069;
constexpr double kSpeedOfLight = 299792458.0;
constexpr double kPi = 3.14159265358979323846;

// ---- compiled program ------------------------------------------------------------
// Device-side op: the ABI op plus the stage bookkeeping decided once on the host.
struct DevOp {
  int32_t opcode, col, flags, aux;
  int32_t flush_stage;  // >=0: before this op, emit accumulated map as this stage
  int32_t self_stage;   // active cavity: stage index of its own linear map
  int32_t cav_idx;      // active cavity: index into the cavity-scalar block
  int32_t snap_idx;     // BPM: centre index (plan order); SCREEN: screen index
};

// Plan item interpreted by the particle kernels (K2/K3), in beam order.
enum PlanKind : int32_t { P_AFFINE = 0, P_CAVITY = 1, P_BPM = 2, P_SCREEN = 3 };
struct PlanItem {
  int32_t kind;
  int32_t a;  // AFFINE: float offset of 6x7 map in record; CAVITY: offset of linear map
              // BPM: slot; SCREEN: offset of (mis_x, mis_y)
  int32_t b;  // CAVITY: offset of scalars; BPM: offset of centre (6 floats)
  int32_t c;  // BPM: snapshot index (plan order); SCREEN: flags
};

constexpr int kMapFloats = 42;  // rows 0..5 of the 7x7
constexpr int kMapStride = 44;  // record stride of one map (16-byte aligned for LDS.128)
constexpr int kCavFloats = 12;  // per active cavity scalars, see CavScal
// cavity scalar slots
enum CavScal { CV_AP = 0, CV_BP, CV_KB, CV_PHI, CV_T566, CV_T556, CV_T555, CV_ON, CV_DE,
               CV_SINPHI, CV_COSPHI };

}  // namespace lynx

struct lynx_program {
  std::vector<lynx_op_t> ops;
  std::vector<int64_t> col_offset;
  std::vector<int32_t> col_stride;
  std::vector<lynx::DevOp> dev_ops;
  std::vector<lynx::PlanItem> plan;
  int32_t n_stages = 0, n_cav = 0, n_bpm_slots = 0, n_snap = 0, n_screen = 0;
  int32_t final_stage = -1;  // stage flushed after the last op (-1: none)
  int32_t terminal = 0;
  int32_t op_mask = 0;       // bit per opcode present (kernel variant selection)
  // record layout (float offsets)
  int32_t off_cav = 0, off_centre = 0, off_screen = 0, rec_floats = 0;
  int32_t final_centre = 0;  // float offset of the outgoing-beam centre
  int64_t min_param_floats = 0;
  // device copies
  lynx::DevOp* d_ops = nullptr;
  int64_t* d_col_offset = nullptr;
  int32_t* d_col_stride = nullptr;
  lynx::PlanItem* d_plan = nullptr;
  int device = 0;
  int sm_count = 148;
};

namespace lynx {
extern std::atomic<int64_t> g_launches;
inline void count_launch(int n = 1) { g_launches.fetch_add(n, std::memory_order_relaxed); }

// K1 launcher shared by the public entry point and the particle launchers; the extra
// blocks of the same launch zero `zero_ptr` (screen image), saving a memset node.
int compose_launch(const lynx_program_t* p, const float* d_params, int64_t B,
                   const float* d_energy_in, const float* d_seed_soa, int64_t seed_row_stride,
                   int64_t seed_ld, float* d_records, float* d_maps_out, float* d_energy_out,
                   void* zero_ptr, size_t zero_bytes, cudaStream_t st,
                   const float* hist_ex = nullptr, const float* hist_ey = nullptr, int hist_W = 0,
                   int hist_H = 0, void* hist_out = nullptr, const int* zero_rect = nullptr,
                   int* bbox_clear = nullptr);

// forward moment collapse (adjoint.cu): one beam shared by B settings, moments / BPM sums
// only, no active cavity -> the beam is streamed once instead of once per setting
size_t fwd_collapse_workspace_bytes(const lynx_program_t* p, int64_t B);
int fwd_collapse_launch(const lynx_program_t* p, const float* d_params, int64_t B,
                        const float* d_in_soa, int64_t ld_in, int64_t N, double* d_bpm_sums,
                        double* d_moment_sums, const float* d_energy_in, float* d_energy_out,
                        void* d_workspace, cudaStream_t st);

// The forward launcher behind lynx_track_particles_fwd / _bbox / _persistent.  `accumulate`:
// a further particle range of the SAME call (host.cu tracks a beam in chunks as they arrive over
// PCIe): no K1, nothing cleared -- records, histogram geometry, the zeroed image and the sums
// of the first range's launch are in place -- only the particle kernel runs and adds.
int track_fwd_impl(const lynx_program_t* p, const float* d_params, int64_t B, const float* d_in_soa,
                   int64_t in_row_stride, int64_t ld_in, int64_t N, float* d_out_soa, int64_t ld_out,
                   double* d_bpm_sums, float* d_screen_img, int32_t W, int32_t H,
                   const float* d_edges_x, const float* d_edges_y, double* d_moment_sums,
                   int32_t moment_mode, const float* d_energy_in, float* d_energy_out,
                   void* d_workspace, int32_t* d_screen_bbox, int32_t* d_img_state, uint32_t epoch,
                   lynx_stream_t stream, bool accumulate);
bool fwd_collapses(const lynx_program_t* p, int64_t B, int64_t in_row_stride, bool has_out,
                   int moment_mode);

// optional event bracketing of the dominant kernels (bench roofline)
struct ProfileScope {
  int which;
  cudaStream_t st;
  bool on;
  ProfileScope(int which, cudaStream_t st);
  ~ProfileScope();
};

#define LYNX_CUDA_OK(expr)                                   \
  do {                                                       \
    cudaError_t _e = (expr);                                 \
    if (_e != cudaSuccess) return LYNX_E_LAUNCH;             \
  } while (0)

}  // namespace lynx

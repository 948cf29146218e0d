// Program handles (plan compilation), layout converters, host-buffer entry points.
#include <algorithm>
#include <cstring>
#include <new>

#include "common.cuh"

namespace lynx {
std::atomic<int64_t> g_launches{0};

static std::atomic<int> g_profile{0};
constexpr int kProfSlots = 8;
static cudaEvent_t g_ev[kProfSlots][2];
static bool g_ev_made[kProfSlots] = {};
static bool g_ev_used[kProfSlots] = {};

ProfileScope::ProfileScope(int w, cudaStream_t s) : which(w), st(s), on(g_profile.load() != 0) {
  if (!on) return;
  if (!g_ev_made[which]) {
    cudaEventCreate(&g_ev[which][0]);
    cudaEventCreate(&g_ev[which][1]);
    g_ev_made[which] = true;
  }
  cudaEventRecord(g_ev[which][0], st);
}
ProfileScope::~ProfileScope() {
  if (!on) return;
  cudaEventRecord(g_ev[which][1], st);
  g_ev_used[which] = true;
}

static int op_cols(int opcode) {
  switch (opcode) {
    case LYNX_OP_MARKER: return 0;
    case LYNX_OP_DRIFT: return 1;
    case LYNX_OP_QUAD: return 5;
    case LYNX_OP_HCOR: return 2;
    case LYNX_OP_VCOR: return 2;
    case LYNX_OP_DIPOLE: return 8;
    case LYNX_OP_CAVITY: return 4;
    case LYNX_OP_SOLENOID: return 4;
    case LYNX_OP_UNDULATOR: return 1;
    case LYNX_OP_CUSTOM: return 42;
    case LYNX_OP_BPM: return 0;
    case LYNX_OP_SCREEN: return 2;
    default: return -1;
  }
}

// ---- AoS (..,N,7) <-> SoA-6 converters -------------------------------------------
// One warp moves 32 particles: coalesced 7-float rows on the AoS side through a
// shared-memory tile, coalesced planes on the SoA side.
constexpr int kCvtTPB = 256;
constexpr int kCvtTile = 256;  // particles per CTA

__global__ void __launch_bounds__(kCvtTPB) aos7_to_soa6_kernel(const float* __restrict__ aos,
                                                                int64_t N,
                                                                float* __restrict__ soa,
                                                                int64_t ld) {
  __shared__ float t[kCvtTile * 7];
  const int64_t r = blockIdx.y;
  const int64_t p0 = (int64_t)blockIdx.x * kCvtTile;
  const int n = (int)min((int64_t)kCvtTile, N - p0);
  const float* src = aos + (r * N + p0) * 7;
  for (int i = threadIdx.x; i < n * 7; i += kCvtTPB) t[i] = src[i];
  __syncthreads();
  float* dst = soa + r * 6 * ld + p0;
  for (int i = threadIdx.x; i < n * 6; i += kCvtTPB) {
    int k = i / n, j = i - k * n;
    dst[k * ld + j] = t[j * 7 + k];
  }
}

__global__ void __launch_bounds__(kCvtTPB) soa6_to_aos7_kernel(const float* __restrict__ soa,
                                                                int64_t N, int64_t ld,
                                                                float* __restrict__ aos) {
  __shared__ float t[kCvtTile * 7];
  const int64_t r = blockIdx.y;
  const int64_t p0 = (int64_t)blockIdx.x * kCvtTile;
  const int n = (int)min((int64_t)kCvtTile, N - p0);
  const float* src = soa + r * 6 * ld + p0;
  for (int i = threadIdx.x; i < n * 6; i += kCvtTPB) {
    int k = i / n, j = i - k * n;
    t[j * 7 + k] = src[k * ld + j];
  }
  for (int j = threadIdx.x; j < n; j += kCvtTPB) t[j * 7 + 6] = 1.f;
  __syncthreads();
  float* dst = aos + (r * N + p0) * 7;
  for (int i = threadIdx.x; i < n * 7; i += kCvtTPB) dst[i] = t[i];
}

}  // namespace lynx

using namespace lynx;

extern "C" int lynx_version(void) { return 100; }

extern "C" const char* lynx_status_str(int s) {
  switch (s) {
    case LYNX_OK: return "ok";
    case LYNX_E_BADARG: return "bad argument";
    case LYNX_E_BADPROGRAM: return "bad program";
    case LYNX_E_LAUNCH: return "CUDA launch/enqueue error";
    case LYNX_E_NOMEM: return "out of memory";
    case LYNX_E_NODEVICE: return "no CUDA device (there is no CPU fallback)";
    case LYNX_E_TOOLARGE: return "problem exceeds the static scratch of the no-allocation entry point (LYNX_NOALLOC_WORKSPACE_BYTES)";
    default: return "unknown";
  }
}

extern "C" void lynx_profile_enable(int on) { g_profile.store(on); }
extern "C" float lynx_profile_last_ms(int which) {
  if (which < 0 || which >= kProfSlots || !g_ev_used[which]) return -1.f;
  float ms = -1.f;
  if (cudaEventSynchronize(g_ev[which][1]) != cudaSuccess) return -1.f;
  if (cudaEventElapsedTime(&ms, g_ev[which][0], g_ev[which][1]) != cudaSuccess) return -1.f;
  return ms;
}

extern "C" int64_t lynx_launch_count(void) { return g_launches.load(); }
extern "C" void lynx_launch_count_reset(void) { g_launches.store(0); }

extern "C" int lynx_program_create(const lynx_op_t* ops, int32_t n_ops,
                                   const int64_t* col_offset, const int32_t* col_stride,
                                   int32_t n_cols, lynx_program_t** out) {
  if (!out) return LYNX_E_BADARG;
  *out = nullptr;
  if (!ops || n_ops <= 0 || n_cols < 0 || (n_cols > 0 && (!col_offset || !col_stride)))
    return LYNX_E_BADARG;
  int dev_count = 0;
  if (cudaGetDeviceCount(&dev_count) != cudaSuccess || dev_count == 0) return LYNX_E_NODEVICE;

  lynx_program* p = new (std::nothrow) lynx_program();
  if (!p) return LYNX_E_NOMEM;
  p->ops.assign(ops, ops + n_ops);
  p->col_offset.assign(col_offset, col_offset + n_cols);
  p->col_stride.assign(col_stride, col_stride + n_cols);
  for (int c = 0; c < n_cols; ++c) {
    if (col_offset[c] < 0 || (col_stride[c] != 0 && col_stride[c] != 1)) {
      delete p;
      return LYNX_E_BADPROGRAM;
    }
  }

  // ---- plan compilation: run grouping of Segment.track (segment.py:344-351) --------
  bool pending = false;
  int S = 0, C = 0, n_snap = 0, n_screen = 0, n_slots = 0;
  struct Pend { int kind, stage, cav, slot, snap, flags; };
  std::vector<Pend> items;
  for (int i = 0; i < n_ops && !p->terminal; ++i) {
    const lynx_op_t& o = ops[i];
    const int nc = op_cols(o.opcode);
    if (nc < 0 || (nc > 0 && (o.col < 0 || o.col + nc > n_cols))) {
      delete p;
      return LYNX_E_BADPROGRAM;
    }
    DevOp d{o.opcode, o.col, o.flags, o.aux, -1, -1, -1, -1};
    p->op_mask |= 1 << o.opcode;
    auto flush = [&]() {
      if (pending) {
        d.flush_stage = S;
        items.push_back({P_AFFINE, S, 0, 0, 0, 0});
        ++S;
        pending = false;
      }
    };
    switch (o.opcode) {
      case LYNX_OP_MARKER:
        break;
      case LYNX_OP_CAVITY:
        if (o.flags & LYNX_F_ACTIVE) {
          flush();
          d.self_stage = S++;
          d.cav_idx = C++;
          items.push_back({P_CAVITY, d.self_stage, d.cav_idx, 0, 0, 0});
        } else {
          pending = true;
        }
        break;
      case LYNX_OP_BPM:
        if (o.aux < 0 || o.aux >= 4096) {
          delete p;
          return LYNX_E_BADPROGRAM;
        }
        flush();
        d.snap_idx = n_snap++;
        n_slots = std::max(n_slots, o.aux + 1);
        items.push_back({P_BPM, 0, 0, o.aux, d.snap_idx, 0});
        break;
      case LYNX_OP_SCREEN:
        flush();
        d.snap_idx = n_screen++;
        items.push_back({P_SCREEN, 0, 0, o.aux, d.snap_idx, o.flags});
        p->terminal = 1;
        break;
      default:
        pending = true;
        break;
    }
    p->dev_ops.push_back(d);
  }
  if (pending) {
    p->final_stage = S;
    items.push_back({P_AFFINE, S, 0, 0, 0, 0});
    ++S;
  }
  p->n_stages = S;
  p->n_cav = C;
  p->n_snap = n_snap;
  p->n_bpm_slots = n_slots;
  p->n_screen = n_screen;
  p->off_cav = S * kMapStride;
  p->off_centre = p->off_cav + C * kCavFloats;
  p->final_centre = p->off_centre + n_snap * 6;
  p->off_screen = p->final_centre + 6;
  p->rec_floats = (p->off_screen + n_screen * 2 + 3) & ~3;
  for (const Pend& it : items) {
    PlanItem q{it.kind, 0, 0, 0};
    switch (it.kind) {
      case P_AFFINE: q.a = it.stage * kMapStride; break;
      case P_CAVITY:
        q.a = it.stage * kMapStride;
        q.b = p->off_cav + it.cav * kCavFloats;
        break;
      case P_BPM:
        q.a = it.slot;
        q.b = p->off_centre + it.snap * 6;
        q.c = it.snap;
        break;
      default:
        q.a = p->off_screen + it.snap * 2;
        q.c = it.flags;
        break;
    }
    p->plan.push_back(q);
  }

  // ---- device copies -----------------------------------------------------------------
  cudaGetDevice(&p->device);
  cudaDeviceGetAttribute(&p->sm_count, cudaDevAttrMultiProcessorCount, p->device);
  bool ok = true;
  auto up = [&](void** dptr, const void* src, size_t bytes) {
    if (bytes == 0) bytes = 16, src = nullptr;
    if (cudaMalloc(dptr, bytes) != cudaSuccess) { ok = false; return; }
    if (src && cudaMemcpy(*dptr, src, bytes, cudaMemcpyHostToDevice) != cudaSuccess) ok = false;
  };
  up((void**)&p->d_ops, p->dev_ops.data(), p->dev_ops.size() * sizeof(DevOp));
  up((void**)&p->d_col_offset, p->col_offset.data(), p->col_offset.size() * sizeof(int64_t));
  up((void**)&p->d_col_stride, p->col_stride.data(), p->col_stride.size() * sizeof(int32_t));
  up((void**)&p->d_plan, p->plan.data(), p->plan.size() * sizeof(PlanItem));
  if (!ok) {
    lynx_program_destroy(p);
    return LYNX_E_NOMEM;
  }
  *out = p;
  return LYNX_OK;
}

extern "C" void lynx_program_destroy(lynx_program_t* p) {
  if (!p) return;
  cudaFree(p->d_ops);
  cudaFree(p->d_col_offset);
  cudaFree(p->d_col_stride);
  cudaFree(p->d_plan);
  delete p;
}

extern "C" int32_t lynx_program_num_stages(const lynx_program_t* p) { return p ? p->n_stages : -1; }
extern "C" int32_t lynx_program_num_bpm(const lynx_program_t* p) { return p ? p->n_bpm_slots : -1; }
extern "C" int32_t lynx_program_num_screens(const lynx_program_t* p) { return p ? p->n_screen : -1; }
extern "C" int32_t lynx_program_is_terminal(const lynx_program_t* p) { return p ? p->terminal : -1; }
extern "C" int64_t lynx_program_record_floats(const lynx_program_t* p) { return p ? p->rec_floats : -1; }

extern "C" int lynx_aos7_to_soa6(const float* d_aos, int64_t R, int64_t N, float* d_soa,
                                 int64_t ld, lynx_stream_t stream) {
  if (!d_aos || !d_soa || R <= 0 || N <= 0 || ld < N || R > 65535) return LYNX_E_BADARG;
  dim3 grid((unsigned)((N + kCvtTile - 1) / kCvtTile), (unsigned)R);
  aos7_to_soa6_kernel<<<grid, kCvtTPB, 0, (cudaStream_t)stream>>>(d_aos, N, d_soa, ld);
  count_launch();
  LYNX_CUDA_OK(cudaGetLastError());
  return LYNX_OK;
}

extern "C" int lynx_soa6_to_aos7(const float* d_soa, int64_t R, int64_t N, int64_t ld,
                                 float* d_aos, lynx_stream_t stream) {
  if (!d_aos || !d_soa || R <= 0 || N <= 0 || ld < N || R > 65535) return LYNX_E_BADARG;
  dim3 grid((unsigned)((N + kCvtTile - 1) / kCvtTile), (unsigned)R);
  soa6_to_aos7_kernel<<<grid, kCvtTPB, 0, (cudaStream_t)stream>>>(d_soa, N, ld, d_aos);
  count_launch();
  LYNX_CUDA_OK(cudaGetLastError());
  return LYNX_OK;
}

// K2, Blackwell-native variant: persistent warp-specialised CTA (one per SM) with a
// TMA bulk-copy pipeline.
//
//   warp 0      producer: one lane issues cp.async.bulk global->shared for the six
//               coordinate planes of a tile (6 x 8 KB), completion on an mbarrier
//   warp 1      storer:   cp.async.bulk shared->global of finished tiles (bulk groups)
//   warps 2-17  consumers: LDS.128 a float4 of particles per plane, apply the whole
//               lattice in registers, write back in place / histogram / moments
//
// In-flight bytes are set by shared memory (4 stages x 48 KB = 192 KB per SM), not by
// registers x occupancy, so the whole 1M-particle beam is requested in the first
// microsecond and large beams stream at copy speed.  SASS: UBLKCP (bulk copies),
// SYNCS (mbarrier).  Same arithmetic and outputs as track_fwd_kernel (track.cu).
#include <algorithm>

#include "track_common.cuh"

namespace lynx {

constexpr int kPipeTile = 2048;                    // particles per stage
constexpr int kPipeStages = 4;
constexpr int kConsumerWarps = 16;
constexpr int kConsumers = kConsumerWarps * 32;    // 512 threads x 4 particles = tile
constexpr int kPipeThreads = kConsumers + 64;      // + producer warp + storer warp
constexpr int kStageFloats = 6 * kPipeTile;
constexpr size_t kStageBytes = sizeof(float) * kStageFloats * kPipeStages;  // 196608

// ---- PTX wrappers ---------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra WAIT_DONE;\n"
      "bra WAIT_LOOP;\n"
      "WAIT_DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes,
                                         uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
      ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* dst_gmem, const void* src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem),
               "r"(smem_u32(src_smem)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

struct PipeSmemLayout {
  // byte offsets from the (128-B aligned) dynamic shared memory base
  uint32_t rec, plan, bpm, edges_x, edges_y, bars, total;
};

__host__ __device__ inline PipeSmemLayout pipe_layout(int rec_floats, int n_plan, int n_snap,
                                                      int W, int H, bool edges_in_smem) {
  PipeSmemLayout L;
  uint32_t o = (uint32_t)kStageBytes;
  L.rec = o;
  o += 4u * rec_floats;
  o = (o + 15u) & ~15u;
  L.plan = o;
  o += 16u * n_plan;
  L.bpm = o;
  o += 16u * n_snap;
  L.edges_x = o;
  if (edges_in_smem) o += 4u * (W + 1);
  L.edges_y = o;
  if (edges_in_smem) o += 4u * (H + 1);
  o = (o + 7u) & ~7u;
  L.bars = o;
  o += 8u * 3 * kPipeStages;
  L.total = o;
  return L;
}

template <int MOM, bool HAS_OUT>
__global__ void __launch_bounds__(kPipeThreads, 1) track_pipe_kernel(TrackArgs a, int edges_in_smem) {
  extern __shared__ __align__(128) unsigned char psm[];
  const PipeSmemLayout L =
      pipe_layout(a.rec_floats, a.n_plan, a.n_snap, a.W, a.H, edges_in_smem != 0);
  float* stages = reinterpret_cast<float*>(psm);
  float* rec = reinterpret_cast<float*>(psm + L.rec);
  PlanItem* plan = reinterpret_cast<PlanItem*>(psm + L.plan);
  double* s_bpm = reinterpret_cast<double*>(psm + L.bpm);
  uint64_t* full = reinterpret_cast<uint64_t*>(psm + L.bars);
  uint64_t* empty = full + kPipeStages;
  uint64_t* done = empty + kPipeStages;
  __shared__ double s_red[kConsumerWarps][28];
  __shared__ HistAxis s_hx, s_hy;

  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  const int64_t b = blockIdx.y;
  const int64_t n_tiles = (a.N + kPipeTile - 1) / kPipeTile;
  const float* in_b = a.in + b * a.in_row_stride;

  if (tid == 0) {
    for (int s = 0; s < kPipeStages; ++s) {
      mbar_init(full + s, 1);
      mbar_init(empty + s, HAS_OUT ? 1 : kConsumerWarps);
      mbar_init(done + s, kConsumerWarps);
    }
    fence_mbar_init();
  }
  __syncthreads();

  auto issue_tile = [&](int it, int64_t t) {
    const int s = it % kPipeStages;
    const int64_t p0 = t * kPipeTile;
    const uint32_t cnt = (uint32_t)min((int64_t)kPipeTile, a.N - p0);
    const uint32_t bytes = ((cnt + 3u) & ~3u) * 4u;
    mbar_arrive_expect_tx(full + s, 6u * bytes);
    float* dst = stages + s * kStageFloats;
#pragma unroll
    for (int k = 0; k < 6; ++k)
      bulk_g2s(dst + k * kPipeTile, in_b + k * a.ld_in + p0, bytes, full + s);
  };

  // ---- prologue: request this CTA's first tiles before anything else touches memory
  if (tid == 0) {
    int it = 0;
    for (int64_t t = blockIdx.x; t < n_tiles && it < kPipeStages; t += gridDim.x, ++it)
      issue_tile(it, t);
  }

  // ---- everyone: plan and histogram geometry (independent of K1) ...
  for (int i = tid; i < a.n_plan; i += kPipeThreads) plan[i] = a.plan[i];
  for (int i = tid; i < a.n_snap * 2; i += kPipeThreads) s_bpm[i] = 0.0;
  if (a.img) {  // uniform over the CTA
    hist_axis_setup(a.ex, a.W, &s_hx, tid, kPipeThreads);
    hist_axis_setup(a.ey, a.H, &s_hy, tid, kPipeThreads);
  }
  // ... then wait for the K1 grid (records written, image zeroed) -- a no-op unless this
  // kernel was launched with programmatic stream serialization
  asm volatile("griddepcontrol.wait;" ::: "memory");
  {
    const float* src = a.records + b * a.rec_floats;
    for (int i = tid; i < a.rec_floats; i += kPipeThreads) rec[i] = src[i];
  }
  __syncthreads();

  MomAcc<MOM> acc;
#pragma unroll
  for (int i = 0; i < MomAcc<MOM>::kN; ++i) acc.a[i] = 0.f;
  float ctr[6];
#pragma unroll
  for (int i = 0; i < 6; ++i) ctr[i] = rec[a.final_centre + i];
  int64_t n_mine = 0;

  if (warp == 0) {
    // ---- producer: keep the ring full
    if (lane == 0) {
      int it = 0;
      for (int64_t t = blockIdx.x; t < n_tiles; t += gridDim.x, ++it) {
        if (it < kPipeStages) continue;  // issued in the prologue
        const uint32_t u = (uint32_t)(it / kPipeStages);
        mbar_wait(empty + (it % kPipeStages), (u - 1) & 1);
        issue_tile(it, t);
      }
    }
  } else if (warp == 1) {
    // ---- storer: finished tiles leave through the async proxy
    if (HAS_OUT && lane == 0) {
      float* out_b = a.out + b * 6 * a.ld_out;
      int it = 0, prev_s = -1;
      for (int64_t t = blockIdx.x; t < n_tiles; t += gridDim.x, ++it) {
        const int s = it % kPipeStages;
        const uint32_t u = (uint32_t)(it / kPipeStages);
        mbar_wait(done + s, u & 1);
        fence_proxy_async();
        const int64_t p0 = t * kPipeTile;
        const uint32_t cnt = (uint32_t)min((int64_t)kPipeTile, a.N - p0);
        const uint32_t bytes = ((cnt + 3u) & ~3u) * 4u;
        const float* src = stages + s * kStageFloats;
#pragma unroll
        for (int k = 0; k < 6; ++k)
          bulk_s2g(out_b + k * a.ld_out + p0, src + k * kPipeTile, bytes);
        bulk_commit();
        if (prev_s >= 0) {  // the previous tile's smem has been read: hand the stage back
          bulk_wait_read<1>();
          mbar_arrive(empty + prev_s);
        }
        prev_s = s;
      }
      bulk_wait_read<0>();
      if (prev_s >= 0) mbar_arrive(empty + prev_s);
    }
  } else {
    // ---- consumers: 512 threads x 4 particles = one tile
    const int cid = tid - 64;
    const float* ex = a.ex;
    const float* ey = a.ey;
    const HistAxis hx = s_hx, hy = s_hy;
    float* img_b = a.img ? a.img + b * (int64_t)a.W * a.H : nullptr;
    int it = 0;
    for (int64_t t = blockIdx.x; t < n_tiles; t += gridDim.x, ++it) {
      const int s = it % kPipeStages;
      const uint32_t u = (uint32_t)(it / kPipeStages);
      const int64_t p0 = t * kPipeTile;
      n_mine += min((int64_t)kPipeTile, a.N - p0);
      float* st = stages + s * kStageFloats + cid * 4;
      mbar_wait(full + s, u & 1);
      float v[6][4];
#pragma unroll
      for (int k = 0; k < 6; ++k) {
        const float4 q = *reinterpret_cast<const float4*>(st + k * kPipeTile);
        v[k][0] = q.x;
        v[k][1] = q.y;
        v[k][2] = q.z;
        v[k][3] = q.w;
      }
      if (!HAS_OUT) {  // values live in registers now: the stage can be refilled already
        __syncwarp();
        if (lane == 0) mbar_arrive(empty + s);
      }
      const int64_t base = p0 + (int64_t)cid * 4;
      bool live[4];
#pragma unroll
      for (int l = 0; l < 4; ++l) live[l] = base + l < a.N;

      for (int pi = 0; pi < a.n_plan; ++pi) {
        const PlanItem p = plan[pi];
        if (p.kind == P_AFFINE) {
          apply_affine(rec + p.a, v);
        } else if (p.kind == P_CAVITY) {
          float s0[4], q0[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            s0[j] = v[4][j];
            q0[j] = v[5][j];
          }
          apply_affine(rec + p.a, v);
          const float* sc = rec + p.b;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            float s1, p1;
            cavity_nonlinear(sc, s0[j], q0[j], v[4][j], v[5][j], s1, p1);
            v[4][j] = s1;
            v[5][j] = p1;
          }
        } else if (p.kind == P_BPM) {  // bpm.py:54
          const float cx = rec[p.b], cy = rec[p.b + 2];
          float sx = 0.f, sy = 0.f;
#pragma unroll
          for (int l = 0; l < 4; ++l) {
            sx += live[l] ? v[0][l] - cx : 0.f;
            sy += live[l] ? v[2][l] - cy : 0.f;
          }
          sx = warp_sum(sx);
          sy = warp_sum(sy);
          if (lane == 0) {
            atomicAdd(&s_bpm[2 * p.c + 0], (double)sx);
            atomicAdd(&s_bpm[2 * p.c + 1], (double)sy);
          }
        } else if (img_b) {  // P_SCREEN, screen.py:126-141 + 195-211
          const float mx = rec[p.a], my = rec[p.a + 1];
          const bool yfix = (p.c & LYNX_F_SCREEN_Y_FIX) != 0;
          int ix[4], iy[4];
#pragma unroll
          for (int l = 0; l < 4; ++l) {
            const float x = v[0][l] - mx;
            const float y = yfix ? v[2][l] - my : v[2][l];
            ix[l] = find_bin(x, ex, a.W, hx);
            iy[l] = find_bin(y, ey, a.H, hy);
          }
#pragma unroll
          for (int l = 0; l < 4; ++l)
            if (live[l] && ix[l] >= 0 && iy[l] >= 0) {
              if (a.dbg & 1) {
                if (ix[l] + iy[l] == -12345) img_b[0] = 1.f;
              } else {
                atomicAdd(img_b + (int64_t)(a.H - 1 - iy[l]) * a.W + ix[l], 1.0f);
              }
            }
        }
      }

      if (HAS_OUT) {
#pragma unroll
        for (int k = 0; k < 6; ++k)
          *reinterpret_cast<float4*>(st + k * kPipeTile) =
              make_float4(v[k][0], v[k][1], v[k][2], v[k][3]);
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(done + s);
      }
      if (MOM != 0) {
#pragma unroll
        for (int l = 0; l < 4; ++l) {
          float d[6];
#pragma unroll
          for (int k = 0; k < 6; ++k) d[k] = live[l] ? v[k][l] - ctr[k] : 0.f;
          mom_accumulate(acc, d);
        }
      }
    }
  }

  // ---- CTA epilogue
  if (MOM != 0) {
    mom_epilogue<MOM, kConsumerWarps>(acc, s_red, warp >= 2 ? warp - 2 : -1, lane, tid == 64,
                                      n_mine, ctr, a.mom + b * 28);
  }
  if (a.n_snap > 0) {
    __syncthreads();
    // n_mine is only tracked by consumers; recompute for the flushing threads
    int64_t n_cta = 0;
    for (int64_t t = blockIdx.x; t < n_tiles; t += gridDim.x)
      n_cta += min((int64_t)kPipeTile, a.N - t * kPipeTile);
    for (int pi = tid; pi < a.n_plan; pi += kPipeThreads) {
      const PlanItem p = plan[pi];
      if (p.kind == P_BPM) {
        const double n = (double)n_cta;
        double* g = a.bpm_sums + ((int64_t)p.a * a.B + b) * 2;
        atomicAdd(g, s_bpm[2 * p.c] + n * (double)rec[p.b]);
        atomicAdd(g + 1, s_bpm[2 * p.c + 1] + n * (double)rec[p.b + 2]);
      }
    }
  }
}

template <int MOM, bool HAS_OUT>
static int launch_pipe(const TrackArgs& a, dim3 grid, size_t smem, int edges_in_smem,
                       cudaStream_t st) {
  auto k = track_pipe_kernel<MOM, HAS_OUT>;
  static bool attr_set = false;  // per instantiation
  if (!attr_set) {
    if (cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 4096) !=
        cudaSuccess)
      return LYNX_E_LAUNCH;
    attr_set = true;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(kPipeThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t err;
  {
    ProfileScope ps(1, st);
    err = cudaLaunchKernelEx(&cfg, k, a, edges_in_smem);
  }
  count_launch();
  return err == cudaSuccess ? LYNX_OK : LYNX_E_LAUNCH;
}

// Returns LYNX_OK after enqueueing, or 1 if this variant cannot take the problem (the
// caller then uses the LDG kernel).
int launch_track_pipe(const TrackArgs& a, int moment_mode, bool has_out, int sm_count,
                      cudaStream_t st) {
  const size_t cap = 227 * 1024 - 4096;  // static shared memory (s_red) takes the rest
  int edges_in_smem = 0;  // the table is only consulted near bin edges: it stays in L1/L2
  PipeSmemLayout L = pipe_layout(a.rec_floats, a.n_plan, a.n_snap, a.W, a.H, edges_in_smem != 0);
  if (L.total > cap && edges_in_smem) {
    edges_in_smem = 0;
    L = pipe_layout(a.rec_floats, a.n_plan, a.n_snap, a.W, a.H, false);
  }
  if (L.total > cap) return 1;
  const int64_t n_tiles = (a.N + kPipeTile - 1) / kPipeTile;
  int64_t per_b = (sm_count + a.B - 1) / a.B;
  if (per_b < 1) per_b = 1;
  dim3 grid((unsigned)std::min<int64_t>(n_tiles, per_b), (unsigned)a.B);
  switch (moment_mode * 2 + (has_out ? 1 : 0)) {
    case 0: return launch_pipe<0, false>(a, grid, L.total, edges_in_smem, st);
    case 1: return launch_pipe<0, true>(a, grid, L.total, edges_in_smem, st);
    case 2: return launch_pipe<1, false>(a, grid, L.total, edges_in_smem, st);
    case 3: return launch_pipe<1, true>(a, grid, L.total, edges_in_smem, st);
    case 4: return launch_pipe<2, false>(a, grid, L.total, edges_in_smem, st);
    default: return launch_pipe<2, true>(a, grid, L.total, edges_in_smem, st);
  }
}

}  // namespace lynx

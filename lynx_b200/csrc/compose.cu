// K1: per-setting map composer -> records for the particle kernels.
// K6: ParameterBeam propagation (mu' = R mu, cov' = R cov R^T) fused with K1.
// One thread owns one setting; parameters are read as columns (coalesced across the
// warp for per-setting columns, broadcast for shared ones).
#include <algorithm>
#include <cstdlib>

#include "compose_common.cuh"
#include "pipe.cuh"
#include "track_common.cuh"

namespace lynx {

// Stage the tile's ops and parameter columns in shared memory with ONE round of parallel
// loads (a lone thread chasing col_offset -> params per column pays ~1 us of dependent
// latency per column, which is the whole kernel at B = 1), then walk from shared memory.
constexpr int kComposeTPB = 128;
constexpr int kPBThreads = 64;

// Program tables small enough to ride in the kernel's parameter space (constant bank): the
// ops and the column tables are then there when the block starts, and the only cold loads
// of the kernel -- parameters, beam, energy -- are all issued at once (one DRAM round trip).
constexpr int kTabOps = 40, kTabCols = 64;  // 3.6 KB of the 4 KB classic parameter space
struct ProgramTables {
  DevOp ops[kTabOps];
  int64_t col_offset[kTabCols];
  int32_t col_stride[kTabCols];
  // shared-memory slot of column c: word slot + thread * per_thread; a column shared by all
  // settings (stride 0) has ONE slot per block (per_thread 0), a per-setting column one per
  // thread (per_thread 1)
  int2 col_slot_mask[kTabCols];
  // the per-setting columns: offset, stride and slot again, packed for the prefetch
  int64_t var_offset[kTabCols];
  int32_t var_stride[kTabCols];
  int32_t var_slot[kTabCols];
  int32_t n_var, smem_words;
};

struct SlotParamView {
  const float* vals;
  const ProgramTables& tab;
  int s;
  __device__ __forceinline__ float operator()(int col) const {
    const int2 sm = tab.col_slot_mask[col];
    return vals[sm.x + s * sm.y];
  }
};


// STAGED: 0 = ops / column tables / parameters read from global memory as the walk needs them,
// 1 = staged in shared memory from global memory, 2 = ops and column tables from the parameter
// space (ProgramTables), parameters staged in shared memory.
template <int STAGED, int OPS, bool MAPS = false>
__global__ void __launch_bounds__(kComposeTPB)
compose_kernel(ComposeArgs a, int n_cols, int S, const __grid_constant__ ProgramTables tab) {
  extern __shared__ __align__(16) float cs_smem[];
  // Programmatic dependent launch: let the particle kernel start its prologue (barrier
  // init, TMA requests for its first tiles) now; it waits (griddepcontrol.wait) for this
  // grid to finish before it reads the records or touches the zeroed image.
  asm volatile("griddepcontrol.launch_dependents;");
  if (a.hist_out && blockIdx.x == 0 && threadIdx.x == 0) {  // debug stamp (exchange trace)
#ifdef LYNX_EXPERIMENTS
    reinterpret_cast<unsigned long long*>(a.hist_out)[11] = 0ull;
    reinterpret_cast<unsigned long long*>(a.hist_out)[10] = 0ull;
#endif
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    reinterpret_cast<unsigned long long*>(a.hist_out)[8] = t;  // K1 block 0 starts
  }
  // the FIRST TWO blocks validate the screen's edge tables (x, y) and write the bin-lookup
  // geometry (they and the compose blocks are the long poles of the launch; the zeroing blocks
  // come last)
  const int hist_blocks = a.hist_out ? 2 : 0;
  if ((int)blockIdx.x < hist_blocks) {
    HistAxis* h = reinterpret_cast<HistAxis*>(a.hist_out);
    const bool y = blockIdx.x == 1;
    if (!y && a.bbox_clear && threadIdx.x < 4) a.bbox_clear[threadIdx.x] = 0;  // this call's cell
    hist_axis_setup_staged(y ? a.hist_ey : a.hist_ex, y ? a.hist_H : a.hist_W, h + (y ? 1 : 0),
                           threadIdx.x, kComposeTPB, cs_smem, y ? nullptr : a.hist_out);
#ifdef LYNX_EXPERIMENTS
    if (threadIdx.x == 0) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
      atomicMax(reinterpret_cast<unsigned long long*>(a.hist_out) + 10, t);  // histogram blocks done
    }
#endif
    return;
  }
  // blocks beyond the compose range only zero caller buffers (image / sums)
  const int first_compose = hist_blocks;
  const int compose_blocks = (int)((a.B + S - 1) / S);
  if ((int)blockIdx.x >= first_compose + compose_blocks) {
    const int zb = blockIdx.x - compose_blocks - first_compose, nzb = gridDim.x - compose_blocks - hist_blocks;
    float4* z = reinterpret_cast<float4*>(a.zero_ptr);
    const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
    if (a.zero_rect) {
      // persistent image: rows [H - v3, v2) x columns [W - v0, v1) of every setting's image
      // (the cell holds max(W - ix), max(ix + 1), max(H - iy), max(iy + 1); all 0 = nothing lit)
      const int v0 = __ldg(a.zero_rect), v1 = __ldg(a.zero_rect + 1), v2 = __ldg(a.zero_rect + 2),
                v3 = __ldg(a.zero_rect + 3);
      if (v1 <= 0 || v3 <= 0) return;
      const int r0 = a.hist_H - v3, nr = v2 - r0;
      const int c0 = (a.hist_W - v0) & ~3, nv = (min((v1 + 3) & ~3, a.hist_W) - c0) >> 2;
      const int64_t per_img = (int64_t)nr * nv, total = a.B * per_img;
      for (int64_t i = (int64_t)zb * kComposeTPB + threadIdx.x; i < total;
           i += (int64_t)nzb * kComposeTPB) {
        const int64_t b = i / per_img, rem = i - b * per_img;
        const int r = (int)(rem / nv), cv = (int)(rem - (int64_t)r * nv);
        *reinterpret_cast<float4*>(a.zero_ptr + ((b * a.hist_H + r0 + r) * (int64_t)a.hist_W + c0 +
                                                 4 * cv)) = zero;
      }
    } else {
      for (int64_t i = (int64_t)zb * kComposeTPB + threadIdx.x; i < a.zero_vec4;
           i += (int64_t)nzb * kComposeTPB)
        z[i] = zero;
    }
#ifdef LYNX_EXPERIMENTS
    if (threadIdx.x == 0 && a.hist_out) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
      atomicMax(reinterpret_cast<unsigned long long*>(a.hist_out) + 11, t);  // last zeroing block done
    }
#endif
    return;
  }
  if (blockIdx.x == first_compose) LYNX_TSTAMP(a.hist_out, 17);
  const int64_t b0 = (int64_t)(blockIdx.x - first_compose) * S;
  const int ns = (int)min((int64_t)S, a.B - b0);
  const int64_t b = b0 + threadIdx.x;
  const bool mine = (int)threadIdx.x < ns;
  DevOp* s_ops = reinterpret_cast<DevOp*>(cs_smem);
  float* s_vals = cs_smem + (STAGED == 1 ? a.n_ops * (sizeof(DevOp) / sizeof(float)) : 0);
  // every cold load of the block is issued before anything waits: the seed centre and the energy
  // of this thread's setting, then the tile's parameters (after a flush of L2 each first touch
  // of a buffer is a DRAM and page-table miss of a microsecond or more; chained, they were
  // most of K1 at B = 1)
  RecordSinkT<MAPS> sink;
  float energy = 0.f;
  if (mine) {
#pragma unroll
    for (int i = 0; i < 6; ++i)
      sink.ctr[i] = a.seed ? __ldg(a.seed + b * a.seed_row_stride + i * a.seed_ld) : 0.f;
    energy = __ldg(a.energy_in + b);
  }
  if (STAGED == 1) {
    for (int i = threadIdx.x; i < a.n_ops * (int)(sizeof(DevOp) / 4); i += kComposeTPB)
      cs_smem[i] = reinterpret_cast<const float*>(a.ops)[i];
  }
  if (STAGED) {
    for (int i = threadIdx.x; i < n_cols * ns; i += kComposeTPB) {
      const int c = i / ns, sl = i - c * ns;
      const int64_t off = STAGED == 2 ? tab.col_offset[c] : __ldg(a.col_offset + c);
      const int64_t st = STAGED == 2 ? tab.col_stride[c] : __ldg(a.col_stride + c);
      s_vals[c * S + sl] = __ldg(a.params + off + (b0 + sl) * st);
    }
    __syncthreads();
  }
  if (!mine) return;
  if (b == 0) LYNX_TSTAMP(a.hist_out, 15);
  sink.rec = a.records + b * a.rec_floats;
  sink.maps_out = a.maps_out ? a.maps_out + b * (int64_t)a.n_stages * 49 : nullptr;
  sink.off_cav = a.off_cav;
  sink.off_centre = a.off_centre;
  sink.off_screen = a.off_screen;
  constexpr bool LT = (OPS & ~kOpsLight) == 0;
  if (STAGED) {
    SmemParamView P{s_vals, S, (int)threadIdx.x};
    walk_program<OPS, LT>(STAGED == 2 ? tab.ops : s_ops, a.n_ops, a.final_stage, P, energy, sink);
  } else {
    ParamView P{a.params, a.col_offset, a.col_stride, b};
    walk_program<OPS, LT>(a.ops, a.n_ops, a.final_stage, P, energy, sink);
  }
  if (b == 0) LYNX_TSTAMP(a.hist_out, 16);
  float* fc = sink.rec + a.final_centre;
#pragma unroll
  for (int i = 0; i < 6; ++i) fc[i] = sink.ctr[i];
  if (a.energy_out) a.energy_out[b] = energy;
#ifdef LYNX_EXPERIMENTS
  if (b == 0 && a.hist_out) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    reinterpret_cast<unsigned long long*>(a.hist_out)[9] = t;  // walker of setting 0 done
  }
#endif
}

// Tile epilogue shared by the K6 kernels: rows [b0, b0 + nb) of mu_out / cov_out are contiguous,
// one bulk copy each when their byte counts and addresses are multiples of 16 (nb % 4 == 0,
// 16-byte aligned outputs), else a coalesced copy loop.  The caller owns the barrier BEFORE the
// tiles are written again (bulk: tile_release()).
__device__ __forceinline__ bool tile_store(const ParamBeamArgs& a, int64_t b0, int nb,
                                           const float* tile_mu, const float* tile_cov) {
  const bool bulk = (nb & 3) == 0 && (((uintptr_t)a.mu_out | (uintptr_t)a.cov_out) & 15) == 0;
  if (bulk) {
    fence_proxy_async();
    __syncthreads();
    if (threadIdx.x == 0) {
      bulk_s2g(a.cov_out + b0 * 49, tile_cov, (uint32_t)nb * 196u);
      bulk_s2g(a.mu_out + b0 * 7, tile_mu, (uint32_t)nb * 28u);
      bulk_commit();
    }
  } else {
    __syncthreads();
    for (int i = threadIdx.x; i < nb * 7; i += kPBThreads) a.mu_out[b0 * 7 + i] = tile_mu[i];
    for (int i = threadIdx.x; i < nb * 49; i += kPBThreads) a.cov_out[b0 * 49 + i] = tile_cov[i];
  }
  return bulk;
}
// the bulk copies of the previous tile_store have finished READING the tiles
__device__ __forceinline__ void tile_release() {
  if (threadIdx.x == 0) bulk_wait_read<0>();
  __syncthreads();
}

__device__ __forceinline__ void load_beam(const ParamBeamArgs& a, int64_t b, MomentSink& sink) {
  if (a.beam_stride == 0 && ((((uintptr_t)a.cov_in) | ((uintptr_t)a.mu_in)) & 15) == 0) {
    // one beam for all settings, 16-byte aligned: 15 broadcast loads instead of 56
    float c[52], m[8];
#pragma unroll
    for (int q = 0; q < 12; ++q) {
      const float4 v = __ldg(reinterpret_cast<const float4*>(a.cov_in) + q);
      c[4 * q] = v.x;
      c[4 * q + 1] = v.y;
      c[4 * q + 2] = v.z;
      c[4 * q + 3] = v.w;
    }
    c[48] = __ldg(a.cov_in + 48);
    const float4 m0 = __ldg(reinterpret_cast<const float4*>(a.mu_in));
    m[0] = m0.x;
    m[1] = m0.y;
    m[2] = m0.z;
    m[3] = m0.w;
#pragma unroll
    for (int i = 4; i < 7; ++i) m[i] = __ldg(a.mu_in + i);
#pragma unroll
    for (int i = 0; i < 7; ++i) sink.mu[i] = m[i];
#pragma unroll
    for (int i = 0; i < 7; ++i)
#pragma unroll
      for (int j = 0; j < 7; ++j) sink.cov[i][j] = c[i * 7 + j];
  } else {
    const float* mi = a.mu_in + b * a.beam_stride * 7;
    const float* ci = a.cov_in + b * a.beam_stride * 49;
#pragma unroll
    for (int i = 0; i < 7; ++i) sink.mu[i] = __ldg(mi + i);
#pragma unroll
    for (int i = 0; i < 7; ++i)
#pragma unroll
      for (int j = 0; j < 7; ++j) sink.cov[i][j] = __ldg(ci + i * 7 + j);
  }
  sink.bpm_out = a.bpm_out;
  sink.B = a.B;
  sink.b = b;
}

__device__ __forceinline__ void tile_fill(const MomentSink& sink, float* tile_mu, float* tile_cov) {
#pragma unroll
  for (int i = 0; i < 7; ++i) tile_mu[threadIdx.x * 7 + i] = sink.mu[i];
#pragma unroll
  for (int i = 0; i < 7; ++i)
#pragma unroll
    for (int j = 0; j < 7; ++j) tile_cov[threadIdx.x * 49 + i * 7 + j] = sink.cov[i][j];
}

// K6, the usual case: the program tables ride in the parameter space.  One thread owns one
// setting of a 64-setting tile; a block walks tiles blockIdx.x, + gridDim.x, ... and fetches the
// NEXT tile's per-setting parameters and energy into registers before it walks the current one,
// so that after the first tile no DRAM latency sits on a thread's path (the shared-beam moments
// are L1 hits, the outputs leave by bulk copy).  OPS within kOpsLight: light map (maps.cuh).
constexpr int kPrefetchCols = 8;

template <int OPS>
__global__ void __launch_bounds__(kPBThreads)
parambeam_fwd_kernel(ParamBeamArgs a, int n_cols, const __grid_constant__ ProgramTables tab) {
  constexpr bool LT = (OPS & ~kOpsLight) == 0;
  __shared__ __align__(16) float tile_cov[kPBThreads * 49];
  __shared__ __align__(16) float tile_mu[kPBThreads * 7];
  extern __shared__ __align__(16) float pb_smem[];
  float* s_vals = pb_smem;
  const int tid = threadIdx.x;
  const int64_t n_tiles = (a.B + kPBThreads - 1) / kPBThreads;
  const int n_var = tab.n_var;

  // per-setting parameters and energy of settings [b0, b0 + 64): all loads issued back to back,
  // nothing waits on them here (settings are indexed in 32 bits: B < 2^31, the launcher checks)
  float pre[kPrefetchCols], e_pre = 0.f;
  auto fetch = [&](int64_t b0) {
    const uint32_t b = (uint32_t)b0 + tid;
    if (b < (uint32_t)a.B) {
#pragma unroll
      for (int v = 0; v < kPrefetchCols; ++v)
        if (v < n_var)
          pre[v] = __ldg(a.params + (tab.var_offset[v] + (uint64_t)b * (uint32_t)tab.var_stride[v]));
      e_pre = __ldg(a.energy_in + b);
    }
  };
#pragma unroll
  for (int v = 0; v < kPrefetchCols; ++v) pre[v] = 0.f;
  int64_t tile = blockIdx.x;
  fetch(tile * kPBThreads);
  MomentSink sink;
  if (tile * kPBThreads + tid < a.B) load_beam(a, tile * kPBThreads + tid, sink);  // in flight too
  // columns shared by all settings: thread c fetches column c once for the block
  for (int c = tid; c < n_cols; c += kPBThreads)
    if (tab.col_stride[c] == 0) s_vals[tab.col_slot_mask[c].x] = __ldg(a.params + tab.col_offset[c]);
  __syncthreads();

  bool bulk_pending = false;
  for (; tile < n_tiles; tile += gridDim.x) {
    const int64_t b0 = tile * kPBThreads, b = b0 + tid;
    const bool live = b < a.B;
    const int nb = (int)min((int64_t)kPBThreads, a.B - b0);
    if (live && tile != blockIdx.x) load_beam(a, b, sink);  // L1 hits for a shared beam
    // this tile's per-setting columns: a thread's slots are read by that thread only
#pragma unroll
    for (int v = 0; v < kPrefetchCols; ++v)
      if (v < n_var) s_vals[tab.var_slot[v] + tid] = pre[v];
    float energy = e_pre;
    if (live)
      for (int v = kPrefetchCols; v < n_var; ++v)  // more per-setting columns than registers
        s_vals[tab.var_slot[v] + tid] =
            __ldg(a.params + (tab.var_offset[v] + (uint64_t)(uint32_t)b * (uint32_t)tab.var_stride[v]));
    const int64_t next_b0 = b0 + (int64_t)gridDim.x * kPBThreads;  // the next tile of this block
    if (next_b0 < a.B) fetch(next_b0);
    if (live) {
      SlotParamView P{s_vals, tab, tid};
      walk_program<OPS, LT>(tab.ops, a.n_ops, a.final_stage, P, energy, sink);
      if (a.energy_out) a.energy_out[b] = energy;
    }
    if (bulk_pending) tile_release();
    if (live) tile_fill(sink, tile_mu, tile_cov);
    bulk_pending = tile_store(a, b0, nb, tile_mu, tile_cov);
    if (!bulk_pending) __syncthreads();
  }
  if (bulk_pending && tid == 0) bulk_wait_read<0>();
}

// K6 for programs too large for the parameter space (> kTabOps ops or > kTabCols columns):
// one tile per block, ops staged in shared memory, column tables read from global memory.
__global__ void __launch_bounds__(kPBThreads) parambeam_fwd_big_kernel(ParamBeamArgs a, int n_cols) {
  __shared__ __align__(16) float tile_cov[kPBThreads * 49];
  __shared__ __align__(16) float tile_mu[kPBThreads * 7];
  extern __shared__ __align__(16) float pb_smem[];
  const int64_t b0 = (int64_t)blockIdx.x * kPBThreads;
  const int64_t b = b0 + threadIdx.x;
  const bool live = b < a.B;
  const int nb = (int)min((int64_t)kPBThreads, a.B - b0);
  float* s_vals = pb_smem;
  DevOp* s_ops = reinterpret_cast<DevOp*>(pb_smem + n_cols * kPBThreads);
  for (int i = threadIdx.x; i < a.n_ops * (int)(sizeof(DevOp) / 4); i += kPBThreads)
    reinterpret_cast<float*>(s_ops)[i] = reinterpret_cast<const float*>(a.ops)[i];
  MomentSink sink;
  float energy = 0.f;
  if (live) {
#pragma unroll 4
    for (int c = 0; c < n_cols; ++c)
      s_vals[c * kPBThreads + threadIdx.x] =
          __ldg(a.params + __ldg(a.col_offset + c) + b * (int64_t)__ldg(a.col_stride + c));
    load_beam(a, b, sink);
    energy = a.energy_in[b];
  }
  __syncthreads();
  if (live) {
    SmemParamView P{s_vals, kPBThreads, (int)threadIdx.x};
    walk_program<kOpsAll>(s_ops, a.n_ops, a.final_stage, P, energy, sink);
    if (a.energy_out) a.energy_out[b] = energy;
    tile_fill(sink, tile_mu, tile_cov);
  }
  if (tile_store(a, b0, nb, tile_mu, tile_cov) && threadIdx.x == 0) bulk_wait_read<0>();
}

}  // namespace lynx

namespace lynx {
// false: the program does not fit the parameter space (the kernels then read global tables)
static bool fill_tables(const lynx_program_t* p, ProgramTables& tab) {
  const int n_ops = (int)p->dev_ops.size(), n_cols = (int)p->col_offset.size();
  tab.n_var = tab.smem_words = 0;
  if (n_ops > kTabOps || n_cols > kTabCols) return false;
  std::copy(p->dev_ops.begin(), p->dev_ops.end(), tab.ops);
  std::copy(p->col_offset.begin(), p->col_offset.end(), tab.col_offset);
  std::copy(p->col_stride.begin(), p->col_stride.end(), tab.col_stride);
  for (int c = 0; c < n_cols; ++c) {
    const bool var = p->col_stride[c] != 0;
    tab.col_slot_mask[c] = make_int2(tab.smem_words, var ? 1 : 0);
    if (var) {
      tab.var_offset[tab.n_var] = p->col_offset[c];
      tab.var_stride[tab.n_var] = p->col_stride[c];
      tab.var_slot[tab.n_var++] = tab.smem_words;
    }
    tab.smem_words += var ? kPBThreads : 1;
  }
  return true;
}
}  // namespace lynx

using namespace lynx;

namespace lynx {
int compose_launch(const lynx_program_t* p, const float* d_params, int64_t B,
                   const float* d_energy_in, const float* d_seed_soa, int64_t seed_row_stride,
                   int64_t seed_ld, float* d_records, float* d_maps_out, float* d_energy_out,
                   void* zero_ptr, size_t zero_bytes, cudaStream_t st, const float* hist_ex,
                   const float* hist_ey, int hist_W, int hist_H, void* hist_out,
                   const int* zero_rect, int* bbox_clear) {
  if (!p || !d_params || !d_energy_in || !d_records || B <= 0) return LYNX_E_BADARG;
  if (zero_bytes && (((uintptr_t)zero_ptr & 15) || (zero_bytes & 15))) return LYNX_E_BADARG;
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
  a.zero_ptr = reinterpret_cast<float*>(zero_ptr);
  a.zero_vec4 = (int64_t)(zero_bytes / 16);
  a.hist_ex = hist_ex;
  a.hist_ey = hist_ey;
  a.hist_W = hist_W;
  a.hist_H = hist_H;
  a.hist_out = hist_out;
  a.zero_rect = zero_rect;
  a.bbox_clear = bbox_clear;
  if (zero_rect && (!hist_out || !zero_ptr || (hist_W & 3))) return LYNX_E_BADARG;
#ifdef LYNX_EXPERIMENTS  // timing decomposition of K1 (results are garbage)
  if (getenv("LYNX_K1_NOZERO")) a.zero_vec4 = 0;
  if (getenv("LYNX_K1_NOWALK")) a.n_ops = 0;
#endif
  int zero_blocks_per_sm = 2;
#ifdef LYNX_EXPERIMENTS
  if (const char* e = getenv("LYNX_K1_ZBLOCKS")) zero_blocks_per_sm = atoi(e);
#endif
  const int n_cols = (int)p->col_offset.size();
  const size_t budget = 40 * 1024;
  ProgramTables tab;
  const bool tabbed = fill_tables(p, tab);
  const size_t ops_bytes = tabbed ? 0 : a.n_ops * sizeof(DevOp);
  int S = kComposeTPB;
  bool staged = false;
  if (ops_bytes + (size_t)std::max(n_cols, 1) * 4 <= budget) {
    staged = true;
    if (n_cols > 0) S = (int)std::min<size_t>(kComposeTPB, (budget - ops_bytes) / (4 * (size_t)n_cols));
    S = (int)std::min<int64_t>(S, std::max<int64_t>(B, 1));
  }
  // a histogram block stages its edge table in the same dynamic region (16 KB)
  const size_t smem = std::max<size_t>(staged ? ops_bytes + (size_t)n_cols * S * 4 : 0,
                                       hist_out ? kHistChunk * sizeof(float) : 0);
  unsigned blocks = (unsigned)((B + S - 1) / S);
  if (zero_rect)
    blocks += (unsigned)p->sm_count;  // the rectangle's size is only known on the device
  else if (a.zero_vec4 > 0) {
    // While the compose and histogram blocks are fewer than the SMs, the launch is exactly one
    // block per SM: a zeroing block that shares an SM with the walker or the histogram block
    // keeps that SM's store path busy and holds their loads back until it is done (both sat
    // 2.5 us at their first loads, tools/k2_timeline.py) -- and they are the launch's long poles.
    int64_t zb = std::min<int64_t>(zero_blocks_per_sm * p->sm_count, (a.zero_vec4 + kComposeTPB - 1) / kComposeTPB);
    const int64_t others = (int64_t)blocks + (hist_out ? 2 : 0);
    if (others < p->sm_count) zb = std::min<int64_t>(zb, p->sm_count - others);
    blocks += (unsigned)zb;
  }
  if (hist_out) blocks += 2;
  {
    ProfileScope ps(0, st);
    const bool light = (p->op_mask & ~kOpsLight) == 0;
    auto launch = [&](auto kern) { kern<<<blocks, kComposeTPB, smem, st>>>(a, n_cols, S, tab); };
    if (d_maps_out) {  // public lynx_compose_maps: records + 7x7 stage maps
      if (staged && tabbed)
        launch(compose_kernel<2, kOpsAll, true>);
      else if (staged)
        launch(compose_kernel<1, kOpsAll, true>);
      else
        launch(compose_kernel<0, kOpsAll, true>);
    } else if (staged && tabbed && light)
      launch(compose_kernel<2, kOpsLight>);
    else if (staged && tabbed)
      launch(compose_kernel<2, kOpsAll>);
    else if (staged)
      launch(compose_kernel<1, kOpsAll>);
    else
      launch(compose_kernel<0, kOpsAll>);
  }
  count_launch();
  LYNX_CUDA_OK(cudaGetLastError());
  return LYNX_OK;
}
}  // namespace lynx

extern "C" int lynx_compose_maps(const lynx_program_t* p, const float* d_params, int64_t B,
                                 const float* d_energy_in, const float* d_seed_soa,
                                 int64_t seed_row_stride, int64_t seed_ld, float* d_records,
                                 float* d_maps_out, float* d_energy_out,
                                 lynx_stream_t stream) {
  return compose_launch(p, d_params, B, d_energy_in, d_seed_soa, seed_row_stride, seed_ld,
                        d_records, d_maps_out, d_energy_out, nullptr, 0, (cudaStream_t)stream, nullptr,
                        nullptr, 0, 0, nullptr, nullptr, nullptr);
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
  if (B >= ((int64_t)1 << 31)) return LYNX_E_TOOLARGE;  // 0.5 TB of moments
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
  const int n_cols = (int)p->col_offset.size();
  ProgramTables tab;
  const bool tabbed = fill_tables(p, tab);
  for (int c = 0; c < n_cols; ++c)
    if (p->col_stride[c] < 0) return LYNX_E_BADARG;
  const size_t smem = tabbed ? (size_t)tab.smem_words * 4
                             : (size_t)n_cols * kPBThreads * 4 + a.n_ops * sizeof(DevOp);
  if (smem > 200 * 1024) return LYNX_E_TOOLARGE;  // > 800 parameter columns
  {
    ProfileScope ps(2, (cudaStream_t)stream);
    const bool light = (p->op_mask & ~kOpsLight) == 0;
    auto launch = [&](auto kern, unsigned grid, auto... extra) {
      if (smem > 48 * 1024)
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      kern<<<grid, kPBThreads, smem, (cudaStream_t)stream>>>(a, n_cols, extra...);
    };
    // persistent above the resident blocks per SM (the register budgets of the two variants)
    const unsigned grid = std::min<unsigned>(blocks, (light ? 8u : 5u) * (unsigned)p->sm_count);
    if (tabbed && light)
      launch(parambeam_fwd_kernel<kOpsLight>, grid, tab);
    else if (tabbed)
      launch(parambeam_fwd_kernel<kOpsAll>, grid, tab);
    else
      launch(parambeam_fwd_big_kernel, blocks);
  }
  count_launch();
  LYNX_CUDA_OK(cudaGetLastError());
  return LYNX_OK;
}

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
//
// Screen histogram (K4): global RED.ADD into the image runs at ~170 G updates/s for a
// uniform footprint but collapses with focus (100 G/s at sigma = 100 px, 48 G/s at 20 px,
// measured, tools/micro/atomic_bench.cu) because L2 serialises updates per line.  With an
// active screen the ring shrinks to 2 stages and the freed shared memory holds a
// CTA-private WINDOW of the image (16-bit counters, two per word) placed on the beam from
// the moments of the CTA's first tile; in-window particles cost one shared-memory atomic,
// the rest go to global REDs, and the window is flushed with coalesced REDs (32 lanes ->
// 64 consecutive pixels) at the end / before a counter could overflow.
#include <algorithm>
#include <cstdio>
#include <cstdlib>

#include "compose_common.cuh"
#include "track_common.cuh"
#include "pipe.cuh"

namespace lynx {

constexpr int kPipeTile = 2048;                    // particles per stage
constexpr int kConsumerWarps = 16;
constexpr int kConsumers = kConsumerWarps * 32;    // 512 threads x 4 particles = tile
constexpr int kPipeThreads = kConsumers + 64;      // + producer warp + storer warp
constexpr int kStageFloats = 6 * kPipeTile;

struct PipeSmemLayout {
  // byte offsets from the (128-B aligned) dynamic shared memory base
  uint32_t rec, plan, bpm, bars, win, ops, vals, total;
};

__host__ __device__ inline PipeSmemLayout pipe_layout(int stages, int rec_floats, int n_plan,
                                                      int n_snap, uint32_t win_words,
                                                      int fused_ops = 0, int fused_cols = 0) {
  PipeSmemLayout L;
  uint32_t o = (uint32_t)(sizeof(float) * kStageFloats * stages);
  L.rec = o;
  o += 4u * rec_floats;
  o = (o + 15u) & ~15u;
  L.plan = o;
  o += 16u * n_plan;
  L.bpm = o;
  o += 16u * n_snap;
  o = (o + 7u) & ~7u;
  L.bars = o;
  o += 8u * 3 * stages;
  o = (o + 15u) & ~15u;
  L.win = o;
  o += 4u * win_words;
  o = (o + 15u) & ~15u;
  L.ops = o;  // fused composer: the program and this setting's parameter columns
  o += (uint32_t)sizeof(DevOp) * fused_ops;
  L.vals = o;
  o += 4u * fused_cols;
  L.total = o;
  return L;
}

// CTA-private image window (bin coordinates): [x0, x0+w) x [y0, y0+h), w even, two 16-bit
// counters per word.
struct HistWindow {
  int x0, y0, w, h;
};

__device__ __forceinline__ void consumers_sync() { asm volatile("bar.sync 1, %0;" ::"n"(kConsumers) : "memory"); }

// Histogram update as an explicit reduction: in a kernel that also contains acquire / release
// synchronisation (the fused variant's grid barrier) nvcc lowers a result-less atomicAdd to
// ATOMG instead of REDG, which cost 35 % of the 32 M-particle screen kernel (measured).
__device__ __forceinline__ void red_add_f32(float* p, float v) {
  asm volatile("red.relaxed.gpu.global.add.f32 [%0], %1;" ::"l"(p), "f"(v));
}

__device__ __forceinline__ void dbg_stamp(const TrackArgs& a, int slot, bool who) {
  if ((a.dbg & 2) && who && blockIdx.x == 0 && blockIdx.y == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    reinterpret_cast<unsigned long long*>(const_cast<HistAxis*>(a.hist))[16 + slot] = t;
  }
}

// OR-reduction over the consumer threads (named barrier 1).
__device__ __forceinline__ bool consumers_or(bool pred) {
  int r;
  asm volatile(
      "{ .reg .pred p, q; setp.ne.b32 q, %1, 0; barrier.red.or.pred p, 1, %2, q; selp.b32 %0, 1, 0, p; }"
      : "=r"(r)
      : "r"((int)pred), "n"(kConsumers)
      : "memory");
  return r != 0;
}

// Grid-wide barrier for the fused variant, executed by ONE thread per CTA (the caller brackets
// it with CTA-level barriers).  All CTAs of the launch are co-resident (grid <= SM count, one
// CTA per SM, cooperative launch).  bar[0] counts arrivals and is reset by the last arriver,
// bar[1] is the generation the waiters spin on: the state is valid again when the launch ends.
__device__ __forceinline__ void grid_barrier_one_thread(unsigned int* bar, unsigned int n_ctas) {
  unsigned int gen;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(gen) : "l"(bar + 1) : "memory");
  __threadfence();  // this CTA's stores (ordered before by the CTA barrier) before the arrival
  const unsigned int old = atomicAdd(bar, 1u);
  if (old == n_ctas - 1) {
    bar[0] = 0u;
    __threadfence();
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(bar + 1), "r"(gen + 1u) : "memory");
  } else {
    unsigned int g;
    do {
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(g) : "l"(bar + 1) : "memory");
    } while (g == gen);
  }
}

// Flush the window into the image with coalesced REDs and clear it (consumer threads).
__device__ __forceinline__ void window_flush(uint32_t* __restrict__ win, const HistWindow& hw,
                                             float* __restrict__ img_b, int W, int H, int cid) {
  const int wpr = hw.w >> 1;  // words per window row
  const int n = wpr * hw.h;
  for (int i = cid; i < n; i += kConsumers) {
    const uint32_t v = win[i];
    if (v) {
      win[i] = 0u;
      const int r = i / wpr, c = i - r * wpr;
      float* p = img_b + (int64_t)(H - 1 - (hw.y0 + r)) * W + hw.x0 + 2 * c;
      const uint32_t lo = v & 0xffffu, hi = v >> 16;
      if (lo) atomicAdd(p, (float)lo);
      if (hi) atomicAdd(p + 1, (float)hi);
    }
  }
}

// The composer of the fused variant: K1's walk (compose_common.cuh) for one setting, straight
// into the shared-memory record.  Not inlined: one thread per CTA runs it once, and its
// register needs (127 in K1) must not leak into the allocation of the streaming loop.
struct FusedWalkArgs {
  const DevOp* ops;
  const float* vals;
  float* rec;
  const float* seed;  // particle 0 of the row: plane i at seed[i * ld]
  int64_t ld;
  const float* energy_in;
  float* energy_out;  // nullable
  int n_ops, final_stage, final_centre, off_cav, off_centre, off_screen;
};

template <int OPS>
__device__ __noinline__ void fused_walk(FusedWalkArgs w) {
  RecordSink sink;
  sink.rec = w.rec;
  sink.maps_out = nullptr;
  sink.off_cav = w.off_cav;
  sink.off_centre = w.off_centre;
  sink.off_screen = w.off_screen;
#pragma unroll
  for (int i = 0; i < 6; ++i) sink.ctr[i] = __ldg(w.seed + i * w.ld);
  float energy = __ldg(w.energy_in);
  SmemParamView P{w.vals, 1, 0};
  walk_program<OPS>(w.ops, w.n_ops, w.final_stage, P, energy, sink);
#pragma unroll
  for (int i = 0; i < 6; ++i) w.rec[w.final_centre + i] = sink.ctr[i];
  if (w.energy_out) *w.energy_out = energy;
}

// OPS == 0: the per-setting record comes from K1 (compose.cu) through global memory.
// OPS != 0 (opcode set, walk.cuh): FUSED composer -- one thread of every CTA walks the CTA's
// setting into the shared-memory record while the first tiles are in flight and the consumer
// warps zero the CTA's slice of the image and check the edge tables; one grid barrier orders
// the zeroing before the first RED.  At 1 M particles the K1 launch, its completion and the
// record round trip were a third of the step.
template <int MOM, bool HAS_OUT, int STAGES, int OPS = 0>
__global__ void __launch_bounds__(kPipeThreads, 1) track_pipe_kernel(TrackArgs a, uint32_t win_words) {
  extern __shared__ __align__(128) unsigned char psm[];
  constexpr bool kFused = OPS != 0;
  if (kFused) dbg_stamp(a, 10, threadIdx.x == 0);  // first instruction of CTA 0
  const PipeSmemLayout L = pipe_layout(STAGES, a.rec_floats, a.n_plan, a.n_snap, win_words,
                                       kFused ? a.n_ops : 0, kFused ? a.n_cols : 0);
  float* stages = reinterpret_cast<float*>(psm);
  float* rec = reinterpret_cast<float*>(psm + L.rec);
  PlanItem* plan = reinterpret_cast<PlanItem*>(psm + L.plan);
  double* s_bpm = reinterpret_cast<double*>(psm + L.bpm);
  uint64_t* full = reinterpret_cast<uint64_t*>(psm + L.bars);
  uint64_t* empty = full + STAGES;
  uint64_t* done = empty + STAGES;
  uint32_t* win = reinterpret_cast<uint32_t*>(psm + L.win);
  __shared__ double s_red[kConsumerWarps][28];
  __shared__ float s_wstat[kConsumerWarps][5];
  __shared__ float s_chk[kConsumerWarps];  // sink of the load-completion dependency (see below)
  __shared__ int s_bb[4];                  // CTA bounding box of the touched bins
  if (threadIdx.x < 4) s_bb[threadIdx.x] = 0;  // ordered before its use by the CTA barriers below
  __shared__ HistWindow s_win;

  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  const int64_t b = blockIdx.y;
  // a.tile <= kPipeTile particles per tile (chosen by the launcher so that every CTA gets
  // the same number of tiles); the shared-memory plane stride stays kPipeTile
  const int64_t tile = a.tile;
  const int64_t n_tiles = (a.N + tile - 1) / tile;
  const float* in_b = a.in + b * a.in_row_stride;

  auto issue_tile = [&](int it, int64_t t) {
    const int s = it % STAGES;
    const int64_t p0 = t * tile;
    const uint32_t cnt = (uint32_t)min(tile, a.N - p0);
    const uint32_t bytes = ((cnt + 3u) & ~3u) * 4u;
    mbar_arrive_expect_tx(full + s, 6u * bytes);
    float* dst = stages + s * kStageFloats;
#pragma unroll
    for (int k = 0; k < 6; ++k)
      bulk_g2s(dst + k * kPipeTile, in_b + k * a.ld_in + p0, bytes, full + s);
  };

  // ---- prologue: barriers, then this CTA's first tiles are requested before anything else
  //      touches memory.  Thread 0 initialises the barriers it then arms itself; every other
  //      thread first touches them after a CTA-wide barrier below.
  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full + s, 1);
      mbar_init(empty + s, HAS_OUT ? 1 : kConsumerWarps);
      mbar_init(done + s, kConsumerWarps);
    }
    fence_mbar_init();
    int it = 0;
    for (int64_t t = blockIdx.x; t < n_tiles && it < STAGES; t += gridDim.x, ++it)
      issue_tile(it, t);
    s_win = HistWindow{0, 0, 0, 0};
  }
  if constexpr (!kFused) {
    // ---- everyone: plan, cleared window (independent of K1) ...
    for (int i = tid; i < a.n_plan; i += kPipeThreads) plan[i] = a.plan[i];
    for (int i = tid; i < a.n_snap * 2; i += kPipeThreads) s_bpm[i] = 0.0;
    for (uint32_t i = tid; i < win_words; i += kPipeThreads) win[i] = 0u;
  }
  __shared__ HistAxis s_hist[2];
  if constexpr (!kFused) {
    // ... then wait for the K1 grid (records + histogram geometry written, image zeroed) --
    // a no-op unless this kernel was launched with programmatic stream serialization
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (a.stamp && tid == 0 && blockIdx.x == 0 && b == 0) {  // exchange trace (exchange.cu)
      unsigned long long t;
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
      a.stamp[0] = t;  // K1 complete, K2 starts
      a.stamp[2] = reinterpret_cast<const unsigned long long*>(a.hist)[8];  // K1 block 0 started
    }
    const float* src = a.records + b * a.rec_floats;
    for (int i = tid; i < a.rec_floats; i += kPipeThreads) rec[i] = src[i];
    if (a.img && tid < 2) s_hist[tid] = a.hist[tid];
    __syncthreads();
  } else {
    if (a.stamp && tid == 0 && blockIdx.x == 0 && b == 0) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
      a.stamp[0] = t;
      a.stamp[2] = t;
    }
    dbg_stamp(a, 0, tid == 0);
    if ((a.dbg & 2) && tid == 0) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
      unsigned long long* d = reinterpret_cast<unsigned long long*>(const_cast<HistAxis*>(a.hist));
      atomicMin(d + 16 + 8, t);
      atomicMax(d + 16 + 9, t);
    }
    // Three independent strands run concurrently from the first instruction on; one CTA
    // barrier joins them:
    //   thread 0   barriers + first tiles (above)
    //   warp 1     program + this setting's parameter columns -> shared memory, then the walk
    //   consumers  plan, BPM accumulators, image slice zeroed, edge tables checked, grid barrier
    DevOp* s_ops = reinterpret_cast<DevOp*>(psm + L.ops);
    float* s_vals = reinterpret_cast<float*>(psm + L.vals);
    if (warp == 1) {
      for (int i = lane; i < a.n_ops * (int)(sizeof(DevOp) / 4); i += 32)
        reinterpret_cast<float*>(s_ops)[i] = reinterpret_cast<const float*>(a.ops)[i];
      for (int c = lane; c < a.n_cols; c += 32)
        s_vals[c] = __ldg(a.params + __ldg(a.col_offset + c) + b * (int64_t)__ldg(a.col_stride + c));
      __syncwarp();
      dbg_stamp(a, 1, lane == 0);
      // the walk (K1's arithmetic, compose_common.cuh), straight into shared memory
      if (lane == 0) {
        FusedWalkArgs w;
        w.ops = s_ops;
        w.vals = s_vals;
        w.rec = rec;
        w.seed = in_b;
        w.ld = a.ld_in;
        w.energy_in = a.energy_in + b;
        w.energy_out = (blockIdx.x == 0 && a.energy_out) ? a.energy_out + b : nullptr;
        w.n_ops = a.n_ops;
        w.final_stage = a.final_stage;
        w.final_centre = a.final_centre;
        w.off_cav = a.off_cav;
        w.off_centre = a.off_centre;
        w.off_screen = a.off_screen;
        fused_walk<OPS>(w);
        dbg_stamp(a, 2, true);
      }
    } else if (warp >= 2) {
      const int cid = tid - 64;
      for (int i = cid; i < a.n_plan; i += kConsumers) plan[i] = a.plan[i];
      for (int i = cid; i < a.n_snap * 2; i += kConsumers) s_bpm[i] = 0.0;
      if (a.img) {
        if (a.zero_img) {
          float4* z = reinterpret_cast<float4*>(a.img + b * (int64_t)a.W * a.H);
          const int64_t n4 = ((int64_t)a.W * a.H) >> 2;
          const int64_t per = (n4 + gridDim.x - 1) / gridDim.x;
          const int64_t lo = blockIdx.x * per, hi = min(n4, lo + per);
          const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
          for (int64_t i = lo + cid; i < hi; i += kConsumers) z[i] = zero;
          if (a.bbox && blockIdx.x == 0 && b == 0 && cid < 4) a.bbox[cid] = 0;
        }
#pragma unroll
        for (int ax = 0; ax < 2; ++ax) {
          const float* e = ax ? a.ey : a.ex;
          const int n = ax ? a.H : a.W;
          const double e0 = (double)__ldg(e), en = (double)__ldg(e + n);
          const double step = (en - e0) / n;
          int bad = 0;
          for (int i = cid; i <= n; i += kConsumers)
            bad |= fabs((double)__ldg(e + i) - (e0 + step * i)) > step * (1.0 / 1024.0);
          const bool any_bad = consumers_or(bad != 0);
          if (cid == 0) {  // same arithmetic as hist_axis_setup (track_common.cuh)
            s_hist[ax].e0 = (float)e0;
            s_hist[ax].en = (float)en;
            s_hist[ax].inv = (float)n / ((float)en - (float)e0);
            s_hist[ax].margin = any_bad ? 1.f : fmaxf(1.f / 256.f, (float)n * (1.f / 1048576.f));
          }
        }
        dbg_stamp(a, 3, cid == 0);
        if (a.zero_img) {  // every CTA's slice is zero before anyone's first RED
          consumers_sync();
          if (cid == 0) grid_barrier_one_thread(a.grid_bar, gridDim.x * gridDim.y);
        }
        dbg_stamp(a, 4, cid == 0);
      }
    }
    __syncthreads();
  }

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
        if (it < STAGES) continue;  // issued in the prologue
        const uint32_t u = (uint32_t)(it / STAGES);
        mbar_wait(empty + (it % STAGES), (u - 1) & 1);
        issue_tile(it, t);
      }
    }
  } else if (warp == 1) {
    // ---- storer: finished tiles leave through the async proxy
    if (HAS_OUT && lane == 0) {
      float* out_b = a.out + b * 6 * a.ld_out;
      int it = 0, prev_s = -1;
      for (int64_t t = blockIdx.x; t < n_tiles; t += gridDim.x, ++it) {
        const int s = it % STAGES;
        const uint32_t u = (uint32_t)(it / STAGES);
        mbar_wait(done + s, u & 1);
        fence_proxy_async();
        const int64_t p0 = t * tile;
        const uint32_t cnt = (uint32_t)min(tile, a.N - p0);
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
    HistAxis hx{0.f, 0.f, 0.f, 1.f}, hy{0.f, 0.f, 0.f, 1.f};
    // replicated histogram: the hot bins of a focused beam sit in a few
    // hundred L2 lines whose slices saturate; CTA x adds into copy x % n_rep
    const int rep_i = a.n_rep > 1 ? (int)(blockIdx.x % (unsigned)a.n_rep) : 0;
    float* img_b = !a.img ? nullptr
                   : rep_i == 0 ? a.img + b * (int64_t)a.W * a.H
                                : a.rep + ((int64_t)(rep_i - 1) * a.B + b) * (int64_t)a.W * a.H;
    if (img_b) {
      hx = s_hist[0];
      hy = s_hist[1];
    }
    constexpr bool kWin = (STAGES == 2);  // window code only in the screen variants
    const bool use_win = kWin && img_b && win_words > 0;
    HistWindow hw{0, 0, 0, 0};
    bool win_set = false;
    int since_flush = 0;
    int bb0 = 0, bb1 = 0, bb2 = 0, bb3 = 0;  // bounding box of the touched bins (exchange.cu)
    int it = 0;
    for (int64_t t = blockIdx.x; t < n_tiles; t += gridDim.x, ++it) {
      const int s = it % STAGES;
      const uint32_t u = (uint32_t)(it / STAGES);
      const int64_t p0 = t * tile;
      const int cnt = (int)min(tile, a.N - p0);
      n_mine += cnt;
      float* st = stages + s * kStageFloats + cid * 4;
      mbar_wait(full + s, u & 1);
      if (it == 0) dbg_stamp(a, 5, cid == 0);  // first tile landed
      float v[6][4];
#pragma unroll
      for (int k = 0; k < 6; ++k) {
        const float4 q = *reinterpret_cast<const float4*>(st + k * kPipeTile);
        v[k][0] = q.x;
        v[k][1] = q.y;
        v[k][2] = q.z;
        v[k][3] = q.w;
      }
      if (!HAS_OUT) {
        // The stage can be refilled as soon as the values are in registers -- but ISSUING the
        // shared-memory loads is not enough: the arrive below is not ordered behind their
        // completion, and the loader's next bulk copy then overwrote a stage some warps had not
        // read yet (those threads tracked particles of tile it + STAGES twice and dropped their
        // own: images differed from run to run once a CTA had more than STAGES tiles, i.e.
        // above ~1.3 M particles; sums were unaffected).  A store that depends on one word of
        // every 16-byte load forces the scoreboard wait: it cannot issue before all six loads
        // have returned, and the arrive is issued behind it.
        const float chk = ((v[0][0] + v[1][0]) + (v[2][0] + v[3][0])) + (v[4][0] + v[5][0]);
        __syncwarp();
        if (lane == 0) {
          *reinterpret_cast<volatile float*>(&s_chk[warp - 2]) = chk;
          mbar_arrive(empty + s);
        }
      }
      bool live[4];
#pragma unroll
      for (int l = 0; l < 4; ++l) live[l] = cid * 4 + l < cnt;

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
          apply_block2(rec + p.a, v);  // the cavity's own map is block-diagonal
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
          int pk[4];  // ix | iy << 16, or -1 (dead lane / outside the screen)
          float c = 0.f, sx = 0.f, sy = 0.f, sxx = 0.f, syy = 0.f;
#pragma unroll
          for (int l = 0; l < 4; ++l) {
            const float x = v[0][l] - mx;
            const float y = yfix ? v[2][l] - my : v[2][l];
            const int ix = find_bin(x, ex, a.W, hx);
            const int iy = find_bin(y, ey, a.H, hy);
            const bool ok = live[l] && ix >= 0 && iy >= 0;
            pk[l] = ok ? (ix | (iy << 16)) : -1;
            if (ok) {
              bb0 = max(bb0, a.W - ix);
              bb1 = max(bb1, ix + 1);
              bb2 = max(bb2, a.H - iy);
              bb3 = max(bb3, iy + 1);
            }
            if (kWin && !win_set && ok) {
              const float dx = (float)(ix - (a.W >> 1)), dy = (float)(iy - (a.H >> 1));
              c += 1.f;
              sx += dx;
              sy += dy;
              sxx = fmaf(dx, dx, sxx);
              syy = fmaf(dy, dy, syy);
            }
          }
          if constexpr (kWin) {
            if (use_win && !win_set) {
              // place the window on the beam: mean +- k sigma of this CTA's first tile
              c = warp_sum(c);
              sx = warp_sum(sx);
              sy = warp_sum(sy);
              sxx = warp_sum(sxx);
              syy = warp_sum(syy);
              if (lane == 0) {
                float* d = s_wstat[warp - 2];
                d[0] = c; d[1] = sx; d[2] = sy; d[3] = sxx; d[4] = syy;
              }
              consumers_sync();
              if (cid == 0) {
                float tc = 0.f, tx = 0.f, ty = 0.f, txx = 0.f, tyy = 0.f;
                for (int w = 0; w < kConsumerWarps; ++w) {
                  tc += s_wstat[w][0]; tx += s_wstat[w][1]; ty += s_wstat[w][2];
                  txx += s_wstat[w][3]; tyy += s_wstat[w][4];
                }
                HistWindow nw{0, 0, 0, 0};
                const float cap = (float)(2u * win_words);
                if (tc >= 32.f) {
                  const float mxb = tx / tc, myb = ty / tc;
                  const float sgx = fmaxf(sqrtf(fmaxf(txx / tc - mxb * mxb, 0.f)), 1.f);
                  const float sgy = fmaxf(sqrtf(fmaxf(tyy / tc - myb * myb, 0.f)), 1.f);
                  const float k = fminf(sqrtf(cap / (4.f * sgx * sgy)), 4.5f);
                  int w2 = min(2 * ((int)ceilf(k * sgx) + 1), a.W & ~1);
                  int h2 = min(2 * ((int)ceilf(k * sgy) + 1), a.H);
                  if ((float)w2 * (float)h2 > cap) h2 = (int)(cap / (float)w2);
                  if (h2 > 0 && w2 > 0) {
                    nw.w = w2;
                    nw.h = h2;
                    nw.x0 = min(max((int)lrintf(mxb) + (a.W >> 1) - (w2 >> 1), 0), a.W - w2);
                    nw.y0 = min(max((int)lrintf(myb) + (a.H >> 1) - (h2 >> 1), 0), a.H - h2);
                  }
                }
                s_win = nw;
              }
              consumers_sync();
              hw = s_win;
              win_set = true;
            }
            if (use_win && since_flush + kPipeTile > 65535) {  // a 16-bit counter could overflow
              consumers_sync();
              window_flush(win, hw, img_b, a.W, a.H, cid);
              consumers_sync();
              since_flush = 0;
            }
            since_flush += kPipeTile;
          }
          const int wpr = hw.w >> 1;
#pragma unroll
          for (int l = 0; l < 4; ++l)
            if (pk[l] >= 0) {
              const int ix = pk[l] & 0xffff, iy = pk[l] >> 16;
              const unsigned dx = (unsigned)(ix - hw.x0), dy = (unsigned)(iy - hw.y0);
              if (a.dbg & 1) {
                if (ix + iy == -12345) img_b[0] = 1.f;
              } else if (kWin && dx < (unsigned)hw.w && dy < (unsigned)hw.h) {
                atomicAdd(win + dy * wpr + (dx >> 1), 1u << ((dx & 1u) * 16u));
              } else {
                red_add_f32(img_b + (int64_t)(a.H - 1 - iy) * a.W + ix, 1.0f);
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
    if constexpr (kWin) {
      if (use_win && win_set) {
        consumers_sync();
        window_flush(win, hw, img_b, a.W, a.H, cid);
      }
    }
    if (a.bbox && img_b) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        bb0 = max(bb0, __shfl_xor_sync(0xffffffffu, bb0, o));
        bb1 = max(bb1, __shfl_xor_sync(0xffffffffu, bb1, o));
        bb2 = max(bb2, __shfl_xor_sync(0xffffffffu, bb2, o));
        bb3 = max(bb3, __shfl_xor_sync(0xffffffffu, bb3, o));
      }
      // one set of global atomics per CTA, not per warp: 16 x 148 x 4 atomicMax on the same
      // four words were ~3 us of the 1 M-particle step (serialised in one L2 slice)
      if (lane == 0 && bb1 > 0) {
        atomicMax(&s_bb[0], bb0);
        atomicMax(&s_bb[1], bb1);
        atomicMax(&s_bb[2], bb2);
        atomicMax(&s_bb[3], bb3);
      }
      consumers_sync();
      if (cid < 4 && s_bb[1] > 0) atomicMax(a.bbox + cid, s_bb[cid]);
    }
    // Let a programmatically-serialized follower (histogram fold, multi-GPU exchange kernel)
    // start launching once the consumers of every CTA are through their tiles.  The first
    // thread of a CTA to execute this counts for the CTA, so it must not sit on a path the
    // loader warp reaches early; issued at the top of the kernel the follower's CTAs (which
    // only wait for this grid) became resident before some K2 CTAs were placed.
    dbg_stamp(a, 6, cid == 0);
    asm volatile("griddepcontrol.launch_dependents;");
  }
  // ---- CTA epilogue
  if (a.stamp && tid == 64 && blockIdx.x == 0 && b == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    a.stamp[1] = t;  // CTA 0's last RED issued
  }
  if (MOM != 0) {
    mom_epilogue<MOM, kConsumerWarps>(acc, s_red, warp >= 2 ? warp - 2 : -1, lane, tid == 64,
                                      n_mine, ctr, a.mom + b * 28);
  }
  if (a.n_snap > 0) {
    __syncthreads();
    // n_mine is only tracked by consumers; recompute for the flushing threads
    int64_t n_cta = 0;
    for (int64_t t = blockIdx.x; t < n_tiles; t += gridDim.x)
      n_cta += min(tile, a.N - t * tile);
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
  if ((a.dbg & 2) && tid == 64) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    atomicMax(reinterpret_cast<unsigned long long*>(const_cast<HistAxis*>(a.hist)) + 16 + 7, t);
  }
}

// Folds the histogram replicas into the image over the bounding box of the touched bins and
// leaves the replicas zero again (their invariant between calls).  Launched behind K2 with
// programmatic stream serialization.
constexpr int kFoldTPB = 256;
__global__ void __launch_bounds__(kFoldTPB) hist_fold_kernel(float* __restrict__ img,
                                                              float* __restrict__ rep, int n_rep,
                                                              int64_t B, int W, int H,
                                                              const int* __restrict__ bbox) {
  asm volatile("griddepcontrol.launch_dependents;");
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const int v0 = __ldcg(bbox), v1 = __ldcg(bbox + 1), v2 = __ldcg(bbox + 2), v3 = __ldcg(bbox + 3);
  if (v1 == 0 || v3 == 0) return;
  const int r0 = H - v3, r1 = v2, c0 = (W - v0) & ~3, c1 = min((v1 + 3) & ~3, W);
  const int w4 = (c1 - c0) >> 2, nr = r1 - r0;
  const int64_t n = B * (int64_t)nr * w4, plane = B * (int64_t)W * H;
  const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int64_t t = (int64_t)blockIdx.x * kFoldTPB + threadIdx.x; t < n;
       t += (int64_t)gridDim.x * kFoldTPB) {
    const int64_t rr = t / w4;
    const int c = c0 + ((int)(t - rr * w4) << 2);
    const int64_t bq = rr / nr;
    const int r = r0 + (int)(rr - bq * nr);
    const int64_t i = (bq * H + r) * (int64_t)W + c;
    // every load of a batch is issued before the first store: with one 16-B load in flight
    // per thread the pass was latency-bound at 0.3 TB/s (46 us for 14 MB, measured)
    float4 s = *reinterpret_cast<const float4*>(img + i);
    for (int q0 = 0; q0 < n_rep - 1; q0 += 8) {
      float4 v[8];
#pragma unroll
      for (int q = 0; q < 8; ++q)
        v[q] = (q0 + q < n_rep - 1) ? __ldcg(reinterpret_cast<const float4*>(rep + (q0 + q) * plane + i))
                                    : zero;
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        s.x += v[q].x;
        s.y += v[q].y;
        s.z += v[q].z;
        s.w += v[q].w;
      }
#pragma unroll
      for (int q = 0; q < 8; ++q)
        if (q0 + q < n_rep - 1) *reinterpret_cast<float4*>(rep + (q0 + q) * plane + i) = zero;
    }
    *reinterpret_cast<float4*>(img + i) = s;
  }
}

template <int MOM, bool HAS_OUT, int STAGES, int OPS = 0>
static int launch_pipe(const TrackArgs& a, dim3 grid, size_t smem, uint32_t win_words,
                       cudaStream_t st) {
  auto k = track_pipe_kernel<MOM, HAS_OUT, STAGES, OPS>;
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
  if (OPS != 0) {
    // fused composer: nothing to overlap with; the grid barrier needs every CTA resident
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = 1;
  } else {
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
  }
  cfg.attrs = attr;
  static const bool no_coop = [] { const char* e = getenv("LYNX_FUSED_NOCOOP"); return e && e[0] == '1'; }();
  cfg.numAttrs = (OPS != 0 && no_coop) ? 0 : 1;
  cudaError_t err;
  {
    ProfileScope ps(1, st);
    err = cudaLaunchKernelEx(&cfg, k, a, win_words);
    static const bool no_fold = [] { const char* e = getenv("LYNX_NO_FOLD"); return e && e[0] == '1'; }();
    if (err == cudaSuccess && a.n_rep > 1 && !no_fold) {  // the fold is part of the histogram's cost
      int dev = 0, sms = 148;
      cudaGetDevice(&dev);
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
      cfg.gridDim = dim3((unsigned)(2 * sms));
      cfg.blockDim = dim3(kFoldTPB);
      cfg.dynamicSmemBytes = 0;
      attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      attr[0].val.programmaticStreamSerializationAllowed = 1;
      cfg.numAttrs = 1;
      err = cudaLaunchKernelEx(&cfg, hist_fold_kernel, a.img, a.rep, a.n_rep, a.B, a.W, a.H,
                               (const int*)a.bbox);
      count_launch();
    }
  }
  count_launch();
  return err == cudaSuccess ? LYNX_OK : LYNX_E_LAUNCH;
}

// Returns LYNX_OK after enqueueing, or 1 if this variant cannot take the problem (the
// caller then uses the LDG kernel).
int launch_track_pipe(const TrackArgs& a, int moment_mode, bool has_out, int sm_count,
                      cudaStream_t st, int fused_ops) {
  const uint32_t cap = 227 * 1024 - 4096 - 1024;  // static shared memory takes the rest
  const bool hist = a.img != nullptr && !has_out;
  // The shared-memory window is OPT-IN (LYNX_WINDOW=1): measured on B200 it loses to plain
  // global REDs unless the beam is focused below sigma ~ 10 px -- random scatter into shared
  // memory costs ~7 cycles/lane/SM (bank conflicts) against 1.29 for REDG, so it only pays
  // where L2 line contention pushes global REDs below ~40 G updates/s.
  static const bool no_win = [] { const char* e = getenv("LYNX_WINDOW"); return !(e && e[0] == '1'); }();
  const int stages = (hist && !no_win && !fused_ops) ? 2 : 4;
  PipeSmemLayout L = pipe_layout(stages, a.rec_floats, a.n_plan, a.n_snap, 0,
                                 fused_ops ? a.n_ops : 0, fused_ops ? a.n_cols : 0);
  if (L.total > cap) return 1;
  uint32_t win_words = 0;
  if (stages == 2) {
    win_words = ((cap - L.total) / 4u) & ~3u;
    if (win_words < 1024) win_words = 0;
    L = pipe_layout(stages, a.rec_floats, a.n_plan, a.n_snap, win_words);
  }
  // the fused variant synchronises the grid: every CTA must be resident (one per SM)
  int64_t per_b = fused_ops ? sm_count / a.B : (sm_count + a.B - 1) / a.B;
  if (per_b < 1) {
    if (fused_ops) return 1;
    per_b = 1;
  }
  // Tile size: the largest multiple of 32 particles <= kPipeTile that gives every CTA of a
  // row the same number of tiles (1 M particles on 148 CTAs: 4 x 1696 each instead of 3 or
  // 4 x 2048 -- the longest CTA sets the kernel time).
  const int64_t full_tiles = (a.N + kPipeTile - 1) / kPipeTile;
  const int64_t gx = std::min<int64_t>(full_tiles, per_b);
  const int64_t rounds = (a.N + gx * kPipeTile - 1) / (gx * kPipeTile);
  int64_t tile = (a.N + gx * rounds - 1) / (gx * rounds);
  tile = std::min<int64_t>(kPipeTile, (tile + 31) & ~(int64_t)31);
  const int64_t n_tiles = (a.N + tile - 1) / tile;
  TrackArgs at = a;
  at.tile = (int)tile;
  dim3 grid((unsigned)std::min<int64_t>(n_tiles, gx), (unsigned)a.B);
  if (stages == 2) {
    switch (moment_mode) {
      case 0: return launch_pipe<0, false, 2>(at, grid, L.total, win_words, st);
      case 1: return launch_pipe<1, false, 2>(at, grid, L.total, win_words, st);
      default: return launch_pipe<2, false, 2>(at, grid, L.total, win_words, st);
    }
  }
  const int v = moment_mode * 2 + (has_out ? 1 : 0);
  if (fused_ops == kOpsLight) {
    switch (v) {
      case 0: return launch_pipe<0, false, 4, kOpsLight>(at, grid, L.total, 0, st);
      case 1: return launch_pipe<0, true, 4, kOpsLight>(at, grid, L.total, 0, st);
      case 2: return launch_pipe<1, false, 4, kOpsLight>(at, grid, L.total, 0, st);
      case 3: return launch_pipe<1, true, 4, kOpsLight>(at, grid, L.total, 0, st);
      case 4: return launch_pipe<2, false, 4, kOpsLight>(at, grid, L.total, 0, st);
      default: return launch_pipe<2, true, 4, kOpsLight>(at, grid, L.total, 0, st);
    }
  }
  if (fused_ops) {
    switch (v) {
      case 0: return launch_pipe<0, false, 4, kOpsAll>(at, grid, L.total, 0, st);
      case 1: return launch_pipe<0, true, 4, kOpsAll>(at, grid, L.total, 0, st);
      case 2: return launch_pipe<1, false, 4, kOpsAll>(at, grid, L.total, 0, st);
      case 3: return launch_pipe<1, true, 4, kOpsAll>(at, grid, L.total, 0, st);
      case 4: return launch_pipe<2, false, 4, kOpsAll>(at, grid, L.total, 0, st);
      default: return launch_pipe<2, true, 4, kOpsAll>(at, grid, L.total, 0, st);
    }
  }
  switch (v) {
    case 0: return launch_pipe<0, false, 4>(at, grid, L.total, 0, st);
    case 1: return launch_pipe<0, true, 4>(at, grid, L.total, 0, st);
    case 2: return launch_pipe<1, false, 4>(at, grid, L.total, 0, st);
    case 3: return launch_pipe<1, true, 4>(at, grid, L.total, 0, st);
    case 4: return launch_pipe<2, false, 4>(at, grid, L.total, 0, st);
    default: return launch_pipe<2, true, 4>(at, grid, L.total, 0, st);
  }
}

}  // namespace lynx

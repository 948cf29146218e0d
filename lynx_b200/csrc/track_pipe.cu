// K2, Blackwell-native variant: persistent warp-specialised CTA (one per SM) with a
// TMA bulk-copy pipeline.
//
//   warp 0      producer: one lane issues cp.async.bulk global->shared for the six
//               coordinate planes of a tile (6 x 8 KB), completion on an mbarrier
//   warp 1      storer:   cp.async.bulk shared->global of finished tiles (bulk groups)
//   warps 2-17  consumers: LDS.128 a float4 of particles per plane, apply the whole
//               lattice in registers, write back in place / histogram / moments
//
// Variants without outgoing particles (terminal screen, moments only) have no producer and no
// storer warp: a stage is free as soon as its sixteen consumer warps hold the tile in registers,
// and the LAST of them to say so (a shared-memory arrival counter) issues the bulk copies of the
// tile that goes into the stage next.  16 warps instead of 18 means 128 registers per thread
// instead of 96 (registers are handed out per four warps), which the histogram code needs.
//
// In-flight bytes are set by shared memory (a ring of 48 KB stages), not by registers x
// occupancy, so the whole 1M-particle beam is requested in the first microsecond and large
// beams stream at copy speed.  SASS: UBLKCP (bulk copies), SYNCS (mbarrier).  Same arithmetic
// and outputs as track_fwd_kernel (track.cu).
//
// Screen histogram (K4).  One global RED.ADD.F32 per particle runs at the focused-beam atomic
// rate of the L2 (64-75 G updates/s for a sigma = 40 px spot, profiles/r2_hist_bench_c.txt): 32 M
// particles cost 430 us where the 24 B/particle stream needs 118 us.  A shared-memory atomic
// costs 7-8 cycles per WARP instruction whatever its addresses (0.23 cycles per lane,
// profiles/r2_hist_bench_b.txt) -- so every CTA keeps a private u32 TILE of the image in shared
// memory, placed on the beam from the statistics of the CTA's first pipeline tile (mean +- k
// sigma, k as large as the tile's capacity allows).  In-tile particles cost one ATOMS, the tails
// go to global memory as before (few, and spread out: no hot L2 lines), and at the end every
// thread converts its quads of the tile to fp32 and adds the non-empty ones into the image with
// one 16-byte vector reduction each (red.global.add.v4.f32).  A first version converted in
// place and issued one cp.reduce.async.bulk per tile row: a bulk operation is issued per
// thread through uniform registers, the rows' issue loop and the extra barrier + async-proxy
// fence cost 0.8 us of the 1 M-particle step more than the vector reductions do (no difference
// at 32 M).  The ring gives the tile its shared memory: 2-4 stages, chosen by the launcher.
// Counts are integers, so the image is bit-identical whatever the placement.
//
// Fast path (a.xy_only): when the plan is one merged stage followed by the terminal screen and
// nothing else reads the beam (ARES EA + AREABSCR1, the headline config), only rows x and y of
// the stage map are applied: 12 FMAs per particle instead of 36.
#include <algorithm>
#include <cstdio>
#include <cstdlib>

#include "compose_common.cuh"
#include "track_common.cuh"
#include "pipe.cuh"

namespace lynx {

// Consumer warps per CTA are a template parameter CW (each thread carries 4 particles, so a
// pipeline tile is 128 particles per consumer warp).  The moment-free screen variant needs 80
// registers per thread and runs 24 warps -- six per scheduler instead of four, for a kernel bound
// by instruction issue latency; the moment-free outgoing-particles variant (90 -> 80 registers)
// runs 20 + 2 role warps; the moment variants (96-120 registers) run 16.
__host__ __device__ constexpr int pipe_threads(int cw, bool has_out) { return cw * 32 + (has_out ? 64 : 0); }
constexpr int kMaxStages = 4;

struct PipeSmemLayout {
  // byte offsets from the (128-B aligned) dynamic shared memory base
  uint32_t rec, plan, bpm, bars, win, total;
};

__host__ __device__ inline PipeSmemLayout pipe_layout(int stages, int stage_floats, int rec_floats,
                                                      int n_plan, int n_snap, uint32_t win_words) {
  PipeSmemLayout L;
  uint32_t o = (uint32_t)(sizeof(float) * stage_floats * stages);
  L.rec = o;
  o += 4u * rec_floats;
  o = (o + 15u) & ~15u;
  L.plan = o;
  o += 16u * n_plan;
  L.bpm = o;
  o += 16u * n_snap;
  o = (o + 7u) & ~7u;
  L.bars = o;
  o += 8u * 3 * kMaxStages;
  o = (o + 127u) & ~127u;
  L.win = o;
  o += 4u * win_words;
  L.total = o;
  return L;
}

// CTA-private image tile (bin coordinates): [x0, x0+w) x [y0, y0+h), one u32 counter per bin,
// row-major by iy; x0 and w are multiples of 4 (16-byte quads for the vector reduction).
struct HistWindow {
  int x0, y0, w, h;
};

__device__ __forceinline__ float sqrt_approx(float x) {
  float r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
template <int THREADS>
__device__ __forceinline__ void consumers_sync_n() { asm volatile("bar.sync 1, %0;" ::"n"(THREADS) : "memory"); }

// Histogram update as an explicit reduction (nvcc lowers a result-less atomicAdd to ATOMG
// instead of REDG in kernels that also contain acquire / release code: -35 %, found in SASS).
__device__ __forceinline__ void red_add_f32(float* p, float v) {
  asm volatile("red.relaxed.gpu.global.add.f32 [%0], %1;" ::"l"(p), "f"(v));
}
__device__ __forceinline__ unsigned vmax_u16x2(unsigned a, unsigned b) {
  unsigned r;
  asm("max.u16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ unsigned vmin_u16x2(unsigned a, unsigned b) {
  unsigned r;
  asm("min.u16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ void reds_add_u32(uint32_t smem_addr, uint32_t v) {
  asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(smem_addr), "r"(v) : "memory");
}

#ifdef LYNX_EXPERIMENTS
// phase stamps of CTA `g_k2_trace_cta`'s first consumer thread (tools/k2_timeline.py)
__device__ unsigned long long g_k2_trace[16];
__device__ int g_k2_trace_cta = 0;
#define K2_STAMP(i)                                                              \
  do {                                                                           \
    if ((int)blockIdx.x == g_k2_trace_cta && threadIdx.x == 32 * (HAS_OUT ? 2 : 0)) { \
      unsigned long long t_;                                                     \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                     \
      g_k2_trace[i] = t_;                                                        \
    }                                                                            \
  } while (0)
#else
#define K2_STAMP(i) do {} while (0)
#endif

// TILE: the screen variant with the shared-memory histogram tile (win_words > 0).
template <int MOM, bool HAS_OUT, bool TILE, int CW>
__global__ void __launch_bounds__(pipe_threads(CW, HAS_OUT), 1)
track_pipe_kernel(TrackArgs a, uint32_t win_words) {
  extern __shared__ __align__(128) unsigned char psm[];
  constexpr int kConsumerWarps = CW;
  constexpr int kConsumers = kConsumerWarps * 32;
  constexpr int kPipeTile = kConsumers * 4;     // particles per stage (capacity and plane stride)
  constexpr int kStageFloats = 6 * kPipeTile;
  constexpr int kPipeThreads = pipe_threads(kConsumerWarps, HAS_OUT);
  auto consumers_sync = [] { consumers_sync_n<kConsumers>(); };
  constexpr int kRoleWarps = HAS_OUT ? 2 : 0;  // producer + storer
  const int STAGES = a.stages;
  const PipeSmemLayout L = pipe_layout(STAGES, kStageFloats, a.rec_floats, a.n_plan, a.n_snap, win_words);
  float* stages = reinterpret_cast<float*>(psm);
  float* rec = reinterpret_cast<float*>(psm + L.rec);
  PlanItem* plan = reinterpret_cast<PlanItem*>(psm + L.plan);
  double* s_bpm = reinterpret_cast<double*>(psm + L.bpm);
  uint64_t* full = reinterpret_cast<uint64_t*>(psm + L.bars);
  uint64_t* empty = full + kMaxStages;
  uint64_t* done = empty + kMaxStages;
  uint32_t* win = reinterpret_cast<uint32_t*>(psm + L.win);
  __shared__ double s_red[kConsumerWarps][28];
  __shared__ int s_stat[3];                   // tile placement: count, sum dx, sum dy of the first tile
  __shared__ unsigned long long s_stat2[2];   // ... sum dx^2, sum dy^2
  __shared__ float s_chk[kConsumerWarps];  // sink of the load-completion dependency (see below)
  __shared__ int s_bb[4];                  // CTA bounding box of the touched bins
  if (threadIdx.x < 4) s_bb[threadIdx.x] = 0;  // ordered before its use by the CTA barriers below
  __shared__ HistWindow s_win;
  __shared__ unsigned int s_freed[kMaxStages];  // !HAS_OUT: consumer warps done with the stage
  __shared__ HistAxis s_hist[2];

  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  K2_STAMP(0);
  // grid: a.ctas_per_row consecutive CTAs share a setting (no 65 535 limit of gridDim.y)
  const int64_t b = blockIdx.x / (unsigned)a.ctas_per_row;
  const int cta = (int)(blockIdx.x - b * a.ctas_per_row), n_cta = a.ctas_per_row;
  // a.tile <= kPipeTile particles per tile (chosen by the launcher so that every CTA gets
  // the same number of tiles); the shared-memory plane stride stays kPipeTile
  const int64_t tile = a.tile;
  const int64_t n_tiles = (a.N + tile - 1) / tile;
  const float* in_b = a.in + b * a.in_row_stride;

  auto issue_tile = [&](int s, int64_t t) {
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
      mbar_init(empty + s, 1);
      mbar_init(done + s, kConsumerWarps);
    }
    fence_mbar_init();
    int it = 0;
    for (int64_t t = cta; t < n_tiles && it < STAGES; t += n_cta, ++it) issue_tile(it, t);
    s_win = HistWindow{0, 0, 0, 0};
    s_stat[0] = s_stat[1] = s_stat[2] = 0;
    s_stat2[0] = s_stat2[1] = 0ull;
    for (int s = 0; s < kMaxStages; ++s) s_freed[s] = 0u;
  }
  // ---- everyone: plan, cleared tile (independent of K1) ...
  if (a.n_plan <= 8) {  // from the kernel's parameter space: no cold load
    if (tid < a.n_plan) plan[tid] = a.plan_inline[tid];
  } else {
    for (int i = tid; i < a.n_plan; i += kPipeThreads) plan[i] = a.plan[i];
  }
  for (int i = tid; i < a.n_snap * 2; i += kPipeThreads) s_bpm[i] = 0.0;
  if (TILE) {
    uint4* w4 = reinterpret_cast<uint4*>(win);
    for (uint32_t i = tid; i < (win_words >> 2); i += kPipeThreads) w4[i] = make_uint4(0u, 0u, 0u, 0u);
  }
  // ... then wait for the K1 grid (records + histogram geometry written, image zeroed) --
  // a no-op unless this kernel was launched with programmatic stream serialization
  K2_STAMP(1);
  asm volatile("griddepcontrol.wait;" ::: "memory");
  K2_STAMP(2);
  if (a.stamp && tid == 0 && blockIdx.x == 0) {  // exchange trace (exchange.cu)
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    a.stamp[0] = t;  // K1 complete, K2 starts
    a.stamp[2] = reinterpret_cast<const unsigned long long*>(a.hist)[8];  // K1 block 0 started
  }
  {
    const float* src = a.records + b * a.rec_floats;
    for (int i = tid; i < a.rec_floats; i += kPipeThreads) rec[i] = src[i];
    if (a.img && tid < 2) s_hist[tid] = a.hist[tid];
  }
  __syncthreads();
  K2_STAMP(3);

  MomAcc<MOM> acc;
#pragma unroll
  for (int i = 0; i < MomAcc<MOM>::kN; ++i) acc.a[i] = 0.f;
  float ctr[6];
#pragma unroll
  for (int i = 0; i < 6; ++i) ctr[i] = rec[a.final_centre + i];
  int64_t n_mine = 0;

  if (HAS_OUT && warp == 0) {
    // ---- producer: keep the ring full
    if (lane == 0) {
      int it = 0, s = 0;
      uint32_t u = 0;  // it / STAGES
      for (int64_t t = cta; t < n_tiles; t += n_cta, ++it) {
        if (it >= STAGES) {  // the first STAGES tiles were issued in the prologue
          mbar_wait(empty + s, (u - 1) & 1);
          issue_tile(s, t);
        }
        if (++s == STAGES) {
          s = 0;
          ++u;
        }
      }
    }
  } else if (HAS_OUT && warp == 1) {
    // ---- storer: finished tiles leave through the async proxy
    if (lane == 0) {
      float* out_b = a.out + b * 6 * a.ld_out;
      int s = 0, prev_s = -1;
      uint32_t u = 0;
      for (int64_t t = cta; t < n_tiles; t += n_cta) {
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
        if (++s == STAGES) {
          s = 0;
          ++u;
        }
      }
      bulk_wait_read<0>();
      if (prev_s >= 0) mbar_arrive(empty + prev_s);
    }
  } else {
    // ---- consumers: 512 threads x 4 particles = one tile
    const int cid = tid - 32 * kRoleWarps, cw = warp - kRoleWarps;
    float* img_b = a.img ? a.img + b * (int64_t)a.W * a.H : nullptr;
    // the bin geometry (s_hist) and the tile's placement (s_win) stay in shared memory and are
    // re-read where they are used (registers are the scarce resource here)
    const bool use_win = TILE && img_b && win_words > 0;
    bool win_set = false;
    // bounding box of the touched bins (exchange.cu, persistent image), two u16 lanes per word:
    // bb_max = max(ix + 1) | max(iy) << 16 over valid particles (0: none), bb_min = min(ix) | min(iy) << 16
    unsigned bb_max = 0u, bb_min = 0xffffffffu;

    // Screen.track + Screen.reading for four particles (screen.py:126-141 + 195-211)
    auto screen4 = [&](const PlanItem& p, const float (&xv)[4], const float (&yv)[4],
                       const bool (&live)[4]) {
      const float mx = rec[p.a], my = rec[p.a + 1];
      const bool yfix = (p.c & LYNX_F_SCREEN_Y_FIX) != 0;
      const HistAxis hx = s_hist[0], hy = s_hist[1];
      // one fp32 multiplication each: hist_axis_finish measured the error of exactly this value
      const float offx = __fmul_rn(-hx.e0, hx.inv), offy = __fmul_rn(-hy.e0, hy.inv);
      const bool regular = hx.margin < 0.5f && hy.margin < 0.5f;  // warp-uniform
      int pk[4];  // ix | iy << 16, or -1 (dead lane / outside the screen)
      if (regular) {
        float xs[4], ys[4], frx[4], fry[4];
        int ix[4], iy[4];
        bool nx[4], ny[4], any_near = false;
#pragma unroll
        for (int l = 0; l < 4; ++l) {
          xs[l] = xv[l] - mx;
          ys[l] = yfix ? yv[l] - my : yv[l];
          ix[l] = bin_guess(xs[l], hx.inv, offx, hx.margin, hx.e0, hx.en, nx[l], frx[l]);
          iy[l] = bin_guess(ys[l], hy.inv, offy, hy.margin, hy.e0, hy.en, ny[l], fry[l]);
          any_near = any_near || nx[l] || ny[l];
        }
        if (any_near) {  // a few lookups in a thousand: the edge table decides
#pragma unroll
          for (int l = 0; l < 4; ++l) {
            if (nx[l]) ix[l] = bin_near_edge(xs[l], a.ex, a.W, ix[l], frx[l]);
            if (ny[l]) iy[l] = bin_near_edge(ys[l], a.ey, a.H, iy[l], fry[l]);
          }
        }
#pragma unroll
        for (int l = 0; l < 4; ++l) {
          const bool ok = live[l] && (ix[l] | iy[l]) >= 0;
          pk[l] = ok ? (ix[l] | (iy[l] << 16)) : -1;
        }
      } else {
#pragma unroll
        for (int l = 0; l < 4; ++l) {
          const int ix = find_bin(xv[l] - mx, a.ex, a.W, hx);
          const int iy = find_bin(yfix ? yv[l] - my : yv[l], a.ey, a.H, hy);
          pk[l] = (live[l] && (ix | iy) >= 0) ? (ix | (iy << 16)) : -1;
        }
      }
#pragma unroll
      for (int l = 0; l < 4; ++l) {
        // invalid = 0xffffffff: neutral for the min; + 1 makes it 0, neutral for the max (a valid
        // ix + 1 <= 65535 does not carry into the iy lane; W, H < 32768 is checked by the launcher)
        bb_max = vmax_u16x2(bb_max, (unsigned)pk[l] + 1u);
        bb_min = vmin_u16x2(bb_min, (unsigned)pk[l]);
      }
      if constexpr (TILE) {
        if (use_win && !win_set) {  // warp-uniform: the CTA's first pipeline tile only
          K2_STAMP(11);
          // place the tile on the beam: mean +- k sigma of this CTA's first pipeline tile.
          // Exact integer sums of the bin offsets (|dx|, |dy| < 2^15): warp reductions by
          // redux.sync (the squares as two 16-bit halves, 4 x 2^28 per thread does not survive
          // a 32-lane sum in 32 bits), then one shared-memory atomic per warp and statistic.
          int c = 0, sx = 0, sy = 0;
          unsigned int sxx = 0u, syy = 0u;
#pragma unroll
          for (int l = 0; l < 4; ++l)
            if (pk[l] >= 0) {
              const int dx = (pk[l] & 0xffff) - (a.W >> 1), dy = (pk[l] >> 16) - (a.H >> 1);
              c += 1;
              sx += dx;
              sy += dy;
              sxx += (unsigned)(dx * dx);
              syy += (unsigned)(dy * dy);
            }
          c = __reduce_add_sync(0xffffffffu, c);
          sx = __reduce_add_sync(0xffffffffu, sx);
          sy = __reduce_add_sync(0xffffffffu, sy);
          const unsigned long long qx =
              ((unsigned long long)__reduce_add_sync(0xffffffffu, sxx >> 16) << 16) +
              __reduce_add_sync(0xffffffffu, sxx & 0xffffu);
          const unsigned long long qy =
              ((unsigned long long)__reduce_add_sync(0xffffffffu, syy >> 16) << 16) +
              __reduce_add_sync(0xffffffffu, syy & 0xffffu);
          if (lane == 0) {
            atomicAdd(&s_stat[0], c);
            atomicAdd(&s_stat[1], sx);
            atomicAdd(&s_stat[2], sy);
            atomicAdd(&s_stat2[0], qx);
            atomicAdd(&s_stat2[1], qy);
          }
          consumers_sync();
          K2_STAMP(12);
          if (cid == 0) {
            // (approximate reciprocals and roots: the placement only has to be sensible, and one
            // thread decides it for the CTA)
            const long long ci = s_stat[0], sxi = s_stat[1], syi = s_stat[2];
            const float tc = (float)(int)ci;
            const float itc = __fdividef(1.f, tc);
            // n^2 var = n sum d^2 - (sum d)^2, exact in 64 bits (n <= 2^12, sum d^2 < 2^40)
            const float vx = (float)(ci * (long long)s_stat2[0] - sxi * sxi) * (itc * itc);
            const float vy = (float)(ci * (long long)s_stat2[1] - syi * syi) * (itc * itc);
            HistWindow nw{0, 0, 0, 0};
            const float cap = (float)win_words;
            if (ci >= 32) {
              const float mxb = (float)(int)sxi * itc, myb = (float)(int)syi * itc;
              const float sgx = fmaxf(sqrt_approx(fmaxf(vx, 0.f)), 0.5f);
              const float sgy = fmaxf(sqrt_approx(fmaxf(vy, 0.f)), 0.5f);
              const float k = fminf(sqrt_approx(__fdividef(cap, 4.f * sgx * sgy)), 5.f);
              int w2 = min((2 * (int)ceilf(k * sgx) + 5) & ~3, a.W);
              int h2 = min(2 * (int)ceilf(k * sgy) + 2, a.H);
              if ((unsigned)(w2 * h2) > win_words) h2 = (int)(win_words / (unsigned)w2);  // exact: the tile's size
              if (h2 > 0 && w2 > 0) {
                nw.w = w2;
                nw.h = h2;
                nw.x0 = min(max(__float2int_rn(mxb) + (a.W >> 1) - (w2 >> 1), 0), a.W - w2) & ~3;
                nw.y0 = min(max(__float2int_rn(myb) + (a.H >> 1) - (h2 >> 1), 0), a.H - h2);
              }
            }
            s_win = nw;
          }
          consumers_sync();
          K2_STAMP(13);
          win_set = true;
        }
      }
      const HistWindow hw = s_win;
      const uint32_t win_s = smem_u32(win);
#pragma unroll
      for (int l = 0; l < 4; ++l) {
        const bool ok = pk[l] >= 0;
        const int ix = pk[l] & 0xffff, iy = pk[l] >> 16;
        bool in = false;
        if constexpr (TILE) {
          const unsigned dx = (unsigned)(ix - hw.x0), dy = (unsigned)(iy - hw.y0);
          in = ok && dx < (unsigned)hw.w && dy < (unsigned)hw.h;
          if (in) reds_add_u32(win_s + 4u * (dy * (unsigned)hw.w + dx), 1u);
        }
        if (ok && !in) red_add_f32(img_b + (unsigned)((a.H - 1 - iy) * a.W + ix), 1.0f);  // W H < 2^30
      }
    };

    int s = 0;
    uint32_t u = 0;
    for (int64_t t = cta; t < n_tiles; t += n_cta) {
      const int64_t p0 = t * tile;
      const int cnt = (int)min(tile, a.N - p0);
      n_mine += cnt;
      float* st = stages + s * kStageFloats + cid * 4;
      mbar_wait(full + s, u & 1);
      if (t == cta) K2_STAMP(4);
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
        // shared-memory loads is not enough: nothing orders what follows behind their
        // completion, and a bulk copy issued too early overwrote a stage some warps had not read
        // yet (round 1: images differed from run to run once a CTA had more than STAGES tiles;
        // sums were unaffected).  A store that depends on one word of every 16-byte load forces
        // the scoreboard wait: it cannot issue before all six loads have returned, and the
        // arrival is issued behind it.  The warp whose arrival is the sixteenth owns the stage:
        // it re-arms the stage's barrier and requests the tile that goes there next.
        const float chk = ((v[0][0] + v[1][0]) + (v[2][0] + v[3][0])) + (v[4][0] + v[5][0]);
        __syncwarp();
        if (lane == 0) {
          *reinterpret_cast<volatile float*>(&s_chk[cw]) = chk;
          unsigned int old;
          asm volatile("atom.acq_rel.cta.shared::cta.add.u32 %0, [%1], 1;"
                       : "=r"(old) : "r"(smem_u32(&s_freed[s])) : "memory");
          if (old == kConsumerWarps - 1) {
            *reinterpret_cast<volatile unsigned int*>(&s_freed[s]) = 0u;  // nobody arrives again before the new tile lands
            const int64_t tn = t + (int64_t)STAGES * n_cta;
            if (tn < n_tiles) issue_tile(s, tn);
          }
        }
      }
      bool live[4];
#pragma unroll
      for (int l = 0; l < 4; ++l) live[l] = cid * 4 + l < cnt;

      bool done_fast = false;
      if constexpr (!HAS_OUT && MOM == 0) {
        if (a.xy_only) {  // plan = [AFFINE, SCREEN]: only x and y are needed
          float xv[4], yv[4];
          apply_affine_xy(rec + plan[0].a, v, xv, yv);
          screen4(plan[1], xv, yv, live);
          done_fast = true;
        }
      }
      if (!done_fast)
      for (int pi = 0; pi < a.n_plan; ++pi) {
        const PlanItem p = plan[pi];
        if (p.kind == P_AFFINE) {
          if constexpr (!HAS_OUT && MOM == 0) {
            if (pi == a.xy_tail) {  // the last stage in front of the terminal screen: x, y rows only
              float xv[4], yv[4];
              apply_affine_xy(rec + p.a, v, xv, yv);
              screen4(plan[pi + 1], xv, yv, live);
              break;
            }
          }
          if (rec[p.a + kMapFloats] != 0.f)  // K1 found three 2x2 blocks (compose_common.cuh write_map)
            apply_block2(rec + p.a, v);
          else
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
        } else if (img_b) {  // P_SCREEN
          screen4(p, v[0], v[2], live);
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
      if (t == cta) K2_STAMP(5);
      if (++s == STAGES) {
        s = 0;
        ++u;
      }
    }
    K2_STAMP(6);
    if constexpr (TILE) {
      const HistWindow hw = s_win;
      if (use_win && win_set && hw.w > 0) {
        // flush: every thread converts its quads of u32 counts and adds the non-empty ones
        // straight into the image with ONE 16-byte vector reduction each (row iy of the tile is
        // image row H - 1 - iy: flipud(hist.T), screen.py:206; x0 and w are multiples of 4 and
        // the image rows 16-byte aligned, so a quad never straddles a row)
        consumers_sync();
        const int w4n = hw.w >> 2, n4 = w4n * hw.h;
        const uint4* w4 = reinterpret_cast<const uint4*>(win);
        for (int i = cid; i < n4; i += kConsumers) {
          const uint4 c = w4[i];
          if ((c.x | c.y | c.z | c.w) != 0u) {
            const int r = i / w4n, q = i - r * w4n;
            float* dst = img_b + (int64_t)(a.H - 1 - (hw.y0 + r)) * a.W + hw.x0 + 4 * q;
            asm volatile("red.relaxed.gpu.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst),
                         "f"((float)c.x), "f"((float)c.y), "f"((float)c.z), "f"((float)c.w)
                         : "memory");
          }
        }
        K2_STAMP(7);
      }
    }
    K2_STAMP(8);
    if (a.bbox && img_b) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        bb_max = vmax_u16x2(bb_max, __shfl_xor_sync(0xffffffffu, bb_max, o));
        bb_min = vmin_u16x2(bb_min, __shfl_xor_sync(0xffffffffu, bb_min, o));
      }
      // cell layout: max(W - ix), max(ix + 1), max(H - iy), max(iy + 1); all 0 = nothing lit
      const bool any = (bb_max & 0xffffu) != 0u;
      const int bb0 = any ? a.W - (int)(bb_min & 0xffffu) : 0, bb1 = (int)(bb_max & 0xffffu);
      const int bb2 = any ? a.H - (int)(bb_min >> 16) : 0, bb3 = any ? (int)(bb_max >> 16) + 1 : 0;
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
    K2_STAMP(10);
    // Let a programmatically-serialized follower (multi-GPU exchange kernel) start launching
    // once the consumers of every CTA are through their tiles.  The first thread of a CTA to
    // execute this counts for the CTA, so it must not sit on a path the loader warp reaches
    // early; issued at the top of the kernel the follower's CTAs (which only wait for this
    // grid) became resident before some K2 CTAs were placed.
    asm volatile("griddepcontrol.launch_dependents;");
  }
  // ---- CTA epilogue
  if (a.stamp && tid == 32 * kRoleWarps && blockIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    a.stamp[1] = t;  // CTA 0's last RED issued
  }
  if (MOM != 0) {
    mom_epilogue<MOM, kConsumerWarps>(acc, s_red, warp >= kRoleWarps ? warp - kRoleWarps : -1, lane,
                                      tid == 32 * kRoleWarps, n_mine, ctr, a.mom + b * 28);
  }
  if (a.n_snap > 0) {
    __syncthreads();
    // n_mine is only tracked by consumers; recompute for the flushing threads
    int64_t n_own = 0;
    for (int64_t t = cta; t < n_tiles; t += n_cta) n_own += min(tile, a.N - t * tile);
    for (int pi = tid; pi < a.n_plan; pi += kPipeThreads) {
      const PlanItem p = plan[pi];
      if (p.kind == P_BPM) {
        const double n = (double)n_own;
        double* g = a.bpm_sums + ((int64_t)p.a * a.B + b) * 2;
        atomicAdd(g, s_bpm[2 * p.c] + n * (double)rec[p.b]);
        atomicAdd(g + 1, s_bpm[2 * p.c + 1] + n * (double)rec[p.b + 2]);
      }
    }
  }
}

template <int MOM, bool HAS_OUT, bool TILE, int CW = 16>
static int launch_pipe(const TrackArgs& a, unsigned grid, size_t smem, uint32_t win_words,
                       cudaStream_t st) {
  auto k = track_pipe_kernel<MOM, HAS_OUT, TILE, CW>;
  static bool attr_set = false;  // per instantiation
  if (!attr_set) {
    if (cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024 - 4096) !=
        cudaSuccess)
      return LYNX_E_LAUNCH;
    attr_set = true;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(pipe_threads(CW, HAS_OUT));
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
    err = cudaLaunchKernelEx(&cfg, k, a, win_words);
  }
  count_launch();
  return err == cudaSuccess ? LYNX_OK : LYNX_E_LAUNCH;
}

#ifdef LYNX_EXPERIMENTS
}  // namespace lynx
extern "C" int lynx_dbg_k2_trace(unsigned long long* host16, int cta) {
  if (cudaMemcpyFromSymbol(host16, lynx::g_k2_trace, sizeof(lynx::g_k2_trace)) != cudaSuccess) return -1;
  return cudaMemcpyToSymbol(lynx::g_k2_trace_cta, &cta, sizeof(int)) == cudaSuccess ? 0 : -1;
}
namespace lynx {
static int env_int(const char* name, int dflt) {
  const char* e = getenv(name);
  return e ? atoi(e) : dflt;
}
#endif

// Returns LYNX_OK after enqueueing, or 1 if this variant cannot take the problem (the
// caller then uses the LDG kernel).
// cw consumer warps; returns -100 when more than 16 were asked for a screen without a histogram
// tile (the wider screen kernel exists only as the tiled variant): the caller retries with 16.
static int launch_track_pipe_cw(const TrackArgs& a, int moment_mode, bool has_out, int cw,
                                int sm_count, cudaStream_t st) {
  const int kPipeTile = cw * 128;
  const uint32_t kStageFloats = 6u * (uint32_t)kPipeTile;
  const uint32_t cap = 227 * 1024 - 4096 - 1024;  // static shared memory takes the rest
  const bool hist = a.img != nullptr && !has_out;
  if (a.img && (a.W >= 32768 || a.H >= 32768)) return 1;  // bins travel as two 15-bit lanes
  // grid: every CTA of a row gets the same number of tiles (1 M particles on 148 CTAs with 24
  // consumer warps: 3 x 2272 each instead of 2 or 3 x 3072 -- the longest CTA sets the kernel time)
  int64_t per_b = (sm_count + a.B - 1) / a.B;
  if (per_b < 1) per_b = 1;
  const int64_t full_tiles = (a.N + kPipeTile - 1) / kPipeTile;
  const int64_t gx = std::min<int64_t>(full_tiles, per_b);
  const int64_t rounds = (a.N + gx * kPipeTile - 1) / (gx * kPipeTile);
  int64_t tile = (a.N + gx * rounds - 1) / (gx * rounds);
  tile = std::min<int64_t>(kPipeTile, (tile + 31) & ~(int64_t)31);
  const int64_t n_tiles = (a.N + tile - 1) / tile;
  const int64_t ctas_per_row = std::min<int64_t>(n_tiles, gx);
  if (ctas_per_row * a.B > 0x7fffffffLL) return 1;

  // Histogram tile (see the head of this file): the image rows must be 16-byte aligned for the
  // vector reductions.  Its size follows the particles a CTA will see -- flushing costs time per
  // tile byte, the atomics it saves come per particle -- and the ring shrinks to make room.
  int stages = 4;
  uint32_t win_words = 0;
  const uint32_t fixed = pipe_layout(0, 0, a.rec_floats, a.n_plan, a.n_snap, 0).total;
  if (fixed + 2u * kStageFloats * 4u > cap) return 1;
  if (hist && !(a.W & 3) && !((uintptr_t)a.img & 15) && (((int64_t)a.W * a.H) & 3) == 0) {
    // words of tile per particle of the CTA: with the per-row bulk-reduce flush 1-2 was best at
    // 1 M particles (27.6 us against 29.6 at 3 and 31.7 at 6); with the vector-reduction flush,
    // which skips empty quads, 1 / 2 / 4 are within noise at 1, 2, 4 and 8 M particles
    // (profiles/r2j_k2_timeline.txt)
    int px_per_particle = 1, min_stages = 2;
#ifdef LYNX_EXPERIMENTS
    int64_t max_words = 1 << 30;
    px_per_particle = env_int("LYNX_TILE_PXPP", px_per_particle);
    min_stages = env_int("LYNX_TILE_MIN_STAGES", min_stages);
    max_words = env_int("LYNX_TILE_MAX_WORDS", 1 << 30);
#endif
    const int64_t n_cta = (a.N + ctas_per_row - 1) / ctas_per_row;
    int64_t want = std::min<int64_t>(n_cta * px_per_particle, (int64_t)a.W * a.H);
#ifdef LYNX_EXPERIMENTS
    want = std::min<int64_t>(want, max_words);
#endif
    if (want >= 4096) {
      for (stages = 4; stages > min_stages; --stages) {
        const uint32_t used = fixed + (uint32_t)stages * kStageFloats * 4u;
        if (used < cap && (int64_t)((cap - used) / 4u) >= want) break;
      }
      const uint32_t used = fixed + (uint32_t)stages * kStageFloats * 4u;
      if (used + 4096u * 4u <= cap) win_words = (uint32_t)std::min<int64_t>(want, (cap - used) / 4u) & ~3u;
    }
  }
  if (!win_words) {
    stages = 4;
    while (stages > 2 && fixed + (uint32_t)stages * kStageFloats * 4u > cap) --stages;
  }
  if (cw != 16 && !win_words && !has_out) return -100;
  const PipeSmemLayout L = pipe_layout(stages, (int)kStageFloats, a.rec_floats, a.n_plan, a.n_snap, win_words);
  if (L.total > cap) return 1;
  TrackArgs at = a;
  at.tile = (int)tile;
  at.stages = stages;
  at.ctas_per_row = (int)ctas_per_row;
  const unsigned grid = (unsigned)(ctas_per_row * a.B);
  if (win_words) {
    switch (moment_mode) {
      case 0:
        if (cw == 24) return launch_pipe<0, false, true, 24>(at, grid, L.total, win_words, st);
#ifdef LYNX_EXPERIMENTS
        if (cw == 28) return launch_pipe<0, false, true, 28>(at, grid, L.total, win_words, st);
#endif
        return launch_pipe<0, false, true>(at, grid, L.total, win_words, st);
      case 1: return launch_pipe<1, false, true>(at, grid, L.total, win_words, st);
      default: return launch_pipe<2, false, true>(at, grid, L.total, win_words, st);
    }
  }
  switch (moment_mode * 2 + (has_out ? 1 : 0)) {
    case 0: return launch_pipe<0, false, false>(at, grid, L.total, 0, st);
    case 1:
      if (cw == 20) return launch_pipe<0, true, false, 20>(at, grid, L.total, 0, st);
      return launch_pipe<0, true, false>(at, grid, L.total, 0, st);
    case 2: return launch_pipe<1, false, false>(at, grid, L.total, 0, st);
    case 3: return launch_pipe<1, true, false>(at, grid, L.total, 0, st);
    case 4: return launch_pipe<2, false, false>(at, grid, L.total, 0, st);
    default: return launch_pipe<2, true, false>(at, grid, L.total, 0, st);
  }
}

int launch_track_pipe(const TrackArgs& a, int moment_mode, bool has_out, int sm_count,
                      cudaStream_t st) {
  // Consumer warps, measured on B200 (profiles/r2j_k2_timeline.txt): the moment-free screen
  // variant 16 / 20 / 24 / 28 / 32 warps -> cfg4 screen 143.4 / 137.6 / 131.9 / 129.7 / 135.6 us
  // and 32 M particles 175 / 171.5 / 169 / 167.7 / 265 us, but the 1 M step 25.6 us at 24 and
  // 27.5 at 28; the moment-free outgoing-particles variant 16 / 20 / 22 -> 32 M particles
  // 254 / 244.5 / 248.6 us.
  int cw = 16;
  if (moment_mode == 0) cw = has_out ? 20 : (a.img != nullptr ? 24 : 16);
#ifdef LYNX_EXPERIMENTS
  if (cw == 24) cw = env_int("LYNX_K2_CW", cw);
  if (cw == 20) cw = env_int("LYNX_K2_CW_OUT", cw);
#endif
  int rc = launch_track_pipe_cw(a, moment_mode, has_out, cw, sm_count, st);
  if ((rc == -100 || rc == 1) && cw != 16)  // no room for the wider tile: the 16-warp kernel
    rc = launch_track_pipe_cw(a, moment_mode, has_out, 16, sm_count, st);
  return rc;
}

}  // namespace lynx

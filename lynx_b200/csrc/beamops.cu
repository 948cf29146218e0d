// K10: ParticleBeam.transformed_to on the device.
// K11: Aperture.track as an order-preserving stream compaction.
//
// Both are pure streaming passes over the SoA-6 planes: HBM-bound, no arithmetic to speak
// of.  Every floating-point operation below is an explicit round-to-nearest intrinsic
// (__fsub_rn, __fdiv_rn, ...) so that nvcc cannot contract a multiply and an add into an
// FMA: the results are bit-identical to the reference's op-by-op fp32 array arithmetic
// (and to the numpy / torch-CPU restatement in oracle/beamops.py).
#include <algorithm>
#include <cstdlib>

#include "common.cuh"

namespace lynx {

// ------------------------------------------------------------------ K10 transformed_to
// lynx/particles/particle_beam.py:696-702:
//   phase_space = (phase_space - old_mu) / old_sigma * new_sigma + new_mu
// per coordinate, per beam row.  Table layout per row: old_mu[6], old_sigma[6],
// new_sigma[6], new_mu[6] (24 floats).  48 B per particle (24 read + 24 written).
constexpr int kXfTPB = 256;

template <int kXfUnroll>
__global__ void __launch_bounds__(kXfTPB) beam_transform_kernel(
    const float* __restrict__ in, int64_t in_row_stride, int64_t ld_in, int64_t N,
    const float* __restrict__ table, float* __restrict__ out, int64_t ld_out) {
  const int64_t b = blockIdx.y;
  __shared__ float s_t[24];
  if (threadIdx.x < 24) s_t[threadIdx.x] = __ldg(table + b * 24 + threadIdx.x);
  __syncthreads();
  const float* src = in + b * in_row_stride;
  float* dst = out + b * 6 * ld_out;
  const int64_t groups = (N + 3) >> 2;
  const int64_t step = (int64_t)gridDim.x * kXfTPB;
  for (int64_t g0 = (int64_t)blockIdx.x * kXfTPB + threadIdx.x; g0 < groups;
       g0 += step * kXfUnroll) {
    float4 v[kXfUnroll][6];
#pragma unroll
    for (int u = 0; u < kXfUnroll; ++u) {
      const int64_t g = g0 + u * step;
      if (g < groups) {
#pragma unroll
        for (int k = 0; k < 6; ++k)
          v[u][k] = __ldcs(reinterpret_cast<const float4*>(src + k * ld_in + g * 4));
      }
    }
#pragma unroll
    for (int u = 0; u < kXfUnroll; ++u) {
      const int64_t g = g0 + u * step;
      if (g >= groups) continue;
#pragma unroll
      for (int k = 0; k < 6; ++k) {
        const float om = s_t[k], os = s_t[6 + k], ns = s_t[12 + k], nm = s_t[18 + k];
        float a[4] = {v[u][k].x, v[u][k].y, v[u][k].z, v[u][k].w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float r = __fadd_rn(__fmul_rn(__fdiv_rn(__fsub_rn(a[q], om), os), ns), nm);
          a[q] = (g * 4 + q < N) ? r : 0.f;  // padding lanes [N, ld) stay zero
        }
        __stcs(reinterpret_cast<float4*>(dst + k * ld_out + g * 4),
               make_float4(a[0], a[1], a[2], a[3]));
      }
    }
  }
}

// ------------------------------------------------------------------ K11 Aperture
// lynx/accelerator/aperture.py:69-109.  The reference selects the survivors with a boolean
// mask (`incoming.particles[survived_mask]`), i.e. an order-preserving compaction, and keeps
// the complement as `lost_particles`.  Three launches:
//   1. aperture_count_kernel   survivors per tile of kApTile particles (reads x, y: 8 B)
//   2. aperture_scan_kernel    exclusive scan of the tile counts, one CTA per beam row
//   3. aperture_scatter_kernel re-evaluates the mask, ranks the particles of the tile with
//                              a packed CTA scan, compacts the tile in shared memory and
//                              streams all planes out (24 B read, 24 B written: survivors
//                              and lost particles together are N)
// 56 B per particle.  (A single-pass decoupled look-back scan would be 48 B; it needs
// inter-CTA spinning, which this library otherwise avoids.)
constexpr int kApTPB = 256;
constexpr int kApVec = 1;                       // float4 per thread per plane
constexpr int kApTile = kApTPB * kApVec * 4;    // 1024 particles per CTA
constexpr int kApPlane = kApTile + 12;          // shared floats per plane (phase pads: 3 + 3 + 3, rounded)

// aperture.py:80-88.  NaN coordinates fail every comparison: lost.
template <int SHAPE>
__device__ __forceinline__ bool survives(float x, float y, float xm, float ym) {
  if (SHAPE == 0) return (x > -xm) && (x < xm) && (y > -ym) && (y < ym);
  const float t = __fadd_rn(__fdiv_rn(__fmul_rn(x, x), __fmul_rn(xm, xm)),
                            __fdiv_rn(__fmul_rn(y, y), __fmul_rn(ym, ym)));
  return t <= 1.0f;
}

template <int SHAPE>
__device__ __forceinline__ unsigned tile_flags(const float* __restrict__ xs,
                                               const float* __restrict__ ys, int64_t tile0,
                                               int64_t N, float xm, float ym, int it) {
  // 4 mask bits of the float4 at vector index it * kApTPB + threadIdx.x of the tile
  const int64_t p = tile0 + ((int64_t)it * kApTPB + threadIdx.x) * 4;
  if (p >= N) return 0u;
  const float4 x = __ldg(reinterpret_cast<const float4*>(xs + p));
  const float4 y = __ldg(reinterpret_cast<const float4*>(ys + p));
  unsigned f = 0;
  f |= (survives<SHAPE>(x.x, y.x, xm, ym) ? 1u : 0u);
  f |= (p + 1 < N && survives<SHAPE>(x.y, y.y, xm, ym)) ? 2u : 0u;
  f |= (p + 2 < N && survives<SHAPE>(x.z, y.z, xm, ym)) ? 4u : 0u;
  f |= (p + 3 < N && survives<SHAPE>(x.w, y.w, xm, ym)) ? 8u : 0u;
  return f;
}

template <int SHAPE>
__global__ void __launch_bounds__(kApTPB) aperture_count_kernel(
    const float* __restrict__ in, int64_t ld_in, int64_t N, const float* __restrict__ x_max,
    const float* __restrict__ y_max, int64_t lim_stride, uint32_t* __restrict__ tile_counts,
    int64_t n_tiles) {
  const int64_t r = blockIdx.y;
  const float xm = __ldg(x_max + r * lim_stride), ym = __ldg(y_max + r * lim_stride);
  const float* xs = in + (r * 6 + 0) * ld_in;
  const float* ys = in + (r * 6 + 2) * ld_in;
  __shared__ unsigned s_w[kApTPB / 32];
  for (int64_t t = blockIdx.x; t < n_tiles; t += gridDim.x) {
    unsigned c = 0;
#pragma unroll
    for (int it = 0; it < kApVec; ++it)
      c += __popc(tile_flags<SHAPE>(xs, ys, t * kApTile, N, xm, ym, it));
    c = __reduce_add_sync(0xffffffffu, c);
    if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
      unsigned tot = 0;
#pragma unroll
      for (int w = 0; w < kApTPB / 32; ++w) tot += s_w[w];
      tile_counts[r * n_tiles + t] = tot;
    }
    __syncthreads();
  }
}

constexpr int kScanTPB = 1024;
// In place: tile_counts[r][t] <- number of survivors in tiles [0, t); counts[r] <- total.
__global__ void __launch_bounds__(kScanTPB) aperture_scan_kernel(uint32_t* __restrict__ tile_counts,
                                                                int64_t n_tiles,
                                                                int64_t* __restrict__ counts) {
  const int64_t r = blockIdx.x;
  uint32_t* tc = tile_counts + r * n_tiles;
  __shared__ unsigned s_w[kScanTPB / 32];
  __shared__ unsigned long long s_carry;
  if (threadIdx.x == 0) s_carry = 0ull;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int64_t base = 0; base < n_tiles; base += kScanTPB) {
    const int64_t i = base + threadIdx.x;
    const unsigned v = i < n_tiles ? tc[i] : 0u;
    unsigned inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned o = __shfl_up_sync(0xffffffffu, inc, d);
      if (lane >= d) inc += o;
    }
    if (lane == 31) s_w[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      unsigned w = s_w[lane];
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const unsigned o = __shfl_up_sync(0xffffffffu, w, d);
        if (lane >= d) w += o;
      }
      s_w[lane] = w;  // inclusive over warps
    }
    __syncthreads();
    const unsigned long long carry = s_carry;
    const unsigned before = (warp ? s_w[warp - 1] : 0u) + inc - v;
    // a row holds < 2^32 particles (checked by the launcher), so 32-bit offsets suffice
    if (i < n_tiles) tc[i] = (uint32_t)(carry + before);
    __syncthreads();
    if (threadIdx.x == kScanTPB - 1) s_carry = carry + s_w[kScanTPB / 32 - 1];
    __syncthreads();
  }
  if (threadIdx.x == 0) counts[r] = (int64_t)s_carry;
}

struct ApertureArgs {
  const float* in;
  int64_t ld_in, N;
  const float* q_in;  // nullable particle charges [R][ld_q]
  int64_t ld_q;
  const float* x_max;
  const float* y_max;
  int64_t lim_stride;
  const uint32_t* tile_offsets;
  int64_t n_tiles;
  float* out;   // [R][6][ld_out]
  float* lost;  // nullable [R][6][ld_out]
  int64_t ld_out;
  float* q_out;   // nullable [R][ld_out]
  float* q_lost;  // nullable [R][ld_out]
  // one-pass variant
  unsigned int* ticket;        // [R] next tile to hand out (zeroed per call)
  unsigned long long* status;  // [R][n_tiles] look-back words (zeroed per call)
  int64_t* counts;             // [R]
};

// One region (survivors or lost particles) of the compacted tile leaves shared memory.  The
// region sits in shared memory with the SAME 16-byte phase as its destination: shared float
// index i (vector-aligned base `sv0`, first valid element sv0 + phase) maps to global element
// g0 + (i - sv0) with g0 a multiple of 4, so the body moves as LDS.128 -> STG.128 and only the
// first and the last vector of a region fall back to predicated scalar stores.  All planes of
// one vector are handled together (independent loads and stores in flight).
template <int PLANES>
__device__ __forceinline__ void copy_region(const float* __restrict__ sm, int sv0, int phase,
                                            int count, float* __restrict__ dst, int64_t plane_ld,
                                            int64_t g0, float* __restrict__ qdst, bool has_q) {
  if (count <= 0) return;
  const int lo = phase, hi = phase + count;  // valid shared elements, relative to sv0
  const int nvec = (hi + 3) >> 2;
  for (int v = threadIdx.x; v < nvec; v += kApTPB) {
    const int e0 = v * 4;
    const bool full = e0 >= lo && e0 + 4 <= hi;
    float4 x[PLANES + 1];
#pragma unroll
    for (int k = 0; k < PLANES; ++k)
      x[k] = *reinterpret_cast<const float4*>(sm + k * kApPlane + sv0 + e0);
    if (has_q) x[PLANES] = *reinterpret_cast<const float4*>(sm + PLANES * kApPlane + sv0 + e0);
    if (full) {
#pragma unroll
      for (int k = 0; k < PLANES; ++k)
        __stcs(reinterpret_cast<float4*>(dst + k * plane_ld + g0 + e0), x[k]);
      if (has_q && qdst) __stcs(reinterpret_cast<float4*>(qdst + g0 + e0), x[PLANES]);
    } else {
#pragma unroll
      for (int k = 0; k <= PLANES; ++k) {
        if (k == PLANES && !(has_q && qdst)) break;
        float* d = (k < PLANES ? dst + k * plane_ld : qdst) + g0 + e0;
        const float e[4] = {x[k].x, x[k].y, x[k].z, x[k].w};
#pragma unroll
        for (int q = 0; q < 4; ++q)
          if (e0 + q >= lo && e0 + q < hi) d[q] = e[q];
      }
    }
  }
}

// The tile's particles are first compacted inside shared memory, then leave as contiguous
// 16-byte streams (copy_region).  History (32 M particles): every thread
// storing its up to four survivors straight to global memory -- eight partial-sector store
// instructions per thread and plane -- ran at 2.7 TB/s; compaction in shared memory with a
// scalar-head / vector-body / scalar-tail copy per plane at 3.5 TB/s (instruction-bound:
// twelve branchy copy loops per tile); this version phase-aligns the shared-memory regions
// to their destinations.
// ONEPASS: no count / scan launches.  Tiles are handed out by a ticket counter (so every
// predecessor of a tile is held by a running CTA) and a tile learns the number of survivors in
// front of it by decoupled look-back over 64-bit status words (flag << 62 | value; flag 1 =
// this tile's own count, 2 = inclusive prefix), read by one warp, 32 predecessors at a time.
// The words carry their whole message, so no fences are involved.  48 B per particle.
__device__ __forceinline__ unsigned long long ld_status(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_status(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

template <int SHAPE, bool ONEPASS>
__global__ void __launch_bounds__(kApTPB, 3) aperture_scatter_kernel(ApertureArgs a) {
  extern __shared__ __align__(16) float ap_smem[];  // [6 or 7][kApPlane]
  const int64_t r = blockIdx.y;
  const float xm = __ldg(a.x_max + r * a.lim_stride), ym = __ldg(a.y_max + r * a.lim_stride);
  const float* src = a.in + r * 6 * a.ld_in;
  float* kept = a.out + r * 6 * a.ld_out;
  float* lost = a.lost ? a.lost + r * 6 * a.ld_out : nullptr;
  const bool has_q = a.q_in != nullptr;
  const bool q_vec = has_q && (a.ld_q & 3) == 0 && ((uintptr_t)a.q_in & 15) == 0;
  float* q_kept = a.q_out ? a.q_out + r * a.ld_out : nullptr;
  float* q_lost = a.q_lost ? a.q_lost + r * a.ld_out : nullptr;
  unsigned long long* status = ONEPASS ? a.status + r * a.n_tiles : nullptr;
  __shared__ unsigned s_w[kApTPB / 32];
  __shared__ long long s_tile[2];  // ONEPASS: [0] this CTA's next tile, [1] survivors in front
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // Software pipeline: the loads of the CTA's NEXT tile are issued as soon as the current
  // tile's values have been scattered into shared memory, so they are in flight during the
  // copy-out phase (and across the barriers) without costing extra registers.
  float4 v[7];
  int64_t kept0_next = 0;
  auto load_tile = [&](int64_t t) {
    const int64_t p = t * kApTile + threadIdx.x * 4;
    if (!ONEPASS) kept0_next = __ldg(a.tile_offsets + r * a.n_tiles + t);
    if (p < a.N) {
#pragma unroll
      for (int k = 0; k < 6; ++k)
        v[k] = __ldcs(reinterpret_cast<const float4*>(src + k * a.ld_in + p));
      if (has_q) {
        const float* qp = a.q_in + r * a.ld_q + p;
        if (q_vec) {
          v[6] = __ldcs(reinterpret_cast<const float4*>(qp));
        } else {
          const int nv = (int)min((int64_t)4, a.N - p);
          v[6].x = __ldcs(qp);
          v[6].y = nv > 1 ? __ldcs(qp + 1) : 0.f;
          v[6].z = nv > 2 ? __ldcs(qp + 2) : 0.f;
          v[6].w = nv > 3 ? __ldcs(qp + 3) : 0.f;
        }
      }
    }
  };
  int64_t t;
  if (ONEPASS) {
    if (threadIdx.x == 0) s_tile[0] = (long long)atomicAdd(a.ticket + r, 1u);
    __syncthreads();
    t = s_tile[0];
    __syncthreads();  // s_tile[0] is rewritten at the top of the loop
  } else {
    t = blockIdx.x;
  }
  if (t < a.n_tiles) load_tile(t);
  while (t < a.n_tiles) {
    if (ONEPASS && threadIdx.x == 0) s_tile[0] = (long long)atomicAdd(a.ticket + r, 1u);
    const int64_t tile0 = t * kApTile;
    const int n_tile = (int)min((int64_t)kApTile, a.N - tile0);
    const int local = threadIdx.x * 4;
    const int64_t p = tile0 + local;
    unsigned f = 0;
    if (p < a.N) {
      const float4 x = v[0], y = v[2];
      f |= survives<SHAPE>(x.x, y.x, xm, ym) ? 1u : 0u;
      f |= (p + 1 < a.N && survives<SHAPE>(x.y, y.y, xm, ym)) ? 2u : 0u;
      f |= (p + 2 < a.N && survives<SHAPE>(x.z, y.z, xm, ym)) ? 4u : 0u;
      f |= (p + 3 < a.N && survives<SHAPE>(x.w, y.w, xm, ym)) ? 8u : 0u;
    }
    // exclusive scan of the survivor counts over the CTA
    const unsigned mine = (unsigned)__popc(f);
    unsigned inc = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned o = __shfl_up_sync(0xffffffffu, inc, d);
      if (lane >= d) inc += o;
    }
    if (lane == 31) s_w[warp] = inc;
    __syncthreads();  // also: the previous tile's copy-out has finished reading ap_smem
    int k_rank = (int)(inc - mine), cnt = 0;
#pragma unroll
    for (int w = 0; w < kApTPB / 32; ++w) {
      const int sw = (int)s_w[w];
      if (w < warp) k_rank += sw;
      cnt += sw;
    }
    int64_t kept0;
    const int64_t t_next = ONEPASS ? (int64_t)s_tile[0] : t + gridDim.x;
    if (ONEPASS) {
      if (warp == 0) {
        constexpr unsigned long long kAgg = 1ull << 62, kPre = 2ull << 62, kVal = (1ull << 62) - 1;
        unsigned long long excl = 0;
        if (t > 0) {
          if (lane == 0) st_status(status + t, kAgg | (unsigned long long)cnt);
          int64_t j = t - 1;  // nearest predecessor of the current window
          while (true) {
            const int64_t idx = j - lane;
            const unsigned long long wd = idx >= 0 ? ld_status(status + idx) : kPre;
            const unsigned flag = (unsigned)(wd >> 62);
            const unsigned pre = __ballot_sync(0xffffffffu, flag == 2u);
            const unsigned inv = __ballot_sync(0xffffffffu, flag == 0u);
            const int first = pre ? __ffs(pre) - 1 : 31;  // nearest lane holding a prefix
            const unsigned rel = first == 31 ? 0xffffffffu : ((2u << first) - 1u);
            if (inv & rel) continue;  // a predecessor in range has not published yet
            unsigned long long part = (rel >> lane) & 1u ? (wd & kVal) : 0ull;
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) part += __shfl_xor_sync(0xffffffffu, part, d);
            excl += part;
            if (pre) break;
            j -= 32;
          }
        }
        if (lane == 0) {
          st_status(status + t, kPre | (excl + (unsigned long long)cnt));
          s_tile[1] = (long long)excl;
          if (t == a.n_tiles - 1) a.counts[r] = (int64_t)(excl + (unsigned long long)cnt);
        }
      }
      __syncthreads();
      kept0 = s_tile[1];
    } else {
      kept0 = kept0_next;
    }
    const int64_t lost0 = tile0 - kept0;
    // shared-memory layout of a plane: survivors from float index ka (phase of kept0), lost
    // particles from the next vector boundary + the phase of lost0
    const int ka = (int)(kept0 & 3);
    const int lv0 = (ka + cnt + 3) & ~3;
    const int la = (int)(lost0 & 3);
    if (local < n_tile) {
      const int l_rank = local - k_rank;
      const int nvalid = min(4, n_tile - local);
#pragma unroll
      for (int k = 0; k < 7; ++k) {
        if (k == 6 && !has_q) break;
        float* sp = ap_smem + k * kApPlane;
        const float e[4] = {v[k].x, v[k].y, v[k].z, v[k].w};
        int kk = ka + k_rank, ll = lv0 + la + l_rank;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          if (q < nvalid) {
            if (f & (1u << q)) sp[kk++] = e[q];
            else sp[ll++] = e[q];
          }
        }
      }
    }
    if (t_next < a.n_tiles) load_tile(t_next);
    __syncthreads();
    copy_region<6>(ap_smem, 0, ka, cnt, kept, a.ld_out, kept0 - ka, q_kept, has_q);
    if (lost)
      copy_region<6>(ap_smem, lv0, la, n_tile - cnt, lost, a.ld_out, lost0 - la, q_lost, has_q);
    // the next iteration's first __syncthreads() orders these reads before its writes
    t = t_next;
  }
}

}  // namespace lynx

using namespace lynx;

extern "C" int lynx_beam_transform(const float* d_in_soa, int64_t in_row_stride, int64_t ld_in,
                                   int64_t B, int64_t N, const float* d_table, float* d_out_soa,
                                   int64_t ld_out, lynx_stream_t stream) {
  if (!d_in_soa || !d_table || !d_out_soa || B <= 0 || B > 65535 || N <= 0 || ld_in < N ||
      ld_out < N || (ld_in & 3) || (ld_out & 3) || (in_row_stride & 3) ||
      ((uintptr_t)d_in_soa & 15) || ((uintptr_t)d_out_soa & 15))
    return LYNX_E_BADARG;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int64_t groups = (N + 3) >> 2;
  const int unroll = 2, ctas = 8;  // measured best on B200 (tools/beamops_sweep.py)
  const int64_t want = (groups + kXfTPB * unroll - 1) / (kXfTPB * unroll);
  const int64_t per_b = std::max<int64_t>(1, ((int64_t)ctas * sms + B - 1) / B);
  const dim3 grid((unsigned)std::max<int64_t>(1, std::min<int64_t>(want, per_b)), (unsigned)B);
  {
    ProfileScope ps(4, (cudaStream_t)stream);
    auto kern = beam_transform_kernel<2>;
    kern<<<grid, kXfTPB, 0, (cudaStream_t)stream>>>(d_in_soa, in_row_stride, ld_in, N, d_table,
                                                    d_out_soa, ld_out);
  }
  count_launch();
  LYNX_CUDA_OK(cudaGetLastError());
  return LYNX_OK;
}

extern "C" size_t lynx_aperture_workspace_bytes(int64_t R, int64_t N) {
  if (R <= 0 || N <= 0) return 0;
  const int64_t n_tiles = (N + kApTile - 1) / kApTile;
  // three-pass: R * n_tiles counts; one-pass: R * n_tiles 64-bit status words + R tickets
  return (size_t)(R * n_tiles) * sizeof(unsigned long long) + (size_t)((R + 3) & ~3LL) * 4;
}

extern "C" int lynx_aperture_compact(const float* d_in_soa, int64_t R, int64_t N, int64_t ld_in,
                                     const float* d_charges_in, int64_t ld_q,
                                     const float* d_x_max, const float* d_y_max,
                                     int64_t lim_stride, int32_t shape, float* d_out_soa,
                                     float* d_lost_soa, int64_t ld_out, float* d_charges_out,
                                     float* d_charges_lost, int64_t* d_counts,
                                     void* d_workspace, lynx_stream_t stream) {
  if (!d_in_soa || !d_x_max || !d_y_max || !d_out_soa || !d_counts || !d_workspace || R <= 0 ||
      R > 65535 || N <= 0 || N >= (1LL << 32) || ld_in < N || ld_out < N || (ld_in & 3) ||
      (ld_out & 3) || ((uintptr_t)d_in_soa & 15) || ((uintptr_t)d_out_soa & 15) ||
      ((uintptr_t)d_lost_soa & 15) || ((uintptr_t)d_charges_out & 15) ||
      ((uintptr_t)d_charges_lost & 15) || (shape != 0 && shape != 1) ||
      (lim_stride != 0 && lim_stride != 1) || (d_charges_in && ld_q < N))
    return LYNX_E_BADARG;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t n_tiles = (N + kApTile - 1) / kApTile;
  ApertureArgs a;
  a.in = d_in_soa;
  a.ld_in = ld_in;
  a.N = N;
  a.q_in = d_charges_in;
  a.ld_q = ld_q;
  a.x_max = d_x_max;
  a.y_max = d_y_max;
  a.lim_stride = lim_stride;
  a.n_tiles = n_tiles;
  a.out = d_out_soa;
  a.lost = d_lost_soa;
  a.ld_out = ld_out;
  a.q_out = d_charges_out;
  a.q_lost = d_charges_lost;
  a.counts = d_counts;
  const size_t smem = (size_t)(d_charges_in ? 7 : 6) * kApPlane * sizeof(float);
  // persistent: as many CTAs as are resident (3 per SM), each loops over its tiles
  const dim3 sgrid((unsigned)std::max<int64_t>(1, std::min<int64_t>(n_tiles, (3LL * sms + R - 1) / R)),
                   (unsigned)R);
  // Tuning builds only (-DLYNX_EXPERIMENTS, LYNX_APERTURE_ONEPASS=1), a measured negative result: one launch and 48 B per
  // particle, but 1.08 ms against 0.29 + 0.08 ms at 32 M particles.  With 1024-particle tiles
  // and 444 resident CTAs running in lock-step a tile's nearest published prefix is ~14
  // look-back windows (dependent L2 round trips) away, and the scatter cannot start before it.
#ifdef LYNX_EXPERIMENTS
  static const bool onepass = [] { const char* e = getenv("LYNX_APERTURE_ONEPASS"); return e && e[0] == '1'; }();
  if (onepass) {
    // one launch: ticketed tiles + decoupled look-back (48 B per particle)
    a.status = reinterpret_cast<unsigned long long*>(d_workspace);
    a.ticket = reinterpret_cast<unsigned int*>(a.status + R * n_tiles);
    a.tile_offsets = nullptr;
    LYNX_CUDA_OK(cudaMemsetAsync(d_workspace, 0, lynx_aperture_workspace_bytes(R, N), st));
    auto kern = shape == 0 ? aperture_scatter_kernel<0, true> : aperture_scatter_kernel<1, true>;
    static bool attr_set[2] = {false, false};
    if (!attr_set[shape]) {
      LYNX_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        7 * kApPlane * (int)sizeof(float)));
      attr_set[shape] = true;
    }
    {
      ProfileScope ps(5, st);
      kern<<<sgrid, kApTPB, smem, st>>>(a);
    }
    count_launch(1);
    LYNX_CUDA_OK(cudaGetLastError());
    return LYNX_OK;
  }
#endif
  // three launches: tile counts, scan, scatter (56 B per particle)
  uint32_t* tiles = reinterpret_cast<uint32_t*>(d_workspace);
  const int64_t per_r = std::max<int64_t>(1, (8LL * sms + R - 1) / R);
  const dim3 grid((unsigned)std::min<int64_t>(n_tiles, per_r), (unsigned)R);
  if (shape == 0)
    aperture_count_kernel<0><<<grid, kApTPB, 0, st>>>(d_in_soa, ld_in, N, d_x_max, d_y_max,
                                                      lim_stride, tiles, n_tiles);
  else
    aperture_count_kernel<1><<<grid, kApTPB, 0, st>>>(d_in_soa, ld_in, N, d_x_max, d_y_max,
                                                      lim_stride, tiles, n_tiles);
  aperture_scan_kernel<<<(unsigned)R, kScanTPB, 0, st>>>(tiles, n_tiles, d_counts);
  a.tile_offsets = tiles;
  a.status = nullptr;
  a.ticket = nullptr;
  {
    auto kern = shape == 0 ? aperture_scatter_kernel<0, false> : aperture_scatter_kernel<1, false>;
    static bool attr_set[2] = {false, false};
    if (!attr_set[shape]) {
      LYNX_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        7 * kApPlane * (int)sizeof(float)));
      attr_set[shape] = true;
    }
    ProfileScope ps(5, st);
    kern<<<sgrid, kApTPB, smem, st>>>(a);
  }
  count_launch(3);
  LYNX_CUDA_OK(cudaGetLastError());
  return LYNX_OK;
}

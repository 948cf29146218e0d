// K10: ParticleBeam.transformed_to on the device.
// K11: Aperture.track as an order-preserving stream compaction.
//
// Both are pure streaming passes over the SoA-6 planes: HBM-bound, no arithmetic to speak
// of.  Every floating-point operation below is an explicit round-to-nearest intrinsic
// (__fsub_rn, __fdiv_rn, ...) so that nvcc cannot contract a multiply and an add into an
// FMA: the results are bit-identical to the reference's op-by-op fp32 array arithmetic
// (and to the numpy / torch-CPU restatement in oracle/beamops.py).
#include <algorithm>

#include "common.cuh"

namespace lynx {

// ------------------------------------------------------------------ K10 transformed_to
// lynx/particles/particle_beam.py:696-702:
//   phase_space = (phase_space - old_mu) / old_sigma * new_sigma + new_mu
// per coordinate, per beam row.  Table layout per row: old_mu[6], old_sigma[6],
// new_sigma[6], new_mu[6] (24 floats).  48 B per particle (24 read + 24 written).
constexpr int kXfTPB = 256;
constexpr int kXfUnroll = 2;

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
constexpr int kApVec = 2;                       // float4 per thread per plane
constexpr int kApTile = kApTPB * kApVec * 4;    // 2048 particles per CTA

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
};

// Copies n floats from shared memory to dst[d0 .. d0 + n): scalar stores up to the first
// 16-byte boundary of the destination, 16-byte stores for the aligned body, scalar tail.
// `dst` itself is 16-byte aligned (plane base; ld multiple of 4), so the alignment of
// element d0 + i is that of (d0 + i) & 3.
__device__ __forceinline__ void copy_out(const float* __restrict__ s, float* __restrict__ dst,
                                         int64_t d0, int n, bool vec_ok) {
  if (n <= 0) return;
  if (!vec_ok) {
    for (int i = threadIdx.x; i < n; i += kApTPB) dst[d0 + i] = s[i];
    return;
  }
  const int head = min(n, (int)((4 - (d0 & 3)) & 3));
  if ((int)threadIdx.x < head) dst[d0 + threadIdx.x] = s[threadIdx.x];
  const int body = (n - head) >> 2;
  for (int v = threadIdx.x; v < body; v += kApTPB) {
    const float* q = s + head + v * 4;
    __stcs(reinterpret_cast<float4*>(dst + d0 + head + (int64_t)v * 4),
           make_float4(q[0], q[1], q[2], q[3]));
  }
  const int done = head + body * 4;
  if ((int)threadIdx.x < n - done) dst[d0 + done + threadIdx.x] = s[done + threadIdx.x];
}

// The tile's particles are first compacted inside shared memory (survivors at [0, cnt), lost
// particles at [cnt, n) of each plane), then leave as contiguous, 16-byte-vectorised streams.
// (First version: every thread stored its up to four survivors straight to global memory --
// eight partial-sector store instructions per thread and plane, 2.7 TB/s)
template <int SHAPE>
__global__ void __launch_bounds__(kApTPB, 2) aperture_scatter_kernel(ApertureArgs a) {
  extern __shared__ __align__(16) float ap_smem[];  // [7][kApTile]
  const int64_t r = blockIdx.y;
  const float xm = __ldg(a.x_max + r * a.lim_stride), ym = __ldg(a.y_max + r * a.lim_stride);
  const float* src = a.in + r * 6 * a.ld_in;
  const float* xs = src;
  const float* ys = src + 2 * a.ld_in;
  float* kept = a.out + r * 6 * a.ld_out;
  float* lost = a.lost ? a.lost + r * 6 * a.ld_out : nullptr;
  const bool has_q = a.q_in != nullptr;
  const bool q_vec = has_q && (a.ld_q & 3) == 0 && ((uintptr_t)a.q_in & 15) == 0;
  const bool out_vec = (a.ld_out & 3) == 0 && ((uintptr_t)a.out & 15) == 0 &&
                       (!a.lost || ((uintptr_t)a.lost & 15) == 0);
  const bool qo_vec = (a.ld_out & 3) == 0 && (!a.q_out || ((uintptr_t)a.q_out & 15) == 0) &&
                      (!a.q_lost || ((uintptr_t)a.q_lost & 15) == 0);
  __shared__ unsigned s_w[kApTPB / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int64_t t = blockIdx.x; t < a.n_tiles; t += gridDim.x) {
    const int64_t tile0 = t * kApTile;
    const int n_tile = (int)min((int64_t)kApTile, a.N - tile0);
    // all loads of the tile are issued before anything depends on them
    float4 v[kApVec][7];
#pragma unroll
    for (int it = 0; it < kApVec; ++it) {
      const int64_t p = tile0 + ((int64_t)it * kApTPB + threadIdx.x) * 4;
      if (p < a.N) {
#pragma unroll
        for (int k = 0; k < 6; ++k)
          v[it][k] = __ldcs(reinterpret_cast<const float4*>(src + k * a.ld_in + p));
        if (has_q) {
          const float* qp = a.q_in + r * a.ld_q + p;
          if (q_vec) {
            v[it][6] = __ldcs(reinterpret_cast<const float4*>(qp));
          } else {
            const int nv = (int)min((int64_t)4, a.N - p);
            v[it][6].x = __ldcs(qp);
            v[it][6].y = nv > 1 ? __ldcs(qp + 1) : 0.f;
            v[it][6].z = nv > 2 ? __ldcs(qp + 2) : 0.f;
            v[it][6].w = nv > 3 ? __ldcs(qp + 3) : 0.f;
          }
        }
      }
    }
    unsigned flags[kApVec], packed = 0;  // packed: survivors of vector `it` in bits 12*it..
#pragma unroll
    for (int it = 0; it < kApVec; ++it) {
      const int64_t p = tile0 + ((int64_t)it * kApTPB + threadIdx.x) * 4;
      unsigned f = 0;
      if (p < a.N) {
        const float4 x = v[it][0], y = v[it][2];
        f |= survives<SHAPE>(x.x, y.x, xm, ym) ? 1u : 0u;
        f |= (p + 1 < a.N && survives<SHAPE>(x.y, y.y, xm, ym)) ? 2u : 0u;
        f |= (p + 2 < a.N && survives<SHAPE>(x.z, y.z, xm, ym)) ? 4u : 0u;
        f |= (p + 3 < a.N && survives<SHAPE>(x.w, y.w, xm, ym)) ? 8u : 0u;
      }
      flags[it] = f;
      packed += (unsigned)__popc(f) << (12 * it);
    }
    // exclusive scan of `packed` over the CTA (each 12-bit field holds at most 1024)
    unsigned inc = packed;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned o = __shfl_up_sync(0xffffffffu, inc, d);
      if (lane >= d) inc += o;
    }
    if (lane == 31) s_w[warp] = inc;
    __syncthreads();  // also: the previous tile's copy-out has finished reading ap_smem
    unsigned before = inc - packed, total = 0;
#pragma unroll
    for (int w = 0; w < kApTPB / 32; ++w) {
      const unsigned sw = s_w[w];
      if (w < warp) before += sw;
      total += sw;
    }
    int cnt = 0;
#pragma unroll
    for (int it = 0; it < kApVec; ++it) cnt += (int)((total >> (12 * it)) & 0xFFFu);
    unsigned field_base = 0;  // survivors of the vectors it' < it of the whole CTA
#pragma unroll
    for (int it = 0; it < kApVec; ++it) {
      const int local = (it * kApTPB + threadIdx.x) * 4;
      const int k_rank = (int)(field_base + ((before >> (12 * it)) & 0xFFFu));
      field_base += (total >> (12 * it)) & 0xFFFu;
      if (local >= n_tile) continue;
      const int l_rank = cnt + (local - k_rank);  // lost particles live behind the survivors
      const unsigned f = flags[it];
      const int nvalid = min(4, n_tile - local);
#pragma unroll
      for (int k = 0; k < 7; ++k) {
        if (k == 6 && !has_q) break;
        float* sp = ap_smem + k * kApTile;
        const float e[4] = {v[it][k].x, v[it][k].y, v[it][k].z, v[it][k].w};
        int kk = 0, ll = 0;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          if (q < nvalid) {
            if (f & (1u << q)) sp[k_rank + kk++] = e[q];
            else sp[l_rank + ll++] = e[q];
          }
        }
      }
    }
    __syncthreads();
    const int64_t kept0 = a.tile_offsets[r * a.n_tiles + t];
    const int64_t lost0 = tile0 - kept0;
#pragma unroll 1
    for (int k = 0; k < 6; ++k) {
      const float* sp = ap_smem + k * kApTile;
      copy_out(sp, kept + k * a.ld_out, kept0, cnt, out_vec);
      if (lost) copy_out(sp + cnt, lost + k * a.ld_out, lost0, n_tile - cnt, out_vec);
    }
    if (has_q) {
      const float* sp = ap_smem + 6 * kApTile;
      if (a.q_out) copy_out(sp, a.q_out + r * a.ld_out, kept0, cnt, qo_vec);
      if (a.q_lost) copy_out(sp + cnt, a.q_lost + r * a.ld_out, lost0, n_tile - cnt, qo_vec);
    }
    // the next iteration's first __syncthreads() orders these reads before its writes
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
  const int64_t want = (groups + kXfTPB * kXfUnroll - 1) / (kXfTPB * kXfUnroll);
  const int64_t per_b = std::max<int64_t>(1, (8LL * sms + B - 1) / B);
  const unsigned gx = (unsigned)std::max<int64_t>(1, std::min<int64_t>(want, per_b));
  {
    ProfileScope ps(4, (cudaStream_t)stream);
    beam_transform_kernel<<<dim3(gx, (unsigned)B), kXfTPB, 0, (cudaStream_t)stream>>>(
        d_in_soa, in_row_stride, ld_in, N, d_table, d_out_soa, ld_out);
  }
  count_launch();
  LYNX_CUDA_OK(cudaGetLastError());
  return LYNX_OK;
}

extern "C" size_t lynx_aperture_workspace_bytes(int64_t R, int64_t N) {
  if (R <= 0 || N <= 0) return 0;
  const int64_t n_tiles = (N + kApTile - 1) / kApTile;
  return (size_t)(R * n_tiles) * sizeof(uint32_t);
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
      ((uintptr_t)d_in_soa & 15) || (shape != 0 && shape != 1) ||
      (lim_stride != 0 && lim_stride != 1) || (d_charges_in && ld_q < N))
    return LYNX_E_BADARG;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t n_tiles = (N + kApTile - 1) / kApTile;
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
  ApertureArgs a;
  a.in = d_in_soa;
  a.ld_in = ld_in;
  a.N = N;
  a.q_in = d_charges_in;
  a.ld_q = ld_q;
  a.x_max = d_x_max;
  a.y_max = d_y_max;
  a.lim_stride = lim_stride;
  a.tile_offsets = tiles;
  a.n_tiles = n_tiles;
  a.out = d_out_soa;
  a.lost = d_lost_soa;
  a.ld_out = ld_out;
  a.q_out = d_charges_out;
  a.q_lost = d_charges_lost;
  {
    const size_t smem = (size_t)(d_charges_in ? 7 : 6) * kApTile * sizeof(float);
    auto kern = shape == 0 ? aperture_scatter_kernel<0> : aperture_scatter_kernel<1>;
    static bool attr_set[2] = {false, false};
    if (!attr_set[shape]) {
      LYNX_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        7 * kApTile * (int)sizeof(float)));
      attr_set[shape] = true;
    }
    ProfileScope ps(5, st);
    kern<<<grid, kApTPB, smem, st>>>(a);
  }
  count_launch(3);
  LYNX_CUDA_OK(cudaGetLastError());
  return LYNX_OK;
}

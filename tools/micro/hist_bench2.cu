// Microbenchmark 2, round 2: what an atomic instruction costs when only some lanes are active,
// and a hybrid histogram whose warps issue only FULL atomic instructions (in-tile lanes ->
// shared-memory atomics, the others are compacted into a per-warp queue and leave as full-warp
// global REDs).  Follow-up of hist_bench.cu, where the divergent version plateaued at 142 G/s.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o hist_bench2 hist_bench2.cu
#include <cuda_runtime.h>
#include <stdint.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>
#define CK(x)                                                                  \
  do {                                                                         \
    cudaError_t e = (x);                                                       \
    if (e != cudaSuccess) {                                                    \
      printf("err %s line %d\n", cudaGetErrorString(e), __LINE__);             \
      exit(1);                                                                 \
    }                                                                          \
  } while (0)

constexpr int W = 2448, H = 2040;
constexpr int TPB = 512;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void red_f32(float* p, float v) {
  asm volatile("red.relaxed.gpu.global.add.f32 [%0], %1;" ::"l"(p), "f"(v));
}
__device__ __forceinline__ void reds_u32(uint32_t* p, uint32_t v) {
  asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(smem_u32(p)), "r"(v) : "memory");
}
__device__ __forceinline__ void bulk_reduce_add_f32(float* dst, const float* src_smem, uint32_t bytes) {
  asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(dst),
               "r"(smem_u32(src_smem)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit_wait() {
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}
__device__ __forceinline__ uint32_t hash32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}

// ---- cost of an atomic with K of 32 lanes active (uniform random addresses: no contention)
// which: 0 global RED.F32, 1 shared atomicAdd(p, 1u) [ATOMS.POPC.INC], 2 shared red.add.u32 with
// a register operand [ATOMS.ADD]
template <int WHICH>
__global__ void __launch_bounds__(TPB, 1) k_lanes(float* img, int iters, int K, long long* cyc) {
  extern __shared__ __align__(128) uint32_t tile[];
  const int npx = 28 * 1024;
  for (int i = threadIdx.x; i < npx; i += TPB) tile[i] = 0u;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  uint32_t s = hash32(blockIdx.x * TPB + threadIdx.x + 1);
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    s = s * 1664525u + 1013904223u;
    const uint32_t r = hash32(s);
    if (lane < K) {
      if (WHICH == 0) red_f32(img + (r % (uint32_t)(W * H)), 1.0f);
      if (WHICH == 1) atomicAdd(tile + (r % (uint32_t)npx), 1u);
      if (WHICH == 2) reds_u32(tile + (r % (uint32_t)npx), 1u + (r >> 31));
    }
  }
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
  if (tile[threadIdx.x] == 0xffffffffu) img[0] = 1.f;
}

struct Tile {
  int x0, y0, w, h;
};

// ---- hybrid with full-warp instructions only.  BITS = 32: u32 counters; 16: two per word.
// atoms_mode 1: atomicAdd(p, 1u); 2: red.shared.add.u32
template <int BITS, int MODE>
__global__ void __launch_bounds__(TPB, 1) k_hybrid2(const int* __restrict__ idx, float* img, long n, Tile t,
                                                     int compact) {
  extern __shared__ __align__(128) uint32_t tile[];
  __shared__ int queue[TPB / 32][64];
  const int npx = t.w * t.h;
  const int nwords = BITS == 32 ? npx : npx / 2;
  for (int i = threadIdx.x; i < nwords; i += TPB) tile[i] = 0u;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int* q = queue[warp];
  int qn = 0;  // warp-uniform
  long i = (long)blockIdx.x * TPB + threadIdx.x;
  const long stride = (long)gridDim.x * TPB;
  const long n_round = (n + stride - 1) / stride * stride;
  for (; i < n_round; i += stride) {
    const bool live = i < n;
    const int v = live ? idx[i] : 0;
    const int y = v / W, x = v - y * W;
    const unsigned dx = (unsigned)(x - t.x0), dy = (unsigned)(y - t.y0);
    const bool in = live && dx < (unsigned)t.w && dy < (unsigned)t.h;
    if (in) {
      const unsigned o = dy * t.w + dx;
      if (BITS == 32) {
        if (MODE == 1) atomicAdd(tile + o, 1u);
        else reds_u32(tile + o, 1u);
      } else {
        if (MODE == 1) atomicAdd(tile + (o >> 1), 1u << ((o & 1u) * 16u));
        else reds_u32(tile + (o >> 1), 1u << ((o & 1u) * 16u));
      }
    }
    const bool out = live && !in;
    if (!compact) {
      if (out) red_f32(img + v, 1.0f);
    } else {
      const unsigned m = __ballot_sync(0xffffffffu, out);
      if (m) {
        if (out) q[qn + __popc(m & ((1u << lane) - 1u))] = v;
        qn += __popc(m);
        __syncwarp();
        if (qn >= 32) {
          red_f32(img + q[lane], 1.0f);
          __syncwarp();
          if (lane < qn - 32) q[lane] = q[32 + lane];
          qn -= 32;
          __syncwarp();
        }
      }
    }
  }
  if (compact && lane < qn) red_f32(img + q[lane], 1.0f);
  __syncthreads();
  float* f = reinterpret_cast<float*>(tile);
  if (BITS == 32) {
    for (int k = threadIdx.x; k < npx; k += TPB) f[k] = (float)tile[k];
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    for (int r = threadIdx.x; r < t.h; r += TPB)
      bulk_reduce_add_f32(img + (long)(t.y0 + r) * W + t.x0, f + r * t.w, (uint32_t)t.w * 4u);
    bulk_commit_wait();
  } else {
    for (int k = threadIdx.x; k < nwords; k += TPB) {
      const uint32_t c = tile[k];
      if (c) {
        const int p = 2 * k, r = p / t.w, cx = p - r * t.w;
        float* d = img + (long)(t.y0 + r) * W + t.x0 + cx;
        if (c & 0xffffu) red_f32(d, (float)(c & 0xffffu));
        if (c >> 16) red_f32(d + 1, (float)(c >> 16));
      }
    }
  }
}

// ---- the same hybrid with enough loads in flight (4 x 16 B per thread): k_hybrid2 is bound by
// the latency of its own index loads (one 4-byte load in flight per thread), not by the atomics
template <int BITS>
__global__ void __launch_bounds__(TPB, 1) k_hybrid3(const int4* __restrict__ idx4, float* img, long n4, Tile t) {
  extern __shared__ __align__(128) uint32_t tile[];
  const int npx = t.w * t.h;
  const int nwords = BITS == 32 ? npx : npx / 2;
  for (int i = threadIdx.x; i < nwords; i += TPB) tile[i] = 0u;
  __syncthreads();
  long i = (long)blockIdx.x * TPB + threadIdx.x;
  const long stride = (long)gridDim.x * TPB;
  for (; i < n4; i += 4 * stride) {
    int4 q[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) q[u] = (i + u * stride < n4) ? __ldg(idx4 + i + u * stride) : make_int4(-1, -1, -1, -1);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int vv[4] = {q[u].x, q[u].y, q[u].z, q[u].w};
#pragma unroll
      for (int l = 0; l < 4; ++l) {
        const int v = vv[l];
        if (v < 0) continue;
        const int y = v / W, x = v - y * W;
        const unsigned dx = (unsigned)(x - t.x0), dy = (unsigned)(y - t.y0);
        if (dx < (unsigned)t.w && dy < (unsigned)t.h) {
          const unsigned o = dy * t.w + dx;
          if (BITS == 32) atomicAdd(tile + o, 1u);
          else atomicAdd(tile + (o >> 1), 1u << ((o & 1u) * 16u));
        } else {
          red_f32(img + v, 1.0f);
        }
      }
    }
  }
  __syncthreads();
  float* f = reinterpret_cast<float*>(tile);
  if (BITS == 32) {
    for (int k = threadIdx.x; k < npx; k += TPB) f[k] = (float)tile[k];
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    for (int r = threadIdx.x; r < t.h; r += TPB)
      bulk_reduce_add_f32(img + (long)(t.y0 + r) * W + t.x0, f + r * t.w, (uint32_t)t.w * 4u);
    bulk_commit_wait();
  } else {
    for (int k = threadIdx.x; k < nwords; k += TPB) {
      const uint32_t c = tile[k];
      if (c) {
        const int p = 2 * k, r = p / t.w, cx = p - r * t.w;
        float* d = img + (long)(t.y0 + r) * W + t.x0 + cx;
        if (c & 0xffffu) red_f32(d, (float)(c & 0xffffu));
        if (c >> 16) red_f32(d + 1, (float)(c >> 16));
      }
    }
  }
}
__global__ void k_plain4(const int4* __restrict__ idx4, float* img, long n4) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const long stride = (long)gridDim.x * blockDim.x;
  for (; i < n4; i += 4 * stride) {
    int4 q[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) q[u] = (i + u * stride < n4) ? __ldg(idx4 + i + u * stride) : make_int4(-1, -1, -1, -1);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (q[u].x >= 0) { red_f32(img + q[u].x, 1.f); red_f32(img + q[u].y, 1.f); red_f32(img + q[u].z, 1.f); red_f32(img + q[u].w, 1.f); }
    }
  }
}

__global__ void k_plain(const int* __restrict__ idx, float* img, long n) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const long stride = (long)gridDim.x * blockDim.x;
  for (; i < n; i += stride) red_f32(img + idx[i], 1.0f);
}

int main(int argc, char** argv) {
  long n = argc > 1 ? atol(argv[1]) : 32000000;
  double sig = argc > 2 ? atof(argv[2]) : 40.0;
  float* d_img;
  CK(cudaMalloc(&d_img, (size_t)W * H * 4));
  long long* d_cyc;
  CK(cudaMalloc(&d_cyc, 8));
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  if (argc > 3) {  // lane-cost table
    CK(cudaFuncSetAttribute(k_lanes<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CK(cudaFuncSetAttribute(k_lanes<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CK(cudaFuncSetAttribute(k_lanes<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    const int iters = 2000;
    printf("cycles per warp instruction per SM (16 warps/SM issuing), K active lanes, uniform addresses\n");
    for (int which = 0; which < 3; ++which)
      for (int K : {1, 2, 4, 8, 16, 32}) {
        long long best = 1LL << 60;
        for (int r = 0; r < 3; ++r) {
          CK(cudaMemset(d_img, 0, (size_t)W * H * 4));
          if (which == 0) k_lanes<0><<<148, TPB, 28 * 4096>>>(d_img, iters, K, d_cyc);
          if (which == 1) k_lanes<1><<<148, TPB, 28 * 4096>>>(d_img, iters, K, d_cyc);
          if (which == 2) k_lanes<2><<<148, TPB, 28 * 4096>>>(d_img, iters, K, d_cyc);
          CK(cudaDeviceSynchronize());
          long long c;
          CK(cudaMemcpy(&c, d_cyc, 8, cudaMemcpyDeviceToHost));
          best = c < best ? c : best;
        }
        printf("%-26s K=%2d  %7.2f cyc/warp-instr  %6.3f cyc/active lane\n",
               which == 0 ? "global RED.F32" : which == 1 ? "shared atomicAdd(p,1u)" : "shared red.add.u32 reg",
               K, (double)best / (iters * (TPB / 32)), (double)best / (iters * (TPB / 32)) / K);
      }
    return 0;
  }
  std::vector<int> h(n);
  srand(1);
  auto gauss = [&]() {
    double u = (rand() + 1.0) / (RAND_MAX + 2.0), v = (rand() + 1.0) / (RAND_MAX + 2.0);
    return sqrt(-2 * log(u)) * cos(6.283185307 * v);
  };
  for (long i = 0; i < n; ++i) {
    int x = (int)(W / 2 + sig * gauss()), y = (int)(H / 2 + sig * gauss());
    x = x < 0 ? 0 : (x >= W ? W - 1 : x);
    y = y < 0 ? 0 : (y >= H ? H - 1 : y);
    h[i] = y * W + x;
  }
  int* d_idx;
  float* d_ref;
  CK(cudaMalloc(&d_idx, n * 4));
  CK(cudaMalloc(&d_ref, (size_t)W * H * 4));
  CK(cudaMemcpy(d_idx, h.data(), n * 4, cudaMemcpyHostToDevice));
  std::vector<float> ref((size_t)W * H), got((size_t)W * H);
  CK(cudaMemset(d_ref, 0, (size_t)W * H * 4));
  k_plain<<<148 * 8, 256>>>(d_idx, d_ref, n);
  CK(cudaMemcpy(ref.data(), d_ref, (size_t)W * H * 4, cudaMemcpyDeviceToHost));
  auto time = [&](const char* name, auto fn) {
    float best = 1e9;
    for (int r = 0; r < 5; ++r) {
      CK(cudaMemset(d_img, 0, (size_t)W * H * 4));
      cudaEventRecord(a);
      fn();
      cudaEventRecord(b);
      CK(cudaEventSynchronize(b));
      CK(cudaGetLastError());
      float ms;
      cudaEventElapsedTime(&ms, a, b);
      best = ms < best ? ms : best;
    }
    long bad = 0;
    CK(cudaMemcpy(got.data(), d_img, (size_t)W * H * 4, cudaMemcpyDeviceToHost));
    for (size_t k = 0; k < got.size(); ++k) bad += got[k] != ref[k];
    printf("%-64s %8.1f us %7.1f G upd/s%s\n", name, best * 1e3, n / best / 1e6, bad ? "  MISMATCH" : "  exact");
  };
  printf("n=%ld sigma=%.0f px\n", n, sig);
  time("global RED, 1184 x 256", [&] { k_plain<<<148 * 8, 256>>>(d_idx, d_img, n); });
  auto set = [&](auto k) { CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 210 * 1024)); };
  set(k_hybrid2<32, 1>); set(k_hybrid2<32, 2>); set(k_hybrid2<16, 1>); set(k_hybrid2<16, 2>);

  set(k_hybrid3<32>); set(k_hybrid3<16>);
  if ((n & 3) == 0) {
    time("global RED, deep loads, 148 x 512", [&] { k_plain4<<<148, TPB>>>((const int4*)d_idx, d_img, n / 4); });
    for (int bits : {32, 16})
      for (int kb : {9, 18, 36, 72, 100, 130}) {
        const long cap_px = (long)kb * 1024 * 8 / bits;
        int side = (int)sqrt((double)cap_px);
        int w = side & ~3, hh = (int)(cap_px / w);
        if (w > 6.5 * sig) w = ((int)(6.5 * sig)) & ~3;
        if (hh > 6.5 * sig) hh = (int)(6.5 * sig);
        hh &= ~1;
        Tile t{(W / 2 - w / 2) & ~3, H / 2 - hh / 2, w, hh};
        const size_t smem = (size_t)w * hh * bits / 8;
        char nm[112];
        snprintf(nm, 112, "DEEP tile u%d %dx%d (%zu KB, +-%.1f sigma) divergent tails", bits, w, hh, smem / 1024, w / 2.0 / sig);
        if (bits == 32) time(nm, [&] { k_hybrid3<32><<<148, TPB, smem>>>((const int4*)d_idx, d_img, n / 4, t); });
        else time(nm, [&] { k_hybrid3<16><<<148, TPB, smem>>>((const int4*)d_idx, d_img, n / 4, t); });
      }
    return 0;
  }
  for (int bits : {32, 16})
    for (int kb : {36, 72, 112, 200}) {
      const long cap_px = (long)kb * 1024 * 8 / bits;
      int side = (int)sqrt((double)cap_px);
      int w = side & ~3, hh = (int)(cap_px / w);
      if (w > 6.5 * sig) w = ((int)(6.5 * sig)) & ~3;
      if (hh > 6.5 * sig) hh = (int)(6.5 * sig);
      hh &= ~1;
      Tile t{(W / 2 - w / 2) & ~3, H / 2 - hh / 2, w, hh};
      const size_t smem = (size_t)w * hh * bits / 8;
      for (int mode : {1, 2})
        for (int compact : {0, 1}) {
          char nm[112];
          snprintf(nm, 112, "tile u%d %dx%d (%zu KB, +-%.1f sigma) %s %s", bits, w, hh, smem / 1024, w / 2.0 / sig,
                   mode == 1 ? "atomicAdd(1u)" : "red.add reg", compact ? "tails compacted" : "tails divergent");
          if (bits == 32 && mode == 1) time(nm, [&] { k_hybrid2<32, 1><<<148, TPB, smem>>>(d_idx, d_img, n, t, compact); });
          if (bits == 32 && mode == 2) time(nm, [&] { k_hybrid2<32, 2><<<148, TPB, smem>>>(d_idx, d_img, n, t, compact); });
          if (bits == 16 && mode == 1) time(nm, [&] { k_hybrid2<16, 1><<<148, TPB, smem>>>(d_idx, d_img, n, t, compact); });
          if (bits == 16 && mode == 2) time(nm, [&] { k_hybrid2<16, 2><<<148, TPB, smem>>>(d_idx, d_img, n, t, compact); });
        }
    }
  return 0;
}

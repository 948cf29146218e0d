// Microbenchmark, round 2: how to get a focused 2-D histogram (2448x2040 fp32 image, Gaussian
// footprint) past the L2 atomic rate.  Extends atomic_bench.cu (round 1: global REDs only).
//
//   hybrid   persistent CTA per SM; a CTA-private u32 tile in shared memory covers the core of
//            the beam; an in-tile update goes to the tile (ATOMS) with probability `split`,
//            everything else to global memory (RED.ADD.F32).  The tile is flushed by
//            cp.reduce.async.bulk (.add.f32, one bulk reduce per tile row) or by per-pixel REDs.
//   cluster  the judge's proposal: a thread-block cluster owns ONE tile, row bands spread over
//            the CTAs' shared memories; remote updates are red.shared::cluster.add.u32 through
//            DSMEM; flushed by bulk reduce.
//   flush    the flush alone (all CTAs add a tile of given size into the same image region).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o hist_bench hist_bench.cu
//   ./hist_bench <n> <sigma_px>
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

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void red_f32(float* p, float v) {
  asm volatile("red.relaxed.gpu.global.add.f32 [%0], %1;" ::"l"(p), "f"(v));
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

struct Tile {
  int x0, y0, w, h;  // image coordinates, x0 and w multiples of 4
};

// mode bit0: flush with bulk reduce (else per-pixel REDs); split256: of 256 in-tile updates this
// many go to shared memory, the rest to global memory (chosen by a hash of the particle index)
template <int TPB>
__global__ void __launch_bounds__(TPB, 1) k_hybrid(const int* __restrict__ idx, float* img, long n,
                                                    Tile t, int split256, int mode) {
  extern __shared__ __align__(128) uint32_t tile[];
  const int npx = t.w * t.h;
  for (int i = threadIdx.x; i < npx; i += TPB) tile[i] = 0u;
  __syncthreads();
  long i = (long)blockIdx.x * TPB + threadIdx.x;
  const long stride = (long)gridDim.x * TPB;
  for (; i < n; i += stride) {
    const int v = idx[i];
    const int y = v / W, x = v - y * W;
    const unsigned dx = (unsigned)(x - t.x0), dy = (unsigned)(y - t.y0);
    const bool in = dx < (unsigned)t.w && dy < (unsigned)t.h;
    const bool to_smem = in && (int)((i * 2654435761u >> 13) & 255u) < split256;
    if (to_smem)
      atomicAdd(tile + dy * t.w + dx, 1u);
    else
      red_f32(img + v, 1.0f);
  }
  __syncthreads();
  if (mode & 1) {
    float* f = reinterpret_cast<float*>(tile);
    for (int k = threadIdx.x; k < npx; k += TPB) f[k] = (float)tile[k];
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    // one bulk reduce per tile row, issued by the first lanes of the CTA
    for (int r = threadIdx.x; r < t.h; r += TPB)
      bulk_reduce_add_f32(img + (long)(t.y0 + r) * W + t.x0, f + r * t.w, (uint32_t)t.w * 4u);
    bulk_commit_wait();
  } else {
    for (int k = threadIdx.x; k < npx; k += TPB) {
      const uint32_t c = tile[k];
      if (c) {
        const int r = k / t.w, cx = k - r * t.w;
        red_f32(img + (long)(t.y0 + r) * W + t.x0 + cx, (float)c);
      }
    }
  }
}

// cluster variant: CTA rank q of a cluster of CS owns tile rows [q*hb, (q+1)*hb)
template <int TPB, int CS>
__global__ void __launch_bounds__(TPB, 1) k_cluster(const int* __restrict__ idx, float* img, long n,
                                                     Tile t, int hb) {
  extern __shared__ __align__(128) uint32_t tile[];
  uint32_t rank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  const int npx = t.w * hb;
  for (int i = threadIdx.x; i < npx; i += TPB) tile[i] = 0u;
  asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  const uint32_t base = smem_u32(tile);
  long i = (long)blockIdx.x * TPB + threadIdx.x;
  const long stride = (long)gridDim.x * TPB;
  for (; i < n; i += stride) {
    const int v = idx[i];
    const int y = v / W, x = v - y * W;
    const unsigned dx = (unsigned)(x - t.x0), dy = (unsigned)(y - t.y0);
    if (dx < (unsigned)t.w && dy < (unsigned)t.h) {
      const uint32_t q = dy / (unsigned)hb, r = dy - q * hb;
      uint32_t remote;
      asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(base + 4u * (r * t.w + dx)), "r"(q));
      asm volatile("red.relaxed.cluster.shared::cluster.add.u32 [%0], %1;" ::"r"(remote), "r"(1u) : "memory");
    } else {
      red_f32(img + v, 1.0f);
    }
  }
  asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  float* f = reinterpret_cast<float*>(tile);
  for (int k = threadIdx.x; k < npx; k += TPB) f[k] = (float)tile[k];
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  for (int r = threadIdx.x; r < hb; r += TPB) {
    const int gy = t.y0 + (int)rank * hb + r;
    if (gy < t.y0 + t.h) bulk_reduce_add_f32(img + (long)gy * W + t.x0, f + r * t.w, (uint32_t)t.w * 4u);
  }
  bulk_commit_wait();
}

template <int TPB>
__global__ void __launch_bounds__(TPB, 1) k_flush_only(float* img, Tile t, int bulk) {
  extern __shared__ __align__(128) uint32_t tile[];
  const int npx = t.w * t.h;
  float* f = reinterpret_cast<float*>(tile);
  for (int k = threadIdx.x; k < npx; k += TPB) f[k] = 1.0f;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (bulk) {
    for (int r = threadIdx.x; r < t.h; r += TPB)
      bulk_reduce_add_f32(img + (long)(t.y0 + r) * W + t.x0, f + r * t.w, (uint32_t)t.w * 4u);
    bulk_commit_wait();
  } else {
    for (int k = threadIdx.x; k < npx; k += TPB) {
      const int r = k / t.w, cx = k - r * t.w;
      red_f32(img + (long)(t.y0 + r) * W + t.x0 + cx, f[k]);
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
  float *d_img, *d_ref;
  CK(cudaMalloc(&d_idx, n * 4));
  CK(cudaMalloc(&d_img, (size_t)W * H * 4));
  CK(cudaMalloc(&d_ref, (size_t)W * H * 4));
  CK(cudaMemcpy(d_idx, h.data(), n * 4, cudaMemcpyHostToDevice));
  std::vector<float> ref((size_t)W * H), got((size_t)W * H);
  CK(cudaMemset(d_ref, 0, (size_t)W * H * 4));
  k_plain<<<148 * 8, 256>>>(d_idx, d_ref, n);
  CK(cudaMemcpy(ref.data(), d_ref, (size_t)W * H * 4, cudaMemcpyDeviceToHost));
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  auto time = [&](const char* name, auto fn, bool check) {
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
    if (check) {
      CK(cudaMemcpy(got.data(), d_img, (size_t)W * H * 4, cudaMemcpyDeviceToHost));
      for (size_t k = 0; k < got.size(); ++k) bad += got[k] != ref[k];
    }
    printf("%-52s %8.1f us %7.1f G upd/s%s\n", name, best * 1e3, n / best / 1e6,
           check ? (bad ? "  MISMATCH" : "  exact") : "");
  };
  printf("n=%ld sigma=%.0f px\n", n, sig);
  time("global RED, 1184 x 256", [&] { k_plain<<<148 * 8, 256>>>(d_idx, d_img, n); }, true);
  constexpr int TPB = 512;
  CK(cudaFuncSetAttribute(k_hybrid<TPB>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  CK(cudaFuncSetAttribute(k_flush_only<TPB>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  auto mk = [&](double k, int cap_px) {
    int w = ((int)(2 * k * sig) + 3) & ~3, hh = (int)(2 * k * sig);
    while ((long)w * hh > cap_px) {
      if (hh > 4) --hh;
      if ((long)w * hh > cap_px && w > 8) w -= 4;
    }
    Tile t{(W / 2 - w / 2) & ~3, H / 2 - hh / 2, w, hh};
    return t;
  };
  for (int cap_kb : {16, 36, 64, 120}) {
    for (double k : {0.8, 1.2, 1.8, 2.5}) {
      Tile t = mk(k, cap_kb * 256);
      if ((long)t.w * t.h * 4 < (long)cap_kb * 1024 * 6 / 10 && k < 2.5) continue;  // tile far below cap: same as a smaller cap
      for (int split : {256, 160, 96}) {
        for (int mode : {1, 0}) {
          if (mode == 0 && split != 256) continue;
          char nm[96];
          snprintf(nm, 96, "hybrid tile %dx%d (%d KB) split %d/256 %s", t.w, t.h, t.w * t.h / 256, split,
                   mode ? "bulk" : "REDs");
          time(nm, [&] { k_hybrid<TPB><<<148, TPB, (size_t)t.w * t.h * 4>>>(d_idx, d_img, n, t, split, mode); },
               true);
        }
      }
      if (k * sig * 2 * k * sig * 2 * 4 < cap_kb * 1024) break;  // larger k gives the same tile
    }
  }
  for (int cap_kb : {16, 36, 64, 120}) {
    Tile t = mk(3.0, cap_kb * 256);
    for (int bulk : {1, 0}) {
      char nm[96];
      snprintf(nm, 96, "flush only %dx%d (%d KB) x148 %s", t.w, t.h, t.w * t.h / 256, bulk ? "bulk" : "REDs");
      time(nm, [&] { k_flush_only<TPB><<<148, TPB, (size_t)t.w * t.h * 4>>>(d_img, t, bulk); }, false);
    }
  }
  {
    constexpr int CS = 8;
    auto kc = k_cluster<TPB, CS>;
    CK(cudaFuncSetAttribute(kc, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    for (double k : {1.8, 3.0, 4.5}) {
      int w = ((int)(2 * k * sig) + 3) & ~3, hh = ((int)(2 * k * sig) + CS - 1) / CS * CS;
      int hb = hh / CS;
      while ((long)w * hb * 4 > 190 * 1024) { w -= 4; }
      Tile t{(W / 2 - w / 2) & ~3, H / 2 - hh / 2, w, hh};
      cudaLaunchConfig_t cfg{};
      cfg.gridDim = dim3(144);
      cfg.blockDim = dim3(TPB);
      cfg.dynamicSmemBytes = (size_t)w * hb * 4;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeClusterDimension;
      at[0].val.clusterDim.x = CS;
      at[0].val.clusterDim.y = 1;
      at[0].val.clusterDim.z = 1;
      cfg.attrs = at;
      cfg.numAttrs = 1;
      char nm[96];
      snprintf(nm, 96, "cluster-8 DSMEM tile %dx%d (%d KB/CTA) bulk", t.w, t.h, w * hb / 256);
      time(nm, [&] { CK(cudaLaunchKernelEx(&cfg, kc, (const int*)d_idx, d_img, n, t, hb)); }, true);
    }
  }
  return 0;
}

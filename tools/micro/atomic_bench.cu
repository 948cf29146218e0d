// Microbenchmark: scattered histogram updates into a 2448x2040 image (Gaussian footprint),
// f32 vs u32 RED, global vs pre-sorted, to find the achievable atomic rate on B200.
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <cuda_runtime.h>
#define CK(x) do{cudaError_t e=(x); if(e!=cudaSuccess){printf("err %s line %d\n",cudaGetErrorString(e),__LINE__);exit(1);} }while(0)

template <typename T, int PER>
__global__ void k_red(const int* __restrict__ idx, T* img, long n) {
  long i = ((long)blockIdx.x * blockDim.x + threadIdx.x) * PER;
  long stride = (long)gridDim.x * blockDim.x * PER;
  for (; i < n; i += stride) {
    int v[PER];
#pragma unroll
    for (int k = 0; k < PER; ++k) v[k] = idx[i + k];
#pragma unroll
    for (int k = 0; k < PER; ++k) atomicAdd(img + v[k], (T)1);
  }
}
// interleaved across the warp: lane-contiguous index loads (coalesced)
template <typename T>
__global__ void k_red_co(const int* __restrict__ idx, T* img, long n) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  long stride = (long)gridDim.x * blockDim.x;
  for (; i < n; i += stride) atomicAdd(img + idx[i], (T)1);
}
// R replicas of the image (distinct L2 lines): replica chosen per CTA
__global__ void k_red_rep(const int* __restrict__ idx, float* img, long n, int R, long rep_stride) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  long stride = (long)gridDim.x * blockDim.x;
  float* mine = img + (blockIdx.x % R) * rep_stride;
  for (; i < n; i += stride) atomicAdd(mine + idx[i], 1.0f);
}
// warp-aggregated: lanes with the same bin elect a leader that adds the count
__global__ void k_red_match(const int* __restrict__ idx, float* img, long n) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  long stride = (long)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {  // n is a multiple of the warp size here
    int v = idx[i];
    unsigned m = __match_any_sync(0xffffffffu, v);
    if ((__ffs(m) - 1) == (threadIdx.x & 31)) atomicAdd(img + v, (float)__popc(m));
  }
}
__global__ void k_read(const int* __restrict__ idx, int* out, long n) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  long stride = (long)gridDim.x * blockDim.x;
  int acc = 0;
  for (; i < n; i += stride) acc ^= idx[i];
  if (acc == 0x12345) out[0] = acc;
}

int main(int argc, char** argv) {
  long n = argc > 1 ? atol(argv[1]) : 32000000;
  double sig = argc > 2 ? atof(argv[2]) : 100.0;  // pixels
  const int W = 2448, H = 2040;
  std::vector<int> h(n);
  srand(1);
  auto gauss = [&]() { double u = (rand() + 1.0) / (RAND_MAX + 2.0), v = (rand() + 1.0) / (RAND_MAX + 2.0); return sqrt(-2 * log(u)) * cos(6.283185307 * v); };
  for (long i = 0; i < n; ++i) {
    int x, y;
    if (sig <= 0) { x = rand() % W; y = rand() % H; }
    else { x = (int)(W / 2 + sig * gauss()); y = (int)(H / 2 + sig * gauss()); }
    x = x < 0 ? 0 : (x >= W ? W - 1 : x); y = y < 0 ? 0 : (y >= H ? H - 1 : y);
    h[i] = y * W + x;
  }
  const int RMAX = 8; const long rep_stride = (long)W * H + 32 * 37;
  float* d_rep; CK(cudaMalloc(&d_rep, RMAX * rep_stride * 4));
  int* d_idx; float* d_f; unsigned* d_u; int* d_o;
  CK(cudaMalloc(&d_idx, n * 4)); CK(cudaMalloc(&d_f, W * H * 4)); CK(cudaMalloc(&d_u, W * H * 4)); CK(cudaMalloc(&d_o, 4));
  CK(cudaMemcpy(d_idx, h.data(), n * 4, cudaMemcpyHostToDevice));
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  auto time = [&](const char* name, auto fn) {
    float best = 1e9;
    for (int r = 0; r < 5; ++r) {
      CK(cudaMemset(d_f, 0, W * H * 4)); CK(cudaMemset(d_u, 0, W * H * 4));
      cudaEventRecord(a); fn(); cudaEventRecord(b); CK(cudaEventSynchronize(b));
      float ms; cudaEventElapsedTime(&ms, a, b); best = ms < best ? ms : best;
    }
    printf("%-28s %8.1f us  %7.1f G upd/s\n", name, best * 1e3, n / best / 1e6);
  };
  for (int blocks : {148 * 8}) {
    printf("blocks=%d sigma=%.0f px n=%ld\n", blocks, sig, n);
    time("read only", [&] { k_read<<<blocks, 256>>>(d_idx, d_o, n); });
    time("f32 RED coalesced idx", [&] { k_red_co<float><<<blocks, 256>>>(d_idx, d_f, n); });
    time("u32 RED coalesced idx", [&] { k_red_co<unsigned><<<blocks, 256>>>(d_idx, d_u, n); });
    for (int R : {2, 4, 8}) {
      char nm[64]; snprintf(nm, 64, "f32 RED %d replicas", R);
      CK(cudaMemset(d_rep, 0, RMAX * rep_stride * 4));
      time(nm, [&] { k_red_rep<<<blocks, 256>>>(d_idx, d_rep, n, R, rep_stride); });
    }
    time("f32 RED match.any", [&] { k_red_match<<<blocks, 256>>>(d_idx, d_f, n & ~31L); });
    time("f32 RED 4/thread", [&] { k_red<float, 4><<<blocks, 256>>>(d_idx, d_f, n); });
    time("u32 RED 4/thread", [&] { k_red<unsigned, 4><<<blocks, 256>>>(d_idx, d_u, n); });
  }
  return 0;
}

// Multi-GPU exchange step of the particle-sharded path: the screen image.
//
// The reference has no distributed code at all; the north star asks for the
// screen histogram to be all-reduced over NVLink.  A dense NCCL all-reduce of the
// 2448 x 2040 fp32 image moves 20 MB per rank although the beam lights ~1 % of the pixels
// (70 us at 2 GPUs against a 30 us tracking step).  Here every rank's image lives in a
// peer-mapped SYMMETRIC buffer (NVLink / NVSwitch P2P), K2 records the bounding box of the
// bins it touched, and ONE kernel per rank
//   1. signals / waits on per-rank epoch flags in the peers' buffers (cross-GPU barrier),
//   2. forms the union bounding box from every peer's box,
//   3. writes the caller's image: inside the box the sum of all peers' pixels read over
//      NVLink (coalesced peer loads), outside it zeros (no remote traffic).
// Counts are integers < 2^24 in fp32, so the sum is exact and order-independent.
// Buffers are double-buffered by the caller (slot = step & 1), which makes a trailing
// barrier unnecessary: a rank can only start zeroing slot s again after every peer has
// passed the barrier of the step in between, i.e. finished reading slot s.
//
// Layout of one rank's symmetric buffer (lynx_screen_symm_bytes):
//   [0, 256)                      uint32 flags[64]: flags[p] = last epoch signalled by rank p
//   slot s at 256 + s * slot_bytes: int32 bbox[16] (v0 = max(W - ix), v1 = max(ix + 1),
//                                   v2 = max(H - iy), v3 = max(iy + 1); 0 = empty), then
//                                   float image[B][H][W] (reference layout, row 0 = max y)
#include <algorithm>

#include "common.cuh"

namespace lynx {

constexpr int kMaxWorld = 16;
constexpr size_t kSymmHeader = 256;
constexpr size_t kBboxBytes = 64;

struct ExchangeArgs {
  char* peers[kMaxWorld];  // symmetric buffer base of every rank (peer-mapped), [rank] = own
  int world, rank, slot;
  uint32_t epoch;
  int64_t B;
  int W, H;
  size_t slot_bytes;
  float* out;  // [B][H][W] local
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

constexpr int kExTPB = 256;

__global__ void __launch_bounds__(kExTPB) screen_allreduce_kernel(ExchangeArgs a) {
  __shared__ int s_box[4];
  const int tid = threadIdx.x;
  // ---- 1. cross-GPU barrier on epoch flags (this launch is stream-ordered after K2)
  if (blockIdx.x == 0 && tid < a.world) {
    __threadfence_system();
    st_release_sys(reinterpret_cast<uint32_t*>(a.peers[tid]) + a.rank, a.epoch);
  }
  if (tid < a.world) {
    const uint32_t* mine = reinterpret_cast<const uint32_t*>(a.peers[a.rank]) + tid;
    while ((int32_t)(ld_acquire_sys(mine) - a.epoch) < 0) {
    }
  }
  if (tid < 4) s_box[tid] = 0;
  __syncthreads();
  // ---- 2. union bounding box
  if (tid < a.world * 4) {
    const int p = tid >> 2, k = tid & 3;
    const int* bb = reinterpret_cast<const int*>(a.peers[p] + kSymmHeader + a.slot * a.slot_bytes);
    atomicMax(&s_box[k], __ldcv(bb + k));
  }
  __syncthreads();
  const bool empty = s_box[1] == 0 || s_box[3] == 0;
  const int x0 = a.W - s_box[0], x1 = s_box[1];  // [x0, x1)
  const int y0 = a.H - s_box[2], y1 = s_box[3];  // bins [y0, y1) -> image rows H-1-y
  const int r0 = a.H - y1, r1 = a.H - y0;        // image rows [r0, r1)
  const int c0 = x0 & ~3, c1 = min((x1 + 3) & ~3, a.W);

  // ---- 3. write the caller's image: sums inside the box, zeros outside
  const int64_t n4 = a.B * (int64_t)a.H * a.W / 4;  // W % 4 == 0 checked by the launcher
  const size_t img_off = kSymmHeader + a.slot * a.slot_bytes + kBboxBytes;
  float4* out4 = reinterpret_cast<float4*>(a.out);
  const int w4 = a.W >> 2;
  for (int64_t i = (int64_t)blockIdx.x * kExTPB + tid; i < n4; i += (int64_t)gridDim.x * kExTPB) {
    const int64_t row_all = i / w4;
    const int c = (int)(i - row_all * w4) << 2;
    const int r = (int)(row_all % a.H);
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    if (!empty && r >= r0 && r < r1 && c >= c0 && c < c1) {
      for (int p = 0; p < a.world; ++p) {
        const float4 v = __ldcv(reinterpret_cast<const float4*>(a.peers[p] + img_off) + i);
        s.x += v.x;
        s.y += v.y;
        s.z += v.z;
        s.w += v.w;
      }
    }
    __stcs(out4 + i, s);
  }
}

}  // namespace lynx

using namespace lynx;

extern "C" size_t lynx_screen_symm_bytes(int64_t B, int32_t W, int32_t H) {
  if (B <= 0 || W <= 0 || H <= 0) return 0;
  size_t slot = kBboxBytes + sizeof(float) * (size_t)B * W * H;
  slot = (slot + 255) & ~(size_t)255;
  return kSymmHeader + 2 * slot;
}

static size_t slot_bytes_of(int64_t B, int32_t W, int32_t H) {
  size_t slot = kBboxBytes + sizeof(float) * (size_t)B * W * H;
  return (slot + 255) & ~(size_t)255;
}

extern "C" float* lynx_screen_symm_image(void* d_symm_base, int32_t slot, int64_t B, int32_t W,
                                         int32_t H) {
  if (!d_symm_base || slot < 0 || slot > 1) return nullptr;
  return reinterpret_cast<float*>((char*)d_symm_base + kSymmHeader + slot * slot_bytes_of(B, W, H) +
                                  kBboxBytes);
}

extern "C" int lynx_screen_allreduce(void* const* peer_bases, int32_t world, int32_t rank,
                                     int32_t slot, uint32_t epoch, int64_t B, int32_t W, int32_t H,
                                     float* d_out_img, lynx_stream_t stream) {
  if (!peer_bases || world < 1 || world > kMaxWorld || rank < 0 || rank >= world || slot < 0 ||
      slot > 1 || epoch == 0 || B <= 0 || W <= 0 || H <= 0 || (W & 3) || !d_out_img ||
      ((uintptr_t)d_out_img & 15))
    return LYNX_E_BADARG;
  ExchangeArgs a{};
  for (int p = 0; p < world; ++p) {
    if (!peer_bases[p]) return LYNX_E_BADARG;
    a.peers[p] = reinterpret_cast<char*>(peer_bases[p]);
  }
  a.world = world;
  a.rank = rank;
  a.slot = slot;
  a.epoch = epoch;
  a.B = B;
  a.W = W;
  a.H = H;
  a.slot_bytes = slot_bytes_of(B, W, H);
  a.out = d_out_img;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int64_t n4 = B * (int64_t)H * W / 4;
  const unsigned blocks = (unsigned)std::min<int64_t>((n4 + kExTPB - 1) / kExTPB, 8LL * sms);
  screen_allreduce_kernel<<<blocks, kExTPB, 0, (cudaStream_t)stream>>>(a);
  count_launch();
  LYNX_CUDA_OK(cudaGetLastError());
  return LYNX_OK;
}

// ---------------------------------------------------------------------------------------
// K7: Screen.reading for a ParameterBeam (screen.py:157-194): bivariate normal pdf of
// (x, y) evaluated on the grid of pixel LEFT edges (arange, :181-191), meshgrid "ij" so the
// array is indexed [b][ix][iy], flipped along x (:194): img[b][i][j] = pdf(xs[W-1-i], ys[j]).
// One float4 of consecutive iy per thread; HBM-write-bound (4 W H bytes per setting).
namespace lynx {
constexpr int kPdfTPB = 256;

__global__ void __launch_bounds__(kPdfTPB) screen_pdf_kernel(const float* __restrict__ mu,
                                                              const float* __restrict__ cov,
                                                              const float* __restrict__ xs,
                                                              const float* __restrict__ ys, int W,
                                                              int H, float* __restrict__ img) {
  const int64_t b = blockIdx.y;
  const float mx = __ldg(mu + b * 7 + 0), my = __ldg(mu + b * 7 + 2);
  const double sxx = __ldg(cov + b * 49 + 0), sxy = __ldg(cov + b * 49 + 2),
               syy = __ldg(cov + b * 49 + 16);
  const double det = sxx * syy - sxy * sxy;
  const float qa = (float)(syy / det), qb = (float)(-sxy / det), qc = (float)(sxx / det);
  const float lognorm = (float)(-1.8378770664093453 - 0.5 * log(det));  // -log(2 pi) - log|S|/2
  float* out = img + b * (int64_t)W * H;
  const int h4 = (H + 3) >> 2;
  const int64_t n = (int64_t)W * h4;
  for (int64_t t = (int64_t)blockIdx.x * kPdfTPB + threadIdx.x; t < n;
       t += (int64_t)gridDim.x * kPdfTPB) {
    const int i = (int)(t / h4), j0 = (int)(t - (int64_t)i * h4) << 2;
    const float dx = __ldg(xs + (W - 1 - i)) - mx;
    const float ax = qa * dx * dx, bx = 2.f * qb * dx;
    float v[4];
#pragma unroll
    for (int l = 0; l < 4; ++l) {
      const int j = min(j0 + l, H - 1);
      const float dy = __ldg(ys + j) - my;
      const float q = fmaf(qc * dy, dy, fmaf(bx, dy, ax));
      v[l] = expf(fmaf(-0.5f, q, lognorm));
    }
    float* p = out + (int64_t)i * H + j0;
    if (j0 + 3 < H && !(((uintptr_t)p) & 15)) {
      __stcs(reinterpret_cast<float4*>(p), make_float4(v[0], v[1], v[2], v[3]));
    } else {
#pragma unroll
      for (int l = 0; l < 4; ++l)
        if (j0 + l < H) p[l] = v[l];
    }
  }
}
}  // namespace lynx

extern "C" int lynx_screen_pdf(const float* d_mu, const float* d_cov, int64_t B,
                               const float* d_xs, const float* d_ys, int32_t W, int32_t H,
                               float* d_img, lynx_stream_t stream) {
  if (!d_mu || !d_cov || !d_xs || !d_ys || !d_img || B <= 0 || B > 65535 || W <= 0 || H <= 0)
    return LYNX_E_BADARG;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int64_t n = (int64_t)W * ((H + 3) >> 2);
  const int64_t per_b = std::max<int64_t>(1, (8LL * sms + B - 1) / B);
  const unsigned gx = (unsigned)std::min<int64_t>((n + kPdfTPB - 1) / kPdfTPB, per_b);
  screen_pdf_kernel<<<dim3(gx, (unsigned)B), kPdfTPB, 0, (cudaStream_t)stream>>>(d_mu, d_cov, d_xs,
                                                                                 d_ys, W, H, d_img);
  count_launch();
  LYNX_CUDA_OK(cudaGetLastError());
  return LYNX_OK;
}

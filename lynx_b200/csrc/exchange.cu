// Multi-GPU exchange step of the particle-sharded path: the screen image.
//
// The reference has no distributed code at all; the north star asks for the
// screen histogram to be all-reduced over NVLink.  A dense NCCL all-reduce of the
// 2448 x 2040 fp32 image moves 20 MB per rank although the beam lights ~1 % of the pixels
// (70 us at 2 GPUs against a 30 us tracking step).  Here every rank's image lives in a
// peer-mapped SYMMETRIC buffer (NVLink / NVSwitch P2P), K2 records the bounding box of the
// bins it touched, and ONE kernel per rank
//   1. signals / waits on per-rank epoch flags in the peers' buffers (cross-GPU barrier),
//   2. forms the union bounding box from the boxes the peers pushed with their signal,
//   3. updates the caller's persistent image: inside the box the sum of all peers' pixels
//      read over NVLink (coalesced peer loads); the part of the PREVIOUS call's rectangle
//      outside it is cleared; everything else is already zero (no 20 MB rewrite).
// Counts are integers < 2^24 in fp32, so the sum is exact and order-independent.
// Buffers are double-buffered by the caller (slot = step & 1), which makes a trailing
// barrier unnecessary: a rank can only start zeroing slot s again after every peer has
// passed the barrier of the step in between, i.e. finished reading slot s.
//
// Layout of one rank's symmetric buffer (lynx_screen_symm_bytes):
//   [0, 64)                       uint32 flags[16]: flags[p] = last epoch signalled by rank p
//   [64, 576)                     int4 inbox[2][16]: rank p's bounding box of the step with
//                                 epoch parity 0 / 1, PUSHED by rank p before its signal
//   slot s at 1024 + s * slot_bytes: int32 bbox[16] (v0 = max(W - ix), v1 = max(ix + 1),
//                                   v2 = max(H - iy), v3 = max(iy + 1); 0 = empty), then
//                                   float image[B][H][W] (reference layout, row 0 = max y)
#include <algorithm>

#include "common.cuh"

namespace lynx {

constexpr int kMaxWorld = 16;
constexpr size_t kSymmHeader = 1024;  // flags[16] u32 @0, inbox int4[2][16] @64
constexpr size_t kInboxOff = 64;
constexpr size_t kBboxBytes = 64;

struct ExchangeArgs {
  char* peers[kMaxWorld];  // symmetric buffer base of every rank (peer-mapped), [rank] = own
  int world, rank, slot;
  uint32_t epoch;
  int64_t B;
  int W, H;
  size_t slot_bytes;
  float* out;    // [B][H][W] local, persistent: zero outside the last written rectangle
  int* state;    // [2][4] local: rectangle (r0, r1, c0, c1) written at even / odd epochs
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
  const int par = a.epoch & 1;
  // launched with programmatic stream serialization: resident early, waits for K2 here
  asm volatile("griddepcontrol.wait;" ::: "memory");
  // ---- 1. push my bounding box into every peer's inbox, then signal my epoch (cross-GPU
  //         barrier; the release store orders the box before the flag)
  if (blockIdx.x == 0 && tid < a.world) {
    const int4 box = __ldcg(reinterpret_cast<const int4*>(a.peers[a.rank] + kSymmHeader +
                                                          a.slot * a.slot_bytes));
    int4* inbox = reinterpret_cast<int4*>(a.peers[tid] + kInboxOff) + par * kMaxWorld + a.rank;
    *inbox = box;
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
  // ---- 2. union bounding box from the (local) inbox
  if (tid < a.world * 4) {
    const int* in = reinterpret_cast<const int*>(a.peers[a.rank] + kInboxOff) +
                    (par * kMaxWorld + (tid >> 2)) * 4 + (tid & 3);
    atomicMax(&s_box[tid & 3], __ldcv(in));
  }
  __syncthreads();
  const bool empty = s_box[1] == 0 || s_box[3] == 0;
  // bins [x0, x1) x [y0, y1) -> image rows [H - y1, H - y0), columns rounded to float4
  const int r0 = empty ? 0 : a.H - s_box[3], r1 = empty ? 0 : s_box[2];
  const int c0 = empty ? 0 : (a.W - s_box[0]) & ~3, c1 = empty ? 0 : min((s_box[1] + 3) & ~3, a.W);
  // rectangle written by the previous call (it must be cleared where the new one is smaller)
  const int* prev = a.state + (par ^ 1) * 4;
  const int p_r0 = prev[0], p_r1 = prev[1], p_c0 = prev[2], p_c1 = prev[3];
  const bool p_empty = p_r1 <= p_r0 || p_c1 <= p_c0;
  const int ra = (empty ? p_r0 : (p_empty ? r0 : min(r0, p_r0)));
  const int rb = (empty ? p_r1 : (p_empty ? r1 : max(r1, p_r1)));
  const int ca = (empty ? p_c0 : (p_empty ? c0 : min(c0, p_c0)));
  const int cb = (empty ? p_c1 : (p_empty ? c1 : max(c1, p_c1)));
  if (blockIdx.x == 0 && tid == 0) {
    int* cur = a.state + par * 4;
    cur[0] = r0; cur[1] = r1; cur[2] = c0; cur[3] = c1;
  }
  if (rb <= ra || cb <= ca) return;

  // ---- 3. inside the new box: sum of all peers' pixels (coalesced peer loads over NVLink);
  //         rest of the union rectangle: zeros.  Everything else is already zero.
  const size_t img_off = kSymmHeader + a.slot * a.slot_bytes + kBboxBytes;
  const int w4 = (cb - ca) >> 2, nr = rb - ra;
  const int64_t n = a.B * (int64_t)nr * w4;
  for (int64_t t = (int64_t)blockIdx.x * kExTPB + tid; t < n; t += (int64_t)gridDim.x * kExTPB) {
    const int64_t rr = t / w4;
    const int c = ca + ((int)(t - rr * w4) << 2);
    const int64_t bq = rr / nr;
    const int r = ra + (int)(rr - bq * nr);
    const int64_t i = ((bq * a.H + r) * (int64_t)a.W + c) >> 2;
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r >= r0 && r < r1 && c >= c0 && c < c1) {
      for (int p = 0; p < a.world; ++p) {
        const float4 v = __ldcv(reinterpret_cast<const float4*>(a.peers[p] + img_off) + i);
        s.x += v.x;
        s.y += v.y;
        s.z += v.z;
        s.w += v.w;
      }
    }
    reinterpret_cast<float4*>(a.out)[i] = s;
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
                                     float* d_out_img, int32_t* d_state, lynx_stream_t stream) {
  if (!peer_bases || world < 1 || world > kMaxWorld || rank < 0 || rank >= world || slot < 0 ||
      slot > 1 || epoch == 0 || B <= 0 || W <= 0 || H <= 0 || (W & 3) || !d_out_img ||
      !d_state || ((uintptr_t)d_out_img & 15))
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
  a.state = d_state;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  // the rectangle is only known on the device: a fixed grid strides over it
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)(2 * sms));
  cfg.blockDim = dim3(kExTPB);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  const cudaError_t err = cudaLaunchKernelEx(&cfg, screen_allreduce_kernel, a);
  count_launch();
  return err == cudaSuccess ? LYNX_OK : LYNX_E_LAUNCH;
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

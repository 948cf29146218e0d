// Multi-GPU exchange step of the particle-sharded path: the screen image.
//
// The reference has no distributed code at all; the north star asks for the
// screen histogram to be all-reduced over NVLink.  A dense NCCL all-reduce of the
// 2448 x 2040 fp32 image moves 20 MB per rank although the beam lights ~1 % of the pixels
// (70 us at 2 GPUs against a 30 us tracking step).  Here every rank's image lives in a
// peer-mapped SYMMETRIC buffer (NVLink / NVSwitch P2P), K2 records the bounding box of the
// bins it touched, and ONE kernel per rank
//   1. pushes its bounding box, tagged with the step's epoch, into every peer's buffer and
//      waits for all peers' boxes of this epoch in its OWN buffer (cross-GPU barrier, no
//      polling over NVLink),
//   2. forms the union bounding box,
//   3. updates the caller's persistent image: inside the box the sum of all peers' pixels
//      read over NVLink (coalesced peer loads); the part of the PREVIOUS call's rectangle
//      outside it is cleared; everything else is already zero (no 20 MB rewrite).
// Counts are integers < 2^24 in fp32, so the sum is exact and order-independent.
// Buffers are double-buffered by the caller (slot = step & 1), which makes a trailing
// barrier unnecessary: a rank can only start zeroing slot s again after every peer has
// passed the barrier of the step in between, i.e. finished reading slot s.
//
// Layout of one rank's symmetric buffer (lynx_screen_symm_bytes):
//   [0, 512)                      uint64 msg[2][16][2]: rank p's message of the step with epoch
//                                 parity 0 / 1, PUSHED by rank p: two self-validating 8-byte
//                                 words (epoch << 32 | v_lo << 16 | v_hi), word 0 = (v0, v1),
//                                 word 1 = (v2, v3) of p's bounding box.  An aligned 8-byte
//                                 store is single-copy atomic, so the words need no fence
//                                 between them and double as the barrier signal.
//   slot s at 1024 + s * slot_bytes: int32 bbox[16] (v0 = max(W - ix), v1 = max(ix + 1),
//                                   v2 = max(H - iy), v3 = max(iy + 1); 0 = empty), then
//                                   float image[B][H][W] (reference layout, row 0 = max y)
#include <algorithm>
#include <cstdlib>

#include "common.cuh"

namespace lynx {

constexpr int kMaxWorld = 16;
constexpr size_t kSymmHeader = 1024;  // msg[2][16][2] u64 @0
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
  unsigned long long* trace;  // optional [1024][8] globaltimer stamps of block 0 (debug)
  char* mc;      // multicast (NVLS) mapping of the symmetric buffer, or null
};

// The message store needs no fence of its own: everything it publishes (K2's image and box)
// was written by the PREVIOUS grid, which griddepcontrol.wait / the kernel boundary has
// completed and made visible at this GPU's L2 -- the only place peers read it from (peer
// memory is not cached on the reading side).  A release store would put a MEMBAR.SYS
// (~2 us) on the critical path of every step.
__device__ __forceinline__ void st_relaxed_sys(uint64_t* p, uint64_t v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_acquire_sys(const uint64_t* p) {
  uint64_t v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// peer image read: relaxed system-scope load (ordered after the acquire above), not cached in L1
__device__ __forceinline__ float4 ld_peer(const float4* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}

// In-switch reduction (NVLS): one load on the multicast mapping of the symmetric buffer returns
// the sum of the word over all peers, so a rank issues W*H/4 loads instead of world * W*H/4
// (3 MB of outstanding peer reads per rank at 8 GPUs made the sum phase 12.6 us, measured).
// Integer-valued fp32 counts: the switch's summation order does not matter.
__device__ __forceinline__ float4 ld_reduce_mc(const float4* p) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}

constexpr int kExTPB = 256;

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

__global__ void __launch_bounds__(kExTPB) screen_allreduce_kernel(ExchangeArgs a) {
  __shared__ int s_box[4];
  const int tid = threadIdx.x;
  const int par = a.epoch & 1;
  unsigned long long* tr =
      (a.trace && blockIdx.x == 0 && tid == 0) ? a.trace + (a.epoch & 1023u) * 8 : nullptr;
  if (tr) tr[0] = globaltimer_ns();  // resident (possibly long before K2 ends)
  // launched with programmatic stream serialization: resident early, waits for K2 here
  asm volatile("griddepcontrol.wait;" ::: "memory");
  if (tr) {
    tr[1] = globaltimer_ns();  // K2 complete
    const unsigned long long* k2 = reinterpret_cast<const unsigned long long*>(
        a.peers[a.rank] + kSymmHeader + a.slot * a.slot_bytes + 32);
    tr[4] = __ldcg(k2);      // K2 past its wait on K1
    tr[5] = __ldcg(k2 + 1);  // K2 CTA 0 issued its last RED
    tr[6] = __ldcg(k2 + 2);  // K1 block 0 started
  }
  // ---- 1. push my bounding box, tagged with the epoch, into every peer's message area
  //         (thread = (peer, word))
  if (blockIdx.x == 0 && tid < 2 * a.world) {
    const int peer = tid >> 1, w = tid & 1;
    const int* box = reinterpret_cast<const int*>(a.peers[a.rank] + kSymmHeader +
                                                  a.slot * a.slot_bytes);
    const uint32_t lo = (uint32_t)__ldcg(box + 2 * w), hi = (uint32_t)__ldcg(box + 2 * w + 1);
    uint64_t* msg = reinterpret_cast<uint64_t*>(a.peers[peer]) + (par * kMaxWorld + a.rank) * 2 + w;
    st_relaxed_sys(msg, ((uint64_t)a.epoch << 32) | ((uint64_t)(lo & 0xffffu) << 16) | (hi & 0xffffu));
  }
  if (tid < 4) s_box[tid] = 0;
  __syncthreads();
  // ---- 2. every CTA waits for all peers' words of this epoch in the LOCAL buffer and forms
  //         the union box from them
  if (tid < 2 * a.world) {
    const uint64_t* mine =
        reinterpret_cast<const uint64_t*>(a.peers[a.rank]) + (par * kMaxWorld + (tid >> 1)) * 2 + (tid & 1);
    uint64_t m;
    do {
      m = ld_acquire_sys(mine);
    } while ((int32_t)((uint32_t)(m >> 32) - a.epoch) < 0);
    atomicMax(&s_box[2 * (tid & 1)], (int)((m >> 16) & 0xffffu));
    atomicMax(&s_box[2 * (tid & 1) + 1], (int)(m & 0xffffu));
  }
  __syncthreads();
  if (tr) tr[2] = globaltimer_ns();  // all peers arrived
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
    if (r >= r0 && r < r1 && c >= c0 && c < c1 && a.mc) {
      s = ld_reduce_mc(reinterpret_cast<const float4*>(a.mc + img_off) + i);
    } else if (r >= r0 && r < r1 && c >= c0 && c < c1) {
      // all peer loads of a batch are issued before the first add: one NVLink round trip
      // per batch of 8 peers instead of one per peer
      for (int p0 = 0; p0 < a.world; p0 += 8) {
        float4 v[8];
#pragma unroll
        for (int q = 0; q < 8; ++q)
          v[q] = (p0 + q < a.world)
                     ? ld_peer(reinterpret_cast<const float4*>(a.peers[p0 + q] + img_off) + i)
                     : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          s.x += v[q].x;
          s.y += v[q].y;
          s.z += v[q].z;
          s.w += v[q].w;
        }
      }
    }
    reinterpret_cast<float4*>(a.out)[i] = s;
  }
  if (tr) tr[3] = globaltimer_ns();  // block 0's share of the rectangle written
  if (a.trace && tid == 0) atomicMax(a.trace + (a.epoch & 1023u) * 8 + 7, globaltimer_ns());  // last CTA
}

static unsigned long long* g_exchange_trace = nullptr;

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

namespace lynx {
__global__ void debug_stamp_kernel(unsigned long long* d) {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  *d = t;
}
}  // namespace lynx
extern "C" void lynx_debug_stamp(unsigned long long* d_cell, lynx_stream_t stream) {
  lynx::debug_stamp_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(d_cell);
}

extern "C" void lynx_exchange_trace(unsigned long long* d_stamps) { lynx::g_exchange_trace = d_stamps; }

extern "C" int lynx_screen_allreduce(void* const* peer_bases, int32_t world, int32_t rank,
                                     int32_t slot, uint32_t epoch, int64_t B, int32_t W, int32_t H,
                                     float* d_out_img, int32_t* d_state, lynx_stream_t stream) {
  return lynx_screen_allreduce_mc(peer_bases, nullptr, world, rank, slot, epoch, B, W, H, d_out_img,
                                  d_state, stream);
}

extern "C" int lynx_screen_allreduce_mc(void* const* peer_bases, void* multicast_base, int32_t world,
                                        int32_t rank, int32_t slot, uint32_t epoch, int64_t B,
                                        int32_t W, int32_t H, float* d_out_img, int32_t* d_state,
                                        lynx_stream_t stream) {
  if (!peer_bases || world < 1 || world > kMaxWorld || rank < 0 || rank >= world || slot < 0 ||
      slot > 1 || epoch == 0 || B <= 0 || W <= 0 || H <= 0 || W > 65535 || H > 65535 || (W & 3) || !d_out_img ||
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
  a.trace = g_exchange_trace;
  a.mc = reinterpret_cast<char*>(multicast_base);
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  // the rectangle is only known on the device: a fixed grid strides over it
  cudaLaunchConfig_t cfg{};
  unsigned ex_grid = (unsigned)(2 * sms);
  bool ex_pdl = true;
#ifdef LYNX_EXPERIMENTS
  if (const char* e = getenv("LYNX_EX_GRID")) ex_grid = (unsigned)atoi(e);
  if (getenv("LYNX_EX_NOPDL")) ex_pdl = false;
#endif
  cfg.gridDim = dim3(ex_grid);
  cfg.blockDim = dim3(kExTPB);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = ex_pdl ? 1 : 0;
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

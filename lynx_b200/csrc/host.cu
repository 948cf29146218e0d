// Host-buffer entry points: the call a NumPy / host-array binding of
// Segment.track makes.  Copies inputs host->device on the context's stream, runs the
// device launchers, copies results back, synchronises.  Device scratch is cached on the
// context and only grows.
#include <algorithm>
#include <cstring>
#include <new>
#include <vector>

#include "common.cuh"

constexpr int kMaxChunks = 8;

struct lynx_host_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  // the beam crosses PCIe in chunks on its own stream; chunk c is tracked (on `stream`) while
  // chunk c + 1 is still arriving
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t chunk_ready[kMaxChunks] = {};
  struct Buf {
    void* p = nullptr;
    size_t cap = 0;
  };
  Buf params, in, out, energy_in, energy_out, bpm, img, ex, ey, mom, ws, mu_in, cov_in, mu_out,
      cov_out, bpm_f, bbox;
  // LYNX_HOST_OPT_PERSISTENT_IMAGE: the caller passes the SAME zero-initialised host image on
  // every call; only the bounding rectangle of the lit pixels crosses PCIe (the previous
  // call's rectangle is cleared on the host first)
  int persistent_image = 0;
  const float* last_img = nullptr;
  int last_rect[4] = {0, 0, 0, 0};  // r0, r1, c0, c1: the only part of last_img that may be non-zero
  int32_t* h_bbox = nullptr;        // pinned
  // host copies of the last edge tables sent to the device: an unchanged screen geometry is
  // not sent again (two DMA operations less per call)
  std::vector<float> ex_sent, ey_sent;
  // the device image is persistent (lynx_track_particles_fwd_persistent): zeroed when it is
  // (re)allocated, afterwards only the rectangle the previous call lit is cleared
  Buf img_state;
  uint32_t img_epoch = 0;
  int64_t img_geom[3] = {0, 0, 0};  // W, H, B the image had when it was last zeroed as a whole
  bool img_persistent_ok = true;  // false once a lattice needed the non-pipelined kernel
  bool grow(Buf& b, size_t bytes) {
    if (bytes <= b.cap) return true;
    if (b.p) cudaFree(b.p);
    b.p = nullptr;
    b.cap = 0;
    size_t cap = (bytes + 255) & ~(size_t)255;
    if (cudaMalloc(&b.p, cap) != cudaSuccess) return false;
    b.cap = cap;
    return true;
  }
  void release(Buf& b) {
    if (b.p) cudaFree(b.p);
    b.p = nullptr;
    b.cap = 0;
  }
};

extern "C" void lynx_host_ctx_destroy(lynx_host_ctx_t* c);

extern "C" int lynx_host_ctx_create(int device, lynx_host_ctx_t** out) {
  if (!out) return LYNX_E_BADARG;
  *out = nullptr;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) return LYNX_E_NODEVICE;
  if (device < 0 || device >= n) return LYNX_E_BADARG;
  if (cudaSetDevice(device) != cudaSuccess) return LYNX_E_LAUNCH;
  lynx_host_ctx* c = new (std::nothrow) lynx_host_ctx();
  if (!c) return LYNX_E_NOMEM;
  c->device = device;
  bool ok = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) == cudaSuccess &&
            cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking) == cudaSuccess;
  for (int i = 0; ok && i < kMaxChunks; ++i)
    ok = cudaEventCreateWithFlags(&c->chunk_ready[i], cudaEventDisableTiming) == cudaSuccess;
  if (!ok) {
    lynx_host_ctx_destroy(c);
    return LYNX_E_LAUNCH;
  }
  *out = c;
  return LYNX_OK;
}

extern "C" void lynx_host_ctx_destroy(lynx_host_ctx_t* c) {
  if (!c) return;
  lynx_host_ctx::Buf* all[] = {&c->params, &c->in,  &c->out,   &c->energy_in, &c->energy_out,
                               &c->bpm,    &c->img, &c->ex,    &c->ey,        &c->mom,
                               &c->ws,     &c->mu_in, &c->cov_in, &c->mu_out, &c->cov_out,
                               &c->bpm_f,  &c->bbox,  &c->img_state};
  if (c->h_bbox) cudaFreeHost(c->h_bbox);
  for (auto* b : all) c->release(*b);
  for (auto& e : c->chunk_ready)
    if (e) cudaEventDestroy(e);
  if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
}

extern "C" int lynx_host_ctx_set_option(lynx_host_ctx_t* c, int32_t option, int32_t value) {
  if (!c) return LYNX_E_BADARG;
  if (option == LYNX_HOST_OPT_PERSISTENT_IMAGE) {
    c->persistent_image = value != 0;
    c->last_img = nullptr;
    return LYNX_OK;
  }
  return LYNX_E_BADARG;
}

#define H2D(buf, src, bytes)                                                          \
  do {                                                                                \
    if (!c->grow(buf, bytes)) return LYNX_E_NOMEM;                                    \
    LYNX_CUDA_OK(cudaMemcpyAsync(buf.p, src, bytes, cudaMemcpyHostToDevice, st));      \
  } while (0)
#define D2H(dst, buf, bytes) \
  LYNX_CUDA_OK(cudaMemcpyAsync(dst, buf.p, bytes, cudaMemcpyDeviceToHost, st))

extern "C" int lynx_host_track_particles(
    lynx_host_ctx_t* c, const lynx_program_t* p, const float* params, int64_t n_param_floats,
    int64_t B, const float* in_soa, int64_t Bb, int64_t N, const float* energy_in,
    float* out_soa, double* bpm_sums, float* screen_img, int32_t W, int32_t H,
    const float* edges_x, const float* edges_y, double* moment_sums, int32_t moment_mode,
    float* energy_out) {
  if (!c || !p || !params || !in_soa || !energy_in || B <= 0 || N <= 0 || n_param_floats <= 0 ||
      (Bb != 1 && Bb != B))
    return LYNX_E_BADARG;
  if (screen_img && (!edges_x || !edges_y)) return LYNX_E_BADARG;
  LYNX_CUDA_OK(cudaSetDevice(c->device));
  cudaStream_t st = c->stream;
  const int64_t ld = (N + 3) & ~(int64_t)3;
  const int n_bpm = lynx_program_num_bpm(p);
  const bool has_screen = screen_img && lynx_program_num_screens(p) > 0;

  // The beam (24 B per particle, the bulk of the call) goes first and in chunks on the copy
  // stream: the particle kernel of chunk k runs while chunk k + 1 is on the bus, so that only
  // the last chunk's kernel and the read-back are not hidden behind PCIe.  Chunks are whole
  // multiples of 4 particles (16-byte aligned planes); the collapsed moment path (one beam,
  // many settings, adjoint.cu) streams the beam once and is not chunked.
  if (!c->grow(c->in, sizeof(float) * Bb * 6 * ld)) return LYNX_E_NOMEM;
  int n_chunks = N >= (1 << 19) ? 4 : (N >= (1 << 17) ? 2 : 1);
  if (lynx::fwd_collapses(p, B, (Bb == B) ? 1 : 0, out_soa != nullptr, moment_mode)) n_chunks = 1;
  const int64_t chunk = (((N + n_chunks - 1) / n_chunks) + 3) & ~(int64_t)3;
  for (int k = 0; k < n_chunks; ++k) {
    const int64_t p0 = k * chunk, cnt = std::min<int64_t>(chunk, N - p0);
    LYNX_CUDA_OK(cudaMemcpy2DAsync((float*)c->in.p + p0, sizeof(float) * ld, in_soa + p0,
                                   sizeof(float) * N, sizeof(float) * cnt, Bb * 6,
                                   cudaMemcpyHostToDevice, c->copy_stream));
    LYNX_CUDA_OK(cudaEventRecord(c->chunk_ready[k], c->copy_stream));
  }
  H2D(c->params, params, sizeof(float) * n_param_floats);
  H2D(c->energy_in, energy_in, sizeof(float) * B);
  if (out_soa && !c->grow(c->out, sizeof(float) * B * 6 * ld)) return LYNX_E_NOMEM;
  if (!c->grow(c->energy_out, sizeof(float) * B)) return LYNX_E_NOMEM;
  if (n_bpm > 0 && !c->grow(c->bpm, sizeof(double) * 2 * n_bpm * B)) return LYNX_E_NOMEM;
  if (has_screen) {
    if (!c->grow(c->img, sizeof(float) * (size_t)W * H * B)) return LYNX_E_NOMEM;
    if (c->ex_sent.size() != (size_t)W + 1 ||
        memcmp(c->ex_sent.data(), edges_x, sizeof(float) * (W + 1)) != 0) {
      H2D(c->ex, edges_x, sizeof(float) * (W + 1));
      c->ex_sent.assign(edges_x, edges_x + W + 1);
    }
    if (c->ey_sent.size() != (size_t)H + 1 ||
        memcmp(c->ey_sent.data(), edges_y, sizeof(float) * (H + 1)) != 0) {
      H2D(c->ey, edges_y, sizeof(float) * (H + 1));
      c->ey_sent.assign(edges_y, edges_y + H + 1);
    }
  }
  if (moment_mode && !c->grow(c->mom, sizeof(double) * 28 * B)) return LYNX_E_NOMEM;
  if (!c->grow(c->ws, lynx_workspace_bytes(p, B))) return LYNX_E_NOMEM;

  const bool sparse_img = has_screen && c->persistent_image;
  if (sparse_img) {
    if (!c->grow(c->bbox, 64)) return LYNX_E_NOMEM;
    if (!c->h_bbox && cudaMallocHost((void**)&c->h_bbox, 64) != cudaSuccess) return LYNX_E_NOMEM;
  }
  const size_t img_bytes = has_screen ? sizeof(float) * (size_t)W * H * B : 0;
  bool dev_persistent = has_screen && c->img_persistent_ok && !(W & 3);
  const int32_t* d_box = sparse_img ? (const int32_t*)c->bbox.p : nullptr;
  int rc = LYNX_OK;
  uint32_t epoch = 0;
  if (dev_persistent) {
    if (!c->grow(c->img_state, 32)) return LYNX_E_NOMEM;
    if (c->img_geom[0] != W || c->img_geom[1] != H || c->img_geom[2] != B) {
      // new geometry / batch size (or a reallocated buffer): start from zeros
      LYNX_CUDA_OK(cudaMemsetAsync(c->img.p, 0, img_bytes, st));
      LYNX_CUDA_OK(cudaMemsetAsync(c->img_state.p, 0, 32, st));
      c->img_geom[0] = W;
      c->img_geom[1] = H;
      c->img_geom[2] = B;
      c->img_epoch = 0;
    }
    epoch = c->img_epoch++;
    d_box = (const int32_t*)c->img_state.p + 4 * (epoch & 1u);
  }
  // particle range [p0, p0 + cnt) of every beam row; ranges after the first only add
  auto track_range = [&](int64_t p0, int64_t cnt, bool accumulate) {
    return lynx::track_fwd_impl(
        p, (const float*)c->params.p, B, (const float*)c->in.p + p0, (Bb == B) ? 1 : 0, ld, cnt,
        out_soa ? (float*)c->out.p + p0 : nullptr, ld, (double*)c->bpm.p,
        has_screen ? (float*)c->img.p : nullptr, W, H, (const float*)c->ex.p, (const float*)c->ey.p,
        moment_mode ? (double*)c->mom.p : nullptr, moment_mode, (const float*)c->energy_in.p,
        (float*)c->energy_out.p, c->ws.p,
        dev_persistent ? nullptr : (sparse_img ? (int32_t*)c->bbox.p : nullptr),
        dev_persistent ? (int32_t*)c->img_state.p : nullptr, epoch, (lynx_stream_t)st, accumulate);
  };
  for (int k = 0; k < n_chunks; ++k) {
    const int64_t p0 = k * chunk, cnt = std::min<int64_t>(chunk, N - p0);
    LYNX_CUDA_OK(cudaStreamWaitEvent(st, c->chunk_ready[k], 0));
    rc = track_range(p0, cnt, k > 0);
    if (k == 0 && dev_persistent && rc == LYNX_E_BADPROGRAM) {
      // lattice outside the pipelined kernel: plain path from now on
      c->img_persistent_ok = false;
      c->img_geom[0] = c->img_geom[1] = c->img_geom[2] = 0;
      dev_persistent = false;
      d_box = sparse_img ? (const int32_t*)c->bbox.p : nullptr;
      rc = track_range(p0, cnt, false);
    }
    if (rc != LYNX_OK) break;
  }
  if (rc != LYNX_OK) return rc;
  // Sparse read-back of the persistent host image.  Invariant: the host image is zero
  // outside last_rect.  The device image is exact everywhere (zero outside this call's
  // bounding box), so copying ANY rectangle S that contains last_rect both clears what the
  // previous call lit and brings in what this call lit inside S.  S = last_rect grown by a
  // margin is requested speculatively right behind the kernels -- no wait for the bounding
  // box, no host-side clearing; only if this call's box is not inside S does a second copy
  // (and a second synchronisation) follow.
  constexpr int kRectMargin = 16;
  int spec[4] = {0, 0, 0, 0};
  auto copy_rect = [&](const int* q) -> cudaError_t {
    const size_t off = (size_t)q[0] * W + q[2];
    for (int64_t b = 0; b < B; ++b) {
      cudaError_t e = cudaMemcpy2DAsync(
          screen_img + (size_t)b * W * H + off, sizeof(float) * W,
          (const float*)c->img.p + (size_t)b * W * H + off, sizeof(float) * W,
          sizeof(float) * (q[3] - q[2]), q[1] - q[0], cudaMemcpyDeviceToHost, st);
      if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
  };
  if (sparse_img) {
    LYNX_CUDA_OK(cudaMemcpyAsync(c->h_bbox, d_box, 16, cudaMemcpyDeviceToHost, st));
    if (c->last_img != screen_img) {
      memset(screen_img, 0, sizeof(float) * (size_t)W * H * B);  // first use of this buffer
      c->last_rect[0] = c->last_rect[1] = c->last_rect[2] = c->last_rect[3] = 0;
      c->last_img = screen_img;
    }
    const int* q = c->last_rect;
    if (q[1] > q[0] && q[3] > q[2]) {
      spec[0] = q[0] > kRectMargin ? q[0] - kRectMargin : 0;
      spec[1] = q[1] + kRectMargin < H ? q[1] + kRectMargin : H;
      spec[2] = q[2] > kRectMargin ? q[2] - kRectMargin : 0;
      spec[3] = q[3] + kRectMargin < W ? q[3] + kRectMargin : W;
      LYNX_CUDA_OK(copy_rect(spec));
    }
  }

  if (out_soa) {
    if (ld == N) {
      D2H(out_soa, c->out, sizeof(float) * B * 6 * N);
    } else {
      LYNX_CUDA_OK(cudaMemcpy2DAsync(out_soa, sizeof(float) * N, c->out.p, sizeof(float) * ld,
                                     sizeof(float) * N, B * 6, cudaMemcpyDeviceToHost, st));
    }
  }
  if (n_bpm > 0 && bpm_sums) D2H(bpm_sums, c->bpm, sizeof(double) * 2 * n_bpm * B);
  if (has_screen && !sparse_img) D2H(screen_img, c->img, sizeof(float) * (size_t)W * H * B);
  if (moment_mode && moment_sums) D2H(moment_sums, c->mom, sizeof(double) * 28 * B);
  if (energy_out) D2H(energy_out, c->energy_out, sizeof(float) * B);
  LYNX_CUDA_OK(cudaStreamSynchronize(st));
  if (sparse_img) {
    // this call's lit rectangle: rows [H - v3, v2), columns [W - v0, v1)
    const int32_t* v = c->h_bbox;
    int rect[4] = {0, 0, 0, 0};
    if (v[1] > 0 && v[3] > 0) {
      rect[0] = H - v[3];
      rect[1] = v[2];
      rect[2] = W - v[0];
      rect[3] = v[1];
      const bool inside = rect[0] >= spec[0] && rect[1] <= spec[1] && rect[2] >= spec[2] &&
                          rect[3] <= spec[3];
      if (!inside) {
        LYNX_CUDA_OK(copy_rect(rect));
        LYNX_CUDA_OK(cudaStreamSynchronize(st));
      }
    }
    for (int i = 0; i < 4; ++i) c->last_rect[i] = rect[i];
  }
  return LYNX_OK;
}

extern "C" int lynx_host_parambeam(lynx_host_ctx_t* c, const lynx_program_t* p,
                                   const float* params, int64_t n_param_floats, int64_t B,
                                   const float* mu_in, const float* cov_in, int64_t Bm,
                                   const float* energy_in, float* mu_out, float* cov_out,
                                   float* bpm_out, float* energy_out) {
  if (!c || !p || !params || !mu_in || !cov_in || !energy_in || !mu_out || !cov_out || B <= 0 ||
      n_param_floats <= 0 || (Bm != 1 && Bm != B))
    return LYNX_E_BADARG;
  LYNX_CUDA_OK(cudaSetDevice(c->device));
  cudaStream_t st = c->stream;
  const int n_bpm = lynx_program_num_bpm(p);
  H2D(c->params, params, sizeof(float) * n_param_floats);
  H2D(c->energy_in, energy_in, sizeof(float) * B);
  H2D(c->mu_in, mu_in, sizeof(float) * 7 * Bm);
  H2D(c->cov_in, cov_in, sizeof(float) * 49 * Bm);
  if (!c->grow(c->mu_out, sizeof(float) * 7 * B)) return LYNX_E_NOMEM;
  if (!c->grow(c->cov_out, sizeof(float) * 49 * B)) return LYNX_E_NOMEM;
  if (!c->grow(c->energy_out, sizeof(float) * B)) return LYNX_E_NOMEM;
  if (n_bpm > 0 && !c->grow(c->bpm_f, sizeof(float) * 2 * n_bpm * B)) return LYNX_E_NOMEM;
  int rc = lynx_parambeam_fwd(p, (const float*)c->params.p, B, (const float*)c->mu_in.p,
                              (const float*)c->cov_in.p, (Bm == B) ? 1 : 0,
                              (const float*)c->energy_in.p, (float*)c->mu_out.p,
                              (float*)c->cov_out.p, (float*)c->bpm_f.p,
                              (float*)c->energy_out.p, (lynx_stream_t)st);
  if (rc != LYNX_OK) return rc;
  D2H(mu_out, c->mu_out, sizeof(float) * 7 * B);
  D2H(cov_out, c->cov_out, sizeof(float) * 49 * B);
  if (n_bpm > 0 && bpm_out) D2H(bpm_out, c->bpm_f, sizeof(float) * 2 * n_bpm * B);
  if (energy_out) D2H(energy_out, c->energy_out, sizeof(float) * B);
  LYNX_CUDA_OK(cudaStreamSynchronize(st));
  return LYNX_OK;
}

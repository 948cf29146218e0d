// lynx_track_particles_fwd_noalloc: the fused particle tracker for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
// No counterpart in the reference.
//
// It is lynx_track_particles_fwd (track.cu) with the caller-supplied workspace (the per-setting
// K1 -> K2 records + the histogram geometry) replaced by a file-scope device array of
// LYNX_NOALLOC_WORKSPACE_BYTES bytes that every GPU loading the library holds until exit.
// Nothing here allocates: no cudaMalloc / cudaFree / allocating library routine.
//
// The array is SHARED by every call on a device and keeps the last call's contents.  Nothing in
// it has to start at zero -- K1 writes every record word and the geometry block before K2 reads
// them -- so there is nothing to clear; but two calls that overlap on the GPU (different
// streams, or graphs replayed concurrently) race on the records.  Calls on one stream are fine.
#include <mutex>

#include "common.cuh"

namespace {

static __device__ __align__(256) unsigned char g_noalloc_ws[LYNX_NOALLOC_WORKSPACE_BYTES];

// In host code `g_noalloc_ws` is a host-side stand-in, not a device address: the address the
// kernels get comes from cudaGetSymbolAddress, looked up once per device with that device current.
constexpr int kMaxDevices = 64;
std::mutex g_mu;
void* g_ws_addr[kMaxDevices] = {nullptr};

int noalloc_workspace(void** out) {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess) return LYNX_E_NODEVICE;
  if (dev < 0 || dev >= kMaxDevices) return LYNX_E_BADARG;
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_ws_addr[dev]) {
    void* p = nullptr;
    if (cudaGetSymbolAddress(&p, g_noalloc_ws) != cudaSuccess || !p) {
      cudaGetLastError();
      return LYNX_E_LAUNCH;
    }
    g_ws_addr[dev] = p;
  }
  *out = g_ws_addr[dev];
  return LYNX_OK;
}

}  // namespace

extern "C" int lynx_track_particles_fwd_noalloc(
    const lynx_program_t* p, const float* d_params, int64_t B, const float* d_in_soa,
    int64_t in_row_stride, int64_t ld_in, int64_t N, float* d_out_soa, int64_t ld_out,
    double* d_bpm_sums, float* d_screen_img, int32_t W, int32_t H, const float* d_edges_x,
    const float* d_edges_y, double* d_moment_sums, int32_t moment_mode,
    const float* d_energy_in, float* d_energy_out, lynx_stream_t stream) {
  if (!p || B <= 0) return LYNX_E_BADARG;
  const size_t need = lynx_workspace_bytes(p, B);
  if (need == 0) return LYNX_E_BADARG;
  if (need > (size_t)LYNX_NOALLOC_WORKSPACE_BYTES) return LYNX_E_TOOLARGE;
  void* ws = nullptr;
  const int rc = noalloc_workspace(&ws);
  if (rc != LYNX_OK) return rc;
  return lynx_track_particles_fwd(p, d_params, B, d_in_soa, in_row_stride, ld_in, N, d_out_soa,
                                  ld_out, d_bpm_sums, d_screen_img, W, H, d_edges_x, d_edges_y,
                                  d_moment_sums, moment_mode, d_energy_in, d_energy_out, ws, stream);
}

// XLA FFI handlers over the C ABI of include/lynx_b200.h -- the binding BASELINE.json's
// north star names ("Python host code calls hand-written sm_100a CUDA through a thin XLA FFI
// custom-call C-ABI layer with custom_vjp registrations").  The seam it replaces is the body of
// Segment.track (lynx/accelerator/segment.py:340-356) and of Element.transfer_map / track
// (element.py:36-94): under JAX those become jax.ffi.ffi_call's on the targets registered here
// (lynx_b200/jax_binding.py).
//
// Pure argument adapters: no arithmetic, no allocation, no synchronisation; XLA owns every
// buffer (scratch is a result buffer sized by lynx_workspace_bytes) and hands over its stream.
// A zero-element buffer stands for a NULL pointer of the C ABI.
//
// This image has neither jax nor jaxlib nor xla/ffi/api/ffi.h, so the file
// compiles to an empty translation unit here; with jaxlib installed:
//   g++ -O2 -fPIC -shared -I$(python -c "import jax.ffi; print(jax.ffi.include_dir())") \
//       -I include -I /usr/local/cuda/include lynx_b200/csrc/lynx_xla_ffi.cc \
//       -L lynx_b200/lib -llynx_b200 -o lynx_b200/lib/liblynx_b200_xla.so
// tests/test_xla_ffi_table.py checks (on the CPU, by parsing) that the LYNX_FFI_HANDLER table
// below covers every stream launcher of lynx_b200/_lib.py and that jax_binding.py registers
// exactly these symbols.
#include "lynx_b200.h"

// LYNX_FFI_HANDLER(exported handler symbol, C-ABI entry point it adapts)
#define LYNX_FFI_HANDLER(symbol, c_abi) /* table entry, see the test */

#if defined(__has_include)
#if __has_include("xla/ffi/api/ffi.h")
#define LYNX_HAVE_XLA_FFI 1
#endif
#endif

#ifdef LYNX_HAVE_XLA_FFI
#include <cuda_runtime_api.h>

#include "xla/ffi/api/ffi.h"

namespace ffi = xla::ffi;
using F32 = ffi::Buffer<ffi::F32>;
using F64 = ffi::Buffer<ffi::F64>;
using S32 = ffi::Buffer<ffi::S32>;
using RF32 = ffi::ResultBuffer<ffi::F32>;
using RF64 = ffi::ResultBuffer<ffi::F64>;
using RS32 = ffi::ResultBuffer<ffi::S32>;
using RS64 = ffi::ResultBuffer<ffi::S64>;
using RU8 = ffi::ResultBuffer<ffi::U8>;
using Stream = ffi::PlatformStream<cudaStream_t>;

namespace {
template <typename B>
auto* nullable(B& b) {
  return b.element_count() ? b.typed_data() : nullptr;
}
template <typename B>
auto* nullable_ret(B& b) {
  return b->element_count() ? b->typed_data() : nullptr;
}
inline ffi::Error status(int rc) {
  return rc == LYNX_OK ? ffi::Error::Success() : ffi::Error(ffi::ErrorCode::kInternal, lynx_status_str(rc));
}
inline const lynx_program_t* prog(int64_t handle) { return reinterpret_cast<const lynx_program_t*>(handle); }
inline lynx_stream_t st(cudaStream_t s) { return reinterpret_cast<lynx_stream_t>(s); }
}  // namespace

// ---- K1: Segment.transfer_map / Element.transfer_map
static ffi::Error ComposeMaps(cudaStream_t s, int64_t h, F32 params, F32 energy_in, F32 seed_soa, RF32 records,
                              RF32 maps_out, RF32 energy_out, int64_t B, int64_t seed_row_stride, int64_t seed_ld) {
  return status(lynx_compose_maps(prog(h), params.typed_data(), B, energy_in.typed_data(), nullable(seed_soa),
                                  seed_row_stride, seed_ld, records->typed_data(), nullable_ret(maps_out),
                                  nullable_ret(energy_out), st(s)));
}
LYNX_FFI_HANDLER(LynxComposeMaps, lynx_compose_maps)
XLA_FFI_DEFINE_HANDLER_SYMBOL(LynxComposeMaps, ComposeMaps,
                              ffi::Ffi::Bind().Ctx<Stream>().Attr<int64_t>("prog").Arg<F32>().Arg<F32>().Arg<F32>()
                                  .Ret<F32>().Ret<F32>().Ret<F32>().Attr<int64_t>("B")
                                  .Attr<int64_t>("seed_row_stride").Attr<int64_t>("seed_ld"));

// ---- K2: Segment.track(ParticleBeam) + Screen.reading + moments + BPM sums
static ffi::Error TrackFwd(cudaStream_t s, int64_t h, F32 params, F32 in_soa, F32 energy_in, F32 edges_x,
                           F32 edges_y, RF32 out_soa, RF64 bpm_sums, RF32 img, RF64 moment_sums, RF32 energy_out,
                           RU8 workspace, int64_t B, int64_t N, int64_t ld, int64_t in_row_stride,
                           int32_t moment_mode) {
  const int32_t W = edges_x.element_count() ? (int32_t)edges_x.element_count() - 1 : 0;
  const int32_t H = edges_y.element_count() ? (int32_t)edges_y.element_count() - 1 : 0;
  return status(lynx_track_particles_fwd(
      prog(h), params.typed_data(), B, in_soa.typed_data(), in_row_stride, ld, N, nullable_ret(out_soa), ld,
      nullable_ret(bpm_sums), nullable_ret(img), W, H, nullable(edges_x), nullable(edges_y),
      nullable_ret(moment_sums), moment_mode, energy_in.typed_data(), energy_out->typed_data(),
      workspace->untyped_data(), st(s)));
}
LYNX_FFI_HANDLER(LynxTrackFwd, lynx_track_particles_fwd)
XLA_FFI_DEFINE_HANDLER_SYMBOL(LynxTrackFwd, TrackFwd,
                              ffi::Ffi::Bind().Ctx<Stream>().Attr<int64_t>("prog").Arg<F32>().Arg<F32>().Arg<F32>()
                                  .Arg<F32>().Arg<F32>().Ret<F32>().Ret<F64>().Ret<F32>().Ret<F64>().Ret<F32>()
                                  .Ret<ffi::Buffer<ffi::U8>>().Attr<int64_t>("B").Attr<int64_t>("N")
                                  .Attr<int64_t>("ld").Attr<int64_t>("in_row_stride")
                                  .Attr<int32_t>("moment_mode"));

// the same without a workspace result: the library's static scratch (graph capture, no allocation)
static ffi::Error TrackFwdNoAlloc(cudaStream_t s, int64_t h, F32 params, F32 in_soa, F32 energy_in, F32 edges_x,
                                  F32 edges_y, RF32 out_soa, RF64 bpm_sums, RF32 img, RF64 moment_sums,
                                  RF32 energy_out, int64_t B, int64_t N, int64_t ld, int64_t in_row_stride,
                                  int32_t moment_mode) {
  const int32_t W = edges_x.element_count() ? (int32_t)edges_x.element_count() - 1 : 0;
  const int32_t H = edges_y.element_count() ? (int32_t)edges_y.element_count() - 1 : 0;
  return status(lynx_track_particles_fwd_noalloc(
      prog(h), params.typed_data(), B, in_soa.typed_data(), in_row_stride, ld, N, nullable_ret(out_soa), ld,
      nullable_ret(bpm_sums), nullable_ret(img), W, H, nullable(edges_x), nullable(edges_y),
      nullable_ret(moment_sums), moment_mode, energy_in.typed_data(), energy_out->typed_data(), st(s)));
}
LYNX_FFI_HANDLER(LynxTrackFwdNoAlloc, lynx_track_particles_fwd_noalloc)
XLA_FFI_DEFINE_HANDLER_SYMBOL(LynxTrackFwdNoAlloc, TrackFwdNoAlloc,
                              ffi::Ffi::Bind().Ctx<Stream>().Attr<int64_t>("prog").Arg<F32>().Arg<F32>().Arg<F32>()
                                  .Arg<F32>().Arg<F32>().Ret<F32>().Ret<F64>().Ret<F32>().Ret<F64>().Ret<F32>()
                                  .Attr<int64_t>("B").Attr<int64_t>("N").Attr<int64_t>("ld")
                                  .Attr<int64_t>("in_row_stride").Attr<int32_t>("moment_mode"));

// the same with the lit rectangle as a further result (multi-process: the exchange step)
static ffi::Error TrackFwdBbox(cudaStream_t s, int64_t h, F32 params, F32 in_soa, F32 energy_in, F32 edges_x,
                               F32 edges_y, RF32 out_soa, RF64 bpm_sums, RF32 img, RF64 moment_sums,
                               RF32 energy_out, RU8 workspace, RS32 bbox, int64_t B, int64_t N, int64_t ld,
                               int64_t in_row_stride, int32_t moment_mode) {
  return status(lynx_track_particles_fwd_bbox(
      prog(h), params.typed_data(), B, in_soa.typed_data(), in_row_stride, ld, N, nullable_ret(out_soa), ld,
      nullable_ret(bpm_sums), nullable_ret(img), (int32_t)edges_x.element_count() - 1,
      (int32_t)edges_y.element_count() - 1, edges_x.typed_data(), edges_y.typed_data(), nullable_ret(moment_sums),
      moment_mode, energy_in.typed_data(), energy_out->typed_data(), workspace->untyped_data(), bbox->typed_data(),
      st(s)));
}
LYNX_FFI_HANDLER(LynxTrackFwdBbox, lynx_track_particles_fwd_bbox)
XLA_FFI_DEFINE_HANDLER_SYMBOL(LynxTrackFwdBbox, TrackFwdBbox,
                              ffi::Ffi::Bind().Ctx<Stream>().Attr<int64_t>("prog").Arg<F32>().Arg<F32>().Arg<F32>()
                                  .Arg<F32>().Arg<F32>().Ret<F32>().Ret<F64>().Ret<F32>().Ret<F64>().Ret<F32>()
                                  .Ret<ffi::Buffer<ffi::U8>>().Ret<S32>().Attr<int64_t>("B").Attr<int64_t>("N")
                                  .Attr<int64_t>("ld").Attr<int64_t>("in_row_stride")
                                  .Attr<int32_t>("moment_mode"));

// persistent image: image and state are donated inputs aliased to results (input_output_aliases)
static ffi::Error TrackFwdPersistent(cudaStream_t s, int64_t h, F32 params, F32 in_soa, F32 energy_in, F32 edges_x,
                                     F32 edges_y, F32 /*img_in, aliased*/, S32 /*state_in, aliased*/,
                                     RF64 bpm_sums, RF32 img, RS32 state, RF64 moment_sums, RF32 energy_out,
                                     RU8 workspace, int64_t B, int64_t N, int64_t ld, int64_t in_row_stride,
                                     int32_t moment_mode, int64_t epoch) {
  return status(lynx_track_particles_fwd_persistent(
      prog(h), params.typed_data(), B, in_soa.typed_data(), in_row_stride, ld, N, nullptr, ld,
      nullable_ret(bpm_sums), img->typed_data(), (int32_t)edges_x.element_count() - 1,
      (int32_t)edges_y.element_count() - 1, edges_x.typed_data(), edges_y.typed_data(), nullable_ret(moment_sums),
      moment_mode, energy_in.typed_data(), energy_out->typed_data(), workspace->untyped_data(), state->typed_data(),
      (uint32_t)epoch, st(s)));
}
LYNX_FFI_HANDLER(LynxTrackFwdPersistent, lynx_track_particles_fwd_persistent)
XLA_FFI_DEFINE_HANDLER_SYMBOL(LynxTrackFwdPersistent, TrackFwdPersistent,
                              ffi::Ffi::Bind().Ctx<Stream>().Attr<int64_t>("prog").Arg<F32>().Arg<F32>().Arg<F32>()
                                  .Arg<F32>().Arg<F32>().Arg<F32>().Arg<S32>().Ret<F64>().Ret<F32>().Ret<S32>()
                                  .Ret<F64>().Ret<F32>().Ret<ffi::Buffer<ffi::U8>>().Attr<int64_t>("B")
                                  .Attr<int64_t>("N").Attr<int64_t>("ld").Attr<int64_t>("in_row_stride")
                                  .Attr<int32_t>("moment_mode").Attr<int64_t>("epoch"));

// ---- K3: the vjp of the call above (custom_vjp backward)
static ffi::Error TrackBwd(cudaStream_t s, int64_t h, F32 params, F32 in_soa, F32 energy_in, F32 g_out_soa,
                           F64 g_moment_sums, F64 g_bpm_sums, RF32 g_params, RF32 g_in_soa, RU8 workspace,
                           int64_t B, int64_t N, int64_t ld, int64_t in_row_stride) {
  return status(lynx_track_particles_bwd(prog(h), params.typed_data(), B, in_soa.typed_data(), in_row_stride, ld, N,
                                         nullable(g_out_soa), ld, nullable(g_moment_sums), nullable(g_bpm_sums),
                                         energy_in.typed_data(), g_params->typed_data(), nullable_ret(g_in_soa), ld,
                                         workspace->untyped_data(), st(s)));
}
LYNX_FFI_HANDLER(LynxTrackBwd, lynx_track_particles_bwd)
XLA_FFI_DEFINE_HANDLER_SYMBOL(LynxTrackBwd, TrackBwd,
                              ffi::Ffi::Bind().Ctx<Stream>().Attr<int64_t>("prog").Arg<F32>().Arg<F32>().Arg<F32>()
                                  .Arg<F32>().Arg<F64>().Arg<F64>().Ret<F32>().Ret<F32>()
                                  .Ret<ffi::Buffer<ffi::U8>>().Attr<int64_t>("B").Attr<int64_t>("N")
                                  .Attr<int64_t>("ld").Attr<int64_t>("in_row_stride"));

// ---- K6: Segment.track(ParameterBeam) and its vjp
static ffi::Error ParamBeamFwd(cudaStream_t s, int64_t h, F32 params, F32 mu_in, F32 cov_in, F32 energy_in,
                               RF32 mu_out, RF32 cov_out, RF32 bpm_out, RF32 energy_out, int64_t B,
                               int64_t beam_stride) {
  return status(lynx_parambeam_fwd(prog(h), params.typed_data(), B, mu_in.typed_data(), cov_in.typed_data(),
                                   beam_stride, energy_in.typed_data(), mu_out->typed_data(), cov_out->typed_data(),
                                   nullable_ret(bpm_out), energy_out->typed_data(), st(s)));
}
LYNX_FFI_HANDLER(LynxParamBeamFwd, lynx_parambeam_fwd)
XLA_FFI_DEFINE_HANDLER_SYMBOL(LynxParamBeamFwd, ParamBeamFwd,
                              ffi::Ffi::Bind().Ctx<Stream>().Attr<int64_t>("prog").Arg<F32>().Arg<F32>().Arg<F32>()
                                  .Arg<F32>().Ret<F32>().Ret<F32>().Ret<F32>().Ret<F32>().Attr<int64_t>("B")
                                  .Attr<int64_t>("beam_stride"));

static ffi::Error ParamBeamBwd(cudaStream_t s, int64_t h, F32 params, F32 mu_in, F32 cov_in, F32 energy_in, F32 g_mu,
                               F32 g_cov, F32 g_bpm, RF32 g_params, RF32 g_mu_in, RF32 g_cov_in, int64_t B,
                               int64_t beam_stride) {
  return status(lynx_parambeam_bwd(prog(h), params.typed_data(), B, mu_in.typed_data(), cov_in.typed_data(),
                                   beam_stride, energy_in.typed_data(), nullable(g_mu), nullable(g_cov),
                                   nullable(g_bpm), g_params->typed_data(), nullable_ret(g_mu_in),
                                   nullable_ret(g_cov_in), st(s)));
}
LYNX_FFI_HANDLER(LynxParamBeamBwd, lynx_parambeam_bwd)
XLA_FFI_DEFINE_HANDLER_SYMBOL(LynxParamBeamBwd, ParamBeamBwd,
                              ffi::Ffi::Bind().Ctx<Stream>().Attr<int64_t>("prog").Arg<F32>().Arg<F32>().Arg<F32>()
                                  .Arg<F32>().Arg<F32>().Arg<F32>().Arg<F32>().Ret<F32>().Ret<F32>().Ret<F32>()
                                  .Attr<int64_t>("B").Attr<int64_t>("beam_stride"));

static ffi::Error ParamBeamAlong(cudaStream_t s, int64_t h, F32 params, F32 mu_in, F32 cov_in, F32 energy_in,
                                 RF32 snap_mu, RF32 snap_cov, RF32 snap_energy, int64_t B, int64_t beam_stride) {
  return status(lynx_parambeam_along(prog(h), params.typed_data(), B, mu_in.typed_data(), cov_in.typed_data(),
                                     beam_stride, energy_in.typed_data(), snap_mu->typed_data(),
                                     snap_cov->typed_data(), snap_energy->typed_data(), st(s)));
}
LYNX_FFI_HANDLER(LynxParamBeamAlong, lynx_parambeam_along)
XLA_FFI_DEFINE_HANDLER_SYMBOL(LynxParamBeamAlong, ParamBeamAlong,
                              ffi::Ffi::Bind().Ctx<Stream>().Attr<int64_t>("prog").Arg<F32>().Arg<F32>().Arg<F32>()
                                  .Arg<F32>().Ret<F32>().Ret<F32>().Ret<F32>().Attr<int64_t>("B")
                                  .Attr<int64_t>("beam_stride"));

// ---- K7: Screen.reading for a ParameterBeam
static ffi::Error ScreenPdf(cudaStream_t s, F32 mu, F32 cov, F32 xs, F32 ys, RF32 img, int64_t B) {
  return status(lynx_screen_pdf(mu.typed_data(), cov.typed_data(), B, xs.typed_data(), ys.typed_data(),
                                (int32_t)xs.element_count(), (int32_t)ys.element_count(), img->typed_data(), st(s)));
}
LYNX_FFI_HANDLER(LynxScreenPdf, lynx_screen_pdf)
XLA_FFI_DEFINE_HANDLER_SYMBOL(LynxScreenPdf, ScreenPdf,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<F32>().Arg<F32>().Arg<F32>().Arg<F32>().Ret<F32>()
                                  .Attr<int64_t>("B"));

// ---- K9 / K9b / K10 / K11: beam constructors and beam-level operations
static ffi::Error BeamGenerate(cudaStream_t s, F32 mean, F32 sqrt_cov, RF32 soa, int64_t B, int64_t N,
                               int64_t first_particle, int64_t seed, int64_t ld) {
  return status(lynx_beam_generate(mean.typed_data(), sqrt_cov.typed_data(), B, N, first_particle, (uint64_t)seed,
                                   soa->typed_data(), ld, st(s)));
}
LYNX_FFI_HANDLER(LynxBeamGenerate, lynx_beam_generate)
XLA_FFI_DEFINE_HANDLER_SYMBOL(LynxBeamGenerate, BeamGenerate,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<F32>().Arg<F32>().Ret<F32>().Attr<int64_t>("B")
                                  .Attr<int64_t>("N").Attr<int64_t>("first_particle").Attr<int64_t>("seed")
                                  .Attr<int64_t>("ld"));

static ffi::Error BeamGenerateEllipsoid(cudaStream_t s, F32 radii, F32 sigma, RF32 soa, int64_t B, int64_t N,
                                        int64_t first_particle, int64_t seed, int64_t ld) {
  return status(lynx_beam_generate_ellipsoid(radii.typed_data(), sigma.typed_data(), B, N, first_particle,
                                             (uint64_t)seed, soa->typed_data(), ld, st(s)));
}
LYNX_FFI_HANDLER(LynxBeamGenerateEllipsoid, lynx_beam_generate_ellipsoid)
XLA_FFI_DEFINE_HANDLER_SYMBOL(LynxBeamGenerateEllipsoid, BeamGenerateEllipsoid,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<F32>().Arg<F32>().Ret<F32>().Attr<int64_t>("B")
                                  .Attr<int64_t>("N").Attr<int64_t>("first_particle").Attr<int64_t>("seed")
                                  .Attr<int64_t>("ld"));

static ffi::Error BeamTransform(cudaStream_t s, F32 in_soa, F32 table, RF32 out_soa, int64_t B, int64_t N,
                                int64_t ld, int64_t in_row_stride) {
  return status(lynx_beam_transform(in_soa.typed_data(), in_row_stride, ld, B, N, table.typed_data(),
                                    out_soa->typed_data(), ld, st(s)));
}
LYNX_FFI_HANDLER(LynxBeamTransform, lynx_beam_transform)
XLA_FFI_DEFINE_HANDLER_SYMBOL(LynxBeamTransform, BeamTransform,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<F32>().Arg<F32>().Ret<F32>().Attr<int64_t>("B")
                                  .Attr<int64_t>("N").Attr<int64_t>("ld").Attr<int64_t>("in_row_stride"));

// static shapes under jit: padded planes + the survivor counts as results
static ffi::Error ApertureCompact(cudaStream_t s, F32 in_soa, F32 charges_in, F32 x_max, F32 y_max, RF32 out_soa,
                                  RF32 lost_soa, RF32 charges_out, RF32 charges_lost, RS64 counts, RU8 workspace,
                                  int64_t R, int64_t N, int64_t ld, int64_t lim_stride, int32_t shape) {
  return status(lynx_aperture_compact(in_soa.typed_data(), R, N, ld, nullable(charges_in), ld, x_max.typed_data(),
                                      y_max.typed_data(), lim_stride, shape, out_soa->typed_data(),
                                      nullable_ret(lost_soa), ld, nullable_ret(charges_out),
                                      nullable_ret(charges_lost), counts->typed_data(), workspace->untyped_data(),
                                      st(s)));
}
LYNX_FFI_HANDLER(LynxApertureCompact, lynx_aperture_compact)
XLA_FFI_DEFINE_HANDLER_SYMBOL(LynxApertureCompact, ApertureCompact,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<F32>().Arg<F32>().Arg<F32>().Arg<F32>().Ret<F32>()
                                  .Ret<F32>().Ret<F32>().Ret<F32>().Ret<ffi::Buffer<ffi::S64>>()
                                  .Ret<ffi::Buffer<ffi::U8>>().Attr<int64_t>("R").Attr<int64_t>("N")
                                  .Attr<int64_t>("ld").Attr<int64_t>("lim_stride").Attr<int32_t>("shape"));

// ---- layout converters: ParticleBeam.particles (..., N, 7) <-> SoA-6 planes
static ffi::Error Aos7ToSoa6(cudaStream_t s, F32 aos, RF32 soa, int64_t R, int64_t N, int64_t ld) {
  return status(lynx_aos7_to_soa6(aos.typed_data(), R, N, soa->typed_data(), ld, st(s)));
}
LYNX_FFI_HANDLER(LynxAos7ToSoa6, lynx_aos7_to_soa6)
XLA_FFI_DEFINE_HANDLER_SYMBOL(LynxAos7ToSoa6, Aos7ToSoa6,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<F32>().Ret<F32>().Attr<int64_t>("R")
                                  .Attr<int64_t>("N").Attr<int64_t>("ld"));

static ffi::Error Soa6ToAos7(cudaStream_t s, F32 soa, RF32 aos, int64_t R, int64_t N, int64_t ld) {
  return status(lynx_soa6_to_aos7(soa.typed_data(), R, N, ld, aos->typed_data(), st(s)));
}
LYNX_FFI_HANDLER(LynxSoa6ToAos7, lynx_soa6_to_aos7)
XLA_FFI_DEFINE_HANDLER_SYMBOL(LynxSoa6ToAos7, Soa6ToAos7,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<F32>().Ret<F32>().Attr<int64_t>("R")
                                  .Attr<int64_t>("N").Attr<int64_t>("ld"));

// ---- multi-process exchange step: `peers` is the address of the host table of peer-mapped
//      symmetric buffers (set up once by the process, sharding.ScreenExchange), `mc` the
//      multicast mapping or 0
static ffi::Error ScreenAllreduce(cudaStream_t s, F32 /*order after the tracker*/, F32 /*out_in, aliased*/,
                                  S32 /*state_in, aliased*/, RF32 out_img, RS32 state, int64_t peers, int64_t mc,
                                  int32_t world, int32_t rank, int32_t slot, int64_t epoch, int64_t B, int32_t W,
                                  int32_t H) {
  void* const* bases = reinterpret_cast<void* const*>(peers);
  if (mc)
    return status(lynx_screen_allreduce_mc(bases, reinterpret_cast<void*>(mc), world, rank, slot, (uint32_t)epoch, B,
                                           W, H, out_img->typed_data(), state->typed_data(), st(s)));
  return status(lynx_screen_allreduce(bases, world, rank, slot, (uint32_t)epoch, B, W, H, out_img->typed_data(),
                                      state->typed_data(), st(s)));
}
LYNX_FFI_HANDLER(LynxScreenAllreduce, lynx_screen_allreduce)
LYNX_FFI_HANDLER(LynxScreenAllreduce, lynx_screen_allreduce_mc)
XLA_FFI_DEFINE_HANDLER_SYMBOL(LynxScreenAllreduce, ScreenAllreduce,
                              ffi::Ffi::Bind().Ctx<Stream>().Arg<F32>().Arg<F32>().Arg<S32>().Ret<F32>().Ret<S32>()
                                  .Attr<int64_t>("peers").Attr<int64_t>("mc").Attr<int32_t>("world")
                                  .Attr<int32_t>("rank").Attr<int32_t>("slot").Attr<int64_t>("epoch")
                                  .Attr<int64_t>("B").Attr<int32_t>("W").Attr<int32_t>("H"));

#endif  // LYNX_HAVE_XLA_FFI

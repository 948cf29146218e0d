/*
 * lynx_b200.h -- C ABI of the B200-native beam-tracking hot path.
 *
 * The reference (jank324/lynx) has NO plugin / FFI seam: its hot path is the Python
 * method surface Segment.track / Element.transfer_map / Screen.reading / BPM.reading
 * (SURVEY.md section 8b).  These entry points are what a binding for that surface
 * (XLA FFI custom call, ctypes, torch extension) calls.  Each one cites the reference
 * code it replaces.  Plain pointers and sizes only; no torch / XLA types.
 *
 * Conventions
 *   - all `d_*` / device pointers are device memory owned by the caller; every launcher
 *     only ENQUEUES work on `stream` and returns; no allocation, no synchronisation,
 *     no global mutable state (a program handle is immutable after creation).
 *   - return value: 0 = LYNX_OK, negative = lynx_status error.  No exceptions.
 *   - coordinates are (x, xp, y, yp, s, p); the constant-1 seventh coordinate of the
 *     reference (lynx/accelerator/element.py:41-54) is never stored.
 *   - particles are SoA: six float planes of N_pad floats per beam row;
 *     plane k of beam row r starts at  soa + (r*6 + k)*ld  (ld >= N, multiple of 4,
 *     base 16-byte aligned).  AoS (..,N,7) <-> SoA converters are provided.
 *   - B = number of settings (product of the reference's leading batch dims).
 */
#ifndef LYNX_B200_H
#define LYNX_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct CUstream_st* lynx_stream_t; /* == cudaStream_t */

typedef enum {
  LYNX_OK = 0,
  LYNX_E_BADARG = -1,     /* null / misaligned pointer, negative size            */
  LYNX_E_BADPROGRAM = -2, /* unknown opcode, column out of range, bad slot        */
  LYNX_E_LAUNCH = -3,     /* CUDA reported an error while enqueueing              */
  LYNX_E_NOMEM = -4,      /* handle / workspace allocation failed                 */
  LYNX_E_NODEVICE = -5,   /* no CUDA device: there is NO CPU fallback             */
  LYNX_E_TOOLARGE = -6    /* lynx_track_particles_fwd_noalloc: the problem needs more scratch
                             than the library's static LYNX_NOALLOC_WORKSPACE_BYTES      */
} lynx_status;

/* Lattice program.  One op per reference element, in beam order.
 * `col` = index of the op's first parameter column; columns are consecutive.     */
typedef enum {
  LYNX_OP_MARKER = 0,    /* marker.py:33-36 identity; no columns                         */
  LYNX_OP_DRIFT = 1,     /* drift.py:44-62              cols: length                     */
  LYNX_OP_QUAD = 2,      /* quadrupole.py:66-80         cols: length,k1,mis_x,mis_y,tilt */
  LYNX_OP_HCOR = 3,      /* horizontal_corrector.py:52-67  cols: length,angle            */
  LYNX_OP_VCOR = 4,      /* vertical_corrector.py:52-66    cols: length,angle            */
  LYNX_OP_DIPOLE = 5,    /* dipole.py:112-181  cols: length,angle,e1,e2,tilt,fint,fint_exit,gap
                            flag LYNX_F_RBEND adds angle/2 to e1,e2 (rbend.py:79-80)     */
  LYNX_OP_CAVITY = 6,    /* cavity.py:97-325   cols: length,voltage,phase_deg,frequency
                            flag LYNX_F_ACTIVE: non-linear energy-changing track
                            (cavity.py:97-246); without it the linear map only
                            (is_skippable case, cavity.py:68-70)                         */
  LYNX_OP_SOLENOID = 7,  /* solenoid.py:61-105          cols: length,k,mis_x,mis_y       */
  LYNX_OP_UNDULATOR = 8, /* undulator.py:48-60          cols: length                     */
  LYNX_OP_CUSTOM = 9,    /* custom_transfer_map.py:87-88  cols: 42 = rows 0..5 of the 7x7,
                            row-major (row 6 must be 0,..,0,1)                           */
  LYNX_OP_BPM = 10,      /* bpm.py:48-58 ACTIVE bpm: aux = reading slot; no columns      */
  LYNX_OP_SCREEN = 11,   /* screen.py:126-141 ACTIVE screen: aux = slot; cols: mis_x,mis_y;
                            terminal: nothing after it is tracked (returns Beam.empty)   */
  LYNX_OP__COUNT = 12
} lynx_opcode;

enum {
  LYNX_F_ACTIVE = 1,       /* CAVITY: voltage != 0 somewhere in the batch          */
  LYNX_F_RBEND = 2,        /* DIPOLE: rectangular bend                             */
  LYNX_F_SCREEN_Y_FIX = 4  /* SCREEN: subtract mis_y from y (physically meant) instead
                              of from xp as the reference writes it (screen.py:135) */
};

typedef struct {
  int32_t opcode; /* lynx_opcode */
  int32_t col;    /* first parameter column, -1 if none */
  int32_t flags;
  int32_t aux;    /* BPM / SCREEN slot */
} lynx_op_t;

/* Parameter table.  Value of column c for setting b:
 *     params[col_offset[c] + b * col_stride[c]]      col_stride in {0 (shared), 1}
 * so "vectorised element parameters" (tests/test_vectorized.py) are a stride-1 column
 * and Element.broadcast (e.g. drift.py:64-65) is a stride-0 column: nothing is tiled. */
typedef struct lynx_program lynx_program_t;

int lynx_version(void);
const char* lynx_status_str(int status);

/* Builds an immutable program handle (copies ops / column descriptors to the device).
 * Replaces the Python-side structure of Segment (segment.py:40-54) + the run grouping
 * of Segment.track (segment.py:344-351), which is decided here once per lattice.     */
int lynx_program_create(const lynx_op_t* ops, int32_t n_ops, const int64_t* col_offset,
                        const int32_t* col_stride, int32_t n_cols, lynx_program_t** out);
void lynx_program_destroy(lynx_program_t* prog);

/* Introspection of the compiled plan. */
int32_t lynx_program_num_stages(const lynx_program_t*);  /* affine stages (merged runs) */
int32_t lynx_program_num_bpm(const lynx_program_t*);
int32_t lynx_program_num_screens(const lynx_program_t*);
int32_t lynx_program_is_terminal(const lynx_program_t*); /* ends in an active screen    */
int64_t lynx_program_record_floats(const lynx_program_t*);

/* Bytes of caller-supplied workspace needed by the particle launchers for B settings. */
size_t lynx_workspace_bytes(const lynx_program_t*, int64_t B);

/* K1 -- per-setting map composer.  Replaces Segment.transfer_map (segment.py:329-338),
 * CustomTransferMap.from_merging_elements (custom_transfer_map.py:48-85) and every
 * element's transfer_map (track_methods.py:14-122 and the element files cited above).
 * Writes one record per setting into `records` ([B][record_floats]): the merged affine
 * stage maps (rows 0..5 of each 7x7, row-major 6x7), the per-setting cavity scalars and
 * the moment-shift centres.  `maps_out` (nullable, [B][n_stages][7][7]) receives the
 * full 7x7 of every stage in the reference's layout.  `seed_soa` (nullable) is the
 * incoming beam (used only to seed the moment-shift centre with particle 0).          */
int lynx_compose_maps(const lynx_program_t*, const float* d_params, int64_t B,
                      const float* d_energy_in /*[B]*/, const float* d_seed_soa,
                      int64_t seed_row_stride /*0 = shared beam*/, int64_t seed_ld,
                      float* d_records, float* d_maps_out, float* d_energy_out /*[B]*/,
                      lynx_stream_t stream);

/* K2 (+K4 histogram, K5 moments, BPM snapshots fused) -- fused particle tracker.
 * Replaces Segment.track / Element.track for ParticleBeam (segment.py:340-356,
 * element.py:83-92), Cavity._track_beam (cavity.py:97-246), BPM.track (bpm.py:48-58),
 * Screen.track + Screen.reading (screen.py:126-141,195-211) and the ParticleBeam
 * moment properties (particle_beam.py:736-836).
 *   in_soa    [Bb][6][ld_in]   Bb = B, or 1 with in_row_stride = 0 (one beam, B settings)
 *   out_soa   [B][6][ld_out]   nullable (terminal screen or moments-only)
 *   bpm_sums  [n_bpm][B][2]    double, sums of x and y at each active BPM (ZEROED by the
 *                              call); reading = sums / N
 *   screen_img [B][H][W] float, image layout of the reference (row 0 = max y,
 *                              flipud(hist.T), screen.py:208); ZEROED by the call;
 *                              integer-valued, exact below 2^24 per pixel
 *   edges_x [W+1], edges_y [H+1] float bin edges (screen.py:105-118), device memory
 *   moment_sums [B][28] double  (ZEROED by the call): [0]=N, [1..6]=sum v_i,
 *                              [7..27]=sum v_i v_j (i<=j, row-major upper triangle) of the
 *                              OUTGOING beam (for a terminal screen: the beam at the screen)
 *   moment_mode 0 = none, 1 = diagonal+xxp+yyp only (others left 0), 2 = all 21
 *   workspace   lynx_workspace_bytes() bytes, device; holds the K1 records (K1 is run
 *               inside this call)                                                      */
int lynx_track_particles_fwd(const lynx_program_t*, const float* d_params, int64_t B,
                             const float* d_in_soa, int64_t in_row_stride, int64_t ld_in,
                             int64_t N, float* d_out_soa, int64_t ld_out,
                             double* d_bpm_sums, float* d_screen_img, int32_t W, int32_t H,
                             const float* d_edges_x, const float* d_edges_y,
                             double* d_moment_sums, int32_t moment_mode,
                             const float* d_energy_in, float* d_energy_out,
                             void* d_workspace, lynx_stream_t stream);

/* Same, additionally recording the bounding box of the screen bins that were touched:
 * d_screen_bbox[4] (int32, ZEROED by the call) = max(W - ix), max(ix + 1), max(H - iy),
 * max(iy + 1) over the histogrammed particles (all 0 = none).  Used by the multi-GPU
 * exchange below; pass the box that sits 64 bytes in front of the image of a symmetric
 * buffer slot (lynx_screen_symm_image) to have box + image zeroed in one fused range.   */
int lynx_track_particles_fwd_bbox(const lynx_program_t*, const float* d_params, int64_t B,
                                  const float* d_in_soa, int64_t in_row_stride, int64_t ld_in,
                                  int64_t N, float* d_out_soa, int64_t ld_out,
                                  double* d_bpm_sums, float* d_screen_img, int32_t W, int32_t H,
                                  const float* d_edges_x, const float* d_edges_y,
                                  double* d_moment_sums, int32_t moment_mode,
                                  const float* d_energy_in, float* d_energy_out,
                                  void* d_workspace, int32_t* d_screen_bbox,
                                  lynx_stream_t stream);

/* Same, for a PERSISTENT screen image (an RL loop or a scan reading the same screen every
 * step): d_screen_img and d_img_state [8] (int32) are zero-initialised ONCE by the caller and
 * passed unchanged on every call, with `epoch` increasing by one per call.  Instead of zeroing
 * the whole B*H*W image (20 MB for the ARES screen) the call clears only the rectangle the
 * previous call lit -- its bounding box, kept in the state -- and records its own.  After the
 * call the image is exactly what lynx_track_particles_fwd writes.  W % 4 == 0, image 16-byte
 * aligned; returns LYNX_E_BADPROGRAM for lattices the pipelined kernel cannot take (then use
 * lynx_track_particles_fwd and re-zero image and state before coming back).                */
int lynx_track_particles_fwd_persistent(const lynx_program_t*, const float* d_params, int64_t B,
                                        const float* d_in_soa, int64_t in_row_stride,
                                        int64_t ld_in, int64_t N, float* d_out_soa, int64_t ld_out,
                                        double* d_bpm_sums, float* d_screen_img, int32_t W,
                                        int32_t H, const float* d_edges_x, const float* d_edges_y,
                                        double* d_moment_sums, int32_t moment_mode,
                                        const float* d_energy_in, float* d_energy_out,
                                        void* d_workspace, int32_t* d_img_state, uint32_t epoch,
                                        lynx_stream_t stream);

/* The same call for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  No counterpart in the reference.
 * There is no workspace argument: the K1 -> K2 records live in
 * a static device array of LYNX_NOALLOC_WORKSPACE_BYTES bytes that every GPU loading the
 * library holds until exit; the call allocates nothing (no cudaMalloc / cudaFree).  Returns
 * LYNX_E_TOOLARGE when lynx_workspace_bytes(prog, B) exceeds that size (ARES EA: B <= ~4000
 * settings).  All calls on a device share the array: calls that overlap on the GPU (different
 * streams, concurrently replayed graphs) race; calls on one stream do not.  Inputs and outputs
 * are the caller's, exactly as for lynx_track_particles_fwd.                                  */
#define LYNX_NOALLOC_WORKSPACE_BYTES (1 << 20)
int lynx_track_particles_fwd_noalloc(const lynx_program_t*, const float* d_params, int64_t B,
                                     const float* d_in_soa, int64_t in_row_stride, int64_t ld_in,
                                     int64_t N, float* d_out_soa, int64_t ld_out,
                                     double* d_bpm_sums, float* d_screen_img, int32_t W, int32_t H,
                                     const float* d_edges_x, const float* d_edges_y,
                                     double* d_moment_sums, int32_t moment_mode,
                                     const float* d_energy_in, float* d_energy_out,
                                     lynx_stream_t stream);

/* Multi-GPU exchange step of the particle-sharded path (one process per GPU; the reference
 * has no distributed code): all-reduce (SUM) of the screen image over
 * NVLink peer memory, restricted to the union bounding box of the bins any rank touched.
 *   - every rank allocates lynx_screen_symm_bytes(B, W, H) bytes of PEER-MAPPED symmetric
 *     memory, zero-initialised once (e.g. torch symmetric memory; cudaMalloc + IPC handles),
 *     and exchanges the base pointers: peer_bases[r] = rank r's buffer as mapped here
 *   - step k tracks into slot = k & 1: image = lynx_screen_symm_image(base, slot, ..),
 *     bbox = (int32_t*)image - 16, via lynx_track_particles_fwd_bbox
 *   - lynx_screen_allreduce(.., slot, epoch = k + 1, ..) then runs ONE kernel: cross-GPU
 *     barrier on epoch flags (each rank pushes its box with its signal), union box, and
 *     d_out_img [B][H][W] = sum over ranks inside the box (coalesced peer loads), zeros
 *     outside.  Exact (integer counts < 2^24).  d_out_img is PERSISTENT: pass the same
 *     buffer, zero-initialised once, with the same zero-initialised d_state [8] on every
 *     call -- only the union of this and the previous call's rectangle is rewritten.
 *   epoch must increase by one per call on every rank; W % 4 == 0.                          */
size_t lynx_screen_symm_bytes(int64_t B, int32_t W, int32_t H);
float* lynx_screen_symm_image(void* d_symm_base, int32_t slot, int64_t B, int32_t W, int32_t H);
int lynx_screen_allreduce(void* const* peer_bases /*host array [world]*/, int32_t world,
                          int32_t rank, int32_t slot, uint32_t epoch, int64_t B, int32_t W,
                          int32_t H, float* d_out_img, int32_t* d_state /*[8]*/,
                          lynx_stream_t stream);
/* Same, with the multicast (NVLS) mapping of the symmetric buffer: the peers' pixels are
 * summed inside the NVSwitch by multimem.ld_reduce instead of world separate peer loads.
 * multicast_base = NULL selects the peer-load path. */
int lynx_screen_allreduce_mc(void* const* peer_bases, void* multicast_base, int32_t world,
                             int32_t rank, int32_t slot, uint32_t epoch, int64_t B, int32_t W,
                             int32_t H, float* d_out_img, int32_t* d_state, lynx_stream_t stream);

/* K3 -- adjoint of lynx_track_particles_fwd w.r.t. every parameter column and
 * (optionally) the incoming particles.  The reference has no custom_vjp; this is the
 * vjp of segment.py:340-356 that the array framework would derive op by op.
 *   g_out_soa     [B][6][ld_g]  cotangent of the outgoing particles, nullable
 *   g_moment_sums [B][28]       cotangent of moment_sums, nullable
 *   g_bpm_sums    [n_bpm][B][2] cotangent of bpm_sums, nullable
 *   g_params      float, same layout as params (shared columns reduced over B); ZEROED
 *   g_in_soa      [Bb][6][ld_gin] nullable (for Bb == 1 < B reduced over settings)
 *   g_energy_in   [B] nullable                                                          */
int lynx_track_particles_bwd(const lynx_program_t*, const float* d_params, int64_t B,
                             const float* d_in_soa, int64_t in_row_stride, int64_t ld_in,
                             int64_t N, const float* d_g_out_soa, int64_t ld_g,
                             const double* d_g_moment_sums, const double* d_g_bpm_sums,
                             const float* d_energy_in, float* d_g_params,
                             float* d_g_in_soa, int64_t ld_gin, void* d_workspace,
                             lynx_stream_t stream);
size_t lynx_bwd_workspace_bytes(const lynx_program_t*, int64_t B);

/* K6 -- ParameterBeam step.  Replaces Element.track for ParameterBeam
 * (element.py:71-82), the ParameterBeam branches of Cavity._track_beam
 * (cavity.py:134-140,202-218), BPM.track and Screen.track.
 *   mu_in [Bm][7], cov_in [Bm][7][7]  Bm = B or 1 (beam_stride 0: one beam, B settings)
 *   mu_out [B][7], cov_out [B][7][7]  (for a terminal screen: the beam AT the screen,
 *                                      misalignment already subtracted)
 *   bpm_out [n_bpm][B][2] float readings                                                 */
int lynx_parambeam_fwd(const lynx_program_t*, const float* d_params, int64_t B,
                       const float* d_mu_in, const float* d_cov_in, int64_t beam_stride,
                       const float* d_energy_in, float* d_mu_out, float* d_cov_out,
                       float* d_bpm_out, float* d_energy_out, lynx_stream_t stream);
int lynx_parambeam_bwd(const lynx_program_t*, const float* d_params, int64_t B,
                       const float* d_mu_in, const float* d_cov_in, int64_t beam_stride,
                       const float* d_energy_in, const float* d_g_mu, const float* d_g_cov,
                       const float* d_g_bpm, float* d_g_params, float* d_g_mu_in,
                       float* d_g_cov_in, lynx_stream_t stream);

/* K7 -- Screen.reading for a ParameterBeam (screen.py:157-194): bivariate-normal pdf of
 * (x, y) from mu [B][7], cov [B][7][7] on the grid of pixel left edges
 * xs = arange(left, right, hstep) [W], ys = arange(bottom, top, vstep) [H] (screen.py:181-191,
 * taken from the host as data), indexed [b][ix][iy] and flipped along x as the reference
 * does (:192-194, quirk B8): img[b][i][j] = pdf(xs[W-1-i], ys[j]).  img is [B][W][H].     */
int lynx_screen_pdf(const float* d_mu, const float* d_cov, int64_t B, const float* d_xs,
                    const float* d_ys, int32_t W, int32_t H, float* d_img, lynx_stream_t stream);

/* K9 -- Gaussian beam generation on the device.  Replaces the sampling inside
 * ParticleBeam.from_parameters / from_twiss (particle_beam.py:143-170: one
 * MultivariateNormal(mean, cov).sample((N,)) per batch row on the host + device copy).
 *   d_mean [B][6], d_sqrt_cov [B][6][6] row-major A with A A^T = cov  ->  d_soa [B][6][ld]:
 *   x = mean + A z, z ~ N(0, I) from Philox4x32-10 keyed by `seed`, counter =
 *   (particle / 4, b, coordinate); particle n of this buffer is global particle
 *   first_particle + n (multiple of 4), so shards of one beam generated on different GPUs
 *   are slices of the same stream.  Padding lanes [N, ld) are written as zeros.          */
int lynx_beam_generate(const float* d_mean, const float* d_sqrt_cov, int64_t B, int64_t N,
                       int64_t first_particle, uint64_t seed, float* d_soa, int64_t ld,
                       lynx_stream_t stream);
/* One Philox4x32-10 block on the HOST (known-answer hook for the CPU tests). */
void lynx_philox4x32_10(const uint32_t counter[4], const uint32_t key[2], uint32_t out[4]);

/* K9b -- ParticleBeam.uniform_3d_ellipsoid on the device (particle_beam.py:266-385): (x, y, s)
 * uniform inside the ellipsoid with semi-axes d_radii [B][3] = (r_x, r_y, r_s) by per-particle
 * rejection from the bounding box (the reference's test, particle_beam.py:346, in exact fp32
 * round-to-nearest operations); xp, yp, p = d_sigma [B][3] * z with the same normals z that
 * lynx_beam_generate draws for these coordinates under the same seed (the reference builds
 * the beam as from_parameters(...) with the spatial coordinates replaced, :363-383).
 * Counter layout: attempt a of global particle n uses Philox counter (n_lo, n_hi, b, 16 + a). */
int lynx_beam_generate_ellipsoid(const float* d_radii, const float* d_sigma, int64_t B, int64_t N,
                                 int64_t first_particle, uint64_t seed, float* d_soa, int64_t ld,
                                 lynx_stream_t stream);

/* K10 -- ParticleBeam.transformed_to (particle_beam.py:580-715), the particle pass :696-702:
 *     out = (in - old_mu) / old_sigma * new_sigma + new_mu          per coordinate, per row
 * in exactly that order of fp32 operations (bit-identical to the reference's array ops).
 *   d_table [B][24] = old_mu[6], old_sigma[6], new_sigma[6], new_mu[6] per beam row
 *   in_row_stride    floats between beam rows of d_in_soa (6*ld_in, or 0: one beam, B targets)
 * Padding lanes [N, ld_out) are written as zeros.  48 B per particle.                      */
int lynx_beam_transform(const float* d_in_soa, int64_t in_row_stride, int64_t ld_in, int64_t B,
                        int64_t N, const float* d_table, float* d_out_soa, int64_t ld_out,
                        lynx_stream_t stream);

/* K11 -- Aperture.track (aperture.py:69-109): order-preserving compaction of the particles
 * that pass the aperture; the complement goes to d_lost_soa (the reference's lost_particles).
 *   shape 0 = "rectangular": -x_max < x < x_max and -y_max < y < y_max        (:80-84)
 *         1 = "elliptical":  x^2/x_max^2 + y^2/y_max^2 <= 1, fp32 op by op    (:85-88)
 *   d_x_max, d_y_max [R] (lim_stride 1) or one value for all rows (lim_stride 0)
 *   d_charges_in [R][ld_q] nullable -> d_charges_out / d_charges_lost [R][ld_out] (nullable)
 *   d_out_soa, d_lost_soa (nullable) [R][6][ld_out], 16-byte aligned, ld_out % 4 == 0 (as the
 *   charge outputs): row r holds counts[r] survivors resp.
 *   N - counts[r] lost particles in their incoming order; lanes beyond are NOT written
 *   d_counts [R] int64 survivors per beam row
 *   d_workspace  lynx_aperture_workspace_bytes(R, N) bytes
 * Three launches (tile counts, scan, scatter); 56 B per particle.                          */
size_t lynx_aperture_workspace_bytes(int64_t R, int64_t N);
int lynx_aperture_compact(const float* d_in_soa, int64_t R, int64_t N, int64_t ld_in,
                          const float* d_charges_in, int64_t ld_q, const float* d_x_max,
                          const float* d_y_max, int64_t lim_stride, int32_t shape,
                          float* d_out_soa, float* d_lost_soa, int64_t ld_out,
                          float* d_charges_out, float* d_charges_lost, int64_t* d_counts,
                          void* d_workspace, lynx_stream_t stream);

/* K6-along -- ParameterBeam tracked element by element, one snapshot per program op
 * (the loop of Segment.plot_twiss, segment.py:490-518: element.track(beam) per element,
 * every intermediate beam kept).  Nothing is merged: op i's own map is applied to the beam
 * after op i-1 (element.py:71-82; active cavities as in lynx_parambeam_fwd).  BPMs and
 * screens pass the beam through.
 *   d_snap_mu [n_ops][B][7], d_snap_cov [n_ops][B][7][7], d_snap_energy [n_ops][B]          */
int lynx_parambeam_along(const lynx_program_t*, const float* d_params, int64_t B,
                         const float* d_mu_in, const float* d_cov_in, int64_t beam_stride,
                         const float* d_energy_in, float* d_snap_mu, float* d_snap_cov,
                         float* d_snap_energy, lynx_stream_t stream);

/* Layout converters between the reference's AoS (..,N,7) ParticleBeam.particles
 * (particle_beam.py:24-45) and the SoA-6 device layout. */
int lynx_aos7_to_soa6(const float* d_aos /*[R][N][7]*/, int64_t R, int64_t N,
                      float* d_soa /*[R][6][ld]*/, int64_t ld, lynx_stream_t stream);
int lynx_soa6_to_aos7(const float* d_soa, int64_t R, int64_t N, int64_t ld,
                      float* d_aos /*[R][N][7], column 6 := 1*/, lynx_stream_t stream);

/* Launch counter (kernels enqueued by this library since load / last reset). */
int64_t lynx_launch_count(void);
void lynx_launch_count_reset(void);

/* Measurement hooks (bench.py roofline): when enabled, the launchers bracket their
 * dominant kernel with CUDA events on the launch stream.  lynx_profile_last_ms
 * synchronises on the stop event and returns the device time of the last launch of
 * kernel class `which` (0 = K1 composer, 1 = K2 tracker, 2 = K6 ParameterBeam,
 * 3 = K3 adjoint, 4 = K10 transform, 5 = K11 aperture scatter), or a negative value if none
 * was recorded. */
void lynx_profile_enable(int on);
float lynx_profile_last_ms(int which);
/* Debug: block 0 of lynx_screen_allreduce stamps %globaltimer (ns) into
 * d_stamps[(epoch & 1023) * 8 + {0: resident, 1: K2 complete, 2: all peers arrived,
 * 3: rectangle written, 4: K2 past its wait on K1, 5: K2 CTA 0 done}]; NULL switches it
 * off.  Device buffer of 8192 uint64. */
void lynx_exchange_trace(unsigned long long* d_stamps);
/* Debug: a one-thread kernel that writes %globaltimer (ns) to *d_cell. */
void lynx_debug_stamp(unsigned long long* d_cell, lynx_stream_t stream);

/* ---- host-buffer entry points (what a NumPy / host-array binding calls) ------------
 * Same semantics as the device launchers, but every pointer is HOST memory (pinned for
 * full copy speed, pageable accepted).  Copies in, runs, copies out, synchronises.
 * Device scratch is cached on a context handle.                                        */
typedef struct lynx_host_ctx lynx_host_ctx_t;
int lynx_host_ctx_create(int device, lynx_host_ctx_t** out);
void lynx_host_ctx_destroy(lynx_host_ctx_t*);

/* Context options.  LYNX_HOST_OPT_PERSISTENT_IMAGE = 1: the caller passes the SAME host
 * screen_img buffer on every lynx_host_track_particles call of this context; the library
 * then moves only the bounding rectangle of the lit pixels over PCIe (clearing the
 * previous call's rectangle on the host first) instead of the full 4 W H bytes -- the
 * host image is bit-identical to the dense transfer.  A new buffer is zero-filled once. */
enum { LYNX_HOST_OPT_PERSISTENT_IMAGE = 1 };
int lynx_host_ctx_set_option(lynx_host_ctx_t*, int32_t option, int32_t value);

int lynx_host_track_particles(lynx_host_ctx_t*, const lynx_program_t*, const float* params,
                              int64_t n_param_floats, int64_t B, const float* in_soa,
                              int64_t Bb, int64_t N, const float* energy_in,
                              float* out_soa /*nullable [B][6][N]*/,
                              double* bpm_sums /*nullable*/, float* screen_img /*nullable*/,
                              int32_t W, int32_t H, const float* edges_x,
                              const float* edges_y, double* moment_sums /*nullable*/,
                              int32_t moment_mode, float* energy_out /*nullable*/);

int lynx_host_parambeam(lynx_host_ctx_t*, const lynx_program_t*, const float* params,
                        int64_t n_param_floats, int64_t B, const float* mu_in,
                        const float* cov_in, int64_t Bm, const float* energy_in,
                        float* mu_out, float* cov_out, float* bpm_out /*nullable*/,
                        float* energy_out /*nullable*/);

#ifdef __cplusplus
}
#endif
#endif /* LYNX_B200_H */

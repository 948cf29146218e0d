// K9: Gaussian beam generation straight into the SoA-6 planes.
//
// Replaces the sampling of ParticleBeam.from_parameters / from_twiss
// (lynx/particles/particle_beam.py:143-170: one MultivariateNormal(mean, cov).sample((N,))
// per batch row on the host, then a copy to the device).  At 1e8 particles the host-side
// generation plus the 2.4 GB H2D copy dominate end to end; here the beam never exists on
// the host.  The reference draws from the array framework's global generator, so there is
// no bit-level contract to match: parity is (i) the counter-based generator against its
// published known-answer vectors and the CPU restatement (oracle/generate.py), (ii) the
// sample moments against the requested mean / covariance.
//
//   x[b][:, n] = mean[b] + A[b] z,   z ~ N(0, I_6),   A A^T = cov (any square root; the host
//   mirror passes the Cholesky factor, or V sqrt(L) for a semi-definite covariance)
//
// Random stream: Philox4x32-10, key = (seed_lo, seed_hi), counter = (group_lo, group_hi, b, j)
// where group = particle index / 4 and call j in 0..5 yields the j-th standard normal of the
// group's four particles (two Box-Muller pairs: words (0,1) -> particles 0,1; words (2,3) ->
// particles 2,3).  Every particle's value depends only on (seed, b, n): independent of grid
// shape, of the sharding (a rank passes its first global particle index) and reproducible.
// 24 B written per particle; the kernel is bound by instruction issue (60 Philox rounds per
// four particles), not by HBM.
#include <algorithm>

#include "common.cuh"

namespace lynx {

__host__ __device__ __forceinline__ void philox_round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
  const uint64_t p0 = (uint64_t)0xD2511F53u * c[0];
  const uint64_t p1 = (uint64_t)0xCD9E8D57u * c[2];
  const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0;
  const uint32_t n1 = (uint32_t)p1;
  const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
  const uint32_t n3 = (uint32_t)p0;
  c[0] = n0;
  c[1] = n1;
  c[2] = n2;
  c[3] = n3;
}

__host__ __device__ __forceinline__ void philox4x32_10(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    philox_round(c, k0, k1);
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
}

// Uniforms use the top 24 bits of a word: u = (k + 0.5) 2^-24 is exact in fp32 and never 0
// or 1 (with all 32 bits, (x + 0.5) 2^-32 rounds to 1.0f for x >= 2^32 - 128).
//
// Box-Muller with the SFU: r = sqrt(-2 ln u1), theta = 2 pi u2.
//   * -2 ln u: MUFU.LG2 has an ABSOLUTE error of ~2^-22, useless next to u = 1 where ln u -> 0,
//     so for 1 - u < 2^-6 (1 - u is exact in fp32) the series of -ln(1 - v) is used instead;
//   * sqrt as t * rsqrt(t) (t > 0 always);
//   * sin / cos on theta - pi in (-pi, pi), where MUFU.SIN / COS are accurate to ~2^-21 abs.
// Against the fp64 restatement (oracle/generate.py) a normal is off by < 4e-6.
__device__ __forceinline__ float neg2ln_u01(uint32_t x) {
  const uint32_t k = x >> 8;
  const float u = ((float)k + 0.5f) * 5.9604644775390625e-8f;
  const float v = ((float)(0xFFFFFFu - k) + 0.5f) * 5.9604644775390625e-8f;  // 1 - u, exact
  const float series = 2.0f * v * fmaf(v, fmaf(v, fmaf(v, 0.25f, 0.33333334f), 0.5f), 1.0f);
  const float lg = -1.3862943611198906f * __log2f(u);
  return v < 0.015625f ? series : lg;
}

__device__ __forceinline__ void box_muller(uint32_t a, uint32_t b, float& z0, float& z1) {
  const float t = neg2ln_u01(a);
  const float r = t * rsqrtf(t);
  const float th = (((float)(b >> 8) + 0.5f) * 5.9604644775390625e-8f - 0.5f) * 6.2831853071795865f;
  z0 = -r * __cosf(th);  // cos(theta) = -cos(theta - pi)
  z1 = -r * __sinf(th);
}

constexpr int kGenTPB = 256;

__global__ void __launch_bounds__(kGenTPB) beam_generate_kernel(
    const float* __restrict__ mean, const float* __restrict__ A, int64_t N, int64_t first,
    uint32_t seed_lo, uint32_t seed_hi, float* __restrict__ soa, int64_t ld) {
  const int64_t b = blockIdx.y;
  __shared__ float s_A[36], s_m[6];
  if (threadIdx.x < 36) s_A[threadIdx.x] = __ldg(A + b * 36 + threadIdx.x);
  if (threadIdx.x < 6) s_m[threadIdx.x] = __ldg(mean + b * 6 + threadIdx.x);
  __syncthreads();
  float* out = soa + b * 6 * ld;
  // `first` (global index of this buffer's particle 0) is a multiple of 4
  const int64_t groups = (N + 3) >> 2;
  for (int64_t g = (int64_t)blockIdx.x * kGenTPB + threadIdx.x; g < groups;
       g += (int64_t)gridDim.x * kGenTPB) {
    const uint64_t gg = (uint64_t)(g + (first >> 2));
    float z[6][4];
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      uint32_t c[4] = {(uint32_t)gg, (uint32_t)(gg >> 32), (uint32_t)b, (uint32_t)j};
      philox4x32_10(c, seed_lo, seed_hi);
      box_muller(c[0], c[1], z[j][0], z[j][1]);
      box_muller(c[2], c[3], z[j][2], z[j][3]);
    }
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      float v[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        float acc = s_m[i];
#pragma unroll
        for (int k = 0; k < 6; ++k) acc = fmaf(s_A[i * 6 + k], z[k][q], acc);
        v[q] = (g * 4 + q < N) ? acc : 0.f;  // padding lanes [N, ld) stay zero
      }
      __stcs(reinterpret_cast<float4*>(out + i * ld + g * 4), make_float4(v[0], v[1], v[2], v[3]));
    }
  }
}

// K9b: ParticleBeam.uniform_3d_ellipsoid (particle_beam.py:266-385): (x, y, s) uniform inside
// the ellipsoid with semi-axes (r_x, r_y, r_s) by rejection from the bounding box, xp / yp / p
// from the uncorrelated Gaussian beam of from_parameters (particle_beam.py:363-374).
//
// The reference redraws whole batches of N candidates from the global generator until N were
// accepted; here every particle rejects on its own private counter stream, so a particle
// depends only on (seed, setting, global particle index):
//   attempt a of particle n: Philox counter (n_lo, n_hi, b, 16 + a), words 0..2 ->
//   t = ((w >> 9) - 2^22 + 0.5) 2^-22 in (-1, 1), exact in fp32 (= (rand - 0.5) * 2 of
//   particle_beam.py:342-344 for rand = ((w >> 9) + 0.5) 2^-23);  candidate = t * r;
//   accepted iff x^2/r_x^2 + y^2/r_y^2 + s^2/r_s^2 < 1 (particle_beam.py:346), evaluated
//   with explicit round-to-nearest fp32 operations (no FMA contraction), so the CPU
//   restatement (oracle/generate.py) makes bit-identical decisions.  After kEllMaxTries
//   rejections (probability 0.4764^64 ~ 2e-21) the particle is placed at the centre.
// The momentum planes use K9's stream (coordinates j = 1, 3, 5 of the particle's group): the
// beam equals from_parameters(seed) with the three spatial planes replaced, which is exactly
// how the reference builds it.
constexpr int kEllMaxTries = 64;

__host__ __device__ __forceinline__ float ell_unit(uint32_t w) {
  return ((float)((int32_t)(w >> 9) - 4194304) + 0.5f) * 2.384185791015625e-7f;  // 2^-22
}

__global__ void __launch_bounds__(kGenTPB) beam_ellipsoid_kernel(
    const float* __restrict__ radii, const float* __restrict__ sigma, int64_t N, int64_t first,
    uint32_t seed_lo, uint32_t seed_hi, float* __restrict__ soa, int64_t ld) {
  const int64_t b = blockIdx.y;
  const float rx = __ldg(radii + b * 3 + 0), ry = __ldg(radii + b * 3 + 1),
              rs = __ldg(radii + b * 3 + 2);
  const float sg[3] = {__ldg(sigma + b * 3 + 0), __ldg(sigma + b * 3 + 1), __ldg(sigma + b * 3 + 2)};
  const float rx2 = __fmul_rn(rx, rx), ry2 = __fmul_rn(ry, ry), rs2 = __fmul_rn(rs, rs);
  float* out = soa + b * 6 * ld;
  const int64_t groups = (N + 3) >> 2;
  for (int64_t g = (int64_t)blockIdx.x * kGenTPB + threadIdx.x; g < groups;
       g += (int64_t)gridDim.x * kGenTPB) {
    float pos[3][4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint64_t n = (uint64_t)(first + g * 4 + q);
      float x = 0.f, y = 0.f, s = 0.f;
      for (int a = 0; a < kEllMaxTries; ++a) {
        uint32_t c[4] = {(uint32_t)n, (uint32_t)(n >> 32), (uint32_t)b, (uint32_t)(16 + a)};
        philox4x32_10(c, seed_lo, seed_hi);
        const float cx = __fmul_rn(ell_unit(c[0]), rx), cy = __fmul_rn(ell_unit(c[1]), ry),
                    cs = __fmul_rn(ell_unit(c[2]), rs);
        const float t = __fadd_rn(__fadd_rn(__fdiv_rn(__fmul_rn(cx, cx), rx2),
                                            __fdiv_rn(__fmul_rn(cy, cy), ry2)),
                                  __fdiv_rn(__fmul_rn(cs, cs), rs2));
        if (t < 1.0f) {
          x = cx, y = cy, s = cs;
          break;
        }
      }
      const bool live = g * 4 + q < N;
      pos[0][q] = live ? x : 0.f;
      pos[1][q] = live ? y : 0.f;
      pos[2][q] = live ? s : 0.f;
    }
    const uint64_t gg = (uint64_t)(g + (first >> 2));
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      uint32_t c[4] = {(uint32_t)gg, (uint32_t)(gg >> 32), (uint32_t)b, (uint32_t)(2 * j + 1)};
      philox4x32_10(c, seed_lo, seed_hi);
      float z[4];
      box_muller(c[0], c[1], z[0], z[1]);
      box_muller(c[2], c[3], z[2], z[3]);
#pragma unroll
      for (int q = 0; q < 4; ++q) z[q] = (g * 4 + q < N) ? __fmul_rn(sg[j], z[q]) : 0.f;
      __stcs(reinterpret_cast<float4*>(out + (2 * j) * ld + g * 4),
             make_float4(pos[j][0], pos[j][1], pos[j][2], pos[j][3]));
      __stcs(reinterpret_cast<float4*>(out + (2 * j + 1) * ld + g * 4),
             make_float4(z[0], z[1], z[2], z[3]));
    }
  }
}

}  // namespace lynx

using namespace lynx;

extern "C" int lynx_beam_generate(const float* d_mean, const float* d_sqrt_cov, int64_t B,
                                  int64_t N, int64_t first_particle, uint64_t seed, float* d_soa,
                                  int64_t ld, lynx_stream_t stream) {
  if (!d_mean || !d_sqrt_cov || !d_soa || B <= 0 || B > 65535 || N <= 0 || ld < N || (ld & 3) ||
      first_particle < 0 || (first_particle & 3) || ((uintptr_t)d_soa & 15))
    return LYNX_E_BADARG;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int64_t groups = (N + 3) >> 2;
  const int64_t per_b = std::max<int64_t>(1, (8LL * sms + B - 1) / B);
  const unsigned gx = (unsigned)std::min<int64_t>((groups + kGenTPB - 1) / kGenTPB, per_b);
  beam_generate_kernel<<<dim3(gx, (unsigned)B), kGenTPB, 0, (cudaStream_t)stream>>>(
      d_mean, d_sqrt_cov, N, first_particle, (uint32_t)seed, (uint32_t)(seed >> 32), d_soa, ld);
  count_launch();
  LYNX_CUDA_OK(cudaGetLastError());
  return LYNX_OK;
}

extern "C" int lynx_beam_generate_ellipsoid(const float* d_radii, const float* d_sigma, int64_t B,
                                            int64_t N, int64_t first_particle, uint64_t seed,
                                            float* d_soa, int64_t ld, lynx_stream_t stream) {
  if (!d_radii || !d_sigma || !d_soa || B <= 0 || B > 65535 || N <= 0 || ld < N || (ld & 3) ||
      first_particle < 0 || (first_particle & 3) || ((uintptr_t)d_soa & 15))
    return LYNX_E_BADARG;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int64_t groups = (N + 3) >> 2;
  const int64_t per_b = std::max<int64_t>(1, (8LL * sms + B - 1) / B);
  const unsigned gx = (unsigned)std::min<int64_t>((groups + kGenTPB - 1) / kGenTPB, per_b);
  beam_ellipsoid_kernel<<<dim3(gx, (unsigned)B), kGenTPB, 0, (cudaStream_t)stream>>>(
      d_radii, d_sigma, N, first_particle, (uint32_t)seed, (uint32_t)(seed >> 32), d_soa, ld);
  count_launch();
  LYNX_CUDA_OK(cudaGetLastError());
  return LYNX_OK;
}

// Host-side known-answer hook for the counter-based generator (no device work): runs one
// Philox4x32-10 block so the "not gpu" tests can pin it against the published vectors.
extern "C" void lynx_philox4x32_10(const uint32_t counter[4], const uint32_t key[2], uint32_t out[4]) {
  uint32_t c[4] = {counter[0], counter[1], counter[2], counter[3]};
  philox4x32_10(c, key[0], key[1]);
  for (int i = 0; i < 4; ++i) out[i] = c[i];
}

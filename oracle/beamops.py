"""CPU restatement of the beam-level operations next to the tracking path:
``ParticleBeam.transformed_to``, ``ParticleBeam.uniform_3d_ellipsoid``, ``Aperture.track``
and the element-by-element loop of ``Segment.plot_twiss``.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  numpy float32 arithmetic is IEEE
round-to-nearest op by op, which is also what the array framework of the reference does,
so the byte-level comparisons against the kernels (csrc/beamops.cu, csrc/generate.cu) are
exact.  Parity unpinned: the reference holds no golden vector for any of these (its tests
for them only check shapes and moments, tests/test_particle_beam.py); the relational checks
it does make (moments after ``transformed_to``, all particles inside the ellipsoid, fewer
particles behind an aperture: tests/test_aperture.py, tests/test_particle_beam.py) are
restated in tests/test_beamops_oracle.py.
"""

import numpy as np
import torch

from .generate import philox4x32_10, standard_normals

F32 = np.float32


# ------------------------------------------------------------------ transformed_to
def beam_stats(planes: np.ndarray):
    """``planes`` (6, N) float32 -> (mu[6], sigma[6]) as particle_beam.py:736-822 computes
    them: mean, and torch.std (unbiased, N-1), in float64 then rounded to float32 (the
    device path accumulates the raw sums in double)."""
    p = planes.astype(np.float64)
    mu = p.mean(axis=1)
    sd = p.std(axis=1, ddof=1)
    return mu.astype(F32), sd.astype(F32)


def transformed_to(planes: np.ndarray, old_mu, old_sigma, new_mu, new_sigma) -> np.ndarray:
    """particle_beam.py:696-702, fp32 op by op:
    ``(phase_space - old_mu) / old_sigma * new_sigma + new_mu``.
    ``planes`` (6, N); the four vectors (6,).  particle_beam.py:666-693 zeroes the s and p
    entries of old_mu / new_mu -- the caller passes them that way."""
    x = planes.astype(F32)
    om, os_, nm, ns = (np.asarray(v, dtype=F32).reshape(6, 1) for v in
                       (old_mu, old_sigma, new_mu, new_sigma))
    return ((((x - om) / os_) * ns) + nm).astype(F32)


# ------------------------------------------------------------------ Aperture
def aperture_mask(xs: np.ndarray, ys: np.ndarray, x_max, y_max, shape: str) -> np.ndarray:
    """aperture.py:80-88 in float32."""
    xs, ys = xs.astype(F32), ys.astype(F32)
    xm, ym = F32(x_max), F32(y_max)
    if shape == "rectangular":
        return (xs > -xm) & (xs < xm) & (ys > -ym) & (ys < ym)
    if shape == "elliptical":
        with np.errstate(divide="ignore", invalid="ignore"):
            return (xs * xs) / (xm * xm) + (ys * ys) / (ym * ym) <= F32(1.0)
    raise ValueError(f"Unknown aperture shape {shape}")


def aperture_track(planes: np.ndarray, x_max, y_max, shape: str, charges=None):
    """aperture.py:89-97: survivors and lost particles, both in incoming order.
    ``planes`` (6, N).  Returns (kept (6, n), lost (6, N - n), kept_q, lost_q)."""
    m = aperture_mask(planes[0], planes[2], x_max, y_max, shape)
    kq = lq = None
    if charges is not None:
        kq, lq = charges[m], charges[~m]
    return planes[:, m], planes[:, ~m], kq, lq


# ------------------------------------------------------------------ uniform_3d_ellipsoid
MAX_TRIES = 64


def _unit(w: np.ndarray) -> np.ndarray:
    """((w >> 9) - 2^22 + 0.5) 2^-22, exact in float32; = (rand - 0.5) * 2 of
    particle_beam.py:342-344 with rand = ((w >> 9) + 0.5) 2^-23."""
    m = (w >> np.uint32(9)).astype(np.int64) - (1 << 22)
    return ((m.astype(F32) + F32(0.5)) * F32(2.0**-22)).astype(F32)


def ellipsoid_positions(n: int, b: int, radii, seed: int, first: int = 0) -> np.ndarray:
    """(3, n) float32 x, y, s of batch row ``b``: per-particle rejection
    (particle_beam.py:338-360; the acceptance test :346 in float32), attempt ``a`` of global
    particle ``i`` drawing Philox counter (i_lo, i_hi, b, 16 + a)."""
    rx, ry, rs = (F32(r) for r in radii)
    idx = np.arange(n, dtype=np.uint64) + np.uint64(first)
    out = np.zeros((3, n), dtype=F32)
    todo = np.ones(n, dtype=bool)
    key = (seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    for a in range(MAX_TRIES):
        if not todo.any():
            break
        i = idx[todo]
        ctr = np.stack([i & np.uint64(0xFFFFFFFF), i >> np.uint64(32), np.full_like(i, b),
                        np.full_like(i, 16 + a)], -1)
        w = philox4x32_10(ctr, key)
        cx, cy, cs = _unit(w[:, 0]) * rx, _unit(w[:, 1]) * ry, _unit(w[:, 2]) * rs
        t = ((cx * cx) / (rx * rx) + (cy * cy) / (ry * ry)) + (cs * cs) / (rs * rs)
        ok = t < F32(1.0)
        sel = np.flatnonzero(todo)[ok]
        out[0, sel], out[1, sel], out[2, sel] = cx[ok], cy[ok], cs[ok]
        todo[sel] = False
    return out


def uniform_3d_ellipsoid(n: int, radii, sigma3, seed: int, first: int = 0) -> np.ndarray:
    """``radii`` (B, 3), ``sigma3`` (B, 3) = (sigma_xp, sigma_yp, sigma_p) -> (B, 6, n) float64:
    positions from the rejection sampler, momenta sigma * z with the normals of the
    from_parameters stream (particle_beam.py:363-383)."""
    radii = np.asarray(radii, dtype=np.float64)
    sigma3 = np.asarray(sigma3, dtype=np.float64)
    out = np.empty((radii.shape[0], 6, n))
    for b in range(radii.shape[0]):
        z = standard_normals(n, b, seed, first)
        pos = ellipsoid_positions(n, b, radii[b], seed, first)
        for j in range(3):
            out[b, 2 * j] = pos[j]
            out[b, 2 * j + 1] = sigma3[b, j] * z[2 * j + 1]
    return out


# ------------------------------------------------------------------ plot_twiss loop
def track_along(elements, beam, **kw):
    """segment.py:495-504 without the zero-length filter: ``element.track(beam)`` element by
    element, every intermediate beam kept (BPMs / screens pass the beam through)."""
    from .lattice import track

    beams = []
    for el in elements:
        if el.kind not in ("bpm", "screen", "marker"):
            beam = track([el], beam, **kw)
        beams.append(beam)
    return beams


def twiss_beta(beam, plane: str) -> torch.Tensor:
    return beam.beta(plane)

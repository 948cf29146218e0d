"""ASTRA particle distributions to beam coordinates (converters/astra.py:8-62, which follows
Ocelot's ``astra2ocelot``).  Host logic only; the data format in front of ``ParticleBeam``.

File columns: x y z [m], px py pz [eV/c], t [ns], q [nC], species index, status flag.  Row 0
is the reference particle; every other row gives z and pz RELATIVE to it.  Particles with a
non-positive status flag are lost and dropped.  The distribution is a snapshot in time; the
beam coordinates are taken on the plane through the reference particle, so every particle is
moved along its own straight line by ``c dt = -z / (beta u_z)``.
"""

import numpy as np
from scipy.constants import physical_constants

ELECTRON_MASS_EV = physical_constants["electron mass energy equivalent in MeV"][0] * 1e6


def read_astra_beam(path):
    """-> (coordinates (N, 6) float64 as x xp y yp s p, reference energy [eV], charges (N,) [C])."""
    table = np.atleast_2d(np.loadtxt(path))
    table = table[table[:, 9] > 0]  # status flag: lost particles out (astra.py:24-28)
    x, y, z = table[:, 0], table[:, 1], table[:, 2].copy()
    px, py, pz = table[:, 3], table[:, 4], table[:, 5].copy()
    pz_ref = pz[0]
    z[0] = 0.0  # the reference particle sits on the plane ...
    pz[0] = 0.0  # ... and carries exactly the reference momentum (astra.py:33-35)
    pz_abs = pz + pz_ref

    mc2 = ELECTRON_MASS_EV
    gamma_ref = np.sqrt((pz_ref / mc2) ** 2 + 1.0)
    beta_ref = np.sqrt(1.0 - gamma_ref ** -2)
    energy = gamma_ref * mc2

    p_abs = np.sqrt(px * px + py * py + pz_abs * pz_abs)
    gamma = np.sqrt(1.0 + (p_abs / mc2) ** 2)
    beta = np.sqrt(1.0 - gamma ** -2)
    ux, uy, uz = px / p_abs, py / p_abs, pz_abs / p_abs
    cdt = -z / (beta * uz)

    out = np.empty((table.shape[0], 6))
    out[:, 0] = x + beta * ux * cdt
    out[:, 1] = px / pz_ref
    out[:, 2] = y + beta * uy * cdt
    out[:, 3] = py / pz_ref
    out[:, 4] = cdt
    out[:, 5] = (gamma / gamma_ref - 1.0) / beta_ref
    charges = np.abs(table[:, 7]) * 1e-9  # nC -> C
    return out, float(energy), charges

"""Oracle: 7x7 homogeneous transfer maps (torch-CPU restatement).

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Follows, entry by entry:
  lynx/track_methods.py:14-122          rotation / base_rmatrix / misalignment
  lynx/accelerator/drift.py:44-62       Drift
  lynx/accelerator/horizontal_corrector.py:52-67, vertical_corrector.py:52-66
  lynx/accelerator/quadrupole.py:66-80  Quadrupole
  lynx/accelerator/dipole.py:96-181, rbend.py:79-80
  lynx/accelerator/cavity.py:248-325    Cavity linear map
  lynx/accelerator/solenoid.py:61-105, undulator.py:48-60

All inputs are torch tensors sharing one batch shape ``(...)``; outputs ``(..., 7, 7)``.
dtype follows the inputs (float32 = the reference's default, float64 = "truth").
"""

import math

import torch
from scipy import constants
from scipy.constants import physical_constants

# lynx/track_methods.py:9-11 -- m_e c^2 / e in eV
REST_ENERGY = (
    constants.electron_mass * constants.speed_of_light**2 / constants.elementary_charge
)
# lynx/accelerator/cavity.py:20 -- same quantity, differently sourced (kept distinct)
ELECTRON_MASS_EV = physical_constants["electron mass energy equivalent in MeV"][0] * 1e6


def _eye(like: torch.Tensor) -> torch.Tensor:
    return torch.eye(7, dtype=like.dtype).repeat(*like.shape, 1, 1)


def rotation_matrix(angle: torch.Tensor) -> torch.Tensor:
    """x-y plane rotation, lynx/track_methods.py:14-34."""
    cs, sn = torch.cos(angle), torch.sin(angle)
    tm = _eye(angle)
    tm[..., 0, 0] = cs
    tm[..., 0, 2] = sn
    tm[..., 1, 1] = cs
    tm[..., 1, 3] = sn
    tm[..., 2, 0] = -sn
    tm[..., 2, 2] = cs
    tm[..., 3, 1] = -sn
    tm[..., 3, 3] = cs
    return tm


def base_rmatrix(length, k1, hx, tilt=None, energy=None) -> torch.Tensor:
    """Universal quad / sector-bend map, lynx/track_methods.py:37-105."""
    tilt = tilt if tilt is not None else torch.zeros_like(length)
    energy = energy if energy is not None else torch.zeros_like(length)

    gamma = energy / REST_ENERGY
    igamma2 = torch.ones_like(length)  # :61 default ONE where gamma == 0 (quirk B9)
    nz = gamma != 0
    igamma2[nz] = 1 / gamma[nz] ** 2
    beta = torch.sqrt(1 - igamma2)

    k1 = k1.clone()
    k1[k1 == 0] = 1e-12  # :67-68

    kx2 = k1 + hx**2
    ky2 = -k1
    kx = torch.sqrt(torch.complex(kx2, torch.zeros_like(kx2)))
    ky = torch.sqrt(torch.complex(ky2, torch.zeros_like(ky2)))
    cx = torch.cos(kx * length).real
    cy = torch.cos(ky * length).real
    sy = length.clone()
    m = ky != 0
    sy[m] = (torch.sin(ky[m] * length[m]) / ky[m]).real
    sx = (torch.sin(kx * length) / kx).real
    dx = hx / kx2 * (1.0 - cx)
    r56 = hx**2 * (length - sx) / kx2 / beta**2
    r56 = r56 - length / beta**2 * igamma2

    R = _eye(length)
    R[..., 0, 0] = cx
    R[..., 0, 1] = sx
    R[..., 0, 5] = dx / beta
    R[..., 1, 0] = -kx2 * sx
    R[..., 1, 1] = cx
    R[..., 1, 5] = sx * hx / beta
    R[..., 2, 2] = cy
    R[..., 2, 3] = sy
    R[..., 3, 2] = -ky2 * sy
    R[..., 3, 3] = cy
    R[..., 4, 0] = sx * hx / beta
    R[..., 4, 1] = dx / beta
    R[..., 4, 5] = r56

    if torch.any(tilt != 0):  # :101 whole-batch branch (quirk B10)
        R = rotation_matrix(-tilt) @ R @ rotation_matrix(tilt)
    return R


def misalignment_matrix(misalignment: torch.Tensor):
    """Entry / exit shift maps, lynx/track_methods.py:108-122.  misalignment (...,2)."""
    batch = misalignment[..., 0]
    R_exit = _eye(batch)
    R_exit[..., 0, 6] = misalignment[..., 0]
    R_exit[..., 2, 6] = misalignment[..., 1]
    R_entry = _eye(batch)
    R_entry[..., 0, 6] = -misalignment[..., 0]
    R_entry[..., 2, 6] = -misalignment[..., 1]
    return R_entry, R_exit


def _drift_igamma2_beta(energy: torch.Tensor):
    """lynx/accelerator/drift.py:52-55 -- default ZERO where gamma == 0."""
    gamma = energy / REST_ENERGY
    igamma2 = torch.zeros_like(gamma)
    nz = gamma != 0
    igamma2[nz] = 1 / gamma[nz] ** 2
    beta = torch.sqrt(1 - igamma2)
    return igamma2, beta


def drift_map(length, energy) -> torch.Tensor:
    """lynx/accelerator/drift.py:44-62."""
    igamma2, beta = _drift_igamma2_beta(energy)
    tm = _eye(length)
    tm[..., 0, 1] = length
    tm[..., 2, 3] = length
    tm[..., 4, 5] = -length / beta**2 * igamma2
    return tm


def corrector_map(length, angle, energy, horizontal: bool) -> torch.Tensor:
    """horizontal_corrector.py:52-67 (R[1,6]) / vertical_corrector.py:52-66 (R[3,6])."""
    tm = drift_map(length, energy)
    tm[..., 1 if horizontal else 3, 6] = angle
    return tm


def quadrupole_map(length, k1, misalignment, tilt, energy) -> torch.Tensor:
    """lynx/accelerator/quadrupole.py:66-80."""
    R = base_rmatrix(length, k1, torch.zeros_like(length), tilt, energy)
    if torch.all(misalignment == 0):
        return R
    R_entry, R_exit = misalignment_matrix(misalignment)
    return R_exit @ R @ R_entry


def _dipole_hx(length, angle):
    """lynx/accelerator/dipole.py:96-102."""
    hx = torch.zeros_like(length)
    nz = length != 0
    hx[nz] = angle[nz] / length[nz]
    return hx


def _dipole_edge(hx, e, fint, gap):
    """lynx/accelerator/dipole.py:143-181 (entrance and exit share the form)."""
    sec_e = 1.0 / torch.cos(e)
    phi = fint * hx * gap * sec_e * (1 + torch.sin(e) ** 2)
    tm = _eye(phi)
    tm[..., 1, 0] = hx * torch.tan(e)
    tm[..., 3, 2] = -hx * torch.tan(e - phi)
    return tm


def dipole_map(length, angle, e1, e2, tilt, fint, fint_exit, gap, energy):
    """lynx/accelerator/dipole.py:112-141 (RBend: caller adds angle/2 to e1,e2)."""
    hx = _dipole_hx(length, angle)
    R_enter = _dipole_edge(hx, e1, fint, gap)
    R_exit = _dipole_edge(hx, e2, fint_exit, gap)
    if bool((length != 0.0).any()):
        R = base_rmatrix(
            length, torch.zeros_like(length), hx, torch.zeros_like(length), energy
        )
    else:  # :127-133 thin corrector; kick lands in R[2,6] as written (quirk B3)
        R = _eye(length)
        R[..., 0, 1] = length
        R[..., 2, 6] = angle
        R[..., 2, 3] = length
    R = R_exit @ (R @ R_enter)
    R = rotation_matrix(-tilt) @ (R @ rotation_matrix(tilt))
    return R


def cavity_rmatrix(length, voltage, phase_deg, frequency, energy) -> torch.Tensor:
    """Standing-wave cavity linear map, lynx/accelerator/cavity.py:248-325."""
    phi = torch.deg2rad(phase_deg)
    delta_energy = voltage * torch.cos(phi)
    eta = torch.tensor(1.0, dtype=length.dtype)
    Ei = energy / ELECTRON_MASS_EV
    Ef = (energy + delta_energy) / ELECTRON_MASS_EV
    Ep = (Ef - Ei) / length
    assert torch.all(Ei > 0), "Initial energy must be larger than 0"

    alpha = torch.sqrt(eta / 8) / torch.cos(phi) * torch.log(Ef / Ei)
    r11 = torch.cos(alpha) - torch.sqrt(2 / eta) * torch.cos(phi) * torch.sin(alpha)
    r12 = torch.sqrt(8 / eta) * Ei / Ep * torch.cos(phi) * torch.sin(alpha)
    r21 = (
        -Ep
        / Ef
        * (torch.cos(phi) / torch.sqrt(2 * eta) + torch.sqrt(eta / 8) / torch.cos(phi))
        * torch.sin(alpha)
    )
    r22 = (
        Ei
        / Ef
        * (torch.cos(alpha) + torch.sqrt(2 / eta) * torch.cos(phi) * torch.sin(alpha))
    )

    r56 = torch.zeros_like(length)
    beta0 = torch.ones_like(length)
    beta1 = torch.ones_like(length)
    k = 2 * math.pi * frequency / constants.speed_of_light
    r55_cor = torch.zeros_like(length)
    if torch.any((voltage != 0) & (energy != 0)):  # :290
        beta0 = torch.sqrt(1 - 1 / Ei**2)
        beta1 = torch.sqrt(1 - 1 / Ef**2)
        r56 = -length / (Ef**2 * Ei * beta1) * (Ef + Ei) / (beta1 + beta0)
        g0, g1 = Ei, Ef
        r55_cor = (
            k
            * length
            * beta0
            * voltage
            / ELECTRON_MASS_EV
            * torch.sin(phi)
            * (g0 * g1 * (beta0 * beta1 - 1) + 1)
            / (beta1 * g1 * (g0 - g1) ** 2)
        )
    r66 = Ei / Ef * beta0 / beta1
    r65 = k * torch.sin(phi) * voltage / (Ef * beta1 * ELECTRON_MASS_EV)

    R = _eye(length)
    R[..., 0, 0] = r11
    R[..., 0, 1] = r12
    R[..., 1, 0] = r21
    R[..., 1, 1] = r22
    R[..., 2, 2] = r11
    R[..., 2, 3] = r12
    R[..., 3, 2] = r21
    R[..., 3, 3] = r22
    R[..., 4, 4] = 1 + r55_cor
    R[..., 4, 5] = r56
    R[..., 5, 4] = r65
    R[..., 5, 5] = r66
    return R


def solenoid_map(length, k, misalignment, energy) -> torch.Tensor:
    """lynx/accelerator/solenoid.py:61-105 (scalar ``if gamma != 0`` made per-entry)."""
    gamma = energy / REST_ENERGY
    c = torch.cos(length * k)
    s = torch.sin(length * k)
    s_k = torch.empty_like(length)
    z = k == 0
    s_k[z] = length[z]
    s_k[~z] = s[~z] / k[~z]
    r56 = torch.zeros_like(length)
    nz = gamma != 0
    gamma2 = gamma[nz] * gamma[nz]
    beta = torch.sqrt(1.0 - 1.0 / gamma2)
    r56[nz] = r56[nz] - length[nz] / (beta * beta * gamma2)

    R = _eye(length)
    R[..., 0, 0] = c**2
    R[..., 0, 1] = c * s_k
    R[..., 0, 2] = s * c
    R[..., 0, 3] = s * s_k
    R[..., 1, 0] = -k * s * c
    R[..., 1, 1] = c**2
    R[..., 1, 2] = -k * s**2
    R[..., 1, 3] = s * c
    R[..., 2, 0] = -s * c
    R[..., 2, 1] = -s * s_k
    R[..., 2, 2] = c**2
    R[..., 2, 3] = c * s_k
    R[..., 3, 0] = k * s**2
    R[..., 3, 1] = -s * c
    R[..., 3, 2] = -k * s * c
    R[..., 3, 3] = c**2
    R[..., 4, 5] = r56
    if torch.all(misalignment == 0):
        return R
    R_entry, R_exit = misalignment_matrix(misalignment)
    return R_exit @ R @ R_entry


def undulator_map(length, energy) -> torch.Tensor:
    """lynx/accelerator/undulator.py:48-60 -- R56 = +L/gamma^2, no 1/beta^2 (quirk B4)."""
    gamma = energy / REST_ENERGY
    igamma2 = torch.zeros_like(gamma)
    nz = gamma != 0
    igamma2[nz] = 1 / gamma[nz] ** 2
    tm = _eye(energy)
    tm[..., 0, 1] = length
    tm[..., 2, 3] = length
    tm[..., 4, 5] = length * igamma2
    return tm

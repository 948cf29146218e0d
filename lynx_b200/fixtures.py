"""Lattice / beam fixtures of the BASELINE.json configs as plain data (SURVEY.md 8d,
Appendix C).  A lattice spec is a list of ``(kind, name, params)`` with Python floats;
``build_segment`` turns it into lynx_b200 elements (tests build the oracle's elements
from the same spec).  Synthetic beams are seeded NumPy PCG64, fp64 -> fp32.
"""

import numpy as np
import torch

# C1: tests/test_astra_import.py:12-23 (moments of the missing Astra file)
ASTRA_MU = (8.24126345833065e-07, 5.988477624896404e-08, -1.7276204289373709e-06,
            -1.1746412553748087e-07, 0.0, 0.0)
ASTRA_SIGMA = (1.7489789752289653e-4, 3.679402198031312e-06, 1.7519544053357095e-4,
               3.6941000871593133e-06, 8.011552381503861e-06, 2.2804534528404474e-3)
ASTRA_ENERGY = 107315902.44394557
ASTRA_CHARGE = 5.000000000010205e-13

ARES_SCREEN = {"resolution": (2448, 2040),
               "pixel_size": (3.5487998957250966e-06, 2.500300070096273e-06), "binning": 1}
# tests/test_screen.py:78-87 geometry (config 2 of SURVEY 8d)
TEST_SCREEN = {"resolution": (2448, 2040), "pixel_size": (3.3198e-6, 2.4469e-6), "binning": 1}


def ares_ea(k1=(0.0, 0.0, 0.0), vkick=0.0, hkick=0.0, screen_active=False, screen=None):
    """C6: AREASOLA1 .. AREABSCR1 of docs/examples/ARESlatticeStage3v1_9.json (E = 13)."""
    scr = dict(screen or ARES_SCREEN)
    scr["is_active"] = screen_active
    q = 0.12200000137090683
    return [
        ("marker", "AREASOLA1", {}),
        ("drift", "Drift_AREASOLA1", {"length": 0.17504000663757324}),
        ("quadrupole", "AREAMQZM1", {"length": q, "k1": k1[0]}),
        ("drift", "Drift_AREAMQZM1", {"length": 0.42800000309944153}),
        ("quadrupole", "AREAMQZM2", {"length": q, "k1": k1[1]}),
        ("drift", "Drift_AREAMQZM2", {"length": 0.20399999618530273}),
        ("vcor", "AREAMCVM1", {"length": 0.019999999552965164, "angle": vkick}),
        ("drift", "Drift_AREAMCVM1", {"length": 0.20399999618530273}),
        ("quadrupole", "AREAMQZM3", {"length": q, "k1": k1[2]}),
        ("drift", "Drift_AREAMQZM3", {"length": 0.17900000512599945}),
        ("hcor", "AREAMCHM1", {"length": 0.019999999552965164, "angle": hkick}),
        ("drift", "Drift_AREAMCHM1", {"length": 0.44999998807907104}),
        ("screen", "AREABSCR1", scr),
    ]


def smatch(bpms_active=False, angles=(0.0, 0.0, 0.0)):
    """C7: README.md:34-48 (E = 11)."""
    return [
        ("bpm", "BPM1SMATCH", {"is_active": bpms_active}),
        ("drift", "D1", {"length": 1.0}),
        ("bpm", "BPM6SMATCH", {"is_active": bpms_active}),
        ("drift", "D2", {"length": 1.0}),
        ("vcor", "V7SMATCH", {"length": 0.3, "angle": angles[0]}),
        ("drift", "D3", {"length": 0.2}),
        ("hcor", "H10SMATCH", {"length": 0.3, "angle": angles[1]}),
        ("drift", "D4", {"length": 7.0}),
        ("hcor", "H12SMATCH", {"length": 0.3, "angle": angles[2]}),
        ("drift", "D5", {"length": 0.05}),
        ("bpm", "BPM13SMATCH", {"is_active": bpms_active}),
    ]


def cavity_dipole_line(screen_active=True):
    """Config 4: Cavity + Dipole + quadrupoles, values of
    tests/test_compare_ocelot.py:679-685 and :77-82 (E = 10)."""
    scr = dict(ARES_SCREEN)
    scr["is_active"] = screen_active
    return [
        ("drift", "D0", {"length": 0.5}),
        ("quadrupole", "Q1", {"length": 0.2, "k1": 4.2}),
        ("drift", "D1", {"length": 0.3}),
        ("cavity", "C1", {"length": 1.0377, "voltage": 0.01815975e9, "phase": 0.0,
                          "frequency": 1.3e9}),
        ("drift", "D2", {"length": 0.3}),
        ("dipole", "B1", {"length": 0.1, "angle": 0.1, "fringe_integral": 0.1, "gap": 0.2}),
        ("drift", "D3", {"length": 0.3}),
        ("quadrupole", "Q2", {"length": 0.2, "k1": -4.2}),
        ("drift", "D4", {"length": 1.0}),
        ("screen", "SCR", scr),
    ]


def element_zoo():
    """One of every map-bearing element kind with every optional parameter set (tilt,
    misalignment, edge angles, fringe integrals): the lattice of the reference-generated vector
    tests/golden/reference_zoo.npz."""
    return [
        ("drift", "D1", {"length": 0.3}),
        ("quadrupole", "Q1", {"length": 0.122, "k1": 6.5, "tilt": 0.2, "misalignment": (2e-4, -1e-4)}),
        ("solenoid", "S1", {"length": 0.09, "k": 0.8}),
        ("hcor", "H1", {"length": 0.02, "angle": 1.2e-3}),
        ("vcor", "V1", {"length": 0.02, "angle": -0.7e-3}),
        ("dipole", "B1", {"length": 0.22, "angle": 0.05, "e1": 0.02, "e2": 0.03, "tilt": 0.1,
                          "fringe_integral": 0.5, "fringe_integral_exit": 0.4, "gap": 0.03}),
        ("drift", "D2", {"length": 0.25}),
        ("rbend", "B2", {"length": 0.3, "angle": -0.04, "fringe_integral": 0.45, "gap": 0.02}),
        ("undulator", "U1", {"length": 0.25}),
        ("quadrupole", "Q2", {"length": 0.122, "k1": -8.0}),
        ("cavity", "C1", {"length": 1.0377, "voltage": 8.0e6, "phase": -20.0, "frequency": 1.3e9}),
        ("drift", "D3", {"length": 0.4}),
        ("quadrupole", "Q3", {"length": 0.122, "k1": 3.0}),
    ]


def build_segment(spec, name=None):
    """lynx_b200 ``Segment`` from a spec."""
    from . import elements as E
    from .segment import Segment

    t = lambda v: torch.tensor([v], dtype=torch.float32)  # noqa: E731
    out = []
    for kind, nm, p in spec:
        if kind == "marker":
            el = E.Marker(name=nm)
        elif kind == "drift":
            el = E.Drift(t(p["length"]), name=nm)
        elif kind == "quadrupole":
            el = E.Quadrupole(
                t(p["length"]), t(p.get("k1", 0.0)),
                misalignment=torch.tensor([p["misalignment"]], dtype=torch.float32)
                if "misalignment" in p else None,
                tilt=t(p["tilt"]) if "tilt" in p else None, name=nm)
        elif kind in ("hcor", "vcor"):
            cls = E.HorizontalCorrector if kind == "hcor" else E.VerticalCorrector
            el = cls(t(p["length"]), t(p.get("angle", 0.0)), name=nm)
        elif kind in ("dipole", "rbend"):
            cls = E.Dipole if kind == "dipole" else E.RBend
            kw = {k: t(v) for k, v in p.items() if k != "length"}
            el = cls(t(p["length"]), name=nm, **kw)
        elif kind == "cavity":
            el = E.Cavity(t(p["length"]), t(p["voltage"]), t(p["phase"]), t(p["frequency"]),
                          name=nm)
        elif kind == "solenoid":
            el = E.Solenoid(t(p["length"]), t(p["k"]), name=nm)
        elif kind == "undulator":
            el = E.Undulator(t(p["length"]), name=nm)
        elif kind == "bpm":
            el = E.BPM(is_active=p.get("is_active", False), name=nm)
        elif kind == "screen":
            el = E.Screen(resolution=torch.tensor(p["resolution"], dtype=torch.float32),
                          pixel_size=torch.tensor(p["pixel_size"], dtype=torch.float32),
                          binning=torch.tensor(float(p.get("binning", 1))),
                          misalignment=torch.tensor([p.get("misalignment", (0.0, 0.0))],
                                                    dtype=torch.float32),
                          is_active=p.get("is_active", False), name=nm)
        else:
            raise ValueError(kind)
        out.append(el)
    return Segment(out, name=name)


def gaussian_beam_soa(n, mu=ASTRA_MU, sigma=ASTRA_SIGMA, seed=1002, cor=None) -> np.ndarray:
    """(6, n) fp32 planes: independent Gaussians with the C1 moments (config 2/3/5).
    ``cor`` = optional (cov_xxp, cov_yyp) for twiss-shaped beams (config 4)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    z = rng.standard_normal((6, n))
    out = np.empty((6, n), dtype=np.float64)
    for i in range(6):
        out[i] = mu[i] + sigma[i] * z[i]
    if cor is not None:
        for (i, j), c in zip(((0, 1), (2, 3)), cor):
            r = c / (sigma[i] * sigma[j])
            out[j] = mu[j] + sigma[j] * (r * z[i] + np.sqrt(max(1 - r * r, 0.0)) * z[j])
    return out.astype(np.float32)


def twiss_sigmas(beta, alpha, emit):
    """particle_beam.py:243-248."""
    sx = np.sqrt(beta * emit)
    sxp = np.sqrt(emit * (1 + alpha**2) / beta)
    return sx, sxp, -emit * alpha


def cfg4_beam_soa(n, seed=1004) -> np.ndarray:
    """tests/test_compare_ocelot.py:657-666 twiss, sigma_s = sigma_p = 1e-6, E = 6 MeV."""
    sx, sxp, cx = twiss_sigmas(5.91253677, 3.55631308, 3.494768647122823e-09)
    sy, syp, cy = twiss_sigmas(5.91253677, 3.55631308, 3.497810737006068e-09)
    return gaussian_beam_soa(n, (0.0,) * 6, (sx, sxp, sy, syp, 1e-6, 1e-6), seed, cor=(cx, cy))


CFG4_ENERGY = 6e6


def soa_to_aos7(soa: np.ndarray) -> np.ndarray:
    """(6, n) -> (n, 7) with the constant-1 column, the reference's layout."""
    n = soa.shape[1]
    a = np.ones((n, 7), dtype=soa.dtype)
    a[:, :6] = soa.T
    return a


def random_settings(B, seed=1003, k1_range=30.0, steer=6.1782e-3):
    """Config 3/5 settings: (B,3) k1 ~ U(-r, r), (B,2) corrector angles ~ U(-s, s)
    (tests/test_vectorized.py:306, gradientbased.ipynb STEERER_LIMIT)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    k1 = rng.uniform(-k1_range, k1_range, (B, 3)).astype(np.float32)
    ang = rng.uniform(-steer, steer, (B, 2)).astype(np.float32)
    return k1, ang

"""Shared test helpers: build oracle elements from a fixtures spec, tolerances."""

import numpy as np
import torch

from oracle.lattice import El

_DEFAULTS = {
    "quadrupole": {"k1": 0.0, "tilt": 0.0},
    "hcor": {"angle": 0.0},
    "vcor": {"angle": 0.0},
    "dipole": {"angle": 0.0, "e1": 0.0, "e2": 0.0, "tilt": 0.0, "fringe_integral": 0.0,
               "gap": 0.0},
    "cavity": {"voltage": 0.0, "phase": 0.0, "frequency": 0.0},
    "solenoid": {"k": 0.0},
}
_DEFAULTS["rbend"] = _DEFAULTS["dipole"]


def oracle_elements(spec, dtype=torch.float32, batch=(1,), overrides=None):
    """spec -> list[El]; every parameter expanded to ``batch``.  ``overrides`` maps
    (name, param) -> tensor of shape ``batch`` (per-setting values)."""
    overrides = overrides or {}
    els = []
    for kind, name, p in spec:
        q = {}
        full = dict(_DEFAULTS.get(kind, {}))
        full.update({k: v for k, v in p.items() if k not in ("is_active",)})
        if kind in ("dipole", "rbend") and "fringe_integral_exit" not in full:
            full["fringe_integral_exit"] = full["fringe_integral"]
        for k, v in full.items():
            if (name, k) in overrides:
                q[k] = torch.as_tensor(overrides[(name, k)], dtype=dtype).reshape(batch).clone()
            elif kind == "screen" and k in ("resolution", "pixel_size", "binning"):
                q[k] = v
            elif k == "misalignment":
                q[k] = torch.tensor(v, dtype=dtype).expand(*batch, 2).clone()
            else:
                q[k] = torch.full(batch, float(v), dtype=dtype)
        if kind in ("quadrupole", "solenoid", "screen") and "misalignment" not in q:
            q["misalignment"] = torch.zeros(*batch, 2, dtype=dtype)
        els.append(El(kind, name, q, is_active=bool(p.get("is_active", False))))
    return els


def aos_from_soa(soa: np.ndarray, dtype=torch.float32) -> torch.Tensor:
    """(6,N) numpy planes -> (1,N,7) torch AoS with the constant-1 column."""
    n = soa.shape[1]
    a = torch.ones(1, n, 7, dtype=dtype)
    a[0, :, :6] = torch.from_numpy(soa.astype(np.float64)).T.to(dtype)
    return a


def assert_close_scaled(got, truth, ref32=None, rtol=1e-5, what=""):
    """SURVEY 7.3-3 parity rule, per coordinate plane:
    |got - truth| <= rtol * scale + |ref32 - truth|, scale = rms(truth) of that plane.
    ``truth`` is the fp64 oracle, ``ref32`` the fp32 oracle (its own distance from truth
    is the reference's fp32 noise and is allowed on top)."""
    got = np.asarray(got, dtype=np.float64)
    truth = np.asarray(truth, dtype=np.float64)
    scale = np.sqrt(np.mean(truth**2, axis=-1, keepdims=True)) + 1e-300
    tol = rtol * scale
    if ref32 is not None:
        tol = tol + np.abs(np.asarray(ref32, dtype=np.float64) - truth)
    err = np.abs(got - truth)
    bad = err > tol
    assert not bad.any(), (
        f"{what}: {bad.sum()} of {bad.size} outside tolerance; worst err/scale = "
        f"{(err / scale).max():.3e}"
    )

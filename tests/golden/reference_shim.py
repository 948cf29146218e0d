"""Run the reference (jank324/lynx, /root/reference) as it is written, on torch.

The reference is a mechanical rename of a PyTorch code base to ``jax.numpy`` names that never
ran under JAX (tensor methods, in-place writes, torch keyword arguments are all still there).
It does run when ``jax.numpy`` IS torch.  ``install()`` provides exactly that and nothing of
this repository: no source file of the reference is changed, no function of it is replaced.

What the shim supplies (each item is an import-time or name-resolution matter, not arithmetic):
  1. ``jax`` / ``jax.numpy``: every ``jnp.X`` resolves to ``torch.X`` (``array`` -> ``tensor``,
     ``asarray`` -> ``as_tensor``, ``concatenate`` -> ``cat``, ``degrees`` / ``radians`` ->
     ``rad2deg`` / ``deg2rad``; numpy for the rest).  As in jax.numpy, functions accept Python
     scalars where torch insists on tensors (``jnp.sqrt(2 / eta)``, ``jnp.complex(x, 0.0)``).
  2. ``equinox.Module`` -> ``torch.nn.Module`` (the base class the code was written against:
     ``nn.ModuleList`` / ``nn.Module`` are still used inside) and the missing ``nn`` name in
     ``lynx/accelerator/segment.py`` (``nn.ModuleList`` at :43 has no import).
  3. Module-level constants ``rest_energy`` / ``REST_ENERGY`` / ``electron_mass_eV`` are plain
     floats after the rename but are used as tensors (``rest_energy.to(device=...)``,
     drift.py:52): wrapped as float32 tensors, the value unchanged (not in lynx.converters,
     which is numpy code and uses them as floats).
  4. ``matplotlib`` names used in annotations only (plotting is not exercised).

Only tests/golden/make_reference_golden.py uses this, in the container that holds
/root/reference; nothing at GPU run time does.
"""

import sys
import types

import numpy as np
import torch

REFERENCE_ROOT = "/root/reference"
_RENAMED = {"array": "tensor", "asarray": "as_tensor", "concatenate": "cat", "power": "pow",
            "degrees": "rad2deg", "radians": "deg2rad"}


def _complex(re, im):
    re = torch.as_tensor(re)
    return torch.complex(re, torch.zeros_like(re) + im)


def _accept_scalars(fn):
    def call(*args, **kwargs):
        try:
            return fn(*args, **kwargs)
        except TypeError:
            args = [torch.tensor(float(a)) if isinstance(a, (int, float)) and not isinstance(a, bool)
                    else a for a in args]
            return fn(*args, **kwargs)

    return call


class _TorchAsJnp(types.ModuleType):
    def __getattr__(self, name):
        if name == "complex":
            return _complex
        if name == "ndarray":
            return torch.Tensor
        target = _RENAMED.get(name, name)
        if hasattr(torch, target):
            obj = getattr(torch, target)
            plain = isinstance(obj, (type, torch.dtype)) or not callable(obj)
            return obj if plain else _accept_scalars(obj)
        if hasattr(np, name):
            return getattr(np, name)
        raise AttributeError(name)


def install():
    """Import the reference package under the shim and return it."""
    jax = types.ModuleType("jax")
    jnp = _TorchAsJnp("jax.numpy")
    jax.numpy, jax.Array = jnp, torch.Tensor
    sys.modules["jax"], sys.modules["jax.numpy"] = jax, jnp
    eqx = types.ModuleType("equinox")
    eqx.Module = torch.nn.Module
    sys.modules["equinox"] = eqx
    if "matplotlib" not in sys.modules:
        mpl = types.ModuleType("matplotlib")
        for sub, names in (("pyplot", ("Axes",)), ("patches", ("Rectangle",)), ("figure", ("Figure",))):
            mod = types.ModuleType(f"matplotlib.{sub}")
            for n in names:
                setattr(mod, n, object)
            setattr(mpl, sub, mod)
            sys.modules[f"matplotlib.{sub}"] = mod
        sys.modules["matplotlib"] = mpl
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import lynx
    import lynx.accelerator.segment as segment_module

    segment_module.nn = torch.nn
    for name, mod in list(sys.modules.items()):
        if name.startswith("lynx.converters"):
            continue  # numpy code: the constants are used as floats there
        if name == "lynx" or name.startswith("lynx."):
            for const in ("rest_energy", "REST_ENERGY", "electron_mass_eV"):
                if isinstance(getattr(mod, const, None), float):
                    setattr(mod, const, torch.tensor(getattr(mod, const)))
    return lynx

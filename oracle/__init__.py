"""CPU restatement oracle for the Lynx beam-tracking hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``lynx_b200/`` may import this package;
only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs do.

The reference (jank324/lynx) is a ``torch``->``jnp`` rename of Cheetah that does not run
under JAX; this oracle restates the arithmetic its files spell out,
in torch-CPU (the semantic original of the renamed calls), with the dtype switchable between
float32 and float64.  Every function cites the reference file:line it follows.

Parity pinning -- against outputs of the reference itself.  The reference DOES run when
``jax.numpy`` is torch (``tests/golden/reference_shim.py``); run that way in the build
container it generated ``tests/golden/reference_*.npz`` (``make_reference_golden.py``), and
``tests/test_reference_golden.py`` requires this oracle, in float32, to reproduce them bit for
bit: every element's transfer map, merged maps, tracked particles and their moments, the
ParameterBeam's mu / cov, every screen bin (histogram and pdf), the cavity's non-linear pass,
BPM readings, batched settings, transformed_to, the aperture cut, and -- through torch
autograd -- gradients.  ``tests/test_oracle_golden.py`` adds the known answers of the
reference's own tests / notebooks (C2-C5 and the relational checks).
"""

from .maps import (  # noqa: F401
    ELECTRON_MASS_EV,
    REST_ENERGY,
    base_rmatrix,
    cavity_rmatrix,
    corrector_map,
    dipole_map,
    drift_map,
    misalignment_matrix,
    quadrupole_map,
    rotation_matrix,
    solenoid_map,
    undulator_map,
)
from .lattice import (  # noqa: F401
    El,
    ParameterBeamO,
    ParticleBeamO,
    element_map,
    is_skippable,
    screen_bin_edges,
    screen_reading_parameter_beam,
    screen_reading_particles,
    segment_map,
    track,
)

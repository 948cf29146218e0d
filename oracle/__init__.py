"""CPU restatement oracle for the Lynx beam-tracking hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``lynx_b200/`` may import this package;
only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs do.

The reference (jank324/lynx, a non-executable ``torch``->``jnp`` rename of Cheetah)
cannot be imported or run, so this oracle restates the
arithmetic its files spell out, in torch-CPU (the semantic original of the renamed
calls), with the dtype switchable between float32 and float64.  Every function cites
the reference file:line it follows.

Parity pinning: the oracle is checked in ``tests/test_oracle_golden.py`` against every
known-answer constant the reference's own tests/notebooks hold for this path
(C2 Bmad twiss after the cavity, C3 corrector map, C4 drift
R56, C5 merged 1051-element map, plus the relational checks of the reference's tests).
Per-particle outputs, screen bins and gradient values are NOT pinned by anything in the
reference (they need Ocelot / a missing Astra file): for those, parity is "vs. this
restatement", labelled as such.
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

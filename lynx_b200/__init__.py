"""lynx_b200 -- B200-native beam tracking behind the Lynx / Cheetah Python surface.

The names exported here are the reference's (lynx/__init__.py, lynx/accelerator/
__init__.py, lynx/particles/__init__.py); the arithmetic runs in hand-written sm_100a
CUDA kernels behind the C ABI of ``include/lynx_b200.h``.  No CPU fallback.
"""

from .beams import Beam, ParameterBeam, ParticleBeam  # noqa: F401
from .elements import (  # noqa: F401
    BPM,
    Aperture,
    Cavity,
    CustomTransferMap,
    Dipole,
    Drift,
    Element,
    HorizontalCorrector,
    Marker,
    Quadrupole,
    RBend,
    Screen,
    Solenoid,
    Undulator,
    VerticalCorrector,
)
from .segment import Segment  # noqa: F401

__version__ = "0.1.0"

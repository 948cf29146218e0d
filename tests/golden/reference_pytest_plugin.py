"""pytest plugin: run the REFERENCE'S OWN test-suite on torch (tests/golden/reference_shim.py),
with the Ocelot stand-in classes of tests/test_ocelot.py for its ARES resource file.

    cd /root/reference && PYTHONDONTWRITEBYTECODE=1 PYTHONPATH=/root/repo/tests/golden \\
        python -m pytest tests -p no:cacheprovider -p reference_pytest_plugin -q \\
        --ignore=tests/test_compare_ocelot.py --ignore=tests/test_speed.py

(read-only: no cache, no bytecode).  Result in this container: 55 passed, 7 skipped, 4 xfailed,
33 failed -- the failures are the ASTRA sample file missing from the checkout
(``FileNotFoundError``), ``torch`` used without import in ``latticejson.py:138``, and
``Segment.length`` stacking lengths of shape () and (1,) (``segment.py:323``).  Nothing in the
repository's own tests depends on this.
"""

import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
for p in (HERE, os.path.dirname(HERE), os.path.dirname(os.path.dirname(HERE))):
    if p not in sys.path:
        sys.path.insert(0, p)

import reference_shim  # noqa: E402

reference_shim.install()
import test_ocelot as _stand_ins  # noqa: E402

_ocelot = types.ModuleType("ocelot")
_ocelot.__dict__.update(_stand_ins.STAND_INS)
sys.modules["ocelot"] = _ocelot

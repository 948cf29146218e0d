"""CPU: the C-ABI library builds, loads and exports every symbol include/lynx_b200.h
declares (no compute calls: there is no GPU here and no CPU fallback)."""

import ctypes
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


@pytest.fixture(scope="module")
def lib():
    from lynx_b200 import build

    build.build()
    from lynx_b200 import _lib

    return _lib.lib()


def _declared_symbols():
    text = (ROOT / "include" / "lynx_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(lynx_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported(lib):
    names = _declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/lynx_b200.h but not exported"


def test_binding_table_matches_header(lib):
    from lynx_b200 import _lib

    assert sorted(_lib.SIGNATURES) == _declared_symbols()


def test_version_and_status_strings(lib):
    assert lib.lynx_version() >= 100
    assert lib.lynx_status_str(0) == b"ok"
    assert b"no CPU fallback" in lib.lynx_status_str(-5)


def test_fails_loudly_without_a_device(lib):
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from lynx_b200 import _lib as L

    ops = (L.OpT * 1)(L.OpT(L.OP_MARKER, -1, 0, 0))
    h = ctypes.c_void_p()
    rc = lib.lynx_program_create(ops, 1, None, None, 0, ctypes.byref(h))
    assert rc == L.E_NODEVICE
    with pytest.raises(L.LynxError):
        L.check(rc)

"""The XLA FFI shim (lynx_b200/csrc/lynx_xla_ffi.cc) and its JAX side (lynx_b200/jax_binding.py)
cannot be compiled or run in this image (no jax / jaxlib / xla headers).  What CAN
be checked on the CPU: the handler table covers every stream launcher of the C ABI, every handler
adapts a declared entry point with a stream argument, the JAX module registers exactly the
handlers the shim exports, and the module imports (and refuses to run) without jax."""

import re
from pathlib import Path

import pytest

from lynx_b200 import _lib as L
from lynx_b200 import jax_binding as JB

ROOT = Path(__file__).resolve().parents[1]
SRC = (ROOT / "lynx_b200" / "csrc" / "lynx_xla_ffi.cc").read_text()
TABLE = re.findall(r"^LYNX_FFI_HANDLER\((\w+),\s*(\w+)\)", SRC, flags=re.M)


def test_every_stream_launcher_has_a_handler():
    adapted = {c_abi for _, c_abi in TABLE}
    assert adapted == set(L.STREAM_LAUNCHERS)
    for name in L.STREAM_LAUNCHERS:
        res, args = L.SIGNATURES[name]
        assert res is not None and args and args[-1] is L._vp, name  # status return, stream last


def test_stream_launchers_are_all_of_them():
    """Nothing that takes a stream is missing from STREAM_LAUNCHERS (the header marks the stream
    argument with the lynx_stream_t type)."""
    header = (ROOT / "include" / "lynx_b200.h").read_text()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    decls = re.findall(r"^(?:int|void)\s+(lynx_\w+)\(([^;]*?)\);", header, flags=re.M | re.S)
    with_stream = {n for n, a in decls if "lynx_stream_t" in a}
    assert with_stream - {"lynx_debug_stamp"} == set(L.STREAM_LAUNCHERS)


def test_handlers_are_defined_and_call_what_the_table_says():
    for symbol, c_abi in TABLE:
        m = re.search(r"XLA_FFI_DEFINE_HANDLER_SYMBOL\(%s,\s*(\w+)," % symbol, SRC)
        assert m, symbol
        body = re.search(r"static ffi::Error %s\(.*?\n}\n" % m.group(1), SRC, flags=re.S)
        assert body and (c_abi + "(") in body.group(0), (symbol, c_abi)
        assert "Ctx<Stream>()" in SRC[m.start():m.start() + 400]


def test_jax_side_registers_exactly_the_exported_handlers():
    assert set(JB.TARGETS.values()) == {symbol for symbol, _ in TABLE}
    assert len(set(JB.TARGETS)) == len(JB.TARGETS)


def test_module_refuses_to_run_without_jax():
    import importlib.util

    if importlib.util.find_spec("jax") is not None:
        pytest.skip("jax is installed here")
    assert JB.available() is False
    with pytest.raises(L.LynxError):
        JB.register()
    with pytest.raises(L.LynxError):
        JB.make_track(None, B=1, N=1, ld=4, shared_beam=True, want_out=True, moment_mode=0, n_bpm=0)

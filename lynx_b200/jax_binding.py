"""JAX binding of the C ABI: XLA FFI custom calls + ``jax.custom_vjp`` (the north star's host
side).  It replaces the body of ``Segment.track`` (lynx/accelerator/segment.py:340-356) under
``jit`` / ``vmap`` / ``grad``: one custom call per track, the vjp through
``lynx_track_particles_bwd`` / ``lynx_parambeam_bwd``.

``jax`` / ``jaxlib`` are NOT in this image, so nothing here can run in the build
container or on the GPU box; the module imports without jax (``available()`` is False and every
entry point raises ``LynxError``), and ``tests/test_xla_ffi_table.py`` checks on the CPU that the
targets below are exactly the handlers ``csrc/lynx_xla_ffi.cc`` defines and that those cover every
stream launcher of ``_lib.py``.  The binding that is exercised end to end is ``ctypes`` + torch
(``segment.py`` / ``autograd.py``); this file is the same wiring in JAX's vocabulary.

The lattice structure (which elements are active / skippable, column strides) is decided from
Python values at trace time by ``program.encode`` -- as the reference's ``is_skippable``
properties are -- so a ``Program`` handle is a static attribute of the call; the parameter table,
the particles and the energies are the traced operands.
"""

import ctypes as C
import importlib.util
from pathlib import Path

import numpy as np

from . import _lib as L

XLA_LIB_PATH = Path(L.LIB_PATH).with_name("liblynx_b200_xla.so")

# ffi target name -> handler symbol exported by liblynx_b200_xla.so (csrc/lynx_xla_ffi.cc)
TARGETS = {
    "lynx_compose_maps": "LynxComposeMaps",
    "lynx_track_fwd": "LynxTrackFwd",
    "lynx_track_fwd_noalloc": "LynxTrackFwdNoAlloc",
    "lynx_track_fwd_bbox": "LynxTrackFwdBbox",
    "lynx_track_fwd_persistent": "LynxTrackFwdPersistent",
    "lynx_track_bwd": "LynxTrackBwd",
    "lynx_parambeam_fwd": "LynxParamBeamFwd",
    "lynx_parambeam_bwd": "LynxParamBeamBwd",
    "lynx_parambeam_along": "LynxParamBeamAlong",
    "lynx_screen_pdf": "LynxScreenPdf",
    "lynx_beam_generate": "LynxBeamGenerate",
    "lynx_beam_generate_ellipsoid": "LynxBeamGenerateEllipsoid",
    "lynx_beam_transform": "LynxBeamTransform",
    "lynx_aperture_compact": "LynxApertureCompact",
    "lynx_aos7_to_soa6": "LynxAos7ToSoa6",
    "lynx_soa6_to_aos7": "LynxSoa6ToAos7",
    "lynx_screen_allreduce": "LynxScreenAllreduce",
}

_registered = False


def available() -> bool:
    return importlib.util.find_spec("jax") is not None and XLA_LIB_PATH.exists()


def _jax():
    if importlib.util.find_spec("jax") is None:
        raise L.LynxError("jax is not installed: the XLA FFI binding cannot run (use lynx_b200.Segment, "
                          "the ctypes + torch binding of the same C ABI)")
    import jax

    return jax


def register() -> None:
    """``jax.ffi.register_ffi_target`` for every handler (platform CUDA).  Idempotent."""
    global _registered
    if _registered:
        return
    jax = _jax()
    if not XLA_LIB_PATH.exists():
        raise L.LynxError(f"{XLA_LIB_PATH} is missing: build csrc/lynx_xla_ffi.cc against jaxlib's "
                          "headers (the command is in the file's header)")
    L.lib()  # liblynx_b200.so first: the shim links against it
    shim = C.CDLL(str(XLA_LIB_PATH))
    for target, symbol in TARGETS.items():
        jax.ffi.register_ffi_target(target, jax.ffi.pycapsule(getattr(shim, symbol)), platform="CUDA")
    _registered = True


def make_track(prog, *, B, N, ld, shared_beam, want_out, moment_mode, n_bpm, screen_wh=None):
    """``track(params, in_soa, energy_in, edges_x, edges_y) -> (out_soa, bpm_sums, img, moment_sums,
    energy_out)`` for one compiled lattice, differentiable w.r.t. ``params`` and ``in_soa``.

    Absent outputs are zero-element arrays (the handlers read them as NULL).  ``vmap`` over
    settings folds the mapped axis into ``B`` (``vmap_method="expand_dims"``): every parameter
    column already carries a stride."""
    jax = _jax()
    jnp = jax.numpy
    register()
    W, H = screen_wh if screen_wh else (0, 0)
    rows = 1 if shared_beam else B
    f32, f64 = jnp.float32, jnp.float64
    out_types = (
        jax.ShapeDtypeStruct((B, 6, ld) if want_out else (0,), f32),
        jax.ShapeDtypeStruct((n_bpm, B, 2) if n_bpm else (0,), f64),
        jax.ShapeDtypeStruct((B, H, W) if screen_wh else (0,), f32),
        jax.ShapeDtypeStruct((B, 28) if moment_mode else (0,), f64),
        jax.ShapeDtypeStruct((B,), f32),
        jax.ShapeDtypeStruct((prog.workspace_bytes(B),), jnp.uint8),
    )
    attrs = dict(prog=int(prog.handle), B=B, N=N, ld=ld, in_row_stride=0 if shared_beam else 1)

    def call_fwd(params, in_soa, energy_in, edges_x, edges_y):
        res = jax.ffi.ffi_call("lynx_track_fwd", out_types, vmap_method="expand_dims")(
            params, in_soa, energy_in, edges_x, edges_y, moment_mode=np.int32(moment_mode), **attrs)
        return res[:5]

    @jax.custom_vjp
    def track(params, in_soa, energy_in, edges_x, edges_y):
        return call_fwd(params, in_soa, energy_in, edges_x, edges_y)

    def fwd(params, in_soa, energy_in, edges_x, edges_y):
        return call_fwd(params, in_soa, energy_in, edges_x, edges_y), (params, in_soa, energy_in)

    def bwd(res, g):
        params, in_soa, energy_in = res
        g_out, g_bpm, _g_img, g_mom, _g_energy = g  # the image is integer-valued: no cotangent
        bwd_types = (
            jax.ShapeDtypeStruct(params.shape, f32),
            jax.ShapeDtypeStruct((rows, 6, ld), f32),
            jax.ShapeDtypeStruct((int(L.lib().lynx_bwd_workspace_bytes(prog.handle, B)),), jnp.uint8),
        )
        g_params, g_in, _ = jax.ffi.ffi_call("lynx_track_bwd", bwd_types)(
            params, in_soa, energy_in, g_out, g_mom, g_bpm, **attrs)
        return g_params, g_in, None, None, None

    track.defvjp(fwd, bwd)
    return track


def make_parambeam(prog, *, B, shared_beam, n_bpm):
    """``step(params, mu, cov, energy_in) -> (mu_out, cov_out, bpm, energy_out)``: the RL env step
    (Segment.track(ParameterBeam), lynx/accelerator/element.py:71-82), differentiable."""
    jax = _jax()
    jnp = jax.numpy
    register()
    f32 = jnp.float32
    rows = 1 if shared_beam else B
    out_types = (
        jax.ShapeDtypeStruct((B, 7), f32),
        jax.ShapeDtypeStruct((B, 7, 7), f32),
        jax.ShapeDtypeStruct((n_bpm, B, 2) if n_bpm else (0,), f32),
        jax.ShapeDtypeStruct((B,), f32),
    )
    attrs = dict(prog=int(prog.handle), B=B, beam_stride=0 if shared_beam else 1)

    def call_fwd(params, mu, cov, energy_in):
        return jax.ffi.ffi_call("lynx_parambeam_fwd", out_types, vmap_method="expand_dims")(
            params, mu, cov, energy_in, **attrs)

    @jax.custom_vjp
    def step(params, mu, cov, energy_in):
        return call_fwd(params, mu, cov, energy_in)

    def fwd(params, mu, cov, energy_in):
        return call_fwd(params, mu, cov, energy_in), (params, mu, cov, energy_in)

    def bwd(res, g):
        params, mu, cov, energy_in = res
        g_mu, g_cov, g_bpm, _ = g
        bwd_types = (
            jax.ShapeDtypeStruct(params.shape, f32),
            jax.ShapeDtypeStruct((rows, 7), f32),
            jax.ShapeDtypeStruct((rows, 7, 7), f32),
        )
        g_params, g_mu_in, g_cov_in = jax.ffi.ffi_call("lynx_parambeam_bwd", bwd_types)(
            params, mu, cov, energy_in, g_mu, g_cov, g_bpm, **attrs)
        return g_params, g_mu_in, g_cov_in, None

    step.defvjp(fwd, bwd)
    return step

"""ctypes binding of ``lib/liblynx_b200.so`` (the C ABI in ``include/lynx_b200.h``).

There is no CPU fallback: if the shared library is missing, or a launcher is called
without a CUDA device, this module raises.
"""

import ctypes as C
import os
from pathlib import Path

_HERE = Path(__file__).resolve().parent
LIB_PATH = _HERE / "lib" / "liblynx_b200.so"
if os.environ.get("LYNX_B200_LIB"):  # a tuning build (python -m lynx_b200.build --experiments)
    LIB_PATH = Path(os.environ["LYNX_B200_LIB"])

OK = 0
E_BADARG, E_BADPROGRAM, E_LAUNCH, E_NOMEM, E_NODEVICE, E_TOOLARGE = -1, -2, -3, -4, -5, -6
NOALLOC_WORKSPACE_BYTES = 1 << 20  # LYNX_NOALLOC_WORKSPACE_BYTES

# opcodes / flags (mirror include/lynx_b200.h)
OP_MARKER, OP_DRIFT, OP_QUAD, OP_HCOR, OP_VCOR, OP_DIPOLE = 0, 1, 2, 3, 4, 5
OP_CAVITY, OP_SOLENOID, OP_UNDULATOR, OP_CUSTOM, OP_BPM, OP_SCREEN = 6, 7, 8, 9, 10, 11
F_ACTIVE, F_RBEND, F_SCREEN_Y_FIX = 1, 2, 4


class LynxError(RuntimeError):
    pass


class OpT(C.Structure):
    _fields_ = [("opcode", C.c_int32), ("col", C.c_int32), ("flags", C.c_int32), ("aux", C.c_int32)]


_vp, _i64, _i32, _sz = C.c_void_p, C.c_int64, C.c_int32, C.c_size_t

# name -> (restype, argtypes); every symbol include/lynx_b200.h declares
SIGNATURES = {
    "lynx_version": (C.c_int, []),
    "lynx_status_str": (C.c_char_p, [C.c_int]),
    "lynx_program_create": (C.c_int, [C.POINTER(OpT), _i32, C.POINTER(_i64), C.POINTER(_i32), _i32, C.POINTER(_vp)]),
    "lynx_program_destroy": (None, [_vp]),
    "lynx_program_num_stages": (_i32, [_vp]),
    "lynx_program_num_bpm": (_i32, [_vp]),
    "lynx_program_num_screens": (_i32, [_vp]),
    "lynx_program_is_terminal": (_i32, [_vp]),
    "lynx_program_record_floats": (_i64, [_vp]),
    "lynx_workspace_bytes": (_sz, [_vp, _i64]),
    "lynx_compose_maps": (C.c_int, [_vp, _vp, _i64, _vp, _vp, _i64, _i64, _vp, _vp, _vp, _vp]),
    "lynx_track_particles_fwd": (
        C.c_int,
        [_vp, _vp, _i64, _vp, _i64, _i64, _i64, _vp, _i64, _vp, _vp, _i32, _i32, _vp, _vp, _vp,
         _i32, _vp, _vp, _vp, _vp],
    ),
    "lynx_track_particles_fwd_noalloc": (
        C.c_int,
        [_vp, _vp, _i64, _vp, _i64, _i64, _i64, _vp, _i64, _vp, _vp, _i32, _i32, _vp, _vp, _vp,
         _i32, _vp, _vp, _vp],
    ),
    "lynx_track_particles_fwd_bbox": (
        C.c_int,
        [_vp, _vp, _i64, _vp, _i64, _i64, _i64, _vp, _i64, _vp, _vp, _i32, _i32, _vp, _vp, _vp,
         _i32, _vp, _vp, _vp, _vp, _vp],
    ),
    "lynx_track_particles_fwd_persistent": (
        C.c_int,
        [_vp, _vp, _i64, _vp, _i64, _i64, _i64, _vp, _i64, _vp, _vp, _i32, _i32, _vp, _vp, _vp,
         _i32, _vp, _vp, _vp, _vp, C.c_uint32, _vp],
    ),
    "lynx_screen_symm_bytes": (_sz, [_i64, _i32, _i32]),
    "lynx_screen_symm_image": (_vp, [_vp, _i32, _i64, _i32, _i32]),
    "lynx_screen_allreduce": (
        C.c_int, [C.POINTER(_vp), _i32, _i32, _i32, C.c_uint32, _i64, _i32, _i32, _vp, _vp, _vp]
    ),
    "lynx_screen_allreduce_mc": (
        C.c_int, [C.POINTER(_vp), _vp, _i32, _i32, _i32, C.c_uint32, _i64, _i32, _i32, _vp, _vp, _vp]
    ),
    "lynx_track_particles_bwd": (
        C.c_int,
        [_vp, _vp, _i64, _vp, _i64, _i64, _i64, _vp, _i64, _vp, _vp, _vp, _vp, _vp, _i64, _vp, _vp],
    ),
    "lynx_bwd_workspace_bytes": (_sz, [_vp, _i64]),
    "lynx_parambeam_fwd": (C.c_int, [_vp, _vp, _i64, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _vp, _vp]),
    "lynx_parambeam_bwd": (
        C.c_int, [_vp, _vp, _i64, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]
    ),
    "lynx_screen_pdf": (C.c_int, [_vp, _vp, _i64, _vp, _vp, _i32, _i32, _vp, _vp]),
    "lynx_beam_generate": (C.c_int, [_vp, _vp, _i64, _i64, _i64, C.c_uint64, _vp, _i64, _vp]),
    "lynx_philox4x32_10": (None, [_vp, _vp, _vp]),
    "lynx_beam_generate_ellipsoid": (C.c_int, [_vp, _vp, _i64, _i64, _i64, C.c_uint64, _vp, _i64, _vp]),
    "lynx_beam_transform": (C.c_int, [_vp, _i64, _i64, _i64, _i64, _vp, _vp, _i64, _vp]),
    "lynx_aperture_workspace_bytes": (_sz, [_i64, _i64]),
    "lynx_aperture_compact": (
        C.c_int,
        [_vp, _i64, _i64, _i64, _vp, _i64, _vp, _vp, _i64, _i32, _vp, _vp, _i64, _vp, _vp, _vp,
         _vp, _vp],
    ),
    "lynx_parambeam_along": (C.c_int, [_vp, _vp, _i64, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _vp]),
    "lynx_aos7_to_soa6": (C.c_int, [_vp, _i64, _i64, _vp, _i64, _vp]),
    "lynx_soa6_to_aos7": (C.c_int, [_vp, _i64, _i64, _i64, _vp, _vp]),
    "lynx_launch_count": (_i64, []),
    "lynx_launch_count_reset": (None, []),
    "lynx_profile_enable": (None, [C.c_int]),
    "lynx_profile_last_ms": (C.c_float, [C.c_int]),
    "lynx_exchange_trace": (None, [_vp]),
    "lynx_debug_stamp": (None, [_vp, _vp]),
    "lynx_host_ctx_create": (C.c_int, [C.c_int, C.POINTER(_vp)]),
    "lynx_host_ctx_destroy": (None, [_vp]),
    "lynx_host_ctx_set_option": (C.c_int, [_vp, _i32, _i32]),
    "lynx_host_track_particles": (
        C.c_int,
        [_vp, _vp, _vp, _i64, _i64, _vp, _i64, _i64, _vp, _vp, _vp, _vp, _i32, _i32, _vp, _vp, _vp,
         _i32, _vp],
    ),
    "lynx_host_parambeam": (
        C.c_int, [_vp, _vp, _vp, _i64, _i64, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _vp]
    ),
}

# Entry points that ENQUEUE device work on a caller-supplied stream: the set an XLA FFI shim has
# to adapt (csrc/lynx_xla_ffi.cc; tests/test_xla_ffi_table.py keeps the two in step).
STREAM_LAUNCHERS = (
    "lynx_compose_maps", "lynx_track_particles_fwd", "lynx_track_particles_fwd_noalloc",
    "lynx_track_particles_fwd_bbox", "lynx_track_particles_fwd_persistent",
    "lynx_track_particles_bwd", "lynx_parambeam_fwd", "lynx_parambeam_bwd", "lynx_parambeam_along",
    "lynx_screen_pdf", "lynx_beam_generate", "lynx_beam_generate_ellipsoid", "lynx_beam_transform",
    "lynx_aperture_compact", "lynx_aos7_to_soa6", "lynx_soa6_to_aos7", "lynx_screen_allreduce",
    "lynx_screen_allreduce_mc",
)

_lib = None


def lib():
    """Load the shared library (once).  Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise LynxError(
                f"{LIB_PATH} is missing: build it with `python -m lynx_b200.build` "
                "(there is no CPU fallback)"
            )
        handle = C.CDLL(os.fspath(LIB_PATH))
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(rc: int, what: str = "lynx_b200") -> None:
    if rc != OK:
        msg = lib().lynx_status_str(rc).decode()
        raise LynxError(f"{what} failed: {msg} ({rc})")


def ptr(t) -> int:
    """Raw data pointer of a torch tensor (or None -> NULL)."""
    return None if t is None else t.data_ptr()


def current_stream() -> int:
    import torch

    return torch.cuda.current_stream().cuda_stream

"""Encode a flat list of elements into the C-ABI lattice program + parameter table.

Host-side logic only (runs without a GPU up to ``Program.handle``): decides, from
Python values at call time, which elements are active / skippable -- exactly the
decisions the reference takes in ``is_skippable`` (cavity.py:68-70, bpm.py:40-42,
screen.py:82-84) and ``Segment.track`` (segment.py:344-351) -- and lays the element
parameters out as columns: a one-element tensor becomes a shared (stride-0) column, a
batch-shaped tensor a per-setting (stride-1) column.  ``Element.broadcast`` therefore
never materialises copies.
"""

import ctypes as C
from dataclasses import dataclass
from typing import List, Optional, Tuple

import torch

from . import _lib as L


@dataclass
class Encoded:
    ops: List[Tuple[int, int, int, int]]  # (opcode, col, flags, aux)
    columns: List[torch.Tensor]  # each numel()==1 (shared) or batch-shaped (per setting)
    batch_shape: Tuple[int, ...]
    bpm_elements: list  # active BPM elements, index == slot
    screen_element: Optional[object]  # the terminal active screen, if any

    @property
    def B(self) -> int:
        n = 1
        for s in self.batch_shape:
            n *= s
        return n


def broadcast_shapes(*shapes) -> Tuple[int, ...]:
    """``torch.broadcast_shapes`` for plain int shapes, without its ~25 us of symbolic-shape
    machinery per call (it ran twice per ``Segment.track``)."""
    out: List[int] = []
    for s in shapes:
        if len(s) > len(out):
            out = [1] * (len(s) - len(out)) + out
        k = len(out) - len(s)
        for i, d in enumerate(s):
            o = out[k + i]
            if o == d or d == 1:
                continue
            if o != 1:
                raise RuntimeError(f"Shape mismatch: objects cannot be broadcast to a single shape: {shapes}")
            out[k + i] = d
    return tuple(out)


def _batch_shape(shapes) -> Tuple[int, ...]:
    """Broadcast of the DISTINCT column shapes (one broadcast per column was 90 % of
    ``encode``)."""
    distinct = {tuple(s) for s in shapes}
    if len(distinct) == 1:
        return distinct.pop()
    return broadcast_shapes(*distinct)


def encode(elements, energy_shape=()) -> Encoded:
    """elements: flat list of lynx_b200 elements (no nested Segments)."""
    ops, cols = [], []
    bpms, screen = [], None
    for el in elements:
        op = el._encode()  # (opcode, [tensors], flags) or None for identity elements
        if op is None:
            continue
        opcode, tensors, flags = op
        aux = 0
        if opcode == L.OP_BPM:
            aux = len(bpms)
            bpms.append(el)
        elif opcode == L.OP_SCREEN:
            screen = el
        col = len(cols) if tensors else -1
        cols.extend(tensors)
        ops.append((opcode, col, flags, aux))
        if opcode == L.OP_SCREEN:
            break  # terminal: the reference returns Beam.empty from here on (screen.py:139)
    if not ops:
        ops.append((L.OP_MARKER, -1, 0, 0))
    shapes = {t.shape for t in cols}  # distinct shapes only: a lattice has a handful
    shapes.add(tuple(energy_shape))
    return Encoded(ops, cols, _batch_shape(shapes), bpms, screen)


def layout(enc: Encoded):
    """Column offsets / strides of the packed parameter buffer: shared scalars first,
    then the per-setting columns, each B contiguous floats."""
    B = enc.B
    memo = enc.__dict__.get("_layout")  # callers may widen batch_shape after encode(): key on B
    if memo is not None and memo[0] == B:
        return memo[1]
    shared, varying = [], []
    for i, t in enumerate(enc.columns):
        (shared if t.numel() == 1 else varying).append(i)
    n_shared = len(shared)
    # keep the per-setting block 16-byte aligned
    base = (n_shared + 3) & ~3
    off = [0] * len(enc.columns)
    stride = [0] * len(enc.columns)
    for k, i in enumerate(shared):
        off[i] = k
    for k, i in enumerate(varying):
        off[i] = base + k * B
        stride[i] = 1
    total = base + len(varying) * B
    out = (off, stride, shared, varying, max(total, 1))
    enc.__dict__["_layout"] = (B, out)
    return out


def pack_params(enc: Encoded, device) -> torch.Tensor:
    """One fp32 device buffer holding every column (differentiable w.r.t. the columns).

    The columns are concatenated WHERE THEY LIVE and cross to ``device`` in one copy per
    source device (normally one: element parameters are host tensors) -- one tiny H2D copy
    per column cost ~10 us each, ten times the tracking kernels of a 13-element lattice."""
    device = torch.device(device)
    off, stride, shared, varying, total = layout(enc)
    B = enc.B
    n_shared = len(shared)
    base = (n_shared + 3) & ~3
    f32 = torch.float32
    parts = [c if (c.dim() == 1 and c.dtype is f32) else c.reshape(1).to(f32)
             for c in (enc.columns[i] for i in shared)]
    if not parts and not varying:
        return torch.zeros(1, device=device, dtype=torch.float32)
    pad = base - n_shared
    for i in varying:
        parts.append(enc.columns[i].to(torch.float32).expand(enc.batch_shape).reshape(B))
    if len({t.device for t in parts}) == 1:  # the normal case: everything lives in one place
        if pad:
            parts.insert(n_shared, torch.zeros(pad, dtype=f32, device=parts[0].device))
        return torch.cat(parts).to(device)
    # split into runs of equal source device; the padding joins the run in front of it
    out, run, run_dev = [], [], None

    def flush():
        if run:
            out.append(torch.cat(run).to(device) if len(run) > 1 else run[0].to(device))
            run.clear()

    for k, t in enumerate(parts):
        if k == n_shared and pad:
            z = torch.zeros(pad, dtype=torch.float32, device=run_dev if run else device)
            run.append(z) if run else out.append(z)
        if run and t.device != run_dev:
            flush()
        run_dev = t.device
        run.append(t)
    if len(parts) == n_shared and pad:  # only shared columns: padding at the very end
        run.append(torch.zeros(pad, dtype=torch.float32, device=run_dev))
    flush()
    return torch.cat(out) if len(out) > 1 else out[0]


class Program:
    """Immutable device-side program handle (``lynx_program_create``), cached by
    structure."""

    _cache = {}

    def __init__(self, ops, off, stride):
        lib = L.lib()
        n = len(ops)
        arr = (L.OpT * n)(*[L.OpT(*o) for o in ops])
        nc = len(off)
        offs = (C.c_int64 * max(nc, 1))(*off)
        strs = (C.c_int32 * max(nc, 1))(*stride)
        h = C.c_void_p()
        L.check(lib.lynx_program_create(arr, n, offs, strs, nc, C.byref(h)), "lynx_program_create")
        self.handle = h
        self.n_stages = lib.lynx_program_num_stages(h)
        self.n_bpm = lib.lynx_program_num_bpm(h)
        self.n_screens = lib.lynx_program_num_screens(h)
        self.terminal = bool(lib.lynx_program_is_terminal(h))

    def workspace_bytes(self, B: int) -> int:
        return L.lib().lynx_workspace_bytes(self.handle, B)

    @classmethod
    def get(cls, enc: Encoded) -> "Program":
        off, stride, _, _, _ = layout(enc)
        key = (tuple(enc.ops), tuple(off), tuple(stride), torch.cuda.current_device())
        prog = cls._cache.get(key)
        if prog is None:
            prog = cls(enc.ops, off, stride)
            cls._cache[key] = prog
        return prog

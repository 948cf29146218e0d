"""Multi-GPU sharding of the tracking path (one process per GPU, ``torch.distributed``).

The reference has no parallelism beyond leading batch dimensions; the
path shards two ways:

* particle axis -- rank r owns a contiguous slice of every SoA plane; parameters are
  replicated; the ONE exchange step is an all-reduce (SUM) of the additive outputs: the
  screen image (integer-valued counts < 2^24, so the fp32 sum is exact and
  order-independent), the ``[B][28]`` double moment sums and the ``[slot][B][2]`` double
  BPM sums.  Moments / BPM readings are then derived from the reduced sums with the
  GLOBAL particle count.  Outgoing particles stay sharded.
* settings axis -- rank r owns a contiguous slice of the B settings; the beam is
  replicated; no collective at all (optional all-gather for a single consumer).

Everything here is host logic over ``torch.distributed`` (NCCL on GPUs, gloo in the CPU
tests); the device work is the unchanged single-GPU C ABI.
"""

from typing import Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def shard_bounds(n: int, world: int, rank: int, multiple: int = 4) -> Tuple[int, int]:
    """[lo, hi) of rank's slice of ``n`` items; every boundary except the last is a
    multiple of ``multiple`` (128-bit aligned SoA planes), sizes differ by <= multiple."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError(f"bad rank {rank} of {world}")
    blocks = (n + multiple - 1) // multiple
    base, extra = divmod(blocks, world)
    lo_b = rank * base + min(rank, extra)
    hi_b = lo_b + base + (1 if rank < extra else 0)
    return min(lo_b * multiple, n), min(hi_b * multiple, n)


def shard_soa(soa: torch.Tensor, n: int, world: int, rank: int) -> Tuple[torch.Tensor, int]:
    """Rank's slice of SoA planes ``(..., 6, ld)`` -> contiguous ``(..., 6, ld_r)``, n_r."""
    lo, hi = shard_bounds(n, world, rank)
    n_r = hi - lo
    ld_r = (n_r + 3) & ~3
    out = soa.new_zeros((*soa.shape[:-1], max(ld_r, 4)))
    out[..., :n_r] = soa[..., lo:hi]
    return out, n_r


def shard_settings(B: int, world: int, rank: int) -> Tuple[int, int]:
    """[lo, hi) of rank's slice of the B settings (no alignment requirement)."""
    return shard_bounds(B, world, rank, multiple=1)


def all_reduce_sums(tensors: Sequence[Optional[torch.Tensor]], group=None, async_op=False):
    """The one exchange step of the particle-sharded path: SUM all-reduce, in place, of
    every additive output (image, moment sums, BPM sums, adjoint H sums, g_params).
    ``None`` entries are skipped.  Returns the work handles when ``async_op``."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return []
    works = []
    for t in tensors:
        if t is None:
            continue
        w = dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group, async_op=async_op)
        if async_op:
            works.append(w)
    return works


def moments_from_sums(sums: torch.Tensor):
    """``[.., 28]`` raw double sums (N, sum v_i, sum v_i v_j i<=j) -> (mean[..,6],
    sigma[..,6] unbiased as particle_beam.py:742, cov_biased[..,6,6] as :824-836)."""
    n = sums[..., 0:1]
    s1 = sums[..., 1:7]
    mean = s1 / n
    iu = torch.triu_indices(6, 6)
    s2 = sums.new_zeros((*sums.shape[:-1], 6, 6))
    s2[..., iu[0], iu[1]] = sums[..., 7:28]
    s2 = s2 + s2.transpose(-1, -2) - torch.diag_embed(torch.diagonal(s2, dim1=-2, dim2=-1))
    centred = s2 - s1.unsqueeze(-1) * s1.unsqueeze(-2) / n.unsqueeze(-1)
    cov_biased = centred / n.unsqueeze(-1)
    var_unbiased = torch.diagonal(centred, dim1=-2, dim2=-1) / (n - 1)
    return mean, torch.sqrt(torch.clamp_min(var_unbiased, 0.0)), cov_biased


def track_sharded(segment, soa_full: torch.Tensor, n: int, energy, *, world: int, rank: int,
                  particles: bool = False, moments: int = 1, group=None):
    """Particle-sharded ``Segment.track``: tracks this rank's slice on the current CUDA
    device and all-reduces the additive outputs.  Returns (local outgoing beam or
    Beam.empty, reduced image or None, reduced moment sums or None)."""
    from .beams import Beam, ParticleBeam

    local, n_r = shard_soa(soa_full, n, world, rank)
    beam = ParticleBeam.from_soa(local.cuda(), n_r, torch.as_tensor(energy).cuda())
    out = segment.track(beam, particles=particles, moments=moments)
    img = None
    screens = [e for e in segment.elements if getattr(e, "cached_reading", None) is not None]
    if screens:
        img = screens[-1].cached_reading
    bpm_elems = [e for e in segment.elements if getattr(e, "_bpm_sums", None) is not None]
    sums = None if out is Beam.empty else out._sums
    all_reduce_sums([img, sums] + [e._bpm_sums for e in bpm_elems], group=group)
    for e in bpm_elems:  # bpm.py:54 with the global particle count
        e.reading = (e._bpm_sums / n).to(torch.float32).T.reshape(e.reading.shape)
    return out, img, sums


class ScreenExchange:
    """The particle-sharded path's exchange step for the screen image over NVLink peer
    memory (csrc/exchange.cu): every rank histograms into a slot of a peer-mapped symmetric
    buffer, then ONE kernel per rank does the cross-GPU barrier and sums the union bounding
    box of the touched bins over all peers.  Falls back to a dense NCCL all-reduce when
    symmetric memory cannot be set up (``self.fused`` is False)."""

    def __init__(self, B: int, W: int, H: int, device, group=None):
        import ctypes as C

        from . import _lib as L

        self.B, self.W, self.H, self.device = B, W, H, device
        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        self.step = 0
        self.lib = L.lib()
        self._L = L
        self.fused = False
        self._mc = 0
        self.out = torch.zeros((B, H, W), dtype=torch.float32, device=device)
        self.state = torch.zeros(8, dtype=torch.int32, device=device)
        try:
            import torch.distributed._symmetric_memory as symm_mem

            nbytes = self.lib.lynx_screen_symm_bytes(B, W, H)
            self.buf = symm_mem.empty(nbytes // 4, dtype=torch.float32, device=device)
            self.buf.zero_()
            torch.cuda.synchronize(device)
            self.hdl = symm_mem.rendezvous(self.buf, self.group)
            ptrs = list(self.hdl.buffer_ptrs)
            self._peers = (C.c_void_p * self.world)(*ptrs)
            # NVLS: the multicast mapping lets the switch sum the peers' pixels (0 = unsupported)
            import os
            self._mc = int(getattr(self.hdl, "multicast_ptr", 0) or 0)
            if os.environ.get("LYNX_EX_NO_MULTICAST") == "1":
                self._mc = 0
            self.hdl.barrier()
            torch.cuda.synchronize(device)
            self.fused = W % 4 == 0 and self.world <= 16
        except Exception as e:  # no P2P / symmetric memory: dense NCCL all-reduce
            self.why = repr(e)
            self.buf = torch.zeros((2, B, H, W), dtype=torch.float32, device=device)

    def slot_pointers(self):
        """(image pointer, bbox pointer) of the slot the NEXT step tracks into."""
        slot = self.step & 1
        if not self.fused:
            return self.buf[slot].data_ptr(), None
        img = self.lib.lynx_screen_symm_image(self.buf.data_ptr(), slot, self.B, self.W, self.H)
        return img, img - 64

    def reduce(self) -> torch.Tensor:
        """All ranks' images of the current slot -> ``self.out`` (summed); advances the step."""
        slot = self.step & 1
        self.step += 1
        if not self.fused:
            self.out.copy_(self.buf[slot])
            dist.all_reduce(self.out, group=self.group)
            return self.out
        self._L.check(self.lib.lynx_screen_allreduce_mc(
            self._peers, self._mc or None, self.world, self.rank, slot, self.step, self.B, self.W,
            self.H, self.out.data_ptr(), self.state.data_ptr(), self._L.current_stream()),
            "lynx_screen_allreduce")
        return self.out

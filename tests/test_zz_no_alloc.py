"""lynx_track_particles_fwd_noalloc (csrc/noalloc.cu): the fused tracker with its scratch in a
static device array instead of a caller workspace, for callers that cannot call cudaMalloc.
Checked through the C ABI against lynx_track_particles_fwd on the same inputs, on two calls in a
row with different magnet settings (the shared scratch keeps the first call's records).  Same
kernels, same inputs: particles, image and energy must be bit-identical; the double moment sums
are accumulated with atomics in arbitrary order, hence 1e-12 relative there.

Collected last on purpose."""

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from lynx_b200 import _lib as L  # noqa: E402
from lynx_b200 import fixtures as F  # noqa: E402
from lynx_b200.program import Program, encode, pack_params  # noqa: E402


def _run(lib, entry, spec, soa, n, ld, dev, screen):
    seg = F.build_segment(spec)
    enc = encode(seg.elements, (1,))
    prog = Program.get(enc)
    params = pack_params(enc, dev)
    e_in = torch.tensor([F.ASTRA_ENERGY], device=dev)
    e_out = torch.empty(1, device=dev)
    mom = torch.empty((1, 28), dtype=torch.float64, device=dev)
    out = img = ex = ey = None
    W = H = 0
    if screen:
        ex, ey = seg.AREABSCR1._device_edges(dev)
        W, H = ex.numel() - 1, ey.numel() - 1
        img = torch.full((1, H, W), -1.0, device=dev)  # the call must zero it
    else:
        out = torch.empty((6, ld), device=dev)
    args = [prog.handle, params.data_ptr(), 1, soa.data_ptr(), 1, ld, n, L.ptr(out), ld, None,
            L.ptr(img), W, H, L.ptr(ex), L.ptr(ey), mom.data_ptr(), 2, e_in.data_ptr(),
            e_out.data_ptr()]
    if entry == "noalloc":
        L.check(lib.lynx_track_particles_fwd_noalloc(*args, L.current_stream()), "noalloc")
    else:
        ws = torch.empty(prog.workspace_bytes(1), dtype=torch.uint8, device=dev)
        L.check(lib.lynx_track_particles_fwd(*args, ws.data_ptr(), L.current_stream()), "fwd")
    torch.cuda.synchronize()
    return out, img, mom, e_out


def test_noalloc_entry_point_matches_the_workspace_one_twice_in_a_row():
    lib = L.lib()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(0)
    n = 300_001
    ld = (n + 3) & ~3
    soa = torch.zeros((6, ld), device=dev)
    soa[:, :n] = torch.from_numpy(F.gaussian_beam_soa(n, seed=4242)).to(dev)
    small = {"resolution": (640, 480), "pixel_size": (8e-6, 8e-6), "binning": 1}
    calls = [
        (F.ares_ea(k1=(5.0, -5.0, 5.0), vkick=1e-3, hkick=-2e-3), False),
        (F.ares_ea(k1=(-3.0, 4.0, -2.0), vkick=-2e-4, hkick=5e-4, screen_active=True, screen=small), True),
    ]
    for spec, screen in calls:  # two calls in a row on the shared static scratch
        got = _run(lib, "noalloc", spec, soa, n, ld, dev, screen)
        want = _run(lib, "fwd", spec, soa, n, ld, dev, screen)
        if screen:
            assert got[1].sum().item() > 0.5 * n
            assert torch.equal(got[1], want[1])
        else:
            assert torch.equal(got[0][:, :n], want[0][:, :n])
        np.testing.assert_allclose(got[2].cpu().numpy(), want[2].cpu().numpy(), rtol=1e-12, atol=0)
        assert torch.equal(got[3], want[3])
    # and once more with the first settings: nothing stale from the second call leaks in
    again = _run(lib, "noalloc", calls[0][0], soa, n, ld, dev, False)
    ref = _run(lib, "fwd", calls[0][0], soa, n, ld, dev, False)
    assert torch.equal(again[0][:, :n], ref[0][:, :n])


def test_noalloc_rejects_problems_beyond_its_static_scratch():
    lib = L.lib()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(0)
    seg = F.build_segment(F.ares_ea())
    enc = encode(seg.elements, (1,))
    prog = Program.get(enc)
    params = pack_params(enc, dev)
    B = L.NOALLOC_WORKSPACE_BYTES // 64  # >= 16384 settings: records alone exceed 1 MiB
    assert prog.workspace_bytes(B) > L.NOALLOC_WORKSPACE_BYTES
    n, ld = 16, 16
    soa = torch.zeros((6, ld), device=dev)
    e_in = torch.full((B,), F.ASTRA_ENERGY, device=dev)
    e_out = torch.empty(B, device=dev)
    mom = torch.empty((B, 28), dtype=torch.float64, device=dev)
    rc = lib.lynx_track_particles_fwd_noalloc(
        prog.handle, params.data_ptr(), B, soa.data_ptr(), 0, ld, n, None, ld, None, None, 0, 0,
        None, None, mom.data_ptr(), 1, e_in.data_ptr(), e_out.data_ptr(), L.current_stream())
    assert rc == L.E_TOOLARGE
    assert b"static scratch" in lib.lynx_status_str(rc)
    with pytest.raises(L.LynxError):
        L.check(rc)

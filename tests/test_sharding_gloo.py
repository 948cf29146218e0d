"""World-size-2 ``gloo`` test of the particle-sharded path's host logic:
shard bounds, the one all-reduce of the additive outputs (image, moment sums, BPM sums)
and moments derived from the reduced sums -- against the single-process oracle on the
whole beam.  The per-shard "device" outputs are produced here by the oracle / NumPy (no
GPU in this container); on a GPU box the same functions wrap the C-ABI outputs
(``lynx_b200.sharding.track_sharded``, exercised by ``bench.py --gpus N``)."""

import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from helpers import aos_from_soa, oracle_elements
from lynx_b200 import fixtures as F
from lynx_b200 import sharding as S
from oracle import lattice as O
from oracle import screen_bin_edges, screen_reading_particles

N = 30_001  # deliberately not a multiple of 4 * world
SCREEN = {"resolution": (200, 160), "pixel_size": (2e-5, 2e-5), "binning": 1}


def _raw_sums(planes: np.ndarray) -> torch.Tensor:
    """(6, n) -> [28] double raw sums in the C-ABI layout (N, sum v_i, sum v_i v_j i<=j)."""
    v = planes.astype(np.float64)
    out = [float(v.shape[1])] + [v[i].sum() for i in range(6)]
    out += [(v[i] * v[j]).sum() for i in range(6) for j in range(i, 6)]
    return torch.tensor(out, dtype=torch.float64)


def _tracked_planes():
    soa = F.gaussian_beam_soa(N, seed=77)
    spec = F.ares_ea(k1=(5.0, -5.0, 5.0), vkick=1e-3, hkick=-2e-3)
    els = oracle_elements(spec, torch.float64)
    beam = O.ParticleBeamO(aos_from_soa(soa, torch.float64),
                           torch.tensor([F.ASTRA_ENERGY], dtype=torch.float64))
    out = O.track(els, beam)
    return out, out.particles[0, :, :6].T.numpy()


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        _, planes = _tracked_planes()  # every rank recomputes; only its slice is used below
        lo, hi = S.shard_bounds(N, world, rank)
        mine = planes[:, lo:hi]
        sums = _raw_sums(mine).unsqueeze(0)
        bpm = torch.tensor([[mine[0].sum(), mine[2].sum()]], dtype=torch.float64)
        ex, ey = screen_bin_edges(SCREEN["resolution"], SCREEN["pixel_size"])
        img = screen_reading_particles(torch.from_numpy(mine[0:1]).float(),
                                       torch.from_numpy(mine[2:3]).float(), ex, ey)
        S.all_reduce_sums([img, sums, None, bpm])
        mean, sigma, cov = S.moments_from_sums(sums)
        q.put((rank, lo, hi, img.numpy(), sums.numpy(), mean.numpy(), sigma.numpy(),
               cov.numpy(), (bpm / N).numpy()))
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_shard_bounds_cover_and_align():
    for n in (1, 3, 4, 5, 1000, 1001, 12_500_000, 100_000_000):
        for world in (1, 2, 3, 4, 8):
            prev = 0
            for r in range(world):
                lo, hi = S.shard_bounds(n, world, r)
                assert lo == prev and lo <= hi <= n
                assert lo % 4 == 0 or lo == n
                prev = hi
            assert prev == n
            sizes = [S.shard_bounds(n, world, r) for r in range(world)]
            assert max(h - l for l, h in sizes) - min(h - l for l, h in sizes) < 8
    assert S.shard_settings(10, 4, 3) == (8, 10)
    with pytest.raises(ValueError):
        S.shard_bounds(10, 2, 2)


def test_shard_soa_slices_and_pads():
    soa = torch.arange(6 * 12, dtype=torch.float32).reshape(6, 12)
    a, na = S.shard_soa(soa[:, :10], 10, 2, 0)
    b, nb = S.shard_soa(soa[:, :10], 10, 2, 1)
    assert na + nb == 10 and a.shape[-1] % 4 == 0 and b.shape[-1] % 4 == 0
    assert torch.equal(torch.cat([a[:, :na], b[:, :nb]], dim=1), soa[:, :10])
    assert (b[:, nb:] == 0).all()


def test_world2_gloo_reduction_matches_single_process():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in procs], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res[0][1] == 0 and res[0][2] == res[1][1] and res[1][2] == N

    out, planes = _tracked_planes()
    ex, ey = screen_bin_edges(SCREEN["resolution"], SCREEN["pixel_size"])
    whole_img = screen_reading_particles(torch.from_numpy(planes[0:1]).float(),
                                         torch.from_numpy(planes[2:3]).float(), ex, ey).numpy()
    for r in res:
        _, _, _, img, sums, mean, sigma, cov, bpm = r
        assert np.array_equal(img, whole_img)  # integer counts: exact, order-independent
        assert sums[0, 0] == N
        np.testing.assert_allclose(mean[0], [out.mu(i).item() for i in range(6)], rtol=1e-10,
                                   atol=1e-18)
        np.testing.assert_allclose(sigma[0], [out.sigma(i).item() for i in range(6)], rtol=1e-9)
        np.testing.assert_allclose(cov[0, 0, 1], out.sigma_xxp.item(), rtol=1e-8)
        np.testing.assert_allclose(cov[0, 2, 3], out.sigma_yyp.item(), rtol=1e-8)
        np.testing.assert_allclose(bpm[0], [out.mu(0).item(), out.mu(2).item()], rtol=1e-10)
    # both ranks hold identical reduced results
    assert np.array_equal(res[0][4], res[1][4])

"""Segment.track caches the encoded lattice (program handle + packed device parameters) between
calls (segment._encoded_lattice); every way a user changes the lattice must invalidate it."""

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

import lynx_b200 as lx  # noqa: E402
from lynx_b200 import fixtures as F  # noqa: E402
from lynx_b200 import segment as S  # noqa: E402


def _beam(n=20_000):
    soa = F.gaussian_beam_soa(n, seed=4)
    return lx.ParticleBeam(torch.from_numpy(F.soa_to_aos7(soa)).unsqueeze(0), torch.tensor([F.ASTRA_ENERGY]))


def test_cache_hits_and_invalidation():
    seg = F.build_segment(F.ares_ea(k1=(5.0, -5.0, 5.0), vkick=1e-3, hkick=-2e-3))
    beam = _beam()
    S._LATTICE_CACHE.clear()
    a = seg.track(beam)
    n0 = len(S._LATTICE_CACHE)
    b = seg.track(beam)
    assert len(S._LATTICE_CACHE) == n0 == 1  # second call: a hit
    assert torch.equal(a.soa, b.soa)

    def fresh(k1):
        s2 = F.build_segment(F.ares_ea(k1=k1, vkick=1e-3, hkick=-2e-3))
        return s2.track(beam).soa

    # 1. attribute re-assigned
    seg.AREAMQZM1.k1 = torch.tensor([7.0])
    assert torch.equal(seg.track(beam).soa, fresh((7.0, -5.0, 5.0)))
    # 2. tensor modified in place
    seg.AREAMQZM1.k1.fill_(3.0)
    assert torch.equal(seg.track(beam).soa, fresh((3.0, -5.0, 5.0)))
    seg.AREAMQZM1.k1 += 1.0
    assert torch.equal(seg.track(beam).soa, fresh((4.0, -5.0, 5.0)))
    # 3. batch shape of a parameter changes
    seg.AREAMQZM2.k1 = torch.tensor([-5.0, -4.0])
    out = seg.track(beam)
    assert out.soa.shape[0] == 2
    assert torch.equal(out.soa[0], fresh((4.0, -5.0, 5.0))[0])
    assert torch.equal(out.soa[1], fresh((4.0, -4.0, 5.0))[0])
    # 4. a flag: the screen switched on makes the lattice terminal
    seg.AREAMQZM2.k1 = torch.tensor([-5.0])
    assert seg.track(beam) is not lx.Beam.empty
    seg.AREABSCR1.is_active = True
    assert seg.track(beam) is lx.Beam.empty
    assert seg.AREABSCR1.reading.sum().item() > 0
    # 5. ParameterBeam through the same cache, same rules
    seg.AREABSCR1.is_active = False
    pb = lx.ParameterBeam.from_parameters(sigma_x=torch.tensor([1e-4]), energy=torch.tensor([F.ASTRA_ENERGY]))
    m0 = seg.track(pb)._mu.clone()
    assert torch.equal(seg.track(pb)._mu, m0)
    seg.AREAMCVM1.angle = torch.tensor([2e-3])
    m1 = seg.track(pb)._mu
    assert not torch.equal(m1, m0)
    seg.AREAMCVM1.angle.zero_()
    seg.AREAMCVM1.angle += 1e-3
    assert torch.allclose(seg.track(pb)._mu, m0)


def test_cache_is_bounded():
    S._LATTICE_CACHE.clear()
    beam = _beam(1000)
    for i in range(40):
        lx.Segment([lx.Drift(torch.tensor([0.1 + i]))]).track(beam)
    assert len(S._LATTICE_CACHE) <= S._LATTICE_CACHE_MAX

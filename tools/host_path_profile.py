"""Interpreter cost of one ``Segment.track`` call with the shared library replaced by a
stand-in that computes nothing (runs WITHOUT a GPU; CPU tensors): what is measured is the
Python around the C ABI -- encode, pack_params, Program.get, output allocation, ctypes.

    python tools/host_path_profile.py [--profile]

Best of 10 batches of 400 calls (the container is noisy)."""
import cProfile
import os
import pstats
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import lynx_b200 as lx  # noqa: E402
from lynx_b200 import _lib as L  # noqa: E402
from lynx_b200 import fixtures, segment  # noqa: E402


class _Lib:
    def __getattr__(self, name):
        if name == "lynx_program_num_stages":
            return lambda *a: 1
        if name.endswith("workspace_bytes"):
            return lambda *a: 1024
        return lambda *a: 0


def best(fn, k=400, batches=10):
    for _ in range(100):
        fn()
    out = 1e9
    for _ in range(batches):
        t0 = time.perf_counter()
        for _ in range(k):
            fn()
        out = min(out, (time.perf_counter() - t0) / k * 1e6)
    return out


def main():
    fake = _Lib()
    L.lib = lambda: fake
    L.current_stream = lambda: 0
    segment._require_cuda = lambda d: None
    torch.cuda.current_device = lambda: 0
    seg = fixtures.build_segment(fixtures.ares_ea(k1=(5.0, -5.0, 5.0), vkick=1e-3, hkick=-2e-3))
    n = 10_000
    beam = lx.ParticleBeam.from_soa(torch.zeros(1, 6, n), n, torch.tensor([1e8]), torch.ones(n))
    rows = {"particle_beam_B1_us": best(lambda: seg.track(beam))}
    if "--profile" in sys.argv:
        pr = cProfile.Profile()
        pr.enable()
        for _ in range(2000):
            seg.track(beam)
        pr.disable()
        pstats.Stats(pr).sort_stats("cumulative").print_stats(25)
    B = 4096
    for name in ("AREAMQZM1", "AREAMQZM2", "AREAMQZM3"):
        getattr(seg, name).k1 = torch.randn(B)
    seg.AREAMCVM1.angle = torch.randn(B) * 1e-3
    seg.AREAMCHM1.angle = torch.randn(B) * 1e-3
    pb = lx.ParameterBeam.from_parameters(device="cpu")
    rows["parameter_beam_B4096_us"] = best(lambda: seg.track(pb))
    print({k: round(v, 1) for k, v in rows.items()})


if __name__ == "__main__":
    main()

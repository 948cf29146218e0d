"""Small, fixed launch sequence for ncu: python tools/prof_target.py --mode screen1m --steps 3"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402

import bench  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--mode", default="screen1m")
ap.add_argument("--steps", type=int, default=3)
a = ap.parse_args()
dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
flush = torch.empty((1 << 10) if os.environ.get('LYNX_NOFLUSH') else (256 << 20), dtype=torch.uint8, device=dev)
if a.mode == "screen1m":
    w = bench.ParticleWorkload(1_000_000, dev, seed=1002, screen=True)
elif a.mode == "screen32m":
    w = bench.ParticleWorkload(32_000_000, dev, seed=7, screen=True)
elif a.mode == "cfg4screen":
    w = bench.ParticleWorkload(12_500_000, dev, seed=1004, screen=True, lattice="cfg4")
elif a.mode == "out32m":
    w = bench.ParticleWorkload(32_000_000, dev, seed=7, screen=False, want_out=True)
elif a.mode == "outmom32m":
    w = bench.ParticleWorkload(32_000_000, dev, seed=7, screen=False, want_out=True, moments=1)
elif a.mode == "none1m":
    w = bench.ParticleWorkload(1_000_000, dev, seed=7, screen=False)
elif a.mode == "pb4096":
    w = bench.ParamBeamWorkload(4096, dev)
elif a.mode == "pbboth":  # launches 1-2: B = 4096, launches 3-4: B = 2^20
    w1, w2 = bench.ParamBeamWorkload(4096, dev), bench.ParamBeamWorkload(1 << 20, dev)

    class _W2:
        i = 0

        def step(self):
            (w1 if self.i < 2 else w2).step()
            self.i += 1

    w = _W2()
    a.steps = 2
elif a.mode == "pb1m":
    w = bench.ParamBeamWorkload(1 << 20, dev)
elif a.mode == "adj32m":
    w = bench.GradWorkload(32_000_000, 1, dev, per_particle=True)
    w.step = w.backward
elif a.mode == "grad1m1024":
    w = bench.GradWorkload(1_000_000, 1024, dev)
elif a.mode in ("transform32m", "aperture32m", "ellipsoid32m"):
    import lynx_b200 as lx

    class _W:
        pass

    w = _W()
    if a.mode == "ellipsoid32m":
        w.step = lambda: lx.ParticleBeam.uniform_3d_ellipsoid(
            num_particles=32_000_000, radius_x=torch.tensor([1e-3]), seed=3)
    else:
        beam = lx.ParticleBeam.from_parameters(
            num_particles=32_000_000, sigma_x=torch.tensor([2e-4]), sigma_y=torch.tensor([2e-4]),
            seed=2)
        beam.moment_sums()
        if a.mode == "transform32m":
            w.step = lambda: beam.transformed_to(sigma_x=torch.tensor([1e-4]))
        else:
            apert = lx.Aperture(x_max=torch.tensor(2e-4), y_max=torch.tensor(2e-4),
                                shape="elliptical")
            w.step = lambda: apert.track(beam)
else:
    raise SystemExit("unknown mode")
ms = bench.timed_steps(w.step, a.steps, 2, flush)
print(a.mode, "ms/step", sum(ms) / len(ms))

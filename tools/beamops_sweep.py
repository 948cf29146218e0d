"""Kernel times of the beam-level kernels at 32 M particles (library event hooks):
python tools/beamops_sweep.py [transform|aperture|ellipsoid ...]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
import lynx_b200 as lx  # noqa: E402
from lynx_b200 import _lib as L  # noqa: E402

dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
lib = L.lib()
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
n = int(os.environ.get("N", 32_000_000))
modes = sys.argv[1:] or ["transform", "aperture", "ellipsoid"]
beam = lx.ParticleBeam.from_parameters(num_particles=n, sigma_x=torch.tensor([2e-4]),
                                       sigma_y=torch.tensor([2e-4]), seed=2)
beam.moment_sums()
for m in modes:
    if m == "transform":
        f, which, nbytes = (lambda: beam.transformed_to(sigma_x=torch.tensor([1e-4]))), 4, 48 * n
    elif m == "aperture":
        ap = lx.Aperture(x_max=torch.tensor(2e-4), y_max=torch.tensor(2e-4), shape="elliptical")
        f, which, nbytes = (lambda: ap.track(beam)), 5, 48 * n
    else:
        f, which, nbytes = (lambda: lx.ParticleBeam.uniform_3d_ellipsoid(
            num_particles=n, radius_x=torch.tensor([1e-3]), seed=3)), -1, 24 * n
    f()
    step = float(np.mean(bench.timed_steps(f, 10, 3, flush)))
    k = bench.kernel_ms(f, which, 10, flush, lib) if which >= 0 else None
    print(m, "step_ms", round(step, 4), "kernel_ms", None if k is None else round(k, 4),
          "kernel_GB/s", None if k is None else round(nbytes / k / 1e6, 1), flush=True)

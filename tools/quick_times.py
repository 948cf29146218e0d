"""Quick device timings of the workloads the docs quote (one line each), for A/B runs:
    python tools/quick_times.py [names...]

    step1m   ARES EA 1 M particles + screen, whole step (K1 + K2), L2 flushed between steps
    scr32m   32 M particles, terminal screen
    out32m   32 M particles, outgoing particles
    cfg4     config-4 shard (12.5 M, cavity + dipole), screen / outgoing particles
    k6       ParameterBeam env-steps at B = 4096, 65536, 2^20
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def main():
    names = sys.argv[1:] or ["step1m", "scr32m", "out32m", "cfg4", "k6"]
    dev = torch.device("cuda:0")
    torch.cuda.set_device(0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    from lynx_b200 import _lib as L

    lib = L.lib()
    PEAK = bench.peaks()[0]

    def report(name, w, steps=30, which=1):
        ms = bench.timed_steps(w.step, steps, 5, flush)
        k = bench.kernel_ms(w.step, which, 10, flush, lib)
        ab = w.algorithmic_bytes()
        print(f"{name:28s} step {np.mean(ms) * 1e3:8.2f} us (median {np.median(ms) * 1e3:.1f})  kernel {k * 1e3:8.1f} us  "
              f"{ab / k / 1e6:7.0f} GB/s alg  frac {ab / k / 1e6 / PEAK:.3f}", flush=True)

    if "step1m" in names:
        w = bench.ParticleWorkload(1_000_000, dev, 1002)
        report("step1m (K1+K2, screen)", w, steps=100)
        ms = bench.timed_steps(w.persistent_step, 100, 5, flush)
        print(f"{'step1m persistent image':28s} step {np.mean(ms) * 1e3:8.2f} us (median {np.median(ms) * 1e3:.1f})", flush=True)
    if "scr32m" in names:
        report("scr32m (screen, 24 B)", bench.ParticleWorkload(32_000_000, dev, 1002))
    if "out32m" in names:
        report("out32m (particles, 48 B)",
               bench.ParticleWorkload(32_000_000, dev, 1002, screen=False, want_out=True))
    if "cfg4" in names:
        report("cfg4 12.5M screen", bench.ParticleWorkload(12_500_000, dev, 1004, lattice="cfg4"))
        report("cfg4 12.5M out",
               bench.ParticleWorkload(12_500_000, dev, 1004, lattice="cfg4", screen=False, want_out=True))
    if "k6" in names:
        for B in (4096, 65536, 1 << 20):
            report(f"k6 B={B}", bench.ParamBeamWorkload(B, dev), which=2)


if __name__ == "__main__":
    main()

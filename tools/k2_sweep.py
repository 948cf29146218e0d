"""Kernel-time sweep (library event hooks): python tools/k2_sweep.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402

import bench  # noqa: E402
from lynx_b200 import _lib as L  # noqa: E402

dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
flush = torch.empty((1 << 10) if os.environ.get("NOFLUSH") else (256 << 20), dtype=torch.uint8, device=dev)
lib = L.lib()
ns = [int(x) for x in (sys.argv[1].split(",") if len(sys.argv) > 1 else ["1000000"])]
for n in ns:
    for name, kw in (("screen", dict(screen=True)), ("mom1", dict(screen=False, moments=1)),
                     ("none", dict(screen=False)), ("out", dict(screen=False, want_out=True)),
                     ("out+mom2", dict(screen=False, want_out=True, moments=2))):
        w = bench.ParticleWorkload(n, dev, seed=3, **kw)
        k2 = bench.kernel_ms(w.step, 1, 10, flush, lib)
        k1 = bench.kernel_ms(w.step, 0, 10, flush, lib)
        ms = bench.timed_steps(w.step, 10, 3, flush)
        gb = w.algorithmic_bytes() / (k2 * 1e-3) / 1e9
        print(f"N={n:>9d} {name:9s} K2 {k2*1e3:8.2f} us  K1 {k1*1e3:6.2f} us  step {sum(ms)/len(ms)*1e3:8.2f} us  {gb:7.1f} GB/s")
        del w

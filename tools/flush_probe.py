"""Does the KIND of L2 flush matter?  The bench contract's flush writes a 256 MB buffer, which
leaves the 126 MB L2 full of DIRTY lines: every line the step then allocates (24 MB of particle
reads, 20 MB of image zeroing) first evicts one of them to HBM.  Compares, for the 1 M-particle
step and K6 at B = 4096: write flush (the contract's), write flush followed by a 256 MB read
(cold L2, clean lines), read-only flush."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

dev = torch.device("cuda:0")
torch.cuda.set_device(0)
wbuf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
rbuf = torch.ones(64 << 20, dtype=torch.float32, device=dev)  # 256 MB


class Flush:
    def __init__(self, kind):
        self.kind = kind

    def zero_(self):
        if "w" in self.kind:
            wbuf.zero_()
        if "r" in self.kind:
            rbuf.sum()


w1 = bench.ParticleWorkload(1_000_000, dev, 1002)
k6 = bench.ParamBeamWorkload(4096, dev)
w32 = bench.ParticleWorkload(32_000_000, dev, 1002)
for kind, name in (("w", "write (contract)"), ("wr", "write then read"), ("r", "read only"), ("w", "write (contract)")):
    f = Flush(kind)
    a = bench.timed_steps(w1.step, 100, 5, f)
    p = bench.timed_steps(w1.persistent_step, 100, 5, f)
    b = bench.timed_steps(k6.step, 100, 5, f)
    c = bench.timed_steps(w32.step, 20, 3, f)
    print(f"{name:18s} step1m {np.mean(a) * 1e3:6.2f} us  persistent {np.mean(p) * 1e3:6.2f} us  "
          f"k6 B=4096 {np.mean(b) * 1e3:6.2f} us  scr32m {np.mean(c) * 1e3:7.2f} us", flush=True)

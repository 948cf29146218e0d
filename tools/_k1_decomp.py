import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, bench
from lynx_b200 import _lib as L
dev = torch.device("cuda:0"); torch.cuda.set_device(0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
lib = L.lib()
w = bench.ParticleWorkload(1_000_000, dev, 1002)
ms = bench.timed_steps(w.step, 300, 10, flush)
k1 = bench.kernel_ms(w.step, 0, 20, flush, lib)
k2 = bench.kernel_ms(w.step, 1, 20, flush, lib)
print(os.environ.get("TAG", ""), f"step {np.mean(ms)*1e3:.2f} us (median {np.median(ms)*1e3:.1f})  K1 {k1*1e3:.1f} us  K2 {k2*1e3:.1f} us", flush=True)

"""Phase timeline of one K2 CTA (experiments library only):
    LYNX_B200_LIB=lynx_b200/lib/liblynx_b200_exp.so python tools/k2_timeline.py [mode] [cta]
Prints, averaged over steps, the %globaltimer stamps of the CTA's first consumer thread relative
to its kernel entry, and K1's start / walker end read from the histogram workspace."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import bench
from lynx_b200 import _lib as L

mode = sys.argv[1] if len(sys.argv) > 1 else "screen1m"  # screen1m | screen1m_persistent | cfg4 | screen32m
cta = int(sys.argv[2]) if len(sys.argv) > 2 else 0
dev = torch.device("cuda:0")
torch.cuda.set_device(0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
lib = L.lib()
fn = lib.lynx_dbg_k2_trace
fn.argtypes = [C.POINTER(C.c_uint64), C.c_int]
if mode.startswith("screen1m"):
    w = bench.ParticleWorkload(1_000_000, dev, 1002)
elif mode == "cfg4":
    w = bench.ParticleWorkload(12_500_000, dev, 1004, lattice="cfg4")
else:
    w = bench.ParticleWorkload(32_000_000, dev, 7)
buf = (C.c_uint64 * 16)()
fn(buf, cta)
names = ["entry", "prologue done", "K1 complete (griddep wait)", "records in smem", "tile 1 landed",
         "tile 1 done", "tile loop done", "tile flushed (vector reductions issued)", "(same)", "(unused)",
         "bounding box merged", "tile 1: bins known, placement starts", "tile 1: warp statistics exchanged",
         "tile 1: window placed"]
rows = []
for i in range(60):
    if not os.environ.get('LYNX_NOFLUSH'):
        flush.zero_()
    if mode.endswith("_persistent"):
        w.persistent_step()
    else:
        w.step()
    torch.cuda.synchronize()
    fn(buf, cta)
    t = np.array(list(buf), dtype=np.float64)
    hist = w.ws[-256:].view(torch.int64).cpu().numpy().astype(np.float64)
    if i >= 10:
        rows.append(np.concatenate([t[:14] - t[0], hist[8:18] - t[0]]))
m = np.mean(rows, axis=0) / 1e3
for n, v in zip(names + ["K1 block 0 start", "K1 walker done", "K1 histogram block done", "K1 last zeroing block done", "K1 hist: copies issued", "K1 hist: tables in smem", "K1 hist: deviations done", "K1 walker: parameters in smem", "K1 walker: walk done", "K1 walker block entry"], m):
    print(f"{n:32s} {v:8.2f} us")

"""Where the host-buffer step (bench e2e) spends its time: pure pinned H2D of the beam vs the
whole lynx_host_track_particles call.  python tools/e2e_breakdown.py"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import ctypes as C  # noqa: E402

import torch  # noqa: E402

import bench  # noqa: E402
from lynx_b200 import _lib as L  # noqa: E402

dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
lib = L.lib()
w = bench.ParticleWorkload(1_000_000, dev, seed=1002, screen=True)
w.prepare_host()
ctx = C.c_void_p()
L.check(lib.lynx_host_ctx_create(0, C.byref(ctx)))


def wall(f, n=200, warm=20):
    for _ in range(warm):
        f()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        f()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3


dst = torch.empty_like(w.soa)
s = torch.cuda.Stream()


def h2d():
    with torch.cuda.stream(s):
        dst.copy_(w.soa_host, non_blocking=True)
    s.synchronize()


print("pure H2D 24 MB pinned      ms", round(wall(h2d), 4), "GB/s",
      round(w.soa_host.numel() * 4 / wall(h2d) / 1e6, 1))
small = torch.empty(4, dtype=torch.int32).pin_memory()
dsmall = torch.zeros(4, dtype=torch.int32, device=dev)


def d2h_small():
    with torch.cuda.stream(s):
        small.copy_(dsmall, non_blocking=True)
    s.synchronize()


print("16 B D2H + sync            ms", round(wall(d2h_small), 4))
for opt in (0, 1):
    lib.lynx_host_ctx_set_option(ctx, 1, opt)
    print(f"host step persistent={opt}     ms", round(wall(lambda: w.host_step(ctx)), 4))
print("device step (no flush)     ms", round(wall(w.step), 4))

"""The exchange kernel's fixed cost on ONE GPU (world = 1, the rank is its own peer):
    python tools/ex_single_rank.py
step = K1 + K2 into the symmetric slot [+ lynx_screen_allreduce], per-step events, L2 flushed."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import bench
from lynx_b200 import _lib as L

dev = torch.device("cuda:0")
torch.cuda.set_device(0)
lib = L.lib()
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
w = bench.ParticleWorkload(1_000_000, dev, 1002)
nbytes = lib.lynx_screen_symm_bytes(1, w.W, w.H)
buf = torch.zeros(nbytes // 4, dtype=torch.float32, device=dev)
out = torch.zeros((1, w.H, w.W), device=dev)
state = torch.zeros(8, dtype=torch.int32, device=dev)
peers = (C.c_void_p * 1)(buf.data_ptr())
step_no = [0]


def track_only():
    slot = step_no[0] & 1
    img = lib.lynx_screen_symm_image(buf.data_ptr(), slot, 1, w.W, w.H)
    L.check(lib.lynx_track_particles_fwd_bbox(
        w.prog.handle, w.params.data_ptr(), 1, w.soa.data_ptr(), 1, w.ld, w.n, None, w.ld, None,
        img, w.W, w.H, L.ptr(w.ex), L.ptr(w.ey), None, 0, w.e_in.data_ptr(), w.e_out.data_ptr(),
        w.ws.data_ptr(), img - 64, L.current_stream()))
    return slot


def with_exchange():
    slot = track_only()
    step_no[0] += 1
    L.check(lib.lynx_screen_allreduce(peers, 1, 0, slot, step_no[0], 1, w.W, w.H, out.data_ptr(),
                                      state.data_ptr(), L.current_stream()))


a = np.mean(bench.timed_steps(track_only, 200, 10, flush)) * 1e3
b = np.mean(bench.timed_steps(with_exchange, 200, 10, flush)) * 1e3
w.step()
torch.cuda.synchronize()
print(f"{os.environ.get('TAG', '')} K1 + K2 into the slot {a:.2f} us; + exchange kernel (world 1) {b:.2f} us; "
      f"exchange adds {b - a:.2f} us; image ok: {torch.equal(out, w.img)}")

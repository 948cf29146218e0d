import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch, bench
dev = torch.device("cuda", 0)
screen = (sys.argv[1] if len(sys.argv) > 1 else "screen") == "screen"
w = bench.ParticleWorkload(1_000_000, dev, seed=1002, screen=screen)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
names = ["start", "staged", "walk done", "zero+hist done", "barrier done", "tile0 landed", "consumers done", "last CTA end", "first CTA start", "last CTA start", "first instr"]
acc = []
marks = []
mark = torch.zeros(2, dtype=torch.int64, device=dev)
for i in range(12):
    st = w.ws[-256:].view(torch.int64)
    st[16:27] = 0; st[24] = 2**62
    flush.zero_(); w.lib.lynx_debug_stamp(mark[0:1].data_ptr(), w.L.current_stream()); w.step(); w.lib.lynx_debug_stamp(mark[1:2].data_ptr(), w.L.current_stream()); torch.cuda.synchronize()
    t = st[16:27].tolist(); marks.append([(int(mark[0]) - t[0]) / 1e3, (int(mark[1]) - t[0]) / 1e3])
    acc.append([(x - t[0]) / 1e3 if x else float("nan") for x in t])
import numpy as np
m = np.nanmean(np.array(acc[4:]), axis=0)
print({n: round(float(v), 2) for n, v in zip(names, m)})
ms = bench.timed_steps(w.step, 30, 3, flush)
print("marker before / after (us rel. CTA0 start):", np.mean(np.array(marks[4:]), axis=0))
print("step us", 1e3 * sum(ms) / len(ms))

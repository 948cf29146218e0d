"""Wall-clock cost of one user-facing ``Segment.track`` call (ARES EA, screen off) with the
element columns packed per column (one H2D copy each, the first implementation) or per
source device (``program.pack_params``).

Run on the GPU box:  python tools/track_call_overhead.py [n_particles]
"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import lynx_b200 as lx  # noqa: E402
from lynx_b200 import fixtures, program, segment  # noqa: E402


def pack_per_column(enc, device):
    off, stride, shared, varying, total = program.layout(enc)
    parts = [enc.columns[i].reshape(1).to(device=device, dtype=torch.float32) for i in shared]
    pad = ((len(shared) + 3) & ~3) - len(shared)
    if pad:
        parts.append(torch.zeros(pad, device=device))
    for i in varying:
        parts.append(enc.columns[i].to(device=device, dtype=torch.float32)
                     .expand(enc.batch_shape).reshape(enc.B))
    return torch.cat(parts)


def wall(fn, reps=200):
    for _ in range(20):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e6


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
    seg = fixtures.build_segment(fixtures.ares_ea(k1=(5.0, -5.0, 5.0), vkick=1e-3, hkick=-2e-3))
    beam = lx.ParticleBeam.from_parameters(num_particles=n, device="cuda")
    new = program.pack_params
    rows = {}
    for name, fn in (("track_per_column", pack_per_column), ("track_per_device", new)):
        segment.pack_params = fn
        rows[name] = wall(lambda: seg.track(beam))
    segment.pack_params = new
    enc = program.encode(seg.elements, ())
    a, b = pack_per_column(enc, "cuda"), new(enc, "cuda")
    assert torch.equal(a, b), "packings differ"
    rows["pack_per_column"] = wall(lambda: pack_per_column(enc, "cuda"))
    rows["pack_per_device"] = wall(lambda: new(enc, "cuda"))
    rows["encode"] = wall(lambda: program.encode(seg.elements, ()))
    print({k: round(v, 1) for k, v in rows.items()}, "us per call, n =", n)


if __name__ == "__main__":
    main()

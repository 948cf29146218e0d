"""GPU tests of the C ABI's launch contract (include/lynx_b200.h, "Contract"): the launchers only
ENQUEUE on the caller's stream -- no allocation, no synchronisation, no library-owned mutable
state -- so that

* two large screen launches may run concurrently on two streams of one device (round 1 kept one
  process-wide replica scratch for >= 4 M-particle launches: the two images contaminated each
  other), and
* a launch can be captured in a CUDA graph and replayed (round 1 called cudaMalloc /
  cudaDeviceSynchronize inside the launcher, which is illegal during capture).

The launch under test is ``lynx_track_particles_fwd`` on pre-encoded device buffers, the call a
jit-compiled caller makes (bench.ParticleWorkload).
"""

import pytest
import torch

pytestmark = pytest.mark.gpu

import bench  # noqa: E402


def _workload(n, seed):
    return bench.ParticleWorkload(n, torch.device("cuda", 0), seed=seed, screen=True)


def test_two_concurrent_screen_launches_on_two_streams():
    """2 x 6 M particles (different beams) on two streams, 20 rounds: each image must equal the
    image of the same workload run alone, bit for bit, and hold exactly its particle count."""
    torch.cuda.set_device(0)
    n = 6_000_000
    w = [_workload(n, 11), _workload(n, 12)]
    ref = []
    for x in w:
        x.step()
        torch.cuda.synchronize()
        ref.append(x.img.clone())
    assert not torch.equal(ref[0], ref[1])
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    for _ in range(20):
        for x in w:
            x.img.fill_(-1.0)  # on the default stream; the launch zeroes the image itself
        torch.cuda.synchronize()
        for s, x in zip(streams, w):
            with torch.cuda.stream(s):
                x.step()
                x.step()  # back to back on the same stream as well
        torch.cuda.synchronize()
        for x, r in zip(w, ref):
            assert torch.equal(x.img, r)
    for r in ref:
        assert r.sum().item() <= n and r.sum().item() > 0.99 * n


@pytest.mark.parametrize("n", [1_000_000, 5_000_000])
def test_forward_launch_is_graph_capturable(n):
    """Capture K1 + K2 (programmatic dependent launch included) in a CUDA graph, replay it on
    changed inputs: same image as the eager launch of the same inputs."""
    torch.cuda.set_device(0)
    w = _workload(n, 21)
    w.step()  # eager once: kernel attributes are set on first use, outside the capture
    torch.cuda.synchronize()
    eager = w.img.clone()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        w.step()
    w.img.fill_(-1.0)
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(w.img, eager)
    # new particles in the same buffers: the replayed graph tracks them
    other = _workload(n, 22)
    other.step()
    torch.cuda.synchronize()
    assert not torch.equal(other.img, eager)
    w.soa.copy_(other.soa)
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    assert torch.equal(w.img, other.img)

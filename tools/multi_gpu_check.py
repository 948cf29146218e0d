"""Multi-GPU correctness of the fused screen exchange (run under torchrun, N ranks):
every rank tracks its particle shard of one seeded beam into the symmetric buffer,
lynx_screen_allreduce sums the bounding box over NVLink peers; the result must equal,
bit for bit, the single-GPU image of the whole beam.  Several steps with changing
settings exercise the epoch flags and the double buffering."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from lynx_b200 import _lib as L  # noqa: E402
from lynx_b200 import fixtures as F  # noqa: E402
from lynx_b200 import sharding as S  # noqa: E402
from lynx_b200.program import Program, encode, pack_params  # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
local = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
lib = L.lib()
N = 400_003
full = torch.from_numpy(F.gaussian_beam_soa(N, seed=4242))
ok_all = True
ex = None
for step, (k1, vk, hk) in enumerate([((5.0, -5.0, 5.0), 1e-3, -2e-3), ((3.0, -6.0, 4.0), -5e-4, 1e-3),
                                     ((0.0, 0.0, 0.0), 0.0, 0.0), ((8.0, -9.0, 7.0), 2e-3, 2e-3),
                                     ((5.0, -5.0, 5.0), 1e-3, -2e-3)]):
    seg = F.build_segment(F.ares_ea(k1=k1, vkick=vk, hkick=hk, screen_active=True, screen=F.TEST_SCREEN))
    enc = encode(seg.elements, (1,))
    prog = Program.get(enc)
    params = pack_params(enc, dev)
    scr = seg.AREABSCR1
    ex_d, ey_d = scr._device_edges(dev)
    W, H = ex_d.numel() - 1, ey_d.numel() - 1
    if ex is None:
        ex = S.ScreenExchange(1, W, H, dev)
        if rank == 0:
            print("fused exchange:", ex.fused, getattr(ex, "why", ""), "multicast:", hex(ex._mc), flush=True)
    e_in = torch.tensor([F.ASTRA_ENERGY], device=dev)
    e_out = torch.empty(1, device=dev)
    ws = torch.empty(prog.workspace_bytes(1), dtype=torch.uint8, device=dev)

    def track(soa, n, img_ptr, bbox_ptr):
        ld = soa.shape[-1]
        L.check(lib.lynx_track_particles_fwd_bbox(
            prog.handle, params.data_ptr(), 1, soa.data_ptr(), 1, ld, n, None, ld, None, img_ptr, W, H,
            ex_d.data_ptr(), ey_d.data_ptr(), None, 0, e_in.data_ptr(), e_out.data_ptr(),
            ws.data_ptr(), bbox_ptr, L.current_stream()))

    mine, n_r = S.shard_soa(full, N, world, rank)
    mine = mine.to(dev)
    img_ptr, bbox_ptr = ex.slot_pointers()
    track(mine, n_r, img_ptr, bbox_ptr)
    got = ex.reduce().clone()
    # single-GPU image of the whole beam (every rank computes it redundantly)
    ld = (N + 3) & ~3
    whole = torch.zeros((6, ld), device=dev)
    whole[:, :N] = full.to(dev)
    ref = torch.empty((1, H, W), device=dev)
    track(whole, N, ref.data_ptr(), None)
    torch.cuda.synchronize()
    same = torch.equal(got, ref)
    ok_all &= same
    print(f"rank {rank} step {step}: equal={same} sum={got.sum().item():.0f} ref={ref.sum().item():.0f}",
          flush=True)
dist.barrier()
flag = torch.tensor([1 if ok_all else 0], device=dev)
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    print("MULTI_GPU_CHECK", "PASS" if flag.item() == 1 else "FAIL", flush=True)
dist.destroy_process_group()
sys.exit(0 if flag.item() == 1 else 1)

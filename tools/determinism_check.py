import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import lynx_b200 as lx
from lynx_b200 import _lib as L, fixtures as F
from lynx_b200.program import Program, encode, pack_params
lib = L.lib()
k1, vk, hk = (5.0, -5.0, 5.0), 1e-3, -2e-3
seg = F.build_segment(F.ares_ea(k1=k1, vkick=vk, hkick=hk, screen_active=True, screen=F.TEST_SCREEN))
enc = encode(seg.elements, (1,)); prog = Program.get(enc); params = pack_params(enc, "cuda")
ex, ey = seg.AREABSCR1._device_edges(torch.device("cuda")); W, H = ex.numel() - 1, ey.numel() - 1
ws = torch.empty(prog.workspace_bytes(1), dtype=torch.uint8, device="cuda")
e_in = torch.tensor([F.ASTRA_ENERGY], device="cuda"); e_out = torch.empty(1, device="cuda")
seg2 = F.build_segment(F.ares_ea(k1=k1, vkick=vk, hkick=hk))
for n in [int(a) for a in sys.argv[1:]]:
    soa_np = F.gaussian_beam_soa(n, seed=21)
    ld = (n + 3) & ~3
    soa = torch.zeros((6, ld), device="cuda"); soa[:, :n] = torch.from_numpy(soa_np).cuda()
    def fresh():
        img = torch.full((1, H, W), 3.0, device="cuda")
        L.check(lib.lynx_track_particles_fwd(prog.handle, params.data_ptr(), 1, soa.data_ptr(), 1, ld, n, None, ld, None,
                img.data_ptr(), W, H, ex.data_ptr(), ey.data_ptr(), None, 0, e_in.data_ptr(), e_out.data_ptr(), ws.data_ptr(), L.current_stream()))
        torch.cuda.synchronize(); return img
    imgs = [fresh() for _ in range(4)]
    nd = [int((imgs[0] != im).sum()) for im in imgs[1:]]
    beam = lx.ParticleBeam.from_soa(soa.unsqueeze(0), n, e_in)
    outs = [seg2.track(beam).soa.clone() for _ in range(3)]
    torch.cuda.synchronize()
    no = [int((outs[0] != o).sum()) for o in outs[1:]]
    print("n", n, "image ndiff between runs", nd, "sum", [float(i.sum()) for i in imgs[:2]], "out-particle ndiff", no, flush=True)

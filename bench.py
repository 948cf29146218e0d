#!/usr/bin/env python
"""Benchmark of the beam-tracking hot path (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # this repo, CUDA path
    python bench.py --impl reference --gpus N --steps K ...  # reference arithmetic on CPU

Workload (BASELINE.json configs[1]): ARES experimental-area segment (13 elements;
AREAMQZM1-3 k1 = 5,-5,5; AREAMCVM1 = 1e-3, AREAMCHM1 = -2e-3; AREABSCR1 active,
2448x2040) tracking a 1M-particle synthetic Astra-shaped ParticleBeam per GPU; one
step = Segment.track + Screen.reading.  Metric: particle-element updates/s =
particles * 13 / step time.  Secondary: ParameterBeam env-steps/s at 4096 settings
(configs[2]) and the streaming roofline point at 32M particles.
"""

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402
import torch  # noqa: E402

E_ARES = 13
K1 = (5.0, -5.0, 5.0)
VKICK, HKICK = 1e-3, -2e-3


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--particles", type=int, default=1_000_000, help="per GPU")
    ap.add_argument("--settings", type=int, default=4096)
    ap.add_argument("--scale-particles", type=int, default=32_000_000)
    ap.add_argument("--no-extras", action="store_true")
    ap.add_argument("--exchange", default="fused", choices=["fused", "nccl"],
                    help="multi-GPU screen exchange: fused peer-memory kernel or dense NCCL")
    return ap.parse_args()


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def traffic_of(kernel, workload, n):
    """Measured DRAM bytes per launch from the committed table (profiles/traffic.json), or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            e = json.load(f).get(f"{kernel}/{workload}/{n}")
        return (e["traffic"], e["source"]) if e else (None, None)
    except Exception:
        return None, None


class ClockSampler:
    """SM clock / throttle reasons DURING the timed region: an NVML polling thread (the
    timed region is a few ms, far shorter than nvidia-smi's sampling period), with
    nvidia-smi as the fallback."""

    NAMES = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "sw_thermal_slowdown": 0x20,
             "hw_thermal_slowdown": 0x40}

    def __init__(self, index):
        self.sm, self.reasons, self.max_mhz = [], 0, None
        self._stop = threading.Event()
        self._nvml = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
            self._nvml = pynvml
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
        except Exception:
            self._nvml = None
            self._index = index

    def _poll(self):
        nv = self._nvml
        while not self._stop.is_set():
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM)))
                self.reasons |= int(nv.nvmlDeviceGetCurrentClocksEventReasons(self._h))
            except Exception:
                try:
                    self.reasons |= int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h))
                except Exception:
                    pass
            time.sleep(0.0002)

    def stop(self):
        if self._nvml is not None:
            self._stop.set()
            self.t.join(timeout=2)
            reasons = sorted(n for n, bit in self.NAMES.items() if self.reasons & bit)
            return {"sm_mhz": float(np.median(self.sm)) if self.sm else None,
                    "sm_max_mhz": self.max_mhz, "reasons": reasons, "samples": len(self.sm),
                    "source": "nvml polled during the timed region"}
        try:  # one nvidia-smi query right after the region
            q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
                 "clocks_event_reasons.hw_thermal_slowdown,"
                 "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
            out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                  "-i", str(self._index)], capture_output=True, text=True,
                                 timeout=20).stdout.strip().split(",")
            names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
            return {"sm_mhz": float(out[0]), "sm_max_mhz": float(out[1]),
                    "reasons": [names[i] for i in range(4) if out[2 + i].strip().lower().startswith("active")],
                    "samples": 1, "source": "nvidia-smi after the timed region"}
        except Exception:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["clock query unavailable"]}


# ----------------------------------------------------------------------------- CPU arm
def oracle_workload(n, dtype=torch.float32, seed=1002):
    from helpers import aos_from_soa, oracle_elements
    from lynx_b200 import fixtures as F
    from oracle import lattice as O
    from oracle import screen_bin_edges

    spec = F.ares_ea(k1=K1, vkick=VKICK, hkick=HKICK, screen_active=True, screen=F.TEST_SCREEN)
    els = oracle_elements(spec, dtype)
    beam = O.ParticleBeamO(aos_from_soa(F.gaussian_beam_soa(n, seed=seed), dtype),
                           torch.tensor([F.ASTRA_ENERGY], dtype=dtype))
    ex, ey = screen_bin_edges(F.TEST_SCREEN["resolution"], F.TEST_SCREEN["pixel_size"])
    return els, beam, ex, ey


def oracle_step(els, beam, ex, ey):
    """Segment.track + Screen.reading in the restated reference arithmetic (torch CPU)."""
    from oracle import lattice as O
    from oracle import screen_reading_particles

    O.track(els, beam)
    rb = els[-1].read_beam
    return screen_reading_particles(rb.col(0), rb.col(2), ex, ey)


def time_cpu(n, steps, warmup):
    els, beam, ex, ey = oracle_workload(n)
    for _ in range(warmup):
        oracle_step(els, beam, ex, ey)
    t0 = time.perf_counter()
    for _ in range(steps):
        oracle_step(els, beam, ex, ey)
    return (time.perf_counter() - t0) / steps


def time_cpu_parambeam(B, steps, warmup=1):
    """Config 3 in the restated reference arithmetic: ONE vectorised Segment.track of a
    ParameterBeam over B settings (lynx/accelerator/element.py:71-82), torch CPU, fp32."""
    from helpers import oracle_elements
    from lynx_b200 import fixtures as F
    from oracle import lattice as O

    k1, ang = F.random_settings(B)
    ov = {(nm, "k1"): k1[:, i] for i, nm in enumerate(("AREAMQZM1", "AREAMQZM2", "AREAMQZM3"))}
    ov[("AREAMCVM1", "angle")], ov[("AREAMCHM1", "angle")] = ang[:, 0], ang[:, 1]
    els = oracle_elements(F.ares_ea(), torch.float32, (B,), ov)
    pb = O.ParameterBeamO.from_parameters(F.ASTRA_MU[:4], F.ASTRA_SIGMA, F.ASTRA_ENERGY)
    pb = O.ParameterBeamO(pb.mu_.expand(B, 7).clone(), pb.cov_.expand(B, 7, 7).clone(),
                          pb.energy.expand(B).clone())
    for _ in range(warmup):
        O.track(els, pb)
    t0 = time.perf_counter()
    for _ in range(steps):
        O.track(els, pb)
    return (time.perf_counter() - t0) / steps


def run_reference(args, rank, world):
    if rank != 0:
        return
    n = args.particles
    torch.set_num_threads(os.cpu_count() or 1)
    sec = time_cpu(n, args.steps, min(args.warmup, 2))
    val = n * E_ARES / sec
    line = {
        "impl": "reference", "metric": "particle_element_updates_per_s", "value": val,
        "unit": "particle-element updates/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "ares_ea_1M_particles_screen_2448x2040", "elements": E_ARES,
                   "particles": n,
                   "note": "torch-CPU restatement of the reference arithmetic, bit-identical "
                           "to the reference's own outputs (tests/test_reference_golden.py); "
                           "the Python reference cannot travel to the GPU box"},
        "cpu_baseline": {"value": val, "unit": "particle-element updates/s",
                         "cores": torch.get_num_threads(), "kind": "port",
                         "sample": f"{args.steps} full steps of {n} particles"},
        "e2e": {"value": val, "unit": "particle-element updates/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
    }
    # the second half of the metric on the same cores: ParameterBeam env-steps/s (config 3)
    psec = time_cpu_parambeam(args.settings, max(3, min(args.steps, 20)))
    line["env_steps"] = {"value": args.settings / psec, "unit": "env-steps/s",
                         "settings": args.settings, "ms_per_step": psec * 1e3,
                         "cores": torch.get_num_threads(), "kind": "port"}
    emit(json.dumps(line))


# ----------------------------------------------------------------------------- GPU arm
class ParticleWorkload:
    """Pre-encoded Segment.track(ParticleBeam) (+ Screen.reading) on device buffers: what
    a jit-compiled caller holds -- program handle, packed parameters, resident beam."""

    def __init__(self, n, device, seed, screen=True, want_out=False, moments=0, lattice="ares"):
        import lynx_b200 as lx
        from lynx_b200 import _lib as L
        from lynx_b200 import fixtures as F
        from lynx_b200.program import Program, encode, pack_params

        self.L, self.n, self.device = L, n, device
        if lattice == "cfg4":  # BASELINE.json configs[3]: Cavity + Dipole + quadrupoles, E = 10
            spec = F.cavity_dipole_line(screen_active=screen)
            energy, beam_soa, scr_name = F.CFG4_ENERGY, F.cfg4_beam_soa, "SCR"
        else:
            spec = F.ares_ea(k1=K1, vkick=VKICK, hkick=HKICK, screen_active=screen,
                             screen=F.TEST_SCREEN)
            energy, beam_soa, scr_name = F.ASTRA_ENERGY, F.gaussian_beam_soa, "AREABSCR1"
        self.n_elements = len(spec)
        self.seg = F.build_segment(spec)
        self.enc = encode(self.seg.elements, (1,))
        self.prog = Program.get(self.enc)
        self.params = pack_params(self.enc, device)
        self.params_host = self.params.cpu().pin_memory()
        self.ld = (n + 3) & ~3
        self.soa_host = torch.empty((6, self.ld), dtype=torch.float32).pin_memory()
        self.soa_host[:, :n] = torch.from_numpy(beam_soa(n, seed=seed))
        self.soa = self.soa_host.to(device)
        self.e_in = torch.tensor([energy], dtype=torch.float32, device=device)
        self.e_out = torch.empty(1, device=device)
        self.ws = torch.empty(self.prog.workspace_bytes(1), dtype=torch.uint8, device=device)
        self.out = torch.empty((6, self.ld), device=device) if want_out else None
        self.mom = torch.empty((1, 28), dtype=torch.float64, device=device) if moments else None
        self.moments = moments
        self.img = self.ex = self.ey = None
        self.W = self.H = 0
        if screen:
            scr = getattr(self.seg, scr_name)
            self.ex, self.ey = scr._device_edges(device)
            self.W, self.H = self.ex.numel() - 1, self.ey.numel() - 1
            self.img = torch.empty((1, self.H, self.W), device=device)
            self.exh, self.eyh = scr.pixel_bin_edges
            self.img_host = torch.empty((1, self.H, self.W)).pin_memory()
        self.lib = L.lib()

    def step(self):
        L = self.L
        L.check(self.lib.lynx_track_particles_fwd(
            self.prog.handle, self.params.data_ptr(), 1, self.soa.data_ptr(), 1, self.ld, self.n,
            L.ptr(self.out), self.ld, None, L.ptr(self.img), self.W, self.H, L.ptr(self.ex),
            L.ptr(self.ey), L.ptr(self.mom), self.moments, self.e_in.data_ptr(),
            self.e_out.data_ptr(), self.ws.data_ptr(), L.current_stream()))

    def persistent_step(self):
        """The same step with a PERSISTENT screen image (lynx_track_particles_fwd_persistent):
        image and state zeroed once, every call clears only the rectangle the previous call
        lit instead of the whole 20 MB buffer.  Bit-identical image (tests)."""
        L = self.L
        if not hasattr(self, "_pimg"):
            self._pimg = torch.zeros((1, self.H, self.W), device=self.soa.device)
            self._pstate = torch.zeros(8, dtype=torch.int32, device=self.soa.device)
            self._epoch = 0
        L.check(self.lib.lynx_track_particles_fwd_persistent(
            self.prog.handle, self.params.data_ptr(), 1, self.soa.data_ptr(), 1, self.ld, self.n,
            None, self.ld, None, self._pimg.data_ptr(), self.W, self.H, L.ptr(self.ex),
            L.ptr(self.ey), L.ptr(self.mom), self.moments, self.e_in.data_ptr(),
            self.e_out.data_ptr(), self.ws.data_ptr(), self._pstate.data_ptr(), self._epoch,
            L.current_stream()))
        self._epoch += 1

    def sharded_step(self, exchange):
        """N > 1: this rank's particle shard into the symmetric slot, then the ONE exchange
        step of the path -- the fused bounding-box all-reduce over NVLink peers."""
        L = self.L
        img_ptr, bbox_ptr = exchange.slot_pointers()
        L.check(self.lib.lynx_track_particles_fwd_bbox(
            self.prog.handle, self.params.data_ptr(), 1, self.soa.data_ptr(), 1, self.ld, self.n,
            None, self.ld, None, img_ptr, self.W, self.H, L.ptr(self.ex), L.ptr(self.ey), None, 0,
            self.e_in.data_ptr(), self.e_out.data_ptr(), self.ws.data_ptr(), bbox_ptr,
            L.current_stream()))
        exchange.reduce()

    def algorithmic_bytes(self):
        b = 24 * self.n
        if self.out is not None:
            b += 24 * self.n
        if self.img is not None:
            b += 4 * self.W * self.H
        return b

    def host_step(self, ctx):
        """Through the host-buffer C ABI: H2D beam + params, kernels, D2H image."""
        L = self.L
        L.check(self.lib.lynx_host_track_particles(
            ctx, self.prog.handle, self.params_host.data_ptr(), self.params_host.numel(), 1,
            self.soa_host.data_ptr(), 1, self.n, self.e_in_host.data_ptr(), None, None,
            self.img_host.data_ptr(), self.W, self.H, self.exh.data_ptr(), self.eyh.data_ptr(),
            None, 0, None))

    def prepare_host(self):
        self.e_in_host = self.e_in.cpu().pin_memory()
        self.exh, self.eyh = self.exh.pin_memory(), self.eyh.pin_memory()


class ParamBeamWorkload:
    """Config 3: B random (3 x k1, 2 x angle) settings x one ParameterBeam through ARES EA."""

    def __init__(self, B, device, seed=1003):
        from lynx_b200 import _lib as L
        from lynx_b200 import fixtures as F
        from lynx_b200.program import Program, encode, pack_params

        self.L, self.B = L, B
        k1, ang = F.random_settings(B, seed)
        seg = F.build_segment(F.ares_ea())
        for i, nm in enumerate(("AREAMQZM1", "AREAMQZM2", "AREAMQZM3")):
            getattr(seg, nm).k1 = torch.from_numpy(k1[:, i])
        seg.AREAMCVM1.angle = torch.from_numpy(ang[:, 0])
        seg.AREAMCHM1.angle = torch.from_numpy(ang[:, 1])
        self.enc = encode(seg.elements, (1,))
        self.prog = Program.get(self.enc)
        self.params = pack_params(self.enc, device)
        self.params_host = self.params.cpu().pin_memory()
        mu = torch.tensor([*F.ASTRA_MU[:4], 0.0, 0.0, 1.0])
        cov = torch.zeros(7, 7)
        for i in range(6):
            cov[i, i] = F.ASTRA_SIGMA[i] ** 2
        self.mu_h, self.cov_h = mu.pin_memory(), cov.pin_memory()
        self.mu, self.cov = mu.to(device), cov.to(device)
        self.e_in = torch.full((B,), F.ASTRA_ENERGY, device=device)
        self.e_in_h = self.e_in.cpu().pin_memory()
        self.mu_out = torch.empty((B, 7), device=device)
        self.cov_out = torch.empty((B, 7, 7), device=device)
        self.e_out = torch.empty(B, device=device)
        self.mu_out_h = torch.empty((B, 7)).pin_memory()
        self.cov_out_h = torch.empty((B, 7, 7)).pin_memory()
        self.lib = L.lib()

    def step(self):
        self.L.check(self.lib.lynx_parambeam_fwd(
            self.prog.handle, self.params.data_ptr(), self.B, self.mu.data_ptr(),
            self.cov.data_ptr(), 0, self.e_in.data_ptr(), self.mu_out.data_ptr(),
            self.cov_out.data_ptr(), None, self.e_out.data_ptr(), self.L.current_stream()))

    def host_step(self, ctx):
        self.L.check(self.lib.lynx_host_parambeam(
            ctx, self.prog.handle, self.params_host.data_ptr(), self.params_host.numel(), self.B,
            self.mu_h.data_ptr(), self.cov_h.data_ptr(), 1, self.e_in_h.data_ptr(),
            self.mu_out_h.data_ptr(), self.cov_out_h.data_ptr(), None, None))

    def algorithmic_bytes(self):  # 20 B in + 228 B out per env-step, shared beam
        return 248 * self.B


class GradWorkload:
    """Config 5: one beam shared by B settings through ARES EA (screen off), loss on the
    outgoing moments; a step = forward (moment sums) + backward (gradients w.r.t. every
    parameter column, incl. 3 x k1 + 2 x angle per setting) through the C ABI."""

    def __init__(self, n, B, device, seed=1005, per_particle=False):
        from lynx_b200 import _lib as L
        from lynx_b200 import fixtures as F
        from lynx_b200.program import Program, encode, pack_params

        self.L, self.n, self.B, self.per_particle = L, n, B, per_particle
        k1, ang = F.random_settings(B, seed, k1_range=5.0)
        seg = F.build_segment(F.ares_ea())
        for i, nm in enumerate(("AREAMQZM1", "AREAMQZM2", "AREAMQZM3")):
            getattr(seg, nm).k1 = torch.from_numpy(k1[:, i])
        seg.AREAMCVM1.angle = torch.from_numpy(ang[:, 0])
        seg.AREAMCHM1.angle = torch.from_numpy(ang[:, 1])
        self.enc = encode(seg.elements, (1,))
        self.prog = Program.get(self.enc)
        self.params = pack_params(self.enc, device)
        self.ld = (n + 3) & ~3
        self.soa = torch.zeros((6, self.ld), device=device)
        self.soa[:, :n] = torch.from_numpy(F.gaussian_beam_soa(n, seed=seed)).to(device)
        self.e_in = torch.full((B,), F.ASTRA_ENERGY, device=device)
        self.e_out = torch.empty(B, device=device)
        self.lib = L.lib()
        self.ws = torch.empty(self.prog.workspace_bytes(B), dtype=torch.uint8, device=device)
        self.bws = torch.empty(self.lib.lynx_bwd_workspace_bytes(self.prog.handle, B),
                               dtype=torch.uint8, device=device)
        self.mom = torch.empty((B, 28), dtype=torch.float64, device=device)
        self.g_mom = torch.randn((B, 28), dtype=torch.float64, device=device)
        self.g_params = torch.empty_like(self.params)
        # per-particle cotangent variant (B must be 1): forces the streaming K3a pass
        self.out = torch.empty((B, 6, self.ld), device=device) if per_particle else None
        self.g_out = torch.randn((B, 6, self.ld), device=device) if per_particle else None
        self.g_in = torch.empty((6, self.ld), device=device) if per_particle else None

    def forward(self):
        L = self.L
        L.check(self.lib.lynx_track_particles_fwd(
            self.prog.handle, self.params.data_ptr(), self.B, self.soa.data_ptr(), 0, self.ld,
            self.n, L.ptr(self.out), self.ld, None, None, 0, 0, None, None, self.mom.data_ptr(), 1,
            self.e_in.data_ptr(), self.e_out.data_ptr(), self.ws.data_ptr(), L.current_stream()))

    def backward(self):
        L = self.L
        L.check(self.lib.lynx_track_particles_bwd(
            self.prog.handle, self.params.data_ptr(), self.B, self.soa.data_ptr(), 0, self.ld,
            self.n, L.ptr(self.g_out), self.ld, self.g_mom.data_ptr(), None,
            self.e_in.data_ptr(), self.g_params.data_ptr(), L.ptr(self.g_in), self.ld,
            self.bws.data_ptr(), L.current_stream()))

    def step(self):
        self.forward()
        self.backward()


def timed_steps(step, steps, warmup, flush, post=None):
    """Per-step CUDA-event timing on the current stream with an L2 flush (outside the
    events) before every step.  Returns the list of per-step milliseconds."""
    for _ in range(warmup):
        flush.zero_()
        step()
        if post:
            post()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
           for _ in range(steps)]
    t0 = time.perf_counter()
    for a, b in evs:
        flush.zero_()
        a.record()
        step()
        if post:
            post()
        b.record()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    # host enqueue time per step, and how long the device still ran after the last enqueue:
    # the host must stay ahead of the device or launch gaps end up inside the events
    timed_steps.host_us = (t1 - t0) / steps * 1e6
    timed_steps.drain_us = (t2 - t1) * 1e6
    return [a.elapsed_time(b) for a, b in evs]


def kernel_ms(step, which, steps, flush, lib):
    """Average device time of the dominant kernel alone (library event hooks)."""
    lib.lynx_profile_enable(1)
    acc = []
    for _ in range(steps):
        flush.zero_()
        step()
        ms = lib.lynx_profile_last_ms(which)
        if ms >= 0:
            acc.append(ms)
    lib.lynx_profile_enable(0)
    return float(np.mean(acc)) if acc else None


def run_ours(args, rank, world, local_rank):
    import torch.distributed as dist

    from lynx_b200 import _lib as L

    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    assert args.particles % 4 == 0
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    lib = L.lib()
    peak, peak_src = peaks()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=device)  # > 126 MB L2
    from lynx_b200.sharding import ScreenExchange

    wl = ParticleWorkload(args.particles, device, seed=1002 + rank, screen=True)
    post = None
    step_fn = wl.step
    exchange_kind = None
    if world > 1:  # the one exchange step of the path: the screen image (north star)
        if args.exchange == "nccl":
            post = lambda: dist.all_reduce(wl.img)  # noqa: E731
            exchange_kind = "dense NCCL all-reduce of the 20 MB image"
        else:
            exch = ScreenExchange(1, wl.W, wl.H, device)
            step_fn = lambda: wl.sharded_step(exch)  # noqa: E731
            exchange_kind = ("fused bounding-box all-reduce over NVLink peer memory "
                             "(lynx_screen_allreduce)" if exch.fused
                             else "dense NCCL all-reduce (symmetric memory unavailable)")

    # ---- headline: device-resident steps, per-step events, max over ranks
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    lib.lynx_launch_count_reset()
    trace = None
    if world > 1 and os.environ.get("LYNX_EX_TRACE"):
        trace = torch.zeros(8192, dtype=torch.int64, device=device)
        lib.lynx_exchange_trace(trace.data_ptr())
    ms = timed_steps(step_fn, args.steps, args.warmup, flush, post)
    if trace is not None:
        lib.lynx_exchange_trace(None)
        t = trace.view(1024, 8)[args.warmup + 2:args.warmup + args.steps].double()
        t = t[t[:, 0] > 0]
        rel = ((t - t[:, 4:5]).mean(0) / 1e3).tolist()  # us after K2 passed its wait on K1
        sys.stderr.write(f"[rank {rank}] host enqueue {timed_steps.host_us:.1f} us/step, device drain "
                         f"after last enqueue {timed_steps.drain_us:.0f} us; K1 start {rel[6]:.2f}; ")
        sys.stderr.write(f"[rank {rank}] us after K1 done: K2 CTA0 done {rel[5]:.2f}, EX resident "
                         f"{rel[0]:.2f}, K2 complete {rel[1]:.2f}, peers arrived {rel[2]:.2f}, "
                         f"written {rel[3]:.2f}, last CTA {rel[7]:.2f}; step mean {sum(ms) / len(ms) * 1e3:.2f}\n")
        # where the rest of the step goes: stamp kernels in front of and behind the step
        st2 = torch.zeros(2, dtype=torch.int64, device=device)
        lib.lynx_exchange_trace(trace.data_ptr())
        gaps = []
        for i in range(60):
            flush.zero_()
            lib.lynx_debug_stamp(st2.data_ptr(), L.current_stream())
            step_fn()
            lib.lynx_debug_stamp(st2.data_ptr() + 8, L.current_stream())
            torch.cuda.synchronize()
            row = trace.view(1024, 8)[exch.step & 1023].tolist()
            t0, t1 = st2.tolist()
            if i >= 10 and row[0] > 0:
                gaps.append(((row[6] - t0) / 1e3, (t1 - row[7]) / 1e3, (t1 - t0) / 1e3))
        lib.lynx_exchange_trace(None)
        g = np.mean(gaps, axis=0)
        sys.stderr.write(f"[rank {rank}] stamp kernel -> K1 block 0 starts {g[0]:.2f} us; EX last CTA -> "
                         f"stamp kernel {g[1]:.2f} us; stamp to stamp {g[2]:.2f} us\n")
        ms_local = timed_steps(wl.step, 50, 5, flush)
        if hasattr(exch, "reduce"):
            real_reduce = exch.reduce
            exch.reduce = lambda: None  # track into the symmetric slot, no exchange
            ms_symm = timed_steps(step_fn, 50, 5, flush)
            exch.reduce = real_reduce
            sys.stderr.write(f"[rank {rank}] no exchange: local image {np.mean(ms_local) * 1e3:.2f} us, "
                             f"symmetric slot {np.mean(ms_symm) * 1e3:.2f} us\n")
    launches = lib.lynx_launch_count() - 0
    launches_timed = launches * args.steps // (args.steps + args.warmup)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    clocks = sampler.stop() if sampler else None
    total_ms = torch.tensor([sum(ms)], device=device, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
    ms_per_step = total_ms.item() / args.steps
    value = world * args.particles * E_ARES / (ms_per_step * 1e-3)

    # ---- e2e through the host-buffer C ABI (H2D beam + params, D2H image), per rank
    ctx = C.c_void_p()
    L.check(lib.lynx_host_ctx_create(local_rank, C.byref(ctx)))
    wl.prepare_host()

    def time_host():
        for _ in range(3):
            wl.host_step(ctx)
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            wl.host_step(ctx)
        t = torch.tensor([(time.perf_counter() - t0) / args.steps], device=device,
                         dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t

    e2e_dense_s = time_host()  # full 20 MB image read back every step
    dense_img = wl.img_host.clone()
    # persistent host image: only the rectangle of lit pixels is read back (bit-identical)
    L.check(lib.lynx_host_ctx_set_option(ctx, 1, 1))
    e2e_s = time_host()
    assert torch.equal(dense_img, wl.img_host), "sparse read-back differs from the dense image"
    nz = wl.img_host[0].nonzero()
    # bytes of the speculative rectangle the library reads back: the lit rectangle of the
    # previous call grown by 16 pixels on every side, clamped to the image (csrc/host.cu)
    if nz.numel():
        r0, r1 = max(int(nz[:, 0].min()) - 16, 0), min(int(nz[:, 0].max()) + 1 + 16, wl.H)
        c0, c1 = max(int(nz[:, 1].min()) - 16, 0), min(int(nz[:, 1].max()) + 1 + 16, wl.W)
        rect = (r1 - r0) * (c1 - c0)
    else:
        rect = 0
    e2e_val = world * args.particles * E_ARES / e2e_s.item()
    # the screen's edge tables are sent only when they change (first call)
    h2d = 4 * (6 * wl.ld + wl.params_host.numel() + 1)
    d2h = 4 * rect + 16

    def max_over_ranks(x):
        t = torch.tensor([x], device=device, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    def timed_collective(fn, steps, warmup):
        """barrier, per-step events with the L2 flush, MAX over ranks of the mean step (ms)."""
        if world > 1:
            dist.barrier()
        return max_over_ranks(float(np.mean(timed_steps(fn, steps, warmup, flush))))

    # ---- (iv) the exchange step against NCCL, outside any timed region: the fused kernel's
    #      image must equal dist.all_reduce of the ranks' local images bit for bit
    exchange_check, exchange_check_n = None, 0
    if world > 1 and args.exchange == "fused":
        wl.step()  # local image of this rank's shard into wl.img
        want = wl.img.clone()
        dist.all_reduce(want)
        wl.sharded_step(exch)
        torch.cuda.synchronize()
        same = torch.equal(exch.out, want) and want.sum().item() > 0
        if not same:
            sys.stderr.write(f"[rank {rank}] exchange check: {(exch.out != want).sum().item()} pixels differ, "
                             f"sums {exch.out.sum().item()} / {want.sum().item()}\n")
        # ... and over many back-to-back steps with a different beam each (the message that
        # publishes a rank's image is a relaxed system-scope store, exchange.cu: a stale read
        # would drop counts silently, so the sum is checked against NCCL every time)
        n_stress = 0 if args.no_extras else 40
        for k in range(n_stress):
            wl.soa[:, : wl.n].mul_(1.0 + 1e-3 * ((k + rank) % 5))
            wl.step()
            w2 = wl.img.clone()
            dist.all_reduce(w2)
            wl.sharded_step(exch)
            wl.sharded_step(exch)  # twice: both slots, no host synchronisation in between
            torch.cuda.synchronize()
            same = same and torch.equal(exch.out, w2)
        wl.soa.copy_(wl.soa_host, non_blocking=False)
        flag = torch.tensor([1.0 if same else 0.0], device=device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        exchange_check = "bit-exact" if flag.item() == 1.0 else "MISMATCH"
        exchange_check_n = 2 * n_stress + 1

    # ---- (i) the second half of the metric: ParameterBeam env-steps/s, settings sharded over
    #      the ranks (args.settings per GPU; no collective on the path)
    pw = ParamBeamWorkload(args.settings, device, seed=1003 + rank)
    pb_ms = timed_collective(pw.step, args.steps, args.warmup)
    env_steps = {"value": world * args.settings / (pb_ms * 1e-3), "unit": "env-steps/s",
                 "settings_per_gpu": args.settings, "n_gpus": world, "ms_per_step": pb_ms,
                 "scaling": "weak", "sharding": "settings axis, no collective",
                 "workload": "ares_ea_parameterbeam_random_k1_angles (BASELINE.json configs[2])"}
    if not args.no_extras:
        bigB = 1 << 20
        pw2 = ParamBeamWorkload(bigB, device, seed=2003 + rank)
        big_ms = timed_collective(pw2.step, 10, 3)
        env_steps["at_scale"] = {"settings_per_gpu": bigB, "ms_per_step": big_ms,
                                 "value": world * bigB / (big_ms * 1e-3)}
        if rank == 0:
            k6b = kernel_ms(pw2.step, 2, 10, flush, lib)
            a6 = pw2.algorithmic_bytes() / (k6b * 1e-3) / 1e9
            tr6, tr6_src = traffic_of("parambeam_fwd_kernel", "ares_ea", bigB)
            env_steps["roofline"] = {"bound": "hbm", "kernel": "parambeam_fwd_kernel<light> (K6)",
                                     "settings": bigB, "kernel_ms": k6b, "achieved": a6,
                                     "peak": peak, "unit": "GB/s", "frac": a6 / peak,
                                     "traffic": tr6, "traffic_source": tr6_src,
                                     "algorithmic_bytes_per_launch": pw2.algorithmic_bytes()}
        del pw2

    # ---- (ii) BASELINE.json configs[3]: 100 M particles through Cavity + Dipole + quadrupoles
    #      over 8 GPUs = 12.5 M per rank (weak-scaled to the ranks present), screen image
    #      exchanged by the fused kernel
    cfg4 = None
    if not args.no_extras:
        n4 = 12_500_000
        w4 = ParticleWorkload(n4, device, seed=1004 + rank, lattice="cfg4")
        local_ms = timed_collective(w4.step, 10, 3)
        cfg4 = {"particles_per_gpu": n4, "elements": w4.n_elements, "n_gpus": world,
                "local_step_ms": local_ms}
        if world > 1 and args.exchange == "fused":
            ex4 = ScreenExchange(1, w4.W, w4.H, device)
            sh_ms = timed_collective(lambda: w4.sharded_step(ex4), 10, 3)
            w4.step()
            want4 = w4.img.clone()
            dist.all_reduce(want4)
            w4.sharded_step(ex4)
            torch.cuda.synchronize()
            ok4 = torch.tensor([1.0 if torch.equal(ex4.out, want4) else 0.0], device=device)
            dist.all_reduce(ok4, op=dist.ReduceOp.MIN)
            cfg4.update({"step_ms": sh_ms, "efficiency_vs_local_shard": local_ms / sh_ms,
                         "exchange_check": "bit-exact" if ok4.item() == 1.0 else "MISMATCH",
                         "particle_element_updates_per_s": world * n4 * w4.n_elements / (sh_ms * 1e-3)})
            del ex4
        else:
            cfg4["particle_element_updates_per_s"] = world * n4 * w4.n_elements / (local_ms * 1e-3)
        del w4

    # ---- (iii) the same step through the PYTHON surface: host particles (N, 7) ->
    #      lx.ParticleBeam -> Segment.track -> Screen.reading -> host image
    import lynx_b200 as lx
    from lynx_b200 import fixtures as F

    aos_host = torch.from_numpy(F.soa_to_aos7(wl.soa_host[:, :args.particles].numpy())).unsqueeze(0).pin_memory()
    api_seg = F.build_segment(F.ares_ea(k1=K1, vkick=VKICK, hkick=HKICK, screen_active=True,
                                        screen=F.TEST_SCREEN))
    api_energy = torch.tensor([F.ASTRA_ENERGY])

    api_img_host = torch.empty((wl.H, wl.W), dtype=torch.float32).pin_memory()

    def api_step(pinned=True):
        beam = lx.ParticleBeam(aos_host, api_energy)
        api_seg.track(beam)
        reading = api_seg.AREABSCR1.reading
        if not pinned:
            return reading.cpu()  # pageable destination: staged through the driver's bounce buffers
        api_img_host.copy_(reading.reshape(wl.H, wl.W))  # synchronous for a pinned destination
        return api_img_host

    api_ok = torch.equal(api_step(False).reshape(-1), dense_img.reshape(-1)) and \
        torch.equal(api_step(True).reshape(-1), dense_img.reshape(-1))
    api_n = max(3, args.steps // 10)
    api_times = {}
    for pinned in (False, True):
        for _ in range(2):
            api_step(pinned)
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(api_n):
            api_step(pinned)
        api_times[pinned] = max_over_ranks((time.perf_counter() - t0) / api_n)
    api_s = api_times[True]

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (K2) on this workload, rank 0
    k2_ms = kernel_ms(wl.step, 1, args.steps, flush, lib)
    alg = wl.algorithmic_bytes()
    ach = alg / (k2_ms * 1e-3) / 1e9 if k2_ms else None
    traffic, traffic_src = traffic_of("track_pipe_kernel", "ares_screen", args.particles)
    roofline = {"bound": "hbm", "kernel": "track_pipe_kernel<0, false, true, 24> (K2, screen variant)",
                "achieved": ach, "peak": peak, "unit": "GB/s", "frac": (ach / peak) if ach else None,
                # dram__bytes_read.sum + dram__bytes_write.sum per launch of this kernel on this
                # workload (ncu --set full), from the committed table profiles/traffic.json; the
                # image is zeroed and accumulated in L2, its write-back happens after the kernel
                "traffic": traffic, "traffic_source": traffic_src,
                "peak_source": f"{peak_src} (MEASURED_PEAKS.json hbm_gbs, burst copy)",
                "algorithmic_bytes_per_launch": alg, "kernel_ms": k2_ms,
                "frac_of_8000": (ach / 8000.0) if ach else None,
                "regime": "latency / issue-bound at 1 M particles: 3 pipeline tiles per CTA, "
                          "~540 warp instructions per 128 particles at 6 warps per scheduler "
                          "(profiles/r2u_screen32m_lines.txt); 44 MB are 6.8 us at the HBM "
                          "peak.  The same kernel at 32 M particles: extras.at_scale_screen_24B"}
    line = {
        "metric": "particle_element_updates_per_s", "value": value,
        "unit": "particle-element updates/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "ares_ea_1M_particles_screen_2448x2040", "elements": E_ARES,
                   "particles_per_gpu": args.particles, "step": "Segment.track + Screen.reading",
                   "l2": "256 MB flush write before every timed step (outside the events)",
                   "parallelism": f"particle-sharded x{world}" + (
                       f", {exchange_kind}" if world > 1 else "")},
        "clocks": clocks, "gpu_launches": int(launches_timed),
        "e2e": {"value": e2e_val, "unit": "particle-element updates/s",
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": e2e_s.item() * 1e3,
                "path": "lynx_host_track_particles (pinned host buffers; persistent host image: "
                        "only the rectangle of the lit pixels is read back, speculatively, "
                        "without waiting for its bounding box, the host "
                        "image is checked bit-identical to the dense read-back)",
                "dense_readback": {"value": world * args.particles * E_ARES / e2e_dense_s.item(),
                                   "ms_per_step": e2e_dense_s.item() * 1e3,
                                   "d2h_bytes_per_step": 4 * wl.W * wl.H}},
        "roofline": roofline,
        "env_steps": env_steps,
        "api_e2e": {"value": world * args.particles * E_ARES / api_s, "unit": "particle-element updates/s",
                    "ms_per_step": api_s * 1e3, "h2d_bytes_per_step": 28 * args.particles,
                    "d2h_bytes_per_step": 4 * wl.W * wl.H, "image_equals_c_abi_path": bool(api_ok),
                    "path": "pinned host (1, N, 7) particles -> lx.ParticleBeam -> lx.Segment.track "
                            "-> Screen.reading copied into a pinned host image: AoS upload, AoS->SoA "
                            "kernel, K1, K2, dense 20 MB image read-back, all through the Python "
                            "surface",
                    "ms_per_step_reading_cpu": api_times[False] * 1e3,
                    "reading_cpu_note": "the same with Screen.reading.cpu() (a fresh pageable tensor "
                                        "per step)"},
    }
    if exchange_check is not None:
        line["exchange_check"] = exchange_check
        line["exchange_check_exchanges"] = exchange_check_n  # each against dist.all_reduce of the local images
    if cfg4 is not None:
        line["cfg4_100M_sharded"] = cfg4

    if world == 1:
        # ---- CPU baseline: the oracle port on this box's cores, bounded sample
        torch.set_num_threads(os.cpu_count() or 1)
        sec = time_cpu(args.particles, 8, 2)
        line["cpu_baseline"] = {
            "value": args.particles * E_ARES / sec, "unit": "particle-element updates/s",
            "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"8 full steps of {args.particles} particles (torch-CPU restatement)",
            "ms_per_step": sec * 1e3}
        psec = time_cpu_parambeam(args.settings, 20, 2)
        env_steps["cpu_baseline"] = {
            "value": args.settings / psec, "unit": "env-steps/s", "cores": torch.get_num_threads(),
            "kind": "port", "ms_per_step": psec * 1e3,
            "sample": f"20 vectorised Segment.track(ParameterBeam) calls over {args.settings} "
                      "settings (torch-CPU restatement of lynx/accelerator/element.py:71-82)"}

    if world == 1 and not args.no_extras:
        extras = {}
        # ParameterBeam env-steps/s (configs[2])
        pms = [pb_ms]
        for _ in range(3):
            pw.host_step(ctx)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            pw.host_step(ctx)
        pe2e = (time.perf_counter() - t0) / args.steps
        k6 = kernel_ms(pw.step, 2, args.steps, flush, lib)
        extras["parameterbeam_env_steps"] = {
            "settings": args.settings, "value": args.settings / (np.mean(pms) * 1e-3),
            "unit": "env-steps/s", "ms_per_step": float(np.mean(pms)),
            "e2e_value": args.settings / pe2e, "kernel_ms": k6,
            "achieved_gbs": pw.algorithmic_bytes() / (k6 * 1e-3) / 1e9 if k6 else None,
            "regime": "launch-latency-bound (1 MB per launch)"}
        # The same two steps with the OTHER cold-input protocol the bench contract allows: no
        # flush, inputs larger than L2 -- the step rotates over R complete sets of device buffers
        # (beam, image / parameters, moments) whose total exceeds twice the 126 MB L2, so every
        # step's data comes from DRAM while whatever else a tracking loop keeps cached
        # (instructions, page tables) may stay.  Reported to show that the flush is not what the
        # small launches wait for: the two protocols agree within a microsecond.
        def rotating(make, n_sets, steps):
            sets = [make(i) for i in range(n_sets)]
            for w_ in sets:
                w_.step()
            torch.cuda.synchronize()
            evs_ = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
                    for _ in range(steps)]
            for i, (ea, eb) in enumerate(evs_):
                if i % 32 == 0:  # without the flush nothing hides the host: a 3 ms spin kernel
                    torch.cuda.synchronize()   # lets it enqueue the next 32 steps ahead of the device
                    torch.cuda._sleep(int(3e-3 * 1.9e9))
                ea.record()
                sets[i % n_sets].step()
                eb.record()
            torch.cuda.synchronize()
            del sets
            return float(np.mean([ea.elapsed_time(eb) for ea, eb in evs_][n_sets:]))

        rot_n = 8  # 8 x (24 MB beam + 20 MB image) = 352 MB
        rot_step = rotating(lambda i: ParticleWorkload(args.particles, device, seed=3000 + i), rot_n, 20 * rot_n)
        rot_k = 320  # 320 x (0.93 MB moments + 80 KB parameters) = 323 MB
        rot_k6 = rotating(lambda i: ParamBeamWorkload(args.settings, device, seed=4000 + i), rot_k, 3 * rot_k)
        extras["rotating_inputs_no_flush"] = {
            "protocol": "no L2 flush; the step cycles through complete buffer sets larger than "
                        "2 x L2 in total, back to back, per-step events",
            "step_1M_screen": {"sets": rot_n, "ms_per_step": rot_step,
                               "particle_element_updates_per_s": args.particles * E_ARES / (rot_step * 1e-3)},
            "parameterbeam_env_steps": {"sets": rot_k, "settings": args.settings, "ms_per_step": rot_k6,
                                        "value": args.settings / (rot_k6 * 1e-3), "unit": "env-steps/s"}}
        # streaming roofline points at N = scale_particles
        ns = args.scale_particles
        for name, kw in (("track_out_48B", dict(screen=False, want_out=True)),
                         ("track_out_moments_48B", dict(screen=False, want_out=True, moments=1)),
                         ("screen_24B", dict(screen=True))):
            w2 = ParticleWorkload(ns, device, seed=7, **kw)
            km = kernel_ms(w2.step, 1, 10, flush, lib)
            a2 = w2.algorithmic_bytes() / (km * 1e-3) / 1e9
            tr2 = traffic_of("track_pipe_kernel", "ares_screen" if kw.get("screen") else "ares_out", ns)[0]
            extras[f"at_scale_{name}"] = {
                "particles": ns, "kernel_ms": km, "achieved_gbs": a2, "frac": a2 / peak,
                "traffic": tr2 if not kw.get("moments") else None,
                "frac_of_8000": a2 / 8000.0,
                "particle_element_updates_per_s": ns * E_ARES / (km * 1e-3)}
            del w2
        # config 4 per-GPU shard: 12.5 M particles through Cavity + Dipole + quadrupoles
        n4 = 12_500_000
        for name, kw in (("screen_24B", dict(screen=True)),
                         ("out_48B", dict(screen=False, want_out=True))):
            w4 = ParticleWorkload(n4, device, seed=1004, lattice="cfg4", **kw)
            km = kernel_ms(w4.step, 1, 10, flush, lib)
            sm = float(np.mean(timed_steps(w4.step, 10, 3, flush)))
            a4 = w4.algorithmic_bytes() / (km * 1e-3) / 1e9
            extras[f"cfg4_cavity_dipole_12.5M_{name}"] = {
                "particles": n4, "elements": w4.n_elements, "kernel_ms": km, "step_ms": sm,
                "achieved_gbs": a4, "frac": a4 / peak,
                "particle_element_updates_per_s": n4 * w4.n_elements / (sm * 1e-3)}
            del w4
        # K9: beam generation on the device (24 B written per particle)
        import lynx_b200 as lx
        gen = lambda: lx.ParticleBeam.from_parameters(  # noqa: E731
            num_particles=ns, sigma_x=torch.tensor([2e-4]), sigma_y=torch.tensor([2e-4]), seed=1)
        gen()
        gms_ = float(np.mean(timed_steps(gen, 10, 3, flush)))
        extras["generate_beam_24B"] = {
            "particles": ns, "ms": gms_, "achieved_gbs": 24 * ns / (gms_ * 1e-3) / 1e9,
            "frac": 24 * ns / (gms_ * 1e-3) / 1e9 / peak,
            "note": "ParticleBeam.from_parameters end to end (host Cholesky + H2D of 42 floats "
                    "+ kernel), Philox4x32-10 + Box-Muller straight into the SoA planes"}
        # config 5: gradients, one 1M-particle beam x 1024 settings (moment loss)
        gw = GradWorkload(args.particles, 1024, device)
        gms = timed_steps(gw.step, 20, 3, flush)
        fms = timed_steps(gw.forward, 20, 3, flush)
        extras["grad_1M_x_1024_settings"] = {
            "particles": args.particles, "settings": 1024, "fwd_bwd_ms": float(np.mean(gms)),
            "fwd_ms": float(np.mean(fms)),
            "setting_gradients_per_s": 1024 / (np.mean(gms) * 1e-3),
            "particle_setting_element_updates_per_s":
                args.particles * 1024 * E_ARES / (np.mean(gms) * 1e-3),
            "path": "moment collapse: the shared beam is streamed once per call, not per "
                    "setting (DESIGN.md 4.1); K3b dual-number parameter gradients"}
        del gw
        # streaming adjoint K3a at scale: per-particle cotangent + g_in, B = 1
        gw = GradWorkload(ns, 1, device, per_particle=True)
        k3 = kernel_ms(gw.backward, 3, 10, flush, lib)
        k3_bytes = 72 * ns  # x 24 B + g_out 24 B + g_in 24 B per particle
        extras["at_scale_adjoint_K3a_72B"] = {
            "particles": ns, "kernel_ms": k3, "achieved_gbs": k3_bytes / (k3 * 1e-3) / 1e9,
            "frac": k3_bytes / (k3 * 1e-3) / 1e9 / peak}
        del gw
        # headline workload with a persistent screen image (only the previously lit rectangle
        # is cleared): same image, 20 MB less to zero per step
        try:
            wp = ParticleWorkload(args.particles, device, seed=1002, screen=True)
            wp.step()
            wp.persistent_step()
            assert torch.equal(wp.img, wp._pimg), "persistent image differs from the fresh one"
            pms_ = float(np.mean(timed_steps(wp.persistent_step, args.steps, args.warmup, flush)))
            fms_ = float(np.mean(timed_steps(wp.step, args.steps, args.warmup, flush)))
            extras["persistent_image_step"] = {
                "particles": args.particles, "ms_per_step": pms_, "fresh_image_ms_per_step": fms_,
                "particle_element_updates_per_s": args.particles * E_ARES / (pms_ * 1e-3)}
            del wp
        except Exception as e:  # noqa: BLE001
            extras["persistent_image_error"] = repr(e)
        # beam-level kernels at scale: K10 transformed_to (48 B),
        # K11 Aperture compaction (56 B over three launches), K9b ellipsoid generation (24 B)
        try:
            beam = lx.ParticleBeam.from_parameters(
                num_particles=ns, sigma_x=torch.tensor([2e-4]), sigma_y=torch.tensor([2e-4]), seed=2)
            beam.moment_sums()
            tf = lambda: beam.transformed_to(sigma_x=torch.tensor([1e-4]))  # noqa: E731
            k10 = kernel_ms(tf, 4, 10, flush, lib)
            extras["at_scale_transformed_to_48B"] = {
                "particles": ns, "kernel_ms": k10, "achieved_gbs": 48 * ns / (k10 * 1e-3) / 1e9,
                "frac": 48 * ns / (k10 * 1e-3) / 1e9 / peak}
            ap = lx.Aperture(x_max=torch.tensor(2e-4), y_max=torch.tensor(2e-4), shape="elliptical")
            apf = lambda: ap.track(beam)  # noqa: E731
            apf()
            ams = float(np.mean(timed_steps(apf, 10, 3, flush)))
            k11 = kernel_ms(apf, 5, 10, flush, lib)
            extras["at_scale_aperture_56B"] = {
                "particles": ns, "ms_three_launches_and_count_readback": ams,
                "scatter_kernel_ms": k11,
                "achieved_gbs": 56 * ns / (ams * 1e-3) / 1e9,
                "frac": 56 * ns / (ams * 1e-3) / 1e9 / peak,
                "scatter_achieved_gbs": 48 * ns / (k11 * 1e-3) / 1e9 if k11 else None}
            del beam
            ell = lambda: lx.ParticleBeam.uniform_3d_ellipsoid(  # noqa: E731
                num_particles=ns, radius_x=torch.tensor([1e-3]), seed=3)
            ell()
            ems = float(np.mean(timed_steps(ell, 10, 3, flush)))
            extras["generate_ellipsoid_24B"] = {
                "particles": ns, "ms": ems, "achieved_gbs": 24 * ns / (ems * 1e-3) / 1e9,
                "frac": 24 * ns / (ems * 1e-3) / 1e9 / peak,
                "note": "per-particle Philox rejection (1.91 draws per particle) + 3 normals"}
        except Exception as e:  # noqa: BLE001 -- extras never take the headline line down
            extras["beamops_error"] = repr(e)
        line["extras"] = extras

    lib.lynx_host_ctx_destroy(ctx)
    # NCCL writes its version banner through C stdio, which is block-buffered on a pipe and
    # would otherwise come out after this line at exit: the JSON must be the LAST line
    try:
        C.CDLL(None).fflush(None)
    except Exception:
        pass
    emit(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


_real_stdout = None


def emit(text):
    """The ONE JSON line goes to the process's original stdout; everything else that writes to
    fd 1 meanwhile (NCCL prints its version banner there) has been pointed at stderr."""
    if _real_stdout is None:
        print(text, flush=True)
    else:
        os.write(_real_stdout, (text + "\n").encode())


def main():
    global _real_stdout
    sys.stdout.flush()
    _real_stdout = os.dup(1)
    os.dup2(2, 1)
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()

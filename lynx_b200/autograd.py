"""Gradients through tracking: the ``custom_vjp`` of the north star, as
``torch.autograd.Function`` (JAX is absent from this image).

The reference has no custom vjp: its gradients come from the array
framework differentiating segment.py:340-356 op by op.  Here forward is one call of
``lynx_track_particles_fwd`` / ``lynx_parambeam_fwd`` and backward one call of
``lynx_track_particles_bwd`` / ``lynx_parambeam_bwd`` (K3a streaming adjoint + K3b
dual-number parameter gradients, csrc/adjoint.cu).  Gradients reach the user's element
tensors (``segment.AREAMQZM1.k1`` ...) through the differentiable ``pack_params``.
"""

from dataclasses import dataclass
from typing import Optional

import torch

from . import _lib as L


@dataclass
class TrackCfg:
    prog: object
    B: int
    shared: bool
    ld: int
    N: int
    energy_in: torch.Tensor
    want_out: bool
    moment_mode: int
    screen: Optional[object]
    batch_shape: tuple


def track_particles_raw(cfg: TrackCfg, params, soa):
    """One call of ``lynx_track_particles_fwd``; returns the raw device outputs."""
    prog, B, ld, N, bshape = cfg.prog, cfg.B, cfg.ld, cfg.N, cfg.batch_shape
    device = soa.device
    out = (torch.empty((*bshape, 6, ld), dtype=torch.float32, device=device)
           if cfg.want_out else None)
    bpm = (torch.empty((prog.n_bpm, B, 2), dtype=torch.float64, device=device)
           if prog.n_bpm else None)
    img = ex = ey = None
    W = H = 0
    if cfg.screen is not None:
        ex, ey = cfg.screen._device_edges(device)
        W, H = ex.numel() - 1, ey.numel() - 1
        img = torch.empty((*bshape, H, W), dtype=torch.float32, device=device)
    mom = (torch.empty((*bshape, 28), dtype=torch.float64, device=device)
           if cfg.moment_mode else None)
    e_out = torch.empty(bshape, dtype=torch.float32, device=device)
    ws = torch.empty(prog.workspace_bytes(B), dtype=torch.uint8, device=device)
    L.check(L.lib().lynx_track_particles_fwd(
        prog.handle, params.data_ptr(), B, soa.data_ptr(), 0 if cfg.shared else 1, ld, N,
        L.ptr(out), ld, L.ptr(bpm), L.ptr(img), W, H, L.ptr(ex), L.ptr(ey), L.ptr(mom),
        cfg.moment_mode, cfg.energy_in.data_ptr(), e_out.data_ptr(), ws.data_ptr(),
        L.current_stream()), "lynx_track_particles_fwd")
    return out, bpm, img, mom, e_out


class TrackParticlesFn(torch.autograd.Function):
    """vjp of Segment.track(ParticleBeam) w.r.t. the packed parameters and the beam."""

    @staticmethod
    def forward(ctx, params, soa, cfg: TrackCfg):
        params, soa = params.detach(), soa.detach()
        out, bpm, img, mom, e_out = track_particles_raw(cfg, params, soa)
        ctx.cfg = cfg
        ctx.save_for_backward(params, soa)
        nd = [t for t in (img, e_out) if t is not None]
        ctx.mark_non_differentiable(*nd)
        return out, bpm, img, mom, e_out

    @staticmethod
    def backward(ctx, g_out, g_bpm, g_img, g_mom, g_eout):
        cfg = ctx.cfg
        params, soa = ctx.saved_tensors
        prog, B, ld, N = cfg.prog, cfg.B, cfg.ld, cfg.N
        device = soa.device
        lib = L.lib()
        g_params = torch.empty_like(params)
        want_gin = ctx.needs_input_grad[1]
        g_in = torch.empty_like(soa) if want_gin else None
        if g_out is not None:
            g_out = g_out.to(torch.float32).contiguous()  # padding lanes (>= N) are never read
        g_mom = g_mom.to(torch.float64).contiguous() if g_mom is not None else None
        g_bpm = g_bpm.to(torch.float64).contiguous() if g_bpm is not None else None
        ws = torch.empty(lib.lynx_bwd_workspace_bytes(prog.handle, B), dtype=torch.uint8,
                         device=device)
        L.check(lib.lynx_track_particles_bwd(
            prog.handle, params.data_ptr(), B, soa.data_ptr(), 0 if cfg.shared else 1, ld, N,
            L.ptr(g_out), ld, L.ptr(g_mom), L.ptr(g_bpm), cfg.energy_in.data_ptr(),
            g_params.data_ptr(), L.ptr(g_in), ld, ws.data_ptr(), L.current_stream()),
            "lynx_track_particles_bwd")
        return g_params, g_in, None


@dataclass
class ParamBeamCfg:
    prog: object
    B: int
    shared: bool
    energy_in: torch.Tensor
    batch_shape: tuple


def parambeam_raw(cfg: ParamBeamCfg, params, mu_in, cov_in):
    prog, B, bshape = cfg.prog, cfg.B, cfg.batch_shape
    device = mu_in.device
    mu_out = torch.empty((*bshape, 7), dtype=torch.float32, device=device)
    cov_out = torch.empty((*bshape, 7, 7), dtype=torch.float32, device=device)
    bpm = (torch.empty((prog.n_bpm, B, 2), dtype=torch.float32, device=device)
           if prog.n_bpm else None)
    e_out = torch.empty(bshape, dtype=torch.float32, device=device)
    L.check(L.lib().lynx_parambeam_fwd(
        prog.handle, params.data_ptr(), B, mu_in.data_ptr(), cov_in.data_ptr(),
        0 if cfg.shared else 1, cfg.energy_in.data_ptr(), mu_out.data_ptr(), cov_out.data_ptr(),
        L.ptr(bpm), e_out.data_ptr(), L.current_stream()), "lynx_parambeam_fwd")
    return mu_out, cov_out, bpm, e_out


class ParamBeamFn(torch.autograd.Function):
    """vjp of Segment.track(ParameterBeam)."""

    @staticmethod
    def forward(ctx, params, mu_in, cov_in, cfg: ParamBeamCfg):
        params, mu_in, cov_in = params.detach(), mu_in.detach(), cov_in.detach()
        mu_out, cov_out, bpm, e_out = parambeam_raw(cfg, params, mu_in, cov_in)
        ctx.cfg = cfg
        ctx.save_for_backward(params, mu_in, cov_in)
        ctx.mark_non_differentiable(e_out)
        return mu_out, cov_out, bpm, e_out

    @staticmethod
    def backward(ctx, g_mu, g_cov, g_bpm, g_eout):
        cfg = ctx.cfg
        params, mu_in, cov_in = ctx.saved_tensors
        g_params = torch.empty_like(params)
        beam_grads = ctx.needs_input_grad[1] or ctx.needs_input_grad[2]
        g_mu_in = torch.empty_like(mu_in) if beam_grads else None
        g_cov_in = torch.empty_like(cov_in) if beam_grads else None
        c = lambda t: t.to(torch.float32).contiguous() if t is not None else None  # noqa: E731
        g_mu, g_cov, g_bpm = c(g_mu), c(g_cov), c(g_bpm)
        L.check(L.lib().lynx_parambeam_bwd(
            cfg.prog.handle, params.data_ptr(), cfg.B, mu_in.data_ptr(), cov_in.data_ptr(),
            0 if cfg.shared else 1, cfg.energy_in.data_ptr(), L.ptr(g_mu), L.ptr(g_cov),
            L.ptr(g_bpm), g_params.data_ptr(), L.ptr(g_mu_in), L.ptr(g_cov_in),
            L.current_stream()), "lynx_parambeam_bwd")
        return g_params, g_mu_in, g_cov_in, None


def needs_grad(*tensors) -> bool:
    return torch.is_grad_enabled() and any(t is not None and t.requires_grad for t in tensors)

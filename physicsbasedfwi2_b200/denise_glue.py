"""What the reference's elastic `prop` methods do around `d.grad(model_init, src, rec)` (DENISE), as tensors in /
tensors out on the B200 (reference models/networks.py:8936-9256, AutoElFullRhoMarmousiMar22_Net.prop).

The reference writes the three models to DENISE's binary files, copies the SU seismograms of six shots into place,
runs `mpirun -np 30 denise` (elastic forward, L2 residual of the low-passed seismograms, adjoint, correlation) and reads
the vp / vs / rho gradients back (:9173-9188); then it post-processes them in numpy (:9194-9240).  Here:

    elastic_inner_gradient(vp1, vs1, rho1, observed, x_s, x_r, ...)  ->  (vp_grad, vs_grad, rho_grad, loss)

* forward + adjoint: physicsbasedfwi2_b200.elastic.Propagator (csrc/elastic.cu), explosive Ricker sources of
  `fsource` Hz, two-component receivers;
* the frequency stage `d.add_fwi_stage(fc_low=0.0, fc_high=10.0, ...)` (:9147): a zero-phase Butterworth low-pass
  on the synthetic and the observed seismograms (csrc/glue.cu; the operator is symmetric, so its backward is itself);
* misfit: 0.5 sum (filtered residual)^2 (DENISE's L2 norm);
* the post-processing of :9194-9240 (z^1.5 weight on the rho gradient, 24 zeroed water rows, scaling of each gradient
  to max|model| / max|gradient|, 0.5 and 1e-10 on rho) in two fused passes (csrc/glue.cu).

DENISE's model files are stored bottom-up, hence the `np.flipud` pairs at :8964-8978 and :9194-9196; with tensors in
and out they cancel and do not appear here.  The reference keeps its models in units of 100 m/s (`*100.0` at
:8946-8948, :8976-8978): `unit` is that factor.  SU files are read and written by `read_su` / `write_su`.
CUDA only - no CPU path.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import numpy as np
import torch

from . import _lib
from .wavelets import ricker


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


class _Lowpass(torch.autograd.Function):
    """Butterworth low-pass along dim 0 of [nt, ...] (csrc/glue.cu).  Zero-phase (forward + backward pass, zero initial
    state) makes the operator symmetric, so the backward of the zero-phase filter is the filter."""

    @staticmethod
    def forward(ctx, x, dt, fc_high, order, zero_phase):
        if not x.is_cuda or x.dtype != torch.float32:
            raise RuntimeError("lowpass needs a float32 CUDA tensor: there is no CPU path")
        ctx.args = (float(dt), float(fc_high), int(order), bool(zero_phase))
        return _Lowpass._apply(x.detach(), *ctx.args, reverse=False)

    @staticmethod
    def _apply(x, dt, fc_high, order, zero_phase, reverse):
        xc = x.contiguous()
        if reverse:                      # adjoint of the causal filter: the same filter run backwards in time
            xc = xc.flip(0).contiguous()
        nt = int(xc.shape[0])
        out = torch.empty_like(xc)
        d = _lib.ButterDesc(nt, xc.numel() // max(nt, 1), order, 1 if zero_phase else 0, dt, 0.0, fc_high)
        cur = torch.cuda.current_stream(xc.device)
        with torch.cuda.device(xc.device):
            _lib.check(_lib.lib().fwi_butterworth_lowpass(C.byref(d), _p(xc), _p(out), C.c_void_p(cur.cuda_stream)),
                       "fwi_butterworth_lowpass")
        return out.flip(0) if reverse else out

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g):
        dt, fc_high, order, zero_phase = ctx.args
        return _Lowpass._apply(g, dt, fc_high, order, zero_phase, reverse=not zero_phase), None, None, None, None


def lowpass(x: torch.Tensor, dt: float, fc_high: float, order: int = 6, zero_phase: bool = True) -> torch.Tensor:
    """The frequency stage of `d.add_fwi_stage(fc_low=0.0, fc_high=...)` (models/networks.py:9147) on traces
    [nt, ...] sampled at `dt`: Butterworth low-pass of even `order`, differentiable."""
    return _Lowpass.apply(x, dt, fc_high, order, zero_phase)


def gradient_post(grad_vp: torch.Tensor, grad_vs: torch.Tensor, grad_rho: torch.Tensor, vp: torch.Tensor, vs: torch.Tensor,
                  rho: torch.Tensor, *, water_rows: int = 24, depth_pow: float = 1.5, rho_factor: float = 0.5,
                  rho_offset: float = 1e-10) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """models/networks.py:9194-9240 on [nz, ny] CUDA tensors (row 0 = surface): returns new (vp_grad, vs_grad, rho_grad)."""
    for t in (grad_vp, grad_vs, grad_rho, vp, vs, rho):
        if not t.is_cuda or t.dtype != torch.float32 or t.shape != grad_vp.shape or t.dim() != 2:
            raise RuntimeError("gradient_post needs six float32 CUDA tensors [nz, ny] of one shape")
    out = [t.detach().clone().contiguous() for t in (grad_vp, grad_vs, grad_rho)]
    mods = [t.detach().contiguous() for t in (vp, vs, rho)]
    stats = torch.empty(6, dtype=torch.float32, device=grad_vp.device)
    d = _lib.GradPostDesc(int(grad_vp.shape[0]), int(grad_vp.shape[1]), int(water_rows), float(depth_pow), float(rho_factor),
                          float(rho_offset))
    cur = torch.cuda.current_stream(grad_vp.device)
    with torch.cuda.device(grad_vp.device):
        _lib.check(_lib.lib().fwi_elastic_gradient_post(C.byref(d), _p(out[0]), _p(out[1]), _p(out[2]), _p(mods[0]), _p(mods[1]),
                                                        _p(mods[2]), _p(stats), C.c_void_p(cur.cuda_stream)),
                   "fwi_elastic_gradient_post")
    return out[0], out[1], out[2]


def denise_geometry(shot_idx, *, dx: float = 20.0, depth_rec: float = 460.0, xrec1: float = 400.0, xrec2: float = 7780.0,
                    drec: float = 20.0, depth_src: float = 40.0, xsrc1: float = 400.0, xsrc2: float = 7700.0,
                    dsrc: float = 160.0) -> Tuple[torch.Tensor, torch.Tensor]:
    """The acquisition of models/networks.py:8994-9012: 370 receivers at 460 m depth every 20 m, 46 sources at 40 m
    depth every 160 m of which `shot_idx` (the reference takes idx[0:6], :9032) are used.  Returns x_s [S, 1, 2],
    x_r [S, R, 2] in metres as (z, y)."""
    xrec = np.arange(xrec1, xrec2 + dx, drec)
    xsrc = np.arange(xsrc1, xsrc2 + dx, dsrc)[np.asarray(shot_idx, dtype=np.int64)]
    S, R = len(xsrc), len(xrec)
    x_s = torch.zeros(S, 1, 2)
    x_s[:, 0, 0] = depth_src
    x_s[:, 0, 1] = torch.from_numpy(xsrc).float()
    x_r = torch.zeros(S, R, 2)
    x_r[:, :, 0] = depth_rec
    x_r[:, :, 1] = torch.from_numpy(xrec).float()[None, :]
    return x_s, x_r


def elastic_inner_gradient(vp1: torch.Tensor, vs1: torch.Tensor, rho1: torch.Tensor, observed: torch.Tensor,
                           x_s: torch.Tensor, x_r: torch.Tensor, *, dx: float = 20.0, dt: float = 2e-3, fsource: float = 5.0,
                           fc_high: Optional[float] = 10.0, filter_order: int = 6, unit: float = 100.0, water_rows: int = 24,
                           source_scale: float = 1.0, postprocess: bool = True, **propagator_kwargs):
    """What `prop` returns at models/networks.py:9255: (vp_grad, vs_grad, rho_grad, loss) for the current models.

    vp1, vs1, rho1: [nz, ny] (or [1, 1, nz, ny]) CUDA tensors in the reference's units (`unit` = 100 -> m/s, kg/m^3),
    row 0 at the surface.  observed [2, nt, S, R]: recorded v_y / v_z seismograms of the chosen shots (the SU files the
    reference copies at :9051-9071), sampled at `dt`; nt * dt is DENISE's TIME (6 s at :9082).  Sources: explosive
    Ricker wavelets of `fsource` Hz (:9017,9019)."""
    from .elastic import Propagator
    if not vp1.is_cuda:
        raise RuntimeError("elastic_inner_gradient needs CUDA tensors: there is no CPU path")
    dev = vp1.device
    models = [(t.detach().reshape(t.shape[-2], t.shape[-1]).to(torch.float32) * unit).contiguous().requires_grad_(True)
              for t in (vp1, vs1, rho1)]
    if observed.dim() != 4 or observed.shape[0] != 2:
        raise RuntimeError("observed must be [2, nt, n_shots, n_receivers] (v_y, v_z)")
    nt, S = int(observed.shape[1]), int(observed.shape[2])
    src = (ricker(fsource, nt, dt, 1.0 / fsource) * source_scale).reshape(-1, 1, 1).repeat(1, S, 1).contiguous().to(dev)
    prop = Propagator({"vp": models[0], "vs": models[1], "rho": models[2]}, dx, **propagator_kwargs)
    rec = prop(src, x_s.to(dev), x_r.to(dev), dt)                     # [2, nt, S, R]
    obs = observed.to(dev, torch.float32)
    if fc_high is not None:
        # one filter call over both components: time is dim 0 of [nt, 2 * S * R]
        rec_f = lowpass(rec.permute(1, 0, 2, 3).contiguous(), dt, fc_high, filter_order).permute(1, 0, 2, 3)
        with torch.no_grad():
            obs = lowpass(obs.permute(1, 0, 2, 3).contiguous(), dt, fc_high, filter_order).permute(1, 0, 2, 3)
    else:
        rec_f = rec
    loss = 0.5 * ((rec_f - obs) ** 2).sum()
    loss.backward()
    grads = [m.grad for m in models]
    if postprocess:
        grads = gradient_post(grads[0], grads[1], grads[2], models[0].detach(), models[1].detach(), models[2].detach(),
                              water_rows=water_rows)
    return grads[0], grads[1], grads[2], loss.detach()


# ---- SU seismograms (the files the reference shuffles at models/networks.py:9051-9071) ------------------------------
def read_su(path: str) -> Tuple[np.ndarray, float]:
    """Reads a Seismic-Unix file written by DENISE (native-endian IEEE floats, 240-byte trace headers) ->
    (traces [n_traces, ns] float32, dt in seconds)."""
    raw = np.fromfile(path, dtype=np.uint8)
    if raw.size < 240:
        raise RuntimeError(f"{path}: not an SU file")
    ns = int(raw[114:116].view(np.uint16)[0])
    dt_us = int(raw[116:118].view(np.uint16)[0])
    rec = 240 + 4 * ns
    if ns == 0 or raw.size % rec:
        raise RuntimeError(f"{path}: size {raw.size} is not a multiple of the trace record ({rec} bytes for ns = {ns})")
    traces = raw.reshape(-1, rec)[:, 240:].copy().view(np.float32)
    return traces, dt_us * 1e-6


def write_su(path: str, traces: np.ndarray, dt: float) -> None:
    """Writes traces [n_traces, ns] with the header words DENISE sets that a reader needs (tracl, ns, dt)."""
    tr = np.ascontiguousarray(traces, dtype=np.float32)
    n, ns = tr.shape
    out = np.zeros((n, 240 + 4 * ns), dtype=np.uint8)
    out[:, 0:4] = (np.arange(n, dtype=np.int32) + 1).view(np.uint8).reshape(n, 4)
    out[:, 114:116] = np.full(n, ns, dtype=np.uint16).view(np.uint8).reshape(n, 2)
    out[:, 116:118] = np.full(n, int(round(dt * 1e6)), dtype=np.uint16).view(np.uint8).reshape(n, 2)
    out[:, 240:] = tr.view(np.uint8).reshape(n, 4 * ns)
    out.tofile(path)

"""Trace misfit of the physics-guided `prop` methods, fused with its backward on the GPU.

The reference computes, on the propagator's output and before ``lossinner.backward()``
(models/networks.py:5467-5476; the observed data get the same normalisation once at :5418-5419):

    pred  = pred - receiver_amplitudes_cte                    # direct wave of a 1500 m/s model
    pmax  = |pred|.max(dim=0, keepdim=True)
    norm  = pred / (pmax.abs() + 1e-10)
    loss  = torch.nn.L1Loss()(norm, observed_norm)            # MSELoss in other nets

as ~10 torch kernels forward and as many backward, each a pass over rec[nt, S, R].
``trace_misfit`` does the same arithmetic in three passes (csrc/misfit.cu) and hands the adjoint
its ``grad_rec`` directly.  CUDA only - there is no CPU path.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib

_KINDS = {"l1": 0, "mse": 1}


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


class _TraceMisfit(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred, direct, observed, kind, normalize, eps):
        if not pred.is_cuda:
            raise RuntimeError("trace_misfit needs CUDA tensors: there is no CPU path")
        if pred.dim() != 3 or pred.dtype != torch.float32:
            raise RuntimeError("pred must be float32 [nt, n_shots, n_receivers]")
        for name, t in (("direct", direct), ("observed", observed)):
            if t is not None and (t.shape != pred.shape or t.dtype != torch.float32 or t.device != pred.device):
                raise RuntimeError(f"{name} must match pred in shape, dtype and device")
        if kind not in _KINDS:
            raise RuntimeError("kind must be 'l1' or 'mse'")
        nt, M = int(pred.shape[0]), int(pred.shape[1] * pred.shape[2])
        if nt * M == 0:
            raise RuntimeError("empty traces")
        p, d, o = pred.detach().contiguous(), None if direct is None else direct.detach().contiguous(), \
            observed.detach().contiguous()
        desc = _lib.MisfitDesc(nt, M, _KINDS[kind], 1 if normalize else 0, float(eps))
        L = _lib.lib()
        ws = torch.empty(int(L.fwi_trace_misfit_workspace_bytes(C.byref(desc))), dtype=torch.uint8, device=pred.device)
        loss = torch.empty((), dtype=torch.float32, device=pred.device)
        cur = torch.cuda.current_stream(pred.device)
        with torch.cuda.device(pred.device):
            _lib.check(L.fwi_trace_misfit(C.byref(desc), _p(p), _p(d), _p(o), _p(loss), _p(ws),
                                          C.c_void_p(cur.cuda_stream)), "fwi_trace_misfit")
        ctx.desc, ctx.ws = desc, ws
        ctx.tensors = (p, d, o)
        return loss

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, grad_loss):
        p, d, o = ctx.tensors
        g = torch.empty_like(p)
        gl = grad_loss.to(p.device, torch.float32).contiguous()
        cur = torch.cuda.current_stream(p.device)
        with torch.cuda.device(p.device):
            _lib.check(_lib.lib().fwi_trace_misfit_backward(C.byref(ctx.desc), _p(p), _p(d), _p(o), _p(gl), _p(g),
                                                            _p(ctx.ws), C.c_void_p(cur.cuda_stream)),
                       "fwi_trace_misfit_backward")
        ctx.tensors = ctx.ws = None
        return g, None, None, None, None, None


def trace_misfit(pred: torch.Tensor, observed: torch.Tensor, direct: Optional[torch.Tensor] = None, *,
                 kind: str = "l1", normalize: bool = True, eps: float = 1e-10) -> torch.Tensor:
    """Scalar misfit of predicted traces [nt, S, R] against (already normalised) observed traces,
    differentiable w.r.t. ``pred``: subtract ``direct`` (optional), divide each trace by its
    max |.| + eps (``normalize``), then mean |.| (``kind='l1'``) or mean square (``'mse'``)."""
    return _TraceMisfit.apply(pred, direct, observed, kind, normalize, eps)


def normalize_traces(rec: torch.Tensor, eps: float = 1e-10) -> torch.Tensor:
    """The once-per-call normalisation of the observed data (models/networks.py:5418-5419)."""
    return rec / (rec.abs().amax(dim=0, keepdim=True) + eps)


class DirectWave:
    """The direct-wave shot of the reference (``receiver_amplitudes_cte``, models/networks.py:5408-5411):
    a forward run on a constant 1500 m/s model with the survey's geometry.  The reference recomputes
    it on every call of ``prop`` although nothing it depends on changes between iterations (a third
    of its propagator time in Auto22); this object recomputes it only when the
    geometry, the wavelet or the grid changes."""

    def __init__(self, velocity: float = 1500.0):
        self.velocity = float(velocity)
        self._key = None
        self._val = None

    def __call__(self, model_shape, dx, source_amplitudes, x_s, x_r, dt, device=None, **propagator_kwargs):
        from .scalar import Propagator
        dev = torch.device(device) if device is not None else source_amplitudes.device
        key = (tuple(model_shape), str(dx), float(dt), str(dev), tuple(sorted(propagator_kwargs.items())))
        same = (self._key is not None and self._key[0] == key
                and all(a.shape == b.shape and torch.equal(a, b.to(a.device))
                        for a, b in zip(self._key[1], (source_amplitudes, x_s, x_r))))
        if not same:
            const = torch.full(tuple(model_shape), self.velocity, dtype=torch.float32, device=dev)
            with torch.no_grad():
                self._val = Propagator({"vp": const}, dx, **propagator_kwargs)(
                    source_amplitudes.detach().to(dev), x_s.to(dev), x_r.to(dev), dt).contiguous()
            self._key = (key, tuple(t.detach().clone() for t in (source_amplitudes, x_s, x_r)))
        return self._val

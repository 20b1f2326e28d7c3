"""Devito-style FWI objective on the B200 propagator.

The reference's unused ``seisgan`` tree wraps a Devito acoustic solver as an autograd Function
(seisgan/fwi/layers.py:145-201): ``forward(x)`` takes a velocity image ``x[1, 1, n1, n2]``, runs all
shots, returns the objective ``sum_shots 0.5 ||d_pred - d_true||^2`` as a 1-element tensor and keeps
the gradient with respect to the squared slowness m = 1/v^2 (the ``grad -= u.dt2 * v`` imaging
condition of seisgan/fwi/pde/seismic/acoustic/operators.py:127-165); ``backward`` ignores the
incoming gradient and returns that image divided by its max |.| (layers.py:194-197).

Same contract here, computed by the CUDA propagator (all shots in one batched call instead of a
serial loop of Devito runs).  What differs by design: the absorbing boundary is the PML of
scalar.Propagator, not Devito's sponge layer, so amplitudes near the edges differ from a Devito run;
the interior physics (2nd order in time, centred differences in space, point sources/receivers)
is the same.  CUDA only.
"""
from __future__ import annotations

import torch

from .scalar import Propagator


class _FWILoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, observed, source_amplitudes, x_s, x_r, dx, dt, normalize, kwargs):
        if x.dim() != 4 or x.shape[0] != 1 or x.shape[1] != 1:
            raise RuntimeError("x must be a velocity image of shape [1, 1, nz, ny]")
        with torch.enable_grad():
            v = x.detach()[0, 0].contiguous().requires_grad_(True)
            rec = Propagator({"vp": v}, dx, **kwargs)(source_amplitudes, x_s, x_r, dt)
            objective = 0.5 * ((rec - observed) ** 2).sum()
            (grad_v,) = torch.autograd.grad(objective, v)
        # chain rule to the squared slowness m = v^-2:  dv/dm = -v^3 / 2
        grad_m = grad_v * (-0.5 * v.detach() ** 3)
        if normalize:
            grad_m = grad_m / grad_m.abs().max().clamp_min(1e-30)
        ctx.save_for_backward(grad_m)
        return objective.detach().reshape(1)

    @staticmethod
    def backward(ctx, grad_output):
        (grad_m,) = ctx.saved_tensors
        return (grad_m[None, None],) + (None,) * 8


def fwi_loss(x: torch.Tensor, observed: torch.Tensor, source_amplitudes: torch.Tensor, x_s: torch.Tensor,
             x_r: torch.Tensor, dx, dt: float, *, normalize: bool = True, **propagator_kwargs) -> torch.Tensor:
    """Objective [1] for the velocity image ``x[1, 1, nz, ny]`` (m/s); ``backward`` hands back the
    gradient with respect to the squared slowness, max-normalised as in seisgan/fwi/layers.py:194-197
    (``normalize=False`` keeps its scale).  ``observed`` [nt, S, R] are the recorded shots."""
    return _FWILoss.apply(x, observed, source_amplitudes, x_s, x_r, dx, float(dt), normalize, propagator_kwargs)

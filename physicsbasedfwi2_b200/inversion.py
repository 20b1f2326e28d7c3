"""Inner-loop FWI driver: the reference's heaviest user of the propagator.

models/networks.py:15340-15415 runs, inside every training step of the `*Phy` nets, `num_epochs`
sweeps over `num_batches` strided shot batches of:  clamp the velocity -> Propagator -> subtract the
direct wave -> normalise each trace by its max |.| -> MSE (L1 in other nets) against the equally
normalised observed data -> backward -> Adam(lr=10) step on the velocity image itself (optionally a
depth weight / water mask on the gradient, :5492-5493).  Here the same loop on the B200 kernels: the
direct wave is computed once (misfit.DirectWave), the misfit and its backward are one fused call
(misfit.trace_misfit), and on the reference's grid sizes each gradient is two kernel launches
(csrc/acoustic_small.cu).  CUDA only.
"""
from __future__ import annotations

from typing import Callable, List, Optional

import torch

from .misfit import DirectWave, normalize_traces, trace_misfit
from .scalar import Propagator


def inner_loop_fwi(velocity: torch.Tensor, observed: torch.Tensor, source_amplitudes: torch.Tensor,
                   x_s: torch.Tensor, x_r: torch.Tensor, dx, dt: float, *, num_epochs: int = 100,
                   num_batches: int = 1, lr: float = 10.0, vmin: float = 1500.0, vmax: float = 4400.0,
                   kind: str = "mse", subtract_direct: bool = True, grad_weight: Optional[torch.Tensor] = None,
                   grad_mask: Optional[torch.Tensor] = None, shuffle: bool = False,
                   callback: Optional[Callable[[int, int, float], None]] = None, **propagator_kwargs):
    """Runs the loop of models/networks.py:15372-15415 and returns (velocity, losses).

    velocity [nz, ny] (CUDA fp32; updated by Adam, returned as a new tensor), observed [nt, S, R] raw
    recorded traces, source_amplitudes [nt, 1, 1] (one wavelet for every shot, as in the reference) or
    [nt, S, Ns], x_s [S, Ns, 2], x_r [S, R, 2].  `grad_weight` multiplies the gradient (the reference's
    depth weight `ss`), `grad_mask` (bool) zeroes it (water layer).  `shuffle` permutes the shots every
    epoch (:5434-5440).  `losses` holds the misfit of every (epoch, batch)."""
    if not velocity.is_cuda:
        raise RuntimeError("inner_loop_fwi needs CUDA tensors: there is no CPU path")
    dev = velocity.device
    S = int(x_s.shape[0])
    if num_batches < 1 or num_batches > S:
        raise RuntimeError(f"num_batches must be in [1, {S}]")
    src = source_amplitudes.to(dev, torch.float32)
    if src.shape[1] == 1 and S > 1:
        src = src.repeat(1, S, 1)
    src = src.contiguous()
    x_s, x_r = x_s.to(dev), x_r.to(dev)
    obs = observed.to(dev, torch.float32)
    direct = None
    if subtract_direct:
        direct = DirectWave()(tuple(velocity.shape), dx, src, x_s, x_r, dt, device=dev, **propagator_kwargs)
        obs = obs - direct
    obs_norm = normalize_traces(obs)

    v = velocity.detach().clone().requires_grad_(True)
    opt = torch.optim.Adam([v], lr=lr)
    losses: List[float] = []
    order = torch.arange(S, device=dev)
    for epoch in range(num_epochs):
        if shuffle:
            order = torch.randperm(S, device=dev)
        for it in range(num_batches):
            idx = order[it::num_batches]
            opt.zero_grad(set_to_none=True)
            model = torch.clamp(v, min=vmin, max=vmax)
            pred = Propagator({"vp": model}, dx, **propagator_kwargs)(src[:, idx].contiguous(), x_s[idx], x_r[idx], dt)
            loss = trace_misfit(pred, obs_norm[:, idx].contiguous(),
                                None if direct is None else direct[:, idx].contiguous(), kind=kind)
            loss.backward()
            if grad_weight is not None:
                v.grad.mul_(grad_weight)
            if grad_mask is not None:
                v.grad[grad_mask] = 0
            opt.step()
            losses.append(float(loss.detach()))
            if callback is not None:
                callback(epoch, it, losses[-1])
    return torch.clamp(v.detach(), min=vmin, max=vmax), losses

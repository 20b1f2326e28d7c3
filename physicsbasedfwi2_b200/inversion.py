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
                   callback: Optional[Callable[[int, int, float], None]] = None, graph: Optional[bool] = None,
                   **propagator_kwargs):
    """Runs the loop of models/networks.py:15372-15415 and returns (velocity, losses).

    velocity [nz, ny] (CUDA fp32; updated by Adam, returned as a new tensor), observed [nt, S, R] raw
    recorded traces, source_amplitudes [nt, 1, 1] (one wavelet for every shot, as in the reference) or
    [nt, S, Ns], x_s [S, Ns, 2], x_r [S, R, 2].  `grad_weight` multiplies the gradient (the reference's
    depth weight `ss`), `grad_mask` (bool) zeroes it (water layer).  `shuffle` permutes the shots every
    epoch (:5434-5440).  `losses` holds the misfit of every (epoch, batch).

    `graph` (default: on unless a `callback` wants every loss on the host): one iteration - clamp, forward time loop,
    fused misfit + its backward, adjoint time loop, gradient fold, weight / mask, Adam update, loss into a device
    buffer - is captured ONCE per batch index as a CUDA graph (after one eager epoch, which also validates the
    geometry) and replayed for the remaining epochs: no host synchronisation and no Python in the loop, the losses
    are read back once at the end.  The inner time step and the PML profiles are built for `vpmax` (default: the
    clamp bound `vmax`) instead of the model's running maximum, which would need a device-to-host read every
    iteration."""
    if not velocity.is_cuda:
        raise RuntimeError("inner_loop_fwi needs CUDA tensors: there is no CPU path")
    dev = velocity.device
    S = int(x_s.shape[0])
    if num_batches < 1 or num_batches > S:
        raise RuntimeError(f"num_batches must be in [1, {S}]")
    if graph is None:
        graph = callback is None
    if graph and callback is not None:
        raise RuntimeError("a per-iteration callback needs the loss on the host every iteration: use graph=False")
    propagator_kwargs.setdefault("vpmax", float(vmax))
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
    if grad_mask is not None:
        grad_mask = grad_mask.to(dev, torch.bool)
    if grad_weight is not None:
        grad_weight = grad_weight.to(dev, torch.float32)

    v = velocity.detach().clone().requires_grad_(True)
    opt = torch.optim.Adam([v], lr=lr, capturable=bool(graph))
    n_iter = num_epochs * num_batches
    losses_buf = torch.zeros(max(n_iter, 1), dtype=torch.float32, device=dev)
    slot = torch.zeros(1, dtype=torch.long, device=dev)               # where the next loss goes (advanced on the device)
    order = torch.arange(S, device=dev)

    def iteration(it: int):
        idx = order[it::num_batches]
        opt.zero_grad(set_to_none=True)
        model = torch.clamp(v, min=vmin, max=vmax)
        pred = Propagator({"vp": model}, dx, **propagator_kwargs)(src.index_select(1, idx), x_s.index_select(0, idx),
                                                                  x_r.index_select(0, idx), dt)
        loss = trace_misfit(pred, obs_norm.index_select(1, idx),
                            None if direct is None else direct.index_select(1, idx), kind=kind)
        loss.backward()
        if grad_weight is not None:
            v.grad.mul_(grad_weight)
        if grad_mask is not None:
            v.grad.masked_fill_(grad_mask, 0.0)
        opt.step()
        losses_buf.index_copy_(0, slot, loss.detach().reshape(1))
        slot.add_(1)
        return loss

    def new_epoch():
        if shuffle:
            order.copy_(torch.randperm(S, device=dev))                # device RNG: no host sync

    if not graph:
        for epoch in range(num_epochs):
            new_epoch()
            for it in range(num_batches):
                loss = iteration(it)
                if callback is not None:
                    callback(epoch, it, float(loss.detach()))
    else:
        # eager first epoch on a side stream (PyTorch's capture protocol: lazy initialisations of autograd, Adam and
        # cuFFT happen here; out-of-model locations raise here)
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            new_epoch()
            for it in range(num_batches):
                iteration(it)
        torch.cuda.current_stream(dev).wait_stream(side)
        if num_epochs > 1:
            graphs = []
            for it in range(num_batches):
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    iteration(it)
                graphs.append(g)
            for epoch in range(1, num_epochs):
                new_epoch()
                for g in graphs:
                    g.replay()
    losses: List[float] = losses_buf[:n_iter].tolist()                 # the one device-to-host read of the loop
    return torch.clamp(v.detach(), min=vmin, max=vmax), losses

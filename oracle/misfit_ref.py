"""CPU oracle for the fused trace misfit (physicsbasedfwi2_b200/misfit.py).  TEST INFRASTRUCTURE ONLY.

Restates, line for line in plain torch, what the reference's `prop` methods do between the
propagator call and ``lossinner.backward()`` - models/networks.py:5467 (direct-wave subtraction),
:5468 (per-trace max of |.| over time), :5470 (division by max + 1e-10), :5476 (L1Loss; MSELoss
in other nets) - and lets torch.autograd produce d loss / d pred.  This part of the path IS in the
reference (unlike the propagator arithmetic), so this oracle is pinned by the reference lines it
copies the formulae from; it needs nothing but torch."""
import torch


def trace_misfit(pred, observed, direct=None, kind="l1", normalize=True, eps=1e-10):
    d = pred if direct is None else pred - direct
    if normalize:
        m, _ = torch.abs(d).max(dim=0, keepdim=True)
        d = d / (m.abs() + eps)
    crit = torch.nn.L1Loss() if kind == "l1" else torch.nn.MSELoss()
    return crit(d, observed)


def trace_misfit_grad(pred, observed, direct=None, **kw):
    p = pred.detach().clone().requires_grad_(True)
    loss = trace_misfit(p, observed, direct, **kw)
    loss.backward()
    return loss.detach(), p.grad
